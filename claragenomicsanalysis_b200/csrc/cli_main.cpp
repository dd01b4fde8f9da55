// cudapoa command line tool over the B200 engine.
//
// Same options, input formats, stdout format and batching flow as the reference's tool (cudapoa/src/main.cpp:129-293,
// application_parameters.cpp:29-121, utils.hpp:107-157): windows come from one cudapoa-format file or from one or
// more FASTA files (one window per file), get_multi_batch_sizes plans the batches, every batch is filled until
// exceeded_maximum_poas, processed, printed (one consensus per line, or the MSA rows of each group), optionally
// dumped as DOT graphs, and reset.  Written against the public C++ interface only, so it is also a usage example.
//
// Extensions: -A/--adaptive-band (BandMode::adaptive_band of this engine), -D/--device.

#include <claraparabricks/genomeworks/cudapoa/batch.hpp>
#include <claraparabricks/genomeworks/cudapoa/cudapoa.hpp>
#include <claraparabricks/genomeworks/cudapoa/utils.hpp>

#include <cuda_runtime_api.h>
#include <getopt.h>

#include <cstdint>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace poa = claraparabricks::genomeworks::cudapoa;
namespace gw  = claraparabricks::genomeworks;

namespace
{

struct Parameters
{
    std::vector<std::string> input_paths;
    std::string graph_output_path;
    bool all_fasta            = true;
    bool msa                  = false;
    bool banded               = true;
    bool adaptive             = false;
    int32_t band_width        = 256;
    int32_t max_groups        = -1;
    int32_t device            = 0;
    double gpu_mem_allocation = 0.9;
    int32_t match_score       = 8;
    int32_t mismatch_score    = -6;
    int32_t gap_score         = -8;
};

[[noreturn]] void usage(int code)
{
    std::cerr << "Usage: cudapoa [options ...]\n"
                 "     options:\n"
                 "        -i, --input <file>\n"
                 "            input in fasta/cudapoa format, can be used multiple times for multiple fasta files, but "
                 "supports only one cudapoa file\n"
                 "        -a, --msa\n"
                 "            generates msa if this flag is passed [default: consensus]\n"
                 "        -f, --full-alignment\n"
                 "            uses full alignment if this flag is passed [banded alignment]\n"
                 "        -b, --band-width <int>\n"
                 "            band-width for banded alignment (must be multiple of 128) [256]\n"
                 "        -A, --adaptive-band\n"
                 "            banded alignment whose band doubles (up to 4x) where the static band is too narrow "
                 "[static band]\n"
                 "        -d, --dot <file>\n"
                 "            output path for printing graph in DOT format [disabled]\n"
                 "        -M, --max-groups  <int>\n"
                 "            maximum number of POA groups to create from file (-1 for all, > 0 for limited) [-1]\n"
                 "            repeats groups if less groups are present than specified\n"
                 "        -R, --gpu-mem-alloc <double>\n"
                 "            fraction of available GPU memory to be used for cudapoa [0.9]\n"
                 "        -D, --device <int>\n"
                 "            CUDA device to run on [0]\n"
                 "        -m, --match  <int>\n"
                 "            score for matching bases (must be positive) [8]\n"
                 "        -n, --mismatch  <int>\n"
                 "            score for mismatching bases (must be non-positive) [-6]\n"
                 "        -g, --gap  <int>\n"
                 "            score for gaps (must be non-positive) [-8]\n"
                 "        -v, --version\n"
                 "            version information\n"
                 "        -h, --help\n"
                 "            prints usage"
              << std::endl;
    std::exit(code);
}

Parameters parse(int argc, char* argv[])
{
    Parameters p;
    static const option options[] = {
        {"input", required_argument, nullptr, 'i'},      {"msa", no_argument, nullptr, 'a'},
        {"full-alignment", no_argument, nullptr, 'f'},   {"band-width", required_argument, nullptr, 'b'},
        {"adaptive-band", no_argument, nullptr, 'A'},    {"dot", required_argument, nullptr, 'd'},
        {"max-groups", required_argument, nullptr, 'M'}, {"gpu-mem-alloc", required_argument, nullptr, 'R'},
        {"device", required_argument, nullptr, 'D'},     {"match", required_argument, nullptr, 'm'},
        {"mismatch", required_argument, nullptr, 'n'},   {"gap", required_argument, nullptr, 'g'},
        {"version", no_argument, nullptr, 'v'},          {"help", no_argument, nullptr, 'h'},
        {nullptr, 0, nullptr, 0},
    };
    int c = 0;
    while ((c = getopt_long(argc, argv, "i:afb:Ad:M:R:D:m:n:g:vh", options, nullptr)) != -1)
    {
        switch (c)
        {
        case 'i': p.input_paths.emplace_back(optarg); break;
        case 'a': p.msa = true; break;
        case 'f': p.banded = false; break;
        case 'b': p.band_width = std::stoi(optarg); break;
        case 'A': p.adaptive = true; break;
        case 'd': p.graph_output_path = optarg; break;
        case 'M': p.max_groups = std::stoi(optarg); break;
        case 'R': p.gpu_mem_allocation = std::stod(optarg); break;
        case 'D': p.device = std::stoi(optarg); break;
        case 'm': p.match_score = std::stoi(optarg); break;
        case 'n': p.mismatch_score = std::stoi(optarg); break;
        case 'g': p.gap_score = std::stoi(optarg); break;
        case 'v': std::cerr << "cudapoa (B200 engine), interface of GenomeWorks 0.5.0" << std::endl; std::exit(1);
        case 'h': usage(0);
        default: std::exit(1);
        }
    }
    // application_parameters.cpp:92-120
    if (p.gpu_mem_allocation <= 0 || p.gpu_mem_allocation > 1.0)
        throw std::runtime_error("gpu-mem-alloc should be greater than 0 and less than or equal to 1.0");
    if (p.banded && p.band_width < 1)
        throw std::runtime_error("band-width must be positive");
    if (p.match_score < 0)
        throw std::runtime_error("match score must be positive");
    if (p.max_groups == 0)
        throw std::runtime_error("max-groups cannot be 0");
    if (p.mismatch_score > 0)
        throw std::runtime_error("mismatch score must be non-positive");
    if (p.gap_score > 0)
        throw std::runtime_error("gap score must be non-positive");
    if (p.adaptive && !p.banded)
        throw std::runtime_error("adaptive-band and full-alignment exclude each other");
    // application_parameters.cpp:125-150: all FASTA, or exactly one cudapoa file
    for (const std::string& path : p.input_paths)
    {
        std::ifstream in(path);
        if (!in.good())
            throw std::runtime_error("Invalid input file: " + path);
        std::string first;
        std::getline(in, first);
        if (first.empty() || first[0] != '>')
            p.all_fasta = false;
    }
    if (p.input_paths.empty() || (!p.all_fasta && p.input_paths.size() > 1))
    {
        std::cerr << "Invalid input. cudapoa needs input in either one cudapoa format file or in one/multiple fasta "
                     "files."
                  << std::endl;
        usage(1);
    }
    return p;
}

void process_batch(poa::Batch& batch, bool msa)
{
    batch.generate_poa();
    std::vector<poa::StatusType> output_status;
    if (msa)
    {
        std::vector<std::vector<std::string>> rows;
        const poa::StatusType st = batch.get_msa(rows, output_status);
        if (st != poa::StatusType::success)
            std::cerr << "Could not generate MSA for batch : " << st << std::endl;
        for (size_t g = 0; g < rows.size(); g++)
        {
            if (output_status[g] != poa::StatusType::success)
                std::cerr << "Error generating  MSA for POA group " << g << ". Error type " << output_status[g]
                          << std::endl;
            else
                for (const std::string& row : rows[g])
                    std::cout << row << '\n';
        }
    }
    else
    {
        std::vector<std::string> consensus;
        std::vector<std::vector<uint16_t>> coverage;
        const poa::StatusType st = batch.get_consensus(consensus, coverage, output_status);
        if (st != poa::StatusType::success)
            std::cerr << "Could not generate consensus for batch : " << st << std::endl;
        for (size_t g = 0; g < consensus.size(); g++)
        {
            if (output_status[g] != poa::StatusType::success)
                std::cerr << "Error generating consensus for POA group " << g << ". Error type " << output_status[g]
                          << std::endl;
            else
                std::cout << consensus[g] << '\n';
        }
    }
    std::cout.flush();
}

int run(const Parameters& p)
{
    std::vector<std::vector<std::string>> windows;
    if (p.all_fasta)
        poa::parse_fasta_files(windows, p.input_paths, p.max_groups);
    else
        poa::parse_cudapoa_file(windows, p.input_paths[0], p.max_groups);

    std::ofstream graph_output;
    if (!p.graph_output_path.empty())
    {
        graph_output.open(p.graph_output_path);
        if (!graph_output)
        {
            std::cerr << "Error opening " << p.graph_output_path << " for graph output" << std::endl;
            return -1;
        }
    }

    std::vector<poa::Group> groups(windows.size());
    for (size_t i = 0; i < windows.size(); i++)
        for (const std::string& seq : windows[i])
            groups[i].push_back(poa::Entry{seq.c_str(), nullptr, static_cast<int32_t>(seq.length())});

    if (cudaSetDevice(p.device) != cudaSuccess)
        throw std::runtime_error("cannot select CUDA device " + std::to_string(p.device));
    poa::Init();

    std::vector<poa::BatchSize> batch_sizes;
    std::vector<std::vector<int32_t>> groups_per_batch;
    poa::get_multi_batch_sizes(batch_sizes, groups_per_batch, groups, p.banded, p.msa, p.band_width, nullptr,
                               static_cast<float>(p.gpu_mem_allocation), p.mismatch_score, p.gap_score, p.match_score);

    const poa::BandMode mode = !p.banded ? poa::BandMode::full_band
                                         : (p.adaptive ? poa::BandMode::adaptive_band : poa::BandMode::static_band);
    int32_t offset = 0;
    for (size_t b = 0; b < batch_sizes.size(); b++)
    {
        size_t free_mem = 0, total_mem = 0;
        cudaMemGetInfo(&free_mem, &total_mem);
        std::unique_ptr<poa::Batch> batch =
            poa::create_batch(p.device, nullptr, static_cast<size_t>(p.gpu_mem_allocation * static_cast<double>(free_mem)),
                              p.msa ? poa::OutputType::msa : poa::OutputType::consensus, batch_sizes[b],
                              static_cast<int16_t>(p.gap_score), static_cast<int16_t>(p.mismatch_score),
                              static_cast<int16_t>(p.match_score), mode);
        const std::vector<int32_t>& ids = groups_per_batch[b];
        const int32_t n_ids            = static_cast<int32_t>(ids.size());
        int32_t first                  = 0;
        for (int32_t i = 0; i < n_ids;)
        {
            std::vector<poa::StatusType> seq_status;
            const poa::StatusType status = batch->add_poa_group(seq_status, groups[static_cast<size_t>(ids[i])]);
            if (status == poa::StatusType::exceeded_maximum_poas || i == n_ids - 1)
            {
                if (batch->get_total_poas() > 0)
                {
                    process_batch(*batch, p.msa);
                    if (graph_output.is_open())
                    {
                        if (!graph_output.good())
                            throw std::runtime_error("Error writing dot file");
                        std::vector<gw::DirectedGraph> graphs;
                        std::vector<poa::StatusType> graph_status;
                        batch->get_graphs(graphs, graph_status);
                        for (const gw::DirectedGraph& g : graphs)
                            graph_output << g.serialize_to_dot() << std::endl;
                    }
                    batch->reset();
                    const int32_t last = (status == poa::StatusType::success) ? i : i - 1;
                    std::cerr << "Processed groups " << first + offset << " - " << last + offset << " (batch " << b << ")"
                              << std::endl;
                }
                else
                {
                    std::cerr << "Could not add POA group " << ids[i] << " to batch " << b << std::endl;
                    i++;
                }
                first = i;
            }
            if (status == poa::StatusType::success)
            {
                for (const poa::StatusType s : seq_status)
                    if (s == poa::StatusType::exceeded_maximum_sequence_size)
                        std::cerr << "Dropping sequence because sequence exceeded maximum size" << std::endl;
                i++;
            }
            else if (status != poa::StatusType::exceeded_maximum_poas)
            {
                std::cerr << "Could not add POA group " << ids[i] << " to batch " << b << ". Error code " << status
                          << std::endl;
                i++;
            }
        }
        offset += n_ids;
    }
    return 0;
}

} // namespace

int main(int argc, char* argv[])
{
    try
    {
        return run(parse(argc, argv));
    }
    catch (const std::exception& e)
    {
        std::cerr << "cudapoa: " << e.what() << std::endl;
        return 1;
    }
}
