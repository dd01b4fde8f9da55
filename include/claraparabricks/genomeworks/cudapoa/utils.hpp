// B200 cudapoa engine -- batch-size planning and window file readers.
// API-compatible with GenomeWorks cudapoa/include/claraparabricks/genomeworks/cudapoa/utils.hpp:48-176
// (get_multi_batch_sizes, resize_windows, parse_cudapoa_file, parse_fasta_files, parse_golden_value_file).
// parse_fasta_files reads plain-text FASTA/FASTQ with a small built-in reader instead of kseqpp.
#pragma once

#include <claraparabricks/genomeworks/cudapoa/batch.hpp>
#include <claraparabricks/genomeworks/utils/signed_integer_utils.hpp>

#include <cassert>
#include <fstream>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace claraparabricks
{
namespace genomeworks
{
namespace cudapoa
{

/// Plan a small set of BatchSize values covering all groups: groups are binned by how many of them fit in one
/// batch (bins 1, 2, 4, ... by default), and bins of smaller groups are merged into a larger-group batch when
/// they fit.  list_of_groups_per_batch[i] holds the indices of the groups to run with list_of_batch_sizes[i].
void get_multi_batch_sizes(std::vector<BatchSize>& list_of_batch_sizes,
                           std::vector<std::vector<int32_t>>& list_of_groups_per_batch,
                           const std::vector<Group>& poa_groups, bool banded_alignment = true, bool msa_flag = false,
                           int32_t band_width = 256, std::vector<int32_t>* bins_capacity = nullptr,
                           float gpu_memory_usage_quota = 0.9, int32_t mismatch_score = -6, int32_t gap_score = -8,
                           int32_t match_score = 8);

/// Truncate, or cyclically extend, `windows` to total_windows entries (no-op when total_windows < 0).
inline void resize_windows(std::vector<std::vector<std::string>>& windows, const int32_t total_windows)
{
    if (total_windows < 0)
    {
        return;
    }
    const int32_t available = get_size<int32_t>(windows);
    if (available > total_windows)
    {
        windows.resize(static_cast<size_t>(total_windows));
    }
    else if (available > 0)
    {
        windows.reserve(static_cast<size_t>(total_windows));
        for (int32_t i = available; i < total_windows; i++)
        {
            windows.push_back(windows[static_cast<size_t>(i - available)]);
        }
    }
    assert(get_size<int32_t>(windows) == total_windows || available == 0);
}

/// Read the cudapoa window text format: a line with the number of sequences, then that many sequence lines,
/// repeated per window.
inline void parse_cudapoa_file(std::vector<std::vector<std::string>>& windows, const std::string& filename,
                               int32_t total_windows)
{
    std::ifstream infile(filename);
    if (!infile.good())
    {
        throw std::runtime_error("Cannot read file " + filename);
    }
    std::string line;
    int32_t remaining = 0;
    while (std::getline(infile, line))
    {
        if (remaining == 0)
        {
            std::istringstream iss(line);
            iss >> remaining;
            windows.emplace_back();
        }
        else
        {
            windows.back().push_back(line);
            remaining--;
        }
    }
    resize_windows(windows, total_windows);
}

/// Read FASTA or FASTQ files, one window per file (all its sequences, in file order).
inline void parse_fasta_files(std::vector<std::vector<std::string>>& windows,
                              const std::vector<std::string>& input_paths, const int32_t total_windows)
{
    windows.resize(input_paths.size());
    for (size_t f = 0; f < input_paths.size(); f++)
    {
        std::ifstream infile(input_paths[f]);
        if (!infile.good())
        {
            throw std::runtime_error("Cannot read file " + input_paths[f]);
        }
        std::string line;
        bool fastq         = false;
        bool in_quality    = false;
        size_t quality_len = 0;
        while (std::getline(infile, line))
        {
            if (!line.empty() && line.back() == '\r')
            {
                line.pop_back();
            }
            if (line.empty())
            {
                continue;
            }
            if (in_quality)
            {
                quality_len += line.size();
                if (quality_len >= windows[f].back().size())
                {
                    in_quality = false;
                }
                continue;
            }
            if (line[0] == '>' || (line[0] == '@' && (windows[f].empty() || fastq)))
            {
                fastq = (line[0] == '@');
                windows[f].emplace_back();
            }
            else if (fastq && line[0] == '+')
            {
                in_quality  = true;
                quality_len = 0;
            }
            else if (!windows[f].empty())
            {
                windows[f].back() += line;
            }
        }
    }
    resize_windows(windows, total_windows);
}

/// First line of a text file (the expected genome of the end-to-end test).
inline std::string parse_golden_value_file(const std::string& filename)
{
    std::ifstream infile(filename);
    if (!infile.good())
    {
        throw std::runtime_error("Cannot read file " + filename);
    }
    std::string line;
    std::getline(infile, line);
    return line;
}

} // namespace cudapoa
} // namespace genomeworks
} // namespace claraparabricks
