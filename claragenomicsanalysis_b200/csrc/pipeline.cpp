// Window pipeline: the stream scheduler / sharding layer above the cudapoa Batch API.
//
// The reference has no such layer; its benchmarks drive several Batch objects from std::async tasks, one stream
// each (cudapoa/benchmarks/multi_batch.hpp:157-170).  This component does the same thing as a library service:
// every worker thread owns one stream and one Batch on one device, pulls fixed-size chunks of windows from a shared
// counter, stages them (add_poa_group), launches (generate_poa) and collects (get_consensus / get_msa) into result
// slots indexed by the global window id -- so the output order is deterministic whatever the interleaving, staging
// of one chunk overlaps the kernel of another, and several devices share one queue (tail balancing for free).
// Only the public C++ interface is used.

#include "cudapoa_b200.h"

#include <claraparabricks/genomeworks/cudapoa/batch.hpp>
#include <claraparabricks/genomeworks/cudapoa/cudapoa.hpp>

#include "engine_stats.hpp"

#include <cuda_runtime_api.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <exception>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

namespace poa = claraparabricks::genomeworks::cudapoa;

struct cudapoa_b200_pipeline
{
    struct Worker
    {
        int32_t device   = 0;
        cudaStream_t stream = nullptr;
        std::unique_ptr<poa::Batch> batch;
        int32_t capacity = 0;
    };
    std::vector<Worker> workers;
    int8_t output_mask = 1;
    // results of the last run, indexed by window
    std::vector<int32_t> status;
    std::vector<std::string> consensus;
    std::vector<std::vector<uint16_t>> coverage;
    std::vector<std::vector<std::string>> msa;
    cudapoa_b200_stats stats{};
    double wall_ms = 0;
    std::string error;
};

namespace
{
thread_local std::string g_pipeline_error;

inline void fnv(uint64_t& h, const void* data, size_t n)
{
    const unsigned char* p = static_cast<const unsigned char*>(data);
    for (size_t i = 0; i < n; i++)
    {
        h ^= p[i];
        h *= 1099511628211ull;
    }
}
} // namespace

extern "C" {

const char* cudapoa_b200_pipeline_last_error(void)
{
    return g_pipeline_error.c_str();
}

cudapoa_b200_pipeline* cudapoa_b200_pipeline_create(const int32_t* devices, int32_t n_devices,
                                                    int32_t workers_per_device, size_t mem_per_worker,
                                                    int8_t output_mask, const cudapoa_b200_batch_size* bs,
                                                    int16_t gap_score, int16_t mismatch_score, int16_t match_score,
                                                    int32_t cuda_banded_alignment)
{
    try
    {
        if (n_devices <= 0 || workers_per_device <= 0)
            throw std::invalid_argument("pipeline needs at least one device and one worker");
        std::unique_ptr<cudapoa_b200_pipeline> p(new cudapoa_b200_pipeline());
        p->output_mask = output_mask;
        poa::BatchSize size(bs->max_sequence_size, bs->max_sequences_per_poa, bs->alignment_band_width);
        size.max_sequence_size                 = bs->max_sequence_size;
        size.max_consensus_size                = bs->max_consensus_size;
        size.max_nodes_per_window              = bs->max_nodes_per_window;
        size.max_nodes_per_window_banded       = bs->max_nodes_per_window_banded;
        size.max_matrix_graph_dimension        = bs->max_matrix_graph_dimension;
        size.max_matrix_graph_dimension_banded = bs->max_matrix_graph_dimension_banded;
        size.max_matrix_sequence_dimension     = bs->max_matrix_sequence_dimension;
        size.alignment_band_width              = bs->alignment_band_width;
        size.max_sequences_per_poa             = bs->max_sequences_per_poa;
        for (int32_t d = 0; d < n_devices; d++)
        {
            for (int32_t k = 0; k < workers_per_device; k++)
            {
                cudapoa_b200_pipeline::Worker w;
                w.device = devices[d];
                if (cudaSetDevice(w.device) != cudaSuccess ||
                    cudaStreamCreateWithFlags(&w.stream, cudaStreamNonBlocking) != cudaSuccess)
                    throw std::runtime_error("pipeline: cannot create a stream on device " + std::to_string(w.device));
                w.batch = poa::create_batch(w.device, w.stream, mem_per_worker, output_mask, size, gap_score,
                                            mismatch_score, match_score,
                                            static_cast<poa::BandMode>(std::min(std::max(cuda_banded_alignment, 0), 2)));
                auto* src  = dynamic_cast<poa::EngineStatsSource*>(w.batch.get());
                w.capacity = src != nullptr ? src->max_poas() : 0;
                if (src != nullptr)
                    src->set_overlapped(workers_per_device > 1); // several batches share the device
                p->workers.push_back(std::move(w));
            }
        }
        return p.release();
    }
    catch (const std::exception& e)
    {
        g_pipeline_error = e.what();
        return nullptr;
    }
}

void cudapoa_b200_pipeline_destroy(cudapoa_b200_pipeline* p)
{
    if (p == nullptr)
        return;
    for (auto& w : p->workers)
    {
        w.batch.reset();
        if (w.stream != nullptr)
        {
            cudaSetDevice(w.device);
            cudaStreamDestroy(w.stream);
        }
    }
    delete p;
}

// Process n_windows windows given as flat host buffers (layout of cudapoa_b200_add_poa_groups_flat).  Blocking.
// chunk_windows: windows per batch round (0: a quarter of the smallest batch capacity, at least 256).
// Returns 0, or a negative value with cudapoa_b200_pipeline_last_error() set.
int32_t cudapoa_b200_pipeline_run(cudapoa_b200_pipeline* p, const uint8_t* bases, const int8_t* weights,
                                  const int32_t* read_lengths, const int32_t* window_n_reads, int32_t n_windows,
                                  int32_t chunk_windows)
{
    try
    {
        const bool want_msa = (p->output_mask & poa::OutputType::msa) != 0;
        p->status.assign(static_cast<size_t>(n_windows), static_cast<int32_t>(poa::StatusType::generic_error));
        p->consensus.assign(want_msa ? 0 : static_cast<size_t>(n_windows), std::string());
        p->coverage.assign(want_msa ? 0 : static_cast<size_t>(n_windows), std::vector<uint16_t>());
        p->msa.assign(want_msa ? static_cast<size_t>(n_windows) : 0, std::vector<std::string>());
        p->stats = cudapoa_b200_stats{};

        // prefix offsets: first read and first base of every window
        std::vector<int64_t> read_off(static_cast<size_t>(n_windows) + 1, 0), base_off(static_cast<size_t>(n_windows) + 1, 0);
        {
            int64_t r = 0, b = 0;
            for (int32_t w = 0; w < n_windows; w++)
            {
                read_off[w] = r;
                base_off[w] = b;
                for (int32_t k = 0; k < window_n_reads[w]; k++)
                    b += read_lengths[r + k];
                r += window_n_reads[w];
            }
            read_off[n_windows] = r;
            base_off[n_windows] = b;
        }
        int32_t min_cap = INT32_MAX;
        for (const auto& w : p->workers)
            min_cap = std::min(min_cap, w.capacity);
        if (min_cap <= 0)
            throw std::runtime_error("pipeline: a batch has no capacity");
        int32_t chunk = chunk_windows > 0 ? chunk_windows : std::max(256, min_cap / 4);
        chunk         = std::min(chunk, min_cap);

        std::atomic<int32_t> next(0);
        std::mutex mu;
        std::string first_error;
        const auto t0 = std::chrono::steady_clock::now();
        std::vector<std::thread> threads;
        for (size_t wi = 0; wi < p->workers.size(); wi++)
        {
            threads.emplace_back([&, wi]() {
                auto& wk = p->workers[wi];
                cudapoa_b200_stats local{};
                try
                {
                    cudaSetDevice(wk.device);
                    poa::Group group;
                    std::vector<poa::StatusType> seq_status, out_status;
                    std::vector<std::string> cons;
                    std::vector<std::vector<uint16_t>> cov;
                    std::vector<std::vector<std::string>> msa;
                    for (;;)
                    {
                        int32_t a = next.fetch_add(chunk);
                        if (a >= n_windows)
                            break;
                        const int32_t b = std::min(n_windows, a + chunk);
                        while (a < b)
                        {
                            wk.batch->reset();
                            int32_t added = 0;
                            for (int32_t w = a; w < b; w++)
                            {
                                group.clear();
                                int64_t off = base_off[w];
                                for (int32_t k = 0; k < window_n_reads[w]; k++)
                                {
                                    poa::Entry e{};
                                    e.seq     = reinterpret_cast<const char*>(bases + off);
                                    e.weights = weights != nullptr ? weights + off : nullptr;
                                    e.length  = read_lengths[read_off[w] + k];
                                    off += e.length;
                                    group.push_back(e);
                                }
                                if (wk.batch->add_poa_group(seq_status, group) != poa::StatusType::success)
                                    break; // batch full: launch what it holds, continue with the rest
                                added++;
                            }
                            if (added == 0)
                                throw std::runtime_error("pipeline: window " + std::to_string(a) + " does not fit a batch");
                            wk.batch->generate_poa();
                            out_status.clear();
                            if (want_msa)
                            {
                                msa.clear();
                                wk.batch->get_msa(msa, out_status);
                                for (int32_t k = 0; k < added; k++)
                                    p->msa[a + k] = std::move(msa[k]);
                            }
                            else
                            {
                                cons.clear();
                                cov.clear();
                                wk.batch->get_consensus(cons, cov, out_status);
                                for (int32_t k = 0; k < added; k++)
                                {
                                    p->consensus[a + k] = std::move(cons[k]);
                                    p->coverage[a + k]  = std::move(cov[k]);
                                }
                            }
                            for (int32_t k = 0; k < added; k++)
                                p->status[a + k] = static_cast<int32_t>(out_status[k]);
                            if (auto* src = dynamic_cast<poa::EngineStatsSource*>(wk.batch.get()))
                            {
                                cudapoa_b200_stats s{};
                                src->fill_stats(&s);
                                local.windows += s.windows;
                                local.reads_aligned += s.reads_aligned;
                                local.dp_cells += s.dp_cells;
                                local.graph_rows += s.graph_rows;
                                local.exact_path += s.exact_path;
                                local.widened += s.widened;
                                local.kernel_launches += s.kernel_launches;
                                local.h2d_bytes += s.h2d_bytes;
                                local.d2h_bytes += s.d2h_bytes;
                                local.kernel_ms += s.kernel_ms;
                                local.graph_kernel_ms += s.graph_kernel_ms;
                            }
                            a += added;
                        }
                    }
                }
                catch (const std::exception& e)
                {
                    std::lock_guard<std::mutex> lock(mu);
                    if (first_error.empty())
                        first_error = e.what();
                    next.store(n_windows); // stop the other workers
                }
                std::lock_guard<std::mutex> lock(mu);
                p->stats.windows += local.windows;
                p->stats.reads_aligned += local.reads_aligned;
                p->stats.dp_cells += local.dp_cells;
                p->stats.graph_rows += local.graph_rows;
                p->stats.exact_path += local.exact_path;
                p->stats.widened += local.widened;
                p->stats.kernel_launches += local.kernel_launches;
                p->stats.h2d_bytes += local.h2d_bytes;
                p->stats.d2h_bytes += local.d2h_bytes;
                p->stats.kernel_ms += local.kernel_ms;
                p->stats.graph_kernel_ms += local.graph_kernel_ms;
            });
        }
        for (auto& t : threads)
            t.join();
        p->wall_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        if (!first_error.empty())
            throw std::runtime_error(first_error);
        return 0;
    }
    catch (const std::exception& e)
    {
        g_pipeline_error = e.what();
        return -2;
    }
}

int32_t cudapoa_b200_pipeline_workers(const cudapoa_b200_pipeline* p)
{
    return static_cast<int32_t>(p->workers.size());
}

int32_t cudapoa_b200_pipeline_capacity(const cudapoa_b200_pipeline* p)
{
    int32_t c = INT32_MAX;
    for (const auto& w : p->workers)
        c = std::min(c, w.capacity);
    return p->workers.empty() ? 0 : c;
}

double cudapoa_b200_pipeline_wall_ms(const cudapoa_b200_pipeline* p)
{
    return p->wall_ms;
}

int32_t cudapoa_b200_pipeline_stats(const cudapoa_b200_pipeline* p, cudapoa_b200_stats* out)
{
    *out = p->stats;
    return 0;
}

int32_t cudapoa_b200_pipeline_status(const cudapoa_b200_pipeline* p, int32_t i)
{
    if (i < 0 || i >= static_cast<int32_t>(p->status.size()))
        return -1;
    return p->status[i];
}

int32_t cudapoa_b200_pipeline_consensus(const cudapoa_b200_pipeline* p, int32_t i, const char** seq,
                                        const uint16_t** coverage)
{
    if (i < 0 || i >= static_cast<int32_t>(p->consensus.size()))
        return -1;
    if (seq != nullptr)
        *seq = p->consensus[i].c_str();
    if (coverage != nullptr)
        *coverage = p->coverage[i].data();
    return static_cast<int32_t>(p->consensus[i].size());
}

int32_t cudapoa_b200_pipeline_msa_rows(const cudapoa_b200_pipeline* p, int32_t i)
{
    if (i < 0 || i >= static_cast<int32_t>(p->msa.size()))
        return -1;
    return static_cast<int32_t>(p->msa[i].size());
}

int32_t cudapoa_b200_pipeline_msa_row(const cudapoa_b200_pipeline* p, int32_t i, int32_t row, const char** seq)
{
    if (i < 0 || i >= static_cast<int32_t>(p->msa.size()) || row < 0 || row >= static_cast<int32_t>(p->msa[i].size()))
        return -1;
    if (seq != nullptr)
        *seq = p->msa[i][row].c_str();
    return static_cast<int32_t>(p->msa[i][row].size());
}

int64_t cudapoa_b200_pipeline_total_chars(const cudapoa_b200_pipeline* p)
{
    int64_t total = 0;
    for (const auto& s : p->consensus)
        total += static_cast<int64_t>(s.size());
    for (const auto& rows : p->msa)
        for (const auto& s : rows)
            total += static_cast<int64_t>(s.size());
    return total;
}

// FNV-1a over the windows [first, first + count) in order: the same bytes cudapoa_b200_result_digest hashes for a
// batch holding exactly those windows (statuses first, then per window length | bytes | coverage).
uint64_t cudapoa_b200_pipeline_digest(const cudapoa_b200_pipeline* p, int32_t first, int32_t count)
{
    uint64_t h        = 14695981039346656037ull;
    const int32_t end = std::min<int32_t>(first + count, static_cast<int32_t>(p->status.size()));
    for (int32_t i = first; i < end; i++)
    {
        int32_t st = p->status[i];
        fnv(h, &st, sizeof(st));
    }
    if (!p->consensus.empty())
    {
        for (int32_t i = first; i < end; i++)
        {
            int32_t len = static_cast<int32_t>(p->consensus[i].size());
            fnv(h, &len, sizeof(len));
            fnv(h, p->consensus[i].data(), p->consensus[i].size());
            fnv(h, p->coverage[i].data(), p->coverage[i].size() * sizeof(uint16_t));
        }
    }
    else
    {
        for (int32_t i = first; i < end; i++)
        {
            int32_t n = static_cast<int32_t>(p->msa[i].size());
            fnv(h, &n, sizeof(n));
            for (const auto& s : p->msa[i])
            {
                int32_t len = static_cast<int32_t>(s.size());
                fnv(h, &len, sizeof(len));
                fnv(h, s.data(), s.size());
            }
        }
    }
    return h;
}

} // extern "C"
