// Window pipeline: the stream scheduler / sharding layer above the cudapoa Batch API.
//
// The reference has no such layer; its benchmarks drive several Batch objects from std::async tasks, one stream
// each (cudapoa/benchmarks/multi_batch.hpp:157-170).  This component does the same thing as a library service:
// a pool of worker threads that lives as long as the pipeline; every worker owns one stream and one Batch on one
// device, pulls chunks of windows from a shared schedule (most expensive windows first), stages them (add_poa_group), launches (generate_poa) and collects (get_consensus /
// get_msa) into result slots indexed by the global window id -- so the output order is deterministic whatever the
// interleaving, staging and collecting of one chunk overlap the kernels of the others (the pool of batches is the
// multi-buffering), and several devices share one queue (tail balancing for free).
// Only the public C++ interface is used.

#include "cudapoa_b200.h"

#include <claraparabricks/genomeworks/cudapoa/batch.hpp>
#include <claraparabricks/genomeworks/cudapoa/cudapoa.hpp>

#include "engine_stats.hpp"
#include "nvtx_ranges.hpp"

#include <cuda_runtime_api.h>

#include <algorithm>
#include <array>
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <chrono>
#include <condition_variable>
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
        std::thread thread; // lives as long as the pipeline; sleeps between runs
    };
    // One run's inputs and schedule, shared by the workers.
    struct Job
    {
        const uint8_t* bases          = nullptr;
        const int8_t* weights         = nullptr;
        const int32_t* read_lengths   = nullptr;
        const int32_t* window_n_reads = nullptr;
        std::vector<int64_t> read_off, base_off; // first read / first base of every window
        std::vector<int32_t> order;              // window ids, most expensive first
        std::vector<int32_t> chunk_end;          // chunk c covers order[chunk_end[c - 1] .. chunk_end[c])
        std::atomic<int32_t> next_chunk{0};
        bool want_msa = false;
        std::string first_error;
        // CUDAPOA_B200_PIPELINE_TRACE=1: per chunk {worker, windows, stage start, launch start, collect start, end} in ms
        // since the run began, printed to stderr (development: where a run's wall time goes, tools/pipeline_trace.py)
        bool trace = false;
        std::chrono::steady_clock::time_point t0;
        std::vector<std::array<double, 6>> trace_rows;
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
    // worker pool
    std::mutex mu;
    std::condition_variable wake, done;
    Job* job           = nullptr;
    uint64_t generation = 0; // bumped by run(): a worker takes part in every generation once
    int32_t busy        = 0; // workers still inside the current job
    bool quit           = false;
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

void add_stats(cudapoa_b200_stats& to, const cudapoa_b200_stats& s)
{
    to.windows += s.windows;
    to.reads_aligned += s.reads_aligned;
    to.dp_cells += s.dp_cells;
    to.graph_rows += s.graph_rows;
    to.exact_path += s.exact_path;
    to.widened += s.widened;
    to.kernel_launches += s.kernel_launches;
    to.h2d_bytes += s.h2d_bytes;
    to.d2h_bytes += s.d2h_bytes;
    to.kernel_ms += s.kernel_ms;
    to.graph_kernel_ms += s.graph_kernel_ms;
}

// One worker's share of a job: pull chunks until none is left.  Stage (add_poa_group), launch (generate_poa), collect
// (get_consensus / get_msa) -- while this worker stages or collects, the other workers' kernels keep the device busy:
// the pool of batches is the multi-buffering.
void work(cudapoa_b200_pipeline* p, cudapoa_b200_pipeline::Worker& wk, cudapoa_b200_pipeline::Job& job)
{
    cudapoa_b200_stats local{};
    try
    {
        cudaSetDevice(wk.device);
        poa::Group group;
        std::vector<poa::StatusType> seq_status, out_status;
        std::vector<std::string> cons;
        std::vector<std::vector<uint16_t>> cov;
        std::vector<std::vector<std::string>> msa;
        const int32_t n_chunks = static_cast<int32_t>(job.chunk_end.size());
        for (;;)
        {
            const int32_t c = job.next_chunk.fetch_add(1);
            if (c >= n_chunks)
                break;
            int32_t a       = c == 0 ? 0 : job.chunk_end[c - 1];
            const int32_t b = job.chunk_end[c];
            while (a < b)
            {
                int32_t added = 0;
                auto now_ms   = [&] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - job.t0).count(); };
                const double t_stage = job.trace ? now_ms() : 0.0;
                {
                    NvtxRange range("pipeline: stage");
                    wk.batch->reset();
                    for (int32_t k = a; k < b; k++)
                    {
                        const int32_t w = job.order[k];
                        group.clear();
                        int64_t off = job.base_off[w];
                        for (int32_t r = 0; r < job.window_n_reads[w]; r++)
                        {
                            poa::Entry e{};
                            e.seq     = reinterpret_cast<const char*>(job.bases + off);
                            e.weights = job.weights != nullptr ? job.weights + off : nullptr;
                            e.length  = job.read_lengths[job.read_off[w] + r];
                            off += e.length;
                            group.push_back(e);
                        }
                        if (wk.batch->add_poa_group(seq_status, group) != poa::StatusType::success)
                            break; // batch full: launch what it holds, continue with the rest
                        added++;
                    }
                }
                if (added == 0)
                    throw std::runtime_error("pipeline: window " + std::to_string(job.order[a]) + " does not fit a batch");
                const double t_launch = job.trace ? now_ms() : 0.0;
                {
                    NvtxRange range("pipeline: launch");
                    wk.batch->generate_poa();
                }
                const double t_collect = job.trace ? now_ms() : 0.0;
                NvtxRange range("pipeline: collect");
                out_status.clear();
                if (job.want_msa)
                {
                    msa.clear();
                    wk.batch->get_msa(msa, out_status);
                    for (int32_t k = 0; k < added; k++)
                        p->msa[job.order[a + k]] = std::move(msa[k]);
                }
                else
                {
                    cons.clear();
                    cov.clear();
                    wk.batch->get_consensus(cons, cov, out_status);
                    for (int32_t k = 0; k < added; k++)
                    {
                        p->consensus[job.order[a + k]] = std::move(cons[k]);
                        p->coverage[job.order[a + k]]  = std::move(cov[k]);
                    }
                }
                for (int32_t k = 0; k < added; k++)
                    p->status[job.order[a + k]] = static_cast<int32_t>(out_status[k]);
                if (auto* src = dynamic_cast<poa::EngineStatsSource*>(wk.batch.get()))
                {
                    cudapoa_b200_stats s{};
                    src->fill_stats(&s);
                    add_stats(local, s);
                }
                if (job.trace)
                {
                    std::lock_guard<std::mutex> lock(p->mu);
                    job.trace_rows.push_back({static_cast<double>(&wk - p->workers.data()), static_cast<double>(added), t_stage,
                                              t_launch, t_collect, now_ms()});
                }
                a += added;
            }
        }
    }
    catch (const std::exception& e)
    {
        std::lock_guard<std::mutex> lock(p->mu);
        if (job.first_error.empty())
            job.first_error = e.what();
        job.next_chunk.store(static_cast<int32_t>(job.chunk_end.size())); // stop the other workers
    }
    std::lock_guard<std::mutex> lock(p->mu);
    add_stats(p->stats, local);
}

void worker_main(cudapoa_b200_pipeline* p, size_t wi)
{
    uint64_t seen = 0;
    for (;;)
    {
        cudapoa_b200_pipeline::Job* job = nullptr;
        {
            std::unique_lock<std::mutex> lock(p->mu);
            p->wake.wait(lock, [&] { return p->quit || p->generation != seen; });
            if (p->quit)
                return;
            seen = p->generation;
            job  = p->job;
        }
        work(p, p->workers[wi], *job);
        {
            std::lock_guard<std::mutex> lock(p->mu);
            if (--p->busy == 0)
                p->done.notify_all();
        }
    }
}

// Chunk schedule of a run over `n` windows: equal chunks of `chunk` windows.  (Measured on a B200 with cfg 2 windows:
// shrinking chunks towards the end, or a small first round, cost more in per-chunk launch overhead and under-filled
// launches than they gain at the tail -- 57 K windows/s against 75 K with equal chunks of 1 184.)
std::vector<int32_t> plan_chunks(int32_t n, int32_t chunk)
{
    std::vector<int32_t> ends;
    for (int32_t at = 0; at < n;)
    {
        at += std::min(chunk, n - at);
        ends.push_back(at);
    }
    return ends;
}
} // namespace

extern "C" {

const char* cudapoa_b200_pipeline_last_error(void)
{
    return g_pipeline_error.c_str();
}

void cudapoa_b200_pipeline_destroy(cudapoa_b200_pipeline* p);

cudapoa_b200_pipeline* cudapoa_b200_pipeline_create(const int32_t* devices, int32_t n_devices,
                                                    int32_t workers_per_device, size_t mem_per_worker,
                                                    int8_t output_mask, const cudapoa_b200_batch_size* bs,
                                                    int16_t gap_score, int16_t mismatch_score, int16_t match_score,
                                                    int32_t cuda_banded_alignment)
{
    cudapoa_b200_pipeline* p = nullptr;
    try
    {
        if (n_devices <= 0 || workers_per_device <= 0)
            throw std::invalid_argument("pipeline needs at least one device and one worker");
        p              = new cudapoa_b200_pipeline();
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
        p->workers.resize(static_cast<size_t>(n_devices) * static_cast<size_t>(workers_per_device));
        size_t wi = 0;
        // worker k of every device first, then worker k + 1: chunk c goes to device c mod n_devices at the start
        for (int32_t k = 0; k < workers_per_device; k++)
        {
            for (int32_t d = 0; d < n_devices; d++)
            {
                cudapoa_b200_pipeline::Worker& w = p->workers[wi++];
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
            }
        }
        for (size_t i = 0; i < p->workers.size(); i++)
            p->workers[i].thread = std::thread(worker_main, p, i);
        return p;
    }
    catch (const std::exception& e)
    {
        g_pipeline_error = e.what();
        cudapoa_b200_pipeline_destroy(p);
        return nullptr;
    }
}

void cudapoa_b200_pipeline_destroy(cudapoa_b200_pipeline* p)
{
    if (p == nullptr)
        return;
    {
        std::lock_guard<std::mutex> lock(p->mu);
        p->quit = true;
    }
    p->wake.notify_all();
    for (auto& w : p->workers)
        if (w.thread.joinable())
            w.thread.join();
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
// chunk_windows: windows per batch round (0: a quarter of the smallest batch capacity, at least 256).  Windows are dealt most expensive first (cost: bases of the window, i.e. reads x length at a fixed
// band), results are indexed by window whatever the order.
// Returns 0, or a negative value with cudapoa_b200_pipeline_last_error() set.
int32_t cudapoa_b200_pipeline_run(cudapoa_b200_pipeline* p, const uint8_t* bases, const int8_t* weights,
                                  const int32_t* read_lengths, const int32_t* window_n_reads, int32_t n_windows,
                                  int32_t chunk_windows)
{
    try
    {
        NvtxRange range("cudapoa_b200_pipeline_run");
        cudapoa_b200_pipeline::Job job;
        const auto t_call = std::chrono::steady_clock::now();
        {
            const char* env = std::getenv("CUDAPOA_B200_PIPELINE_TRACE");
            job.trace       = env != nullptr && env[0] == '1';
        }
        job.bases          = bases;
        job.weights        = weights;
        job.read_lengths   = read_lengths;
        job.window_n_reads = window_n_reads;
        job.want_msa       = (p->output_mask & poa::OutputType::msa) != 0;
        p->status.assign(static_cast<size_t>(n_windows), static_cast<int32_t>(poa::StatusType::generic_error));
        p->consensus.assign(job.want_msa ? 0 : static_cast<size_t>(n_windows), std::string());
        p->coverage.assign(job.want_msa ? 0 : static_cast<size_t>(n_windows), std::vector<uint16_t>());
        p->msa.assign(job.want_msa ? static_cast<size_t>(n_windows) : 0, std::vector<std::string>());
        p->stats = cudapoa_b200_stats{};
        if (n_windows <= 0)
            return 0;

        job.read_off.assign(static_cast<size_t>(n_windows) + 1, 0);
        job.base_off.assign(static_cast<size_t>(n_windows) + 1, 0);
        {
            int64_t r = 0, b = 0;
            for (int32_t w = 0; w < n_windows; w++)
            {
                job.read_off[w] = r;
                job.base_off[w] = b;
                for (int32_t k = 0; k < window_n_reads[w]; k++)
                    b += read_lengths[r + k];
                r += window_n_reads[w];
            }
            job.read_off[n_windows] = r;
            job.base_off[n_windows] = b;
        }
        // most expensive first, in 4 K-base classes (a stable sort: windows of one class keep the input order)
        job.order.resize(static_cast<size_t>(n_windows));
        for (int32_t w = 0; w < n_windows; w++)
            job.order[w] = w;
        std::stable_sort(job.order.begin(), job.order.end(), [&](int32_t x, int32_t y) {
            return ((job.base_off[x + 1] - job.base_off[x]) >> 12) > ((job.base_off[y + 1] - job.base_off[y]) >> 12);
        });
        int32_t min_cap = INT32_MAX;
        for (const auto& w : p->workers)
            min_cap = std::min(min_cap, w.capacity);
        if (min_cap <= 0)
            throw std::runtime_error("pipeline: a batch has no capacity");
        int32_t chunk = chunk_windows > 0 ? chunk_windows : std::max(256, min_cap / 4);
        chunk         = std::min(chunk, min_cap);
        job.chunk_end = plan_chunks(n_windows, chunk);

        const auto t0 = std::chrono::steady_clock::now();
        job.t0        = t0;
        {
            std::unique_lock<std::mutex> lock(p->mu);
            p->job  = &job;
            p->busy = static_cast<int32_t>(p->workers.size());
            p->generation++;
            p->wake.notify_all();
            p->done.wait(lock, [&] { return p->busy == 0; });
            p->job = nullptr;
        }
        p->wall_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        if (job.trace)
        {
            std::fprintf(stderr, "pipeline trace: %d windows, schedule built in %.2f ms, workers ran %.2f ms\n", n_windows,
                         std::chrono::duration<double, std::milli>(t0 - t_call).count(), p->wall_ms);
            std::sort(job.trace_rows.begin(), job.trace_rows.end(),
                      [](const std::array<double, 6>& x, const std::array<double, 6>& y) { return x[2] < y[2]; });
            for (const auto& r : job.trace_rows)
                std::fprintf(stderr, "  worker %2d  %5d windows  stage %7.2f  launch %7.2f  collect %7.2f  end %7.2f   (stage %.2f, enqueue %.2f, wait+collect %.2f)\n",
                             static_cast<int>(r[0]), static_cast<int>(r[1]), r[2], r[3], r[4], r[5], r[3] - r[2], r[4] - r[3], r[5] - r[4]);
        }
        if (!job.first_error.empty())
            throw std::runtime_error(job.first_error);
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
