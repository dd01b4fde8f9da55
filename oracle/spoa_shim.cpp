// Test/benchmark infrastructure: a C ABI over the UNMODIFIED reference SPOA (3rdparty/spoa v3.0.0),
// compiled by oracle/Makefile into oracle/_ref/libspoa_ref.so.  SPOA is the reference repository's own
// CPU implementation of partial-order alignment + consensus (used by cudapoa/tests/Test_CudapoaGenerateMSA2.cu:55-74
// exactly like run_window() below), so it is the "reference CPU path" timed by bench.py and the
// agreement baseline.  Never linked into the product library.

#include "spoa/spoa.hpp"

#include <atomic>
#include <chrono>
#include <cstdint>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

namespace
{

struct SpoaResult
{
    std::vector<std::string> consensus;
    std::vector<std::vector<std::string>> msa;
    int64_t cells   = 0; // sum over aligned reads of (graph nodes x read length): full-alignment DP cells
    double seconds  = 0;
    int32_t threads = 0;
};

struct Flat
{
    const uint8_t* bases;
    const int32_t* read_lengths;
    const int32_t* window_n_reads;
    int32_t n_windows;
    std::vector<int64_t> base_off; // per window
    std::vector<int64_t> read_off; // per window
};

void run_window(const Flat& in, int32_t w, spoa::AlignmentEngine* engine, bool want_msa, SpoaResult& out,
                int64_t& cells)
{
    auto graph      = spoa::createGraph();
    int64_t off     = in.base_off[w];
    int64_t ridx    = in.read_off[w];
    int32_t n_reads = in.window_n_reads[w];
    for (int32_t r = 0; r < n_reads; r++)
    {
        int32_t len = in.read_lengths[ridx + r];
        std::string seq(reinterpret_cast<const char*>(in.bases + off), static_cast<size_t>(len));
        off += len;
        if (r > 0)
            cells += static_cast<int64_t>(graph->nodes().size()) * len;
        auto alignment = engine->align(seq, graph);
        graph->add_alignment(alignment, seq);
    }
    out.consensus[w] = graph->generate_consensus();
    if (want_msa)
        graph->generate_multiple_sequence_alignment(out.msa[w]);
}

} // namespace

extern "C" {

// Runs SPOA (kNW, linear gaps) on every window with n_threads host threads (0 = hardware concurrency).
void* spoa_ref_run(const uint8_t* bases, const int32_t* read_lengths, const int32_t* window_n_reads,
                   int32_t n_windows, int32_t match, int32_t mismatch, int32_t gap, int32_t want_msa,
                   int32_t n_threads)
{
    Flat in{bases, read_lengths, window_n_reads, n_windows, {}, {}};
    in.base_off.resize(n_windows);
    in.read_off.resize(n_windows);
    int64_t bo = 0, ro = 0;
    int32_t max_len = 1;
    for (int32_t w = 0; w < n_windows; w++)
    {
        in.base_off[w] = bo;
        in.read_off[w] = ro;
        for (int32_t r = 0; r < window_n_reads[w]; r++)
        {
            bo += read_lengths[ro + r];
            if (read_lengths[ro + r] > max_len)
                max_len = read_lengths[ro + r];
        }
        ro += window_n_reads[w];
    }
    if (n_threads <= 0)
        n_threads = static_cast<int32_t>(std::thread::hardware_concurrency());
    if (n_threads <= 0)
        n_threads = 1;
    if (n_threads > n_windows && n_windows > 0)
        n_threads = n_windows;

    auto* res = new SpoaResult();
    res->consensus.resize(n_windows);
    res->msa.resize(n_windows);
    res->threads = n_threads;

    std::atomic<int32_t> next(0);
    std::vector<int64_t> cells(static_cast<size_t>(n_threads), 0);
    auto t0 = std::chrono::steady_clock::now();
    std::vector<std::thread> pool;
    for (int32_t t = 0; t < n_threads; t++)
    {
        pool.emplace_back([&, t]() {
            auto engine = spoa::createAlignmentEngine(spoa::AlignmentType::kNW, static_cast<int8_t>(match),
                                                      static_cast<int8_t>(mismatch), static_cast<int8_t>(gap));
            engine->prealloc(static_cast<uint32_t>(max_len), 4);
            for (;;)
            {
                int32_t w = next.fetch_add(1);
                if (w >= n_windows)
                    break;
                run_window(in, w, engine.get(), want_msa != 0, *res, cells[static_cast<size_t>(t)]);
            }
        });
    }
    for (auto& th : pool)
        th.join();
    auto t1      = std::chrono::steady_clock::now();
    res->seconds = std::chrono::duration<double>(t1 - t0).count();
    for (int64_t c : cells)
        res->cells += c;
    return res;
}

void spoa_ref_free(void* h)
{
    delete static_cast<SpoaResult*>(h);
}

double spoa_ref_seconds(void* h)
{
    return static_cast<SpoaResult*>(h)->seconds;
}

int64_t spoa_ref_cells(void* h)
{
    return static_cast<SpoaResult*>(h)->cells;
}

int32_t spoa_ref_threads(void* h)
{
    return static_cast<SpoaResult*>(h)->threads;
}

int32_t spoa_ref_consensus(void* h, int32_t w, const char** seq)
{
    auto* r = static_cast<SpoaResult*>(h);
    *seq    = r->consensus[w].c_str();
    return static_cast<int32_t>(r->consensus[w].size());
}

int32_t spoa_ref_msa_rows(void* h, int32_t w)
{
    return static_cast<int32_t>(static_cast<SpoaResult*>(h)->msa[w].size());
}

int32_t spoa_ref_msa_row(void* h, int32_t w, int32_t row, const char** seq)
{
    auto* r = static_cast<SpoaResult*>(h);
    *seq    = r->msa[w][row].c_str();
    return static_cast<int32_t>(r->msa[w][row].size());
}

} // extern "C"
