// cudapoa_b200_consensus_noalloc: banded int16 consensus for a few windows without any allocation.
//
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  It runs the single-launch POA kernel
// (poa_window.cuh, the same device code create_batch() uses for small batches) with every piece of per-window state --
// graph arena, score rows, window descriptors, counters, raw outputs -- in ONE file-scope device array of
// kNoAllocScratchBytes.  Every GPU that loads the library holds that array until the process exits.
//
// Rules that follow from the shared array:
//   * the scratch keeps the last call's contents; what must start at zero (the window counter) is cleared by the
//     set-up kernel of every call, everything else is written before it is read;
//   * calls that overlap on the GPU (different streams, or different host threads) race on it: serialise them;
//   * requests whose state does not fit are refused with an error that names the bytes needed and the maximum.
// In host code `g_noalloc_scratch` is a host-side stand-in, not a device address: the address the kernels receive comes
// from cudaGetSymbolAddress, looked up once per device with that device current.
#include "cudapoa_b200.h"
#include "poa_layout.hpp"
#include "poa_window.cuh"

#include <claraparabricks/genomeworks/cudapoa/batch.hpp>
#include <claraparabricks/genomeworks/utils/cudautils.hpp>

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <mutex>

namespace
{

using namespace poa_b200;
namespace poa = claraparabricks::genomeworks::cudapoa;

constexpr size_t kNoAllocScratchBytes = static_cast<size_t>(CUDAPOA_B200_NOALLOC_SCRATCH_BYTES);
constexpr int32_t kMaxDevices         = 64;

static __device__ __align__(256) uint8_t g_noalloc_scratch[kNoAllocScratchBytes];

// Byte offsets of one call's state inside the scratch array.
struct ScratchPlan
{
    int64_t windows, counter, status, node_count, stat_rows, stat_exact, stat_cells, raw_consensus, raw_coverage, arena,
        scores, total;
};

ScratchPlan plan_scratch(int64_t n, int64_t C, int64_t arena_stride, int64_t score_bytes_per_window)
{
    ScratchPlan p{};
    int64_t off = 0;
    auto take   = [&](int64_t bytes) {
        const int64_t at = off;
        off              = align_up64(off + bytes, 256);
        return at;
    };
    p.windows       = take(n * static_cast<int64_t>(sizeof(WindowDesc)));
    p.counter       = take(256);
    p.status        = take(n);
    p.node_count    = take(n * 4);
    p.stat_rows     = take(n * 4);
    p.stat_exact    = take(n * 4);
    p.stat_cells    = take(n * 8);
    p.raw_consensus = take(n * C);
    p.raw_coverage  = take(n * C * 2);
    p.arena         = take(n * arena_stride);
    p.scores        = take(n * score_bytes_per_window);
    p.total         = off;
    return p;
}

// Set-up of one call: window descriptors from the reads-per-window and read-length arrays (a serial prefix sum: the
// entry point serves a handful of windows), the window counter back to zero.  A window that breaks the preconditions
// (no reads, more reads than max_reads, a read longer than max_seq) gets n_reads = 0, which the POA kernel reports as
// generic_error.
__global__ void noalloc_setup_kernel(WindowDesc* windows, int32_t* counter, const int32_t* read_len,
                                     const int32_t* window_n_reads, int32_t n_windows, int32_t max_reads, int32_t max_seq)
{
    if (blockIdx.x != 0 || threadIdx.x != 0)
        return;
    int64_t base_off = 0;
    int32_t read_off = 0;
    bool trusted     = true; // false once a read count was out of range: the offsets of later windows are unknown
    for (int32_t w = 0; w < n_windows; w++)
    {
        const int32_t nr = window_n_reads[w];
        bool ok          = trusted && nr > 0 && nr <= max_reads;
        if (nr < 0 || nr > max_reads)
            trusted = false; // never use a count that was not validated as a loop bound or an offset
        int64_t bases = 0;
        if (ok)
        {
            for (int32_t r = 0; r < nr; r++)
            {
                const int32_t len = read_len[read_off + r];
                ok                = ok && len >= 0 && len <= max_seq;
                bases += len > 0 ? len : 0;
            }
        }
        WindowDesc wd{};
        wd.base_off     = base_off;
        wd.score_off    = 0;
        wd.read_off     = read_off;
        wd.n_reads      = ok ? nr : 0;
        wd.scores_width = 0;
        windows[w]      = wd;
        if (trusted)
        {
            base_off += bases;
            read_off += nr;
        }
    }
    *counter = 0;
}

// The POA kernel leaves consensus and coverage end -> start, NUL terminated (the host getters of the Batch interface
// turn them round); here a warp per window writes them start -> end into the caller's buffers.
__global__ void noalloc_emit_kernel(const uint8_t* raw_consensus, const uint16_t* raw_coverage, const uint8_t* status,
                                    int32_t n_windows, int32_t C, uint8_t* consensus_out, uint16_t* coverage_out,
                                    int32_t* length_out, int32_t* status_out)
{
    const int32_t lane = threadIdx.x & 31;
    for (int32_t w = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); w < n_windows; w += gridDim.x * (blockDim.x >> 5))
    {
        const uint8_t* rc  = raw_consensus + static_cast<int64_t>(w) * C;
        const uint16_t* rv = raw_coverage + static_cast<int64_t>(w) * C;
        uint8_t* oc        = consensus_out + static_cast<int64_t>(w) * C;
        uint16_t* ov       = coverage_out + static_cast<int64_t>(w) * C;
        const int32_t st   = status[w];
        int32_t len        = 0;
        if (st == 0)
        {
            // length = index of the first NUL (warp-parallel strnlen)
            len = C;
            for (int32_t k0 = 0; k0 < C && len == C; k0 += 32)
            {
                const int32_t k     = k0 + lane;
                const unsigned mask = __ballot_sync(0xffffffffu, k < C && rc[k] == 0);
                if (mask != 0u)
                    len = k0 + __ffs(mask) - 1;
            }
        }
        for (int32_t k = lane; k < len; k += 32)
        {
            oc[k] = rc[len - 1 - k];
            ov[k] = rv[len - 1 - k];
        }
        if (lane == 0)
        {
            if (len < C)
                oc[len] = 0;
            length_out[w] = len;
            status_out[w] = st;
        }
    }
}

struct DeviceSetup
{
    bool ready        = false;
    uint8_t* scratch  = nullptr; // device address of g_noalloc_scratch on this device
    int32_t sm_count  = 0;
    int32_t blocks_sm = 1;
};
DeviceSetup g_setup[kMaxDevices];
std::mutex g_setup_mutex;

int32_t fail(char* msg, int32_t cap, int32_t code, const char* text)
{
    if (msg != nullptr && cap > 0)
    {
        std::strncpy(msg, text, static_cast<size_t>(cap) - 1);
        msg[cap - 1] = 0;
    }
    return code;
}

} // namespace

extern "C" int32_t cudapoa_b200_consensus_noalloc(int32_t device_id, void* stream, const uint8_t* bases_d,
                                                  const int32_t* read_lengths_d, const int32_t* window_n_reads_d,
                                                  int32_t n_windows, int32_t max_sequences_per_poa,
                                                  int32_t max_sequence_size, int32_t alignment_band_width,
                                                  int16_t gap_score, int16_t mismatch_score, int16_t match_score,
                                                  uint8_t* consensus_d, uint16_t* coverage_d, int32_t* length_d,
                                                  int32_t* status_d, char* error_msg, int32_t error_msg_cap)
{
    char text[320];
    if (n_windows < 0 || max_sequences_per_poa <= 0 || max_sequence_size <= 0 || alignment_band_width <= 0)
        return fail(error_msg, error_msg_cap, -1, "cudapoa_b200_consensus_noalloc: sizes must be positive");
    if (n_windows == 0)
        return 0;
    if (bases_d == nullptr || read_lengths_d == nullptr || window_n_reads_d == nullptr || consensus_d == nullptr ||
        coverage_d == nullptr || length_d == nullptr || status_d == nullptr)
        return fail(error_msg, error_msg_cap, -1, "cudapoa_b200_consensus_noalloc: null input or output pointer");
    if (device_id < 0 || device_id >= kMaxDevices)
        return fail(error_msg, error_msg_cap, -1, "cudapoa_b200_consensus_noalloc: device id out of range");

    // the sizes create_batch() derives from BatchSize(max_sequence_size, max_sequences_per_poa, band) (batch.hpp:53-79)
    poa::BatchSize bs;
    try
    {
        bs = poa::BatchSize(max_sequence_size, max_sequences_per_poa, alignment_band_width);
    }
    catch (const std::exception& ex)
    {
        return fail(error_msg, error_msg_cap, -1, ex.what());
    }
    // int16 scores only: the reference's choice of score type (cudapoa_limits.hpp:28-37) must come out as int16
    {
        const int32_t upper = bs.max_sequence_size * match_score;
        const int32_t lower = bs.max_sequence_size * std::max<int32_t>(gap_score, mismatch_score) +
                              (bs.max_matrix_graph_dimension - bs.max_sequence_size) * gap_score;
        if (upper > INT16_MAX || (-lower) > (INT16_MAX + 1) || bs.max_matrix_graph_dimension_banded > INT16_MAX)
        {
            std::snprintf(text, sizeof(text),
                          "cudapoa_b200_consensus_noalloc: max_sequence_size %d needs 32-bit scores or node ids with these "
                          "scores; this entry point serves the int16 configurations only",
                          max_sequence_size);
            return fail(error_msg, error_msg_cap, -2, text);
        }
    }
    const int32_t M          = bs.max_nodes_per_window_banded;
    const int32_t G          = std::max(bs.max_matrix_graph_dimension_banded, M + 1);
    const int32_t C          = bs.max_consensus_size;
    const ArenaLayout layout = make_arena_layout(M, M + bs.max_sequence_size + 8, max_sequences_per_poa,
                                                 bs.max_sequence_size, 2, false);
    const int64_t score_stride = static_cast<int64_t>(bs.alignment_band_width + kBandRowPad) * G; // int16 cells
    const ScratchPlan sp       = plan_scratch(n_windows, C, layout.stride, score_stride * 2);
    if (sp.total > static_cast<int64_t>(kNoAllocScratchBytes))
    {
        std::snprintf(text, sizeof(text),
                      "cudapoa_b200_consensus_noalloc: %d windows of BatchSize(%d, %d, %d) need %lld bytes of scratch, the "
                      "maximum is %lld (%lld bytes per window): pass fewer windows per call or use create_batch()",
                      n_windows, max_sequence_size, max_sequences_per_poa, bs.alignment_band_width,
                      static_cast<long long>(sp.total), static_cast<long long>(kNoAllocScratchBytes),
                      static_cast<long long>(layout.stride + score_stride * 2));
        return fail(error_msg, error_msg_cap, -3, text);
    }

    // an error some earlier call of this thread left pending must not be reported as this call's
    if (cudaPeekAtLastError() != cudaSuccess)
        return fail(error_msg, error_msg_cap, -4, "a CUDA error was already pending when the call was made");
    int previous = 0;
    cudaError_t e = cudaGetDevice(&previous);
    if (e == cudaSuccess)
        e = cudaSetDevice(device_id);
    if (e != cudaSuccess)
        return fail(error_msg, error_msg_cap, -4, cudaGetErrorString(e));

    auto kernel = poa_window_kernel<int16_t, int16_t, true, false>;
    const size_t smem = static_cast<size_t>(kWarpsPerBlock) * static_cast<size_t>(kSmemPerWarp);
    DeviceSetup setup;
    {
        // once per device, with that device current: the device address of the scratch array and the launch limits
        std::lock_guard<std::mutex> lock(g_setup_mutex);
        DeviceSetup& s = g_setup[device_id];
        if (!s.ready)
        {
            void* ptr = nullptr;
            e         = cudaGetSymbolAddress(&ptr, g_noalloc_scratch);
            if (e == cudaSuccess)
                e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
            int sm = 0, per_sm = 1;
            if (e == cudaSuccess)
                e = cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, device_id);
            if (e == cudaSuccess)
                e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, kWarpsPerBlock * 32, smem);
            if (e != cudaSuccess || ptr == nullptr)
            {
                cudaSetDevice(previous);
                return fail(error_msg, error_msg_cap, -4,
                            e != cudaSuccess ? cudaGetErrorString(e) : "cudaGetSymbolAddress returned a null address");
            }
            s.scratch   = static_cast<uint8_t*>(ptr);
            s.sm_count  = sm;
            s.blocks_sm = std::max(per_sm, 1);
            s.ready     = true;
        }
        setup = s;
    }

    uint8_t* base      = setup.scratch;
    cudaStream_t st    = static_cast<cudaStream_t>(stream);
    WindowDesc* wins   = reinterpret_cast<WindowDesc*>(base + sp.windows);
    int32_t* counter   = reinterpret_cast<int32_t*>(base + sp.counter);
    uint8_t* status    = base + sp.status;
    uint8_t* raw_cons  = base + sp.raw_consensus;
    uint16_t* raw_cov  = reinterpret_cast<uint16_t*>(base + sp.raw_coverage);

    noalloc_setup_kernel<<<1, 32, 0, st>>>(wins, counter, read_lengths_d, window_n_reads_d, n_windows,
                                           max_sequences_per_poa, bs.max_sequence_size);

    KernelParams P{};
    P.bases               = bases_d;
    P.weights             = nullptr;
    P.read_len            = read_lengths_d;
    P.windows             = wins;
    P.n_windows           = n_windows;
    P.arena               = base + sp.arena;
    P.layout              = layout;
    P.scores              = base + sp.scores;
    P.banded_score_stride = score_stride;
    P.status              = status;
    P.node_count          = reinterpret_cast<int32_t*>(base + sp.node_count);
    P.consensus           = raw_cons;
    P.coverage            = raw_cov;
    P.msa                 = nullptr;
    P.stat_rows           = reinterpret_cast<uint32_t*>(base + sp.stat_rows);
    P.stat_exact          = reinterpret_cast<uint32_t*>(base + sp.stat_exact);
    P.stat_cells          = reinterpret_cast<unsigned long long*>(base + sp.stat_cells);
    P.max_nodes           = M;
    P.max_consensus       = C;
    P.max_reads           = max_sequences_per_poa;
    P.band                = bs.alignment_band_width;
    P.gap                 = gap_score;
    P.gap_x2              = (static_cast<uint32_t>(P.gap) & 0xffffu) * 0x10001u;
    P.mismatch            = mismatch_score;
    P.match               = match_score;
    P.spoa_accurate       = 0;
    P.next_window         = counter;
    P.adaptive_max_band   = 0;
    P.smem_per_warp       = kSmemPerWarp;
    const int32_t needed  = (n_windows + kWarpsPerBlock - 1) / kWarpsPerBlock;
    const int32_t grid    = std::max(1, std::min(needed, setup.sm_count * setup.blocks_sm));
    kernel<<<grid, kWarpsPerBlock * 32, smem, st>>>(P);

    noalloc_emit_kernel<<<std::max(1, std::min((n_windows + 3) / 4, 1024)), 128, 0, st>>>(
        raw_cons, raw_cov, status, n_windows, C, consensus_d, coverage_d, length_d, status_d);
    e = cudaPeekAtLastError();
    cudaSetDevice(previous);
    if (e != cudaSuccess)
        return fail(error_msg, error_msg_cap, -4, cudaGetErrorString(e));
    return 0;
}
