// Phase-split schedule of the banded POA build: the per-read loop of poa_window_kernel (poa_window.cuh) cut at the
// boundary between the score fill and the graph work, one launch per phase and read:
//
//   for s = 1 .. max reads - 1:   poa_fill_kernel(s)    descriptors, fill, sink selection          (all lanes busy)
//                                 poa_graph_kernel(s)   traceback, fusion, topological sort, and -- for windows whose
//                                                       last read this is -- consensus / MSA       (mostly one lane)
//
// Same device functions, same order per window, hence the same results as the single launch; what changes is what
// shares an SM.  The fill is a chain of packed integer instructions that needs 128 registers and the row ring; the
// graph phases are latency-bound walks that need few registers and the traceback tile.  In one kernel every warp holds
// the larger of the two footprints all the time (16 windows per SM); apart, the graph launches keep 24 windows per SM
// in flight and the fill launches run with every resident warp in the fill.  Windows carry their state between
// launches in PhaseState; the graph, the score rows and the row descriptors already live in global memory.
// Follows generatePOAKernel's control flow and error protocol (cudapoa_kernels.cuh:171-359) like poa_window.cuh.
#pragma once

#include "poa_device.cuh"
#include "poa_fast.cuh"
#include "poa_trace2.cuh"

namespace poa_b200
{

constexpr int32_t kFillWarpsPerBlock  = 4;
constexpr int32_t kGraphWarpsPerBlock = 4;
constexpr int32_t kTileBytes          = kTileAndDescBytes; // traceback tile + its row descriptors
#ifndef POA_FILL_MIN_BLOCKS
#define POA_FILL_MIN_BLOCKS 8
#endif

// First launch of a batch: per-window state and the backbone from read 0 (cudapoa_kernels.cuh:171-209).
template <typename IdT, bool MSA>
__global__ void __launch_bounds__(kFillWarpsPerBlock * 32) poa_init_kernel(const KernelParams P)
{
    const int32_t lane = threadIdx.x & 31;
    for (int32_t w = blockIdx.x * kFillWarpsPerBlock + (threadIdx.x >> 5); w < P.n_windows;
         w += gridDim.x * kFillWarpsPerBlock)
    {
        const WindowDesc wd = P.windows[w];
        const int32_t* rlen = P.read_len + wd.read_off;
        PhaseState st       = PhaseState{};
        st.status           = kStSuccess;
        if (wd.n_reads <= 0 || rlen[0] <= 0)
            st.status = kStGenericError; // the reference has no defined behaviour for an empty backbone
        else if (rlen[0] > P.layout.M)
            st.status = kStNodeCountExceeded;
        else
        {
            const Graph<IdT> g = make_graph<IdT>(P.arena + static_cast<int64_t>(w) * P.layout.stride, P.layout);
            st.node_count      = rlen[0];
            init_backbone<IdT, MSA>(g, P.bases + wd.base_off, P.weights ? P.weights + wd.base_off : nullptr,
                                    st.node_count, lane);
        }
        if (lane == 0)
            P.phase[w] = st;
    }
}

template <typename ScoreT, typename IdT>
__global__ void __launch_bounds__(kFillWarpsPerBlock * 32, POA_FILL_MIN_BLOCKS) poa_fill_kernel(const KernelParams P)
{
    POA_DYNAMIC_SMEM(smem_block);
    int32_t lane  = threadIdx.x & 31;
    uint32_t smem = smem_u32(smem_block) + static_cast<uint32_t>(threadIdx.x >> 5) * static_cast<uint32_t>(P.smem_per_warp);
    asm volatile("" : "+r"(lane));
    asm volatile("" : "+r"(smem));
    // the block's operand table of the match-bit profile (poa_fill2.cuh), behind the warps' scratch
    uint32_t lut = smem_u32(smem_block) + static_cast<uint32_t>(kFillWarpsPerBlock) * static_cast<uint32_t>(P.smem_per_warp);
    asm volatile("" : "+r"(lut));
    fill2_build_lut(lut, P.match - P.gap, P.mismatch - P.gap, lane);
    const int32_t s = P.step;

    for (;;)
    {
        int32_t w = 0;
        if (lane == 0)
            w = atomicAdd(P.next_window, 1);
        w = __shfl_sync(kFull, w, 0);
        if (w >= P.n_windows)
            break;

        const WindowDesc wd = P.windows[w];
        const Graph<IdT> g  = make_graph<IdT>(P.arena + static_cast<int64_t>(w) * P.layout.stride, P.layout);
        const int32_t* rlen = P.read_len + wd.read_off;
        PhaseState* ps      = P.phase + w;
        PhaseState st       = *ps;
        if (st.status == kStSuccess && s < wd.n_reads)
        {
            st.seq_off += rlen[s - 1];
            const uint8_t* seq = P.bases + wd.base_off + st.seq_off;
            const int32_t len  = rlen[s];
            if (st.node_count >= P.max_nodes) // cudapoa_kernels.cuh:220-226
            {
                st.status = kStNodeCountExceeded;
            }
            else
            {
                ScoreT* scores      = static_cast<ScoreT*>(P.scores) + static_cast<int64_t>(w) * P.banded_score_stride;
                int32_t aln_len     = kFastNotTaken;
                int32_t start_row   = 0;
                int32_t start_score = 0;
                // This kernel holds the packed fill of the short windows only.  Long windows (rebased rows) and the
                // exact-arithmetic alignment run in launches of their own (poa_fill_rebased_kernel, poa_exact_kernel):
                // registers and instruction cache of this one belong to the loop the benchmark spends half its time in.
                int32_t beta = fast_path_beta(st.node_count, len, P.band, P.gap, P.mismatch, P.match);
                if (beta == 0 && P.smem_per_warp < fill2_smem_bytes(len)) // the read outgrows this launch's profile
                    beta = fast_path_rebased_beta(st.node_count, P.band, P.gap, P.mismatch, P.match);
                if (beta > 0)
                    aln_len = kAlnNeedsRebased;
                else if (beta == 0 &&
                         fill2_impl<IdT, false>(g, st.node_count, seq, len, reinterpret_cast<uint8_t*>(scores),
                                                P.banded_score_stride * static_cast<int64_t>(sizeof(ScoreT)), P.max_nodes,
                                                P.gap, P.mismatch, P.match, smem, lane, start_row, start_score, 0, P.gap_x2, lut) == 0)
                    aln_len = kAlnPendingTraceback;
                else
                    aln_len = kAlnNeedsExact;
                st.aln_len   = aln_len;
                st.start_row = start_row;
                st.aln_score = start_score;
                st.beta      = 0;
                st.rows += static_cast<uint32_t>(st.node_count);
                st.cells += static_cast<unsigned long long>(st.node_count) * static_cast<unsigned long long>(P.band);
            }
        }
        if (lane == 0)
            *ps = st;
        __syncwarp(); // the scratch is reused by the next window
    }
}

// Full alignment (the reference's default mode, cudapoa_nw.cuh): this launch takes the place of poa_fill_kernel.  The
// packed fill runs the matrix in column chunks of 256 cells (fill2_impl<.., FULL>); a window's score area
// (scores_width x score_rows cells at its own offset) holds the chunks' move codes and the spill slots.
template <typename ScoreT, typename IdT>
__global__ void __launch_bounds__(kFillWarpsPerBlock * 32, POA_FILL_MIN_BLOCKS) poa_fill_full_kernel(const KernelParams P)
{
    POA_DYNAMIC_SMEM(smem_block);
    int32_t lane  = threadIdx.x & 31;
    uint32_t smem = smem_u32(smem_block) + static_cast<uint32_t>(threadIdx.x >> 5) * static_cast<uint32_t>(P.smem_per_warp);
    asm volatile("" : "+r"(lane));
    asm volatile("" : "+r"(smem));
    // the block's operand table of the match-bit profile (poa_fill2.cuh), behind the warps' scratch
    uint32_t lut = smem_u32(smem_block) + static_cast<uint32_t>(kFillWarpsPerBlock) * static_cast<uint32_t>(P.smem_per_warp);
    asm volatile("" : "+r"(lut));
    fill2_build_lut(lut, P.match - P.gap, P.mismatch - P.gap, lane);
    const int32_t s = P.step;
    for (;;)
    {
        int32_t w = 0;
        if (lane == 0)
            w = atomicAdd(P.next_window, 1);
        w = __shfl_sync(kFull, w, 0);
        if (w >= P.n_windows)
            break;
        const WindowDesc wd = P.windows[w];
        const Graph<IdT> g  = make_graph<IdT>(P.arena + static_cast<int64_t>(w) * P.layout.stride, P.layout);
        const int32_t* rlen = P.read_len + wd.read_off;
        PhaseState* ps      = P.phase + w;
        PhaseState st       = *ps;
        if (st.status == kStSuccess && s < wd.n_reads)
        {
            st.seq_off += rlen[s - 1];
            const uint8_t* seq = P.bases + wd.base_off + st.seq_off;
            const int32_t len  = rlen[s];
            if (st.node_count >= P.max_nodes) // cudapoa_kernels.cuh:220-226
            {
                st.status = kStNodeCountExceeded;
            }
            else
            {
                uint8_t* area       = reinterpret_cast<uint8_t*>(static_cast<ScoreT*>(P.scores) + wd.score_off);
                int64_t area_bytes  = static_cast<int64_t>(wd.scores_width) * P.score_rows * static_cast<int64_t>(sizeof(ScoreT));
                const int64_t skew  = (16 - (reinterpret_cast<uintptr_t>(area) & 15)) & 15; // code rows want 16-byte alignment
                area += skew;
                area_bytes -= skew;
                int32_t start_row   = 0;
                int32_t start_score = 0;
                const int32_t beta  = fast_path_beta(st.node_count, len, kFastBand, P.gap, P.mismatch, P.match);
                if (beta == 0 && P.smem_per_warp >= fill2_smem_bytes(len) &&
                    fill2_impl<IdT, false, true>(g, st.node_count, seq, len, area, area_bytes, P.max_nodes, P.gap, P.mismatch,
                                                 P.match, smem, lane, start_row, start_score, 0, P.gap_x2, lut) == 0)
                    st.aln_len = kAlnPendingTraceback;
                else
                    st.aln_len = kAlnNeedsExact;
                st.start_row = start_row;
                st.aln_score = start_score;
                st.beta      = 0;
                st.rows += static_cast<uint32_t>(st.node_count);
                st.cells += static_cast<unsigned long long>(st.node_count) * static_cast<unsigned long long>(len);
            }
        }
        if (lane == 0)
            *ps = st;
        __syncwarp(); // the scratch is reused by the next window
    }
}

// The packed fill of long windows (rows stored relative to a per-row offset, poa_fast.cuh fast_path_beta), for the
// reads poa_fill_kernel marked kAlnNeedsRebased.
template <typename ScoreT, typename IdT>
__global__ void __launch_bounds__(kFillWarpsPerBlock * 32, 6) poa_fill_rebased_kernel(const KernelParams P)
{
    POA_DYNAMIC_SMEM(smem_block);
    int32_t lane  = threadIdx.x & 31;
    uint32_t smem = smem_u32(smem_block) + static_cast<uint32_t>(threadIdx.x >> 5) * static_cast<uint32_t>(P.smem_per_warp);
    asm volatile("" : "+r"(lane));
    asm volatile("" : "+r"(smem));
    const int32_t s = P.step;
    for (;;)
    {
        int32_t w = 0;
        if (lane == 0)
            w = atomicAdd(P.next_window, 1);
        w = __shfl_sync(kFull, w, 0);
        if (w >= P.n_windows)
            break;
        PhaseState* ps = P.phase + w;
        if (ps->aln_len != kAlnNeedsRebased)
            continue;
        PhaseState st       = *ps;
        const WindowDesc wd = P.windows[w];
        if (st.status != kStSuccess || st.finalized || s >= wd.n_reads)
            continue;
        const Graph<IdT> g  = make_graph<IdT>(P.arena + static_cast<int64_t>(w) * P.layout.stride, P.layout);
        const int32_t len   = P.read_len[wd.read_off + s];
        const uint8_t* seq  = P.bases + wd.base_off + st.seq_off;
        ScoreT* scores      = static_cast<ScoreT*>(P.scores) + static_cast<int64_t>(w) * P.banded_score_stride;
        int32_t start_row   = 0;
        int32_t start_score = 0;
        int32_t beta        = fast_path_beta(st.node_count, len, P.band, P.gap, P.mismatch, P.match);
        if (beta == 0) // sent here for its length (poa_fill_kernel): rebased rows are exact for every size
            beta = fast_path_rebased_beta(st.node_count, P.band, P.gap, P.mismatch, P.match);
        if (beta > 0 && fill2_impl<IdT, true>(g, st.node_count, seq, len, reinterpret_cast<uint8_t*>(scores),
                                              P.banded_score_stride * static_cast<int64_t>(sizeof(ScoreT)), P.max_nodes,
                                              P.gap, P.mismatch, P.match, smem, lane, start_row, start_score, beta, P.gap_x2) == 0)
        {
            st.aln_len   = kAlnPendingTraceback;
            st.start_row = start_row;
            st.aln_score = start_score;
            st.beta      = beta;
        }
        else
        {
            st.aln_len = kAlnNeedsExact;
        }
        if (lane == 0)
            *ps = st;
        __syncwarp(); // the scratch is reused by the next window
    }
}

// The reads the packed fill declined (other band widths, scores outside its range, rows decided by sentinel
// arithmetic): the exact-arithmetic alignment, traceback included.  Launched after every fill launch; a warp that
// finds nothing to do for a window moves on after one 4-byte load.
template <typename ScoreT, typename IdT, bool FULL = false>
__global__ void __launch_bounds__(kFillWarpsPerBlock * 32, 4) poa_exact_kernel(const KernelParams P)
{
    int32_t lane = threadIdx.x & 31;
    asm volatile("" : "+r"(lane));
    const int32_t s = P.step;
    for (;;)
    {
        int32_t w = 0;
        if (lane == 0)
            w = atomicAdd(P.next_window, 1);
        w = __shfl_sync(kFull, w, 0);
        if (w >= P.n_windows)
            break;
        PhaseState* ps = P.phase + w;
        if (ps->aln_len != kAlnNeedsExact)
            continue;
        PhaseState st       = *ps;
        const WindowDesc wd = P.windows[w];
        if (st.status != kStSuccess || st.finalized || s >= wd.n_reads)
            continue;
        const Graph<IdT> g  = make_graph<IdT>(P.arena + static_cast<int64_t>(w) * P.layout.stride, P.layout);
        const int32_t* rlen = P.read_len + wd.read_off;
        const uint8_t* seq  = P.bases + wd.base_off + st.seq_off;
        ScoreT* scores      = static_cast<ScoreT*>(P.scores) +
                         (FULL ? wd.score_off : static_cast<int64_t>(w) * P.banded_score_stride);
        int32_t aln_score   = 0;
        st.aln_len = align_exact<ScoreT, IdT, !FULL>(g, st.node_count, seq, rlen[s], scores,
                                                     FULL ? wd.scores_width : P.band + kBandRowPad, P.band, P.gap,
                                                     P.mismatch, P.match, lane, &aln_score);
        st.aln_score = aln_score;
        st.exact++;
        if (lane == 0)
            *ps = st;
        __syncwarp();
    }
}

// ---- graph work of a step, one launch per phase ------------------------------------------------------------------------
// The phases after the fill are walks by one lane per window (traceback, topological sort) or gathers over the graph
// (fusion): bound by dependent latency times the windows an SM keeps in flight, not by issue slots.  Each phase gets
// a kernel of its own so that each runs at ITS occupancy -- registers and scratch of the traceback do not cap the
// sort, and the other way round.

#ifndef POA_TRACE_MIN_BLOCKS
#define POA_TRACE_MIN_BLOCKS 6
#endif
#ifndef POA_FUSE_MIN_BLOCKS
#define POA_FUSE_MIN_BLOCKS 8
#endif
#ifndef POA_SORT_MIN_BLOCKS
#define POA_SORT_MIN_BLOCKS 10
#endif

// Traceback over the move codes of the read the fill launch has just aligned (poa_trace2.cuh).
template <typename ScoreT, typename IdT, bool FULL = false>
__global__ void __launch_bounds__(kGraphWarpsPerBlock * 32, POA_TRACE_MIN_BLOCKS) poa_trace_kernel(const KernelParams P)
{
    POA_DYNAMIC_SMEM(smem_block);
    int32_t lane  = threadIdx.x & 31;
    uint32_t smem = smem_u32(smem_block) + static_cast<uint32_t>(threadIdx.x >> 5) * static_cast<uint32_t>(P.smem_per_warp);
    asm volatile("" : "+r"(lane));
    asm volatile("" : "+r"(smem));
    for (;;)
    {
        int32_t w = 0;
        if (lane == 0)
            w = atomicAdd(P.next_window, 1);
        w = __shfl_sync(kFull, w, 0);
        if (w >= P.n_windows)
            break;
        PhaseState* ps = P.phase + w;
        const PhaseState st = *ps;
        __syncwarp(); // every lane has read the state before lane 0 changes it
        if (st.aln_len != kAlnPendingTraceback || st.status != kStSuccess || st.finalized)
            continue;
        const WindowDesc wd = P.windows[w];
        const Graph<IdT> g  = make_graph<IdT>(P.arena + static_cast<int64_t>(w) * P.layout.stride, P.layout);
        const int32_t len   = P.read_len[wd.read_off + P.step];
        const uint8_t* area = reinterpret_cast<const uint8_t*>(
            static_cast<ScoreT*>(P.scores) + (FULL ? wd.score_off : static_cast<int64_t>(w) * P.banded_score_stride));
        if (FULL)
            area += (16 - (reinterpret_cast<uintptr_t>(area) & 15)) & 15; // as poa_fill_full_kernel aligned it
        const int32_t aln_len = trace2_impl<IdT, FULL>(g, st.node_count, len, area, smem, P.smem_per_warp, st.start_row, lane);
        if (lane == 0)
            ps->aln_len = aln_len;
        __syncwarp(); // the scratch is reused by the next window
    }
}

// The traceback with one thread per window (trace2_lane): no shared memory, no window counter -- thread t of the grid
// takes window t.
template <typename ScoreT, typename IdT, bool FULL = false>
__global__ void __launch_bounds__(kGraphWarpsPerBlock * 32) poa_trace_lanes_kernel(const KernelParams P)
{
    const int32_t w = static_cast<int32_t>(blockIdx.x * blockDim.x + threadIdx.x);
    if (w >= P.n_windows)
        return;
    PhaseState* ps      = P.phase + w;
    const PhaseState st = *ps;
    if (st.aln_len != kAlnPendingTraceback || st.status != kStSuccess || st.finalized)
        return;
    const WindowDesc wd = P.windows[w];
    const Graph<IdT> g  = make_graph<IdT>(P.arena + static_cast<int64_t>(w) * P.layout.stride, P.layout);
    const int32_t len   = P.read_len[wd.read_off + P.step];
    const uint8_t* area = reinterpret_cast<const uint8_t*>(
        static_cast<ScoreT*>(P.scores) + (FULL ? wd.score_off : static_cast<int64_t>(w) * P.banded_score_stride));
    if (FULL)
        area += (16 - (reinterpret_cast<uintptr_t>(area) & 15)) & 15; // as poa_fill_full_kernel aligned it
    ps->aln_len = trace2_lane<IdT, FULL>(g, st.node_count, len, area, st.start_row);
}

// Adaptive banding (extension, no counterpart in the reference; the rule is poa_window.cuh's): after the traceback of a
// step, an alignment that scores below 3/4 of a perfect match or that runs within kAdaptiveBandMargin cells of a band
// edge is repeated with the band doubled, while that raises the score or the path still hugs an edge, up to
// P.adaptive_max_band.  The first alignment comes from the packed fill like every other; repeats (rare on ordinary
// windows) use the exact-arithmetic alignment, whose band is a parameter.  Launched only for adaptive batches.
template <typename ScoreT, typename IdT>
__global__ void __launch_bounds__(kFillWarpsPerBlock * 32, 4) poa_widen_kernel(const KernelParams P)
{
    int32_t lane = threadIdx.x & 31;
    asm volatile("" : "+r"(lane));
    const int32_t s = P.step;
    for (;;)
    {
        int32_t w = 0;
        if (lane == 0)
            w = atomicAdd(P.next_window, 1);
        w = __shfl_sync(kFull, w, 0);
        if (w >= P.n_windows)
            break;
        PhaseState* ps = P.phase + w;
        PhaseState st  = *ps;
        __syncwarp(); // every lane has read the state before lane 0 changes it
        const WindowDesc wd = P.windows[w];
        if (st.status != kStSuccess || st.finalized || s >= wd.n_reads || st.aln_len < 0)
            continue;
        const Graph<IdT> g = make_graph<IdT>(P.arena + static_cast<int64_t>(w) * P.layout.stride, P.layout);
        const int32_t len  = P.read_len[wd.read_off + s];
        const uint8_t* seq = P.bases + wd.base_off + st.seq_off;
        ScoreT* scores     = static_cast<ScoreT*>(P.scores) + static_cast<int64_t>(w) * P.banded_score_stride;
        int32_t band_now   = P.band;
        int32_t aln_len    = st.aln_len;
        int32_t aln_score  = st.aln_score;
        bool again = 4 * aln_score < 3 * P.match * len || alignment_touches_band<IdT>(g, st.node_count, len, band_now, lane);
        if (!again)
            continue;
        while (again && band_now < P.adaptive_max_band)
        {
            band_now *= 2;
            const int32_t before = aln_score;
            __syncwarp();
            aln_len = align_exact<ScoreT, IdT, true>(g, st.node_count, seq, len, scores, band_now + kBandRowPad, band_now,
                                                     P.gap, P.mismatch, P.match, lane, &aln_score);
            st.exact += 0x10000u; // high half: widened alignments
            st.cells += static_cast<unsigned long long>(st.node_count) * static_cast<unsigned long long>(band_now);
            again = aln_len >= 0 &&
                    (aln_score > before || alignment_touches_band<IdT>(g, st.node_count, len, band_now, lane));
        }
        st.aln_len   = aln_len;
        st.aln_score = aln_score;
        if (lane == 0)
            *ps = st;
        __syncwarp();
    }
}

// Fusion of the alignment into the graph (add_alignment).
template <typename IdT, bool MSA>
__global__ void __launch_bounds__(kGraphWarpsPerBlock * 32, POA_FUSE_MIN_BLOCKS) poa_fuse_kernel(const KernelParams P)
{
    int32_t lane = threadIdx.x & 31;
    asm volatile("" : "+r"(lane));
    const int32_t s = P.step;
    for (;;)
    {
        int32_t w = 0;
        if (lane == 0)
            w = atomicAdd(P.next_window, 1);
        w = __shfl_sync(kFull, w, 0);
        if (w >= P.n_windows)
            break;
        PhaseState* ps = P.phase + w;
        const PhaseState st = *ps;
        __syncwarp(); // every lane has read the state before lane 0 changes it
        if (st.status != kStSuccess || st.finalized)
            continue;
        const WindowDesc wd = P.windows[w];
        if (s >= wd.n_reads)
            continue;
        int32_t node_count = st.node_count;
        int32_t status     = kStSuccess;
        if (st.aln_len == -1) // cudapoa_kernels.cuh:281-289
        {
            status = kStLoopCountExceeded;
        }
        else
        {
            const Graph<IdT> g = make_graph<IdT>(P.arena + static_cast<int64_t>(w) * P.layout.stride, P.layout);
            const int64_t off  = st.seq_off;
            const uint8_t* seq = P.bases + wd.base_off + off;
            const int8_t* wts  = P.weights ? P.weights + wd.base_off + off : nullptr;
            const int32_t len  = P.read_len[wd.read_off + s];
            status = add_alignment<IdT, MSA>(g, node_count, seq, len, wts,
                                             MSA ? g.path + static_cast<int64_t>(s) * P.layout.max_seq : nullptr,
                                             P.max_nodes, lane); // cudapoa_kernels.cuh:314-319 on failure
        }
        if (lane == 0)
        {
            ps->node_count = node_count;
            ps->status     = status;
        }
    }
}

// Topological sort of the graph the fusion left (two launches per step that differ in scratch: a window goes to the one
// its graph fits; outgrowing both is legal, the sort then runs out of global memory).  LEAN: the instance of the
// small-scratch launch; it leaves the SPOA_ACCURATE order to the other one and does not carry its code.
template <typename IdT, bool LEAN>
__global__ void __launch_bounds__(kGraphWarpsPerBlock * 32, POA_SORT_MIN_BLOCKS) poa_sort_kernel(const KernelParams P)
{
    POA_DYNAMIC_SMEM(smem_block);
    int32_t lane  = threadIdx.x & 31;
    uint32_t smem = smem_u32(smem_block) + static_cast<uint32_t>(threadIdx.x >> 5) * static_cast<uint32_t>(P.smem_per_warp);
    asm volatile("" : "+r"(lane));
    asm volatile("" : "+r"(smem));
    const int32_t s = P.step;
    for (;;)
    {
        int32_t w = 0;
        if (lane == 0)
            w = atomicAdd(P.next_window, 1);
        w = __shfl_sync(kFull, w, 0);
        if (w >= P.n_windows)
            break;
        PhaseState* ps = P.phase + w;
        const PhaseState st = *ps;
        __syncwarp(); // every lane has read the state before lane 0 changes it
        if (st.status != kStSuccess || st.finalized || st.graph_step == s)
            continue;
        const WindowDesc wd = P.windows[w];
        if (s >= wd.n_reads)
            continue;
        const int32_t node_count = st.node_count;
        // class 0: fits the small scratch; class 1: the 14 KB scratch or, beyond it, global memory
        const int32_t cls = ((P.lean_class0 && P.spoa_accurate) || node_count > P.class_cap) ? 1 : 0;
        if (cls != P.phase_class)
            continue;
        const Graph<IdT> g = make_graph<IdT>(P.arena + static_cast<int64_t>(w) * P.layout.stride, P.layout);
        if (!LEAN && P.spoa_accurate)
            topsort_spoa_order<IdT>(g, node_count, lane);
        else if ((P.debug_flags & 1) || !topsort_kahn_smem<IdT>(g, node_count, smem, P.smem_per_warp, lane))
            topsort_kahn_warp<IdT>(g, node_count, lane);
        if (lane == 0)
            ps->graph_step = s;
        __syncwarp(); // the scratch is reused by the next window
    }
}

// Outputs of the windows that are complete after step P.step (their last read, or a failure): consensus + coverage or
// MSA rows, status, counters.  Launched after the steps at which some window of the batch ends, and after the last one.
template <typename IdT, bool MSA>
__global__ void __launch_bounds__(kGraphWarpsPerBlock * 32, 4) poa_final_kernel(const KernelParams P)
{
    POA_DYNAMIC_SMEM(smem_block);
    int32_t lane  = threadIdx.x & 31;
    uint32_t smem = smem_u32(smem_block) + static_cast<uint32_t>(threadIdx.x >> 5) * static_cast<uint32_t>(P.smem_per_warp);
    asm volatile("" : "+r"(lane));
    asm volatile("" : "+r"(smem));
    const int32_t s = P.step;
    for (;;)
    {
        int32_t w = 0;
        if (lane == 0)
            w = atomicAdd(P.next_window, 1);
        w = __shfl_sync(kFull, w, 0);
        if (w >= P.n_windows)
            break;
        PhaseState* ps = P.phase + w;
        const PhaseState st = *ps;
        __syncwarp(); // every lane has read the state before lane 0 changes it
        if (st.finalized)
            continue;
        const WindowDesc wd = P.windows[w];
        if (st.status == kStSuccess && s < wd.n_reads - 1)
            continue; // more reads to come
        int32_t status = st.status;
        if (status == kStSuccess)
        {
            const Graph<IdT> g  = make_graph<IdT>(P.arena + static_cast<int64_t>(w) * P.layout.stride, P.layout);
            const int32_t* rlen = P.read_len + wd.read_off;
            if (!MSA)
            {
                uint8_t* cons_out = P.consensus + static_cast<int64_t>(w) * P.max_consensus;
                uint16_t* cov_out = P.coverage + static_cast<int64_t>(w) * P.max_consensus;
                status = generate_consensus<IdT>(g, st.node_count, cons_out, cov_out, P.max_consensus, smem,
                                                 P.smem_per_warp, lane);
            }
            else
            {
                topsort_spoa_order<IdT>(g, st.node_count, lane);
                const int32_t msa_len = msa_columns<IdT>(g, st.node_count, lane);
                if (msa_len >= P.max_consensus) // cudapoa_generate_msa.cuh:200-204
                    status = kStSeqSizeExceeded;
                if (status == kStSuccess)
                    msa_rows<IdT>(g, P.bases + wd.base_off, rlen, wd.n_reads,
                                  P.msa + static_cast<int64_t>(w) * P.max_reads * P.max_consensus, msa_len,
                                  P.max_consensus, P.layout.max_seq, lane);
            }
        }
        if (lane == 0)
        {
            P.status[w]     = static_cast<uint8_t>(status);
            P.node_count[w] = st.node_count;
            P.stat_rows[w]  = st.rows;
            P.stat_exact[w] = st.exact;
            P.stat_cells[w] = st.cells;
            ps->status      = status;
            ps->finalized   = 1;
        }
        __syncwarp(); // the scratch is reused by the next window
    }
}

} // namespace poa_b200
