// Per-window driver kernel: backbone -> per read {align, fuse, sort} -> consensus or MSA, one warp per window.
// Follows the control flow and error protocol of generatePOAKernel + generateConsensusKernel /
// generateMSAKernel (cudapoa_kernels.cuh:66-359, cudapoa_generate_consensus.cuh:278-347,
// cudapoa_generate_msa.cuh:120-224), fused into a single launch.
#pragma once

#include "poa_device.cuh"
#include "poa_fast.cuh"

namespace poa_b200
{

constexpr int32_t kWarpsPerBlock = 4;
constexpr int32_t kSmemPerWarp   = 14 * 1024; // 4 blocks x 4 warps x 14 KB per SM

template <typename ScoreT, typename IdT, bool BANDED, bool MSA>
__global__ void __launch_bounds__(kWarpsPerBlock * 32, 4) poa_window_kernel(const KernelParams P)
{
    POA_DYNAMIC_SMEM(smem_block);
    int32_t lane = threadIdx.x & 31;
    // this warp's scratch as a 32-bit shared address; both values are made opaque so that they live in registers
    // instead of being recomputed from the special registers all over the kernel
    uint32_t smem = smem_u32(smem_block) + static_cast<uint32_t>(threadIdx.x >> 5) * static_cast<uint32_t>(P.smem_per_warp);
    asm volatile("" : "+r"(lane));
    asm volatile("" : "+r"(smem));

    // persistent warp: windows differ in length and depth, so they are handed out one at a time
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
    const uint8_t* seq  = P.bases + wd.base_off;
    const int8_t* wts   = P.weights ? P.weights + wd.base_off : nullptr;
    const int32_t* rlen = P.read_len + wd.read_off;
    ScoreT* scores      = static_cast<ScoreT*>(P.scores) +
                     (BANDED ? static_cast<int64_t>(w) * P.banded_score_stride : wd.score_off);
    const int32_t width = BANDED ? P.band + kBandRowPad : wd.scores_width;

    int32_t status     = kStSuccess;
    int32_t node_count = 0;
    uint32_t rows      = 0;
    uint32_t exact     = 0;
    unsigned long long cells = 0;

    if (wd.n_reads <= 0 || rlen[0] <= 0)
    {
        status = kStGenericError; // the reference has no defined behaviour for an empty backbone
    }
    else if (rlen[0] > P.layout.M)
    {
        status = kStNodeCountExceeded;
    }
    else
    {
        node_count = rlen[0];
        init_backbone<IdT, MSA>(g, seq, wts, node_count, lane);
        __syncwarp();

        for (int32_t s = 1; s < wd.n_reads; s++)
        {
            seq += rlen[s - 1];
            if (wts)
                wts += rlen[s - 1];
            const int32_t len = rlen[s];
            if (node_count >= P.max_nodes) // cudapoa_kernels.cuh:220-226
            {
                status = kStNodeCountExceeded;
                break;
            }
            int32_t aln_len   = kFastNotTaken;
            int32_t aln_score = 0;
            // The packed path computes exact integer scores whenever it accepts a window, so it also serves batches
            // whose score type is int32 because of their largest window; it only borrows the window's score area.
            if (BANDED)
                aln_len = align_fast_banded16<IdT>(g, node_count, seq, len, reinterpret_cast<int16_t*>(scores),
                                                   P.band, P.gap, P.mismatch, P.match, smem, P.smem_per_warp, lane,
                                                   &aln_score);
            if (aln_len == kFastNotTaken)
            {
                aln_len = align_exact<ScoreT, IdT, BANDED>(g, node_count, seq, len, scores, width, P.band, P.gap,
                                                           P.mismatch, P.match, lane, &aln_score);
                exact++;
            }
            rows += static_cast<uint32_t>(node_count);
            cells += static_cast<unsigned long long>(node_count) * static_cast<unsigned long long>(BANDED ? P.band : len);
            if (BANDED && P.adaptive_max_band > 0)
            {
                // Adaptive banding (extension, no counterpart in the reference).  A static band can miss the optimum
                // in two ways: the path found runs along a band edge, or it stays on the diagonal as a run of
                // mismatches because the true path left the band.  So the alignment is repeated with the band
                // doubled when an aligned pair lies within kAdaptiveBandMargin cells of a band edge that is not a
                // matrix edge, or when the score is below 3/4 of a perfect match; doubling goes on while it raises
                // the score or the path still hugs an edge.  Repeats use the exact path (its band is a parameter).
                int32_t band_now = P.band;
                bool again       = aln_len >= 0 && (4 * aln_score < 3 * P.match * len ||
                                              alignment_touches_band<IdT>(g, node_count, len, band_now, lane));
                while (again && band_now < P.adaptive_max_band)
                {
                    band_now *= 2;
                    const int32_t before = aln_score;
                    __syncwarp();
                    aln_len = align_exact<ScoreT, IdT, true>(g, node_count, seq, len, scores, band_now + kBandRowPad,
                                                             band_now, P.gap, P.mismatch, P.match, lane, &aln_score);
                    exact += 0x10000u; // high half: widened alignments
                    cells += static_cast<unsigned long long>(node_count) * static_cast<unsigned long long>(band_now);
                    again = aln_len >= 0 &&
                            (aln_score > before || alignment_touches_band<IdT>(g, node_count, len, band_now, lane));
                }
            }
            if (aln_len == -1) // cudapoa_kernels.cuh:281-289
            {
                status = kStLoopCountExceeded;
                break;
            }
            __syncwarp(); // lane 0's traceback output becomes visible to the warp
            int32_t rc = add_alignment<IdT, MSA>(g, node_count, seq, len, wts,
                                                 MSA ? g.path + static_cast<int64_t>(s) * P.layout.max_seq : nullptr,
                                                 P.max_nodes, lane);
            __syncwarp();
            if (rc == kStSuccess)
            {
                if (P.spoa_accurate)
                    topsort_spoa_order<IdT>(g, node_count, lane);
                else if (!topsort_kahn_smem<IdT>(g, node_count, smem, P.smem_per_warp, lane))
                    topsort_kahn_warp<IdT>(g, node_count, lane);
            }
            __syncwarp();
            if (rc != kStSuccess) // cudapoa_kernels.cuh:314-319
            {
                status = rc;
                break;
            }
        }
    }

    if (status == kStSuccess)
    {
        if (!MSA)
        {
            uint8_t* cons_out  = P.consensus + static_cast<int64_t>(w) * P.max_consensus;
            uint16_t* cov_out  = P.coverage + static_cast<int64_t>(w) * P.max_consensus;
            status = generate_consensus<IdT>(g, node_count, cons_out, cov_out, P.max_consensus, smem, P.smem_per_warp, lane);
        }
        else
        {
            topsort_spoa_order<IdT>(g, node_count, lane);
            const int32_t msa_len = msa_columns<IdT>(g, node_count, lane);
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
        P.node_count[w] = node_count;
        P.stat_rows[w]  = rows;
        P.stat_exact[w] = exact;
        P.stat_cells[w] = cells;
    }
    __syncwarp(); // the scratch is reused by the next window
    }
}

} // namespace poa_b200
