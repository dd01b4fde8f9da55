// Traceback over the move codes written by fill2_impl (poa_fill2.cuh).
//
// The reference re-derives every move from stored scores (cudapoa_nw_banded.cuh:386-474); here the fill has already
// decided, per cell, which move that derivation takes (kind + predecessor slot), so a step of the walk is: load the
// cell's code, load the row's descriptor, move.  Lane 0 walks; the warp stages the codes of 64 rows x 64 columns around
// the walk into shared memory (2 KB, column-aligned so that a cell's nibble is addressed by (row, column) alone) with
// the rows' records (predecessor rows, in-degree, band start).  Rows with more than kWidePreds predecessors keep a byte per cell in global
// memory and are read from there.
//
// Output as fast_traceback_impl's: g.aln_node[k] = node paired with read position k, or -1 for an inserted base.
#pragma once

#include "poa_fill2.cuh"

#include <type_traits>

namespace poa_b200
{

constexpr int32_t kT2Rows      = 64;
constexpr int32_t kT2Cols      = 64;                     // columns per tile, a multiple of 8
constexpr int32_t kT2RowBytes  = kT2Cols / 2;            // nibbles
constexpr int32_t kT2TileBytes = kT2Rows * kT2RowBytes;  // 2 KB
constexpr int32_t kT2Bytes     = kT2TileBytes + kT2Rows * 16; // + the rows' records

// Shared memory the traceback wants: the tile, and -- when it fits -- 2 bytes per read position for the rows the walk
// records (the serial lane then never stores to global memory).
__host__ __device__ inline int32_t trace2_smem_bytes(int32_t read_len)
{
    return kT2Bytes + ((2 * read_len + 15) & ~15);
}

// Code of cell `cell` of a row that stores a byte per cell (more than kWidePreds predecessors).  Out of line on purpose:
// the walk's loop should not carry the address arithmetic of a load it almost never issues.
__device__ POA_NOINLINE uint32_t wide_row_code(const uint8_t* codes, int64_t row, int32_t cell)
{
    POA_COUNT("trace2.wide_steps");
    return __ldcg(codes + row * kCodeRowBytes + cell);
}

// Returns the alignment length (> 0), or -1 if a code is not a move (cannot happen for codes the fill wrote; kept as the
// reference's loop-count guard).  All 32 lanes.  `smem_bytes` of scratch at `tile`.
// FULL (full alignment, fill2_impl<.., FULL>): the codes of column chunk c, row r are code row c * (node_count + 1) + r
// and every row's "band start" is 256 c; the walk stops at a chunk's left edge and the next tile is staged from the
// chunk before.
template <typename IdT, bool FULL = false>
__device__ int32_t trace2_impl(const Graph<IdT>& g, int32_t node_count, int32_t read_len, const uint8_t* area,
                               uint32_t tile, int32_t smem_bytes, int32_t start_row, int32_t lane)
{
    const uint8_t* codes = area;
    const uint4* records = reinterpret_cast<const uint4*>(g.rowinfo);
    const uint32_t tile_desc = tile + kT2TileBytes;
    const uint32_t rows_s    = tile + kT2Bytes;                        // u16[read_len]: row paired with each base
    const bool in_smem       = smem_bytes >= trace2_smem_bytes(read_len);
    int32_t i = start_row, j = read_len;
    int32_t steps = 0;
    bool broken   = false;
    while (i > 0 && j > 0 && !broken) // uniform: i, j are broadcast at the end of every round
    {
        // tile: rows tile_r0 .. tile_r0 + 63 (the walk's row is the last one), columns tile_c0 .. tile_c0 + 63 with
        // tile_c0 = 1 (mod 8) and j among the last eight
        POA_COUNT("trace2.tiles");
        const int32_t tile_r0 = i - (kT2Rows - 1);
        const int32_t tile_c0 = ((j - 1) & ~7) + 1 - (kT2Cols - 8);
        const int32_t chunk_lo = FULL ? ((j - 1) & ~(kFastBand - 1)) : 0;          // the walk stays right of this column
        const int64_t chunk_row0 = FULL ? static_cast<int64_t>(chunk_lo / kFastBand) * (node_count + 1) : 0;
#pragma unroll
        for (int32_t u = 0; u < kT2Rows / 32; u++)
        {
            const int32_t tr = lane + 32 * u;
            const int32_t r  = tile_r0 + tr;
            uint32_t w[kT2Cols / 8 + 1];
            uint4 dsc = make_uint4(0u, 0u, 0u, 0u);
            int32_t c0 = 0;
            if (r >= 1)
            {
                dsc              = records[r];
                const int32_t b  = FULL ? chunk_lo : static_cast<int32_t>(dsc.w & 0xffffu);
                c0               = tile_c0 - b - 1;       // band cell under the tile's first column, a multiple of 4
                const int32_t w0 = c0 >> 3;               // first word of the row the tile needs (floor division)
                const uint32_t* row = reinterpret_cast<const uint32_t*>(codes + (chunk_row0 + r) * kCodeRowBytes);
#pragma unroll
                for (int32_t q = 0; q <= kT2Cols / 8; q++)
                {
                    const int32_t wi = w0 + q;
                    w[q]             = (static_cast<uint32_t>(wi) < 32u) ? __ldcg(row + wi) : 0xEEEEEEEEu; // outside the band
                }
            }
            else
            {
#pragma unroll
                for (int32_t q = 0; q <= kT2Cols / 8; q++)
                    w[q] = 0xEEEEEEEEu;
            }
            // column-align: when the tile's first column falls in the middle of a word, shift by half a word
            const uint32_t sel = (c0 & 4) ? 0x5432u : 0x3210u;
            const uint32_t at  = tile + static_cast<uint32_t>(tr * kT2RowBytes);
            sts_v4(at, make_uint4(__byte_perm(w[0], w[1], sel), __byte_perm(w[1], w[2], sel), __byte_perm(w[2], w[3], sel),
                                  __byte_perm(w[3], w[4], sel)));
            sts_v4(at + 16u, make_uint4(__byte_perm(w[4], w[5], sel), __byte_perm(w[5], w[6], sel),
                                        __byte_perm(w[6], w[7], sel), __byte_perm(w[7], w[8], sel)));
            sts_v4(tile_desc + 16u * static_cast<uint32_t>(tr), dsc);
        }
        __syncwarp();
        if (lane == 0)
        {
            const int32_t row_floor = max(tile_r0, 1);
            // shared addresses of record(row) and code word(row, column) as base + 16 * row / base + 32 * row + ..
            // (32-bit arithmetic; it may wrap)
            const uint32_t desc_base = tile_desc - 16u * static_cast<uint32_t>(tile_r0);
            const uint32_t code_base = tile - static_cast<uint32_t>(kT2RowBytes) * static_cast<uint32_t>(tile_r0);
            // the walk proper, instantiated for both homes of the recorded rows so that the choice costs nothing per step
            auto walk = [&](auto rows_in_smem) {
                while (i >= row_floor && j >= tile_c0 && j > chunk_lo)
                {
                    // descriptor and code word are fetched together: one shared-memory latency per step
                    const int32_t x   = j - tile_c0; // 0 .. 63
                    const uint4 dd    = lds_v4(desc_base + 16u * static_cast<uint32_t>(i));
                    const uint32_t cw = lds_u32(code_base + static_cast<uint32_t>(kT2RowBytes) * static_cast<uint32_t>(i) +
                                                static_cast<uint32_t>((x >> 3) << 2));
                    const int32_t pc  = static_cast<int32_t>((dd.z >> 16) & 0x3fu);
                    uint32_t code     = (cw >> ((x & 7) * 4)) & 0xfu;
                    if (pc > kWidePreds) // rare: such rows keep a byte per cell in global memory
                        code = wide_row_code(codes, chunk_row0 + i, j - (FULL ? chunk_lo : static_cast<int32_t>(dd.w & 0xffffu)) - 1);
                    steps++;
                    const uint32_t at = rows_s + 2u * static_cast<uint32_t>(j - 1); // where read position j - 1 is recorded
                    if (code & 2u) // horizontal: an inserted base
                    {
                        if (rows_in_smem.value)
                            sts_u16(at, 0u);
                        else
                            g.aln_node[j - 1] = 0;
                        j--;
                        continue;
                    }
                    const int32_t p = static_cast<int32_t>(code >> 2);
                    int32_t pi      = static_cast<int32_t>(dd.x & 0xffffu);
                    if (p != 0)
                    {
                        if (p >= pc)
                        {
                            broken = true; // not a code the fill writes
                            break;
                        }
                        pi = record_pred_row<IdT>(g, dd.x, dd.y, i, p);
                    }
                    if ((code & 1u) == 0u) // diagonal: the base pairs with this row's node
                    {
                        if (rows_in_smem.value)
                            sts_u16(at, static_cast<uint32_t>(i));
                        else
                            g.aln_node[j - 1] = static_cast<IdT>(i);
                        j--;
                    }
                    i = pi;
                }
            };
            if (in_smem)
                walk(std::true_type{});
            else
                walk(std::false_type{});
        }
        i      = __shfl_sync(kFull, i, 0);
        j      = __shfl_sync(kFull, j, 0);
        broken = __shfl_sync(kFull, static_cast<int32_t>(broken), 0) != 0;
        __syncwarp();
    }
    steps = __shfl_sync(kFull, steps, 0);
    if (broken)
        return -1;
    // the walk reached row 0 (the remaining bases are inserted before the graph's first node) or column 0.
    // Recorded ROW indices (0: inserted base) become node ids here.
    __syncwarp();
    for (int32_t k = lane; k < read_len; k += 32)
    {
        int32_t row = 0;
        if (k >= j)
            row = in_smem ? static_cast<int32_t>(lds_u16(rows_s + 2u * static_cast<uint32_t>(k)))
                          : static_cast<int32_t>(g.aln_node[k]);
        g.aln_node[k] = static_cast<IdT>(row > 0 ? static_cast<int32_t>(g.sorted[row - 1]) : -1);
    }
    return max(steps + j, 1);
}

// The same traceback with ONE THREAD PER WINDOW (poa_trace_lanes_kernel): a step is a serial decision whatever is done, so
// the warp-per-window walk above keeps 31 lanes idle and at most a few dozen windows per SM in flight; here every
// window of the batch walks at once, each thread following its own codes out of global memory (the records of
// neighbouring rows share cache lines; the code row a few steps ahead is prefetched), the 32 threads of a warp share
// the instruction stream of 32 walks, and only what a walk touches is read: one 32-byte sector per row instead of the
// tile's two or three.  Moves are taken exactly as trace2_impl takes them; node ids are written as the walk goes.
// Returns the alignment length (> 0) or -1.
#ifndef POA_TRACE_PREFETCH
#define POA_TRACE_PREFETCH 1 // 0: none, 1: into L2, 2: into L1
#endif
#ifndef POA_TRACE_AHEAD
#define POA_TRACE_AHEAD 6
#endif
template <typename IdT, bool FULL = false>
__device__ int32_t trace2_lane(const Graph<IdT>& g, int32_t node_count, int32_t read_len, const uint8_t* area,
                               int32_t start_row)
{
    const uint8_t* codes = area;
    const uint4* records = reinterpret_cast<const uint4*>(g.rowinfo);
    // A row's band start follows from its rank (band_start, as the fill placed it): the code word's address then does
    // not wait for the row's record, and the two loads of a step are in flight together.
    const int32_t max_column = read_len + 1;
    const float gradient     = band_gradient(read_len, node_count);
    int32_t i = start_row, j = read_len;
    int32_t steps = 0;
    // the node id of a paired base is loaded in one step and stored in the next: the store never waits for its load
    int32_t pend_at = -1;
    IdT pend_id     = static_cast<IdT>(-1);
    while (i > 0 && j > 0)
    {
        const int32_t chunk_lo   = FULL ? ((j - 1) & ~(kFastBand - 1)) : 0;
        const int64_t chunk_row0 = FULL ? static_cast<int64_t>(chunk_lo / kFastBand) * (node_count + 1) : 0;
        const int32_t bs         = FULL ? chunk_lo : band_start(i, gradient, kFastBand, max_column);
        const int32_t cell       = j - bs - 1; // 0-based cell of the row's band
        const bool inside        = static_cast<uint32_t>(cell) < static_cast<uint32_t>(kFastBand);
        const uint8_t* row       = codes + (chunk_row0 + i) * kCodeRowBytes;
        const uint4 dd           = __ldg(records + i);
        uint32_t word            = 0xEEEEEEEEu; // outside the band: an inserted base, as the tile's padding says
        if (inside)
            word = __ldg(reinterpret_cast<const uint32_t*>(row) + (cell >> 3));
#if !defined(POA_EMU) && POA_TRACE_PREFETCH > 0
        // the walk climbs about a row per step: the record and the code sector POA_TRACE_AHEAD rows up
        if (i > POA_TRACE_AHEAD)
        {
#if POA_TRACE_PREFETCH == 1
            asm volatile("prefetch.global.L2 [%0];" ::"l"(row - POA_TRACE_AHEAD * kCodeRowBytes + (max(cell - 4, 0) >> 1)));
            asm volatile("prefetch.global.L2 [%0];" ::"l"(records + i - POA_TRACE_AHEAD));
#else
            asm volatile("prefetch.global.L1 [%0];" ::"l"(row - POA_TRACE_AHEAD * kCodeRowBytes + (max(cell - 4, 0) >> 1)));
            asm volatile("prefetch.global.L1 [%0];" ::"l"(records + i - POA_TRACE_AHEAD));
#endif
        }
#endif
        if (pend_at >= 0)
        {
            g.aln_node[pend_at] = pend_id;
            pend_at             = -1;
        }
        const int32_t pc = static_cast<int32_t>((dd.z >> 16) & 0x3fu);
        uint32_t code    = (word >> ((cell & 7) * 4)) & 0xfu;
        if (pc > kWidePreds && inside) // rare: such rows keep a byte per cell
            code = __ldg(row + cell);
        steps++;
        if (code & 2u) // horizontal: an inserted base
        {
            g.aln_node[j - 1] = static_cast<IdT>(-1);
            j--;
            continue;
        }
        const int32_t p = static_cast<int32_t>(code >> 2);
        int32_t pi      = static_cast<int32_t>(dd.x & 0xffffu);
        if (p != 0)
        {
            if (p >= pc)
                return -1; // not a code the fill writes
            pi = record_pred_row<IdT>(g, dd.x, dd.y, i, p);
        }
        if ((code & 1u) == 0u) // diagonal: the base pairs with this row's node
        {
            pend_id = g.sorted[i - 1];
            pend_at = j - 1;
            j--;
        }
        i = pi;
    }
    if (pend_at >= 0)
        g.aln_node[pend_at] = pend_id;
    // the walk reached row 0 (the remaining bases are inserted before the graph's first node) or column 0
    for (int32_t k = 0; k < j; k++)
        g.aln_node[k] = static_cast<IdT>(-1);
    return max(steps + j, 1);
}

} // namespace poa_b200
