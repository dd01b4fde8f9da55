// DPX (s16x2) fast path of the banded int16 alignment, band width 256.
//
// Same alignment as align_exact<int16_t, IdT, true> whenever it returns a length; it returns kFastNotTaken when it
// cannot guarantee that (the caller then runs the exact path):
//   * a priori, when scores could leave the range the packed arithmetic keeps exact (fast_path_eligible);
//   * during the fill, on the first row none of whose predecessors overlaps the band -- the only situation in
//     which the reference's sentinel arithmetic (cudapoa_nw_banded.cuh:122-174, :199-200) shapes stored values.
//
// Formulation ("fill kernel"):
//   * tilted scores T[r][j] = H[r][j] - gap * j.  The horizontal term H[j-1] + gap becomes T[j-1], so the
//     horizontal pass is a plain prefix maximum; the diagonal term uses (match|mismatch) - gap, the vertical gap.
//   * one warp per row, 8 consecutive cells per lane held as four s16x2 registers R_i = (cell i, cell i + 4), so
//     the in-lane part of the scan runs on both halves at once and three of the four diagonal operands are the
//     vertical operands of the neighbouring register.
//   * rows are stored in that register order, 35 blocks of 16 bytes (560 B): a lead block {band start, -, -,
//     left boundary in the high half}, 32 cell blocks (lane l owns block 1 + l), two pad blocks at the sentinel.
//     Band starts move in multiples of 4 cells, so a predecessor row is read with two aligned 16 B loads per lane
//     and four byte-permutes whose selector depends on the shift; block indices are clamped into the pad blocks and
//     the lead block doubles as "the cell left of cell 0", so the row loop has no range branches.  Row 0 (T == 0)
//     is stored like any other row, its first pad block being (0, sentinel) to reproduce the reference's band + 4
//     rule for it.
//   * cell update: two DPX instructions per register and predecessor,
//       S = __viaddmax_s16x2(D, Q, __viaddmax_s16x2(V, gap, S))   (VIADDMNMX.S16x2 on sm_100a).
//   * memory: every row goes to the window's score matrix in global memory (the traceback needs all of them), and
//     the most recent kRingRows rows also stay in a shared-memory ring, which is where predecessor rows are read
//     from (graph predecessors are almost always a few ranks back).  Per-row descriptors (base, in-degree, up to
//     five predecessor rows) are built by a warp-parallel pre-pass, fetched 32 rows at a time with coalesced loads
//     one chunk ahead, and handed to the row loop by shuffles: the serial loop never chases graph pointers.
//   * traceback: score re-derivation as in the reference, on lane 0, out of shared-memory tiles (scores +
//     descriptors) that the whole warp loads around the current cell.
//   * shared memory per window is the ring plus the zero-padded read (under 10 KB), which is what lets 16 windows
//     share an SM; the code is a chain of dependent integer operations, so resident warps are what hides latency.
#pragma once

#include "poa_device.cuh"

namespace poa_b200
{

constexpr int32_t kFastNotTaken = -2; // the caller must run align_exact instead
constexpr int32_t kFastBand     = 256;
constexpr int32_t kFastRowWords = (kFastBand + kBandRowPad) / 2; // lead block + 32 cell blocks + 2 pad blocks
constexpr int32_t kRowBlocks    = kFastRowWords / 4;             // 35
constexpr int32_t kPadBlock     = kRowBlocks - 1;                // last block of a row, always at the sentinel
constexpr int32_t kNegT         = -30000; // tilted-domain stand-in for the reference's sentinel class
constexpr int32_t kNegThreshold = -25000; // anything at or below is sentinel class
#ifndef POA_RING_ROWS
#define POA_RING_ROWS 8
#endif
constexpr int32_t kRingRows     = POA_RING_ROWS; // power of two; predecessors further back are read through L2
constexpr int32_t kRingBytes    = kRingRows * kFastRowWords * 4;
constexpr int32_t kTileRows     = 64; // traceback tile: rows x columns of plain int16 tilted scores
constexpr int32_t kTileCols     = 48; // columns the walk may use; a tile row also has 8 cells of slack on either side
constexpr int32_t kTileRowBytes = (kTileCols + 16) * 2;
constexpr int32_t kTileItems    = kTileCols / 8 + 1; // 16-byte blocks of a score row that can overlap the columns
static_assert(kRowBlocks == 35, "row layout");
constexpr int32_t kTileAndDescBytes = kTileRows * kTileRowBytes + kTileRows * 8;
static_assert((kRingRows & (kRingRows - 1)) == 0 && kRingRows >= 4, "ring rows are addressed with a mask");

__device__ __forceinline__ uint32_t pack2(int32_t lo, int32_t hi)
{
    return (static_cast<uint32_t>(lo) & 0xffffu) | (static_cast<uint32_t>(hi) << 16);
}
__device__ __forceinline__ int32_t lo16(uint32_t v)
{
    return static_cast<int16_t>(v & 0xffffu);
}
__device__ __forceinline__ int32_t hi16(uint32_t v)
{
    return static_cast<int16_t>(v >> 16);
}

// Which packed formulation serves this alignment.
//   0    real tilted scores provably stay inside [-20000, 30000]; sentinel-class values stay below -25000.
//   > 0  longer windows: scores of row r are stored relative to beta * band_start(r) ("rebased"), which follows the
//        score's growth along the band; the range [kRealLo, kRealHi] is then checked row by row during the fill and
//        the exact path takes over when a row leaves it.  beta is a slope between the tilted score's extremes
//        (-gap per column for a run of poor matches, match - gap for perfect ones).
//   < 0  not eligible.
constexpr int32_t kRealLo = -16000;
constexpr int32_t kRealHi = 28000;
constexpr int32_t kNegAbs = -(1 << 30); // sentinel class once the row offset has been added back (rebased rows)

__device__ __forceinline__ int32_t fast_path_beta(int32_t node_count, int32_t read_len, int32_t band, int32_t gap,
                                                  int32_t mismatch, int32_t match)
{
    if (band != kFastBand || gap >= 0 || match <= 0)
        return -1;
    const int32_t a = max(max(-gap, abs(mismatch)), match);
    if (a > 127 || node_count + 1 >= 32768)
        return -1;
    if (a * (node_count + read_len + 2) <= 20000 && (match - gap) * (read_len + 2) <= 30000)
        return 0;
    const int32_t beta = -gap + (5 * match) / 8;
    // the offset between two rows whose bands overlap must leave real scores real, and a row's own spread must fit
    if (beta > 30 || (match - gap) * (kFastBand + 16) > 12000)
        return -1;
    return beta;
}

// The slope of the rebased formulation for these scores (> 0), or -1 if it cannot serve them.  Rebased rows are exact
// for every window size, so a window whose scores would also fit the plain formulation may use them (the phase-split
// schedule does when the read is longer than the plain fill's shared-memory profile).
__device__ __forceinline__ int32_t fast_path_rebased_beta(int32_t node_count, int32_t band, int32_t gap, int32_t mismatch,
                                                          int32_t match)
{
    if (band != kFastBand || gap >= 0 || match <= 0)
        return -1;
    const int32_t a = max(max(-gap, abs(mismatch)), match);
    if (a > 127 || node_count + 1 >= 32768)
        return -1;
    const int32_t beta = -gap + (5 * match) / 8;
    if (beta > 30 || (match - gap) * (kFastBand + 16) > 12000)
        return -1;
    return beta;
}

// Per-row descriptors, three 32-bit words in three global arrays of stride g.M + 40:
//   d0 = base | in-degree << 8 | sink << 14 | row of predecessor 0 << 16
//   d1 = row of predecessor 1 | row of predecessor 2 << 16
//   d2 = row of predecessor 3 | row of predecessor 4 << 16
struct RowDesc
{
    uint32_t* d0;
    uint32_t* d1;
    uint32_t* d2;
};

// Row of predecessor p of row `row`; slots beyond the fifth are looked up in the graph.
template <typename IdT>
__device__ __forceinline__ int32_t pred_row(const Graph<IdT>& g, uint32_t d0, uint32_t d1, uint32_t d2, int32_t row,
                                            int32_t p)
{
    if (p == 0)
        return static_cast<int32_t>(d0 >> 16);
    const uint32_t w = (p <= 2) ? d1 : d2;
    if (p <= 4)
        return static_cast<int32_t>((p & 1) ? (w & 0xffffu) : (w >> 16));
    return g.pos[g.in_edge(g.sorted[row - 1], p)] + 1;
}

// Reads of the tilted score matrix from global memory with the reference's get_score semantics
// (cudapoa_nw_banded.cuh:106-120), in the tilted domain: out-of-band cells come back as kNegT, which stands for the
// reference's sentinel.  Real values are above kNegThreshold by construction, so a comparison "score == neighbour +
// cost" between a real score and a sentinel-class neighbour is false here exactly as it is in the reference.
__device__ __forceinline__ int32_t tilted_score_global(const uint32_t* rows, int32_t r, int32_t column)
{
    const uint32_t* grow = rows + r * kFastRowWords;
    const uint4 lead     = __ldcg(reinterpret_cast<const uint4*>(grow)); // {band start, -, -, left boundary << 16}
    const int32_t c      = column - static_cast<int32_t>(lead.x) - 1;    // 0-based cell of the band
    int32_t t            = kNegT;
    if (static_cast<uint32_t>(c) < static_cast<uint32_t>(kFastBand) && column != 0)
    {
        const uint32_t w = __ldcg(grow + 4 + ((c >> 3) << 2) + (c & 3));
        t                = (c & 4) ? hi16(w) : lo16(w);
    }
    else if (column == 0 || c == -1)
    {
        t = hi16(lead.w);
    }
    return (t <= kNegThreshold) ? kNegT : t;
}

// The same cell with the row's offset added back (rebased rows, beta > 0): an int32 that compares across rows; the
// sentinel class becomes kNegAbs.  beta == 0: tilted_score_global.
__device__ __forceinline__ int32_t tilted_score_abs(const uint32_t* rows, int32_t r, int32_t column, int32_t beta)
{
    const int32_t t = tilted_score_global(rows, r, column);
    if (beta == 0)
        return t;
    if (t == kNegT)
        return kNegAbs;
    return t + beta * static_cast<int32_t>(__ldcg(rows + r * kFastRowWords));
}

// Load the traceback tile: rows tile_r0 .. tile_r0 + kTileRows - 1, absolute columns tile_c0 .. tile_c0 + kTileCols - 1
// (tile_c0 = 1 mod 4), as plain int16 tilted scores indexed by absolute column: cell (row, column) lives at
//   tile + (row - tile_r0) * kTileRowBytes + 2 * (column - tile_c0 + 8).
// The row-register order of the score matrix is undone here, once per tile, so that the walk reads a cell with one
// address computation and one load.  Cells outside a row's band are written as kNegT, the left-boundary value goes
// to column == band start, which reproduces get_score for every column of the tile.  Row descriptors (d0, d1) go to
// tile_desc.  All lanes; loads are issued in groups of seven per lane before any of them is consumed.
__device__ __forceinline__ void load_tile(const uint32_t* rows, const RowDesc& rd, uint32_t tile, uint32_t tile_desc,
                                          int32_t tile_r0, int32_t tile_c0, float gradient, int32_t max_column,
                                          int32_t lane, int32_t beta)
{
    constexpr int32_t kPerLane = kTileRows * kTileItems / 32; // 14
    constexpr int32_t kGroup   = 7;
    static_assert(kPerLane % kGroup == 0 && kTileRows % 32 == 0, "tile shape");
    const uint32_t NEG2 = pack2(kNegT, kNegT);
    uint2 dsc[kTileRows / 32];
    uint4 lead[kTileRows / 32];
#pragma unroll
    for (int32_t u = 0; u < kTileRows / 32; u++)
    {
        const int32_t r = tile_r0 + lane + 32 * u;
        dsc[u]          = (r >= 1) ? make_uint2(rd.d0[r], rd.d1[r]) : make_uint2(0u, 0u);
        lead[u]         = __ldcg(reinterpret_cast<const uint4*>(rows + max(r, 0) * kFastRowWords));
    }
#pragma unroll
    for (int32_t g0 = 0; g0 < kPerLane; g0 += kGroup)
    {
        uint4 v[kGroup];
        uint32_t at[kGroup];
#pragma unroll
        for (int32_t u = 0; u < kGroup; u++)
        {
            const int32_t idx = lane + 32 * (g0 + u);
            const int32_t tr  = idx / kTileItems;
            const int32_t kb  = idx - tr * kTileItems;
            const int32_t r   = max(tile_r0 + tr, 0);
            const int32_t b   = band_start(r, gradient, kFastBand, max_column);
            const int32_t c0  = tile_c0 - b - 1;  // band cell under the tile's first column, a multiple of 4
            const int32_t blk = (c0 >> 3) + kb;   // block of the row this item moves (floor division)
            const int32_t x   = 8 * blk - c0 + 8; // its first cell in the tile row: 4 .. 56
            at[u]             = tile + static_cast<uint32_t>(tr * kTileRowBytes + 2 * x);
            v[u]              = make_uint4(NEG2, NEG2, NEG2, NEG2);
            if (static_cast<uint32_t>(blk) < 32u)
                v[u] = __ldcg(reinterpret_cast<const uint4*>(rows + r * kFastRowWords) + 1 + blk);
        }
#pragma unroll
        for (int32_t u = 0; u < kGroup; u++)
        {
            // words (c0,c4) (c1,c5) (c2,c6) (c3,c7) -> cells in column order
            sts_v2(at[u], make_uint2(__byte_perm(v[u].x, v[u].y, 0x5410), __byte_perm(v[u].z, v[u].w, 0x5410)));
            sts_v2(at[u] + 8u, make_uint2(__byte_perm(v[u].x, v[u].y, 0x7632), __byte_perm(v[u].z, v[u].w, 0x7632)));
        }
    }
    __syncwarp();
#pragma unroll
    for (int32_t u = 0; u < kTileRows / 32; u++)
    {
        sts_v2(tile_desc + 8u * (lane + 32 * u), dsc[u]);
        // the row's score offset (rebased rows) in the first, otherwise unused, bytes of the tile row
        sts_u32(tile + static_cast<uint32_t>((lane + 32 * u) * kTileRowBytes), static_cast<uint32_t>(beta) * lead[u].x);
        // left boundary: the cell at column == band start
        const int32_t x = static_cast<int32_t>(lead[u].x) - tile_c0;
        if (static_cast<uint32_t>(x) < static_cast<uint32_t>(kTileCols))
            sts_u16(tile + static_cast<uint32_t>((lane + 32 * u) * kTileRowBytes + 2 * (x + 8)), lead[u].w >> 16);
    }
    __syncwarp();
}

// Shared memory the packed fill needs for a read of this length: the ring plus the zero-padded read.
__device__ __forceinline__ int32_t fast_fill_smem_bytes(int32_t read_len)
{
    return kRingBytes + ((read_len + kFastBand + 8 + 15) & ~15);
}

// First half of the packed alignment: row descriptors, the fill (every row streamed to the score matrix) and the
// sink selection.  Returns kFastNotTaken, or 0 with the row the traceback starts from in start_row (uniform over
// the warp).  The read stays staged at smem + kRingBytes.
template <typename IdT, bool REBASED>
__device__ int32_t fast_fill_impl(const Graph<IdT>& g, int32_t node_count, const uint8_t* read, int32_t read_len,
                                  int16_t* scores, int32_t gap, int32_t mismatch, int32_t match, uint32_t smem,
                                  int32_t lane, int32_t& start_row, int32_t beta_arg, int32_t* score_out)
{
    // shared-memory scratch: [ring of recent rows | read, zero padded]
    const int32_t rd_bytes = (read_len + kFastBand + 8 + 15) & ~15;
    const int32_t beta     = REBASED ? beta_arg : 0; // a compile-time zero in the instance the short windows use
    uint32_t* rows      = reinterpret_cast<uint32_t*>(scores);
    const uint32_t ring = smem;            // shared addresses from here on
    const uint32_t rd   = ring + kRingBytes;
    for (int32_t k = lane; k < rd_bytes; k += 32)
        sts_u8(rd + k, (k < read_len) ? read[k] : 0u);
    RowDesc desc;
    desc.d0 = g.rowinfo;
    desc.d1 = desc.d0 + (g.M + 40);
    desc.d2 = desc.d1 + (g.M + 40);

    const int32_t max_column = read_len + 1;
    const float gradient     = band_gradient(read_len, node_count);
    const uint32_t NEG2      = pack2(kNegT, kNegT);
    const uint4 NEG4         = make_uint4(NEG2, NEG2, NEG2, NEG2);

    // pre-pass: row descriptors.  Each lane takes kPre rows per pass; the three dependent gather levels
    // (rank -> node -> in-edges -> ranks of the predecessors) are each issued for all of them before the next.
    {
        constexpr int32_t kPre = 4;
        for (int32_t rb = 1; rb <= node_count; rb += 32 * kPre)
        {
            int32_t node[kPre];
            uint32_t pc[kPre], word[kPre];
            int32_t pe[kPre][5];
#pragma unroll
            for (int u = 0; u < kPre; u++)
            {
                const int32_t r = min(rb + lane + 32 * u, node_count);
                node[u]         = g.sorted[r - 1];
            }
#pragma unroll
            for (int u = 0; u < kPre; u++)
            {
                pc[u]   = g.in_cnt[node[u]];
                word[u] = static_cast<uint32_t>(g.nodes[node[u]]) | ((g.out_cnt[node[u]] == 0) ? (1u << 14) : 0u);
#pragma unroll
                for (int k = 0; k < 5; k++)
                    pe[u][k] = g.in_edge(node[u], k); // slots beyond the in-degree: stale but readable
            }
#pragma unroll
            for (int u = 0; u < kPre; u++)
            {
                uint32_t pr[5];
#pragma unroll
                for (int k = 0; k < 5; k++)
                    pr[k] = (static_cast<uint32_t>(k) < pc[u]) ? static_cast<uint32_t>(g.pos[pe[u][k]] + 1) : 0u;
                const int32_t r = rb + lane + 32 * u;
                if (r <= node_count)
                {
                    desc.d0[r] = word[u] | (pc[u] << 8) | (pr[0] << 16);
                    desc.d1[r] = pr[1] | (pr[2] << 16);
                    desc.d2[r] = pr[3] | (pr[4] << 16);
                }
            }
        }
    }
    // Row 0: T == 0 in columns 0 .. band + 3, sentinel beyond (cudapoa_nw_banded.cuh:138-143, :208-212); lead block
    // {band start 0, -, -, left boundary 0}.  It lives in global memory only: a row whose predecessor is row 0 (a
    // node without in-edges) takes the far path below, which knows row 0's padding.  Ring slots carry their two pad
    // blocks at the sentinel once and for all; global rows are never read beyond their cell blocks.
    {
        const uint4 z = make_uint4(0u, 0u, 0u, 0u);
        __stcg(reinterpret_cast<uint4*>(rows) + lane, z);
        if (lane == 0)
            __stcg(reinterpret_cast<uint4*>(rows) + 32, z);
        if (lane < kRingRows)
        {
            sts_v4(ring + lane * (kFastRowWords * 4) + 16u * 33, NEG4);
            sts_v4(ring + lane * (kFastRowWords * 4) + 16u * 34, NEG4);
        }
    }
    __syncwarp();

    const uint32_t gapv  = pack2(gap, gap);
    const int32_t pd_m   = match - gap;    // diagonal step in the tilted domain, bases equal
    const int32_t pd_x   = mismatch - gap; // bases differ
    const uint32_t pdx2  = pack2(pd_x, pd_x);
    const uint32_t pdmx2 = pack2(pd_m ^ pd_x, pd_m ^ pd_x);

    uint32_t rb_lo = 0, rb_hi = 0;
    uint4 prev_out      = make_uint4(0u, 0u, 0u, 0u); // cells of the row written last, its band start and left boundary
    int32_t prev_bs     = -1;
    int32_t prev_left_t = 0;
    // Descriptors travel in registers: chunk c* holds rows r0 .. r0 + 31 (one row per lane), chunk n* the next 32
    // rows, loaded a chunk ahead; band starts likewise.  The row loop receives its descriptor by shuffle, issued one
    // iteration early.
    uint32_t c0 = 0, c1 = 0, c2 = 0, n0, n1, n2;
    int32_t cbs = 0, nbs;
    {
        const int32_t rr = min(1 + lane, node_count);
        n0  = desc.d0[rr];
        n1  = desc.d1[rr];
        n2  = desc.d2[rr];
        nbs = band_start(rr, gradient, kFastBand, max_column);
    }
    uint32_t meta_next = __shfl_sync(kFull, n0, 0);
    int32_t bs_next    = __shfl_sync(kFull, nbs, 0);
    {
        for (int32_t r = 1; r <= node_count; r++)
        {
            const int32_t k = (r - 1) & 31;
            if (k == 0) // next chunk becomes current, fetch the one after
            {
                c0  = n0;
                c1  = n1;
                c2  = n2;
                cbs = nbs;
                const int32_t rr = min(r + 32 + lane, node_count);
                n0  = desc.d0[rr];
                n1  = desc.d1[rr];
                n2  = desc.d2[rr];
                nbs = band_start(rr, gradient, kFastBand, max_column);
            }
            const uint32_t meta   = meta_next;
            const int32_t bs      = bs_next;
            const int32_t pc      = (meta >> 8) & 0x3f;
            const uint32_t base   = meta & 0xff;
            const int32_t ring_lo = max(r - (kRingRows - 1), 1); // oldest row the ring still holds (never row 0)

            // profile registers Q_i = (pd(cell i), pd(cell i + 4)) from the lane's 8 read bases (positions
            // bs + 8*lane .., zero beyond the read)
            {
                const uint32_t ra = rd + bs + 8 * lane; // bs is a multiple of 4: 4-byte aligned, not always 8
                rb_lo             = lds_u32(ra);
                rb_hi             = lds_u32(ra + 4);
            }
            uint32_t Q[4];
            {
                const uint32_t base4 = base * 0x01010101u;
                const uint32_t eq_lo = __vcmpeq4(rb_lo, base4); // 0xFF per matching byte (a base is never 0)
                const uint32_t eq_hi = __vcmpeq4(rb_hi, base4);
                Q[0] = (__byte_perm(eq_lo, eq_hi, 0x4400) & pdmx2) ^ pdx2;
                Q[1] = (__byte_perm(eq_lo, eq_hi, 0x5511) & pdmx2) ^ pdx2;
                Q[2] = (__byte_perm(eq_lo, eq_hi, 0x6622) & pdmx2) ^ pdx2;
                Q[3] = (__byte_perm(eq_lo, eq_hi, 0x7733) & pdmx2) ^ pdx2;
            }

            uint32_t S[4]     = {NEG2, NEG2, NEG2, NEG2};
            int32_t min_d     = 1 << 20;   // smallest band shift over the predecessors
            int32_t left_best = INT16_MIN; // max over predecessors of their column-0 value (rows with bs == 0)

            // one predecessor row: S = max(S, diagonal and vertical candidates).  The vertical operands V_i are the
            // predecessor's cells under this lane's cells, the diagonal operands D_i = V_{i-1}, D_0 = (corner, V_3.lo).
            // All branches below are warp-uniform (band shifts and row numbers are).
            auto step = [&](int32_t prow) {
                uint32_t V0, V1, V2, V3, D0;
                int32_t d;
                if (prow == r - 1 && bs == prev_bs && r > 1)
                {
                    // the row just written, same band start: its cells are still in this lane's registers; the
                    // corner is the last cell of the lane to the left (lane 0: that row's left boundary)
                    d  = 0;
                    V0 = prev_out.x;
                    V1 = prev_out.y;
                    V2 = prev_out.z;
                    V3 = prev_out.w;
                    uint32_t X = __shfl_up_sync(kFull, prev_out.w, 1);
                    if (lane == 0)
                        X = static_cast<uint32_t>(prev_left_t) << 16;
                    D0        = __byte_perm(X, V3, 0x5432);
                    left_best = max(left_best, prev_left_t);
                }
                else if (prow >= ring_lo) // shared-memory ring
                {
                    const uint32_t pa = ring + static_cast<uint32_t>((prow & (kRingRows - 1)) * (kFastRowWords * 4));
                    const uint4 lead  = lds_v4(pa);
                    d                 = bs - static_cast<int32_t>(lead.x); // >= 0, multiple of 4
                    const int32_t su  = lane + 1 + (d >> 3);               // block of the first vertical operand
                    const uint4 a     = lds_v4(pa + 16u * min(su, kPadBlock));
                    if (d & 4)
                    {
                        // shift by an odd multiple of 4: operands straddle two blocks -> (hi of this block, lo of the
                        // next); the corner is the low half of this block's last word
                        const uint4 b = lds_v4(pa + 16u * min(su + 1, kPadBlock));
                        V0            = __byte_perm(a.x, b.x, 0x5432);
                        V1            = __byte_perm(a.y, b.y, 0x5432);
                        V2            = __byte_perm(a.z, b.z, 0x5432);
                        V3            = __byte_perm(a.w, b.w, 0x5432);
                        D0            = __byte_perm(a.w, V3, 0x5410);
                    }
                    else
                    {
                        // whole blocks: the corner is the high half of the previous block's last word (the lead
                        // block for lane 0: the predecessor's left boundary)
                        const uint32_t wv = lds_u32(pa + 16u * min(su - 1, kPadBlock) + 12u);
                        V0                = a.x;
                        V1                = a.y;
                        V2                = a.z;
                        V3                = a.w;
                        D0                = __byte_perm(wv, V3, 0x5432);
                    }
                    left_best = max(left_best, hi16(lead.w));
                }
                else // far predecessor: the global score matrix (through L2); pad blocks are implied
                {
                    const uint4* prow4 = reinterpret_cast<const uint4*>(rows + prow * kFastRowWords);
                    const uint4 lead   = __ldcg(prow4);
                    d                  = bs - static_cast<int32_t>(lead.x);
                    const int32_t sl   = lane + 1 + (d >> 3);
                    const int32_t sc   = (8 * lane + d + 7) >> 3;
                    const uint32_t p33 = (prow == 0) ? pack2(0, kNegT) : NEG2; // row 0's first pad block
                    uint4 a = NEG4, b = NEG4;
                    uint32_t wv = NEG2;
                    if (sl <= 32)
                        a = __ldcg(prow4 + sl);
                    else if (sl == 33)
                        a = make_uint4(p33, p33, p33, p33);
                    if (sl + 1 <= 32)
                        b = __ldcg(prow4 + sl + 1);
                    else if (sl + 1 == 33)
                        b = make_uint4(p33, p33, p33, p33);
                    if (sc <= 32)
                        wv = __ldcg(reinterpret_cast<const uint32_t*>(prow4 + sc) + 3);
                    else if (sc == 33)
                        wv = p33;
                    const bool odd      = (d & 4) != 0;
                    const uint32_t selV = odd ? 0x5432u : 0x3210u;
                    V0                  = __byte_perm(a.x, b.x, selV);
                    V1                  = __byte_perm(a.y, b.y, selV);
                    V2                  = __byte_perm(a.z, b.z, selV);
                    V3                  = __byte_perm(a.w, b.w, selV);
                    D0                  = __byte_perm(odd ? a.w : wv, V3, odd ? 0x5410u : 0x5432u);
                    left_best           = max(left_best, hi16(lead.w));
                }
                min_d = min(min_d, d);
                if (beta != 0 && d != 0)
                {
                    // rebased rows: the predecessor's scores are relative to its own band start, beta * d below this
                    // row's.  Sentinel-class operands are lifted first so that the subtraction cannot wrap; real ones
                    // (>= kRealLo) are above the lift for every shift at which the bands still overlap.
                    const int32_t bd  = min(beta * d, 8000);
                    const uint32_t fl = pack2(kNegT + bd, kNegT + bd);
                    const uint32_t c2 = pack2(-bd, -bd);
                    V0                = __vadd2(__vmaxs2(V0, fl), c2);
                    V1                = __vadd2(__vmaxs2(V1, fl), c2);
                    V2                = __vadd2(__vmaxs2(V2, fl), c2);
                    V3                = __vadd2(__vmaxs2(V3, fl), c2);
                    D0                = __vadd2(__vmaxs2(D0, fl), c2);
                }
                S[0]  = __viaddmax_s16x2(D0, Q[0], __viaddmax_s16x2(V0, gapv, S[0]));
                S[1]  = __viaddmax_s16x2(V0, Q[1], __viaddmax_s16x2(V1, gapv, S[1]));
                S[2]  = __viaddmax_s16x2(V1, Q[2], __viaddmax_s16x2(V2, gapv, S[2]));
                S[3]  = __viaddmax_s16x2(V2, Q[3], __viaddmax_s16x2(V3, gapv, S[3]));
            };
            step(static_cast<int32_t>(meta >> 16)); // predecessor 0 (row 0 for a node without in-edges)
            {
                // descriptor of the next row, needed first thing next iteration
                const int32_t kn = r & 31;
                meta_next        = __shfl_sync(kFull, kn ? c0 : n0, kn);
                bs_next          = __shfl_sync(kFull, kn ? cbs : nbs, kn);
            }
            if (pc > 1)
            {
                const uint32_t p12 = __shfl_sync(kFull, c1, k);
                const uint32_t p34 = (pc > 3) ? __shfl_sync(kFull, c2, k) : 0u;
                for (int32_t p = 1; p < pc; p++)
                    step(pred_row<IdT>(g, meta, p12, p34, r, p));
            }
            // every predecessor wholly left of this row's band: sentinel arithmetic would decide the row
            if (min_d > kFastBand)
                return kFastNotTaken;

            // left boundary of this row in the tilted domain
            int32_t left_t = kNegT;
            if (bs == 0)
                left_t = (pc == 0) ? gap : static_cast<int16_t>(left_best + gap);

            // horizontal pass = prefix maximum, both halves at once
            S[1] = __vmaxs2(S[1], S[0]);
            S[2] = __vmaxs2(S[2], S[1]);
            S[3] = __vmaxs2(S[3], S[2]);
            const uint32_t tl = __byte_perm(S[3], NEG2, 0x1054); // (NEG, S3.lo): cells 0..3 feed cells 4..7
            uint32_t tot      = __vmaxs2(S[3], tl);         // hi half = maximum over this lane's 8 cells
            // prefix maximum over the lanes in three rounds (radix 4, 4, 2): the shuffles of a round are independent,
            // so the chain is three shuffle latencies instead of five.  Lanes below the shift receive their own value.
            {
                const uint32_t u1 = __shfl_up_sync(kFull, tot, 1);
                const uint32_t u2 = __shfl_up_sync(kFull, tot, 2);
                const uint32_t u3 = __shfl_up_sync(kFull, tot, 3);
                tot               = __vmaxs2(__vimax3_s16x2(tot, u1, u2), u3);
                const uint32_t v1 = __shfl_up_sync(kFull, tot, 4);
                const uint32_t v2 = __shfl_up_sync(kFull, tot, 8);
                const uint32_t v3 = __shfl_up_sync(kFull, tot, 12);
                tot               = __vmaxs2(__vimax3_s16x2(tot, v1, v2), v3);
                tot               = __vmaxs2(tot, __shfl_up_sync(kFull, tot, 16));
            }
            uint32_t excl = __shfl_up_sync(kFull, tot, 1);
            if (lane == 0)
                excl = static_cast<uint32_t>(left_t) << 16;
            const uint32_t cb = __byte_perm(excl, excl, 0x3232);
            uint4 out;
            out.x = __vimax3_s16x2(S[0], cb, tl);
            out.y = __vimax3_s16x2(S[1], cb, tl);
            out.z = __vimax3_s16x2(S[2], cb, tl);
            out.w = __vimax3_s16x2(S[3], cb, tl);
            if (beta != 0)
            {
                // rebased rows carry no a-priori range: every stored cell must be sentinel class or inside
                // [kRealLo, kRealHi].  Cells grow along the row, so the lane's first and last cell decide the common case.
                const bool lane_ok = lo16(out.x) >= kRealLo && hi16(out.w) <= kRealHi;
                if (!__all_sync(kFull, lane_ok))
                {
                    auto bad = [](uint32_t reg) {
                        const int32_t lo = lo16(reg), hi = hi16(reg);
                        return (lo > kNegThreshold && lo < kRealLo) || lo > kRealHi || (hi > kNegThreshold && hi < kRealLo) ||
                               hi > kRealHi;
                    };
                    if (__any_sync(kFull, bad(out.x) || bad(out.y) || bad(out.z) || bad(out.w)))
                        return kFastNotTaken;
                }
            }
            uint4* this_row      = reinterpret_cast<uint4*>(rows + r * kFastRowWords);
            const uint32_t rrow  = ring + static_cast<uint32_t>((r & (kRingRows - 1)) * (kFastRowWords * 4));
            __stcg(this_row + 1 + lane, out);
            sts_v4(rrow + 16u + 16u * lane, out);
            if (lane == 0)
            {
                const uint4 lead = make_uint4(static_cast<uint32_t>(bs), 0u, 0u, static_cast<uint32_t>(left_t) << 16);
                __stcg(this_row, lead);
                sts_v4(rrow, lead);
            }
            prev_out    = out;
            prev_bs     = bs;
            prev_left_t = static_cast<int16_t>(left_t);
            __syncwarp();
        }
    }

    // sink selection: first strict maximum in row order over rows of nodes without out-edges, column read_len.
    // All candidates share the column, so tilted scores compare like the scores themselves.
    // (rebased rows: offsets added back; a sink row whose band misses that column cannot be ranked here)
    const int32_t best_floor = (beta == 0) ? INT16_MIN : INT32_MIN;
    int32_t best = best_floor, best_row = 0;
    bool unranked = false;
    for (int32_t r = 1 + lane; r <= node_count; r += 32)
    {
        if (desc.d0[r] & (1u << 14))
        {
            const int32_t s = tilted_score_abs(rows, r, read_len, beta);
            unranked |= (beta != 0 && s == kNegAbs);
            if (s > best)
            {
                best     = s;
                best_row = r;
            }
        }
    }
#pragma unroll
    for (int32_t k = 16; k > 0; k >>= 1)
    {
        const int32_t ob   = __shfl_xor_sync(kFull, best, k);
        const int32_t orow = __shfl_xor_sync(kFull, best_row, k);
        // keep the larger score; on ties the smaller (non-zero) row
        if (ob > best || (ob == best && orow != 0 && (best_row == 0 || orow < best_row)))
        {
            best     = ob;
            best_row = orow;
        }
    }

    if (__any_sync(kFull, unranked))
        return kFastNotTaken;
    const int32_t i0 = (best > best_floor) ? best_row : 0;
    if (tilted_score_global(rows, i0, read_len) <= kNegThreshold)
        return kFastNotTaken;
    if (score_out != nullptr) // alignment score, un-tilted (the start cell is a real score here)
        *score_out = tilted_score_abs(rows, i0, read_len, beta) + gap * read_len;
    start_row = i0;
    return 0;
}

template <typename IdT>
__device__ __forceinline__ int32_t fast_fill(const Graph<IdT>& g, int32_t node_count, const uint8_t* read,
                                             int32_t read_len, int16_t* scores, int32_t band, int32_t gap,
                                             int32_t mismatch, int32_t match, uint32_t smem, int32_t smem_bytes,
                                             int32_t lane, int32_t& start_row, int32_t& beta_out,
                                             int32_t* score_out = nullptr)
{
    const int32_t beta = fast_path_beta(node_count, read_len, band, gap, mismatch, match);
    if (beta < 0 || smem_bytes < fast_fill_smem_bytes(read_len))
        return kFastNotTaken;
    beta_out = beta;
    if (beta != 0)
        return fast_fill_impl<IdT, true>(g, node_count, read, read_len, scores, gap, mismatch, match, smem, lane,
                                         start_row, beta, score_out);
    return fast_fill_impl<IdT, false>(g, node_count, read, read_len, scores, gap, mismatch, match, smem, lane,
                                      start_row, 0, score_out);
}

// Second half: traceback (cudapoa_nw_banded.cuh:386-480) from (start_row, read_len): lane 0 walks, the warp reloads
// the tile when the walk leaves it.  `tile` is kTileRows * (kTileRowBytes + 8) bytes of shared memory, `rd` the read
// in shared memory (staged here from `read` unless the caller's fill left it there: read == nullptr).
// The walk compares tilted scores: with T = H - gap * j, "H[i][j] == H[p][j-1] + cost" reads
// "T[i][j] == T[p][j-1] + cost - gap", the vertical move "T[i][j] == T[p][j] + gap", the horizontal move
// "T[i][j] == T[i][j-1]".  This is the reference's decision as long as the cell the walk stands on is a real
// score (see tilted_score_global); it is from the first cell on, or the exact path took over.
// g.aln_node[] first receives ROW indices (0: inserted base); they become node ids in a parallel pass below.
// REBASED (beta > 0, see fast_path_beta): stored scores are relative to beta * band start of their row; the walk adds
// the offset back (kept in the first bytes of every tile row) and compares int32 values, sentinel class = kNegAbs.
// Returns the alignment length, or -1 when the walk exceeds the reference's step limit.
template <typename IdT, bool REBASED>
__device__ int32_t fast_traceback_impl(const Graph<IdT>& g, int32_t node_count, const uint8_t* read, int32_t read_len,
                                       int16_t* scores, int32_t gap, int32_t mismatch, int32_t match, uint32_t tile,
                                       uint32_t rd, int32_t start_row, int32_t lane, int32_t beta)
{
    uint32_t* rows = reinterpret_cast<uint32_t*>(scores);
#ifndef POA_EMU
    asm volatile("" : "+l"(rows)); // keep the pointer in registers: recomputing it inside the walk costs more
#endif
    RowDesc desc;
    desc.d0 = g.rowinfo;
    desc.d1 = desc.d0 + (g.M + 40);
    desc.d2 = desc.d1 + (g.M + 40);
    const int32_t max_column = read_len + 1;
    const float gradient     = band_gradient(read_len, node_count);
    const int32_t pd_m       = match - gap;
    const int32_t pd_x       = mismatch - gap;
    const uint32_t tile_desc = tile + kTileRows * kTileRowBytes;
    if (read != nullptr)
    {
        for (int32_t k = lane; k < read_len; k += 32)
            sts_u8(rd + k, read[k]);
        __syncwarp();
    }
    int32_t i = start_row, j = read_len;
    int32_t aligned = 0, loops = 0;
    int32_t prev_i = 0, prev_j = 0; // lane 0's copies carry over between tiles, as in the reference's single loop
    const int32_t limit = read_len + node_count + 2;
    while (!(i == 0 && j == 0) && loops < limit) // uniform: i, j, loops are broadcast at the end of every round
    {
        const int32_t tile_r0 = i - (kTileRows - 1);
        const int32_t tile_c0 = ((j + 3) & ~3) + 1 - kTileCols; // columns tile_c0 .. tile_c0 + 47 include j
        load_tile(rows, desc, tile, tile_desc, tile_r0, tile_c0, gradient, max_column, lane, REBASED ? beta : 0);
        if (lane == 0)
        {
            const int32_t row_floor = (tile_r0 <= 0) ? 0 : tile_r0 + 12; // leave before predecessors fall out
            // tilted score of (row, column), column within the tile's columns
            auto cell = [&](int32_t row, int32_t column) -> int32_t {
                const int32_t tr = row - tile_r0;
                if (tr >= 0)
                {
                    const int32_t raw = lds_s16(tile + static_cast<uint32_t>(tr * kTileRowBytes + 2 * (column - tile_c0 + 8)));
                    if (!REBASED)
                        return raw;
                    return (raw <= kNegThreshold) ? kNegAbs
                                                  : raw + static_cast<int32_t>(lds_u32(tile + static_cast<uint32_t>(tr * kTileRowBytes)));
                }
                return tilted_score_abs(rows, row, column, REBASED ? beta : 0);
            };
            int32_t sij = cell(i, j);
            // addresses of descriptor(row) and cell(row, column) as base + 8 * row and base + kTileRowBytes * row +
            // 2 * column (32-bit shared addresses: the arithmetic may wrap)
            const uint32_t desc_base = tile_desc - 8u * static_cast<uint32_t>(tile_r0);
            const uint32_t cell_base = tile + 2u * static_cast<uint32_t>(8 - tile_c0) -
                                       static_cast<uint32_t>(kTileRowBytes) * static_cast<uint32_t>(tile_r0);
            const uint32_t row_base = tile - static_cast<uint32_t>(kTileRowBytes) * static_cast<uint32_t>(tile_r0);
            const int32_t fast_row_floor = max(row_floor, 1);
            const int32_t fast_col_floor = max(tile_c0, 0);
            while (!(i == 0 && j == 0) && loops < limit && i >= row_floor && j > tile_c0)
            {
                // Runs of diagonal moves through predecessor 0 -- most of the walk -- in a loop of their own: the
                // first test of the general step below, with nothing else around it.  Overshooting `limit` is
                // harmless: the result is -1 as soon as loops reaches it.
                {
                    int32_t fi = i, fj = j, fs = sij;
                    IdT* out = g.aln_node + (fj - 1);
                    while (fi >= fast_row_floor && fj > fast_col_floor)
                    {
                        const uint32_t meta = lds_u32(desc_base + 8u * static_cast<uint32_t>(fi));
                        const int32_t p0    = static_cast<int32_t>(meta >> 16);
                        if (p0 < tile_r0)
                            break;
                        const int32_t cost = ((meta & 0xff) == lds_u8(rd + fj - 1)) ? pd_m : pd_x;
                        int32_t h          = lds_s16(cell_base + static_cast<uint32_t>(kTileRowBytes) * static_cast<uint32_t>(p0) +
                                            2u * static_cast<uint32_t>(fj - 1));
                        if (REBASED)
                        {
                            if (h <= kNegThreshold)
                                break;
                            h += static_cast<int32_t>(lds_u32(row_base + static_cast<uint32_t>(kTileRowBytes) * static_cast<uint32_t>(p0)));
                        }
                        if (fs != h + cost)
                            break;
                        *out-- = static_cast<IdT>(fi);
                        fi     = p0;
                        fj--;
                        fs = h;
                    }
                    const int32_t n = j - fj;
                    if (n > 0)
                    {
                        loops += n;
                        aligned += n;
                        prev_i = i = fi;
                        prev_j = j = fj;
                        sij    = fs;
                        continue;
                    }
                }
                loops++;
                bool found = false;
                int32_t h  = 0;
                if (i != 0)
                {
                    const uint2 dd      = lds_v2(tile_desc + 8u * static_cast<uint32_t>(i - tile_r0));
                    const uint32_t meta = dd.x;
                    const uint32_t p12  = dd.y;
                    const int32_t pc    = (meta >> 8) & 0x3f;
                    const int32_t p0    = static_cast<int32_t>(meta >> 16);
                    uint32_t p34        = 0;
                    if (j != 0)
                    {
                        // diagonal moves, predecessor 0 first: the common step of the walk
                        const int32_t cost = ((meta & 0xff) == lds_u8(rd + j - 1)) ? pd_m : pd_x;
                        h                  = cell(p0, j - 1);
                        if (sij == h + cost)
                        {
                            prev_i = p0;
                            prev_j = j - 1;
                            found  = true;
                        }
                        else if (pc > 1)
                        {
                            if (pc > 3)
                                p34 = desc.d2[i];
                            for (int32_t p = 1; p < pc && !found; p++)
                            {
                                const int32_t pi = pred_row<IdT>(g, meta, p12, p34, i, p);
                                h                = cell(pi, j - 1);
                                if (sij == h + cost)
                                {
                                    prev_i = pi;
                                    prev_j = j - 1;
                                    found  = true;
                                }
                            }
                        }
                    }
                    if (!found)
                    {
                        if (pc > 3)
                            p34 = desc.d2[i];
                        const int32_t np = max(pc, 1);
                        for (int32_t p = 0; p < np && !found; p++)
                        {
                            const int32_t pi = pred_row<IdT>(g, meta, p12, p34, i, p);
                            h                = cell(pi, j);
                            if (sij == h + gap)
                            {
                                prev_i = pi;
                                prev_j = j;
                                found  = true;
                            }
                        }
                    }
                }
                if (!found && j > 0)
                {
                    h = cell(i, j - 1);
                    if (sij == h)
                    {
                        prev_i = i;
                        prev_j = j - 1;
                        found  = true;
                    }
                }
                if (j != prev_j) // a read base was consumed: remember the row it pairs with (0: inserted base)
                    g.aln_node[j - 1] = static_cast<IdT>((i == prev_i) ? 0 : i);
                aligned++;
                if (found)
                    sij = h; // the cell just compared against is the next cell of the walk
                i = prev_i;
                j = prev_j;
            }
        }
        i       = __shfl_sync(kFull, i, 0);
        j       = __shfl_sync(kFull, j, 0);
        loops   = __shfl_sync(kFull, loops, 0);
        aligned = __shfl_sync(kFull, aligned, 0);
    }
    __syncwarp();
    for (int32_t k = lane; k < read_len; k += 32)
    {
        const int32_t row = g.aln_node[k];
        g.aln_node[k]     = static_cast<IdT>(row > 0 ? static_cast<int32_t>(g.sorted[row - 1]) : -1);
    }
    if (loops >= limit)
        aligned = -1;
    return aligned;
}

template <typename IdT>
__device__ __forceinline__ int32_t fast_traceback(const Graph<IdT>& g, int32_t node_count, const uint8_t* read,
                                                  int32_t read_len, int16_t* scores, int32_t gap, int32_t mismatch,
                                                  int32_t match, uint32_t tile, uint32_t rd, int32_t start_row,
                                                  int32_t lane, int32_t beta)
{
    if (beta != 0)
        return fast_traceback_impl<IdT, true>(g, node_count, read, read_len, scores, gap, mismatch, match, tile, rd,
                                              start_row, lane, beta);
    return fast_traceback_impl<IdT, false>(g, node_count, read, read_len, scores, gap, mismatch, match, tile, rd,
                                           start_row, lane, 0);
}

// Both halves back to back (the single-launch kernel): the tile reuses the ring's memory.
template <typename IdT>
__device__ int32_t align_fast_banded16(const Graph<IdT>& g, int32_t node_count, const uint8_t* read,
                                       int32_t read_len, int16_t* scores, int32_t band, int32_t gap,
                                       int32_t mismatch, int32_t match, uint32_t smem, int32_t smem_bytes,
                                       int32_t lane, int32_t* score_out = nullptr)
{
    int32_t start_row = 0, beta = 0;
    if (fast_fill<IdT>(g, node_count, read, read_len, scores, band, gap, mismatch, match, smem, smem_bytes, lane,
                       start_row, beta, score_out) == kFastNotTaken)
        return kFastNotTaken;
    // the tile reuses the ring's memory; the read stays where the fill staged it unless the tile would cover it
    if (kTileAndDescBytes <= kRingBytes)
        return fast_traceback<IdT>(g, node_count, nullptr, read_len, scores, gap, mismatch, match, smem,
                                   smem + kRingBytes, start_row, lane, beta);
    if (smem_bytes < kTileAndDescBytes + read_len)
        return kFastNotTaken;
    __syncwarp();
    return fast_traceback<IdT>(g, node_count, read, read_len, scores, gap, mismatch, match, smem,
                               smem + kTileAndDescBytes, start_row, lane, beta);
}

} // namespace poa_b200
