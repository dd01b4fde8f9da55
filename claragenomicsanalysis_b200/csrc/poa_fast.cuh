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
//   * rows are stored in that register order: lane l owns 16 bytes, a row is 32 * 16 B plus one 16 B pad block at
//     the sentinel, 528 B in all (the same footprint as the reference's band + 8 cells).  Band starts move in
//     multiples of 4 cells, so a predecessor row is read with one aligned 16 B load per lane, or two plus four
//     byte-permutes when the shift is an odd multiple of 4.
//   * cell update: two DPX instructions per register and predecessor,
//       S = __viaddmax_s16x2(D, Q, __viaddmax_s16x2(V, gap, S))   (VIADDMNMX.S16x2 on sm_100a).
//   * memory: every row goes to the window's score matrix in global memory (the traceback needs all of them), and
//     the most recent kRingRows rows also stay in a shared-memory ring, which is where predecessor rows are read
//     from (graph predecessors are almost always a few ranks back).  Per-row descriptors (base, in-degree,
//     predecessor rows, band start, left boundary) are built by a warp-parallel pre-pass into shared memory, so the
//     serial row loop and the traceback never chase graph pointers through global memory.
//   * traceback: score re-derivation as in the reference, on lane 0, out of shared-memory tiles of the score
//     matrix that the whole warp loads around the current cell.
#pragma once

#include "poa_device.cuh"

namespace poa_b200
{

constexpr int32_t kFastNotTaken = -2; // the caller must run align_exact instead
constexpr int32_t kFastBand     = 256;
constexpr int32_t kFastRowWords = 132;    // 128 cell words + 4 pad words
constexpr int32_t kNegT         = -30000; // tilted-domain stand-in for the reference's sentinel class
constexpr int32_t kNegThreshold = -25000; // anything at or below is sentinel class
constexpr int32_t kRingRows     = 16;     // power of two
constexpr int32_t kRingBytes    = kRingRows * kFastRowWords * 4;
constexpr int32_t kTileRows     = 64;     // traceback tile: rows x 16-byte blocks (8 cells each)
constexpr int32_t kTileBlocks   = 6;
static_assert(kTileRows * kTileBlocks * 16 + kTileRows * 4 <= kRingBytes, "the tile reuses the ring's memory");

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

// Real tilted scores stay inside [-20000, 30000] for this alignment; sentinel-class values stay below -25000.
__device__ __forceinline__ bool fast_path_eligible(int32_t node_count, int32_t read_len, int32_t band, int32_t gap,
                                                   int32_t mismatch, int32_t match)
{
    if (band != kFastBand || gap >= 0 || match <= 0)
        return false;
    const int32_t a = max(max(-gap, abs(mismatch)), match);
    if (a > 127)
        return false;
    if (a * (node_count + read_len + 2) > 20000)
        return false;
    if ((match - gap) * (read_len + 2) > 30000)
        return false;
    if (node_count + 1 >= 65536)
        return false;
    return true;
}

// Per-row descriptor, three 32-bit words kept as three arrays (shared memory when the graph fits, else global):
//   w0 = base | in-degree << 8 | sink << 14 | row of predecessor 0 << 16
//   w1 = band start | left-boundary value (tilted, int16) << 16
//   w2 = row of predecessor 1 | row of predecessor 2 << 16
struct RowInfo
{
    uint32_t* w0;
    uint32_t* w1;
    uint32_t* w2;
};

// Reads of the tilted score matrix with the reference's get_score semantics (cudapoa_nw_banded.cuh:106-120),
// returning un-tilted H values; sentinel-class cells come back as the reference's sentinel.  Cells are taken from
// the shared-memory tile when it holds them, else from the global score matrix.
struct FastMatrix
{
    const uint32_t* rows;   // row r at rows + r * kFastRowWords (row 0 is implicit: H = j * gap)
    RowInfo ri;
    const uint4* tile;      // [kTileRows][kTileBlocks]
    const int32_t* tile_b0; // first block held for each tile row
    int32_t tile_r0;        // first row of the tile (tile rows are tile_r0 .. tile_r0 + kTileRows - 1)
    int32_t gap;
    int32_t sentinel;

    __device__ __forceinline__ int32_t cell_t(int32_t r, int32_t c) const // c: 0-based cell of the band
    {
        const int32_t blk = c >> 3;
        const int32_t tr  = r - tile_r0;
        uint32_t w;
        if (tr >= 0 && tr < kTileRows)
        {
            const int32_t tb = blk - tile_b0[tr];
            if (tb >= 0 && tb < kTileBlocks)
            {
                w = reinterpret_cast<const uint32_t*>(tile + tr * kTileBlocks + tb)[c & 3];
                return (c & 4) ? hi16(w) : lo16(w);
            }
        }
        w = __ldcg(rows + static_cast<int64_t>(r) * kFastRowWords + (blk << 2) + (c & 3));
        return (c & 4) ? hi16(w) : lo16(w);
    }
    __device__ __forceinline__ int32_t get(int32_t r, int32_t column) const
    {
        if (r == 0)
            return (column >= 0 && column <= kFastBand) ? column * gap : sentinel;
        const uint32_t w1 = ri.w1[r];
        const int32_t b   = static_cast<int32_t>(w1 & 0xffffu);
        int32_t t;
        if (column == 0 || column == b)
            t = hi16(w1);
        else if (column > b + kFastBand || column < b)
            return sentinel;
        else
            t = cell_t(r, column - b - 1);
        return (t <= kNegThreshold) ? sentinel : t + gap * column;
    }
};

// Load the tile of the score matrix that ends at (row tile_r0 + kTileRows - 1, column j): for every row the
// kTileBlocks blocks ending with the block that holds column j.  All lanes.
__device__ __forceinline__ void load_tile(const uint32_t* rows, const RowInfo& ri, uint4* tile, int32_t* tile_b0,
                                          int32_t tile_r0, int32_t j, int32_t lane)
{
    for (int32_t tr = lane; tr < kTileRows; tr += 32)
    {
        const int32_t r = tile_r0 + tr;
        int32_t b0      = 0;
        if (r >= 1)
        {
            const int32_t b  = static_cast<int32_t>(ri.w1[r] & 0xffffu);
            const int32_t hi = min(max((j - b - 1) >> 3, 0), 32); // block of column j, clamped to the row
            b0               = max(hi - (kTileBlocks - 1), 0);
        }
        tile_b0[tr] = b0;
    }
    __syncwarp();
    for (int32_t idx = lane; idx < kTileRows * kTileBlocks; idx += 32)
    {
        const int32_t tr  = idx / kTileBlocks;
        const int32_t tb  = idx - tr * kTileBlocks;
        const int32_t r   = tile_r0 + tr;
        const int32_t blk = tile_b0[tr] + tb;
        if (r >= 1 && blk <= 32)
            tile[idx] = __ldcg(reinterpret_cast<const uint4*>(rows + static_cast<int64_t>(r) * kFastRowWords) + blk);
    }
    __syncwarp();
}

template <typename IdT>
__device__ int32_t align_fast_banded16(const Graph<IdT>& g, int32_t node_count, const uint8_t* read,
                                       int32_t read_len, int16_t* scores, int32_t band, int32_t gap,
                                       int32_t mismatch, int32_t match, uint8_t* smem, int32_t smem_bytes,
                                       int32_t lane)
{
    if (!fast_path_eligible(node_count, read_len, band, gap, mismatch, match) || smem_bytes < kRingBytes)
        return kFastNotTaken;

    uint32_t* rows = reinterpret_cast<uint32_t*>(scores);
    // shared-memory scratch: [ring of recent rows | row descriptors]; descriptors spill to global when they do not fit
    uint32_t* ring = reinterpret_cast<uint32_t*>(smem);
    RowInfo ri;
    {
        const int32_t cap = ((smem_bytes - kRingBytes) / 12) & ~3;
        if (node_count + 1 <= cap)
        {
            ri.w0 = reinterpret_cast<uint32_t*>(smem + kRingBytes);
            ri.w1 = ri.w0 + cap;
            ri.w2 = ri.w1 + cap;
        }
        else
        {
            ri.w0 = g.rowinfo;
            ri.w1 = ri.w0 + (g.M + 8);
            ri.w2 = ri.w1 + (g.M + 8);
        }
    }
    const int32_t max_column = read_len + 1;
    const float gradient     = band_gradient(read_len, node_count);
    const int32_t sentinel   = band_sentinel<int16_t>(gap, mismatch, match);

    // pre-pass, one row per lane: band start and row descriptor (independent gather chains in flight)
#pragma unroll 2
    for (int32_t r = lane; r <= node_count; r += 32)
    {
        const uint32_t bs = static_cast<uint32_t>(band_start(r, gradient, kFastBand, max_column));
        uint32_t w0 = 0, w2 = 0;
        if (r > 0)
        {
            const int32_t node = g.sorted[r - 1];
            const uint32_t pc  = g.in_cnt[node];
            const uint32_t p0  = (pc == 0) ? 0u : static_cast<uint32_t>(g.pos[g.in_edge(node, 0)] + 1);
            const uint32_t p1  = (pc > 1) ? static_cast<uint32_t>(g.pos[g.in_edge(node, 1)] + 1) : 0u;
            const uint32_t p2  = (pc > 2) ? static_cast<uint32_t>(g.pos[g.in_edge(node, 2)] + 1) : 0u;
            w0 = static_cast<uint32_t>(g.nodes[node]) | (pc << 8) | ((g.out_cnt[node] == 0) ? (1u << 14) : 0u) |
                 (p0 << 16);
            w2 = p1 | (p2 << 16);
        }
        ri.w0[r] = w0;
        ri.w1[r] = bs; // row 0: left boundary T = 0
        ri.w2[r] = w2;
    }
    __syncwarp();

    const uint32_t NEG2  = pack2(kNegT, kNegT);
    const uint32_t gapv  = pack2(gap, gap);
    const int32_t pd_m   = match - gap;    // diagonal step in the tilted domain, bases equal
    const int32_t pd_x   = mismatch - gap; // bases differ
    const uint32_t pdx2  = pack2(pd_x, pd_x);
    const uint32_t pdmx2 = pack2(pd_m ^ pd_x, pd_m ^ pd_x);

    int32_t cached_bs = -1;
    uint32_t rb_lo = 0, rb_hi = 0;
    uint32_t w0n = ri.w0[1], w1n = ri.w1[1], w2n = ri.w2[1];
    for (int32_t r = 1; r <= node_count; r++)
    {
        const uint32_t meta   = w0n;
        const int32_t bs      = static_cast<int32_t>(w1n & 0xffffu);
        const uint32_t pred12 = w2n;
        const int32_t pc      = (meta >> 8) & 0x3f;
        const uint32_t base   = meta & 0xff;
        {
            // descriptor of the next row: issued now, consumed an iteration later
            const int32_t rn = min(r + 1, node_count);
            w0n = ri.w0[rn];
            w1n = ri.w1[rn];
            w2n = ri.w2[rn];
        }

        // profile registers Q_i = (pd(cell i), pd(cell i + 4)); the lane's 8 read bases (positions bs + 8*lane ..)
        // are re-read only when the band start moves
        if (bs != cached_bs)
        {
            const int32_t rp = bs + 8 * lane;
            rb_lo = rb_hi = 0;
#pragma unroll
            for (int k = 0; k < 4; k++)
            {
                if (rp + k < read_len)
                    rb_lo |= static_cast<uint32_t>(read[rp + k]) << (8 * k);
                if (rp + 4 + k < read_len)
                    rb_hi |= static_cast<uint32_t>(read[rp + 4 + k]) << (8 * k);
            }
            cached_bs = bs;
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

        uint32_t S[4] = {NEG2, NEG2, NEG2, NEG2};
        bool any_in_band  = false;
        int32_t left_best = INT16_MIN; // max over predecessors of their column-0 value (rows with bs == 0)
        for (int32_t p = 0; p < max(pc, 1); p++)
        {
            int32_t prow;
            if (p == 0)
                prow = static_cast<int32_t>(meta >> 16);
            else if (p == 1)
                prow = static_cast<int32_t>(pred12 & 0xffffu);
            else if (p == 2)
                prow = static_cast<int32_t>(pred12 >> 16);
            else
                prow = g.pos[g.in_edge(g.sorted[r - 1], p)] + 1;
            const uint32_t pw1  = ri.w1[prow];
            const int32_t d     = bs - static_cast<int32_t>(pw1 & 0xffffu);
            const int32_t pleft = hi16(pw1);
            if (bs == 0)
                left_best = max(left_best, pleft);
            if (d > kFastBand)
                continue; // wholly outside this predecessor's band: raw sentinel, contributes nothing
            any_in_band = true;
            if (prow == 0)
            {
                // row 0 is T == 0 in every cell it has (columns 0 .. band + 7); a 4-cell chunk at cells 4q.. is
                // in band while 4q + d < band + 4 (cudapoa_nw_banded.cuh:138-143)
                const bool in_lo = (8 * lane + d) < kFastBand + 4;
                const bool in_hi = (8 * lane + 4 + d) < kFastBand + 4;
                const uint32_t keep = (in_lo ? 0x0000ffffu : 0u) | (in_hi ? 0xffff0000u : 0u);
#pragma unroll
                for (int i = 0; i < 4; i++)
                {
                    const uint32_t c = __vmaxs2(Q[i], gapv); // max(0 + pd, 0 + gap)
                    S[i]             = __vmaxs2(S[i], (c & keep) | (NEG2 & ~keep));
                }
                continue;
            }
            // the predecessor row: shared-memory ring when recent, else the global score matrix (through L2)
            const int32_t sl = lane + (d >> 3);     // block of this lane's first vertical operand
            const int32_t c  = 8 * lane + d - 1;    // cell before this lane's first one (the diagonal corner)
            uint4 a          = make_uint4(NEG2, NEG2, NEG2, NEG2);
            uint4 b          = a;
            uint32_t wv      = NEG2;
            if (r - prow < kRingRows)
            {
                const uint32_t* prev_row = ring + (prow & (kRingRows - 1)) * kFastRowWords;
                if (sl <= 32)
                    a = *reinterpret_cast<const uint4*>(prev_row + 4 * sl);
                if ((d & 4) && sl + 1 <= 32)
                    b = *reinterpret_cast<const uint4*>(prev_row + 4 * (sl + 1));
                if (c >= 0 && (c >> 3) <= 32)
                    wv = prev_row[((c >> 3) << 2) + 3];
            }
            else
            {
                const uint32_t* prev_row = rows + static_cast<int64_t>(prow) * kFastRowWords;
                if (sl <= 32)
                    a = __ldcg(reinterpret_cast<const uint4*>(prev_row + 4 * sl));
                if ((d & 4) && sl + 1 <= 32)
                    b = __ldcg(reinterpret_cast<const uint4*>(prev_row + 4 * (sl + 1)));
                if (c >= 0 && (c >> 3) <= 32)
                    wv = __ldcg(prev_row + ((c >> 3) << 2) + 3);
            }
            uint32_t V[4] = {a.x, a.y, a.z, a.w};
            if (d & 4)
            {
                V[0] = __byte_perm(a.x, b.x, 0x5432); // (hi of this block, lo of the next)
                V[1] = __byte_perm(a.y, b.y, 0x5432);
                V[2] = __byte_perm(a.z, b.z, 0x5432);
                V[3] = __byte_perm(a.w, b.w, 0x5432);
            }
            // corner value wanted in the high half of `before`
            uint32_t before = (d & 4) ? (wv << 16) : wv;
            if (c < 0)
                before = static_cast<uint32_t>(pleft) << 16;
            // diagonal operands: D_i = V_{i-1}; D_0 = (cell before this lane's first, V_3.lo)
            const uint32_t D0 = __byte_perm(before, V[3], 0x5432);
            S[0] = __viaddmax_s16x2(D0, Q[0], __viaddmax_s16x2(V[0], gapv, S[0]));
            S[1] = __viaddmax_s16x2(V[0], Q[1], __viaddmax_s16x2(V[1], gapv, S[1]));
            S[2] = __viaddmax_s16x2(V[1], Q[2], __viaddmax_s16x2(V[2], gapv, S[2]));
            S[3] = __viaddmax_s16x2(V[2], Q[3], __viaddmax_s16x2(V[3], gapv, S[3]));
        }
        if (!any_in_band)
            return kFastNotTaken; // sentinel arithmetic would decide this row: let the exact path do it

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
#pragma unroll
        for (int32_t k = 1; k < 32; k <<= 1)
            tot = __vmaxs2(tot, __shfl_up_sync(kFull, tot, k)); // lanes < k receive their own value back
        uint32_t excl = __shfl_up_sync(kFull, tot, 1);
        if (lane == 0)
            excl = static_cast<uint32_t>(left_t) << 16;
        const uint32_t cb = __byte_perm(excl, excl, 0x3232);
        uint4 out;
        out.x = __vimax3_s16x2(S[0], cb, tl);
        out.y = __vimax3_s16x2(S[1], cb, tl);
        out.z = __vimax3_s16x2(S[2], cb, tl);
        out.w = __vimax3_s16x2(S[3], cb, tl);
        uint32_t* this_row = rows + static_cast<int64_t>(r) * kFastRowWords;
        uint32_t* ring_row = ring + (r & (kRingRows - 1)) * kFastRowWords;
        __stcg(reinterpret_cast<uint4*>(this_row + 4 * lane), out);
        *reinterpret_cast<uint4*>(ring_row + 4 * lane) = out;
        if (lane < 4)
        {
            __stcg(this_row + 128 + lane, NEG2);
            ring_row[128 + lane] = NEG2;
        }
        if (lane == 0)
            ri.w1[r] = static_cast<uint32_t>(bs) | (static_cast<uint32_t>(left_t) << 16);
        __syncwarp();
    }

    // sink selection: first strict maximum in row order over rows of nodes without out-edges, column read_len
    uint4* tile      = reinterpret_cast<uint4*>(smem);
    int32_t* tile_b0 = reinterpret_cast<int32_t*>(smem + kTileRows * kTileBlocks * 16);
    FastMatrix M{rows, ri, tile, tile_b0, -(1 << 30), gap, sentinel}; // no tile yet
    int32_t best = INT16_MIN, best_row = 0;
    for (int32_t r = 1 + lane; r <= node_count; r += 32)
    {
        if (ri.w0[r] & (1u << 14))
        {
            const int32_t s = M.get(r, read_len);
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

    // traceback (cudapoa_nw_banded.cuh:386-480): lane 0 walks, the warp reloads the tile when the walk leaves it
    int32_t i = (best > INT16_MIN) ? best_row : 0, j = read_len;
    int32_t aligned = 0, loops = 0;
    int32_t prev_i = 0, prev_j = 0; // lane 0's copies carry over between tiles, as in the reference's single loop
    const int32_t limit = read_len + node_count + 2;
    while (!(i == 0 && j == 0) && loops < limit) // uniform: i, j, loops are broadcast at the end of every round
    {
        const int32_t tile_r0 = i - (kTileRows - 1);
        load_tile(rows, ri, tile, tile_b0, tile_r0, j, lane);
        M.tile_r0 = tile_r0;
        if (lane == 0)
        {
            const int32_t row_floor = (tile_r0 <= 1) ? 0 : tile_r0 + 12; // leave before predecessors fall out
            const int32_t col_floor = j - 8 * (kTileBlocks - 2);         // ... or column j - 1 does
            while (!(i == 0 && j == 0) && loops < limit && i >= row_floor && j > col_floor)
            {
                loops++;
                const int32_t sij = M.get(i, j);
                bool found        = false;
                if (i != 0)
                {
                    const uint32_t meta = ri.w0[i];
                    const uint32_t p12  = ri.w2[i];
                    const int32_t pc    = (meta >> 8) & 0x3f;
                    if (j != 0)
                    {
                        const int32_t cost = ((meta & 0xff) == read[j - 1]) ? match : mismatch;
                        for (int32_t p = 0; p < max(pc, 1) && !found; p++)
                        {
                            const int32_t pi = (p == 0)   ? static_cast<int32_t>(meta >> 16)
                                               : (p == 1) ? static_cast<int32_t>(p12 & 0xffffu)
                                               : (p == 2) ? static_cast<int32_t>(p12 >> 16)
                                                          : g.pos[g.in_edge(g.sorted[i - 1], p)] + 1;
                            if (sij == M.get(pi, j - 1) + cost)
                            {
                                prev_i = pi;
                                prev_j = j - 1;
                                found  = true;
                            }
                        }
                    }
                    for (int32_t p = 0; p < max(pc, 1) && !found; p++)
                    {
                        const int32_t pi = (p == 0)   ? static_cast<int32_t>(meta >> 16)
                                           : (p == 1) ? static_cast<int32_t>(p12 & 0xffffu)
                                           : (p == 2) ? static_cast<int32_t>(p12 >> 16)
                                                      : g.pos[g.in_edge(g.sorted[i - 1], p)] + 1;
                        if (sij == M.get(pi, j) + gap)
                        {
                            prev_i = pi;
                            prev_j = j;
                            found  = true;
                        }
                    }
                }
                if (!found && j > 0 && sij == M.get(i, j - 1) + gap)
                {
                    prev_i = i;
                    prev_j = j - 1;
                    found  = true;
                }
                if (j != prev_j) // a read base was consumed: remember the node it pairs with (-1: inserted base)
                    g.aln_node[j - 1] = static_cast<IdT>((i == prev_i) ? -1 : g.sorted[i - 1]);
                aligned++;
                i = prev_i;
                j = prev_j;
            }
        }
        i       = __shfl_sync(kFull, i, 0);
        j       = __shfl_sync(kFull, j, 0);
        loops   = __shfl_sync(kFull, loops, 0);
        aligned = __shfl_sync(kFull, aligned, 0);
    }
    if (loops >= limit)
        aligned = -1;
    return aligned;
}

} // namespace poa_b200
