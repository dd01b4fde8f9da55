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
//   * rows are stored in that register order: lane l owns 16 bytes, a row is 32 * 16 B plus two 16 B pad blocks at
//     the sentinel, 544 B in all.  Band starts move in multiples of 4 cells, so a predecessor row is read with two
//     aligned 16 B loads per lane and four byte-permutes whose selector depends on the shift; block indices are
//     clamped into the pad blocks, so the row loop needs no range branches.  Row 0 (T == 0) is stored like any
//     other row, its first pad block being (0, sentinel) to reproduce the reference's band + 4 rule for it.
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
constexpr int32_t kFastRowWords = (kFastBand + kBandRowPad) / 2; // 128 cell words + 2 pad blocks of 4 words
constexpr int32_t kPadBlock     = 33;                            // last block of a row, always at the sentinel
constexpr int32_t kNegT         = -30000; // tilted-domain stand-in for the reference's sentinel class
constexpr int32_t kNegThreshold = -25000; // anything at or below is sentinel class
constexpr int32_t kRingRows     = 16;     // power of two
constexpr int32_t kRingBytes    = kRingRows * kFastRowWords * 4;
constexpr int32_t kTileRows     = 64;     // traceback tile: rows x 16-byte blocks (8 cells each)
constexpr int32_t kTileBlocks   = 6;
static_assert(kTileRows * kTileBlocks * 16 <= kRingBytes, "the tile reuses the ring's memory");

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
    if (node_count + 1 >= 32768)
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

// Row of predecessor p of the row described by (meta, p12); slots beyond the third are looked up in the graph.
template <typename IdT>
__device__ __forceinline__ int32_t pred_row(const Graph<IdT>& g, uint32_t meta, uint32_t p12, int32_t row, int32_t p)
{
    if (p == 0)
        return static_cast<int32_t>(meta >> 16);
    if (p == 1)
        return static_cast<int32_t>(p12 & 0xffffu);
    if (p == 2)
        return static_cast<int32_t>(p12 >> 16);
    return g.pos[g.in_edge(g.sorted[row - 1], p)] + 1;
}

// Reads of the tilted score matrix with the reference's get_score semantics (cudapoa_nw_banded.cuh:106-120),
// returning un-tilted H values; sentinel-class cells come back as the reference's sentinel.  Cells are taken from
// the shared-memory tile when it holds them, else from the global score matrix.  The tile holds, for each of its
// rows, the kTileBlocks blocks that end with the block of column tile_j.
struct FastMatrix
{
    const uint32_t* rows; // row r at rows + r * kFastRowWords
    const uint32_t* w1;   // RowInfo::w1
    const uint32_t* tile; // [kTileRows][kTileBlocks][4]
    int32_t tile_r0;      // first row of the tile
    int32_t tile_j;
    int32_t gap;
    int32_t sentinel;

    static __device__ __forceinline__ int32_t first_block(int32_t tile_j, int32_t b)
    {
        const int32_t hi = min(max((tile_j - b - 1) >> 3, 0), 31); // block of column tile_j, clamped to the row
        return max(hi - (kTileBlocks - 1), 0);
    }
    __device__ __forceinline__ int32_t get(int32_t r, int32_t column) const
    {
        const uint32_t d = w1[r];
        const int32_t b  = static_cast<int32_t>(d & 0xffffu);
        const int32_t c  = column - b - 1; // 0-based cell of the band
        int32_t t;
        if (column == 0 || c == -1)
        {
            t = hi16(d);
        }
        else if (static_cast<uint32_t>(c) >= static_cast<uint32_t>(kFastBand))
        {
            return sentinel;
        }
        else
        {
            const int32_t blk = c >> 3;
            const int32_t tr  = r - tile_r0;
            const int32_t tb  = blk - first_block(tile_j, b);
            uint32_t w;
            if (tr >= 0 && static_cast<uint32_t>(tb) < static_cast<uint32_t>(kTileBlocks))
                w = tile[((tr * kTileBlocks + tb) << 2) + (c & 3)];
            else
                w = __ldcg(rows + r * kFastRowWords + (blk << 2) + (c & 3));
            t = (c & 4) ? hi16(w) : lo16(w);
        }
        return (t <= kNegThreshold) ? sentinel : t + gap * column;
    }
};

// Load the tile that ends at (row tile_r0 + kTileRows - 1, column j).  All lanes.
__device__ __forceinline__ void load_tile(const uint32_t* rows, const uint32_t* w1, uint4* tile, int32_t tile_r0,
                                          int32_t j, int32_t lane)
{
#pragma unroll 4
    for (int32_t idx = lane; idx < kTileRows * kTileBlocks; idx += 32)
    {
        const int32_t tr = idx / kTileBlocks;
        const int32_t tb = idx - tr * kTileBlocks;
        const int32_t r  = tile_r0 + tr;
        if (r >= 0)
        {
            const int32_t b   = static_cast<int32_t>(w1[r] & 0xffffu);
            const int32_t blk = FastMatrix::first_block(j, b) + tb;
            tile[idx] = __ldcg(reinterpret_cast<const uint4*>(rows + r * kFastRowWords) + blk);
        }
    }
    __syncwarp();
}

template <typename IdT, bool RI_SMEM>
__device__ int32_t align_fast_impl(const Graph<IdT>& g, int32_t node_count, int32_t read_len, uint32_t* rows,
                                   int32_t gap, int32_t mismatch, int32_t match, uint32_t* ring, const uint8_t* rd,
                                   RowInfo ri, int32_t lane)
{
    const int32_t max_column = read_len + 1;
    const float gradient     = band_gradient(read_len, node_count);
    const int32_t sentinel   = band_sentinel<int16_t>(gap, mismatch, match);
    const uint32_t NEG2      = pack2(kNegT, kNegT);

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
        ri.w1[r] = bs; // row 0: band start 0, left boundary T = 0
        ri.w2[r] = w2;
    }
    // row 0: T == 0 in columns 0 .. band + 3, sentinel beyond (cudapoa_nw_banded.cuh:138-143, :208-212)
    {
        const uint4 z = make_uint4(0u, 0u, 0u, 0u);
        __stcg(reinterpret_cast<uint4*>(rows) + lane, z);
        reinterpret_cast<uint4*>(ring)[lane] = z;
        if (lane < 8)
        {
            const uint32_t v = (lane < 4) ? pack2(0, kNegT) : NEG2;
            __stcg(rows + 128 + lane, v);
            ring[128 + lane] = v;
        }
    }
    __syncwarp();

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

        // profile registers Q_i = (pd(cell i), pd(cell i + 4)); the lane's 8 read bases (positions bs + 8*lane ..,
        // zero beyond the read) are re-read only when the band start moves
        if (bs != cached_bs)
        {
            const uint32_t* src = reinterpret_cast<const uint32_t*>(rd + bs + 8 * lane);
            rb_lo     = src[0];
            rb_hi     = src[1];
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

        uint32_t S[4]     = {NEG2, NEG2, NEG2, NEG2};
        bool any_in_band  = false;
        int32_t left_best = INT16_MIN; // max over predecessors of their column-0 value (used by rows with bs == 0)
        const int32_t np  = max(pc, 1);
        for (int32_t p = 0; p < np; p++)
        {
            const int32_t prow  = pred_row<IdT>(g, meta, pred12, r, p);
            const uint32_t pw1  = ri.w1[prow];
            const int32_t d     = bs - static_cast<int32_t>(pw1 & 0xffffu); // >= 0, multiple of 4
            const int32_t pleft = hi16(pw1);
            left_best           = max(left_best, pleft);
            // d > band: wholly outside this predecessor's band; every operand below then comes from the pad block
            any_in_band |= (d <= kFastBand);
            const int32_t c  = 8 * lane + d - 1;                  // cell before this lane's first one (the corner)
            const int32_t sl = min(lane + (d >> 3), kPadBlock);   // block of this lane's first vertical operand
            const int32_t sn = min(sl + 1, kPadBlock);
            const int32_t sc = min(max(c >> 3, 0), kPadBlock);
            uint4 a, b;
            uint32_t wv;
            if (r - prow < kRingRows)
            {
                const uint32_t* prev_row = ring + (prow & (kRingRows - 1)) * kFastRowWords;
                a  = *reinterpret_cast<const uint4*>(prev_row + 4 * sl);
                b  = *reinterpret_cast<const uint4*>(prev_row + 4 * sn);
                wv = prev_row[4 * sc + 3];
            }
            else
            {
                const uint32_t* prev_row = rows + prow * kFastRowWords;
                a  = __ldcg(reinterpret_cast<const uint4*>(prev_row + 4 * sl));
                b  = __ldcg(reinterpret_cast<const uint4*>(prev_row + 4 * sn));
                wv = __ldcg(prev_row + 4 * sc + 3);
            }
            // shift by an odd multiple of 4: operands straddle two blocks -> (hi of this block, lo of the next);
            // the corner is then the low half of this block's last word, else the high half of the previous block's
            const bool odd      = (d & 4) != 0;
            const uint32_t selV = odd ? 0x5432u : 0x3210u;
            const uint32_t selD = odd ? 0x5410u : 0x5432u;
            uint32_t X          = odd ? a.w : wv;
            if (c < 0)
                X = static_cast<uint32_t>(pleft) << 16; // lane 0, no shift: the predecessor's left boundary
            const uint32_t V0 = __byte_perm(a.x, b.x, selV);
            const uint32_t V1 = __byte_perm(a.y, b.y, selV);
            const uint32_t V2 = __byte_perm(a.z, b.z, selV);
            const uint32_t V3 = __byte_perm(a.w, b.w, selV);
            // diagonal operands: D_i = V_{i-1}; D_0 = (corner, V_3.lo)
            const uint32_t D0 = __byte_perm(X, V3, selD);
            S[0] = __viaddmax_s16x2(D0, Q[0], __viaddmax_s16x2(V0, gapv, S[0]));
            S[1] = __viaddmax_s16x2(V0, Q[1], __viaddmax_s16x2(V1, gapv, S[1]));
            S[2] = __viaddmax_s16x2(V1, Q[2], __viaddmax_s16x2(V2, gapv, S[2]));
            S[3] = __viaddmax_s16x2(V2, Q[3], __viaddmax_s16x2(V3, gapv, S[3]));
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
        uint32_t* this_row = rows + r * kFastRowWords;
        uint32_t* ring_row = ring + (r & (kRingRows - 1)) * kFastRowWords;
        __stcg(reinterpret_cast<uint4*>(this_row) + lane, out);
        reinterpret_cast<uint4*>(ring_row)[lane] = out;
        if (lane < 8)
        {
            __stcg(this_row + 128 + lane, NEG2);
            ring_row[128 + lane] = NEG2;
        }
        if (lane == 0)
            ri.w1[r] = static_cast<uint32_t>(bs) | (static_cast<uint32_t>(left_t) << 16);
        __syncwarp();
    }

    // sink selection: first strict maximum in row order over rows of nodes without out-edges, column read_len
    uint32_t* tile = ring;
    FastMatrix M{rows, ri.w1, tile, 1 << 30, 0, gap, sentinel}; // no tile yet: every cell comes from global memory
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

    // traceback (cudapoa_nw_banded.cuh:386-480): lane 0 walks, the warp reloads the tile when the walk leaves it.
    // g.aln_node[] first receives ROW indices (0: inserted base); they become node ids in a parallel pass below.
    int32_t i = (best > INT16_MIN) ? best_row : 0, j = read_len;
    int32_t aligned = 0, loops = 0;
    int32_t prev_i = 0, prev_j = 0; // lane 0's copies carry over between tiles, as in the reference's single loop
    const int32_t limit = read_len + node_count + 2;
    while (!(i == 0 && j == 0) && loops < limit) // uniform: i, j, loops are broadcast at the end of every round
    {
        const int32_t tile_r0 = i - (kTileRows - 1);
        load_tile(rows, ri.w1, reinterpret_cast<uint4*>(tile), tile_r0, j, lane);
        M.tile_r0 = tile_r0;
        M.tile_j  = j;
        if (lane == 0)
        {
            const int32_t row_floor = (tile_r0 <= 0) ? 0 : tile_r0 + 12; // leave before predecessors fall out
            const int32_t col_floor = j - 8 * (kTileBlocks - 2);         // ... or column j - 1 does
            int32_t sij             = M.get(i, j);
            while (!(i == 0 && j == 0) && loops < limit && i >= row_floor && j > col_floor)
            {
                loops++;
                bool found = false;
                int32_t h  = 0;
                if (i != 0)
                {
                    const uint32_t meta = ri.w0[i];
                    const uint32_t p12  = ri.w2[i];
                    const int32_t np    = max(static_cast<int32_t>((meta >> 8) & 0x3f), 1);
                    if (j != 0)
                    {
                        const int32_t cost = ((meta & 0xff) == rd[j - 1]) ? match : mismatch;
                        for (int32_t p = 0; p < np && !found; p++)
                        {
                            const int32_t pi = pred_row<IdT>(g, meta, p12, i, p);
                            h                = M.get(pi, j - 1);
                            if (sij == h + cost)
                            {
                                prev_i = pi;
                                prev_j = j - 1;
                                found  = true;
                            }
                        }
                    }
                    for (int32_t p = 0; p < np && !found; p++)
                    {
                        const int32_t pi = pred_row<IdT>(g, meta, p12, i, p);
                        h                = M.get(pi, j);
                        if (sij == h + gap)
                        {
                            prev_i = pi;
                            prev_j = j;
                            found  = true;
                        }
                    }
                }
                if (!found && j > 0)
                {
                    h = M.get(i, j - 1);
                    if (sij == h + gap)
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
__device__ int32_t align_fast_banded16(const Graph<IdT>& g, int32_t node_count, const uint8_t* read,
                                       int32_t read_len, int16_t* scores, int32_t band, int32_t gap,
                                       int32_t mismatch, int32_t match, uint8_t* smem, int32_t smem_bytes,
                                       int32_t lane)
{
    // shared-memory scratch: [ring of recent rows | read, zero padded | row descriptors]
    const int32_t rd_bytes = (read_len + kFastBand + 8 + 15) & ~15;
    if (!fast_path_eligible(node_count, read_len, band, gap, mismatch, match) || smem_bytes < kRingBytes + rd_bytes)
        return kFastNotTaken;
    uint32_t* ring = reinterpret_cast<uint32_t*>(smem);
    uint8_t* rd    = smem + kRingBytes;
    for (int32_t k = lane; k < rd_bytes; k += 32)
        rd[k] = (k < read_len) ? read[k] : static_cast<uint8_t>(0);
    const int32_t cap = ((smem_bytes - kRingBytes - rd_bytes) / 12) & ~3;
    RowInfo ri;
    if (node_count + 1 <= cap)
    {
        ri.w0 = reinterpret_cast<uint32_t*>(smem + kRingBytes + rd_bytes);
        ri.w1 = ri.w0 + cap;
        ri.w2 = ri.w1 + cap;
        return align_fast_impl<IdT, true>(g, node_count, read_len, reinterpret_cast<uint32_t*>(scores), gap, mismatch,
                                          match, ring, rd, ri, lane);
    }
    ri.w0 = g.rowinfo; // descriptors do not fit: keep them in global memory
    ri.w1 = ri.w0 + (g.M + 8);
    ri.w2 = ri.w1 + (g.M + 8);
    return align_fast_impl<IdT, false>(g, node_count, read_len, reinterpret_cast<uint32_t*>(scores), gap, mismatch,
                                       match, ring, rd, ri, lane);
}

} // namespace poa_b200
