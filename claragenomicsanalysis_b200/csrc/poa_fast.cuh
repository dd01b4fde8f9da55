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
#pragma once

#include "poa_device.cuh"

namespace poa_b200
{

constexpr int32_t kFastNotTaken = -2; // the caller must run align_exact instead
constexpr int32_t kFastBand     = 256;
constexpr int32_t kFastRowWords = 132;    // 128 cell words + 4 pad words
constexpr int32_t kNegT         = -30000; // tilted-domain stand-in for the reference's sentinel class
constexpr int32_t kNegThreshold = -25000; // anything at or below is sentinel class

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
    return true;
}

// Reads of the tilted score matrix with the reference's get_score semantics (cudapoa_nw_banded.cuh:106-120),
// returning un-tilted H values; sentinel-class cells come back as the reference's sentinel.
struct FastMatrix
{
    const uint32_t* rows;  // row r at rows + r * kFastRowWords (row 0 is implicit: H = j * gap)
    const int32_t* row_bs;
    const uint16_t* left;  // tilted left-boundary value per row (as int16)
    int32_t gap;
    int32_t sentinel;

    __device__ __forceinline__ int32_t cell_t(int32_t r, int32_t c) const // c: 0-based cell of the band
    {
        const uint32_t w = rows[static_cast<int64_t>(r) * kFastRowWords + ((c >> 3) << 2) + (c & 3)];
        return (c & 4) ? hi16(w) : lo16(w);
    }
    __device__ __forceinline__ int32_t get(int32_t r, int32_t column) const
    {
        if (r == 0)
            return (column >= 0 && column <= kFastBand) ? column * gap : sentinel;
        const int32_t b = row_bs[r];
        int32_t t;
        if (column == 0 || column == b)
            t = static_cast<int16_t>(left[r]);
        else if (column > b + kFastBand || column < b)
            return sentinel;
        else
            t = cell_t(r, column - b - 1);
        return (t <= kNegThreshold) ? sentinel : t + gap * column;
    }
};

// Row descriptor packed by the pre-pass: base | in-degree << 8 | sink << 14 | first-predecessor row << 16.
__device__ __forceinline__ uint32_t make_row_meta(uint32_t base, uint32_t pc, bool sink, uint32_t pred0_row)
{
    return base | (pc << 8) | (sink ? (1u << 14) : 0u) | (pred0_row << 16);
}

template <typename IdT>
__device__ int32_t align_fast_banded16(const Graph<IdT>& g, int32_t node_count, const uint8_t* read,
                                       int32_t read_len, int16_t* scores, int32_t band, int32_t gap,
                                       int32_t mismatch, int32_t match, int32_t lane)
{
    if (!fast_path_eligible(node_count, read_len, band, gap, mismatch, match))
        return kFastNotTaken;

    uint32_t* rows       = reinterpret_cast<uint32_t*>(scores);
    uint32_t* row_meta   = reinterpret_cast<uint32_t*>(g.cscore); // free until the consensus stage
    uint16_t* left       = g.kahn;                                // free until the next topological sort
    const int32_t max_column = read_len + 1;
    const float gradient     = band_gradient(read_len, node_count);
    const int32_t sentinel   = band_sentinel<int16_t>(gap, mismatch, match);

    // pre-pass, one row per lane: band start and row descriptor (32 independent gather chains in flight)
    for (int32_t r = lane; r <= node_count; r += 32)
    {
        g.row_bs[r] = band_start(r, gradient, kFastBand, max_column);
        if (r > 0)
        {
            const int32_t node = g.sorted[r - 1];
            const uint32_t pc  = g.in_cnt[node];
            const uint32_t p0  = (pc == 0) ? 0u : static_cast<uint32_t>(g.pos[g.in_edge(node, 0)] + 1);
            row_meta[r]        = make_row_meta(g.nodes[node], pc, g.out_cnt[node] == 0, p0);
        }
    }
    if (lane == 0)
        left[0] = 0;
    __syncwarp();

    const uint32_t NEG2  = pack2(kNegT, kNegT);
    const uint32_t gapv  = pack2(gap, gap);
    const int32_t pd_m   = match - gap;    // diagonal step in the tilted domain, bases equal
    const int32_t pd_x   = mismatch - gap; // bases differ
    const uint32_t pdx2  = pack2(pd_x, pd_x);
    const uint32_t pdmx2 = pack2(pd_m ^ pd_x, pd_m ^ pd_x);

    int32_t cached_bs = -1;
    uint32_t rb_lo = 0, rb_hi = 0;
    for (int32_t r = 1; r <= node_count; r++)
    {
        const uint32_t meta = row_meta[r];
        const int32_t bs    = g.row_bs[r];
        const int32_t pc    = (meta >> 8) & 0x3f;
        const uint32_t base = meta & 0xff;

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
        bool any_in_band = false;
        int32_t left_best = INT16_MIN; // max over predecessors of their column-0 value (rows with bs == 0)
        const int32_t node = (pc > 1) ? static_cast<int32_t>(g.sorted[r - 1]) : 0;
        for (int32_t p = 0; p < max(pc, 1); p++)
        {
            const int32_t prow = (p == 0) ? static_cast<int32_t>(meta >> 16) : g.pos[g.in_edge(node, p)] + 1;
            const int32_t d    = bs - g.row_bs[prow];
            if (bs == 0)
                left_best = max(left_best, static_cast<int32_t>(static_cast<int16_t>(left[prow])));
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
            const uint32_t* prev_row = rows + static_cast<int64_t>(prow) * kFastRowWords;
            const int32_t sl         = lane + (d >> 3);
            uint32_t V[4];
            if (sl <= 32)
            {
                const uint4 a = *reinterpret_cast<const uint4*>(prev_row + 4 * sl);
                V[0] = a.x; V[1] = a.y; V[2] = a.z; V[3] = a.w;
            }
            else
            {
                V[0] = V[1] = V[2] = V[3] = NEG2;
            }
            if (d & 4)
            {
                uint32_t W[4] = {NEG2, NEG2, NEG2, NEG2};
                if (sl + 1 <= 32)
                {
                    const uint4 b = *reinterpret_cast<const uint4*>(prev_row + 4 * (sl + 1));
                    W[0] = b.x; W[1] = b.y; W[2] = b.z; W[3] = b.w;
                }
#pragma unroll
                for (int i = 0; i < 4; i++)
                    V[i] = __byte_perm(V[i], W[i], 0x5432); // (hi of this block, lo of the next)
            }
            // diagonal operands: D_i = V_{i-1}; D_0 = (cell before this lane's first, V_3.lo)
            uint32_t before = __shfl_up_sync(kFull, V[3], 1);
            if (lane == 0)
            {
                int32_t t;
                if (d == 0)
                    t = static_cast<int16_t>(left[prow]);
                else
                {
                    const int32_t c = d - 1;
                    const uint32_t w = prev_row[((c >> 3) << 2) + (c & 3)];
                    t = (c & 4) ? hi16(w) : lo16(w);
                }
                before = static_cast<uint32_t>(t) << 16;
            }
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
        *reinterpret_cast<uint4*>(this_row + 4 * lane) = out;
        if (lane < 4)
            this_row[128 + lane] = NEG2;
        if (lane == 0)
            left[r] = static_cast<uint16_t>(left_t);
        __syncwarp();
    }

    // sink selection: first strict maximum in row order over rows of nodes without out-edges, column read_len
    FastMatrix M{rows, g.row_bs, left, gap, sentinel};
    int32_t best = INT16_MIN, best_row = 0;
    for (int32_t r = 1 + lane; r <= node_count; r += 32)
    {
        if (row_meta[r] & (1u << 14))
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
        const int32_t ob = __shfl_xor_sync(kFull, best, k);
        const int32_t orow = __shfl_xor_sync(kFull, best_row, k);
        // keep the larger score; on ties the smaller (non-zero) row
        if (ob > best || (ob == best && orow != 0 && (best_row == 0 || orow < best_row)))
        {
            best     = ob;
            best_row = orow;
        }
    }

    int32_t aligned = 0;
    if (lane == 0)
    {
        int32_t i = (best > INT16_MIN) ? best_row : 0, j = read_len;
        int32_t prev_i = 0, prev_j = 0, loops = 0;
        const int32_t limit = read_len + node_count + 2;
        while (!(i == 0 && j == 0) && loops < limit)
        {
            loops++;
            const int32_t sij = M.get(i, j);
            bool found        = false;
            if (i != 0)
            {
                const uint32_t meta = row_meta[i];
                const int32_t pc    = (meta >> 8) & 0x3f;
                const int32_t node  = (pc > 1) ? static_cast<int32_t>(g.sorted[i - 1]) : 0;
                if (j != 0)
                {
                    const int32_t cost = ((meta & 0xff) == read[j - 1]) ? match : mismatch;
                    for (int32_t p = 0; p < max(pc, 1) && !found; p++)
                    {
                        const int32_t pi = (p == 0) ? static_cast<int32_t>(meta >> 16) : g.pos[g.in_edge(node, p)] + 1;
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
                    const int32_t pi = (p == 0) ? static_cast<int32_t>(meta >> 16) : g.pos[g.in_edge(node, p)] + 1;
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
        if (loops >= limit)
            aligned = -1;
    }
    return __shfl_sync(kFull, aligned, 0);
}

} // namespace poa_b200
