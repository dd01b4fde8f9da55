// Packed (DPX s16x2) banded fill, second formulation: the fill emits a MOVE CODE per cell instead of streaming every
// score row to HBM, and the traceback (poa_trace2.cuh) follows the codes.
//
// Same alignment as align_exact<int16_t, IdT, true> / the first formulation (poa_fast.cuh, whose exactness argument,
// tilted scores, row layout and band rules it keeps -- cudapoa_nw_banded.cuh:28-174,199-206,301-358 are the reference
// rules): band 256, one warp per row, 8 cells per lane in four s16x2 registers (cell i, cell i + 4).  What changes:
//   * per cell, the move the reference's traceback would take from it (cudapoa_nw_banded.cuh:386-474: diagonal through
//     the first predecessor that explains the score, else vertical through the first such predecessor, else
//     horizontal) is decided here, where all candidates are in registers: 4 bits per cell (2-bit kind, 2-bit
//     predecessor slot), 128 B per row; rows with more than four predecessors store a byte per cell.  The score matrix
//     itself never reaches HBM: score rows live in the 8-row shared-memory ring, and only rows that a LATER row names
//     as a predecessor from beyond the ring (known from the descriptor pre-pass) are also spilled to global memory.
//   * the substitution profile comes from shared memory as MATCH BITS: per distinct base of the read (at most four,
//     else the exact path takes the read) one bit per read position, in two copies four positions apart, so that the
//     eight positions under a lane's cells are one byte; a 256-entry table of 16 bytes per block turns the byte into
//     the lane's four operand registers (the rebased fill of long windows slides a 512-position int8 profile instead).
//   * a row's descriptor is one 16-byte record {predecessor rows 0..3, profile address, band shifts and ring blocks of
//     predecessors 0 and 1, in-degree, flags, band start} built by a warp-parallel pre-pass, which also finds the rows
//     the packed arithmetic cannot serve and hands out spill slots (a side array); the row loop gets the record with
//     one broadcast shared-memory load out of a 32-row buffer that the lanes refill together every 32 rows, and
//     branches on its flags only (chain / ring / spilled predecessor, one / several / more than four predecessors).
//   * several predecessors cost one maximum and one slot update per cell: the candidates of a cell follow from the
//     element-wise maximum of the predecessors' aligned rows (see the row loop).
//   * sink selection happens as the sink rows are produced.
// Window's score area: [code rows: (M + 1) x 256 B][spill slots: 528 B each].
#pragma once

#include "poa_fast.cuh"

namespace poa_b200
{

constexpr int32_t kCodeRowBytes  = 256; // nibble codes use the first 128 B, byte codes (wide rows) all of it
constexpr int32_t kSpillRowBytes = 33 * 16; // lead block + 32 cell blocks
constexpr int32_t kWidePreds     = 4;  // rows with more predecessors than this store one byte per cell
constexpr int32_t kProfRows      = 4;  // distinct read bases the profile holds
constexpr uint32_t kRecChain     = 1u << 24; // record.z: predecessor 0 is the previous row and the band did not move
constexpr uint32_t kRecSink      = 1u << 25; //           the row's node has no out-edge
constexpr uint32_t kRecSpill     = 1u << 26; //           a later row reads this row from beyond the ring
constexpr uint32_t kRecMulti     = 1u << 27; //           more than one predecessor
constexpr uint32_t kRecWide      = 1u << 28; //           more than kWidePreds predecessors (a code byte per cell)
constexpr uint32_t kRecFar0      = 1u << 29; //           predecessor 0 / 1 / 2 is read from its spilled copy (beyond the
constexpr uint32_t kRecFar1      = 1u << 30; //           ring, or row 0); predecessor 3: bit 22
constexpr uint32_t kRecFar2      = 1u << 31;
constexpr uint32_t kRecFar3      = 1u << 22;
// record.z, low half: bits 0-10 the row's profile address; for predecessors 0 and 1 that sit in the ring no more than 8
// columns to the left (every ring predecessor unless the read is much longer than the graph) the band shift in units
// of 4 (bits 11-12, 13-14) and a flag (bit 15, bit 23) that says so: their block in the ring is precomputed in
// record.w (bits 16-23, 24-31), and the row loop neither reads the predecessor's band start nor clamps block indices.
constexpr uint32_t kRecProfMask  = 0x7ffu;
constexpr int32_t kRecShift0     = 11, kRecShift1 = 13;
constexpr uint32_t kRecFast0     = 1u << 15;
constexpr uint32_t kRecFast1     = 1u << 23;
constexpr int32_t kDescBufBytes  = 32 * 16;  // descriptor records of 32 rows
constexpr int32_t kMoveDiag      = 0;
constexpr int32_t kMoveVert      = 1;
constexpr int32_t kMoveHoriz     = 2;  // bit 1 set (3 also means horizontal)

// The plain formulation keeps the profile as MATCH BITS: per distinct base of the read one bit per read position, twice
// (copy c starts at position 4 c, so that the eight positions under a lane's cells -- band starts are multiples of 4 --
// are one byte of one of the copies), and turns a lane's byte into its four operand registers through a 256-entry
// table of 16 bytes that the block shares (kLutBytes behind the warps' scratch).  Bytes of one copy of one base:
__host__ __device__ inline int32_t fill2_bits_stride(int32_t read_len)
{
    const int32_t reach = ((read_len > kFastBand ? read_len : kFastBand) + kFastBand - 1) / kFastBand * kFastBand; // band end, full-alignment chunks included
    return (((reach + 8 + 31) / 32) * 4 + 15) & ~15;
}
constexpr int32_t kLutBytes = 256 * 16;
// Shared memory the fill needs per warp for a read of this length: the ring, the descriptor buffer and the profile.
__host__ __device__ inline int32_t fill2_smem_bytes(int32_t read_len)
{
    return kRingBytes + kDescBufBytes + kProfRows * 2 * fill2_bits_stride(read_len);
}
// The block's table: entry b = the operand registers Q_i = (pd(bit i), pd(bit i + 4)) of a lane whose eight positions
// match as the bits of b say.  Every warp of the block writes all of it (the same values: no block barrier needed)
// before its first window.
__device__ __forceinline__ void fill2_build_lut(uint32_t lut, int32_t pd_m, int32_t pd_x, int32_t lane)
{
    for (int32_t b = lane; b < 256; b += 32)
    {
        uint32_t q[4];
#pragma unroll
        for (int i = 0; i < 4; i++)
            q[i] = (static_cast<uint32_t>(((b >> i) & 1) ? pd_m : pd_x) & 0xffffu) |
                   (static_cast<uint32_t>(((b >> (i + 4)) & 1) ? pd_m : pd_x) << 16);
        sts_v4(lut + 16u * static_cast<uint32_t>(b), make_uint4(q[0], q[1], q[2], q[3]));
    }
    __syncwarp();
}
// Long windows (the REBASED instance) keep a sliding window of the profile instead: kProfWindow read positions per base,
// addressed modulo the window and refilled 256 positions at a time as the band moves right -- shared memory does not
// grow with the read, so a 6 400-base read leaves as many windows per SM as a 640-base one.
constexpr int32_t kProfWindow = 512;
__host__ __device__ inline int32_t fill2_smem_bytes_sliding()
{
    return kRingBytes + kDescBufBytes + kProfRows * kProfWindow;
}

__device__ __forceinline__ uint32_t vminu2_1(uint32_t x) // per half: 0 if zero, else 1
{
    return __vminu2(x, 0x00010001u);
}
// a * b + c as an integer multiply-add: the fill's row loop is bound by the integer ALU pipe (logic, permutes, packed
// min / max), so additions and subtractions that can be phrased this way go to the other pipe (IMAD).
__device__ __forceinline__ uint32_t mad_u32(uint32_t a, uint32_t b, uint32_t c)
{
#ifdef POA_EMU
    return a * b + c;
#else
    uint32_t d;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
#endif
}
// Per half: 1 where the two differ, else 0.
__device__ __forceinline__ uint32_t grew2(uint32_t hi, uint32_t lo)
{
    return vminu2_1(hi ^ lo);
}
// Byte codes of a row with more than kWidePreds predecessors (rare; out of line so that the row loop does not carry
// its instructions under a predicate).
POA_NOINLINE __device__ void store_wide_codes(uint8_t* at, uint32_t K0, uint32_t K1, uint32_t K2, uint32_t K3)
{
    const uint32_t t01 = __byte_perm(K0, K1, 0x6240); // c0 c1 c4 c5
    const uint32_t t23 = __byte_perm(K2, K3, 0x6240); // c2 c3 c6 c7
    __stcg(reinterpret_cast<uint2*>(at), make_uint2(__byte_perm(t01, t23, 0x5410), __byte_perm(t01, t23, 0x7632)));
}

// Row of predecessor p of row `row` from its record; slots beyond the fourth are looked up in the graph.
template <typename IdT>
__device__ __forceinline__ int32_t record_pred_row(const Graph<IdT>& g, uint32_t x, uint32_t y, int32_t row, int32_t p)
{
    if (p < 4)
    {
        const uint32_t w = (p < 2) ? x : y;
        return static_cast<int32_t>((p & 1) ? (w >> 16) : (w & 0xffffu));
    }
    return g.pos[g.in_edge(g.sorted[row - 1], p)] + 1;
}

// Returns 0 and the row the traceback starts from in start_row, or kFastNotTaken (nothing the caller needs was
// written: the exact path redoes the read).  All 32 lanes.
//
// FULL: full alignment (cudapoa_nw.cuh:41-466; no band): the matrix is cut into column chunks of 256 cells and the row
// loop runs once per chunk -- chunk c is a "band" whose start is 256 c in every row, so rows never shift against each
// other, the row above is the previous row of the same pass, and the 8-row ring and the profile serve as they are.
// What a chunk hands to the next is one value per row, the cell at its right edge (the horizontal carry and, for the
// rows below, the diagonal corner): int32 per row in g.cscore, written by lane 31, read 32 rows at a time.  Codes of
// chunk c, row r live in code row c * (node_count + 1) + r.
template <typename IdT, bool REBASED, bool FULL = false>
__device__ int32_t fill2_impl(const Graph<IdT>& g, int32_t node_count, const uint8_t* read, int32_t read_len,
                              uint8_t* area, int64_t area_bytes, int32_t max_nodes, int32_t gap, int32_t mismatch,
                              int32_t match, uint32_t smem, int32_t lane, int32_t& start_row, int32_t& start_score,
                              int32_t beta_arg, uint32_t gapv, uint32_t lut = 0u)
{
    constexpr bool BITS = !REBASED; // match-bit profile + table (plain and full); the rebased fill slides a byte profile
    const int32_t beta = REBASED ? beta_arg : 0;
    const int32_t pd_m = match - gap;    // diagonal step in the tilted domain, bases equal
    const int32_t pd_x = mismatch - gap; // bases differ
    const int32_t pstride = REBASED ? kProfWindow : 2 * fill2_bits_stride(read_len); // bytes per base in the profile
    if (pd_m > 127 || pd_m < -128 || pd_x > 127 || pd_x < -128 || kProfRows * pstride > static_cast<int32_t>(kRecProfMask) + 1)
        return kFastNotTaken;
    static_assert(!(FULL && REBASED), "full alignment is served by the plain formulation only");
    const int32_t n_chunks    = FULL ? max((read_len + kFastBand - 1) / kFastBand, 1) : 1;
    const int64_t code_bytes  = FULL ? static_cast<int64_t>(n_chunks) * (node_count + 1) * kCodeRowBytes
                                     : static_cast<int64_t>(max_nodes + 1) * kCodeRowBytes;
    const int32_t spill_slots = static_cast<int32_t>(min(static_cast<long long>(65535),
                                                         static_cast<long long>((area_bytes - code_bytes) / kSpillRowBytes)));
    if (spill_slots < 2)
        return kFastNotTaken;
    uint8_t* codes    = area;
    uint32_t* spill   = reinterpret_cast<uint32_t*>(area + code_bytes);
    uint4* records    = reinterpret_cast<uint4*>(g.rowinfo);
    uint8_t* far_mark = g.marks; // per row: a later row reads it from beyond the ring
    uint32_t* slot_of = reinterpret_cast<uint32_t*>(g.row_bs); // ... and its spill slot (the exact path's array, free here)

    const uint32_t ring = smem;
    const uint32_t dbuf = ring + kRingBytes;
    const uint32_t prof = dbuf + kDescBufBytes;
    const unsigned lt   = (1u << lane) - 1u;
    const uint32_t NEG2 = pack2(kNegT, kNegT);
    const uint4 NEG4    = make_uint4(NEG2, NEG2, NEG2, NEG2);

    // ---- the read's alphabet (at most kProfRows distinct bytes) and the profile ------------------------------------
    uint32_t alpha = 0; // byte a = base of profile row a
    int32_t nalpha = 0;
    for (int32_t k0 = 0; k0 < read_len; k0 += 32)
    {
        const int32_t k  = k0 + lane;
        const bool have  = k < read_len;
        const uint32_t b = have ? read[k] : 0u;
        for (;;)
        {
            const uint32_t valid = (nalpha >= 4) ? 0xffffffffu : ((1u << (8 * nalpha)) - 1u);
            const bool known     = (__vcmpeq4(alpha, b * 0x01010101u) & valid) != 0u;
            const unsigned fresh = __ballot_sync(kFull, have && !known);
            if (fresh == 0u)
                break;
            if (nalpha == kProfRows)
                return kFastNotTaken; // a fifth distinct base: the exact path compares bytes
            const uint32_t nb = __shfl_sync(kFull, b, __ffs(fresh) - 1);
            alpha |= nb << (8 * nalpha);
            nalpha++;
        }
    }
    const uint32_t alpha_valid = (nalpha >= 4) ? 0xffffffffu : ((1u << (8 * nalpha)) - 1u);
    // rows beyond the alphabet hold "mismatch" everywhere: the first of them serves graph bases the read does not have.
    // profile_fill(first, count): read positions first .. first + count - 1 (slot = position modulo the stride for the
    // sliding window; the whole padded read otherwise)
    auto profile_fill = [&](int32_t first, int32_t count) {
        for (int32_t k = first + lane; k < first + count; k += 32)
        {
            const uint32_t b  = (k < read_len) ? read[k] : 0u;
            const uint32_t eq = (k < read_len) ? (__vcmpeq4(alpha, b * 0x01010101u) & alpha_valid) : 0u;
            const int32_t at  = REBASED ? (k & (kProfWindow - 1)) : k;
#pragma unroll
            for (int a = 0; a < kProfRows; a++)
                sts_u8(prof + static_cast<uint32_t>(a * pstride + at), static_cast<uint32_t>(((eq >> (8 * a)) & 1u) ? pd_m : pd_x) & 0xffu);
        }
    };
    int32_t prof_loaded = pstride; // sliding window: positions below this are in the window (the last kProfWindow of them)
    if (!BITS)
    {
        profile_fill(0, pstride);
    }
    else
    {
        // copy 0: one ballot per base and 32 positions; copy 1: the same bits four positions on
        const int32_t bstride = pstride / 2;
        const int32_t nwords  = bstride / 4;
        for (int32_t w0 = 0; w0 < nwords; w0++)
        {
            const int32_t k   = 32 * w0 + lane;
            const uint32_t b  = (k < read_len) ? read[k] : 0u;
            const uint32_t eq = (k < read_len) ? (__vcmpeq4(alpha, b * 0x01010101u) & alpha_valid) : 0u;
            uint32_t mine     = 0u;
#pragma unroll
            for (int a = 0; a < kProfRows; a++)
            {
                const uint32_t m = __ballot_sync(kFull, ((eq >> (8 * a)) & 1u) != 0u);
                if (lane == a)
                    mine = m;
            }
            if (lane < kProfRows)
                sts_u32(prof + static_cast<uint32_t>(lane * pstride + 4 * w0), mine);
        }
        __syncwarp();
        for (int32_t a = 0; a < kProfRows; a++)
        {
            const uint32_t c0 = prof + static_cast<uint32_t>(a * pstride);
            for (int32_t wi = lane; wi < nwords; wi += 32)
            {
                const uint32_t lo = lds_u32(c0 + 4u * static_cast<uint32_t>(wi));
                const uint32_t hi = (wi + 1 < nwords) ? lds_u32(c0 + 4u * static_cast<uint32_t>(wi + 1)) : 0u;
                sts_u32(c0 + static_cast<uint32_t>(bstride) + 4u * static_cast<uint32_t>(wi), (lo >> 4) | (hi << 28));
            }
        }
        __syncwarp();
    }

    const int32_t max_column = read_len + 1;
    const float gradient     = band_gradient(read_len, node_count);

    // ---- pre-pass: records, far marks, spill slots ---------------------------------------------------------------------
    for (int32_t r = lane; r <= node_count; r += 32)
        far_mark[r] = 0;
    __syncwarp();
    bool foreign = false; // a graph base the read does not have, and no spare profile row for it
    bool apart   = false; // a row none of whose predecessors' bands overlaps its own: sentinel arithmetic would decide it
    {
        constexpr int32_t kPre = 4;
        for (int32_t rb = 1; rb <= node_count; rb += 32 * kPre)
        {
            int32_t node[kPre];
            uint32_t pc[kPre], zw[kPre];
            int32_t pe[kPre][4];
#pragma unroll
            for (int u = 0; u < kPre; u++)
            {
                const int32_t r = min(rb + lane + 32 * u, node_count);
                node[u]         = g.sorted[r - 1];
            }
#pragma unroll
            for (int u = 0; u < kPre; u++)
            {
                pc[u]               = g.in_cnt[node[u]];
                const uint32_t base = g.nodes[node[u]];
                const uint32_t eq   = __vcmpeq4(alpha, base * 0x01010101u) & alpha_valid;
                uint32_t a          = eq ? static_cast<uint32_t>((__ffs(static_cast<int32_t>(eq)) - 1) >> 3)
                                         : static_cast<uint32_t>(nalpha);
                if (a >= static_cast<uint32_t>(kProfRows))
                {
                    foreign = true;
                    a       = 0;
                }
                zw[u] = a * static_cast<uint32_t>(pstride) + (pc[u] << 16) | ((g.out_cnt[node[u]] == 0) ? kRecSink : 0u);
#pragma unroll
                for (int k = 0; k < 4; k++)
                    pe[u][k] = g.in_edge(node[u], k); // slots beyond the in-degree: stale but readable
            }
#pragma unroll
            for (int u = 0; u < kPre; u++)
            {
                uint32_t pr[4];
#pragma unroll
                for (int k = 0; k < 4; k++)
                    pr[k] = (static_cast<uint32_t>(k) < pc[u]) ? static_cast<uint32_t>(g.pos[pe[u][k]] + 1) : 0u;
                const int32_t r = rb + lane + 32 * u;
                if (r <= node_count)
                {
                    const int32_t bs  = FULL ? 0 : band_start(r, gradient, kFastBand, max_column);
                    const int32_t bsp = FULL ? 0 : band_start(r - 1, gradient, kFastBand, max_column);
                    // profile address of the row: its base's bits, the copy that starts at bs mod 8, the byte of cell 0
                    // (sliding byte window: bs is added at use)
                    uint32_t z        = zw[u] + (REBASED ? 0u : static_cast<uint32_t>(((bs >> 2) & 1) * (pstride / 2) + (bs >> 3)));
                    if (r > 1 && static_cast<int32_t>(pr[0]) == r - 1 && bs == bsp)
                        z |= kRecChain;
                    if (pc[u] > 1u)
                        z |= kRecMulti;
                    if (pc[u] > static_cast<uint32_t>(kWidePreds))
                        z |= kRecWide;
                    // predecessors the ring (rows r - 8 .. r - 1) will no longer hold; row 0 is always in slot 0.
                    // min_d: the smallest band shift against a predecessor (a node without in-edges reads row 0).
                    int32_t min_d = (pc[u] == 0u) ? bs : INT32_MAX;
                    uint32_t blocks = 0u; // record.w, high half: ring blocks of predecessors 0 and 1
#pragma unroll
                    for (int k = 0; k < 4; k++)
                    {
                        const bool have = static_cast<uint32_t>(k) < pc[u] || k == 0;
                        const bool far  = static_cast<int32_t>(pr[k]) < r - kRingRows || pr[k] == 0u;
                        if (have && far)
                        {
                            z |= (k == 0) ? kRecFar0 : (k == 1) ? kRecFar1 : (k == 2) ? kRecFar2 : kRecFar3;
                            if (pr[k] != 0u)
                                far_mark[pr[k]] = 1;
                        }
                        if (static_cast<uint32_t>(k) < pc[u])
                        {
                            const int32_t d = bs - (FULL ? 0 : band_start(static_cast<int32_t>(pr[k]), gradient, kFastBand, max_column));
                            min_d           = min(min_d, d);
                            if (k < 2 && !far && d >= 0 && d <= 8)
                            {
                                // block of the ring that lies under lane 0's cells: ring row, lead block, whole-block shift
                                const uint32_t blk = (pr[k] & static_cast<uint32_t>(kRingRows - 1)) * kRowBlocks + 1u + static_cast<uint32_t>(d >> 3);
                                z |= (k == 0) ? (kRecFast0 | (static_cast<uint32_t>(d >> 2) << kRecShift0))
                                              : (kRecFast1 | (static_cast<uint32_t>(d >> 2) << kRecShift1));
                                blocks |= blk << (16 + 8 * k);
                            }
                        }
                    }
                    for (uint32_t k = 4; k < pc[u]; k++)
                    {
                        const int32_t p = g.pos[g.in_edge(node[u], static_cast<int32_t>(k))] + 1;
                        if (p < r - kRingRows)
                            far_mark[p] = 1;
                        min_d = min(min_d, bs - (FULL ? 0 : band_start(p, gradient, kFastBand, max_column)));
                    }
                    apart |= min_d > kFastBand;
                    records[r] = make_uint4(pr[0] | (pr[1] << 16), pr[2] | (pr[3] << 16), z, static_cast<uint32_t>(bs) | blocks);
                }
            }
        }
    }
    if (__any_sync(kFull, foreign || apart))
        return kFastNotTaken;
    {
        int32_t used = 1; // slot 0: row 0
        if (lane == 0)
            slot_of[0] = 0u;
        for (int32_t r0 = 1; r0 <= node_count; r0 += 32)
        {
            const int32_t r     = r0 + lane;
            const bool m        = r <= node_count && far_mark[r] != 0;
            const unsigned mask = __ballot_sync(kFull, m);
            if (m)
            {
                const uint32_t slot = static_cast<uint32_t>(used + __popc(mask & lt));
                records[r].z |= kRecSpill;
                slot_of[r] = slot;
            }
            used += __popc(mask);
        }
        if (used > spill_slots)
            return kFastNotTaken;
    }
    // Row 0: T == 0 in columns 0 .. band + 3, sentinel beyond (cudapoa_nw_banded.cuh:138-143, :208-212); lead block
    // {band start 0, -, -, left boundary 0}.  It lives in spill slot 0 only: a row whose predecessor is row 0 (a node
    // without in-edges) takes the far path, which knows row 0's padding.  Ring slots carry their two pad blocks at the
    // sentinel once and for all.
    {
        const uint4 z = make_uint4(0u, 0u, 0u, 0u);
        __stcg(reinterpret_cast<uint4*>(spill) + lane, z);
        if (lane == 0)
            __stcg(reinterpret_cast<uint4*>(spill) + 32, z);
        if (lane < kRingRows)
        {
            sts_v4(ring + lane * (kFastRowWords * 4) + 16u * 33, NEG4);
            sts_v4(ring + lane * (kFastRowWords * 4) + 16u * 34, NEG4);
        }
    }
    __syncwarp();

    const uint32_t prof_l = prof + (BITS ? 1u : 8u) * static_cast<uint32_t>(lane);
    const uint32_t ring_l = ring + 16u + 16u * static_cast<uint32_t>(lane); // this lane's cell block in ring row 0
    const uint32_t ring_l0 = ring + 16u * static_cast<uint32_t>(lane);      // block `lane` of the ring

    const int32_t best_floor = REBASED ? INT32_MIN : INT16_MIN;
    int32_t best = best_floor, best_row = 0;
    bool unranked = false;
    int32_t* edge_cell = g.cscore; // FULL: per row, the cell at the right edge of the chunk before

    for (int32_t chunk = 0; chunk < n_chunks; chunk++)
    {
    const int32_t chunk_bs = chunk * kFastBand; // FULL: the band start of every row of this pass
    if (FULL && chunk > 0)
    {
        // row 0 holds T == 0 in every column: only its band start moves
        if (lane == 0)
            __stcg(reinterpret_cast<uint4*>(spill), make_uint4(static_cast<uint32_t>(chunk_bs), 0u, 0u, 0u));
        __syncwarp();
    }
    uint4 prev_out      = make_uint4(0u, 0u, 0u, 0u); // cells of the row written last and its left boundary (high half)
    uint32_t prev_left  = 0u;
    // records of the next 32 rows, one per lane, loaded a chunk ahead (FULL: and their left-edge cells)
    uint4 nrec    = records[min(1 + lane, node_count)];
    int32_t nedge = (FULL && chunk > 0) ? __ldcg(edge_cell + min(1 + lane, node_count)) : 0;
    int32_t cedge = 0;

    // codes of the row in hand: nibble rows one word per lane (byte rows: two words, at + 4 * lane from here)
    uint8_t* code_at = codes + (FULL ? static_cast<int64_t>(chunk) * (node_count + 1) + 1 : static_cast<int64_t>(1)) * kCodeRowBytes + 4 * lane;

    for (int32_t r0 = 1; r0 <= node_count; r0 += 32)
    {
        // the lanes hand over the records of rows r0 .. r0 + 31 and fetch the next 32
        sts_v4(dbuf + 16u * static_cast<uint32_t>(lane), nrec);
        __syncwarp();
        nrec = records[min(r0 + 32 + lane, node_count)];
        if (FULL && chunk > 0)
        {
            cedge = nedge;
            nedge = __ldcg(edge_cell + min(r0 + 32 + lane, node_count));
        }
        const int32_t r_end = min(r0 + 31, node_count);
        uint32_t rec_at     = dbuf;
    for (int32_t r = r0; r <= r_end; r++, rec_at += 16u, code_at += kCodeRowBytes)
    {
        const uint4 rec = lds_v4(rec_at); // broadcast
        const int32_t bs = FULL ? chunk_bs : static_cast<int32_t>(rec.w & 0xffffu);

        // profile registers Q_i = (pd(cell i), pd(cell i + 4)): read positions bs + 8 * lane ..
        uint32_t Q0, Q1, Q2, Q3;
        if (BITS)
        {
            uint32_t at = prof_l + (rec.z & kRecProfMask); // the byte of match bits under this lane's cells
            if (FULL)
                at += static_cast<uint32_t>(chunk_bs >> 3); // the records carry band start 0
            const uint4 q = lds_v4(lut + 16u * lds_u8(at));
            Q0            = q.x;
            Q1            = q.y;
            Q2            = q.z;
            Q3            = q.w;
        }
        else
        {
            while (bs + kFastBand > prof_loaded) // uniform; the band moved past what the window holds
            {
                profile_fill(prof_loaded, kFastBand);
                prof_loaded += kFastBand;
                __syncwarp();
            }
            const uint32_t row = prof + (rec.z & kRecProfMask);
            const uint32_t off = static_cast<uint32_t>(bs) + 8u * static_cast<uint32_t>(lane);
            const uint32_t lo  = lds_u32(row + (off & (kProfWindow - 1)));
            const uint32_t hi  = lds_u32(row + ((off + 4u) & (kProfWindow - 1)));
            Q0                 = prmt(lo, hi, 0xC480u);
            Q1                 = prmt(lo, hi, 0xD591u);
            Q2                 = prmt(lo, hi, 0xE6A2u);
            Q3                 = prmt(lo, hi, 0xF7B3u);
        }

        // A ring predecessor whose block and band shift the pre-pass worked out (kRecFast0 / kRecFast1): one aligned
        // 16-byte load, plus the neighbouring block (a shift by an odd multiple of 4 cells) or the word before (the corner).
        auto load_fast = [&](uint32_t blk, uint32_t shift4, uint32_t& V0, uint32_t& V1, uint32_t& V2, uint32_t& V3, uint32_t& D0) {
            POA_COUNT("fill2.pred_ring_fast");
            const uint32_t at = ring_l0 + 16u * blk;
            const uint4 a     = lds_v4(at);
            if (shift4 & 1u)
            {
                const uint4 b = lds_v4(at + 16u);
                V0            = __byte_perm(a.x, b.x, 0x5432);
                V1            = __byte_perm(a.y, b.y, 0x5432);
                V2            = __byte_perm(a.z, b.z, 0x5432);
                V3            = __byte_perm(a.w, b.w, 0x5432);
                D0            = __byte_perm(a.w, V3, 0x5410);
            }
            else
            {
                const uint32_t wv = lds_u32(at - 4u);
                V0                = a.x;
                V1                = a.y;
                V2                = a.z;
                V3                = a.w;
                D0                = __byte_perm(wv, V3, 0x5432);
            }
            if (REBASED && shift4 != 0u)
            {
                const int32_t bd   = min(beta * static_cast<int32_t>(4u * shift4), 8000);
                const uint32_t fl  = pack2(kNegT + bd, kNegT + bd);
                const uint32_t c2_ = pack2(-bd, -bd);
                V0                 = __vadd2(__vmaxs2(V0, fl), c2_);
                V1                 = __vadd2(__vmaxs2(V1, fl), c2_);
                V2                 = __vadd2(__vmaxs2(V2, fl), c2_);
                V3                 = __vadd2(__vmaxs2(V3, fl), c2_);
                D0                 = __vadd2(__vmaxs2(D0, fl), c2_);
            }
        };
        // One predecessor row from the ring or from its spilled copy: its cells under this lane's cells (V_i) and the
        // corner (D0 = (corner, V3.lo)); the diagonal operands are D_i = V_{i-1}.  All branches are warp-uniform.
        // (The corner of lane 0 is the predecessor's left boundary while the bands have not moved: the maximum over the
        // predecessors' corners is what this row's own left boundary follows from.)
        auto load_pred = [&](int32_t prow, bool far, uint32_t& V0, uint32_t& V1, uint32_t& V2, uint32_t& V3, uint32_t& D0) {
            int32_t d;
            if (!far) // shared-memory ring: rows r - 8 .. r - 1
            {
                POA_COUNT("fill2.pred_ring");
                const uint32_t pa_ = ring + static_cast<uint32_t>((prow & (kRingRows - 1)) * (kFastRowWords * 4));
                d                  = bs - static_cast<int32_t>(lds_u32(pa_)); // lead block: band start; >= 0, multiple of 4
                const int32_t su   = lane + 1 + (d >> 3);               // block of the first vertical operand
                const uint4 a      = lds_v4(pa_ + 16u * min(su, kPadBlock));
                if (d & 4)
                {
                    // shift by an odd multiple of 4: operands straddle two blocks -> (hi of this block, lo of the
                    // next); the corner is the low half of this block's last word
                    const uint4 b = lds_v4(pa_ + 16u * min(su + 1, kPadBlock));
                    V0            = __byte_perm(a.x, b.x, 0x5432);
                    V1            = __byte_perm(a.y, b.y, 0x5432);
                    V2            = __byte_perm(a.z, b.z, 0x5432);
                    V3            = __byte_perm(a.w, b.w, 0x5432);
                    D0            = __byte_perm(a.w, V3, 0x5410);
                }
                else
                {
                    // whole blocks: the corner is the high half of the previous block's last word (the lead block
                    // for lane 0: the predecessor's left boundary)
                    const uint32_t wv = lds_u32(pa_ + 16u * min(su - 1, kPadBlock) + 12u);
                    V0                = a.x;
                    V1                = a.y;
                    V2                = a.z;
                    V3                = a.w;
                    D0                = __byte_perm(wv, V3, 0x5432);
                }
            }
            else // far predecessor: its spilled copy (through L2); pad blocks are implied
            {
                POA_COUNT("fill2.pred_spilled");
                const uint32_t slot = slot_of[prow]; // row 0: slot 0
                const uint4* prow4  = reinterpret_cast<const uint4*>(spill + static_cast<int64_t>(slot) * (kSpillRowBytes / 4));
                const uint4 lead    = __ldcg(prow4);
                d                   = bs - static_cast<int32_t>(lead.x);
                const int32_t sl    = lane + 1 + (d >> 3);
                const int32_t sc    = (8 * lane + d + 7) >> 3;
                const uint32_t p33  = (prow == 0) ? pack2(0, kNegT) : NEG2; // row 0's first pad block
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
            }
            if (REBASED && d != 0)
            {
                // rebased rows: the predecessor's scores are relative to its own band start, beta * d below this
                // row's.  Sentinel-class operands are lifted first so that the subtraction cannot wrap; real ones
                // (>= kRealLo) are above the lift for every shift at which the bands still overlap.
                const int32_t bd   = min(beta * d, 8000);
                const uint32_t fl  = pack2(kNegT + bd, kNegT + bd);
                const uint32_t c2_ = pack2(-bd, -bd);
                V0                 = __vadd2(__vmaxs2(V0, fl), c2_);
                V1                 = __vadd2(__vmaxs2(V1, fl), c2_);
                V2                 = __vadd2(__vmaxs2(V2, fl), c2_);
                V3                 = __vadd2(__vmaxs2(V3, fl), c2_);
                D0                 = __vadd2(__vmaxs2(D0, fl), c2_);
            }
        };

        // S_i: best diagonal / vertical candidate per cell; K_i: its move code (kind | predecessor slot << 2) under the
        // assumption that the cell is not reached horizontally.
        //
        // Several predecessors: every predecessor p offers cell j the candidates V^p[j-1] + pd[j] (diagonal) and
        // V^p[j] + gap (vertical), V^p being its row aligned to this row's band.  Both additions are the same for every
        // p, so the best candidates follow from the element-wise maximum M = max_p V^p alone, and the slot the
        // reference's traceback would pick (the first predecessor that explains the cell, cudapoa_nw_banded.cuh:
        // 404-455) is the first p that attains M at the operand's column: one maximum and one index update per cell
        // and predecessor instead of two candidates and two updates.  (Candidates that the clamp at the sentinel
        // makes equal can get another slot than a candidate-wise comparison would give them; a walk never enters
        // sentinel cells.)
        uint32_t S0, S1, S2, S3, K0, K1, K2, K3;
        uint32_t left_hi; // this row's left boundary in the tilted domain, in the HIGH half; LANE 0 holds it, the other
                          // lanes do not use theirs
        {
            uint32_t V0, V1, V2, V3, D0;
            if (rec.z & kRecChain)
            {
                // the row just written, same band start: its cells are still in this lane's registers; the corner is
                // the last cell of the lane to the left (lane 0: that row's left boundary)
                POA_COUNT("fill2.pred_registers");
                V0 = prev_out.x;
                V1 = prev_out.y;
                V2 = prev_out.z;
                V3 = prev_out.w;
                uint32_t X = __shfl_up_sync(kFull, prev_out.w, 1);
                if (lane == 0)
                    X = prev_left;
                D0        = __byte_perm(X, V3, 0x5432);
            }
            else if (rec.z & kRecFast0)
            {
                load_fast((rec.w >> 16) & 0xffu, (rec.z >> kRecShift0) & 3u, V0, V1, V2, V3, D0);
            }
            else
            {
                load_pred(static_cast<int32_t>(rec.x & 0xffffu), (rec.z & kRecFar0) != 0u, V0, V1, V2, V3, D0); // row 0 for a source node
            }
            // candidates, vertical-or-diagonal kind (0: a diagonal move explains the cell, 1: only a vertical move does)
            // and the left boundary (a vertical move in the column left of the band while the band starts at column 0,
            // the sentinel afterwards: lane 0's corner is the best boundary among the predecessors, row 0's being 0)
            auto cells = [&]() {
                left_hi = (bs == 0) ? (__viaddmax_s16x2(D0, gapv, NEG2) << 16) : (static_cast<uint32_t>(kNegT) << 16);
                const uint32_t dq0 = __viaddmax_s16x2(D0, Q0, NEG2);
                const uint32_t dq1 = __viaddmax_s16x2(V0, Q1, NEG2);
                const uint32_t dq2 = __viaddmax_s16x2(V1, Q2, NEG2);
                const uint32_t dq3 = __viaddmax_s16x2(V2, Q3, NEG2);
                S0 = __viaddmax_s16x2(V0, gapv, dq0);
                S1 = __viaddmax_s16x2(V1, gapv, dq1);
                S2 = __viaddmax_s16x2(V2, gapv, dq2);
                S3 = __viaddmax_s16x2(V3, gapv, dq3);
                K0 = grew2(S0, dq0);
                K1 = grew2(S1, dq1);
                K2 = grew2(S2, dq2);
                K3 = grew2(S3, dq3);
            };
            if (!(rec.z & kRecMulti))
            {
                cells();
            }
            else
            {
                const int32_t pc = static_cast<int32_t>((rec.z >> 16) & 0x3fu);
                uint32_t I0, I1, I2, I3, ID; // slot of the first predecessor that attains V_i / D0
                uint32_t W0, W1, W2, W3, WD;
                {
                    // predecessor 1: the slot is just "strictly better than predecessor 0"
                    if (rec.z & kRecFast1)
                        load_fast(rec.w >> 24, (rec.z >> kRecShift1) & 3u, W0, W1, W2, W3, WD);
                    else
                        load_pred(static_cast<int32_t>(rec.x >> 16), (rec.z & kRecFar1) != 0u, W0, W1, W2, W3, WD);
                    auto second = [&](uint32_t& bestv, uint32_t& idx, uint32_t cand) {
                        const uint32_t nb = __vmaxs2(bestv, cand);
                        idx               = grew2(nb, bestv);
                        bestv             = nb;
                    };
                    second(V0, I0, W0);
                    second(V1, I1, W1);
                    second(V2, I2, W2);
                    second(V3, I3, W3);
                    second(D0, ID, WD);
                }
                for (int32_t p = 2; p < pc; p++)
                {
                    const int32_t prow = record_pred_row<IdT>(g, rec.x, rec.y, r, p);
                    load_pred(prow, prow < r - kRingRows || prow == 0, W0, W1, W2, W3, WD);
                    const uint32_t pp = pack2(p, p);
                    auto better = [&](uint32_t& bestv, uint32_t& idx, uint32_t cand) {
                        const uint32_t nb  = __vmaxs2(bestv, cand);
                        const uint32_t chg = grew2(nb, bestv) * 0xffffu; // 0xffff where cand is strictly better
                        idx                = (idx & ~chg) | (pp & chg);   // (first slot wins ties)
                        bestv              = nb;
                    };
                    better(V0, I0, W0);
                    better(V1, I1, W1);
                    better(V2, I2, W2);
                    better(V3, I3, W3);
                    better(D0, ID, WD);
                }
                cells();
                // the slot: that of the diagonal operand (the cell to the left: register i - 1, the corner register for
                // i == 0) or of the vertical operand
                auto slot = [&](uint32_t& K, uint32_t iD, uint32_t iV) {
                    const uint32_t vm = K * 0xffffu;
                    K                 = mad_u32((iD & ~vm) | (iV & vm), 4u, K);
                };
                slot(K0, ID, I0);
                slot(K1, I0, I1);
                slot(K2, I1, I2);
                slot(K3, I2, I3);
            }
        }
        // (a row every predecessor of which lies wholly left of its band was found by the pre-pass)
        if (FULL && chunk > 0)
            left_hi = static_cast<uint32_t>(__shfl_sync(kFull, cedge, r - r0)) << 16; // this row's cell at column 256 * chunk, from the pass before

        // horizontal pass = prefix maximum, both halves at once
        const uint32_t P1 = __vmaxs2(S1, S0);
        const uint32_t P2 = __vmaxs2(S2, P1);
        const uint32_t P3 = __vmaxs2(S3, P2);
        const uint32_t tl = __byte_perm(P3, NEG2, 0x1054); // (NEG, P3.lo): cells 0..3 feed cells 4..7
        uint32_t tot      = __vmaxs2(P3, tl);              // hi half = maximum over this lane's 8 cells
        // exclusive prefix maximum over the lanes (Kogge-Stone on the values shifted by one lane)
        {
            uint32_t e = __shfl_up_sync(kFull, tot, 1);
            if (lane == 0)
                e = left_hi;
            uint32_t u;
            u = __shfl_up_sync(kFull, e, 1);
            e = __vmaxs2(e, u); // lanes below the shift receive their own value: harmless for a maximum
            u = __shfl_up_sync(kFull, e, 2);
            e = __vmaxs2(e, u);
            u = __shfl_up_sync(kFull, e, 4);
            e = __vmaxs2(e, u);
            u = __shfl_up_sync(kFull, e, 8);
            e = __vmaxs2(e, u);
            u = __shfl_up_sync(kFull, e, 16);
            tot = __vmaxs2(e, u);
        }
        const uint32_t cb = __byte_perm(tot, tot, 0x3232);
        uint4 out;
        out.x = __vimax3_s16x2(S0, cb, tl);
        out.y = __vimax3_s16x2(P1, cb, tl);
        out.z = __vimax3_s16x2(P2, cb, tl);
        out.w = __vimax3_s16x2(P3, cb, tl);
        if (REBASED)
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
        // move codes: a cell whose value exceeds its own candidate was reached horizontally
        K0 = mad_u32(grew2(out.x, S0), 2u, K0);
        K1 = mad_u32(grew2(out.y, S1), 2u, K1);
        K2 = mad_u32(grew2(out.z, S2), 2u, K2);
        K3 = mad_u32(grew2(out.w, S3), 2u, K3);
        POA_COUNT(!(rec.z & kRecMulti) ? "fill2.rows_1pred" : !(rec.z & kRecWide) ? "fill2.rows_2to4pred" : "fill2.rows_wide");
        if (!(rec.z & kRecWide))
        {
            // nibble k of the lane's word = cell k (low half-word: cells 0..3, high: cells 4..7)
            const uint32_t w = mad_u32(K3, 4096u, mad_u32(K2, 256u, mad_u32(K1, 16u, K0)));
            __stcg(reinterpret_cast<uint32_t*>(code_at), w);
        }
        else
        {
            store_wide_codes(code_at + 4 * lane, K0, K1, K2, K3);
        }

        const uint32_t rrow = static_cast<uint32_t>((r & (kRingRows - 1)) * (kFastRowWords * 4));
        sts_v4(ring_l + rrow, out);
        const uint4 lead = make_uint4(static_cast<uint32_t>(bs), 0u, 0u, left_hi);
        if (lane == 0)
            sts_v4(ring + rrow, lead);
        if (rec.z & (kRecSpill | kRecSink))
        {
            if (rec.z & kRecSpill) // a later row reads this one from beyond the ring
            {
                POA_COUNT("fill2.rows_spilled");
                uint4* dst = reinterpret_cast<uint4*>(spill + static_cast<int64_t>(slot_of[r]) * (kSpillRowBytes / 4));
                __stcg(dst + 1 + lane, out);
                if (lane == 0)
                    __stcg(dst, lead);
            }
            if ((rec.z & kRecSink) && (!FULL || chunk == n_chunks - 1))
            {
                // sink row: its score at column read_len (all candidates share the column, so tilted scores compare
                // like the scores themselves); first strict maximum in row order (cudapoa_nw_banded.cuh:372-383)
                const int32_t c = read_len - bs - 1; // 0-based cell of the band
                int32_t t       = kNegT;
                if (static_cast<uint32_t>(c) < static_cast<uint32_t>(kFastBand) && read_len != 0)
                {
                    const int32_t q  = c & 3;
                    const uint32_t w = (q == 0) ? out.x : (q == 1) ? out.y : (q == 2) ? out.z : out.w;
                    const int32_t v  = (c & 4) ? hi16(w) : lo16(w);
                    t                = __shfl_sync(kFull, v, c >> 3);
                }
                else if (read_len == 0 || c == -1)
                {
                    t = hi16(__shfl_sync(kFull, left_hi, 0));
                }
                if (t <= kNegThreshold)
                    t = kNegT;
                int32_t s = t;
                if (REBASED)
                {
                    s = (t == kNegT) ? kNegAbs : t + beta * bs;
                    unranked |= (t == kNegT);
                }
                if (s > best)
                {
                    best     = s;
                    best_row = r;
                }
            }
        }
        if (FULL && chunk + 1 < n_chunks && lane == 31)
            __stcg(edge_cell + r, hi16(out.w)); // the row's cell at column 256 * (chunk + 1)
        prev_out    = out;
        prev_left   = left_hi;
        __syncwarp();
    }
    } // 32 rows
    } // chunk

    if (unranked)
        return kFastNotTaken; // a sink row whose band misses the last column cannot be ranked on rebased scores
    if (best > best_floor)
    {
        // the start cell must be a real score (the walk compares nothing, but the exact path decides sentinel rows)
        if (!REBASED && best <= kNegThreshold)
            return kFastNotTaken;
        start_row   = best_row;
        start_score = best + gap * read_len; // un-tilted: the alignment's score
    }
    else
    {
        // no sink row: the reference starts from row 0, whose cell at column read_len is inside its band up to column band
        if (read_len > kFastBand)
            return kFastNotTaken;
        start_row   = 0;
        start_score = gap * read_len;
    }
    return 0;
}

} // namespace poa_b200
