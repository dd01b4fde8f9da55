// Device functions of the B200 POA engine: exact-arithmetic alignment (all modes), graph update,
// topological sorts, consensus and MSA.  One warp owns one window.
//
// "Exact" = same stored values, comparisons and tie-breaks as the reference kernels
// (cudapoa_nw.cuh, cudapoa_nw_banded.cuh, cudapoa_add_alignment.cuh, cudapoa_topsort.cuh,
// cudapoa_generate_consensus.cuh, cudapoa_generate_msa.cuh), re-formulated: slot-major graph arrays,
// prefix-max scans instead of the data-dependent fix-point loop, band geometry cached per row,
// per-read node paths instead of per-edge read lists for the MSA.
#pragma once

#include "poa_layout.hpp"

#include <cuda_runtime.h>
#include <stdint.h>

// Kernel launch: `kernel` is a function pointer or the name of a non-template kernel.
#ifdef POA_EMU
#define POA_LAUNCH(kernel, grid, block, smem, stream, ...) POA_EMU_LAUNCH(kernel, grid, block, smem, stream, __VA_ARGS__)
#else
#define POA_LAUNCH(kernel, grid, block, smem, stream, ...) kernel<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__)
#endif

namespace poa_b200
{

constexpr unsigned kFull = 0xffffffffu;

// ---- explicit shared-memory accessors -----------------------------------------------------------------------
// The per-warp scratch is reached through a pointer computed at run time, which the compiler treats as a generic
// 64-bit address; these wrappers keep shared-memory traffic on 32-bit shared addresses (LDS/STS).
#ifdef POA_EMU
// Emulator build (tests/emu, test infrastructure): shared addresses are bounds- and alignment-checked offsets into the
// emulated block's shared memory.
#define POA_DYNAMIC_SMEM(name) uint8_t* name = poa_emu::smem_base()
#define POA_COUNT(name) poa_emu::count(name)
#define POA_NOINLINE __attribute__((noinline))
#define POA_COUNT_N(name, n) poa_emu::count(name, n)
template <typename T>
inline T* emu_shared(uint32_t a, int line)
{
    const uint32_t off = a - 0x100u;
    if (a < 0x100u || off + sizeof(T) > poa_emu::smem_size() || (a % sizeof(T)) != 0)
        poa_emu::fail("shared-memory access out of bounds or misaligned", line);
    return reinterpret_cast<T*>(poa_emu::smem_base() + off);
}
inline uint32_t smem_u32(const void* p)
{
    return static_cast<uint32_t>(static_cast<const uint8_t*>(p) - poa_emu::smem_base()) + 0x100u;
}
inline uint32_t lds_u32(uint32_t a, int line = __builtin_LINE()) { return *emu_shared<uint32_t>(a, line); }
inline uint32_t lds_u16(uint32_t a, int line = __builtin_LINE()) { return *emu_shared<uint16_t>(a, line); }
inline uint32_t lds_u8(uint32_t a, int line = __builtin_LINE()) { return *emu_shared<uint8_t>(a, line); }
inline int32_t lds_s16(uint32_t a, int line = __builtin_LINE()) { return *emu_shared<int16_t>(a, line); }
inline uint2 lds_v2(uint32_t a, int line = __builtin_LINE()) { return *emu_shared<uint2>(a, line); }
inline uint4 lds_v4(uint32_t a, int line = __builtin_LINE()) { return *emu_shared<uint4>(a, line); }
inline void sts_u32(uint32_t a, uint32_t v, int line = __builtin_LINE()) { *emu_shared<uint32_t>(a, line) = v; }
inline void sts_u16(uint32_t a, uint32_t v, int line = __builtin_LINE())
{
    *emu_shared<uint16_t>(a, line) = static_cast<uint16_t>(v);
}
inline void sts_u8(uint32_t a, uint32_t v, int line = __builtin_LINE())
{
    *emu_shared<uint8_t>(a, line) = static_cast<uint8_t>(v);
}
inline void sts_u16_if(uint32_t a, uint32_t v, bool on, int line = __builtin_LINE())
{
    if (on)
        sts_u16(a, v, line);
}
inline void sts_u8_if(uint32_t a, uint32_t v, bool on, int line = __builtin_LINE())
{
    if (on)
        sts_u8(a, v, line);
}
inline void sts_v2(uint32_t a, uint2 v, int line = __builtin_LINE()) { *emu_shared<uint2>(a, line) = v; }
inline void sts_v4(uint32_t a, uint4 v, int line = __builtin_LINE()) { *emu_shared<uint4>(a, line) = v; }
// prmt.b32 with the sign-replication bit of the selector nibbles honoured
inline uint32_t prmt(uint32_t x, uint32_t y, uint32_t s) { return poa_emu_prmt(x, y, s); }
#else
#define POA_DYNAMIC_SMEM(name) extern __shared__ __align__(16) uint8_t name[]
#define POA_COUNT(name)
#define POA_COUNT_N(name, n)
#define POA_NOINLINE __noinline__
__device__ __forceinline__ uint32_t smem_u32(const void* p)
{
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ uint32_t lds_u32(uint32_t a)
{
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a));
    return v;
}
__device__ __forceinline__ uint32_t lds_u16(uint32_t a)
{
    uint16_t v;
    asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(a));
    return v;
}
__device__ __forceinline__ uint32_t lds_u8(uint32_t a)
{
    uint32_t v;
    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(a));
    return v;
}
__device__ __forceinline__ uint2 lds_v2(uint32_t a)
{
    uint2 v;
    asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(a));
    return v;
}
__device__ __forceinline__ uint4 lds_v4(uint32_t a)
{
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a));
    return v;
}
__device__ __forceinline__ void sts_u32(uint32_t a, uint32_t v)
{
    asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(v));
}
__device__ __forceinline__ void sts_u16(uint32_t a, uint32_t v)
{
    asm volatile("st.shared.u16 [%0], %1;" ::"r"(a), "h"(static_cast<uint16_t>(v)));
}
__device__ __forceinline__ void sts_u8(uint32_t a, uint32_t v)
{
    asm volatile("st.shared.u8 [%0], %1;" ::"r"(a), "r"(v));
}
__device__ __forceinline__ int32_t lds_s16(uint32_t a)
{
    int32_t v;
    asm volatile("{\n\t.reg .s16 t;\n\tld.shared.s16 t, [%1];\n\tcvt.s32.s16 %0, t;\n\t}" : "=r"(v) : "r"(a));
    return v;
}
// predicated stores: straight-line code where a branch would cost more than the store
__device__ __forceinline__ void sts_u16_if(uint32_t a, uint32_t v, bool on)
{
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %2, 0;\n\t@p st.shared.u16 [%0], %1;\n\t}" ::"r"(a),
                 "h"(static_cast<uint16_t>(v)), "r"(static_cast<uint32_t>(on)));
}
__device__ __forceinline__ void sts_u8_if(uint32_t a, uint32_t v, bool on)
{
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %2, 0;\n\t@p st.shared.u8 [%0], %1;\n\t}" ::"r"(a), "r"(v),
                 "r"(static_cast<uint32_t>(on)));
}
__device__ __forceinline__ void sts_v2(uint32_t a, uint2 v)
{
    asm volatile("st.shared.v2.u32 [%0], {%1, %2};" ::"r"(a), "r"(v.x), "r"(v.y));
}
__device__ __forceinline__ void sts_v4(uint32_t a, uint4 v)
{
    asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w));
}
__device__ __forceinline__ uint32_t prmt(uint32_t x, uint32_t y, uint32_t s)
{
    uint32_t r;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(x), "r"(y), "r"(s));
    return r;
}
#endif // POA_EMU

enum : int32_t
{
    kStSuccess           = 0,
    kStSeqSizeExceeded   = 2,
    kStNodeCountExceeded = 4,
    kStEdgeCountExceeded = 5,
    kStLoopCountExceeded = 7,
    kStGenericError      = 9
};

// Per-window view of the arena (pointers into global memory).
template <typename IdT>
struct Graph
{
    uint8_t* nodes;
    uint16_t *in_cnt, *out_cnt, *aln_cnt, *cov, *kahn;
    IdT *sorted, *pos;
    IdT *in_edges, *out_edges, *aligned;
    uint16_t* in_w;
    IdT* aln_node; // per read position: graph node the base pairs with, or -1 (the alignment, see align_exact)
    int32_t* row_bs;
    uint32_t* rowinfo;
    int32_t* cscore;
    IdT* cpred;
    uint8_t *marks, *check;
    IdT* msa_pos;
    IdT* stack;
    IdT* path;
    int32_t M;

    __device__ __forceinline__ IdT& in_edge(int32_t node, int32_t slot) const { return in_edges[slot * M + node]; }
    __device__ __forceinline__ IdT& out_edge(int32_t node, int32_t slot) const { return out_edges[slot * M + node]; }
    __device__ __forceinline__ IdT& aln(int32_t node, int32_t slot) const { return aligned[slot * M + node]; }
    __device__ __forceinline__ uint16_t& weight(int32_t node, int32_t slot) const { return in_w[slot * M + node]; }
};

template <typename IdT>
__device__ __forceinline__ Graph<IdT> make_graph(uint8_t* a, const ArenaLayout& L)
{
    Graph<IdT> g;
    g.nodes     = a + L.nodes;
    g.in_cnt    = reinterpret_cast<uint16_t*>(a + L.in_cnt);
    g.out_cnt   = reinterpret_cast<uint16_t*>(a + L.out_cnt);
    g.aln_cnt   = reinterpret_cast<uint16_t*>(a + L.aln_cnt);
    g.cov       = reinterpret_cast<uint16_t*>(a + L.cov);
    g.kahn      = reinterpret_cast<uint16_t*>(a + L.kahn);
    g.sorted    = reinterpret_cast<IdT*>(a + L.sorted);
    g.pos       = reinterpret_cast<IdT*>(a + L.pos);
    g.in_edges  = reinterpret_cast<IdT*>(a + L.in_edges);
    g.out_edges = reinterpret_cast<IdT*>(a + L.out_edges);
    g.aligned   = reinterpret_cast<IdT*>(a + L.aligned);
    g.in_w      = reinterpret_cast<uint16_t*>(a + L.in_w);
    g.aln_node  = reinterpret_cast<IdT*>(a + L.aln_node);
    g.row_bs    = reinterpret_cast<int32_t*>(a + L.row_bs);
    g.rowinfo   = reinterpret_cast<uint32_t*>(a + L.rowinfo);
    g.cscore    = reinterpret_cast<int32_t*>(a + L.cscore);
    g.cpred     = reinterpret_cast<IdT*>(a + L.cpred);
    g.marks     = a + L.marks;
    g.check     = a + L.check;
    g.msa_pos   = reinterpret_cast<IdT*>(a + L.msa_pos);
    g.stack     = reinterpret_cast<IdT*>(a + L.stack);
    g.path      = reinterpret_cast<IdT*>(a + L.path);
    g.M         = L.M;
    return g;
}

// ---- score arithmetic with the reference's integer semantics -------------------------------------------
// int16 scores: sums are formed in int and truncated when stored; int32 scores: sums wrap.
template <typename ScoreT>
struct ScoreOps;

template <>
struct ScoreOps<int16_t>
{
    static __device__ __forceinline__ int32_t add(int16_t a, int32_t b) { return static_cast<int32_t>(a) + b; }
    static __device__ __forceinline__ int16_t narrow(int32_t v) { return static_cast<int16_t>(v); }
    static __device__ __forceinline__ int16_t type_min() { return INT16_MIN; }
};

template <>
struct ScoreOps<int32_t>
{
    static __device__ __forceinline__ int32_t add(int32_t a, int32_t b)
    {
        return static_cast<int32_t>(static_cast<uint32_t>(a) + static_cast<uint32_t>(b));
    }
    static __device__ __forceinline__ int32_t narrow(int32_t v) { return v; }
    static __device__ __forceinline__ int32_t type_min() { return INT32_MIN; }
};

// Band sentinel, cudapoa_nw_banded.cuh:199-200.
template <typename ScoreT>
__device__ __forceinline__ ScoreT band_sentinel(int32_t gap, int32_t mismatch, int32_t match)
{
    int32_t m = min(min(gap, mismatch), -match) - 1;
    return ScoreOps<ScoreT>::narrow(ScoreOps<ScoreT>::add(ScoreOps<ScoreT>::type_min(), 2 * abs(m)));
}

// Band gradient exactly as the reference's fast-math build computes it (cudapoa_nw_banded.cuh:206):
// FMUL.FTZ(float(L + 1), MUFU.RCP(float(N + 1))).
__device__ __forceinline__ float band_gradient(int32_t read_len, int32_t node_count)
{
    float r;
    float d = static_cast<float>(node_count + 1);
    float q;
    float n = static_cast<float>(read_len + 1);
#ifdef POA_EMU
    r = poa_emu::rcp_approx(d);
    q = n * r;
#else
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(d));
    asm("mul.ftz.f32 %0, %1, %2;" : "=f"(q) : "f"(n), "f"(r));
#endif
    return q;
}

// cudapoa_nw_banded.cuh:28-48
__device__ __forceinline__ int32_t band_start(int32_t row, float gradient, int32_t band, int32_t max_column)
{
    float p;
    float fr = static_cast<float>(row);
#ifdef POA_EMU
    p = fr * gradient;
#else
    asm("mul.ftz.f32 %0, %1, %2;" : "=f"(p) : "f"(fr), "f"(gradient));
#endif
    int32_t start = __float2int_rz(p) - band / 2;
    start         = max(start, 0);
    if (start + band > max_column)
        start = max_column - band + 4;
    start = max(start, 0);
    return start - (start % 4);
}

__device__ __forceinline__ int32_t wrap_add(int32_t a, int32_t b)
{
    return static_cast<int32_t>(static_cast<uint32_t>(a) + static_cast<uint32_t>(b));
}

// Inclusive warp prefix maximum.
__device__ __forceinline__ int32_t warp_prefix_max(int32_t v, int32_t lane)
{
#pragma unroll
    for (int32_t d = 1; d < 32; d <<= 1)
    {
        int32_t o = __shfl_up_sync(kFull, v, d);
        if (lane >= d)
            v = max(v, o);
    }
    return v;
}

// Horizontal pass over one 128-cell chunk (4 cells per lane): H[j] = max(S[j], H[j-1] + gap) with H[-1] = carry.
// Computed as a prefix maximum of S[j] - j*gap; equal to the reference's warp fix-point
// (cudapoa_nw_banded.cuh:301-351, cudapoa_nw.cuh:257-313), which never truncates in this step.
template <typename ScoreT>
__device__ __forceinline__ void horizontal_pass(ScoreT (&c)[4], ScoreT& carry, int32_t gap, int32_t lane)
{
    using O      = ScoreOps<ScoreT>;
    int32_t base = lane * 4;
    int32_t x0   = O::add(c[0], -(base + 0) * gap);
    int32_t x1   = max(x0, O::add(c[1], -(base + 1) * gap));
    int32_t x2   = max(x1, O::add(c[2], -(base + 2) * gap));
    int32_t x3   = max(x2, O::add(c[3], -(base + 3) * gap));
    int32_t incl = warp_prefix_max(x3, lane);
    int32_t prev = __shfl_up_sync(kFull, incl, 1);
    int32_t cin  = O::add(carry, gap); // position -1 of the chunk
    prev         = (lane == 0) ? cin : max(prev, cin);
    c[0]         = O::narrow(wrap_add(max(x0, prev), (base + 0) * gap));
    c[1]         = O::narrow(wrap_add(max(x1, prev), (base + 1) * gap));
    c[2]         = O::narrow(wrap_add(max(x2, prev), (base + 2) * gap));
    c[3]         = O::narrow(wrap_add(max(x3, prev), (base + 3) * gap));
    carry        = __shfl_sync(kFull, c[3], 31);
}

// Score matrix accessor with the reference's effective semantics.
//  banded: row stride band + 8; cell [0] is the left boundary (column 0 value while band_start == 0, the
//          sentinel afterwards), cells [1..band] are columns band_start+1.., cells [band+1..band+7] stay at the
//          sentinel; row 0 holds j*gap in all band + 8 cells (cudapoa_nw_banded.cuh:208-212).
//  full:   row stride = the window's scores_width; cell [j] is column j.
template <typename ScoreT, bool BANDED>
struct ScoreMatrix
{
    ScoreT* s;
    const int32_t* row_bs; // banded only
    int32_t width;
    int32_t band;
    ScoreT sentinel;

    __device__ __forceinline__ ScoreT* row(int32_t r) const { return s + static_cast<int64_t>(r) * width; }

    // get_score (cudapoa_nw_banded.cuh:106-120) / direct read (cudapoa_nw.cuh)
    __device__ __forceinline__ ScoreT get(int32_t r, int32_t column) const
    {
        if (!BANDED)
            return row(r)[column];
        if (column == 0)
            return row(r)[0];
        int32_t b = row_bs[r];
        if (column > b + band || column < b)
            return sentinel;
        return row(r)[column - b];
    }
};

// Needleman-Wunsch of one read against the graph, exact arithmetic.  All 32 lanes of the warp call this.
// Returns the alignment length, or -1.  The alignment itself is left in g.aln_node[0 .. read_len): the node each
// read base pairs with, or -1 for an inserted base.  (The reference's traceback emits (node, read position) pairs
// end -> start, cudapoa_nw_banded.cuh:386-474; the pairs without a read position are never looked at by
// addAlignmentToGraph, cudapoa_add_alignment.cuh:84-90, and the others carry read positions len-1 .. 0 in order, so
// this array holds the same information.)
template <typename ScoreT, typename IdT, bool BANDED>
__device__ int32_t align_exact(const Graph<IdT>& g, int32_t node_count, const uint8_t* read, int32_t read_len,
                               ScoreT* scores, int32_t width, int32_t band, int32_t gap, int32_t mismatch,
                               int32_t match, int32_t lane, int32_t* score_out = nullptr)
{
    using O = ScoreOps<ScoreT>;
    ScoreMatrix<ScoreT, BANDED> M;
    M.s        = scores;
    M.row_bs   = g.row_bs;
    M.width    = width;
    M.band     = band;
    M.sentinel = BANDED ? band_sentinel<ScoreT>(gap, mismatch, match) : 0;

    const int32_t max_column = read_len + 1;
    float gradient           = 0.f;
    if (BANDED)
    {
        gradient = band_gradient(read_len, node_count);
        for (int32_t r = lane; r <= node_count; r += 32)
            g.row_bs[r] = band_start(r, gradient, band, max_column);
    }
    // row 0
    const int32_t row0_cells = BANDED ? width : read_len + 1;
    for (int32_t j = lane; j < row0_cells; j += 32)
        scores[j] = O::narrow(j * gap);
    __syncwarp();

    const int32_t span = BANDED ? band : ((read_len + 127) / 128) * 128;
    for (int32_t r = 1; r <= node_count; r++)
    {
        const int32_t node = g.sorted[r - 1];
        const int32_t pc   = g.in_cnt[node];
        const uint8_t base = g.nodes[node];
        const int32_t bs   = BANDED ? g.row_bs[r] : 0;
        ScoreT* out        = M.row(r);

        // left boundary (cudapoa_nw_banded.cuh:214-249 / cudapoa_nw.cuh:181-210)
        ScoreT left;
        if (BANDED && bs != 0)
        {
            left = M.sentinel;
        }
        else if (pc == 0)
        {
            left = O::narrow(gap);
        }
        else
        {
            ScoreT best = O::type_min();
            for (int32_t p = 0; p < pc; p++)
            {
                int32_t prow = g.pos[g.in_edge(node, p)] + 1;
                best         = max(best, M.row(prow)[0]);
            }
            left = O::narrow(O::add(best, gap));
        }
        if (lane == 0)
            out[0] = left;
        if (BANDED && lane < kBandPad - 1)
            out[band + 1 + lane] = M.sentinel;

        ScoreT carry = left;
        for (int32_t c0 = 0; c0 < span; c0 += 128)
        {
            const int32_t col_idx0 = c0 + lane * 4;   // cells col_idx0+1 .. col_idx0+4 of this row
            const int32_t read_pos = bs + col_idx0;   // 0-based read position of the first cell
            int32_t prof[4];
#pragma unroll
            for (int k = 0; k < 4; k++)
                prof[k] = (read_pos + k < read_len && read[read_pos + k] == base) ? match : mismatch;

            ScoreT cell[4];
            const bool active = BANDED || (read_pos < read_len);
            if (active)
            {
                for (int32_t p = 0; p < max(pc, 1); p++)
                {
                    const int32_t prow = (pc == 0) ? 0 : g.pos[g.in_edge(node, p)] + 1;
                    ScoreT cand[4];
                    bool in_band = true;
                    int32_t off  = read_pos;
                    if (BANDED)
                    {
                        const int32_t pb = g.row_bs[prow];
                        // get_scores, cudapoa_nw_banded.cuh:122-174
                        in_band = !((read_pos + 1 > pb + band + 4) || (read_pos + 1 < pb));
                        off     = read_pos - pb;
                    }
                    if (!in_band)
                    {
#pragma unroll
                        for (int k = 0; k < 4; k++)
                            cand[k] = M.sentinel;
                    }
                    else
                    {
                        const ScoreT* ps = M.row(prow) + off;
                        ScoreT pv[5];
#pragma unroll
                        for (int k = 0; k < 5; k++)
                            pv[k] = ps[k];
#pragma unroll
                        for (int k = 0; k < 4; k++)
                            cand[k] = O::narrow(max(O::add(pv[k], prof[k]), O::add(pv[k + 1], gap)));
                    }
#pragma unroll
                    for (int k = 0; k < 4; k++)
                        cell[k] = (p == 0) ? cand[k] : max(cell[k], cand[k]);
                }
            }
            else
            {
#pragma unroll
                for (int k = 0; k < 4; k++)
                    cell[k] = O::type_min(); // right of the read: never stored, never propagates leftwards
            }
            horizontal_pass<ScoreT>(cell, carry, gap, lane);
            if (active)
            {
#pragma unroll
                for (int k = 0; k < 4; k++)
                    out[col_idx0 + 1 + k] = cell[k];
            }
        }
        __syncwarp();
    }

    // sink selection + traceback on lane 0 (cudapoa_nw_banded.cuh:369-480 / cudapoa_nw.cuh:328-459)
    int32_t aligned = 0;
    int32_t start_score = 0; // score of the cell the traceback starts from
    if (lane == 0)
    {
        int32_t i = 0, j = read_len;
        ScoreT mscore = O::type_min();
        for (int32_t idx = 1; idx <= node_count; idx++)
        {
            if (g.out_cnt[g.sorted[idx - 1]] == 0)
            {
                ScoreT s = M.get(idx, j);
                if (mscore < s)
                {
                    mscore = s;
                    i      = idx;
                }
            }
        }
        int32_t prev_i = 0, prev_j = 0, loops = 0;
        const int32_t limit = read_len + node_count + 2;
        start_score         = static_cast<int32_t>(M.get(i, j));
        while (!(i == 0 && j == 0) && loops < limit)
        {
            loops++;
            const ScoreT sij = M.get(i, j);
            bool found       = false;
            if (i != 0)
            {
                const int32_t node = g.sorted[i - 1];
                const int32_t pc   = g.in_cnt[node];
                if (j != 0)
                {
                    const int32_t cost = (g.nodes[node] == read[j - 1]) ? match : mismatch;
                    for (int32_t p = 0; p < max(pc, 1) && !found; p++)
                    {
                        const int32_t pi = (pc == 0) ? 0 : g.pos[g.in_edge(node, p)] + 1;
                        if (static_cast<int32_t>(sij) == O::add(M.get(pi, j - 1), cost))
                        {
                            prev_i = pi;
                            prev_j = j - 1;
                            found  = true;
                        }
                    }
                }
                for (int32_t p = 0; p < max(pc, 1) && !found; p++)
                {
                    const int32_t pi = (pc == 0) ? 0 : g.pos[g.in_edge(node, p)] + 1;
                    if (static_cast<int32_t>(sij) == O::add(M.get(pi, j), gap))
                    {
                        prev_i = pi;
                        prev_j = j;
                        found  = true;
                    }
                }
            }
            if (!found && j > 0 && static_cast<int32_t>(sij) == O::add(M.get(i, j - 1), gap))
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
    if (score_out != nullptr)
        *score_out = __shfl_sync(kFull, start_score, 0);
    return __shfl_sync(kFull, aligned, 0);
}

// Backbone from read 0 (cudapoa_kernels.cuh:171-209), one node per lane and step.
template <typename IdT, bool MSA>
__device__ void init_backbone(const Graph<IdT>& g, const uint8_t* seq, const int8_t* w, int32_t len, int32_t lane)
{
    for (int32_t i = lane; i < len; i += 32)
    {
        g.nodes[i]   = seq[i];
        g.sorted[i]  = static_cast<IdT>(i);
        g.pos[i]     = static_cast<IdT>(i);
        g.aln_cnt[i] = 0;
        g.cov[i]     = 1;
        const int32_t wi = w ? w[i] : 1;
        if (i == 0)
        {
            g.in_cnt[0]    = 0;
            g.weight(0, 0) = static_cast<uint16_t>(wi);
        }
        else
        {
            g.in_cnt[i]     = 1;
            g.in_edge(i, 0) = static_cast<IdT>(i - 1);
            g.weight(i, 0)  = static_cast<uint16_t>((w ? w[i - 1] : 1) + wi);
        }
        if (i + 1 < len)
        {
            g.out_cnt[i]     = 1;
            g.out_edge(i, 0) = static_cast<IdT>(i + 1);
        }
        else
        {
            g.out_cnt[i] = 0;
        }
        if (MSA)
            g.path[i] = static_cast<IdT>(i); // read 0 visits node i with base i
    }
}

// Fuse one alignment into the graph (cudapoa_add_alignment.cuh:59-279), 32 read positions per step.
//
// The reference walks the alignment serially.  Its per-position work only touches (a) the node the base pairs
// with and that node's aligned set, (b) the in-edge list of the position's node and (c) the out-edge list of the
// previous position's node.  An alignment path visits a node at most once and never visits two nodes of one
// aligned set (aligned nodes are mutually unreachable: every edge the fusion adds either shortcuts an existing
// path or attaches a new node beside the one it is aligned to), so distinct positions touch disjoint state and
// the lanes can proceed independently.  Order-dependent pieces are reproduced explicitly: new node ids are handed
// out in position order (ballot + popc), and the first failing position decides the status.
// path (MSA only): node visited by each base of this read, replacing the reference's per-edge read lists.
template <typename IdT, bool MSA>
__device__ int32_t add_alignment(const Graph<IdT>& g, int32_t& node_count, const uint8_t* read, int32_t read_len,
                                 const int8_t* w, IdT* path, int32_t max_nodes, int32_t lane)
{
    int32_t n          = node_count;
    int32_t carry_curr = -1; // node of read position k0 - 1
    const unsigned lt  = (1u << lane) - 1u;
    for (int32_t k0 = 0; k0 < read_len; k0 += 32)
    {
        const int32_t k = k0 + lane;
        bool active     = k < read_len;
        int32_t curr    = -1;
        int32_t gn      = -1;
        int32_t bw      = 1;
        uint8_t base    = 0;
        bool make_new   = false;
        if (active)
        {
            base = read[k];
            bw   = w ? w[k] : 1;
            gn   = g.aln_node[k];
            if (gn == -1)
            {
                make_new = true;
            }
            else if (g.nodes[gn] == base)
            {
                curr = gn;
            }
            else
            {
                const int32_t na = g.aln_cnt[gn];
                for (int32_t a = 0; a < na; a++)
                {
                    const int32_t aid = g.aln(gn, a);
                    if (g.nodes[aid] == base)
                    {
                        curr = aid;
                        break;
                    }
                }
                make_new = (curr == -1);
            }
        }
        const unsigned new_mask = __ballot_sync(kFull, make_new);
        if (make_new)
            curr = n + __popc(new_mask & lt);
        // `curr = n++; if (n >= max_nodes) return error` (:113-117, :169-173): the first creation that fills the graph
        const unsigned node_err = __ballot_sync(kFull, make_new && (curr + 1 >= max_nodes));
        const int32_t node_err_lane = node_err ? (__ffs(node_err) - 1) : 32;
        if (lane >= node_err_lane)
            active = false;
        n += __popc(new_mask);
        if (active && make_new)
        {
            g.nodes[curr]   = base;
            g.out_cnt[curr] = 0;
            g.in_cnt[curr]  = 0;
            g.cov[curr]     = 0;
            int32_t k2      = 0;
            if (gn != -1)
            {
                // cross-link the new node with gn and everything already aligned to gn (:150-170)
                const int32_t na = g.aln_cnt[gn];
                for (int32_t a = 0; a < na; a++)
                {
                    const int32_t aid = g.aln(gn, a);
                    const int32_t c   = g.aln_cnt[aid];
                    g.aln(aid, c)     = static_cast<IdT>(curr);
                    g.aln_cnt[aid]    = static_cast<uint16_t>(c + 1);
                    g.aln(curr, k2++) = static_cast<IdT>(aid);
                }
                g.aln(gn, na)     = static_cast<IdT>(curr);
                g.aln_cnt[gn]     = static_cast<uint16_t>(na + 1);
                g.aln(curr, k2++) = static_cast<IdT>(gn);
            }
            g.aln_cnt[curr] = static_cast<uint16_t>(k2);
        }
        if (MSA && active)
            path[k] = static_cast<IdT>(curr);

        int32_t head = __shfl_up_sync(kFull, curr, 1);
        if (lane == 0)
            head = carry_curr;
        bool edge_err = false;
        if (active)
        {
            if (head != -1)
            {
                const int32_t prev_weight = w ? w[k - 1] : 1;
                const uint16_t add        = static_cast<uint16_t>(prev_weight + bw);
                bool exists               = false;
                const int32_t inc         = make_new ? 0 : static_cast<int32_t>(g.in_cnt[curr]);
                // the first three slots are fetched together (independent loads), the rest one by one; an edge
                // occurs at most once per list, so "update every match" (:221-229) is "update the match"
                const int32_t s0 = g.in_edge(curr, 0), s1 = g.in_edge(curr, 1), s2 = g.in_edge(curr, 2);
                int32_t hit      = -1;
                if (inc > 0 && s0 == head)
                    hit = 0;
                else if (inc > 1 && s1 == head)
                    hit = 1;
                else if (inc > 2 && s2 == head)
                    hit = 2;
                for (int32_t e = 3; e < inc && hit < 0; e++)
                {
                    if (g.in_edge(curr, e) == head)
                        hit = e;
                }
                if (hit >= 0)
                {
                    exists              = true;
                    g.weight(curr, hit) = static_cast<uint16_t>(g.weight(curr, hit) + add);
                }
                if (!exists)
                {
                    g.in_edge(curr, inc)   = static_cast<IdT>(head);
                    g.weight(curr, inc)    = add;
                    g.in_cnt[curr]         = static_cast<uint16_t>(inc + 1);
                    const int32_t outc     = g.out_cnt[head];
                    g.out_edge(head, outc) = static_cast<IdT>(curr);
                    g.out_cnt[head]        = static_cast<uint16_t>(outc + 1);
                    edge_err               = (outc + 1 >= kMaxNodeEdges || inc + 1 >= kMaxNodeEdges); // :245-249
                }
            }
            g.cov[curr]++;
        }
        const unsigned edge_mask = __ballot_sync(kFull, edge_err);
        if (edge_mask != 0) // an edge failure sits at an earlier position than any node failure of this step
            return kStEdgeCountExceeded;
        if (node_err != 0)
            return kStNodeCountExceeded;
        carry_curr = __shfl_sync(kFull, curr, 31);
        __syncwarp();
    }
    node_count = n;
    return kStSuccess;
}

// Adaptive banding: does the alignment just found (g.aln_node[k] = node paired with read position k, or < 0) come
// within kAdaptiveBandMargin cells of an edge of the band of width `band` that is not an edge of the matrix?
// Band placement as in the alignment itself (band_gradient / band_start).  All lanes; uniform result.
template <typename IdT>
__device__ bool alignment_touches_band(const Graph<IdT>& g, int32_t node_count, int32_t read_len, int32_t band,
                                       int32_t lane)
{
    const int32_t max_column = read_len + 1;
    const float gradient     = band_gradient(read_len, node_count);
    bool touch               = false;
    for (int32_t k = lane; k < read_len; k += 32)
    {
        const int32_t node = g.aln_node[k];
        if (node < 0)
            continue;
        const int32_t row = static_cast<int32_t>(g.pos[node]) + 1;
        const int32_t bs  = band_start(row, gradient, band, max_column);
        const int32_t c   = k - bs; // 0-based cell of column k + 1 inside the band
        if ((bs > 0 && c < kAdaptiveBandMargin) || (bs + band < max_column && c >= band - kAdaptiveBandMargin))
            touch = true;
    }
    return __any_sync(kFull, touch);
}

// Kahn order with the reference's FIFO discipline (cudapoa_topsort.cuh:38-88: sources in id order, children in out-edge
// order), out of global memory: the path of graphs that outgrow the shared-memory sort below and of 32-bit node ids.
// Edge-parallel: sorted[] doubles as the queue; the warp pops one node at a time, lane e satisfies one in-edge of the
// node's e-th child (the children of a node are distinct, so the decrements do not collide) and the children that
// became ready are appended in lane order by a ballot.  The sources are compacted the same way, 32 ids at a time.
template <typename IdT>
__device__ void topsort_kahn_warp(const Graph<IdT>& g, int32_t node_count, int32_t lane)
{
    const unsigned below = (1u << lane) - 1u;
    int32_t tail         = 0;
    for (int32_t n0 = 0; n0 < node_count; n0 += 32)
    {
        const int32_t n = n0 + lane;
        bool source     = false;
        if (n < node_count)
        {
            const uint16_t pending = g.in_cnt[n];
            g.kahn[n]              = pending;
            source                 = pending == 0;
        }
        const unsigned mask = __ballot_sync(kFull, source);
        if (source)
        {
            const int32_t at = tail + __popc(mask & below);
            g.sorted[at]     = static_cast<IdT>(n);
            g.pos[n]         = static_cast<IdT>(at);
        }
        tail += __popc(mask);
    }
    __syncwarp();
    for (int32_t head = 0; head < tail; head++)
    {
        const int32_t node   = g.sorted[head]; // one address for the warp: a broadcast
        const int32_t fanout = g.out_cnt[node];
        for (int32_t e0 = 0; e0 < fanout; e0 += 32)
        {
            const int32_t e = e0 + lane;
            bool ready      = false;
            int32_t child   = 0;
            if (e < fanout)
            {
                child                  = g.out_edge(node, e);
                const uint16_t pending = static_cast<uint16_t>(g.kahn[child] - 1);
                g.kahn[child]          = pending;
                ready                  = pending == 0;
            }
            const unsigned mask = __ballot_sync(kFull, ready);
            if (ready)
            {
                const int32_t at = tail + __popc(mask & below);
                g.sorted[at]     = static_cast<IdT>(child);
                g.pos[child]     = static_cast<IdT>(at);
            }
            tail += __popc(mask);
        }
        __syncwarp(); // the appended ids are read by every lane
    }
}

// The same order computed in the warp's shared-memory scratch.  All 32 lanes call this.
//   * the warp stages one 4-byte record per node (coalesced reads of the slot-major graph arrays):
//       out-edge 0 | out-edge 1 << 12 | min(out-degree, 3) << 24 | remaining in-degree << 26
//     (node ids must fit 12 bits); the third and later out-edges of the few nodes that have them stay in global
//     memory and are fetched as soon as such a node is popped, under the visits of its first two children;
//   * lane 0 runs the FIFO.  The queue is a 256-entry ring (a POA graph's frontier is a handful of nodes), sorted[] and
//     node_id_to_pos[] are written as nodes leave it: the walk is a chain of dependent shared-memory accesses, so
//     the stores ride along for free, and 4 bytes per node is what lets ~48 windows per SM sort at the same time.
//     A POA graph is mostly chains: when the node in hand has one child and nothing else is queued, the child -- if
//     this was its last open in-edge -- is the next node and its record is already in a register: such a step is one
//     shared-memory load and no store.  Otherwise the next queued node and its record are fetched before the
//     children are visited, so the two dependent loads of a pop hide under the visits.
// Returns false when the graph does not fit the scratch or the ring is overrun (sorted[] / node_id_to_pos[] may then
// be partly written); the caller then runs topsort_kahn_warp, which rewrites both.
constexpr int32_t kKahnBytesPerNode = 4;
constexpr int32_t kKahnQueueEntries = 256;
constexpr int32_t kKahnMaxNodes     = 4095;

// Nodes a graph may have for the sort to run out of `smem_bytes` of scratch.
__host__ __device__ inline int32_t kahn_smem_nodes(int32_t smem_bytes)
{
    const int32_t n = ((smem_bytes - 2 * kKahnQueueEntries) / kKahnBytesPerNode) & ~7;
    return n < kKahnMaxNodes ? n : (kKahnMaxNodes & ~7);
}

template <typename IdT>
__device__ bool topsort_kahn_smem(const Graph<IdT>& g, int32_t node_count, uint32_t smem, int32_t smem_bytes,
                                  int32_t lane)
{
    if (sizeof(IdT) != 2)
        return false;
    const int32_t cap = kahn_smem_nodes(smem_bytes);
    if (node_count > cap)
        return false;
    const uint32_t rec   = smem;           // u32[cap]
    const uint32_t queue = rec + 4u * cap; // u16[kKahnQueueEntries], a ring indexed by pop / push counts
    constexpr uint32_t kQMask = kKahnQueueEntries - 1;
    const unsigned lt    = (1u << lane) - 1u;
    int32_t tail         = 0;
#pragma unroll 2
    for (int32_t n0 = 0; n0 < node_count; n0 += 32)
    {
        const int32_t n = n0 + lane;
        bool source     = false;
        if (n < node_count)
        {
            const uint32_t ic = g.in_cnt[n];
            const uint32_t oc = g.out_cnt[n];
            // slots beyond the out-degree hold stale but readable values: load unconditionally, never used
            const uint32_t e0 = static_cast<uint16_t>(g.out_edge(n, 0)) & 0xfffu;
            const uint32_t e1 = static_cast<uint16_t>(g.out_edge(n, 1)) & 0xfffu;
            sts_u32(rec + 4u * n, e0 | (e1 << 12) | (min(oc, 3u) << 24) | (ic << 26));
            source = (ic == 0);
        }
        const unsigned mask = __ballot_sync(kFull, source);
        if (tail + __popc(mask) > kKahnQueueEntries) // more sources than the ring holds (uniform; nothing written yet)
            return false;
        if (source)
            sts_u16(queue + 2u * (tail + __popc(mask & lt)), static_cast<uint32_t>(n));
        tail += __popc(mask);
    }
    __syncwarp();
    int32_t ok = 1;
    if (lane == 0)
    {
        constexpr uint32_t kOne = 1u << 26; // one in-edge
        uint32_t head   = 0;                // nodes popped so far == rank of the node in hand
        uint32_t pushed = static_cast<uint32_t>(tail);
        uint32_t node   = lds_u16(queue);   // a graph always has a source
        uint32_t v      = lds_u32(rec + 4u * node);
        IdT* const sorted_out = g.sorted;
        IdT* const pos_out    = g.pos;
        // visit(child): one in-edge of child is satisfied; queue it when it was the last one
#define POA_KAHN_VISIT(child_expr)                                   \
    {                                                                \
        const uint32_t child_ = (child_expr);                        \
        const uint32_t at_    = rec + 4u * child_;                   \
        const uint32_t left_  = lds_u32(at_) - kOne;                 \
        sts_u32(at_, left_);                                         \
        if (left_ < kOne)                                            \
        {                                                            \
            sts_u16(queue + 2u * (pushed & kQMask), child_);         \
            pushed++;                                                \
        }                                                            \
    }
        for (;;)
        {
            sorted_out[head] = static_cast<IdT>(node);
            pos_out[node]    = static_cast<IdT>(head);
            head++;
            const uint32_t oc = (v >> 24) & 3u;
            if (oc == 1 && head == pushed)
            {
                // a chain step with nothing else waiting (most of a POA graph): if the only child becomes ready it is
                // the next node, and its record is in hand -- no queue traffic, nothing stored
                const uint32_t child = v & 0xfffu;
                const uint32_t at    = rec + 4u * child;
                const uint32_t left  = lds_u32(at) - kOne;
                if (left < kOne)
                {
                    pushed++;
                    node = child;
                    v    = left;
                    continue;
                }
                sts_u32(at, left);
                break; // the queue is empty and nothing became ready: every node is out
            }
            // the next queued node and its record are fetched under the visits (a queued node's record is final)
            const bool have_next = head < pushed;
            uint32_t next_node = 0, next_v = 0;
            if (have_next)
            {
                next_node = lds_u16(queue + 2u * (head & kQMask));
                next_v    = lds_u32(rec + 4u * next_node);
            }
            if (oc > 2) // rare: the out-degree and the further children come from global memory
            {
                const uint32_t more = static_cast<uint32_t>(g.out_cnt[node]);
                const uint32_t e2   = static_cast<uint16_t>(g.out_edge(node, 2));
                POA_KAHN_VISIT(v & 0xfffu)
                POA_KAHN_VISIT((v >> 12) & 0xfffu)
                POA_KAHN_VISIT(e2)
                for (uint32_t e = 3; e < more; e++)
                    POA_KAHN_VISIT(static_cast<uint16_t>(g.out_edge(node, static_cast<int32_t>(e))))
            }
            else
            {
                if (oc > 0)
                    POA_KAHN_VISIT(v & 0xfffu)
                if (oc > 1)
                    POA_KAHN_VISIT((v >> 12) & 0xfffu)
            }
            if (head == pushed)
                break;
            if (pushed - head > kKahnQueueEntries) // the ring was overrun: give up, the caller sorts again
            {
                ok = 0;
                break;
            }
            if (have_next)
            {
                node = next_node;
                v    = next_v;
            }
            else
            {
                node = lds_u16(queue + 2u * (head & kQMask));
                v    = lds_u32(rec + 4u * node);
            }
        }
#undef POA_KAHN_VISIT
    }
    ok = __shfl_sync(kFull, ok, 0);
    __syncwarp();
    return ok != 0;
}

// Topological order that keeps aligned nodes next to each other (the SPOA_ACCURATE build of the reference and every
// MSA: cudapoa_topsort.cuh:94-189).  The reference's order is that of a depth-first walk towards predecessors in which
// a node first visits its aligned nodes (last to first), then its predecessors (last to first), and a group of
// aligned nodes is emitted -- leader, then the leader's aligned list -- when the member the walk met first (the leader)
// completes.  Here the walk keeps one frame per open node, (node, candidates still to visit), instead of a stack of
// pending ids; the warp looks at 32 candidates of the open node at a time (lane k: the k-th from the end), and the
// group is written out by the lanes together.  check[n] stays 1 exactly for the leaders, which is what msa_columns
// counts.  All 32 lanes.
template <typename IdT>
__device__ void topsort_spoa_order(const Graph<IdT>& g, int32_t node_count, int32_t lane)
{
    constexpr uint8_t kOpen = 1, kDone = 2;
    for (int32_t n = lane; n < node_count; n += 32)
    {
        g.marks[n] = 0;
        g.check[n] = 1;
    }
    __syncwarp();
    IdT* frames = g.stack; // frames[2 d] = node, frames[2 d + 1] = candidates left; depth <= node_count <= M
    int32_t out = 0;
    for (int32_t root = 0; root < node_count; root++)
    {
        const bool visited = g.marks[root] != 0; // uniform
        __syncwarp();                            // .. and read by every lane before lane 0 marks the root
        if (visited)
            continue;
        int32_t depth = 0;
        int32_t node  = root;
        int32_t left  = -1; // -1: not looked at yet
        for (;;)
        {
            const int32_t n_in  = g.in_cnt[node];
            const int32_t n_aln = g.check[node] ? static_cast<int32_t>(g.aln_cnt[node]) : 0;
            if (left < 0)
            {
                // first look: this node leads its group unless an aligned node got here first; the aligned nodes
                // that are still open follow it (check = 0: they neither emit nor look at their own aligned lists)
                for (int32_t a = lane; a < n_aln; a += 32)
                {
                    const int32_t peer = g.aln(node, a);
                    if (g.marks[peer] != kDone)
                        g.check[peer] = 0;
                }
                if (lane == 0)
                    g.marks[node] = kOpen;
                left = n_in + n_aln;
            }
            // next candidate that is not done, from the end: aligned nodes occupy [n_in, n_in + n_aln)
            int32_t next = -1;
            while (left > 0)
            {
                const int32_t c = left - 1 - lane;
                int32_t cand    = -1;
                if (c >= 0)
                {
                    cand = c >= n_in ? static_cast<int32_t>(g.aln(node, c - n_in)) : static_cast<int32_t>(g.in_edge(node, c));
                    if (g.marks[cand] == kDone)
                        cand = -1;
                }
                const unsigned mask = __ballot_sync(kFull, cand >= 0);
                if (mask != 0)
                {
                    const int32_t first = __ffs(mask) - 1;
                    next                = __shfl_sync(kFull, cand, first);
                    left                = left - 1 - first;
                    break;
                }
                left = max(left - 32, 0);
            }
            if (next >= 0)
            {
                if (lane == 0)
                {
                    frames[2 * depth]     = static_cast<IdT>(node);
                    frames[2 * depth + 1] = static_cast<IdT>(left);
                }
                depth++;
                node = next;
                left = -1;
                __syncwarp();
                continue;
            }
            // every candidate is done: the node is, too; a leader emits its group
            if (g.check[node])
            {
                const int32_t members = 1 + g.aln_cnt[node];
                for (int32_t m = lane; m < members; m += 32)
                {
                    const int32_t id = m == 0 ? node : static_cast<int32_t>(g.aln(node, m - 1));
                    g.sorted[out + m] = static_cast<IdT>(id);
                    g.pos[id]         = static_cast<IdT>(out + m);
                }
                out += members;
            }
            if (lane == 0)
                g.marks[node] = kDone;
            __syncwarp();
            if (depth == 0)
                break;
            depth--;
            node = frames[2 * depth];
            left = frames[2 * depth + 1];
            __syncwarp(); // the frame's slot is free for the next push
        }
    }
}

// ---- heaviest-bundle consensus (cudapoa_generate_consensus.cuh:28-276) ---------------------------------
// One formulation for every graph size.  All 32 lanes call these.
//   * the node scores and predecessors of the recurrence live in the warp's scratch when the graph fits it (6 bytes
//     per node, 16-bit ids) and in the window's arena otherwise (ConsensusStore);
//   * scoring pass (consensus_scores): the warp stages a block of ranks at a time into shared memory -- node, in-degree,
//     first three in-edges and their weights; two gather levels, eight ranks per lane in flight -- and lane 0 runs the
//     recurrence over the staged records.  The first pass scores every rank; after the heaviest node turned out not to
//     be a sink, its competitors (the other predecessors of its children) are struck out by the lanes together and
//     the same pass re-scores the ranks after it, now ignoring struck-out predecessors (the reference's
//     branchCompletion);
//   * backtrack: lane 0 follows the predecessor chain and records the node ids, then the warp emits bases and
//     coverages for all of them in parallel.
// Tie-breaks as in the reference: among the in-edges of a node the heaviest wins, then the one whose source scores
// higher, then the later one (`<=`, :179-181); among nodes the later rank wins (`<=`, :196).
constexpr int32_t kConsStageBytes = 4096;

template <typename IdT, bool SMEM>
struct ConsensusStore
{
    const Graph<IdT>& g;
    uint32_t score_s; // int32[cap], by node id
    uint32_t pred_s;  // u16[cap], by node id (0xffff: none)
    __device__ __forceinline__ int32_t score(int32_t n) const
    {
        if (SMEM)
            return static_cast<int32_t>(lds_u32(score_s + 4u * static_cast<uint32_t>(n)));
        return g.cscore[n];
    }
    __device__ __forceinline__ void set_score(int32_t n, int32_t v) const
    {
        if (SMEM)
            sts_u32(score_s + 4u * static_cast<uint32_t>(n), static_cast<uint32_t>(v));
        else
            g.cscore[n] = v;
    }
    __device__ __forceinline__ int32_t pred(int32_t n) const // -1: none
    {
        if (SMEM)
        {
            const uint32_t p = lds_u16(pred_s + 2u * static_cast<uint32_t>(n));
            return p == 0xffffu ? -1 : static_cast<int32_t>(p);
        }
        return static_cast<int32_t>(g.cpred[n]);
    }
    __device__ __forceinline__ void set_pred(int32_t n, int32_t p) const
    {
        if (SMEM)
            sts_u16(pred_s + 2u * static_cast<uint32_t>(n), static_cast<uint32_t>(p) & 0xffffu);
        else
            g.cpred[n] = static_cast<IdT>(p);
    }
};

// Scores ranks first_rank .. node_count - 1 and returns the heaviest node among them (`best_id` if none reaches
// `best_score`).  `skip_struck`: predecessors whose score is -1 do not count.
template <typename IdT, bool SMEM>
__device__ int32_t consensus_scores(const ConsensusStore<IdT, SMEM>& store, int32_t first_rank, int32_t node_count,
                                    bool skip_struck, int32_t best_score, int32_t best_id, uint32_t stage, int32_t lane)
{
    const Graph<IdT>& g = store.g;
    // staged record of a rank: 16 bytes with 16-bit ids, 32 with 32-bit ids
    constexpr int32_t kRecBytes = sizeof(IdT) == 2 ? 16 : 32;
    constexpr int32_t kRanks    = kConsStageBytes / kRecBytes;
    constexpr int32_t kPerLane  = kRanks / 32;
    for (int32_t c0 = first_rank; c0 < node_count; c0 += kRanks)
    {
        int32_t node[kPerLane];
#pragma unroll
        for (int u = 0; u < kPerLane; u++)
            node[u] = g.sorted[min(c0 + lane + 32 * u, node_count - 1)];
#pragma unroll
        for (int u = 0; u < kPerLane; u++)
        {
            const int32_t n   = node[u];
            const uint32_t ic = g.in_cnt[n];
            // slots beyond the in-degree hold stale but readable values
            const uint32_t w0 = g.weight(n, 0), w1 = g.weight(n, 1), w2 = g.weight(n, 2);
            const uint32_t at = stage + static_cast<uint32_t>(kRecBytes * (lane + 32 * u));
            if (sizeof(IdT) == 2)
            {
                const uint32_t e0 = static_cast<uint16_t>(g.in_edge(n, 0)), e1 = static_cast<uint16_t>(g.in_edge(n, 1)),
                               e2 = static_cast<uint16_t>(g.in_edge(n, 2));
                sts_v4(at, make_uint4(static_cast<uint32_t>(n) | (ic << 16), e0 | (e1 << 16), w0 | (w1 << 16), e2 | (w2 << 16)));
            }
            else
            {
                sts_v4(at, make_uint4(static_cast<uint32_t>(n), ic, static_cast<uint32_t>(g.in_edge(n, 0)),
                                      static_cast<uint32_t>(g.in_edge(n, 1))));
                sts_v4(at + 16u, make_uint4(static_cast<uint32_t>(g.in_edge(n, 2)), w0 | (w1 << 16), w2, 0u));
            }
        }
        __syncwarp();
        if (lane == 0)
        {
            const int32_t cend = min(kRanks, node_count - c0);
            for (int32_t q = 0; q < cend; q++)
            {
                const uint4 ra = lds_v4(stage + static_cast<uint32_t>(kRecBytes * q));
                uint4 rb       = make_uint4(0u, 0u, 0u, 0u);
                int32_t n, ic;
                if (sizeof(IdT) == 2)
                {
                    n  = static_cast<int32_t>(ra.x & 0xffffu);
                    ic = static_cast<int32_t>(ra.x >> 16);
                }
                else
                {
                    rb = lds_v4(stage + static_cast<uint32_t>(kRecBytes * q) + 16u);
                    n  = static_cast<int32_t>(ra.x);
                    ic = static_cast<int32_t>(ra.y);
                }
                int32_t pred       = -1;
                int32_t score      = -1;
                int32_t pred_score = 0; // score of `pred`
                for (int32_t e = 0; e < ic; e++)
                {
                    int32_t begin, w;
                    if (e > 2)
                    {
                        begin = static_cast<int32_t>(g.in_edge(n, e));
                        w     = g.weight(n, e);
                    }
                    else if (sizeof(IdT) == 2)
                    {
                        begin = static_cast<int32_t>(e == 0 ? (ra.y & 0xffffu) : e == 1 ? (ra.y >> 16) : (ra.w & 0xffffu));
                        w     = static_cast<int32_t>(e == 0 ? (ra.z & 0xffffu) : e == 1 ? (ra.z >> 16) : (ra.w >> 16));
                    }
                    else
                    {
                        begin = static_cast<int32_t>(e == 0 ? ra.z : e == 1 ? ra.w : rb.x);
                        w     = static_cast<int32_t>(e == 0 ? (rb.y & 0xffffu) : e == 1 ? (rb.y >> 16) : rb.z);
                    }
                    const int32_t bsc = store.score(begin);
                    if (skip_struck && bsc == -1)
                        continue;
                    if (score < w || (score == w && pred_score <= bsc))
                    {
                        score      = w;
                        pred       = begin;
                        pred_score = bsc;
                    }
                }
                if (pred != -1)
                    score += pred_score;
                if (best_score <= score)
                {
                    best_id    = n;
                    best_score = score;
                }
                store.set_score(n, score);
                store.set_pred(n, pred);
            }
        }
        __syncwarp();
    }
    return __shfl_sync(kFull, best_id, 0);
}

template <typename IdT, bool SMEM>
__device__ int32_t consensus_with_store(const ConsensusStore<IdT, SMEM>& store, int32_t node_count, uint8_t* consensus,
                                        uint16_t* coverage, int32_t max_consensus, uint32_t stage, int32_t lane)
{
    const Graph<IdT>& g = store.g;
    int32_t best        = consensus_scores<IdT, SMEM>(store, 0, node_count, false, -1, 0, stage, lane);
    int32_t rounds      = 0;
    while (g.out_cnt[best] != 0 && rounds < node_count) // uniform
    {
        // the heaviest node is not a sink: strike out the other predecessors of its children, score what follows it again
        const int32_t fanout = g.out_cnt[best];
        for (int32_t oe = 0; oe < fanout; oe++)
        {
            const int32_t child = g.out_edge(best, oe);
            const int32_t ic    = g.in_cnt[child];
            for (int32_t ie = lane; ie < ic; ie += 32)
            {
                const int32_t rival = g.in_edge(child, ie);
                if (rival != best)
                    store.set_score(rival, -1);
            }
        }
        __syncwarp();
        best = consensus_scores<IdT, SMEM>(store, static_cast<int32_t>(g.pos[best]) + 1, node_count, true, 0, 0, stage, lane);
        rounds++;
    }
    if (rounds >= node_count)
        return kStLoopCountExceeded;
    // backtrack: node ids of the path, end -> start, into the arena's stack area (4 M ids; a path has at most M)
    IdT* path     = g.stack;
    int32_t count = 0;
    if (lane == 0)
    {
        int32_t id = best;
        for (;;)
        {
            if (count < max_consensus)
                path[count] = static_cast<IdT>(id);
            const int32_t pred = store.pred(id);
            if (pred == -1)
                break;
            id = pred;
            count++;
        }
    }
    count = __shfl_sync(kFull, count, 0);
    if (count >= max_consensus - 1)
        return kStSeqSizeExceeded;
    __syncwarp();
    for (int32_t k = lane; k <= count; k += 32)
    {
        const int32_t id = path[k];
        consensus[k]     = g.nodes[id];
        uint16_t cov     = g.cov[id];
        const int32_t na = g.aln_cnt[id];
        for (int32_t a = 0; a < na; a++)
            cov = static_cast<uint16_t>(cov + g.cov[g.aln(id, a)]);
        coverage[k] = cov;
    }
    if (lane == 0)
        consensus[count + 1] = 0;
    return kStSuccess;
}

// Nodes a graph may have for its consensus scores to live in `smem_bytes` of scratch.
__host__ __device__ inline int32_t consensus_smem_nodes(int32_t smem_bytes)
{
    return ((smem_bytes - kConsStageBytes) / 6) & ~7;
}

// Writes the consensus and its coverage end -> start, NUL-terminated; returns a StatusType value.  Needs at least
// kConsStageBytes of scratch at `smem`.
template <typename IdT>
__device__ int32_t generate_consensus(const Graph<IdT>& g, int32_t node_count, uint8_t* consensus, uint16_t* coverage,
                                      int32_t max_consensus, uint32_t smem, int32_t smem_bytes, int32_t lane)
{
    const int32_t cap = consensus_smem_nodes(smem_bytes);
    if (sizeof(IdT) == 2 && node_count <= cap)
    {
        const ConsensusStore<IdT, true> store{g, smem + kConsStageBytes, smem + kConsStageBytes + 4u * static_cast<uint32_t>(cap)};
        return consensus_with_store<IdT, true>(store, node_count, consensus, coverage, max_consensus, smem, lane);
    }
    const ConsensusStore<IdT, false> store{g, 0u, 0u};
    return consensus_with_store<IdT, false>(store, node_count, consensus, coverage, max_consensus, smem, lane);
}

// MSA column of every node (cudapoa_generate_msa.cuh:27-46): the nodes of an aligned group share a column, columns
// count the groups in sorted order.  After topsort_spoa_order a group's ranks are contiguous and start with its leader,
// the one member with check[] == 1, so the column of a rank is the number of leaders up to it: a ballot per 32 ranks.
template <typename IdT>
__device__ int32_t msa_columns(const Graph<IdT>& g, int32_t node_count, int32_t lane)
{
    const unsigned upto = 0xffffffffu >> (31 - lane);
    int32_t columns     = 0;
    for (int32_t r0 = 0; r0 < node_count; r0 += 32)
    {
        const int32_t r = r0 + lane;
        int32_t node    = 0;
        bool leader     = false;
        if (r < node_count)
        {
            node   = g.sorted[r];
            leader = g.check[node] != 0;
        }
        const unsigned mask = __ballot_sync(kFull, leader);
        if (r < node_count)
            g.msa_pos[node] = static_cast<IdT>(columns + __popc(mask & upto) - 1);
        columns += __popc(mask);
    }
    __syncwarp();
    return columns;
}

// One MSA row per read (cudapoa_generate_msa.cuh:48-118): the whole warp writes row s -- '-' everywhere, then
// each base into the column of the node it was fused into.
template <typename IdT>
__device__ void msa_rows(const Graph<IdT>& g, const uint8_t* bases, const int32_t* read_len, int32_t n_reads,
                         uint8_t* msa, int32_t msa_len, int32_t max_consensus, int32_t max_seq, int32_t lane)
{
    const uint8_t* seq = bases;
    for (int32_t s = 0; s < n_reads; s++)
    {
        uint8_t* row      = msa + static_cast<int64_t>(s) * max_consensus;
        const int32_t len = read_len[s];
        const IdT* path   = g.path + static_cast<int64_t>(s) * max_seq;
        for (int32_t c = lane; c < msa_len; c += 32)
            row[c] = '-';
        if (lane == 0)
            row[msa_len] = 0;
        __syncwarp();
        for (int32_t k = lane; k < len; k += 32)
            row[g.msa_pos[path[k]]] = seq[k];
        seq += len;
    }
}



} // namespace poa_b200
