// Device data layout of the B200 POA engine (shared by host and device code).
//
// Every window owns one "arena" slice of device memory holding its graph and scratch arrays, plus a slice of
// the score pool.  Edge-like arrays are SLOT-MAJOR: element (node, slot) lives at [slot * M + node], so that
// the first few slots of consecutive nodes are contiguous (real POA graphs use 1-3 of the 50 slots the
// reference semantics require, cudapoa_structs.cuh:18-21) and warp-parallel passes over nodes coalesce.
#pragma once

#include <cstddef>
#include <cstdint>

namespace poa_b200
{

constexpr int32_t kMaxNodeEdges      = 50; // same limits as the reference: they are observable through
constexpr int32_t kMaxNodeAlignments = 50; // edge_count_exceeded_maximum_graph_size
constexpr int32_t kBandPad           = 8;  // cells right of the band kept at the sentinel (reference layout)
constexpr int32_t kBandRowPad        = 24; // cells of padding in this engine's banded row stride (three 16-byte blocks)
constexpr int32_t kWarp              = 32;
constexpr int32_t kAdaptiveBandFactor = 4; // adaptive banding: the band may double up to this multiple
constexpr int32_t kAdaptiveBandMargin = 4; // ... while an aligned pair lies within this many cells of a band edge

struct WindowDesc
{
    int64_t base_off;     // first base of the window in the flat bases/weights buffers
    int64_t score_off;    // first score cell of the window in the score pool (elements)
    int32_t read_off;     // first read of the window in read_len[]
    int32_t n_reads;      // reads accepted into the window
    int32_t scores_width; // full alignment: row stride of this window's score matrix
    int32_t aln_score;  // score of the pending alignment (adaptive banding decides on it)
};

// Byte offsets of the per-window arrays inside the window's arena slice.
struct ArenaLayout
{
    int64_t stride; // bytes per window
    int64_t nodes;
    int64_t in_cnt, out_cnt, aln_cnt, cov, kahn; // uint16[M]
    int64_t sorted, pos;                         // IdT[M]
    int64_t in_edges, out_edges, aligned;        // IdT[50][M]
    int64_t in_w;                                // uint16[50][M]
    int64_t aln_node;                            // IdT[max_seq + 32]  the alignment: node per read position
    int64_t row_bs;                              // int32[M + 1]   band start per score row
    int64_t rowinfo;                             // uint4[M + 40] per-row records of the packed fill (poa_fill2.cuh);
                                                 // the first formulation reads it as uint32[3][M + 40]
    int64_t cscore;                              // int32[M]       consensus scores / racon marks scratch
    int64_t cpred;                               // IdT[M]         consensus predecessors / racon stack
    int64_t marks, check;                        // uint8[M]
    int64_t msa_pos;                             // IdT[M]
    int64_t stack;                               // IdT[4 * M]     DFS stack of the SPOA-order topological sort
    int64_t path;                                // IdT[max_reads][max_seq]  node visited by each base (MSA only)
    int32_t M;                                   // node capacity (max_nodes_per_window[_banded])
    int32_t aln_cap;                             // alignment capacity
    int32_t max_reads;
    int32_t max_seq;
};

// Per-window state carried between the launches of the phase-split schedule (poa_phases.cuh).
struct PhaseState
{
    int32_t status;     // kSt* value; anything but success stops the window
    int32_t node_count;
    int32_t finalized;  // outputs written
    int32_t aln_len;    // kAlnPendingTraceback, or the result of the exact path
    int32_t start_row;  // row the packed traceback starts from
    int32_t graph_step; // last step whose graph phase ran (a step has two graph launches, a window takes one)
    uint32_t rows, exact;
    unsigned long long cells;
    int64_t seq_off;    // offset of the current read's first base from the window's first base
    int32_t beta;       // row-offset slope of the packed fill that produced the pending alignment (0: none)
    int32_t aln_score;  // score of the pending alignment (adaptive banding decides on it)
};
constexpr int32_t kAlnPendingTraceback = -3;
constexpr int32_t kAlnNeedsExact       = -4;  // the packed fill declined: poa_exact_kernel aligns this read
constexpr int32_t kAlnNeedsRebased     = -5;  // long window: poa_fill_rebased_kernel fills this read
constexpr int32_t kClassMargin         = 64;  // nodes one fusion is allowed to add before the scratch is outgrown
constexpr int32_t kPhaseCounters       = 4096; // one window counter per launch of a batch

struct KernelParams
{
    const uint8_t* bases;
    const int8_t* weights; // nullptr: every weight is 1
    const int32_t* read_len;
    const WindowDesc* windows;
    int32_t n_windows;

    uint8_t* arena;
    ArenaLayout layout;
    void* scores;                 // ScoreT pool
    int64_t banded_score_stride;  // elements per window in banded mode

    uint8_t* status;       // per window: 0 or a StatusType value
    int32_t* node_count;   // per window: nodes in the final graph
    uint8_t* consensus;    // per window max_consensus bytes (written end -> start, NUL terminated)
    uint16_t* coverage;    // per window max_consensus entries
    uint8_t* msa;          // per window max_reads * max_consensus bytes
    uint32_t* stat_rows;   // per window: sum of graph rows over its alignments
    uint32_t* stat_exact;  // per window: alignments that took the exact path
    unsigned long long* stat_cells; // per window: sum of N_s * Wc (Wc = band, or the read length in full mode)

    int32_t max_nodes;
    int32_t max_consensus;
    int32_t max_reads;
    int32_t band;
    int32_t gap, mismatch, match;
    uint32_t gap_x2;       // the gap score in both 16-bit halves (operand of the packed fill)
    int32_t spoa_accurate;
    int32_t* next_window; // zeroed before the launch; warps take window ids from it
    int32_t adaptive_max_band; // 0: static band; else the widest band the score rows can hold
    int32_t smem_per_warp; // bytes of shared-memory scratch each warp (= window) owns
    PhaseState* phase;     // phase-split schedule: per-window state, else nullptr
    int32_t step;          // phase-split schedule: the read (1 .. n_reads - 1) this launch works on
    int32_t phase_class;   // graph launches come in two shared-memory sizes: 0 takes the windows whose graph fits
    int32_t class_cap;     // ... the smaller scratch (node_count + kClassMargin <= class_cap), 1 the others
    int32_t score_rows;    // rows of a window's score matrix (full alignment: its area is scores_width x score_rows cells)
    int32_t class_cap_final; // the same bound for a window's last read, whose launch also runs the consensus
    int32_t lean_class0;   // launch 0 is the lean instance: rebased rows and the SPOA_ACCURATE order go to launch 1
    int32_t debug_flags;   // development switches (CUDAPOA_B200_DEBUG_FLAGS), 0 in production
};

inline int64_t align_up64(int64_t v, int64_t a)
{
    return (v + a - 1) / a * a;
}

// id_bytes: sizeof(IdT); msa: whether the per-read node paths are needed.
inline ArenaLayout make_arena_layout(int32_t M, int32_t aln_cap, int32_t max_reads, int32_t max_seq, int32_t id_bytes,
                                     bool msa)
{
    ArenaLayout L{};
    L.M         = M;
    L.aln_cap   = aln_cap;
    L.max_reads = max_reads;
    L.max_seq   = max_seq;
    int64_t off = 0;
    auto take   = [&](int64_t bytes) {
        int64_t at = off;
        off        = align_up64(off + bytes, 16);
        return at;
    };
    const int64_t m = M;
    L.nodes         = take(m);
    L.in_cnt        = take(2 * m);
    L.out_cnt       = take(2 * m);
    L.aln_cnt       = take(2 * m);
    L.cov           = take(2 * m);
    L.kahn          = take(2 * m);
    L.sorted        = take(id_bytes * m);
    L.pos           = take(id_bytes * m);
    L.in_edges      = take(id_bytes * m * kMaxNodeEdges);
    L.out_edges     = take(id_bytes * m * kMaxNodeEdges);
    L.aligned       = take(id_bytes * m * kMaxNodeAlignments);
    L.in_w          = take(2 * m * kMaxNodeEdges);
    L.aln_node      = take(static_cast<int64_t>(id_bytes) * (max_seq + 32));
    L.row_bs        = take(4 * (m + 1));
    L.rowinfo       = take(16 * (m + 40));
    L.cscore        = take(4 * m);
    L.cpred         = take(id_bytes * (m + 8));
    L.marks         = take(m);
    L.check         = take(m);
    L.msa_pos       = take(id_bytes * m);
    L.stack         = take(id_bytes * m * 4);
    L.path          = take(msa ? static_cast<int64_t>(id_bytes) * max_reads * max_seq : 0);
    L.stride        = align_up64(off, 256);
    return L;
}

} // namespace poa_b200
