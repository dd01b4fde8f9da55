// TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference cudapoa algorithm.
//
// A sequential, single-threaded C++ restatement of GenomeWorks 0.5.0 cudapoa/src (the kernels are
// warp-parallel CUDA; here every warp-wide step is written as the plain loop it is equivalent to).
// Each function cites the reference file:line it follows.  Parity pins:
//   * the reference's own known-answer vectors (cudapoa/tests/Test_Cudapoa{NW,TopSort,AddAlignment,
//     GenerateConsensus,Batch}.cu), replayed by tests/test_oracle_golden.py;
//   * outputs of the UNMODIFIED reference built for sm_100a (oracle/_ref/libcudapoa_ref.so) on seeded
//     windows, stored under tests/golden/ by tests/golden/make_reference_golden.py.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may use this file's library.

#include "cudapoa_oracle.h"

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <string>
#include <tuple>
#include <vector>

namespace
{

constexpr int32_t kMaxEdges      = 50; // CUDAPOA_MAX_NODE_EDGES, cudapoa_structs.cuh:18
constexpr int32_t kMaxAlignments = 50; // CUDAPOA_MAX_NODE_ALIGNMENTS, cudapoa_structs.cuh:21
constexpr int32_t kCellsPerLane  = 4;  // CELLS_PER_THREAD, cudapoa_structs.cuh:25
constexpr int32_t kAdaptiveBandFactor = 4; // adaptive banding (engine extension): the band doubles up to 4x ...
constexpr int32_t kAdaptiveBandMargin = 4; // ... while an aligned pair is this close to a band edge
constexpr int32_t kRightPad      = 8;  // CUDAPOA_BANDED_MATRIX_RIGHT_PADDING, cudapoa_structs.cuh:27

// cudapoa::StatusType values (cudapoa.hpp:26-38)
enum Status : int32_t
{
    kSuccess            = 0,
    kExceededSeqSize    = 2,
    kNodeCountExceeded  = 4,
    kEdgeCountExceeded  = 5,
    kLoopCountExceeded  = 7,
};

inline int32_t align_up(int32_t v, int32_t a)
{
    return ((v + a - 1) / a) * a;
}

// ScoreT arithmetic as the device code performs it: operands promoted to int, results converted back
// to ScoreT on assignment (two's-complement truncation for int16; int32 sums wrap).
template <typename ScoreT>
inline ScoreT narrow(int64_t v)
{
    return static_cast<ScoreT>(static_cast<typename std::make_unsigned<ScoreT>::type>(v));
}

template <typename ScoreT>
inline int64_t add_as_device(ScoreT a, int32_t b)
{
    // int16 + int16 is an int addition on the device; int32 + int32 wraps.
    if (sizeof(ScoreT) == 2)
        return static_cast<int64_t>(a) + b;
    return static_cast<int32_t>(static_cast<uint32_t>(a) + static_cast<uint32_t>(b));
}

// ------------------------------------------------------------------------------------------------
// Band geometry, cudapoa_nw_banded.cuh:28-48.  gradient comes from banded_gradient() below.
// ------------------------------------------------------------------------------------------------
inline int32_t band_start_for_row(int32_t row, float gradient, int32_t band_width, int32_t max_column)
{
    // SizeT(row_idx * gradient): FMUL.FTZ then F2I.TRUNC (fast-math build; operands are far from denormals)
    float prod        = static_cast<float>(row) * gradient;
    int32_t start_pos = static_cast<int32_t>(prod) - band_width / 2;
    start_pos         = std::max(start_pos, 0);
    int32_t end_pos   = start_pos + band_width;
    if (end_pos > max_column)
        start_pos = max_column - band_width + kCellsPerLane;
    start_pos = std::max(start_pos, 0);
    start_pos = start_pos - (start_pos % kCellsPerLane);
    return start_pos;
}

struct RcpTable
{
    const float* v = nullptr;
    int32_t n      = 0;
};

// cudapoa_nw_banded.cuh:206 under -use_fast_math: FMUL.FTZ(float(L+1), MUFU.RCP(float(N+1))).
inline float banded_gradient(int32_t read_length, int32_t graph_count, const RcpTable& rcp, int64_t* misses)
{
    float num = static_cast<float>(read_length + 1);
    int32_t d = graph_count + 1;
    if (rcp.v != nullptr && d >= 0 && d < rcp.n)
        return num * rcp.v[d];
    if (misses != nullptr)
        (*misses)++;
    return num / static_cast<float>(d);
}

// Score matrix in the reference's banded layout: row r occupies W = band + 8 cells.
template <typename ScoreT>
struct BandedMatrix
{
    ScoreT* s;
    float gradient;
    int32_t band;
    int32_t max_column;
    const int32_t* bs_cache; // band_start_for_row(row) for row = 0..bs_rows-1 (pure function of the row)
    int32_t bs_rows;
    int32_t width() const { return band + kRightPad; }

    int32_t bs(int32_t row) const
    {
        if (row >= 0 && row < bs_rows)
            return bs_cache[row];
        return band_start_for_row(row, gradient, band, max_column);
    }

    // get_score_ptr, cudapoa_nw_banded.cuh:50-68
    int64_t index_for_read(int32_t row, int32_t column) const
    {
        int32_t col_idx = (column == 0) ? 0 : column - bs(row);
        return static_cast<int64_t>(col_idx) + static_cast<int64_t>(row) * width();
    }
    // set_score, cudapoa_nw_banded.cuh:70-87 (column 0 is stored at offset band_start, a reference quirk)
    void set(int32_t row, int32_t column, ScoreT v)
    {
        int32_t b       = bs(row);
        int32_t col_idx = (column == 0) ? b : column - b;
        s[static_cast<int64_t>(col_idx) + static_cast<int64_t>(row) * width()] = v;
    }
    // get_score, cudapoa_nw_banded.cuh:106-120
    ScoreT get(int32_t row, int32_t column, ScoreT min_score_value) const
    {
        int32_t b   = bs(row);
        int32_t end = b + band;
        if (((column > end) || (column < b)) && column != 0)
            return min_score_value;
        return s[index_for_read(row, column)];
    }
    // initialize_band, cudapoa_nw_banded.cuh:89-104
    void init_band(int32_t row, ScoreT v)
    {
        int32_t b   = bs(row);
        int32_t end = b + band;
        set(row, (b == 0) ? 1 : b, v);
        for (int32_t j = end; j < end + kRightPad; j++)
            set(row, j, v);
    }
};

template <typename ScoreT>
inline ScoreT banded_min_score(int32_t gap, int32_t mismatch, int32_t match)
{
    // cudapoa_nw_banded.cuh:199-200
    int32_t m = std::min(std::min(gap, mismatch), -match) - 1;
    return narrow<ScoreT>(static_cast<int64_t>(2) * std::abs(m) + std::numeric_limits<ScoreT>::min());
}

// runNeedlemanWunschBanded, cudapoa_nw_banded.cuh:176-487.  Returns the alignment length or -1.
template <typename ScoreT, typename IdT>
int32_t nw_banded(const uint8_t* nodes, const IdT* graph, const IdT* node_id_to_pos, int32_t graph_count,
                  const uint16_t* in_cnt, const IdT* in_edges, const uint16_t* out_cnt, const uint8_t* read,
                  int32_t read_length, ScoreT* scores, IdT* aln_graph, IdT* aln_read, int32_t band_width,
                  int32_t gap, int32_t mismatch, int32_t match, float gradient, int32_t* score_out = nullptr)
{
    const ScoreT type_min  = std::numeric_limits<ScoreT>::min();
    const ScoreT min_score = banded_min_score<ScoreT>(gap, mismatch, match);
    std::vector<int32_t> band_starts(static_cast<size_t>(graph_count) + 1);
    for (int32_t r = 0; r <= graph_count; r++)
        band_starts[static_cast<size_t>(r)] = band_start_for_row(r, gradient, band_width, read_length + 1);
    BandedMatrix<ScoreT> M{scores, gradient, band_width, read_length + 1, band_starts.data(), graph_count + 1};
    const int32_t W = M.width();

    // horizontal boundary, :208-212
    for (int32_t j = 0; j < W; j++)
        M.set(0, j, narrow<ScoreT>(static_cast<int64_t>(j) * gap));

    // vertical boundary, :214-249 (lane 0)
    for (int32_t graph_pos = 0; graph_pos < graph_count; graph_pos++)
    {
        M.set(0, 0, 0);
        int32_t node_id = graph[graph_pos];
        int32_t i       = graph_pos + 1;
        uint16_t pc     = in_cnt[node_id];
        if (pc == 0)
        {
            M.set(i, 0, narrow<ScoreT>(gap));
        }
        else
        {
            ScoreT penalty = type_min;
            for (uint16_t p = 0; p < pc; p++)
            {
                int32_t pred     = in_edges[node_id * kMaxEdges + p];
                int32_t pred_pos = node_id_to_pos[pred] + 1;
                penalty          = std::max(penalty, M.get(pred_pos, 0, min_score));
            }
            M.set(i, 0, narrow<ScoreT>(add_as_device<ScoreT>(penalty, gap)));
        }
    }

    // row fill, :253-366
    for (int32_t graph_pos = 0; graph_pos < graph_count; graph_pos++)
    {
        int32_t node_id = graph[graph_pos];
        int32_t row     = graph_pos + 1;
        int32_t b       = M.bs(row);
        M.init_band(row, min_score);
        ScoreT carry    = M.get(row, 0, min_score); // first_element_prev_score
        uint16_t pc     = in_cnt[node_id];
        uint8_t base    = nodes[node_id];

        // one 4-cell group at a time; groups are independent before the horizontal pass
        for (int32_t read_pos = b; read_pos < b + band_width; read_pos += kCellsPerLane)
        {
            ScoreT cell[4];
            int32_t profile[4];
            for (int k = 0; k < 4; k++)
                profile[k] = (read_pos + k < read_length && base == read[read_pos + k]) ? match : mismatch;
            // (the device reads whatever bytes follow the read here; those cells lie right of column L and
            //  nothing left of them, the sink column or the traceback ever depends on them)

            for (uint16_t p = 0; p < std::max<uint16_t>(pc, 1); p++)
            {
                int32_t pred_row = (pc == 0) ? 0 : node_id_to_pos[in_edges[node_id * kMaxEdges + p]] + 1;
                // get_scores, :122-174
                int32_t pb   = M.bs(pred_row);
                int32_t pend = pb + band_width + kCellsPerLane;
                ScoreT cand[4];
                if (((read_pos + 1 > pend) || (read_pos + 1 < pb)) && (read_pos + 1 != 0))
                {
                    for (int k = 0; k < 4; k++)
                        cand[k] = min_score;
                }
                else
                {
                    const ScoreT* ps = &scores[M.index_for_read(pred_row, read_pos)];
                    for (int k = 0; k < 4; k++)
                    {
                        int64_t diag = add_as_device<ScoreT>(ps[k], profile[k]);
                        int64_t vert = add_as_device<ScoreT>(ps[k + 1], gap);
                        cand[k]      = narrow<ScoreT>(std::max(diag, vert));
                    }
                }
                for (int k = 0; k < 4; k++)
                    cell[k] = (p == 0) ? cand[k] : std::max(cell[k], cand[k]);
            }

            // horizontal pass, :301-351: the warp fix-point converges to the prefix recurrence
            // H[j] = max(H[j], H[j-1] + gap), carried across groups and 128-cell chunks.
            for (int k = 0; k < 4; k++)
            {
                int64_t t = std::max<int64_t>(add_as_device<ScoreT>(carry, gap), cell[k]);
                ScoreT ts = narrow<ScoreT>(t);
                if (ts > cell[k])
                    cell[k] = ts;
                carry = cell[k];
            }
            int64_t idx = static_cast<int64_t>(read_pos + 1 - b) + static_cast<int64_t>(row) * W;
            for (int k = 0; k < 4; k++)
                scores[idx + k] = cell[k];
        }
    }

    // sink selection, :369-384
    int32_t i = 0, j = read_length;
    ScoreT mscore = type_min;
    for (int32_t idx = 1; idx <= graph_count; idx++)
    {
        if (out_cnt[graph[idx - 1]] == 0)
        {
            ScoreT s = M.get(idx, j, min_score);
            if (mscore < s)
            {
                mscore = s;
                i      = idx;
            }
        }
    }

    if (score_out != nullptr) // score of the cell the traceback starts from (adaptive banding, engine extension)
        *score_out = static_cast<int32_t>(M.get(i, j, min_score));

    // traceback, :386-480
    int32_t aligned = 0;
    int32_t prev_i = 0, prev_j = 0;
    int32_t loop_count = 0;
    while (!(i == 0 && j == 0) && loop_count < read_length + graph_count + 2)
    {
        loop_count++;
        ScoreT sij      = M.get(i, j, min_score);
        bool pred_found = false;
        if (i != 0 && j != 0)
        {
            int32_t node_id = graph[i - 1];
            int32_t cost    = (nodes[node_id] == read[j - 1]) ? match : mismatch;
            uint16_t pc     = in_cnt[node_id];
            for (uint16_t p = 0; p < std::max<uint16_t>(pc, 1) && !pred_found; p++)
            {
                int32_t pred_i = (pc == 0) ? 0 : node_id_to_pos[in_edges[node_id * kMaxEdges + p]] + 1;
                if (static_cast<int64_t>(sij) == add_as_device<ScoreT>(M.get(pred_i, j - 1, min_score), cost))
                {
                    prev_i     = pred_i;
                    prev_j     = j - 1;
                    pred_found = true;
                }
            }
        }
        if (!pred_found && i != 0)
        {
            int32_t node_id = graph[i - 1];
            uint16_t pc     = in_cnt[node_id];
            for (uint16_t p = 0; p < std::max<uint16_t>(pc, 1) && !pred_found; p++)
            {
                int32_t pred_i = (pc == 0) ? 0 : node_id_to_pos[in_edges[node_id * kMaxEdges + p]] + 1;
                if (static_cast<int64_t>(sij) == add_as_device<ScoreT>(M.get(pred_i, j, min_score), gap))
                {
                    prev_i     = pred_i;
                    prev_j     = j;
                    pred_found = true;
                }
            }
        }
        if (!pred_found &&
            static_cast<int64_t>(sij) == add_as_device<ScoreT>(M.get(i, j - 1, min_score), gap))
        {
            prev_i     = i;
            prev_j     = j - 1;
            pred_found = true;
        }
        aln_graph[aligned] = static_cast<IdT>((i == prev_i) ? -1 : graph[i - 1]);
        aln_read[aligned]  = static_cast<IdT>((j == prev_j) ? -1 : j - 1);
        aligned++;
        i = prev_i;
        j = prev_j;
    }
    if (loop_count >= read_length + graph_count + 2)
        aligned = -1;
    return aligned;
}

// runNeedlemanWunsch + computeScore, cudapoa_nw.cuh:41-466 (unbanded).  scores is row-major with stride
// scores_width.  Returns the alignment length or -1.
template <typename ScoreT, typename IdT>
int32_t nw_full(const uint8_t* nodes, const IdT* graph, const IdT* node_id_to_pos, int32_t graph_count,
                const uint16_t* in_cnt, const IdT* in_edges, const uint16_t* out_cnt, const uint8_t* read,
                int32_t read_length, ScoreT* scores, int32_t scores_width, IdT* aln_graph, IdT* aln_read,
                int32_t gap, int32_t mismatch, int32_t match)
{
    const ScoreT type_min = std::numeric_limits<ScoreT>::min();
    auto at               = [&](int32_t row, int32_t col) -> ScoreT& {
        return scores[static_cast<int64_t>(row) * scores_width + col];
    };

    // boundaries, :175-210
    for (int32_t j = 0; j < read_length + 1; j++)
        at(0, j) = narrow<ScoreT>(static_cast<int64_t>(j) * gap);
    for (int32_t graph_pos = 0; graph_pos < graph_count; graph_pos++)
    {
        int32_t node_id = graph[graph_pos];
        int32_t i       = graph_pos + 1;
        uint16_t pc     = in_cnt[node_id];
        if (pc == 0)
        {
            at(i, 0) = narrow<ScoreT>(gap);
        }
        else
        {
            ScoreT penalty = type_min;
            for (uint16_t p = 0; p < pc; p++)
            {
                int32_t pred_pos = node_id_to_pos[in_edges[node_id * kMaxEdges + p]] + 1;
                penalty          = std::max(penalty, at(pred_pos, 0));
            }
            at(i, 0) = narrow<ScoreT>(add_as_device<ScoreT>(penalty, gap));
        }
    }

    // row fill, :222-326.  Only columns 1..read_length are restated: the device also computes up to three
    // cells right of the read in the last group, which nothing to their left ever reads.
    for (int32_t graph_pos = 0; graph_pos < graph_count; graph_pos++)
    {
        int32_t node_id = graph[graph_pos];
        int32_t row     = graph_pos + 1;
        ScoreT carry    = at(row, 0);
        uint16_t pc     = in_cnt[node_id];
        uint8_t base    = nodes[node_id];
        for (int32_t j = 1; j <= read_length; j++)
        {
            int32_t profile = (base == read[j - 1]) ? match : mismatch;
            int64_t best    = 0;
            for (uint16_t p = 0; p < std::max<uint16_t>(pc, 1); p++)
            {
                int32_t pred_row = (pc == 0) ? 0 : node_id_to_pos[in_edges[node_id * kMaxEdges + p]] + 1;
                int64_t diag     = add_as_device<ScoreT>(at(pred_row, j - 1), profile);
                int64_t vert     = add_as_device<ScoreT>(at(pred_row, j), gap);
                int64_t cand     = std::max(diag, vert);
                // computeScore keeps a ScoreT running value between predecessors (:85-117)
                best = (p == 0) ? static_cast<int64_t>(narrow<ScoreT>(cand))
                                : static_cast<int64_t>(narrow<ScoreT>(std::max(best, cand)));
            }
            ScoreT cell = narrow<ScoreT>(best);
            int64_t t   = std::max<int64_t>(add_as_device<ScoreT>(carry, gap), cell);
            ScoreT ts   = narrow<ScoreT>(t);
            if (ts > cell)
                cell = ts;
            carry     = cell;
            at(row, j) = cell;
        }
    }

    // sink selection, :330-348
    int32_t i = 0, j = read_length;
    ScoreT mscore = type_min;
    for (int32_t idx = 1; idx <= graph_count; idx++)
    {
        if (out_cnt[graph[idx - 1]] == 0)
        {
            ScoreT s = at(idx, j);
            if (mscore < s)
            {
                mscore = s;
                i      = idx;
            }
        }
    }

    // traceback, :350-457
    int32_t aligned = 0, prev_i = 0, prev_j = 0, loop_count = 0;
    while (!(i == 0 && j == 0) && loop_count < read_length + graph_count + 2)
    {
        loop_count++;
        ScoreT sij      = at(i, j);
        bool pred_found = false;
        if (i != 0 && j != 0)
        {
            int32_t node_id = graph[i - 1];
            int32_t cost    = (nodes[node_id] == read[j - 1]) ? match : mismatch;
            uint16_t pc     = in_cnt[node_id];
            for (uint16_t p = 0; p < std::max<uint16_t>(pc, 1) && !pred_found; p++)
            {
                int32_t pred_i = (pc == 0) ? 0 : node_id_to_pos[in_edges[node_id * kMaxEdges + p]] + 1;
                if (static_cast<int64_t>(sij) == add_as_device<ScoreT>(at(pred_i, j - 1), cost))
                {
                    prev_i     = pred_i;
                    prev_j     = j - 1;
                    pred_found = true;
                }
            }
        }
        if (!pred_found && i != 0)
        {
            int32_t node_id = graph[i - 1];
            uint16_t pc     = in_cnt[node_id];
            for (uint16_t p = 0; p < std::max<uint16_t>(pc, 1) && !pred_found; p++)
            {
                int32_t pred_i = (pc == 0) ? 0 : node_id_to_pos[in_edges[node_id * kMaxEdges + p]] + 1;
                if (static_cast<int64_t>(sij) == add_as_device<ScoreT>(at(pred_i, j), gap))
                {
                    prev_i     = pred_i;
                    prev_j     = j;
                    pred_found = true;
                }
            }
        }
        // horizontal move, :437-443.  At j == 0 the device reads scores[i*w - 1], the cell before the row;
        // it can only matter on a row whose column-0 value has no vertical explanation, which the boundary
        // initialisation rules out for i > 0.
        if (!pred_found && j > 0 && static_cast<int64_t>(sij) == add_as_device<ScoreT>(at(i, j - 1), gap))
        {
            prev_i     = i;
            prev_j     = j - 1;
            pred_found = true;
        }
        aln_graph[aligned] = static_cast<IdT>((i == prev_i) ? -1 : graph[i - 1]);
        aln_read[aligned]  = static_cast<IdT>((j == prev_j) ? -1 : j - 1);
        aligned++;
        i = prev_i;
        j = prev_j;
    }
    if (loop_count >= read_length + graph_count + 2)
        aligned = -1;
    return aligned;
}

// topologicalSortDeviceUtil, cudapoa_topsort.cuh:38-88 (Kahn; sorted_poa doubles as the FIFO).
template <typename IdT>
void topsort_kahn(IdT* sorted, IdT* node_id_to_pos, int32_t node_count, const uint16_t* in_cnt,
                  const IdT* out_edges, const uint16_t* out_cnt, uint16_t* local_in)
{
    int32_t tail = 0;
    for (int32_t n = 0; n < node_count; n++)
    {
        local_in[n] = in_cnt[n];
        if (local_in[n] == 0)
        {
            node_id_to_pos[n] = static_cast<IdT>(tail);
            sorted[tail++]    = static_cast<IdT>(n);
        }
    }
    for (int32_t head = 0; head < tail; head++)
    {
        int32_t node = sorted[head];
        for (uint16_t e = 0; e < out_cnt[node]; e++)
        {
            int32_t child = out_edges[node * kMaxEdges + e];
            local_in[child]--;
            if (local_in[child] == 0)
            {
                node_id_to_pos[child] = static_cast<IdT>(tail);
                sorted[tail++]        = static_cast<IdT>(child);
            }
        }
    }
}

// raconTopologicalSortDeviceUtil, cudapoa_topsort.cuh:94-189 (SPOA's iterative DFS; aligned nodes contiguous).
template <typename IdT>
void topsort_racon(IdT* sorted, IdT* node_id_to_pos, int32_t node_count, const uint16_t* in_cnt,
                   const IdT* in_edges, const uint16_t* aligned_cnt, const IdT* aligned, int32_t max_nodes)
{
    std::vector<uint8_t> marks(static_cast<size_t>(max_nodes), 0);
    std::vector<uint8_t> check_aligned(static_cast<size_t>(max_nodes), 1);
    std::vector<int32_t> stack;
    int32_t out = 0;
    for (int32_t i = 0; i < node_count; i++)
    {
        if (marks[i] != 0)
            continue;
        stack.push_back(i);
        while (!stack.empty())
        {
            int32_t node_id = stack.back();
            bool valid      = true;
            if (marks[node_id] != 2)
            {
                for (uint16_t e = 0; e < in_cnt[node_id]; e++)
                {
                    int32_t begin = in_edges[node_id * kMaxEdges + e];
                    if (marks[begin] != 2)
                    {
                        stack.push_back(begin);
                        valid = false;
                    }
                }
                if (check_aligned[node_id])
                {
                    for (uint16_t a = 0; a < aligned_cnt[node_id]; a++)
                    {
                        int32_t aid = aligned[node_id * kMaxAlignments + a];
                        if (marks[aid] != 2)
                        {
                            stack.push_back(aid);
                            check_aligned[aid] = 0;
                            valid              = false;
                        }
                    }
                }
                if (valid)
                {
                    marks[node_id] = 2;
                    if (check_aligned[node_id])
                    {
                        sorted[out]             = static_cast<IdT>(node_id);
                        node_id_to_pos[node_id] = static_cast<IdT>(out);
                        out++;
                        for (uint16_t a = 0; a < aligned_cnt[node_id]; a++)
                        {
                            int32_t aid         = aligned[node_id * kMaxAlignments + a];
                            sorted[out]         = static_cast<IdT>(aid);
                            node_id_to_pos[aid] = static_cast<IdT>(out);
                            out++;
                        }
                    }
                }
                else
                {
                    marks[node_id] = 1;
                }
            }
            if (valid)
            {
                // the device pops by decrementing its stack index; entries pushed above node_id were
                // already consumed, so node_id is on top here
                stack.pop_back();
            }
        }
    }
}

// addAlignmentToGraph<msa>, cudapoa_add_alignment.cuh:59-279.  Returns a StatusType value.
// edge_reads / edge_read_cnt are the MSA per-edge read lists (outgoing_edges_coverage[_count]); pass
// nullptr when msa is off.
template <typename IdT>
int32_t add_alignment(int32_t& new_node_count, uint8_t* nodes, int32_t node_count, IdT* aligned,
                      uint16_t* aligned_cnt, IdT* in_edges, uint16_t* in_cnt, IdT* out_edges, uint16_t* out_cnt,
                      uint16_t* in_w, int32_t aln_len, const IdT* aln_graph, const uint8_t* read,
                      const IdT* aln_read, uint16_t* coverage, const int8_t* base_weights, bool msa,
                      IdT* begin_node_of_read, uint16_t* edge_reads, uint16_t* edge_read_cnt, uint16_t read_id,
                      int32_t max_reads, int32_t max_nodes)
{
    int32_t head         = -1;
    int32_t curr         = -1;
    uint16_t prev_weight = 0;
    for (int32_t pos = aln_len - 1; pos >= 0; pos--)
    {
        int32_t read_pos = aln_read[pos];
        if (read_pos == -1)
            continue;
        int8_t w          = base_weights ? base_weights[read_pos] : static_cast<int8_t>(1);
        uint8_t read_base = read[read_pos];
        int32_t gnode     = aln_graph[pos];
        auto new_node     = [&]() -> bool {
            curr = node_count++;
            if (node_count >= max_nodes)
                return false;
            nodes[curr]       = read_base;
            out_cnt[curr]     = 0;
            in_cnt[curr]      = 0;
            aligned_cnt[curr] = 0;
            coverage[curr]    = 0;
            return true;
        };
        if (gnode == -1)
        {
            if (!new_node())
                return kNodeCountExceeded;
        }
        else if (nodes[gnode] == read_base)
        {
            curr = gnode;
        }
        else
        {
            uint16_t n_aligned = aligned_cnt[gnode];
            int32_t match_id   = -1;
            for (uint16_t n = 0; n < n_aligned; n++)
            {
                int32_t aid = aligned[gnode * kMaxAlignments + n];
                if (nodes[aid] == read_base)
                {
                    match_id = aid;
                    break;
                }
            }
            if (match_id != -1)
            {
                curr = match_id;
            }
            else
            {
                if (!new_node())
                    return kNodeCountExceeded;
                int32_t k = 0;
                for (uint16_t n = 0; n < n_aligned; n++)
                {
                    int32_t aid                                = aligned[gnode * kMaxAlignments + n];
                    uint16_t c                                 = aligned_cnt[aid];
                    aligned[aid * kMaxAlignments + c]          = static_cast<IdT>(curr);
                    aligned_cnt[aid]                           = c + 1;
                    aligned[curr * kMaxAlignments + k]         = static_cast<IdT>(aid);
                    k++;
                }
                aligned[gnode * kMaxAlignments + n_aligned] = static_cast<IdT>(curr);
                aligned_cnt[gnode]                          = n_aligned + 1;
                aligned[curr * kMaxAlignments + k]          = static_cast<IdT>(gnode);
                k++;
                aligned_cnt[curr] = static_cast<uint16_t>(k);
            }
        }

        if (msa && read_pos == 0)
            *begin_node_of_read = static_cast<IdT>(curr);

        if (head != -1)
        {
            bool exists  = false;
            uint16_t inc = in_cnt[curr];
            for (uint16_t e = 0; e < inc; e++)
            {
                if (in_edges[curr * kMaxEdges + e] == head)
                {
                    exists = true;
                    in_w[curr * kMaxEdges + e] = static_cast<uint16_t>(in_w[curr * kMaxEdges + e] + (prev_weight + w));
                }
            }
            if (!exists)
            {
                in_edges[curr * kMaxEdges + inc]  = static_cast<IdT>(head);
                in_w[curr * kMaxEdges + inc]      = static_cast<uint16_t>(prev_weight + w);
                in_cnt[curr]                      = inc + 1;
                uint16_t outc                     = out_cnt[head];
                out_edges[head * kMaxEdges + outc] = static_cast<IdT>(curr);
                if (msa)
                {
                    edge_read_cnt[head * kMaxEdges + outc]                                  = 1;
                    edge_reads[(static_cast<int64_t>(head) * kMaxEdges + outc) * max_reads] = read_id;
                }
                out_cnt[head] = outc + 1;
                if (outc + 1 >= kMaxEdges || inc + 1 >= kMaxEdges)
                    return kEdgeCountExceeded;
            }
            else if (msa)
            {
                uint16_t outc = out_cnt[head];
                for (uint16_t e = 0; e < outc; e++)
                {
                    if (out_edges[head * kMaxEdges + e] == curr)
                    {
                        uint16_t c = edge_read_cnt[head * kMaxEdges + e];
                        edge_reads[(static_cast<int64_t>(head) * kMaxEdges + e) * max_reads + c] = read_id;
                        edge_read_cnt[head * kMaxEdges + e]                                      = c + 1;
                        break;
                    }
                }
            }
        }
        head = curr;
        coverage[head]++;
        prev_weight = static_cast<uint16_t>(w);
    }
    new_node_count = node_count;
    return kSuccess;
}

// branchCompletion, cudapoa_generate_consensus.cuh:28-112
template <typename IdT>
int32_t branch_completion(int32_t max_pos, int32_t node_count, const IdT* graph, const IdT* in_edges,
                          const uint16_t* in_cnt, const IdT* out_edges, const uint16_t* out_cnt,
                          const uint16_t* in_w, int32_t* scores, int32_t* preds)
{
    int32_t node_id = graph[max_pos];
    for (uint16_t oe = 0; oe < out_cnt[node_id]; oe++)
    {
        int32_t out_node = out_edges[node_id * kMaxEdges + oe];
        for (uint16_t ie = 0; ie < in_cnt[out_node]; ie++)
        {
            int32_t id = in_edges[out_node * kMaxEdges + ie];
            if (id != node_id)
                scores[id] = -1;
        }
    }
    int32_t max_score = 0, max_id = 0;
    for (int32_t pos = max_pos + 1; pos < node_count; pos++)
    {
        node_id        = graph[pos];
        preds[node_id] = -1;
        int32_t score  = -1;
        for (uint16_t e = 0; e < in_cnt[node_id]; e++)
        {
            int32_t begin = in_edges[node_id * kMaxEdges + e];
            if (scores[begin] == -1)
                continue;
            int32_t w = in_w[node_id * kMaxEdges + e];
            if (score < w || (score == w && scores[preds[node_id]] <= scores[begin]))
            {
                score          = w;
                preds[node_id] = begin;
            }
        }
        if (preds[node_id] != -1)
            score += scores[preds[node_id]];
        if (max_score <= score)
        {
            max_score = score;
            max_id    = node_id;
        }
        scores[node_id] = score;
    }
    return max_id;
}

// generateConsensus, cudapoa_generate_consensus.cuh:134-276.  consensus/coverage are written reversed
// (end -> start) exactly like the device buffers; returns 0 or the StatusType written to consensus[1].
template <typename IdT>
int32_t generate_consensus(const uint8_t* nodes, int32_t node_count, const IdT* graph, const IdT* node_id_to_pos,
                           const IdT* in_edges, const uint16_t* in_cnt, const IdT* out_edges,
                           const uint16_t* out_cnt, const uint16_t* in_w, uint8_t* consensus, uint16_t* cov_out,
                           const uint16_t* node_cov, const IdT* aligned, const uint16_t* aligned_cnt,
                           int32_t max_consensus_size)
{
    std::vector<int32_t> preds(static_cast<size_t>(std::max(node_count, 1)), -1);
    std::vector<int32_t> scores(static_cast<size_t>(std::max(node_count, 1)), -1);
    int32_t max_id = 0, max_score = -1;
    for (int32_t pos = 0; pos < node_count; pos++)
    {
        int32_t node_id = graph[pos];
        int32_t score   = scores[node_id];
        for (uint16_t e = 0; e < in_cnt[node_id]; e++)
        {
            int32_t w     = in_w[node_id * kMaxEdges + e];
            int32_t begin = in_edges[node_id * kMaxEdges + e];
            if (score < w || (score == w && scores[preds[node_id]] <= scores[begin]))
            {
                score          = w;
                preds[node_id] = begin;
            }
        }
        if (preds[node_id] != -1)
            score += scores[preds[node_id]];
        if (max_score <= score)
        {
            max_id    = node_id;
            max_score = score;
        }
        scores[node_id] = score;
    }
    int32_t loop_count = 0;
    while (out_cnt[max_id] != 0 && loop_count < node_count)
    {
        max_id = branch_completion<IdT>(node_id_to_pos[max_id], node_count, graph, in_edges, in_cnt, out_edges,
                                        out_cnt, in_w, scores.data(), preds.data());
        loop_count++;
    }
    if (loop_count >= node_count)
    {
        consensus[0] = 0xFF;
        consensus[1] = kLoopCountExceeded;
        return kLoopCountExceeded;
    }
    auto cov_of = [&](int32_t id) {
        uint16_t c = node_cov[id];
        for (uint16_t a = 0; a < aligned_cnt[id]; a++)
            c = static_cast<uint16_t>(c + node_cov[aligned[id * kMaxAlignments + a]]);
        return c;
    };
    int32_t pos = 0, count = 0;
    while (preds[max_id] != -1)
    {
        consensus[pos] = nodes[max_id];
        cov_out[pos]   = cov_of(max_id);
        max_id         = preds[max_id];
        pos            = std::min(pos + 1, max_consensus_size - 1);
        count++;
    }
    consensus[pos] = nodes[max_id];
    cov_out[pos]   = cov_of(max_id);
    if (count >= max_consensus_size - 1)
    {
        consensus[0] = 0xFF;
        consensus[1] = kExceededSeqSize;
        return kExceededSeqSize;
    }
    pos++;
    consensus[pos] = '\0';
    return kSuccess;
}

// getNodeIDToMSAPosDevice, cudapoa_generate_msa.cuh:27-46
template <typename IdT>
int32_t msa_columns(int32_t node_count, const IdT* sorted, int32_t* msa_pos, const uint16_t* aligned_cnt)
{
    int32_t col = 0;
    for (int32_t rank = 0; rank < node_count; rank++)
    {
        int32_t node_id  = sorted[rank];
        msa_pos[node_id] = col;
        uint16_t n       = aligned_cnt[node_id];
        for (uint16_t k = 0; k < n; k++)
            msa_pos[sorted[++rank]] = col;
        col++;
    }
    return col;
}

// generateMSADevice, cudapoa_generate_msa.cuh:48-118, for one read s.
template <typename IdT>
std::string msa_row(const uint8_t* nodes, uint16_t s, const uint16_t* out_cnt, const IdT* out_edges,
                    const uint16_t* edge_reads, const uint16_t* edge_read_cnt, const int32_t* msa_pos,
                    int32_t begin_node, int32_t msa_length, int32_t max_reads)
{
    std::string row(static_cast<size_t>(msa_length), '-');
    int32_t node_id = begin_node;
    while (true)
    {
        row[static_cast<size_t>(msa_pos[node_id])] = static_cast<char>(nodes[node_id]);
        bool end_node                              = true;
        for (uint16_t n = 0; n < out_cnt[node_id] && end_node; n++)
        {
            int32_t to = out_edges[node_id * kMaxEdges + n];
            for (uint16_t m = 0; m < edge_read_cnt[node_id * kMaxEdges + n]; m++)
            {
                if (edge_reads[(static_cast<int64_t>(node_id) * kMaxEdges + n) * max_reads + m] == s)
                {
                    end_node = false;
                    node_id  = to;
                    break;
                }
            }
        }
        if (end_node)
            break;
    }
    return row;
}

struct WindowResult
{
    int32_t status = 0;
    std::string consensus;
    std::vector<uint16_t> coverage;
    std::vector<std::string> msa;
    std::string node_labels;
    std::vector<std::tuple<int32_t, int32_t, int32_t>> edges; // (sink, source, weight)
};

struct RunResult
{
    std::vector<WindowResult> windows;
    int32_t score_bits  = 16;
    int32_t size_bits   = 16;
    int64_t dp_cells    = 0;
    int64_t graph_rows  = 0;
    int64_t rcp_misses  = 0;
    int64_t widened     = 0; // adaptive banding: alignments repeated with a wider band
};

// cudapoa_limits.hpp:28-53
bool use_32bit_score(const oracle_params& p)
{
    int32_t upper = p.max_sequence_size * p.match_score;
    int32_t lower = p.max_sequence_size * std::max(p.gap_score, p.mismatch_score) +
                    (p.max_matrix_graph_dimension - p.max_sequence_size) * p.gap_score;
    return (upper > INT16_MAX || (-lower) > (INT16_MAX + 1));
}

bool use_32bit_size(const oracle_params& p)
{
    int32_t m = p.max_consensus_size;
    m         = std::max(m, p.banded ? p.max_matrix_graph_dimension_banded : p.max_matrix_graph_dimension);
    m         = std::max(m, p.max_matrix_sequence_dimension);
    return m > INT16_MAX;
}

// generatePOAKernel + generateConsensusKernel / generateMSAKernel for one window,
// cudapoa_kernels.cuh:66-359, cudapoa_generate_consensus.cuh:278-347, cudapoa_generate_msa.cuh:120-224.
template <typename ScoreT>
void run_window(const oracle_params& P, const std::vector<const uint8_t*>& reads,
                const std::vector<const int8_t*>& weights, const std::vector<int32_t>& lens, WindowResult& out,
                RunResult& totals, const RcpTable& rcp)
{
    using IdT              = int32_t;
    const bool msa         = (P.output_mask & 2) != 0;
    const int32_t max_n    = P.banded ? P.max_nodes_per_window_banded : P.max_nodes_per_window;
    const int32_t max_g    = P.banded ? P.max_matrix_graph_dimension_banded : P.max_matrix_graph_dimension;
    const int32_t n_reads  = static_cast<int32_t>(reads.size());
    const int32_t max_rds  = std::max(P.max_sequences_per_poa, 1);
    const size_t cap       = static_cast<size_t>(std::max(max_n, lens.empty() ? 1 : lens[0] + 1)) + 8;

    std::vector<uint8_t> nodes(cap, 0);
    std::vector<IdT> in_edges(cap * kMaxEdges, 0), out_edges(cap * kMaxEdges, 0), aligned(cap * kMaxAlignments, 0);
    std::vector<uint16_t> in_cnt(cap, 0), out_cnt(cap, 0), aligned_cnt(cap, 0), in_w(cap * kMaxEdges, 0);
    std::vector<uint16_t> coverage(cap, 0), local_in(cap, 0);
    std::vector<IdT> sorted(cap, 0), pos(cap, 0);
    std::vector<IdT> aln_graph(static_cast<size_t>(max_g) + cap, 0), aln_read(static_cast<size_t>(max_g) + cap, 0);
    std::vector<IdT> begin_nodes(static_cast<size_t>(max_rds) + 1, 0);
    std::vector<uint16_t> edge_reads, edge_read_cnt;
    if (msa)
    {
        edge_reads.assign(cap * kMaxEdges * static_cast<size_t>(max_rds), 0);
        edge_read_cnt.assign(cap * kMaxEdges, 0);
    }

    // score buffer: reference layout, never cleared between reads (stale_fill stands in for whatever the
    // device memory held before)
    int32_t full_width = 0;
    for (int32_t l : lens)
        full_width = std::max(full_width, align_up(l + 1 + kCellsPerLane, 4)); // cudapoa_batch.cuh:498-503
    const int32_t band_cap = P.alignment_band_width * ((P.banded && P.adaptive_band) ? kAdaptiveBandFactor : 1);
    const int32_t width    = P.banded ? band_cap + kRightPad : full_width;
    // a few hundred spare rows: the reference's column-0 quirk writes at row*W + band_start(row)
    std::vector<ScoreT> scores((static_cast<size_t>(max_g) + 64) * static_cast<size_t>(std::max(width, 1)) +
                               static_cast<size_t>(P.max_sequence_size) + 4096);
    std::memset(scores.data(), P.stale_fill & 0xFF, scores.size() * sizeof(ScoreT));

    uint8_t err[2] = {0, 0};

    // backbone, cudapoa_kernels.cuh:171-209
    int32_t node_count = lens[0];
    {
        const uint8_t* seq = reads[0];
        const int8_t* w    = weights[0];
        auto wt            = [&](int32_t i) -> int32_t { return w ? w[i] : 1; };
        if (node_count > 0)
        {
            nodes[0]       = seq[0];
            in_w[0]        = static_cast<uint16_t>(wt(0));
            coverage[0]    = 1;
            out_cnt[node_count - 1] = 0;
        }
        for (int32_t i = 1; i < node_count; i++)
        {
            nodes[i]                        = seq[i];
            sorted[i]                       = i;
            out_edges[(i - 1) * kMaxEdges]  = i;
            out_cnt[i - 1]                  = 1;
            in_edges[i * kMaxEdges]         = i - 1;
            in_w[i * kMaxEdges]             = static_cast<uint16_t>(wt(i - 1) + wt(i));
            in_cnt[i]                       = 1;
            aligned_cnt[i]                  = 0;
            pos[i]                          = i;
            coverage[i]                     = 1;
            if (msa)
            {
                edge_reads[static_cast<size_t>(i - 1) * kMaxEdges * max_rds] = 0;
                edge_read_cnt[(i - 1) * kMaxEdges]                           = 1;
            }
        }
    }

    bool failed = false;
    for (int32_t s = 1; s < n_reads && !failed; s++)
    {
        if (node_count >= max_n) // :220-226
        {
            err[0] = 0xFF;
            err[1] = kNodeCountExceeded;
            failed = true;
            break;
        }
        int32_t aln_len;
        int32_t aln_score = 0;
        if (P.banded)
        {
            float gradient = banded_gradient(lens[s], node_count, rcp, &totals.rcp_misses);
            aln_len = nw_banded<ScoreT, IdT>(nodes.data(), sorted.data(), pos.data(), node_count, in_cnt.data(),
                                             in_edges.data(), out_cnt.data(), reads[s], lens[s], scores.data(),
                                             aln_graph.data(), aln_read.data(), P.alignment_band_width, P.gap_score,
                                             P.mismatch_score, P.match_score, gradient, &aln_score);
            totals.dp_cells += static_cast<int64_t>(node_count) * P.alignment_band_width;
            if (P.adaptive_band)
            {
                // engine extension (poa_window.cuh): repeat with the band doubled when the path hugs a band edge or
                // the score is below 3/4 of a perfect match, and go on while that raises the score or the path still
                // hugs an edge
                int32_t band_now = P.alignment_band_width;
                auto touches     = [&]() {
                    const int32_t max_column = lens[s] + 1;
                    for (int32_t a = 0; a < aln_len; a++)
                    {
                        const int32_t node = aln_graph[static_cast<size_t>(a)], k = aln_read[static_cast<size_t>(a)];
                        if (node < 0 || k < 0)
                            continue;
                        const int32_t row = pos[static_cast<size_t>(node)] + 1;
                        const int32_t bs  = band_start_for_row(row, gradient, band_now, max_column);
                        const int32_t c   = k - bs;
                        if ((bs > 0 && c < kAdaptiveBandMargin) ||
                            (bs + band_now < max_column && c >= band_now - kAdaptiveBandMargin))
                            return true;
                    }
                    return false;
                };
                bool again = aln_len >= 0 && (4 * aln_score < 3 * P.match_score * lens[s] || touches());
                while (again && band_now < band_cap)
                {
                    band_now *= 2;
                    const int32_t before = aln_score;
                    aln_len = nw_banded<ScoreT, IdT>(nodes.data(), sorted.data(), pos.data(), node_count,
                                                     in_cnt.data(), in_edges.data(), out_cnt.data(), reads[s], lens[s],
                                                     scores.data(), aln_graph.data(), aln_read.data(), band_now,
                                                     P.gap_score, P.mismatch_score, P.match_score, gradient,
                                                     &aln_score);
                    totals.dp_cells += static_cast<int64_t>(node_count) * band_now;
                    totals.widened++;
                    again = aln_len >= 0 && (aln_score > before || touches());
                }
            }
        }
        else
        {
            aln_len = nw_full<ScoreT, IdT>(nodes.data(), sorted.data(), pos.data(), node_count, in_cnt.data(),
                                           in_edges.data(), out_cnt.data(), reads[s], lens[s], scores.data(), width,
                                           aln_graph.data(), aln_read.data(), P.gap_score, P.mismatch_score,
                                           P.match_score);
            totals.dp_cells += static_cast<int64_t>(node_count) * lens[s];
        }
        totals.graph_rows += node_count;
        if (aln_len == -1) // :281-289
        {
            err[0] = 0xFF;
            err[1] = kLoopCountExceeded;
            failed = true;
            break;
        }
        int32_t new_count = node_count;
        int32_t rc = add_alignment<IdT>(new_count, nodes.data(), node_count, aligned.data(), aligned_cnt.data(),
                                        in_edges.data(), in_cnt.data(), out_edges.data(), out_cnt.data(), in_w.data(),
                                        aln_len, aln_graph.data(), reads[s], aln_read.data(), coverage.data(),
                                        weights[s], msa, &begin_nodes[static_cast<size_t>(s)],
                                        msa ? edge_reads.data() : nullptr, msa ? edge_read_cnt.data() : nullptr,
                                        static_cast<uint16_t>(s), max_rds, max_n);
        if (rc != kSuccess) // :314-319
        {
            err[0] = 0xFF;
            err[1] = static_cast<uint8_t>(rc);
            failed = true;
            break;
        }
        node_count = new_count;
        if (P.spoa_accurate)
            topsort_racon<IdT>(sorted.data(), pos.data(), node_count, in_cnt.data(), in_edges.data(),
                               aligned_cnt.data(), aligned.data(), static_cast<int32_t>(cap));
        else
            topsort_kahn<IdT>(sorted.data(), pos.data(), node_count, in_cnt.data(), out_edges.data(), out_cnt.data(),
                              local_in.data());
    }

    if (failed)
    {
        out.status = err[1];
        return;
    }

    // graph snapshot (what get_graphs decodes, cudapoa_batch.cuh:359-384)
    out.node_labels.assign(reinterpret_cast<const char*>(nodes.data()), static_cast<size_t>(node_count));
    for (int32_t n = 0; n < node_count; n++)
        for (uint16_t e = 0; e < in_cnt[n]; e++)
            out.edges.emplace_back(n, in_edges[n * kMaxEdges + e], in_w[n * kMaxEdges + e]);
    std::sort(out.edges.begin(), out.edges.end());

    if (P.output_mask & 1)
    {
        std::vector<uint8_t> cons(static_cast<size_t>(P.max_consensus_size) + 4, 0);
        std::vector<uint16_t> cov(static_cast<size_t>(P.max_consensus_size) + 4, 0);
        int32_t rc = generate_consensus<IdT>(nodes.data(), node_count, sorted.data(), pos.data(), in_edges.data(),
                                             in_cnt.data(), out_edges.data(), out_cnt.data(), in_w.data(), cons.data(),
                                             cov.data(), coverage.data(), aligned.data(), aligned_cnt.data(),
                                             P.max_consensus_size);
        if (rc != kSuccess)
        {
            out.status = rc;
            return;
        }
        // host decode, cudapoa_batch.cuh:236-247: C string, reversed; coverage reversed
        std::string c(reinterpret_cast<const char*>(cons.data()));
        out.coverage.assign(cov.begin(), cov.begin() + static_cast<long>(c.size()));
        std::reverse(c.begin(), c.end());
        std::reverse(out.coverage.begin(), out.coverage.end());
        out.consensus = c;
    }
    if (msa)
    {
        topsort_racon<IdT>(sorted.data(), pos.data(), node_count, in_cnt.data(), in_edges.data(), aligned_cnt.data(),
                           aligned.data(), static_cast<int32_t>(cap));
        std::vector<int32_t> mpos(cap, 0);
        int32_t msa_len = msa_columns<IdT>(node_count, sorted.data(), mpos.data(), aligned_cnt.data());
        if (msa_len >= P.max_consensus_size) // cudapoa_generate_msa.cuh:200-204
        {
            out.status = kExceededSeqSize;
            return;
        }
        for (int32_t s = 0; s < n_reads; s++)
            out.msa.push_back(msa_row<IdT>(nodes.data(), static_cast<uint16_t>(s), out_cnt.data(), out_edges.data(),
                                           edge_reads.data(), edge_read_cnt.data(), mpos.data(),
                                           begin_nodes[static_cast<size_t>(s)], msa_len, max_rds));
    }
    out.status = kSuccess;
}

} // namespace

extern "C" {

void oracle_default_params(oracle_params* p, int32_t max_seq_sz, int32_t max_seq_per_poa, int32_t band_width)
{
    // BatchSize 3-argument constructor, batch.hpp:75-96
    std::memset(p, 0, sizeof(*p));
    p->max_sequence_size                 = max_seq_sz;
    p->max_consensus_size                = 2 * max_seq_sz;
    p->max_nodes_per_window              = align_up(3 * max_seq_sz, 4);
    p->max_nodes_per_window_banded       = align_up(4 * max_seq_sz, 4);
    p->max_matrix_graph_dimension        = align_up(p->max_nodes_per_window, 4);
    p->max_matrix_graph_dimension_banded = align_up(p->max_nodes_per_window_banded, 4);
    p->max_matrix_sequence_dimension     = align_up(max_seq_sz, 4);
    p->alignment_band_width              = align_up(band_width, 128);
    p->max_sequences_per_poa             = max_seq_per_poa;
    p->gap_score                         = -8;
    p->mismatch_score                    = -6;
    p->match_score                       = 8;
    p->banded                            = 1;
    p->output_mask                       = 1;
    p->spoa_accurate                     = 0;
    p->stale_fill                        = 0;
    p->rcp_table                         = nullptr;
    p->rcp_table_len                     = 0;
}

void* oracle_run(const uint8_t* bases, const int8_t* weights, const int32_t* read_lengths,
                 const int32_t* window_n_reads, int32_t n_windows, const oracle_params* params)
{
    auto* res = new RunResult();
    res->windows.resize(static_cast<size_t>(n_windows));
    const oracle_params& P = *params;
    const bool s32         = use_32bit_score(P);
    res->score_bits        = s32 ? 32 : 16;
    res->size_bits         = (s32 && use_32bit_size(P)) ? 32 : 16;
    RcpTable rcp{P.rcp_table, P.rcp_table_len};

    int64_t base_off = 0, read_idx = 0;
    for (int32_t w = 0; w < n_windows; w++)
    {
        std::vector<const uint8_t*> reads;
        std::vector<const int8_t*> wts;
        std::vector<int32_t> lens;
        for (int32_t r = 0; r < window_n_reads[w]; r++)
        {
            int32_t len = read_lengths[read_idx + r];
            // add_seq_to_poa, cudapoa_batch.cuh:490-539: too-long reads and reads beyond the per-POA limit are
            // skipped, the rest are packed back to back
            bool keep = (len <= P.max_sequence_size) && (static_cast<int32_t>(reads.size()) < P.max_sequences_per_poa);
            if (keep)
            {
                reads.push_back(bases + base_off);
                wts.push_back(weights ? weights + base_off : nullptr);
                lens.push_back(len);
            }
            base_off += len;
        }
        read_idx += window_n_reads[w];
        if (reads.empty() || lens[0] <= 0)
        {
            res->windows[static_cast<size_t>(w)].status = 9; // no defined reference behaviour; flagged generic_error
            continue;
        }
        if (s32)
            run_window<int32_t>(P, reads, wts, lens, res->windows[static_cast<size_t>(w)], *res, rcp);
        else
            run_window<int16_t>(P, reads, wts, lens, res->windows[static_cast<size_t>(w)], *res, rcp);
    }
    return res;
}

void oracle_free(void* h)
{
    delete static_cast<RunResult*>(h);
}
int32_t oracle_score_bits(void* h)
{
    return static_cast<RunResult*>(h)->score_bits;
}
int32_t oracle_size_bits(void* h)
{
    return static_cast<RunResult*>(h)->size_bits;
}
int32_t oracle_status(void* h, int32_t w)
{
    return static_cast<RunResult*>(h)->windows[static_cast<size_t>(w)].status;
}
int32_t oracle_consensus(void* h, int32_t w, const char** seq, const uint16_t** coverage)
{
    auto& r = static_cast<RunResult*>(h)->windows[static_cast<size_t>(w)];
    if (seq)
        *seq = r.consensus.c_str();
    if (coverage)
        *coverage = r.coverage.data();
    return static_cast<int32_t>(r.consensus.size());
}
int32_t oracle_msa_rows(void* h, int32_t w)
{
    return static_cast<int32_t>(static_cast<RunResult*>(h)->windows[static_cast<size_t>(w)].msa.size());
}
int32_t oracle_msa_row(void* h, int32_t w, int32_t row, const char** seq)
{
    auto& r = static_cast<RunResult*>(h)->windows[static_cast<size_t>(w)];
    *seq    = r.msa[static_cast<size_t>(row)].c_str();
    return static_cast<int32_t>(r.msa[static_cast<size_t>(row)].size());
}
int32_t oracle_graph_nodes(void* h, int32_t w, char* labels, int32_t cap)
{
    auto& r   = static_cast<RunResult*>(h)->windows[static_cast<size_t>(w)];
    int32_t n = static_cast<int32_t>(r.node_labels.size());
    for (int32_t i = 0; i < n && i < cap; i++)
        labels[i] = r.node_labels[static_cast<size_t>(i)];
    return n;
}
int32_t oracle_graph_edges(void* h, int32_t w, int32_t* src, int32_t* dst, int32_t* weight, int32_t cap)
{
    auto& r   = static_cast<RunResult*>(h)->windows[static_cast<size_t>(w)];
    int32_t n = static_cast<int32_t>(r.edges.size());
    for (int32_t i = 0; i < n && i < cap; i++)
    {
        dst[i]    = std::get<0>(r.edges[static_cast<size_t>(i)]);
        src[i]    = std::get<1>(r.edges[static_cast<size_t>(i)]);
        weight[i] = std::get<2>(r.edges[static_cast<size_t>(i)]);
    }
    return n;
}
int64_t oracle_dp_cells(void* h)
{
    return static_cast<RunResult*>(h)->dp_cells;
}
int64_t oracle_graph_rows(void* h)
{
    return static_cast<RunResult*>(h)->graph_rows;
}
int64_t oracle_widened(void* h)
{
    return static_cast<RunResult*>(h)->widened;
}

int64_t oracle_rcp_misses(void* h)
{
    return static_cast<RunResult*>(h)->rcp_misses;
}

// ---- device-function level entry points (reference known-answer tests) ----------------------------------

int32_t oracle_nw_full(const uint8_t* nodes, const int16_t* sorted, const int16_t* node_id_to_pos, int32_t node_count,
                       const uint16_t* in_cnt, const int16_t* in_edges, const uint16_t* out_cnt,
                       const uint8_t* read, int32_t read_len, int32_t scores_width, int32_t gap, int32_t mismatch,
                       int32_t match, int16_t* aln_graph, int16_t* aln_read)
{
    std::vector<int16_t> scores(static_cast<size_t>(node_count + 1) * static_cast<size_t>(scores_width) + 16, 0);
    return nw_full<int16_t, int16_t>(nodes, sorted, node_id_to_pos, node_count, in_cnt, in_edges, out_cnt, read,
                                     read_len, scores.data(), scores_width, aln_graph, aln_read, gap, mismatch,
                                     match);
}

void oracle_topsort_kahn(int16_t* sorted, int16_t* node_id_to_pos, int32_t node_count, const uint16_t* in_cnt,
                         const int16_t* out_edges, const uint16_t* out_cnt)
{
    std::vector<uint16_t> local(static_cast<size_t>(node_count) + 1, 0);
    topsort_kahn<int16_t>(sorted, node_id_to_pos, node_count, in_cnt, out_edges, out_cnt, local.data());
}

int32_t oracle_add_alignment(uint8_t* nodes, int32_t* node_count, int16_t* aligned, uint16_t* aligned_cnt,
                             int16_t* in_edges, uint16_t* in_cnt, int16_t* out_edges, uint16_t* out_cnt,
                             uint16_t* in_w, int32_t aln_len, const int16_t* aln_graph, const uint8_t* read,
                             const int16_t* aln_read, uint16_t* coverage, const int8_t* base_weights,
                             int32_t max_nodes)
{
    int32_t new_count = *node_count;
    int16_t begin     = 0;
    int32_t rc = add_alignment<int16_t>(new_count, nodes, *node_count, aligned, aligned_cnt, in_edges, in_cnt,
                                        out_edges, out_cnt, in_w, aln_len, aln_graph, read, aln_read, coverage,
                                        base_weights, false, &begin, nullptr, nullptr, 0, 1, max_nodes);
    *node_count = new_count;
    return rc;
}

int32_t oracle_consensus_from_graph(const uint8_t* nodes, int32_t node_count, const int16_t* sorted,
                                    const int16_t* node_id_to_pos, const int16_t* in_edges, const uint16_t* in_cnt,
                                    const int16_t* out_edges, const uint16_t* out_cnt, const uint16_t* in_w,
                                    const uint16_t* coverage_in, const int16_t* aligned, const uint16_t* aligned_cnt,
                                    int32_t max_consensus_size, uint8_t* consensus, uint16_t* coverage_out)
{
    return generate_consensus<int16_t>(nodes, node_count, sorted, node_id_to_pos, in_edges, in_cnt, out_edges,
                                       out_cnt, in_w, consensus, coverage_out, coverage_in, aligned, aligned_cnt,
                                       max_consensus_size);
}

} // extern "C"
