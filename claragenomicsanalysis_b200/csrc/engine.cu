// Host side of the B200 POA engine: the cudapoa::Batch implementation behind create_batch().
//
// Observable behaviour follows the reference host code (cudapoa/src/cudapoa_batch.cuh, allocate_block.hpp,
// batch.cu, cudapoa_limits.hpp, utils.cu, cudapoa.cpp); memory layout, staging and kernels are this engine's own.
#include "nvtx_ranges.hpp"
#include "engine_stats.hpp"
#include "poa_layout.hpp"
#include "poa_window.cuh"
#include "poa_phases.cuh"

#include <claraparabricks/genomeworks/cudapoa/batch.hpp>
#include <claraparabricks/genomeworks/cudapoa/cudapoa.hpp>
#include <claraparabricks/genomeworks/cudapoa/utils.hpp>

#include <algorithm>
#include <cstdio>
#include <atomic>
#include <cstring>
#include <iomanip>
#include <iostream>
#include <map>
#include <mutex>
#include <tuple>
#include <string>
#include <vector>

namespace claraparabricks
{
namespace genomeworks
{
namespace cudapoa
{

using namespace poa_b200;

namespace
{

// cudapoa_limits.hpp:28-37.  Always uses the NON-banded graph dimension, as the reference does.
bool use_32bit_score(const BatchSize& b, int32_t gap, int32_t mismatch, int32_t match)
{
    const int32_t upper = b.max_sequence_size * match;
    const int32_t lower = b.max_sequence_size * std::max(gap, mismatch) +
                          (b.max_matrix_graph_dimension - b.max_sequence_size) * gap;
    return (upper > INT16_MAX || (-lower) > (INT16_MAX + 1));
}

// cudapoa_limits.hpp:39-53
bool use_32bit_size(const BatchSize& b, bool banded)
{
    int32_t m = b.max_consensus_size;
    m         = std::max(m, banded ? b.max_matrix_graph_dimension_banded : b.max_matrix_graph_dimension);
    m         = std::max(m, b.max_matrix_sequence_dimension);
    return m > INT16_MAX;
}

struct MemoryPlan
{
    bool score32 = false;
    bool size32  = false;
    int32_t M    = 0; // node capacity
    int32_t G    = 0; // score matrix rows
    int32_t band_cells = 0; // widest band a score row must hold (adaptive mode: kAdaptiveBandFactor x the band)
    ArenaLayout layout{};
    int64_t input_per_poa  = 0;
    int64_t output_per_poa = 0;
    int64_t fixed_per_poa  = 0; // arena + inputs + outputs
    int64_t score_per_poa  = 0; // banded: exact; full: for the widest allowed read
};

MemoryPlan make_plan(const BatchSize& b, bool banded, bool msa, int32_t gap, int32_t mismatch, int32_t match,
                     bool adaptive = false)
{
    MemoryPlan p;
    p.band_cells = b.alignment_band_width * ((banded && adaptive) ? kAdaptiveBandFactor : 1);
    p.score32 = use_32bit_score(b, gap, mismatch, match);
    p.size32  = p.score32 && use_32bit_size(b, banded);
    p.M       = banded ? b.max_nodes_per_window_banded : b.max_nodes_per_window;
    p.G       = banded ? b.max_matrix_graph_dimension_banded : b.max_matrix_graph_dimension;
    p.G       = std::max(p.G, p.M + 1); // one score row per node plus row 0
    const int32_t S = std::max(b.max_sequences_per_poa, 1);
    const int32_t L = std::max(b.max_sequence_size, 1);
    p.layout  = make_arena_layout(p.M, p.M + L + 8, S, L, p.size32 ? 4 : 2, msa);
    p.input_per_poa  = static_cast<int64_t>(S) * L * 2 + static_cast<int64_t>(S) * 4 + sizeof(WindowDesc);
    p.output_per_poa = 1 + 4 + 8 +
                       (msa ? static_cast<int64_t>(S) * b.max_consensus_size
                            : static_cast<int64_t>(b.max_consensus_size) * 3);
    p.fixed_per_poa = p.layout.stride + align_up64(p.input_per_poa, 16) + align_up64(p.output_per_poa, 16) + 64;
    const int64_t width = banded ? p.band_cells + kBandRowPad : b.max_matrix_sequence_dimension + 8;
    p.score_per_poa     = width * static_cast<int64_t>(p.G) * (p.score32 ? 4 : 2);
    return p;
}

std::atomic<int32_t> g_batches_16_16{0}, g_batches_32_16{0}, g_batches_32_32{0};

template <typename T>
T* carve(uint8_t*& cursor, int64_t count)
{
    T* p   = reinterpret_cast<T*>(cursor);
    cursor += align_up64(count * static_cast<int64_t>(sizeof(T)), 256);
    return p;
}

} // namespace

struct PhaseTimeline;

// 0: phase-split schedule wherever it applies (unless CUDAPOA_B200_SCHEDULE=fused), 1: always the single launch,
// 2: phase-split regardless of the environment.  Process-wide; set through cudapoa_b200_debug_set_schedule.
static std::atomic<int32_t> g_schedule_override{0};

class B200Batch : public Batch, public EngineStatsSource
{
public:
    B200Batch(int32_t device_id, cudaStream_t stream, size_t max_mem, int8_t output_mask, const BatchSize& batch_size,
              int32_t gap, int32_t mismatch, int32_t match, bool banded, bool adaptive = false)
        : max_sequences_per_poa_(throw_on_negative(batch_size.max_sequences_per_poa,
                                                   "Maximum sequences per POA has to be non-negative"))
        , device_id_(throw_on_negative(device_id, "Device ID has to be non-negative"))
        , stream_(stream)
        , output_mask_(output_mask)
        , batch_size_(batch_size)
        , gap_(gap)
        , mismatch_(mismatch)
        , match_(match)
        , banded_(banded)
        , adaptive_(banded && adaptive)
        , msa_((output_mask & OutputType::msa) != 0)
    {
        plan_ = make_plan(batch_size_, banded_, msa_, gap_, mismatch_, match_, adaptive_);
        // allocate_block.hpp:66-76: refuse budgets that cannot hold a single POA
        const size_t minimum = static_cast<size_t>(plan_.fixed_per_poa);
        if (max_mem < minimum + static_cast<size_t>(plan_.score_per_poa))
        {
            throw std::runtime_error(std::string("Require at least ")
                                         .append(std::to_string(minimum + plan_.score_per_poa))
                                         .append(" bytes of device memory per CUDAPOA batch to process correctly."));
        }
        scoped_device_switch dev(device_id_);
        max_poas_ = static_cast<int32_t>(
            std::min<int64_t>(max_mem / static_cast<size_t>(plan_.fixed_per_poa + plan_.score_per_poa), INT32_MAX / 4));
        bid_ = (plan_.score32 ? (plan_.size32 ? g_batches_32_32 : g_batches_32_16) : g_batches_16_16)++;
        allocate(max_mem);
        GW_CU_CHECK_ERR(cudaEventCreate(&ev_start_));
        GW_CU_CHECK_ERR(cudaEventCreate(&ev_stop_));
        reset();
    }

    ~B200Batch() override
    {
        scoped_device_switch dev(device_id_);
        cudaEventDestroy(ev_start_);
        cudaEventDestroy(ev_stop_);
        cudaFree(block_d_);
        cudaFreeHost(block_h_);
        if (weights_h_ != nullptr)
            cudaFreeHost(weights_h_);
    }

    StatusType add_poa_group(std::vector<StatusType>& per_seq_status, const Group& poa_group) override;
    int32_t get_total_poas() const override { return poa_count_; }
    void generate_poa() override;
    StatusType get_consensus(std::vector<std::string>& consensus, std::vector<std::vector<uint16_t>>& coverage,
                             std::vector<StatusType>& output_status) override;
    StatusType get_msa(std::vector<std::vector<std::string>>& msa, std::vector<StatusType>& output_status) override;
    void get_graphs(std::vector<DirectedGraph>& graphs, std::vector<StatusType>& output_status) override;
    int32_t batch_id() const override { return bid_; }

    void reset() override
    {
        poa_count_       = 0;
        num_bases_       = 0;
        num_reads_       = 0;
        next_score_off_  = 0;
        avail_score_mem_ = score_pool_bytes_;
        has_weights_     = false;
        launched_        = false;
        timed_           = false;
    }

    void fill_stats(cudapoa_b200_stats* out) const override;
    int32_t max_poas() const override { return max_poas_; }
    double relaunch(int32_t iterations) override;
    int32_t relaunch_phases(double* fill_ms, double* graph_ms) override;
    double relaunch_timed(int32_t iterations, double* fill_ms, double* graph_ms, int32_t* launches) override;
    int32_t last_phase_ms(double* out, int32_t cap) const override
    {
        const int32_t n = std::min<int32_t>(cap, kPhaseKinds);
        for (int32_t i = 0; i < n; i++)
            out[i] = phase_ms_[i];
        return n;
    }
    void set_overlapped(bool overlapped) override { overlapped_ = overlapped; }
    int64_t debug_read_window(int32_t window, int32_t region, void* out, int64_t cap) override
    {
        scoped_device_switch dev(device_id_);
        if (window < 0 || window >= poa_count_)
            return -1;
        GW_CU_CHECK_ERR(cudaStreamSynchronize(stream_));
        const int64_t score_bytes = static_cast<int64_t>(plan_.band_cells + kBandRowPad) * plan_.G * (plan_.score32 ? 4 : 2);
        const int64_t size        = (region == 0) ? plan_.layout.stride : score_bytes;
        const uint8_t* src        = (region == 0) ? arena_d_ + static_cast<int64_t>(window) * plan_.layout.stride
                                                  : scores_d_ + static_cast<int64_t>(window) * score_bytes;
        if (out != nullptr && cap > 0)
            GW_CU_CHECK_ERR(cudaMemcpy(out, src, static_cast<size_t>(std::min(size, cap)), cudaMemcpyDeviceToHost));
        return size;
    }

private:
    void allocate(size_t max_mem);
    bool reserve_scores(int32_t max_seq_length);
    StatusType add_sequence(const Entry& e);
    void ensure_weights();
    void launch();
    bool launch_split(KernelParams& P);
    void sync_and_collect();
    void push_status(std::vector<StatusType>& out, int32_t poa) const
    {
        out.emplace_back(static_cast<StatusType>(status_h_[poa]));
    }

    int32_t max_sequences_per_poa_;
    int32_t device_id_;
    cudaStream_t stream_;
    int8_t output_mask_;
    BatchSize batch_size_;
    int32_t gap_, mismatch_, match_;
    bool banded_;
    bool adaptive_; // static band first, widened on demand (BandMode::adaptive_band)
    bool msa_;
    MemoryPlan plan_;
    int32_t max_poas_ = 0;
    int32_t bid_      = 0;

    // device block and its carve-up
    uint8_t* block_d_      = nullptr;
    uint8_t* bases_d_      = nullptr;
    int8_t* weights_d_     = nullptr;
    int32_t* read_len_d_   = nullptr;
    WindowDesc* windows_d_ = nullptr;
    uint8_t* status_d_     = nullptr;
    int32_t* node_count_d_ = nullptr;
    uint32_t* stat_rows_d_ = nullptr;
    uint32_t* stat_exact_d_ = nullptr;
    unsigned long long* stat_cells_d_ = nullptr;
    int32_t* next_window_d_           = nullptr; // windows are handed to warps through this counter
    int32_t* phase_counters_d_        = nullptr; // phase-split schedule: one such counter per launch
    PhaseState* phase_d_              = nullptr; // phase-split schedule: per-window state between launches
    int32_t sm_count_                 = 0;
    uint8_t* consensus_d_  = nullptr;
    uint16_t* coverage_d_  = nullptr;
    uint8_t* msa_d_        = nullptr;
    uint8_t* arena_d_      = nullptr;
    uint8_t* scores_d_     = nullptr;
    size_t score_pool_bytes_ = 0;

    // pinned host block
    uint8_t* block_h_       = nullptr;
    uint8_t* bases_h_       = nullptr;
    int8_t* weights_h_      = nullptr; // allocated on the first weighted entry
    int32_t* read_len_h_    = nullptr;
    WindowDesc* windows_h_  = nullptr;
    uint8_t* status_h_      = nullptr;
    int32_t* node_count_h_  = nullptr;
    uint32_t* stat_rows_h_  = nullptr;
    uint32_t* stat_exact_h_ = nullptr;
    unsigned long long* stat_cells_h_ = nullptr;
    uint8_t* consensus_h_   = nullptr;
    uint16_t* coverage_h_   = nullptr;
    uint8_t* msa_h_         = nullptr;

    // staging state
    int32_t poa_count_      = 0;
    int64_t num_bases_      = 0;
    int32_t num_reads_      = 0;
    int64_t next_score_off_ = 0; // full alignment: score elements handed out so far
    size_t avail_score_mem_ = 0;
    bool has_weights_       = false;
    bool launched_          = false;
    int64_t kernel_launches_ = 0; // kernels enqueued by the last generate_poa / relaunch
    PhaseTimeline* timeline_ = nullptr; // set while relaunch_phases() is timing the launches
    static constexpr int kPhaseKinds = 5;
    double phase_ms_[kPhaseKinds]    = {0, 0, 0, 0, 0}; // last relaunch_timed(): fill, traceback, fusion, sort, consensus
    bool overlapped_         = false;   // other batches are in flight on this device (set_overlapped)
    bool timed_             = false;
    int32_t timed_iters_    = 1;

    cudaEvent_t ev_start_{}, ev_stop_{};
    mutable cudapoa_b200_stats stats_{};
};

void B200Batch::allocate(size_t max_mem)
{
    const int64_t S = std::max(max_sequences_per_poa_, 1);
    const int64_t L = std::max(batch_size_.max_sequence_size, 1);
    const int64_t C = std::max(batch_size_.max_consensus_size, 2);
    GW_CU_CHECK_ERR(cudaMalloc(reinterpret_cast<void**>(&block_d_), max_mem));

    // Shrink max_poas_ until the fixed arrays plus one banded score matrix per POA fit (alignment padding).
    auto fixed_bytes = [&](int64_t n) {
        int64_t b = 0;
        auto a    = [&](int64_t x) { b += align_up64(x, 256); };
        a(n * S * L);                 // bases
        a(n * S * L);                 // weights
        a(n * S * 4);                 // read_len
        a(n * static_cast<int64_t>(sizeof(WindowDesc)));
        a(n);                         // status
        a(n * 4);                     // node_count
        a(n * 4);                     // stat_rows
        a(n * 4);                     // stat_exact
        a((n + 1) * 8);               // stat_cells + the launch's window counter
        a(kPhaseCounters * 4);                              // phase-split schedule: window counters
        a(n * static_cast<int64_t>(sizeof(PhaseState)));    // ... and per-window state
        a(msa_ ? 0 : n * C);          // consensus
        a(msa_ ? 0 : n * C * 2);      // coverage
        a(msa_ ? n * S * C : 0);      // msa
        a(n * plan_.layout.stride);   // arena
        return b;
    };
    while (max_poas_ > 1 &&
           fixed_bytes(max_poas_) + static_cast<int64_t>(max_poas_) * plan_.score_per_poa > static_cast<int64_t>(max_mem))
        max_poas_--;

    const int64_t n = max_poas_;
    uint8_t* cur    = block_d_;
    bases_d_        = carve<uint8_t>(cur, n * S * L);
    weights_d_      = carve<int8_t>(cur, n * S * L);
    read_len_d_     = carve<int32_t>(cur, n * S);
    windows_d_      = carve<WindowDesc>(cur, n);
    status_d_       = carve<uint8_t>(cur, n);
    node_count_d_   = carve<int32_t>(cur, n);
    stat_rows_d_    = carve<uint32_t>(cur, n);
    stat_exact_d_   = carve<uint32_t>(cur, n);
    GW_CU_CHECK_ERR(cudaDeviceGetAttribute(&sm_count_, cudaDevAttrMultiProcessorCount, device_id_));
    stat_cells_d_   = carve<unsigned long long>(cur, n + 1);
    next_window_d_  = reinterpret_cast<int32_t*>(stat_cells_d_ + n);
    phase_counters_d_ = carve<int32_t>(cur, kPhaseCounters);
    phase_d_          = reinterpret_cast<PhaseState*>(carve<uint8_t>(cur, n * static_cast<int64_t>(sizeof(PhaseState))));
    consensus_d_    = carve<uint8_t>(cur, msa_ ? 0 : n * C);
    coverage_d_     = carve<uint16_t>(cur, msa_ ? 0 : n * C);
    msa_d_          = carve<uint8_t>(cur, msa_ ? n * S * C : 0);
    arena_d_        = carve<uint8_t>(cur, n * plan_.layout.stride);
    scores_d_       = cur;
    const int64_t used = cur - block_d_;
    score_pool_bytes_  = (static_cast<int64_t>(max_mem) > used) ? max_mem - static_cast<size_t>(used) : 0;

    // pinned staging: inputs (weights are added lazily) and outputs
    int64_t hb = 0;
    auto h     = [&](int64_t x) {
        int64_t at = hb;
        hb += align_up64(x, 256);
        return at;
    };
    const int64_t o_bases = h(n * S * L), o_len = h(n * S * 4), o_win = h(n * sizeof(WindowDesc)), o_st = h(n),
                  o_nc = h(n * 4), o_sr = h(n * 4), o_se = h(n * 4), o_sc = h(n * 8), o_cons = h(msa_ ? 0 : n * C),
                  o_cov = h(msa_ ? 0 : n * C * 2), o_msa = h(msa_ ? n * S * C : 0);
    GW_CU_CHECK_ERR(cudaHostAlloc(reinterpret_cast<void**>(&block_h_), static_cast<size_t>(hb + 256), cudaHostAllocDefault));
    bases_h_      = block_h_ + o_bases;
    read_len_h_   = reinterpret_cast<int32_t*>(block_h_ + o_len);
    windows_h_    = reinterpret_cast<WindowDesc*>(block_h_ + o_win);
    status_h_     = block_h_ + o_st;
    node_count_h_ = reinterpret_cast<int32_t*>(block_h_ + o_nc);
    stat_rows_h_  = reinterpret_cast<uint32_t*>(block_h_ + o_sr);
    stat_exact_h_ = reinterpret_cast<uint32_t*>(block_h_ + o_se);
    stat_cells_h_ = reinterpret_cast<unsigned long long*>(block_h_ + o_sc);
    consensus_h_  = block_h_ + o_cons;
    coverage_h_   = reinterpret_cast<uint16_t*>(block_h_ + o_cov);
    msa_h_        = block_h_ + o_msa;
}

// reserve_buf, cudapoa_batch.cuh:542-564
bool B200Batch::reserve_scores(int32_t max_seq_length)
{
    const int64_t width = banded_ ? plan_.band_cells + kBandRowPad
                                  : cudautils::align<int32_t, 4>(max_seq_length + 1 + 4);
    const size_t bytes  = static_cast<size_t>(width) * static_cast<size_t>(plan_.G) * (plan_.score32 ? 4 : 2);
    if (bytes > avail_score_mem_)
    {
        if (get_total_poas() == 0)
        {
            std::cout << "Memory available " << std::fixed << std::setprecision(2)
                      << static_cast<double>(avail_score_mem_) / 1024. / 1024. / 1024.;
            std::cout << "GB, Memory required " << static_cast<double>(bytes) / 1024. / 1024. / 1024.;
            std::cout << "GB (sequence length " << max_seq_length << ", graph length " << plan_.G << ")" << std::endl;
        }
        return false;
    }
    avail_score_mem_ -= bytes;
    return true;
}

void B200Batch::ensure_weights()
{
    if (has_weights_)
        return;
    if (weights_h_ == nullptr)
    {
        const size_t bytes = static_cast<size_t>(max_poas_) * std::max(max_sequences_per_poa_, 1) *
                             std::max(batch_size_.max_sequence_size, 1);
        GW_CU_CHECK_ERR(cudaHostAlloc(reinterpret_cast<void**>(&weights_h_), bytes, cudaHostAllocDefault));
    }
    std::memset(weights_h_, 1, static_cast<size_t>(num_bases_)); // everything staged so far had weight 1
    has_weights_ = true;
}

// add_seq_to_poa, cudapoa_batch.cuh:490-539
StatusType B200Batch::add_sequence(const Entry& e)
{
    if (e.length > batch_size_.max_sequence_size)
        return StatusType::exceeded_maximum_sequence_size;
    WindowDesc& wd      = windows_h_[poa_count_ - 1];
    const int32_t width = cudautils::align<int32_t, 4>(e.length + 1 + 4);
    if (width > wd.scores_width)
    {
        if (!banded_)
            next_score_off_ += static_cast<int64_t>(width - wd.scores_width) * plan_.G;
        wd.scores_width = width;
    }
    if (wd.n_reads >= max_sequences_per_poa_)
        return StatusType::exceeded_maximum_sequences_per_poa;
    if (e.weights != nullptr)
    {
        for (int32_t i = 0; i < e.length; i++)
            throw_on_negative(e.weights[i], "Base weights need to be non-negative");
        ensure_weights();
    }
    if (e.length > 0)
    {
        std::memcpy(bases_h_ + num_bases_, e.seq, static_cast<size_t>(e.length));
        if (has_weights_)
        {
            if (e.weights != nullptr)
                std::memcpy(weights_h_ + num_bases_, e.weights, static_cast<size_t>(e.length));
            else
                std::memset(weights_h_ + num_bases_, 1, static_cast<size_t>(e.length));
        }
    }
    read_len_h_[num_reads_] = e.length;
    wd.n_reads++;
    num_bases_ += e.length;
    num_reads_++;
    return StatusType::success;
}

// add_poa_group, cudapoa_batch.cuh:103-143
StatusType B200Batch::add_poa_group(std::vector<StatusType>& per_seq_status, const Group& poa_group)
{
    int32_t max_len = 0;
    for (const Entry& e : poa_group)
        max_len = std::max(max_len, e.length);
    if (!reserve_scores(max_len))
        return StatusType::exceeded_maximum_poas;
    per_seq_status.clear();
    if (poa_count_ == max_poas_)
        return StatusType::exceeded_maximum_poas;

    WindowDesc wd{};
    wd.base_off     = num_bases_;
    wd.score_off    = next_score_off_;
    wd.read_off     = num_reads_;
    wd.n_reads      = 0;
    wd.scores_width = 0;
    windows_h_[poa_count_++] = wd;
    for (const Entry& e : poa_group)
        per_seq_status.push_back(add_sequence(e));
    return StatusType::success;
}

static int32_t phase_blocks_per_sm(const void* kernel, int32_t threads, size_t smem);

template <typename ScoreT, typename IdT, bool BANDED, bool MSA>
static void launch_one(const KernelParams& P, int32_t sm_count, cudaStream_t stream)
{
    const dim3 block(kWarpsPerBlock * 32);
    const size_t smem = static_cast<size_t>(kWarpsPerBlock) * static_cast<size_t>(P.smem_per_warp);
    auto kernel       = poa_window_kernel<ScoreT, IdT, BANDED, MSA>;
    // persistent warps: at most one resident wave of CTAs, windows are pulled from P.next_window.  The function
    // attribute and the occupancy query are paid once per kernel and device (phase_blocks_per_sm).
    const int32_t needed = (P.n_windows + kWarpsPerBlock - 1) / kWarpsPerBlock;
    const int32_t per_sm = phase_blocks_per_sm(reinterpret_cast<const void*>(kernel), static_cast<int>(block.x), smem);
    const dim3 grid(std::max(1, std::min(needed, sm_count * per_sm)));
    POA_LAUNCH(kernel, grid, block, smem, stream, P);
}

template <typename ScoreT, typename IdT>
static void launch_typed(const KernelParams& P, bool banded, bool msa, int32_t sm_count, cudaStream_t stream)
{
    if (banded)
    {
        if (msa)
            launch_one<ScoreT, IdT, true, true>(P, sm_count, stream);
        else
            launch_one<ScoreT, IdT, true, false>(P, sm_count, stream);
    }
    else
    {
        if (msa)
            launch_one<ScoreT, IdT, false, true>(P, sm_count, stream);
        else
            launch_one<ScoreT, IdT, false, false>(P, sm_count, stream);
    }
}

// One launch of a phase kernel: persistent warps, at most one resident wave of CTAs.
// The dynamic shared-memory limit is an attribute of the function, shared by every batch and host thread of the
// process: it is only ever set to one value, the most any launch uses.
constexpr int kPhaseSmemLimit = 200 * 1024;

// Resident CTAs per SM of a phase kernel at a given scratch size, looked up once per (kernel, device, size): a batch
// makes well over a hundred launches and several host threads launch batches, so the two driver calls per launch this
// replaces were a measurable share of a staging thread's time.
static int32_t phase_blocks_per_sm(const void* kernel, int32_t threads, size_t smem)
{
    static std::mutex mu;
    static std::map<std::tuple<const void*, int, size_t>, int32_t> cache;
    int device = 0;
    GW_CU_CHECK_ERR(cudaGetDevice(&device));
    const auto key = std::make_tuple(kernel, device, smem);
    std::lock_guard<std::mutex> lock(mu);
    auto it = cache.find(key);
    if (it != cache.end())
        return it->second;
    GW_CU_CHECK_ERR(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kPhaseSmemLimit));
    int per_sm = 1;
    GW_CU_CHECK_ERR(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, smem));
    cache.emplace(key, std::max(per_sm, 1));
    return std::max(per_sm, 1);
}

// Development switch: CUDAPOA_B200_BLOCKS_<FILL|TRACE|FUSE|SORT|FINAL>=n caps the resident blocks per SM of a phase
// (the walks of the graph phases run faster below their maximum occupancy; the defaults below were measured).
static int32_t blocks_cap_from_env(const char* name, int32_t fallback)
{
    const char* v = std::getenv(name);
    const int32_t n = v ? std::atoi(v) : 0;
    return n > 0 ? n : fallback;
}

template <typename Kernel>
static void launch_phase(Kernel kernel, const KernelParams& P, int32_t warps_per_block, int32_t sm_count,
                         cudaStream_t stream, int32_t blocks_cap = INT32_MAX, size_t block_smem = 0)
{
    const dim3 block(warps_per_block * 32);
    const size_t smem    = static_cast<size_t>(warps_per_block) * static_cast<size_t>(P.smem_per_warp) + block_smem;
    const int32_t per_sm = std::min(blocks_cap, phase_blocks_per_sm(reinterpret_cast<const void*>(kernel), static_cast<int>(block.x), smem));
    const int32_t needed = (P.n_windows + warps_per_block - 1) / warps_per_block;
    const dim3 grid(std::max(1, std::min(needed, sm_count * per_sm)));
    POA_LAUNCH(kernel, grid, block, smem, stream, P);
#ifndef POA_EMU
    // development: CUDAPOA_B200_DEBUG_SYNC=1 waits for every phase launch and names the one that failed
    static const bool debug_sync = [] {
        const char* env = std::getenv("CUDAPOA_B200_DEBUG_SYNC");
        return env != nullptr && env[0] == '1';
    }();
    if (debug_sync)
    {
        const cudaError_t e = cudaStreamSynchronize(stream);
        if (e != cudaSuccess)
        {
            cudaFuncAttributes attr{};
            std::fprintf(stderr, "phase launch failed: %s; step %d, grid %u, block %u, shared memory %zu, scratch per warp %d, class %d\n",
                         cudaGetErrorString(e), P.step, grid.x, block.x, smem, P.smem_per_warp, P.phase_class);
            (void)attr;
        }
    }
#endif
}

// Nodes a graph may have for the consensus (staging area + 6 bytes per node) to run out of `smem` bytes of scratch.
static int32_t consensus_scratch_nodes(int32_t smem)
{
    return ((smem - kConsStageBytes) / 6) & ~7;
}

// Optional CUDA events between the launches of a batch (relaunch_timed): kinds[i] tells what ran between events[i - 1]
// and events[i] (0 fill launches, 1 traceback, 2 fusion, 3 topological sort, 4 consensus / MSA, -1 anything else: counter
// reset, the gap between two steps).
struct PhaseTimeline
{
    std::vector<cudaEvent_t> events;
    std::vector<int8_t> kinds;
    size_t used = 0;
    void mark(cudaStream_t stream, int kind)
    {
        if (used == events.size())
        {
            cudaEvent_t e;
            GW_CU_CHECK_ERR(cudaEventCreate(&e));
            events.push_back(e);
        }
        GW_CU_CHECK_ERR(cudaEventRecord(events[used++], stream));
        kinds.push_back(static_cast<int8_t>(kind));
    }
    ~PhaseTimeline()
    {
        for (cudaEvent_t e : events)
            cudaEventDestroy(e);
    }
};

// Scratch sizes of the launches of one step (bytes per warp == per window).
struct PhaseScratch
{
    int32_t fill, fill_long, trace, sort_small, sort_large, final_;
};
constexpr int32_t kCountersPerStep = 9;

template <typename ScoreT, typename IdT, bool MSA>
static int32_t launch_split_typed(KernelParams& P, int32_t steps, const PhaseScratch& scratch,
                                  int32_t* counters, int32_t sm_count,
                                  cudaStream_t stream, PhaseTimeline* timeline, bool full)
{
    int32_t launches = 0;
    if (timeline)
        timeline->mark(stream, -1);
    static const int32_t cap_fill  = blocks_cap_from_env("CUDAPOA_B200_BLOCKS_FILL", INT32_MAX);
    static const int32_t cap_trace = blocks_cap_from_env("CUDAPOA_B200_BLOCKS_TRACE", 6);
    static const int32_t cap_fuse  = blocks_cap_from_env("CUDAPOA_B200_BLOCKS_FUSE", INT32_MAX);
    static const int32_t cap_sort  = blocks_cap_from_env("CUDAPOA_B200_BLOCKS_SORT", INT32_MAX);
    const bool two = scratch.sort_large > scratch.sort_small;
    P.class_cap    = two ? kahn_smem_nodes(scratch.sort_small) : INT32_MAX;
    P.lean_class0  = two ? 1 : 0;
    {
        const int32_t blocks = std::max(1, std::min((P.n_windows + kFillWarpsPerBlock - 1) / kFillWarpsPerBlock, sm_count * 16));
        auto init_kernel = poa_init_kernel<IdT, MSA>;
        POA_LAUNCH(init_kernel, dim3(blocks), dim3(kFillWarpsPerBlock * 32), 0, stream, P);
        launches++;
    }
    for (int32_t s = 1; s <= steps; s++)
    {
        int32_t* c      = counters + kCountersPerStep * (s - 1);
        P.step          = s;
        P.next_window   = c + 0;
        P.smem_per_warp = scratch.fill;
        if (full)
        {
            // full alignment: the chunked packed fill, the exact full alignment for what it declines
            launch_phase(poa_fill_full_kernel<ScoreT, IdT>, P, kFillWarpsPerBlock, sm_count, stream, cap_fill, kLutBytes);
            P.next_window   = c + 2;
            P.smem_per_warp = 0;
            launch_phase(poa_exact_kernel<ScoreT, IdT, true>, P, kFillWarpsPerBlock, sm_count, stream);
            launches += 2;
        }
        else
        {
            launch_phase(poa_fill_kernel<ScoreT, IdT>, P, kFillWarpsPerBlock, sm_count, stream, cap_fill, kLutBytes);
            P.next_window   = c + 1;
            P.smem_per_warp = scratch.fill_long;
            launch_phase(poa_fill_rebased_kernel<ScoreT, IdT>, P, kFillWarpsPerBlock, sm_count, stream);
            P.next_window   = c + 2;
            P.smem_per_warp = 0;
            launch_phase(poa_exact_kernel<ScoreT, IdT>, P, kFillWarpsPerBlock, sm_count, stream);
            launches += 3;
        }
        if (timeline)
            timeline->mark(stream, 0);
        P.next_window   = c + 3;
        P.smem_per_warp = scratch.trace;
        // traceback: one warp per window over shared-memory tiles (trace2_impl), or -- large batches -- one thread per
        // window (trace2_lane): its launch takes as long as one walk out of L2 whatever the batch size (0.61 ms per
        // read at cfg 2), the warp-per-window launch 0.049 ms per thousand windows, which is more from about 12 500
        // windows (18 944 windows: 17.7 against 26.7 ms per step; CUDAPOA_B200_TRACE=warp / lane forces either)
        static const int trace_choice = [] {
            const char* env = std::getenv("CUDAPOA_B200_TRACE");
            return env == nullptr ? 0 : std::strcmp(env, "warp") == 0 ? 1 : std::strcmp(env, "lane") == 0 ? 2 : 0;
        }();
        constexpr int32_t kTraceLanesMinWindows = 12288;
        const int32_t trace_forced = (g_schedule_override.load() >> 2) & 3; // tests: 1 warp per window, 2 thread per window
        const int32_t trace_mode   = trace_forced != 0 ? trace_forced : trace_choice;
        if (trace_mode == 2 || (trace_mode == 0 && P.n_windows >= kTraceLanesMinWindows))
        {
            const dim3 block(kGraphWarpsPerBlock * 32);
            const dim3 grid((P.n_windows + block.x - 1) / block.x);
            if (full)
                POA_LAUNCH((poa_trace_lanes_kernel<ScoreT, IdT, true>), grid, block, 0, stream, P);
            else
                POA_LAUNCH((poa_trace_lanes_kernel<ScoreT, IdT, false>), grid, block, 0, stream, P);
        }
        else if (full)
            launch_phase(poa_trace_kernel<ScoreT, IdT, true>, P, kGraphWarpsPerBlock, sm_count, stream, cap_trace);
        else
            launch_phase(poa_trace_kernel<ScoreT, IdT>, P, kGraphWarpsPerBlock, sm_count, stream, cap_trace);
        if (P.adaptive_max_band > 0)
        {
            P.next_window   = c + 8;
            P.smem_per_warp = 0;
            launch_phase(poa_widen_kernel<ScoreT, IdT>, P, kFillWarpsPerBlock, sm_count, stream);
            launches++;
        }
        if (timeline)
            timeline->mark(stream, 1);
        P.next_window   = c + 4;
        P.smem_per_warp = 0;
        launch_phase(poa_fuse_kernel<IdT, MSA>, P, kGraphWarpsPerBlock, sm_count, stream, cap_fuse);
        if (timeline)
            timeline->mark(stream, 2);
        P.next_window   = c + 5;
        P.smem_per_warp = scratch.sort_small;
        P.phase_class   = 0;
        if (two)
            launch_phase(poa_sort_kernel<IdT, true>, P, kGraphWarpsPerBlock, sm_count, stream, cap_sort);
        else
            launch_phase(poa_sort_kernel<IdT, false>, P, kGraphWarpsPerBlock, sm_count, stream, cap_sort);
        launches += 3;
        if (two)
        {
            P.next_window   = c + 6;
            P.smem_per_warp = scratch.sort_large;
            P.phase_class   = 1;
            launch_phase(poa_sort_kernel<IdT, false>, P, kGraphWarpsPerBlock, sm_count, stream);
            launches++;
        }
        if (timeline)
            timeline->mark(stream, 3);
        if (s == steps) // windows that ended earlier keep their graph untouched until here: one launch finalises all
        {
            P.next_window   = c + 7;
            P.smem_per_warp = scratch.final_;
            launch_phase(poa_final_kernel<IdT, MSA>, P, kGraphWarpsPerBlock, sm_count, stream);
            launches++;
            if (timeline)
                timeline->mark(stream, 4);
        }
    }
    return launches;
}

// Phase-split schedule (poa_phases.cuh) for banded batches: one launch per phase and read instead of one per batch.
// Returns false when the batch has to take the single launch (full alignment, 32-bit node ids, very deep windows,
// or CUDAPOA_B200_SCHEDULE=fused).
bool B200Batch::launch_split(KernelParams& P)
{
    static const bool fused_by_env = [] {
        const char* env = std::getenv("CUDAPOA_B200_SCHEDULE");
        return env != nullptr && std::strcmp(env, "fused") == 0;
    }();
    const int32_t mode = g_schedule_override.load() & 3;
    if (plan_.size32 || mode == 1 || (mode == 0 && fused_by_env))
        return false;
    // full alignment: the chunked packed fill keeps a read's whole profile in shared memory (reads of up to 768 bases:
    // three column chunks); longer full alignments stay with the single launch
    if (!banded_ && batch_size_.max_sequence_size > 768)
        return false;
    int32_t max_reads = 1;
    for (int32_t i = 0; i < poa_count_; i++)
        max_reads = std::max(max_reads, windows_h_[i].n_reads);
    const int32_t steps = std::max(max_reads - 1, 1);
    // A lone batch that keeps fewer than about 15 windows per SM busy per step is bound by the latency of a step's kernel
    // chain, and then the single launch (no barrier between the steps of different windows) is as fast or faster.
    // Work per step in units of a 500-base read: ragged batches (windows with few reads drop out early) and long reads
    // are what the plain window count misses.  Measured on a B200 (profiles/r2_fill_cap_costream_and_crossover.txt,
    // profiles/r2_other_shapes_both_schedules.txt): cfg 2 2 368 windows 35.9 against 37.3 ms, 3 552 windows 45 against
    // 68 ms (the single launch holds 16 windows per SM: a second wave); cfg 5b 2 368 windows (1 970 units) 444 against
    // 402 ms, 9 472 windows 583 against 766 ms; cfg 4 592 windows (4 190 units) 2 045 against 2 131 ms.
    // With the round's last fill kernel the crossover moved down for polishing-size windows, whose graphs sort out of
    // the small shared-memory class (lone cfg 2 batches, profiles/r2s_pipeline_trace_and_crossover.txt: 592 windows 25.1
    // against 26.5 ms, 1 184 27.0 against 29.8, 2 368 32.1 against 37.1); batches with long windows keep the old bound
    // (cfg 5b, 2 368 windows, 1 970 units: 423 against 391 ms -- their graphs sort out of global memory).
    const int64_t units     = num_bases_ / (static_cast<int64_t>(500) * steps);
    const int64_t min_units = static_cast<int64_t>(batch_size_.max_sequence_size <= 1024 ? 4 : 15) * sm_count_;
    if (mode == 0 && !overlapped_ && units < min_units)
        return false;
    if (kCountersPerStep * steps > kPhaseCounters)
        return false;
    // shared memory per window and launch: the fill needs the row ring, the record buffer and the profile; the
    // traceback its tile and the rows it records; the sort 4 bytes per node (small class: graphs of up to 1 264 nodes -- a 500-base polishing window ends at 1 130 +- 25,
    // large class: the single launch's scratch); the consensus its staging area and 6 bytes per node
    const int32_t L = std::max(batch_size_.max_sequence_size, 1);
    PhaseScratch scratch{};
    // (the plain packed fill keeps the whole profile of a read in shared memory: reads of up to 768 bases, which leaves
    // six blocks per SM; longer reads take the rebased fill, which slides a 512-column window of the profile)
    scratch.fill          = (fill2_smem_bytes(std::min(L, 768)) + 255) & ~255;
    scratch.fill_long     = (fill2_smem_bytes_sliding() + 255) & ~255;
    scratch.trace      = (trace2_smem_bytes(std::min(L, 1024)) + 255) & ~255; // longer reads record their rows in global memory
    scratch.sort_large = kSmemPerWarp;
    // (1 264 nodes: 5.5 KB per window, the most that leaves ten blocks per SM next to the 1 KB the system reserves per
    // block -- the sort is bound by windows in flight: 36 ms per cfg 2 step at eight blocks, 41 at six, 53 at four)
    static const int32_t sort_nodes = blocks_cap_from_env("CUDAPOA_B200_SORT_NODES", 1264);
    scratch.sort_small = std::min(kSmemPerWarp, (kKahnBytesPerNode * std::min(plan_.M, sort_nodes) + 2 * kKahnQueueEntries + 15) & ~15);
    if (kahn_smem_nodes(scratch.sort_small) >= plan_.M)
        scratch.sort_large = scratch.sort_small; // one class serves every graph of this batch
    // (measured dead end, profiles/r2_wide_sort_class_and_pipeline_connections.txt: a third class with 5-byte records
    // and one warp per block for graphs of 3 456 .. 10 240 nodes -- cfg 4's sort got slower, 841 against 682 ms at 592
    // windows: four such windows per SM lose to sixteen walking global memory)
    scratch.final_ = kSmemPerWarp;
    if (4 * std::max(std::max(std::max(scratch.fill, scratch.fill_long), scratch.trace), scratch.sort_large) > kPhaseSmemLimit)
        return false;
    P.phase = phase_d_;
    int32_t launches = 0;
    GW_CU_CHECK_ERR(cudaMemsetAsync(phase_counters_d_, 0, sizeof(int32_t) * kPhaseCounters, stream_));
    if (!plan_.score32)
    {
        if (msa_)
            launches = launch_split_typed<int16_t, int16_t, true>(P, steps, scratch, phase_counters_d_, sm_count_, stream_, timeline_, !banded_);
        else
            launches = launch_split_typed<int16_t, int16_t, false>(P, steps, scratch, phase_counters_d_, sm_count_, stream_, timeline_, !banded_);
    }
    else
    {
        if (msa_)
            launches = launch_split_typed<int32_t, int16_t, true>(P, steps, scratch, phase_counters_d_, sm_count_, stream_, timeline_, !banded_);
        else
            launches = launch_split_typed<int32_t, int16_t, false>(P, steps, scratch, phase_counters_d_, sm_count_, stream_, timeline_, !banded_);
    }
    kernel_launches_ += launches;
    return true;
}

void B200Batch::launch()
{
    KernelParams P{};
    P.bases               = bases_d_;
    P.weights             = has_weights_ ? weights_d_ : nullptr;
    P.read_len            = read_len_d_;
    P.windows             = windows_d_;
    P.n_windows           = poa_count_;
    P.arena               = arena_d_;
    P.layout              = plan_.layout;
    P.scores              = scores_d_;
    P.banded_score_stride = static_cast<int64_t>(plan_.band_cells + kBandRowPad) * plan_.G;
    P.adaptive_max_band   = adaptive_ ? plan_.band_cells : 0;
    P.score_rows          = plan_.G;
    P.status              = status_d_;
    P.node_count          = node_count_d_;
    P.consensus           = consensus_d_;
    P.coverage            = coverage_d_;
    P.msa                 = msa_d_;
    P.stat_rows           = stat_rows_d_;
    P.stat_exact          = stat_exact_d_;
    P.stat_cells          = stat_cells_d_;
    P.next_window         = next_window_d_;
    P.max_nodes           = plan_.M;
    P.max_consensus       = batch_size_.max_consensus_size;
    P.max_reads           = std::max(max_sequences_per_poa_, 1);
    P.band                = batch_size_.alignment_band_width;
    P.gap                 = gap_;
    P.gap_x2              = (static_cast<uint32_t>(P.gap) & 0xffffu) * 0x10001u;
    P.mismatch            = mismatch_;
    P.match               = match_;
    P.smem_per_warp       = kSmemPerWarp;
#ifdef SPOA_ACCURATE
    P.spoa_accurate = 1;
#else
    const char* env = std::getenv("CUDAPOA_B200_SPOA_ACCURATE");
    P.spoa_accurate = (env != nullptr && env[0] == '1') ? 1 : 0;
#endif
    {
        const char* dbg = std::getenv("CUDAPOA_B200_DEBUG_FLAGS");
        P.debug_flags   = dbg ? std::atoi(dbg) : 0;
    }
    if (launch_split(P))
    {
        GW_CU_CHECK_ERR(cudaPeekAtLastError());
        return;
    }
    GW_CU_CHECK_ERR(cudaMemsetAsync(next_window_d_, 0, sizeof(int32_t), stream_));
    kernel_launches_ += 1;
    if (!plan_.score32)
        launch_typed<int16_t, int16_t>(P, banded_, msa_, sm_count_, stream_);
    else if (!plan_.size32)
        launch_typed<int32_t, int16_t>(P, banded_, msa_, sm_count_, stream_);
    else
        launch_typed<int32_t, int32_t>(P, banded_, msa_, sm_count_, stream_);
    GW_CU_CHECK_ERR(cudaPeekAtLastError());
}

// generate_poa, cudapoa_batch.cuh:152-193: copies and launch are enqueued, nothing is waited for.
void B200Batch::generate_poa()
{
    NvtxRange range("cudapoa_b200::generate_poa (h2d + launches)");
    scoped_device_switch dev(device_id_);
    if (poa_count_ == 0)
        return;
    stats_           = cudapoa_b200_stats{};
    stats_.windows   = poa_count_;
    stats_.reads_aligned = num_reads_ - poa_count_;
    GW_CU_CHECK_ERR(cudaMemcpyAsync(bases_d_, bases_h_, static_cast<size_t>(num_bases_), cudaMemcpyHostToDevice, stream_));
    stats_.h2d_bytes += num_bases_;
    if (has_weights_)
    {
        GW_CU_CHECK_ERR(cudaMemcpyAsync(weights_d_, weights_h_, static_cast<size_t>(num_bases_), cudaMemcpyHostToDevice, stream_));
        stats_.h2d_bytes += num_bases_;
    }
    GW_CU_CHECK_ERR(cudaMemcpyAsync(read_len_d_, read_len_h_, sizeof(int32_t) * num_reads_, cudaMemcpyHostToDevice, stream_));
    GW_CU_CHECK_ERR(cudaMemcpyAsync(windows_d_, windows_h_, sizeof(WindowDesc) * poa_count_, cudaMemcpyHostToDevice, stream_));
    stats_.h2d_bytes += static_cast<int64_t>(sizeof(int32_t)) * num_reads_ + static_cast<int64_t>(sizeof(WindowDesc)) * poa_count_;
    GW_CU_CHECK_ERR(cudaEventRecord(ev_start_, stream_));
    kernel_launches_ = 0;
    launch();
    GW_CU_CHECK_ERR(cudaEventRecord(ev_stop_, stream_));
    launched_    = true;
    timed_       = false;
    timed_iters_ = 1;
}

void B200Batch::sync_and_collect()
{
    NvtxRange range("cudapoa_b200::sync_and_collect (d2h + wait)");
    const size_t n = static_cast<size_t>(poa_count_);
    GW_CU_CHECK_ERR(cudaMemcpyAsync(status_h_, status_d_, n, cudaMemcpyDeviceToHost, stream_));
    GW_CU_CHECK_ERR(cudaMemcpyAsync(node_count_h_, node_count_d_, n * 4, cudaMemcpyDeviceToHost, stream_));
    GW_CU_CHECK_ERR(cudaMemcpyAsync(stat_rows_h_, stat_rows_d_, n * 4, cudaMemcpyDeviceToHost, stream_));
    GW_CU_CHECK_ERR(cudaMemcpyAsync(stat_exact_h_, stat_exact_d_, n * 4, cudaMemcpyDeviceToHost, stream_));
    GW_CU_CHECK_ERR(cudaMemcpyAsync(stat_cells_h_, stat_cells_d_, n * 8, cudaMemcpyDeviceToHost, stream_));
    stats_.d2h_bytes += static_cast<int64_t>(n) * 21;
    GW_CU_CHECK_ERR(cudaStreamSynchronize(stream_));
    if (launched_ && !timed_)
    {
        float ms = 0.f;
        GW_CU_CHECK_ERR(cudaEventElapsedTime(&ms, ev_start_, ev_stop_));
        stats_.kernel_ms       = ms / static_cast<float>(timed_iters_);
        stats_.graph_kernel_ms = stats_.kernel_ms;
        int64_t rows = 0, exact = 0, cells = 0, widened = 0;
        for (size_t i = 0; i < n; i++)
        {
            rows += stat_rows_h_[i];
            exact += stat_exact_h_[i] & 0xffffu;
            widened += stat_exact_h_[i] >> 16;
            cells += static_cast<int64_t>(stat_cells_h_[i]);
        }
        stats_.graph_rows = rows;
        stats_.exact_path = exact;
        stats_.widened    = widened;
        stats_.kernel_launches = kernel_launches_;
        stats_.dp_cells   = cells; // sum of N_s * Wc
        timed_          = true;
    }
}

double B200Batch::relaunch(int32_t iterations)
{
    scoped_device_switch dev(device_id_);
    if (!launched_ || poa_count_ == 0 || iterations <= 0)
        return -1.0;
    GW_CU_CHECK_ERR(cudaEventRecord(ev_start_, stream_));
    kernel_launches_ = 0;
    for (int32_t i = 0; i < iterations; i++)
        launch();
    GW_CU_CHECK_ERR(cudaEventRecord(ev_stop_, stream_));
    GW_CU_CHECK_ERR(cudaStreamSynchronize(stream_));
    float ms = 0.f;
    GW_CU_CHECK_ERR(cudaEventElapsedTime(&ms, ev_start_, ev_stop_));
    timed_       = false; // the next get_* refreshes kernel_ms from these events
    timed_iters_ = iterations;
    return static_cast<double>(ms);
}

double B200Batch::relaunch_timed(int32_t iterations, double* fill_ms, double* graph_ms, int32_t* launches)
{
    scoped_device_switch dev(device_id_);
    if (!launched_ || poa_count_ == 0 || iterations <= 0)
        return -1.0;
    PhaseTimeline timeline;
    timeline_        = &timeline;
    kernel_launches_ = 0;
    GW_CU_CHECK_ERR(cudaEventRecord(ev_start_, stream_));
    for (int32_t i = 0; i < iterations; i++)
        launch();
    GW_CU_CHECK_ERR(cudaEventRecord(ev_stop_, stream_));
    timeline_ = nullptr;
    GW_CU_CHECK_ERR(cudaStreamSynchronize(stream_));
    float total = 0.f;
    GW_CU_CHECK_ERR(cudaEventElapsedTime(&total, ev_start_, ev_stop_));
    double fill = 0.0, graph = 0.0;
    std::fill(phase_ms_, phase_ms_ + kPhaseKinds, 0.0);
    if (timeline.used < 2) // single launch: everything is one kernel
        fill = phase_ms_[0] = total;
    for (size_t i = 1; i < timeline.used; i++)
    {
        if (timeline.kinds[i] < 0)
            continue;
        float ms = 0.f;
        GW_CU_CHECK_ERR(cudaEventElapsedTime(&ms, timeline.events[i - 1], timeline.events[i]));
        (timeline.kinds[i] == 0 ? fill : graph) += ms;
        phase_ms_[std::min<int>(timeline.kinds[i], kPhaseKinds - 1)] += ms;
    }
    if (fill_ms)
        *fill_ms = fill;
    if (graph_ms)
        *graph_ms = graph;
    if (launches)
        *launches = static_cast<int32_t>(kernel_launches_);
    timed_       = false; // the next get_* refreshes kernel_ms from these events
    timed_iters_ = iterations;
    return static_cast<double>(total);
}

int32_t B200Batch::relaunch_phases(double* fill_ms, double* graph_ms)
{
    int32_t launches = 0;
    return relaunch_timed(1, fill_ms, graph_ms, &launches) < 0 ? -1 : launches;
}

void B200Batch::fill_stats(cudapoa_b200_stats* out) const
{
    *out = stats_;
}

// get_consensus, cudapoa_batch.cuh:196-252: blocking; outputs are appended to the caller's vectors.
StatusType B200Batch::get_consensus(std::vector<std::string>& consensus, std::vector<std::vector<uint16_t>>& coverage,
                                    std::vector<StatusType>& output_status)
{
    if (!(OutputType::consensus & output_mask_))
        return StatusType::output_type_unavailable;
    NvtxRange range("cudapoa_b200::get_consensus");
    scoped_device_switch dev(device_id_);
    const int64_t C = batch_size_.max_consensus_size;
    if (poa_count_ > 0 && launched_)
    {
        const size_t n = static_cast<size_t>(poa_count_) * static_cast<size_t>(C);
        GW_CU_CHECK_ERR(cudaMemcpyAsync(consensus_h_, consensus_d_, n, cudaMemcpyDeviceToHost, stream_));
        GW_CU_CHECK_ERR(cudaMemcpyAsync(coverage_h_, coverage_d_, n * 2, cudaMemcpyDeviceToHost, stream_));
        stats_.d2h_bytes = static_cast<int64_t>(n) * 3;
        sync_and_collect();
    }
    consensus.reserve(consensus.size() + poa_count_);
    coverage.reserve(coverage.size() + poa_count_);
    output_status.reserve(output_status.size() + poa_count_);
    for (int32_t poa = 0; poa < poa_count_; poa++)
    {
        push_status(output_status, poa);
        if (status_h_[poa] != 0)
        {
            consensus.emplace_back();
            coverage.emplace_back();
            continue;
        }
        // the device wrote the path end -> start, NUL terminated
        const char* c      = reinterpret_cast<const char*>(consensus_h_ + poa * C);
        const uint16_t* cv = coverage_h_ + poa * C;
        const size_t len   = strnlen(c, static_cast<size_t>(C));
        consensus.emplace_back(std::string(std::reverse_iterator<const char*>(c + len),
                                           std::reverse_iterator<const char*>(c)));
        coverage.emplace_back(std::vector<uint16_t>(std::reverse_iterator<const uint16_t*>(cv + len),
                                                    std::reverse_iterator<const uint16_t*>(cv)));
    }
    return StatusType::success;
}

// get_msa, cudapoa_batch.cuh:255-307
StatusType B200Batch::get_msa(std::vector<std::vector<std::string>>& msa, std::vector<StatusType>& output_status)
{
    if (!(OutputType::msa & output_mask_))
        return StatusType::output_type_unavailable;
    NvtxRange range("cudapoa_b200::get_msa");
    scoped_device_switch dev(device_id_);
    const int64_t C = batch_size_.max_consensus_size;
    const int64_t S = std::max(max_sequences_per_poa_, 1);
    if (poa_count_ > 0 && launched_)
    {
        const size_t n = static_cast<size_t>(poa_count_) * static_cast<size_t>(S * C);
        GW_CU_CHECK_ERR(cudaMemcpyAsync(msa_h_, msa_d_, n, cudaMemcpyDeviceToHost, stream_));
        stats_.d2h_bytes = static_cast<int64_t>(n);
        sync_and_collect();
    }
    for (int32_t poa = 0; poa < poa_count_; poa++)
    {
        msa.emplace_back();
        push_status(output_status, poa);
        if (status_h_[poa] != 0)
            continue;
        const int32_t n_reads = windows_h_[poa].n_reads;
        msa.back().reserve(n_reads);
        for (int32_t r = 0; r < n_reads; r++)
            msa.back().emplace_back(reinterpret_cast<const char*>(msa_h_ + (poa * S + r) * C));
    }
    return StatusType::success;
}

// get_graphs, cudapoa_batch.cuh:309-387: node label = base, one edge source -> sink per in-edge slot.
void B200Batch::get_graphs(std::vector<DirectedGraph>& graphs, std::vector<StatusType>& output_status)
{
    NvtxRange range("cudapoa_b200::get_graphs");
    scoped_device_switch dev(device_id_);
    graphs.resize(poa_count_);
    if (poa_count_ == 0)
        return;
    if (launched_)
        sync_and_collect();
    const ArenaLayout& L = plan_.layout;
    const size_t idb     = plan_.size32 ? 4 : 2;
    const size_t M       = static_cast<size_t>(L.M);
    // Groups of windows, four strided copies per group (cudaMemcpy2DAsync, pitch = a window's arena) and two waits:
    // bases + in-degrees first, then the in-edge and weight planes up to the group's largest in-degree (slot-major
    // planes: the first k planes of a window are contiguous).  The host staging stays below ~256 MB.
    const size_t per_window_guess = M * (3 + 4 * (idb + 2));
    const int32_t group = static_cast<int32_t>(std::max<size_t>(1, std::min<size_t>(poa_count_, (size_t(256) << 20) / per_window_guess)));
    std::vector<uint8_t> nodes, in_cnt, in_edge, in_w;
    for (int32_t first = 0; first < poa_count_; first += group)
    {
        const int32_t count  = std::min(group, poa_count_ - first);
        const uint8_t* arena = arena_d_ + static_cast<int64_t>(first) * L.stride;
        nodes.resize(static_cast<size_t>(count) * M);
        in_cnt.resize(static_cast<size_t>(count) * M * 2);
        GW_CU_CHECK_ERR(cudaMemcpy2DAsync(nodes.data(), M, arena + L.nodes, static_cast<size_t>(L.stride), M, count,
                                          cudaMemcpyDeviceToHost, stream_));
        GW_CU_CHECK_ERR(cudaMemcpy2DAsync(in_cnt.data(), M * 2, arena + L.in_cnt, static_cast<size_t>(L.stride), M * 2, count,
                                          cudaMemcpyDeviceToHost, stream_));
        GW_CU_CHECK_ERR(cudaStreamSynchronize(stream_));
        size_t max_slots = 0;
        for (int32_t k = 0; k < count; k++)
        {
            if (status_h_[first + k] != 0)
                continue;
            const uint16_t* cnt = reinterpret_cast<const uint16_t*>(in_cnt.data()) + static_cast<size_t>(k) * M;
            for (int32_t i = 0; i < node_count_h_[first + k]; i++)
                max_slots = std::max<size_t>(max_slots, cnt[i]);
        }
        in_edge.resize(static_cast<size_t>(count) * max_slots * M * idb);
        in_w.resize(static_cast<size_t>(count) * max_slots * M * 2);
        if (max_slots > 0)
        {
            GW_CU_CHECK_ERR(cudaMemcpy2DAsync(in_edge.data(), max_slots * M * idb, arena + L.in_edges,
                                              static_cast<size_t>(L.stride), max_slots * M * idb, count,
                                              cudaMemcpyDeviceToHost, stream_));
            GW_CU_CHECK_ERR(cudaMemcpy2DAsync(in_w.data(), max_slots * M * 2, arena + L.in_w, static_cast<size_t>(L.stride),
                                              max_slots * M * 2, count, cudaMemcpyDeviceToHost, stream_));
            GW_CU_CHECK_ERR(cudaStreamSynchronize(stream_));
        }
        for (int32_t k = 0; k < count; k++)
        {
            const int32_t poa = first + k;
            push_status(output_status, poa);
            if (status_h_[poa] != 0)
                continue;
            const int32_t n      = node_count_h_[poa];
            const uint8_t* base  = nodes.data() + static_cast<size_t>(k) * M;
            const uint16_t* cnt  = reinterpret_cast<const uint16_t*>(in_cnt.data()) + static_cast<size_t>(k) * M;
            const uint8_t* edges = in_edge.data() + static_cast<size_t>(k) * max_slots * M * idb;
            const uint16_t* wts  = reinterpret_cast<const uint16_t*>(in_w.data()) + static_cast<size_t>(k) * max_slots * M;
            DirectedGraph& graph = graphs[poa];
            for (int32_t sink = 0; sink < n; sink++)
            {
                graph.set_node_label(sink, std::string(1, static_cast<char>(base[sink])));
                for (int32_t e = 0; e < cnt[sink]; e++)
                {
                    const size_t at   = static_cast<size_t>(e) * M + static_cast<size_t>(sink);
                    const int32_t src = (idb == 2) ? reinterpret_cast<const int16_t*>(edges)[at]
                                                   : reinterpret_cast<const int32_t*>(edges)[at];
                    graph.add_edge(src, sink, wts[at]);
                }
            }
        }
    }
}

// batch.cu:27-77
std::unique_ptr<Batch> create_batch(int32_t device_id, cudaStream_t stream, size_t max_mem, int8_t output_mask,
                                    const BatchSize& batch_size, int16_t gap_score, int16_t mismatch_score,
                                    int16_t match_score, bool cuda_banded_alignment)
{
    return std::unique_ptr<Batch>(new B200Batch(device_id, stream, max_mem, output_mask, batch_size, gap_score,
                                                mismatch_score, match_score, cuda_banded_alignment));
}

// Extension (BandMode as in later GenomeWorks releases): adaptive_band aligns with the static band first and repeats
// the alignment with a doubled band while the path found runs along a band edge (poa_window.cuh).
std::unique_ptr<Batch> create_batch(int32_t device_id, cudaStream_t stream, size_t max_mem, int8_t output_mask,
                                    const BatchSize& batch_size, int16_t gap_score, int16_t mismatch_score,
                                    int16_t match_score, BandMode band_mode)
{
    return std::unique_ptr<Batch>(new B200Batch(device_id, stream, max_mem, output_mask, batch_size, gap_score,
                                                mismatch_score, match_score, band_mode != BandMode::full_band,
                                                band_mode == BandMode::adaptive_band));
}

// cudapoa.cpp:23-29.  The engine has no logger to start; the call stays for source compatibility.
StatusType Init()
{
    return StatusType::success;
}

// Groups -> batches (cudapoa/include/.../utils.hpp:48-58; the reference's binning is cudapoa/src/utils.cu:24-138).
// Same contract: groups are classed by how many windows of their shape one batch holds under the memory quota, in
// power-of-two classes (or the caller's `bins_capacity`); a class of large windows also takes the following classes of
// smaller windows while they are no more numerous than its capacity.  Differences, deliberate:
//   * capacity comes from this engine's footprint (make_plan), and cudaMemGetInfo is asked once, not once per group;
//   * a batch's BatchSize covers the longest read AND the largest read count among ALL the groups it received (the
//     reference takes the group with the largest length x reads product of the first class only, so a merged group can
//     be rejected by its own batch: allocate_block.hpp:364-401, utils.cu:86-91,117).
void get_multi_batch_sizes(std::vector<BatchSize>& list_of_batch_sizes,
                           std::vector<std::vector<int32_t>>& list_of_groups_per_batch,
                           const std::vector<Group>& poa_groups, bool banded_alignment, bool msa_flag,
                           int32_t band_width, std::vector<int32_t>* bins_capacity, float gpu_memory_usage_quota,
                           int32_t mismatch_score, int32_t gap_score, int32_t match_score)
{
    size_t total = 0, free_mem = 0;
    GW_CU_CHECK_ERR(cudaMemGetInfo(&free_mem, &total));
    const double budget = static_cast<double>(gpu_memory_usage_quota) * static_cast<double>(free_mem);

    std::vector<int32_t> pow2;
    if (bins_capacity == nullptr)
    {
        for (int32_t j = 0; j < 20; j++)
            pow2.push_back(1 << j);
        bins_capacity = &pow2;
    }
    const std::vector<int32_t>& cap = *bins_capacity;
    const int32_t n_classes         = get_size<int32_t>(cap);
    if (n_classes == 0)
        return;

    struct Class
    {
        std::vector<int32_t> groups;
        int32_t longest = 0, most_reads = 0;
    };
    std::vector<Class> classes(n_classes);
    // windows of one shape share one footprint: shapes repeat massively in real inputs
    std::map<std::pair<int32_t, int32_t>, int32_t> class_of_shape;
    const int32_t n_groups = get_size<int32_t>(poa_groups);
    for (int32_t i = 0; i < n_groups; i++)
    {
        int32_t longest = 0;
        for (const Entry& e : poa_groups[i])
            longest = std::max(longest, e.length);
        const int32_t reads = get_size<int32_t>(poa_groups[i]);
        auto it             = class_of_shape.find({longest, reads});
        if (it == class_of_shape.end())
        {
            const MemoryPlan plan = make_plan(BatchSize(longest, reads, band_width), banded_alignment, msa_flag,
                                              gap_score, mismatch_score, match_score);
            const int64_t fit = static_cast<int64_t>(budget / static_cast<double>(plan.fixed_per_poa + plan.score_per_poa));
            // first class whose capacity covers `fit` (capacities ascend); the last class takes the rest
            int32_t c = static_cast<int32_t>(std::lower_bound(cap.begin(), cap.end(), fit,
                                                              [](int32_t capacity, int64_t v) { return capacity < v; }) -
                                             cap.begin());
            c  = std::min(c, n_classes - 1);
            it = class_of_shape.emplace(std::make_pair(longest, reads), c).first;
        }
        Class& k = classes[it->second];
        k.groups.push_back(i);
        k.longest    = std::max(k.longest, longest);
        k.most_reads = std::max(k.most_reads, reads);
    }
    for (int32_t j = 0; j < n_classes; j++)
    {
        Class& head = classes[j];
        if (head.groups.empty())
            continue;
        for (int32_t k = j + 1; k < n_classes; k++)
        {
            Class& next = classes[k];
            if (next.groups.empty())
                continue;
            if (cap[j] < get_size<int32_t>(next.groups))
                break;
            head.groups.insert(head.groups.end(), next.groups.begin(), next.groups.end());
            head.longest    = std::max(head.longest, next.longest);
            head.most_reads = std::max(head.most_reads, next.most_reads);
            next.groups.clear();
        }
        list_of_batch_sizes.emplace_back(head.longest, head.most_reads, band_width);
        list_of_groups_per_batch.push_back(std::move(head.groups));
        head.groups.clear();
    }
}

} // namespace cudapoa
} // namespace genomeworks
} // namespace claraparabricks

#ifndef POA_EMU // device micro-benchmarks and tables: nothing to emulate
namespace
{
__global__ void rcp_table_kernel(float* out, int32_t n)
{
    int32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n)
    {
        float r;
        float d = static_cast<float>(i);
        asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(d));
        out[i] = r;
    }
}
} // namespace

extern "C" int32_t cudapoa_b200_debug_rcp_table(int32_t device_id, float* out, int32_t n)
{
    cudaError_t e = cudaSetDevice(device_id);
    if (e != cudaSuccess)
        return static_cast<int32_t>(e);
    float* d = nullptr;
    e        = cudaMalloc(reinterpret_cast<void**>(&d), sizeof(float) * static_cast<size_t>(n));
    if (e != cudaSuccess)
        return static_cast<int32_t>(e);
    rcp_table_kernel<<<(n + 255) / 256, 256>>>(d, n);
    e = cudaMemcpy(out, d, sizeof(float) * static_cast<size_t>(n), cudaMemcpyDeviceToHost);
    cudaFree(d);
    return static_cast<int32_t>(e);
}

namespace
{
template <int KIND>
__global__ void __launch_bounds__(256) dpx_peak_kernel(uint32_t* out, int32_t iters, uint32_t seed)
{
    uint32_t a[8];
#pragma unroll
    for (int k = 0; k < 8; k++)
        a[k] = seed + threadIdx.x * 8u + k;
    const uint32_t b = seed | 0x00010001u;
    for (int32_t i = 0; i < iters; i++)
    {
#pragma unroll
        for (int k = 0; k < 8; k++)
        {
            if (KIND == 0)
                a[k] = __viaddmax_s16x2(a[k], b, a[(k + 1) & 7]);
            else
                a[k] = static_cast<uint32_t>(__viaddmax_s32(static_cast<int32_t>(a[k]), static_cast<int32_t>(b),
                                                            static_cast<int32_t>(a[(k + 1) & 7])));
        }
    }
    uint32_t x = 0;
#pragma unroll
    for (int k = 0; k < 8; k++)
        x ^= a[k];
    if (x == 0x12345678u)
        out[blockIdx.x] = x;
}
} // namespace

extern "C" int32_t cudapoa_b200_debug_dpx_peak(int32_t device_id, int32_t kind, double* instr_per_s, double* ms_out)
{
    cudaError_t e = cudaSetDevice(device_id);
    if (e != cudaSuccess)
        return static_cast<int32_t>(e);
    cudaDeviceProp prop{};
    cudaGetDeviceProperties(&prop, device_id);
    const int32_t blocks = prop.multiProcessorCount * 8; // 8 x 256 threads = 2048 threads per SM
    const int32_t iters  = 1 << 16;
    uint32_t* d          = nullptr;
    cudaMalloc(reinterpret_cast<void**>(&d), sizeof(uint32_t) * blocks);
    cudaEvent_t t0, t1;
    cudaEventCreate(&t0);
    cudaEventCreate(&t1);
    float best = 1e30f;
    for (int rep = 0; rep < 4; rep++)
    {
        cudaEventRecord(t0);
        if (kind == 0)
            dpx_peak_kernel<0><<<blocks, 256>>>(d, iters, 12345u + rep);
        else
            dpx_peak_kernel<1><<<blocks, 256>>>(d, iters, 12345u + rep);
        cudaEventRecord(t1);
        e = cudaEventSynchronize(t1);
        float ms = 0.f;
        cudaEventElapsedTime(&ms, t0, t1);
        if (rep > 0 && ms < best)
            best = ms;
    }
    cudaEventDestroy(t0);
    cudaEventDestroy(t1);
    cudaFree(d);
    const double instr = static_cast<double>(blocks) * 256.0 * 8.0 * static_cast<double>(iters);
    if (instr_per_s != nullptr)
        *instr_per_s = instr / (static_cast<double>(best) * 1e-3);
    if (ms_out != nullptr)
        *ms_out = best;
    return static_cast<int32_t>(e);
}

#endif // POA_EMU

extern "C" void cudapoa_b200_debug_set_schedule(int32_t mode)
{
    claraparabricks::genomeworks::cudapoa::g_schedule_override.store(mode);
}
