// C ABI over the cudapoa C++ plugin interface (create_batch / Batch).  See include/cudapoa_b200.h.
//
// This translation unit only uses the public C++ API, so the very same file is compiled twice:
//   * against this repository's headers and engine       -> libcudapoa_b200.so  (prefix cudapoa_b200_)
//   * against the reference's headers and library, on the
//     machine that has /root/reference (oracle/Makefile)  -> oracle/_ref/libcudapoa_ref.so (prefix cudapoa_ref_)
// which gives the parity tests two engines behind one calling convention.

#include "cudapoa_b200.h"

#include <claraparabricks/genomeworks/cudapoa/batch.hpp>
#include <claraparabricks/genomeworks/cudapoa/cudapoa.hpp>
#include <claraparabricks/genomeworks/cudapoa/utils.hpp>

#include <algorithm>
#include <cstring>
#include <exception>
#include <memory>
#include <stdexcept>
#include <string>
#include <tuple>
#include <vector>

#ifdef CUDAPOA_B200_NATIVE
#include "engine_stats.hpp"
#endif

namespace gw  = claraparabricks::genomeworks;
namespace poa = claraparabricks::genomeworks::cudapoa;

struct cudapoa_b200_batch
{
    std::unique_ptr<poa::Batch> impl;
    // outputs of the last get_*
    std::vector<std::string> consensus;
    std::vector<std::vector<uint16_t>> coverage;
    std::vector<std::vector<std::string>> msa;
    std::vector<gw::DirectedGraph> graphs;
    std::vector<poa::StatusType> status;
    int32_t result_kind = 0; // 0 none, 1 consensus, 2 msa, 3 graphs
    // scratch for add_poa_group
    poa::Group group;
    std::vector<poa::StatusType> seq_status;
};

namespace
{

thread_local std::string g_last_error;

template <typename F>
int32_t guarded(F&& f)
{
    try
    {
        return f();
    }
    catch (const std::invalid_argument& e)
    {
        g_last_error = e.what();
        return -1;
    }
    catch (const std::runtime_error& e)
    {
        g_last_error = e.what();
        return -2;
    }
    catch (const std::exception& e)
    {
        g_last_error = e.what();
        return -3;
    }
    catch (...)
    {
        g_last_error = "unknown exception";
        return -3;
    }
}

void to_c(const poa::BatchSize& in, cudapoa_b200_batch_size* out)
{
    out->max_sequence_size                 = in.max_sequence_size;
    out->max_consensus_size                = in.max_consensus_size;
    out->max_nodes_per_window              = in.max_nodes_per_window;
    out->max_nodes_per_window_banded       = in.max_nodes_per_window_banded;
    out->max_matrix_graph_dimension        = in.max_matrix_graph_dimension;
    out->max_matrix_graph_dimension_banded = in.max_matrix_graph_dimension_banded;
    out->max_matrix_sequence_dimension     = in.max_matrix_sequence_dimension;
    out->alignment_band_width              = in.alignment_band_width;
    out->max_sequences_per_poa             = in.max_sequences_per_poa;
}

poa::BatchSize from_c(const cudapoa_b200_batch_size& in)
{
    // Start from a valid object, then take the caller's fields verbatim: callers of the C++ API may
    // also edit the public fields after construction (cudapoa/tests/Test_CudapoaGenerateMSA2.cu:131).
    poa::BatchSize bs(in.max_sequence_size, in.max_sequences_per_poa, in.alignment_band_width);
    bs.max_sequence_size                 = in.max_sequence_size;
    bs.max_consensus_size                = in.max_consensus_size;
    bs.max_nodes_per_window              = in.max_nodes_per_window;
    bs.max_nodes_per_window_banded       = in.max_nodes_per_window_banded;
    bs.max_matrix_graph_dimension        = in.max_matrix_graph_dimension;
    bs.max_matrix_graph_dimension_banded = in.max_matrix_graph_dimension_banded;
    bs.max_matrix_sequence_dimension     = in.max_matrix_sequence_dimension;
    bs.alignment_band_width              = in.alignment_band_width;
    bs.max_sequences_per_poa             = in.max_sequences_per_poa;
    return bs;
}

void clear_results(cudapoa_b200_batch* b)
{
    b->consensus.clear();
    b->coverage.clear();
    b->msa.clear();
    b->graphs.clear();
    b->status.clear();
    b->result_kind = 0;
}

inline void fnv(uint64_t& h, const void* data, size_t n)
{
    const unsigned char* p = static_cast<const unsigned char*>(data);
    for (size_t i = 0; i < n; i++)
    {
        h ^= p[i];
        h *= 1099511628211ull;
    }
}

} // namespace

extern "C" {

const char* CUDAPOA_CABI(last_error)(void)
{
    return g_last_error.c_str();
}

const char* CUDAPOA_CABI(implementation)(void)
{
#ifdef CUDAPOA_B200_NATIVE
    return "b200-native";
#else
    return "genomeworks-reference";
#endif
}

int32_t CUDAPOA_CABI(init)(void)
{
    return guarded([&]() { return static_cast<int32_t>(poa::Init()); });
}

int32_t CUDAPOA_CABI(batch_size_default)(cudapoa_b200_batch_size* out, int32_t max_seq_sz, int32_t max_seq_per_poa,
                                         int32_t band_width)
{
    return guarded([&]() {
        poa::BatchSize bs(max_seq_sz, max_seq_per_poa, band_width);
        to_c(bs, out);
        return 0;
    });
}

int32_t CUDAPOA_CABI(batch_size_explicit)(cudapoa_b200_batch_size* out, int32_t max_seq_sz, int32_t max_consensus_sz,
                                          int32_t max_nodes_per_w, int32_t max_nodes_per_w_banded, int32_t band_width,
                                          int32_t max_seq_per_poa)
{
    return guarded([&]() {
        poa::BatchSize bs(max_seq_sz, max_consensus_sz, max_nodes_per_w, max_nodes_per_w_banded, band_width,
                          max_seq_per_poa);
        to_c(bs, out);
        return 0;
    });
}

cudapoa_b200_batch* CUDAPOA_CABI(create_batch)(int32_t device_id, void* stream, size_t max_mem, int8_t output_mask,
                                               const cudapoa_b200_batch_size* batch_size, int16_t gap_score,
                                               int16_t mismatch_score, int16_t match_score,
                                               int32_t cuda_banded_alignment)
{
    cudapoa_b200_batch* handle = nullptr;
    int32_t rc                 = guarded([&]() {
        std::unique_ptr<cudapoa_b200_batch> h(new cudapoa_b200_batch());
#ifdef CUDAPOA_B200_NATIVE
        if (cuda_banded_alignment == 2) // extension: BandMode::adaptive_band
            h->impl = poa::create_batch(device_id, static_cast<cudaStream_t>(stream), max_mem, output_mask,
                                        from_c(*batch_size), gap_score, mismatch_score, match_score,
                                        poa::BandMode::adaptive_band);
        else
#else
        if (cuda_banded_alignment == 2)
            throw std::invalid_argument("adaptive banding is not part of this cudapoa build");
#endif
            h->impl = poa::create_batch(device_id, static_cast<cudaStream_t>(stream), max_mem, output_mask,
                                        from_c(*batch_size), gap_score, mismatch_score, match_score,
                                        cuda_banded_alignment != 0);
        handle  = h.release();
        return 0;
    });
    return rc == 0 ? handle : nullptr;
}

void CUDAPOA_CABI(destroy_batch)(cudapoa_b200_batch* batch)
{
    delete batch;
}

int32_t CUDAPOA_CABI(add_poa_group)(cudapoa_b200_batch* batch, const char* const* seqs, const int8_t* const* weights,
                                    const int32_t* lengths, int32_t n_entries, int32_t* per_seq_status,
                                    int32_t* n_status)
{
    return guarded([&]() {
        batch->group.clear();
        for (int32_t i = 0; i < n_entries; i++)
        {
            poa::Entry e{};
            e.seq     = seqs[i];
            e.weights = (weights != nullptr) ? weights[i] : nullptr;
            e.length  = lengths[i];
            batch->group.push_back(e);
        }
        batch->seq_status.clear();
        poa::StatusType st = batch->impl->add_poa_group(batch->seq_status, batch->group);
        int32_t n          = static_cast<int32_t>(batch->seq_status.size());
        if (per_seq_status != nullptr)
        {
            for (int32_t i = 0; i < n && i < n_entries; i++)
                per_seq_status[i] = static_cast<int32_t>(batch->seq_status[i]);
        }
        if (n_status != nullptr)
            *n_status = n;
        return static_cast<int32_t>(st);
    });
}

int32_t CUDAPOA_CABI(add_poa_groups_flat)(cudapoa_b200_batch* batch, const uint8_t* bases, const int8_t* weights,
                                          const int32_t* read_lengths, const int32_t* window_n_reads,
                                          int32_t n_windows, int32_t* n_added, int32_t* per_seq_status)
{
    return guarded([&]() {
        int64_t base_off = 0;
        int64_t read_idx = 0;
        int32_t added    = 0;
        int32_t last     = static_cast<int32_t>(poa::StatusType::success);
        for (int32_t w = 0; w < n_windows; w++)
        {
            batch->group.clear();
            int64_t off = base_off;
            for (int32_t r = 0; r < window_n_reads[w]; r++)
            {
                poa::Entry e{};
                e.seq     = reinterpret_cast<const char*>(bases + off);
                e.weights = (weights != nullptr) ? (weights + off) : nullptr;
                e.length  = read_lengths[read_idx + r];
                off += e.length;
                batch->group.push_back(e);
            }
            batch->seq_status.clear();
            poa::StatusType st = batch->impl->add_poa_group(batch->seq_status, batch->group);
            last               = static_cast<int32_t>(st);
            if (st != poa::StatusType::success)
                break;
            if (per_seq_status != nullptr)
            {
                for (size_t i = 0; i < batch->seq_status.size(); i++)
                    per_seq_status[read_idx + static_cast<int64_t>(i)] = static_cast<int32_t>(batch->seq_status[i]);
            }
            base_off = off;
            read_idx += window_n_reads[w];
            added++;
        }
        if (n_added != nullptr)
            *n_added = added;
        return last;
    });
}

int32_t CUDAPOA_CABI(get_total_poas)(const cudapoa_b200_batch* batch)
{
    return guarded([&]() { return batch->impl->get_total_poas(); });
}

int32_t CUDAPOA_CABI(generate_poa)(cudapoa_b200_batch* batch)
{
    return guarded([&]() {
        batch->impl->generate_poa();
        return 0;
    });
}

int32_t CUDAPOA_CABI(batch_id)(const cudapoa_b200_batch* batch)
{
    return guarded([&]() { return batch->impl->batch_id(); });
}

int32_t CUDAPOA_CABI(reset)(cudapoa_b200_batch* batch)
{
    return guarded([&]() {
        batch->impl->reset();
        return 0;
    });
}

int32_t CUDAPOA_CABI(get_consensus)(cudapoa_b200_batch* batch)
{
    return guarded([&]() {
        clear_results(batch);
        poa::StatusType st = batch->impl->get_consensus(batch->consensus, batch->coverage, batch->status);
        batch->result_kind = 1;
        return static_cast<int32_t>(st);
    });
}

int32_t CUDAPOA_CABI(get_msa)(cudapoa_b200_batch* batch)
{
    return guarded([&]() {
        clear_results(batch);
        poa::StatusType st = batch->impl->get_msa(batch->msa, batch->status);
        batch->result_kind = 2;
        return static_cast<int32_t>(st);
    });
}

int32_t CUDAPOA_CABI(get_graphs)(cudapoa_b200_batch* batch)
{
    return guarded([&]() {
        clear_results(batch);
        batch->impl->get_graphs(batch->graphs, batch->status);
        batch->result_kind = 3;
        return 0;
    });
}

int32_t CUDAPOA_CABI(result_count)(const cudapoa_b200_batch* batch)
{
    switch (batch->result_kind)
    {
    case 1: return static_cast<int32_t>(batch->consensus.size());
    case 2: return static_cast<int32_t>(batch->msa.size());
    case 3: return static_cast<int32_t>(batch->graphs.size());
    default: return 0;
    }
}

int32_t CUDAPOA_CABI(result_status_count)(const cudapoa_b200_batch* batch)
{
    return static_cast<int32_t>(batch->status.size());
}

int32_t CUDAPOA_CABI(result_status)(const cudapoa_b200_batch* batch, int32_t i)
{
    if (i < 0 || i >= static_cast<int32_t>(batch->status.size()))
        return -1;
    return static_cast<int32_t>(batch->status[i]);
}

int32_t CUDAPOA_CABI(result_consensus)(const cudapoa_b200_batch* batch, int32_t i, const char** seq,
                                       const uint16_t** coverage)
{
    if (batch->result_kind != 1 || i < 0 || i >= static_cast<int32_t>(batch->consensus.size()))
        return -1;
    if (seq != nullptr)
        *seq = batch->consensus[i].c_str();
    if (coverage != nullptr)
        *coverage = batch->coverage[i].data();
    return static_cast<int32_t>(batch->consensus[i].size());
}

int32_t CUDAPOA_CABI(result_msa_rows)(const cudapoa_b200_batch* batch, int32_t i)
{
    if (batch->result_kind != 2 || i < 0 || i >= static_cast<int32_t>(batch->msa.size()))
        return -1;
    return static_cast<int32_t>(batch->msa[i].size());
}

int32_t CUDAPOA_CABI(result_msa_row)(const cudapoa_b200_batch* batch, int32_t i, int32_t row, const char** seq)
{
    if (batch->result_kind != 2 || i < 0 || i >= static_cast<int32_t>(batch->msa.size()))
        return -1;
    if (row < 0 || row >= static_cast<int32_t>(batch->msa[i].size()))
        return -1;
    if (seq != nullptr)
        *seq = batch->msa[i][row].c_str();
    return static_cast<int32_t>(batch->msa[i][row].size());
}

int32_t CUDAPOA_CABI(result_graph_nodes)(const cudapoa_b200_batch* batch, int32_t i, char* labels, int32_t cap)
{
    if (batch->result_kind != 3 || i < 0 || i >= static_cast<int32_t>(batch->graphs.size()))
        return -1;
    const gw::DirectedGraph& g = batch->graphs[i];
    // Node ids are dense 0..n-1 in cudapoa graphs; find n as the first id without a label.
    int32_t n = 0;
    while (!g.get_node_label(n).empty())
        n++;
    for (int32_t k = 0; k < n && k < cap; k++)
        labels[k] = g.get_node_label(k)[0];
    return n;
}

int32_t CUDAPOA_CABI(result_graph_edges)(const cudapoa_b200_batch* batch, int32_t i, int32_t* src, int32_t* dst,
                                         int32_t* weight, int32_t cap)
{
    if (batch->result_kind != 3 || i < 0 || i >= static_cast<int32_t>(batch->graphs.size()))
        return -1;
    auto edges = batch->graphs[i].get_edges();
    std::vector<std::tuple<int32_t, int32_t, int32_t>> flat;
    flat.reserve(edges.size());
    for (const auto& e : edges)
        flat.emplace_back(e.first.second, e.first.first, e.second);
    std::sort(flat.begin(), flat.end());
    int32_t n = static_cast<int32_t>(flat.size());
    for (int32_t k = 0; k < n && k < cap; k++)
    {
        dst[k]    = std::get<0>(flat[k]);
        src[k]    = std::get<1>(flat[k]);
        weight[k] = std::get<2>(flat[k]);
    }
    return n;
}

int64_t CUDAPOA_CABI(result_total_chars)(const cudapoa_b200_batch* batch)
{
    int64_t total = 0;
    if (batch->result_kind == 1)
    {
        for (const auto& s : batch->consensus)
            total += static_cast<int64_t>(s.size());
    }
    else if (batch->result_kind == 2)
    {
        for (const auto& rows : batch->msa)
            for (const auto& s : rows)
                total += static_cast<int64_t>(s.size());
    }
    return total;
}

uint64_t CUDAPOA_CABI(result_digest)(const cudapoa_b200_batch* batch)
{
    uint64_t h = 14695981039346656037ull;
    for (size_t i = 0; i < batch->status.size(); i++)
    {
        int32_t st = static_cast<int32_t>(batch->status[i]);
        fnv(h, &st, sizeof(st));
    }
    if (batch->result_kind == 1)
    {
        for (size_t i = 0; i < batch->consensus.size(); i++)
        {
            int32_t len = static_cast<int32_t>(batch->consensus[i].size());
            fnv(h, &len, sizeof(len));
            fnv(h, batch->consensus[i].data(), batch->consensus[i].size());
            fnv(h, batch->coverage[i].data(), batch->coverage[i].size() * sizeof(uint16_t));
        }
    }
    else if (batch->result_kind == 2)
    {
        for (const auto& rows : batch->msa)
        {
            int32_t n = static_cast<int32_t>(rows.size());
            fnv(h, &n, sizeof(n));
            for (const auto& s : rows)
            {
                int32_t len = static_cast<int32_t>(s.size());
                fnv(h, &len, sizeof(len));
                fnv(h, s.data(), s.size());
            }
        }
    }
    return h;
}

int32_t CUDAPOA_CABI(get_multi_batch_sizes)(const int32_t* read_lengths, const int32_t* window_n_reads,
                                            int32_t n_windows, int32_t banded_alignment, int32_t msa_flag,
                                            int32_t band_width, float gpu_memory_usage_quota, int32_t mismatch_score,
                                            int32_t gap_score, int32_t match_score, cudapoa_b200_batch_size* out,
                                            int32_t cap, int32_t* window_to_batch)
{
    return guarded([&]() {
        // Only Entry::length is consulted by the binning (cudapoa/src/utils.cu:40-53).
        std::vector<poa::Group> groups(static_cast<size_t>(n_windows));
        int64_t read_idx = 0;
        for (int32_t w = 0; w < n_windows; w++)
        {
            for (int32_t r = 0; r < window_n_reads[w]; r++)
            {
                poa::Entry e{};
                e.seq     = nullptr;
                e.weights = nullptr;
                e.length  = read_lengths[read_idx++];
                groups[static_cast<size_t>(w)].push_back(e);
            }
        }
        std::vector<poa::BatchSize> sizes;
        std::vector<std::vector<int32_t>> ids;
        poa::get_multi_batch_sizes(sizes, ids, groups, banded_alignment != 0, msa_flag != 0, band_width, nullptr,
                                   gpu_memory_usage_quota, mismatch_score, gap_score, match_score);
        if (window_to_batch != nullptr)
        {
            for (int32_t w = 0; w < n_windows; w++)
                window_to_batch[w] = -1;
            for (size_t b = 0; b < ids.size(); b++)
                for (int32_t w : ids[b])
                    window_to_batch[w] = static_cast<int32_t>(b);
        }
        int32_t n = static_cast<int32_t>(sizes.size());
        for (int32_t b = 0; b < n && b < cap; b++)
            to_c(sizes[static_cast<size_t>(b)], &out[b]);
        return n;
    });
}

int32_t CUDAPOA_CABI(get_stats)(const cudapoa_b200_batch* batch, cudapoa_b200_stats* out)
{
#ifdef CUDAPOA_B200_NATIVE
    return guarded([&]() {
        const auto* src = dynamic_cast<const poa::EngineStatsSource*>(batch->impl.get());
        if (src == nullptr || out == nullptr)
            return -1;
        src->fill_stats(out);
        return 0;
    });
#else
    (void)batch;
    if (out != nullptr)
        std::memset(out, 0, sizeof(*out));
    return -1;
#endif
}

int32_t CUDAPOA_CABI(max_poas)(const cudapoa_b200_batch* batch)
{
#ifdef CUDAPOA_B200_NATIVE
    return guarded([&]() {
        const auto* src = dynamic_cast<const poa::EngineStatsSource*>(batch->impl.get());
        return src == nullptr ? -1 : src->max_poas();
    });
#else
    (void)batch;
    return -1;
#endif
}

#ifdef CUDAPOA_B200_NATIVE
double cudapoa_b200_bench_relaunch(cudapoa_b200_batch* batch, int32_t iterations)
{
    double ms  = -1.0;
    int32_t rc = guarded([&]() {
        auto* src = dynamic_cast<poa::EngineStatsSource*>(batch->impl.get());
        if (src == nullptr)
            return -1;
        ms = src->relaunch(iterations);
        return 0;
    });
    return rc == 0 ? ms : -1.0;
}

double cudapoa_b200_bench_relaunch_timed(cudapoa_b200_batch* batch, int32_t iterations, double* fill_ms,
                                         double* graph_ms, int32_t* launches)
{
    double ms  = -1.0;
    int32_t rc = guarded([&]() {
        auto* src = dynamic_cast<poa::EngineStatsSource*>(batch->impl.get());
        if (src == nullptr)
            return -1;
        ms = src->relaunch_timed(iterations, fill_ms, graph_ms, launches);
        return 0;
    });
    return rc == 0 ? ms : -1.0;
}

int32_t cudapoa_b200_bench_phase_ms(cudapoa_b200_batch* batch, double* out, int32_t cap)
{
    int32_t n  = -1;
    int32_t rc = guarded([&]() {
        auto* src = dynamic_cast<poa::EngineStatsSource*>(batch->impl.get());
        if (src == nullptr || out == nullptr || cap <= 0)
            return -1;
        n = src->last_phase_ms(out, cap);
        return 0;
    });
    return rc == 0 ? n : -1;
}

int32_t cudapoa_b200_bench_relaunch_phases(cudapoa_b200_batch* batch, double* fill_ms, double* graph_ms)
{
    int32_t n  = -1;
    int32_t rc = guarded([&]() {
        auto* src = dynamic_cast<poa::EngineStatsSource*>(batch->impl.get());
        if (src == nullptr)
            return -1;
        n = src->relaunch_phases(fill_ms, graph_ms);
        return 0;
    });
    return rc == 0 ? n : -1;
}

int64_t cudapoa_b200_debug_read_window(cudapoa_b200_batch* batch, int32_t window, int32_t region, void* out, int64_t cap)
{
    int64_t n  = -1;
    int32_t rc = guarded([&]() {
        auto* src = dynamic_cast<poa::EngineStatsSource*>(batch->impl.get());
        if (src == nullptr)
            return -1;
        n = src->debug_read_window(window, region, out, cap);
        return 0;
    });
    return rc == 0 ? n : -1;
}
#endif

} // extern "C"
