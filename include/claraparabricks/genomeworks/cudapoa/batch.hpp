// B200 cudapoa engine -- batch interface.
// API-compatible with GenomeWorks cudapoa/include/claraparabricks/genomeworks/cudapoa/batch.hpp:
// Entry/Group :39-50, BatchSize :53-129 (same nine public fields, same two constructors, same validation and
// messages), Batch :133-205 (same pure virtuals), create_batch :220-228.  `BatchConfig` is offered as an alias.
#pragma once

#include <claraparabricks/genomeworks/cudapoa/cudapoa.hpp>
#include <claraparabricks/genomeworks/utils/cudautils.hpp>
#include <claraparabricks/genomeworks/utils/graph.hpp>
#include <claraparabricks/genomeworks/utils/signed_integer_utils.hpp>

#include <cuda_runtime_api.h>
#include <stdint.h>

#include <iostream>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace claraparabricks
{
namespace genomeworks
{
namespace cudapoa
{

/// One sequence of a POA group.  The caller owns the memory; it is copied during add_poa_group().
struct Entry
{
    const char* seq;       ///< bases
    const int8_t* weights; ///< per-base weights (>= 0), or nullptr for all ones
    int32_t length;        ///< number of bases
};

/// Sequences of one window, in the order they are added to the graph (entry 0 is the backbone).
typedef std::vector<Entry> Group;

/// Upper limits a batch is sized for.
struct BatchSize
{
    int32_t max_sequence_size;                 ///< longest sequence accepted
    int32_t max_consensus_size;                ///< longest consensus / MSA row
    int32_t max_nodes_per_window;              ///< graph node limit, full alignment
    int32_t max_nodes_per_window_banded;       ///< graph node limit, banded alignment
    int32_t max_matrix_graph_dimension;        ///< score matrix rows, full alignment
    int32_t max_matrix_graph_dimension_banded; ///< score matrix rows, banded alignment
    int32_t max_matrix_sequence_dimension;     ///< score matrix columns
    int32_t alignment_band_width;              ///< band width (multiple of 128)
    int32_t max_sequences_per_poa;             ///< most sequences in one group

    /// Derive every limit from the maximum sequence size.
    BatchSize(int32_t max_seq_sz = 1024, int32_t max_seq_per_poa = 100, int32_t band_width = 256)
        : max_sequence_size(max_seq_sz)
        , max_consensus_size(2 * max_sequence_size)
        , max_nodes_per_window(cudautils::align<int32_t, 4>(3 * max_sequence_size))
        , max_nodes_per_window_banded(cudautils::align<int32_t, 4>(4 * max_sequence_size))
        , max_matrix_graph_dimension(cudautils::align<int32_t, 4>(max_nodes_per_window))
        , max_matrix_graph_dimension_banded(cudautils::align<int32_t, 4>(max_nodes_per_window_banded))
        , max_matrix_sequence_dimension(cudautils::align<int32_t, 4>(max_sequence_size))
        , alignment_band_width(cudautils::align<int32_t, 128>(band_width))
        , max_sequences_per_poa(max_seq_per_poa)
    {
        throw_on_negative(max_seq_sz, "max_sequence_size cannot be negative.");
        throw_on_negative(max_seq_per_poa, "max_sequences_per_poa cannot be negative.");
        throw_on_negative(band_width, "alignment_band_width cannot be negative.");
        note_band_rounding(band_width);
    }

    /// Set the limits individually.
    BatchSize(int32_t max_seq_sz, int32_t max_consensus_sz, int32_t max_nodes_per_w, int32_t max_nodes_per_w_banded,
              int32_t band_width, int32_t max_seq_per_poa)
        : max_sequence_size(max_seq_sz)
        , max_consensus_size(max_consensus_sz)
        , max_nodes_per_window(cudautils::align<int32_t, 4>(max_nodes_per_w))
        , max_nodes_per_window_banded(cudautils::align<int32_t, 4>(max_nodes_per_w_banded))
        , max_matrix_graph_dimension(cudautils::align<int32_t, 4>(max_nodes_per_window))
        , max_matrix_graph_dimension_banded(cudautils::align<int32_t, 4>(max_nodes_per_window_banded))
        , max_matrix_sequence_dimension(cudautils::align<int32_t, 4>(max_sequence_size))
        , alignment_band_width(cudautils::align<int32_t, 128>(band_width))
        , max_sequences_per_poa(max_seq_per_poa)
    {
        throw_on_negative(max_seq_sz, "max_sequence_size cannot be negative.");
        throw_on_negative(max_consensus_sz, "max_consensus_size cannot be negative.");
        throw_on_negative(max_nodes_per_w, "max_nodes_per_window cannot be negative.");
        throw_on_negative(max_nodes_per_w_banded, "max_nodes_per_window_banded cannot be negative.");
        throw_on_negative(max_seq_per_poa, "max_sequences_per_poa cannot be negative.");
        throw_on_negative(band_width, "alignment_band_width cannot be negative.");
        if (max_nodes_per_window < max_sequence_size)
            throw std::invalid_argument("max_nodes_per_window should be greater than or equal to max_sequence_size.");
        if (max_nodes_per_window_banded < max_sequence_size)
            throw std::invalid_argument("max_nodes_per_window should be greater than or equal to max_sequence_size.");
        if (max_consensus_size < max_sequence_size)
            throw std::invalid_argument("max_consensus_size should be greater than or equal to max_sequence_size.");
        if (max_sequence_size < alignment_band_width)
            throw std::invalid_argument("alignment_band_width should not be greater than max_sequence_size.");
    }

private:
    void note_band_rounding(int32_t requested) const
    {
        if (alignment_band_width != requested)
        {
            std::cerr << "Band-width should be multiple of 128. The input was changed from " << requested << " to "
                      << alignment_band_width << std::endl;
        }
    }
};

/// Name used by the B200 north star for the same structure.
using BatchConfig = BatchSize;

/// A batch of POA groups processed together on one GPU stream.  Not thread-safe; use one batch per host thread.
class Batch
{
public:
    virtual ~Batch() = default;

    /// Add one group.  per_seq_status is cleared and then receives one status per entry of the group; entries
    /// that do not fit the batch limits are skipped and reported there.  Returns exceeded_maximum_poas when the
    /// batch is full (per_seq_status untouched in that case).
    virtual StatusType add_poa_group(std::vector<StatusType>& per_seq_status, const Group& poa_group) = 0;

    /// Number of groups added since the last reset().
    virtual int32_t get_total_poas() const = 0;

    /// Enqueue the POA of all groups on the batch's stream (asynchronous).
    virtual void generate_poa() = 0;

    /// Append one consensus string, its per-base coverage and a status per group.
    virtual StatusType get_consensus(std::vector<std::string>& consensus,
                                     std::vector<std::vector<uint16_t>>& coverage,
                                     std::vector<genomeworks::cudapoa::StatusType>& output_status) = 0;

    /// Append the multiple sequence alignment (one row per sequence) and a status per group.
    virtual StatusType get_msa(std::vector<std::vector<std::string>>& msa, std::vector<StatusType>& output_status) = 0;

    /// Resize graphs to the number of groups and fill in the POA graph of each; append a status per group.
    virtual void get_graphs(std::vector<DirectedGraph>& graphs, std::vector<StatusType>& output_status) = 0;

    /// Id of this batch (unique per process among batches of the same kind).
    virtual int32_t batch_id() const = 0;

    /// Forget all groups; must be called before the batch is reused.
    virtual void reset() = 0;
};

/// Create a batch on `device_id` that enqueues its work on `stream` and uses at most `max_mem` bytes of device
/// memory.  `output_mask` selects OutputType::consensus or OutputType::msa.
std::unique_ptr<Batch> create_batch(int32_t device_id, cudaStream_t stream, size_t max_mem, int8_t output_mask,
                                    const BatchSize& batch_size, int16_t gap_score, int16_t mismatch_score,
                                    int16_t match_score, bool cuda_banded_alignment);

/// Extension of this engine (names follow later GenomeWorks releases): how the score matrix is banded.
enum class BandMode : int8_t
{
    full_band     = 0, ///< cuda_banded_alignment == false
    static_band   = 1, ///< cuda_banded_alignment == true
    adaptive_band = 2  ///< static band first; the band doubles (up to 4x) while the alignment runs along a band edge
};

/// Same as create_batch above with the band mode spelled out; adaptive_band has no counterpart in the reference.
std::unique_ptr<Batch> create_batch(int32_t device_id, cudaStream_t stream, size_t max_mem, int8_t output_mask,
                                    const BatchSize& batch_size, int16_t gap_score, int16_t mismatch_score,
                                    int16_t match_score, BandMode band_mode);

} // namespace cudapoa
} // namespace genomeworks
} // namespace claraparabricks
