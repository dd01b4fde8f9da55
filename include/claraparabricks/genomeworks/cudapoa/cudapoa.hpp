// B200 cudapoa engine -- package-level declarations.
// API-compatible with GenomeWorks cudapoa/include/claraparabricks/genomeworks/cudapoa/cudapoa.hpp:26-48:
// same enumerators, same numeric values (the Python binding depends on them).
#pragma once

namespace claraparabricks
{
namespace genomeworks
{
namespace cudapoa
{

/// Status / error codes of the cudapoa API.
enum StatusType
{
    success                                   = 0,
    exceeded_maximum_poas                     = 1,
    exceeded_maximum_sequence_size            = 2,
    exceeded_maximum_sequences_per_poa        = 3,
    node_count_exceeded_maximum_graph_size    = 4,
    edge_count_exceeded_maximum_graph_size    = 5,
    seq_len_exceeded_maximum_nodes_per_window = 6,
    loop_count_exceeded_upper_bound           = 7,
    output_type_unavailable                   = 8,
    generic_error                             = 9
};

/// Initialise the package (idempotent).
StatusType Init();

/// Output selection bit mask for create_batch().
enum OutputType
{
    consensus = 0x1,
    msa       = 0x1 << 1
};

} // namespace cudapoa
} // namespace genomeworks
} // namespace claraparabricks
