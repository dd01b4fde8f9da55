// DPX (s16x2) fast path of the banded int16 alignment.  See DESIGN.md "fill kernel".
#pragma once

#include "poa_device.cuh"

namespace poa_b200
{

constexpr int32_t kFastNotTaken = -2; // the caller must run align_exact instead

template <typename IdT>
__device__ __forceinline__ int32_t align_fast_banded16(const Graph<IdT>& g, int32_t node_count, const uint8_t* read,
                                                       int32_t read_len, int16_t* scores, int32_t band, int32_t gap,
                                                       int32_t mismatch, int32_t match, int32_t lane)
{
    (void)g;
    (void)node_count;
    (void)read;
    (void)read_len;
    (void)scores;
    (void)band;
    (void)gap;
    (void)mismatch;
    (void)match;
    (void)lane;
    return kFastNotTaken;
}

} // namespace poa_b200
