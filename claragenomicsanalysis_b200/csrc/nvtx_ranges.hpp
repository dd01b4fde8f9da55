// NVTX ranges around the host-side stages (stage / launch / collect) so that a timeline tool shows where a batch's
// wall time goes (the reference offers the same through GW_NVTX_RANGE,
// common/base/include/claraparabricks/genomeworks/utils/cudautils.hpp:150-154, without using it in cudapoa).  Header-only NVTX v3: the calls are no-ops
// unless a tool injects itself.
#pragma once

#if defined(POA_EMU) || defined(CUDAPOA_B200_NO_NVTX)
struct NvtxRange
{
    explicit NvtxRange(const char*) {}
};
#else
#include <nvtx3/nvToolsExt.h>
struct NvtxRange
{
    explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
    ~NvtxRange() { nvtxRangePop(); }
    NvtxRange(const NvtxRange&)            = delete;
    NvtxRange& operator=(const NvtxRange&) = delete;
};
#endif
