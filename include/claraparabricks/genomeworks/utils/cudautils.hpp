// B200 cudapoa engine -- CUDA helper subset used by the cudapoa public headers and engine.
// API-compatible subset of GenomeWorks common/base/include/claraparabricks/genomeworks/utils/cudautils.hpp
// (GW_CU_CHECK_ERR :30-88, cudautils::align :91-105, scoped_device_switch :186-210).
#pragma once

#include <cuda_runtime_api.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>

#ifndef __CUDACC__
#ifndef __host__
#define __host__
#endif
#ifndef __device__
#define __device__
#endif
#endif

/// Abort with a message when a CUDA runtime call fails (the reference logs and calls std::abort()).
#define GW_CU_CHECK_ERR(call)                                                                            \
    do                                                                                                   \
    {                                                                                                    \
        ::claraparabricks::genomeworks::cudautils::gpu_assert((call), __FILE__, __LINE__);               \
    } while (0)
#define GW_CU_ABORT_ON_ERR(call) GW_CU_CHECK_ERR(call)

namespace claraparabricks
{
namespace genomeworks
{
namespace cudautils
{

inline void gpu_assert(cudaError_t code, const char* file, int line)
{
    if (code != cudaSuccess)
    {
        std::fprintf(stderr, "GPU Error:: %s %s %d\n", cudaGetErrorString(code), file, line);
        std::fflush(stderr);
        std::abort();
    }
}

/// Round value up to a multiple of boundary (a power of two).
template <typename IntType, int32_t boundary>
__host__ __device__ inline IntType align(const IntType& value)
{
    static_assert((boundary & (boundary - 1)) == 0, "boundary must be a power of two");
    return (value + boundary - 1) & ~(static_cast<IntType>(boundary) - 1);
}

} // namespace cudautils

/// RAII: make device_id current for the lifetime of the object, then restore the previous device.
class scoped_device_switch
{
public:
    explicit scoped_device_switch(int32_t device_id)
    {
        GW_CU_CHECK_ERR(cudaGetDevice(&previous_));
        GW_CU_CHECK_ERR(cudaSetDevice(device_id));
    }
    ~scoped_device_switch() { cudaSetDevice(previous_); }
    scoped_device_switch()                            = delete;
    scoped_device_switch(const scoped_device_switch&) = delete;
    scoped_device_switch& operator=(const scoped_device_switch&) = delete;

private:
    int32_t previous_ = 0;
};

} // namespace genomeworks
} // namespace claraparabricks
