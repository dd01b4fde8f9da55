// TEST INFRASTRUCTURE ONLY -- never part of the product library.
//
// A small SIMT emulator: lets the engine's device code (csrc/*.cuh) and its host side (csrc/engine.cu) be compiled
// with g++ and run on a machine without a GPU, so that kernel logic can be checked bit for bit against the oracle
// before GPU time is spent on it.  The product (libcudapoa_b200.so) is built by nvcc from the same sources and has no
// reference to this file; tests/emu/build.py compiles libcudapoa_emu.so with -DPOA_EMU and this directory in front
// of the include path, where cuda_runtime.h / cuda_runtime_api.h are stand-ins that include this header.
//
// Model: one OS thread.  A kernel launch runs block after block, warp after warp; the 32 lanes of a warp are fibers
// (hand-written x86-64 context switch) scheduled round-robin.  A lane runs until it reaches a warp collective
// (__shfl*_sync, __ballot_sync, __any/__all_sync, __syncwarp), where it waits for the other lanes named in the mask;
// the last lane to arrive computes everybody's result.  Lanes that reach different collectives with the same mask at
// the same time (divergent code around a collective) are reported with the source lines, and so are deadlocks.
// Shared memory is a per-block byte array; the 32-bit "shared addresses" the device code computes are offsets into
// it and every access is bounds-checked.  Global memory is host memory (cudaMalloc = malloc).
#pragma once

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>

#define __global__
#define __device__
#define __host__
#define __forceinline__ inline __attribute__((always_inline))
#define __launch_bounds__(...)
#define __align__(n) __attribute__((aligned(n)))
#define __restrict__

// ---- CUDA vector types ----------------------------------------------------------------------------------------
struct uint2
{
    unsigned int x, y;
};
struct uint3
{
    unsigned int x, y, z;
};
struct uint4
{
    unsigned int x, y, z, w;
};
struct int2
{
    int x, y;
};
struct int4
{
    int x, y, z, w;
};
static inline uint2 make_uint2(unsigned x, unsigned y) { return uint2{x, y}; }
static inline uint4 make_uint4(unsigned x, unsigned y, unsigned z, unsigned w) { return uint4{x, y, z, w}; }
static inline int2 make_int2(int x, int y) { return int2{x, y}; }
struct dim3
{
    unsigned int x, y, z;
    dim3(unsigned int x_ = 1, unsigned int y_ = 1, unsigned int z_ = 1) : x(x_), y(y_), z(z_) {}
};

// ---- runtime API stand-ins ------------------------------------------------------------------------------------
typedef int cudaError_t;
enum : int
{
    cudaSuccess = 0,
    cudaErrorInvalidValue = 1,
    cudaErrorMemoryAllocation = 2
};
typedef struct poa_emu_stream* cudaStream_t;
typedef struct poa_emu_event* cudaEvent_t;
enum cudaMemcpyKind
{
    cudaMemcpyHostToHost = 0,
    cudaMemcpyHostToDevice = 1,
    cudaMemcpyDeviceToHost = 2,
    cudaMemcpyDeviceToDevice = 3,
    cudaMemcpyDefault = 4
};
enum cudaFuncAttribute
{
    cudaFuncAttributeMaxDynamicSharedMemorySize = 8
};
enum cudaDeviceAttr
{
    cudaDevAttrMultiProcessorCount = 16
};
enum : unsigned
{
    cudaHostAllocDefault = 0
};
struct cudaDeviceProp
{
    int multiProcessorCount;
    char name[64];
};

static inline const char* cudaGetErrorString(cudaError_t e) { return e == 0 ? "no error" : "emulated CUDA error"; }
static inline cudaError_t cudaGetDevice(int* d)
{
    *d = 0;
    return cudaSuccess;
}
static inline cudaError_t cudaSetDevice(int d) { return d == 0 ? cudaSuccess : cudaErrorInvalidValue; }
static inline cudaError_t cudaMalloc(void** p, size_t n)
{
    *p = std::malloc(n ? n : 1);
    return *p ? cudaSuccess : cudaErrorMemoryAllocation;
}
static inline cudaError_t cudaFree(void* p)
{
    std::free(p);
    return cudaSuccess;
}
static inline cudaError_t cudaHostAlloc(void** p, size_t n, unsigned) { return cudaMalloc(p, n); }
static inline cudaError_t cudaFreeHost(void* p) { return cudaFree(p); }
static inline cudaError_t cudaMemcpy(void* d, const void* s, size_t n, cudaMemcpyKind)
{
    std::memcpy(d, s, n);
    return cudaSuccess;
}
static inline cudaError_t cudaMemcpyAsync(void* d, const void* s, size_t n, cudaMemcpyKind, cudaStream_t = nullptr)
{
    std::memcpy(d, s, n);
    return cudaSuccess;
}
static inline cudaError_t cudaMemcpy2D(void* d, size_t dpitch, const void* s, size_t spitch, size_t width, size_t height,
                                       cudaMemcpyKind)
{
    for (size_t r = 0; r < height; r++)
        std::memcpy(static_cast<char*>(d) + r * dpitch, static_cast<const char*>(s) + r * spitch, width);
    return cudaSuccess;
}
static inline cudaError_t cudaMemcpy2DAsync(void* d, size_t dpitch, const void* s, size_t spitch, size_t width,
                                            size_t height, cudaMemcpyKind kind, cudaStream_t = nullptr)
{
    return cudaMemcpy2D(d, dpitch, s, spitch, width, height, kind);
}
static inline cudaError_t cudaMemsetAsync(void* d, int v, size_t n, cudaStream_t = nullptr)
{
    std::memset(d, v, n);
    return cudaSuccess;
}
static inline cudaError_t cudaMemset(void* d, int v, size_t n)
{
    std::memset(d, v, n);
    return cudaSuccess;
}
static inline cudaError_t cudaStreamSynchronize(cudaStream_t) { return cudaSuccess; }
constexpr unsigned cudaStreamNonBlocking = 1u;
static inline cudaError_t cudaStreamCreateWithFlags(cudaStream_t* s, unsigned)
{
    *s = nullptr; // every launch runs to completion inside the call: streams carry no state
    return cudaSuccess;
}
static inline cudaError_t cudaStreamDestroy(cudaStream_t) { return cudaSuccess; }
static inline cudaError_t cudaDeviceSynchronize() { return cudaSuccess; }
static inline cudaError_t cudaEventCreate(cudaEvent_t* e)
{
    *e = nullptr;
    return cudaSuccess;
}
static inline cudaError_t cudaEventDestroy(cudaEvent_t) { return cudaSuccess; }
static inline cudaError_t cudaEventRecord(cudaEvent_t, cudaStream_t = nullptr) { return cudaSuccess; }
static inline cudaError_t cudaEventSynchronize(cudaEvent_t) { return cudaSuccess; }
static inline cudaError_t cudaEventElapsedTime(float* ms, cudaEvent_t, cudaEvent_t)
{
    *ms = 1.0f;
    return cudaSuccess;
}
static inline cudaError_t cudaPeekAtLastError() { return cudaSuccess; }
static inline cudaError_t cudaGetLastError() { return cudaSuccess; }
static inline cudaError_t cudaMemGetInfo(size_t* free_b, size_t* total_b)
{
    *free_b  = size_t(4) << 30;
    *total_b = size_t(4) << 30;
    return cudaSuccess;
}
static inline cudaError_t cudaDeviceGetAttribute(int* v, cudaDeviceAttr, int)
{
    *v = 2; // "SMs": keeps the emulated grids small
    return cudaSuccess;
}
static inline cudaError_t cudaGetDeviceProperties(cudaDeviceProp* p, int)
{
    std::memset(p, 0, sizeof(*p));
    p->multiProcessorCount = 2;
    std::strcpy(p->name, "poa-emu");
    return cudaSuccess;
}
template <typename K>
static inline cudaError_t cudaFuncSetAttribute(K, cudaFuncAttribute, int)
{
    return cudaSuccess;
}
template <typename K>
static inline cudaError_t cudaOccupancyMaxActiveBlocksPerMultiprocessor(int* n, K, int, size_t)
{
    *n = 1;
    return cudaSuccess;
}

// ---- the SIMT machine -----------------------------------------------------------------------------------------
namespace poa_emu
{

struct Machine;
Machine& machine();

void launch(dim3 grid, dim3 block, size_t smem_bytes, const std::function<void()>& body);
int lane_id();
uint8_t* smem_base();
size_t smem_size();
[[noreturn]] void fail(const char* what, int line);

enum Op
{
    kOpSync,
    kOpShflIdx,
    kOpShflUp,
    kOpShflDown,
    kOpShflXor,
    kOpBallot,
};
// One warp collective: every lane named in `mask` calls this with the same op; returns this lane's result.
uint32_t collective(unsigned mask, Op op, uint32_t value, int arg, int width, int line);

float rcp_approx(float d); // MUFU.RCP as measured on a B200 (tests/golden/b200_mufu_rcp_u16.npy)
void count(const char* name, long long n = 1); // coverage counters: which device-code paths a test exercised

} // namespace poa_emu

extern uint3 threadIdx, blockIdx;
extern dim3 blockDim, gridDim;

// ---- warp intrinsics --------------------------------------------------------------------------------------------
static inline void __syncwarp(unsigned mask = 0xffffffffu, int line = __builtin_LINE())
{
    poa_emu::collective(mask, poa_emu::kOpSync, 0, 0, 32, line);
}
#define POA_EMU_SHFL(NAME, OP)                                                                                       \
    static inline int NAME(unsigned mask, int v, int arg, int width = 32, int line = __builtin_LINE())              \
    {                                                                                                                \
        return static_cast<int>(poa_emu::collective(mask, poa_emu::OP, static_cast<uint32_t>(v), arg, width, line)); \
    }                                                                                                                \
    static inline unsigned NAME(unsigned mask, unsigned v, int arg, int width = 32, int line = __builtin_LINE())    \
    {                                                                                                                \
        return poa_emu::collective(mask, poa_emu::OP, v, arg, width, line);                                          \
    }                                                                                                                \
    static inline float NAME(unsigned mask, float v, int arg, int width = 32, int line = __builtin_LINE())          \
    {                                                                                                                \
        uint32_t b;                                                                                                  \
        std::memcpy(&b, &v, 4);                                                                                      \
        b = poa_emu::collective(mask, poa_emu::OP, b, arg, width, line);                                             \
        std::memcpy(&v, &b, 4);                                                                                      \
        return v;                                                                                                    \
    }
POA_EMU_SHFL(__shfl_sync, kOpShflIdx)
POA_EMU_SHFL(__shfl_up_sync, kOpShflUp)
POA_EMU_SHFL(__shfl_down_sync, kOpShflDown)
POA_EMU_SHFL(__shfl_xor_sync, kOpShflXor)
#undef POA_EMU_SHFL
static inline unsigned __ballot_sync(unsigned mask, int pred, int line = __builtin_LINE())
{
    return poa_emu::collective(mask, poa_emu::kOpBallot, pred ? 1u : 0u, 0, 32, line);
}
static inline int __any_sync(unsigned mask, int pred, int line = __builtin_LINE())
{
    return (poa_emu::collective(mask, poa_emu::kOpBallot, pred ? 1u : 0u, 0, 32, line) & mask) != 0;
}
static inline int __all_sync(unsigned mask, int pred, int line = __builtin_LINE())
{
    return (poa_emu::collective(mask, poa_emu::kOpBallot, pred ? 1u : 0u, 0, 32, line) & mask) == mask;
}

// ---- scalar intrinsics --------------------------------------------------------------------------------------
static inline int __popc(unsigned v) { return __builtin_popcount(v); }
static inline int __ffs(int v) { return __builtin_ffs(v); }
static inline int __clz(int v) { return v == 0 ? 32 : __builtin_clz(static_cast<unsigned>(v)); }
static inline int __float2int_rz(float f) { return static_cast<int>(f); } // values here are small and finite
template <typename T>
static inline T __ldcg(const T* p)
{
    return *p;
}
template <typename T>
static inline T __ldg(const T* p)
{
    return *p;
}
template <typename T>
static inline void __stcg(T* p, T v)
{
    *p = v;
}
static inline int atomicAdd(int* p, int v)
{
    const int old = *p;
    *p            = old + v;
    return old;
}
static inline unsigned atomicAdd(unsigned* p, unsigned v)
{
    const unsigned old = *p;
    *p                 = old + v;
    return old;
}
static inline unsigned atomicOr(unsigned* p, unsigned v)
{
    const unsigned old = *p;
    *p                 = old | v;
    return old;
}

// prmt.b32 (default mode): selector nibble = byte index 0..7 of {x, y}; bit 3 set = replicate that byte's sign bit
static inline unsigned poa_emu_prmt(unsigned x, unsigned y, unsigned s)
{
    const uint64_t src = (static_cast<uint64_t>(y) << 32) | x;
    unsigned out       = 0;
    for (int i = 0; i < 4; i++)
    {
        const unsigned sel = (s >> (4 * i)) & 0xfu;
        unsigned b         = static_cast<unsigned>(src >> (8 * (sel & 7u))) & 0xffu;
        if (sel & 8u)
            b = (b & 0x80u) ? 0xffu : 0u;
        out |= b << (8 * i);
    }
    return out;
}
static inline unsigned __byte_perm(unsigned x, unsigned y, unsigned s) { return poa_emu_prmt(x, y, s & 0x7777u); }

static inline int poa_emu_lo(unsigned v) { return static_cast<int16_t>(v & 0xffffu); }
static inline int poa_emu_hi(unsigned v) { return static_cast<int16_t>(v >> 16); }
static inline unsigned poa_emu_pack(int lo, int hi)
{
    return (static_cast<unsigned>(lo) & 0xffffu) | (static_cast<unsigned>(hi) << 16);
}
static inline unsigned __vadd2(unsigned a, unsigned b)
{
    return poa_emu_pack(poa_emu_lo(a) + poa_emu_lo(b), poa_emu_hi(a) + poa_emu_hi(b));
}
static inline unsigned __vsub2(unsigned a, unsigned b)
{
    return poa_emu_pack(poa_emu_lo(a) - poa_emu_lo(b), poa_emu_hi(a) - poa_emu_hi(b));
}
static inline unsigned __vmaxs2(unsigned a, unsigned b)
{
    return poa_emu_pack(std::max(poa_emu_lo(a), poa_emu_lo(b)), std::max(poa_emu_hi(a), poa_emu_hi(b)));
}
static inline unsigned __vmins2(unsigned a, unsigned b)
{
    return poa_emu_pack(std::min(poa_emu_lo(a), poa_emu_lo(b)), std::min(poa_emu_hi(a), poa_emu_hi(b)));
}
static inline unsigned __vminu2(unsigned a, unsigned b)
{
    return std::min(a & 0xffffu, b & 0xffffu) | (std::min(a >> 16, b >> 16) << 16);
}
static inline unsigned __vimax_s16x2(unsigned a, unsigned b) { return __vmaxs2(a, b); }
static inline unsigned __vimin_s16x2(unsigned a, unsigned b) { return __vmins2(a, b); }
static inline unsigned __vimin_u16x2(unsigned a, unsigned b) { return __vminu2(a, b); }
static inline unsigned __vibmax_s16x2(unsigned a, unsigned b, bool* pred_hi, bool* pred_lo)
{
    *pred_hi = poa_emu_hi(a) >= poa_emu_hi(b);
    *pred_lo = poa_emu_lo(a) >= poa_emu_lo(b);
    return __vmaxs2(a, b);
}
static inline unsigned __vimax3_s16x2(unsigned a, unsigned b, unsigned c) { return __vmaxs2(__vmaxs2(a, b), c); }
static inline unsigned __viaddmax_s16x2(unsigned a, unsigned b, unsigned c)
{
    // the sum wraps per half-word before the maximum (DPX semantics)
    return __vmaxs2(poa_emu_pack(static_cast<int16_t>(poa_emu_lo(a) + poa_emu_lo(b)),
                                 static_cast<int16_t>(poa_emu_hi(a) + poa_emu_hi(b))),
                    c);
}
static inline int __viaddmax_s32(int a, int b, int c)
{
    return std::max(static_cast<int>(static_cast<unsigned>(a) + static_cast<unsigned>(b)), c);
}
static inline unsigned __vcmpeq4(unsigned a, unsigned b)
{
    unsigned out = 0;
    for (int i = 0; i < 4; i++)
        if (((a >> (8 * i)) & 0xffu) == ((b >> (8 * i)) & 0xffu))
            out |= 0xffu << (8 * i);
    return out;
}
static inline unsigned __vcmpeq2(unsigned a, unsigned b)
{
    return (((a & 0xffffu) == (b & 0xffffu)) ? 0xffffu : 0u) | (((a >> 16) == (b >> 16)) ? 0xffff0000u : 0u);
}
static inline unsigned __vcmpgts2(unsigned a, unsigned b)
{
    return ((poa_emu_lo(a) > poa_emu_lo(b)) ? 0xffffu : 0u) | ((poa_emu_hi(a) > poa_emu_hi(b)) ? 0xffff0000u : 0u);
}

// CUDA's global min / max / abs overloads
static inline int min(int a, int b) { return a < b ? a : b; }
static inline int max(int a, int b) { return a > b ? a : b; }
static inline unsigned min(unsigned a, unsigned b) { return a < b ? a : b; }
static inline unsigned max(unsigned a, unsigned b) { return a > b ? a : b; }
static inline long long min(long long a, long long b) { return a < b ? a : b; }
static inline long long max(long long a, long long b) { return a > b ? a : b; }
static inline short min(short a, short b) { return a < b ? a : b; }
static inline short max(short a, short b) { return a > b ? a : b; }

// ---- kernel launch ------------------------------------------------------------------------------------------------
#define POA_EMU_LAUNCH(kernel, grid, block, smem, stream, ...) \
    poa_emu::launch((grid), (block), (smem), [&]() { kernel(__VA_ARGS__); })
