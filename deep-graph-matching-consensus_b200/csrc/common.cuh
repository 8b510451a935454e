// Shared helpers for the libdgmc_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#define DGMC_OK 0
#define DGMC_ERR_INVALID (-1)   // bad argument (shape, alignment, null pointer)
#define DGMC_ERR_CUDA (-2)      // a CUDA runtime call or launch failed
#define DGMC_ERR_UNSUPPORTED (-3)  // shape outside what the kernel was built for
#define DGMC_ERR_WORKSPACE (-4)    // caller-provided workspace too small

namespace dgmc {

void set_error(const char* fmt, ...);  // thread-local message, read by dgmc_last_error()

void count_launches(int n);  // process-wide tally of kernels launched by this library

inline int check_launch(const char* what, int kernels = 1) {
    count_launches(kernels);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        set_error("%s: %s", what, cudaGetErrorString(e));
        return DGMC_ERR_CUDA;
    }
    return DGMC_OK;
}

constexpr int MAX_DEVICES = 64;

inline int current_device() {
    int dev = 0;
    cudaGetDevice(&dev);
    return (dev >= 0 && dev < MAX_DEVICES) ? dev : 0;
}

// SM count of the CURRENT device (cached per device: one process may drive several GPUs)
inline int sm_count() {
    static int cached[MAX_DEVICES] = {};
    const int dev = current_device();
    if (!cached[dev]) {
        int n = 0;
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        cached[dev] = n > 0 ? n : 148;
    }
    return cached[dev];
}

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) is a per-DEVICE setting: a kernel configured on one
// GPU is not configured on the next one this process uses.  One flag array per kernel instantiation:
//   static DeviceOnce once;  if (!once.done()) { ...set the attribute...; once.mark(); }
struct DeviceOnce {
    bool flag[MAX_DEVICES] = {};
    bool done() const { return flag[current_device()]; }
    void mark() { flag[current_device()] = true; }
};

// cuTensorMapEncodeTiled is a DRIVER call and needs a current context on the calling thread.  A thread
// that has not yet made a runtime call that binds the primary context (e.g. PyTorch's autograd worker
// running the first backward pass of the process) has none: bind it once per thread.
inline void ensure_context() {
    static thread_local bool bound = false;
    if (!bound) {
        cudaFree(nullptr);
        bound = true;
    }
}

// grid for a grid-stride kernel: a whole number of waves of `per_sm` resident CTAs per SM
inline int wave_grid(long long work_items, int items_per_block, int per_sm) {
    long long need = (work_items + items_per_block - 1) / items_per_block;
    long long cap = (long long)sm_count() * per_sm;
    if (need < 1) need = 1;
    return (int)(need < cap ? need : cap);
}

#define DGMC_REQUIRE(cond, ...)                    \
    do {                                           \
        if (!(cond)) {                             \
            dgmc::set_error(__VA_ARGS__);          \
            return DGMC_ERR_INVALID;               \
        }                                          \
    } while (0)

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
// sum over an aligned group of G lanes (G power of two <= 32)
template <int G>
__device__ __forceinline__ float group_sum(float v) {
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__device__ __forceinline__ float4 ldg4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }

}  // namespace dgmc
