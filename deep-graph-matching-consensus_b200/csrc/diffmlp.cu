// K6: difference gather + consensus MLP update of S_hat (dgmc.py:219-223), first layer hoisted:
//   s_out[m, j] = s_in[m, j] + b2 + sum_c w2[c] * relu(ps[m, c] - pt[idx[m, j], c])
// ps = o_s W1^T + b1 and pt = o_t W1^T are per-node GEMMs done by the host; the per-pair work
// here is O(r) and D [B, N_s, k, R] is never materialised.  HBM/L2-bound gather kernels:
// one warp per source row, groups of G lanes per candidate, float4 rows.
#include "common.cuh"
#include "../../include/dgmc_b200.h"

namespace dgmc {

constexpr int DM_MAX_BLOCKS = 1024;  // rows of the partial-sum workspace of the backward

__device__ __forceinline__ float relu_dot(const float4 w, const float4 p, const float4 t) {
    return w.x * fmaxf(p.x - t.x, 0.f) + w.y * fmaxf(p.y - t.y, 0.f) +
           w.z * fmaxf(p.z - t.z, 0.f) + w.w * fmaxf(p.w - t.w, 0.f);
}

template <int G, int NV>
__global__ void __launch_bounds__(256) diff_mlp_fwd_kernel(
    const float* __restrict__ s_in, const float* __restrict__ ps, const float* __restrict__ pt,
    const int* __restrict__ idx, const float* __restrict__ w2, const float* __restrict__ b2,
    float* __restrict__ s_out, long long m, int k, int r4) {
    constexpr int NG = 32 / G;
    const int lane = threadIdx.x & 31, gl = lane % G, g = lane / G;
    const long long warps = (long long)gridDim.x * (blockDim.x >> 5);
    float4 wv[NV];
#pragma unroll
    for (int v = 0; v < NV; ++v) {
        const int q = gl + v * G;
        wv[v] = q < r4 ? ldg4(w2 + 4 * q) : make_float4(0, 0, 0, 0);
    }
    const float bias = __ldg(b2);
    for (long long row = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); row < m;
         row += warps) {
        float4 pv[NV];
#pragma unroll
        for (int v = 0; v < NV; ++v) {
            const int q = gl + v * G;
            pv[v] = q < r4 ? ldg4(ps + (size_t)row * r4 * 4 + 4 * q) : make_float4(0, 0, 0, 0);
        }
        const int* ir = idx + (size_t)row * k;
        for (int j0 = 0; j0 < k; j0 += 2 * NG) {
            const int ja = j0 + g, jb = j0 + NG + g;
            const int ta = ja < k ? __ldg(ir + ja) : 0, tb = jb < k ? __ldg(ir + jb) : 0;
            float sa = 0.f, sb = 0.f;
#pragma unroll
            for (int v = 0; v < NV; ++v) {
                const int q = gl + v * G;
                if (q < r4) {
                    const float4 qa = ldg4(pt + (size_t)ta * r4 * 4 + 4 * q);
                    const float4 qb = ldg4(pt + (size_t)tb * r4 * 4 + 4 * q);
                    sa += relu_dot(wv[v], pv[v], qa);
                    sb += relu_dot(wv[v], pv[v], qb);
                }
            }
            sa = group_sum<G>(sa);
            sb = group_sum<G>(sb);
            if (gl == 0) {
                if (ja < k) s_out[(size_t)row * k + ja] = __ldg(s_in + (size_t)row * k + ja) + bias + sa;
                if (jb < k) s_out[(size_t)row * k + jb] = __ldg(s_in + (size_t)row * k + jb) + bias + sb;
            }
        }
    }
}

// row side of the backward: d_ps, and per-block partials of d_w2 / d_b2
template <int G, int NV>
__global__ void __launch_bounds__(256) diff_mlp_bwd_rows_kernel(
    const float* __restrict__ gr, const float* __restrict__ ps, const float* __restrict__ pt,
    const int* __restrict__ idx, const float* __restrict__ w2, float* __restrict__ d_ps,
    float* __restrict__ partial, long long m, int k, int r4) {
    constexpr int NG = 32 / G;
    extern __shared__ float red[];  // [warps][r + 1]
    const int lane = threadIdx.x & 31, gl = lane % G, g = lane / G, warp = threadIdx.x >> 5;
    const int r = r4 * 4;
    const long long warps = (long long)gridDim.x * (blockDim.x >> 5);
    float4 wv[NV], dw[NV];
#pragma unroll
    for (int v = 0; v < NV; ++v) {
        const int q = gl + v * G;
        wv[v] = q < r4 ? ldg4(w2 + 4 * q) : make_float4(0, 0, 0, 0);
        dw[v] = make_float4(0, 0, 0, 0);
    }
    float db = 0.f;
    for (long long row = (long long)blockIdx.x * (blockDim.x >> 5) + warp; row < m; row += warps) {
        float4 pv[NV], dp[NV];
#pragma unroll
        for (int v = 0; v < NV; ++v) {
            const int q = gl + v * G;
            pv[v] = q < r4 ? ldg4(ps + (size_t)row * r + 4 * q) : make_float4(0, 0, 0, 0);
            dp[v] = make_float4(0, 0, 0, 0);
        }
        for (int j = g; j < k; j += NG) {
            const float gj = __ldg(gr + (size_t)row * k + j);
            const int t = __ldg(idx + (size_t)row * k + j);
            if (gl == 0) db += gj;
#pragma unroll
            for (int v = 0; v < NV; ++v) {
                const int q = gl + v * G;
                if (q < r4) {
                    const float4 tv = ldg4(pt + (size_t)t * r + 4 * q);
                    const float dx = pv[v].x - tv.x, dy = pv[v].y - tv.y, dz = pv[v].z - tv.z,
                                dwv = pv[v].w - tv.w;
                    dp[v].x += dx > 0.f ? gj : 0.f; dw[v].x += gj * fmaxf(dx, 0.f);
                    dp[v].y += dy > 0.f ? gj : 0.f; dw[v].y += gj * fmaxf(dy, 0.f);
                    dp[v].z += dz > 0.f ? gj : 0.f; dw[v].z += gj * fmaxf(dz, 0.f);
                    dp[v].w += dwv > 0.f ? gj : 0.f; dw[v].w += gj * fmaxf(dwv, 0.f);
                }
            }
        }
        // fold the NG candidate groups of this warp together (lanes gl, gl+G, ... hold the same c)
#pragma unroll
        for (int v = 0; v < NV; ++v) {
#pragma unroll
            for (int o = G; o < 32; o <<= 1) {
                dp[v].x += __shfl_xor_sync(0xffffffffu, dp[v].x, o);
                dp[v].y += __shfl_xor_sync(0xffffffffu, dp[v].y, o);
                dp[v].z += __shfl_xor_sync(0xffffffffu, dp[v].z, o);
                dp[v].w += __shfl_xor_sync(0xffffffffu, dp[v].w, o);
            }
            const int q = gl + v * G;
            if (g == 0 && q < r4) {
                float4 o4 = make_float4(dp[v].x * wv[v].x, dp[v].y * wv[v].y, dp[v].z * wv[v].z,
                                        dp[v].w * wv[v].w);
                *reinterpret_cast<float4*>(d_ps + (size_t)row * r + 4 * q) = o4;
            }
        }
    }
    // block reduction of dw / db in a fixed order -> partial[blockIdx][0..r] (r = slot of d_b2)
#pragma unroll
    for (int v = 0; v < NV; ++v) {
#pragma unroll
        for (int o = G; o < 32; o <<= 1) {
            dw[v].x += __shfl_xor_sync(0xffffffffu, dw[v].x, o);
            dw[v].y += __shfl_xor_sync(0xffffffffu, dw[v].y, o);
            dw[v].z += __shfl_xor_sync(0xffffffffu, dw[v].z, o);
            dw[v].w += __shfl_xor_sync(0xffffffffu, dw[v].w, o);
        }
        const int q = gl + v * G;
        if (g == 0 && q < r4) {
            float* dst = red + warp * (r + 1) + 4 * q;
            dst[0] = dw[v].x; dst[1] = dw[v].y; dst[2] = dw[v].z; dst[3] = dw[v].w;
        }
    }
    db = warp_sum(db);
    if (lane == 0) red[warp * (r + 1) + r] = db;
    __syncthreads();
    for (int c = threadIdx.x; c <= r; c += blockDim.x) {
        float s = 0.f;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) s += red[w * (r + 1) + c];
        partial[(size_t)blockIdx.x * (r + 1) + c] = s;
    }
}

// one warp per output column: lanes stride over the per-block partials, fixed shuffle tree
__global__ void __launch_bounds__(256) diff_mlp_bwd_finish_kernel(const float* __restrict__ partial,
                                                                  int blocks, int r,
                                                                  float* __restrict__ d_w2,
                                                                  float* __restrict__ d_b2) {
    const int lane = threadIdx.x & 31;
    for (int c = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); c <= r; c += gridDim.x * (blockDim.x >> 5)) {
        float s = 0.f;
        for (int b = lane; b < blocks; b += 32) s += partial[(size_t)b * (r + 1) + c];
        s = warp_sum(s);
        if (lane == 0) { if (c < r) d_w2[c] = s; else *d_b2 = s; }
    }
}

// target side of the backward: one group of G lanes per target node, by-target CSR
template <int G, int NV>
__global__ void __launch_bounds__(256) diff_mlp_bwd_targets_kernel(
    const float* __restrict__ gr, const float* __restrict__ ps, const float* __restrict__ pt,
    const int* __restrict__ tptr, const int* __restrict__ perm, const float* __restrict__ w2,
    float* __restrict__ d_pt, long long n_t, int k, int r4) {
    const int gl = threadIdx.x % G;
    const int r = r4 * 4;
    const long long groups = (long long)gridDim.x * (blockDim.x / G);
    float4 wv[NV];
#pragma unroll
    for (int v = 0; v < NV; ++v) {
        const int q = gl + v * G;
        wv[v] = q < r4 ? ldg4(w2 + 4 * q) : make_float4(0, 0, 0, 0);
    }
    for (long long t = (long long)blockIdx.x * (blockDim.x / G) + threadIdx.x / G; t < n_t;
         t += groups) {
        float4 tv[NV], acc[NV];
#pragma unroll
        for (int v = 0; v < NV; ++v) {
            const int q = gl + v * G;
            tv[v] = q < r4 ? ldg4(pt + (size_t)t * r + 4 * q) : make_float4(0, 0, 0, 0);
            acc[v] = make_float4(0, 0, 0, 0);
        }
        const int e1 = __ldg(tptr + t + 1);
        for (int e = __ldg(tptr + t); e < e1; e += 2) {
            const bool two = e + 1 < e1;
            const int qa = __ldg(perm + e), qb = two ? __ldg(perm + e + 1) : qa;
            const float ga = __ldg(gr + qa), gb = two ? __ldg(gr + qb) : 0.f;
            const int ra = qa / k, rb = qb / k;
#pragma unroll
            for (int v = 0; v < NV; ++v) {
                const int q = gl + v * G;
                if (q < r4) {
                    const float4 pa = ldg4(ps + (size_t)ra * r + 4 * q);
                    const float4 pb = ldg4(ps + (size_t)rb * r + 4 * q);
                    acc[v].x += (pa.x > tv[v].x ? ga : 0.f) + (pb.x > tv[v].x ? gb : 0.f);
                    acc[v].y += (pa.y > tv[v].y ? ga : 0.f) + (pb.y > tv[v].y ? gb : 0.f);
                    acc[v].z += (pa.z > tv[v].z ? ga : 0.f) + (pb.z > tv[v].z ? gb : 0.f);
                    acc[v].w += (pa.w > tv[v].w ? ga : 0.f) + (pb.w > tv[v].w ? gb : 0.f);
                }
            }
        }
#pragma unroll
        for (int v = 0; v < NV; ++v) {
            const int q = gl + v * G;
            if (q < r4) {
                float4 o4 = make_float4(-wv[v].x * acc[v].x, -wv[v].y * acc[v].y,
                                        -wv[v].z * acc[v].z, -wv[v].w * acc[v].w);
                *reinterpret_cast<float4*>(d_pt + (size_t)t * r + 4 * q) = o4;
            }
        }
    }
}

}  // namespace dgmc

using namespace dgmc;

#define DM_DISPATCH(KERNEL, GRID, SMEM, ...)                                                   \
    do {                                                                                       \
        if (r4 <= 8) KERNEL<8, 1><<<GRID, 256, SMEM, stream>>>(__VA_ARGS__);                  \
        else if (r4 <= 16) KERNEL<16, 1><<<GRID, 256, SMEM, stream>>>(__VA_ARGS__);           \
        else if (r4 <= 32) KERNEL<32, 1><<<GRID, 256, SMEM, stream>>>(__VA_ARGS__);           \
        else if (r4 <= 64) KERNEL<32, 2><<<GRID, 256, SMEM, stream>>>(__VA_ARGS__);           \
        else KERNEL<32, 4><<<GRID, 256, SMEM, stream>>>(__VA_ARGS__);                         \
    } while (0)

static int check_r(const char* who, int r) {
    if (r < 4 || r % 4 != 0 || r > 512) {
        set_error("%s: r=%d unsupported (need a multiple of 4 in [4, 512]; the host pads)", who, r);
        return DGMC_ERR_UNSUPPORTED;
    }
    return DGMC_OK;
}

extern "C" int dgmc_diff_mlp_fwd(const float* s_in, const float* ps, const float* pt,
                                 const int32_t* idx, const float* w2, const float* b2, float* s_out,
                                 int64_t m, int k, int r, dgmc_stream_t stream) {
    if (m == 0 && k >= 1) return DGMC_OK;
    DGMC_REQUIRE(s_in && ps && pt && idx && w2 && b2 && s_out && m >= 0 && k >= 1,
                 "diff_mlp_fwd: bad arguments");
    if (int e = check_r("diff_mlp_fwd", r)) return e;
    if (m == 0) return DGMC_OK;
    const int r4 = r / 4;
    const int grid = wave_grid(m, 8, 8);
    DM_DISPATCH(diff_mlp_fwd_kernel, grid, 0, s_in, ps, pt, idx, w2, b2, s_out, m, k, r4);
    return check_launch("diff_mlp_fwd");
}

extern "C" size_t dgmc_diff_mlp_bwd_workspace(int64_t m, int r) {
    (void)m;
    return (size_t)DM_MAX_BLOCKS * (size_t)(r + 1) * sizeof(float);
}

extern "C" int dgmc_diff_mlp_bwd_rows(const float* g, const float* ps, const float* pt,
                                      const int32_t* idx, const float* w2, float* d_ps, float* d_w2,
                                      float* d_b2, int64_t m, int k, int r, void* workspace,
                                      size_t workspace_bytes, dgmc_stream_t stream) {
    DGMC_REQUIRE(g && ps && pt && idx && w2 && d_ps && d_w2 && d_b2 && workspace && m >= 0 && k >= 1,
                 "diff_mlp_bwd_rows: bad arguments");
    if (int e = check_r("diff_mlp_bwd_rows", r)) return e;
    if (workspace_bytes < dgmc_diff_mlp_bwd_workspace(m, r)) {
        set_error("diff_mlp_bwd_rows: workspace too small");
        return DGMC_ERR_WORKSPACE;
    }
    const int r4 = r / 4;
    int grid = wave_grid(m > 0 ? m : 1, 8, 4);
    if (grid > DM_MAX_BLOCKS) grid = DM_MAX_BLOCKS;
    const size_t smem = 8 * (size_t)(r + 1) * sizeof(float);
    float* partial = (float*)workspace;
    DM_DISPATCH(diff_mlp_bwd_rows_kernel, grid, smem, g, ps, pt, idx, w2, d_ps, partial, m, k, r4);
    diff_mlp_bwd_finish_kernel<<<(r + 8) / 8, 256, 0, stream>>>(partial, grid, r, d_w2, d_b2);
    return check_launch("diff_mlp_bwd_rows", 2);
}

extern "C" int dgmc_diff_mlp_bwd_targets(const float* g, const float* ps, const float* pt,
                                         const int32_t* tptr, const int32_t* perm, const float* w2,
                                         float* d_pt, int64_t n_t, int k, int r,
                                         dgmc_stream_t stream) {
    DGMC_REQUIRE(g && ps && pt && tptr && perm && w2 && d_pt && n_t >= 0 && k >= 1,
                 "diff_mlp_bwd_targets: bad arguments");
    if (int e = check_r("diff_mlp_bwd_targets", r)) return e;
    if (n_t == 0) return DGMC_OK;
    const int r4 = r / 4;
    const int G = r4 <= 8 ? 8 : (r4 <= 16 ? 16 : 32);
    const int grid = wave_grid(n_t, 256 / G, 8);
    DM_DISPATCH(diff_mlp_bwd_targets_kernel, grid, 0, g, ps, pt, tptr, perm, w2, d_pt, n_t, k, r4);
    return check_launch("diff_mlp_bwd_targets");
}

// ---------------------------------------------------------------------------------------------
// Allocation-free variant of dgmc_diff_mlp_bwd_rows.
//
// Some programs that bind this library cannot hand it a workspace: a C/C++ serving process that
// reserved the whole device for its own arena at start-up and gives plug-ins no allocator, or a
// latency-critical loop in which cudaMalloc / cudaFree (which serialise the device) are banned.
// For them the per-block partial sums of d_w2 / d_b2 live in a fixed array inside the library's
// device image instead of in caller memory.
//
//  * limit: r <= DGMC_NOALLOC_MAX_R (128).  The array is 1024 x 129 floats = 528,384 bytes and is
//    resident on EVERY GPU that loads the library until the process exits.
//  * every call on a device uses the same array, so two calls that overlap on the GPU (different
//    streams, or different host threads) race on it; serialise them.
//  * nothing in it has to start at zero: each launched block overwrites its own row before the
//    finishing kernel reads exactly those rows, so left-over contents of an earlier call are
//    never observed.
//  * the first call on each device looks the array's device address up (cudaGetSymbolAddress —
//    the symbol's name is only a host-side handle, not an address a kernel can use).  Make that
//    first call outside any stream capture.
// ---------------------------------------------------------------------------------------------
#include <mutex>

#define DGMC_NOALLOC_MAX_R 128
#define DGMC_NOALLOC_MAX_DEVICES 64

namespace dgmc {
static __device__ __align__(256) float g_noalloc_partials[DM_MAX_BLOCKS * (DGMC_NOALLOC_MAX_R + 1)];

// device address of g_noalloc_partials on the CURRENT device (looked up once per device)
static int noalloc_scratch(float** out) {
    static std::mutex mu;
    static float* cached[DGMC_NOALLOC_MAX_DEVICES] = {nullptr};
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= DGMC_NOALLOC_MAX_DEVICES) {
        set_error("diff_mlp_bwd_rows_noalloc: cudaGetDevice failed or device %d out of range: %s", dev,
                  cudaGetErrorString(e));
        return DGMC_ERR_CUDA;
    }
    std::lock_guard<std::mutex> lock(mu);
    if (!cached[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_noalloc_partials);
        if (e != cudaSuccess || !p) {
            set_error("diff_mlp_bwd_rows_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev,
                      cudaGetErrorString(e));
            return DGMC_ERR_CUDA;
        }
        cached[dev] = static_cast<float*>(p);
    }
    *out = cached[dev];
    return DGMC_OK;
}
}  // namespace dgmc

extern "C" int dgmc_diff_mlp_bwd_rows_noalloc(const float* g, const float* ps, const float* pt,
                                              const int32_t* idx, const float* w2, float* d_ps,
                                              float* d_w2, float* d_b2, int64_t m, int k, int r,
                                              dgmc_stream_t stream) {
    if (r > DGMC_NOALLOC_MAX_R) {
        set_error("diff_mlp_bwd_rows_noalloc: r=%d exceeds the built-in scratch (max r=%d); use "
                  "dgmc_diff_mlp_bwd_rows with a caller-provided workspace", r, DGMC_NOALLOC_MAX_R);
        return DGMC_ERR_UNSUPPORTED;
    }
    float* scratch = nullptr;
    if (int e = noalloc_scratch(&scratch)) return e;
    return dgmc_diff_mlp_bwd_rows(g, ps, pt, idx, w2, d_ps, d_w2, d_b2, m, k, r, scratch,
                                  sizeof(float) * DM_MAX_BLOCKS * (DGMC_NOALLOC_MAX_R + 1), stream);
}
