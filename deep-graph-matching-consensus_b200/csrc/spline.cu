// K5 (SplineConv): degree-1 open B-spline continuous convolution of PyG SplineConv(kernel_size=5)
// as the reference builds it (spline.py:21, [3p] torch_spline_conv).
//
//   message_e = sum_{s < 2^dim} basis[e, s] * x[src_e] @ W[widx[e, s]]        out_i = mean_{e -> i}
//
// "Transform first": the host computes Y[n, kk, :] = x[n] @ W[kk] for all K^dim kernels with ONE
// library GEMM ([N, in] x [in, K*out]); the kernels here do the data-dependent part:
//   spline_basis        pseudo -> (basis, widx), once per edge_attr (cached by the host)
//   spline_gather_fwd   out[i] (+)= inv_deg[i] * sum_{e -> i} sum_s basis * Y[src_e, widx, :]
//   spline_gather_bwd   dY[j, kk, :] = sum_{e: src_e = j} sum_{s: widx = kk} basis * inv_deg[dst_e] * g[dst_e]
// Both are CSR segmented gathers in a fixed order (deterministic, no atomics).
#include "common.cuh"
#include "../../include/dgmc_b200.h"

namespace dgmc {

__global__ void __launch_bounds__(256) spline_basis_kernel(const float* __restrict__ pseudo,
                                                           long long E, int dim, int ksize,
                                                           float* __restrict__ basis,
                                                           int* __restrict__ widx) {
    const int S = 1 << dim;
    const long long total = E * S;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += stride) {
        const long long e = t / S;
        const int s = (int)(t % S);
        float b = 1.f;
        int w = 0, off = 1;
        for (int d = 0; d < dim; ++d) {
            const float v = __ldg(pseudo + e * dim + d) * (float)(ksize - 1);  // open spline, degree 1
            const float fl = floorf(v);
            const float frac = v - fl;
            const int bit = (s >> d) & 1;
            b *= bit ? frac : 1.f - frac;
            int cell = ((int)fl + bit) % ksize;
            if (cell < 0) cell += ksize;
            w += cell * off;
            off *= ksize;
        }
        basis[t] = b;
        widx[t] = w;
    }
}

template <int G, int NV>
__global__ void __launch_bounds__(256) spline_gather_fwd_kernel(
    const float* __restrict__ y, const int* __restrict__ ptr, const int* __restrict__ nbr,
    const int* __restrict__ eid, const float* __restrict__ basis, const int* __restrict__ widx, int S,
    const float* __restrict__ inv_deg, float* __restrict__ out, long long ldo, long long n, int K,
    int c4, int accumulate) {
    const int gl = threadIdx.x % G;
    const long long groups = (long long)gridDim.x * (blockDim.x / G);
    const int c = c4 * 4;
    for (long long i = (long long)blockIdx.x * (blockDim.x / G) + threadIdx.x / G; i < n; i += groups) {
        float4 acc[NV];
#pragma unroll
        for (int v = 0; v < NV; ++v) acc[v] = make_float4(0, 0, 0, 0);
        const int e1 = __ldg(ptr + i + 1);
        for (int e = __ldg(ptr + i); e < e1; ++e) {
            const long long j = __ldg(nbr + e);
            const long long ed = __ldg(eid + e);
            for (int s = 0; s < S; ++s) {
                const float b = __ldg(basis + ed * S + s);
                const int w = __ldg(widx + ed * S + s);
                const float* yr = y + ((size_t)j * K + w) * c;
#pragma unroll
                for (int v = 0; v < NV; ++v) {
                    const int q = gl + v * G;
                    if (q < c4) {
                        const float4 t = ldg4(yr + 4 * q);
                        acc[v].x += b * t.x; acc[v].y += b * t.y; acc[v].z += b * t.z; acc[v].w += b * t.w;
                    }
                }
            }
        }
        const float sc = __ldg(inv_deg + i);
#pragma unroll
        for (int v = 0; v < NV; ++v) {
            const int q = gl + v * G;
            if (q < c4) {
                float4* dst = reinterpret_cast<float4*>(out + (size_t)i * ldo + 4 * q);
                float4 o = make_float4(acc[v].x * sc, acc[v].y * sc, acc[v].z * sc, acc[v].w * sc);
                if (accumulate) { const float4 p = *dst; o.x += p.x; o.y += p.y; o.z += p.z; o.w += p.w; }
                *dst = o;
            }
        }
    }
}

// one group per SOURCE node j: owns dY[j, :, :], zeroes it, then adds edge by edge (fixed order)
template <int G, int NV>
__global__ void __launch_bounds__(256) spline_gather_bwd_kernel(
    const float* __restrict__ g, long long ldg_, const int* __restrict__ ptr,
    const int* __restrict__ nbr, const int* __restrict__ eid, const float* __restrict__ basis,
    const int* __restrict__ widx, int S, const float* __restrict__ inv_deg, float* __restrict__ dy,
    long long n, int K, int c4) {
    const int gl = threadIdx.x % G;
    const long long groups = (long long)gridDim.x * (blockDim.x / G);
    const int c = c4 * 4;
    for (long long j = (long long)blockIdx.x * (blockDim.x / G) + threadIdx.x / G; j < n; j += groups) {
        float* block = dy + (size_t)j * K * c;
        for (int kk = 0; kk < K; ++kk)
#pragma unroll
            for (int v = 0; v < NV; ++v) {
                const int q = gl + v * G;
                if (q < c4) *reinterpret_cast<float4*>(block + (size_t)kk * c + 4 * q) = make_float4(0, 0, 0, 0);
            }
        const int e1 = __ldg(ptr + j + 1);
        for (int e = __ldg(ptr + j); e < e1; ++e) {
            const long long i = __ldg(nbr + e);  // destination node of this out-edge
            const long long ed = __ldg(eid + e);
            const float sc = __ldg(inv_deg + i);
            float4 gv[NV];
#pragma unroll
            for (int v = 0; v < NV; ++v) {
                const int q = gl + v * G;
                gv[v] = q < c4 ? ldg4(g + (size_t)i * ldg_ + 4 * q) : make_float4(0, 0, 0, 0);
            }
            for (int s = 0; s < S; ++s) {
                const float b = __ldg(basis + ed * S + s) * sc;
                const int w = __ldg(widx + ed * S + s);
#pragma unroll
                for (int v = 0; v < NV; ++v) {
                    const int q = gl + v * G;
                    if (q < c4) {
                        float4* dst = reinterpret_cast<float4*>(block + (size_t)w * c + 4 * q);
                        float4 p = *dst;  // same lane wrote it: program order suffices
                        p.x += b * gv[v].x; p.y += b * gv[v].y; p.z += b * gv[v].z; p.w += b * gv[v].w;
                        *dst = p;
                    }
                }
            }
        }
    }
}

}  // namespace dgmc

using namespace dgmc;

extern "C" int dgmc_spline_basis(const float* pseudo, int64_t n_edges, int dim, int kernel_size,
                                 float* basis, int32_t* widx, dgmc_stream_t stream) {
    DGMC_REQUIRE(n_edges >= 0 && dim >= 1 && dim <= 5 && kernel_size >= 2,
                 "spline_basis: bad arguments (dim=%d kernel_size=%d)", dim, kernel_size);
    if (n_edges == 0) return DGMC_OK;
    DGMC_REQUIRE(pseudo && basis && widx, "spline_basis: null pointer");
    spline_basis_kernel<<<wave_grid(n_edges << dim, 256, 8), 256, 0, stream>>>(
        pseudo, n_edges, dim, kernel_size, basis, widx);
    return check_launch("spline_basis");
}

#define SP_DISPATCH(KERNEL, ...)                                                               \
    do {                                                                                       \
        if (c4 <= 8) KERNEL<8, 1><<<wave_grid(n, 32, 8), 256, 0, stream>>>(__VA_ARGS__);      \
        else if (c4 <= 16) KERNEL<16, 1><<<wave_grid(n, 16, 8), 256, 0, stream>>>(__VA_ARGS__); \
        else if (c4 <= 32) KERNEL<32, 1><<<wave_grid(n, 8, 8), 256, 0, stream>>>(__VA_ARGS__); \
        else if (c4 <= 64) KERNEL<32, 2><<<wave_grid(n, 8, 8), 256, 0, stream>>>(__VA_ARGS__); \
        else KERNEL<32, 4><<<wave_grid(n, 8, 8), 256, 0, stream>>>(__VA_ARGS__);               \
    } while (0)

static int sp_check_c(const char* who, int c) {
    if (c < 4 || c % 4 != 0 || c > 512) {
        set_error("%s: out_channels=%d unsupported (multiple of 4 in [4, 512])", who, c);
        return DGMC_ERR_UNSUPPORTED;
    }
    return DGMC_OK;
}

extern "C" int dgmc_spline_gather_fwd(const float* y, const int32_t* in_ptr, const int32_t* in_nbr,
                                      const int32_t* in_edge, const float* basis,
                                      const int32_t* widx, int n_corners, const float* inv_deg,
                                      float* out, int64_t ldo, int64_t n, int n_kernels, int c,
                                      int accumulate, dgmc_stream_t stream) {
    DGMC_REQUIRE(y && in_ptr && inv_deg && out && n >= 0 && n_kernels >= 1 && n_corners >= 1 &&
                     ldo >= c && ldo % 4 == 0, "spline_gather_fwd: bad arguments");
    if (int e = sp_check_c("spline_gather_fwd", c)) return e;
    if (n == 0) return DGMC_OK;
    const int c4 = c / 4;
    SP_DISPATCH(spline_gather_fwd_kernel, y, in_ptr, in_nbr, in_edge, basis, widx, n_corners,
                inv_deg, out, ldo, n, n_kernels, c4, accumulate);
    return check_launch("spline_gather_fwd");
}

extern "C" int dgmc_spline_gather_bwd(const float* g, int64_t ldg, const int32_t* out_ptr,
                                      const int32_t* out_nbr, const int32_t* out_edge,
                                      const float* basis, const int32_t* widx, int n_corners,
                                      const float* inv_deg, float* dy, int64_t n, int n_kernels, int c,
                                      dgmc_stream_t stream) {
    DGMC_REQUIRE(g && out_ptr && inv_deg && dy && n >= 0 && n_kernels >= 1 && n_corners >= 1 &&
                     ldg >= c && ldg % 4 == 0, "spline_gather_bwd: bad arguments");
    if (int e = sp_check_c("spline_gather_bwd", c)) return e;
    if (n == 0) return DGMC_OK;
    const int c4 = c / 4;
    SP_DISPATCH(spline_gather_bwd_kernel, g, ldg, out_ptr, out_nbr, out_edge, basis, widx, n_corners,
                inv_deg, dy, n, n_kernels, c4);
    return check_launch("spline_gather_bwd");
}
