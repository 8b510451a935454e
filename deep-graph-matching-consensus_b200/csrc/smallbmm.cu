// Batched products of SMALL per-graph matrices in the dense branch (dgmc.py:163-183): the indicator
// transport r_t = S^T r_s (dgmc.py:170), its adjoints dS = r_s g^T and dr_s = S g, and the adjoints of
// the score product S_hat = h_s h_t^T (dh_s = dS h_t, dh_t = dS^T h_s).  Per graph these are
// [<= 128 x <= 128] x [<= 128 x C] products (PascalVOC: 20 x 20 x 256), B = 512 of them: a few hundred
// MFLOP in total and a few MB of traffic, i.e. bound by launch latency, not by any pipe — a
// tensor-core tile would be > 90 % padding.  Exact fp32 FFMA, one CTA per (graph, 128-column strip),
// deterministic (no atomics).  The reference runs them as ATen bmm (cuBLAS).
//
//   nn:  out[b] [M, N] = op(A[b]) [M, K] * Bm[b] [K, N]      op = identity (A[b] is [M, K]) or
//                                                            transpose (A[b] is [K, M])
//   nt:  out[b] [M, N] = A[b] [M, K] * Bm[b]^T               (Bm[b] is [N, K]): rows against rows
#include "common.cuh"
#include "../../include/dgmc_b200.h"

namespace dgmc {

constexpr int SB_THREADS = 256;
constexpr int SB_STRIP = 128;      // output columns per CTA (32 float4 lanes)

// A' is staged in shared memory as [M][K + 1] (the +1 breaks bank conflicts of the transposed load);
// thread (mg = t / 32, n4 = t % 32) accumulates rows mg, mg + 8, ... of the strip's float4 column n4.
__global__ void __launch_bounds__(SB_THREADS) small_bmm_nn_kernel(const float* __restrict__ A, int trans_a,
                                                                 const float* __restrict__ Bm,
                                                                 float* __restrict__ out, int M, int K, int N) {
    extern __shared__ float a_s[];
    const int b = blockIdx.x, n0 = blockIdx.y * SB_STRIP;
    const float* Ab = A + (size_t)b * M * K;
    const float* Bb = Bm + (size_t)b * K * N;
    float* ob = out + (size_t)b * M * N;
    const int ldk = K + 1;
    for (int i = threadIdx.x; i < M * K; i += SB_THREADS) {
        // consecutive threads read consecutive global elements in either layout
        const int m = trans_a ? i % M : i / K, k = trans_a ? i / M : i % K;
        a_s[m * ldk + k] = __ldg(Ab + i);
    }
    __syncthreads();
    const int n4 = threadIdx.x & 31, mg = threadIdx.x >> 5;
    const int n = n0 + 4 * n4;
    if (n >= N) return;
    const bool vec = (N % 4 == 0) && (n + 3 < N) && ((reinterpret_cast<uintptr_t>(Bm) | reinterpret_cast<uintptr_t>(out)) % 16 == 0);
    for (int m = mg; m < M; m += SB_THREADS / 32) {
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        const float* am = a_s + m * ldk;
        if (vec) {
#pragma unroll 4
            for (int k = 0; k < K; ++k) {
                const float a = am[k];
                const float4 v = ldg4(Bb + (size_t)k * N + n);
                acc.x = fmaf(a, v.x, acc.x); acc.y = fmaf(a, v.y, acc.y);
                acc.z = fmaf(a, v.z, acc.z); acc.w = fmaf(a, v.w, acc.w);
            }
            *reinterpret_cast<float4*>(ob + (size_t)m * N + n) = acc;
        } else {
            float r[4] = {0.f, 0.f, 0.f, 0.f};
            for (int k = 0; k < K; ++k) {
                const float a = am[k];
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    if (n + u < N) r[u] = fmaf(a, __ldg(Bb + (size_t)k * N + n + u), r[u]);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (n + u < N) ob[(size_t)m * N + n + u] = r[u];
        }
    }
}

// one warp per output row m: lanes stride over K for every n, warp reduction
__global__ void __launch_bounds__(SB_THREADS) small_bmm_nt_kernel(const float* __restrict__ A,
                                                                 const float* __restrict__ Bm,
                                                                 float* __restrict__ out, int M, int N, int K) {
    const int b = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const float* Ab = A + (size_t)b * M * K;
    const float* Bb = Bm + (size_t)b * N * K;
    float* ob = out + (size_t)b * M * N;
    for (int m = blockIdx.y * (SB_THREADS / 32) + warp; m < M; m += gridDim.y * (SB_THREADS / 32)) {
        for (int n = 0; n < N; ++n) {
            float s = 0.f;
            for (int k = lane; k < K; k += 32) s = fmaf(__ldg(Ab + (size_t)m * K + k), __ldg(Bb + (size_t)n * K + k), s);
            s = warp_sum(s);
            if (lane == 0) ob[(size_t)m * N + n] = s;
        }
    }
}

}  // namespace dgmc

using namespace dgmc;

extern "C" int dgmc_small_bmm_nn(const float* a, int trans_a, const float* bm, float* out, int batch, int m,
                                 int k, int n, dgmc_stream_t stream) {
    DGMC_REQUIRE(a && bm && out && batch >= 0 && m >= 1 && k >= 1 && n >= 1, "small_bmm_nn: bad arguments");
    const size_t smem = (size_t)m * (k + 1) * sizeof(float);
    if (m > 128 || k > 128) {
        set_error("small_bmm_nn: per-graph matrices up to 128 x 128 (got %d x %d)", m, k);
        return DGMC_ERR_UNSUPPORTED;
    }
    if (batch == 0) return DGMC_OK;
    static DeviceOnce configured;
    if (!configured.done()) {
        if (cudaFuncSetAttribute(small_bmm_nn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 128 * 129 * (int)sizeof(float)) != cudaSuccess) {
            set_error("small_bmm_nn: cannot set the dynamic shared memory size");
            return DGMC_ERR_CUDA;
        }
        configured.mark();
    }
    dim3 grid(batch, (n + SB_STRIP - 1) / SB_STRIP);
    small_bmm_nn_kernel<<<grid, SB_THREADS, smem, stream>>>(a, trans_a, bm, out, m, k, n);
    return check_launch("small_bmm_nn");
}

extern "C" int dgmc_small_bmm_nt(const float* a, const float* bm, float* out, int batch, int m, int n, int k,
                                 dgmc_stream_t stream) {
    DGMC_REQUIRE(a && bm && out && batch >= 0 && m >= 1 && k >= 1 && n >= 1, "small_bmm_nt: bad arguments");
    if (batch == 0) return DGMC_OK;
    const int rows_per_cta = SB_THREADS / 32;
    int gy = (m + rows_per_cta - 1) / rows_per_cta;
    if (gy > 16) gy = 16;
    dim3 grid(batch, gy);
    small_bmm_nt_kernel<<<grid, SB_THREADS, 0, stream>>>(a, bm, out, m, n, k);
    return check_launch("small_bmm_nt");
}
