// Weight-gradient GEMM for the per-node linear maps of psi_2 / the consensus MLP:
//     dW[m, n] = sum_k G[k, m] * X[k, n]        db[m] = sum_k G[k, m]
// with K = number of nodes (tens of thousands to millions) and a tiny output (M = out_channels,
// N = in_channels: 64 x 192 at the DBP15K shape).  A library SGEMM tiles only the OUTPUT, so it runs
// this on 3 CTAs for 116 us (ncu, profiles/README.md r1e); the work is a long reduction and has to be
// split along K instead.  Grid = (N tiles) x (M tiles) x (K splits) sized to ~2 waves of the SMs;
// each CTA streams its K slice through shared memory (128-bit loads), keeps a 64 x 64 tile in
// registers (8 x 4 per thread, fp32 FFMA), writes its partial tile to the caller's workspace, and a
// second kernel adds the partials in split order (deterministic; no atomics).
// FP32 FFMA throughout.  Since the tensor-core kernel (wgrad_tc.cu: same partial layout, same
// reduction) this partial kernel only serves operands TMA cannot read (base or pitch not 16-byte
// aligned); dgmc_wgrad_reduce below is shared by both.
#include "common.cuh"
#include "../../include/dgmc_b200.h"

namespace dgmc {

constexpr int WG_TM = 64, WG_TN = 64, WG_KC = 32, WG_THREADS = 128;

__global__ void __launch_bounds__(WG_THREADS) wgrad_partial_kernel(
    const float* __restrict__ g, long long ldg_, const float* __restrict__ x, long long ldx, long long K,
    int M, int N, int splits, float* __restrict__ partial, float* __restrict__ partial_b) {
    __shared__ __align__(16) float sg[WG_KC][WG_TM];
    __shared__ __align__(16) float sx[WG_KC][WG_TN];
    const int tid = threadIdx.x;
    const int n0 = blockIdx.x * WG_TN, m0 = blockIdx.y * WG_TM, split = blockIdx.z;
    const long long per = (K + splits - 1) / splits;
    const long long k_begin = per * split, k_end = k_begin + per < K ? k_begin + per : K;
    const int tm = (tid / 16) * 8;   // 8 rows of the tile (m)
    const int tn = (tid % 16) * 4;   // 4 columns (n)
    float acc[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
    float bsum = 0.f;  // thread `tid` < 64 owns column m0 + tid of db (only when blockIdx.x == 0)
    const bool g_vec = (ldg_ % 4 == 0) && ((reinterpret_cast<uintptr_t>(g) & 15) == 0) && (m0 + WG_TM <= M);
    const bool x_vec = (ldx % 4 == 0) && ((reinterpret_cast<uintptr_t>(x) & 15) == 0) && (n0 + WG_TN <= N);
    // software pipeline: the next [KC x 64] slabs of G and X travel global -> registers while the
    // current ones are consumed from shared memory
    constexpr int PF = WG_KC * 16 / WG_THREADS;  // float4 per thread and matrix
    float4 pg[PF], px[PF];
    auto fetch = [&](long long k0) {
#pragma unroll
        for (int p = 0; p < PF; ++p) {
            const int t = tid + p * WG_THREADS;
            const int kk = t / 16, c = (t % 16) * 4;
            const long long k = k0 + kk;
            float4 vg = make_float4(0, 0, 0, 0), vx = make_float4(0, 0, 0, 0);
            if (k < k_end) {
                if (g_vec) vg = ldg4(g + k * ldg_ + m0 + c);
                else {
                    float* q = &vg.x;
                    for (int u = 0; u < 4; ++u) if (m0 + c + u < M) q[u] = __ldg(g + k * ldg_ + m0 + c + u);
                }
                if (x_vec) vx = ldg4(x + k * ldx + n0 + c);
                else {
                    float* q = &vx.x;
                    for (int u = 0; u < 4; ++u) if (n0 + c + u < N) q[u] = __ldg(x + k * ldx + n0 + c + u);
                }
            }
            pg[p] = vg; px[p] = vx;
        }
    };
    if (k_begin < k_end) fetch(k_begin);
    for (long long k0 = k_begin; k0 < k_end; k0 += WG_KC) {
#pragma unroll
        for (int p = 0; p < PF; ++p) {
            const int t = tid + p * WG_THREADS;
            const int kk = t / 16, c = (t % 16) * 4;
            *reinterpret_cast<float4*>(&sg[kk][c]) = pg[p];
            *reinterpret_cast<float4*>(&sx[kk][c]) = px[p];
        }
        __syncthreads();
        if (k0 + WG_KC < k_end) fetch(k0 + WG_KC);
#pragma unroll 8
        for (int kk = 0; kk < WG_KC; ++kk) {
            const float4 a0 = *reinterpret_cast<const float4*>(&sg[kk][tm]);
            const float4 a1 = *reinterpret_cast<const float4*>(&sg[kk][tm + 4]);
            const float4 b = *reinterpret_cast<const float4*>(&sx[kk][tn]);
            const float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            const float bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
        }
        if (blockIdx.x == 0 && tid < WG_TM) {
#pragma unroll 8
            for (int kk = 0; kk < WG_KC; ++kk) bsum += sg[kk][tid];
        }
        __syncthreads();
    }
    float* out = partial + (size_t)split * M * N;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int m = m0 + tm + i;
        if (m < M) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int n = n0 + tn + j;
                if (n < N) out[(size_t)m * N + n] = acc[i][j];
            }
        }
    }
    if (partial_b && blockIdx.x == 0 && tid < WG_TM && m0 + tid < M)
        partial_b[(size_t)split * M + m0 + tid] = bsum;
}

// one warp per output element: lanes stride over the partial tiles (coalesced across the warps of a
// block, which own consecutive elements), fixed shuffle tree -> deterministic
__global__ void __launch_bounds__(256) wgrad_reduce_kernel(const float* __restrict__ partial,
                                                           const float* __restrict__ partial_b, int splits,
                                                           int MN, int M, float* __restrict__ dw,
                                                           float* __restrict__ db) {
    const int total = MN + (db ? M : 0);
    const int lane = threadIdx.x & 31;
    // block = 8 warps x 32 lanes handles 32 consecutive elements: lane l of warp w reads element
    // (base + l) of partial tiles w, w + 8, ...  -> 128-byte coalesced rows; then a shared-memory
    // transpose-free reduction over the 8 warps.
    __shared__ float red[8][33];
    for (int base = blockIdx.x * 32; base < total; base += gridDim.x * 32) {
        const int i = base + lane;
        const int w = threadIdx.x >> 5;
        float s = 0.f;
        if (i < total) {
            if (i < MN) for (int p = w; p < splits; p += 8) s += partial[(size_t)p * MN + i];
            else        for (int p = w; p < splits; p += 8) s += partial_b[(size_t)p * M + (i - MN)];
        }
        red[w][lane] = s;
        __syncthreads();
        if (w == 0 && i < total) {
            float t = 0.f;
#pragma unroll
            for (int q = 0; q < 8; ++q) t += red[q][lane];
            if (i < MN) dw[i] = t; else db[i - MN] = t;
        }
        __syncthreads();
    }
}

static int wgrad_splits(long long K, int M, int N) {
    const int tiles = ((N + WG_TN - 1) / WG_TN) * ((M + WG_TM - 1) / WG_TM);
    int splits = (4 * sm_count() + tiles - 1) / tiles;   // ~4 CTAs of 128 threads per SM
    const long long max_by_k = (K + 4 * WG_KC - 1) / (4 * WG_KC);  // at least 128 rows per split
    if (splits > max_by_k) splits = (int)max_by_k;
    if (splits < 1) splits = 1;
    if (splits > 1024) splits = 1024;
    return splits;
}

}  // namespace dgmc

using namespace dgmc;

// ---- two-stage interface: several partial launches (e.g. one per consensus step, all for the same
// weight) can share ONE reduction -------------------------------------------------------------------
extern "C" int dgmc_wgrad_splits(int64_t k, int m, int n) { return wgrad_splits(k, m, n); }

extern "C" int dgmc_wgrad_partial(const float* g, int64_t ldg, const float* x, int64_t ldx, int64_t k, int m,
                                  int n, float* partial, float* partial_b, dgmc_stream_t stream) {
    DGMC_REQUIRE(g && x && partial && k >= 0 && m >= 1 && n >= 1 && ldg >= m && ldx >= n,
                 "wgrad_partial: bad arguments (k=%lld m=%d n=%d)", (long long)k, m, n);
    const int splits = wgrad_splits(k, m, n);
    dim3 grid((n + WG_TN - 1) / WG_TN, (m + WG_TM - 1) / WG_TM, splits);
    wgrad_partial_kernel<<<grid, WG_THREADS, 0, stream>>>(g, ldg, x, ldx, k, m, n, splits, partial, partial_b);
    return check_launch("wgrad_partial");
}

extern "C" int dgmc_wgrad_reduce(const float* partial, const float* partial_b, int n_partials, int m, int n,
                                 float* dw, float* db, dgmc_stream_t stream) {
    DGMC_REQUIRE(partial && dw && n_partials >= 1 && m >= 1 && n >= 1 && (!db || partial_b),
                 "wgrad_reduce: bad arguments");
    const int total = m * n + (db ? m : 0);
    wgrad_reduce_kernel<<<(total + 31) / 32, 256, 0, stream>>>(partial, partial_b, n_partials, m * n, m, dw, db);
    return check_launch("wgrad_reduce");
}

extern "C" size_t dgmc_wgrad_workspace(int64_t k, int m, int n) {
    const int splits = wgrad_splits(k, m, n);
    return (size_t)splits * ((size_t)m * n + m) * sizeof(float) + 256;
}

extern "C" int dgmc_wgrad(const float* g, int64_t ldg, const float* x, int64_t ldx, int64_t k, int m, int n,
                          float* dw, float* db, void* workspace, size_t workspace_bytes,
                          dgmc_stream_t stream) {
    DGMC_REQUIRE(g && x && dw && workspace && k >= 0 && m >= 1 && n >= 1 && ldg >= m && ldx >= n,
                 "wgrad: bad arguments (k=%lld m=%d n=%d)", (long long)k, m, n);
    if (workspace_bytes < dgmc_wgrad_workspace(k, m, n)) {
        set_error("wgrad: workspace too small");
        return DGMC_ERR_WORKSPACE;
    }
    const int splits = wgrad_splits(k, m, n);
    float* partial = static_cast<float*>(workspace);
    float* partial_b = partial + (size_t)splits * m * n;
    dim3 grid((n + WG_TN - 1) / WG_TN, (m + WG_TM - 1) / WG_TM, splits);
    wgrad_partial_kernel<<<grid, WG_THREADS, 0, stream>>>(g, ldg, x, ldx, k, m, n, splits, partial,
                                                         db ? partial_b : nullptr);
    const int total = m * n + (db ? m : 0);
    wgrad_reduce_kernel<<<(total + 31) / 32, 256, 0, stream>>>(partial, partial_b, splits, m * n, m, dw, db);
    return check_launch("wgrad", 2);
}
