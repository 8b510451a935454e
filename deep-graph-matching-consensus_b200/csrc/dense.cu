// Dense-branch consensus update (dgmc.py:178-179) with the first MLP layer hoisted to the nodes:
//   s_out[b,i,j] = s_in[b,i,j] + mask[b,i,j] * (b2 + sum_c w2[c] * relu(ps[b,i,c] - pt[b,j,c]))
// D [B, N_s, N_t, R] and mlp(D) are never materialised.  Same lane mapping as K6: one warp per
// (b, i) row, groups of G lanes per target column, float4 channel vectors.  pt[b] is small
// (N_t x R floats) and stays in L1/L2 across the rows of a graph.
#include "common.cuh"
#include "../../include/dgmc_b200.h"

namespace dgmc {

constexpr int DC_MAX_BLOCKS = 1024;

template <int G, int NV>
__global__ void __launch_bounds__(256) dense_consensus_fwd_kernel(
    const float* __restrict__ s_in, const float* __restrict__ ps, const float* __restrict__ pt,
    const float* __restrict__ w2, const float* __restrict__ b2, const uint8_t* __restrict__ sm,
    const uint8_t* __restrict__ tm, float* __restrict__ s_out, int B, int n_s, int n_t, int r4) {
    constexpr int NG = 32 / G;
    const int lane = threadIdx.x & 31, gl = lane % G, g = lane / G;
    const int r = r4 * 4;
    const long long rows = (long long)B * n_s;
    const long long warps = (long long)gridDim.x * (blockDim.x >> 5);
    float4 wv[NV];
#pragma unroll
    for (int v = 0; v < NV; ++v) {
        const int q = gl + v * G;
        wv[v] = q < r4 ? ldg4(w2 + 4 * q) : make_float4(0, 0, 0, 0);
    }
    const float bias = __ldg(b2);
    for (long long row = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); row < rows;
         row += warps) {
        const int b = (int)(row / n_s);
        const bool row_on = sm[row] != 0;
        float4 pv[NV];
#pragma unroll
        for (int v = 0; v < NV; ++v) {
            const int q = gl + v * G;
            pv[v] = q < r4 ? ldg4(ps + (size_t)row * r + 4 * q) : make_float4(0, 0, 0, 0);
        }
        const float* ptb = pt + (size_t)b * n_t * r;
        for (int j0 = 0; j0 < n_t; j0 += NG) {
            const int j = j0 + g;
            const bool on = row_on && j < n_t && tm[(size_t)b * n_t + j] != 0;
            float s = 0.f;
            if (on) {
#pragma unroll
                for (int v = 0; v < NV; ++v) {
                    const int q = gl + v * G;
                    if (q < r4) {
                        const float4 t = ldg4(ptb + (size_t)j * r + 4 * q);
                        s += wv[v].x * fmaxf(pv[v].x - t.x, 0.f) + wv[v].y * fmaxf(pv[v].y - t.y, 0.f) +
                             wv[v].z * fmaxf(pv[v].z - t.z, 0.f) + wv[v].w * fmaxf(pv[v].w - t.w, 0.f);
                    }
                }
            }
            s = group_sum<G>(s);
            if (gl == 0 && j < n_t) {
                const size_t o = (size_t)row * n_t + j;
                s_out[o] = __ldg(s_in + o) + (on ? bias + s : 0.f);
            }
        }
    }
}

// side == 0: rows (b,i) of ps against all j   -> d_ps, partial d_w2 / d_b2
// side == 1: rows (b,j) of pt against all i   -> d_pt
template <int G, int NV, int SIDE>
__global__ void __launch_bounds__(256) dense_consensus_bwd_kernel(
    const float* __restrict__ gr, const float* __restrict__ ps, const float* __restrict__ pt,
    const float* __restrict__ w2, const uint8_t* __restrict__ sm, const uint8_t* __restrict__ tm,
    float* __restrict__ d_own, float* __restrict__ partial, int B, int n_s, int n_t, int r4) {
    constexpr int NG = 32 / G;
    extern __shared__ float red[];
    const int lane = threadIdx.x & 31, gl = lane % G, g = lane / G, warp = threadIdx.x >> 5;
    const int r = r4 * 4;
    const int n_own = SIDE == 0 ? n_s : n_t, n_other = SIDE == 0 ? n_t : n_s;
    const float* own = SIDE == 0 ? ps : pt;
    const float* other = SIDE == 0 ? pt : ps;
    const uint8_t* m_own = SIDE == 0 ? sm : tm;
    const uint8_t* m_other = SIDE == 0 ? tm : sm;
    const long long rows = (long long)B * n_own;
    const long long warps = (long long)gridDim.x * (blockDim.x >> 5);
    float4 wv[NV], dw[NV];
#pragma unroll
    for (int v = 0; v < NV; ++v) {
        const int q = gl + v * G;
        wv[v] = q < r4 ? ldg4(w2 + 4 * q) : make_float4(0, 0, 0, 0);
        dw[v] = make_float4(0, 0, 0, 0);
    }
    float db = 0.f;
    for (long long row = (long long)blockIdx.x * (blockDim.x >> 5) + warp; row < rows; row += warps) {
        const int b = (int)(row / n_own), i_own = (int)(row % n_own);
        const bool row_on = m_own[row] != 0;
        float4 pv[NV], dp[NV];
#pragma unroll
        for (int v = 0; v < NV; ++v) {
            const int q = gl + v * G;
            pv[v] = q < r4 ? ldg4(own + (size_t)row * r + 4 * q) : make_float4(0, 0, 0, 0);
            dp[v] = make_float4(0, 0, 0, 0);
        }
        if (row_on) {
            for (int o = g; o < n_other; o += NG) {
                if (!m_other[(size_t)b * n_other + o]) continue;
                const size_t gi = SIDE == 0 ? ((size_t)row * n_t + o)
                                            : (((size_t)b * n_s + o) * n_t + i_own);
                const float gj = __ldg(gr + gi);
                if (SIDE == 0 && gl == 0) db += gj;
#pragma unroll
                for (int v = 0; v < NV; ++v) {
                    const int q = gl + v * G;
                    if (q < r4) {
                        const float4 t = ldg4(other + ((size_t)b * n_other + o) * r + 4 * q);
                        // difference is always ps - pt
                        const float dx = SIDE == 0 ? pv[v].x - t.x : t.x - pv[v].x;
                        const float dy = SIDE == 0 ? pv[v].y - t.y : t.y - pv[v].y;
                        const float dz = SIDE == 0 ? pv[v].z - t.z : t.z - pv[v].z;
                        const float dq = SIDE == 0 ? pv[v].w - t.w : t.w - pv[v].w;
                        dp[v].x += dx > 0.f ? gj : 0.f;
                        dp[v].y += dy > 0.f ? gj : 0.f;
                        dp[v].z += dz > 0.f ? gj : 0.f;
                        dp[v].w += dq > 0.f ? gj : 0.f;
                        if (SIDE == 0) {
                            dw[v].x += gj * fmaxf(dx, 0.f); dw[v].y += gj * fmaxf(dy, 0.f);
                            dw[v].z += gj * fmaxf(dz, 0.f); dw[v].w += gj * fmaxf(dq, 0.f);
                        }
                    }
                }
            }
        }
        const float sign = SIDE == 0 ? 1.f : -1.f;
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
                float4 o4 = make_float4(sign * dp[v].x * wv[v].x, sign * dp[v].y * wv[v].y,
                                        sign * dp[v].z * wv[v].z, sign * dp[v].w * wv[v].w);
                *reinterpret_cast<float4*>(d_own + (size_t)row * r + 4 * q) = o4;
            }
        }
    }
    if (SIDE == 0) {
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
}

__global__ void __launch_bounds__(256) dense_finish_kernel(const float* __restrict__ partial,
                                                           int blocks, int r,
                                                           float* __restrict__ d_w2,
                                                           float* __restrict__ d_b2) {
    for (int c = blockIdx.x * blockDim.x + threadIdx.x; c <= r; c += gridDim.x * blockDim.x) {
        float s = 0.f;
        for (int b = 0; b < blocks; ++b) s += partial[(size_t)b * (r + 1) + c];
        if (c < r) d_w2[c] = s; else *d_b2 = s;
    }
}

}  // namespace dgmc

using namespace dgmc;

#define DC_DISPATCH(KERNEL, GRID, SMEM, ...)                                                    \
    do {                                                                                        \
        if (r4 <= 8) KERNEL(8, 1)<<<GRID, 256, SMEM, stream>>>(__VA_ARGS__);                   \
        else if (r4 <= 16) KERNEL(16, 1)<<<GRID, 256, SMEM, stream>>>(__VA_ARGS__);            \
        else if (r4 <= 32) KERNEL(32, 1)<<<GRID, 256, SMEM, stream>>>(__VA_ARGS__);            \
        else if (r4 <= 64) KERNEL(32, 2)<<<GRID, 256, SMEM, stream>>>(__VA_ARGS__);            \
        else KERNEL(32, 4)<<<GRID, 256, SMEM, stream>>>(__VA_ARGS__);                          \
    } while (0)

static int dc_check_r(const char* who, int r) {
    if (r < 4 || r % 4 != 0 || r > 512) {
        set_error("%s: r=%d unsupported (need a multiple of 4 in [4, 512]; the host pads)", who, r);
        return DGMC_ERR_UNSUPPORTED;
    }
    return DGMC_OK;
}

extern "C" int dgmc_dense_consensus_fwd(const float* s_in, const float* ps, const float* pt,
                                        const float* w2, const float* b2, const uint8_t* s_mask,
                                        const uint8_t* t_mask, float* s_out, int B, int n_s, int n_t,
                                        int r, dgmc_stream_t stream) {
    DGMC_REQUIRE(s_in && ps && pt && w2 && b2 && s_mask && t_mask && s_out && B >= 0 && n_s >= 0 &&
                     n_t >= 1, "dense_consensus_fwd: bad arguments");
    if (int e = dc_check_r("dense_consensus_fwd", r)) return e;
    const long long rows = (long long)B * n_s;
    if (rows == 0) return DGMC_OK;
    const int r4 = r / 4;
    const int grid = wave_grid(rows, 8, 8);
#define K_FWD(G, NV) dense_consensus_fwd_kernel<G, NV>
    DC_DISPATCH(K_FWD, grid, 0, s_in, ps, pt, w2, b2, s_mask, t_mask, s_out, B, n_s, n_t, r4);
#undef K_FWD
    return check_launch("dense_consensus_fwd");
}

extern "C" size_t dgmc_dense_consensus_bwd_workspace(int r) {
    return (size_t)DC_MAX_BLOCKS * (size_t)(r + 1) * sizeof(float);
}

extern "C" int dgmc_dense_consensus_bwd(const float* g, const float* ps, const float* pt,
                                        const float* w2, const uint8_t* s_mask,
                                        const uint8_t* t_mask, float* d_ps, float* d_pt, float* d_w2,
                                        float* d_b2, int B, int n_s, int n_t, int r, void* workspace,
                                        size_t workspace_bytes, dgmc_stream_t stream) {
    DGMC_REQUIRE(g && ps && pt && w2 && s_mask && t_mask && d_ps && d_pt && d_w2 && d_b2 &&
                     workspace && B >= 1 && n_s >= 1 && n_t >= 1, "dense_consensus_bwd: bad arguments");
    if (int e = dc_check_r("dense_consensus_bwd", r)) return e;
    if (workspace_bytes < dgmc_dense_consensus_bwd_workspace(r)) {
        set_error("dense_consensus_bwd: workspace too small");
        return DGMC_ERR_WORKSPACE;
    }
    const int r4 = r / 4;
    float* partial = (float*)workspace;
    int grid_s = wave_grid((long long)B * n_s, 8, 4);
    if (grid_s > DC_MAX_BLOCKS) grid_s = DC_MAX_BLOCKS;
    const int grid_t = wave_grid((long long)B * n_t, 8, 4);
    const size_t smem = 8 * (size_t)(r + 1) * sizeof(float);
#define K_S(G, NV) dense_consensus_bwd_kernel<G, NV, 0>
#define K_T(G, NV) dense_consensus_bwd_kernel<G, NV, 1>
    DC_DISPATCH(K_S, grid_s, smem, g, ps, pt, w2, s_mask, t_mask, d_ps, partial, B, n_s, n_t, r4);
    DC_DISPATCH(K_T, grid_t, smem, g, ps, pt, w2, s_mask, t_mask, d_pt, partial, B, n_s, n_t, r4);
#undef K_S
#undef K_T
    dense_finish_kernel<<<(r + 256) / 256, 256, 0, stream>>>(partial, grid_s, r, d_w2, d_b2);
    return check_launch("dense_consensus_bwd", 3);
}
