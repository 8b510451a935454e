// K3 (row softmax over k candidates, masked row softmax) and K2 (candidate gather-dot).
// All three are HBM/L2-bandwidth kernels: one sub-warp group of G lanes per row (softmax) or
// per candidate (gather-dot), float4 accesses, grid-stride over rows.
#include "common.cuh"
#include "../../include/dgmc_b200.h"

namespace dgmc {

// ---------------------------------------------------------------------------------------------
// softmax over k  (dgmc.py:202,205,225)
// ---------------------------------------------------------------------------------------------
template <int G>
__global__ void __launch_bounds__(256) softmax_k_fwd_kernel(const float* __restrict__ x,
                                                            float* __restrict__ y, long long rows,
                                                            int k) {
    const int gl = threadIdx.x % G;
    const long long groups = (long long)gridDim.x * (blockDim.x / G);
    for (long long row = (long long)blockIdx.x * (blockDim.x / G) + threadIdx.x / G;
         row < (rows + (32 / G) - 1) / (32 / G) * (32 / G); row += groups) {
        const bool live = row < rows;
        const float* xr = x + (size_t)(live ? row : 0) * k;
        float mx = -INFINITY;
        if (live)
            for (int j = gl; j < k; j += G) mx = fmaxf(mx, __ldg(xr + j));
#pragma unroll
        for (int o = G / 2; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        float sum = 0.f;
        if (live)
            for (int j = gl; j < k; j += G) sum += __expf(__ldg(xr + j) - mx);
        sum = group_sum<G>(sum);
        if (live) {
            const float inv = 1.f / sum;
            float* yr = y + (size_t)row * k;
            for (int j = gl; j < k; j += G) yr[j] = __expf(__ldg(xr + j) - mx) * inv;
        }
    }
}

template <int G>
__global__ void __launch_bounds__(256) softmax_k_bwd_kernel(const float* __restrict__ y,
                                                            const float* __restrict__ dy,
                                                            float* __restrict__ dx, long long rows,
                                                            int k) {
    const int gl = threadIdx.x % G;
    const long long groups = (long long)gridDim.x * (blockDim.x / G);
    for (long long row = (long long)blockIdx.x * (blockDim.x / G) + threadIdx.x / G;
         row < (rows + (32 / G) - 1) / (32 / G) * (32 / G); row += groups) {
        const bool live = row < rows;
        const size_t base = (size_t)(live ? row : 0) * k;
        float dot = 0.f;
        if (live)
            for (int j = gl; j < k; j += G) dot += __ldg(y + base + j) * __ldg(dy + base + j);
        dot = group_sum<G>(dot);
        if (live)
            for (int j = gl; j < k; j += G)
                dx[base + j] = __ldg(y + base + j) * (__ldg(dy + base + j) - dot);
    }
}

template <typename F>
static int launch_by_group(int k, F&& f) {
    if (k <= 24) return f(std::integral_constant<int, 8>());
    if (k <= 48) return f(std::integral_constant<int, 16>());
    return f(std::integral_constant<int, 32>());
}

// ---------------------------------------------------------------------------------------------
// masked softmax, dense branch (dgmc.py:15-19).  One warp per (b, i) row of [B, n_s, n_t].
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) masked_softmax_fwd_kernel(
    const float* __restrict__ x, const uint8_t* __restrict__ s_mask,
    const uint8_t* __restrict__ t_mask, float* __restrict__ y, int B, int n_s, int n_t) {
    const int lane = threadIdx.x & 31;
    const long long rows = (long long)B * n_s;
    const long long warps = (long long)gridDim.x * (blockDim.x >> 5);
    for (long long row = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); row < rows;
         row += warps) {
        const int b = (int)(row / n_s);
        const float* xr = x + (size_t)row * n_t;
        float* yr = y + (size_t)row * n_t;
        const uint8_t* tm = t_mask + (size_t)b * n_t;
        if (!s_mask[row]) {  // padded source row: reference gives NaN, then re-zeroes it
            for (int j = lane; j < n_t; j += 32) yr[j] = 0.f;
            continue;
        }
        float mx = -INFINITY;
        for (int j = lane; j < n_t; j += 32)
            if (tm[j]) mx = fmaxf(mx, __ldg(xr + j));
        mx = warp_max(mx);
        float sum = 0.f;
        for (int j = lane; j < n_t; j += 32)
            if (tm[j]) sum += __expf(__ldg(xr + j) - mx);
        sum = warp_sum(sum);
        const float inv = 1.f / sum;
        for (int j = lane; j < n_t; j += 32) yr[j] = tm[j] ? __expf(__ldg(xr + j) - mx) * inv : 0.f;
    }
}

// masked entries have y == 0, so dx = y * (dy - <y, dy>) is already 0 there.
__global__ void __launch_bounds__(256) masked_softmax_bwd_kernel(const float* __restrict__ y,
                                                                 const float* __restrict__ dy,
                                                                 float* __restrict__ dx,
                                                                 long long rows, int n_t) {
    const int lane = threadIdx.x & 31;
    const long long warps = (long long)gridDim.x * (blockDim.x >> 5);
    for (long long row = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); row < rows;
         row += warps) {
        const size_t base = (size_t)row * n_t;
        float dot = 0.f;
        for (int j = lane; j < n_t; j += 32) dot += __ldg(y + base + j) * __ldg(dy + base + j);
        dot = warp_sum(dot);
        for (int j = lane; j < n_t; j += 32)
            dx[base + j] = __ldg(y + base + j) * (__ldg(dy + base + j) - dot);
    }
}

// ---------------------------------------------------------------------------------------------
// gather-dot  (dgmc.py:198-201):  out[m, j] = <a[m], b[idx[m, j]]>
// One warp per row m; the row of `a` stays in registers (NV float4 per lane); groups of G lanes
// take candidates j = g, g + 32/G, ... so that narrow rows (c = 32, 64) still use all 32 lanes.
// ---------------------------------------------------------------------------------------------
template <int G, int NV>
__global__ void __launch_bounds__(256) gather_dot_kernel(const float* __restrict__ a, long long lda,
                                                         const float* __restrict__ b, long long ldb,
                                                         const int* __restrict__ idx,
                                                         float* __restrict__ out, long long m,
                                                         int k, int c4) {
    constexpr int NG = 32 / G;
    const int lane = threadIdx.x & 31, gl = lane % G, g = lane / G;
    const long long warps = (long long)gridDim.x * (blockDim.x >> 5);
    for (long long row = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); row < m;
         row += warps) {
        float4 av[NV];
#pragma unroll
        for (int v = 0; v < NV; ++v) {
            const int q = gl + v * G;
            av[v] = q < c4 ? ldg4(a + (size_t)row * lda + 4 * q) : make_float4(0, 0, 0, 0);
        }
        const int* ir = idx + (size_t)row * k;
        float* orow = out + (size_t)row * k;
        for (int j0 = 0; j0 < k; j0 += 2 * NG) {  // two candidates in flight per group
            const int ja = j0 + g, jb = j0 + NG + g;
            const int ta = ja < k ? __ldg(ir + ja) : 0, tb = jb < k ? __ldg(ir + jb) : 0;
            float sa = 0.f, sb = 0.f;
#pragma unroll
            for (int v = 0; v < NV; ++v) {
                const int q = gl + v * G;
                if (q < c4) {
                    const float4 pa = ldg4(b + (size_t)ta * ldb + 4 * q);
                    const float4 pb = ldg4(b + (size_t)tb * ldb + 4 * q);
                    sa += av[v].x * pa.x + av[v].y * pa.y + av[v].z * pa.z + av[v].w * pa.w;
                    sb += av[v].x * pb.x + av[v].y * pb.y + av[v].z * pb.z + av[v].w * pb.w;
                }
            }
            sa = group_sum<G>(sa);
            sb = group_sum<G>(sb);
            if (gl == 0) {
                if (ja < k) orow[ja] = sa;
                if (jb < k) orow[jb] = sb;
            }
        }
    }
}

// any c (no alignment assumption): one warp per (row, candidate), scalar loads
__global__ void __launch_bounds__(256) gather_dot_generic_kernel(
    const float* __restrict__ a, long long lda, const float* __restrict__ b, long long ldb,
    const int* __restrict__ idx, float* __restrict__ out, long long m, int k, int c) {
    const int lane = threadIdx.x & 31;
    const long long total = m * k;
    const long long warps = (long long)gridDim.x * (blockDim.x >> 5);
    for (long long e = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); e < total;
         e += warps) {
        const long long row = e / k;
        const float* ar = a + (size_t)row * lda;
        const float* br = b + (size_t)__ldg(idx + e) * ldb;
        float s = 0.f;
        for (int q = lane; q < c; q += 32) s += __ldg(ar + q) * __ldg(br + q);
        s = warp_sum(s);
        if (lane == 0) out[e] = s;
    }
}

}  // namespace dgmc

namespace dgmc {
// out[i, :] = y[i, :] > 0 ? g[i, :] : 0 on row-strided operands, one float4 per thread
__global__ void __launch_bounds__(256) relu_bwd_kernel(const float* __restrict__ g, long long ldg,
                                                       const float* __restrict__ y, long long ldy,
                                                       float* __restrict__ out, long long ldo,
                                                       long long rows, int c4) {
    const long long total = rows * c4;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += stride) {
        const long long i = t / c4;
        const int q = (int)(t % c4) * 4;
        const float4 gv = ldg4(g + i * ldg + q), yv = ldg4(y + i * ldy + q);
        *reinterpret_cast<float4*>(out + i * ldo + q) =
            make_float4(yv.x > 0.f ? gv.x : 0.f, yv.y > 0.f ? gv.y : 0.f, yv.z > 0.f ? gv.z : 0.f,
                        yv.w > 0.f ? gv.w : 0.f);
    }
}
}  // namespace dgmc

using namespace dgmc;

extern "C" int dgmc_relu_bwd(const float* g, int64_t ldg, const float* y, int64_t ldy, float* out,
                             int64_t ldo, int64_t rows, int c, dgmc_stream_t stream) {
    if (rows == 0) return DGMC_OK;
    DGMC_REQUIRE(g && y && out && rows > 0 && c >= 4 && c % 4 == 0 && ldg % 4 == 0 && ldy % 4 == 0 &&
                     ldo % 4 == 0 && ldg >= c && ldy >= c && ldo >= c &&
                     ((reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(y) |
                       reinterpret_cast<uintptr_t>(out)) & 15) == 0,
                 "relu_bwd: needs c %% 4 == 0 and 16-byte aligned rows (c=%d)", c);
    relu_bwd_kernel<<<wave_grid(rows * (c / 4), 256, 8), 256, 0, stream>>>(g, ldg, y, ldy, out, ldo, rows, c / 4);
    return check_launch("relu_bwd");
}

extern "C" int dgmc_softmax_k_fwd(const float* x, float* y, int64_t rows, int k,
                                  dgmc_stream_t stream) {
    if (rows == 0 && k >= 1) return DGMC_OK;   // empty input: pointers may be null
    DGMC_REQUIRE(x && y && rows >= 0 && k >= 1, "softmax_k_fwd: bad arguments (rows=%lld k=%d)",
                 (long long)rows, k);
    if (rows == 0) return DGMC_OK;
    return launch_by_group(k, [&](auto G) {
        constexpr int g = decltype(G)::value;
        softmax_k_fwd_kernel<g><<<wave_grid(rows, 256 / g, 8), 256, 0, stream>>>(x, y, rows, k);
        return check_launch("softmax_k_fwd");
    });
}

extern "C" int dgmc_softmax_k_bwd(const float* y, const float* dy, float* dx, int64_t rows, int k,
                                  dgmc_stream_t stream) {
    if (rows == 0 && k >= 1) return DGMC_OK;
    DGMC_REQUIRE(y && dy && dx && rows >= 0 && k >= 1, "softmax_k_bwd: bad arguments");
    if (rows == 0) return DGMC_OK;
    return launch_by_group(k, [&](auto G) {
        constexpr int g = decltype(G)::value;
        softmax_k_bwd_kernel<g><<<wave_grid(rows, 256 / g, 8), 256, 0, stream>>>(y, dy, dx, rows, k);
        return check_launch("softmax_k_bwd");
    });
}

extern "C" int dgmc_masked_softmax_fwd(const float* x, const uint8_t* s_mask, const uint8_t* t_mask,
                                       float* y, int B, int n_s, int n_t, dgmc_stream_t stream) {
    DGMC_REQUIRE(x && y && s_mask && t_mask && B >= 0 && n_s >= 0 && n_t >= 1,
                 "masked_softmax_fwd: bad arguments");
    const long long rows = (long long)B * n_s;
    if (rows == 0) return DGMC_OK;
    masked_softmax_fwd_kernel<<<wave_grid(rows, 8, 8), 256, 0, stream>>>(x, s_mask, t_mask, y, B,
                                                                         n_s, n_t);
    return check_launch("masked_softmax_fwd");
}

extern "C" int dgmc_masked_softmax_bwd(const float* y, const float* dy, float* dx, int B, int n_s,
                                       int n_t, dgmc_stream_t stream) {
    DGMC_REQUIRE(y && dy && dx && B >= 0 && n_s >= 0 && n_t >= 1, "masked_softmax_bwd: bad arguments");
    const long long rows = (long long)B * n_s;
    if (rows == 0) return DGMC_OK;
    masked_softmax_bwd_kernel<<<wave_grid(rows, 8, 8), 256, 0, stream>>>(y, dy, dx, rows, n_t);
    return check_launch("masked_softmax_bwd");
}

extern "C" int dgmc_gather_dot_fwd(const float* a, int64_t lda, const float* b, int64_t ldb,
                                   const int32_t* idx, float* out, int64_t m, int k, int c,
                                   dgmc_stream_t stream) {
    if (m == 0 && k >= 1 && c >= 1) return DGMC_OK;
    DGMC_REQUIRE(a && b && idx && out && m >= 0 && k >= 1 && c >= 1 && lda >= c && ldb >= c,
                 "gather_dot_fwd: bad arguments (m=%lld k=%d c=%d)", (long long)m, k, c);
    if (m == 0) return DGMC_OK;
    const bool vec = (c % 4 == 0) && (lda % 4 == 0) && (ldb % 4 == 0) &&
                     ((uintptr_t)a % 16 == 0) && ((uintptr_t)b % 16 == 0);
    const int grid = wave_grid(m, 8, 8);
    if (vec && c <= 512) {
        const int c4 = c / 4;
#define GD(G, NV)                                                                                \
    gather_dot_kernel<G, NV><<<grid, 256, 0, stream>>>(a, lda, b, ldb, idx, out, m, k, c4)
        if (c4 <= 8) GD(8, 1);
        else if (c4 <= 16) GD(16, 1);
        else if (c4 <= 32) GD(32, 1);
        else if (c4 <= 64) GD(32, 2);
        else if (c4 <= 96) GD(32, 3);
        else GD(32, 4);
#undef GD
    } else {
        gather_dot_generic_kernel<<<wave_grid(m * k, 8, 8), 256, 0, stream>>>(a, lda, b, ldb, idx,
                                                                             out, m, k, c);
    }
    return check_launch("gather_dot_fwd");
}

// ---------------------------------------------------------------------------------------------
// loss / acc / hits_at_k on the sparse side-cars (dgmc.py:246-311), one warp per ground-truth pair.
// For pair g = (row, col): with v = val[row, :], c = idx[row, :]
//   nll[g]     = sum_j [c_j == col] * -log(v_j + eps)        (dgmc.py:264-266: every listed copy counts)
//   hits[g]    = #{j : c_j == col}                            (0: the pair is dropped from the mean)
//   correct[g] = [c_{argmax v} == col]                        (first maximum, as torch.argmax; dgmc.py:281-284)
//   in_top[g]  = #{j : c_j == col and v_j is among the top_k largest}   (stable order; dgmc.py:302-308)
// The host sums the four arrays on the device; only the Python number the reference API returns
// (`.item()`, dgmc.py:285,309) costs a synchronisation.
// ---------------------------------------------------------------------------------------------
namespace dgmc {
__global__ void __launch_bounds__(256) sparse_metrics_kernel(const long long* __restrict__ idx,
                                                             const float* __restrict__ val,
                                                             const long long* __restrict__ y_row,
                                                             const long long* __restrict__ y_col, long long pairs,
                                                             int k, int top_k, float eps, float* __restrict__ nll,
                                                             float* __restrict__ hits, float* __restrict__ correct,
                                                             float* __restrict__ in_top) {
    const int lane = threadIdx.x & 31;
    const long long warps = (long long)gridDim.x * (blockDim.x >> 5);
    for (long long g = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); g < pairs; g += warps) {
        const long long row = y_row[g], col = y_col[g];
        const long long* c = idx + row * k;
        const float* v = val + row * k;
        float my_nll = 0.f, my_hits = 0.f, my_top = 0.f;
        float best = -INFINITY;
        int best_j = k;
        for (int j = lane; j < k; j += 32) {
            const float vj = v[j];
            if (vj > best) { best = vj; best_j = j; }        // ascending j per lane: first maximum kept
            if (c[j] == col) {
                my_nll += -logf(vj + eps);
                my_hits += 1.f;
                int rank = 0;                                 // entries that sort before j (descending, stable)
                for (int i = 0; i < k; ++i) {
                    const float vi = v[i];
                    rank += (vi > vj) || (vi == vj && i < j);
                }
                my_top += rank < top_k ? 1.f : 0.f;
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ob = __shfl_xor_sync(0xffffffffu, best, o);
            const int oj = __shfl_xor_sync(0xffffffffu, best_j, o);
            if (ob > best || (ob == best && oj < best_j)) { best = ob; best_j = oj; }
        }
        my_nll = warp_sum(my_nll);
        my_hits = warp_sum(my_hits);
        my_top = warp_sum(my_top);
        if (lane == 0) {
            nll[g] = my_nll;
            hits[g] = my_hits;
            correct[g] = (best_j < k && c[best_j] == col) ? 1.f : 0.f;
            in_top[g] = my_top;
        }
    }
}
}  // namespace dgmc

extern "C" int dgmc_sparse_metrics(const int64_t* idx, const float* val, const int64_t* y_row, const int64_t* y_col,
                                   int64_t pairs, int k, int top_k, float eps, float* nll, float* hits,
                                   float* correct, float* in_top, dgmc_stream_t stream) {
    DGMC_REQUIRE(idx && val && y_row && y_col && nll && hits && correct && in_top && k >= 1 && pairs >= 0,
                 "sparse_metrics: bad arguments");
    if (pairs == 0) return DGMC_OK;
    dgmc::sparse_metrics_kernel<<<dgmc::wave_grid(pairs, 8, 8), 256, 0, stream>>>(
        reinterpret_cast<const long long*>(idx), val, reinterpret_cast<const long long*>(y_row),
        reinterpret_cast<const long long*>(y_col), pairs, k, top_k, eps, nll, hits, correct, in_top);
    return dgmc::check_launch("sparse_metrics");
}
