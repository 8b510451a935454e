// K4 / K5: segmented weighted gather-reduce (CSR SpMM with fp32 rows of width r).
//   out[s, :] = seg_scale[s] * sum_{e in seg(s)} w_e * x[row_e, :]
// One group of G lanes (G = pow2 >= r/4, <= 32) owns one segment; each lane keeps NV float4
// accumulators; four entries are kept in flight per group to cover L2 latency.  Terms are added
// in CSR order, so results are reproducible (the CSR comes from the stable dgmc_csr_build).
#include "common.cuh"
#include "../../include/dgmc_b200.h"

namespace dgmc {

struct SegArgs {
    const int* ptr; int fixed_len;
    const int* src; int src_div;
    const int* w_index; const float* w; const float* src_scale; const float* seg_scale;
    const float* x; long long ldx; float* out; long long ldo; long long n_seg; int r; int accumulate;
};

template <int G, int NV, bool VEC>
__global__ void __launch_bounds__(256) segment_gather_kernel(const SegArgs a) {
    constexpr int NG = 32 / G;
    const int lane = threadIdx.x & 31, gl = lane % G;
    const long long groups = (long long)gridDim.x * (blockDim.x / G);
    const int width = VEC ? a.r / 4 : a.r;  // columns in units of (float4 | float)
    for (long long seg = (long long)blockIdx.x * (blockDim.x / G) + threadIdx.x / G; seg < a.n_seg;
         seg += groups) {
        long long e0, e1;
        if (a.ptr) { e0 = __ldg(a.ptr + seg); e1 = __ldg(a.ptr + seg + 1); }
        else { e0 = seg * a.fixed_len; e1 = e0 + a.fixed_len; }
        float4 acc[NV];
#pragma unroll
        for (int v = 0; v < NV; ++v) acc[v] = make_float4(0, 0, 0, 0);
        for (long long e = e0; e < e1; e += 4) {
            int rows[4]; float wt[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const bool live = e + u < e1;
                int row = 0; float we = 0.f;
                if (live) {
                    const int sv = __ldg(a.src + e + u);
                    row = a.src_div > 1 ? sv / a.src_div : sv;
                    we = a.w ? __ldg(a.w + (a.w_index ? __ldg(a.w_index + e + u) : (int)(e + u))) : 1.f;
                    if (a.src_scale) we *= __ldg(a.src_scale + row);
                }
                rows[u] = row; wt[u] = we;
            }
#pragma unroll
            for (int v = 0; v < NV; ++v) {
                const int q = gl + v * G;
                if (q < width) {
                    if (VEC) {
                        float4 t[4];
#pragma unroll
                        for (int u = 0; u < 4; ++u) t[u] = ldg4(a.x + (size_t)rows[u] * a.ldx + 4 * q);
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            acc[v].x += wt[u] * t[u].x; acc[v].y += wt[u] * t[u].y;
                            acc[v].z += wt[u] * t[u].z; acc[v].w += wt[u] * t[u].w;
                        }
                    } else {
                        float t[4];
#pragma unroll
                        for (int u = 0; u < 4; ++u) t[u] = __ldg(a.x + (size_t)rows[u] * a.ldx + q);
#pragma unroll
                        for (int u = 0; u < 4; ++u) acc[v].x += wt[u] * t[u];
                    }
                }
            }
        }
        const float sc = a.seg_scale ? __ldg(a.seg_scale + seg) : 1.f;
#pragma unroll
        for (int v = 0; v < NV; ++v) {
            const int q = gl + v * G;
            if (q < width) {
                if (VEC) {
                    float4 o = make_float4(acc[v].x * sc, acc[v].y * sc, acc[v].z * sc, acc[v].w * sc);
                    float4* dst = reinterpret_cast<float4*>(a.out + (size_t)seg * a.ldo + 4 * q);
                    if (a.accumulate) { const float4 p = *dst; o.x += p.x; o.y += p.y; o.z += p.z; o.w += p.w; }
                    *dst = o;
                } else {
                    float* dst = a.out + (size_t)seg * a.ldo + q;
                    *dst = acc[v].x * sc + (a.accumulate ? *dst : 0.f);
                }
            }
        }
    }
    (void)NG;
}

template <bool VEC>
static int launch_segment_gather(const SegArgs& a, cudaStream_t stream) {
    const int width = VEC ? a.r / 4 : a.r;
#define SG(G, NV)                                                                             \
    segment_gather_kernel<G, NV, VEC><<<wave_grid(a.n_seg, 256 / G, 8), 256, 0, stream>>>(a)
    if (width <= 8) SG(8, 1);
    else if (width <= 16) SG(16, 1);
    else if (width <= 32) SG(32, 1);
    else if (width <= 64) SG(32, 2);
    else if (width <= 128) SG(32, 4);
    else if (width <= 256) SG(32, 8);
    else {
        set_error("segment_gather: r=%d too wide (max %d)", a.r, VEC ? 1024 : 256);
        return DGMC_ERR_UNSUPPORTED;
    }
#undef SG
    return check_launch("segment_gather");
}

}  // namespace dgmc

using namespace dgmc;

extern "C" int dgmc_segment_gather(const int32_t* ptr, int fixed_len, const int32_t* src, int src_div,
                                   const int32_t* w_index, const float* w, const float* src_scale,
                                   const float* seg_scale, const float* x, int64_t ldx, float* out,
                                   int64_t ldo, int64_t n_seg, int r, int accumulate, dgmc_stream_t stream) {
    DGMC_REQUIRE(x && out && n_seg >= 0 && r >= 1 && ldx >= r && ldo >= r && src_div >= 1,
                 "segment_gather: bad arguments (n_seg=%lld r=%d)", (long long)n_seg, r);
    DGMC_REQUIRE(ptr || fixed_len >= 0, "segment_gather: need ptr or fixed_len");
    DGMC_REQUIRE(src || (ptr == nullptr && fixed_len == 0), "segment_gather: null src");
    if (n_seg == 0) return DGMC_OK;
    SegArgs a{ptr, fixed_len, src, src_div, w_index, w, src_scale, seg_scale,
              x, ldx, out, ldo, n_seg, r, accumulate};
    const bool vec = (r % 4 == 0) && (ldx % 4 == 0) && (ldo % 4 == 0) &&
                     ((uintptr_t)x % 16 == 0) && ((uintptr_t)out % 16 == 0);
    return vec ? launch_segment_gather<true>(a, stream) : launch_segment_gather<false>(a, stream);
}
