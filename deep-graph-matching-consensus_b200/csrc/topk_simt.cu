// K1 (small-shape variant): exact fp32 top-k of <h_s[b,i], h_t[b,j]> over j, one warp per source
// row, scores never leave the SM.  This is what runs when the pair is too small to fill a
// 128x128 tcgen05 tile (the reference's unit-test graphs, mini-batches of keypoint graphs) or
// when C is not a multiple of 8; it is also the on-device exact cross-check for the tensor-core
// kernel in tests.  Order: descending score, ties by ascending column (dgmc.py:85-94).
#include <float.h>
#include "common.cuh"
#include "../../include/dgmc_b200.h"

namespace dgmc {

constexpr int TS_WARPS = 4;

__global__ void __launch_bounds__(TS_WARPS * 32) topk_simt_kernel(
    const float* __restrict__ h_s, const float* __restrict__ h_t, int n_s, int n_t, int c, int k,
    int* __restrict__ idx_out, float* __restrict__ score_out) {
    extern __shared__ float smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // per-warp layout: a-row [c_pad] | vals [k][32] | idx [k][32]
    const int c_pad = (c + 3) & ~3;
    float* arow = smem + (size_t)warp * (c_pad + 64 * k);
    float* vals = arow + c_pad;
    int* ids = reinterpret_cast<int*>(vals + 32 * k);
    const int b = blockIdx.y;
    const float* hs = h_s + (size_t)b * n_s * c;
    const float* ht = h_t + (size_t)b * n_t * c;
    for (int row = blockIdx.x * TS_WARPS + warp; row < n_s; row += gridDim.x * TS_WARPS) {
        for (int q = lane; q < c; q += 32) arow[q] = __ldg(hs + (size_t)row * c + q);
        __syncwarp();
        int cnt = 0;
        for (int j = lane; j < n_t; j += 32) {
            const float* br = ht + (size_t)j * c;
            float s = 0.f;
            for (int q = 0; q < c; ++q) s = fmaf(arow[q], __ldg(br + q), s);
            if (cnt < k || s > vals[(k - 1) * 32 + lane]) {
                int pos = cnt < k ? cnt : k - 1;
                while (pos > 0 && vals[(pos - 1) * 32 + lane] < s) {
                    vals[pos * 32 + lane] = vals[(pos - 1) * 32 + lane];
                    ids[pos * 32 + lane] = ids[(pos - 1) * 32 + lane];
                    --pos;
                }
                vals[pos * 32 + lane] = s;
                ids[pos * 32 + lane] = j;
                if (cnt < k) ++cnt;
            }
        }
        // k-way merge of the 32 per-lane lists
        int head = 0;
        int* orow = idx_out + ((size_t)b * n_s + row) * k;
        float* srow = score_out ? score_out + ((size_t)b * n_s + row) * k : nullptr;
        for (int round = 0; round < k; ++round) {
            float v = head < cnt ? vals[head * 32 + lane] : -FLT_MAX;
            int id = head < cnt ? ids[head * 32 + lane] : 0x7fffffff;
            float bv = v; int bi = id;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
                const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
                if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
            }
            if (id == bi && head < cnt) ++head;  // column ids are unique across lanes
            if (lane == 0) {
                orow[round] = bi;
                if (srow) srow[round] = bv;
            }
        }
        __syncwarp();
    }
}

}  // namespace dgmc

using namespace dgmc;

extern "C" int dgmc_topk_simt(const float* h_s, const float* h_t, int B, int n_s, int n_t, int c,
                              int k, int32_t* idx_out, float* score_out, dgmc_stream_t stream) {
    DGMC_REQUIRE(h_s && h_t && idx_out && B >= 1 && n_s >= 0 && n_t >= 1 && c >= 1,
                 "topk_simt: bad arguments");
    DGMC_REQUIRE(k >= 1 && k <= n_t, "topk_simt: k=%d must be in [1, n_t=%d]", k, n_t);
    if (n_s == 0) return DGMC_OK;
    const size_t smem = (size_t)TS_WARPS * (((c + 3) & ~3) + 64 * (size_t)k) * sizeof(float);
    if (smem > 200 * 1024) {
        set_error("topk_simt: k=%d, c=%d need %zu B of shared memory (max 200 KB)", k, c, smem);
        return DGMC_ERR_UNSUPPORTED;
    }
    static thread_local size_t configured = 0;
    if (smem > 48 * 1024 && smem > configured) {
        if (cudaFuncSetAttribute(topk_simt_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)smem) != cudaSuccess) {
            set_error("topk_simt: cannot raise dynamic shared memory to %zu", smem);
            return DGMC_ERR_CUDA;
        }
        configured = smem;
    }
    int gx = (n_s + TS_WARPS - 1) / TS_WARPS;
    const int cap = sm_count() * 8;
    if (gx > cap) gx = cap;
    topk_simt_kernel<<<dim3(gx, B), TS_WARPS * 32, smem, stream>>>(h_s, h_t, n_s, n_t, c, k, idx_out,
                                                                  score_out);
    return check_launch("topk_simt");
}
