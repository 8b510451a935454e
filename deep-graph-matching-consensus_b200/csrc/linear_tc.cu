// Per-node linear map  Y[rows, n_out] = act(X[rows, k_in] W[n_out, k_in]^T + bias)  on the 5th-gen
// tensor cores with fp32-level accuracy ("3xTF32"): X = X_hi + X_lo and W = W_hi + W_lo with hi = x
// rounded to tf32 (19 bits) and lo = x - hi rounded to tf32; the product is accumulated as
// X_hi W_hi + X_hi W_lo + X_lo W_hi in fp32 TMEM accumulators (the dropped lo*lo term and the
// rounding of lo are ~2^-24 relative and zero-mean; measured error against fp64 = that of the
// library's IEEE fp32 SGEMM, tests/test_gpu_linear_tc.py).  This is the GEMM behind every RelConv /
// SplineConv / Lin layer of psi_1 and psi_2 (rel.py:27-30, spline.py:40-47, mlp.py) and, with the
// transposed weight, their input gradient; the library SGEMM runs those on the FFMA pipe at ~40
// TFLOP/s.
//
// W is small (<= a few MB): a prologue kernel splits (and, for the input gradient, transposes) it
// ONCE into K-major, zero-padded hi / lo copies in the caller's workspace.  X is large and is split
// on the fly.  One CTA per 128-row x BN-column tile, 6 warps:
//   warp 0      TMA producer: per 32-wide k block, X[128 x 32] (raw fp32), W_hi[BN x 32], W_lo[BN x 32]
//               -> 128-byte-swizzled shared-memory stage, mbarrier complete_tx
//   warps 2-5   split the raw X tile into hi (in place) and lo (second buffer, same swizzled
//               position, so no address math), fence.proxy.async, arrive
//   warp 1      one thread issues, per k8 step, X_hi x [W_hi ; W_lo] (N = 2 BN) and X_lo x W_hi
//               (N = BN) as tcgen05.mma.kind::tf32; tcgen05.commit releases the stage to the producer
//   warps 2-5   epilogue: tcgen05.ld of the accumulator rows (summing the accumulator pairs),
//               + bias, ReLU, 128-bit stores
// Rows and k beyond the matrix are zero-filled by TMA (X) or by the padded workspace (W).
#include "common.cuh"
#include "tc_common.cuh"
#include "../../include/dgmc_b200.h"

namespace dgmc {
namespace tc {

constexpr int LT_BM = 128;                    // rows per tile (UMMA M)
constexpr int LT_KB = 32;                     // k per stage = one 128-byte swizzled row of fp32
constexpr int LT_A_BYTES = LT_BM * LT_KB * 4; // 16 KB
constexpr int LT_THREADS = 192;
constexpr int LT_MAX_STAGES = 6;
constexpr int LT_ACC = 2;                     // accumulator pairs [H | L] the k8 steps alternate between

// x = hi + lo + O(2^-24 x) with hi and lo both EXACT tf32 values (round to nearest, so that the
// tensor core's operand truncation is a no-op).  Rounding - not truncating - matters: truncated
// parts all err towards zero, which biases every product the same way; such an error does not
// average out over the reduction (measured: gradients 100x further from fp64 than with IEEE fp32),
// whereas the rounded split is zero-mean and ends up at fp32 level.
__device__ __forceinline__ float round_tf32(float v) {
    return __uint_as_float((__float_as_uint(v) + 0x1000u) & 0xffffe000u);
}
__device__ __forceinline__ void split_rn(float v, float& hi, float& lo) {
    hi = round_tf32(v);
    lo = round_tf32(v - hi);
}

__global__ void __launch_bounds__(256) split_weight_kernel(const float* __restrict__ w, long long ldw,
                                                           int trans, int n_out, int k_in, int n_pad,
                                                           int k_pad, float* __restrict__ hi,
                                                           float* __restrict__ lo) {
    const long long total = (long long)n_pad * k_pad;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        // trans: consecutive threads walk n (the contiguous axis of w[k_in, n_out]) for coalescing
        const int n = trans ? (int)(i % n_pad) : (int)(i / k_pad);
        const int k = trans ? (int)(i / n_pad) : (int)(i % k_pad);
        float v = 0.f;
        if (n < n_out && k < k_in) v = trans ? __ldg(w + (long long)k * ldw + n) : __ldg(w + (long long)n * ldw + k);
        float h, l;
        split_rn(v, h, l);
        hi[(size_t)n * k_pad + k] = h;
        lo[(size_t)n * k_pad + k] = l;
    }
}

template <int BN>
__global__ void __launch_bounds__(LT_THREADS)
linear_tc_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_whi,
                 const __grid_constant__ CUtensorMap map_wlo, const float* __restrict__ bias, int relu,
                 float* __restrict__ y, long long ldy, long long rows, int n_out, int k_blocks,
                 int n_stages) {
    constexpr int B_BYTES = BN * LT_KB * 4;
    constexpr int STAGE_BYTES = 2 * LT_A_BYTES + 2 * B_BYTES;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (size_t)n_stages * STAGE_BYTES);
    uint64_t* full = bars;                          // TMA bytes landed
    uint64_t* split_done = bars + LT_MAX_STAGES;    // X tile split into hi / lo (4 warp arrivals)
    uint64_t* empty = bars + 2 * LT_MAX_STAGES;     // MMAs of the stage retired
    uint64_t* acc_full = bars + 3 * LT_MAX_STAGES;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_full + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n0 = blockIdx.x * BN;
    const long long m0 = (long long)blockIdx.y * LT_BM;

    if (warp == 0 && lane == 0) {
        for (int s = 0; s < n_stages; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&split_done[s], 4);
            mbar_init(&empty[s], 1);
        }
        mbar_init(acc_full, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                         smem_u32(tmem_slot)), "r"(2 * LT_ACC * BN));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_d = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            // ================= TMA producer =================
            int s = 0; uint32_t phase = 0;
            for (int kb = 0; kb < k_blocks; ++kb) {
                mbar_wait(&empty[s], phase ^ 1u);
                uint8_t* st = smem + (size_t)s * STAGE_BYTES;
                mbar_expect_tx(&full[s], (uint32_t)(LT_A_BYTES + 2 * B_BYTES));
                tma_load_2d(st, &map_x, &full[s], kb * LT_KB, (int)m0);
                tma_load_2d(st + 2 * LT_A_BYTES, &map_whi, &full[s], kb * LT_KB, n0);
                tma_load_2d(st + 2 * LT_A_BYTES + B_BYTES, &map_wlo, &full[s], kb * LT_KB, n0);
                if (++s == n_stages) { s = 0; phase ^= 1u; }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            // ================= MMA issuer =================
            // kind::tf32, D = f32, A = B = tf32, both K-major, M = 128.  Per k8 step two MMAs:
            //   X_hi x [W_hi ; W_lo]  (N = 2 BN: the two W tiles are adjacent in shared memory, so one
            //                          instruction reads X_hi once and writes the column pair [H | L])
            //   X_lo x W_hi           (N = BN, accumulated into L)
            // The tensor core TRUNCATES when it adds an instruction's result to the accumulator, so
            // the error grows with the number of instructions per accumulator: the small lo terms
            // get their own accumulator L, and successive k8 steps alternate between LT_ACC column
            // pairs, which the epilogue adds in round-to-nearest fp32.
            constexpr uint32_t idesc_wide = (1u << 4) | (2u << 7) | (2u << 10) |
                                            ((uint32_t)((2 * BN) >> 3) << 17) | ((uint32_t)(LT_BM >> 4) << 24);
            constexpr uint32_t idesc_lo = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(BN >> 3) << 17) |
                                          ((uint32_t)(LT_BM >> 4) << 24);
            int s = 0; uint32_t phase = 0; int step = 0;
            for (int kb = 0; kb < k_blocks; ++kb) {
                mbar_wait(&full[s], phase);          // W tiles (written by TMA)
                mbar_wait(&split_done[s], phase);    // X tiles (rewritten by warps 2-5)
                tcgen05_fence_after();
                const uint32_t st = smem_u32(smem + (size_t)s * STAGE_BYTES);
                const uint64_t d_ahi = umma_desc_sw128(st), d_alo = umma_desc_sw128(st + LT_A_BYTES);
                const uint64_t d_b = umma_desc_sw128(st + 2 * LT_A_BYTES);   // W_hi rows, then W_lo rows
#pragma unroll
                for (int ks = 0; ks < LT_KB / 8; ++ks, ++step) {   // +32 bytes per K=8 step inside the swizzled row
                    const uint64_t o = (uint64_t)(2 * ks);
                    const uint32_t d = tmem_d + (uint32_t)((step % LT_ACC) * 2 * BN);
                    tcgen05_mma_tf32(d, d_ahi + o, d_b + o, idesc_wide, (uint32_t)(step >= LT_ACC));
                    tcgen05_mma_tf32(d + BN, d_alo + o, d_b + o, idesc_lo, 1u);
                }
                tcgen05_commit(&empty[s]);
                if (++s == n_stages) { s = 0; phase ^= 1u; }
            }
            tcgen05_commit(acc_full);
        }
    } else {
        // ================= split X in place, then epilogue =================
        const int t = threadIdx.x - 64;               // 0..127
        int s = 0; uint32_t phase = 0;
        for (int kb = 0; kb < k_blocks; ++kb) {
            mbar_wait(&full[s], phase);
            float4* a_hi = reinterpret_cast<float4*>(smem + (size_t)s * STAGE_BYTES);
            float4* a_lo = a_hi + LT_A_BYTES / 16;
#pragma unroll
            for (int i = 0; i < LT_A_BYTES / 16 / 128; ++i) {
                const float4 v = a_hi[t + i * 128];
                float4 h, l;
                split_rn(v.x, h.x, l.x); split_rn(v.y, h.y, l.y);
                split_rn(v.z, h.z, l.z); split_rn(v.w, h.w, l.w);
                a_hi[t + i * 128] = h;
                a_lo[t + i * 128] = l;
            }
            fence_proxy_async_smem();     // st.shared above -> UMMA operand reads
            __syncwarp();
            if (lane == 0) mbar_arrive(&split_done[s]);
            if (++s == n_stages) { s = 0; phase ^= 1u; }
        }
        mbar_wait(acc_full, 0);
        tcgen05_fence_after();
        const int quarter = warp & 3;                  // TMEM lanes this warp may read
        const long long row = m0 + quarter * 32 + lane;
        const bool vec = (ldy % 4 == 0) && ((reinterpret_cast<uintptr_t>(y) & 15) == 0);
#pragma unroll 1
        for (int cb = 0; cb < BN; cb += 32) {
            if (n0 + cb >= n_out) break;               // warp-uniform
            float r[32], q[32];
            const uint32_t lane_base = tmem_d + ((uint32_t)(quarter * 32) << 16) + (uint32_t)cb;
            tmem_ld_32x32b_x32(lane_base, r);                                   // H_0
#pragma unroll
            for (int p = 1; p < LT_ACC; ++p) {                                  // + H_1 ...
                tmem_ld_32x32b_x32(lane_base + (uint32_t)(p * 2 * BN), q);
#pragma unroll
                for (int j = 0; j < 32; ++j) r[j] += q[j];
            }
            {                                                                   // + (L_0 + L_1 ...)
                float l[32];
                tmem_ld_32x32b_x32(lane_base + (uint32_t)BN, l);
#pragma unroll
                for (int p = 1; p < LT_ACC; ++p) {
                    tmem_ld_32x32b_x32(lane_base + (uint32_t)(p * 2 * BN + BN), q);
#pragma unroll
                    for (int j = 0; j < 32; ++j) l[j] += q[j];
                }
#pragma unroll
                for (int j = 0; j < 32; ++j) r[j] += l[j];
            }
            const int c0 = n0 + cb;
            if (bias) {
#pragma unroll
                for (int j = 0; j < 32; ++j) r[j] += (c0 + j < n_out) ? __ldg(bias + c0 + j) : 0.f;
            }
            if (relu) {
#pragma unroll
                for (int j = 0; j < 32; ++j) r[j] = fmaxf(r[j], 0.f);
            }
            if (row < rows) {
                float* dst = y + row * ldy + c0;
                if (vec && c0 + 32 <= n_out) {
#pragma unroll
                    for (int j = 0; j < 32; j += 4)
                        *reinterpret_cast<float4*>(dst + j) = make_float4(r[j], r[j + 1], r[j + 2], r[j + 3]);
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        if (c0 + j < n_out) dst[j] = r[j];
                }
            }
        }
    }
    tcgen05_fence_before();
    __syncthreads();
    if (warp == 1) {
        tcgen05_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"(2 * LT_ACC * BN));
    }
}

typedef CUresult (*EncodeTiledFnL)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                   CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                   CUtensorMapFloatOOBfill);

static EncodeTiledFnL lt_encode_tiled() {
    static EncodeTiledFnL fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFnL>(p);
    }
    return fn;
}

// fp32 [rows x cols] row-major with pitch ld (floats); box = 32 columns x box_rows rows
static int lt_make_map(CUtensorMap* map, const void* base, long long rows, long long cols, long long ld,
                       int box_rows) {
    EncodeTiledFnL enc = lt_encode_tiled();
    if (!enc) { set_error("linear_tc: cuTensorMapEncodeTiled not available from the driver"); return DGMC_ERR_CUDA; }
    const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)ld * 4};
    const cuuint32_t box[2] = {(cuuint32_t)LT_KB, (cuuint32_t)box_rows};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(base), dims, strides,
                           box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                           CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("linear_tc: cuTensorMapEncodeTiled failed (%d)", (int)r); return DGMC_ERR_CUDA; }
    return DGMC_OK;
}

static int lt_bn(int n_out) { return n_out <= 64 ? 64 : 128; }
static int lt_round(int v, int m) { return (v + m - 1) / m * m; }

template <int BN>
static int lt_launch(const CUtensorMap& mx, const CUtensorMap& mh, const CUtensorMap& ml, const float* bias,
                     int relu, float* y, long long ldy, long long rows, int n_out, int k_blocks, int n_pad,
                     cudaStream_t stream) {
    constexpr int STAGE = 2 * LT_A_BYTES + 2 * BN * LT_KB * 4;
    // BN = 64: 2 stages of 48 KB -> two CTAs per SM interleave; BN = 128: 3 stages of 64 KB
    int n_stages = BN == 64 ? 2 : 3;
    if (n_stages > k_blocks) n_stages = k_blocks;
    const size_t smem = (size_t)n_stages * STAGE + (3 * LT_MAX_STAGES + 2) * 8 + 1024;
    static size_t configured = 0;
    if (smem > configured) {
        if (cudaFuncSetAttribute(linear_tc_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)smem) != cudaSuccess) {
            set_error("linear_tc: cannot set %zu bytes of dynamic shared memory", smem);
            return DGMC_ERR_CUDA;
        }
        configured = smem;
    }
    dim3 grid(n_pad / BN, (unsigned)((rows + LT_BM - 1) / LT_BM));
    linear_tc_kernel<BN><<<grid, LT_THREADS, smem, stream>>>(mx, mh, ml, bias, relu, y, ldy, rows, n_out,
                                                            k_blocks, n_stages);
    return DGMC_OK;
}

}  // namespace tc
}  // namespace dgmc

using namespace dgmc;

extern "C" int dgmc_linear_tc_supported(int64_t rows, int n_out, int k_in, int64_t ldx, const float* x) {
    // TMA: 16-byte aligned base and pitch; grid.y limit
    return rows >= 1 && rows <= (int64_t)65535 * tc::LT_BM && n_out >= 1 && k_in >= 1 && ldx >= k_in &&
           ldx % 4 == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0;
}

extern "C" size_t dgmc_linear_tc_workspace(int n_out, int k_in) {
    if (n_out < 1 || k_in < 1) return 0;
    const size_t n_pad = tc::lt_round(n_out, tc::lt_bn(n_out)), k_pad = tc::lt_round(k_in, tc::LT_KB);
    return 2 * n_pad * k_pad * sizeof(float);
}

extern "C" int dgmc_linear_tc_split_weight(const float* w, int64_t ldw, int transposed, int n_out, int k_in,
                                           void* workspace, size_t workspace_bytes, dgmc_stream_t stream) {
    DGMC_REQUIRE(w && workspace && n_out >= 1 && k_in >= 1 && ldw >= (transposed ? n_out : k_in),
                 "linear_tc_split_weight: bad arguments (n_out=%d k_in=%d)", n_out, k_in);
    DGMC_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 127) == 0, "linear_tc_split_weight: workspace must be 128-byte aligned");
    if (workspace_bytes < dgmc_linear_tc_workspace(n_out, k_in)) {
        set_error("linear_tc_split_weight: workspace too small");
        return DGMC_ERR_WORKSPACE;
    }
    const int n_pad = tc::lt_round(n_out, tc::lt_bn(n_out)), k_pad = tc::lt_round(k_in, tc::LT_KB);
    float* hi = static_cast<float*>(workspace);
    float* lo = hi + (size_t)n_pad * k_pad;
    const long long total = (long long)n_pad * k_pad;
    int blocks = (int)((total + 255) / 256);
    if (blocks > 8 * sm_count()) blocks = 8 * sm_count();
    tc::split_weight_kernel<<<blocks, 256, 0, stream>>>(w, ldw, transposed, n_out, k_in, n_pad, k_pad, hi, lo);
    return check_launch("linear_tc_split_weight");
}

extern "C" int dgmc_linear_tc(const float* x, int64_t ldx, const void* split_weight, const float* bias,
                              int relu, float* y, int64_t ldy, int64_t rows, int n_out, int k_in,
                              dgmc_stream_t stream) {
    DGMC_REQUIRE(x && split_weight && y && ldy >= n_out, "linear_tc: bad arguments");
    if (!dgmc_linear_tc_supported(rows, n_out, k_in, ldx, x)) {
        set_error("linear_tc: unsupported operand (rows=%lld n_out=%d k_in=%d ldx=%lld; x needs a 16-byte "
                  "aligned base and pitch)", (long long)rows, n_out, k_in, (long long)ldx);
        return DGMC_ERR_UNSUPPORTED;
    }
    const int bn = tc::lt_bn(n_out);
    const int n_pad = tc::lt_round(n_out, bn), k_pad = tc::lt_round(k_in, tc::LT_KB);
    const float* hi = static_cast<const float*>(split_weight);
    const float* lo = hi + (size_t)n_pad * k_pad;
    CUtensorMap mx, mh, ml;
    if (int e = tc::lt_make_map(&mx, x, rows, k_in, ldx, tc::LT_BM)) return e;
    if (int e = tc::lt_make_map(&mh, hi, n_pad, k_pad, k_pad, bn)) return e;
    if (int e = tc::lt_make_map(&ml, lo, n_pad, k_pad, k_pad, bn)) return e;
    const int k_blocks = k_pad / tc::LT_KB;
    int e;
    if (bn == 64) e = tc::lt_launch<64>(mx, mh, ml, bias, relu, y, ldy, rows, n_out, k_blocks, n_pad, stream);
    else          e = tc::lt_launch<128>(mx, mh, ml, bias, relu, y, ldy, rows, n_out, k_blocks, n_pad, stream);
    if (e) return e;
    return check_launch("linear_tc");
}
