// Weight gradient dW[m, n] = sum_k G[k, m] X[k, n] (k = nodes) on the 5th-gen tensor cores with
// fp32-level accuracy ("3xTF32", see linear_tc.cu: hi / lo parts rounded to tf32, X_hi G_hi +
// X_hi G_lo + X_lo G_hi, separate accumulators for the hi and the lo terms).
//
// The reduction runs over the ROW index of both operands, so both are "MN-major" for the tensor core
// (channels contiguous, k strided).  No transpose is needed: a TMA box of 32 channels x 32 k rows
// loaded with CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B IS the canonical MN-major layout for 32-bit operands
// (SWIZZLE_128B_BASE32B: atoms of 4 k rows x 128 bytes); the UMMA descriptor takes LBO = 4 KB (next
// 32-channel box), SBO = 512 B (next 4 k rows), and advances 1 KB per K=8 instruction.
// Grid = (128-wide tiles of input channels) x (K splits), one CTA per SM; per CTA a 4-stage ring:
//   warp 0      TMA producer: X[32 k x 128] and G[32 k x BN] as 32 x 32 boxes
//   warps 2-9   split both tiles into hi (in place) and lo (position-identical second buffer), sum the
//               columns of G for the bias gradient, fence.proxy.async, arrive
//   warp 1      per 8-k step: X_hi x [G_hi ; G_lo] (N = 2 BN) and X_lo x G_hi (N = BN), tcgen05.commit
//   warps 2-5   epilogue: accumulator rows -> partial[split][m][n] (dgmc_wgrad_reduce adds the splits)
#include <cstring>
#include "common.cuh"
#include "tc_common.cuh"
#include "../../include/dgmc_b200.h"

namespace dgmc {
namespace tc {

constexpr int WT_KB = 32;                  // k rows (nodes) per stage
constexpr int WT_TN = 128;                 // input channels per CTA (UMMA M)
constexpr int WT_THREADS = 320;            // TMA producer, MMA issuer, 8 split warps (4 of them also the epilogue)
constexpr int WT_SPLITTERS = 256;
constexpr int WT_A_BYTES = WT_TN * WT_KB * 4;    // 16 KB
constexpr int WT_MAX_M = 128;              // output channels (UMMA N), padded to 64 or 128
constexpr int WT_CHUNK = 32 * WT_KB * 4;   // one TMA box: 32 channels x 32 k rows = 4 KB
constexpr int WT_ACC = 2;                  // accumulator pairs [H | L] (see linear_tc.cu)
constexpr int WT_MAX_STAGES = 4;

__device__ __forceinline__ float wt_round_tf32(float v) {
    return __uint_as_float((__float_as_uint(v) + 0x1000u) & 0xffffe000u);
}
__device__ __forceinline__ void wt_split(float v, float& hi, float& lo) {   // see linear_tc.cu: round, not truncate
    hi = wt_round_tf32(v);
    lo = wt_round_tf32(v - hi);
}
__device__ __forceinline__ void wt_split4(const float4& v, float4& h, float4& l) {
    wt_split(v.x, h.x, l.x); wt_split(v.y, h.y, l.y); wt_split(v.z, h.z, l.z); wt_split(v.w, h.w, l.w);
}

// UMMA shared-memory descriptor for an MN-major tf32 operand.  32-bit MN-major operands only exist in
// the SWIZZLE_128B_BASE32B layout (layout type 1; TMA: CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B): atoms of
// [32 channels (128 bytes) x 4 k rows], 32-byte pieces of a row XOR-ed with (k & 3).  LBO = distance
// between consecutive 32-channel chunks, SBO = distance between consecutive 4-k groups.
__device__ __forceinline__ uint64_t umma_desc_mn_sw128(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)1 << 61;
    return d;
}

template <int BN>   // padded output channels
__global__ void __launch_bounds__(WT_THREADS, 1)
wgrad_tc_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_g,
                long long K, int M, int N, int splits, int n_stages, long long bias_rows,
                float* __restrict__ partial, float* __restrict__ partial_b) {
    constexpr int B_BYTES = BN * WT_KB * 4;
    constexpr int STAGE_BYTES = 2 * WT_A_BYTES + 2 * B_BYTES;
    constexpr int QUADS = BN / 4;                 // 4-channel groups of G per k row
    constexpr int ROWS_PER_PASS = WT_SPLITTERS / QUADS;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    float4* bias_red = reinterpret_cast<float4*>(smem + (size_t)n_stages * STAGE_BYTES);   // [256]
    uint64_t* bars = reinterpret_cast<uint64_t*>(bias_red + WT_SPLITTERS);
    uint64_t* full = bars;
    uint64_t* split_done = bars + WT_MAX_STAGES;
    uint64_t* empty = bars + 2 * WT_MAX_STAGES;
    uint64_t* acc_full = bars + 3 * WT_MAX_STAGES;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_full + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n0 = blockIdx.x * WT_TN, split = blockIdx.y;
    const long long per = ((K + splits - 1) / splits + WT_KB - 1) / WT_KB * WT_KB;
    const long long k_begin = per * split;
    const long long k_end = k_begin + per < K ? k_begin + per : K;
    const int n_chunks = k_end > k_begin ? (int)((k_end - k_begin + WT_KB - 1) / WT_KB) : 0;
    float* out = partial + (size_t)split * M * N;

    if (n_chunks == 0) {                      // empty K slice (only the last splits): a zero partial
        for (int i = threadIdx.x; i < M * WT_TN; i += WT_THREADS) {
            const int m = i / WT_TN, n = n0 + i % WT_TN;
            if (n < N) out[(size_t)m * N + n] = 0.f;
        }
        if (partial_b && blockIdx.x == 0)
            for (int m = threadIdx.x; m < M; m += WT_THREADS) partial_b[(size_t)split * M + m] = 0.f;
        return;
    }

    if (warp == 0 && lane == 0) {
        for (int s = 0; s < n_stages; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&split_done[s], WT_SPLITTERS / 32);
            mbar_init(&empty[s], 1);
        }
        mbar_init(acc_full, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                         smem_u32(tmem_slot)), "r"(2 * WT_ACC * BN));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_d = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            // ===== TMA producer: X[32 k x 128 channels] and G[32 k x BN channels] as 32 x 32 boxes =====
            int s = 0; uint32_t phase = 0;
            for (int c = 0; c < n_chunks; ++c) {
                mbar_wait(&empty[s], phase ^ 1u);
                uint8_t* st = smem + (size_t)s * STAGE_BYTES;
                const int k0 = (int)(k_begin + (long long)c * WT_KB);
                mbar_expect_tx(&full[s], (uint32_t)(WT_A_BYTES + B_BYTES));
#pragma unroll
                for (int q = 0; q < WT_TN / 32; ++q)
                    tma_load_2d(st + q * WT_CHUNK, &map_x, &full[s], n0 + 32 * q, k0);
#pragma unroll
                for (int q = 0; q < BN / 32; ++q)
                    tma_load_2d(st + 2 * WT_A_BYTES + q * WT_CHUNK, &map_g, &full[s], 32 * q, k0);
                if (++s == n_stages) { s = 0; phase ^= 1u; }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            // ===== MMA issuer: both operands MN-major (channels contiguous, k strided) =====
            constexpr uint32_t idesc_wide = (1u << 4) | (2u << 7) | (2u << 10) | (1u << 15) | (1u << 16) |
                                            ((uint32_t)((2 * BN) >> 3) << 17) | ((uint32_t)(WT_TN >> 4) << 24);
            constexpr uint32_t idesc_lo = (1u << 4) | (2u << 7) | (2u << 10) | (1u << 15) | (1u << 16) |
                                          ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(WT_TN >> 4) << 24);
            int s = 0; uint32_t phase = 0; int step = 0;
            for (int c = 0; c < n_chunks; ++c) {
                mbar_wait(&full[s], phase);
                mbar_wait(&split_done[s], phase);
                tcgen05_fence_after();
                const uint32_t st = smem_u32(smem + (size_t)s * STAGE_BYTES);
#pragma unroll
                for (int ks = 0; ks < WT_KB / 8; ++ks, ++step) {   // one 8-k group (1024 bytes) per instruction
                    const uint64_t d_ahi = umma_desc_mn_sw128(st + ks * 1024, WT_CHUNK, 512);
                    const uint64_t d_alo = umma_desc_mn_sw128(st + WT_A_BYTES + ks * 1024, WT_CHUNK, 512);
                    const uint64_t d_b = umma_desc_mn_sw128(st + 2 * WT_A_BYTES + ks * 1024, WT_CHUNK, 512);
                    const uint32_t d = tmem_d + (uint32_t)((step % WT_ACC) * 2 * BN);
                    tcgen05_mma_tf32(d, d_ahi, d_b, idesc_wide, (uint32_t)(step >= WT_ACC));   // X_hi x [G_hi ; G_lo]
                    tcgen05_mma_tf32(d + BN, d_alo, d_b, idesc_lo, 1u);                        // X_lo x G_hi
                }
                tcgen05_commit(&empty[s]);
                if (++s == n_stages) { s = 0; phase ^= 1u; }
            }
            tcgen05_commit(acc_full);
        }
    } else {
        // ===== split both tiles into hi (in place) / lo; column sums of G; then the epilogue =====
        const int t = threadIdx.x - 64;               // 0..255
        const int quad = t % QUADS, krow0 = t / QUADS;
        const uint32_t g_off = (uint32_t)((quad >> 3) * WT_CHUNK);   // 32-channel chunk of this quad
        float4 bsum = make_float4(0.f, 0.f, 0.f, 0.f);
        int s = 0; uint32_t phase = 0;
        for (int c = 0; c < n_chunks; ++c) {
            mbar_wait(&full[s], phase);
            uint8_t* st = smem + (size_t)s * STAGE_BYTES;
            float4* a_hi = reinterpret_cast<float4*>(st);
            float4* a_lo = a_hi + WT_A_BYTES / 16;
#pragma unroll
            for (int i = 0; i < WT_A_BYTES / 16 / WT_SPLITTERS; ++i) {
                float4 h, l;
                wt_split4(a_hi[t + i * WT_SPLITTERS], h, l);
                a_hi[t + i * WT_SPLITTERS] = h;
                a_lo[t + i * WT_SPLITTERS] = l;
            }
            uint8_t* b_hi = st + 2 * WT_A_BYTES;
#pragma unroll
            for (int i = 0; i < WT_KB / ROWS_PER_PASS; ++i) {
                const int k = krow0 + i * ROWS_PER_PASS;
                const int c16 = quad & 7;          // 16-byte piece of the 128-byte row; its 32-byte pair is swizzled
                const uint32_t o = g_off + (uint32_t)(k * 128) +
                                   (uint32_t)((((c16 >> 1) ^ (k & 3)) << 5) | ((c16 & 1) << 4));
                const float4 v = *reinterpret_cast<const float4*>(b_hi + o);
                if (k_begin + (long long)c * WT_KB + k < bias_rows) {   // the bias may belong to the leading rows only
                    bsum.x += v.x; bsum.y += v.y; bsum.z += v.z; bsum.w += v.w;
                }
                float4 h, l;
                wt_split4(v, h, l);
                *reinterpret_cast<float4*>(b_hi + o) = h;
                *reinterpret_cast<float4*>(b_hi + B_BYTES + o) = l;
            }
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) mbar_arrive(&split_done[s]);
            if (++s == n_stages) { s = 0; phase ^= 1u; }
        }
        // ---- column sums of G (db): fixed-order reduction over the 128 / QUADS threads per quad ----
        if (partial_b && blockIdx.x == 0) {
            bias_red[t] = bsum;
            asm volatile("bar.sync 1, 256;" ::: "memory");
            if (t < QUADS) {
                float4 acc4 = bias_red[t];
                for (int i = 1; i < ROWS_PER_PASS; ++i) {
                    const float4 v = bias_red[t + i * QUADS];
                    acc4.x += v.x; acc4.y += v.y; acc4.z += v.z; acc4.w += v.w;
                }
                const float a4[4] = {acc4.x, acc4.y, acc4.z, acc4.w};
                for (int u = 0; u < 4; ++u)
                    if (4 * t + u < M) partial_b[(size_t)split * M + 4 * t + u] = a4[u];
            }
        }
        // ---- partial tile: D[n, m] -> partial[split][m][n0 + n] (lanes = consecutive n: coalesced) ----
        mbar_wait(acc_full, 0);
        tcgen05_fence_after();
        const int quarter = warp & 3;
        const int n = n0 + quarter * 32 + lane;
#pragma unroll 1
        for (int cb = 0; cb < BN; cb += 32) {
            if (cb >= M || warp >= 6) break;      // warps 2-5 cover the four TMEM lane quarters
            const uint32_t lane_base = tmem_d + ((uint32_t)(quarter * 32) << 16) + (uint32_t)cb;
            float r[32], l[32], q[32], m2[32];
            tmem_ld_32x32b_x32_pair(lane_base, lane_base + (uint32_t)BN, r, l);
            tmem_ld_32x32b_x32_pair(lane_base + (uint32_t)(2 * BN), lane_base + (uint32_t)(3 * BN), q, m2);
#pragma unroll
            for (int j = 0; j < 32; ++j) r[j] = (r[j] + q[j]) + (l[j] + m2[j]);
            if (n < N) {
#pragma unroll
                for (int j = 0; j < 32; ++j)
                    if (cb + j < M) out[(size_t)(cb + j) * N + n] = r[j];
            }
        }
    }
    tcgen05_fence_before();
    __syncthreads();
    if (warp == 1) {
        tcgen05_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"(2 * WT_ACC * BN));
    }
}

typedef CUresult (*EncodeTiledFnW)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                   CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                   CUtensorMapFloatOOBfill);

static EncodeTiledFnW wt_encode_tiled() {
    static EncodeTiledFnW fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFnW>(p);
    }
    return fn;
}

// fp32 [rows(k) x cols(channels)] row-major, pitch ld floats; box = 32 channels x 32 k rows
static int wt_make_map(CUtensorMap* map, const void* base, long long rows, long long cols, long long ld) {
    ensure_context();
    EncodeTiledFnW enc = wt_encode_tiled();
    if (!enc) { set_error("wgrad_tc: cuTensorMapEncodeTiled not available from the driver"); return DGMC_ERR_CUDA; }
    memset(map, 0, sizeof(*map));
    const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)ld * 4};
    const cuuint32_t box[2] = {32, (cuuint32_t)WT_KB};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(base), dims, strides,
                           box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B,
                           CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("wgrad_tc: cuTensorMapEncodeTiled failed (%d)", (int)r); return DGMC_ERR_CUDA; }
    return DGMC_OK;
}

static int wt_stages(int bn) { return bn == 64 ? 4 : 3; }   // 4 x 48 KB / 3 x 64 KB
static size_t wt_smem(int bn) {
    return (size_t)wt_stages(bn) * (2 * WT_A_BYTES + 2 * bn * WT_KB * 4) + WT_SPLITTERS * 16 + (3 * WT_MAX_STAGES + 2) * 8 + 1024;
}

template <int BN>
static int wt_launch(const CUtensorMap& mx, const CUtensorMap& mg, long long K, int M, int N, int splits,
                     long long bias_rows, float* partial, float* partial_b, cudaStream_t stream) {
    const size_t smem = wt_smem(BN);
    static DeviceOnce configured;
    if (!configured.done()) {
        if (cudaFuncSetAttribute(wgrad_tc_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)smem) != cudaSuccess) {
            set_error("wgrad_tc_partial: cannot set %zu bytes of dynamic shared memory", smem);
            return DGMC_ERR_CUDA;
        }
        configured.mark();
    }
    dim3 grid((N + WT_TN - 1) / WT_TN, splits);
    wgrad_tc_kernel<BN><<<grid, WT_THREADS, smem, stream>>>(mx, mg, K, M, N, splits, wt_stages(BN), bias_rows, partial, partial_b);
    return DGMC_OK;
}

}  // namespace tc
}  // namespace dgmc

using namespace dgmc;

static int wgrad_tc_splits_impl(long long K, int N) {
    const int tiles = (N + tc::WT_TN - 1) / tc::WT_TN;
    int splits = sm_count() / tiles;                        // one CTA per SM, one wave
    const long long max_by_k = (K + 8 * tc::WT_KB - 1) / (8 * tc::WT_KB);   // >= 256 rows per split
    if (splits > max_by_k) splits = (int)max_by_k;
    if (splits < 1) splits = 1;
    if (splits > 1024) splits = 1024;
    return splits;
}

extern "C" int dgmc_wgrad_tc_supported(int m, int n, int64_t ldg, int64_t ldx, const float* g, const float* x) {
    // TMA: 16-byte aligned bases and pitches
    return m >= 1 && m <= tc::WT_MAX_M && n >= 1 && ldg >= m && ldx >= n && ldg % 4 == 0 && ldx % 4 == 0 &&
           (reinterpret_cast<uintptr_t>(g) & 15) == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0;
}

extern "C" int dgmc_wgrad_tc_splits(int64_t k, int m, int n) {
    (void)m;
    return wgrad_tc_splits_impl(k, n);
}

static int wgrad_tc_partial_impl(const float* g, int64_t ldg, const float* x, int64_t ldx, int64_t k, int m,
                                 int n, int64_t bias_rows, float* partial, float* partial_b,
                                 dgmc_stream_t stream) {
    DGMC_REQUIRE(g && x && partial && k >= 1 && k < ((int64_t)1 << 31), "wgrad_tc_partial: bad arguments");
    if (!dgmc_wgrad_tc_supported(m, n, ldg, ldx, g, x)) {
        set_error("wgrad_tc_partial: unsupported operands (m=%d of 1..%d output channels; g and x need "
                  "16-byte aligned bases and pitches)", m, tc::WT_MAX_M);
        return DGMC_ERR_UNSUPPORTED;
    }
    CUtensorMap mx, mg;
    if (int e = tc::wt_make_map(&mx, x, k, n, ldx)) return e;
    if (int e = tc::wt_make_map(&mg, g, k, m, ldg)) return e;
    const int splits = wgrad_tc_splits_impl(k, n);
    int e;
    if (m <= 64) e = tc::wt_launch<64>(mx, mg, k, m, n, splits, bias_rows, partial, partial_b, stream);
    else         e = tc::wt_launch<128>(mx, mg, k, m, n, splits, bias_rows, partial, partial_b, stream);
    if (e) return e;
    return check_launch("wgrad_tc_partial");
}

extern "C" int dgmc_wgrad_tc_partial(const float* g, int64_t ldg, const float* x, int64_t ldx, int64_t k,
                                     int m, int n, float* partial, float* partial_b, dgmc_stream_t stream) {
    return wgrad_tc_partial_impl(g, ldg, x, ldx, k, m, n, k, partial, partial_b, stream);
}

extern "C" int dgmc_wgrad_tc_partial_rows(const float* g, int64_t ldg, const float* x, int64_t ldx, int64_t k,
                                          int m, int n, int64_t bias_rows, float* partial, float* partial_b,
                                          dgmc_stream_t stream) {
    DGMC_REQUIRE(bias_rows >= 0 && bias_rows <= k, "wgrad_tc_partial_rows: bias_rows out of range");
    return wgrad_tc_partial_impl(g, ldg, x, ldx, k, m, n, bias_rows, partial, partial_b, stream);
}
