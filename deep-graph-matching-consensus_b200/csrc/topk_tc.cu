// K1: streaming top-k of <h_s[i], h_t[j]> on the 5th-gen tensor cores (replaces DGMC.__top_k__,
// dgmc.py:85-94: KeOps argKmin, else broadcast-mul-sum + topk).  The [N_s, N_t] score matrix
// never reaches HBM.
//
//   cast     fp32 -> bf16 copies of h_s / h_t, C zero-padded to a multiple of 64
//   main     per CTA: a block of 128*MH rows of h_s resident in shared memory (TMA, SWIZZLE_128B);
//            128-column tiles of h_t streamed through an NSTAGE ring (TMA);
//            one thread issues tcgen05.mma.cta_group::1.kind::f16 (M=128, N=128, K=16) into
//            double-buffered TMEM accumulators; 4*MH epilogue warps read them back with
//            tcgen05.ld 32x32b.x32 — one thread owns one score row — and keep that row's KP best
//            (score, column) pairs sorted in registers.
//            Work split: stream-K.  The (row-block, column-tile) grid is cut into equal contiguous
//            runs, one per CTA; whenever a run leaves a row block the CTA writes its partial lists
//            to the caller's scratch.
//   merge    one warp per row: unite the <= P partial lists, keep the KP best by tensor-core score,
//            re-score those in exact fp32 (bf16 alone flips ~4 % of top-10 sets; top-(k+6) + fp32
//            rescoring flipped none) and emit the k best, descending, ties by column.
//
// Roofline: tensor pipe.  2*N_s*N_t*C flop per call (0.194 TFLOP at the DBP15K shape).
#include <cuda.h>
#include <cuda_bf16.h>
#include <float.h>
#include "common.cuh"
#include "tc_common.cuh"
#include "../../include/dgmc_b200.h"

namespace dgmc {
namespace tc {

constexpr int TILE_M = 128;        // rows per MMA (TMEM lanes)
constexpr int TILE_N = 128;        // columns of h_t per tile
constexpr int KCHUNK = 64;         // bf16 elements per 128-byte swizzled smem row
constexpr int CHUNK_BYTES = TILE_M * KCHUNK * 2;  // 16 KB: one [128 x 64] bf16 box
constexpr int NSTAGE = 5;          // ring of h_t boxes
constexpr int MAX_A_CHUNKS = 8;    // MH * C_pad / 64 <= 8  (128 KB of resident h_s)
constexpr int TMEM_COLS = 512;

// instruction descriptor, kind::f16: D=f32, A=B=bf16, both K-major, N=128, M=128
constexpr uint32_t IDESC_BF16_128x128 =
    (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(TILE_N >> 3) << 17) | ((uint32_t)(TILE_M >> 4) << 24);

// ---- fp32 -> bf16 with zero padding of the channel dimension -----------------------------------
__global__ void __launch_bounds__(256) cast_bf16_kernel(const float* __restrict__ src, long long ld,
                                                        __nv_bfloat16* __restrict__ dst, long long rows,
                                                        int c, int c_pad) {
    const long long total = rows * (c_pad / 2);
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += stride) {
        const long long row = t / (c_pad / 2);
        const int q = (int)(t % (c_pad / 2)) * 2;
        const float a = q < c ? __ldg(src + row * ld + q) : 0.f;
        const float b = q + 1 < c ? __ldg(src + row * ld + q + 1) : 0.f;
        reinterpret_cast<__nv_bfloat162*>(dst)[t] = __floats2bfloat162_rn(a, b);
    }
}

// ---- stream-K bookkeeping shared by the main and the merge kernel -----------------------------------
struct Split {
    int n_rb, n_ct, grid;       // row blocks, column tiles, CTAs
    long long tiles;            // n_rb * n_ct
    int slots;                  // partial lists per row (upper bound)
    __host__ __device__ long long begin(int cta) const { return tiles * cta / grid; }
    __host__ __device__ int owner(long long tile) const { return (int)(((tile + 1) * grid - 1) / tiles); }
    __host__ __device__ int first_cta(int rb) const { return owner((long long)rb * n_ct); }
    __host__ __device__ int last_cta(int rb) const { return owner((long long)(rb + 1) * n_ct - 1); }
};

// register-resident sorted list (descending) of the KP best (score, column) of one row
template <int KP>
struct RowList {
    float v[KP];
    int id[KP];
    __device__ __forceinline__ void reset() {
#pragma unroll
        for (int j = 0; j < KP; ++j) { v[j] = -INFINITY; id[j] = -1; }
    }
    __device__ __forceinline__ float threshold() const { return v[KP - 1]; }
    __device__ __forceinline__ void insert(float s, int col) {
#pragma unroll
        for (int j = KP - 1; j > 0; --j) {
            const bool shift = s > v[j - 1];     // everything below the insertion point moves down
            const bool here = s > v[j];
            v[j] = shift ? v[j - 1] : (here ? s : v[j]);
            id[j] = shift ? id[j - 1] : (here ? col : id[j]);
        }
        if (s > v[0]) { v[0] = s; id[0] = col; }
    }
};

__device__ __forceinline__ float select32(const float (&r)[32], int i) {
    float a[16], b[8], c[4], d[2];
#pragma unroll
    for (int j = 0; j < 16; ++j) a[j] = (i & 1) ? r[2 * j + 1] : r[2 * j];
#pragma unroll
    for (int j = 0; j < 8; ++j) b[j] = (i & 2) ? a[2 * j + 1] : a[2 * j];
#pragma unroll
    for (int j = 0; j < 4; ++j) c[j] = (i & 4) ? b[2 * j + 1] : b[2 * j];
#pragma unroll
    for (int j = 0; j < 2; ++j) d[j] = (i & 8) ? c[2 * j + 1] : c[2 * j];
    return (i & 16) ? d[1] : d[0];
}

struct MainArgs {
    int n_s, n_t, kc;   // kc = C_pad / 64
    Split sp;
    float* cand_val;    // [rows_pad][slots][KP]
    int* cand_idx;
};

template <int KP, int MH>
__global__ void __launch_bounds__(128 + 128 * MH, 1)
topk_tc_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
               const MainArgs args) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const int KC = args.kc;
    uint8_t* smem_a = smem;                                       // [MH][KC] boxes
    uint8_t* smem_b = smem + (size_t)MH * KC * CHUNK_BYTES;       // [NSTAGE] boxes
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem_b + (size_t)NSTAGE * CHUNK_BYTES);
    uint64_t* b_full = bars;                 // [NSTAGE]
    uint64_t* b_empty = bars + NSTAGE;       // [NSTAGE]
    uint64_t* a_full = bars + 2 * NSTAGE;    // [1]
    uint64_t* a_empty = a_full + 1;          // [1]
    uint64_t* acc_full = a_empty + 1;        // [2]
    uint64_t* acc_empty = acc_full + 2;      // [2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const Split sp = args.sp;
    const long long t_begin = sp.begin(blockIdx.x), t_end = sp.begin(blockIdx.x + 1);

    if (warp == 1 && lane == 0) {
        for (int s = 0; s < NSTAGE; ++s) { mbar_init(&b_full[s], 1); mbar_init(&b_empty[s], 1); }
        mbar_init(a_full, 1);
        mbar_init(a_empty, 1);
        for (int s = 0; s < 2; ++s) { mbar_init(&acc_full[s], 1); mbar_init(&acc_empty[s], 4 * MH); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                         smem_u32(tmem_slot)), "r"(TMEM_COLS));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0 && lane == 0) {
        // ================= TMA producer =================
        int stage = 0; uint32_t phase = 0; int cur_rb = -1; int seg = 0;
        for (long long t = t_begin; t < t_end; ++t) {
            const int rb = (int)(t / sp.n_ct), ct = (int)(t % sp.n_ct);
            if (rb != cur_rb) {
                if (cur_rb >= 0) { mbar_wait(a_empty, (uint32_t)((seg - 1) & 1)); }
                mbar_expect_tx(a_full, (uint32_t)(MH * KC * CHUNK_BYTES));
                for (int h = 0; h < MH; ++h)
                    for (int kc = 0; kc < KC; ++kc)
                        tma_load_2d(smem_a + (size_t)(h * KC + kc) * CHUNK_BYTES, &map_a, a_full,
                                    kc * KCHUNK, (rb * MH + h) * TILE_M);
                cur_rb = rb; ++seg;
            }
            for (int kc = 0; kc < KC; ++kc) {
                mbar_wait(&b_empty[stage], phase ^ 1u);
                mbar_expect_tx(&b_full[stage], (uint32_t)CHUNK_BYTES);
                tma_load_2d(smem_b + (size_t)stage * CHUNK_BYTES, &map_b, &b_full[stage], kc * KCHUNK,
                            ct * TILE_N);
                if (++stage == NSTAGE) { stage = 0; phase ^= 1u; }
            }
        }
    } else if (warp == 1 && lane == 0) {
        // ================= MMA issuer =================
        int stage = 0; uint32_t phase = 0; int cur_rb = -1; int seg = 0; int it = 0;
        for (long long t = t_begin; t < t_end; ++t, ++it) {
            const int rb = (int)(t / sp.n_ct);
            if (rb != cur_rb) { mbar_wait(a_full, (uint32_t)(seg & 1)); cur_rb = rb; ++seg; }
            const int buf = it & 1;
            mbar_wait(&acc_empty[buf], (uint32_t)(((it >> 1) & 1) ^ 1));
            tcgen05_fence_after();
            for (int kc = 0; kc < KC; ++kc) {
                mbar_wait(&b_full[stage], phase);
                tcgen05_fence_after();
                const uint64_t bdesc = umma_desc_sw128(smem_u32(smem_b + (size_t)stage * CHUNK_BYTES));
#pragma unroll
                for (int h = 0; h < MH; ++h) {
                    const uint64_t adesc = umma_desc_sw128(smem_u32(smem_a + (size_t)(h * KC + kc) * CHUNK_BYTES));
                    const uint32_t d = tmem_base + (uint32_t)((buf * MH + h) * TILE_N);
#pragma unroll
                    for (int ks = 0; ks < KCHUNK / 16; ++ks)   // +32 bytes per K=16 step inside the swizzled row
                        tcgen05_mma_bf16(d, adesc + (uint64_t)(2 * ks), bdesc + (uint64_t)(2 * ks),
                                         IDESC_BF16_128x128, (uint32_t)((kc | ks) != 0));
                }
                tcgen05_commit(&b_empty[stage]);   // frees the ring slot when these MMAs retire
                if (++stage == NSTAGE) { stage = 0; phase ^= 1u; }
            }
            tcgen05_commit(&acc_full[buf]);
            const bool last_of_segment = (t + 1 == t_end) || ((int)((t + 1) / sp.n_ct) != rb);
            if (last_of_segment) tcgen05_commit(a_empty);
        }
    } else if (warp >= 4) {
        // ================= epilogue: one thread = one score row =================
        const int e = warp - 4, h = e >> 2, q = e & 3;   // q == warp % 4: the TMEM lane quarter this warp may read
        RowList<KP> best;
        best.reset();
        float thr = -INFINITY;
        int cur_rb = -1; int it = 0;
        auto flush = [&](int rb) {
            const int slot = (int)blockIdx.x - sp.first_cta(rb);
            const long long row = ((long long)rb * MH + h) * TILE_M + q * 32 + lane;
            float* cv = args.cand_val + ((size_t)row * sp.slots + slot) * KP;
            int* ci = args.cand_idx + ((size_t)row * sp.slots + slot) * KP;
#pragma unroll
            for (int j = 0; j < KP; ++j) { cv[j] = best.v[j]; ci[j] = best.id[j]; }
        };
        for (long long t = t_begin; t < t_end; ++t, ++it) {
            const int rb = (int)(t / sp.n_ct), ct = (int)(t % sp.n_ct);
            if (rb != cur_rb) {
                if (cur_rb >= 0) flush(cur_rb);
                best.reset(); thr = -INFINITY; cur_rb = rb;
            }
            const int buf = it & 1;
            mbar_wait(&acc_full[buf], (uint32_t)((it >> 1) & 1));
            tcgen05_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)((buf * MH + h) * TILE_N);
            const int col0 = ct * TILE_N;
            auto scan = [&](const uint32_t (&raw)[32], int c4) {
                float r[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) r[i] = __uint_as_float(raw[i]);
                const int cbase = col0 + c4 * 32;
                const int valid = args.n_t - cbase;   // columns >= n_t are TMA zero fill: ignore them
                unsigned hits = 0;
#pragma unroll
                for (int i = 0; i < 32; ++i) hits |= (r[i] > thr) ? (1u << i) : 0u;
                if (valid < 32) hits &= valid <= 0 ? 0u : ((1u << valid) - 1u);
                while (hits) {
                    const int i = __ffs(hits) - 1;
                    hits &= hits - 1;
                    const float s = select32(r, i);
                    if (s > thr) { best.insert(s, cbase + i); thr = best.threshold(); }
                }
            };
            if (KP <= 16) {
                // two register buffers: the TMEM load of the next 32 columns is in flight while this
                // chunk is scanned (with two epilogue warps per scheduler nothing else hides its latency)
                uint32_t ra[32], rb[32];
                tmem_ld_32x32b_x32_issue(taddr, ra);
                tmem_ld_wait(ra);
                tmem_ld_32x32b_x32_issue(taddr + 32u, rb);
                scan(ra, 0);
                tmem_ld_wait(rb);
                tmem_ld_32x32b_x32_issue(taddr + 64u, ra);
                scan(rb, 1);
                tmem_ld_wait(ra);
                tmem_ld_32x32b_x32_issue(taddr + 96u, rb);
                scan(ra, 2);
                tmem_ld_wait(rb);
                scan(rb, 3);
            } else {
#pragma unroll 1
                for (int c4 = 0; c4 < TILE_N / 32; ++c4) {
                    uint32_t ra[32];
                    tmem_ld_32x32b_x32_issue(taddr + (uint32_t)(c4 * 32), ra);
                    tmem_ld_wait(ra);
                    scan(ra, c4);
                }
            }
            tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&acc_empty[buf]);
        }
        if (cur_rb >= 0) flush(cur_rb);
    }

    tcgen05_fence_before();
    __syncthreads();
    if (warp == 2) {
        tcgen05_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS));
    }
}

}  // namespace tc
}  // namespace dgmc

// ================================================================================================
// merge + exact fp32 rescoring: one warp per source row
// ================================================================================================
namespace dgmc {
namespace tc {

constexpr int MERGE_MAX_PER_LANE = 8;   // slots * KP <= 256

__device__ __forceinline__ void warp_argmax(float& v, int& id) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, v, o);
        const int oi = __shfl_xor_sync(0xffffffffu, id, o);
        // larger score wins; ties go to the smaller column; -1 (empty) always loses
        if (ov > v || (ov == v && (unsigned)oi < (unsigned)id)) { v = ov; id = oi; }
    }
}

template <int KP>
__global__ void __launch_bounds__(256) topk_merge_kernel(
    const float* __restrict__ h_s, long long lds, const float* __restrict__ h_t, long long ldt,
    const float* __restrict__ cand_val, const int* __restrict__ cand_idx, const Split sp, int mh,
    int n_s, int c, int k, int* __restrict__ idx_out, float* __restrict__ score_out) {
    const int lane = threadIdx.x & 31;
    const long long warps = (long long)gridDim.x * (blockDim.x >> 5);
    const bool vec = (c % 4 == 0) && (lds % 4 == 0) && (ldt % 4 == 0) &&
                     ((reinterpret_cast<uintptr_t>(h_s) | reinterpret_cast<uintptr_t>(h_t)) % 16 == 0);
    for (long long row = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); row < n_s;
         row += warps) {
        const int rb = (int)(row / (TILE_M * mh));
        const int nseg = sp.last_cta(rb) - sp.first_cta(rb) + 1;
        const int total = nseg * KP;   // the first nseg slots of this row are the live ones
        const float* cv = cand_val + (size_t)row * sp.slots * KP;
        const int* ci = cand_idx + (size_t)row * sp.slots * KP;
        // ---- 1. the KP best by tensor-core score; lane j ends up holding candidate j -----------
        float my_v = -INFINITY; int my_c = -1;
        if (nseg == 1) {
            if (lane < KP) { my_v = cv[lane]; my_c = ci[lane]; }
        } else {
            float lv[MERGE_MAX_PER_LANE]; int lc[MERGE_MAX_PER_LANE];
#pragma unroll
            for (int u = 0; u < MERGE_MAX_PER_LANE; ++u) {
                const int p = lane + 32 * u;
                lv[u] = p < total ? cv[p] : -INFINITY;
                lc[u] = p < total ? ci[p] : -1;
            }
            for (int round = 0; round < KP; ++round) {
                float bv = -INFINITY; int bc = -1;
#pragma unroll
                for (int u = 0; u < MERGE_MAX_PER_LANE; ++u)
                    if (lv[u] > bv || (lv[u] == bv && (unsigned)lc[u] < (unsigned)bc)) { bv = lv[u]; bc = lc[u]; }
                warp_argmax(bv, bc);
#pragma unroll
                for (int u = 0; u < MERGE_MAX_PER_LANE; ++u)
                    if (lc[u] == bc) { lv[u] = -INFINITY; lc[u] = -1; }   // columns are unique across lists
                if (lane == round) { my_v = bv; my_c = bc; }
            }
        }
        // ---- 2. exact fp32 score of each surviving candidate ------------------------------------
        const float* a = h_s + (size_t)row * lds;
        float exact = -INFINITY;
        for (int j = 0; j < KP; j += 2) {
            const int c0 = __shfl_sync(0xffffffffu, my_c, j);
            const int c1 = __shfl_sync(0xffffffffu, my_c, j + 1);
            const float* b0 = h_t + (size_t)(c0 < 0 ? 0 : c0) * ldt;
            const float* b1 = h_t + (size_t)(c1 < 0 ? 0 : c1) * ldt;
            float s0 = 0.f, s1 = 0.f;
            if (vec) {
                for (int q = lane; q < c / 4; q += 32) {
                    const float4 av = ldg4(a + 4 * q), x0 = ldg4(b0 + 4 * q), x1 = ldg4(b1 + 4 * q);
                    s0 += av.x * x0.x + av.y * x0.y + av.z * x0.z + av.w * x0.w;
                    s1 += av.x * x1.x + av.y * x1.y + av.z * x1.z + av.w * x1.w;
                }
            } else {
                for (int q = lane; q < c; q += 32) {
                    const float av = __ldg(a + q);
                    s0 += av * __ldg(b0 + q);
                    s1 += av * __ldg(b1 + q);
                }
            }
            s0 = warp_sum(s0);
            s1 = warp_sum(s1);
            if (lane == j && c0 >= 0) exact = s0;
            if (lane == j + 1 && c1 >= 0) exact = s1;
        }
        // ---- 3. exact top-k, descending, ties by column ------------------------------------------
        float cur_v = my_c >= 0 ? exact : -INFINITY;
        int cur_c = my_c;
        for (int round = 0; round < k; ++round) {
            float bv = cur_v; int bc = cur_c;
            warp_argmax(bv, bc);
            if (cur_c == bc) { cur_v = -INFINITY; cur_c = -1; }
            if (lane == 0) {
                idx_out[(size_t)row * k + round] = bc;
                if (score_out) score_out[(size_t)row * k + round] = bv;
            }
        }
    }
}

// ================================================================================================
// host side
// ================================================================================================
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_tiled() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

static int make_map(CUtensorMap* map, const void* base, long long rows, int c_pad) {
    ensure_context();
    EncodeTiledFn enc = encode_tiled();
    if (!enc) { set_error("topk_tc: cuTensorMapEncodeTiled not available from the driver"); return DGMC_ERR_CUDA; }
    const cuuint64_t dims[2] = {(cuuint64_t)c_pad, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)c_pad * 2};
    const cuuint32_t box[2] = {(cuuint32_t)KCHUNK, (cuuint32_t)TILE_M};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides,
                           box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                           CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("topk_tc: cuTensorMapEncodeTiled failed (%d)", (int)r); return DGMC_ERR_CUDA; }
    return DGMC_OK;
}

struct Plan {
    int c_pad, kc, mh, kp, rows_pad;
    Split sp;
    size_t off_a, off_b, off_val, off_idx, total, smem;
    int ok;
};

static size_t up256(size_t x) { return (x + 255) & ~(size_t)255; }

static Plan make_plan(int n_s, int n_t, int c, int k) {
    Plan p{};
    p.ok = 0;
    if (n_s < 1 || n_t < 1 || c < 1 || k < 1 || k > n_t) return p;
    p.c_pad = (c + KCHUNK - 1) / KCHUNK * KCHUNK;
    p.kc = p.c_pad / KCHUNK;
    if (p.kc > MAX_A_CHUNKS) return p;                 // C <= 512
    p.mh = 2 * p.kc <= MAX_A_CHUNKS ? 2 : 1;
    p.kp = k + 6 <= 16 ? 16 : (k + 6 <= 32 ? 32 : 0);  // register-resident lists: k <= 26
    if (!p.kp) return p;
    Split& s = p.sp;
    s.n_rb = (n_s + TILE_M * p.mh - 1) / (TILE_M * p.mh);
    s.n_ct = (n_t + TILE_N - 1) / TILE_N;
    s.tiles = (long long)s.n_rb * s.n_ct;
    const int sms = sm_count();
    s.grid = (int)(s.tiles < sms ? s.tiles : sms);
    // the merge kernel unites at most 32 * MERGE_MAX_PER_LANE candidates per row: shrink the grid
    // (longer runs per CTA) until no row block is cut into more partial lists than that
    for (;;) {
        s.slots = 1;
        for (int rb = 0; rb < s.n_rb; ++rb) {
            const int n = s.last_cta(rb) - s.first_cta(rb) + 1;
            if (n > s.slots) s.slots = n;
        }
        if (s.slots * p.kp <= 32 * MERGE_MAX_PER_LANE || s.grid == 1) break;
        s.grid = s.grid * 3 / 4 > 0 ? s.grid * 3 / 4 : 1;
    }
    if (s.slots * p.kp > 32 * MERGE_MAX_PER_LANE) return p;
    p.rows_pad = s.n_rb * TILE_M * p.mh;
    size_t o = 0;
    p.off_a = o;   o += up256((size_t)n_s * p.c_pad * 2);
    p.off_b = o;   o += up256((size_t)n_t * p.c_pad * 2);
    p.off_val = o; o += up256((size_t)p.rows_pad * s.slots * p.kp * 4);
    p.off_idx = o; o += up256((size_t)p.rows_pad * s.slots * p.kp * 4);
    p.total = o;
    p.smem = (size_t)(p.mh * p.kc + NSTAGE) * CHUNK_BYTES + 256 + 1024;
    p.ok = 1;
    return p;
}

template <int KP, int MH>
static int launch_main(const Plan& p, const CUtensorMap& ma, const CUtensorMap& mb, const MainArgs& args,
                       cudaStream_t stream) {
    static size_t configured_on[MAX_DEVICES] = {};      // per device: the attribute is a per-device setting
    size_t& configured = configured_on[current_device()];
    if (p.smem > configured) {
        if (cudaFuncSetAttribute(topk_tc_kernel<KP, MH>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)p.smem) != cudaSuccess) {
            set_error("topk_tc: cannot set %zu bytes of dynamic shared memory", p.smem);
            return DGMC_ERR_CUDA;
        }
        configured = p.smem;
    }
    topk_tc_kernel<KP, MH><<<p.sp.grid, 128 + 128 * MH, p.smem, stream>>>(ma, mb, args);
    return DGMC_OK;
}

}  // namespace tc
}  // namespace dgmc

using namespace dgmc;

extern "C" int dgmc_topk_tc_supported(int n_s, int n_t, int c, int k) {
    return tc::make_plan(n_s, n_t, c, k).ok;
}

extern "C" size_t dgmc_topk_tc_workspace(int n_s, int n_t, int c, int k) {
    const tc::Plan p = tc::make_plan(n_s, n_t, c, k);
    return p.ok ? p.total : 0;
}

extern "C" int dgmc_topk_tc(const float* h_s, int64_t lds, const float* h_t, int64_t ldt, int n_s,
                            int n_t, int c, int k, int32_t* idx_out, float* score_out, void* workspace,
                            size_t workspace_bytes, dgmc_stream_t stream) {
    DGMC_REQUIRE(h_s && h_t && idx_out && workspace && lds >= c && ldt >= c, "topk_tc: bad arguments");
    const tc::Plan p = tc::make_plan(n_s, n_t, c, k);
    if (!p.ok) {
        set_error("topk_tc: unsupported shape n_s=%d n_t=%d c=%d k=%d (need c <= 512, 1 <= k <= 26, k <= n_t)",
                  n_s, n_t, c, k);
        return DGMC_ERR_UNSUPPORTED;
    }
    if (workspace_bytes < p.total || (reinterpret_cast<uintptr_t>(workspace) & 255) != 0) {
        set_error("topk_tc: workspace needs %zu bytes, 256-byte aligned (got %zu)", p.total, workspace_bytes);
        return DGMC_ERR_WORKSPACE;
    }
    char* ws = static_cast<char*>(workspace);
    __nv_bfloat16* a16 = reinterpret_cast<__nv_bfloat16*>(ws + p.off_a);
    __nv_bfloat16* b16 = reinterpret_cast<__nv_bfloat16*>(ws + p.off_b);
    float* cand_val = reinterpret_cast<float*>(ws + p.off_val);
    int* cand_idx = reinterpret_cast<int*>(ws + p.off_idx);

    tc::cast_bf16_kernel<<<wave_grid((long long)n_s * p.c_pad / 2, 256, 8), 256, 0, stream>>>(
        h_s, lds, a16, n_s, c, p.c_pad);
    tc::cast_bf16_kernel<<<wave_grid((long long)n_t * p.c_pad / 2, 256, 8), 256, 0, stream>>>(
        h_t, ldt, b16, n_t, c, p.c_pad);

    CUtensorMap ma, mb;
    if (int e = tc::make_map(&ma, a16, n_s, p.c_pad)) return e;
    if (int e = tc::make_map(&mb, b16, n_t, p.c_pad)) return e;
    tc::MainArgs args{n_s, n_t, p.kc, p.sp, cand_val, cand_idx};
    int e;
    if (p.kp == 16) e = p.mh == 2 ? tc::launch_main<16, 2>(p, ma, mb, args, stream)
                                  : tc::launch_main<16, 1>(p, ma, mb, args, stream);
    else            e = p.mh == 2 ? tc::launch_main<32, 2>(p, ma, mb, args, stream)
                                  : tc::launch_main<32, 1>(p, ma, mb, args, stream);
    if (e) return e;
    const int mgrid = wave_grid(n_s, 8, 8);
    if (p.kp == 16)
        tc::topk_merge_kernel<16><<<mgrid, 256, 0, stream>>>(h_s, lds, h_t, ldt, cand_val, cand_idx, p.sp,
                                                            p.mh, n_s, c, k, idx_out, score_out);
    else
        tc::topk_merge_kernel<32><<<mgrid, 256, 0, stream>>>(h_s, lds, h_t, ldt, cand_val, cand_idx, p.sp,
                                                            p.mh, n_s, c, k, idx_out, score_out);
    return check_launch("topk_tc", 4);
}
