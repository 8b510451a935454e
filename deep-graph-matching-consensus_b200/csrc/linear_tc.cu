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
// on the fly.  Persistent kernel, one CTA per SM looping over 128-row x BN-column tiles, 10 warps:
//   warp 0      TMA producer: per 32-wide k block, X[128 x 32] (raw fp32), W_hi[BN x 32], W_lo[BN x 32]
//               -> 128-byte-swizzled shared-memory stage, mbarrier complete_tx
//   warps 2-5   split the raw X tile into hi (in place) and lo (second buffer, same swizzled
//               position, so no address math), fence.proxy.async, arrive
//   warp 1      one thread issues, per k8 step, X_hi x [W_hi ; W_lo] (N = 2 BN) and X_lo x W_hi
//               (N = BN) as tcgen05.mma.kind::tf32; tcgen05.commit releases the stage to the producer
//   warps 6-9   epilogue: tcgen05.ld of the accumulator rows (summing the accumulator pairs),
//               + bias, ReLU, swizzled staging rows, one TMA store per 32 x 32 box; with two
//               accumulator sets in TMEM the epilogue of one tile overlaps the MMAs of the next.
//               Reductions longer than 384 columns are cut into ranges of 128 whose accumulator sets
//               these warps drain into register sums (the TMEM accumulate truncates: see RANGE)
// Rows and k beyond the matrix are zero-filled by TMA (X) or by the padded workspace (W).
#include <cstring>
#include "common.cuh"
#include "tc_common.cuh"
#include "../../include/dgmc_b200.h"

namespace dgmc {
namespace tc {

constexpr int LT_BM = 128;                    // rows per tile (UMMA M)
constexpr int LT_KB = 32;                     // k per stage = one 128-byte swizzled row of fp32
constexpr int LT_A_BYTES = LT_BM * LT_KB * 4; // 16 KB
constexpr int LT_THREADS = 320;                // producer, MMA issuer, 4 split warps, 4 epilogue warps
constexpr int LT_MAX_STAGES = 6;
constexpr int LT_TMEM_COLS = 512;
constexpr int LT_OUT_BYTES = 32 * 128;         // epilogue staging per warp: 32 rows x 32 fp32, 128-byte swizzled
__host__ __device__ constexpr int lt_stages(int bn) { return bn == 64 ? 4 : 3; }   // 4 x 48 KB / 3 x 64 KB

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

// Dense-branch mode (DENSE = true): the batched score product S_hat[b] = h_s[b] h_t[b]^T of the dense
// variant (dgmc.py:163) with the masked row softmax (dgmc.py:15-19,165) in the epilogue.  "X" = the
// padded source embeddings [B n_s, C], "W" = the (split) padded target embeddings [B n_t, C]; tile t
// covers the g graphs t g .. t g + g - 1 (g n_s <= 128 rows, g n_t <= 128 columns): rows t g n_s ..,
// W rows t g n_t ...  Only the diagonal n_s x n_t blocks of the 128 x 128 tile are meaningful; the
// tensor core computes the rest for free.  One epilogue thread = one row: it walks its graph's column
// window three times (max, sum of exponentials, output) over TMEM and writes S_hat and S_0 itself.
struct DenseArgs {
    const uint8_t* s_mask; const uint8_t* t_mask;   // [B, n_s], [B, n_t]
    float* s_hat; float* s_0;                       // [B, n_s, n_t]
    int B, n_s, n_t, g;                             // g = graphs per tile
};

// Accumulator layout of one set (see the MMA issuer):
//   LT_ACC pairs [H | L] the k8 steps alternate between (1 for k_in <= 128, which leaves room for a
//   second set at BN = 128; 2 up to k_in = 384).
//   RANGE > 0 (long reductions: psi_1, k_in = 768 ... 1068): the reduction is cut into ranges of RANGE
//   k blocks; each range fills one accumulator set [H | L] (LT_ACC = 1: two sets in TMEM) and the
//   epilogue warps drain it into REGISTER running sums (round-to-nearest fp32) while the next range's
//   MMAs fill the other set — 4 RANGE truncating adds per accumulator whatever k_in is, on full
//   128-column tiles.
template <int BN, int LT_ACC, bool DENSE, int RANGE = 0>
__global__ void __launch_bounds__(LT_THREADS, 1)
linear_tc_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_whi,
                 const __grid_constant__ CUtensorMap map_wlo, const __grid_constant__ CUtensorMap map_y,
                 int y_tma, const float* __restrict__ bias, long long bias_rows, int relu, float* __restrict__ y,
                 long long ldy,
                 long long rows, int n_out, int k_blocks, int n_tiles_n, int n_tiles, int bm, const DenseArgs dense) {
    constexpr int B_BYTES = BN * LT_KB * 4;
    constexpr int STAGE_BYTES = 2 * LT_A_BYTES + 2 * B_BYTES;
    constexpr int NSTAGE = lt_stages(BN);
    constexpr int ACC_COLS = 2 * LT_ACC * BN;          // one accumulator set: LT_ACC pairs [H | L]
    // accumulator sets: with two, the epilogue of one tile overlaps the MMAs of the next
    constexpr int NBUF = LT_TMEM_COLS / ACC_COLS >= 2 ? 2 : 1;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* out_stage = smem + (size_t)NSTAGE * STAGE_BYTES;      // 4 epilogue warps x [32 rows x 128 B]
    uint64_t* bars = reinterpret_cast<uint64_t*>(out_stage + 4 * LT_OUT_BYTES);
    uint64_t* full = bars;                          // TMA bytes landed
    uint64_t* split_done = bars + LT_MAX_STAGES;    // X tile split into hi / lo (4 warp arrivals)
    uint64_t* empty = bars + 2 * LT_MAX_STAGES;     // MMAs of the stage retired
    uint64_t* acc_full = bars + 3 * LT_MAX_STAGES;  // [2] accumulator set complete (tcgen05.commit)
    uint64_t* acc_empty = acc_full + 2;             // [2] drained by the 4 epilogue warps
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        for (int s = 0; s < NSTAGE; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&split_done[s], 4);
            mbar_init(&empty[s], 1);
        }
        for (int b = 0; b < 2; ++b) { mbar_init(&acc_full[b], 1); mbar_init(&acc_empty[b], 4); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                         smem_u32(tmem_slot)), "r"(LT_TMEM_COLS));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_d = *tmem_slot;

    // Persistent: CTA b works on tiles b, b + grid, ...; consecutive tiles are the column tiles of one
    // row block, so CTAs running side by side share their X rows in L2.  The shared-memory ring
    // runs on across tile boundaries: the loads of the next tile are in flight while this one's
    // MMAs and epilogue finish.
    if (warp == 0) {
        if (lane == 0) {
            // ================= TMA producer =================
            int s = 0; uint32_t phase = 0;
            for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
                const int m0 = DENSE ? tile * dense.g * dense.n_s : (tile / n_tiles_n) * bm;
                const int n0 = DENSE ? tile * dense.g * dense.n_t : (tile % n_tiles_n) * BN;
                for (int kb = 0; kb < k_blocks; ++kb) {
                    mbar_wait(&empty[s], phase ^ 1u);
                    uint8_t* st = smem + (size_t)s * STAGE_BYTES;
                    mbar_expect_tx(&full[s], (uint32_t)(bm * LT_KB * 4 + 2 * B_BYTES));   // X box: bm rows
                    tma_load_2d(st, &map_x, &full[s], kb * LT_KB, m0);
                    tma_load_2d(st + 2 * LT_A_BYTES, &map_whi, &full[s], kb * LT_KB, n0);
                    tma_load_2d(st + 2 * LT_A_BYTES + B_BYTES, &map_wlo, &full[s], kb * LT_KB, n0);
                    if (++s == NSTAGE) { s = 0; phase ^= 1u; }
                }
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
            // pairs, which the epilogue adds in round-to-nearest fp32.  With RANGE > 0 an accumulator set
            // only ever receives RANGE k blocks (4 RANGE adds): see the epilogue.
            constexpr uint32_t idesc_wide = (1u << 4) | (2u << 7) | (2u << 10) |
                                            ((uint32_t)((2 * BN) >> 3) << 17) | ((uint32_t)(LT_BM >> 4) << 24);
            constexpr uint32_t idesc_lo = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(BN >> 3) << 17) |
                                          ((uint32_t)(LT_BM >> 4) << 24);
            int s = 0; uint32_t phase = 0; int it = 0;
            const int per = RANGE > 0 ? RANGE : k_blocks;       // k blocks per accumulator set
            for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x)
            for (int kb0 = 0; kb0 < k_blocks; kb0 += per, ++it) {
                const int buf = it % NBUF, use = it / NBUF;
                mbar_wait(&acc_empty[buf], (uint32_t)((use & 1) ^ 1));   // epilogue has drained this set
                tcgen05_fence_after();
                const uint32_t acc = tmem_d + (uint32_t)(buf * ACC_COLS);
                int step = 0;
                const int kb1 = kb0 + per < k_blocks ? kb0 + per : k_blocks;
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(&full[s], phase);          // W tiles (written by TMA)
                    mbar_wait(&split_done[s], phase);    // X tiles (rewritten by warps 2-5)
                    tcgen05_fence_after();
                    const uint32_t st = smem_u32(smem + (size_t)s * STAGE_BYTES);
                    const uint64_t d_ahi = umma_desc_sw128(st), d_alo = umma_desc_sw128(st + LT_A_BYTES);
                    const uint64_t d_b = umma_desc_sw128(st + 2 * LT_A_BYTES);   // W_hi rows, then W_lo rows
#pragma unroll
                    for (int ks = 0; ks < LT_KB / 8; ++ks, ++step) {   // +32 bytes per K=8 step inside the swizzled row
                        const uint64_t o = (uint64_t)(2 * ks);
                        const uint32_t d = acc + (uint32_t)((step % LT_ACC) * 2 * BN);
                        tcgen05_mma_tf32(d, d_ahi + o, d_b + o, idesc_wide, (uint32_t)(step >= LT_ACC));
                        tcgen05_mma_tf32(d + BN, d_alo + o, d_b + o, idesc_lo, 1u);
                    }
                    tcgen05_commit(&empty[s]);
                    if (++s == NSTAGE) { s = 0; phase ^= 1u; }
                }
                tcgen05_commit(&acc_full[buf]);
            }
        }
    } else if (warp < 6) {
        // ================= split X into hi (in place) and lo =================
        const int t = threadIdx.x - 64;               // 0..127
        int s = 0; uint32_t phase = 0;
        for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
            for (int kb = 0; kb < k_blocks; ++kb) {
                mbar_wait(&full[s], phase);
                float4* a_hi = reinterpret_cast<float4*>(smem + (size_t)s * STAGE_BYTES);
                float4* a_lo = a_hi + LT_A_BYTES / 16;
#pragma unroll
                for (int i = 0; i < LT_A_BYTES / 16 / 128; ++i) {
                    // all 128 rows, whatever bm is: rows >= bm hold stale values whose products land in
                    // accumulator rows nobody reads — a test on bm here costs the eight independent loads
                    // their overlap (measured: 15.1 -> 18.5 us at 38,960 x 192 -> 64)
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
                if (++s == NSTAGE) { s = 0; phase ^= 1u; }
            }
        }
    } else {
        // ================= epilogue (warps 6-9) =================
        const int quarter = warp & 3;                  // TMEM lanes this warp may read
        uint8_t* my_stage = out_stage + (warp - 6) * LT_OUT_BYTES;
        int it = 0;
        for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, it += (RANGE > 0 ? 0 : 1)) {
            const int buf = it % NBUF, use = it / NBUF;
            const long long m0 = (long long)(tile / n_tiles_n) * bm;
            const bool idle = !DENSE && quarter * 32 >= bm;   // this warp's 32 rows lie beyond the tile
            const int n0 = (tile % n_tiles_n) * BN;
            const long long row0 = m0 + quarter * 32, row = row0 + lane;
            // bias + ReLU on one 32-column chunk of this thread's row, then out through TMA or plain stores
            auto finish = [&](int cb, float (&r)[32]) {
                const int c0 = n0 + cb;
                if (bias && row < bias_rows) {       // the bias may apply to the leading rows only
#pragma unroll
                    for (int j = 0; j < 32; ++j) r[j] += (c0 + j < n_out) ? __ldg(bias + c0 + j) : 0.f;
                }
                if (relu) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) r[j] = fmaxf(r[j], 0.f);
                }
                if (y_tma) {
                    // registers -> swizzled staging rows (conflict-free 128-bit stores) -> one TMA
                    // store of the 32 x 32 box, which writes whole 128-byte row segments and clips
                    // rows / columns beyond the matrix
                    if (row0 < rows) {
                        if (lane == 0) tma_store_wait_read();
                        __syncwarp();
#pragma unroll
                        for (int j = 0; j < 8; ++j)
                            *reinterpret_cast<float4*>(my_stage + lane * 128 + ((j ^ (lane & 7)) << 4)) =
                                make_float4(r[4 * j], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]);
                        fence_proxy_async_smem();
                        __syncwarp();
                        if (lane == 0) {
                            tma_store_2d(&map_y, my_stage, c0, (int)row0);
                            tma_store_commit();
                        }
                    }
                } else if (row < rows) {
                    float* dst = y + row * ldy + c0;
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        if (c0 + j < n_out) dst[j] = r[j];
                }
            };
            if (RANGE > 0) {
                static_assert(RANGE == 0 || (!DENSE && LT_ACC == 1), "ranges drain [H | L] sets");
                float sum[BN / 32][32];
#pragma unroll
                for (int c = 0; c < BN / 32; ++c)
#pragma unroll
                    for (int j = 0; j < 32; ++j) sum[c][j] = 0.f;
                for (int kb0 = 0; kb0 < k_blocks; kb0 += RANGE, ++it) {
                    const int rbuf = it % NBUF, ruse = it / NBUF;
                    mbar_wait(&acc_full[rbuf], (uint32_t)(ruse & 1));
                    tcgen05_fence_after();
                    if (idle) {
                        tcgen05_fence_before();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(&acc_empty[rbuf]);
                        continue;
                    }
                    const uint32_t lane_base = tmem_d + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(rbuf * ACC_COLS);
#pragma unroll
                    for (int c = 0; c < BN / 16; ++c) {      // 16 columns at a time: 128 running sums stay in registers
                        float h[16], l[16];
                        tmem_ld_32x32b_x16_pair(lane_base + (uint32_t)(c * 16), lane_base + (uint32_t)(BN + c * 16), h, l);
#pragma unroll
                        for (int j = 0; j < 16; ++j) sum[c / 2][(c % 2) * 16 + j] += h[j] + l[j];
                    }
                    tcgen05_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&acc_empty[rbuf]);
                }
#pragma unroll
                for (int c = 0; c < BN / 32; ++c)
                    if (!idle && n0 + c * 32 < n_out) finish(c * 32, sum[c]);
                continue;
            }
            mbar_wait(&acc_full[buf], (uint32_t)(use & 1));
            tcgen05_fence_after();
            if (idle) {
                tcgen05_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&acc_empty[buf]);
                continue;
            }
            if (DENSE) {
                // ---- masked row softmax of this row's n_s x n_t block, straight out of TMEM ----
                const int r = quarter * 32 + lane;                 // row of the tile
                const int gi = r / dense.n_s, i = r - gi * dense.n_s;
                const int b = tile * dense.g + gi;
                const bool valid = gi < dense.g && b < dense.B;
                const int w0 = gi * dense.n_t;                     // first column of the graph's window
                const bool on = valid && dense.s_mask[(size_t)b * dense.n_s + i] != 0;
                const uint8_t* tm = dense.t_mask + (size_t)(valid ? b : 0) * dense.n_t;
                const uint32_t lane_base0 = tmem_d + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(buf * ACC_COLS);
                auto chunk = [&](int cb, float (&v)[32]) {         // v = (H_0 + H_1) + (L_0 + L_1)
                    const uint32_t lane_base = lane_base0 + (uint32_t)cb;
                    float l[32];
                    tmem_ld_32x32b_x32_pair(lane_base, lane_base + (uint32_t)BN, v, l);
                    if (LT_ACC == 2) {
                        float q[32], m[32];
                        tmem_ld_32x32b_x32_pair(lane_base + (uint32_t)(2 * BN), lane_base + (uint32_t)(3 * BN), q, m);
#pragma unroll
                        for (int j = 0; j < 32; ++j) { v[j] += q[j]; l[j] += m[j]; }
                    }
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] += l[j];
                };
                float mx = -INFINITY, sum = 0.f;
#pragma unroll 1
                for (int cb = 0; cb < BN; cb += 32) {              // pass 1: row maximum over the live columns
                    float v[32];
                    chunk(cb, v);
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        const int col = cb + j - w0;
                        if (on && col >= 0 && col < dense.n_t && tm[col]) mx = fmaxf(mx, v[j]);
                    }
                }
#pragma unroll 1
                for (int cb = 0; cb < BN; cb += 32) {              // pass 2: sum of exponentials
                    float v[32];
                    chunk(cb, v);
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        const int col = cb + j - w0;
                        if (on && col >= 0 && col < dense.n_t && tm[col]) sum += __expf(v[j] - mx);
                    }
                }
                const float inv = sum > 0.f ? 1.f / sum : 0.f;
#pragma unroll 1
                for (int cb = 0; cb < BN; cb += 32) {              // pass 3: S_hat and S_0
                    float v[32];
                    chunk(cb, v);
                    if (cb + 32 >= BN) {                           // last TMEM read of this accumulator set
                        tcgen05_fence_before();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(&acc_empty[buf]);
                    }
                    if (valid) {
                        float* sh = dense.s_hat + ((size_t)b * dense.n_s + i) * dense.n_t;
                        float* s0 = dense.s_0 + ((size_t)b * dense.n_s + i) * dense.n_t;
#pragma unroll
                        for (int j = 0; j < 32; ++j) {
                            const int col = cb + j - w0;
                            if (col >= 0 && col < dense.n_t) {
                                sh[col] = v[j];
                                s0[col] = (on && tm[col]) ? __expf(v[j] - mx) * inv : 0.f;
                            }
                        }
                    }
                }
                continue;
            }
#pragma unroll 1
            for (int cb = 0; cb < BN; cb += 32) {
                if (n0 + cb >= n_out) break;               // warp-uniform; cb = 0 is always live
                float r[32];
                {
                    const uint32_t lane_base = tmem_d + ((uint32_t)(quarter * 32) << 16) +
                                               (uint32_t)(buf * ACC_COLS + cb);
                    float l[32];
                    tmem_ld_32x32b_x32_pair(lane_base, lane_base + (uint32_t)BN, r, l);      // H_0, L_0
                    if (LT_ACC == 2) {
                        float q[32], m[32];
                        tmem_ld_32x32b_x32_pair(lane_base + (uint32_t)(2 * BN), lane_base + (uint32_t)(3 * BN), q, m);
#pragma unroll
                        for (int j = 0; j < 32; ++j) { r[j] += q[j]; l[j] += m[j]; }         // + H_1, + L_1
                    }
#pragma unroll
                    for (int j = 0; j < 32; ++j) r[j] += l[j];
                }
                if (cb + 32 >= BN || n0 + cb + 32 >= n_out) {
                    // last chunk read: the MMA warp may overwrite this accumulator set while the
                    // stores below drain
                    tcgen05_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&acc_empty[buf]);
                }
                finish(cb, r);
            }
        }
        if (lane == 0) tma_store_wait_all();
    }
    tcgen05_fence_before();
    __syncthreads();
    if (warp == 1) {
        tcgen05_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"(LT_TMEM_COLS));
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
    memset(map, 0, sizeof(*map));
    ensure_context();
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

template <int BN, int ACC, int RANGE = 0>
static int lt_launch(const CUtensorMap& mx, const CUtensorMap& mh, const CUtensorMap& ml,
                     const CUtensorMap& my, int y_tma, const float* bias, long long bias_rows, int relu, float* y, long long ldy, long long rows, int n_out, int k_blocks, int n_pad,
                     int bm, cudaStream_t stream) {
    const DenseArgs no_dense = {};
    constexpr int STAGE = 2 * LT_A_BYTES + 2 * BN * LT_KB * 4;
    // the full ring always (>= 192 KB): one CTA per SM, which owns all 512 TMEM columns
    const size_t smem = (size_t)lt_stages(BN) * STAGE + 4 * LT_OUT_BYTES + (3 * LT_MAX_STAGES + 6) * 8 + 1024;
    static DeviceOnce configured;
    if (!configured.done()) {
        if (cudaFuncSetAttribute(linear_tc_kernel<BN, ACC, false, RANGE>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)smem) != cudaSuccess) {
            set_error("linear_tc: cannot set %zu bytes of dynamic shared memory", smem);
            return DGMC_ERR_CUDA;
        }
        configured.mark();
    }
    const int n_tiles_n = n_pad / BN;
    const long long tiles = (long long)n_tiles_n * ((rows + bm - 1) / bm);
    if (tiles > 0x7fffffffLL) { set_error("linear_tc: too many tiles"); return DGMC_ERR_UNSUPPORTED; }
    const int grid = (int)(tiles < sm_count() ? tiles : sm_count());
    linear_tc_kernel<BN, ACC, false, RANGE><<<grid, LT_THREADS, smem, stream>>>(mx, mh, ml, my, y_tma, bias, bias_rows, relu, y,
                                                                   ldy, rows, n_out, k_blocks, n_tiles_n, (int)tiles,
                                                                   bm, no_dense);
    return DGMC_OK;
}

}  // namespace tc
}  // namespace dgmc

using namespace dgmc;

extern "C" int dgmc_linear_tc_supported(int64_t rows, int n_out, int k_in, int64_t ldx, const float* x) {
    // TMA: 16-byte aligned base and pitch; tile count fits an int
    return rows >= 1 && rows <= (int64_t)1 << 30 && n_out >= 1 && n_out <= 65536 && k_in >= 1 && ldx >= k_in &&
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

static int linear_tc_impl(const float* x, int64_t ldx, const void* split_weight, const float* bias,
                          int64_t bias_rows, int relu, float* y, int64_t ldy, int64_t rows, int n_out, int k_in,
                          dgmc_stream_t stream) {
    DGMC_REQUIRE(x && split_weight && y && ldy >= n_out, "linear_tc: bad arguments");
    if (!dgmc_linear_tc_supported(rows, n_out, k_in, ldx, x)) {
        set_error("linear_tc: unsupported operand (rows=%lld n_out=%d k_in=%d ldx=%lld; x needs a 16-byte "
                  "aligned base and pitch)", (long long)rows, n_out, k_in, (long long)ldx);
        return DGMC_ERR_UNSUPPORTED;
    }
    // Long reductions (more than 12 k blocks: psi_1's 768 ... 1068 columns) run in ranges of 4 k blocks
    // whose accumulator sets the epilogue drains into register sums (RANGE in the kernel): 16 truncating
    // adds per TMEM accumulator whatever k_in is.  Measured on psi_1 at the DBP15K shape against fp64:
    // 6.1e-7 of max|h| (IEEE fp32 on the CPU: 8.4e-7; two accumulator pairs over the whole reduction:
    // 2.4e-6, which put S_0 5x further from fp64 than the reference's own fp32 noise), at 99 us instead
    // of 93 us for 34.5 k x 900 -> 256.
    const int n_pad = tc::lt_round(n_out, tc::lt_bn(n_out)), k_pad = tc::lt_round(k_in, tc::LT_KB);
    const bool long_k = k_pad / tc::LT_KB > 12;
    const int bn = tc::lt_bn(n_out);
    const float* hi = static_cast<const float*>(split_weight);
    const float* lo = hi + (size_t)n_pad * k_pad;
    // Rows per tile: 128, except for matrices too small to give every SM a tile — there a tile holds only
    // 96 or 64 live rows of the M = 128 instruction (the MMA is not the limit of these GEMMs), so that
    // more SMs take part: 257 x 72 -> 50: 8.8 -> 7.0 us.  With more than one round of tiles shorter tiles
    // measured level or slower (38,960 x 192 -> 64: 15.1 us with 128 rows, 15.5 us with 96), so the rule
    // stops there.
    int bm = tc::LT_BM;
    {
        const long long sms = sm_count(), n_tiles_n = n_pad / bn;
        auto tiles = [&](int h) { return n_tiles_n * ((rows + h - 1) / h); };
        if (tiles(tc::LT_BM) < sms) bm = tiles(64) <= sms ? 64 : (tiles(96) <= sms ? 96 : tc::LT_BM);
    }
    CUtensorMap mx, mh, ml;
    if (int e = tc::lt_make_map(&mx, x, rows, k_in, ldx, bm)) return e;
    if (int e = tc::lt_make_map(&mh, hi, n_pad, k_pad, k_pad, bn)) return e;
    if (int e = tc::lt_make_map(&ml, lo, n_pad, k_pad, k_pad, bn)) return e;
    // y through TMA stores when its base and pitch are 16-byte aligned, else plain stores
    const int y_tma = (ldy % 4 == 0) && ((reinterpret_cast<uintptr_t>(y) & 15) == 0);
    CUtensorMap my;
    if (y_tma) { if (int e = tc::lt_make_map(&my, y, rows, n_out, ldy, 32)) return e; }
    else my = mx;
    const int k_blocks = k_pad / tc::LT_KB;
    int e;
    const bool short_k = k_blocks <= 4;
#define DGMC_LT_ARGS mx, mh, ml, my, y_tma, bias, bias_rows, relu, y, ldy, rows, n_out, k_blocks, n_pad, bm, stream
    if (long_k) e = bn == 64 ? tc::lt_launch<64, 1, 4>(DGMC_LT_ARGS) : tc::lt_launch<128, 1, 4>(DGMC_LT_ARGS);
    else if (bn == 64) e = short_k ? tc::lt_launch<64, 1>(DGMC_LT_ARGS) : tc::lt_launch<64, 2>(DGMC_LT_ARGS);
    else          e = short_k ? tc::lt_launch<128, 1>(DGMC_LT_ARGS) : tc::lt_launch<128, 2>(DGMC_LT_ARGS);
#undef DGMC_LT_ARGS
    if (e) return e;
    return check_launch("linear_tc");
}

extern "C" int dgmc_linear_tc(const float* x, int64_t ldx, const void* split_weight, const float* bias, int relu,
                              float* y, int64_t ldy, int64_t rows, int n_out, int k_in, dgmc_stream_t stream) {
    return linear_tc_impl(x, ldx, split_weight, bias, rows, relu, y, ldy, rows, n_out, k_in, stream);
}

extern "C" int dgmc_linear_tc_rows(const float* x, int64_t ldx, const void* split_weight, const float* bias,
                                   int64_t bias_rows, int relu, float* y, int64_t ldy, int64_t rows, int n_out,
                                   int k_in, dgmc_stream_t stream) {
    DGMC_REQUIRE(bias_rows >= 0 && bias_rows <= rows, "linear_tc_rows: bias_rows out of range");
    return linear_tc_impl(x, ldx, split_weight, bias, bias_rows, relu, y, ldy, rows, n_out, k_in, stream);
}

// ---- dense branch: S_hat = h_s h_t^T per graph + masked row softmax ---------------------------------
extern "C" int dgmc_dense_scores_tc_supported(int B, int n_s, int n_t, int c, const float* h_s, const float* h_t) {
    return B >= 1 && n_s >= 1 && n_s <= tc::LT_BM && n_t >= 1 && n_t <= 128 && c >= 1 && c % 4 == 0 &&
           (long long)B * (n_s > n_t ? n_s : n_t) < (1LL << 30) &&
           ((reinterpret_cast<uintptr_t>(h_s) | reinterpret_cast<uintptr_t>(h_t)) & 15) == 0;
}

extern "C" size_t dgmc_dense_scores_tc_workspace(int B, int n_t, int c) {
    if (B < 1 || n_t < 1 || c < 1) return 0;
    const size_t n_pad = tc::lt_round(B * n_t, 128), k_pad = tc::lt_round(c, tc::LT_KB);
    return 2 * n_pad * k_pad * sizeof(float);
}

extern "C" int dgmc_dense_scores_tc(const float* h_s, const float* h_t, const uint8_t* s_mask, const uint8_t* t_mask,
                                    float* s_hat, float* s_0, int B, int n_s, int n_t, int c, void* workspace,
                                    size_t workspace_bytes, dgmc_stream_t stream) {
    DGMC_REQUIRE(h_s && h_t && s_mask && t_mask && s_hat && s_0 && workspace, "dense_scores_tc: null argument");
    if (!dgmc_dense_scores_tc_supported(B, n_s, n_t, c, h_s, h_t)) {
        set_error("dense_scores_tc: unsupported shape (B=%d n_s=%d n_t=%d c=%d; graphs of up to 128 nodes, c %% 4 == 0, "
                  "16-byte aligned embeddings)", B, n_s, n_t, c);
        return DGMC_ERR_UNSUPPORTED;
    }
    DGMC_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 127) == 0, "dense_scores_tc: workspace must be 128-byte aligned");
    if (workspace_bytes < dgmc_dense_scores_tc_workspace(B, n_t, c)) {
        set_error("dense_scores_tc: workspace too small");
        return DGMC_ERR_WORKSPACE;
    }
    // 1. split h_t (the "weight" operand: [B n_t, c]) into tf32 hi / lo, K-major, zero-padded
    const int n_rows_t = B * n_t;
    const int n_pad = tc::lt_round(n_rows_t, 128), k_pad = tc::lt_round(c, tc::LT_KB);
    float* hi = static_cast<float*>(workspace);
    float* lo = hi + (size_t)n_pad * k_pad;
    {
        const long long total = (long long)n_pad * k_pad;
        int blocks = (int)((total + 255) / 256);
        if (blocks > 8 * sm_count()) blocks = 8 * sm_count();
        tc::split_weight_kernel<<<blocks, 256, 0, stream>>>(h_t, c, 0, n_rows_t, c, n_pad, k_pad, hi, lo);
        if (int e = check_launch("dense_scores_tc (split)")) return e;
    }
    // 2. one 128 x 128 tile per g graphs
    const int g_s = tc::LT_BM / n_s, g_t = 128 / n_t;
    tc::DenseArgs d;
    d.s_mask = s_mask; d.t_mask = t_mask; d.s_hat = s_hat; d.s_0 = s_0;
    d.B = B; d.n_s = n_s; d.n_t = n_t; d.g = g_s < g_t ? g_s : g_t;
    const int tiles = (B + d.g - 1) / d.g;
    CUtensorMap mx, mh, ml;
    if (int e = tc::lt_make_map(&mx, h_s, (long long)B * n_s, c, c, tc::LT_BM)) return e;
    if (int e = tc::lt_make_map(&mh, hi, n_pad, k_pad, k_pad, 128)) return e;
    if (int e = tc::lt_make_map(&ml, lo, n_pad, k_pad, k_pad, 128)) return e;
    constexpr int STAGE = 2 * tc::LT_A_BYTES + 2 * 128 * tc::LT_KB * 4;
    const size_t smem = (size_t)tc::lt_stages(128) * STAGE + 4 * tc::LT_OUT_BYTES + (3 * tc::LT_MAX_STAGES + 6) * 8 + 1024;
    static DeviceOnce configured;
    if (!configured.done()) {
        if (cudaFuncSetAttribute(tc::linear_tc_kernel<128, 2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)smem) != cudaSuccess) {
            set_error("dense_scores_tc: cannot set %zu bytes of dynamic shared memory", smem);
            return DGMC_ERR_CUDA;
        }
        configured.mark();
    }
    const int grid = tiles < sm_count() ? tiles : sm_count();
    tc::linear_tc_kernel<128, 2, true><<<grid, tc::LT_THREADS, smem, stream>>>(
        mx, mh, ml, mx, 0, nullptr, 0, 0, nullptr, 0, (long long)B * n_s, 128, k_pad / tc::LT_KB, 1, tiles, tc::LT_BM, d);
    return check_launch("dense_scores_tc");
}
