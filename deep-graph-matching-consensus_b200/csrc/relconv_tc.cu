// Fused RelConv layer: neighbourhood aggregation -> tensor-core GEMM in ONE kernel.
//
//   y[i, :] = act( [ x_i | sum_{e in seg(3i+1)} w_e x_{src_e} | sum_{e in seg(3i+2)} w_e x_{src_e} ] W^T
//                  + bias ) (+ addend[i, :])
//
// replaces, per layer application, the chain  aggregation kernel -> fix-up kernel -> [N, 3C] buffer ->
// GEMM kernel  behind RelConv.forward (rel.py:25-31: root(x) + mean_{j->i} lin1(x_j) + mean_{i->j}
// lin2(x_j); the means commute with the linear maps, so the layer is ONE GEMM of the aggregated
// operand [x | mean_in | mean_out] with the weights side by side).  The operand tile never makes the
// round trip through HBM: the warps that linear_tc.cu uses to split a TMA-loaded X tile into tf32
// hi / lo parts here BUILD the tile — they gather the neighbour rows (L2-resident at the DBP15K shape)
// straight into the 128-byte-swizzled shared-memory stage the UMMA descriptor expects, already split.
// With the operator of the transposed aggregation and the transposed weight the same kernel is the
// layer's input gradient (dx = [dz | P dz | Q dz] [W_root ; W_1 ; W_2], see ops.py), whose epilogue
// adds the gradient that reaches x through the concatenation (`addend`).
//
// Operator: CSR with three segments per node (`ptr` [3 rows + 1], `src`, `w`), segment 3i being the
// node itself (one entry, weight 1; taken through TMA instead of the list), 3i+1 and 3i+2 the two
// weighted neighbour lists.  `order` holds, for every group of 32 consecutive rows, a permutation of
// 0..31 with the longest lists first, so that the rows one warp gathers at the same time need the same
// number of rounds.
//
// Work split: the gather is bound by the L2 -> SM request rate (~7.5 us per 128 rows at C = 64), so
// the rows — not the tiles — are divided evenly: every CTA owns a contiguous range of 32-row groups
// and walks it in tiles of up to 128 rows (the UMMA M); the last tile of a range may be partial.
//
// Shared memory (C <= 64): the split weight is RESIDENT (loaded once per CTA, <= 96 KB), three
// operand stages [in_h | out_h | own_h] of 32 KB (one 32-column slice h of the three parts), two
// buffers for the tile's slice of the operator (prefetched one tile ahead with cp.async).
//
// Persistent kernel, one CTA per SM, 22 warps:
//   warp 0       TMA: the weight once, then per tile and slice the raw own-features tile
//   warp 1       MMA issuer (tcgen05.mma.kind::tf32, 3xTF32 exactly as in linear_tc.cu)
//   warps 2-17   operand builders: 8 lanes per row walk BOTH neighbour lists of the row (contiguous in
//                the CSR) with eight float4 gathers in flight per lane, hi / lo split, swizzled
//                st.shared into the in- and the out-stage; then the own-features tile, which arrived
//                through TMA meanwhile, is split in place; in training the fp32 operand rows also go
//                to `stacked` (the weight gradient reads them)
//   warps 18-21  epilogue: TMEM -> registers (+ addend, + bias, ReLU) -> global (one row per thread)
#include <cstdlib>
#include <cstring>
#include "common.cuh"
#include "tc_common.cuh"
#include "../../include/dgmc_b200.h"

namespace dgmc {
namespace tc {

constexpr int RC_BM = 128;
constexpr int RC_KB = 32;
constexpr int RC_BN = 64;
constexpr int RC_MAX_C = 64;                       // the resident weight must fit: 3 C x 64 x 2 x 4 B
constexpr int RC_A_BYTES = RC_BM * RC_KB * 4;      // 16 KB
constexpr int RC_B_BYTES = RC_BN * RC_KB * 4;      // 8 KB: one k block of W_hi (or W_lo)
constexpr int RC_STAGE_BYTES = 2 * RC_A_BYTES;     // operand stage: hi + lo
constexpr int RC_NSTAGE = 3;                       // = the k blocks of one 32-column slice: [in | out | own]
constexpr int RC_MAX_KB = 3 * RC_MAX_C / RC_KB;    // 6 k blocks
constexpr int RC_W_BYTES = RC_MAX_KB * 2 * RC_B_BYTES;   // 96 KB
constexpr int RC_BUILDERS = 16;                    // operand-builder warps
constexpr int RC_THREADS = 32 * (2 + RC_BUILDERS + 4);
constexpr int RC_ACC = 2;                          // accumulator pairs [H | L] the k8 steps alternate between
constexpr int RC_ACC_COLS = 2 * RC_ACC * RC_BN;    // 256: two sets fit the 512 TMEM columns
constexpr int RC_TMEM_COLS = 512;
constexpr int RC_PTR_INTS = 3 * RC_BM + 1;         // segment boundaries of one tile
constexpr int RC_PTR_PAD = 400;
constexpr int RC_ENT_CAP = 1792;                   // operator entries of one tile kept in shared memory
constexpr size_t RC_SMEM = (size_t)RC_W_BYTES + RC_NSTAGE * RC_STAGE_BYTES +
                           2 * (RC_PTR_PAD * 4 + RC_ENT_CAP * 8) + 256 + 1024;
static_assert(RC_SMEM <= 232448, "relconv_tc: shared memory budget");

__device__ __forceinline__ float rc_round_tf32(float v) {
    return __uint_as_float((__float_as_uint(v) + 0x1000u) & 0xffffe000u);
}
__device__ __forceinline__ void rc_split(const float4 v, float4& h, float4& l) {
    h.x = rc_round_tf32(v.x); l.x = rc_round_tf32(v.x - h.x);
    h.y = rc_round_tf32(v.y); l.y = rc_round_tf32(v.y - h.y);
    h.z = rc_round_tf32(v.z); l.z = rc_round_tf32(v.z - h.z);
    h.w = rc_round_tf32(v.w); l.w = rc_round_tf32(v.w - h.w);
}
__device__ __forceinline__ void rc_cp_async4(void* dst, const void* src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void rc_builders_sync() {
    asm volatile("bar.sync 1, %0;" ::"n"(RC_BUILDERS * 32) : "memory");
}

struct RelConvArgs {
    const int* ptr; const int* src; const float* w;     // operator CSR (3 segments per node)
    const int* order;                                   // per 32-row group: its rows, longest lists first
    const float* x; long long ldx;                      // [rows, c]
    float* stacked; long long lds;                      // nullable: [rows, 3c] operand rows (fp32)
    const float* bias; int relu;
    const float* addend; long long lda;                 // nullable: added to the result (may alias y)
    float* y; long long ldy;
    long long rows; int c, n_out, k_blocks;
    int groups;                                         // 32-row groups in total
    int debug;   // ablation switches for profiles/relconv_microbench.py (DGMC_RELCONV_DEBUG), 0 in production
};

// this CTA's range of 32-row groups [g0, g1)
__device__ __forceinline__ void rc_range(const RelConvArgs& a, int& g0, int& g1) {
    const int base = a.groups / (int)gridDim.x, extra = a.groups % (int)gridDim.x, b = (int)blockIdx.x;
    g0 = b * base + (b < extra ? b : extra);
    g1 = g0 + base + (b < extra ? 1 : 0);
}

__global__ void __launch_bounds__(RC_THREADS, 1)
relconv_tc_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_whi,
                  const __grid_constant__ CUtensorMap map_wlo, const RelConvArgs a) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* w_res = smem;                                            // [k block][W_hi 8 KB | W_lo 8 KB]
    uint8_t* stages = smem + RC_W_BYTES;                              // [3][A_hi 16 KB | A_lo 16 KB]
    int* ptr_s = reinterpret_cast<int*>(stages + RC_NSTAGE * RC_STAGE_BYTES);   // [2][RC_PTR_PAD]
    int* src_s = ptr_s + 2 * RC_PTR_PAD;                              // [2][RC_ENT_CAP]
    float* w_s = reinterpret_cast<float*>(src_s + 2 * RC_ENT_CAP);    // [2][RC_ENT_CAP]
    uint64_t* bars = reinterpret_cast<uint64_t*>(w_s + 2 * RC_ENT_CAP);
    uint64_t* w_full = bars;                     // resident weight landed
    uint64_t* own_full = bars + 1;               // own-features tile landed in stage 2
    uint64_t* built = bars + 2;                  // [3] operand stage complete (one arrival per builder warp)
    uint64_t* empty = bars + 5;                  // [3] MMAs of the stage retired
    uint64_t* acc_full = bars + 8;               // [2]
    uint64_t* acc_empty = bars + 10;             // [2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 12);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int halves = a.c / RC_KB;              // 32-column slices per part
    int g0, g1;
    rc_range(a, g0, g1);
    const int n_tiles = (g1 - g0 + 3) / 4;       // tiles of up to four 32-row groups

    if (warp == 0 && lane == 0) {
        mbar_init(w_full, 1);
        mbar_init(own_full, 1);
        for (int s = 0; s < RC_NSTAGE; ++s) { mbar_init(&built[s], RC_BUILDERS); mbar_init(&empty[s], 1); }
        for (int b = 0; b < 2; ++b) { mbar_init(&acc_full[b], 1); mbar_init(&acc_empty[b], 4); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                         smem_u32(tmem_slot)), "r"(RC_TMEM_COLS));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_d = *tmem_slot;

    if (warp == 0) {
        if (lane == 0 && n_tiles > 0) {
            // ================= TMA =================
            mbar_expect_tx(w_full, (uint32_t)(a.k_blocks * 2 * RC_B_BYTES));
            for (int kb = 0; kb < a.k_blocks; ++kb) {
                tma_load_2d(w_res + (size_t)kb * 2 * RC_B_BYTES, &map_whi, w_full, kb * RC_KB, 0);
                tma_load_2d(w_res + (size_t)kb * 2 * RC_B_BYTES + RC_B_BYTES, &map_wlo, w_full, kb * RC_KB, 0);
            }
            uint32_t phase = 0;
            for (int tile = 0; tile < n_tiles; ++tile) {
                const int m0 = (g0 + 4 * tile) * 32;
                for (int half = 0; half < halves; ++half) {
                    mbar_wait(&empty[2], phase ^ 1u);
                    mbar_expect_tx(own_full, (uint32_t)RC_A_BYTES);
                    tma_load_2d(stages + 2 * RC_STAGE_BYTES, &map_x, own_full, half * RC_KB, m0);
                    phase ^= 1u;
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0 && n_tiles > 0) {
            // ================= MMA issuer (see linear_tc.cu) =================
            constexpr uint32_t idesc_wide = (1u << 4) | (2u << 7) | (2u << 10) |
                                            ((uint32_t)((2 * RC_BN) >> 3) << 17) | ((uint32_t)(RC_BM >> 4) << 24);
            constexpr uint32_t idesc_lo = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(RC_BN >> 3) << 17) |
                                          ((uint32_t)(RC_BM >> 4) << 24);
            mbar_wait(w_full, 0);
            uint32_t phase = 0;
            for (int tile = 0; tile < n_tiles; ++tile) {
                const int buf = tile & 1, use = tile >> 1;
                mbar_wait(&acc_empty[buf], (uint32_t)((use & 1) ^ 1));
                tcgen05_fence_after();
                const uint32_t acc = tmem_d + (uint32_t)(buf * RC_ACC_COLS);
                int step = 0;
                for (int half = 0; half < halves; ++half) {
#pragma unroll 1
                    for (int s = 0; s < RC_NSTAGE; ++s) {
                        // stage s of slice `half`: part (s + 1) % 3 of the operand = k block part * halves + half of W
                        const int part = (s + 1) % 3;
                        mbar_wait(&built[s], phase);
                        tcgen05_fence_after();
                        const uint32_t st = smem_u32(stages + (size_t)s * RC_STAGE_BYTES);
                        const uint64_t d_ahi = umma_desc_sw128(st), d_alo = umma_desc_sw128(st + RC_A_BYTES);
                        const uint64_t d_b = umma_desc_sw128(smem_u32(w_res + (size_t)(part * halves + half) * 2 * RC_B_BYTES));
#pragma unroll
                        for (int ks = 0; ks < RC_KB / 8; ++ks, ++step) {
                            const uint64_t o = (uint64_t)(2 * ks);
                            const uint32_t d = acc + (uint32_t)((step % RC_ACC) * 2 * RC_BN);
                            tcgen05_mma_tf32(d, d_ahi + o, d_b + o, idesc_wide, (uint32_t)(step >= RC_ACC));
                            tcgen05_mma_tf32(d + RC_BN, d_alo + o, d_b + o, idesc_lo, 1u);
                        }
                        tcgen05_commit(&empty[s]);
                    }
                    phase ^= 1u;
                }
                tcgen05_commit(&acc_full[buf]);
            }
        }
    } else if (warp < 2 + RC_BUILDERS) {
        // ================= operand builders =================
        const int t = threadIdx.x - 64;                 // 0 .. 511
        const int grp = t >> 3, l8 = t & 7;             // 64 lane groups of 8
        const long long total_seg = 3 * a.rows;

        // operator entries of a tile -> shared memory (asynchronously; consumed one tile later).  The
        // entry range [e_lo, e_hi) of a tile is read one tile earlier still, so that no load sits
        // between the barrier and the copies.
        auto bounds = [&](int tile, int& e_lo, int& e_hi) {
            e_lo = e_hi = 0;
            if (tile >= n_tiles) return;
            long long seg0 = 3LL * (g0 + 4 * tile) * 32, seg1 = seg0 + 3 * RC_BM;
            if (seg0 > total_seg) seg0 = total_seg;
            if (seg1 > total_seg) seg1 = total_seg;
            e_lo = __ldg(a.ptr + seg0);
            e_hi = __ldg(a.ptr + seg1);
        };
        auto prefetch = [&](int tile, int buf, int e_lo, int e_hi) {
            if (tile >= n_tiles) return;
            const long long seg0 = 3LL * (g0 + 4 * tile) * 32;
            if (t < RC_PTR_INTS) {
                long long sgi = seg0 + t;
                if (sgi > total_seg) sgi = total_seg;
                rc_cp_async4(ptr_s + buf * RC_PTR_PAD + t, a.ptr + sgi);
            }
            if (e_hi - e_lo <= RC_ENT_CAP) {
                for (int e = e_lo + t; e < e_hi; e += RC_BUILDERS * 32) {
                    rc_cp_async4(src_s + buf * RC_ENT_CAP + (e - e_lo), a.src + e);
                    rc_cp_async4(w_s + buf * RC_ENT_CAP + (e - e_lo), a.w + e);
                }
            }
        };
        int nb_lo, nb_hi;
        bounds(0, nb_lo, nb_hi);
        prefetch(0, 0, nb_lo, nb_hi);
        asm volatile("cp.async.commit_group;" ::: "memory");
        bounds(1, nb_lo, nb_hi);

        uint32_t phase = 0;
        for (int tile = 0; tile < n_tiles; ++tile) {
            const int buf = tile & 1;
            const int gt = g0 + 4 * tile;                       // first 32-row group of the tile
            const long long m0 = 32LL * gt;
            const int valid = 32 * (g1 - gt < 4 ? g1 - gt : 4); // rows of the tile that belong to this CTA
            asm volatile("cp.async.wait_group 0;" ::: "memory");
            rc_builders_sync();          // this tile's lists are in shared memory; the other buffer is free
            prefetch(tile + 1, buf ^ 1, nb_lo, nb_hi);
            asm volatile("cp.async.commit_group;" ::: "memory");
            bounds(tile + 2, nb_lo, nb_hi);
            const int* ps = ptr_s + buf * RC_PTR_PAD;
            const int e_base = ps[0];
            const bool staged = ps[3 * RC_BM] - e_base <= RC_ENT_CAP;
            const int* srcp = staged ? src_s + buf * RC_ENT_CAP : a.src + e_base;
            const float* wp = staged ? w_s + buf * RC_ENT_CAP : a.w + e_base;
            // the two rows of this lane group: slots 2 grp and 2 grp + 1 of the tile's order
            int my_row[2];
#pragma unroll
            for (int pass = 0; pass < 2; ++pass) {
                const int u = 2 * grp + pass, q = u >> 5;
                my_row[pass] = 32 * q + (32 * q < valid ? __ldg(a.order + (long long)(gt + q) * 32 + (u & 31)) : (u & 31));
            }

            for (int half = 0; half < halves; ++half) {
                uint8_t* st_in = stages;
                uint8_t* st_out = stages + RC_STAGE_BYTES;
                uint8_t* st_own = stages + 2 * RC_STAGE_BYTES;
                // ---- both neighbour lists of a row in one pass: eight gathers in flight per lane ----
                mbar_wait(&empty[0], phase ^ 1u);    // the MMAs that read these stages last have retired
                mbar_wait(&empty[1], phase ^ 1u);
                const float* xcol = a.x + half * RC_KB + 4 * l8;
#pragma unroll 1
                for (int pass = 0; pass < 2; ++pass) {
                    const int r = my_row[pass];
                    float4 acc_in = make_float4(0.f, 0.f, 0.f, 0.f), acc_out = acc_in;
                    int e = ps[3 * r + 1] - e_base;
                    const int e_mid = ps[3 * r + 2] - e_base;
                    int e_end = ps[3 * r + 3] - e_base;
                    if (r >= valid || (a.debug & 1)) e_end = e;      // rows of another CTA's range
                    while (__any_sync(0xffffffffu, e < e_end)) {
                        float4 v[8]; float wt[8];
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const bool on = e + j < e_end;
                            const int sj = on ? srcp[e + j] : 0;
                            wt[j] = on ? wp[e + j] : 0.f;
                            v[j] = on ? ldg4(xcol + (long long)sj * a.ldx) : make_float4(0.f, 0.f, 0.f, 0.f);
                        }
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const float wi = e + j < e_mid ? wt[j] : 0.f, wo = e + j < e_mid ? 0.f : wt[j];
                            acc_in.x = fmaf(wi, v[j].x, acc_in.x); acc_in.y = fmaf(wi, v[j].y, acc_in.y);
                            acc_in.z = fmaf(wi, v[j].z, acc_in.z); acc_in.w = fmaf(wi, v[j].w, acc_in.w);
                            acc_out.x = fmaf(wo, v[j].x, acc_out.x); acc_out.y = fmaf(wo, v[j].y, acc_out.y);
                            acc_out.z = fmaf(wo, v[j].z, acc_out.z); acc_out.w = fmaf(wo, v[j].w, acc_out.w);
                        }
                        e += 8;
                    }
                    const int pos = r * 8 + (l8 ^ (r & 7));
                    float4 h, l;
                    rc_split(acc_in, h, l);
                    reinterpret_cast<float4*>(st_in)[pos] = h;
                    reinterpret_cast<float4*>(st_in + RC_A_BYTES)[pos] = l;
                    rc_split(acc_out, h, l);
                    reinterpret_cast<float4*>(st_out)[pos] = h;
                    reinterpret_cast<float4*>(st_out + RC_A_BYTES)[pos] = l;
                    if (a.stacked && r < valid && m0 + r < a.rows) {
                        float* dst = a.stacked + (m0 + r) * a.lds + half * RC_KB + 4 * l8;
                        *reinterpret_cast<float4*>(dst + a.c) = acc_in;
                        *reinterpret_cast<float4*>(dst + 2 * a.c) = acc_out;
                    }
                }
                fence_proxy_async_smem();     // st.shared above -> UMMA operand reads
                __syncwarp();
                if (lane == 0) { mbar_arrive(&built[0]); mbar_arrive(&built[1]); }
                // ---- the node's own features arrived through TMA: split in place ----
                mbar_wait(own_full, phase);
                float4* a_hi = reinterpret_cast<float4*>(st_own);
                float4* a_lo = reinterpret_cast<float4*>(st_own + RC_A_BYTES);
#pragma unroll
                for (int i = 0; i < RC_A_BYTES / 16 / (RC_BUILDERS * 32); ++i) {
                    const int f = t + i * RC_BUILDERS * 32;
                    const float4 v = a_hi[f];
                    float4 h, l;
                    rc_split(v, h, l);
                    a_hi[f] = h;
                    a_lo[f] = l;
                    if (a.stacked) {
                        const int row = f >> 3, chunk = (f & 7) ^ (row & 7);
                        if (row < valid && m0 + row < a.rows)
                            *reinterpret_cast<float4*>(a.stacked + (m0 + row) * a.lds + half * RC_KB + 4 * chunk) = v;
                    }
                }
                fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) mbar_arrive(&built[2]);
                phase ^= 1u;
            }
        }
        asm volatile("cp.async.wait_group 0;" ::: "memory");
    } else {
        // ================= epilogue (warps 18-21): one output row per thread =================
        const int quarter = warp & 3;                  // TMEM lanes this warp may read
        for (int tile = 0; tile < n_tiles; ++tile) {
            const int buf = tile & 1, use = tile >> 1;
            const int gt = g0 + 4 * tile;
            const int valid = 32 * (g1 - gt < 4 ? g1 - gt : 4);
            mbar_wait(&acc_full[buf], (uint32_t)(use & 1));
            tcgen05_fence_after();
            const long long row = 32LL * gt + quarter * 32 + lane;
            const bool live = quarter * 32 < valid && row < a.rows;
#pragma unroll 1
            for (int cb = 0; cb < RC_BN; cb += 32) {
                if (cb >= a.n_out) break;
                float r[32];
                {
                    const uint32_t lane_base = tmem_d + ((uint32_t)(quarter * 32) << 16) +
                                               (uint32_t)(buf * RC_ACC_COLS + cb);
                    // (H_0 + H_1) + L_0 + L_1, one 32-column load at a time (the register budget of
                    // this 704-thread kernel is 88 per thread)
                    float q[32];
                    tmem_ld_32x32b_x32(lane_base, r);
                    tmem_ld_32x32b_x32(lane_base + (uint32_t)(2 * RC_BN), q);
#pragma unroll
                    for (int j = 0; j < 32; ++j) r[j] += q[j];
                    tmem_ld_32x32b_x32(lane_base + (uint32_t)RC_BN, q);
#pragma unroll
                    for (int j = 0; j < 32; ++j) r[j] += q[j];
                    tmem_ld_32x32b_x32(lane_base + (uint32_t)(3 * RC_BN), q);
#pragma unroll
                    for (int j = 0; j < 32; ++j) r[j] += q[j];
                }
                if (cb + 32 >= RC_BN || cb + 32 >= a.n_out) {
                    tcgen05_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&acc_empty[buf]);
                }
                if (!live || (a.debug & 8)) continue;
                const bool whole = cb + 32 <= a.n_out;
                if (a.addend) {
                    const float* ad = a.addend + row * a.lda + cb;
                    if (whole) {
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const float4 p = *reinterpret_cast<const float4*>(ad + 4 * j);
                            r[4 * j] += p.x; r[4 * j + 1] += p.y; r[4 * j + 2] += p.z; r[4 * j + 3] += p.w;
                        }
                    } else {
#pragma unroll
                        for (int j = 0; j < 32; ++j) if (cb + j < a.n_out) r[j] += ad[j];
                    }
                }
                if (a.bias) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) r[j] += (cb + j < a.n_out) ? __ldg(a.bias + cb + j) : 0.f;
                }
                if (a.relu) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) r[j] = fmaxf(r[j], 0.f);
                }
                float* dst = a.y + row * a.ldy + cb;
                if (whole && (a.ldy & 3) == 0 && (reinterpret_cast<uintptr_t>(a.y) & 15) == 0) {
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        *reinterpret_cast<float4*>(dst + 4 * j) = make_float4(r[4 * j], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]);
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        if (cb + j < a.n_out) dst[j] = r[j];
                }
            }
        }
    }
    tcgen05_fence_before();
    __syncthreads();
    if (warp == 1) {
        tcgen05_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"(RC_TMEM_COLS));
    }
}

typedef CUresult (*EncodeTiledFnR)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                   CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                   CUtensorMapFloatOOBfill);

static EncodeTiledFnR rc_encode_tiled() {
    static EncodeTiledFnR fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFnR>(p);
    }
    return fn;
}

static int rc_make_map(CUtensorMap* map, const void* base, long long rows, long long cols, long long ld,
                       int box_rows) {
    memset(map, 0, sizeof(*map));
    ensure_context();
    EncodeTiledFnR enc = rc_encode_tiled();
    if (!enc) { set_error("relconv_tc: cuTensorMapEncodeTiled not available from the driver"); return DGMC_ERR_CUDA; }
    const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)ld * 4};
    const cuuint32_t box[2] = {(cuuint32_t)RC_KB, (cuuint32_t)box_rows};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(base), dims, strides,
                           box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                           CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("relconv_tc: cuTensorMapEncodeTiled failed (%d)", (int)r); return DGMC_ERR_CUDA; }
    return DGMC_OK;
}

}  // namespace tc
}  // namespace dgmc

using namespace dgmc;

extern "C" int dgmc_relconv_tc_supported(int64_t rows, int c, int n_out, int64_t ldx, const float* x) {
    return rows >= 1 && rows <= ((int64_t)1 << 30) && c >= 32 && c % 32 == 0 && c <= tc::RC_MAX_C && n_out >= 1 &&
           n_out <= tc::RC_BN && ldx >= c && ldx % 4 == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0;
}

extern "C" int dgmc_relconv_tc(const int32_t* ptr, const int32_t* src, const float* w, const int32_t* order,
                               const float* x, int64_t ldx,
                               const void* split_weight, const float* bias, int relu, const float* addend,
                               int64_t lda, float* stacked, int64_t lds, float* y, int64_t ldy, int64_t rows,
                               int c, int n_out, dgmc_stream_t stream) {
    DGMC_REQUIRE(ptr && src && w && order && x && split_weight && y && ldy >= n_out, "relconv_tc: bad arguments");
    if (!dgmc_relconv_tc_supported(rows, c, n_out, ldx, x)) {
        set_error("relconv_tc: unsupported operand (rows=%lld c=%d n_out=%d ldx=%lld; needs c = 32 or 64, "
                  "n_out <= 64, 16-byte aligned x rows)", (long long)rows, c, n_out, (long long)ldx);
        return DGMC_ERR_UNSUPPORTED;
    }
    DGMC_REQUIRE(!stacked || (lds >= 3 * c && lds % 4 == 0 && (reinterpret_cast<uintptr_t>(stacked) & 15) == 0),
                 "relconv_tc: stacked needs a 16-byte aligned base and pitch");
    DGMC_REQUIRE(!addend || (lda >= n_out && lda % 4 == 0 && (reinterpret_cast<uintptr_t>(addend) & 15) == 0),
                 "relconv_tc: addend needs a 16-byte aligned base and pitch");
    const int k_in = 3 * c, k_pad = k_in;       // c % 32 == 0
    const int n_pad = tc::RC_BN;
    const float* hi = static_cast<const float*>(split_weight);
    const float* lo = hi + (size_t)n_pad * k_pad;
    CUtensorMap mx, mh, ml;
    if (int e = tc::rc_make_map(&mx, x, rows, c, ldx, tc::RC_BM)) return e;
    if (int e = tc::rc_make_map(&mh, hi, n_pad, k_pad, k_pad, tc::RC_BN)) return e;
    if (int e = tc::rc_make_map(&ml, lo, n_pad, k_pad, k_pad, tc::RC_BN)) return e;
    static DeviceOnce configured;
    if (!configured.done()) {
        if (cudaFuncSetAttribute(tc::relconv_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)tc::RC_SMEM) != cudaSuccess) {
            set_error("relconv_tc: cannot set %zu bytes of dynamic shared memory", tc::RC_SMEM);
            return DGMC_ERR_CUDA;
        }
        configured.mark();
    }
    tc::RelConvArgs a;
    a.ptr = ptr; a.src = src; a.w = w; a.order = order; a.x = x; a.ldx = ldx; a.stacked = stacked; a.lds = lds;
    a.bias = bias; a.relu = relu; a.addend = addend; a.lda = lda; a.y = y; a.ldy = ldy; a.rows = rows;
    a.c = c; a.n_out = n_out; a.k_blocks = k_in / tc::RC_KB;
    static const int debug = getenv("DGMC_RELCONV_DEBUG") ? atoi(getenv("DGMC_RELCONV_DEBUG")) : 0;
    a.debug = debug;
    a.groups = (int)((rows + 31) / 32);
    // one CTA per SM; with fewer than four 32-row groups per SM, fewer CTAs (a tile holds four groups)
    const int want = (a.groups + 3) / 4;
    const int grid = want < sm_count() ? want : sm_count();
    tc::relconv_tc_kernel<<<grid, tc::RC_THREADS, tc::RC_SMEM, stream>>>(mx, mh, ml, a);
    return check_launch("relconv_tc");
}
