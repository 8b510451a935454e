// Weight gradient dW[m, n] = sum_k G[k, m] X[k, n] on the 5th-gen tensor cores with fp32-level
// accuracy ("3xTF32"): every fp32 operand is split into hi = x rounded to tf32 (19 bits) and
// lo = x - hi rounded to tf32, and the product is accumulated as hi*hi + hi*lo + lo*hi in the fp32
// TMEM accumulator (the dropped lo*lo term and the rounding of lo are ~2^-24 relative, zero-mean).  Used for the
// BACKWARD pass only: gradients carry a 5e-4 tolerance, the forward GEMMs that decide S_L stay in
// IEEE fp32 (library SGEMM).
//
// The reduction runs over K = #nodes, the output is tiny, so the grid is (128-wide tiles of n) x
// (K splits).  Per CTA and 32-row chunk of its K slice: all threads load G[32 x M] and X[32 x 128]
// from global (float4, coalesced), split them and scatter hi / lo into K-major, 128-byte-swizzled
// shared-memory operand tiles (A = X^T: 128 rows(n) x 32 k, B = G^T: M_pad rows(m) x 32 k; two
// stages), fence to the async proxy, and one thread issues 4 (k8 steps) x 3 tcgen05.mma.kind::tf32
// (M=128, N=M_pad, K=8) into one TMEM accumulator; tcgen05.commit frees the stage.  The staging of
// chunk c+1 overlaps the MMAs of chunk c.  At the end every thread reads its accumulator row
// (tcgen05.ld) and writes the partial tile, which dgmc_wgrad_reduce adds up in split order.
#include "common.cuh"
#include "tc_common.cuh"
#include "../../include/dgmc_b200.h"

namespace dgmc {
namespace tc {

constexpr int WT_KC = 32;                 // K rows per chunk = one 128-byte swizzled row of tf32
constexpr int WT_TN = 128;                // input channels per CTA (UMMA M)
constexpr int WT_THREADS = 128;
constexpr int WT_A_BYTES = WT_TN * WT_KC * 4;   // 16 KB
constexpr int WT_MAX_M = 128;                   // output channels (UMMA N), padded to 64 or 128
constexpr int WT_B_BYTES = WT_MAX_M * WT_KC * 4;  // 16 KB

// byte offset of element (row, k) in a K-major [rows x 32 fp32] tile with 128-byte swizzle
__device__ __forceinline__ uint32_t sw128_offset(int row, int k) {
    return (uint32_t)((row >> 3) * 1024 + (row & 7) * 128 + ((((k >> 2) ^ (row & 7)) & 7) << 4) + ((k & 3) << 2));
}

__device__ __forceinline__ void split_tf32(float v, float& hi, float& lo) {
    // round (not truncate) both parts to tf32: truncation errs towards zero in every product, a
    // bias that does not average out over the reduction (see linear_tc.cu)
    hi = __uint_as_float((__float_as_uint(v) + 0x1000u) & 0xffffe000u);
    lo = __uint_as_float((__float_as_uint(v - hi) + 0x1000u) & 0xffffe000u);
}

__global__ void __launch_bounds__(WT_THREADS, 2)
wgrad_tc_kernel(const float* __restrict__ g, long long ldg_, const float* __restrict__ x, long long ldx,
                long long K, int M, int N, int m_pad, int splits, float* __restrict__ partial,
                float* __restrict__ partial_b) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const int b_bytes = m_pad * WT_KC * 4;                       // 8 KB (m_pad = 64) or 16 KB
    const int stage_bytes = 2 * WT_A_BYTES + 2 * b_bytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 2 * stage_bytes);
    uint64_t* stage_free = bars;      // [2]  arrives via tcgen05.commit
    uint64_t* all_done = bars + 2;    // [1]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 3);
    float* bias_red = reinterpret_cast<float*>(bars + 4);  // [8][64] column sums of G

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int n0 = blockIdx.x * WT_TN, split = blockIdx.y;
    const long long per = ((K + splits - 1) / splits + WT_KC - 1) / WT_KC * WT_KC;
    const long long k_begin = per * split;
    const long long k_end = k_begin + per < K ? k_begin + per : K;

    if (tid == 0) {
        mbar_init(&stage_free[0], 1);
        mbar_init(&stage_free[1], 1);
        mbar_init(all_done, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                         smem_u32(tmem_slot)), "r"(128));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_d = *tmem_slot;
    // kind::tf32, D = f32, A = B = tf32, both K-major, N = m_pad, M = 128
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(m_pad >> 3) << 17) |
                           ((uint32_t)(WT_TN >> 4) << 24);

    // Staging map: thread t owns operand ROW t (A: input channel n0 + t; B: output channel t % m_pad,
    // with the two halves of the block taking alternate k-quadruples when m_pad = 64).  A warp's
    // global loads are 32 consecutive channels of one k row (128 bytes, coalesced); each thread
    // gathers 4 consecutive k into one 16-byte st.shared, which the 128-byte swizzle spreads over
    // the banks (lanes of a quarter-warp hit 8 different 16-byte chunks).
    const int b_row = tid % m_pad;
    const int b_k0 = m_pad == 64 ? (tid / 64) * 4 : 0;     // first k-quadruple of this thread
    const int b_kstep = m_pad == 64 ? 8 : 4;
    const bool a_live = n0 + tid < N, b_live = b_row < M;
    float bsum = 0.f;                                       // column sum of G for channel b_row
    const int n_chunks = (int)((k_end - k_begin + WT_KC - 1) / WT_KC);
    float na[WT_KC], nb[WT_KC];       // next chunk: 32 k of this thread's X channel, up to 32 k of its G channel
    auto fetch = [&](long long k0) {
#pragma unroll
        for (int kk = 0; kk < WT_KC; ++kk)
            na[kk] = (a_live && k0 + kk < k_end) ? __ldg(x + (k0 + kk) * ldx + n0 + tid) : 0.f;
#pragma unroll
        for (int j = 0; j < WT_KC / 4; ++j) {                 // m_pad = 64: only j < 4 is in range
            const int kq = b_k0 + j * b_kstep;
#pragma unroll
            for (int u = 0; u < 4; ++u)
                nb[4 * j + u] = (b_live && kq < WT_KC && k0 + kq + u < k_end)
                                    ? __ldg(g + (k0 + kq + u) * ldg_ + b_row) : 0.f;
        }
    };
    if (n_chunks > 0) fetch(k_begin);
    for (int c = 0; c < (n_chunks > 0 ? n_chunks : 0); ++c) {
        const int stage = c & 1;
        uint8_t* a_hi = smem + stage * stage_bytes;
        uint8_t* a_lo = a_hi + WT_A_BYTES;
        uint8_t* b_hi = a_lo + WT_A_BYTES;
        uint8_t* b_lo = b_hi + b_bytes;
        if (c >= 2) mbar_wait(&stage_free[stage], (uint32_t)(((c >> 1) - 1) & 1));  // MMAs of chunk c-2 retired
        // the values of this chunk were fetched one iteration ago; fetch the next chunk now so that
        // its global-memory latency hides behind this chunk's split / st.shared / MMA issue
        float va[WT_KC], vb[WT_KC];
#pragma unroll
        for (int i = 0; i < WT_KC; ++i) va[i] = na[i];
#pragma unroll
        for (int i = 0; i < WT_KC; ++i) vb[i] = nb[i];
        if (c + 1 < n_chunks) fetch(k_begin + (long long)(c + 1) * WT_KC);
        // ---- X chunk [32 x 128] -> A tiles (row = input channel, column = k) ----
#pragma unroll
        for (int kq = 0; kq < WT_KC; kq += 4) {
            float4 hi, lo;
            split_tf32(va[kq], hi.x, lo.x); split_tf32(va[kq + 1], hi.y, lo.y);
            split_tf32(va[kq + 2], hi.z, lo.z); split_tf32(va[kq + 3], hi.w, lo.w);
            const uint32_t o = sw128_offset(tid, kq);
            *reinterpret_cast<float4*>(a_hi + o) = hi;
            *reinterpret_cast<float4*>(a_lo + o) = lo;
        }
        // ---- G chunk [32 x m_pad] -> B tiles (row = output channel, column = k) ----
#pragma unroll
        for (int j = 0; j < WT_KC / 4; ++j) {                 // k-quadruples of this thread (m_pad = 64: 4 in range)
            const int kq = b_k0 + j * b_kstep;
            if (kq < WT_KC) {
                float4 hi, lo;
                split_tf32(vb[4 * j], hi.x, lo.x); split_tf32(vb[4 * j + 1], hi.y, lo.y);
                split_tf32(vb[4 * j + 2], hi.z, lo.z); split_tf32(vb[4 * j + 3], hi.w, lo.w);
                bsum += (vb[4 * j] + vb[4 * j + 1]) + (vb[4 * j + 2] + vb[4 * j + 3]);
                const uint32_t o = sw128_offset(b_row, kq);
                *reinterpret_cast<float4*>(b_hi + o) = hi;
                *reinterpret_cast<float4*>(b_lo + o) = lo;
            }
        }
        fence_proxy_async_smem();     // st.shared above -> UMMA operand reads below
        __syncthreads();
        if (tid == 0) {
            tcgen05_fence_after();
            const uint64_t d_ahi = umma_desc_sw128(smem_u32(a_hi)), d_alo = umma_desc_sw128(smem_u32(a_lo));
            const uint64_t d_bhi = umma_desc_sw128(smem_u32(b_hi)), d_blo = umma_desc_sw128(smem_u32(b_lo));
#pragma unroll
            for (int ks = 0; ks < WT_KC / 8; ++ks) {     // +32 bytes per K=8 step inside the swizzled row
                const uint64_t o = (uint64_t)(2 * ks);
                tcgen05_mma_tf32(tmem_d, d_ahi + o, d_bhi + o, idesc, (uint32_t)((c | ks) != 0));
                tcgen05_mma_tf32(tmem_d, d_ahi + o, d_blo + o, idesc, 1u);
                tcgen05_mma_tf32(tmem_d, d_alo + o, d_bhi + o, idesc, 1u);
            }
            tcgen05_commit(&stage_free[stage]);
            if (c == n_chunks - 1) tcgen05_commit(all_done);
        }
    }

    // ---- write the partial tile: D'[n, m] -> partial[split][m][n0 + n] ----
    float* out = partial + (size_t)split * M * N;
    if (n_chunks > 0) {
        mbar_wait(all_done, 0);
        tcgen05_fence_after();
        const int n = n0 + warp * 32 + lane;            // accumulator row of this thread
        for (int cb = 0; cb < m_pad; cb += 32) {
            float r[32];
            tmem_ld_32x32b_x32(tmem_d + ((uint32_t)(warp * 32) << 16) + (uint32_t)cb, r);
            if (n < N) {
#pragma unroll
                for (int j = 0; j < 32; ++j)
                    if (cb + j < M) out[(size_t)(cb + j) * N + n] = r[j];
            }
        }
    } else {
        for (int i = tid; i < M * WT_TN; i += WT_THREADS) {
            const int m = i / WT_TN, n = n0 + i % WT_TN;
            if (n < N) out[(size_t)m * N + n] = 0.f;
        }
    }
    // ---- column sums of G (db) ----
    if (partial_b && blockIdx.x == 0) {
        __syncthreads();
        bias_red[tid] = bsum;                             // [128]: for m_pad = 64 two halves per channel
        __syncthreads();
        if (tid < M) partial_b[(size_t)split * M + tid] = m_pad == 64 ? bias_red[tid] + bias_red[tid + 64] : bias_red[tid];
    }
    tcgen05_fence_before();
    __syncthreads();
    if (warp == 1) {
        tcgen05_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"(128));
    }
}

static size_t wt_smem(int m_pad) { return 2 * (size_t)(2 * WT_A_BYTES + 2 * m_pad * WT_KC * 4) + 64 + 128 * 4 + 1024; }

}  // namespace tc
}  // namespace dgmc

using namespace dgmc;

static int wgrad_tc_splits_impl(long long K, int N) {
    const int tiles = (N + tc::WT_TN - 1) / tc::WT_TN;
    int splits = (2 * sm_count() + tiles - 1) / tiles;     // one wave at 2 CTAs per SM
    const long long max_by_k = (K + 8 * tc::WT_KC - 1) / (8 * tc::WT_KC);   // >= 256 rows per split
    if (splits > max_by_k) splits = (int)max_by_k;
    if (splits < 1) splits = 1;
    if (splits > 1024) splits = 1024;
    return splits;
}

extern "C" int dgmc_wgrad_tc_supported(int m, int n) { return m >= 1 && m <= tc::WT_MAX_M && n >= 1; }

extern "C" int dgmc_wgrad_tc_splits(int64_t k, int m, int n) {
    (void)m;
    return wgrad_tc_splits_impl(k, n);
}

extern "C" int dgmc_wgrad_tc_partial(const float* g, int64_t ldg, const float* x, int64_t ldx, int64_t k,
                                     int m, int n, float* partial, float* partial_b, dgmc_stream_t stream) {
    DGMC_REQUIRE(g && x && partial && k >= 0 && n >= 1 && ldg >= m && ldx >= n, "wgrad_tc_partial: bad arguments");
    if (!dgmc_wgrad_tc_supported(m, n)) {
        set_error("wgrad_tc_partial: m=%d unsupported (1..%d output channels)", m, tc::WT_MAX_M);
        return DGMC_ERR_UNSUPPORTED;
    }
    const int m_pad = m <= 64 ? 64 : 128;
    const size_t smem = tc::wt_smem(m_pad);
    static size_t configured = 0;
    if (smem > configured) {
        if (cudaFuncSetAttribute(tc::wgrad_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)smem) != cudaSuccess) {
            set_error("wgrad_tc_partial: cannot set %zu bytes of dynamic shared memory", smem);
            return DGMC_ERR_CUDA;
        }
        configured = smem;
    }
    const int splits = wgrad_tc_splits_impl(k, n);
    dim3 grid((n + tc::WT_TN - 1) / tc::WT_TN, splits);
    tc::wgrad_tc_kernel<<<grid, tc::WT_THREADS, smem, stream>>>(g, ldg, x, ldx, k, m, n, m_pad, splits,
                                                                      partial, partial_b);
    return check_launch("wgrad_tc_partial");
}
