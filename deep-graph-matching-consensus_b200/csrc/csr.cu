// dgmc_csr_build: stable LSD radix sort (8-bit digits) of int32 keys carrying their position,
// then segment boundaries.  Stability makes every segment list ascending in entry id, so the
// segmented reductions that consume it (K4 transport, K5 aggregation, K6 backward) add their
// terms in a fixed order: results are bit-reproducible run to run, unlike atomics.
//
// Per pass: (1) per-tile digit histogram, (2) exclusive scan of the [256][tiles] table in
// digit-major order, (3) stable scatter.  Tiles are 256 threads x 4 keys, each warp owning a
// contiguous 128-key run so that (warp, round, lane) order == key order.
#include "common.cuh"
#include "../../include/dgmc_b200.h"

namespace dgmc {

constexpr int RS_THREADS = 256;
constexpr int RS_ITEMS = 4;
constexpr int RS_TILE = RS_THREADS * RS_ITEMS;  // 1024 keys per block
constexpr int RS_WARPS = RS_THREADS / 32;
constexpr int SCAN_TILE = 1024 * 4;

__device__ __forceinline__ int digit_of(int key, int shift) { return (key >> shift) & 255; }

// counts[w][d] for this tile; also returns each key's stable rank inside (warp, digit)
__device__ __forceinline__ void tile_rank(const int* __restrict__ keys, long long n, long long tile0,
                                          int shift, int (*cnt)[256], int* my_digit, int* my_rank) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < RS_WARPS * 256; i += RS_THREADS) (&cnt[0][0])[i] = 0;
    __syncthreads();
    const long long wbase = tile0 + (long long)warp * (32 * RS_ITEMS);
#pragma unroll
    for (int r = 0; r < RS_ITEMS; ++r) {
        const long long pos = wbase + r * 32 + lane;
        const bool live = pos < n;
        const int d = live ? digit_of(__ldg(keys + pos), shift) : 256 + lane;  // dead lanes: unique
        const unsigned peers = __match_any_sync(0xffffffffu, d);
        const int before = __popc(peers & ((1u << lane) - 1u));
        int prior = 0;
        if (live) prior = cnt[warp][d];
        __syncwarp();
        if (live && before == 0) cnt[warp][d] = prior + __popc(peers);
        __syncwarp();
        my_digit[r] = live ? d : -1;
        my_rank[r] = prior + before;
    }
    __syncthreads();
}

__global__ void __launch_bounds__(RS_THREADS) radix_hist_kernel(const int* __restrict__ keys,
                                                                long long n, int shift,
                                                                int* __restrict__ table,
                                                                int tiles) {
    __shared__ int cnt[RS_WARPS][256];
    int dg[RS_ITEMS], rk[RS_ITEMS];
    tile_rank(keys, n, (long long)blockIdx.x * RS_TILE, shift, cnt, dg, rk);
    const int d = threadIdx.x;  // 256 threads <-> 256 digits
    int total = 0;
#pragma unroll
    for (int w = 0; w < RS_WARPS; ++w) total += cnt[w][d];
    table[(size_t)d * tiles + blockIdx.x] = total;
}

__global__ void __launch_bounds__(RS_THREADS) radix_scatter_kernel(
    const int* __restrict__ keys, const int* __restrict__ vals, long long n, int shift,
    const int* __restrict__ table_scanned, int tiles, int* __restrict__ keys_out,
    int* __restrict__ vals_out) {
    __shared__ int cnt[RS_WARPS][256];
    __shared__ int base[RS_WARPS][256];
    int dg[RS_ITEMS], rk[RS_ITEMS];
    const long long tile0 = (long long)blockIdx.x * RS_TILE;
    tile_rank(keys, n, tile0, shift, cnt, dg, rk);
    {
        const int d = threadIdx.x;
        int run = table_scanned[(size_t)d * tiles + blockIdx.x];
#pragma unroll
        for (int w = 0; w < RS_WARPS; ++w) {
            base[w][d] = run;
            run += cnt[w][d];
        }
    }
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long long wbase = tile0 + (long long)warp * (32 * RS_ITEMS);
#pragma unroll
    for (int r = 0; r < RS_ITEMS; ++r) {
        if (dg[r] >= 0) {
            const long long pos = wbase + r * 32 + lane;
            const int dst = base[warp][dg[r]] + rk[r];
            keys_out[dst] = __ldg(keys + pos);
            vals_out[dst] = vals ? __ldg(vals + pos) : (int)pos;  // first pass: identity payload
        }
    }
}

// ---- exclusive scan of int32 (three-phase) ---------------------------------------------------
__device__ __forceinline__ int block_exclusive_scan(int v, int* total) {
    __shared__ int wsum[32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) wsum[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        int s = lane < (int)(blockDim.x >> 5) ? wsum[lane] : 0;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, s, o);
            if (lane >= o) s += t;
        }
        wsum[lane] = s;
    }
    __syncthreads();
    const int woff = warp ? wsum[warp - 1] : 0;
    *total = wsum[(blockDim.x >> 5) - 1];
    __syncthreads();
    return woff + inc - v;
}

__global__ void __launch_bounds__(1024) scan_tiles_kernel(const int* __restrict__ in,
                                                          int* __restrict__ out, long long n,
                                                          int* __restrict__ tile_sums) {
    const long long base = (long long)blockIdx.x * SCAN_TILE + threadIdx.x * 4;
    int v[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) v[i] = base + i < n ? in[base + i] : 0;
    const int mine = v[0] + v[1] + v[2] + v[3];
    int total;
    int run = block_exclusive_scan(mine, &total);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        if (base + i < n) out[base + i] = run;
        run += v[i];
    }
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}

__global__ void __launch_bounds__(1024) scan_sums_kernel(int* __restrict__ sums, int n) {
    // single block; n <= SCAN_TILE * 1024 tiles is far beyond what the callers need
    int carry = 0;
    for (int start = 0; start < n; start += 1024) {
        const int i = start + threadIdx.x;
        const int v = i < n ? sums[i] : 0;
        int total;
        const int ex = block_exclusive_scan(v, &total);
        if (i < n) sums[i] = carry + ex;
        carry += total;
    }
}

__global__ void __launch_bounds__(1024) scan_add_kernel(int* __restrict__ out, long long n,
                                                        const int* __restrict__ tile_sums) {
    const long long base = (long long)blockIdx.x * SCAN_TILE + threadIdx.x * 4;
    const int add = tile_sums[blockIdx.x];
#pragma unroll
    for (int i = 0; i < 4; ++i)
        if (base + i < n) out[base + i] += add;
}

// ptr[s] = first position whose sorted key is >= s
__global__ void __launch_bounds__(256) boundaries_kernel(const int* __restrict__ sorted,
                                                         long long n, long long n_seg,
                                                         int* __restrict__ ptr) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i <= n; i += stride) {
        long long lo = i == 0 ? -1 : sorted[i - 1];
        long long hi = i == n ? n_seg : sorted[i];
        lo = lo < -1 ? -1 : (lo > n_seg ? n_seg : lo);  // out-of-range keys never write outside ptr
        hi = hi < -1 ? -1 : (hi > n_seg ? n_seg : hi);
        for (long long s = lo + 1; s <= hi; ++s) ptr[s] = (int)i;
    }
}

static inline size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

struct CsrPlan {
    long long tiles, table, scan_tiles;
    size_t off_keys0, off_keys1, off_vals, off_table, off_sums, total;
};

static CsrPlan plan_csr(long long nnz) {
    CsrPlan p;
    p.tiles = (nnz + RS_TILE - 1) / RS_TILE;
    if (p.tiles < 1) p.tiles = 1;
    p.table = 256 * p.tiles;
    p.scan_tiles = (p.table + SCAN_TILE - 1) / SCAN_TILE;
    size_t o = 0;
    p.off_keys0 = o; o += align256((size_t)nnz * 4);
    p.off_keys1 = o; o += align256((size_t)nnz * 4);
    p.off_vals = o;  o += align256((size_t)nnz * 4);
    p.off_table = o; o += align256((size_t)p.table * 4);
    p.off_sums = o;  o += align256((size_t)p.scan_tiles * 4);
    p.total = o;
    return p;
}

}  // namespace dgmc

using namespace dgmc;

extern "C" size_t dgmc_csr_build_workspace(int64_t nnz, int64_t n_seg) {
    (void)n_seg;
    return plan_csr(nnz < 0 ? 0 : nnz).total;
}

extern "C" int dgmc_csr_build(const int32_t* keys, int64_t nnz, int64_t n_seg, int32_t* ptr,
                              int32_t* perm, int32_t* sorted_keys, void* workspace,
                              size_t workspace_bytes, dgmc_stream_t stream) {
    DGMC_REQUIRE(ptr && nnz >= 0 && n_seg >= 0 && nnz < (1ll << 31) && n_seg < (1ll << 31) - 1,
                 "csr_build: bad sizes (nnz=%lld n_seg=%lld)", (long long)nnz, (long long)n_seg);
    DGMC_REQUIRE(nnz == 0 || (keys && perm && workspace), "csr_build: null pointer");
    const CsrPlan p = plan_csr(nnz);
    if (workspace_bytes < p.total) {
        set_error("csr_build: workspace %zu < %zu bytes", workspace_bytes, p.total);
        return DGMC_ERR_WORKSPACE;
    }
    if (nnz == 0) {
        if (cudaMemsetAsync(ptr, 0, (size_t)(n_seg + 1) * 4, stream) != cudaSuccess) {
            set_error("csr_build: memset failed");
            return DGMC_ERR_CUDA;
        }
        return DGMC_OK;
    }
    char* ws = (char*)workspace;
    int* k0 = (int*)(ws + p.off_keys0);
    int* k1 = (int*)(ws + p.off_keys1);
    int* v1 = (int*)(ws + p.off_vals);
    int* table = (int*)(ws + p.off_table);
    int* sums = (int*)(ws + p.off_sums);

    int bits = 1;
    while (bits < 31 && (1ll << bits) < n_seg) ++bits;
    const int passes = (bits + 7) / 8;
    // ping-pong so that the LAST pass writes payload into `perm`
    const int* src_k = keys;
    const int* src_v = nullptr;
    for (int pass = 0; pass < passes; ++pass) {
        const bool last = pass == passes - 1;
        int* dst_k = (last && sorted_keys) ? sorted_keys : ((pass & 1) ? k1 : k0);
        // payload alternates perm <-> v1, arranged so that the final write lands in perm
        int* dst_v = ((passes - 1 - pass) & 1) ? v1 : perm;
        radix_hist_kernel<<<(int)p.tiles, RS_THREADS, 0, stream>>>(src_k, nnz, 8 * pass, table,
                                                                   (int)p.tiles);
        scan_tiles_kernel<<<(int)p.scan_tiles, 1024, 0, stream>>>(table, table, p.table, sums);
        scan_sums_kernel<<<1, 1024, 0, stream>>>(sums, (int)p.scan_tiles);
        scan_add_kernel<<<(int)p.scan_tiles, 1024, 0, stream>>>(table, p.table, sums);
        radix_scatter_kernel<<<(int)p.tiles, RS_THREADS, 0, stream>>>(
            src_k, src_v, nnz, 8 * pass, table, (int)p.tiles, dst_k, dst_v);
        src_k = dst_k;
        src_v = dst_v;
    }
    boundaries_kernel<<<wave_grid(nnz + 1, 256, 4), 256, 0, stream>>>(src_k, nnz, n_seg, ptr);
    return check_launch("csr_build", 5 * passes + 1);
}
