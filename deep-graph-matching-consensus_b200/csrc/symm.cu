// Collectives of the row-sharded path over NVLink peer memory (one process per GPU, one node).
//
// The reference is single-process; the sharded path (dgmc/models/sharded.py) needs, per
// consensus step, the sum over ranks of the partial indicator transport r_t [N_t, R] (dgmc.py:209-212 is
// a sum over SOURCE rows, which are sharded) and, for node-sharded psi_2, an all-gather of a layer's
// features / a reduce-scatter of their gradients.  At the DBP15K shape these are 1-10 MB messages, 100
// of them per training step: latency, not bandwidth.  NCCL's ring costs ~38 us per 5 MB all-reduce
// here; the kernels below take ~10 us because every rank reads / writes its peers' buffers DIRECTLY
// (plain ld / st through the NVLink peer mapping of a symmetric allocation: same offset, every rank),
// synchronising with flag words in the same allocation:
//
//   all-reduce      two-shot: rank r sums slice r of all ranks' partials (fixed rank order, so every
//                   rank ends up with bit-identical sums) and stores the result into every rank's output
//   all-gather      each rank stores its block into every rank's output
//   reduce-scatter  each rank sums its block over all ranks' inputs
//
// Synchronisation is per thread block: block b of the kernel handshakes with block b of every peer
// (a release-store of the block's call count into the peer's flag word [b][my rank], then polling the
// local words [b][peer]: one NVLink crossing, no round trip).  A block starts to touch peer memory only
// after every peer's block b has started, i.e. after everything earlier in that peer's stream has
// finished, and — where the caller asks for the closing handshake — leaves only after every peer's
// block b has finished touching it.  Spins give up after ~2 s and raise a device flag
// (dgmc_sym_timeouts) instead of hanging the GPU.
//
// These entry points read and write ONLY what the caller passes: the table of the ranks' base pointers
// of one symmetric allocation (e.g. torch.distributed._symmetric_memory: handle.buffer_ptrs_dev) and
// byte offsets inside it.  Nothing here uses TMA.
#include "common.cuh"
#include "../../include/dgmc_b200.h"

namespace dgmc {

constexpr int SYM_THREADS = 512;
constexpr int SYM_MAX_BLOCKS = DGMC_SYM_MAX_BLOCKS;
constexpr int SYM_MAX_WORLD = DGMC_SYM_MAX_WORLD;
constexpr long long SYM_TIMEOUT_CYCLES = 4000000000LL;   // ~2 s at 1.9 GHz

__device__ int g_sym_timeouts = 0;

struct SymPeers {
    unsigned char* base[SYM_MAX_WORLD];
};

// flag words of one rank: [phase 0/1][block][sender rank], then this rank's own call counters
// [phase][block] (never touched by peers)
__device__ __forceinline__ unsigned* sym_flag(const SymPeers& p, int owner, long long pad_off, int phase,
                                              int block, int from) {
    return reinterpret_cast<unsigned*>(p.base[owner] + pad_off) +
           ((size_t)phase * SYM_MAX_BLOCKS + block) * SYM_MAX_WORLD + from;
}
__device__ __forceinline__ unsigned* sym_epoch(const SymPeers& p, int rank, long long pad_off, int phase, int block) {
    return reinterpret_cast<unsigned*>(p.base[rank] + pad_off) + (size_t)2 * SYM_MAX_BLOCKS * SYM_MAX_WORLD +
           (size_t)phase * SYM_MAX_BLOCKS + block;
}

// Block-wise handshake with the same block of every peer: ONE-WAY.  Each side keeps a call counter
// per (phase, block); arriving = a release-store of the new count into every peer's flag word (no
// round trip over NVLink), waiting = polling the local flag words until every peer's count has
// caught up.  Counters only grow, so the words never need resetting.
__device__ __forceinline__ void sym_block_barrier(const SymPeers& p, long long pad_off, int phase, int rank,
                                                  int world) {
    __syncthreads();          // the block's stores so far happen-before the release below (cumulativity)
    if ((int)threadIdx.x < world) {
        const int peer = threadIdx.x;
        unsigned* ep = sym_epoch(p, rank, pad_off, phase, blockIdx.x);
        const unsigned target = *reinterpret_cast<volatile unsigned*>(ep) + 1u;
        if (peer != rank) {
            unsigned* theirs = sym_flag(p, peer, pad_off, phase, blockIdx.x, rank);
            asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(theirs), "r"(target) : "memory");
            const unsigned* mine = sym_flag(p, rank, pad_off, phase, blockIdx.x, peer);
            const long long t0 = clock64();
            unsigned seen;
            do {    // relaxed polling, one acquire fence once the peer's count has arrived
                asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(seen) : "l"(mine) : "memory");
                if ((int)(seen - target) >= 0) break;
                if (clock64() - t0 > SYM_TIMEOUT_CYCLES) { atomicAdd(&g_sym_timeouts, 1); break; }
            } while (true);
            asm volatile("fence.acquire.sys;" ::: "memory");
        }
        __syncwarp((1u << world) - 1u);
        if (peer == rank) *reinterpret_cast<volatile unsigned*>(ep) = target;
    }
    __syncthreads();
}

__device__ __forceinline__ float4 add4(float4 a, const float4 b) {
    a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
    return a;
}

// out_r[i] = sum_p in_p[i] on every rank r; n4 float4 elements.  mc != nullptr: the allocation's
// multicast mapping (NVLink SHARP): one multimem.ld_reduce makes the SWITCH add the ranks' copies of an
// element and one multimem.st writes the sum into every rank's output — per rank n / world elements in
// and out instead of (world - 1) / world * n each way through SM loads.
__global__ void __launch_bounds__(SYM_THREADS) sym_all_reduce_kernel(const SymPeers p, unsigned char* mc, int rank,
                                                                     int world, long long in_off, long long out_off,
                                                                     long long pad_off, long long n4) {
    sym_block_barrier(p, pad_off, 0, rank, world);       // every rank's partial is complete
    const long long lo = n4 * rank / world, hi = n4 * (rank + 1) / world;
    const long long stride = (long long)gridDim.x * blockDim.x;
    if (mc) {
        const float4* src = reinterpret_cast<const float4*>(mc + in_off);
        float4* dst = reinterpret_cast<float4*>(mc + out_off);
        for (long long i = lo + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += stride) {
            float4 v;
            asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                         : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(src + i) : "memory");
            asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
                         ::"l"(dst + i), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
        }
    } else {
        // Peers are visited in the order rank + 1, rank + 2, ...: at any moment every rank reads from /
        // writes to a DIFFERENT peer, so no GPU's NVLink ingress is hit by all the others at once.  The
        // sum itself is taken in rank order 0, 1, ... (bit-identical results on every rank).
        for (long long i = lo + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += stride) {
            float4 v[SYM_MAX_WORLD];
#pragma unroll
            for (int j = 0; j < SYM_MAX_WORLD; ++j) {
                if (j < world) {
                    int q = rank + 1 + j;
                    if (q >= world) q -= world;
                    v[j] = __ldcg(reinterpret_cast<const float4*>(p.base[q] + in_off) + i);
                }
            }
            // v[j] belongs to rank (rank + 1 + j) % world, i.e. rank q sits at j = (q - rank - 1 + world) % world
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int q = 0; q < SYM_MAX_WORLD; ++q) {
                if (q < world) {
                    int j = q - rank - 1;
                    if (j < 0) j += world;
                    float4 t = v[0];
#pragma unroll
                    for (int u = 1; u < SYM_MAX_WORLD; ++u) if (u == j) t = v[u];
                    acc = q == 0 ? t : add4(acc, t);
                }
            }
#pragma unroll
            for (int j = 0; j < SYM_MAX_WORLD; ++j) {
                if (j < world) {
                    int q = rank + 1 + j;
                    if (q >= world) q -= world;
                    __stcg(reinterpret_cast<float4*>(p.base[q] + out_off) + i, acc);
                }
            }
        }
    }
    sym_block_barrier(p, pad_off, 1, rank, world);       // every rank's slice has landed everywhere
}

// A rank's part of a gathered tensor: up to two runs (float4 units, relative to the tensor's start);
// the rank's local buffer holds run 0 then run 1, contiguously.  Equal blocks in rank order are the
// special case {rank * block, block, 0, 0}; the node-sharded psi networks use two runs — the rank's
// source nodes and its target nodes in the natural [source ; target] node order.
struct SymRuns {
    long long at0, len0, at1, len1;
    __device__ __forceinline__ long long where(long long i) const { return i < len0 ? at0 + i : at1 + (i - len0); }
};

// out_r[runs(i)] = in[i] on every rank r
__global__ void __launch_bounds__(SYM_THREADS) sym_all_gather_kernel(const SymPeers p, int rank, int world,
                                                                     const float4* __restrict__ in,
                                                                     long long out_off, long long pad_off,
                                                                     const SymRuns runs, int sync_before) {
    // the opening handshake (peers have let go of the output region) is not needed when the caller
    // alternates between two output regions: see dgmc_b200.h
    if (sync_before) sym_block_barrier(p, pad_off, 0, rank, world);
    const long long total = runs.len0 + runs.len1;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const float4 v = in[i];
        const long long j = runs.where(i);
        for (int u = 0; u < world; ++u) {     // staggered: see sym_all_reduce_kernel
            int q = rank + 1 + u;
            if (q >= world) q -= world;
            __stcg(reinterpret_cast<float4*>(p.base[q] + out_off) + j, v);
        }
    }
    sym_block_barrier(p, pad_off, 1, rank, world);
}

// out[i] = sum_p in_p[runs(i)]
__global__ void __launch_bounds__(SYM_THREADS) sym_reduce_scatter_kernel(const SymPeers p, int rank, int world,
                                                                         long long in_off, float4* __restrict__ out,
                                                                         long long pad_off, const SymRuns runs,
                                                                         int sync_after) {
    sym_block_barrier(p, pad_off, 0, rank, world);       // every rank's input is complete
    const long long total = runs.len0 + runs.len1;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const long long j = runs.where(i);
        float4 v[SYM_MAX_WORLD];              // loads staggered over the peers, sum in rank order
#pragma unroll
        for (int u = 0; u < SYM_MAX_WORLD; ++u) {
            if (u < world) {
                int q = rank + 1 + u;
                if (q >= world) q -= world;
                v[u] = __ldcg(reinterpret_cast<const float4*>(p.base[q] + in_off) + j);
            }
        }
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int q = 0; q < SYM_MAX_WORLD; ++q) {
            if (q < world) {
                int u = q - rank - 1;
                if (u < 0) u += world;
                float4 t = v[0];
#pragma unroll
                for (int m = 1; m < SYM_MAX_WORLD; ++m) if (m == u) t = v[m];
                acc = q == 0 ? t : add4(acc, t);
            }
        }
        out[i] = acc;
    }
    // closing handshake (peers have finished reading this rank's input): not needed when the caller
    // alternates between two input regions
    if (sync_after) sym_block_barrier(p, pad_off, 1, rank, world);
}

static int sym_peers(SymPeers* p, const void* const* bases, int rank, int world, const char* what) {
    if (!bases || world < 1 || world > SYM_MAX_WORLD || rank < 0 || rank >= world) {
        set_error("%s: bad rank / world (%d / %d; at most %d ranks)", what, rank, world, SYM_MAX_WORLD);
        return DGMC_ERR_INVALID;
    }
    for (int q = 0; q < SYM_MAX_WORLD; ++q)
        p->base[q] = q < world ? static_cast<unsigned char*>(const_cast<void*>(bases[q])) : nullptr;
    for (int q = 0; q < world; ++q)
        if (!p->base[q]) { set_error("%s: null base pointer for rank %d", what, q); return DGMC_ERR_INVALID; }
    return DGMC_OK;
}

// whole CTAs only where there is work, never more than the flag array holds
static int sym_grid(long long items) {
    long long need = (items + SYM_THREADS - 1) / SYM_THREADS;
    if (need < 1) need = 1;
    const long long cap = sm_count() < SYM_MAX_BLOCKS ? sm_count() : SYM_MAX_BLOCKS;
    return (int)(need < cap ? need : cap);
}

}  // namespace dgmc

using namespace dgmc;

extern "C" size_t dgmc_sym_pad_bytes(void) {
    return ((size_t)2 * SYM_MAX_BLOCKS * SYM_MAX_WORLD + 2 * SYM_MAX_BLOCKS) * sizeof(unsigned);
}

extern "C" int dgmc_sym_all_reduce(const void* const* peer_bases, const void* multicast_base, int rank, int world,
                                   int64_t in_off, int64_t out_off, int64_t pad_off, int64_t n,
                                   dgmc_stream_t stream) {
    SymPeers p;
    if (int e = sym_peers(&p, peer_bases, rank, world, "sym_all_reduce")) return e;
    DGMC_REQUIRE(n >= 0 && n % 4 == 0 && in_off % 16 == 0 && out_off % 16 == 0 && pad_off % 4 == 0,
                 "sym_all_reduce: n must be a multiple of 4 floats, offsets 16-byte aligned");
    const long long n4 = n / 4;
    sym_all_reduce_kernel<<<sym_grid((n4 + world - 1) / world), SYM_THREADS, 0, stream>>>(
        p, static_cast<unsigned char*>(const_cast<void*>(multicast_base)), rank, world, in_off, out_off, pad_off, n4);
    return check_launch("sym_all_reduce");
}

static int sym_runs(SymRuns* r, int64_t at0, int64_t len0, int64_t at1, int64_t len1, int64_t max_len,
                    const char* what) {
    if (at0 < 0 || len0 < 0 || at1 < 0 || len1 < 0 || (at0 | len0 | at1 | len1 | max_len) % 4 != 0 ||
        max_len < len0 + len1) {
        set_error("%s: runs must be non-negative multiples of 4 floats, max_len >= len0 + len1", what);
        return DGMC_ERR_INVALID;
    }
    r->at0 = at0 / 4; r->len0 = len0 / 4; r->at1 = at1 / 4; r->len1 = len1 / 4;
    return DGMC_OK;
}

extern "C" int dgmc_sym_all_gather(const void* const* peer_bases, int rank, int world, const float* in,
                                   int64_t out_off, int64_t pad_off, int64_t at0, int64_t len0, int64_t at1,
                                   int64_t len1, int64_t max_len, int sync_before, dgmc_stream_t stream) {
    SymPeers p;
    SymRuns r;
    if (int e = sym_peers(&p, peer_bases, rank, world, "sym_all_gather")) return e;
    if (int e = sym_runs(&r, at0, len0, at1, len1, max_len, "sym_all_gather")) return e;
    DGMC_REQUIRE(in && out_off % 16 == 0 && pad_off % 4 == 0 && (reinterpret_cast<uintptr_t>(in) & 15) == 0,
                 "sym_all_gather: pointers / offsets must be 16-byte aligned");
    sym_all_gather_kernel<<<sym_grid(max_len / 4), SYM_THREADS, 0, stream>>>(
        p, rank, world, reinterpret_cast<const float4*>(in), out_off, pad_off, r, sync_before);
    return check_launch("sym_all_gather");
}

extern "C" int dgmc_sym_reduce_scatter(const void* const* peer_bases, int rank, int world, int64_t in_off,
                                       float* out, int64_t pad_off, int64_t at0, int64_t len0, int64_t at1,
                                       int64_t len1, int64_t max_len, int sync_after, dgmc_stream_t stream) {
    SymPeers p;
    SymRuns r;
    if (int e = sym_peers(&p, peer_bases, rank, world, "sym_reduce_scatter")) return e;
    if (int e = sym_runs(&r, at0, len0, at1, len1, max_len, "sym_reduce_scatter")) return e;
    DGMC_REQUIRE(out && in_off % 16 == 0 && pad_off % 4 == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0,
                 "sym_reduce_scatter: pointers / offsets must be 16-byte aligned");
    sym_reduce_scatter_kernel<<<sym_grid(max_len / 4), SYM_THREADS, 0, stream>>>(
        p, rank, world, in_off, reinterpret_cast<float4*>(out), pad_off, r, sync_after);
    return check_launch("sym_reduce_scatter");
}

extern "C" int dgmc_sym_timeouts(void) {
    int v = 0;
    if (cudaMemcpyFromSymbol(&v, g_sym_timeouts, sizeof(int)) != cudaSuccess) return -1;
    return v;
}
