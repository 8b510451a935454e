// Load-balanced ("entry-parallel") segmented reductions over a CSR.
//
// The by-target CSR of the candidate lists has hub segments: with freshly initialised embeddings a
// few target nodes sit in thousands of top-k lists, and a kernel that gives one lane group per
// SEGMENT then waits ~340 us on the longest one (ncu, profiles/README.md r1c).  Here the unit of
// work is a block of DGMC_CSR_BLOCK = 32 consecutive ENTRIES: a group of G lanes first prefetches
// the block's metadata (segment id, source row, weight of every entry: 32/G coalesced loads per
// lane, nothing dependent on a previous gather), then streams the 32 source rows eight at a time,
// switching accumulator whenever the segment id changes.  Control flow is the same for every group
// of a warp (4 batches), so groups do not serialise each other.  A segment that lies inside one
// block is written straight to the output; pieces of a segment that spans several blocks go to a
// small partial buffer (at most two pieces per block: one that starts at the block start, one that
// runs to the block end) and a second kernel (one group per BLOCK) adds them up in order —
// deterministic, no atomics.  Segments without entries are zeroed by the block after the gap.
//
//   MODE 0: out[s]  = seg_scale[s] * sum_e w_e * x[row_e]                (K4 transport, K5 aggregation)
//   MODE 1: d_pt[t] = -w2 (.) sum_e g[perm[e]] * [ps[perm[e]/k] > pt[t]]  (K6 backward, target side)
#include "common.cuh"
#include "../../include/dgmc_b200.h"

namespace dgmc {

constexpr int CB = DGMC_CSR_BLOCK;

struct BalArgs {
    const int* ptr; const int* seg_of; long long n_seg, nnz;
    // MODE 0
    const int* src; int src_div; const int* w_index; const float* w; const float* src_scale;
    const float* seg_scale; const float* x; long long ldx;
    // MODE 1
    const float* g; const float* ps; const float* pt; const int* perm; const float* w2; int k;
    // output
    float* out; long long ldo; int r4; int accumulate; float* partial;
    // ReLU backward fused into the store (dgmc_segment_gather_balanced_relu): rows of `mask` have the
    // pitch of `out`; a finished output element is zeroed where mask <= 0
    const float* mask;
    int hub_cap;   // entries of the fix-up pass's per-CTA hub list in use (<= FIX_HUB_CAP)
};

// 5 CTAs per SM for NV = 1 (48 registers, 8-48 bytes of spills): 740 resident CTAs hold the whole
// DBP15K-shape aggregation (723 CTAs of 16 entry blocks) in ONE wave; at 4 per SM (60 registers, no
// spills) the last 131 CTAs ran as a second, mostly empty wave (measured: step 7.67 -> 7.59 ms).
template <int G, int NV, int MODE>
__global__ void __launch_bounds__(256, NV == 1 ? 5 : 1) balanced_main_kernel(const BalArgs a) {
    constexpr int PER = CB / G;  // metadata entries held per lane
    constexpr int BR = NV == 2 ? 8 : 4;  // source rows in flight per lane group
    const int lane = threadIdx.x & 31, gl = lane % G;
    const unsigned gmask = G == 32 ? 0xffffffffu : (((1u << G) - 1u) << (lane - gl));
    const long long groups = (long long)gridDim.x * (blockDim.x / G);
    const long long nblk = (a.nnz + CB - 1) / CB;
    const int r = a.r4 * 4;
    float4 wv[NV];
    if (MODE == 1) {
#pragma unroll
        for (int v = 0; v < NV; ++v) {
            const int q = gl + v * G;
            wv[v] = q < a.r4 ? ldg4(a.w2 + 4 * q) : make_float4(0, 0, 0, 0);
        }
    }
    for (long long b = (long long)blockIdx.x * (blockDim.x / G) + threadIdx.x / G; b < nblk; b += groups) {
        const long long e_begin = b * CB;
        const int count = (int)(a.nnz - e_begin < CB ? a.nnz - e_begin : CB);
        // ---- prefetch the block's metadata ---------------------------------------------------
        int m_seg[PER], m_row[PER]; float m_w[PER];
#pragma unroll
        for (int p = 0; p < PER; ++p) {
            const int j = gl + p * G;
            const long long e = e_begin + j;
            int sg = -1, row = 0; float we = 0.f;
            if (j < count) {
                sg = __ldg(a.seg_of + e);
                if (MODE == 0) {
                    const int sv = __ldg(a.src + e);
                    row = a.src_div > 1 ? sv / a.src_div : sv;
                    we = a.w ? __ldg(a.w + (a.w_index ? __ldg(a.w_index + e) : (int)e)) : 1.f;
                    if (a.src_scale) we *= __ldg(a.src_scale + row);
                } else {
                    const int qe = __ldg(a.perm + e);
                    row = qe / a.k;
                    we = __ldg(a.g + qe);
                }
            }
            m_seg[p] = sg; m_row[p] = row; m_w[p] = we;
        }
        // does the first / last segment of this block continue into the neighbouring blocks?
        const int first_seg = __shfl_sync(gmask, m_seg[0], lane - gl);
        const int before = e_begin > 0 ? __ldg(a.seg_of + e_begin - 1) : -1;
        const int after = e_begin + count < a.nnz ? __ldg(a.seg_of + e_begin + count) : -1;

        float4 acc[NV], tv[NV];
#pragma unroll
        for (int v = 0; v < NV; ++v) { acc[v] = make_float4(0, 0, 0, 0); tv[v] = make_float4(0, 0, 0, 0); }
        int cur = first_seg;
        bool at_block_start = true;
        auto load_own = [&](int s) {
            if (MODE == 1) {
#pragma unroll
                for (int v = 0; v < NV; ++v) {
                    const int q = gl + v * G;
                    if (q < a.r4) tv[v] = ldg4(a.pt + (size_t)s * r + 4 * q);
                }
            }
        };
        auto flush = [&](int s, bool ends_at_block_end) {
            const bool open_left = at_block_start && s == before;
            const bool open_right = ends_at_block_end && s == after;
            const bool complete = !open_left && !open_right;
            float sc = 1.f;
            if (MODE == 0 && complete && a.seg_scale) sc = __ldg(a.seg_scale + s);
#pragma unroll
            for (int v = 0; v < NV; ++v) {
                const int q = gl + v * G;
                if (q < a.r4) {
                    float4 o = acc[v];
                    if (MODE == 1) { o.x *= -wv[v].x; o.y *= -wv[v].y; o.z *= -wv[v].z; o.w *= -wv[v].w; }
                    else { o.x *= sc; o.y *= sc; o.z *= sc; o.w *= sc; }
                    if (complete) {
                        float4* dst = reinterpret_cast<float4*>(a.out + (size_t)s * a.ldo + 4 * q);
                        if (a.accumulate) { const float4 p = *dst; o.x += p.x; o.y += p.y; o.z += p.z; o.w += p.w; }
                        *dst = o;
                    } else {
                        const long long slot = 2 * b + (at_block_start ? 0 : 1);
                        *reinterpret_cast<float4*>(a.partial + (size_t)slot * r + 4 * q) = o;
                    }
                }
            }
        };
        // empty segments produce no entries: the block holding the first entry AFTER a gap zeroes it
        auto zero_rows = [&](long long s0, long long s1) {   // rows s0 .. s1-1
            if (a.accumulate) return;
            for (long long s = s0; s < s1; ++s)
#pragma unroll
                for (int v = 0; v < NV; ++v) {
                    const int q = gl + v * G;
                    if (q < a.r4)
                        *reinterpret_cast<float4*>(a.out + (size_t)s * a.ldo + 4 * q) = make_float4(0, 0, 0, 0);
                }
        };
        if (before != first_seg) zero_rows((long long)before + 1, first_seg);
        load_own(cur);
        // ---- stream the rows, BR at a time (all BR gathers of a lane are in flight together) ------
#pragma unroll
        for (int batch = 0; batch < CB / BR; ++batch) {
            if (batch * BR < count) {
                int sg[BR], rows[BR]; float wt[BR];
#pragma unroll
                for (int u = 0; u < BR; ++u) {
                    const int j = batch * BR + u;          // compile-time after unrolling
                    const int srcl = (lane - gl) + (j % G);
                    sg[u] = __shfl_sync(gmask, m_seg[j / G], srcl);
                    rows[u] = __shfl_sync(gmask, m_row[j / G], srcl);
                    wt[u] = __shfl_sync(gmask, m_w[j / G], srcl);
                }
                float4 t[NV][BR];
#pragma unroll
                for (int v = 0; v < NV; ++v) {
                    const int q = gl + v * G;
                    const float* base = MODE == 0 ? a.x : a.ps;
                    const long long ld = MODE == 0 ? a.ldx : (long long)r;
#pragma unroll
                    for (int u = 0; u < BR; ++u)
                        t[v][u] = q < a.r4 ? ldg4(base + (size_t)rows[u] * ld + 4 * q) : make_float4(0, 0, 0, 0);
                }
#pragma unroll
                for (int u = 0; u < BR; ++u) {
                    if (sg[u] < 0) continue;              // beyond the last entry
                    if (sg[u] != cur) {
                        flush(cur, false);
                        zero_rows((long long)cur + 1, sg[u]);
                        cur = sg[u];
                        at_block_start = false;
#pragma unroll
                        for (int v = 0; v < NV; ++v) acc[v] = make_float4(0, 0, 0, 0);
                        load_own(cur);
                    }
#pragma unroll
                    for (int v = 0; v < NV; ++v) {
                        if (MODE == 0) {
                            acc[v].x += wt[u] * t[v][u].x; acc[v].y += wt[u] * t[v][u].y;
                            acc[v].z += wt[u] * t[v][u].z; acc[v].w += wt[u] * t[v][u].w;
                        } else {
                            acc[v].x += t[v][u].x > tv[v].x ? wt[u] : 0.f;
                            acc[v].y += t[v][u].y > tv[v].y ? wt[u] : 0.f;
                            acc[v].z += t[v][u].z > tv[v].z ? wt[u] : 0.f;
                            acc[v].w += t[v][u].w > tv[v].w ? wt[u] : 0.f;
                        }
                    }
                }
            }
        }
        flush(cur, true);
        if (e_begin + count >= a.nnz) zero_rows((long long)cur + 1, a.n_seg);  // trailing empty segments
    }
}

// ---------------------------------------------------------------------------------------------
// Lean MODE 0 kernel for rows of up to 128 channels (one float4 per lane).  Same blocks, same pieces,
// same partial slots and same summation order as balanced_main_kernel<G, 1, 0> — bit-identical
// results — but everything that depends only on the ENTRY (does a segment end here, where does its
// sum go: output row or partial slot, is there a gap of empty segments behind it) is worked out ONCE
// by the lane that holds the entry's metadata and travels with the shuffles, instead of being
// re-derived by all lanes at every entry.  ncu on the general kernel at the DBP15K shape: 71 warp
// instructions per entry step, 5.6 % of them FFMA, issue slots 69 % busy — the gather was
// instruction-bound, not memory-bound.
// ---------------------------------------------------------------------------------------------
// rows s0 .. s1-1 of `out` (channels 4q .. 4q+3 of this lane) = 0: the empty segments in a gap.  Rare
// and kept out of line so that it costs the hot loop neither registers nor code.
__device__ __noinline__ void lean_zero_rows(float* outq, long long ldo, long long s0, long long s1) {
    for (long long s = s0; s < s1; ++s)
        *reinterpret_cast<float4*>(outq + s * ldo) = make_float4(0, 0, 0, 0);
}

template <int G, bool ACC, bool SCALE, bool MASK = false>
__global__ void __launch_bounds__(256, 5) balanced_lean_kernel(const BalArgs a) {
    constexpr int PER = CB / G;
    constexpr int BR = 4;
    constexpr int GPW = 32 / G;                       // groups per warp
    const int lane = threadIdx.x & 31, gl = lane % G;
    const long long warps = (long long)gridDim.x * (blockDim.x >> 5);
    const long long nblk = (a.nnz + CB - 1) / CB;
    const int r = a.r4 * 4;
    const bool live = gl < a.r4;                      // lanes beyond the row width only take part in shuffles
    const float* __restrict__ xq = a.x + 4 * gl;
    float* __restrict__ outq = a.out + 4 * gl;
    float* __restrict__ partq = a.partial + 4 * gl;
    const float* __restrict__ maskq = MASK ? a.mask + 4 * gl : nullptr;
    const int src_div = a.src_div;
    const long long ldx = a.ldx, ldo = a.ldo;

    // All lanes of a warp run the same number of iterations (full-mask shuffles below): the loop runs
    // over WARP-sized sets of GPW consecutive blocks; a group whose block lies beyond the end works
    // on zero entries.
    for (long long wb = ((long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * GPW; wb < nblk;
         wb += warps * GPW) {
        const long long b = wb + lane / G;
        const long long e_begin = b * CB;
        const int count = b < nblk ? (int)(a.nnz - e_begin < CB ? a.nnz - e_begin : CB) : 0;
        int before = -1, after = -1, first_seg = -1;
        if (count > 0) {
            before = e_begin > 0 ? __ldg(a.seg_of + e_begin - 1) : -1;
            after = e_begin + count < a.nnz ? __ldg(a.seg_of + e_begin + count) : -1;
            first_seg = __ldg(a.seg_of + e_begin);
        }
        // ---- metadata of the entries this lane holds: source row, weight, and what happens AFTER the
        //      entry (flag 1: the segment's piece ends, its sum goes to element offset `off` of the
        //      output (flag 2 clear) or of the partial buffer (flag 2 set); flag 4: a gap of empty
        //      segments follows) ----
        //      packed into one word: bit 31 = piece ends, bit 0 = to the partial buffer, bit 1 = gap; the
        //      offsets are multiples of 4 (float4 rows) and below 2^31, so those bits are free
        int m_row[PER], m_act[PER]; float m_w[PER];
#pragma unroll
        for (int p = 0; p < PER; ++p) {
            const int j = gl + p * G;
            const long long e = e_begin + j;
            int row = 0, act = 0; float we = 0.f;
            if (j < count) {
                const int sg = __ldg(a.seg_of + e);
                const int nx = j + 1 < count ? __ldg(a.seg_of + e + 1) : -2;
                const int sv = __ldg(a.src + e);
                row = src_div > 1 ? sv / src_div : sv;
                we = a.w ? __ldg(a.w + (a.w_index ? __ldg(a.w_index + e) : (int)e)) : 1.f;
                if (a.src_scale) we *= __ldg(a.src_scale + row);
                if (nx != sg) {
                    const bool open_left = sg == first_seg && sg == before;
                    const bool open_right = nx == -2 && sg == after;
                    const bool complete = !open_left && !open_right;
                    const int off = complete ? (int)((long long)sg * ldo)
                                             : (int)((2 * b + (sg == first_seg ? 0 : 1)) * (long long)r);
                    act = (int)(0x80000000u | (unsigned)off | (complete ? 0u : 1u) | ((nx > sg + 1) ? 2u : 0u));
                }
            }
            m_row[p] = row; m_w[p] = we; m_act[p] = act;
        }
        if (!ACC && live && count > 0 && before != first_seg)
            lean_zero_rows(outq, ldo, (long long)before + 1, first_seg);

        float4 acc = make_float4(0, 0, 0, 0);
        // ---- stream the rows, BR at a time; control flow is the same for the whole warp ----
#pragma unroll
        for (int batch = 0; batch < CB / BR; ++batch) {
            int rows[BR], act[BR]; float wt[BR]; float4 t[BR];
#pragma unroll
            for (int u = 0; u < BR; ++u) {
                const int j = batch * BR + u;              // compile-time after unrolling
                rows[u] = __shfl_sync(0xffffffffu, m_row[j / G], j % G, G);
                wt[u] = __shfl_sync(0xffffffffu, m_w[j / G], j % G, G);
                act[u] = __shfl_sync(0xffffffffu, m_act[j / G], j % G, G);
            }
#pragma unroll
            for (int u = 0; u < BR; ++u)                   // entries beyond `count` read row 0 and are ignored
                t[u] = live ? ldg4(xq + (size_t)rows[u] * ldx) : make_float4(0, 0, 0, 0);
#pragma unroll
            for (int u = 0; u < BR; ++u) {
                const int j = batch * BR + u;
                if (j < count) {
                    acc.x += wt[u] * t[u].x; acc.y += wt[u] * t[u].y;
                    acc.z += wt[u] * t[u].z; acc.w += wt[u] * t[u].w;
                }
                if (act[u] < 0) {                          // bit 31: the segment's piece ends here
                    const bool to_partial = act[u] & 1;
                    float4 o = acc;
                    if (SCALE && !to_partial) {
                        const float sc = __ldg(a.seg_scale + __ldg(a.seg_of + e_begin + j));
                        o.x *= sc; o.y *= sc; o.z *= sc; o.w *= sc;
                    }
                    if (live) {
                        float4* dst = reinterpret_cast<float4*>((to_partial ? partq : outq) + (act[u] & 0x7ffffffc));
                        if (ACC && !to_partial) { const float4 q = *dst; o.x += q.x; o.y += q.y; o.z += q.z; o.w += q.w; }
                        if (MASK && !to_partial) {             // ReLU backward of the layer that produced `mask`
                            const float4 m = ldg4(maskq + (act[u] & 0x7ffffffc));
                            o.x = m.x > 0.f ? o.x : 0.f; o.y = m.y > 0.f ? o.y : 0.f;
                            o.z = m.z > 0.f ? o.z : 0.f; o.w = m.w > 0.f ? o.w : 0.f;
                        }
                        *dst = o;
                    }
                    acc = make_float4(0, 0, 0, 0);
                    if (!ACC && (act[u] & 2) && live)      // empty segments behind this one (rare)
                        lean_zero_rows(outq, ldo, (long long)__ldg(a.seg_of + e_begin + j) + 1,
                                       __ldg(a.seg_of + e_begin + j + 1));
                }
            }
        }
        if (!ACC && live && count > 0 && e_begin + count >= a.nnz)     // trailing empty segments
            lean_zero_rows(outq, ldo, (long long)__ldg(a.seg_of + a.nnz - 1) + 1, a.n_seg);
    }
}

// One lane group per entry block: if the block holds the FIRST piece of a segment that runs on into
// later blocks, add that segment's pieces up in block order and write the result.  A HUB (a target node
// that sits in thousands of candidate lists: 100+ pieces) is not walked by its one group — a chain of
// dependent L2 round trips, 34 us per launch at the DBP15K shape and 60 us on a hub-heavy edge list —
// but handed to the whole CTA: group g adds the pieces g, g + GPC, ... and the first group adds the
// GPC sums in group order.  Fixed order, so still deterministic; no atomics on data.
constexpr int FIX_HUB_MIN = 16;   // pieces beyond the first two from which the CTA shares the work
constexpr int FIX_HUB_CAP = 64;   // hubs a CTA can put aside (further ones are walked by their group)

template <int G, int NV>
__global__ void __launch_bounds__(256) balanced_fixup_kernel(const BalArgs a, int apply_seg_scale) {
    constexpr int GPC = 256 / G;                        // lane groups per CTA
    __shared__ int hub_n;
    __shared__ int hub_seg[FIX_HUB_CAP];
    __shared__ long long hub_first[FIX_HUB_CAP];        // 2 * first block of the hub + (its piece is `whole`)
    __shared__ float4 hub_part[GPC][G * NV];
    const int lane = threadIdx.x & 31, gl = threadIdx.x % G, g = threadIdx.x / G;
    const unsigned gmask = G == 32 ? 0xffffffffu : (((1u << G) - 1u) << (lane - gl));
    const long long nblk = (a.nnz + CB - 1) / CB;
    const int r = a.r4 * 4;
    auto store = [&](int s, const float4 (&acc)[NV]) {
        const float sc = (apply_seg_scale && a.seg_scale) ? __ldg(a.seg_scale + s) : 1.f;
#pragma unroll
        for (int v = 0; v < NV; ++v) {
            const int q = gl + v * G;
            if (q < a.r4) {
                float4* dst = reinterpret_cast<float4*>(a.out + (size_t)s * a.ldo + 4 * q);
                float4 o = make_float4(acc[v].x * sc, acc[v].y * sc, acc[v].z * sc, acc[v].w * sc);
                if (a.accumulate) { const float4 p = *dst; o.x += p.x; o.y += p.y; o.z += p.z; o.w += p.w; }
                if (a.mask) {
                    const float4 m = ldg4(a.mask + (size_t)s * a.ldo + 4 * q);
                    o.x = m.x > 0.f ? o.x : 0.f; o.y = m.y > 0.f ? o.y : 0.f;
                    o.z = m.z > 0.f ? o.z : 0.f; o.w = m.w > 0.f ? o.w : 0.f;
                }
                *dst = o;
            }
        }
    };
    auto piece = [&](long long slot, int v) {
        const int q = gl + v * G;
        return q < a.r4 ? *reinterpret_cast<const float4*>(a.partial + (size_t)slot * r + 4 * q)
                        : make_float4(0, 0, 0, 0);
    };
    if (threadIdx.x == 0) hub_n = 0;
    __syncthreads();
    // ---- pass 1: every group on its own (no barrier): segments cut into a few pieces are finished here,
    //      hubs are put on the CTA's list ----
    for (long long b = (long long)blockIdx.x * GPC + g; b < nblk; b += (long long)gridDim.x * GPC) {
        const long long e_begin = b * CB;
        const long long e_end = e_begin + CB < a.nnz ? e_begin + CB : a.nnz;
        if (e_end >= a.nnz) continue;                           // last block: nothing runs on
        // the boundary ids are independent loads: issue them together
        const int s = __ldg(a.seg_of + e_end - 1);
        const int s_next = __ldg(a.seg_of + e_end);
        const int s_first = __ldg(a.seg_of + e_begin);
        const int s_prev = e_begin > 0 ? __ldg(a.seg_of + e_begin - 1) : -1;
        const long long nb2 = e_end + CB;                        // first entry of the block after next
        const int s_next2 = nb2 < a.nnz ? __ldg(a.seg_of + nb2) : -1;
        if (s_next != s) continue;                              // last segment ends with the block
        const bool whole = s_first == s;                        // the piece starts at the block start
        if (whole && s_prev == s) continue;                     // not the first piece
        const long long b_last = s_next2 == s ? ((long long)__ldg(a.ptr + s + 1) - 1) / CB : b + 1;
        if (b_last - (b + 1) > FIX_HUB_MIN) {
            int h = 0;
            if (gl == 0) h = atomicAdd(&hub_n, 1);               // (list order does not matter: hubs are independent)
            h = __shfl_sync(gmask, h, lane - gl);
            if (h < a.hub_cap) {
                if (gl == 0) { hub_seg[h] = s; hub_first[h] = 2 * b + (whole ? 1 : 0); }
                continue;
            }
            // list full (which hubs end up here depends on timing): this group alone, but in the SAME
            // association as pass 2, so that the result does not depend on the path
            float4 acc[NV];
#pragma unroll
            for (int v = 0; v < NV; ++v) {
                const float4 t0 = piece(2 * b + (whole ? 0 : 1), v), t1 = piece(2 * (b + 1), v);
                acc[v] = make_float4(0, 0, 0, 0);
                acc[v].x += t0.x; acc[v].y += t0.y; acc[v].z += t0.z; acc[v].w += t0.w;
                acc[v].x += t1.x; acc[v].y += t1.y; acc[v].z += t1.z; acc[v].w += t1.w;
            }
            for (int gg = 0; gg < GPC; ++gg) {
                float4 sum[NV];
#pragma unroll
                for (int v = 0; v < NV; ++v) sum[v] = make_float4(0, 0, 0, 0);
                for (long long bb = b + 2 + gg; bb <= b_last; bb += 2 * GPC)
#pragma unroll
                    for (int v = 0; v < NV; ++v) {
                        const float4 t = piece(2 * bb, v);
                        const float4 u = bb + GPC <= b_last ? piece(2 * (bb + GPC), v) : make_float4(0, 0, 0, 0);
                        sum[v].x += t.x; sum[v].y += t.y; sum[v].z += t.z; sum[v].w += t.w;
                        sum[v].x += u.x; sum[v].y += u.y; sum[v].z += u.z; sum[v].w += u.w;
                    }
#pragma unroll
                for (int v = 0; v < NV; ++v) {
                    acc[v].x += sum[v].x; acc[v].y += sum[v].y; acc[v].z += sum[v].z; acc[v].w += sum[v].w;
                }
            }
            store(s, acc);
            continue;
        }
        float4 acc[NV];
#pragma unroll
        for (int v = 0; v < NV; ++v) {
            // first two pieces (the common case: a segment cut once) loaded together
            const float4 t0 = piece(2 * b + (whole ? 0 : 1), v), t1 = piece(2 * (b + 1), v);
            acc[v] = make_float4(0, 0, 0, 0);
            acc[v].x += t0.x; acc[v].y += t0.y; acc[v].z += t0.z; acc[v].w += t0.w;
            acc[v].x += t1.x; acc[v].y += t1.y; acc[v].z += t1.z; acc[v].w += t1.w;
        }
        for (long long bb = b + 2; bb <= b_last; ++bb)
#pragma unroll
            for (int v = 0; v < NV; ++v) {
                const float4 t = piece(2 * bb, v);
                acc[v].x += t.x; acc[v].y += t.y; acc[v].z += t.z; acc[v].w += t.w;
            }
        store(s, acc);
    }
    __syncthreads();
    // ---- pass 2: the CTA's hubs, one after the other, all groups on each ----
    const int nh = hub_n < a.hub_cap ? hub_n : a.hub_cap;
    for (int h = 0; h < nh; ++h) {
        const int s = hub_seg[h];
        const long long bf = hub_first[h] >> 1;
        const bool whole = hub_first[h] & 1;
        const long long b_last = ((long long)__ldg(a.ptr + s + 1) - 1) / CB;
        float4 sum[NV];
#pragma unroll
        for (int v = 0; v < NV; ++v) sum[v] = make_float4(0, 0, 0, 0);
        for (long long bb = bf + 2 + g; bb <= b_last; bb += 2 * GPC) {     // two independent loads per pass
#pragma unroll
            for (int v = 0; v < NV; ++v) {
                const float4 t = piece(2 * bb, v);
                const float4 u = bb + GPC <= b_last ? piece(2 * (bb + GPC), v) : make_float4(0, 0, 0, 0);
                sum[v].x += t.x; sum[v].y += t.y; sum[v].z += t.z; sum[v].w += t.w;
                sum[v].x += u.x; sum[v].y += u.y; sum[v].z += u.z; sum[v].w += u.w;
            }
        }
#pragma unroll
        for (int v = 0; v < NV; ++v) hub_part[g][gl + v * G] = sum[v];
        __syncthreads();
        if (g == 0) {
            float4 acc[NV];
#pragma unroll
            for (int v = 0; v < NV; ++v) {
                const float4 t0 = piece(2 * bf + (whole ? 0 : 1), v), t1 = piece(2 * (bf + 1), v);
                acc[v] = make_float4(0, 0, 0, 0);
                acc[v].x += t0.x; acc[v].y += t0.y; acc[v].z += t0.z; acc[v].w += t0.w;
                acc[v].x += t1.x; acc[v].y += t1.y; acc[v].z += t1.z; acc[v].w += t1.w;
                for (int gg = 0; gg < GPC; ++gg) {
                    const float4 t = hub_part[gg][gl + v * G];
                    acc[v].x += t.x; acc[v].y += t.y; acc[v].z += t.z; acc[v].w += t.w;
                }
            }
            store(s, acc);
        }
        __syncthreads();
    }
}

// nnz == 0: every segment is empty
__global__ void __launch_bounds__(256) balanced_zero_kernel(const BalArgs a) {
    const long long total = a.n_seg * a.r4;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += stride)
        *reinterpret_cast<float4*>(a.out + (size_t)(t / a.r4) * a.ldo + 4 * (t % a.r4)) = make_float4(0, 0, 0, 0);
}

static int g_hub_cap = FIX_HUB_CAP;

template <int MODE>
static int launch_balanced(const BalArgs& a_in, cudaStream_t stream) {
    BalArgs a = a_in;
    a.hub_cap = g_hub_cap;
    const long long nblk = (a.nnz + CB - 1) / CB;
    if (nblk == 0) {
        if (!a.accumulate) balanced_zero_kernel<<<wave_grid(a.n_seg * a.r4, 256, 8), 256, 0, stream>>>(a);
        return check_launch("balanced segmented reduction (empty)", a.accumulate ? 0 : 1);
    }
    if (MODE == 0 && a.r4 <= 32 && a.n_seg * a.ldo < (1ll << 31) && 2 * nblk * (a.r4 * 4ll) < (1ll << 31)) {
        // rows of up to 128 channels, element offsets that fit 32 bits: the lean kernel (same pieces
        // and slots, same fix-up pass)
        const int G = a.r4 <= 8 ? 8 : (a.r4 <= 16 ? 16 : 32);
        const int grid = wave_grid(nblk, 256 / G, 8);
#define LEAN(G_, ACC_, SCALE_) balanced_lean_kernel<G_, ACC_, SCALE_><<<grid, 256, 0, stream>>>(a)
#define LEAN_G(G_)                                                                                  \
    do {                                                                                            \
        if (a.mask) balanced_lean_kernel<G_, true, false, true><<<grid, 256, 0, stream>>>(a);      \
        else if (a.accumulate) { if (a.seg_scale) LEAN(G_, true, true); else LEAN(G_, true, false); } \
        else { if (a.seg_scale) LEAN(G_, false, true); else LEAN(G_, false, false); }               \
    } while (0)
        if (G == 8) LEAN_G(8); else if (G == 16) LEAN_G(16); else LEAN_G(32);
#undef LEAN_G
#undef LEAN
        if (nblk > 1) {
            if (G == 8) balanced_fixup_kernel<8, 1><<<wave_grid(nblk, 32, 8), 256, 0, stream>>>(a, 1);
            else if (G == 16) balanced_fixup_kernel<16, 1><<<wave_grid(nblk, 16, 8), 256, 0, stream>>>(a, 1);
            else balanced_fixup_kernel<32, 1><<<wave_grid(nblk, 8, 8), 256, 0, stream>>>(a, 1);
        }
        return check_launch("balanced segmented reduction", nblk > 1 ? 2 : 1);
    }
    if (a.mask) {
        set_error("segment_gather_balanced_relu: rows of up to 128 channels and 32-bit offsets only");
        return DGMC_ERR_UNSUPPORTED;
    }
#define BAL(G, NV)                                                                                  \
    do {                                                                                            \
        balanced_main_kernel<G, NV, MODE><<<wave_grid(nblk, 256 / G, 8), 256, 0, stream>>>(a);     \
        if (nblk > 1)                                                                               \
            balanced_fixup_kernel<G, NV><<<wave_grid(nblk, 256 / G, 8), 256, 0, stream>>>(a, MODE == 0); \
    } while (0)
    if (a.r4 <= 8) BAL(8, 1);
    else if (a.r4 <= 16) BAL(16, 1);
    else if (a.r4 <= 32) BAL(32, 1);
    else if (a.r4 <= 64) BAL(32, 2);
    else BAL(32, 4);
#undef BAL
    return check_launch("balanced segmented reduction", nblk > 1 ? 2 : 1);
}

}  // namespace dgmc

using namespace dgmc;

extern "C" size_t dgmc_balanced_workspace(int64_t nnz, int r) {
    const size_t nblk = (size_t)((nnz + CB - 1) / CB);
    return 2 * nblk * (size_t)((r + 3) / 4 * 4) * sizeof(float) + 256;
}

static int bal_common(const char* who, int r, const void* out, int64_t ldo, void* ws, size_t ws_bytes,
                      int64_t nnz) {
    if (r < 4 || r % 4 != 0 || r > 512 || ldo % 4 != 0 || (reinterpret_cast<uintptr_t>(out) & 15)) {
        set_error("%s: needs r %% 4 == 0, r <= 512, 16-byte aligned rows (r=%d ldo=%lld)", who, r, (long long)ldo);
        return DGMC_ERR_UNSUPPORTED;
    }
    if (!ws || ws_bytes < dgmc_balanced_workspace(nnz, r) || (reinterpret_cast<uintptr_t>(ws) & 15)) {
        set_error("%s: workspace too small or misaligned", who);
        return DGMC_ERR_WORKSPACE;
    }
    return DGMC_OK;
}

extern "C" int dgmc_segment_gather_balanced(const int32_t* ptr, const int32_t* seg_of, const int32_t* src,
                                            int src_div, const int32_t* w_index, const float* w,
                                            const float* src_scale, const float* seg_scale, const float* x,
                                            int64_t ldx, float* out, int64_t ldo, int64_t n_seg, int64_t nnz,
                                            int r, int accumulate, void* workspace, size_t workspace_bytes,
                                            dgmc_stream_t stream) {
    DGMC_REQUIRE(ptr && x && out && n_seg >= 0 && nnz >= 0 && src_div >= 1 && ldx >= r && ldo >= r,
                 "segment_gather_balanced: bad arguments");
    DGMC_REQUIRE(nnz == 0 || (seg_of && src), "segment_gather_balanced: null CSR arrays");
    if (n_seg == 0) return DGMC_OK;
    if (int e = bal_common("segment_gather_balanced", r, out, ldo, workspace, workspace_bytes, nnz)) return e;
    if (ldx % 4 != 0 || (reinterpret_cast<uintptr_t>(x) & 15)) {
        set_error("segment_gather_balanced: x rows must be 16-byte aligned");
        return DGMC_ERR_UNSUPPORTED;
    }
    BalArgs a{};
    a.ptr = ptr; a.seg_of = seg_of; a.n_seg = n_seg; a.nnz = nnz;
    a.src = src; a.src_div = src_div; a.w_index = w_index; a.w = w; a.src_scale = src_scale;
    a.seg_scale = seg_scale; a.x = x; a.ldx = ldx;
    a.out = out; a.ldo = ldo; a.r4 = r / 4; a.accumulate = accumulate; a.partial = (float*)workspace;
    return launch_balanced<0>(a, stream);
}

extern "C" int dgmc_diff_mlp_bwd_targets_balanced(const float* g, const float* ps, const float* pt,
                                                  const int32_t* tptr, const int32_t* seg_of,
                                                  const int32_t* perm, const float* w2, float* d_pt,
                                                  int64_t n_t, int64_t nnz, int k, int r, void* workspace,
                                                  size_t workspace_bytes, dgmc_stream_t stream) {
    DGMC_REQUIRE(g && ps && pt && tptr && w2 && d_pt && n_t >= 0 && nnz >= 0 && k >= 1,
                 "diff_mlp_bwd_targets_balanced: bad arguments");
    DGMC_REQUIRE(nnz == 0 || (seg_of && perm), "diff_mlp_bwd_targets_balanced: null CSR arrays");
    if (n_t == 0) return DGMC_OK;
    if (int e = bal_common("diff_mlp_bwd_targets_balanced", r, d_pt, r, workspace, workspace_bytes, nnz)) return e;
    BalArgs a{};
    a.ptr = tptr; a.seg_of = seg_of; a.n_seg = n_t; a.nnz = nnz;
    a.g = g; a.ps = ps; a.pt = pt; a.perm = perm; a.w2 = w2; a.k = k;
    a.out = d_pt; a.ldo = r; a.r4 = r / 4; a.accumulate = 0; a.partial = (float*)workspace;
    return launch_balanced<1>(a, stream);
}

// The transposed RelConv aggregation of the backward pass with the ReLU backward of the PREVIOUS layer
// fused into its store: out = (out + A x) * [relu_mask > 0], where out is the gradient of that layer's
// output (already holding the contribution of the concatenation) and relu_mask is the output itself,
// stored with the same row pitch as out.  replaces: threshold_backward after the scatter of PyG's
// propagate backward (rel.py:27-30,84-85).  r % 4 == 0, r <= 128.
extern "C" int dgmc_segment_gather_balanced_relu(const int32_t* ptr, const int32_t* seg_of, const int32_t* src,
                                                 const float* w, const float* x, int64_t ldx, float* out,
                                                 int64_t ldo, const float* relu_mask, int64_t n_seg, int64_t nnz,
                                                 int r, void* workspace, size_t workspace_bytes,
                                                 dgmc_stream_t stream) {
    DGMC_REQUIRE(ptr && x && out && relu_mask && n_seg >= 0 && nnz >= 1 && ldx >= r && ldo >= r && seg_of && src,
                 "segment_gather_balanced_relu: bad arguments");
    if (n_seg == 0) return DGMC_OK;
    if (int e = bal_common("segment_gather_balanced_relu", r, out, ldo, workspace, workspace_bytes, nnz)) return e;
    if (r > 128 || ldx % 4 != 0 || ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(relu_mask)) & 15)) {
        set_error("segment_gather_balanced_relu: r <= 128, x and relu_mask rows 16-byte aligned");
        return DGMC_ERR_UNSUPPORTED;
    }
    BalArgs a{};
    a.ptr = ptr; a.seg_of = seg_of; a.n_seg = n_seg; a.nnz = nnz;
    a.src = src; a.src_div = 1; a.w = w; a.x = x; a.ldx = ldx;
    a.out = out; a.ldo = ldo; a.r4 = r / 4; a.accumulate = 1; a.partial = (float*)workspace; a.mask = relu_mask;
    return launch_balanced<0>(a, stream);
}

// Test hook: how many hubs a CTA of the fix-up pass may put on its list (0 .. 64, default 64); hubs beyond
// that are reduced by their own lane group in the same association, so results must not depend on it.
extern "C" int dgmc_set_fixup_hub_list(int entries) {
    DGMC_REQUIRE(entries >= 0 && entries <= FIX_HUB_CAP, "set_fixup_hub_list: 0 .. %d entries", FIX_HUB_CAP);
    g_hub_cap = entries;
    return DGMC_OK;
}
