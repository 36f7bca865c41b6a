// k3_replay.cu — the order of exact-magnitude ties among the seeds, as libstdc++'s std::sort leaves it.
//
// The reference sorts its peak list with std::sort and a comparator on the magnitude only (reference src/filter.cpp:188).
// std::sort is introsort and not stable: two peaks of EXACTLY equal magnitude come out in an order that depends on which
// partitions moved them.  The seed order is the flood priority and — through the component list — the order of the
// output segments, so the order of ties is part of the result.  This kernel reproduces it.
//
// What has to be replayed (bits/stl_algo.h of the reference's toolchain, GCC 13 libstdc++):
//   std::sort            = __introsort_loop(first, last, 2 * floor(log2 n)) ; __final_insertion_sort(first, last)
//   __introsort_loop     : while (last - first > 16) { depth exhausted -> heap sort of the range, done;
//                            cut = __unguarded_partition_pivot(first, last); recurse on [cut, last); last = cut; }
//   partition_pivot      : median of (first+1, mid, last-1) swapped to *first, then the unguarded Hoare partition of
//                          [first+1, last) around *first
//   final insertion sort : stable (an element moves left only past strictly smaller-ranked ones)
// => final order = STABLE sort, by magnitude, of the arrangement the partition phase leaves.  Elements whose magnitude is
// unique have their final rank from K3 already; only the relative order inside a tie group needs the replay, and a
// range that holds fewer than two tied elements cannot change any of it, so it is not followed further.
//
// A Hoare partition step in data-parallel form.  Scanning from the left stops at elements that are not "before" the
// pivot, scanning from the right at elements that are not "after" it; the k-th left stopper is exchanged with the k-th
// right stopper as long as it lies left of it.  With L[k] / R[k] the stoppers of the ORIGINAL arrangement (the scans
// only ever look at untouched positions or at elements they have just exchanged, which stop them):
//   K   = number of leading k with L[k] < R[k]                      (exchanges)
//   cut = K == 0 ? L[0] : min(L[K] (if any), R[K-1])                (what __unguarded_partition returns)
// (checked against std::sort on millions of elements with heavy ties, including the heap-sort branch: tests/test_sort_replay_gpu.py
// drives this kernel through lrb_sort_replay against the oracle's std::sort.)
//
// Work distribution: one CTA (32 warps) per frame, two phases.
//   Phase 1, ranges longer than kSmall: level by level, all ranges of a level at once (they are disjoint).  Every warp owns a
//     contiguous slice of every range: count its stoppers, scan the 32 counts, write the stopper positions, then the whole
//     CTA performs the exchanges.  A level costs a fixed number of barriers and memory round trips, whatever its ranges.
//   Phase 2, ranges of at most kSmall elements: one warp per range, the range staged in shared memory; the warp follows the
//     recursion below it on its own stack and writes the range back.
// Keys are the dense ranks K3 computed (ties share a key; a SMALLER key is a LARGER magnitude), the payload is the K3 rank
// (magnitude desc, raster asc) with the tied flag in bit 31.
#include <climits>
#include "common.cuh"
#include "kernels.h"

namespace lrb {

namespace {

constexpr int RP_THREADS = 1024;
constexpr int kSortThreshold = 16;            // libstdc++ _S_threshold
constexpr unsigned kTiedBit = 0x80000000u;

// the comparator of reference src/filter.cpp:188, `i.v > j.v`, on dense-rank keys
__device__ __forceinline__ bool rp_before(unsigned ka, unsigned kb) { return ka < kb; }

// ---- the heap-sort branch (depth limit exhausted): std::__partial_sort(first, last, last) = make_heap + sort_heap, by one lane ----
__device__ void rp_push_heap(uint2 * first, int hole, int top, uint2 value)
{
    int parent = (hole - 1) / 2;
    while (hole > top && rp_before(first[parent].x, value.x)) { first[hole] = first[parent]; hole = parent; parent = (hole - 1) / 2; }
    first[hole] = value;
}
__device__ void rp_adjust_heap(uint2 * first, int hole, int len, uint2 value)
{
    const int top = hole; int second = hole;
    while (second < (len - 1) / 2) {
        second = 2 * (second + 1);
        if (rp_before(first[second].x, first[second - 1].x)) --second;
        first[hole] = first[second]; hole = second;
    }
    if ((len & 1) == 0 && second == (len - 2) / 2) { second = 2 * (second + 1); first[hole] = first[second - 1]; hole = second - 1; }
    rp_push_heap(first, hole, top, value);
}
__device__ void rp_heap_sort(uint2 * first, int len)
{
    if (len >= 2) {
        int parent = (len - 2) / 2;
        while (true) { const uint2 v = first[parent]; rp_adjust_heap(first, parent, len, v); if (parent == 0) break; --parent; }
    }
    int last = len;
    while (last > 1) { --last; const uint2 v = first[last]; first[last] = first[0]; rp_adjust_heap(first, 0, last, v); }
}

// __move_median_to_first(first, first + 1, mid, last - 1): index of the median of the three
__device__ __forceinline__ int rp_median3(const uint2 * a, int f, int l)
{
    const int ia = f + 1, ib = f + (l - f) / 2, ic = l - 1;
    const unsigned ka = a[ia].x, kb = a[ib].x, kc = a[ic].x;
    if (rp_before(ka, kb)) { if (rp_before(kb, kc)) return ib; if (rp_before(ka, kc)) return ic; return ia; }
    if (rp_before(ka, kc)) return ia;
    if (rp_before(kb, kc)) return ic;
    return ib;
}

// One __unguarded_partition_pivot of [f, l) by one warp on a range in shared memory (phase 2).  Lp / Rp: 16-bit scratch of the
// range.  Returns the cut; tied_left = tied elements in [f, cut) afterwards.
template <class IdxT>
__device__ int rp_partition_small(uint2 * a, IdxT * Lp, IdxT * Rp, int f, int l, int tied_total, int & tied_left)
{
    const int lane = threadIdx.x & 31;
    const unsigned lt = (1u << lane) - 1u;
    const int m = rp_median3(a, f, l);
    __syncwarp();
    if (lane == 0) { const uint2 t = a[f]; a[f] = a[m]; a[m] = t; }
    __syncwarp();
    const unsigned pv = a[f].x;
    // U chunks of 32 keys in flight per memory round trip (the range may live in global memory: phase 2a)
    constexpr int U = sizeof(IdxT) == 4 ? 4 : 2;
    int cL = 0;                                                            // left stoppers, in scan order: !(x before pivot)
    for (int base = f + 1; base < l; base += 32 * U) {
        unsigned kk[U];
#pragma unroll
        for (int u = 0; u < U; ++u) { const int i = base + 32 * u + lane; kk[u] = i < l ? a[i].x : 0u; }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int i = base + 32 * u + lane;
            const bool fl = i < l && !rp_before(kk[u], pv);
            const unsigned bal = __ballot_sync(0xffffffffu, fl);
            if (fl) Lp[cL + __popc(bal & lt)] = (IdxT)i;
            cL += __popc(bal);
        }
    }
    int cR = 0;                                                            // right stoppers, from the right: !(pivot before x)
    for (int base = l - 1; base > f; base -= 32 * U) {
        unsigned kk[U];
#pragma unroll
        for (int u = 0; u < U; ++u) { const int i = base - 32 * u - lane; kk[u] = i > f ? a[i].x : 0u; }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int i = base - 32 * u - lane;
            const bool fl = i > f && !rp_before(pv, kk[u]);
            const unsigned bal = __ballot_sync(0xffffffffu, fl);
            if (fl) Rp[cR + __popc(bal & lt)] = (IdxT)i;
            cR += __popc(bal);
        }
    }
    __syncwarp();
    const int mn = min(cL, cR);
    int K = 0;
    for (int kb0 = 0; kb0 < mn; kb0 += 32 * U) {                           // the k-th left stopper <-> the k-th right stopper while left of it
        int lp[U], rp[U]; bool sw[U]; uint2 x[U], y[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int k = kb0 + 32 * u + lane;
            lp[u] = k < mn ? (int)Lp[k] : 1; rp[u] = k < mn ? (int)Rp[k] : 0;
            sw[u] = k < mn && lp[u] < rp[u];
        }
#pragma unroll
        for (int u = 0; u < U; ++u) if (sw[u]) { x[u] = a[lp[u]]; y[u] = a[rp[u]]; }
        bool done = false;
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (sw[u]) { a[lp[u]] = y[u]; a[rp[u]] = x[u]; }              // exchanges are a prefix in k: what follows a non-exchange stays put
            const unsigned bal = __ballot_sync(0xffffffffu, sw[u]);
            K += __popc(bal);
            done |= bal != 0xffffffffu;
        }
        if (done) break;
    }
    __syncwarp();
    const int cut = (K == 0) ? (int)Lp[0] : min(K < cL ? (int)Lp[K] : INT_MAX, (int)Rp[K - 1]);
    const bool left_small = (cut - f) <= (l - cut);
    const int s0 = left_small ? f : cut, s1 = left_small ? cut : l;
    int cnt = 0;
    for (int i = s0 + lane; i < s1; i += 32) cnt += (a[i].y & kTiedBit) ? 1 : 0;
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, s);
    tied_left = left_small ? cnt : tied_total - cnt;
    return cut;
}

constexpr int kSmall = 512;                   // ranges up to this size are finished by one warp in shared memory
constexpr int kMid = 2048;                    // ranges up to this size leave the level-synchronous phase: one warp follows each of them in global memory
constexpr int kMaxSeg = 64;                   // ranges of one level processed together in phase 1
constexpr int kWorkers = 8;                   // warps that take part in phase 2
constexpr int kStack = 48;                    // >= 2 * log2(kSmall) + slack: pending right parts below one small range

struct RpPhase1 {                             // shared memory of phase 1
    int4 seg[kMaxSeg];                        // first, last, depth left, tied elements
    unsigned piv[kMaxSeg];
    int cntL[kMaxSeg][32], cntR[kMaxSeg][32]; // stoppers per warp slice -> exclusive prefix (L) / suffix (R)
    int totL[kMaxSeg], totR[kMaxSeg], K[kMaxSeg], cut[kMaxSeg], tl[kMaxSeg];
};
struct RpPhase2 {                             // shared memory of phase 2 (one set per worker warp)
    uint2 a[kWorkers][kSmall];
    unsigned short Lp[kWorkers][kSmall], Rp[kWorkers][kSmall];
};
union RpShared { RpPhase1 p1; RpPhase2 p2; };

__global__ void __launch_bounds__(RP_THREADS)
k_seed_replay(Planes P, Seeds S, FrameCtrl * ctrl, int depth_override, int fixup)
{
    extern __shared__ __align__(16) unsigned char rp_smem_raw[];
    RpShared & sh = *reinterpret_cast<RpShared *>(rp_smem_raw);
    __shared__ int n_cur, n_next, n_small, small_head, steps;
    const int slot = blockIdx.x;
    FrameCtrl & C = ctrl[slot];
    const int n = min(C.n_seeds, S.cap);
    const int n_tied = C.n_tied;
    if (n_tied < 2 || n <= kSortThreshold || C.seed_overflow) return;      // no ties, or a plain (stable) insertion sort: K3's order is final
    const size_t so = (size_t)slot * S.cap;
    uint2 * a = S.rp_el + so;
    // work lists in global memory: two level lists (ping-pong) and the list of small ranges; every listed range holds more than
    // kSortThreshold elements and ranges of one list are disjoint, so cap / 3 entries each are ample
    int4 * qbase = S.rp_q + so;
    const int qcap = S.cap / 3;
    int4 * qa = qbase; int4 * qb = qbase + qcap; int4 * qs = qbase + 2 * (size_t)qcap;
    int * Lp = reinterpret_cast<int *>(S.us_mag + so);                     // free after k_seed_scatter
    int * Rp = reinterpret_cast<int *>(S.us_idx + so);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned lt = (1u << lane) - 1u;
    if (tid == 0) {
        int lg = 0; while ((2 << lg) <= n) ++lg;                           // std::__lg(n)
        const int4 root = make_int4(0, n, depth_override >= 0 ? depth_override : 2 * lg, n_tied);
        n_cur = 0; n_next = 0; n_small = 0; small_head = 0; steps = 0;
        if (n > kMid) { qa[0] = root; n_cur = 1; } else { qs[0] = root; n_small = 1; }
    }
    __syncthreads();

#ifdef LRB_FLOOD_STATS
    const long long t_p0 = clock64();
#endif
    // ---- phase 1: ranges longer than kSmall, level by level ---------------------------------------------------------------
    RpPhase1 & A = sh.p1;
    while (true) {
        const int na = n_cur;
        if (na == 0) break;
        for (int b0 = 0; b0 < na; b0 += kMaxSeg) {
            const int ns = min(kMaxSeg, na - b0);
            __syncthreads();
            // (0) depth limit / pivot
            if (tid < ns) {
                int4 g = qa[b0 + tid];
                if (g.z == 0) { rp_heap_sort(a + g.x, g.y - g.x); g.y = g.x; }          // sorted: nothing left to do in this range
                else { const int m = rp_median3(a, g.x, g.y); const uint2 t = a[g.x]; a[g.x] = a[m]; a[m] = t; A.piv[tid] = a[g.x].x; }
                A.seg[tid] = g;
            }
            __syncthreads();
            // (1) stoppers per warp slice
            for (int s = 0; s < ns; ++s) {
                const int f = A.seg[s].x, l = A.seg[s].y;
                if (l <= f) continue;
                const unsigned pv = A.piv[s];
                const int c = ((l - f - 1 + 31) / 32 + 31) / 32 * 32;                    // slice length: a multiple of 32
                const int i0 = f + 1 + warp * c, i1 = min(l, i0 + c);
                int cl = 0, cr = 0;
#pragma unroll 4
                for (int i = i0 + lane; i < i1; i += 32) { const unsigned k = a[i].x; cl += !rp_before(k, pv); cr += !rp_before(pv, k); }
#pragma unroll
                for (int sft = 16; sft > 0; sft >>= 1) { cl += __shfl_xor_sync(0xffffffffu, cl, sft); cr += __shfl_xor_sync(0xffffffffu, cr, sft); }
                if (lane == 0) { A.cntL[s][warp] = cl; A.cntR[s][warp] = cr; }
            }
            __syncthreads();
            // (2) prefix of the left stoppers, suffix of the right stoppers over the 32 slices
            for (int s = warp; s < ns; s += 32) {
                const int cl = A.cntL[s][lane], cr = A.cntR[s][lane];
                int il = cl, ir = cr;
#pragma unroll
                for (int sft = 1; sft < 32; sft <<= 1) {
                    const int v = __shfl_up_sync(0xffffffffu, il, sft); if (lane >= sft) il += v;
                    const int u = __shfl_down_sync(0xffffffffu, ir, sft); if (lane + sft < 32) ir += u;
                }
                A.cntL[s][lane] = il - cl; A.cntR[s][lane] = ir - cr;
                if (lane == 31) A.totL[s] = il;
                if (lane == 0) { A.totR[s] = ir; A.K[s] = 0; }
            }
            __syncthreads();
            // (3) stopper positions, in scan order
            for (int s = 0; s < ns; ++s) {
                const int f = A.seg[s].x, l = A.seg[s].y;
                if (l <= f) continue;
                const unsigned pv = A.piv[s];
                const int c = ((l - f - 1 + 31) / 32 + 31) / 32 * 32;
                const int i0 = f + 1 + warp * c, i1 = min(l, i0 + c);
                int bl = A.cntL[s][warp];
                // right stoppers of this slice that lie right of element i = (all of the slice) - (those up to and including i)
                const int slice_r = (warp == 0 ? A.totR[s] : A.cntR[s][warp - 1]) - A.cntR[s][warp];
                int br = A.cntR[s][warp] + slice_r;                                      // index of a right stopper = br - (right stoppers of the slice up to and including it)
                for (int base0 = i0; base0 < i1; base0 += 128) {             // four chunks of keys in flight per round trip
                    unsigned kk[4];
#pragma unroll
                    for (int u = 0; u < 4; ++u) { const int i = base0 + 32 * u + lane; kk[u] = i < i1 ? a[i].x : 0u; }
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        const int i = base0 + 32 * u + lane;
                        const bool fL = i < i1 && !rp_before(kk[u], pv), fR = i < i1 && !rp_before(pv, kk[u]);
                        const unsigned bL = __ballot_sync(0xffffffffu, fL), bR = __ballot_sync(0xffffffffu, fR);
                        if (fL) Lp[f + 1 + bl + __popc(bL & lt)] = i;
                        if (fR) Rp[f + 1 + br - 1 - __popc(bR & lt)] = i;
                        bl += __popc(bL); br -= __popc(bR);
                    }
                }
            }
            __syncthreads();
            // (4) exchanges: L[k] <-> R[k] while L[k] < R[k] (true for a prefix of k); K = length of that prefix
            for (int s = 0; s < ns; ++s) {
                const int f = A.seg[s].x, l = A.seg[s].y;
                if (l <= f) continue;
                const int mn = min(A.totL[s], A.totR[s]);
                for (int k0 = tid; k0 < mn; k0 += 4 * RP_THREADS) {            // four independent exchanges per thread in flight
                    int lp[4], rp[4], ln[4], rn[4]; bool sw[4];
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        const int k = k0 + u * RP_THREADS; const bool ok = k < mn, okn = k + 1 < mn;
                        lp[u] = ok ? Lp[f + 1 + k] : 0; rp[u] = ok ? Rp[f + 1 + k] : 0;
                        ln[u] = okn ? Lp[f + 2 + k] : 1; rn[u] = okn ? Rp[f + 2 + k] : 0;
                        sw[u] = ok && lp[u] < rp[u];
                    }
                    uint2 x[4], y[4];
#pragma unroll
                    for (int u = 0; u < 4; ++u) if (sw[u]) { x[u] = a[lp[u]]; y[u] = a[rp[u]]; }
#pragma unroll
                    for (int u = 0; u < 4; ++u) if (sw[u]) {
                        a[lp[u]] = y[u]; a[rp[u]] = x[u];
                        if (!(ln[u] < rn[u])) A.K[s] = k0 + u * RP_THREADS + 1;  // the last exchange of the prefix
                    }
                }
            }
            __syncthreads();
            // (5) the cut
            if (tid < ns) {
                const int f = A.seg[tid].x, l = A.seg[tid].y;
                if (l > f) {
                    const int K = A.K[tid];
                    A.cut[tid] = (K == 0) ? Lp[f + 1] : min(K < A.totL[tid] ? Lp[f + 1 + K] : INT_MAX, Rp[f + K]);
                }
                A.tl[tid] = 0;
            }
            __syncthreads();
            // (6) tied elements on the smaller side
            for (int s = 0; s < ns; ++s) {
                const int f = A.seg[s].x, l = A.seg[s].y;
                if (l <= f) continue;
                const int cut = A.cut[s];
                const bool left_small = (cut - f) <= (l - cut);
                const int s0 = left_small ? f : cut, s1 = left_small ? cut : l;
                int cnt = 0;
#pragma unroll 4
                for (int i = s0 + tid; i < s1; i += RP_THREADS) cnt += (a[i].y & kTiedBit) ? 1 : 0;
                if (__any_sync(0xffffffffu, cnt != 0)) {
#pragma unroll
                    for (int sft = 16; sft > 0; sft >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, sft);
                    if (lane == 0 && cnt) atomicAdd(&A.tl[s], cnt);
                }
            }
            __syncthreads();
            // (7) children: [f, cut) continues the loop, [cut, l) is the recursive call; both with the decremented depth limit
            if (tid < ns) {
                const int4 g = A.seg[tid];
                if (g.y > g.x) {
                    const int cut = A.cut[tid];
                    const bool left_small = (cut - g.x) <= (g.y - cut);
                    const int tl = left_small ? A.tl[tid] : g.w - A.tl[tid];
                    const int4 kid[2] = {make_int4(g.x, cut, g.z - 1, tl), make_int4(cut, g.y, g.z - 1, g.w - tl)};
                    for (int c = 0; c < 2; ++c) {
                        const int len = kid[c].y - kid[c].x;
                        if (len <= kSortThreshold || kid[c].w < 2) continue;
                        if (len > kMid) qb[atomicAdd(&n_next, 1)] = kid[c];
                        else qs[atomicAdd(&n_small, 1)] = kid[c];
                    }
                    atomicAdd(&steps, 1);
                }
            }
        }
        __syncthreads();
        if (tid == 0) { n_cur = n_next; n_next = 0; }
        int4 * t = qa; qa = qb; qb = t;
        __syncthreads();
    }
    __syncthreads();

#ifdef LRB_FLOOD_STATS
    const long long t_p1 = clock64();
#endif
    // ---- phase 2a: ranges of kSmall+1 .. kMid elements, one warp each, in place in global memory (all 32 warps) ------------
    // The level-synchronous phase pays its barriers and memory round trips once per RANGE of a level (the CTA walks the ranges
    // one after the other), which is right for a few long ranges and wrong for the dozens of mid-sized ones a frame with 40 tied
    // pairs produces two or three levels down (measured: 300 of the kernel's 400 us).  Here every warp follows one such range
    // down on its own stack; what falls to kSmall elements or below joins the list of small ranges.
    {
        __shared__ int mid_head;
        if (tid == 0) mid_head = 0;
        __syncthreads();
        const int n_mid = n_small;                                        // items listed by phase 1 (any size up to kMid)
        __syncthreads();
        int4 stack[kStack];
        int my_steps = 0;
        while (true) {
            int item = 0;
            if (lane == 0) item = atomicAdd(&mid_head, 1);
            item = __shfl_sync(0xffffffffu, item, 0);
            if (item >= n_mid) break;
            const int4 g = qs[item];
            if (g.y - g.x <= kSmall) continue;                            // phase 2 takes it
            int sp = 0;
            stack[sp++] = g;
            while (sp > 0) {
                const int4 r = stack[--sp];
                int f = r.x, l = r.y, d = r.z, t = r.w;
                while (l - f > kSortThreshold && t >= 2) {
                    if (l - f <= kSmall) { if (lane == 0) qs[atomicAdd(&n_small, 1)] = make_int4(f, l, d, t); break; }
                    if (d == 0) { if (lane == 0) rp_heap_sort(a + f, l - f); __syncwarp(); break; }
                    --d;
                    int tl;
                    const int cut = rp_partition_small<int>(a, Lp + f + 1, Rp + f + 1, f, l, t, tl);
                    ++my_steps;
                    if (l - cut > kSortThreshold && t - tl >= 2 && sp < kStack) stack[sp++] = make_int4(cut, l, d, t - tl);
                    l = cut; t = tl;
                }
            }
            __syncwarp();
        }
        if (lane == 0 && my_steps) atomicAdd(&steps, my_steps);
        __threadfence_block();
        __syncthreads();
    }
    // ---- phase 2: small ranges, one warp each, in shared memory ----------------------------------------------------------
    if (warp < kWorkers) {
        RpPhase2 & B = sh.p2;
        uint2 * sa = B.a[warp]; unsigned short * sL = B.Lp[warp]; unsigned short * sR = B.Rp[warp];
        int4 stack[kStack];
        int my_steps = 0;
        const int nsm = n_small;
        while (true) {
            int item = 0;
            if (lane == 0) item = atomicAdd(&small_head, 1);
            item = __shfl_sync(0xffffffffu, item, 0);
            if (item >= nsm) break;
            const int4 g = qs[item];
            const int len = g.y - g.x;
            if (len > kSmall) continue;                                    // followed in phase 2a
            for (int i = lane; i < len; i += 32) sa[i] = a[g.x + i];
            __syncwarp();
            int sp = 0;
            stack[sp++] = make_int4(0, len, g.z, g.w);
            while (sp > 0) {
                const int4 r = stack[--sp];
                int f = r.x, l = r.y, d = r.z, t = r.w;
                while (l - f > kSortThreshold && t >= 2) {
                    if (d == 0) {                                          // std::__partial_sort(first, last, last): the range ends up sorted
                        if (lane == 0) rp_heap_sort(sa + f, l - f);
                        __syncwarp();
                        break;
                    }
                    --d;
                    int tl;
                    const int cut = rp_partition_small<unsigned short>(sa, sL, sR, f, l, t, tl);
                    ++my_steps;
                    if (l - cut > kSortThreshold && t - tl >= 2 && sp < kStack) stack[sp++] = make_int4(cut, l, d, t - tl);   // __introsort_loop(cut, last, depth_limit)
                    l = cut; t = tl;                                       // last = cut
                }
            }
            __syncwarp();
            for (int i = lane; i < len; i += 32) a[g.x + i] = sa[i];
            __syncwarp();
        }
        if (lane == 0 && my_steps) atomicAdd(&steps, my_steps);
    }
    __syncthreads();
    if (tid == 0) C.replay_steps = steps;
#ifdef LRB_FLOOD_STATS
    const long long t_p2 = clock64();
#endif
    if (!fixup) return;
    // ---- final order: stable by key of the arrangement above.  Only tied peaks move, inside their group's rank range. ----
    // (1) position of every rank
#pragma unroll 4
    for (int p = tid; p < n; p += RP_THREADS) Lp[a[p].y & ~kTiedBit] = p;
    __syncthreads();
    // (2) final rank of a tied peak = first rank of its group + members of the group left of it; remember its pixel
    unsigned * sidx = S.idx + so;
#pragma unroll 4
    for (int p = tid; p < n; p += RP_THREADS) {
        const uint2 el = a[p];
        if (!(el.y & kTiedBit)) continue;
        const int rank = (int)(el.y & ~kTiedBit);
        int before = 0;
        for (int j = (int)el.x; j < n; ++j) {
            const int pj = Lp[j];
            if (a[pj].x != el.x) break;
            before += (pj < p);
        }
        Rp[rank] = (int)el.x + before;
    }
    __syncthreads();
#pragma unroll 4
    for (int p = tid; p < n; p += RP_THREADS) {                            // Lp is free again: park the pixel of every tied rank
        const uint2 el = a[p];
        if (el.y & kTiedBit) { const int rank = (int)(el.y & ~kTiedBit); Lp[rank] = (int)sidx[rank]; }
    }
    __syncthreads();
    // (3) move the per-seed records (pixel, bin, flood threshold; the other per-seed arrays still hold their initial values)
    const size_t pbase = (size_t)slot * P.stride;
#pragma unroll 4
    for (int p = tid; p < n; p += RP_THREADS) {
        const uint2 el = a[p];
        if (!(el.y & kTiedBit)) continue;
        const int rank = (int)(el.y & ~kTiedBit), fr = Rp[rank];
        if (fr == rank) continue;
        const unsigned ix = (unsigned)Lp[rank];
        const int r = ix / P.w, c = ix - r * P.w;
        const size_t po = pbase + (size_t)r * P.pitch + c;
        const int b = P.bin[po];
        sidx[fr] = ix;
        S.bin[so + fr] = (uint8_t)b;
        const float2 gseed = P.dxy[po];
        S.thr[so + fr] = __fmul_rn(kTraceKeep, orient_energy(gseed.x, gseed.y, b));       // reference src/filter.cpp:119-121
    }
#ifdef LRB_FLOOD_STATS
    __syncthreads();
    if (tid == 0) { C.dbg_replay[0] = t_p1 - t_p0; C.dbg_replay[1] = t_p2 - t_p1; C.dbg_replay[2] = clock64() - t_p2; C.dbg_replay[3] = n_small; }
#endif
}

} // namespace

void launch_seed_replay(const Planes & P, const Seeds & S, FrameCtrl * ctrl, int n_slots, int depth_override, bool fixup, cudaStream_t st)
{
    static PerDeviceOnce once;
    once.run([] { cudaFuncSetAttribute(k_seed_replay, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(RpShared)); });
    LRB_CARVEOUT(k_seed_replay);
    k_seed_replay<<<n_slots, RP_THREADS, sizeof(RpShared), st>>>(P, S, ctrl, depth_override, fixup ? 1 : 0);
}

} // namespace lrb
