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
// Work distribution: one CTA per frame, warps pull ranges from a work list; a warp partitions its range with ballots
// (three passes over the range: left stoppers, right stoppers, exchanges), pushes the right part and continues with the
// left one, like the recursion it replays.  Keys are the dense ranks K3 computed (ties share a key; a SMALLER key is a
// LARGER magnitude), the payload is the K3 rank (magnitude desc, raster asc) with the tied flag in bit 31.
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

// One __unguarded_partition_pivot of [f, l) by one warp.  Lp / Rp: scratch, only entries f+1 .. l-1 are used (so warps
// working on disjoint ranges never meet).  Returns the cut; tied_left = tied elements in [f, cut) afterwards.
__device__ int rp_partition(uint2 * a, int * Lp, int * Rp, int f, int l, int tied_total, int & tied_left)
{
    const int lane = threadIdx.x & 31;
    const unsigned lt = (1u << lane) - 1u;
    // __move_median_to_first(first, first + 1, mid, last - 1)
    const int mid = f + (l - f) / 2;
    const int ia = f + 1, ib = mid, ic = l - 1;
    const unsigned ka = a[ia].x, kb = a[ib].x, kc = a[ic].x;
    int m;
    if (rp_before(ka, kb)) { if (rp_before(kb, kc)) m = ib; else if (rp_before(ka, kc)) m = ic; else m = ia; }
    else if (rp_before(ka, kc)) m = ia;
    else if (rp_before(kb, kc)) m = ic;
    else m = ib;
    __syncwarp();
    if (lane == 0) { const uint2 t = a[f]; a[f] = a[m]; a[m] = t; }
    __syncwarp();
    const unsigned pv = a[f].x;
    // left stoppers, in scan order: !(x before pivot)
    int cL = 0;
    for (int base = f + 1; base < l; base += 32) {
        const int i = base + lane;
        const bool fl = i < l && !rp_before(a[i].x, pv);
        const unsigned bal = __ballot_sync(0xffffffffu, fl);
        if (fl) Lp[f + 1 + cL + __popc(bal & lt)] = i;
        cL += __popc(bal);
    }
    // right stoppers, in scan order (from the right): !(pivot before x)
    int cR = 0;
    for (int base = l - 1; base > f; base -= 32) {
        const int i = base - lane;
        const bool fl = i > f && !rp_before(pv, a[i].x);
        const unsigned bal = __ballot_sync(0xffffffffu, fl);
        if (fl) Rp[f + 1 + cR + __popc(bal & lt)] = i;
        cR += __popc(bal);
    }
    __syncwarp();
    // exchanges: the k-th left stopper with the k-th right stopper while it lies left of it
    const int mn = min(cL, cR);
    int K = 0;
    for (int kb0 = 0; kb0 < mn; kb0 += 32) {
        const int k = kb0 + lane;
        bool sw = false;
        if (k < mn) {
            const int lp = Lp[f + 1 + k], rp = Rp[f + 1 + k];
            sw = lp < rp;
            if (sw) { const uint2 x = a[lp], y = a[rp]; a[lp] = y; a[rp] = x; }
        }
        const unsigned bal = __ballot_sync(0xffffffffu, sw);
        K += __popc(bal);
        if (bal != 0xffffffffu) break;
    }
    __syncwarp();
    int cut;
    if (K == 0) cut = Lp[f + 1];
    else cut = min(K < cL ? Lp[f + 1 + K] : INT_MAX, Rp[f + 1 + K - 1]);
    // tied elements on the smaller side
    const bool left_small = (cut - f) <= (l - cut);
    const int s0 = left_small ? f : cut, s1 = left_small ? cut : l;
    int cnt = 0;
    for (int i = s0 + lane; i < s1; i += 32) cnt += (a[i].y & kTiedBit) ? 1 : 0;
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, s);
    tied_left = left_small ? cnt : tied_total - cnt;
    return cut;
}

__global__ void __launch_bounds__(RP_THREADS)
k_seed_replay(Planes P, Seeds S, FrameCtrl * ctrl, int depth_override, int fixup)
{
    __shared__ int q_head, q_tail, q_lock, pending, steps;
    const int slot = blockIdx.x;
    FrameCtrl & C = ctrl[slot];
    const int n = min(C.n_seeds, S.cap);
    const int n_tied = C.n_tied;
    if (n_tied < 2 || n <= kSortThreshold || C.seed_overflow) return;      // no ties, or a plain (stable) insertion sort: K3's order is final
    const size_t so = (size_t)slot * S.cap;
    uint2 * a = S.rp_el + so;
    int4 * q = S.rp_q + so;
    int * Lp = reinterpret_cast<int *>(S.us_mag + so);                     // free after k_seed_scatter
    int * Rp = reinterpret_cast<int *>(S.us_idx + so);
    const int lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        int lg = 0; while ((2 << lg) <= n) ++lg;                           // std::__lg(n)
        q[0] = make_int4(0, n, depth_override >= 0 ? depth_override : 2 * lg, n_tied);
        q_head = 0; q_tail = 1; q_lock = 0; pending = 1; steps = 0;
    }
    __syncthreads();
    volatile int * vhead = &q_head; volatile int * vtail = &q_tail; volatile int * vpend = &pending;
    while (true) {
        int item = -1;
        if (lane == 0) {
            while (true) {
                if (*vpend == 0) { item = -2; break; }
                if (*vhead < *vtail) {
                    while (atomicCAS(&q_lock, 0, 1) != 0) { }
                    __threadfence_block();
                    if (*vhead < *vtail) { item = *vhead; *vhead = item + 1; }
                    __threadfence_block();
                    atomicExch(&q_lock, 0);
                    if (item >= 0) break;
                } else __nanosleep(200);
            }
        }
        item = __shfl_sync(0xffffffffu, item, 0);
        if (item == -2) break;
        const int4 seg = __ldcg(&q[item]);
        int f = seg.x, l = seg.y, d = seg.z, t = seg.w, my_steps = 0;
        while (l - f > kSortThreshold && t >= 2) {
            if (d == 0) {                                                  // std::__partial_sort(first, last, last): the range ends up sorted
                if (lane == 0) rp_heap_sort(a + f, l - f);
                __syncwarp();
                break;
            }
            --d;
            int tl;
            const int cut = rp_partition(a, Lp, Rp, f, l, t, tl);
            ++my_steps;
            const int tr = t - tl;
            if (l - cut > kSortThreshold && tr >= 2) {                     // __introsort_loop(cut, last, depth_limit)
                if (lane == 0) {
                    while (atomicCAS(&q_lock, 0, 1) != 0) { }
                    __threadfence_block();
                    const int at = *vtail;
                    q[at] = make_int4(cut, l, d, tr);
                    __threadfence_block();
                    *vtail = at + 1; atomicAdd(&pending, 1);
                    __threadfence_block();
                    atomicExch(&q_lock, 0);
                }
                __syncwarp();
            }
            l = cut; t = tl;                                               // last = cut
        }
        __threadfence_block();
        if (lane == 0) { atomicSub(&pending, 1); atomicAdd(&steps, my_steps); }
    }
    __syncthreads();
    if (threadIdx.x == 0) C.replay_steps = steps;
    if (!fixup) return;
    // ---- final order: stable by key of the arrangement above.  Only tied peaks move, inside their group's rank range. ----
    // (1) position of every rank
    for (int p = threadIdx.x; p < n; p += RP_THREADS) Lp[a[p].y & ~kTiedBit] = p;
    __syncthreads();
    // (2) final rank of a tied peak = first rank of its group + members of the group left of it; remember its pixel
    unsigned * sidx = S.idx + so;
    for (int p = threadIdx.x; p < n; p += RP_THREADS) {
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
    for (int p = threadIdx.x; p < n; p += RP_THREADS) {                    // Lp is free again: park the pixel of every tied rank
        const uint2 el = a[p];
        if (el.y & kTiedBit) { const int rank = (int)(el.y & ~kTiedBit); Lp[rank] = (int)sidx[rank]; }
    }
    __syncthreads();
    // (3) move the per-seed records (pixel, bin, flood threshold; the other per-seed arrays still hold their initial values)
    const size_t pbase = (size_t)slot * P.stride;
    for (int p = threadIdx.x; p < n; p += RP_THREADS) {
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
        S.thr[so + fr] = __fmul_rn(kTraceKeep, orient_energy(P.dx[po], P.dy[po], b));      // reference src/filter.cpp:119-121
    }
}

} // namespace

void launch_seed_replay(const Planes & P, const Seeds & S, FrameCtrl * ctrl, int n_slots, int depth_override, bool fixup, cudaStream_t st)
{
    LRB_CARVEOUT(k_seed_replay);
    k_seed_replay<<<n_slots, RP_THREADS, 0, st>>>(P, S, ctrl, depth_override, fixup ? 1 : 0);
}

} // namespace lrb
