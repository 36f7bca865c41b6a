// k4_flood.cu — K4 `flood_rounds`: the reference's greedy, order-dependent region growing as speculative rounds.
//
// Replaces flood_init_mask / flood / find_components (reference src/filter.cpp:101-153, src/line_detector.cpp:92-122).
// Reference semantics: seeds are taken in descending magnitude; a seed whose pixel is already visited is skipped;
// otherwise an 8-connected flood over G_b (b = the seed's bin; G_b(p) > 0.75f * G_b(seed)) claims every reached pixel
// in a shared `visited` mask (1-px border pre-visited) — whether or not the component is kept (size > 5) afterwards.
//
// Parallel restatement (hard part 1): per round, every live seed floods TENTATIVELY against the visited
// state of the previous rounds and stamps each reached pixel with atomicMin(owner, epoch:rank).  A seed commits iff no
// higher-priority live seed stamped any of its pixels.  Since a tentative region can only shrink as more pixels become
// visited, a conflict-free region is exactly the region the sequential algorithm would produce.  Each round the
// highest-priority live seed commits, so the rounds terminate; in practice 6-8 rounds suffice.
//
// owner word = [epoch:8 | rank:24].  Epochs DEcrease from round to round (and frame to frame), so a newer stamp always
// wins atomicMin over stale ones and the 4 B/px plane is re-initialised (0xFF..) only when the 8-bit epoch wraps.
// "visited(p)" is never materialised: p is visited iff status[rank(owner[p])] == owner[p], i.e. the seed that stamped
// it committed in exactly that epoch.  Conflicts are detected by whoever performs the second atomicMin on a pixel.
#include "common.cuh"
#include "kernels.h"

namespace lrb {

namespace {

constexpr int FL_WARPS = 8;
constexpr int FL_QCAP  = 1024;    // BFS ring per warp: must hold three consecutive BFS levels (frontier-, not region-bound)

// reference src/filter.cpp:113-116 neighbour order (dr, dc)
static __device__ __constant__ const int kDr[8] = {0, 0, 1, -1, -1, -1, 1, 1};
static __device__ __constant__ const int kDc[8] = {-1, 1, 0, 0, -1, 1, -1, 1};

__device__ __forceinline__ bool is_visited(unsigned o, unsigned epoch_first, const unsigned * status)
{
    if (owner_epoch(o) > epoch_first) return false;            // never stamped in this frame
    return __ldcg(status + owner_rank(o)) == o;
}

// one tentative flood of seed `rank` by one warp
//
// The warp runs a level-synchronous BFS over the seed's tentative region: each step expands up to four pixels of the
// current level against their eight neighbours (32 candidates, one per lane).  The ring `q` keeps every pixel pushed
// for the previous, the current and the next level.  A pixel this seed stamped is recognised by owner == key; a pixel
// stamped by a higher-priority seed (it stays part of the tentative region and is explored through) cannot be, so it
// is looked up in that three-level window instead — a neighbour of a level-L pixel always lies in level L-1, L or L+1.
__device__ __forceinline__ void flood_one(const Planes & P, const Seeds & S, FrameCtrl & C, int slot, int rank,
                                          unsigned epoch, unsigned epoch_first, unsigned * q)
{
    const int lane = threadIdx.x & 31;
    const size_t so = (size_t)slot * S.cap;
    unsigned * status = S.status + so;
    if (__ldcg(status + rank) != kStatusLive) return;

    const size_t pbase = (size_t)slot * P.stride;
    unsigned * owner = P.owner + pbase;
    const float * dx = P.dx + pbase; const float * dy = P.dy + pbase; const uint8_t * mask8 = P.mask8 + pbase;

    const unsigned key = (epoch << 24) | (unsigned)rank;
    const unsigned spx = S.idx[so + rank];
    const int sr = spx / P.w, sc = spx - sr * P.w;
    const size_t spo = (size_t)sr * P.pitch + sc;
    if (is_visited(__ldcg(owner + spo), epoch_first, status)) {    // line_detector.cpp:103: seed already swallowed
        if (lane == 0) status[rank] = kStatusDropped;
        return;
    }
    const int b = S.bin[so + rank];
    const float thr = S.thr[so + rank];
    const float sb = kSin[b], cb = kCos[b];

    // the seed itself: G_b(seed) > 0.75 G_b(seed) holds by construction (filter.cpp:119-131)
    bool failed = false, overflow = false;
    if (lane == 0) {
        const unsigned old = atomicMin(owner + spo, key);
        if (owner_epoch(old) == epoch) {
            if (old < key) failed = true;                                   // a higher-priority seed already reached this pixel
            else if (old != key) S.fail_epoch[so + owner_rank(old)] = epoch;
        }
        q[0] = ((unsigned)sr << 16) | (unsigned)sc;
    }
    __syncwarp();
    int head = 0, tail = 1;                 // absolute positions in the push sequence; ring index = pos & (FL_QCAP-1)
    int s_prev = 0, s_cur = 0, s_next = 1;  // first push of the previous / current / next BFS level

    while (head < tail) {
        if (head >= s_next) { s_prev = s_cur; s_cur = s_next; s_next = tail; }
        const int cnt = min(4, s_next - head);                             // a step never straddles two levels
        const int pslot = lane >> 3, k = lane & 7;
        const bool act = pslot < cnt;
        const unsigned prc = q[(head + (act ? pslot : 0)) & (FL_QCAP - 1)];
        head += cnt;
        const int r = (int)(prc >> 16) + kDr[k], c = (int)(prc & 0xFFFFu) + kDc[k];
        bool elig = act && r >= 1 && r < P.h - 1 && c >= 1 && c < P.w - 1;      // filter.cpp:101-107: border is visited
        const size_t po = (size_t)(elig ? r : 0) * P.pitch + (elig ? c : 0);
        if (elig) elig = (__ldg(mask8 + po) >> b) & 1;                          // dilated bin mask, else G_b = 0
        if (elig) {
            const float g = fabsf(__fadd_rn(__fmul_rn(__ldg(dx + po), sb), __fmul_rn(__ldg(dy + po), cb)));
            elig = g > thr;                                                     // filter.cpp:131
        }
        unsigned o = 0;
        if (elig) {
            o = __ldcg(owner + po);
            if (o == key) elig = false;                                         // already in this region
            else if (is_visited(o, epoch_first, status)) elig = false;
        }
        // the same pixel may be reached from several parents in one step: lowest lane handles it
        const unsigned rcq = ((unsigned)r << 16) | (unsigned)(c & 0xFFFF);
        const unsigned long long mkey = elig ? (unsigned long long)rcq : ((1ull << 40) | (unsigned)lane);
        const unsigned same = __match_any_sync(0xffffffffu, mkey);
        const bool leader = elig && (lane == __ffs(same) - 1);
        bool reached = false, lostcand = false;
        if (leader) {
            if (owner_epoch(o) == epoch && o < key) lostcand = true;            // stamped this round by a higher-priority seed
            else {
                const unsigned old = atomicMin(owner + po, key);
                if (old == key) { /* cannot happen after de-duplication */ }
                else if (owner_epoch(old) == epoch) {
                    if (old < key) lostcand = true;
                    else { S.fail_epoch[so + owner_rank(old)] = epoch; reached = true; }   // the other seed loses
                } else reached = true;
            }
        }
        if (lostcand) {                                                         // not ours: was it pushed in the last three levels?
            failed = true;
            bool seen = false;
            for (int i = s_prev; i < tail; ++i) seen |= (q[i & (FL_QCAP - 1)] == rcq);
            reached = !seen;                                                    // still part of the tentative region: explore through it
        }
        const unsigned pm = __ballot_sync(0xffffffffu, reached);
        const int np = __popc(pm);
        if (tail + np - s_prev > FL_QCAP) { overflow = true; break; }           // the ring must keep three levels
        if (reached) q[(tail + __popc(pm & ((1u << lane) - 1))) & (FL_QCAP - 1)] = rcq;
        tail += np;
        __syncwarp();
    }

    failed = __any_sync(0xffffffffu, failed);
    if (lane == 0) {
        if (overflow) { C.flood_overflow = 1; S.fail_epoch[so + rank] = epoch; }
        else {
            if (failed) S.fail_epoch[so + rank] = epoch;
            S.size[so + rank] = tail;
        }
    }
    __syncwarp();
}

// Rounds work on a WINDOW of the highest-priority live seeds (ranks [win_lo, win_hi)): every live seed outside the
// window has lower priority than every seed inside, so it cannot influence what the window commits, and by the time its
// turn comes most of it has been swallowed by a committed region (seeds along one line are dominated by the line's
// strongest seed) and is dropped without ever flooding.  This keeps the work close to the sequential algorithm's.
__global__ void __launch_bounds__(FL_WARPS * 32)
k_flood_round(Planes P, Seeds S, FrameCtrl * ctrl, int round, unsigned epoch_first)
{
    __shared__ unsigned q[FL_WARPS][FL_QCAP];
    const int slot = blockIdx.y;
    FrameCtrl & C = ctrl[slot];
    if (C.live[round] == 0 || C.flood_overflow) return;
    const int lo = C.win_lo, hi = C.win_hi;
    const int warp = threadIdx.x >> 5;
    const unsigned epoch = epoch_first - (unsigned)round;
    for (int rank = lo + blockIdx.x * FL_WARPS + warp; rank < hi; rank += gridDim.x * FL_WARPS)
        flood_one(P, S, C, slot, rank, epoch, epoch_first, q[warp]);
}

// commit every window seed that nobody out-prioritised in this round; the last block to finish moves the window
__global__ void __launch_bounds__(256)
k_flood_commit(Seeds S, FrameCtrl * ctrl, int round, unsigned epoch_first)
{
    __shared__ int s_last;
    const int slot = blockIdx.y;
    FrameCtrl & C = ctrl[slot];
    if (C.live[round] == 0 || C.flood_overflow) return;
    const int n = C.n_seeds;
    const int lo = C.win_lo, hi = C.win_hi;
    const unsigned epoch = epoch_first - (unsigned)round;
    const size_t so = (size_t)slot * S.cap;
    int live = 0, first = 0x7fffffff;
    for (int i = lo + blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += gridDim.x * blockDim.x) {
        if (S.status[so + i] == kStatusLive) {
            if (S.fail_epoch[so + i] != epoch) S.status[so + i] = (epoch << 24) | (unsigned)i;
            else { ++live; first = min(first, i); }
        }
    }
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) { live += __shfl_xor_sync(0xffffffffu, live, s); first = min(first, __shfl_xor_sync(0xffffffffu, first, s)); }
    if ((threadIdx.x & 31) == 0 && live) { atomicAdd(&C.win_live, live); atomicMin(&C.win_first, first); }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) s_last = (atomicAdd(&C.win_blocks, 1) == (int)gridDim.x - 1);
    __syncthreads();
    if (s_last && threadIdx.x == 0) {
        __threadfence();
        const int failed = atomicAdd(&C.win_live, 0);                 // window seeds that stay live (out-prioritised this round)
        const int firstf = atomicMin(&C.win_first, 0x7fffffff);
        const int budget = max(kFloodWindowMin, n / kFloodWindowDiv);
        const int nlo = failed ? min(firstf, hi) : hi;
        const int nhi = min(n, hi + max(0, budget - failed));
        C.win_lo = nlo; C.win_hi = nhi;
        C.win_live = 0; C.win_first = 0x7fffffff; C.win_blocks = 0;
        C.live[round + 1] = failed + (n - hi);                       // seeds that may still have to flood
        C.rounds_run = round + 1;
    }
}

// Fallback when a speculative flood ran out of on-chip space or the epoch budget: the reference algorithm verbatim,
// one warp per frame, seeds strictly in order, queue in global memory (capacity = one entry per pixel).
__global__ void __launch_bounds__(32)
k_flood_serial(Planes P, Seeds S, Comps K, FrameCtrl * ctrl, unsigned epoch_first, unsigned epoch_serial)
{
    const int slot = blockIdx.x;
    FrameCtrl & C = ctrl[slot];
    const int n = C.n_seeds;
    const int lane = threadIdx.x;
    const size_t so = (size_t)slot * S.cap;
    unsigned * status = S.status + so;
    const size_t pbase = (size_t)slot * P.stride;
    unsigned * owner = P.owner + pbase;
    const float * dx = P.dx + pbase; const float * dy = P.dy + pbase; const uint8_t * mask8 = P.mask8 + pbase;
    unsigned * q = K.px + (size_t)slot * K.cap_px;             // not in use yet at this stage

    for (int rank = 0; rank < n; ++rank) {
        if (__ldcg(status + rank) != kStatusLive) continue;
        const unsigned key = (epoch_serial << 24) | (unsigned)rank;
        const unsigned spx = S.idx[so + rank];
        const int sr = spx / P.w, sc = spx - sr * P.w;
        if (is_visited(__ldcg(owner + (size_t)sr * P.pitch + sc), epoch_first, status)) {
            if (lane == 0) status[rank] = kStatusDropped;
            __syncwarp();
            continue;
        }
        const int b = S.bin[so + rank];
        const float thr = S.thr[so + rank];
        const float sb = kSin[b], cb = kCos[b];
        int head = 0, tail = 0;
        bool first = true;
        while (true) {
            int cnt; bool act; int r, c;
            if (first) { cnt = 1; act = (lane == 0); r = sr; c = sc; }
            else {
                cnt = min(4, tail - head);
                if (cnt == 0) break;
                const int pslot = lane >> 3, k = lane & 7;
                act = pslot < cnt;
                const unsigned prc = __ldcg(q + head + (act ? pslot : 0));
                head += cnt;
                r = (int)(prc >> 16) + kDr[k]; c = (int)(prc & 0xFFFFu) + kDc[k];
            }
            bool elig = act && r >= 1 && r < P.h - 1 && c >= 1 && c < P.w - 1;
            const size_t po = (size_t)(elig ? r : 0) * P.pitch + (elig ? c : 0);
            if (elig && !first) elig = (__ldg(mask8 + po) >> b) & 1;
            if (elig && !first) {
                const float g = fabsf(__fadd_rn(__fmul_rn(__ldg(dx + po), sb), __fmul_rn(__ldg(dy + po), cb)));
                elig = g > thr;
            }
            if (elig) {
                const unsigned o = __ldcg(owner + po);
                if (o == key || is_visited(o, epoch_first, status)) elig = false;
            }
            const unsigned rcq = ((unsigned)r << 16) | (unsigned)(c & 0xFFFF);
            const unsigned long long mkey = elig ? (unsigned long long)rcq : ((1ull << 40) | (unsigned)lane);
            const unsigned same = __match_any_sync(0xffffffffu, mkey);
            const bool leader = elig && (lane == __ffs(same) - 1);
            if (leader) __stcg(owner + po, key);
            const unsigned pm = __ballot_sync(0xffffffffu, leader);
            if (leader) __stcg(q + tail + __popc(pm & ((1u << lane) - 1)), rcq);
            tail += __popc(pm);
            first = false;
            __syncwarp();
        }
        if (lane == 0) { S.size[so + rank] = tail; __stcg(status + rank, key); }
        __syncwarp();
    }
    if (lane == 0) { C.flood_overflow = 0; C.flood_done = 1; }
}

} // namespace

void launch_flood_rounds(const Planes & P, const Seeds & S, FrameCtrl * ctrl, int n_slots, int seeds_hint,
                         int round_begin, int round_end, unsigned epoch_first, cudaStream_t st)
{
    // flood: CTAs striding over the window's ranks; commit: a fixed small grid striding over the window
    const int fblocks = max(1, min((seeds_hint + FL_WARPS - 1) / FL_WARPS, 148 * 4 / max(1, min(n_slots, 4))));
    dim3 gf(fblocks, n_slots), gc(32, n_slots);
    for (int r = round_begin; r < round_end; ++r) {
        k_flood_round<<<gf, FL_WARPS * 32, 0, st>>>(P, S, ctrl, r, epoch_first);
        k_flood_commit<<<gc, 256, 0, st>>>(S, ctrl, r, epoch_first);
    }
}

void launch_flood_serial(const Planes & P, const Seeds & S, const Comps & K, FrameCtrl * ctrl, int n_slots,
                         unsigned epoch_first, unsigned epoch_serial, cudaStream_t st)
{
    k_flood_serial<<<n_slots, 32, 0, st>>>(P, S, K, ctrl, epoch_first, epoch_serial);
}

} // namespace lrb
