// k4_flood.cu — K4 `flood_rounds`: the reference's greedy, order-dependent region growing as speculative rounds.
//
// Replaces flood_init_mask / flood / find_components (reference src/filter.cpp:101-153, src/line_detector.cpp:92-122).
// Reference semantics: seeds are taken in descending magnitude; a seed whose pixel is already visited is skipped;
// otherwise an 8-connected flood over G_b (b = the seed's bin; G_b(p) > 0.75f * G_b(seed)) claims every reached pixel
// in a shared `visited` mask (1-px border pre-visited) — whether or not the component is kept (size > 5) afterwards.
//
// Parallel restatement (hard part 1).  A round takes a WINDOW of the highest-priority live seeds.
//   flood : every window seed floods TENTATIVELY against the visited mask of the previous rounds, stamps each reached
//           pixel with a fire-and-forget atomicMin(owner, epoch:rank) and appends it to a pixel list.
//   verify: a seed commits iff every pixel of its list still carries its own stamp, i.e. no higher-priority seed of
//           the window reached any of them.  A tentative region can only shrink as more pixels become visited, so a
//           conflict-free region IS the region the sequential algorithm produces.  Committed regions are written into
//           the visited mask; the highest-priority window seed always commits, so rounds terminate.
// Live seeds outside the window all have lower priority than every window seed, so they cannot change what the window
// commits — and by the time their turn comes most of them lie inside a committed region (seeds along one line are
// dominated by the line's strongest seed) and are dropped without ever flooding, as in the sequential algorithm.
//
// Latency is what bounds this stage (a 2000-px line is a 1000-level BFS), so the exploration loop touches only data
// that is constant during a round — mask8, dx, dy and the visited mask — through the non-coherent L1 path; the owner
// stamps are write-only here and are inspected by the verify kernel.  owner word = [epoch | rank]; epochs DEcrease from
// round to round and frame to frame, so a newer stamp always wins over stale ones and the plane is re-initialised only
// when the epoch field wraps.
#include "common.cuh"
#include "kernels.h"

namespace lrb {

namespace {

constexpr int FL_WARPS = 4;
constexpr int FL_QCAP  = 1024;    // BFS ring per warp: must hold three consecutive BFS levels (frontier-, not region-bound)

// reference src/filter.cpp:113-116 neighbour order (dr, dc)
static __device__ __constant__ const int kDr[8] = {0, 0, 1, -1, -1, -1, 1, 1};
static __device__ __constant__ const int kDc[8] = {-1, 1, 0, 0, -1, 1, -1, 1};

// Pixel lists live in a per-frame pool of 32-word chunks: word 0 = index of the next chunk + 1 (0: none), words 1..31 =
// pixels as (row << 16 | col).  The pool is recycled every round.
__device__ __forceinline__ int pool_alloc(FrameCtrl & C, int cap_chunks)
{
    int c = 0;
    if ((threadIdx.x & 31) == 0) c = atomicAdd(&C.pool_cursor, 1);
    c = __shfl_sync(0xffffffffu, c, 0);
    return c < cap_chunks ? c : -1;
}

// one tentative flood of seed `rank` by one warp: level-synchronous BFS, up to four pixels of the current level are
// expanded against their eight neighbours per step (one candidate per lane).  The ring keeps every pixel pushed for the
// previous, the current and the next level; a neighbour of a level-L pixel lies in level L-1, L or L+1, so "already
// reached" is a search in that window.
__device__ __forceinline__ void flood_one(const Planes & P, const Seeds & S, const Comps & K, FrameCtrl & C, int slot, int rank,
                                          unsigned epoch, unsigned * q)
{
    const int lane = threadIdx.x & 31;
    const size_t so = (size_t)slot * S.cap;
    if (__ldcg(S.status + so + rank) != kStatusLive) return;

    const size_t pbase = (size_t)slot * P.stride;
    unsigned * owner = P.owner + pbase;
    const float * dx = P.dx + pbase; const float * dy = P.dy + pbase;
    const uint8_t * mask8 = P.mask8 + pbase; const uint8_t * vis = P.vis + pbase;
    unsigned * pool = K.pool + (size_t)slot * K.cap_pool;
    const int cap_chunks = K.cap_pool / 32;

    const unsigned key = (epoch << 24) | (unsigned)rank;
    const unsigned spx = S.idx[so + rank];
    const int sr = spx / P.w, sc = spx - sr * P.w;
    const size_t spo = (size_t)sr * P.pitch + sc;
    if (__ldg(vis + spo)) {                                   // line_detector.cpp:103: seed already swallowed
        if (lane == 0) S.status[so + rank] = kStatusDropped;
        return;
    }
    const int b = S.bin[so + rank];
    const float thr = S.thr[so + rank];
    const float sb = kSin[b], cb = kCos[b];

    // the seed itself: G_b(seed) > 0.75 G_b(seed) holds by construction (filter.cpp:119-131)
    if (lane == 0) { atomicMin(owner + spo, key); q[0] = ((unsigned)sr << 16) | (unsigned)sc; }
    __syncwarp();
    int head = 0, tail = 1;                 // absolute positions in the push sequence; ring index = pos & (FL_QCAP-1)
    int s_prev = 0, s_cur = 0, s_next = 1;  // first push of the previous / current / next BFS level
    int flushed = 0, first_chunk = -1, last_chunk = -1;
    bool overflow = false;

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
        if (elig) {
            const unsigned m8 = __ldg(mask8 + po), v = __ldg(vis + po);
            const float gx = __ldg(dx + po), gy = __ldg(dy + po);
            const float g = fabsf(__fadd_rn(__fmul_rn(gx, sb), __fmul_rn(gy, cb)));
            elig = ((m8 >> b) & 1u) && !v && (g > thr);                         // dilated bin mask, not visited, filter.cpp:131
        }
        const unsigned rcq = ((unsigned)r << 16) | (unsigned)(c & 0xFFFF);
        if (elig) {                                                             // reached before (previous / this / next level)?
            bool seen = false;
            for (int i = s_prev; i < tail; ++i) seen |= (q[i & (FL_QCAP - 1)] == rcq);
            elig = !seen;
        }
        // the same pixel may be reached from several parents in one step: lowest lane handles it
        const unsigned long long mkey = elig ? (unsigned long long)rcq : ((1ull << 40) | (unsigned)lane);
        const unsigned same = __match_any_sync(0xffffffffu, mkey);
        const bool reached = elig && (lane == __ffs(same) - 1);
        if (reached) atomicMin(owner + po, key);                                // result unused: a reduction, off the critical path
        const unsigned pm = __ballot_sync(0xffffffffu, reached);
        const int np = __popc(pm);
        if (tail + np - min(s_prev, flushed) > FL_QCAP) { overflow = true; break; }   // the ring must keep three levels + unflushed pixels
        if (reached) q[(tail + __popc(pm & ((1u << lane) - 1))) & (FL_QCAP - 1)] = rcq;
        tail += np;
        __syncwarp();
        // spill completed groups of 31 pixels to the pool
        while (tail - flushed >= 31) {
            const int ch = pool_alloc(C, cap_chunks);
            if (ch < 0) { overflow = true; break; }
            pool[(size_t)ch * 32 + lane] = lane ? q[(flushed + lane - 1) & (FL_QCAP - 1)] : 0u;
            if (lane == 0 && last_chunk >= 0) pool[(size_t)last_chunk * 32] = (unsigned)ch + 1u;
            if (first_chunk < 0) first_chunk = ch;
            last_chunk = ch; flushed += 31;
        }
        if (overflow) break;
    }
    if (!overflow && tail > flushed) {                                          // the last, partial chunk
        const int ch = pool_alloc(C, cap_chunks);
        if (ch < 0) overflow = true;
        else {
            pool[(size_t)ch * 32 + lane] = (lane && flushed + lane - 1 < tail) ? q[(flushed + lane - 1) & (FL_QCAP - 1)] : 0u;
            if (lane == 0 && last_chunk >= 0) pool[(size_t)last_chunk * 32] = (unsigned)ch + 1u;
            if (first_chunk < 0) first_chunk = ch;
        }
    }
    if (lane == 0) {
        if (overflow) C.flood_overflow = 1;
        S.size[so + rank] = overflow ? -1 : tail;
        S.list_head[so + rank] = first_chunk;
    }
    __syncwarp();
}

__global__ void __launch_bounds__(FL_WARPS * 32)
k_flood_round(Planes P, Seeds S, Comps K, FrameCtrl * ctrl, int round, unsigned epoch_first)
{
    __shared__ unsigned q[FL_WARPS][FL_QCAP];
    const int slot = blockIdx.y;
    FrameCtrl & C = ctrl[slot];
    if (C.live[round] == 0 || C.flood_overflow) return;
    const int lo = C.win_lo, hi = C.win_hi;
    const int warp = threadIdx.x >> 5;
    const unsigned epoch = epoch_first - (unsigned)round;
    for (int rank = lo + blockIdx.x * FL_WARPS + warp; rank < hi; rank += gridDim.x * FL_WARPS)
        flood_one(P, S, K, C, slot, rank, epoch, q[warp]);
}

// verify + commit + mark visited, one warp per window seed; the last block to finish moves the window
__global__ void __launch_bounds__(256)
k_flood_verify(Planes P, Seeds S, Comps K, FrameCtrl * ctrl, int round, unsigned epoch_first)
{
    __shared__ int s_last;
    const int slot = blockIdx.y;
    FrameCtrl & C = ctrl[slot];
    if (C.live[round] == 0 || C.flood_overflow) return;
    const int n = C.n_seeds;
    const int lo = C.win_lo, hi = C.win_hi;
    const unsigned epoch = epoch_first - (unsigned)round;
    const size_t so = (size_t)slot * S.cap, pbase = (size_t)slot * P.stride;
    const unsigned * owner = P.owner + pbase;
    uint8_t * vis = P.vis + pbase;
    const unsigned * pool = K.pool + (size_t)slot * K.cap_pool;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    int live = 0, first = 0x7fffffff;
    for (int rank = lo + blockIdx.x * nw + warp; rank < hi; rank += gridDim.x * nw) {
        if (S.status[so + rank] != kStatusLive) continue;
        const unsigned key = (epoch << 24) | (unsigned)rank;
        const int size = S.size[so + rank];
        // pass 1: is every pixel still stamped with this seed's key?
        bool mine = true;
        int left = size;
        for (int ch = S.list_head[so + rank]; ch >= 0 && left > 0; ) {
            const unsigned w = __ldcg(pool + (size_t)ch * 32 + lane);
            const int here = min(left, 31);
            if (lane >= 1 && lane <= here) mine &= (__ldcg(owner + (size_t)(w >> 16) * P.pitch + (w & 0xFFFFu)) == key);
            left -= here;
            ch = (int)__shfl_sync(0xffffffffu, w, 0) - 1;
        }
        mine = __all_sync(0xffffffffu, mine);
        if (!mine) { ++live; first = min(first, rank); continue; }
        // pass 2: the region is final — it becomes part of the visited mask (whether or not it is kept later)
        left = size;
        for (int ch = S.list_head[so + rank]; ch >= 0 && left > 0; ) {
            const unsigned w = __ldcg(pool + (size_t)ch * 32 + lane);
            const int here = min(left, 31);
            if (lane >= 1 && lane <= here) vis[(size_t)(w >> 16) * P.pitch + (w & 0xFFFFu)] = 1;
            left -= here;
            ch = (int)__shfl_sync(0xffffffffu, w, 0) - 1;
        }
        if (lane == 0) S.status[so + rank] = key;
    }
    if (lane == 0 && live) { atomicAdd(&C.win_live, live); atomicMin(&C.win_first, first); }
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
        C.win_live = 0; C.win_first = 0x7fffffff; C.win_blocks = 0; C.pool_cursor = 0;
        C.live[round + 1] = failed + (n - hi);                       // seeds that may still have to flood
        C.rounds_run = round + 1;
    }
}

// Fallback when a speculative flood ran out of on-chip or pool space, or the round budget: the reference algorithm
// verbatim, one warp per frame, seeds strictly in order, queue in global memory (capacity = one entry per pixel).
__global__ void __launch_bounds__(32)
k_flood_serial(Planes P, Seeds S, Comps K, FrameCtrl * ctrl, unsigned epoch_serial)
{
    const int slot = blockIdx.x;
    FrameCtrl & C = ctrl[slot];
    const int n = C.n_seeds;
    const int lane = threadIdx.x;
    const size_t so = (size_t)slot * S.cap;
    unsigned * status = S.status + so;
    const size_t pbase = (size_t)slot * P.stride;
    unsigned * owner = P.owner + pbase;
    uint8_t * vis = P.vis + pbase;
    const float * dx = P.dx + pbase; const float * dy = P.dy + pbase; const uint8_t * mask8 = P.mask8 + pbase;
    unsigned * q = K.px + (size_t)slot * K.cap_px;             // not in use yet at this stage

    for (int rank = 0; rank < n; ++rank) {
        if (__ldcg(status + rank) != kStatusLive) continue;
        const unsigned key = (epoch_serial << 24) | (unsigned)rank;
        const unsigned spx = S.idx[so + rank];
        const int sr = spx / P.w, sc = spx - sr * P.w;
        if (__ldcg(vis + (size_t)sr * P.pitch + sc)) {
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
            if (elig) elig = (__ldcg(vis + po) == 0) && (__ldcg(owner + po) != key);
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
        for (int i = lane; i < tail; i += 32) {                 // the region is final: visited
            const unsigned w = __ldcg(q + i);
            __stcg(vis + (size_t)(w >> 16) * P.pitch + (w & 0xFFFFu), (uint8_t)1);
        }
        if (lane == 0) { S.size[so + rank] = tail; __stcg(status + rank, key); }
        __syncwarp();
    }
    if (lane == 0) { C.flood_overflow = 0; C.flood_done = 1; }
}

} // namespace

void launch_flood_rounds(const Planes & P, const Seeds & S, const Comps & K, FrameCtrl * ctrl, int n_slots, int seeds_hint,
                         int round_begin, int round_end, unsigned epoch_first, cudaStream_t st)
{
    // flood: CTAs striding over the window's ranks; verify: warps striding over the window
    const int fblocks = max(1, min((seeds_hint + FL_WARPS - 1) / FL_WARPS, 148 * 8 / max(1, min(n_slots, 4))));
    dim3 gf(fblocks, n_slots), gv(148 / max(1, min(n_slots, 4)) + 1, n_slots);
    for (int r = round_begin; r < round_end; ++r) {
        k_flood_round<<<gf, FL_WARPS * 32, 0, st>>>(P, S, K, ctrl, r, epoch_first);
        k_flood_verify<<<gv, 256, 0, st>>>(P, S, K, ctrl, r, epoch_first);
    }
}

void launch_flood_serial(const Planes & P, const Seeds & S, const Comps & K, FrameCtrl * ctrl, int n_slots,
                         unsigned epoch_serial, cudaStream_t st)
{
    k_flood_serial<<<n_slots, 32, 0, st>>>(P, S, K, ctrl, epoch_serial);
}

} // namespace lrb
