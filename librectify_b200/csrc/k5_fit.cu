// k5_fit.cu — component selection, the reference's BFS pixel order, and K5 `fit_segments`.
//
// Replaces the tail of find_components (size > 5 filter, reference src/line_detector.cpp:111-119), the pixel order of
// flood (src/filter.cpp:123-151), fit_lines_to_components (src/line_detector.cpp:66-89) and fit_line_parameters
// (src/geometry.cpp:20-61).
//
// After the speculative rounds each committed seed owns exactly its final region (owner[p] == status[rank]), but the
// rounds discover pixels in arbitrary order.  The reference's weighted sums run in BFS acceptance order, and float
// addition is not associative, so one warp per kept component re-walks its region in the reference's order:
// four accepted pixels x eight neighbours (in the reference's neighbour order) are tested per step, duplicates inside
// the step resolved to the lowest lane, and accepted pixels appended in lane order — which is exactly the FIFO order
// of the sequential flood.  The fit then evaluates every sum sequentially in that order (element terms are produced by
// 32 lanes in parallel, a single lane per running sum adds them up), so raw segments are bit-identical to the CPU code.
#include "common.cuh"
#include "eig.cuh"
#include "kernels.h"

namespace lrb {

namespace {

constexpr int FIT_WARPS = 4;
constexpr int FIT_RING = 256;        // power of two
constexpr int FIT_CT = 8;            // owner-plane tiles cached per warp during the walk of a large component

static __device__ __constant__ const int kDr[8] = {0, 0, 1, -1, -1, -1, 1, 1};
static __device__ __constant__ const int kDc[8] = {-1, 1, 0, 0, -1, 1, -1, 1};

// kept components in seed order: ordered compaction over ranks; one block per frame
__global__ void __launch_bounds__(1024) k_comp_select(Seeds S, Comps K, FrameCtrl * ctrl, int rounds_total)
{
    __shared__ int wcnt[32], wpx[32];
    __shared__ int n_big;
    const int slot = blockIdx.x;
    FrameCtrl & C = ctrl[slot];
    const int n = C.n_seeds;
    const size_t so = (size_t)slot * S.cap, ko = (size_t)slot * K.cap;
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    if (C.fit_done) return;                               // repair pass for another slot: this one is already fitted
    // floods not finished (more rounds or the serial fallback needed): publish nothing; the host re-runs from here
    if ((C.live[rounds_total] != 0 && !C.flood_done) || C.flood_overflow) {
        if (t == 0) { C.n_comp = 0; C.n_px = 0; C.fit_queue = 0; C.fit_queue_small = 0; C.n_big = 0; }
        return;
    }
    int * biglist = S.esc + so;                           // the flood's escalation list is free by now
    if (t == 0) n_big = 0;
    // Ordered compaction with ONE block-wide scan: every WARP owns a contiguous run of ranks and walks it 32 ranks at a time
    // (coalesced; a run per thread touched 4096 cache lines from one SM and missed the L1 on every load: 63 us for a single frame).
    const int per = ((n + 31) / 32 + 31) / 32 * 32;                  // ranks per warp, a multiple of 32
    const int w0 = min(n, warp * per), w1 = min(n, w0 + per);
    int kc = 0, kp = 0;                                               // kept components / pixels of this warp's run (warp-uniform)
#pragma unroll 4
    for (int i = w0 + lane; i - lane < w1; i += 32) {
        const bool in = i < w1;
        const unsigned st = in ? S.status[so + i] : 0u;
        const int sz = in ? S.size[so + i] : 0;
        const bool keep = in && (st > kStatusDropped) && (sz > kComponentMinSize);      // line_detector.cpp:111
        int v = keep ? sz : 0;
#pragma unroll
        for (int sft = 16; sft > 0; sft >>= 1) v += __shfl_xor_sync(0xffffffffu, v, sft);
        kc += __popc(__ballot_sync(0xffffffffu, keep)); kp += v;
    }
    if (lane == 0) { wcnt[warp] = kc; wpx[warp] = kp; }
    __syncthreads();
    if (warp == 0) {
        int a = wcnt[lane], b = wpx[lane];
#pragma unroll
        for (int s = 1; s < 32; s <<= 1) {
            const int x = __shfl_up_sync(0xffffffffu, a, s), y = __shfl_up_sync(0xffffffffu, b, s);
            if (lane >= s) { a += x; b += y; }
        }
        wcnt[lane] = a; wpx[lane] = b;
    }
    __syncthreads();
    int cbase = warp ? wcnt[warp - 1] : 0, pbase = warp ? wpx[warp - 1] : 0;
    for (int i = w0 + lane; i - lane < w1; i += 32) {
        const bool in = i < w1;
        const unsigned st = in ? S.status[so + i] : 0u;
        const int sz = in ? S.size[so + i] : 0;
        const bool keep = in && (st > kStatusDropped) && (sz > kComponentMinSize);
        const unsigned bm = __ballot_sync(0xffffffffu, keep);
        int inc = keep ? sz : 0;                                      // inclusive prefix of the kept sizes inside the 32 ranks
#pragma unroll
        for (int sft = 1; sft < 32; sft <<= 1) { const int v = __shfl_up_sync(0xffffffffu, inc, sft); if (lane >= sft) inc += v; }
        const int tot = __shfl_sync(0xffffffffu, inc, 31);
        if (keep) {
            const int cpos = cbase + __popc(bm & ((1u << lane) - 1u)), ppos = pbase + inc - sz;
            if (cpos < K.cap && ppos + sz <= K.cap_px) {
                K.rank[ko + cpos] = i; K.off[ko + cpos] = ppos; K.size[ko + cpos] = sz;
                if (sz > kFitSmallMax) biglist[atomicAdd(&n_big, 1)] = cpos;     // any order: every component writes its own segment record
            }
            else C.comp_overflow = 1;
        }
        cbase += __popc(bm); pbase += tot;
    }
    __syncthreads();
    if (t == 0) { C.n_comp = min(wcnt[31], K.cap); C.n_px = wpx[31]; C.fit_queue = 0; C.fit_queue_small = 0; C.n_big = n_big; }
}

// running sums evaluated strictly in list order: lanes 0..NS-1 each own one sum; terms come through shared memory
template <int NS>
__device__ __forceinline__ void seq_accumulate(float (*stage)[32], const float (&term)[NS], int valid, int lane, bool first_chunk, float & acc)
{
#pragma unroll
    for (int s = 0; s < NS; ++s) stage[s][lane] = term[s];
    __syncwarp();
    if (lane < NS) {
        // terms are fetched four at a time (the adds stay strictly sequential: that is the reference's summation order)
        const float4 * row = reinterpret_cast<const float4 *>(stage[lane]);
        for (int k0 = 0; k0 < valid; k0 += 4) {
            const float4 v = row[k0 >> 2];
            acc = (first_chunk && k0 == 0) ? v.x : __fadd_rn(acc, v.x);
            if (k0 + 1 < valid) acc = __fadd_rn(acc, v.y);
            if (k0 + 2 < valid) acc = __fadd_rn(acc, v.z);
            if (k0 + 3 < valid) acc = __fadd_rn(acc, v.w);
        }
    }
    __syncwarp();
}

__global__ void __launch_bounds__(FIT_WARPS * 32)
k_fit_segments(Planes P, Seeds S, Comps K, FrameCtrl * ctrl, int use_cache)
{
    // per warp: the large-component path uses stage[3][32], val[FIT_RING], ring[FIT_RING]; the sub-warp path the same memory as
    // four sets of stage[3][kFitSmallMax], val[kFitSmallMax], ring[kFitSmallMax]
    __shared__ __align__(16) float stage_all[FIT_WARPS][4 * 3 * kFitSmallMax];
    __shared__ float val_all[FIT_WARPS][FIT_RING];         // weights of a component that fits the ring (no re-reads from global memory)
    __shared__ unsigned ring_all[FIT_WARPS][FIT_RING];     // the last FIT_RING accepted pixels of the walk (the BFS frontier lives here)
    __shared__ unsigned ctile_all[FIT_WARPS][FIT_CT * 32]; // per warp: 'pixel of this component, not listed yet' bitmaps of the last FIT_CT tiles touched
    __shared__ int cid_all[FIT_WARPS][FIT_CT];
    static_assert(4 * kFitSmallMax <= FIT_RING && 4 * 3 * kFitSmallMax >= 3 * 32, "shared-memory sets of the two paths");
    const int slot = blockIdx.y;
    FrameCtrl & C = ctrl[slot];
    if (C.fit_done) return;
    const int ncomp = C.n_big;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float (*stage)[32] = reinterpret_cast<float (*)[32]>(stage_all[warp]);
    unsigned * ring = ring_all[warp];
    float * vring = val_all[warp];
    const int * biglist = S.esc + (size_t)slot * S.cap;
    unsigned * ctile = ctile_all[warp]; int * cid = cid_all[warp];
    const size_t so = (size_t)slot * S.cap, ko = (size_t)slot * K.cap, pbase = (size_t)slot * P.stride;
    unsigned * owner = P.owner + pbase;
    const float2 * dxy = P.dxy + pbase;

    // ---- large components (more than kFitSmallMax pixels): one warp each, handed out dynamically from the list k_comp_select
    // made; the next one is requested before the current one is walked.  Every warp starts here, so the long walks start first.
    int nxt = 0;
    if (lane == 0) nxt = atomicAdd(&C.fit_queue, 1);
    while (true) {
        const int bpos = __shfl_sync(0xffffffffu, nxt, 0);
        if (bpos >= ncomp) break;
        if (lane == 0) nxt = atomicAdd(&C.fit_queue, 1);
        const int comp = biglist[bpos];
        const int rank = K.rank[ko + comp];
        const int n = K.size[ko + comp];
        unsigned * px = K.px + (size_t)slot * K.cap_px + K.off[ko + comp];
        float * val = K.val + (size_t)slot * K.cap_px + K.off[ko + comp];
        const unsigned ckey = S.status[so + rank];
        const unsigned listed = 0xFF000000u | (unsigned)rank;                  // epoch 0xFF: reads as "never stamped" later
        const int b = S.bin[so + rank];
        const unsigned spx = S.idx[so + rank];

        // ---- pixel list in the reference's acceptance order (filter.cpp:123-151) ----------------------------------
        // The frontier is read from a shared-memory ring (one memory round trip per step: the owner probe); the list in
        // global memory is write-only here unless the frontier outgrows the ring.
        int head = 0, tail = 1, victim = 0;
        if (lane < FIT_CT) cid[lane] = -1;                                      // the cached answers belong to one component
        if (lane == 0) {
            const int sr = spx / P.w, sc = spx - sr * P.w;
            const unsigned rc0 = ((unsigned)sr << 16) | (unsigned)sc;
            __stcg(px, rc0); ring[0] = rc0;
            __stcg(owner + (size_t)sr * P.pitch + sc, listed);
        }
        __syncwarp();
        const int pslot = lane >> 3;
        const int ndr = kDr[lane & 7], ndc = kDc[lane & 7];                     // this lane's neighbour offset (filter.cpp:113-116 order)
        while (head < tail) {
            const int cnt = min(4, tail - head);
            const bool act = pslot < cnt;
            const int pi = head + (act ? pslot : 0);
            const unsigned prc = (tail - pi <= FIT_RING) ? ring[pi & (FIT_RING - 1)] : __ldcg(px + pi);
            head += cnt;
            const int r = (int)(prc >> 16) + ndr, c = (int)(prc & 0xFFFFu) + ndc;
            const size_t po = (size_t)r * P.pitch + c;                          // parents are interior pixels: in range
            // Is the candidate a pixel of this component that is not listed yet?  The answer lives in the owner plane (one memory
            // round trip per step: the whole cost of a long walk), so the warp keeps the answer for the last eight 32 x 32 tiles it
            // touched as bitmaps in shared memory: a step whose candidates all fall into cached tiles costs no global load.
            const int tid_ = ((r >> 5) << 16) | (c >> 5);
            int cslot = -1;
#pragma unroll
            for (int k = 0; k < FIT_CT; ++k) cslot = (cid[k] == tid_) ? k : cslot;
            if (!use_cache) cslot = -1;                                         // LRB_FIT_NOCACHE=1 (experiments): every step probes the owner plane
            const bool miss = act && cslot < 0;
            const bool any_miss = __any_sync(0xffffffffu, miss);
            bool mine;
            if (!any_miss) mine = act && ((ctile[(act ? cslot : 0) * 32 + (r & 31)] >> (c & 31)) & 1u);
            else mine = __ldcg(owner + po) == ckey;
            // A pixel can be the neighbour of several of the (up to four) parents of this step; the reference's FIFO order
            // gives it to the first of them, i.e. the lowest lane.  Two candidates coincide exactly when the pixel lies in
            // the 3x3 neighbourhood of an earlier parent, so three shuffles (independent of the memory round trip above)
            // replace a warp-wide match.
            bool dup = false;
#pragma unroll
            for (int q = 0; q < 3; ++q) {
                const unsigned o = __shfl_sync(0xffffffffu, prc, q * 8);
                const int dr_ = r - (int)(o >> 16), dc_ = c - (int)(o & 0xFFFFu);
                dup |= (q < pslot) && (dr_ >= -1) && (dr_ <= 1) && (dc_ >= -1) && (dc_ <= 1);
            }
            const bool take = act && mine && !dup;
            const unsigned rcq = ((unsigned)r << 16) | (unsigned)(c & 0xFFFF);
            const unsigned tm = __ballot_sync(0xffffffffu, take);
            __syncwarp();                                                       // every lane has read its parent: ring entries may be overwritten
            if (take) {
                const int pos = tail + __popc(tm & ((1u << lane) - 1));
                if (pos < n) __stcg(px + pos, rcq);
                ring[pos & (FIT_RING - 1)] = rcq;
                __stcg(owner + po, listed);
                if (cslot >= 0) atomicAnd(&ctile[cslot * 32 + (r & 31)], ~(1u << (c & 31)));
            }
            tail += __popc(tm);
            __syncwarp();
            if (any_miss && use_cache) {                                        // bring in the tile of the first candidate that missed
                const int leader = __ffs(__ballot_sync(0xffffffffu, miss)) - 1;
                const int ltid = __shfl_sync(0xffffffffu, tid_, leader);
                const int y0 = (ltid >> 16) << 5, x0 = (ltid & 0xFFFF) << 5;
                const int v = victim; victim = (victim + 1) % FIT_CT;
                const int x = x0 + lane;
                unsigned oo[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) oo[i] = (y0 + i < P.h && x < P.pitch) ? __ldcg(owner + (size_t)(y0 + i) * P.pitch + x) : 0u;
                unsigned wordv = 0u;
#pragma unroll
                for (int i = 0; i < 32; ++i) { const unsigned bal = __ballot_sync(0xffffffffu, oo[i] == ckey); wordv = (lane == i) ? bal : wordv; }
                ctile[v * 32 + lane] = wordv;
                if (lane == 0) cid[v] = ltid;
                __syncwarp();
            }
        }
        // tail == n by construction (the committed flood counted the same set)
        const int m = min(tail, n);
        const bool small = m <= FIT_RING;                       // the whole list is still in the ring (position p at ring[p])
#define LRB_PX(i)  (small ? ring[i] : __ldcg(px + (i)))
#define LRB_VAL(i) (small ? vring[i] : val[i])

        // ---- fit_line_parameters (geometry.cpp:20-61); X = (row, col), weights = G_b ------------------------------
        float acc = 0.f;
        // pass 1: weights and their sum
        for (int i0 = 0; i0 < m; i0 += 32) {
            const int i = i0 + lane; const int valid = min(32, m - i0);
            float t[1] = {0.f};
            if (i < m) {
                const unsigned rc = LRB_PX(i);
                const size_t po = (size_t)(rc >> 16) * P.pitch + (rc & 0xFFFFu);
                { const float2 g = __ldg(dxy + po); t[0] = orient_energy(g.x, g.y, b); }
                val[i] = t[0];
                if (small) vring[i] = t[0];
            }
            seq_accumulate<1>(stage, t, valid, lane, i0 == 0, acc);
        }
        const float wsum = __shfl_sync(0xffffffffu, acc, 0);
        __syncwarp();
        // pass 2: weighted mean a = sum X * (w / wsum)
        acc = 0.f;
        for (int i0 = 0; i0 < m; i0 += 32) {
            const int i = i0 + lane; const int valid = min(32, m - i0);
            float t[2] = {0.f, 0.f};
            if (i < m) {
                const unsigned rc = LRB_PX(i);
                const float wt = __fdiv_rn(LRB_VAL(i), wsum);
                t[0] = __fmul_rn((float)(rc >> 16), wt); t[1] = __fmul_rn((float)(rc & 0xFFFFu), wt);
            }
            seq_accumulate<2>(stage, t, valid, lane, i0 == 0, acc);
        }
        const float a0 = __shfl_sync(0xffffffffu, acc, 0), a1 = __shfl_sync(0xffffffffu, acc, 1);
        // pass 3: lower triangle of centered^T diag(wt) centered: (c_i * wt) * c_j
        acc = 0.f;
        for (int i0 = 0; i0 < m; i0 += 32) {
            const int i = i0 + lane; const int valid = min(32, m - i0);
            float t[3] = {0.f, 0.f, 0.f};
            if (i < m) {
                const unsigned rc = LRB_PX(i);
                const float wt = __fdiv_rn(LRB_VAL(i), wsum);
                const float c0 = __fsub_rn((float)(rc >> 16), a0), c1 = __fsub_rn((float)(rc & 0xFFFFu), a1);
                t[0] = __fmul_rn(__fmul_rn(c0, wt), c0);       // cov(0,0)
                t[1] = __fmul_rn(__fmul_rn(c1, wt), c0);       // cov(1,0)
                t[2] = __fmul_rn(__fmul_rn(c1, wt), c1);       // cov(1,1)
            }
            seq_accumulate<3>(stage, t, valid, lane, i0 == 0, acc);
        }
        float cov[4];
        cov[0] = __shfl_sync(0xffffffffu, acc, 0); cov[2] = __shfl_sync(0xffffffffu, acc, 1); cov[3] = __shfl_sync(0xffffffffu, acc, 2);
        cov[1] = cov[2];
        float eval[2], evec[4];
        sym_eig<2>(cov, eval, evec);                           // every lane computes the same thing
        const float d0 = evec[2], d1 = evec[3];                // column 1: largest eigenvalue = direction (geometry.cpp:40)
        const float n0 = evec[0], n1 = evec[1];                // column 0: normal
        // pass 4: projection range (order-free) and mean |normal residual| (ordered)
        acc = 0.f;
        float tmin = INFINITY, tmax = -INFINITY;
        for (int i0 = 0; i0 < m; i0 += 32) {
            const int i = i0 + lane; const int valid = min(32, m - i0);
            float t[1] = {0.f};
            if (i < m) {
                const unsigned rc = LRB_PX(i);
                const float c0 = __fsub_rn((float)(rc >> 16), a0), c1 = __fsub_rn((float)(rc & 0xFFFFu), a1);
                const float tt = __fadd_rn(__fmul_rn(c0, d0), __fmul_rn(c1, d1));
                tmin = fminf(tmin, tt); tmax = fmaxf(tmax, tt);
                t[0] = fabsf(__fadd_rn(__fmul_rn(c0, n0), __fmul_rn(c1, n1)));
            }
            seq_accumulate<1>(stage, t, valid, lane, i0 == 0, acc);
        }
#pragma unroll
        for (int s = 16; s > 0; s >>= 1) {
            tmin = fminf(tmin, __shfl_xor_sync(0xffffffffu, tmin, s));
            tmax = fmaxf(tmax, __shfl_xor_sync(0xffffffffu, tmax, s));
        }
        if (lane == 0) {
            SegRec s;
            s.y1 = __fadd_rn(a0, __fmul_rn(d0, tmin)); s.x1 = __fadd_rn(a1, __fmul_rn(d1, tmin));      // endpoint (row, col) -> (y, x)
            s.y2 = __fadd_rn(a0, __fmul_rn(d0, tmax)); s.x2 = __fadd_rn(a1, __fmul_rn(d1, tmax));
            s.weight = __fdiv_rn(wsum, (float)m);
            s.err = __fdiv_rn(acc, (float)m);
            s.group_id = -1;
            reinterpret_cast<SegRec *>(K.seg)[ko + comp] = s;
        }
#undef LRB_PX
#undef LRB_VAL
        __syncwarp();
    }

    // ---- small components: eight lanes each, four per warp ---------------------------------------------------------------
    // The typical component has 10-20 pixels, and its walk has one or two parents per step: a warp per component leaves most
    // lanes idle.  Here a step pops ONE parent per component and tests its eight neighbours in the reference's order (lane =
    // neighbour) — literally the sequential flood of src/filter.cpp:123-151 — and the sums run over shared-memory rows with one
    // lane per running sum, in list order, exactly as above.
    {
        const int sub = lane & 7, grp = lane >> 3;
        const int ncs = C.n_comp;
        float * sstage = stage_all[warp] + grp * 3 * kFitSmallMax;          // [3][kFitSmallMax]
        unsigned * sring = ring_all[warp] + grp * kFitSmallMax;
        float * sval = val_all[warp] + grp * kFitSmallMax;
        const int ndr = kDr[sub], ndc = kDc[sub];
        // sequential sum of row `r` (list order), by the lane that calls it
        auto row_sum = [&](int r, int m) {
            const float4 * row = reinterpret_cast<const float4 *>(sstage + r * kFitSmallMax);
            float acc = 0.f;
            for (int k0 = 0; k0 < m; k0 += 4) {
                const float4 v = row[k0 >> 2];
                acc = (k0 == 0) ? v.x : __fadd_rn(acc, v.x);
                if (k0 + 1 < m) acc = __fadd_rn(acc, v.y);
                if (k0 + 2 < m) acc = __fadd_rn(acc, v.z);
                if (k0 + 3 < m) acc = __fadd_rn(acc, v.w);
            }
            return acc;
        };
        int nx2 = 0;
        if (lane == 0) nx2 = atomicAdd(&C.fit_queue_small, 4);
        while (true) {
            const int c0 = __shfl_sync(0xffffffffu, nx2, 0);
            if (c0 >= ncs) break;
            if (lane == 0) nx2 = atomicAdd(&C.fit_queue_small, 4);
            const int comp = c0 + grp;
            int n = comp < ncs ? K.size[ko + comp] : 0;
            if (n > kFitSmallMax) n = 0;                                     // a large one: done above
            if (!__any_sync(0xffffffffu, n > 0)) continue;
            const bool on = n > 0;
            const int rank = on ? K.rank[ko + comp] : 0;
            unsigned * px = K.px + (size_t)slot * K.cap_px + (on ? K.off[ko + comp] : 0);
            const unsigned ckey = on ? S.status[so + rank] : 0u;
            const unsigned listed = 0xFF000000u | (unsigned)rank;
            const int b = on ? S.bin[so + rank] : 0;
            const unsigned spx = on ? S.idx[so + rank] : 0u;
            int head = 0, tail = on ? 1 : 0;
            if (on && sub == 0) {
                const int sr = spx / P.w, sc = spx - sr * P.w;
                const unsigned rc0 = ((unsigned)sr << 16) | (unsigned)sc;
                __stcg(px, rc0); sring[0] = rc0;
                __stcg(owner + (size_t)sr * P.pitch + sc, listed);
            }
            __syncwarp();
            while (__any_sync(0xffffffffu, head < tail)) {
                const bool act = head < tail;
                const unsigned prc = act ? sring[head] : 0x00010001u;
                head += act ? 1 : 0;
                const int r = (int)(prc >> 16) + ndr, c = (int)(prc & 0xFFFFu) + ndc;
                const size_t po = (size_t)r * P.pitch + c;
                const unsigned own = act ? __ldcg(owner + po) : 0u;
                const bool take = act && (own == ckey);
                const unsigned tm = (__ballot_sync(0xffffffffu, take) >> (8 * grp)) & 0xFFu;
                if (take) {
                    const int pos = tail + __popc(tm & ((1u << sub) - 1u));
                    const unsigned rcq = ((unsigned)r << 16) | (unsigned)(c & 0xFFFF);
                    if (pos < n) { __stcg(px + pos, rcq); sring[pos] = rcq; }
                    __stcg(owner + po, listed);
                }
                tail += __popc(tm);
                __syncwarp();
            }
            const int m = min(tail, n);
            // pass 1: weights and their sum
            for (int i = sub; i < m; i += 8) {
                const unsigned rc = sring[i];
                const size_t po = (size_t)(rc >> 16) * P.pitch + (rc & 0xFFFFu);
                const float2 g = __ldg(dxy + po);
                const float v = orient_energy(g.x, g.y, b);
                sval[i] = v; sstage[i] = v;
            }
            __syncwarp();
            float acc = (sub == 0 && m > 0) ? row_sum(0, m) : 0.f;
            const float wsum = __shfl_sync(0xffffffffu, acc, 0, 8);
            __syncwarp();
            // pass 2: weighted mean
            for (int i = sub; i < m; i += 8) {
                const unsigned rc = sring[i];
                const float wt = __fdiv_rn(sval[i], wsum);
                sstage[i] = __fmul_rn((float)(rc >> 16), wt); sstage[kFitSmallMax + i] = __fmul_rn((float)(rc & 0xFFFFu), wt);
            }
            __syncwarp();
            acc = (sub < 2 && m > 0) ? row_sum(sub, m) : 0.f;
            const float a0 = __shfl_sync(0xffffffffu, acc, 0, 8), a1 = __shfl_sync(0xffffffffu, acc, 1, 8);
            __syncwarp();
            // pass 3: covariance, lower triangle
            for (int i = sub; i < m; i += 8) {
                const unsigned rc = sring[i];
                const float wt = __fdiv_rn(sval[i], wsum);
                const float c0_ = __fsub_rn((float)(rc >> 16), a0), c1_ = __fsub_rn((float)(rc & 0xFFFFu), a1);
                sstage[i] = __fmul_rn(__fmul_rn(c0_, wt), c0_);
                sstage[kFitSmallMax + i] = __fmul_rn(__fmul_rn(c1_, wt), c0_);
                sstage[2 * kFitSmallMax + i] = __fmul_rn(__fmul_rn(c1_, wt), c1_);
            }
            __syncwarp();
            acc = (sub < 3 && m > 0) ? row_sum(sub, m) : 0.f;
            float cov[4];
            cov[0] = __shfl_sync(0xffffffffu, acc, 0, 8); cov[2] = __shfl_sync(0xffffffffu, acc, 1, 8); cov[3] = __shfl_sync(0xffffffffu, acc, 2, 8);
            cov[1] = cov[2];
            __syncwarp();
            float eval[2], evec[4];
            sym_eig<2>(cov, eval, evec);
            const float d0 = evec[2], d1 = evec[3], n0 = evec[0], n1 = evec[1];
            // pass 4: projection range (order-free) and mean |normal residual| (ordered)
            float tmin = INFINITY, tmax = -INFINITY;
            for (int i = sub; i < m; i += 8) {
                const unsigned rc = sring[i];
                const float c0_ = __fsub_rn((float)(rc >> 16), a0), c1_ = __fsub_rn((float)(rc & 0xFFFFu), a1);
                const float tt = __fadd_rn(__fmul_rn(c0_, d0), __fmul_rn(c1_, d1));
                tmin = fminf(tmin, tt); tmax = fmaxf(tmax, tt);
                sstage[i] = fabsf(__fadd_rn(__fmul_rn(c0_, n0), __fmul_rn(c1_, n1)));
            }
            __syncwarp();
            acc = (sub == 0 && m > 0) ? row_sum(0, m) : 0.f;
#pragma unroll
            for (int sft = 4; sft > 0; sft >>= 1) {
                tmin = fminf(tmin, __shfl_xor_sync(0xffffffffu, tmin, sft));
                tmax = fmaxf(tmax, __shfl_xor_sync(0xffffffffu, tmax, sft));
            }
            if (on && sub == 0) {
                SegRec sg;
                sg.y1 = __fadd_rn(a0, __fmul_rn(d0, tmin)); sg.x1 = __fadd_rn(a1, __fmul_rn(d1, tmin));
                sg.y2 = __fadd_rn(a0, __fmul_rn(d0, tmax)); sg.x2 = __fadd_rn(a1, __fmul_rn(d1, tmax));
                sg.weight = __fdiv_rn(wsum, (float)m);
                sg.err = __fdiv_rn(acc, (float)m);
                sg.group_id = -1;
                reinterpret_cast<SegRec *>(K.seg)[ko + comp] = sg;
            }
            __syncwarp();
        }
    }
}

// debug/test helper: label plane (kept component index, -1 elsewhere) from the pixel lists
__global__ void k_labels(Planes P, Comps K, FrameCtrl * ctrl, int * labels, int slot)
{
    const int ncomp = ctrl[slot].n_comp;
    const size_t ko = (size_t)slot * K.cap;
    for (int comp = blockIdx.x; comp < ncomp; comp += gridDim.x) {
        const unsigned * px = K.px + (size_t)slot * K.cap_px + K.off[ko + comp];
        const int n = K.size[ko + comp];
        for (int i = threadIdx.x; i < n; i += blockDim.x) {
            const unsigned rc = px[i];
            labels[(size_t)(rc >> 16) * P.w + (rc & 0xFFFFu)] = comp;
        }
    }
}

} // namespace

void launch_fit(const Planes & P, const Seeds & S, const Comps & K, FrameCtrl * ctrl, int n_slots, int rounds_total, cudaStream_t st)
{
    LRB_CARVEOUT(k_comp_select); LRB_CARVEOUT(k_fit_segments);
    // 14 KB of shared memory per CTA: without a stated preference the driver would give the SM its largest shared-memory split to
    // fit sixteen CTAs, and the kernels of the other engines on that SM would be left with a 28 KB L1
    static PerDeviceOnce once;
    once.run([&] { cudaFuncSetAttribute(k_fit_segments, cudaFuncAttributePreferredSharedMemoryCarveout, 50); });
    k_comp_select<<<n_slots, 1024, 0, st>>>(S, K, ctrl, rounds_total);
    dim3 grid(max(1, 148 * 8 / max(1, min(n_slots, 8))), n_slots);
    static const int use_cache = getenv("LRB_FIT_NOCACHE") ? 0 : 1;
    k_fit_segments<<<grid, FIT_WARPS * 32, 0, st>>>(P, S, K, ctrl, use_cache);
}

void launch_labels(const Planes & P, const Comps & K, FrameCtrl * ctrl, int * labels, int slot, cudaStream_t st)
{
    k_labels<<<256, 128, 0, st>>>(P, K, ctrl, labels, slot);
}

} // namespace lrb
