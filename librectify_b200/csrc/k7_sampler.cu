// k7_sampler.cu — K7 `knuth_sampler`: the reference's sequential sampling chain, evaluated by 128 speculative workers.
//
// Replaces choice_knuth + the mt19937 draws inside RANSAC_Estimator::solve (reference src/math_utils.cpp:14-39,
// src/estimator.h:46-50).  One RANSAC sample is a 2-subset of [0,n) by Knuth's Algorithm S:
//     t = 0, m = 0;  while m < 2:  u = next uniform;  if (n - t) * u >= 2 - m: skip  else: pick t, ++m;   ++t
// so a sample consumes a data-dependent number of draws (its second pick + 1), and the 10^4 samples of a solve form ONE
// chain through the uniform stream: where sample i+1 starts depends on every draw before it.  The reference runs it
// under an `omp critical`; a single warp needs milliseconds for it.
//
// Parallel restatement (exact):  "select at stream position p" for a sample that started at o depends only on
// (p - o, m, k[p]), so ANY position can be tried as a sample start.  Worker j walks the chain speculatively from
// s_j = o + j*L.  Two chains that ever start a sample at the same position are identical from there on, and chains
// in the same neighbourhood coalesce within a handful of samples (a pick made because k[p] is very small is shared by
// every chain that has not picked yet).  Phase 1: every worker records the samples that start inside its own segment
// [s_j, s_j+1).  Phase 2: every worker keeps walking past the end of its segment until one of its sample starts is
// found among the recorded starts of the segment it has reached — from there the next worker's record continues the
// chain.  Phase 3: one warp follows the links from worker 0 (whose start IS the true chain start), which yields the
// true samples in order, and writes the first T of them.  If some link is missing (tiny n, or no coalescence inside
// the record capacity) the remaining samples are produced by the plain sequential walk — correct in every case.
//
// Integer-exact selection test: the stream stores k = 2^32 * u; select <=> (n-t) * k < (2-m) << 32.
#include "common.cuh"
#include "kernels.h"

namespace lrb {

namespace {

constexpr int SMP_AHEAD = 8;        // 32-position chunks kept in flight per warp

// Walk the chain from sample start `o`, calling f(start, s0, s1, next_start) after every completed sample (warp-uniform
// arguments); stops when f returns false.  Returns false if the stream ran out.
// The loop runs on 32-bit stream positions (the stream has 2^25 entries) and decides a chunk of 32 positions with one high
// multiply per lane: (n - t) * k < (2 - m) * 2^32  <=>  umulhi(n - t, k) < 2 - m  (n - t < 2^31, k < 2^32; exact).  Lanes past the
// end of a sample (t >= n, where n - t wraps) cannot be mistaken for a pick: Algorithm S is forced to pick at n - t == 2 - m, so a
// legitimate pick always lies at an earlier lane and the first set bit of the ballot is taken.  12 instructions per chunk instead
// of the 45 of the 64-bit form (profiles/r2_launches_final.md: the sampler was 20 M of a frame's 115 M warp instructions).
template <class F>
__device__ __forceinline__ bool chain_walk(const unsigned * __restrict__ stream, unsigned long long stream_len, long long n_ll,
                                           unsigned long long o_ull, F f)
{
    const unsigned lane = threadIdx.x & 31;
    const unsigned len = (unsigned)min(stream_len, 0xFFFFFF00ull);
    if (o_ull >= len) return false;
    const unsigned n = (unsigned)n_ll, last = len - 1;
    unsigned o = (unsigned)o_ull;                   // start of the current sample
    unsigned cur = o;                               // first unconsumed stream position
    unsigned base = o & ~31u;
    unsigned no = n + o;                            // n - t = no - p for position p
    unsigned lim = 2;                               // 2 - m
    int s0 = 0;
    unsigned kk[SMP_AHEAD];
#pragma unroll
    for (int a = 0; a < SMP_AHEAD; ++a) kk[a] = __ldg(stream + min(base + 32u * a + lane, last));
    while (true) {
#pragma unroll
        for (int a = 0; a < SMP_AHEAD; ++a) {
            const unsigned k = kk[a];
            const unsigned p = base + lane;
            if (base + 32u > len) return false;
            kk[a] = __ldg(stream + min(p + 32u * SMP_AHEAD, last));
            while (true) {
                const bool sel = (p >= cur) && (__umulhi(no - p, k) < lim);
                const unsigned bm = __ballot_sync(0xffffffffu, sel);
                if (bm == 0) break;
                const unsigned pos = base + (unsigned)(__ffs(bm) - 1);
                if (lim == 2) { s0 = (int)(pos - o); lim = 1; cur = pos + 1; }
                else {
                    const unsigned start = o;
                    const int s1 = (int)(pos - o);
                    lim = 2; o = pos + 1; cur = o; no = n + o;
                    if (!f((unsigned long long)start, s0, s1, (unsigned long long)o)) return true;
                }
            }
            base += 32u;
        }
    }
}

struct SamplerPlan { unsigned long long o0; unsigned long long L; long long n; };

__device__ __forceinline__ SamplerPlan make_plan(const FrameCtrl & C, int solve, int T, int workers)
{
    SamplerPlan pl;
    pl.n = C.n_obs[solve];
    pl.o0 = C.stream_pos;
    // expected draws of the whole solve: T * 2(n+1)/3 (the second pick of a 2-subset of [0,n) is at 2(n+1)/3 - 1 on
    // average); relative spread over 10^4 samples ~0.4 %, so +3 % covers it — the last worker has no upper bound anyway
    const unsigned long long D = (2ull * (unsigned long long)(pl.n + 1) * (unsigned long long)T) / 3ull;
    pl.L = (D + D / 32) / workers + 1;
    return pl;
}

// ---- phase 1: samples that start inside the worker's own segment ----------------------------------------------------------
__global__ void __launch_bounds__(32 * 4) k_sampler_segments(Lines L, FrameCtrl * ctrl, const unsigned * __restrict__ stream,
                                                             unsigned long long stream_len, int solve, int T)
{
    const int slot = blockIdx.y;
    FrameCtrl & C = ctrl[slot];
    if (C.n_solves != solve || C.num_obs < 2) return;
    const int j = blockIdx.x * 4 + (threadIdx.x >> 5);
    if (j >= L.samp_workers) return;
    const int lane = threadIdx.x & 31;
    const SamplerPlan pl = make_plan(C, solve, T, L.samp_workers);
    const size_t ro = ((size_t)slot * L.samp_workers + j) * L.samp_records;
    unsigned * rs = L.rec_start + ro; int2 * rp = reinterpret_cast<int2 *>(L.rec_pair) + ro;
    const unsigned long long sj = pl.o0 + (unsigned long long)j * pl.L;
    const unsigned long long send = sj + pl.L;      // the planned range already carries a 3 % (> 7 sigma) margin; beyond it the tail walk takes over
    int cnt = 0;
    unsigned long long next = sj;
    const bool ok = chain_walk(stream, stream_len, pl.n, sj, [&](unsigned long long start, int s0, int s1, unsigned long long nx) {
        if (lane == 0) { rs[cnt] = (unsigned)(start - pl.o0); rp[cnt] = make_int2(s0, s1); }
        ++cnt; next = nx;
        return (nx < send) && (cnt < L.samp_records - 1);
    });
    if (lane == 0) {
        const size_t wo = (size_t)slot * L.samp_workers + j;
        L.rec_cnt1[wo] = cnt; L.rec_cnt[wo] = cnt; L.rec_next[wo] = (unsigned)(next - pl.o0);
        L.rec_link[2 * wo] = -1; L.rec_link[2 * wo + 1] = 0;
        if (!ok && j == 0) C.stream_overflow = 1;      // only the true chain's own stretch must exist; speculative workers may fall off the end
        if (!ok) L.rec_cnt1[wo] = -1 - cnt;            // mark "ran out of stream": nothing after its records is known
    }
}

// ---- phase 2: walk on until a sample start is found among the starts recorded for the segment reached ----------------------
__global__ void __launch_bounds__(32 * 4) k_sampler_link(Lines L, FrameCtrl * ctrl, const unsigned * __restrict__ stream,
                                                         unsigned long long stream_len, int solve, int T)
{
    const int slot = blockIdx.y;
    FrameCtrl & C = ctrl[slot];
    if (C.n_solves != solve || C.num_obs < 2) return;
    const int j = blockIdx.x * 4 + (threadIdx.x >> 5);
    if (j >= L.samp_workers - 1) return;                 // the last worker has no successor
    const int lane = threadIdx.x & 31;
    const SamplerPlan pl = make_plan(C, solve, T, L.samp_workers);
    const size_t wo = (size_t)slot * L.samp_workers + j;
    if (L.rec_cnt1[wo] < 0) return;                    // ran out of stream in phase 1
    const size_t ro = wo * L.samp_records;
    unsigned * rs = L.rec_start + ro; int2 * rp = reinterpret_cast<int2 *>(L.rec_pair) + ro;
    int cnt = L.rec_cnt[wo];
    if (cnt >= L.samp_records - 1) return;               // record space used up inside its own segment: no link

    // is `start` (relative to o0) a recorded phase-1 sample start of the segment that contains it?
    auto lookup = [&](unsigned rel, int & k_out) -> int {
        int k = (int)min((unsigned long long)rel / pl.L, (unsigned long long)(L.samp_workers - 1));
        k_out = k;
        const size_t wk = (size_t)slot * L.samp_workers + k;
        int c1 = L.rec_cnt1[wk]; if (c1 < 0) c1 = -1 - c1;
        const unsigned * rk = L.rec_start + wk * L.samp_records;
        int found = -1;
        for (int i0 = 0; i0 < c1; i0 += 32) {
            const int i = i0 + lane;
            const unsigned bm = __ballot_sync(0xffffffffu, i < c1 && __ldg(rk + i) == rel);
            if (bm) { found = i0 + __ffs(bm) - 1; break; }
        }
        return found;
    };

    unsigned long long next = pl.o0 + L.rec_next[wo];
    int lk = -1, lb = 0;
    { int k; const int b = lookup((unsigned)(next - pl.o0), k); if (b >= 0 && k != j) { lk = k; lb = b; } }
    if (lk < 0) {
        chain_walk(stream, stream_len, pl.n, next, [&](unsigned long long start, int s0, int s1, unsigned long long nx) {
            if (lane == 0) { rs[cnt] = (unsigned)(start - pl.o0); rp[cnt] = make_int2(s0, s1); }
            ++cnt; next = nx;
            int k; const int b = lookup((unsigned)(nx - pl.o0), k);
            if (b >= 0 && k != j) { lk = k; lb = b; return false; }
            return cnt < L.samp_records - 1;
        });
    }
    if (lane == 0) {
        L.rec_cnt[wo] = cnt; L.rec_next[wo] = (unsigned)(next - pl.o0);
        L.rec_link[2 * wo] = lk; L.rec_link[2 * wo + 1] = lb;
    }
}

// ---- phase 3: follow the links from worker 0 and emit the first T samples in chain order -----------------------------------
__global__ void __launch_bounds__(1024) k_sampler_stitch(Lines L, FrameCtrl * ctrl, int solve, int T)
{
    __shared__ int s_w[kSampWorkersMax], s_b[kSampWorkersMax], s_off[kSampWorkersMax];
    __shared__ int s_n, s_total;
    const int slot = blockIdx.x;
    FrameCtrl & C = ctrl[slot];
    if (C.n_solves != solve || C.num_obs < 2) return;
    const size_t w0 = (size_t)slot * L.samp_workers;
    // one 16-byte record per worker: the serial walk below pays ONE dependent shared-memory load per chain segment
    __shared__ int4 t_node[kSampWorkersMax];             // records made, link worker, link record, continuation
    for (int i = threadIdx.x; i < L.samp_workers; i += blockDim.x)
        t_node[i] = make_int4(L.rec_cnt[w0 + i], L.rec_link[2 * (w0 + i)], L.rec_link[2 * (w0 + i) + 1], (int)L.rec_next[w0 + i]);
    __syncthreads();
    // Fast path: a worker's chain almost always enters the NEXT worker's segment (a segment is ~T / workers samples long and a
    // sample never spans one), so the chain is 0 -> 1 -> 2 -> ... and the offsets are a prefix sum: one block scan instead of a
    // serial walk of up to 512 dependent steps (47 us per solve of a single frame's latency).
    __shared__ int s_end, s_bad, s_wsum[32];
    if (threadIdx.x == 0) { s_end = L.samp_workers - 1; s_bad = 0; }
    __syncthreads();
    const int W = L.samp_workers;
    const int wi = threadIdx.x;
    int4 me = make_int4(0, -1, 0, 0);
    if (wi < W) { me = t_node[wi]; if (me.y < 0) atomicMin(&s_end, wi); }
    __syncthreads();
    const int E = s_end;                                             // the chain 0, 1, 2, ... stops here at the latest
    if (wi < E && me.y != wi + 1) s_bad = 1;
    __syncthreads();
    if (!s_bad) {
        const int bi = (wi > 0 && wi <= E) ? t_node[wi - 1].z : 0;   // where the chain enters this worker's records
        const int ci = (wi <= E) ? max(me.x - bi, 0) : 0;            // samples it contributes
        int inc = ci;
        const int lane = wi & 31, wrp = wi >> 5;
#pragma unroll
        for (int sft = 1; sft < 32; sft <<= 1) { const int v = __shfl_up_sync(0xffffffffu, inc, sft); if (lane >= sft) inc += v; }
        if (lane == 31) s_wsum[wrp] = inc;
        __syncthreads();
        if (wrp == 0) {
            int v = s_wsum[lane];
#pragma unroll
            for (int sft = 1; sft < 32; sft <<= 1) { const int u = __shfl_up_sync(0xffffffffu, v, sft); if (lane >= sft) v += u; }
            s_wsum[lane] = v;
        }
        __syncthreads();
        inc += wrp ? s_wsum[wrp - 1] : 0;
        const int off = inc - ci;                                    // samples before this worker's
        if (wi <= E && off < T) {
            s_w[wi] = wi; s_b[wi] = bi; s_off[wi] = off;
            const bool last_seg = (inc >= T) || (wi == E);           // the chain ends inside / with this worker
            if (last_seg) {
                const int take = min(ci, T - off);
                const int total = off + take;
                unsigned rel;
                if (total >= T) { const int last = bi + take; rel = (last < me.x) ? L.rec_start[(w0 + wi) * L.samp_records + last] : (unsigned)me.w; }
                else rel = (unsigned)me.w;
                s_n = wi + 1; s_total = total;
                C.samp_done = total;
                C.stream_pos = C.stream_pos + rel;
            }
        }
    }
    __syncthreads();
    if (s_bad && threadIdx.x == 0) {
        int cur = 0, b = 0, off = 0, nseg = 0;
        unsigned long long endpos = C.stream_pos;
        while (cur >= 0 && nseg < L.samp_workers) {
            const int4 nd = t_node[cur];
            const int have = nd.x - b;                       // samples worker `cur` contributes from its entry point on
            const int take = min(have, T - off);
            s_w[nseg] = cur; s_b[nseg] = b; s_off[nseg] = off; ++nseg;
            off += take;
            if (off >= T) {
                // start of sample T = start of the record after the last one taken, or the worker's saved continuation
                const int last = b + take;                             // index of the first record NOT taken
                const unsigned rel = (last < nd.x) ? L.rec_start[(w0 + cur) * L.samp_records + last] : (unsigned)nd.w;
                endpos = C.stream_pos + rel; break;
            }
            if (nd.y < 0) { endpos = C.stream_pos + (unsigned)nd.w; break; }
            b = nd.z; cur = nd.y;
        }
        s_n = nseg; s_total = off;
        C.samp_done = off;                                           // < T: the sequential walk below finishes the job
        C.stream_pos = endpos;
    }
    __syncthreads();
    // every output sample finds its chain segment by bisection over the segment offsets (one pass, all loads independent: the
    // per-segment loop it replaces paid one memory round trip per segment, 80 us per solve for a single frame)
    int2 * pairs = reinterpret_cast<int2 *>(L.pairs + (size_t)slot * 2 * T);
    const int nseg = s_n, total = s_total;
    for (int j = threadIdx.x; j < total; j += blockDim.x) {
        int lo_ = 0, hi_ = nseg - 1;
        while (lo_ < hi_) { const int mid = (lo_ + hi_ + 1) >> 1; if (s_off[mid] <= j) lo_ = mid; else hi_ = mid - 1; }
        pairs[j] = reinterpret_cast<const int2 *>(L.rec_pair)[(w0 + s_w[lo_]) * L.samp_records + s_b[lo_] + (j - s_off[lo_])];
    }
}

// ---- sequential completion (and the whole job when called on its own): one warp per frame ----------------------------------
__global__ void __launch_bounds__(32) k_sampler_tail(Lines L, FrameCtrl * ctrl, const unsigned * __restrict__ stream,
                                                     unsigned long long stream_len, int solve, int T)
{
    const int slot = blockIdx.x;
    FrameCtrl & C = ctrl[slot];
    if (C.n_solves != solve || C.num_obs < 2) return;
    int done = C.samp_done;
    if (done >= T) { if (threadIdx.x == 0) C.samp_done = 0; return; }
    const int lane = threadIdx.x;
    int * pairs = L.pairs + (size_t)slot * 2 * T;
    unsigned long long next = C.stream_pos;
    const bool ok = chain_walk(stream, stream_len, (long long)C.n_obs[solve], next, [&](unsigned long long, int s0, int s1, unsigned long long nx) {
        if (lane == 0) { pairs[2 * done] = s0; pairs[2 * done + 1] = s1; }
        ++done; next = nx;
        return done < T;
    });
    if (lane == 0) { C.stream_pos = next; C.samp_done = 0; if (!ok) C.stream_overflow = 1; }
}

} // namespace

void launch_sampler(const Lines & L, FrameCtrl * ctrl, int n_slots, const unsigned * stream, unsigned long long stream_len,
                    int solve, int T, cudaStream_t st)
{
    dim3 g((L.samp_workers + 3) / 4, n_slots);
    LRB_CARVEOUT(k_sampler_segments); LRB_CARVEOUT(k_sampler_link); LRB_CARVEOUT(k_sampler_stitch); LRB_CARVEOUT(k_sampler_tail);
    k_sampler_segments<<<g, 128, 0, st>>>(L, ctrl, stream, stream_len, solve, T);
    k_sampler_link<<<g, 128, 0, st>>>(L, ctrl, stream, stream_len, solve, T);
    k_sampler_stitch<<<n_slots, 1024, 0, st>>>(L, ctrl, solve, T);
    k_sampler_tail<<<n_slots, 32, 0, st>>>(L, ctrl, stream, stream_len, solve, T);
}

// known-answer entry: the plain sequential walk only (slot 0, solve 0)
void launch_sampler_only(const Lines & L, FrameCtrl * ctrl, const unsigned * stream, unsigned long long stream_len, int T, cudaStream_t st)
{
    k_sampler_tail<<<1, 32, 0, st>>>(L, ctrl, stream, stream_len, 0, T);
}

void launch_sampler_parallel_only(const Lines & L, FrameCtrl * ctrl, const unsigned * stream, unsigned long long stream_len, int T, cudaStream_t st)
{
    launch_sampler(L, ctrl, 1, stream, stream_len, 0, T, st);
}

} // namespace lrb
