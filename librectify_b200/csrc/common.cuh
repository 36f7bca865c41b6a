// common.cuh — shared device-side definitions of the librectify hot path on B200 (sm_100a).
//
// Every constant below is carried over from the reference's compile-time configuration
// (reference src/config.h:10-59, src/line_pencil.h:21, src/line_detector.cpp:203) so that the kernels make the same
// decisions as the CPU library.  All translation units that include this header are compiled with -fmad=false:
// the parity contract is bit-exact float32 arithmetic in the order the reference evaluates it, so a multiply and
// the add that follows it must round separately.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <cstdlib>
#include <mutex>

namespace lrb {

// Function attributes (dynamic shared-memory opt-in, carveout) and occupancy answers are PER DEVICE: a launcher that sets
// them once per process leaves the second GPU of a single-process multi-GPU host without the opt-in.  once.run(fn) runs fn
// the first time the calling launcher is used on the current device (thread-safe); dev() is the current device ordinal.
struct PerDeviceOnce {
    static constexpr int kMaxDev = 64;
    std::once_flag flag[kMaxDev];
    static int dev() { int d = 0; cudaGetDevice(&d); return d & (kMaxDev - 1); }
    template <class F> int run(F && fn) { const int d = dev(); std::call_once(flag[d], fn); return d; }
};

// ---- reference configuration (src/config.h) ---------------------------------------------------------------------
constexpr float kEps              = 1e-6f;    // config.h:10
constexpr int   kMaxModels        = 4;        // config.h:23
constexpr int   kRansacIter       = 10000;    // config.h:29
constexpr float kLineMaxErr       = 2.0f;     // config.h:50
constexpr float kLineMinLength    = 5.0f;     // config.h:56
constexpr int   kComponentMinSize = 5;        // config.h:59
constexpr float kDegeneracyTol    = 0.05f;    // line_pencil.h:21
// derived float constants, evaluated as the reference evaluates them (values pinned by tests/golden/kat.npz)
constexpr float kSeedFactor = 0x1.9999ap-5f;  // 1 - 0.95f            (line_detector.cpp:209, config.h:44)
constexpr float kTraceKeep  = 0.75f;          // 1 - 0.25f            (filter.cpp:121, config.h:47)
constexpr float kTolInlier  = 0x1.3f6p-11f;   // 1 - cos(2 deg)       (line_pencil.cpp:143-146, config.h:25)
constexpr float kTolGarbage = 0x1.3f48p-9f;   // 1 - cos(4 deg)       (config.h:27)

// 5x5 Gaussian-derivative taps, sigma = 1 (reference src/filter.cpp:65-78).  Hx(a,b) = sign(b-2) * kTap[min(a,4-a)][min(b,4-b)],
// Hy = Hx^T; the centre column (row) is exactly zero.
#define LRB_K00 0x1.7e13fep-8f
#define LRB_K01 0x1.ac16bap-7f
#define LRB_K10 0x1.ac16bap-6f
#define LRB_K11 0x1.dfa3e4p-5f
#define LRB_K20 0x1.60e648p-5f
#define LRB_K21 0x1.8b6582p-4f

// orientation tables sin/cos(float(i*pi)/8) (reference src/line_detector.cpp:142-146)
static __device__ __constant__ const float kSin[8] = {0.0f, 0x1.87de2cp-2f, 0x1.6a09e6p-1f, 0x1.d906bcp-1f,
                                               1.0f, 0x1.d906bep-1f, 0x1.6a09e6p-1f, 0x1.87de2ep-2f};
static __device__ __constant__ const float kCos[8] = {1.0f, 0x1.d906bcp-1f, 0x1.6a09e6p-1f, 0x1.87de2ap-2f,
                                               -0x1.777a5cp-25f, -0x1.87de28p-2f, -0x1.6a09e6p-1f, -0x1.d906bcp-1f};

// |dx*sin_b + dy*cos_b| exactly as gradient_directions evaluates it (two rounded products, one rounded sum)
__device__ __forceinline__ float orient_energy(float dx, float dy, int b)
{
    return fabsf(__fadd_rn(__fmul_rn(dx, kSin[b]), __fmul_rn(dy, kCos[b])));
}

// ---- seed ordering ------------------------------------------------------------------------------------------------
constexpr int kSeedBuckets = 4096;
constexpr int kSampWorkers = 128, kSampRecords = 512;          // speculative sampler: workers x records per worker (throughput shape)
constexpr int kSampWorkersMax = 512;                          // latency shape (one or two frames): 512 workers x 128 records, same memory
constexpr int kFloodWindowMin = 4096, kFloodWindowDiv = 6;   // steady window: max(4096, n_seeds / 6) seeds per round ...
constexpr int kFloodWindowFirst = 256;                       // ... reached by doubling from this size in round 0
constexpr int kFloodWindowLate = 7;                          // ... and doubling again from this round on: the weak tail of the list (late rounds end with the tail of one slow blob each)
constexpr int kFitSmallMax = 32;                             // components up to this size are fitted by eight lanes (k5_fit.cu)

// ---- per-frame control block (device memory, one per slot; zeroed by k_frame_init) -------------------------------
constexpr int kMaxRounds = 96;
struct FrameCtrl {
    unsigned gmax_bits;        // max gradient magnitude (float bits, mag >= 0 so uint order == float order)
    int n_seeds;               // peaks emitted by the NMS kernel
    int n_tied;                // seeds whose magnitude equals another seed's exactly (their order needs the std::sort replay)
    int replay_steps;          // partition steps replayed for this frame (statistics)
    int seed_overflow;         // more peaks than capacity
    int flood_overflow;        // a flood ran out of queue / lost-list space in the speculative rounds
    int n_comp;                // kept components (size > kComponentMinSize)
    int n_px;                  // pixels in kept components
    int comp_overflow;
    int n_lines;               // filtered segments
    int line_overflow;
    int rounds_run;
    int flood_done;            // set by the serial fallback
    int fit_done;              // components selected and fitted on final floods
    int win_lo, win_hi;        // flood window: ranks processed by the current round
    int pool_cursor;           // next free word of the pixel-list pool
    int win_live, win_first, win_blocks;   // commit-kernel scratch (live seeds left in the window, lowest of them, blocks done)
    int live[kMaxRounds + 1];  // live[r]: seeds still live after the commit of round r-1 (live[0] = n_seeds)
    int queue[kMaxRounds + 1]; // queue[r]: next window position handed out in flood round r (warps take seeds dynamically, in priority order)
    int floods_small, floods_mid, floods_esc;   // statistics over all rounds: floods finished in the 8-lane / 16-lane window / handed to the full-tile path
    int esc2_n, esc2_q;        // the same for the second (16-lane) pass -> full-tile path
    int wl_n;                  // live seeds of the current window as listed by the prune pass (the sub-warp pass floods that list)
    int esc_n, esc_q;          // floods of the current round that outgrew the sub-warp window (escalated to the full-tile path), next one handed out
    int fit_queue;             // next entry of the list of large components handed out by the fit kernel (one warp per component)
    int fit_queue_small;       // next kept component handed out to the sub-warp path (eight lanes per component, four per warp)
    int n_big;                 // kept components with more than kFitSmallMax pixels (their indices: Seeds::esc, free after the flood)
    int n_edges, n_merged, refine_overflow;   // `refine` post-process (k_refine.cu): collinearity edges, merged segments, edge list full
    // grouping state
    int num_obs;               // reference estimator.h:119
    int n_solves;
    int n_obs[kMaxModels];
    int winner[kMaxModels];
    unsigned long long best[kMaxModels];   // (score bits << 32) | ~draw index
    float best_h[kMaxModels][3];
    float refined[kMaxModels][3];
    unsigned long long stream_pos;         // position in the shared uniform stream (carried across solves)
    int stream_overflow;
    int samp_done;                         // samples already emitted by the parallel sampler (the sequential walk finishes the rest)
    float bbox[4], cx, cy, scale;
    int dbg_floods, dbg_big, dbg_tiles, dbg_px, dbg_maxpx, dbg_commit_px;   // flood statistics (LRB_DEBUG_STATS)
    unsigned long long dbg_maxdur;   // (ns << 24) | px of the slowest flood
    unsigned long long dbg_sumdur;
    int dbg_slow[4];
    unsigned long long dbg_rmax[24];  // per round: (ns << 24) | px of the slowest full-tile flood (LRB_FLOOD_STATS builds)
    int dbg_rtiles[24];
    long long dbg_replay[4];          // clock64 spans of the tie-order replay: phase 1, phase 2, fix-up (LRB_FLOOD_STATS builds)
};

// ---- views passed to kernels by value -----------------------------------------------------------------------------
struct Planes {                 // one frame slot = one set of planes; slot s lives at base + s * stride
    int w, h, pitch;            // pitch in elements, multiple of 16 (TMA needs 16-byte row strides, also for the byte planes)
    int wq;                     // w rounded up to 4: columns written by the 4-wide vector stores
    // plane pointers address pixel (0,0); an optional zero margin (kMarginX / kMarginY, currently 0) may precede it
    size_t stride;              // elements per slot plane (pitch * h rounded up)
    const float * img;          // input frames (may alias user memory): img + s * img_stride + r * img_pitch + c
    size_t img_stride; int img_pitch;
    float2 * dxy;               // (dx, dy) interleaved: the flood and the fit always want both, and a pixel's pair shares one sector
    float * mag;
    uint8_t * bin;
    // low byte: mask8 (bit b = some 3x3 neighbour has bin b; written by K2, which thereby also clears the high byte for the frame);
    // high byte: the visited mask of the flood (reference filter.cpp `visited`).  One 16-bit plane instead of two byte planes: a
    // flood window costs one sector per row for both (a TALL sub-warp window: 32 rows x 64-byte bursts less).
    unsigned short * mv;
    unsigned * owner;
};

struct Seeds {                  // per slot: s * cap
    int cap;
    int win_min, win_div, win_first;          // seeds flooded per round: win_first in round 0, doubling up to max(win_min, n_seeds / win_div)
    int win_late;                             // ... and doubling again from this round on (the weak tail of the list)
    unsigned * us_mag; unsigned * us_idx;     // unsorted (emission order)
    unsigned * b_mag;  unsigned * b_idx;      // bucketed
    unsigned * idx;                           // sorted by (mag desc, raster idx asc): rank -> pixel
    float * thr;                              // 0.75f * G_bin(seed)
    uint8_t * bin;
    unsigned * status;                        // 0 live, 1 dropped, else commit key
    unsigned * fail_epoch;
    int * size;
    int * list_head;                          // first pool chunk of the seed's tentative pixel list (this round)
    int * esc; int * esc2;                    // ranks of this round's escalated floods: 8-lane pass -> 16-lane pass -> full-tile path (k_flood_small)
    int * hist; int * start; int * cursor;    // kSeedBuckets each per slot
    // raster order of the peaks and the std::sort replay (k3_replay.cu): the reference sorts its peak list, which is in raster
    // order, with std::sort (reference src/filter.cpp:174-188) — unstable, so the order of exact-magnitude ties is whatever
    // introsort's partitions leave behind
    int rows, roww;                           // frame height (per-slot stride of the three row arrays) and width
    int * rowhist; int * rowstart; int * rowcursor;   // peaks per image row -> first raster position of the row's peaks
    int * rb_col;                             // per slot: cap; columns of the peaks, bucketed by row
    uint2 * rp_el;                            // per slot: cap; the peak list as std::sort sees it: x = key (dense rank of the magnitude, ties share it), y = tied << 31 | rank by (magnitude desc, raster asc)
    int4 * rp_q;                              // per slot: cap; work list of the replay (segment first, last, depth left, tied elements inside)
};

struct Comps {                  // per slot
    int cap; int cap_px;
    int * rank;                 // kept component -> seed rank
    int * off;                  // offset into px lists
    int * size;
    unsigned * px;              // pixel raster indices in the reference's BFS order
    float * val;                // G_bin at those pixels
    float * seg;                // raw LineSegment records (7 x 4 bytes each)
    unsigned * pool; int cap_pool;   // per-round pixel lists of the tentative floods (contiguous, one atomicAdd per list)
    // tile tables in global memory for floods that outgrow the shared-memory table (k4_flood.cu); shared by all slots
    unsigned * big; int * big_lock; int big_slots, big_cap; size_t big_bytes;
};

struct Lines {                  // per slot: filtered segments and the pencil model built on them
    int cap;
    float * seg;                // LineSegment records (7 x 4 bytes), group_id filled in by k_finalize
    float * h;                  // 3 per line: unit homogeneous line through the normalised endpoints
    float * anchor;             // 2 per line: midpoint
    float * dir;                // 2 per line: unit direction
    float * len;
    int * inlier; int * garbage;
    int * obs;                  // current observation list (indices into the lines)
    int * tmp;                  // inlier list of the winning hypothesis
    int * pairs;                // 2 x T per slot: sampled positions in the observation list
    float * scores;             // kMaxModels x T per slot (-1 = rejected by sample_check); kept for the parity tests
    // speculative sampler workers (k7_sampler.cu): samp_workers x samp_records records per slot (their product is fixed)
    int samp_workers, samp_records;
    unsigned * rec_start;       // sample start, relative to the solve's first stream position
    int * rec_pair;             // 2 per record: the two picks
    int * rec_cnt1; int * rec_cnt;   // records made inside the own segment / in total
    unsigned * rec_next;        // start of the sample after the last record
    int * rec_link;             // 2 per worker: (worker, record) where the chain continues, worker < 0: none
};

constexpr int kMarginX = 0, kMarginY = 0;    // no margin needed: TMA accepts negative box origins (they only have to be 16-byte aligned in x)

constexpr unsigned kStatusLive = 0u, kStatusDropped = 1u;
constexpr unsigned kOwnerInit = 0xFFFFFFFFu;          // epoch field 0xFF: never visited, loses every atomicMin
__device__ __forceinline__ unsigned owner_epoch(unsigned o) { return o >> 24; }
__device__ __forceinline__ unsigned owner_rank(unsigned o)  { return o & 0x00FFFFFFu; }

struct SegRec { float x1, y1, x2, y2, weight, err; int group_id; };

// Experiment knob: LRB_CARVEOUT = preferred shared-memory split (percent) for every hot kernel of the library; unset: the
// driver's per-kernel choice.  Kernels of different engines share the SMs, and an SM keeps the largest split any resident
// kernel asked for: LRB_CARVEOUT=100 (28 KB of L1 left) costs the flood 17 % (3415 -> 2830 images/s), 30..65 change nothing.
// This is why no kernel of this library may ask for the maximum split (see launch_nms).
inline int set_carveout(const void * kern)
{
    static const int pct = [] { const char * v = getenv("LRB_CARVEOUT"); return v ? atoi(v) : -1; }();
    if (pct >= 0) cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, pct);
    return 0;
}
#define LRB_CARVEOUT(kern) do { static const int once_ = ::lrb::set_carveout(reinterpret_cast<const void *>(kern)); (void)once_; } while (0)

// warp helpers
__device__ __forceinline__ int lane_id() { return threadIdx.x & 31; }

} // namespace lrb
