// engine.cu — host side of the B200 engine: workspace, TMA descriptors, launch sequence, read-back, fall-backs.
//
// One engine owns one CUDA stream and a workspace of `slots` frame slots; every kernel takes the slot as a grid
// dimension, so a launch group pushes `slots` frames through the whole path
//   ingest -> K1 grad5x5_bin -> K2 nms5x5_seeds -> K3 seed order -> K4 flood rounds -> K5 fit -> K6 filter+model
//          -> 4 x (K7 sampler, K8 scoring grid, K9 refit/assign) -> read-back
// with no host round-trip in between: all data-dependent sizes (seeds, components, segments, observations) live in a
// per-slot control block on the device, kernels early-exit on it, and the host looks at it once per group.
// Only when that single look shows unfinished floods (more speculative rounds needed, or an on-chip capacity was hit)
// or an exhausted uniform stream does the host enqueue the missing work and look again.
#include <cuda.h>
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <functional>
#include <vector>

#include "../../include/librectify_b200.h"
#include "common.cuh"
#include "kernels.h"


using namespace lrb;

namespace {

thread_local std::string g_error;
thread_local int g_status = 0;          // return code of the last C-ABI call on this thread (lrb_last_status)
unsigned g_seed = 0;

int fail(int code, const char * fmt, ...)
{
    char buf[512];
    va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
    g_error = buf; g_status = code;
    return code;
}

// every C-ABI entry point starts clean: a stale message of an earlier failure on this thread must not be mistaken for this call's
inline void begin_call() { g_error.clear(); g_status = 0; }

#define CK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail(-2, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); } while (0)

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn()
{
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void * p = nullptr; cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    });
    return fn;
}

// 3-D tensor map (x = column, y = row, z = frame) over a pitched plane set; out-of-range reads give zeros
int make_tmap(CUtensorMap * m, CUtensorMapDataType dt, int esize, const void * base, int w, int h, int frames,
              size_t pitch_elems, size_t frame_stride_elems, int box_w, int box_h)
{
    EncodeTiledFn fn = encode_fn();
    if (!fn) return fail(-3, "cuTensorMapEncodeTiled not available from the driver");
    cuuint64_t dims[3] = {(cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)frames};
    cuuint64_t strides[2] = {(cuuint64_t)pitch_elems * esize, (cuuint64_t)frame_stride_elems * esize};
    if (frames == 1 && (strides[1] == 0 || strides[1] % 16)) strides[1] = ((cuuint64_t)pitch_elems * esize * h + 15) / 16 * 16;
    cuuint32_t box[3] = {(cuuint32_t)box_w, (cuuint32_t)box_h, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = fn(m, dt, 3, const_cast<void *>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(-3, "cuTensorMapEncodeTiled failed (%d) w=%d h=%d frames=%d pitch=%zu stride=%zu", (int)r, w, h, frames, pitch_elems, frame_stride_elems);
    return 0;
}

constexpr int kT = kRansacIter;
constexpr int kLinesHead = 4096;       // segments per slot copied back together with the control block
constexpr int kInitialRounds = 20;
constexpr int kBigSlots = 64;          // concurrent long floods (global-memory tile tables) per engine

} // namespace

struct lrb_engine {
    int device = 0, slots = 1;
    cudaStream_t st = nullptr;
    // host input: the next group's frames are copied on a second stream as soon as K1 of the current group has consumed
    // the image workspace, so the copy overlaps flood / fit / RANSAC of the current group
    cudaStream_t st_copy = nullptr; cudaEvent_t ev_k1_done = nullptr, ev_copy_done = nullptr; bool pre_copied = false;
    void * aux[7] = {}; size_t aux_bytes[7] = {};   // device scratch of the warp / prescale / autorectify entry points (source, destination, tables, float frame, prescaled frame, Hough work space, Hough generator outputs)
    unsigned char * u8_stage = nullptr; size_t u8_stage_bytes = 0;    // device staging of 8-bit host frames (lrb_*_host_u8)
    // workspace geometry
    int w = 0, h = 0, pitch = 0;
    size_t plane = 0;                  // elements per slot plane
    void * arena = nullptr; size_t arena_bytes = 0;
    Planes P{}; Seeds S{}; Comps K{}; Lines L{};
    float * ws_img = nullptr;          // pixel (0,0) of slot 0 of the workspace image (inside the zero margin)
    float * ws_img_raw = nullptr; unsigned * owner_raw = nullptr;
    FrameCtrl * ctrl = nullptr;        // device
    FrameCtrl * h_ctrl = nullptr;      // pinned
    SegRec * h_lines = nullptr;        // pinned, slots * kLinesHead
    int * labels = nullptr;            // debug plane (lazily used)
    CUtensorMap tm_img{}, tm_mag{}, tm_bin{};
    unsigned next_epoch = 0xFE;
    bool refine = false;               // run the `refine` merge post-process (reference interface.cpp:56-59)
    int seed_div = 8;                  // seed capacity = pixels / seed_div; grown to 1 the first time a frame needs it
    // uniform stream
    std::shared_ptr<void> shared_stream;         // keeps the per-device copy of the uniform stream alive (see ensure_stream)
    unsigned * stream = nullptr; unsigned long long stream_len = 0; unsigned seed = 0;
    // bookkeeping of the last group
    int last_n = 0; std::vector<int> last_rounds, last_fallback;
    // profiling
    bool prof = false; cudaEvent_t ev[LRB_T_COUNT + 1] = {}; cudaEvent_t ev_score[2 * kMaxModels] = {}; double ms[LRB_T_COUNT] = {}; long long groups = 0, kernels = 0;
    long long h2d_bytes = 0, d2h_bytes = 0;      // bytes this engine actually moved over PCIe (input frames / read-back), since the last reset
};

namespace {

template <class T> T * carve(char *& p, size_t n)
{
    T * r = reinterpret_cast<T *>(p);
    p += (n * sizeof(T) + 255) / 256 * 256;
    return r;
}

// The sampler's uniform stream depends on the seed only, so the engines of one device share ONE copy (128 MB at the default
// length, ~28 ms to generate): a registry of weak references per (device, seed); an engine that needs a longer stream
// generates a new one and the registry hands that to everybody who asks later.
struct SharedStream {
    unsigned * data = nullptr; unsigned long long len = 0; unsigned seed = 0; int device = 0;
    ~SharedStream() { if (data) { int cur = 0; cudaGetDevice(&cur); cudaSetDevice(device); cudaFree(data); cudaSetDevice(cur); } }
};
std::mutex g_stream_mu;
std::vector<std::weak_ptr<SharedStream>> g_streams;

int ensure_stream(lrb_engine * e, unsigned long long need)
{
    if (e->stream && e->stream_len >= need && e->seed == g_seed) return 0;
    std::lock_guard<std::mutex> lk(g_stream_mu);
    std::shared_ptr<SharedStream> found;
    for (auto it = g_streams.begin(); it != g_streams.end();) {
        std::shared_ptr<SharedStream> s = it->lock();
        if (!s) { it = g_streams.erase(it); continue; }
        if (s->device == e->device && s->seed == g_seed && s->len >= need && (!found || s->len > found->len)) found = s;
        ++it;
    }
    if (!found) {
        found = std::make_shared<SharedStream>();
        found->device = e->device; found->seed = g_seed; found->len = need;
        CK(cudaMalloc(&found->data, need * sizeof(unsigned)));
        launch_mt_stream(found->seed, found->data, need, e->st);
        CK(cudaGetLastError());
        CK(cudaStreamSynchronize(e->st));          // complete before any other engine's stream may read it
        e->kernels++;
        g_streams.push_back(found);
    }
    if (e->stream) CK(cudaStreamSynchronize(e->st));   // kernels still reading the old copy
    e->shared_stream = found;
    e->stream = found->data; e->stream_len = found->len; e->seed = found->seed;
    return 0;
}

// (re)allocate the workspace for frames of w x h
int ensure_workspace(lrb_engine * e, int w, int h)
{
    if (e->arena && e->w == w && e->h == h && e->S.cap >= (int)std::min<size_t>((size_t)w * h / e->seed_div + 1024, (1u << 24) - 1)) return 0;
    if (w < 5 || h < 5 || w > 65535 || h > 65535) return fail(-1, "unsupported frame size %dx%d", w, h);
    if (e->arena) { CK(cudaStreamSynchronize(e->st)); CK(cudaFree(e->arena)); e->arena = nullptr; }
    const int G = e->slots;
    const int pitch = kMarginX + (w + 15) / 16 * 16;
    const size_t plane = (size_t)pitch * (h + kMarginY);
    const size_t org = (size_t)kMarginY * pitch + kMarginX;
    const size_t npx = (size_t)w * h;
    const int cap_seeds = (int)std::min<size_t>(npx / e->seed_div + 1024, (1u << 24) - 1);
    const int cap_comp = (int)(npx / 6 + 16);
    const size_t cap_px = npx;
    const size_t cap_pool = (3 * npx + 65536 + (size_t)148 * 24 * 256) / 32 * 32;
    const int cap_lines = cap_comp;
    const int big_cap = (int)std::min<size_t>(((size_t)(w + 31) / 32) * ((size_t)(h + 31) / 32), 4096);   // tiles per long flood
    const size_t big_bytes = ((size_t)big_cap * (256 + 12) + 255) / 256 * 256;
    auto layout = [&](char * base) {
        char * p = base;
        e->ws_img_raw = carve<float>(p, plane * G); e->ws_img = e->ws_img_raw + org;
        e->P.dxy = carve<float2>(p, plane * G) + org; e->P.mag = carve<float>(p, plane * G) + org;
        e->P.bin = carve<uint8_t>(p, plane * G) + org; e->P.mv = carve<unsigned short>(p, plane * G) + org;
        e->owner_raw = carve<unsigned>(p, plane * G); e->P.owner = e->owner_raw + org;
        e->S.us_mag = carve<unsigned>(p, (size_t)cap_seeds * G); e->S.us_idx = carve<unsigned>(p, (size_t)cap_seeds * G);
        e->S.b_mag = carve<unsigned>(p, (size_t)cap_seeds * G); e->S.b_idx = carve<unsigned>(p, (size_t)cap_seeds * G);
        e->S.idx = carve<unsigned>(p, (size_t)cap_seeds * G); e->S.thr = carve<float>(p, (size_t)cap_seeds * G);
        e->S.bin = carve<uint8_t>(p, (size_t)cap_seeds * G); e->S.status = carve<unsigned>(p, (size_t)cap_seeds * G);
        e->S.fail_epoch = carve<unsigned>(p, (size_t)cap_seeds * G); e->S.size = carve<int>(p, (size_t)cap_seeds * G);
        e->S.list_head = carve<int>(p, (size_t)cap_seeds * G); e->S.esc = carve<int>(p, (size_t)cap_seeds * G); e->S.esc2 = carve<int>(p, (size_t)cap_seeds * G);
        e->S.hist = carve<int>(p, (size_t)kSeedBuckets * G); e->S.start = carve<int>(p, (size_t)kSeedBuckets * G);
        e->S.cursor = carve<int>(p, (size_t)kSeedBuckets * G);
        e->S.rowhist = carve<int>(p, (size_t)h * G); e->S.rowstart = carve<int>(p, (size_t)h * G); e->S.rowcursor = carve<int>(p, (size_t)h * G);
        e->S.rb_col = carve<int>(p, (size_t)cap_seeds * G);
        e->S.rp_el = carve<uint2>(p, (size_t)cap_seeds * G); e->S.rp_q = carve<int4>(p, (size_t)cap_seeds * G);
        e->K.rank = carve<int>(p, (size_t)cap_comp * G); e->K.off = carve<int>(p, (size_t)cap_comp * G);
        e->K.size = carve<int>(p, (size_t)cap_comp * G);
        e->K.px = carve<unsigned>(p, cap_px * G); e->K.val = carve<float>(p, cap_px * G);
        e->K.seg = carve<float>(p, (size_t)cap_comp * 7 * G);
        e->K.pool = carve<unsigned>(p, cap_pool * G);
        e->K.big = carve<unsigned>(p, big_bytes / 4 * kBigSlots); e->K.big_lock = carve<int>(p, kBigSlots);
        e->L.seg = carve<float>(p, (size_t)cap_lines * 7 * G); e->L.h = carve<float>(p, (size_t)cap_lines * 3 * G);
        e->L.anchor = carve<float>(p, (size_t)cap_lines * 2 * G); e->L.dir = carve<float>(p, (size_t)cap_lines * 2 * G);
        e->L.len = carve<float>(p, (size_t)cap_lines * G);
        e->L.inlier = carve<int>(p, (size_t)cap_lines * G); e->L.garbage = carve<int>(p, (size_t)cap_lines * G);
        e->L.obs = carve<int>(p, (size_t)cap_lines * G); e->L.tmp = carve<int>(p, (size_t)cap_lines * G);
        e->L.pairs = carve<int>(p, (size_t)2 * kT * G); e->L.scores = carve<float>(p, (size_t)kMaxModels * kT * G);
        const size_t nw = (size_t)kSampWorkersMax * G, nr = (size_t)kSampWorkers * kSampRecords * G;
        e->L.samp_workers = kSampWorkers; e->L.samp_records = kSampRecords;
        e->L.rec_start = carve<unsigned>(p, nr); e->L.rec_pair = carve<int>(p, 2 * nr);
        e->L.rec_cnt1 = carve<int>(p, nw); e->L.rec_cnt = carve<int>(p, nw); e->L.rec_next = carve<unsigned>(p, nw);
        e->L.rec_link = carve<int>(p, 2 * nw);
        e->labels = carve<int>(p, npx);
        e->ctrl = carve<FrameCtrl>(p, G);
        return (size_t)(p - base);
    };
    const size_t bytes = layout(nullptr);
    CK(cudaMalloc(&e->arena, bytes));
    e->arena_bytes = bytes;
    layout(static_cast<char *>(e->arena));
    e->w = w; e->h = h; e->pitch = pitch; e->plane = plane;
    e->P.w = w; e->P.h = h; e->P.pitch = pitch; e->P.stride = plane; e->P.wq = (w + 3) / 4 * 4;
    e->P.img = e->ws_img; e->P.img_stride = plane; e->P.img_pitch = pitch;
    e->S.win_min = kFloodWindowMin; e->S.win_div = kFloodWindowDiv;
    if (const char * v = getenv("LRB_WINDOW_MIN")) e->S.win_min = std::max(1, atoi(v));
    if (const char * v = getenv("LRB_WINDOW_DIV")) e->S.win_div = std::max(1, atoi(v));
    e->S.win_first = kFloodWindowFirst;
    e->S.win_late = kFloodWindowLate;
    if (const char * v = getenv("LRB_WINDOW_LATE")) e->S.win_late = std::max(1, atoi(v));
    if (const char * v = getenv("LRB_WINDOW_FIRST")) e->S.win_first = std::max(1, atoi(v));
    e->S.rows = h; e->S.roww = w;
    e->S.cap = cap_seeds; e->K.cap = cap_comp; e->K.cap_px = (int)std::min<size_t>(cap_px, 0x7fffffff); e->K.cap_pool = (int)std::min<size_t>(cap_pool, 0x7fffffe0); e->L.cap = cap_lines;
    e->K.big_slots = kBigSlots; e->K.big_cap = big_cap; e->K.big_bytes = big_bytes;
    CK(cudaMemsetAsync(e->arena, 0, bytes, e->st));
    launch_fill_u32(e->owner_raw, plane * G, kOwnerInit, e->st);
    e->next_epoch = 0xFE;
    int mw, mh, bw, bh; nms_tile_dims(&mw, &mh, &bw, &bh);
    const int tw = kMarginX + w, th = kMarginY + h;      // tensor extents: margin + image; everything beyond reads as zero
    if (int r = make_tmap(&e->tm_img, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, e->ws_img_raw, tw, th, G, pitch, plane, grad_tile_w(), grad_tile_h())) return r;
    if (int r = make_tmap(&e->tm_mag, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, e->P.mag - org, tw, th, G, pitch, plane, mw, mh)) return r;
    if (int r = make_tmap(&e->tm_bin, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, e->P.bin - org, tw, th, G, pitch, plane, bw, bh)) return r;
    if (!e->h_ctrl) { CK(cudaHostAlloc(&e->h_ctrl, sizeof(FrameCtrl) * G, cudaHostAllocDefault)); }
    if (!e->h_lines) { CK(cudaHostAlloc(&e->h_lines, sizeof(SegRec) * kLinesHead * G, cudaHostAllocDefault)); }
    CK(cudaGetLastError());
    return 0;
}

} // namespace

// ---- one launch group ---------------------------------------------------------------------------------------------------
namespace {

struct Tick {
    lrb_engine * e; int cur = -1;
    explicit Tick(lrb_engine * e_) : e(e_) {}
    void at(int stage) { if (e->prof) { cudaEventRecord(e->ev[stage], e->st); cur = stage; } }
    void end() { if (e->prof) cudaEventRecord(e->ev[LRB_T_COUNT], e->st); }
};

// The sampler's shape: a large group wants few workers (less speculative overlap), one or two frames are latency-bound and want
// short chains — four times the workers with a quarter of the records each (same memory, same result).
Lines lines_view(const lrb_engine * e, int n)
{
    Lines L = e->L;
    if (n <= 2) { L.samp_workers = kSampWorkersMax; L.samp_records = kSampWorkers * kSampRecords / kSampWorkersMax; }
    return L;
}

void enqueue_tail(lrb_engine * e, int n, float min_length, int rounds_total, Tick & tk)
{
    tk.at(LRB_T_FIT);
    launch_fit(e->P, e->S, e->K, e->ctrl, n, rounds_total, e->st);
    tk.at(LRB_T_MODEL);
    if (e->refine) e->kernels += launch_refine(e->K, e->ctrl, n, e->st);
    launch_filter_model(e->K, e->L, e->ctrl, n, min_length, rounds_total, e->refine, e->st);
    tk.at(LRB_T_RANSAC);
    launch_ransac(lines_view(e, n), e->ctrl, n, e->stream, e->stream_len, kT, e->st, e->prof ? e->ev_score : nullptr);
    e->kernels += 2 + 1 + 6 * kMaxModels + 1;
}

int readback(lrb_engine * e, int n, Tick & tk)
{
    tk.at(LRB_T_READBACK);
    CK(cudaMemcpyAsync(e->h_ctrl, e->ctrl, sizeof(FrameCtrl) * n, cudaMemcpyDeviceToHost, e->st));
    const int head = std::min(e->L.cap, kLinesHead);
    CK(cudaMemcpy2DAsync(e->h_lines, sizeof(SegRec) * kLinesHead, e->L.seg, sizeof(SegRec) * (size_t)e->L.cap,
                         sizeof(SegRec) * head, n, cudaMemcpyDeviceToHost, e->st));
    tk.end();
    e->d2h_bytes += (long long)(sizeof(FrameCtrl) + sizeof(SegRec) * head) * n;
    CK(cudaStreamSynchronize(e->st));
    return 0;
}

void collect_profile(lrb_engine * e, const std::vector<int> & order)
{
    if (!e->prof) return;
    // order: the stages recorded, in stream order; each lasts until the next recorded event
    for (size_t i = 0; i < order.size(); ++i) {
        cudaEvent_t a = e->ev[order[i]], b = (i + 1 < order.size()) ? e->ev[order[i + 1]] : e->ev[LRB_T_COUNT];
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, a, b) == cudaSuccess) e->ms[order[i]] += ms;
    }
    for (int k = 0; k < kMaxModels; ++k) {           // the scoring grid alone (a sub-interval of the RANSAC stage)
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, e->ev_score[2 * k], e->ev_score[2 * k + 1]) == cudaSuccess) e->ms[LRB_T_SCORE] += ms;
    }
}

// 8-bit input (u8.channels != 0): `frames` is ignored, pitch / fstride are in bytes
struct U8Src { const unsigned char * data = nullptr; int channels = 0; bool on_device = false; };

int run_group(lrb_engine * e, const float * frames, bool on_device, int n, int w, int h, long long pitch,
              long long fstride, float min_length, int n_next, U8Src u8);

int copy_host_u8(lrb_engine * e, const unsigned char * src, int n, int w, int h, int ch, long long pitch, long long fstride, cudaStream_t st)
{
    const size_t row = (size_t)w * ch;
    for (int i = 0; i < n; ++i)
        CK(cudaMemcpy2DAsync(e->u8_stage + (size_t)i * row * h, row, src + (long long)i * fstride, (size_t)pitch, row, h, cudaMemcpyHostToDevice, st));
    e->h2d_bytes += (long long)(row * h) * n;
    return 0;
}

int copy_host_frames(lrb_engine * e, const float * frames, int n, int w, int h, long long pitch, long long fstride, cudaStream_t st)
{
    // dense batch (rows and frames back to back, and the workspace rows need no padding): the whole group is one contiguous
    // block -> ONE plain async copy instead of one 2-D copy per frame (a 32-frame 1080p group is a single 265 MB transfer)
    e->h2d_bytes += (long long)sizeof(float) * w * h * n;
    if (pitch == w && e->pitch == w && fstride == (long long)w * h && e->plane == (size_t)w * h) {
        CK(cudaMemcpyAsync(e->ws_img, frames, sizeof(float) * (size_t)w * h * n, cudaMemcpyHostToDevice, st));
        return 0;
    }
    for (int i = 0; i < n; ++i) {
        const float * base = frames + (long long)i * fstride; long long p = pitch;
        if (p < 0) { base = base + (long long)(h - 1) * p; p = -p; }
        CK(cudaMemcpy2DAsync(e->ws_img + (size_t)i * e->plane, sizeof(float) * e->pitch, base, sizeof(float) * (size_t)p,
                             sizeof(float) * w, h, cudaMemcpyHostToDevice, st));
    }
    return 0;
}

// n_next > 0: the n_next frames that follow this group in the caller's host buffer form the next group (host input only)
int run_group(lrb_engine * e, const float * frames, bool on_device, int n, int w, int h, long long pitch,
              long long fstride, float min_length, int n_next, U8Src u8)
{
    if (n < 1 || n > e->slots) return fail(-1, "group of %d frames does not fit %d slots", n, e->slots);
    CK(cudaSetDevice(e->device));
    if (int r = ensure_workspace(e, w, h)) return r;
    // ~0.67 n draws per sample; start with room for n ~ 3000 over four solves and grow on demand
    if (int r = ensure_stream(e, std::max<unsigned long long>(e->stream_len, 1ull << 25))) return r;
    if (e->next_epoch < (unsigned)kMaxRounds + 4) {
        launch_fill_u32(e->owner_raw, e->plane * e->slots, kOwnerInit, e->st);
        e->next_epoch = 0xFE; e->kernels++;
    }
    const unsigned epoch_first = e->next_epoch;
    Tick tk(e);
    std::vector<int> order;
    static const bool debug_sync = getenv("LRB_DEBUG_SYNC") != nullptr;      // fault isolation: sync after every stage
    static const char * stage_names[] = {"ingest", "grad", "nms", "seeds", "flood", "fit", "model", "ransac", "readback"};
    int prev_stage = -1; cudaError_t stage_err = cudaSuccess;
    auto stage = [&](int s) {
        if (debug_sync && prev_stage >= 0 && stage_err == cudaSuccess) {
            stage_err = cudaStreamSynchronize(e->st);
            if (stage_err != cudaSuccess) fprintf(stderr, "[lrb] stage '%s' failed: %s\n", stage_names[prev_stage], cudaGetErrorString(stage_err));
        }
        prev_stage = s;
        tk.at(s); order.push_back(s);
    };

    // ---- ingest (reference src/image.cpp:11-20: dense copy of a strided, possibly bottom-up buffer) ------------------
    stage(LRB_T_INGEST);
    launch_frame_init(e->S, e->ctrl, n, e->st);
    // (the visited mask needs no memset: K2 writes the whole mask / visited plane of the frame, visited = 0)
    CUtensorMap tm_user; const CUtensorMap * tm = &e->tm_img;
    Planes P1 = e->P; int tma_x = kMarginX, tma_y = kMarginY;      // K1's view: where the input image lives
    if (u8.channels) {
        // 8-bit frames: (host: staged on the device first) -> gray / 256 into the image workspace
        const size_t row = (size_t)w * u8.channels, need = row * h * e->slots;
        const unsigned char * dsrc = u8.data; long long dpitch = pitch, dstride = fstride;
        if (!u8.on_device) {
            if (e->u8_stage_bytes < need) {
                CK(cudaStreamSynchronize(e->st)); CK(cudaStreamSynchronize(e->st_copy));
                if (e->u8_stage) CK(cudaFree(e->u8_stage));
                e->u8_stage = nullptr; e->u8_stage_bytes = 0;
                CK(cudaMalloc(&e->u8_stage, need)); e->u8_stage_bytes = need;
                e->pre_copied = false;
            }
            if (e->pre_copied) { CK(cudaStreamWaitEvent(e->st, e->ev_copy_done, 0)); e->pre_copied = false; }
            else if (int r = copy_host_u8(e, u8.data, n, w, h, u8.channels, pitch, fstride, e->st)) return r;
            dsrc = e->u8_stage; dpitch = (long long)row; dstride = (long long)(row * h);
        }
        launch_ingest_u8(dsrc, dpitch, dstride, u8.channels, e->ws_img, e->pitch, e->plane, w, h, n, e->st); e->kernels++;
        static const bool no_prefetch_u8 = getenv("LRB_NO_PREFETCH") != nullptr;
        if (!u8.on_device && n_next > 0 && !no_prefetch_u8) {             // the staging buffer is free again after the ingest kernel
            CK(cudaEventRecord(e->ev_k1_done, e->st));
            CK(cudaStreamWaitEvent(e->st_copy, e->ev_k1_done, 0));
            if (int r = copy_host_u8(e, u8.data + (long long)n * fstride, n_next, w, h, u8.channels, pitch, fstride, e->st_copy)) return r;
            CK(cudaEventRecord(e->ev_copy_done, e->st_copy));
            e->pre_copied = true;
        }
    } else if (on_device) {
        const bool direct = pitch > 0 && (reinterpret_cast<uintptr_t>(frames) % 16 == 0) && (pitch % 4 == 0) &&
                            (n == 1 || (fstride > 0 && fstride % 4 == 0));
        if (direct) {
            if (int r = make_tmap(&tm_user, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, frames, w, h, n, (size_t)pitch,
                                  (size_t)(n == 1 ? 0 : fstride), grad_tile_w(), grad_tile_h())) return r;
            tm = &tm_user; tma_x = tma_y = 0;
            P1.img = frames; P1.img_pitch = (int)pitch; P1.img_stride = (size_t)(n == 1 ? 0 : fstride);
        } else {
            const float * base = frames; long long p = pitch;
            if (p < 0) { base = frames + (long long)(h - 1) * p; p = -p; }
            launch_ingest(base, p, fstride, e->ws_img, e->pitch, e->plane, w, h, n, e->st); e->kernels++;
        }
    } else if (e->pre_copied) {
        CK(cudaStreamWaitEvent(e->st, e->ev_copy_done, 0));              // copied while the previous group was computing
        e->pre_copied = false;
    } else {
        if (int r = copy_host_frames(e, frames, n, w, h, pitch, fstride, e->st)) return r;
    }
    // ---- detection ---------------------------------------------------------------------------------------------------
    stage(LRB_T_GRAD);
    launch_grad(*tm, P1, e->ctrl, n, tma_x, tma_y, e->st);
    static const bool no_prefetch = getenv("LRB_NO_PREFETCH") != nullptr;
    if (!on_device && !u8.channels && n_next > 0 && !no_prefetch) {       // the image workspace is free again after K1
        CK(cudaEventRecord(e->ev_k1_done, e->st));
        CK(cudaStreamWaitEvent(e->st_copy, e->ev_k1_done, 0));
        if (int r = copy_host_frames(e, frames + (long long)n * fstride, n_next, w, h, pitch, fstride, e->st_copy)) return r;
        CK(cudaEventRecord(e->ev_copy_done, e->st_copy));
        e->pre_copied = true;
    }
    stage(LRB_T_NMS);
    launch_nms(e->tm_mag, e->tm_bin, e->P, e->S, e->ctrl, n, kMarginX, kMarginY, e->st);
    // Window schedule of the flood: a large batch starts with a small window that doubles (fewer redundant floods of the same long
    // line: throughput); one or two frames are latency-bound, every extra round is four dependent launches, so they start at the
    // steady size (three rounds fewer).  The result does not depend on the schedule.
    Seeds Sg = e->S;
    if (n <= 2 && !getenv("LRB_WINDOW_FIRST")) Sg.win_first = 1 << 24;
    stage(LRB_T_SEEDS);
    launch_seed_order(e->P, Sg, e->ctrl, n, e->S.cap, e->st);
    stage(LRB_T_FLOOD);
    // LRB_FORCE_SERIAL (tests): one speculative round, the rest of the seeds by the verbatim serial fallback
    const bool force_serial = getenv("LRB_FORCE_SERIAL") != nullptr;
    int rounds = force_serial ? 1 : (n <= 2) ? 12 : kInitialRounds;       // one or two frames at the steady window need 9-10 rounds; every blind empty round is four launches of latency
    e->kernels += launch_flood_rounds(e->P, Sg, e->K, e->ctrl, n, e->S.cap, 0, rounds, epoch_first, e->st);
    e->kernels += 2 + 1 + 3;
    if (debug_sync) {
        stage(LRB_T_FIT); launch_fit(e->P, e->S, e->K, e->ctrl, n, rounds, e->st);
        stage(LRB_T_MODEL); if (e->refine) e->kernels += launch_refine(e->K, e->ctrl, n, e->st);
        launch_filter_model(e->K, e->L, e->ctrl, n, min_length, rounds, e->refine, e->st);
        stage(LRB_T_RANSAC); launch_ransac(lines_view(e, n), e->ctrl, n, e->stream, e->stream_len, kT, e->st, e->prof ? e->ev_score : nullptr);
        stage(LRB_T_READBACK);
        order.resize(order.size() - 4);
        if (stage_err != cudaSuccess) return fail(-2, "debug sync: a stage failed (see stderr): %s", cudaGetErrorString(stage_err));
    }
    order.push_back(LRB_T_FIT); order.push_back(LRB_T_MODEL); order.push_back(LRB_T_RANSAC); order.push_back(LRB_T_READBACK);
    if (!debug_sync) enqueue_tail(e, n, min_length, rounds, tk);
    if (int r = readback(e, n, tk)) return r;
    CK(cudaGetLastError());
    collect_profile(e, order);
    e->groups++;

    // ---- the single look at the control blocks; enqueue what is missing (rare) ---------------------------------------
    e->last_fallback.assign(n, 0);
    bool prof_saved = e->prof; e->prof = false;          // repair passes are not part of the per-stage profile
    Tick quiet(e);
    for (int guard = 0; guard < 64; ++guard) {
        bool more_rounds = false, serial = false, more_stream = false;
        for (int i = 0; i < n; ++i) {
            const FrameCtrl & C = e->h_ctrl[i];
            if (C.seed_overflow == 2) { e->prof = prof_saved; return fail(-4, "frame %d: more than %d gradient peaks of (nearly) equal magnitude", i, 16384); }
            if (C.seed_overflow) {
                e->prof = prof_saved;
                if (e->seed_div > 1) {                    // plateau-like input: every pixel can be a peak; enlarge the seed arrays once
                    e->seed_div = 1;
                    if (e->pre_copied) { CK(cudaStreamSynchronize(e->st_copy)); e->pre_copied = false; }   // the workspace is rebuilt: copy again
                    return run_group(e, frames, on_device, n, w, h, pitch, fstride, min_length, n_next, u8);
                }
                return fail(-4, "frame %d: more than %d gradient peaks (seed capacity)", i, e->S.cap);
            }
            if (C.comp_overflow || C.line_overflow) { e->prof = prof_saved; return fail(-4, "frame %d: component/segment capacity exceeded", i); }
            if (e->refine && C.refine_overflow) { e->prof = prof_saved; return fail(-4, "frame %d: refine found more collinear segment pairs than the edge list holds", i); }
            if (C.flood_overflow) serial = true;
            else if (!C.flood_done && C.live[rounds] != 0) more_rounds = true;
            if (C.stream_overflow) more_stream = true;
        }
        if (!more_rounds && !serial && !more_stream) break;
        if (more_rounds && !serial) {
            if (force_serial) serial = true;
            else if (rounds + 16 <= kMaxRounds) {
                e->kernels += launch_flood_rounds(e->P, Sg, e->K, e->ctrl, n, e->S.cap, rounds, rounds + 16, epoch_first, e->st);
                rounds += 16;
            } else serial = true;
        }
        if (serial) {
            launch_flood_serial(e->P, e->S, e->K, e->ctrl, n, epoch_first - (unsigned)kMaxRounds - 1, e->st);
            e->kernels++;
            for (int i = 0; i < n; ++i) if (e->h_ctrl[i].flood_overflow || e->h_ctrl[i].live[rounds]) e->last_fallback[i] = 1;
        }
        if (more_stream) { if (int r = ensure_stream(e, e->stream_len * 2)) { e->prof = prof_saved; return r; } }
        enqueue_tail(e, n, min_length, rounds, quiet);
        if (int r = readback(e, n, quiet)) { e->prof = prof_saved; return r; }
        CK(cudaGetLastError());
    }
    if (getenv("LRB_DEBUG_STATS")) for (int i = 0; i < n; ++i) { const FrameCtrl & C = e->h_ctrl[i];
        fprintf(stderr, "[lrb] slot %d replay: tied %d steps %d, cycles phase1 %lld phase2 %lld fixup %lld, small ranges %lld\n", i, C.n_tied, C.replay_steps, C.dbg_replay[0], C.dbg_replay[1], C.dbg_replay[2], C.dbg_replay[3]);
        { fprintf(stderr, "[lrb] slot %d slowest full-tile flood per round (us/px/tiles, by epoch):", i); for (int r = 0; r < 24; ++r) if (C.dbg_rmax[r]) fprintf(stderr, " %d:%.0f/%d/%d", r, (C.dbg_rmax[r] >> 24) * 1e-3, (int)(C.dbg_rmax[r] & 0xFFFFFF), C.dbg_rtiles[r]); fprintf(stderr, "\n"); }
        fprintf(stderr, "[lrb] slot %d: seeds %d rounds %d floods %d big %d tiles %d px %d maxpx %d committed px %d comps %d slowest flood %.1f us (%d px; visits %d iters %d evals %d tiles %d) mean %.2f us\n", i, C.n_seeds, C.rounds_run, C.dbg_floods, C.dbg_big, C.dbg_tiles, C.dbg_px, C.dbg_maxpx, C.dbg_commit_px, C.n_comp, (C.dbg_maxdur >> 24) * 1e-3, (int)(C.dbg_maxdur & 0xFFFFFF), C.dbg_slow[0], C.dbg_slow[1], C.dbg_slow[2], C.dbg_slow[3], C.dbg_sumdur * 1e-3 / std::max(1, C.dbg_floods)); }
    e->prof = prof_saved;
    e->next_epoch = epoch_first - (unsigned)rounds - 2;
    if (std::any_of(e->last_fallback.begin(), e->last_fallback.end(), [](int v) { return v != 0; })) e->next_epoch = 0;   // force a re-init
    e->last_n = n;
    e->last_rounds.assign(n, 0);
    for (int i = 0; i < n; ++i) e->last_rounds[i] = e->h_ctrl[i].rounds_run;
    return 0;
}

void fill_info(const lrb_engine * e, int i, lrb_frame_info * out)
{
    const FrameCtrl & C = e->h_ctrl[i];
    out->n_lines = C.n_lines; out->n_seeds = C.n_seeds; out->n_components = C.n_comp; out->n_pixels = C.n_px;
    out->flood_rounds = C.rounds_run; out->flood_fallback = e->last_fallback[i]; out->n_solves = C.n_solves;
    for (int k = 0; k < 4; ++k) out->n_obs[k] = k < C.n_solves ? C.n_obs[k] : 0;
    unsigned g = C.gmax_bits; float f; memcpy(&f, &g, 4); out->max_magnitude = f;
    out->floods_small = C.floods_small; out->floods_mid = C.floods_mid; out->floods_escalated = C.floods_esc;
    out->n_tied = C.n_tied; out->replay_steps = C.replay_steps;
}

int detect(lrb_engine * e, const float * frames, bool on_device, int n_frames, int w, int h, long long pitch,
           long long fstride, float min_length, lrb_frame_info * info, lrb_segment * lines, int lines_cap,
           const lrb_config * cfg, int n_cfg, lrb_transform * tf, U8Src u8 = U8Src())
{
    begin_call();
    if (!e) return fail(-1, "null engine");
    if (n_frames < 0 || (!frames && !u8.channels)) return fail(-1, "bad arguments");
    if (u8.channels && (!u8.data || (u8.channels != 1 && u8.channels != 3) || pitch < (long long)w * u8.channels))
        return fail(-1, "8-bit input: 1 (gray) or 3 (BGR) channels, pitch in bytes >= width * channels");
    // float input: |pitch| >= width; a negative pitch means every frame pointer addresses its LAST row (reference librectify.h:15-18)
    if (!u8.channels && (pitch < 0 ? -pitch : pitch) < (long long)w) return fail(-1, "pitch (%lld elements) smaller than a row of %d pixels", pitch, w);
    std::vector<lrb_segment> tmp;
    for (int f0 = 0; f0 < n_frames; f0 += e->slots) {
        const int n = std::min(e->slots, n_frames - f0);
        const int n_next = std::min(e->slots, n_frames - f0 - n);
        U8Src g8 = u8; if (u8.channels) g8.data = u8.data + (long long)f0 * fstride;
        if (int r = run_group(e, u8.channels ? nullptr : frames + (long long)f0 * fstride, on_device, n, w, h, pitch, fstride, min_length, n_next, g8)) {
            if (e->pre_copied) { cudaStreamSynchronize(e->st_copy); e->pre_copied = false; }
            return r;
        }
        for (int i = 0; i < n; ++i) {
            const int nl = e->h_ctrl[i].n_lines;
            if (info) fill_info(e, i, &info[f0 + i]);
            const lrb_segment * src = reinterpret_cast<const lrb_segment *>(e->h_lines) + (size_t)i * kLinesHead;
            const int want = tf ? nl : std::min(nl, lines ? lines_cap : 0);
            if (want > kLinesHead) {                       // long tail: fetch the rest of this frame's segments
                tmp.resize(want);
                CK(cudaMemcpy(tmp.data(), e->L.seg + (size_t)i * e->L.cap * 7, sizeof(lrb_segment) * want, cudaMemcpyDeviceToHost));
                src = tmp.data();
            }
            if (lines && lines_cap > 0) memcpy(lines + (size_t)(f0 + i) * lines_cap, src, sizeof(lrb_segment) * std::min(nl, lines_cap));
            if (tf) lrb_compute_transform(src, nl, w, h, &cfg[(f0 + i) % n_cfg], &tf[f0 + i]);
        }
    }
    return 0;
}

} // namespace

#include "engine_api.inc"
#include "pool_api.inc"
