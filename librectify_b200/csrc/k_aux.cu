// k_aux.cu — the reference's two auxiliary API functions on the device, for large line sets.
//
//   fit_vanishing_point   reference src/interface.cpp:211-215 -> src/transform.cpp:24-47 (fit_single_vanishing_points)
//   assign_to_group       reference src/interface.cpp:218-265 (fit_vanishing_points per group, src/transform.cpp:52-81, then the
//                         inclination of every new line towards every group's point, src/geometry.cpp:214-229)
//
// Same arithmetic as the host versions in transform_host.cpp (every operation rounded separately, sums in index order), so the
// results are bit-identical to them: the model (bounding box, normalised homogeneous lines, lengths) is built one thread per
// line, the 3x3 scatter of a group is six sequential running sums (one lane each, members in index order), the eigenvector comes
// from the shared Eigen-convention solver, and the assignment is one thread per new line looping over the groups in ascending id.
#include <cstdio>
#include <mutex>

#include "common.cuh"
#include "eig.cuh"
#include "kernels.h"

namespace lrb {

namespace {

__device__ __forceinline__ float aux_sum3(float a, float b, float c) { return __fadd_rn(a, __fadd_rn(b, c)); }

// bounding box of all endpoints (min / max are exact in any order): out4 = {minx, miny, maxx, maxy} as ordered-uint atomics
__device__ __forceinline__ unsigned f2ord(float f) { const unsigned u = __float_as_uint(f); return (u & 0x80000000u) ? ~u : (u | 0x80000000u); }
__device__ __forceinline__ float ord2f(unsigned o) { return __uint_as_float((o & 0x80000000u) ? (o & 0x7FFFFFFFu) : ~o); }

__global__ void __launch_bounds__(256) k_aux_bbox(const SegRec * seg, int n, unsigned * box)
{
    float mnx = INFINITY, mny = INFINITY, mxx = -INFINITY, mxy = -INFINITY;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const SegRec s = seg[i];
        mnx = fminf(mnx, fminf(s.x1, s.x2)); mxx = fmaxf(mxx, fmaxf(s.x1, s.x2));
        mny = fminf(mny, fminf(s.y1, s.y2)); mxy = fmaxf(mxy, fmaxf(s.y1, s.y2));
    }
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) {
        mnx = fminf(mnx, __shfl_xor_sync(0xffffffffu, mnx, s)); mny = fminf(mny, __shfl_xor_sync(0xffffffffu, mny, s));
        mxx = fmaxf(mxx, __shfl_xor_sync(0xffffffffu, mxx, s)); mxy = fmaxf(mxy, __shfl_xor_sync(0xffffffffu, mxy, s));
    }
    if ((threadIdx.x & 31) == 0) {
        atomicMin(&box[0], f2ord(mnx)); atomicMin(&box[1], f2ord(mny));
        atomicMax(&box[2], f2ord(mxx)); atomicMax(&box[3], f2ord(mxy));
    }
}

// LinePencilModel of the normalised lines (src/line_pencil.cpp:25-32, 148-158): h (3 per line) and len
__global__ void __launch_bounds__(256) k_aux_model(const SegRec * seg, int n, const unsigned * box, float * h, float * len, float * norm3)
{
    const float x0 = ord2f(box[0]), y0 = ord2f(box[1]), x1 = ord2f(box[2]), y1 = ord2f(box[3]);
    const float sx = __fsub_rn(x1, x0), sy = __fsub_rn(y1, y0);
    const float cx = __fadd_rn(x0, __fmul_rn(0.5f, sx)), cy = __fadd_rn(y0, __fmul_rn(0.5f, sy)), scale = fmaxf(sx, sy);
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) { norm3[0] = cx; norm3[1] = cy; norm3[2] = scale; }
    if (i >= n) return;
    const SegRec s = seg[i];
    const float ax = __fdiv_rn(__fsub_rn(s.x1, cx), scale), ay = __fdiv_rn(__fsub_rn(s.y1, cy), scale);
    const float bx = __fdiv_rn(__fsub_rn(s.x2, cx), scale), by = __fdiv_rn(__fsub_rn(s.y2, cy), scale);
    const float hx = __fsub_rn(__fmul_rn(ay, 1.f), __fmul_rn(1.f, by)), hy = __fsub_rn(__fmul_rn(1.f, bx), __fmul_rn(ax, 1.f));
    const float hz = __fsub_rn(__fmul_rn(ax, by), __fmul_rn(ay, bx));
    const float nrm = __fsqrt_rn(aux_sum3(__fmul_rn(hx, hx), __fmul_rn(hy, hy), __fmul_rn(hz, hz)));
    h[3 * i] = __fdiv_rn(hx, nrm); h[3 * i + 1] = __fdiv_rn(hy, nrm); h[3 * i + 2] = __fdiv_rn(hz, nrm);
    const float dx = __fsub_rn(bx, ax), dy = __fsub_rn(by, ay);
    len[i] = __fsqrt_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)));
}

// One block per requested group: fit_optimal over the group's lines (all lines when the group has no member or `all` is set:
// src/line_pencil.cpp:114-117), smallest eigenvector, back to image coordinates (src/geometry.cpp:232-238, src/transform.cpp:41-45).
// groups[g] = group id; out: 3 floats per group.
__global__ void __launch_bounds__(32) k_aux_common_point(const SegRec * seg, int n, const float * h, const float * len, const float * norm3,
                                                         const int * groups, int all, float * out)
{
    __shared__ float s_cov[9];
    __shared__ int s_members;
    const int gid = groups[blockIdx.x];
    const int t = threadIdx.x;
    {   // does the group have members at all?  (empty index list => all lines)
        int cnt = 0;
        if (!all) for (int i = t; i < n; i += 32) cnt += (seg[i].group_id == gid);
#pragma unroll
        for (int s = 16; s > 0; s >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, s);
        if (t == 0) s_members = cnt;
    }
    __syncwarp();
    const bool use_all = all || s_members == 0;
    if (t < 6) {
        const int r = (t == 0) ? 0 : (t < 3 ? 1 : 2);
        const int c = (t == 0) ? 0 : (t == 1 ? 0 : (t == 2 ? 1 : t - 3));
        float acc = 0.f; bool first = true;
        for (int q = 0; q < n; ++q) {
            if (!use_all && seg[q].group_id != gid) continue;
            const float v = __fmul_rn(__fmul_rn(h[3 * q + r], len[q]), h[3 * q + c]);
            acc = first ? v : __fadd_rn(acc, v);
            first = false;
        }
        s_cov[r * 3 + c] = acc; s_cov[c * 3 + r] = acc;
    }
    __syncwarp();
    if (t == 0) {
        float eval[3], evec[9];
        sym_eig<3>(s_cov, eval, evec);
        int k = 0;
        for (int i = 1; i < 3; ++i) if (eval[i] < eval[k]) k = i;
        const float px = evec[3 * k], py = evec[3 * k + 1], pz = evec[3 * k + 2];
        float qx, qy, qz;
        if (fabsf(pz) < kEps) { qx = px; qy = py; qz = 0.f; }
        else { qx = __fdiv_rn(px, pz); qy = __fdiv_rn(py, pz); qz = 1.f; }
        if (qz > 0.f) { qx = __fadd_rn(__fmul_rn(norm3[2], qx), norm3[0]); qy = __fadd_rn(__fmul_rn(norm3[2], qy), norm3[1]); }
        out[3 * blockIdx.x] = qx; out[3 * blockIdx.x + 1] = qy; out[3 * blockIdx.x + 2] = qz;
    }
}

// assign_to_group: one thread per new line, groups in ascending id (std::map order of the reference), first strict maximum wins
__global__ void __launch_bounds__(256) k_aux_assign(SegRec * fresh, int n_fresh, const int * groups, const float * vps, int n_groups, float thr)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_fresh) return;
    const SegRec l = fresh[i];
    const float ax = __fdiv_rn(__fadd_rn(l.x2, l.x1), 2.f), ay = __fdiv_rn(__fadd_rn(l.y2, l.y1), 2.f);
    float dx = __fsub_rn(l.x2, l.x1), dy = __fsub_rn(l.y2, l.y1);
    { const float z = __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)); if (z > 0.f) { const float nr = __fsqrt_rn(z); dx = __fdiv_rn(dx, nr); dy = __fdiv_rn(dy, nr); } }
    float best = 0.f; int best_g = 0;
    for (int g = 0; g < n_groups; ++g) {
        const float p0 = vps[3 * g], p1 = vps[3 * g + 1], p2 = vps[3 * g + 2];
        float vx, vy;
        if (fabsf(p2) < kEps) { vx = p0; vy = p1; }
        else { vx = __fadd_rn(-ax, __fdiv_rn(p0, p2)); vy = __fadd_rn(-ay, __fdiv_rn(p1, p2)); }
        { const float z = __fadd_rn(__fmul_rn(vx, vx), __fmul_rn(vy, vy)); if (z > 0.f) { const float nr = __fsqrt_rn(z); vx = __fdiv_rn(vx, nr); vy = __fdiv_rn(vy, nr); } }
        const float inc = fabsf(__fadd_rn(__fmul_rn(vx, dx), __fmul_rn(vy, dy)));
        if (inc > best) { best = inc; best_g = groups[g]; }
    }
    if (best > thr) fresh[i].group_id = best_g;
}

} // namespace

void launch_aux_model(const float * seg7, int n, unsigned * box4, float * h, float * len, float * norm3, cudaStream_t st)
{
    const SegRec * seg = reinterpret_cast<const SegRec *>(seg7);
    const unsigned init[4] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0u, 0u};
    cudaMemcpyAsync(box4, init, sizeof init, cudaMemcpyHostToDevice, st);
    const int blocks = (n + 255) / 256;
    k_aux_bbox<<<blocks < 1184 ? blocks : 1184, 256, 0, st>>>(seg, n, box4);
    k_aux_model<<<blocks, 256, 0, st>>>(seg, n, box4, h, len, norm3);
}

void launch_aux_common_points(const float * seg7, int n, const float * h, const float * len, const float * norm3,
                              const int * groups, int n_groups, bool all, float * out, cudaStream_t st)
{
    k_aux_common_point<<<n_groups, 32, 0, st>>>(reinterpret_cast<const SegRec *>(seg7), n, h, len, norm3, groups, all ? 1 : 0, out);
}

void launch_aux_assign(float * fresh7, int n_fresh, const int * groups, const float * vps, int n_groups, float thr, cudaStream_t st)
{
    k_aux_assign<<<(n_fresh + 255) / 256, 256, 0, st>>>(reinterpret_cast<SegRec *>(fresh7), n_fresh, groups, vps, n_groups, thr);
}


// ---- allocation-free variant of fit_vanishing_point (lrb_fit_vanishing_point_noalloc) ------------------------------------------
// No counterpart in the reference: for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
// harnesses that forbid it).  Same kernels as launch_aux_model + launch_aux_common_points; the scratch they need (bounding box,
// normalisation, group id, h and len of the model) lives in ONE file-scope device array that every device loading the library
// holds until exit: kAuxNoAllocScratchBytes = 768 + 16 * kAuxNoAllocMaxLines = 1 MiB + 768 B.  All calls on a device share it
// and it keeps the last call's contents: the bounding box is re-initialised on the stream by every call, h / len / norm3 / group
// id are overwritten for the lines of the call before they are read.  Calls that overlap on the GPU (two streams) race.
namespace {

constexpr size_t kNaBox = 0, kNaNorm = 256, kNaGroup = 512, kNaH = 768;
constexpr size_t kNaLen = kNaH + 12 * (size_t)kAuxNoAllocMaxLines;
constexpr size_t kAuxNoAllocScratchBytes = kNaLen + 4 * (size_t)kAuxNoAllocMaxLines;

static __device__ __align__(256) unsigned char g_aux_noalloc_scratch[kAuxNoAllocScratchBytes];

constexpr int kNaMaxDevices = 64;
unsigned char * g_na_ptr[kNaMaxDevices];      // device address of the array above, per device (the symbol itself is a host-side stand-in)
std::mutex g_na_mutex;

// what must not carry over from the previous call: min / max identities of the bounding box; and the group id of this call
__global__ void k_aux_noalloc_init(unsigned * box, int * groups, int group)
{
    box[0] = 0xFFFFFFFFu; box[1] = 0xFFFFFFFFu; box[2] = 0u; box[3] = 0u;
    groups[0] = group;
}

} // namespace

int aux_fit_vanishing_point_noalloc(int device, const float * d_seg7, int n, int group, float * d_out3, cudaStream_t st,
                                    char * err, size_t err_len)
{
#define NA_CK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { snprintf(err, err_len, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); return -2; } } while (0)
    if (device < 0 || device >= kNaMaxDevices) { snprintf(err, err_len, "device %d out of range", device); return -1; }
    NA_CK(cudaSetDevice(device));
    unsigned char * base;
    {   // set-up, once per device, with that device current
        std::lock_guard<std::mutex> lock(g_na_mutex);
        if (!g_na_ptr[device]) {
            void * p = nullptr;
            NA_CK(cudaGetSymbolAddress(&p, g_aux_noalloc_scratch));
            g_na_ptr[device] = static_cast<unsigned char *>(p);
        }
        base = g_na_ptr[device];
    }
    unsigned * box = reinterpret_cast<unsigned *>(base + kNaBox);
    float * norm3 = reinterpret_cast<float *>(base + kNaNorm);
    int * groups = reinterpret_cast<int *>(base + kNaGroup);
    float * h = reinterpret_cast<float *>(base + kNaH);
    float * len = reinterpret_cast<float *>(base + kNaLen);
    const SegRec * seg = reinterpret_cast<const SegRec *>(d_seg7);
    const int blocks = (n + 255) / 256;
    k_aux_noalloc_init<<<1, 1, 0, st>>>(box, groups, group);
    k_aux_bbox<<<blocks < 1184 ? blocks : 1184, 256, 0, st>>>(seg, n, box);
    k_aux_model<<<blocks, 256, 0, st>>>(seg, n, box, h, len, norm3);
    k_aux_common_point<<<1, 32, 0, st>>>(seg, n, h, len, norm3, groups, group <= 0 ? 1 : 0, d_out3);
    NA_CK(cudaGetLastError());
    return 0;
#undef NA_CK
}

} // namespace lrb
