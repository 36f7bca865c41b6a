// k_refine.cu — the `refine` post-process of find_line_segment_groups.
//
// Replaces postprocess_lines_segments / graph_components / dfs / merge_lines (reference src/line_detector.cpp:254-444),
// called between the segment fit and the length/err filter when `refine` is set (src/interface.cpp:56-59):
//   affinity  for i < j: |d_i . d_j| >= 0.99, and the endpoints of the shorter segment, in the frame of the longer one
//             (origin = its first endpoint, axes = unit direction / normal, unit = its length), have |normal offset| < 0.02
//             and some along-offset > -0.5 and some < 1.5  (:357-401; the 0.99 / 0.02 compares are DOUBLE compares)
//   labels    graph_components follows forward edges only (dfs scans u > n, :293), so
//             label(u) = min{ a : a == u or a -> ... -> u along edges i < j }
//   merge     unique labels ascending, members in index order; a group of k > 1 segments is refitted through
//             fit_line_parameters on its 2k endpoints weighted length*weight, weight = sum(l*w) / sum(l)  (:254-271)
//
// The O(n^2) affinity test (3.7e7 pairs for n = 8600 raw segments) is a tiled all-pairs kernel: one thread owns segment
// i, the j-tile sits in shared memory, the cosine test rejects ~99.9 % of the pairs after four multiplies.  Edges are rare
// and go to a list (one atomic each); labels are propagated over the list to a fixed point by one CTA per frame, which
// then compacts the roots in order and refits the (few) multi-member groups, one thread per group, every sum evaluated
// sequentially in the reference's order.  Scratch memory is the flood's pixel-list pool, idle at this point.
#include "common.cuh"
#include "eig.cuh"
#include "kernels.h"

namespace lrb {

namespace {

constexpr int RF_TILE = 256;

// scratch layout inside the per-slot pool (32-bit words); cap = K.cap
struct RefineScratch {
    float * dx; float * dy; float * len;    // unit direction and length per raw segment
    int * label; int * cnt; int * rrank;    // label, members per root, output position of a root
    int * multi; int * goff; int * fill;    // member lists of the multi-member groups (contiguous per group, ascending), their offsets
    SegRec * out;                           // merged segments (read by k_filter_model)
    int * edges; int edge_cap;              // (i, j) pairs
};

__device__ __host__ __forceinline__ RefineScratch refine_scratch(const Comps & K, int slot)
{
    RefineScratch R;
    unsigned * base = K.pool + (size_t)slot * K.cap_pool;
    const size_t c = (size_t)K.cap;
    R.dx = reinterpret_cast<float *>(base); R.dy = R.dx + c; R.len = R.dy + c;
    R.label = reinterpret_cast<int *>(base + 3 * c); R.cnt = R.label + c; R.rrank = R.cnt + c; R.multi = R.rrank + c;
    R.goff = R.multi + c; R.fill = R.goff + c;
    R.out = reinterpret_cast<SegRec *>(base + 9 * c);
    R.edges = reinterpret_cast<int *>(base + 16 * c);
    const long long left = (long long)K.cap_pool - 16ll * (long long)c;
    R.edge_cap = (int)(left > 0 ? left / 2 : 0);
    return R;
}

__global__ void __launch_bounds__(256) k_refine_prep(Comps K, FrameCtrl * ctrl)
{
    const int slot = blockIdx.y;
    FrameCtrl & C = ctrl[slot];
    const int n = C.n_comp;
    const RefineScratch R = refine_scratch(K, slot);
    const SegRec * raw = reinterpret_cast<const SegRec *>(K.seg) + (size_t)slot * K.cap;
    if (blockIdx.x == 0 && threadIdx.x == 0) { C.n_edges = 0; C.refine_overflow = 0; }
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const SegRec s = raw[i];
        const float x = __fsub_rn(s.x2, s.x1), y = __fsub_rn(s.y2, s.y1);                     // geometry.cpp:76-79
        const float l = __fsqrt_rn(__fadd_rn(__fmul_rn(x, x), __fmul_rn(y, y)));             // rowwise().norm()
        R.dx[i] = __fdiv_rn(x, l); R.dy[i] = __fdiv_rn(y, l); R.len[i] = l;                   // rowwise().normalized(): no zero check
        R.label[i] = i; R.cnt[i] = 0; R.fill[i] = 0;
    }
}

// The reference compares floats against the double literals 0.99 and 0.02 (line_detector.cpp:377,391).  For a float x,
// x < 0.99 (as double) <=> x < kCosLimit with kCosLimit the smallest float >= 0.99 (no float lies in between); same for 0.02.
constexpr float kCosLimit = 0x1.fae148p-1f, kOffsetLimit = 0x1.47ae16p-6f;
static_assert((double)kCosLimit >= 0.99 && (double)0x1.fae146p-1f < 0.99, "kCosLimit is not the float successor of 0.99");
static_assert((double)kOffsetLimit >= 0.02 && (double)0x1.47ae14p-6f < 0.02, "kOffsetLimit is not the float successor of 0.02");

// line_detector.cpp:378-401 for one pair that passed the cosine test; (ix..) = segment i, (jx..) = segment j
__device__ __forceinline__ bool refine_affine(float ix1, float iy1, float ix2, float iy2, float idx, float idy, float il,
                                              float jx1, float jy1, float jx2, float jy2, float jdx, float jdy, float jl)
{
    const bool i_short = il < jl;                                                           // project the shorter one
    const float sx1 = i_short ? ix1 : jx1, sy1 = i_short ? iy1 : jy1, sx2 = i_short ? ix2 : jx2, sy2 = i_short ? iy2 : jy2;
    const float ox = i_short ? jx1 : ix1, oy = i_short ? jy1 : iy1;
    const float dx = i_short ? jdx : idx, dy = i_short ? jdy : idy, len = i_short ? jl : il;
    const float nx = -dy, ny = dx;
    const float ax = __fsub_rn(sx1, ox), ay = __fsub_rn(sy1, oy);
    // cheap conservative reject before the four IEEE divisions: |n01| >= 1.01 * 0.02 * len implies |n01 / len| >= 0.02
    // whatever the roundings (each is below 1e-6 relative); parallel but distant segments — most pairs that pass the
    // cosine test in a Manhattan scene — leave here
    const float n01 = __fadd_rn(__fmul_rn(ax, nx), __fmul_rn(ay, ny));
    if (fabsf(n01) >= __fmul_rn(1.01f * kOffsetLimit, len)) return false;
    const float bx = __fsub_rn(sx2, ox), by = __fsub_rn(sy2, oy);
    const float w00 = __fdiv_rn(__fadd_rn(__fmul_rn(ax, dx), __fmul_rn(ay, dy)), len);
    const float w01 = __fdiv_rn(n01, len);
    const float w10 = __fdiv_rn(__fadd_rn(__fmul_rn(bx, dx), __fmul_rn(by, dy)), len);
    const float w11 = __fdiv_rn(__fadd_rn(__fmul_rn(bx, nx), __fmul_rn(by, ny)), len);
    if (!(fmaxf(fabsf(w01), fabsf(w11)) < kOffsetLimit)) return false;
    return (w00 > -0.5f || w10 > -0.5f) && (w00 < 1.5f || w10 < 1.5f);
}

// all pairs i < j: tile pairs (bi <= bj) are handed out over the grid; thread t of a block owns i = bi*RF_TILE + t
__global__ void __launch_bounds__(RF_TILE) k_refine_edges(Comps K, FrameCtrl * ctrl)
{
    __shared__ float2 sd[RF_TILE];              // unit directions of the j tile (all that ~99.9 % of the pairs ever need)
    __shared__ float sj[5][RF_TILE];
    const int slot = blockIdx.y;
    FrameCtrl & C = ctrl[slot];
    const int n = C.n_comp;
    if (n < 2) return;
    const RefineScratch R = refine_scratch(K, slot);
    const SegRec * raw = reinterpret_cast<const SegRec *>(K.seg) + (size_t)slot * K.cap;
    const int nb = (n + RF_TILE - 1) / RF_TILE;
    const int t = threadIdx.x;
    for (int tp = blockIdx.x; tp < nb * nb; tp += gridDim.x) {
        const int bi = tp / nb, bj = tp - bi * nb;
        if (bj < bi) continue;                                                              // block-uniform
        const int i = bi * RF_TILE + t, j0 = bj * RF_TILE;
        __syncthreads();
        {
            const int j = j0 + t;
            if (j < n) {
                const SegRec s = raw[j];
                sj[0][t] = s.x1; sj[1][t] = s.y1; sj[2][t] = s.x2; sj[3][t] = s.y2; sj[4][t] = R.len[j];
                sd[t] = make_float2(R.dx[j], R.dy[j]);
            }
        }
        __syncthreads();
        if (i >= n) continue;
        const SegRec si = raw[i];
        const float idx = R.dx[i], idy = R.dy[i], il = R.len[i];
        const int jn = min(RF_TILE, n - j0);
        for (int k = (bi == bj) ? t + 1 : 0; k < jn; ++k) {
            const float2 dj = sd[k];
            const float dot = __fadd_rn(__fmul_rn(idx, dj.x), __fmul_rn(idy, dj.y));        // (U^T V)(0,0), line_detector.cpp:377
            if (fabsf(dot) < kCosLimit) continue;
            if (refine_affine(si.x1, si.y1, si.x2, si.y2, idx, idy, il,
                              sj[0][k], sj[1][k], sj[2][k], sj[3][k], dj.x, dj.y, sj[4][k])) {
                const int e = atomicAdd(&C.n_edges, 1);
                if (e < R.edge_cap) { R.edges[2 * e] = i; R.edges[2 * e + 1] = j0 + k; }
                else C.refine_overflow = 1;
            }
        }
    }
}

// ordered compaction helper: position of the flagged threads of a 1024-thread block (exclusive), total in `total`
__device__ __forceinline__ int scan_flag_1024(int flag, int * wsum, int & total)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned m = __ballot_sync(0xffffffffu, flag);
    const int inw = __popc(m & ((1u << lane) - 1));
    if (lane == 0) wsum[warp] = __popc(m);
    __syncthreads();
    if (warp == 0) {
        int v = wsum[lane];
#pragma unroll
        for (int s = 1; s < 32; s <<= 1) { const int u = __shfl_up_sync(0xffffffffu, v, s); if (lane >= s) v += u; }
        wsum[lane] = v;
    }
    __syncthreads();
    const int pos = inw + (warp ? wsum[warp - 1] : 0);
    total = wsum[31];
    __syncthreads();
    return pos;
}

// exclusive block scan of a value over 1024 threads, total in `total`
__device__ __forceinline__ int scan_val_1024(int v, int * wsum, int & total)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int inc = v;
#pragma unroll
    for (int s = 1; s < 32; s <<= 1) { const int u = __shfl_up_sync(0xffffffffu, inc, s); if (lane >= s) inc += u; }
    if (lane == 31) wsum[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        int w = wsum[lane];
#pragma unroll
        for (int s = 1; s < 32; s <<= 1) { const int u = __shfl_up_sync(0xffffffffu, w, s); if (lane >= s) w += u; }
        wsum[lane] = w;
    }
    __syncthreads();
    const int pos = inc - v + (warp ? wsum[warp - 1] : 0);
    total = wsum[31];
    __syncthreads();
    return pos;
}

// merge_lines (line_detector.cpp:254-271) for root r: members = mem[0..km), in index order.
// fit_line_parameters (geometry.cpp:20-61) on the 2k endpoints (row, col) = (y, x) with weights length*weight, every
// sum sequential in member order — the same arithmetic as k5_fit.cu, one thread per group (groups are few and small).
__device__ void merge_group(const SegRec * raw, const RefineScratch & R, const int * mem, int km, SegRec & out)
{
    // pass 1: weights and their sums
    float wsum = 0.f, ws = 0.f, ls = 0.f; int k = 0;
    for (int q = 0; q < km; ++q) {
        const int u = mem[q];
        const float l = R.len[u];
        const float w = __fmul_rn(l, raw[u].weight);                                        // l.cwiseProduct(weights)
        wsum = (k == 0) ? w : __fadd_rn(wsum, w);  wsum = __fadd_rn(wsum, w);               // W(2i) and W(2i+1)
        ws = (k == 0) ? w : __fadd_rn(ws, w);
        ls = (k == 0) ? l : __fadd_rn(ls, l);
        ++k;
    }
    const int np = 2 * k;
    // pass 2: weighted mean a = sum X * (w / wsum)
    float a0 = 0.f, a1 = 0.f; bool first = true;
    for (int q = 0; q < km; ++q) {
        const int u = mem[q];
        const SegRec s = raw[u];
        const float wt = __fdiv_rn(__fmul_rn(R.len[u], s.weight), wsum);
        const float t0 = __fmul_rn(s.y1, wt), t1 = __fmul_rn(s.x1, wt), t2 = __fmul_rn(s.y2, wt), t3 = __fmul_rn(s.x2, wt);
        a0 = first ? t0 : __fadd_rn(a0, t0); a1 = first ? t1 : __fadd_rn(a1, t1); first = false;
        a0 = __fadd_rn(a0, t2); a1 = __fadd_rn(a1, t3);
    }
    // pass 3: lower triangle of centered^T diag(wt) centered: (c_i * wt) * c_j
    float c00 = 0.f, c10 = 0.f, c11 = 0.f; first = true;
    for (int q = 0; q < km; ++q) {
        const int u = mem[q];
        const SegRec s = raw[u];
        const float wt = __fdiv_rn(__fmul_rn(R.len[u], s.weight), wsum);
#pragma unroll
        for (int e = 0; e < 2; ++e) {
            const float c0 = __fsub_rn(e ? s.y2 : s.y1, a0), c1 = __fsub_rn(e ? s.x2 : s.x1, a1);
            const float t00 = __fmul_rn(__fmul_rn(c0, wt), c0), t10 = __fmul_rn(__fmul_rn(c1, wt), c0), t11 = __fmul_rn(__fmul_rn(c1, wt), c1);
            c00 = first ? t00 : __fadd_rn(c00, t00); c10 = first ? t10 : __fadd_rn(c10, t10); c11 = first ? t11 : __fadd_rn(c11, t11);
            first = false;
        }
    }
    float cov[4] = {c00, c10, c10, c11}, eval[2], evec[4];
    sym_eig<2>(cov, eval, evec);
    const float d0 = evec[2], d1 = evec[3], n0 = evec[0], n1 = evec[1];                    // column 1 = direction, column 0 = normal
    // pass 4: projection range and mean |normal residual|
    float tmin = INFINITY, tmax = -INFINITY, esum = 0.f; first = true;
    for (int q = 0; q < km; ++q) {
        const int u = mem[q];
        const SegRec s = raw[u];
#pragma unroll
        for (int e = 0; e < 2; ++e) {
            const float c0 = __fsub_rn(e ? s.y2 : s.y1, a0), c1 = __fsub_rn(e ? s.x2 : s.x1, a1);
            const float tt = __fadd_rn(__fmul_rn(c0, d0), __fmul_rn(c1, d1));
            tmin = fminf(tmin, tt); tmax = fmaxf(tmax, tt);
            const float er = fabsf(__fadd_rn(__fmul_rn(c0, n0), __fmul_rn(c1, n1)));
            esum = first ? er : __fadd_rn(esum, er); first = false;
        }
    }
    out.y1 = __fadd_rn(a0, __fmul_rn(d0, tmin)); out.x1 = __fadd_rn(a1, __fmul_rn(d1, tmin));
    out.y2 = __fadd_rn(a0, __fmul_rn(d0, tmax)); out.x2 = __fadd_rn(a1, __fmul_rn(d1, tmax));
    out.err = __fdiv_rn(esum, (float)np);
    out.weight = __fdiv_rn(ws, ls);                                                         // line_detector.cpp:269
    out.group_id = -1;
}

__global__ void __launch_bounds__(1024) k_refine_merge(Comps K, FrameCtrl * ctrl)
{
    __shared__ int wsum[32];
    __shared__ int s_changed, s_base;
    const int slot = blockIdx.x;
    FrameCtrl & C = ctrl[slot];
    const int n = C.n_comp;
    const int t = threadIdx.x;
    const RefineScratch R = refine_scratch(K, slot);
    const SegRec * raw = reinterpret_cast<const SegRec *>(K.seg) + (size_t)slot * K.cap;
    if (n < 2 || C.refine_overflow) { if (t == 0) C.n_merged = 0; return; }
    const int ne = min(C.n_edges, R.edge_cap);
    // ---- labels: min over forward ancestors, propagated along the edge list until nothing changes ---------------------
    volatile int * label = R.label;
    for (int pass = 0; pass < n; ++pass) {
        if (t == 0) s_changed = 0;
        __syncthreads();
        bool ch = false;
        for (int e = t; e < ne; e += 1024) {
            const int i = R.edges[2 * e], j = R.edges[2 * e + 1];
            const int li = label[i];
            if (li < label[j]) { atomicMin(&R.label[j], li); ch = true; }
        }
        if (ch) s_changed = 1;
        __syncthreads();
        const bool again = s_changed != 0;
        __syncthreads();
        if (!again) break;
    }
    // ---- members per root ---------------------------------------------------------------------------------------------
    for (int u = t; u < n; u += 1024) atomicAdd(&R.cnt[R.label[u]], 1);
    __syncthreads();
    // ---- roots in ascending order -> output positions; members of multi-member groups in ascending order ----------------
    if (t == 0) s_base = 0;
    __syncthreads();
    for (int u0 = 0; u0 < n; u0 += 1024) {
        const int u = u0 + t;
        const int flag = (u < n) && (R.label[u] == u);
        int total; const int pos = s_base + scan_flag_1024(flag, wsum, total);
        if (flag) R.rrank[u] = pos;
        if (t == 0) s_base += total;
        __syncthreads();
    }
    const int n_roots = s_base;
    __syncthreads();
    // member lists of the multi-member groups: offsets by an ordered scan over the roots, unordered fill, then every
    // group's (short) list is put into index order by its own thread
    if (t == 0) s_base = 0;
    __syncthreads();
    for (int u0 = 0; u0 < n; u0 += 1024) {
        const int u = u0 + t;
        const int v = (u < n && R.label[u] == u && R.cnt[u] > 1) ? R.cnt[u] : 0;
        int total; const int pos = s_base + scan_val_1024(v, wsum, total);
        if (v) R.goff[u] = pos;
        if (t == 0) s_base += total;
        __syncthreads();
    }
    for (int u = t; u < n; u += 1024) {
        const int r = R.label[u];
        if (R.cnt[r] > 1) R.multi[R.goff[r] + atomicAdd(&R.fill[r], 1)] = u;
    }
    __syncthreads();
    // ---- merged segments (std::set order = ascending root index) ---------------------------------------------------------
    for (int u = t; u < n; u += 1024) {
        if (R.label[u] != u) continue;
        SegRec o;
        const int km = R.cnt[u];
        if (km == 1) o = raw[u];                                                            // line_detector.cpp:256-257
        else {
            int * mem = R.multi + R.goff[u];
            for (int a = 1; a < km; ++a) {                                                  // insertion sort, ascending index
                const int v = mem[a]; int b2 = a - 1;
                while (b2 >= 0 && mem[b2] > v) { mem[b2 + 1] = mem[b2]; --b2; }
                mem[b2 + 1] = v;
            }
            merge_group(raw, R, mem, km, o);
        }
        R.out[R.rrank[u]] = o;
    }
    if (t == 0) C.n_merged = n_roots;
}

} // namespace

int launch_refine(const Comps & K, FrameCtrl * ctrl, int n_slots, cudaStream_t st)
{
    dim3 gp(8, n_slots), ge(max(1, 148 * 4 / max(1, min(n_slots, 8))), n_slots);
    k_refine_prep<<<gp, 256, 0, st>>>(K, ctrl);
    k_refine_edges<<<ge, RF_TILE, 0, st>>>(K, ctrl);
    k_refine_merge<<<n_slots, 1024, 0, st>>>(K, ctrl);
    return 3;
}

const float * refine_output(const Comps & K, int slot) { return reinterpret_cast<const float *>(refine_scratch(K, slot).out); }

} // namespace lrb
