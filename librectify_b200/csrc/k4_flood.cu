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
// Latency is what bounds this stage, so a tentative flood works on 32x32 tiles held as bitmaps (see below): memory
// latency is paid per tile, not per BFS level, and the exploration touches only data that is constant during a round —
// mask8, dx, dy and the visited mask — through the non-coherent path; the owner stamps are write-only here and are
// inspected by the verify kernel.  owner word = [epoch | rank]; epochs DEcrease from round to round and frame to frame,
// so a newer stamp always wins over stale ones and the plane is re-initialised only when the epoch field wraps.
#include <cstdio>
#include <cstdlib>
#include "common.cuh"
#include "kernels.h"

namespace lrb {

namespace {

constexpr int FL_WARPS = 4;
constexpr int kPoolGranule = 256;  // list space a warp takes from the pool at a time (words)
constexpr int kVerifyWide = 64;   // regions longer than this are verified by a whole warp instead of eight lanes
constexpr int FT_SMALL = 16;      // tiles a flood may touch before it moves from shared memory to a global-memory table

// reference src/filter.cpp:113-116 neighbour order (dr, dc) — used by the serial fallback below
static __device__ __constant__ const int kDr[8] = {0, 0, 1, -1, -1, -1, 1, 1};
static __device__ __constant__ const int kDc[8] = {-1, 1, 0, 0, -1, 1, -1, 1};

// ---- tile-bitmap flood ----------------------------------------------------------------------------------------------
// The plane is cut into 32x32 tiles (a grid anchored at the seed).  For one seed a tile is two 32-word bitmaps, lane = row, bit = column:
//   E  eligible pixels: interior (filter.cpp:101-107), dilated bin mask bit b set, not visited, G_b(p) > thr (filter.cpp:131)
//   R  pixels reached so far
// A flood step inside a tile is pure register work: R <- E & dilate3x3(R) with two shuffles for the rows above/below,
// iterated to a fixed point; only what spills over the tile edge is handed to the (up to eight) neighbouring tiles, whose
// E is computed when they are first touched.  Memory latency is paid once per tile (two dependent rounds of wide loads)
// instead of once per BFS level, and the region comes out as the same set the sequential flood reaches — the order in
// which a connected set is discovered does not change the set.
struct FloodCtx {
    const unsigned short * mv; const float2 * dxy;      // mask8 | visited << 8; (dx, dy)
    int w, h, pitch, b;
    float thr, sb, cb;
};

// table of touched tiles: ids (ty << 16 | tx), dirty flags, a ring of dirty entries, and the bitmaps [entry][E 0..31 | R 32..63]
struct TileTab { int * ids; int * dirty; int * ring; unsigned * er; int cap; };
constexpr int kTileBias = 4096;
__device__ __forceinline__ int tile_id(int ty, int tx) { return ((ty + kTileBias) << 16) | (tx + kTileBias); }
__device__ __forceinline__ int tile_origin_y(int sr) { return sr - 16; }
__device__ __forceinline__ int tile_origin_x(int sc) { return ((sc - 8) >> 4) << 4; }      // seed column lands in [8, 24)

// 32 bytes (two 16-byte loads) -> 32 bits: bit i = bit `sh` of byte i
__device__ __forceinline__ unsigned bits_of_bytes(const uint4 & a, const uint4 & b, int sh)
{
#define LRB_F(wd) (((((wd) >> sh) & 0x01010101u) * 0x01020408u) >> 24)
    return LRB_F(a.x) | (LRB_F(a.y) << 4) | (LRB_F(a.z) << 8) | (LRB_F(a.w) << 12) |
           (LRB_F(b.x) << 16) | (LRB_F(b.y) << 20) | (LRB_F(b.z) << 24) | (LRB_F(b.w) << 28);
#undef LRB_F
}

// eligibility bitmap of the 32x32 tile whose top-left pixel is (y0, x0) — x0 a multiple of 16, either may be negative
// or beyond the frame; lane = row of the tile.  All loads are issued up front (one memory round trip per tile).
// Branch-free on purpose: every address is clamped into the plane and what lies outside the interior is masked out
// afterwards, so the function is ~5 KB of straight-line code (the predicated form was 29 KB — with 20+ warps per SM at
// different places of a 70 KB kernel the instruction cache was the bottleneck: ncu "no instruction" stalls).
// Returns (visited bitmap << 32) | eligible bitmap.  RB = rows whose dx/dy loads are in flight together (16: two dependent
// batches, 64 live registers less; 32: a single batch).
template <int RB>
__device__ __forceinline__ unsigned long long tile_eligible(const FloodCtx X, int y0, int x0, int lane)
{
    // (1) mask / visited bytes of this lane's row, two 16-byte halves each
    const int y = y0 + lane;
    const bool rowok = y >= 1 && y < X.h - 1;
    const bool h0 = x0 >= 0 && x0 + 16 <= X.pitch, h1 = x0 + 16 >= 0 && x0 + 32 <= X.pitch;      // warp-uniform
    const ptrdiff_t rbase = (ptrdiff_t)min(max(y, 0), X.h - 1) * X.pitch;
    const ptrdiff_t o0 = rbase + (h0 ? x0 : 0), o1 = rbase + (h1 ? x0 + 16 : 0);
    // 16-bit entries (mask8 | visited << 8): four 16-byte loads per row, split into mask bytes and visited bytes
    const uint4 w0 = __ldg(reinterpret_cast<const uint4 *>(X.mv + o0)), w1 = __ldg(reinterpret_cast<const uint4 *>(X.mv + o0) + 1);
    const uint4 w2 = __ldg(reinterpret_cast<const uint4 *>(X.mv + o1)), w3 = __ldg(reinterpret_cast<const uint4 *>(X.mv + o1) + 1);
#define LRB_LO(q_) __byte_perm((q_).x, (q_).y, 0x6420u), __byte_perm((q_).z, (q_).w, 0x6420u)
#define LRB_HI(q_) __byte_perm((q_).x, (q_).y, 0x7531u), __byte_perm((q_).z, (q_).w, 0x7531u)
    const uint4 m0 = make_uint4(LRB_LO(w0), LRB_LO(w1)), m1 = make_uint4(LRB_LO(w2), LRB_LO(w3));
    const uint4 v0 = make_uint4(LRB_HI(w0), LRB_HI(w1)), v1 = make_uint4(LRB_HI(w2), LRB_HI(w3));
#undef LRB_LO
#undef LRB_HI
    // (2) G_b(p) > thr for every pixel of the tile, row by row with lane = column (coalesced).  Rows outside the
    //     interior are dropped by `rowok` of the lane that keeps them, columns outside it by the `cols` mask below.
    unsigned G = 0;
    if (y0 >= 0 && y0 + 32 <= X.h && x0 >= 0 && x0 + 32 <= X.pitch) {       // the tile lies inside the plane (warp-uniform): no clamping
        const float2 * pg = X.dxy + ((ptrdiff_t)y0 * X.pitch + x0 + lane);
        const int pitch = X.pitch;
#pragma unroll
        for (int rb = 0; rb < 32; rb += RB) {
            float2 gg[RB];
#pragma unroll
            for (int i = 0; i < RB; ++i) gg[i] = __ldg(pg + (rb + i) * pitch);
#pragma unroll
            for (int i = 0; i < RB; ++i) {
                const float g = fabsf(__fadd_rn(__fmul_rn(gg[i].x, X.sb), __fmul_rn(gg[i].y, X.cb)));
                const unsigned bal = __ballot_sync(0xffffffffu, g > X.thr);
                G = (lane == rb + i) ? bal : G;
            }
        }
    } else {                                                                // frame border: clamped addresses, rolled loop
        const int xc = min(max(x0 + lane, 0), X.pitch - 1);
#pragma unroll 1
        for (int i = 0; i < 32; ++i) {
            const ptrdiff_t ro = (ptrdiff_t)min(max(y0 + i, 0), X.h - 1) * X.pitch + xc;
            const float2 gq = __ldg(X.dxy + ro);
            const float g = fabsf(__fadd_rn(__fmul_rn(gq.x, X.sb), __fmul_rn(gq.y, X.cb)));
            const unsigned bal = __ballot_sync(0xffffffffu, g > X.thr);
            G = (lane == i) ? bal : G;
        }
    }
    unsigned cols = 0xFFFFFFFFu;                                            // columns 1 .. w-2 only
    {
        const int first = 1 - x0;                                           // first interior column of this tile
        if (first > 0) cols &= (first >= 32) ? 0u : ~((1u << first) - 1u);
        const int last = X.w - 1 - x0;                                      // first column that is NOT interior
        if (last < 32) cols &= (last <= 0) ? 0u : ((1u << last) - 1u);
    }
    const unsigned lo = h0 ? 0x0000FFFFu : 0u, hi = h1 ? 0xFFFF0000u : 0u;  // halves that lie outside the pitch
    const unsigned V = bits_of_bytes(v0, v1, 0);
    const unsigned E = rowok ? (bits_of_bytes(m0, m1, X.b) & ~V & cols & G & (lo | hi)) : 0u;
    return ((unsigned long long)V << 32) | E;
}

__device__ __forceinline__ int tab_find(const TileTab & T, int n_ent, int id, int lane)
{
    for (int base = 0; base < n_ent; base += 32) {
        const int i = base + lane;
        const unsigned m = __ballot_sync(0xffffffffu, i < n_ent && T.ids[i] == id);
        if (m) return base + __ffs(m) - 1;
    }
    return -1;
}

// Tentative region of one seed by one warp.  Returns the pixel count, -1 when the table is full, or -2 when the seed
// pixel is already visited (line_detector.cpp:103: the seed was swallowed by an earlier region; its visited byte comes
// with the seed tile, so no separate memory round trip is spent on it).
// On success the bitmaps of the n_ent touched tiles are in T (R part).
template <int RB>
__device__ __forceinline__ int flood_tiles(const FloodCtx & X, const TileTab & T, int sr, int sc, int & n_ent_out, int * dbg)
{
    const int lane = threadIdx.x & 31;
    // the tile grid is the seed's own: tile (0,0) is centred on the seed (x origin rounded to 16 for the vector loads),
    // so that the typical small region never leaves its first tile
    const int oy = tile_origin_y(sr), ox = tile_origin_x(sc);
    int n_ent = 1, q_head = 0, q_tail = 1;
    {   // the seed's tile; the seed itself is accepted unconditionally (filter.cpp:119-122)
        const unsigned long long ve = tile_eligible<RB>(X, oy, ox, lane);
        unsigned E = (unsigned)ve;
        const unsigned sbit = (lane == sr - oy) ? (1u << (sc - ox)) : 0u;
        if (__any_sync(0xffffffffu, ((unsigned)(ve >> 32) & sbit) != 0u)) { n_ent_out = 0; return -2; }
        E |= sbit;
        T.er[lane] = E; T.er[32 + lane] = sbit;
        if (lane == 0) { T.ids[0] = tile_id(0, 0); T.dirty[0] = 1; T.ring[0] = 0; }
        __syncwarp();
    }
    while (q_head < q_tail) {
        const int e = T.ring[q_head % T.cap]; ++q_head; ++dbg[0];
        const int id = T.ids[e];
        const int ty = (id >> 16) - kTileBias, tx = (id & 0xFFFF) - kTileBias;
        const unsigned E = T.er[e * 64 + lane];
        unsigned R = T.er[e * 64 + 32 + lane];
        __syncwarp();
        if (lane == 0) T.dirty[e] = 0;
        unsigned hz;
        // Fixed point of R <- E & dilate(R).  Inside a row the flood is completed in one go with the carry trick: for
        // R within E, ((E + R) ^ E) & E | R is R extended to the upper end of every run of E it touches (the carry of the
        // addition ripples through the run); the same on the bit-reversed words extends downwards.  An iteration is then
        // one complete horizontal fill plus one step to the rows above and below, so a near-horizontal line takes a
        // handful of iterations instead of one per pixel of its length.
        const unsigned Erev = __brev(E);
        while (true) {
            const unsigned s1 = E + R;
            const unsigned f1 = ((s1 ^ E) & E) | R;
            const unsigned r1 = __brev(f1);
            const unsigned s2 = Erev + r1;
            const unsigned H = __brev(((s2 ^ Erev) & Erev) | r1);              // R, horizontally closed
            hz = H | (H << 1) | (H >> 1);
            unsigned up = __shfl_up_sync(0xffffffffu, hz, 1), dn = __shfl_down_sync(0xffffffffu, hz, 1);
            if (lane == 0) up = 0u;
            if (lane == 31) dn = 0u;
            const unsigned Rn = H | (E & (up | dn));
            const bool ch = __any_sync(0xffffffffu, Rn != R);
            R = Rn; ++dbg[1];
            if (!ch) break;
        }
        T.er[e * 64 + 32 + lane] = R;
        __syncwarp();
        // what reaches over the tile edge: left/right columns, first/last row
        const bool edge = __any_sync(0xffffffffu, (R & 0x80000001u) != 0u || ((lane == 0 || lane == 31) && R != 0u));
        if (!edge) continue;
        const unsigned hr = R >> 31, hl = (R & 1u) << 31;
#pragma unroll 1
        for (int k = 0; k < 8; ++k) {                                       // rolled: the kernel has to stay inside the instruction cache
            const int dy = (k < 3) ? -1 : (k < 5 ? 0 : 1);
            const int dxk = (k < 3) ? k - 1 : (k < 5 ? (k == 3 ? -1 : 1) : k - 6);
            const unsigned src = dxk == 0 ? hz : (dxk > 0 ? hr : hl);
            unsigned spill;
            if (dy == 0) {
                unsigned up = __shfl_up_sync(0xffffffffu, src, 1), dn = __shfl_down_sync(0xffffffffu, src, 1);
                if (lane == 0) up = 0u;
                if (lane == 31) dn = 0u;
                spill = src | up | dn;
            } else if (dy > 0) {
                const unsigned v = __shfl_sync(0xffffffffu, src, 31);
                spill = lane == 0 ? v : 0u;
            } else {
                const unsigned v = __shfl_sync(0xffffffffu, src, 0);
                spill = lane == 31 ? v : 0u;
            }
            if (!__any_sync(0xffffffffu, spill != 0u)) continue;
            const int ny = ty + dy, nx = tx + dxk;
            const int ny0 = oy + ny * 32, nx0 = ox + nx * 32;
            if (ny0 + 31 < 1 || ny0 > X.h - 2 || nx0 + 31 < 1 || nx0 > X.w - 2) continue;      // no interior pixel there
            const int nid = tile_id(ny, nx);
            int ne = tab_find(T, n_ent, nid, lane);
            if (ne < 0) {
                const unsigned En = (unsigned)tile_eligible<RB>(X, ny0, nx0, lane); ++dbg[2];
                if (!__any_sync(0xffffffffu, (En & spill) != 0u)) continue;       // nothing to reach there: do not spend an entry
                if (n_ent >= T.cap) { n_ent_out = n_ent; return -1; }
                ne = n_ent++;
                T.er[ne * 64 + lane] = En; T.er[ne * 64 + 32 + lane] = 0u;
                if (lane == 0) { T.ids[ne] = nid; T.dirty[ne] = 0; }
                __syncwarp();
            }
            const unsigned nb = spill & T.er[ne * 64 + lane] & ~T.er[ne * 64 + 32 + lane];
            if (__any_sync(0xffffffffu, nb != 0u)) {
                const int was_dirty = T.dirty[ne];                            // same word for every lane
                __syncwarp();
                T.er[ne * 64 + 32 + lane] |= nb;
                if (!was_dirty) {
                    if (lane == 0) { T.dirty[ne] = 1; T.ring[q_tail % T.cap] = ne; }
                    ++q_tail;
                }
                __syncwarp();
            }
        }
    }
    int cnt = 0;
    for (int e = 0; e < n_ent; ++e) cnt += __popc(T.er[e * 64 + 32 + lane]);
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, s);
    n_ent_out = n_ent;
    return cnt;
}

// stamp the region (fire-and-forget atomicMin(owner, epoch:rank)) and write its pixel list (any order: the verify kernel
// only needs the set, the fit re-walks the region in the reference's order).  Per tile either lane = row, every lane
// walking the set bits of its own row, or row by row with lane = column — whichever takes fewer iterations (a vertical
// line has 32 rows of 1-3 pixels, a horizontal one 1-3 rows of 32).
__device__ __forceinline__ void emit_region(const Planes & P, unsigned * owner, const TileTab & T, int n_ent, int oy, int ox, unsigned key, unsigned * list)
{
    const int lane = threadIdx.x & 31;
    int off = 0;
    for (int e = 0; e < n_ent; ++e) {
        const unsigned R = T.er[e * 64 + 32 + lane];
        const int id = T.ids[e];
        const int y0 = oy + ((id >> 16) - kTileBias) * 32, x0 = ox + ((id & 0xFFFF) - kTileBias) * 32;
        const int c = __popc(R);
        unsigned rows = __ballot_sync(0xffffffffu, R != 0u);
        if (rows == 0u) continue;
        const int maxc = __reduce_max_sync(0xffffffffu, c);
        if (maxc <= __popc(rows)) {                                         // lane = row
            int pre = c;
#pragma unroll
            for (int sft = 1; sft < 32; sft <<= 1) { const int v = __shfl_up_sync(0xffffffffu, pre, sft); if (lane >= sft) pre += v; }
            const int total = __shfl_sync(0xffffffffu, pre, 31);
            int i = off + pre - c;
            const int y = y0 + lane;
            unsigned * orow = owner + (size_t)y * P.pitch + x0;
            unsigned m = R;
            while (m) {
                const int bcol = __ffs(m) - 1; m &= m - 1;
                atomicMin(orow + bcol, key);
                list[i++] = ((unsigned)y << 16) | (unsigned)(x0 + bcol);
            }
            off += total;
        } else {                                                            // row by row, lane = column
            const int x = x0 + lane;
            while (rows) {
                const int r = __ffs(rows) - 1; rows &= rows - 1;
                const unsigned m = __shfl_sync(0xffffffffu, R, r);
                if ((m >> lane) & 1u) {
                    const int y = y0 + r;
                    atomicMin(owner + (size_t)y * P.pitch + x, key);
                    list[off + __popc(m & ((1u << lane) - 1u))] = ((unsigned)y << 16) | (unsigned)x;
                }
                off += __popc(m);
            }
        }
    }
}

// one tentative flood of seed `rank` by one warp
template <int RB>
__device__ __forceinline__ void flood_one(const Planes & P, const Seeds & S, const Comps & K, FrameCtrl & C, int slot, int rank,
                                          unsigned epoch, const TileTab & Tsm, int & g_off, int & g_left)
{
    const int lane = threadIdx.x & 31;
    const size_t so = (size_t)slot * S.cap;
    const unsigned st0 = __ldcg(S.status + so + rank);
    const unsigned spx = S.idx[so + rank];                        // independent loads: one round trip for all four
    const int sbin = S.bin[so + rank];
    const float sthr = S.thr[so + rank];
    if (st0 != kStatusLive) return;
    unsigned long long t_begin; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_begin));

    const size_t pbase = (size_t)slot * P.stride;
    unsigned * owner = P.owner + pbase;
    FloodCtx X;
    X.dxy = P.dxy + pbase; X.mv = P.mv + pbase;
    X.w = P.w; X.h = P.h; X.pitch = P.pitch;

    const unsigned key = (epoch << 24) | (unsigned)rank;
    const int sr = spx / P.w, sc = spx - sr * P.w;
    X.b = sbin;
    X.thr = sthr;
    X.sb = kSin[X.b]; X.cb = kCos[X.b];

    int n_ent = 0; int dbg[3] = {0, 0, 0};
    TileTab T = Tsm;
    int big = -1, size = -1;
    for (int attempt = 0; attempt < 2 && size == -1; ++attempt) {
        if (attempt == 1) {                                       // a long region: redo it on a table in global memory
            if (lane == 0) {
                unsigned h = (unsigned)rank * 2654435761u + blockIdx.x;
                while (big < 0) {
                    for (int t = 0; t < K.big_slots && big < 0; ++t) {
                        const int i = (int)((h + t) % (unsigned)K.big_slots);
                        if (atomicCAS(K.big_lock + i, 0, 1) == 0) big = i;
                    }
                    if (big < 0) __nanosleep(500);
                }
                __threadfence();
            }
            big = __shfl_sync(0xffffffffu, big, 0);
            char * base = reinterpret_cast<char *>(K.big) + (size_t)big * K.big_bytes;
            T.cap = K.big_cap;
            T.er = reinterpret_cast<unsigned *>(base);
            T.ids = reinterpret_cast<int *>(base + (size_t)K.big_cap * 256);
            T.dirty = T.ids + K.big_cap; T.ring = T.dirty + K.big_cap;
        }
        size = flood_tiles<RB>(X, T, sr, sc, n_ent, dbg);
    }
    if (size == -2) {                                             // line_detector.cpp:103: seed already swallowed
        if (lane == 0) S.status[so + rank] = kStatusDropped;
        return;
    }
    bool overflow = size < 0;
    int off = 0;
    if (!overflow) {
        if (size > g_left) {                                      // a new granule of list space for this warp
            const int want = max(size, kPoolGranule);
            if (lane == 0) g_off = atomicAdd(&C.pool_cursor, want);
            g_off = __shfl_sync(0xffffffffu, g_off, 0);
            g_left = (g_off + want <= K.cap_pool) ? want : -1;
        }
        off = g_off;
        if (g_left < 0) { overflow = true; g_left = 0; }
        else {
            g_off += size; g_left -= size;
            emit_region(P, owner, T, n_ent, tile_origin_y(sr), tile_origin_x(sc), key, K.pool + (size_t)slot * K.cap_pool + off);
        }
    }
    __syncwarp();
    if (big >= 0 && lane == 0) { __threadfence(); atomicExch(K.big_lock + big, 0); }
    if (lane == 0) {
#ifdef LRB_FLOOD_STATS
        { unsigned long long t_end; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_end)); const unsigned long long d = t_end - t_begin;
          const unsigned long long pk = (d << 24) | (unsigned long long)(max(size, 0) & 0xFFFFFF);
          { const int rr = (int)((0xFEu - epoch) % 24u); if (atomicMax(&C.dbg_rmax[rr], pk) < pk) C.dbg_rtiles[rr] = n_ent; }
          if (atomicMax(&C.dbg_maxdur, pk) < pk) { C.dbg_slow[0] = dbg[0]; C.dbg_slow[1] = dbg[1]; C.dbg_slow[2] = dbg[2]; C.dbg_slow[3] = n_ent; }
          atomicAdd(&C.dbg_sumdur, d); }
        atomicAdd(&C.dbg_floods, 1); atomicAdd(&C.dbg_tiles, n_ent); atomicAdd(&C.dbg_px, max(size, 0)); atomicMax(&C.dbg_maxpx, size);
        if (big >= 0) atomicAdd(&C.dbg_big, 1);
#endif
        if (overflow) C.flood_overflow = 1;
        S.size[so + rank] = overflow ? -1 : size;
        S.list_head[so + rank] = off;
    }
    __syncwarp();
}

template <int MINB, int RB>
__global__ void __launch_bounds__(FL_WARPS * 32, MINB)
k_flood_round(Planes P, Seeds S, Comps K, FrameCtrl * ctrl, int round, unsigned epoch_first, int slot0)
{
    __shared__ unsigned s_er[FL_WARPS][FT_SMALL * 64];
    __shared__ int s_ids[FL_WARPS][FT_SMALL * 3];
    const int slot = slot0 + blockIdx.y;
    FrameCtrl & C = ctrl[slot];
    if (C.live[round] == 0 || C.flood_overflow) return;
    const int n_esc = C.esc2_n;                                 // this round's floods that outgrew both sub-warp windows (k_flood_small<8>, <16>)
    const int warp = threadIdx.x >> 5;
    const unsigned epoch = epoch_first - (unsigned)round;
    TileTab T;
    T.er = s_er[warp]; T.ids = s_ids[warp]; T.dirty = T.ids + FT_SMALL; T.ring = T.dirty + FT_SMALL; T.cap = FT_SMALL;
    int g_off = 0, g_left = 0;
    // Seeds are handed out dynamically, one at a time and in priority order (the strongest seeds own the longest regions,
    // so the long floods start first and the short ones fill the gaps); the next position is requested before the
    // current flood starts so that the round trip of the atomic is hidden behind it.
    const int lane = threadIdx.x & 31;
    int * q = &C.esc2_q;
    const int * esc = S.esc2 + (size_t)slot * S.cap;
    int nxt = 0;
    if (lane == 0) nxt = atomicAdd(q, 1);
    while (true) {
        const int pos = __shfl_sync(0xffffffffu, nxt, 0);
        if (pos >= n_esc) break;
        if (lane == 0) nxt = atomicAdd(q, 1);
        flood_one<RB>(P, S, K, C, slot, esc[pos], epoch, T, g_off, g_left);
    }
}

// ---- sub-warp flood: 8 lanes per seed, four seeds per warp -------------------------------------------------------------
// Most tentative regions are tiny (median 6 pixels, 95 % below 24: profiles/r2_flood_regions.md) and thin, and the seed's
// orientation bin says which way they run.  The fast path therefore floods inside ONE 256-pixel window per seed, laid along the
// expected line direction — WIDE: 8 rows x 32 columns for the bins of near-horizontal edges, TALL: 32 rows x 8 columns
// otherwise — held as eight 32-bit words, one per lane of an 8-lane group: word = one line of the window along its long axis.
// The flood step is the same register-only fixed point as in flood_tiles (complete fill along the long axis by the carry
// trick, one shuffle step across it), 8-connectivity being symmetric under the transposition that distinguishes the layouts.
// A region that reaches the window's outer ring may continue outside: the seed is handed to the full-tile path (k_flood_round)
// through the round's escalation list, as are seeds whose window does not lie inside the plane.  ~12x fewer instructions and
// 4x fewer bytes per typical flood than a 32 x 32 tile per warp.
__device__ __forceinline__ unsigned bits4(unsigned w, int sh) { return (((w >> sh) & 0x01010101u) * 0x01020408u) >> 24; }   // 4 bytes -> 4 bits: bit i = bit `sh` of byte i

// G = lanes per seed (8: first pass over the round's window, 256-pixel windows; 16: second pass over what the first pass
// escalated, 512-pixel windows 16 x 32 / 32 x 16).  Ranks come from the window (src == nullptr) or from an escalation list;
// floods that outgrow this pass's window go to the list `dst` (consumed by the next pass / the full-tile kernel).
template <int G, int MINB>
__global__ void __launch_bounds__(FL_WARPS * 32, MINB)
k_flood_small(Planes P, Seeds S, Comps K, FrameCtrl * ctrl, int round, unsigned epoch_first, int slot0, int mode, int last_pass)
{
    constexpr int NG = 32 / G;                                    // seeds per warp
    constexpr int RUNS_PER_ROW = G / 8;                           // TALL: runs of eight pixels per window row
    const int slot = slot0 + blockIdx.y;
    FrameCtrl & C = ctrl[slot];
    if (C.live[round] == 0 || C.flood_overflow) return;
    const unsigned epoch = epoch_first - (unsigned)round;
    const int lane = threadIdx.x & 31, sub = lane % G, grp = lane / G;
    const size_t so = (size_t)slot * S.cap, pbase = (size_t)slot * P.stride;
    // first pass: the window's ranks; second pass: the first pass's escalation list
    const bool first_pass = (G == 8);
    const bool listed = first_pass && (mode & 8);                 // the window's live seeds as listed by the prune pass (S.esc, C.wl_n)
    const int lo = (first_pass && !listed) ? C.win_lo : 0, hi = first_pass ? (listed ? C.wl_n : C.win_hi) : C.esc_n;
    const int * src = S.esc + so;
    // what this pass cannot finish goes to the next pass's list, from the last sub-warp pass to the full-tile kernel's
    int * dst = last_pass ? S.esc2 + so : S.esc + so;
    int * dst_n = last_pass ? &C.esc2_n : &C.esc_n;
    int * q = first_pass ? &C.queue[round] : &C.esc_q;
    const float2 * pdxy = P.dxy + pbase;
    const unsigned short * pmv = P.mv + pbase;
    unsigned * owner = P.owner + pbase;
    unsigned * pool = K.pool + (size_t)slot * K.cap_pool;
    const int W = P.w, H = P.h, pitch = P.pitch;
    int g_off = 0, g_left = 0;
    int nxt = 0;
    if (lane == 0) nxt = atomicAdd(q, NG);
    while (true) {
        const int base = lo + __shfl_sync(0xffffffffu, nxt, 0);
        if (base >= hi) break;
        if (lane == 0) nxt = atomicAdd(q, NG);                    // the next batch, requested before this one starts
        const int pos = base + grp;
        const bool valid = pos < hi;
        const int rank = !valid ? 0 : ((first_pass && !listed) ? pos : src[pos]);
        const unsigned st0 = valid ? __ldcg(S.status + so + rank) : kStatusDropped;
        const unsigned spx = valid ? S.idx[so + rank] : 0u;
        const int sbin = valid ? S.bin[so + rank] : 0;
        const float sthr = valid ? S.thr[so + rank] : 0.f;
        const bool live = valid && st0 == kStatusLive;
        const int sr = (int)(spx / (unsigned)W), sc = (int)(spx - (unsigned)sr * (unsigned)W);
        // bins 0, 1, 2, 6, 7: the edge runs closer to the x axis (the energy weighs dy more) -> WIDE window
        const bool wide = (sbin <= 2) || (sbin >= 6);
        int y0 = wide ? sr - (G / 2 - 1) : sr - 15;
        int x0 = wide ? ((sc - 14) & ~3) : ((sc - (G / 2 - 2)) & ~3);                   // 16-byte aligned float rows
        const bool inside = !(mode & 1) && y0 >= 0 && y0 + (wide ? G : 32) <= H && x0 >= 0 && x0 + (wide ? 32 : G) <= pitch;   // mode 1 (LRB_FLOOD_SMALL_MODE=1, diagnosis): every flood takes the full-tile path
        const bool work = live && inside;
        if (!__any_sync(0xffffffffu, work)) {                     // nothing to flood here: committed / dropped seeds, or all need the next pass
            const unsigned be = __ballot_sync(0xffffffffu, sub == 0 && live);
            if (sub == 0 && live) dst[atomicAdd(dst_n, 1)] = rank;
            if (lane == 0 && be && last_pass) atomicAdd(&C.floods_esc, __popc(be));
            continue;
        }
        if (!work) { y0 = 0; x0 = 0; }                           // idle groups compute on a harmless window (every shift below stays in range)
        const float sb = kSin[sbin], cb = kCos[sbin];
        // ---- eligibility of the window: every lane evaluates four runs of eight pixels (row-major inside the window)
        unsigned E = 0u, V = 0u;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int row = !work ? 0 : (wide ? y0 + sub : y0 + sub + G * (k / RUNS_PER_ROW));
            const int col = !work ? 0 : (wide ? x0 + 8 * k : x0 + 8 * (k % RUNS_PER_ROW));
            const ptrdiff_t o = (ptrdiff_t)row * pitch + col;     // idle groups read the first pixels of the plane
            // eight pixels: 64 contiguous bytes of (dx, dy) pairs, 16 contiguous bytes of mask | visited entries (8-byte aligned: x0 is a multiple of 4)
            const float4 * qg = reinterpret_cast<const float4 *>(pdxy + o);
            const float4 a0 = __ldg(qg), a1 = __ldg(qg + 1), a2 = __ldg(qg + 2), a3 = __ldg(qg + 3);
            const uint2 e0 = __ldg(reinterpret_cast<const uint2 *>(pmv + o)), e1 = __ldg(reinterpret_cast<const uint2 *>(pmv + o) + 1);
            uint2 m, v;
            m.x = __byte_perm(e0.x, e0.y, 0x6420u); m.y = __byte_perm(e1.x, e1.y, 0x6420u);
            v.x = __byte_perm(e0.x, e0.y, 0x7531u); v.y = __byte_perm(e1.x, e1.y, 0x7531u);
            unsigned g = 0u;
#define LRB_G(val_x, val_y, bit) g |= (fabsf(__fadd_rn(__fmul_rn(val_x, sb), __fmul_rn(val_y, cb))) > sthr) ? (1u << (bit)) : 0u;
            LRB_G(a0.x, a0.y, 0) LRB_G(a0.z, a0.w, 1) LRB_G(a1.x, a1.y, 2) LRB_G(a1.z, a1.w, 3)
            LRB_G(a2.x, a2.y, 4) LRB_G(a2.z, a2.w, 5) LRB_G(a3.x, a3.y, 6) LRB_G(a3.z, a3.w, 7)
#undef LRB_G
            const unsigned mb = bits4(m.x, sbin) | (bits4(m.y, sbin) << 4);
            const unsigned vb = bits4(v.x, 0) | (bits4(v.y, 0) << 4);
            // interior only (reference src/filter.cpp:101-107): rows 1 .. H-2, columns 1 .. W-2
            const int c_lo = max(0, 1 - col), c_hi = min(8, W - 1 - col);
            const unsigned cm = (row >= 1 && row < H - 1 && c_hi > c_lo) ? (((1u << c_hi) - 1u) & ~((1u << c_lo) - 1u)) : 0u;
            E |= (g & mb & ~vb & cm) << (8 * k);
            V |= vb << (8 * k);
        }
        // the seed: lane / bit in the row-major form
        const int sy = work ? sr - y0 : 0, sx = work ? sc - x0 : 0;
        const unsigned sbit_rm = wide ? ((sub == sy) ? (1u << sx) : 0u) : ((sub == (sy % G)) ? (1u << (G * (sy / G) + sx)) : 0u);
        const unsigned gmask = (G == 8 ? 0xFFu : 0xFFFFu) << (G * grp);
        const unsigned seen = __ballot_sync(0xffffffffu, (V & sbit_rm) != 0u);       // evaluated by every lane (no short-circuit around a collective)
        const bool swallowed = work && ((seen & gmask) != 0u);                       // line_detector.cpp:103
        {   // TALL: G x G bit blocks transposed across the group -> lane = column, bit = row (computed by every lane: the
            // groups of a warp may use different layouts and the shuffles must stay convergent)
            unsigned Et = E;
#pragma unroll
            for (int s = G / 2; s > 0; s >>= 1) {
                const unsigned msk = s == 8 ? 0xFF00FF00u : (s == 4 ? 0xF0F0F0F0u : (s == 2 ? 0xCCCCCCCCu : 0xAAAAAAAAu));
                const unsigned t = __shfl_xor_sync(0xffffffffu, Et, s);
                Et = (sub & s) ? ((Et & msk) | ((t >> s) & ~msk)) : ((Et & ~msk) | ((t << s) & msk));
            }
            E = wide ? E : Et;
        }
        const unsigned sbit = wide ? sbit_rm : ((sub == sx) ? (1u << sy) : 0u);
        const bool go = work && !swallowed;
        E = go ? (E | sbit) : 0u;                                 // the seed itself is accepted unconditionally (filter.cpp:119-122)
        unsigned R = go ? sbit : 0u;
        const unsigned Erev = __brev(E);
        int guard_fl = 0;
        while (true) {
            if (++guard_fl > 64 * G) { if (lane == 0) C.dbg_tiles = rank; break; }     // cannot happen: R grows monotonically inside E
            const unsigned s1 = E + R;
            const unsigned f1 = ((s1 ^ E) & E) | R;
            const unsigned r1 = __brev(f1);
            const unsigned s2 = Erev + r1;
            const unsigned Hh = __brev(((s2 ^ Erev) & Erev) | r1);              // R closed along the long axis
            const unsigned hz = Hh | (Hh << 1) | (Hh >> 1);
            unsigned up = __shfl_up_sync(0xffffffffu, hz, 1, G), dn = __shfl_down_sync(0xffffffffu, hz, 1, G);
            if (sub == 0) up = 0u;
            if (sub == G - 1) dn = 0u;
            const unsigned Rn = Hh | (E & (up | dn));
            const bool ch = Rn != R;
            R = Rn;
            if (!__any_sync(0xffffffffu, ch)) break;
        }
        // a region on the outer ring of the window may go on outside it
        const bool ring = (sub == 0 || sub == G - 1) ? (R != 0u) : ((R & 0x80000001u) != 0u);
        const bool spilled = (__ballot_sync(0xffffffffu, ring) & gmask) != 0u;
        const bool escalate = live && (!inside || (go && (spilled || (mode & 4))));
        const bool done = go && !spilled && !(mode & 4);
        int cnt = done ? __popc(R) : 0;
        int pre = cnt;                                            // inclusive prefix inside the group
#pragma unroll
        for (int sft = 1; sft < G; sft <<= 1) { const int v = __shfl_up_sync(0xffffffffu, pre, sft, G); if (sub >= sft) pre += v; }
        const int size = __shfl_sync(0xffffffffu, pre, G - 1, G);
        // list space for the regions of the warp: one granule request at a time
        int total = 0, before = 0;
#pragma unroll
        for (int gi = 0; gi < NG; ++gi) { const int sg = __shfl_sync(0xffffffffu, size, gi * G); total += sg; before += (gi < grp) ? sg : 0; }
        bool overflow = false;
        if (total > g_left) {
            const int want = max(total, kPoolGranule);
            if (lane == 0) g_off = atomicAdd(&C.pool_cursor, want);
            g_off = __shfl_sync(0xffffffffu, g_off, 0);
            g_left = (g_off + want <= K.cap_pool) ? want : -1;
        }
        if (g_left < 0) { overflow = true; g_left = 0; }
        const int off = g_off + before;
        if (!overflow) { g_off += total; g_left -= total; }
        if (done && !overflow && !(mode & 2)) {
            const unsigned key = (epoch << 24) | (unsigned)rank;
            unsigned * list = pool + off + (pre - cnt);
            unsigned mm = R;
            while (mm) {
                const int b = __ffs(mm) - 1; mm &= mm - 1;
                const int y = wide ? y0 + sub : y0 + b, x = wide ? x0 + b : x0 + sub;
                atomicMin(owner + (size_t)y * pitch + x, key);
                *list++ = ((unsigned)y << 16) | (unsigned)x;
            }
        }
        {   // statistics: one atomic per warp and kind
            const unsigned bd = __ballot_sync(0xffffffffu, sub == 0 && done), be = __ballot_sync(0xffffffffu, sub == 0 && escalate);
            if (lane == 0) { if (bd) atomicAdd(first_pass ? &C.floods_small : &C.floods_mid, __popc(bd)); if (be && last_pass) atomicAdd(&C.floods_esc, __popc(be)); }
        }
        if (sub == 0 && live) {
            if (swallowed) S.status[so + rank] = kStatusDropped;
            else if (escalate) dst[atomicAdd(dst_n, 1)] = rank;
            else {
                if (overflow) C.flood_overflow = 1;
                S.size[so + rank] = overflow ? -1 : size;
                S.list_head[so + rank] = off;
            }
        }
        __syncwarp();
    }
}

// Seeds of the coming window whose pixel already lies in a committed region are skipped by the sequential algorithm
// (line_detector.cpp:103).  The flood kernel finds that out only after it has evaluated the seed's whole tile (the visited
// byte arrives with the tile); from the second round on 40 % of the window seeds are in that state (ncu, profiles/), so
// they are dropped here first, one byte load each.  Exact for the same reason as the -2 path of flood_tiles: the visited
// mask only holds committed regions, and those belong to seeds of higher priority.
__global__ void __launch_bounds__(256)
k_flood_prune(Planes P, Seeds S, FrameCtrl * ctrl, int round, int slot0)
{
    const int slot = slot0 + blockIdx.y;
    FrameCtrl & C = ctrl[slot];
    if (C.live[round] == 0 || C.flood_overflow) return;
    const int lo = C.win_lo, hi = C.win_hi;
    const size_t so = (size_t)slot * S.cap;
    const unsigned short * mv = P.mv + (size_t)slot * P.stride;
    // The survivors are listed (S.esc; any order: tentative floods are independent and their stamps carry the rank), so that the
    // sub-warp pass floods a dense list: from the second round on half of a window's seeds are committed, dropped here, or were
    // dropped before, and a batch of four seeds with one live seed costs the same as a full one.
    int * wl = S.esc + so;
    const int lane = threadIdx.x & 31;
    for (int r0 = lo + (blockIdx.x * blockDim.x + threadIdx.x - lane); r0 < hi; r0 += gridDim.x * blockDim.x) {
        const int rank = r0 + lane;
        bool keep = rank < hi && S.status[so + rank] == kStatusLive;
        if (keep) {
            const unsigned spx = S.idx[so + rank];
            const int sr = spx / P.w, sc = spx - sr * P.w;
            if (mv[(size_t)sr * P.pitch + sc] >> 8) { S.status[so + rank] = kStatusDropped; keep = false; }
        }
        const unsigned bm = __ballot_sync(0xffffffffu, keep);
        if (bm == 0u) continue;
        int base = 0;
        if (lane == 0) base = atomicAdd(&C.wl_n, __popc(bm));
        base = __shfl_sync(0xffffffffu, base, 0);
        if (keep) wl[base + __popc(bm & ((1u << lane) - 1u))] = rank;
    }
}

// verify + commit + mark visited, eight lanes per window seed; the last block to finish moves the window
__global__ void __launch_bounds__(256)
k_flood_verify(Planes P, Seeds S, Comps K, FrameCtrl * ctrl, int round, unsigned epoch_first, int slot0)
{
    __shared__ int s_last;
    const int slot = slot0 + blockIdx.y;
    FrameCtrl & C = ctrl[slot];
    if (C.live[round] == 0 || C.flood_overflow) return;
    const int n = C.n_seeds;
    const int lo = C.win_lo, hi = C.win_hi;
    const unsigned epoch = epoch_first - (unsigned)round;
    const size_t so = (size_t)slot * S.cap, pbase = (size_t)slot * P.stride;
    const unsigned * owner = P.owner + pbase;
    uint8_t * vis = reinterpret_cast<uint8_t *>(P.mv + pbase) + 1;      // the visited byte of pixel p is vis[2 * p]
    const unsigned * pool = K.pool + (size_t)slot * K.cap_pool;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    // eight lanes per seed, four seeds per warp (the typical region has 6 pixels: profiles/r2_flood_regions.md); four list entries
    // per lane in flight so that a long region still costs size / 32 dependent round trips
    const int sub = lane & 7, grp = lane >> 3;
    int live = 0, first = 0x7fffffff;
    for (int r0 = lo + (blockIdx.x * nw + warp) * 4; r0 < hi; r0 += gridDim.x * nw * 4) {
        const int rank = r0 + grp;
        const bool valid = rank < hi && S.status[so + rank] == kStatusLive;
        const unsigned key = (epoch << 24) | (unsigned)rank;
        const int size = valid ? S.size[so + rank] : 0;
        const int lhead = valid ? S.list_head[so + rank] : 0;
        const unsigned * list = pool + lhead;
        const bool big = size > kVerifyWide;                      // a long region: the whole warp takes it afterwards
        // pass 1: is every pixel still stamped with this seed's key?
        bool mine = true;
        if (!big) for (int i = sub; i < size; i += 32) {
            unsigned w[4], o[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) w[u] = (i + 8 * u < size) ? __ldcg(list + i + 8 * u) : 0u;
#pragma unroll
            for (int u = 0; u < 4; ++u) o[u] = (i + 8 * u < size) ? __ldcg(owner + (size_t)(w[u] >> 16) * P.pitch + (w[u] & 0xFFFFu)) : key;
#pragma unroll
            for (int u = 0; u < 4; ++u) mine &= (o[u] == key);
        }
        const unsigned bm = __ballot_sync(0xffffffffu, mine);
        unsigned bigm = __ballot_sync(0xffffffffu, big && sub == 0);
        if (valid && !big) {
            if (((bm >> (8 * grp)) & 0xFFu) != 0xFFu) { ++live; first = min(first, rank); }
            else {
                // pass 2: the region is final — it becomes part of the visited mask (whether or not it is kept later)
                for (int i = sub; i < size; i += 8) {
                    const unsigned w = __ldcg(list + i);
                    vis[2 * ((size_t)(w >> 16) * P.pitch + (w & 0xFFFFu))] = 1;
                }
                if (sub == 0) S.status[so + rank] = key;
            }
        }
        while (bigm) {                                            // long regions, one after the other, 32 lanes x 4 entries in flight
            const int src = __ffs(bigm) - 1; bigm &= bigm - 1;
            const int brank = __shfl_sync(0xffffffffu, rank, src), bsize = __shfl_sync(0xffffffffu, size, src);
            const unsigned * bl = pool + __shfl_sync(0xffffffffu, lhead, src);
            const unsigned bkey = (epoch << 24) | (unsigned)brank;
            bool m2 = true;
            for (int i = lane; i < bsize && m2; i += 128) {
                unsigned w[4], o[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) w[u] = (i + 32 * u < bsize) ? __ldcg(bl + i + 32 * u) : 0u;
#pragma unroll
                for (int u = 0; u < 4; ++u) o[u] = (i + 32 * u < bsize) ? __ldcg(owner + (size_t)(w[u] >> 16) * P.pitch + (w[u] & 0xFFFFu)) : bkey;
#pragma unroll
                for (int u = 0; u < 4; ++u) m2 &= (o[u] == bkey);
            }
            if (!__all_sync(0xffffffffu, m2)) { if (lane == src) { ++live; first = min(first, brank); } continue; }
            for (int i = lane; i < bsize; i += 32) {
                const unsigned w = __ldcg(bl + i);
                vis[2 * ((size_t)(w >> 16) * P.pitch + (w & 0xFFFFu))] = 1;
            }
            if (lane == src) S.status[so + brank] = bkey;
        }
    }
    if (sub != 0) live = 0;
    if (live) { atomicAdd(&C.win_live, live); atomicMin(&C.win_first, first); }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) s_last = (atomicAdd(&C.win_blocks, 1) == (int)gridDim.x - 1);
    __syncthreads();
    if (s_last && threadIdx.x == 0) {
        __threadfence();
        const int failed = atomicAdd(&C.win_live, 0);                 // window seeds that stay live (out-prioritised this round)
        const int firstf = atomicMin(&C.win_first, 0x7fffffff);
        // seeds flooded per round: small at first — the strongest seeds own the long lines, and every other seed on such a line
        // would flood the whole line again, only to lose against the line's strongest seed — then doubling up to the steady size
        const long long grown = (long long)S.win_first << min(round + 1, 16);
        long long steady = max(S.win_min, n / S.win_div);
        // late rounds: the seeds left are weak, their regions rarely touch, and every round ends with the tail of its slowest blob:
        // the window doubles again from round win_late on, so that the rest of the list takes two or three rounds instead of five
        if (round + 1 >= S.win_late) steady <<= min(round + 2 - S.win_late, 8);
        const int budget = (int)min(steady, grown);
        const int nlo = failed ? min(firstf, hi) : hi;
        const int nhi = min(n, hi + max(0, budget - failed));
        C.win_lo = nlo; C.win_hi = nhi;
        C.wl_n = 0;
        C.win_live = 0; C.win_first = 0x7fffffff; C.win_blocks = 0; C.pool_cursor = 0; C.esc_n = 0; C.esc_q = 0; C.esc2_n = 0; C.esc2_q = 0;
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
    uint8_t * vis = reinterpret_cast<uint8_t *>(P.mv + pbase) + 1;      // the visited byte of pixel p is vis[2 * p], its mask byte vis[2 * p - 1]
    const float2 * dxy = P.dxy + pbase; const uint8_t * mask8 = reinterpret_cast<const uint8_t *>(P.mv + pbase);
    unsigned * q = K.px + (size_t)slot * K.cap_px;             // not in use yet at this stage

    for (int rank = 0; rank < n; ++rank) {
        if (__ldcg(status + rank) != kStatusLive) continue;
        const unsigned key = (epoch_serial << 24) | (unsigned)rank;
        const unsigned spx = S.idx[so + rank];
        const int sr = spx / P.w, sc = spx - sr * P.w;
        if (__ldcg(vis + 2 * ((size_t)sr * P.pitch + sc))) {
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
            if (elig && !first) elig = (__ldg(mask8 + 2 * po) >> b) & 1;
            if (elig && !first) {
                const float2 gq = __ldg(dxy + po);
                const float g = fabsf(__fadd_rn(__fmul_rn(gq.x, sb), __fmul_rn(gq.y, cb)));
                elig = g > thr;
            }
            if (elig) elig = (__ldcg(vis + 2 * po) == 0) && (__ldcg(owner + po) != key);
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
            __stcg(vis + 2 * ((size_t)(w >> 16) * P.pitch + (w & 0xFFFFu)), (uint8_t)1);
        }
        if (lane == 0) { S.size[so + rank] = tail; __stcg(status + rank, key); }
        __syncwarp();
    }
    if (lane == 0) { C.flood_overflow = 0; C.flood_done = 1; }
}

} // namespace

int launch_flood_rounds(const Planes & P, const Seeds & S, const Comps & K, FrameCtrl * ctrl, int n_slots, int seeds_hint,
                         int round_begin, int round_end, unsigned epoch_first, cudaStream_t st)
{
    // flood: a resident set of CTAs per frame, warps pull seeds from the frame's queue; verify: warps striding over the window
    // register budget of the flood kernel: 4 resident CTAs per SM at 128 registers, a tile's 32 rows loaded in one batch (default), or
    // LRB_FLOOD_MINB=6: 6 CTAs at 80 registers, two batches of 16 rows.  3, 5, 7 and 8 CTAs were measured too and pruned from the build:
    // every smaller budget is slower (profiles/r2c_experiments.md)
    static const int minb = [] { const char * v = getenv("LRB_FLOOD_MINB"); return v ? atoi(v) : 4; }();
    auto kern = (minb == 6) ? k_flood_round<6, 16> : k_flood_round<4, 32>;
    static PerDeviceOnce once; static int per_sm_d[PerDeviceOnce::kMaxDev], n_sm_d[PerDeviceOnce::kMaxDev];
    const int d = once.run([&] {
        const int dv = PerDeviceOnce::dev();
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm_d[dv], kern, FL_WARPS * 32, 0) != cudaSuccess || per_sm_d[dv] < 1) per_sm_d[dv] = 4;
        n_sm_d[dv] = 148; cudaDeviceGetAttribute(&n_sm_d[dv], cudaDevAttrMultiProcessorCount, dv);
    });
    const int per_sm = per_sm_d[d], n_sm = n_sm_d[d];
    // LRB_FLOOD_SUB=k floods k frames at a time (all rounds of one sub-batch, then the next) to keep their planes L2
    // resident.  Measured on B200 (gpurun_out/sweep4.log): slower at every k < n_slots — a round is bound by its longest
    // flood chain, not by bandwidth, so more frames in flight win — hence the default is one batch of all slots.
    static const int sub_env = [] { const char * v = getenv("LRB_FLOOD_SUB"); return v ? atoi(v) : 0; }();
    int sub = sub_env > 0 ? sub_env : n_slots;
    sub = max(1, min(sub, n_slots));
    int launches = 0;
    for (int s0 = 0; s0 < n_slots; s0 += sub) {
        const int ns = min(sub, n_slots - s0);
        const int fblocks = max(1, min((seeds_hint + FL_WARPS - 1) / FL_WARPS, (n_sm * per_sm + ns - 1) / ns));
        dim3 gf(fblocks, ns), gv(148 / max(1, min(ns, 4)) + 1, ns);
        static const bool prune = getenv("LRB_FLOOD_NO_PRUNE") == nullptr;
        dim3 gp(max(1, min((seeds_hint + 255) / 256, 64)), ns);
        LRB_CARVEOUT(kern); LRB_CARVEOUT(k_flood_verify); LRB_CARVEOUT(k_flood_prune);
        // the sub-warp passes take four / two seeds per warp
        const int sblocks = max(1, min((seeds_hint + 4 * FL_WARPS - 1) / (4 * FL_WARPS), (n_sm * 8 + ns - 1) / ns));
        const int mblocks = max(1, min((seeds_hint + 2 * FL_WARPS - 1) / (2 * FL_WARPS), (n_sm * 6 + ns - 1) / ns / 2));
        dim3 gs(sblocks, ns), gm(mblocks, ns);
        LRB_CARVEOUT((k_flood_small<8, 8>)); LRB_CARVEOUT((k_flood_small<16, 6>));
        static const int small_mode = [] { const char * v = getenv("LRB_FLOOD_SMALL_MODE"); return v ? atoi(v) : 0; }();
        for (int r = round_begin; r < round_end; ++r) {
            if (prune && r > 0) { k_flood_prune<<<gp, 256, 0, st>>>(P, S, ctrl, r, s0); ++launches; }
            static const bool trace = getenv("LRB_TRACE") != nullptr;      // diagnosis: synchronise and report after every kernel of a round
#define LRB_TRACE_SYNC(what) do { if (trace) { fprintf(stderr, "[lrb] flood round %d: %s ...", r, what); fflush(stderr); cudaError_t te = cudaStreamSynchronize(st); fprintf(stderr, " %s\n", cudaGetErrorString(te)); fflush(stderr); } } while (0)
            // LRB_FLOOD_MID=1 adds the 16-lane pass (512-pixel windows) between the two: measured on B200 it finishes only ~30 % of
            // what the 8-lane pass escalates (those floods are long lines, not slightly larger blobs) and costs 7 % of the job
            // throughput (profiles/r2_flood_regions.md), so it is off by default
            static const bool mid = [] { const char * v = getenv("LRB_FLOOD_MID"); return v && atoi(v) != 0; }();
            // from the second round on the first pass floods the prune pass's list of survivors (not with the middle pass: it owns S.esc)
            const int list_mode = (prune && r > 0 && !mid) ? 8 : 0;
            k_flood_small<8, 8><<<gs, FL_WARPS * 32, 0, st>>>(P, S, K, ctrl, r, epoch_first, s0, small_mode | list_mode, mid ? 0 : 1);
            LRB_TRACE_SYNC("sub-warp pass (8 lanes)");
            if (mid) {
                k_flood_small<16, 6><<<gm, FL_WARPS * 32, 0, st>>>(P, S, K, ctrl, r, epoch_first, s0, small_mode, 1);
                LRB_TRACE_SYNC("sub-warp pass (16 lanes)");
                ++launches;
            }
            kern<<<gf, FL_WARPS * 32, 0, st>>>(P, S, K, ctrl, r, epoch_first, s0);
            LRB_TRACE_SYNC("full-tile pass");
            k_flood_verify<<<gv, 256, 0, st>>>(P, S, K, ctrl, r, epoch_first, s0);
            LRB_TRACE_SYNC("verify");
#undef LRB_TRACE_SYNC
            launches += 3;
        }
    }
    return launches;                                                        // kernels launched (the engine counts them)
}

void launch_flood_serial(const Planes & P, const Seeds & S, const Comps & K, FrameCtrl * ctrl, int n_slots,
                         unsigned epoch_serial, cudaStream_t st)
{
    k_flood_serial<<<n_slots, 32, 0, st>>>(P, S, K, ctrl, epoch_serial);
}

} // namespace lrb
