// mt_stream.cu — the reference's random uniform stream, generated once per seed.
//
// The reference draws its RANSAC samples with std::mt19937 + std::uniform_real_distribution<float>
// (reference src/estimator.h:35, src/math_utils.cpp:25-28; libstdc++ generate_canonical<float,24>: one 32-bit output
// per uniform, u = float(raw) * 2^-32 with u >= 1 replaced by nextafter(1,0)).  The stream depends on the seed only,
// not on the image, so all frames share one device-resident copy of k = 2^32 * u stored as an exact 32-bit integer
// (float(raw) keeps 24 significant bits, so k is a multiple of 256 and at most 2^32 - 256).
//
// MT19937's 624-word state advances by a "twist" whose element i reads elements i, i+1 and i+397 (mod 624): elements
// [0,227) read only old values, [227,454) read new values produced by the first group, [454,624) new values of the
// second — three data-parallel phases per 624 outputs.  One 256-thread CTA runs the recurrence out of shared memory;
// this is a one-off cost per seed (a few ms for 2^25 draws), off the per-frame path.
#include "common.cuh"
#include "kernels.h"

namespace lrb {

namespace {

__device__ __forceinline__ unsigned mt_twist(unsigned cur, unsigned nxt, unsigned far)
{
    const unsigned y = (cur & 0x80000000u) | (nxt & 0x7fffffffu);
    return far ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
}

__device__ __forceinline__ unsigned canonical_k(unsigned raw)
{
    const float f = __uint2float_rn(raw);                       // round to nearest even at 24 bits, as float(raw) does
    unsigned long long k = __float2ull_rz(f);
    if (k >= (1ull << 32)) k = (1ull << 32) - 256;              // u == 1.0f -> nextafter(1, 0)
    return (unsigned)k;
}

__global__ void __launch_bounds__(256) k_mt_stream(unsigned seed, unsigned * __restrict__ out, unsigned long long n, int raw)
{
    __shared__ unsigned x[624];
    const int t = threadIdx.x;
    if (t == 0) {                                                // std::mt19937 seeding recurrence (sequential, once)
        unsigned v = seed; x[0] = v;
        for (int i = 1; i < 624; ++i) { v = 1812433253u * (v ^ (v >> 30)) + (unsigned)i; x[i] = v; }
    }
    __syncthreads();
    for (unsigned long long base = 0; base < n; base += 624) {
        // every phase reads all its operands, meets at a barrier and only then writes: element i's "next" operand
        // x[i+1] belongs to the neighbouring thread (possibly of another warp) and must still hold its old value
        // phase 1: i in [0, 227): all operands old
        unsigned v = 0;
        if (t < 227) v = mt_twist(x[t], x[t + 1], x[t + 397]);
        __syncthreads();
        if (t < 227) x[t] = v;
        __syncthreads();
        // phase 2: i in [227, 454): far operand x[i-227] is new
        if (t < 227) { const int i = 227 + t; v = mt_twist(x[i], x[i + 1], x[i - 227]); }
        __syncthreads();
        if (t < 227) x[227 + t] = v;
        __syncthreads();
        // phase 3: i in [454, 624): far operand new; element 623 wraps to the new x[0]
        if (t < 170) { const int i = 454 + t; v = mt_twist(x[i], i == 623 ? x[0] : x[i + 1], x[i - 227]); }
        __syncthreads();
        if (t < 170) x[454 + t] = v;
        __syncthreads();
        for (int i = t; i < 624; i += 256) {
            unsigned y = x[i];
            y ^= y >> 11; y ^= (y << 7) & 0x9d2c5680u; y ^= (y << 15) & 0xefc60000u; y ^= y >> 18;
            if (base + i < n) out[base + i] = raw ? y : canonical_k(y);
        }
        __syncthreads();
    }
}

__global__ void k_fill_u32(unsigned * p, size_t n, unsigned v)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = v;
}

// dense copy of frames whose layout TMA cannot address directly (row pitch not a multiple of 16 bytes, unaligned base)
__global__ void k_ingest(const float * __restrict__ src, long long src_pitch, long long src_frame_stride,
                         float * __restrict__ dst, int dst_pitch, size_t dst_stride, int w, int h)
{
    const int slot = blockIdx.z;
    const int y = blockIdx.y;
    const float * s = src + (long long)slot * src_frame_stride + (long long)y * src_pitch;
    float * d = dst + slot * dst_stride + (size_t)y * dst_pitch;
    for (int x = blockIdx.x * blockDim.x + threadIdx.x; x < w; x += gridDim.x * blockDim.x) d[x] = s[x];
}

// 8-bit ingest (what the reference's demo does on the CPU before calling the library):
//   channels == 1: gray / 256                      cv::Mat::convertTo(CV_32F, 1/256.0), reference src/autorectify.cpp:119
//   channels == 3: BGR -> gray first, OpenCV's 8-bit fixed point (b*3735 + g*19235 + r*9798 + 2^14) >> 15
//                  (cv::cvtColor COLOR_BGR2GRAY as the demo calls it; pinned against cv2 in tests/golden/ingest_u8.npz)
// x / 256 is exact in float, so the result is bit-identical to the float frame the CPU path would have passed.
template <int CH>
__global__ void k_ingest_u8(const unsigned char * __restrict__ src, long long src_pitch, long long src_frame_stride,
                            float * __restrict__ dst, int dst_pitch, size_t dst_stride, int w, int h)
{
    const int slot = blockIdx.z;
    const int y = blockIdx.y;
    const unsigned char * s = src + (long long)slot * src_frame_stride + (long long)y * src_pitch;
    float * d = dst + slot * dst_stride + (size_t)y * dst_pitch;
    for (int x = blockIdx.x * blockDim.x + threadIdx.x; x < w; x += gridDim.x * blockDim.x) {
        unsigned g;
        if (CH == 1) g = s[x];
        else g = ((unsigned)s[3 * x] * 3735u + (unsigned)s[3 * x + 1] * 19235u + (unsigned)s[3 * x + 2] * 9798u + (1u << 14)) >> 15;
        d[x] = __fmul_rn((float)g, 0.00390625f);
    }
}

} // namespace

void launch_ingest_u8(const unsigned char * src, long long src_pitch, long long src_frame_stride, int channels, float * dst,
                      int dst_pitch, size_t dst_stride, int w, int h, int n_slots, cudaStream_t st)
{
    dim3 grid((w + 255) / 256, h, n_slots);
    if (channels == 1) k_ingest_u8<1><<<grid, 256, 0, st>>>(src, src_pitch, src_frame_stride, dst, dst_pitch, dst_stride, w, h);
    else               k_ingest_u8<3><<<grid, 256, 0, st>>>(src, src_pitch, src_frame_stride, dst, dst_pitch, dst_stride, w, h);
}

void launch_mt_stream(unsigned seed, unsigned * stream, unsigned long long n, cudaStream_t st)
{
    k_mt_stream<<<1, 256, 0, st>>>(seed, stream, n, 0);
}

// the generator's raw 32-bit outputs (std::mt19937::operator()), for std::uniform_int_distribution (k_hough.cu)
void launch_mt_raw(unsigned seed, unsigned * out, unsigned long long n, cudaStream_t st)
{
    k_mt_stream<<<1, 256, 0, st>>>(seed, out, n, 1);
}

void launch_fill_u32(unsigned * p, size_t n, unsigned v, cudaStream_t st)
{
    k_fill_u32<<<148 * 8, 256, 0, st>>>(p, n, v);
}

void launch_ingest(const float * src, long long src_pitch, long long src_frame_stride, float * dst, int dst_pitch,
                   size_t dst_stride, int w, int h, int n_slots, cudaStream_t st)
{
    dim3 grid((w + 255) / 256, h, n_slots);
    k_ingest<<<grid, 256, 0, st>>>(src, src_pitch, src_frame_stride, dst, dst_pitch, dst_stride, w, h);
}

} // namespace lrb
