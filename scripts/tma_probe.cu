// standalone probe: which TMA box shapes work on this box (debug aid, not part of the product)
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "../librectify_b200/csrc/tma.cuh"
using namespace lrb;

template <int BW, int BH>
__global__ void probe(const __grid_constant__ CUtensorMap tm, float * out, int x, int y)
{
    __shared__ __align__(128) float tile[BW * BH];
    __shared__ __align__(8) uint64_t bar;
    if (threadIdx.x == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
    __syncthreads();
    if (threadIdx.x == 0) { mbar_arrive_expect_tx(&bar, BW * BH * 4); tma_load_3d(tile, &tm, x, y, 0, &bar); }
    mbar_wait(&bar, 0);
    for (int i = threadIdx.x; i < BW * BH; i += blockDim.x) out[i] = tile[i];
}

typedef CUresult (*Fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static int gx = 0, gy = 0;
template <int BW, int BH> int run(Fn fn, float * img, int w, int h, int pitch, float * out)
{
    CUtensorMap tm;
    cuuint64_t dims[3] = {(cuuint64_t)w, (cuuint64_t)h, 1}; cuuint64_t str[2] = {(cuuint64_t)pitch * 4, (cuuint64_t)pitch * 4 * h};
    cuuint32_t box[3] = {BW, BH, 1}, es[3] = {1, 1, 1};
    CUresult r = fn(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, img, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("box %dx%d: encode failed %d\n", BW, BH, (int)r); return 1; }
    probe<BW, BH><<<1, 128>>>(tm, out, gx, gy);
    cudaError_t e = cudaDeviceSynchronize();
    std::vector<float> hbuf(BW * BH);
    cudaMemcpy(hbuf.data(), out, BW * BH * 4, cudaMemcpyDeviceToHost);
    printf("box %dx%d (%d B): %s  tile[0]=%g tile[2*BW+2]=%g (expect 0, 0) tile[3*BW+5]=%g (expect %d)\n", BW, BH, BW * BH * 4, cudaGetErrorString(e), hbuf[0], hbuf[2 * BW + 2], hbuf[3 * BW + 5], 195 * pitch + 275);
    return e != cudaSuccess;
}

int main(int argc, char ** argv)
{
    if (argc > 2) { gx = atoi(argv[1]); gy = atoi(argv[2]); }
    printf("coords %d %d\n", gx, gy);
    void * p = nullptr; cudaDriverEntryPointQueryResult q;
    cudaFree(0);
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || !p) { printf("no entry point\n"); return 1; }
    Fn fn = (Fn)p;
    const int w = 320, h = 240, pitch = 320;
    float * img, * out; cudaMalloc(&img, pitch * h * 4); cudaMalloc(&out, 1 << 20);
    std::vector<float> hi(pitch * h); for (int i = 0; i < pitch * h; ++i) hi[i] = (float)i;
    cudaMemcpy(img, hi.data(), pitch * h * 4, cudaMemcpyHostToDevice);
    if (run<32, 8>(fn, img, w, h, pitch, out)) return 0;
    if (run<64, 36>(fn, img, w, h, pitch, out)) return 0;
    if (run<128, 36>(fn, img, w, h, pitch, out)) return 0;
    if (run<132, 36>(fn, img, w, h, pitch, out)) return 0;
    if (run<132, 68>(fn, img, w, h, pitch, out)) return 0;
    return 0;
}
