#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "../librectify_b200/csrc/tma.cuh"
using namespace lrb;

__global__ void v_bar_only(float * out)
{
    __shared__ __align__(8) uint64_t bar;
    if (threadIdx.x == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
    __syncthreads();
    if (threadIdx.x == 0) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(smem_u32(&bar)) : "memory"); }
    mbar_wait(&bar, 0);
    out[threadIdx.x] = 1.f;
}

__global__ void v_bulk(const float * src, float * out)
{
    __shared__ __align__(128) float tile[256];
    __shared__ __align__(8) uint64_t bar;
    if (threadIdx.x == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
    __syncthreads();
    if (threadIdx.x == 0) {
        mbar_arrive_expect_tx(&bar, 1024);
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     :: "r"(smem_u32(tile)), "l"(src), "r"(1024), "r"(smem_u32(&bar)) : "memory");
    }
    mbar_wait(&bar, 0);
    out[threadIdx.x] = tile[threadIdx.x];
}

template <int RANK>
__global__ void v_tensor(const __grid_constant__ CUtensorMap tm, float * out, int cx)
{
    __shared__ __align__(128) float tile[32 * 8];
    __shared__ __align__(8) uint64_t bar;
    if (threadIdx.x == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
    __syncthreads();
    if (threadIdx.x == 0) {
        mbar_arrive_expect_tx(&bar, 1024);
        if (RANK == 3) tma_load_3d(tile, &tm, cx, cx, 0, &bar);
        else asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                          :: "r"(smem_u32(tile)), "l"(reinterpret_cast<uint64_t>(&tm)), "r"(0), "r"(0), "r"(smem_u32(&bar)) : "memory");
    }
    mbar_wait(&bar, 0);
    out[threadIdx.x] = tile[threadIdx.x];
}

typedef CUresult (*Fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main(int argc, char ** argv)
{
    int v = argc > 1 ? atoi(argv[1]) : 0;
    void * p = nullptr; cudaDriverEntryPointQueryResult q;
    cudaFree(0);
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
    Fn fn = (Fn)p;
    const int w = 320, h = 240, pitch = 320;
    float * img, * out; cudaMalloc(&img, pitch * h * 4); cudaMalloc(&out, 1 << 20);
    std::vector<float> hi(pitch * h); for (int i = 0; i < pitch * h; ++i) hi[i] = (float)i;
    cudaMemcpy(img, hi.data(), pitch * h * 4, cudaMemcpyHostToDevice);
    CUtensorMap tm; memset(&tm, 0, sizeof tm);
    int rank = (v >= 3) ? 3 : 2;
    cuuint64_t dims[3] = {(cuuint64_t)w, (cuuint64_t)h, 1}; cuuint64_t str[2] = {(cuuint64_t)pitch * 4, (cuuint64_t)pitch * 4 * h};
    cuuint32_t box[3] = {32, 8, 1}, es[3] = {1, 1, 1};
    CUresult r = fn(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, rank, img, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, (v == 4 ? CU_TENSOR_MAP_L2_PROMOTION_L2_128B : CU_TENSOR_MAP_L2_PROMOTION_NONE), CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("variant %d encode=%d q=%d: ", v, (int)r, (int)q);
    if (v == 0) v_bar_only<<<1, 128>>>(out);
    if (v == 1) v_bulk<<<1, 128>>>(img, out);
    if (v == 2) v_tensor<2><<<1, 128>>>(tm, out, 0);
    if (v == 3 || v == 4) v_tensor<3><<<1, 128>>>(tm, out, 0);
    if (v == 5) v_tensor<3><<<1, 128>>>(tm, out, -2);
    if (v == 6) v_tensor<3><<<1, 128>>>(tm, out, 300);
    cudaError_t e = cudaDeviceSynchronize();
    float hb[128]; cudaMemcpy(hb, out, 512, cudaMemcpyDeviceToHost);
    printf("%s out[0..3]=%g %g %g %g out[33]=%g\n", cudaGetErrorString(e), hb[0], hb[1], hb[2], hb[3], hb[33]);
    return 0;
}
