// Second probe: the CUDA programming guide's own 2-D TMA example (libcu++ wrappers), plus a descriptor dump.
#include <cstdio>
#include <vector>
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda/barrier>
using barrier = cuda::barrier<cuda::thread_scope_block>;
namespace cde = cuda::device::experimental;
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("FAIL %s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)
constexpr int GW = 256, GH = 64, SW = 32, SH = 8;

__global__ void kernel(const __grid_constant__ CUtensorMap tensor_map, int x, int y, int *out)
{
    __shared__ alignas(128) int smem_buffer[SH][SW];
#pragma nv_diag_suppress static_var_with_dynamic_init
    __shared__ barrier bar;
    if (threadIdx.x == 0) {
        init(&bar, blockDim.x);
        cde::fence_proxy_async_shared_cta();
    }
    __syncthreads();
    barrier::arrival_token token;
    if (threadIdx.x == 0) {
        cde::cp_async_bulk_tensor_2d_global_to_shared(&smem_buffer, &tensor_map, x, y, bar);
        token = cuda::device::barrier_arrive_tx(bar, 1, sizeof(smem_buffer));
    } else {
        token = bar.arrive();
    }
    bar.wait(std::move(token));
    for (int i = threadIdx.x; i < SH * SW; i += blockDim.x) out[i] = smem_buffer[i / SW][i % SW];
}

constexpr int BX3 = 40, BY3 = 18, BZ3 = 16;
__global__ void kernel3(const __grid_constant__ CUtensorMap tensor_map, int x, int y, int z, short *out)
{
    __shared__ alignas(128) short smem_buffer[BZ3 * BY3 * BX3];
#pragma nv_diag_suppress static_var_with_dynamic_init
    __shared__ barrier bar;
    if (threadIdx.x == 0) {
        init(&bar, blockDim.x);
        cde::fence_proxy_async_shared_cta();
    }
    __syncthreads();
    barrier::arrival_token token;
    if (threadIdx.x == 0) {
        cde::cp_async_bulk_tensor_3d_global_to_shared(&smem_buffer, &tensor_map, x, y, z, bar);
        token = cuda::device::barrier_arrive_tx(bar, 1, sizeof(smem_buffer));
    } else {
        token = bar.arrive();
    }
    bar.wait(std::move(token));
    for (int i = threadIdx.x; i < BZ3 * BY3 * BX3; i += blockDim.x) out[i] = smem_buffer[i];
}

int main3(int l2)
{
    const int nx = 64, ny = 40, nz = 24;
    std::vector<short> h((size_t)nx * ny * nz);
    for (size_t i = 0; i < h.size(); ++i) h[i] = (short)(i % 30011 - 15000);
    short *d, *o;
    CK(cudaMalloc(&d, h.size() * 2)); CK(cudaMalloc(&o, BZ3 * BY3 * BX3 * 2));
    CK(cudaMemcpy(d, h.data(), h.size() * 2, cudaMemcpyHostToDevice));
    void *fn = nullptr; cudaDriverEntryPointQueryResult q;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q));
    typedef CUresult (*Enc)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                            const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                            CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    CUtensorMap map{};
    cuuint64_t size[3] = {nx, ny, nz}, stride[2] = {nx * 2, (cuuint64_t)nx * ny * 2};
    cuuint32_t box[3] = {BX3, BY3, BZ3}, es[3] = {1, 1, 1};
    CUresult r = ((Enc)fn)(&map, CU_TENSOR_MAP_DATA_TYPE_UINT16, 3, d, size, stride, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                           CU_TENSOR_MAP_SWIZZLE_NONE, l2 ? CU_TENSOR_MAP_L2_PROMOTION_L2_128B : CU_TENSOR_MAP_L2_PROMOTION_NONE,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode3 rc=%d l2=%d\n", (int)r, l2);
    const unsigned char *b = reinterpret_cast<const unsigned char *>(&map);
    for (int i = 0; i < 128; ++i) printf("%02x%s", b[i], (i % 32 == 31) ? "\n" : " ");
    kernel3<<<1, 128>>>(map, 3, 5, 2, o);
    CK(cudaDeviceSynchronize());
    std::vector<short> ho(BZ3 * BY3 * BX3);
    CK(cudaMemcpy(ho.data(), o, ho.size() * 2, cudaMemcpyDeviceToHost));
    long bad = 0;
    for (int c = 0; c < BZ3; ++c) for (int bb = 0; bb < BY3; ++bb) for (int a = 0; a < BX3; ++a)
        bad += ho[(c * BY3 + bb) * BX3 + a] != h[((size_t)(2 + c) * ny + 5 + bb) * nx + 3 + a];
    printf("3-D u16 libcu++: %ld mismatches\n", bad);
    return 0;
}

int main(int argc, char **argv)
{
    if (argc > 1) return main3(argv[1][0] == '1');

    int dev = 0, drv = 0, rt = 0;
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, dev));
    cudaDriverGetVersion(&drv); cudaRuntimeGetVersion(&rt);
    printf("%s cc %d.%d driver %d runtime %d\n", prop.name, prop.major, prop.minor, drv, rt);
    std::vector<int> h(GW * GH);
    for (int i = 0; i < GW * GH; ++i) h[i] = i;
    int *d, *o;
    CK(cudaMalloc(&d, h.size() * 4)); CK(cudaMalloc(&o, SH * SW * 4));
    CK(cudaMemcpy(d, h.data(), h.size() * 4, cudaMemcpyHostToDevice));
    void *fn = nullptr; cudaDriverEntryPointQueryResult q;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q));
    printf("entry point %p query %d\n", fn, (int)q);
    typedef CUresult (*Enc)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                            const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                            CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    CUtensorMap map{};
    cuuint64_t size[2] = {GW, GH}, stride[1] = {GW * sizeof(int)};
    cuuint32_t box[2] = {SW, SH}, es[2] = {1, 1};
    CUresult r = ((Enc)fn)(&map, CU_TENSOR_MAP_DATA_TYPE_INT32, 2, d, size, stride, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                           CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode rc=%d, map at %p\n", (int)r, (void *)&map);
    const unsigned char *b = reinterpret_cast<const unsigned char *>(&map);
    for (int i = 0; i < 128; ++i) printf("%02x%s", b[i], (i % 32 == 31) ? "\n" : " ");
    kernel<<<1, 128>>>(map, 32, 8, o);
    CK(cudaDeviceSynchronize());
    std::vector<int> ho(SH * SW);
    CK(cudaMemcpy(ho.data(), o, ho.size() * 4, cudaMemcpyDeviceToHost));
    long bad = 0;
    for (int i = 0; i < SH * SW; ++i) bad += ho[i] != h[(8 + i / SW) * GW + 32 + i % SW];
    printf("guide example: %ld mismatches\n", bad);
    return 0;
}
