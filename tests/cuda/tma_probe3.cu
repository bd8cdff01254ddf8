// Third probe: rank x dtype matrix for cp.async.bulk.tensor (raw PTX), tiny boxes.
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("FAIL %s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

template <int RANK>
__global__ void k(const __grid_constant__ CUtensorMap map, unsigned char *out, int bytes, int x, int y, int z)
{
    __shared__ alignas(128) unsigned char buf[16384];
    __shared__ alignas(8) unsigned long long bar;
    const uint32_t b = (uint32_t)__cvta_generic_to_shared(&bar), dst = (uint32_t)__cvta_generic_to_shared(buf);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(b), "r"(1) : "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(bytes) : "memory");
        if (RANK == 2)
            asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                         ::"r"(dst), "l"(&map), "r"(x), "r"(y), "r"(b) : "memory");
        else
            asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                         ::"r"(dst), "l"(&map), "r"(x), "r"(y), "r"(z), "r"(b) : "memory");
    }
    asm volatile("{\n.reg .pred p;\nW:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n@p bra D;\nbra W;\nD:\n}\n" ::"r"(b) : "memory");
    for (int i = threadIdx.x; i < bytes; i += blockDim.x) out[i] = buf[i];
}

int main(int argc, char **argv)
{
    const int rank = atoi(argv[1]), es = atoi(argv[2]);           // es: 2 = u16, 4 = i32
    const int bx = atoi(argv[3]), by = atoi(argv[4]), bz = rank == 3 ? atoi(argv[5]) : 1;
    const int nx = 64, ny = 40, nz = 24;
    const int X = argc > 6 ? atoi(argv[6]) : 8;
    std::vector<unsigned char> h((size_t)nx * ny * nz * es);
    for (size_t i = 0; i < h.size(); ++i) h[i] = (unsigned char)(i * 7 + i / 251);
    unsigned char *d, *o;
    CK(cudaMalloc(&d, h.size())); CK(cudaMalloc(&o, 16384));
    CK(cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice));
    void *fn = nullptr; cudaDriverEntryPointQueryResult q;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q));
    typedef CUresult (*Enc)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                            const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                            CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    CUtensorMap map{};
    cuuint64_t size[3] = {nx, ny, nz}, stride[2] = {(cuuint64_t)nx * es, (cuuint64_t)nx * ny * es};
    cuuint32_t box[3] = {(cuuint32_t)bx, (cuuint32_t)by, (cuuint32_t)bz}, est[3] = {1, 1, 1};
    CUresult r = ((Enc)fn)(&map, es == 2 ? CU_TENSOR_MAP_DATA_TYPE_UINT16 : CU_TENSOR_MAP_DATA_TYPE_INT32, rank, d, size, stride, box, est,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    const int bytes = bx * by * bz * es;
    printf("x0=%d rank %d es %d box %dx%dx%d (%d B) encode rc=%d: ", X, rank, es, bx, by, bz, bytes, (int)r);
    if (r != CUDA_SUCCESS || bytes > 16384) { printf("skip\n"); return 1; }
    if (rank == 2) k<2><<<1, 128>>>(map, o, bytes, X, 5, 0); else k<3><<<1, 128>>>(map, o, bytes, X, 5, 2);
    CK(cudaDeviceSynchronize());
    std::vector<unsigned char> ho(bytes);
    CK(cudaMemcpy(ho.data(), o, bytes, cudaMemcpyDeviceToHost));
    long bad = 0;
    for (int c = 0; c < bz; ++c) for (int b = 0; b < by; ++b) for (int a = 0; a < bx * es; ++a)
        bad += ho[(size_t)(c * by + b) * bx * es + a] != h[(((size_t)((rank == 3 ? 2 : 0) + c) * ny + 5 + b) * nx + X) * es + a];
    printf("%ld mismatches\n", bad);
    return 0;
}
