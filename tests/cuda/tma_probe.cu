// Standalone probe of the inline-PTX primitives used by csrc/mia_tile.cuh (debug aid, not a product path).
// usage: tma_probe <stage>   1 = mbarrier only, 2 = TMA box load, 3 = scalar helpers
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "../../medicalimageanalysis_b200/csrc/mia_tile.cuh"
using namespace mia;

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("FAIL %s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

__global__ void k_mbar(int *out)
{
    __shared__ __align__(8) unsigned long long bar;
    const uint32_t b = smem_u32(&bar);
    if (threadIdx.x == 0) {
        mbar_init(b, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) mbar_arrive(b);
    mbar_wait<20>(b, 0);
    out[threadIdx.x] = threadIdx.x;
}

static int BX = 40, BY = 18, BZ = 16;   // host-side; kernels take the element count as an argument
__global__ void k_tma(const __grid_constant__ CUtensorMap map, short *out, int x, int y, int z, int nelem)
{
    extern __shared__ __align__(1024) unsigned char sm[];
    __shared__ __align__(8) unsigned long long bar;
    const uint32_t b = smem_u32(&bar);
    if (threadIdx.x == 0) {
        mbar_init(b, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        mbar_arrive_expect_tx(b, nelem * 2);
        tma_load_3d(smem_u32(sm), &map, b, x, y, z);
    }
    mbar_wait<20>(b, 0);
    const short *s = reinterpret_cast<const short *>(sm);
    for (int i = threadIdx.x; i < nelem; i += blockDim.x) out[i] = s[i];
}

__global__ void k_scalar(float *f, short *s)
{
    __shared__ short sh[64];
    sh[threadIdx.x] = (short)(threadIdx.x * 100 - 3000);
    __syncthreads();
    const float l = -3.25f + 0.37f * threadIdx.x;
    const float t = add_rm(l, kMagic);
    f[threadIdx.x * 4 + 0] = t - kMagic;
    f[threadIdx.x * 4 + 1] = l - (t - kMagic);
    f[threadIdx.x * 4 + 2] = Tap<int16_t, 2>::ld(smem_u32(sh) + 2 * (threadIdx.x % 60));
    f[threadIdx.x * 4 + 3] = (float)(__float_as_uint(t) - kMagicBits);
    st_out(s + threadIdx.x * 2, tile_to_out<int16_t, true>(l * 9000.0f));
    st_out(s + threadIdx.x * 2 + 1, tile_to_out<int16_t, false>(l * 9000.0f));
}

// variant: descriptor read from global memory instead of a __grid_constant__ parameter
__global__ void k_tma_g(const CUtensorMap *map, short *out, int x, int y, int z, int nelem)
{
    extern __shared__ __align__(1024) unsigned char sm[];
    __shared__ __align__(8) unsigned long long bar;
    const uint32_t b = smem_u32(&bar);
    if (threadIdx.x == 0) {
        mbar_init(b, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        mbar_arrive_expect_tx(b, nelem * 2);
        tma_load_3d(smem_u32(sm), map, b, x, y, z);
    }
    mbar_wait<20>(b, 0);
    const short *s = reinterpret_cast<const short *>(sm);
    for (int i = threadIdx.x; i < nelem; i += blockDim.x) out[i] = s[i];
}

// variant: plain 1-D bulk copy (no tensor map)
__global__ void k_bulk(const short *src, short *out, int bytes)
{
    extern __shared__ __align__(1024) unsigned char sm[];
    __shared__ __align__(8) unsigned long long bar;
    const uint32_t b = smem_u32(&bar);
    if (threadIdx.x == 0) {
        mbar_init(b, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        mbar_arrive_expect_tx(b, bytes);
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(smem_u32(sm)), "l"(src), "r"(bytes), "r"(b) : "memory");
    }
    mbar_wait<20>(b, 0);
    const short *s = reinterpret_cast<const short *>(sm);
    for (int i = threadIdx.x; i < bytes / 2; i += blockDim.x) out[i] = s[i];
}

int main(int argc, char **argv)
{
    const int stage = argc > 1 ? atoi(argv[1]) : 1;
    if (argc > 4) { BX = atoi(argv[2]); BY = atoi(argv[3]); BZ = atoi(argv[4]); }
    if (stage == 1) {
        int *d; CK(cudaMalloc(&d, 64 * 4));
        k_mbar<<<1, 64>>>(d);
        CK(cudaDeviceSynchronize());
        printf("stage1 ok\n");
    } else if (stage == 2 || stage == 4 || stage == 5) {
        const int nx = 64, ny = 40, nz = 24;
        std::vector<short> h((size_t)nx * ny * nz);
        for (size_t i = 0; i < h.size(); ++i) h[i] = (short)(i % 30011 - 15000);
        short *dv, *dout;
        CK(cudaMalloc(&dv, h.size() * 2)); CK(cudaMalloc(&dout, BX * BY * BZ * 2));
        CK(cudaMemcpy(dv, h.data(), h.size() * 2, cudaMemcpyHostToDevice));
        void *fn = nullptr; cudaDriverEntryPointQueryResult q;
        CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q));
        typedef CUresult (*Enc)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
        alignas(64) CUtensorMap map;
        const cuuint64_t gd[3] = {nx, ny, nz}, gs[2] = {nx * 2, (cuuint64_t)nx * ny * 2};
        const cuuint32_t box[3] = {(cuuint32_t)BX, (cuuint32_t)BY, (cuuint32_t)BZ}, es[3] = {1, 1, 1};
        CUresult r = ((Enc)fn)(&map, CU_TENSOR_MAP_DATA_TYPE_UINT16, 3, dv, gd, gs, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                               CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        printf("encode rc=%d\n", (int)r);
        if (r != CUDA_SUCCESS) return 1;
        CK(cudaFuncSetAttribute(k_tma, cudaFuncAttributeMaxDynamicSharedMemorySize, BX * BY * BZ * 2));
        if (stage == 5) {
            k_bulk<<<1, 128, BX * BY * BZ * 2>>>(dv, dout, 4096);
            CK(cudaDeviceSynchronize());
            std::vector<short> o(2048);
            CK(cudaMemcpy(o.data(), dout, 4096, cudaMemcpyDeviceToHost));
            long bad = 0;
            for (int i = 0; i < 2048; ++i) bad += o[i] != h[i];
            printf("stage5 bulk: %ld mismatches\n", bad);
            return 0;
        }
        CUtensorMap *dmap = nullptr;
        CK(cudaMalloc(&dmap, sizeof(CUtensorMap)));
        CK(cudaMemcpy(dmap, &map, sizeof(CUtensorMap), cudaMemcpyHostToDevice));
        for (int t = 0; t < 3; ++t) {
            const int x = t == 0 ? 3 : (t == 1 ? -2 : 40), y = t == 0 ? 5 : (t == 1 ? -1 : 30), z = t == 0 ? 2 : (t == 1 ? -3 : 20);
            if (stage == 4) k_tma_g<<<1, 128, BX * BY * BZ * 2>>>(dmap, dout, x, y, z, BX * BY * BZ);
            else k_tma<<<1, 128, BX * BY * BZ * 2>>>(map, dout, x, y, z, BX * BY * BZ);
            CK(cudaDeviceSynchronize());
            std::vector<short> o(BX * BY * BZ);
            CK(cudaMemcpy(o.data(), dout, o.size() * 2, cudaMemcpyDeviceToHost));
            long bad = 0;
            for (int c = 0; c < BZ; ++c) for (int b = 0; b < BY; ++b) for (int a = 0; a < BX; ++a) {
                const int gx = x + a, gy = y + b, gz = z + c;
                const short want = (gx < 0 || gy < 0 || gz < 0 || gx >= nx || gy >= ny || gz >= nz) ? 0 : h[((size_t)gz * ny + gy) * nx + gx];
                if (o[(c * BY + b) * BX + a] != want) ++bad;
            }
            printf("stage2 box (%d,%d,%d): %ld mismatches\n", x, y, z, bad);
        }
    } else {
        float *f; short *s;
        CK(cudaMalloc(&f, 64 * 16)); CK(cudaMalloc(&s, 64 * 4));
        k_scalar<<<1, 64>>>(f, s);
        CK(cudaDeviceSynchronize());
        std::vector<float> hf(256); std::vector<short> hs(128);
        CK(cudaMemcpy(hf.data(), f, 1024, cudaMemcpyDeviceToHost)); CK(cudaMemcpy(hs.data(), s, 256, cudaMemcpyDeviceToHost));
        for (int i = 0; i < 12; ++i)
            printf("l=%g floor=%g frac=%g tap=%g idx=%g  vtk=%d itk=%d\n", -3.25f + 0.37f * i, hf[i * 4], hf[i * 4 + 1], hf[i * 4 + 2], hf[i * 4 + 3], hs[i * 2], hs[i * 2 + 1]);
        printf("stage3 ok\n");
    }
    return 0;
}
