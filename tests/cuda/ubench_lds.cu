// LDS issue-rate micro-benchmark: loop-invariant addresses, one LOP3 per load.
#include <cstdio>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("FAIL %s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)
constexpr int ITERS = 8192;
// MODE: 0 = LDS.S16 lanes 1 short apart, 1 = LDS.S16 lanes 2 shorts apart, 2 = LDS.32 lanes 1 word apart,
//       3 = LDS.32 two lanes per word, 4 = LDS.U16 1 short apart, 5 = LDS.64 lanes 2 words apart, 6 = LDS.S16 with a 2-row split (lanes 16.. on next row, pitch 24 words)
template <int MODE>
__global__ void __launch_bounds__(256) k(int *out)
{
    __shared__ __align__(16) unsigned sh[8192 + 4096];
    for (int i = threadIdx.x; i < 8192; i += 256) sh[i] = i * 2654435761u;
    __syncthreads();
    const unsigned base = (unsigned)__cvta_generic_to_shared(sh);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned addr;
    if (MODE == 0 || MODE == 4) addr = base + warp * 512 + lane * 2;
    else if (MODE == 1) addr = base + warp * 512 + lane * 4;
    else if (MODE == 2) addr = base + warp * 512 + lane * 4;
    else if (MODE == 3) addr = base + warp * 512 + (lane >> 1) * 4;
    else if (MODE == 5) addr = base + warp * 512 + lane * 8;
    else addr = base + warp * 512 + lane * 2 + (lane >= 16 ? 96 : 0);
    int acc = 0;
    for (int it = 0; it < ITERS; ++it) {
        const unsigned addr_it = addr + ((it & 3) << 10) * 0 + ((it & 1) << 14) ;
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            int v, v2 = 0;
            if (MODE == 0 || MODE == 1 || MODE == 6) asm volatile("ld.shared.s16 %0, [%1];" : "=r"(v) : "r"(addr_it + i * 1024) : "memory");
            else if (MODE == 4) asm volatile("ld.shared.u16 %0, [%1];" : "=r"(v) : "r"(addr_it + i * 1024) : "memory");
            else if (MODE == 5) asm volatile("ld.shared.v2.b32 {%0,%1}, [%2];" : "=r"(v), "=r"(v2) : "r"(addr_it + i * 1024) : "memory");
            else asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(addr_it + i * 1024) : "memory");
            acc ^= v ^ v2;
        }
    }
    out[blockIdx.x * 256 + threadIdx.x] = acc;
}

template <int MODE>
int run(const char *name)
{
    int *out;
    const int blocks = 148 * 4;
    CK(cudaMalloc(&out, blocks * 256 * 4));
    k<MODE><<<blocks, 256>>>(out);
    CK(cudaDeviceSynchronize());
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    k<MODE><<<blocks, 256>>>(out);
    cudaEventRecord(e1);
    CK(cudaDeviceSynchronize());
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double lds_per_sm = 16.0 * ITERS * 8 * 4;
    printf("%-44s %.3f ms  %.2f LDS/clk/SM at 1965 MHz\n", name, ms, lds_per_sm / (ms * 1e-3 * 1.965e9));
    cudaFree(out);
    return 0;
}

int main()
{
    run<0>("LDS.S16 lanes 1 short apart (2 lanes/word)");
    run<4>("LDS.U16 lanes 1 short apart (2 lanes/word)");
    run<1>("LDS.S16 lanes 2 shorts apart (1 lane/word)");
    run<2>("LDS.32 lanes 1 word apart");
    run<3>("LDS.32 two lanes per word");
    run<5>("LDS.64 lanes 2 words apart");
    run<6>("LDS.S16 row split at lane 16, pitch 24 words");
    return 0;
}
