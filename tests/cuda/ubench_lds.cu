// LDS throughput micro-benchmark with un-hoistable addresses (addr += v & runtime_zero).
#include <cstdio>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("FAIL %s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)
constexpr int ITERS = 4096;
template <int WIDTH>   // 2 = LDS.S16, 4 = LDS.32
__global__ void __launch_bounds__(256) k(int *out, int zero, int lane_stride_bytes, int split_lane, int split_extra_bytes)
{
    __shared__ __align__(16) unsigned sh[12288];
    for (int i = threadIdx.x; i < 12288; i += 256) sh[i] = i * 2654435761u;
    __syncthreads();
    const unsigned base = (unsigned)__cvta_generic_to_shared(sh);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned addr[8];
    for (int i = 0; i < 8; ++i)
        addr[i] = base + warp * 256 + i * 4096 + lane * lane_stride_bytes + (lane >= split_lane ? split_extra_bytes : 0);
    int acc = 0;
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            int v;
            if (WIDTH == 2) asm volatile("ld.shared.s16 %0, [%1];" : "=r"(v) : "r"(addr[i]));
            else asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(addr[i]));
            addr[i] += v & zero;
            acc += v;
        }
    }
    out[blockIdx.x * 256 + threadIdx.x] = acc;
}

template <int WIDTH>
int run(const char *name, int stride, int split_lane, int extra)
{
    int *out;
    const int blocks = 148 * 4;
    CK(cudaMalloc(&out, blocks * 256 * 4));
    k<WIDTH><<<blocks, 256>>>(out, 0, stride, split_lane, extra);
    CK(cudaDeviceSynchronize());
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    k<WIDTH><<<blocks, 256>>>(out, 0, stride, split_lane, extra);
    cudaEventRecord(e1);
    CK(cudaDeviceSynchronize());
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double lds_per_sm = 8.0 * ITERS * 8 * 4;
    printf("%-58s %.3f ms  %.2f LDS/clk/SM at 1965 MHz\n", name, ms, lds_per_sm / (ms * 1e-3 * 1.965e9));
    cudaFree(out);
    return 0;
}

int main()
{
    run<4>("LDS.32 lanes 1 word apart", 4, 99, 0);
    run<2>("LDS.S16 lanes 1 short apart (2 lanes per word)", 2, 99, 0);
    run<2>("LDS.S16 lanes 2 shorts apart (1 lane per word)", 4, 99, 0);
    run<2>("LDS.S16 1 short apart, lanes>=14 on next 128-B row", 2, 14, 128);
    run<2>("LDS.S16 1 short apart, lanes>=14 on next 96-B row", 2, 14, 96);
    run<2>("LDS.S16 1 short apart, lanes>=14 on next 80-B row", 2, 14, 80);
    run<2>("LDS.S16 1 short apart, lanes>=14 on next 160-B row", 2, 14, 160);
    run<2>("LDS.S16 1 short apart, lanes>=14 on next 2048-B plane", 2, 14, 2048);
    run<2>("LDS.S16 all lanes same address", 0, 99, 0);
    return 0;
}
