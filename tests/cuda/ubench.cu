// Micro-benchmarks behind DESIGN.md's instruction budget: issue rate of FFMA vs FFMA2 (f32x2), I2FP, LDS.S16,
// shared atomics.  Prints warp-instructions per clock per SM (4 schedulers => 4.0 is the issue ceiling).
#include <cstdio>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("FAIL %s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)
constexpr int ITERS = 4096;

template <int MODE>
__global__ void __launch_bounds__(256) k(float *out, long long *cyc, unsigned *hist_out)
{
    __shared__ unsigned sh[4096];
    for (int i = threadIdx.x; i < 4096; i += 256) sh[i] = i;
    __syncthreads();
    float a[16];
    unsigned long long p[8];
    for (int i = 0; i < 16; ++i) a[i] = threadIdx.x * 0.001f + i;
    for (int i = 0; i < 8; ++i) asm("mov.b64 %0, {%1,%2};" : "=l"(p[i]) : "f"(a[2 * i]), "f"(a[2 * i + 1]));
    const float m = 1.0001f, c = 0.5f;
    unsigned long long m2, c2;
    asm("mov.b64 %0, {%1,%2};" : "=l"(m2) : "f"(m), "f"(m));
    asm("mov.b64 %0, {%1,%2};" : "=l"(c2) : "f"(c), "f"(c));
    int iv[16];
    for (int i = 0; i < 16; ++i) iv[i] = threadIdx.x + i;
    const long long t0 = clock64();
    for (int it = 0; it < ITERS; ++it) {
        if (MODE == 0) {
#pragma unroll
            for (int i = 0; i < 16; ++i) a[i] = fmaf(a[i], m, c);
        } else if (MODE == 1) {
#pragma unroll
            for (int i = 0; i < 8; ++i) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[i]) : "l"(m2), "l"(c2));
        } else if (MODE == 2) {           // 8 FFMA2 + 8 FFMA interleaved
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[i]) : "l"(m2), "l"(c2));
                a[i] = fmaf(a[i], m, c);
            }
        } else if (MODE == 3) {           // 8 FFMA2 + 8 integer ALU ops
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[i]) : "l"(m2), "l"(c2));
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(iv[i]) : "r"(iv[i + 8]), "r"(it));
            }
        } else if (MODE == 4) {           // I2FP: int -> float (LOP3 + I2FP + FADD per element)
#pragma unroll
            for (int i = 0; i < 16; ++i) { a[i] += (float)(iv[i] ^ it); }
        } else if (MODE == 9) {           // LOP3 + FADD only (baseline for mode 4)
#pragma unroll
            for (int i = 0; i < 16; ++i) { a[i] += __int_as_float((iv[i] ^ it) & 0x3fffffff); }
        } else if (MODE == 10 || MODE == 11 || MODE == 12) {   // LDS.S16 lanes 2 shorts apart / LDS.32 / LDS.S16 + same word pairs
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                int v;
                const unsigned base = (unsigned)__cvta_generic_to_shared(sh);
                if (MODE == 10) asm volatile("ld.shared.s16 %0, [%1];" : "=r"(v) : "r"(base + ((iv[i] * 4 + it * 64) & 8188)));
                if (MODE == 11) asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(base + ((iv[i] * 4 + it * 64) & 8188)));
                if (MODE == 12) asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(base + ((iv[i] * 2 + it * 64) & 8188)));
                iv[i] += v & 1;
            }
        } else if (MODE == 13) {          // F2I (XU) + FADD
#pragma unroll
            for (int i = 0; i < 16; ++i) { iv[i] += __float2int_rd(a[i]); a[i] += 0.37f; }
        } else if (MODE == 5) {           // LDS.S16 x16
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                int v;
                asm volatile("ld.shared.s16 %0, [%1];" : "=r"(v) : "r"((unsigned)__cvta_generic_to_shared(sh) + ((iv[i] * 2 + it * 64) & 8190)));
                iv[i] += v & 1;
            }
        } else if (MODE == 6) {           // shared atomics, spread addresses
#pragma unroll
            for (int i = 0; i < 4; ++i) atomicAdd(&sh[(iv[i] * 33 + it * 7) & 4095], 1u);
        } else if (MODE == 7) {           // shared atomics, mostly the same address per warp (air/air bin)
#pragma unroll
            for (int i = 0; i < 4; ++i) atomicAdd(&sh[(it + i) & 4095], 1u);
        } else if (MODE == 8) {           // FADD2.RM + FADD2 chain like the floor trick
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                unsigned long long t;
                asm volatile("add.rm.f32x2 %0, %1, %2;" : "=l"(t) : "l"(p[i]), "l"(c2));
                asm volatile("sub.rn.f32x2 %0, %1, %2;" : "+l"(p[i]) : "l"(t), "l"(m2));
            }
        }
    }
    const long long t1 = clock64();
    float s = 0;
    for (int i = 0; i < 16; ++i) s += a[i] + (float)iv[i];
    for (int i = 0; i < 8; ++i) { float x, y; asm("mov.b64 {%0,%1}, %2;" : "=f"(x), "=f"(y) : "l"(p[i])); s += x + y; }
    out[blockIdx.x * 256 + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
    if (MODE >= 6) hist_out[blockIdx.x * 256 + threadIdx.x] = sh[threadIdx.x];
}

template <int MODE>
int run(const char *name, double warp_instr_per_iter_per_warp, int blocks_per_sm)
{
    int sms = 148;
    float *out; long long *cyc; unsigned *h;
    const int blocks = sms * blocks_per_sm;
    CK(cudaMalloc(&out, blocks * 256 * 4)); CK(cudaMalloc(&cyc, blocks * 8)); CK(cudaMalloc(&h, blocks * 256 * 4));
    k<MODE><<<blocks, 256>>>(out, cyc, h);
    CK(cudaDeviceSynchronize());
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    k<MODE><<<blocks, 256>>>(out, cyc, h);
    cudaEventRecord(e1);
    CK(cudaDeviceSynchronize());
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    long long hc[148 * 8];
    CK(cudaMemcpy(hc, cyc, blocks * 8, cudaMemcpyDeviceToHost));
    double avg = 0; for (int i = 0; i < blocks; ++i) avg += hc[i]; avg /= blocks;
    const double winstr = warp_instr_per_iter_per_warp * ITERS * 8 * blocks_per_sm;   // per SM
    (void)avg;
    printf("%-34s %d CTA/SM: %.3f ms, %.2f warp-instr/clk/SM at 1965 MHz (counted ops only)\n", name, blocks_per_sm, ms,
           winstr / (ms * 1.0e-3 * 1.965e9));
    cudaFree(out); cudaFree(cyc); cudaFree(h);
    return 0;
}

int main()
{
    for (int b = 2; b <= 4; b += 2) {
        run<0>("FFMA x16", 16, b);
        run<1>("FFMA2 x8", 8, b);
        run<2>("FFMA2 x8 + FFMA x8", 16, b);
        run<3>("FFMA2 x8 + LOP3 x8", 16, b);
        run<4>("LOP3+I2FP+FADD x16 (48 ops)", 48, b);
        run<5>("LDS.S16 x16 (+2 alu each)", 16, b);
        run<6>("ATOMS spread x4", 4, b);
        run<7>("ATOMS same-address x4", 4, b);
        run<8>("FADD2.RM + FADD2 x8 (16 ops)", 16, b);
        run<9>("LOP3+LOP3+FADD x16 (48 ops)", 48, b);
        run<10>("LDS.S16 distinct words x16", 16, b);
        run<11>("LDS.32 distinct words x16", 16, b);
        run<12>("LDS.32 two lanes per word x16", 16, b);
        run<13>("F2I+IADD+FADD x16 (48 ops)", 48, b);
    }
    return 0;
}
