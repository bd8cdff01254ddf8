// Micro-benchmark behind DESIGN.md 7 ("what comes next"): how fast is a trilinear resample whose 8 taps come from the
// texture unit as TWO 4-texel gathers (tld4: the 2 x 2 footprint in slice z0 and in slice z0 + 1)
// instead of 8 LDS.S16 + 8 I2FP + 3 IMAD?  The texture unit only FETCHES here (point sampling through gather, int16 -> float through the normalised read mode = an exact scale by 2^-15); the 7
// lerps stay float32 FMAs with the kernel's own weights, so the result is the tile kernels' result -- the filtering
// hardware with its 1.8 fixed-point weights is NOT used.
// The volume has to live in a cudaArray for that (opaque tiled layout, cudaArrayTextureGather): the copy is timed too.
// Prints: the copy, the kernel, output Gvoxel/s, the sum of the output (to compare with the LDS path on the same input).
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tests/cuda/ubench_tex tests/cuda/ubench_tex.cu
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("FAIL %s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

constexpr int NX = 512, NY = 512, NZ = 300;

struct Mat { float a[12]; };

// The slices sit side by side in ONE 2-D array (texture gather exists for plain 2-D arrays only: cudaMalloc3DArray rejects
// cudaArrayLayered | cudaArrayTextureGather): slice z at (z % ATLAS_COLS, z / ATLAS_COLS) * 512.  A footprint that reaches
// one texel into the neighbouring slice (x0 = 511) carries weight 0 there.
constexpr int ATLAS_COLS = 8;
__device__ __forceinline__ float4 gather_layer(cudaTextureObject_t tex, int layer, float x, float y)
{
    // r.w = T[i][j], r.z = T[i+1][j], r.x = T[i][j+1], r.y = T[i+1][j+1]   (i along x, j along y)
    return tex2Dgather<float4>(tex, x + (float)((layer % ATLAS_COLS) * NX), y + (float)((layer / ATLAS_COLS) * NY), 0);
}

// one thread = one output column of 8 planes (lane = x): the layout of the 32-wide tile kernels
template <int PLANES>
__global__ void __launch_bounds__(256) resample_tex_kernel(cudaTextureObject_t tex, short *__restrict__ out, int ox, int oy, int oz, Mat A,
                                                           float bg)
{
    const int i = blockIdx.x * 32 + (threadIdx.x & 31), j = blockIdx.y * 8 + (threadIdx.x >> 5), k0 = blockIdx.z * PLANES;
    if (i >= ox || j >= oy) return;
    const float cx0 = A.a[0] * i + A.a[1] * j + A.a[3], cy0 = A.a[4] * i + A.a[5] * j + A.a[7], cz0 = A.a[8] * i + A.a[9] * j + A.a[11];
    short *o = out + ((size_t)k0 * oy + j) * ox + i;
    const size_t plane = (size_t)oy * ox;
#pragma unroll
    for (int kz = 0; kz < PLANES; ++kz) {
        if (k0 + kz >= oz) break;
        const float k = (float)(k0 + kz);
        const float cx = fmaf(A.a[2], k, cx0), cy = fmaf(A.a[6], k, cy0), cz = fmaf(A.a[10], k, cz0);
        const bool in = cx >= -0.5f && cx <= NX - 0.5f && cy >= -0.5f && cy <= NY - 0.5f && cz >= -0.5f && cz <= NZ - 0.5f;
        const float x = fminf(fmaxf(cx, 0.0f), NX - 1.0f), y = fminf(fmaxf(cy, 0.0f), NY - 1.0f), z = fminf(fmaxf(cz, 0.0f), NZ - 1.0f);
        const float fx = floorf(x), fy = floorf(y), fz = floorf(z);
        const float wx = x - fx, wy = y - fy, wz = z - fz;
        const int z0 = (int)fz, z1 = min(z0 + 1, NZ - 1);
        // the footprint of (fx + 1, fy + 1) in unnormalised texel coordinates is the texels fx, fx + 1 / fy, fy + 1
        const float4 lo = gather_layer(tex, z0, fx + 1.0f, fy + 1.0f), hi = gather_layer(tex, z1, fx + 1.0f, fy + 1.0f);
        const float a00 = fmaf(wx, lo.z - lo.w, lo.w), a10 = fmaf(wx, lo.y - lo.x, lo.x);
        const float a01 = fmaf(wx, hi.z - hi.w, hi.w), a11 = fmaf(wx, hi.y - hi.x, hi.x);
        const float b0 = fmaf(wy, a10 - a00, a00), b1 = fmaf(wy, a11 - a01, a01);
        const float v = fmaf(wz, b1 - b0, b0) * 32768.0f;                 // undo the normalised read (exact)
        o[(size_t)kz * plane] = in ? (short)floorf(v + 0.5f) : (short)bg;
    }
}

int main()
{
    std::vector<short> h((size_t)NX * NY * NZ);
    for (int z = 0; z < NZ; ++z)
        for (int y = 0; y < NY; ++y)
            for (int x = 0; x < NX; ++x)
                h[((size_t)z * NY + y) * NX + x] = (short)(600.0 * sin(x / 17.0) * cos(y / 23.0) * sin(z / 19.0) + ((x * 7 + y * 3 + z) % 31));
    short *d_lin = nullptr, *d_out = nullptr;
    const size_t bytes = h.size() * sizeof(short);
    CK(cudaMalloc(&d_lin, bytes));
    CK(cudaMalloc(&d_out, bytes));
    CK(cudaMemcpy(d_lin, h.data(), bytes, cudaMemcpyHostToDevice));

    cudaChannelFormatDesc desc = cudaCreateChannelDesc<short>();
    cudaArray_t arr = nullptr;
    const int rows = (NZ + ATLAS_COLS - 1) / ATLAS_COLS;
    CK(cudaMallocArray(&arr, &desc, (size_t)ATLAS_COLS * NX, (size_t)rows * NY, cudaArrayTextureGather));
    auto upload = [&]() -> cudaError_t {
        for (int z = 0; z < NZ; ++z) {
            cudaError_t e = cudaMemcpy2DToArrayAsync(arr, (size_t)(z % ATLAS_COLS) * NX * sizeof(short), (size_t)(z / ATLAS_COLS) * NY,
                                                     d_lin + (size_t)z * NX * NY, NX * sizeof(short), NX * sizeof(short), NY,
                                                     cudaMemcpyDeviceToDevice, 0);
            if (e != cudaSuccess) return e;
        }
        return cudaSuccess;
    };
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    CK(upload());
    CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(e0));
    for (int r = 0; r < 3; ++r) CK(upload());
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms = 0.0f;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    printf("linear -> 2-D atlas array copy (157 MB, 300 slice copies): %.1f us\n", 1e3 * ms / 3);

    cudaResourceDesc rd = {};
    rd.resType = cudaResourceTypeArray;
    rd.res.array.array = arr;
    cudaTextureDesc td = {};
    td.addressMode[0] = td.addressMode[1] = cudaAddressModeClamp;
    td.filterMode = cudaFilterModePoint;
    td.readMode = cudaReadModeNormalizedFloat;
    td.normalizedCoords = 0;
    cudaTextureObject_t tex = 0;
    CK(cudaCreateTextureObject(&tex, &rd, &td, nullptr));

    // (3, -2, 4) degrees about the volume centre, anisotropic 0.977 / 0.977 / 2.5 mm voxels, like config 2 onto its own lattice
    const double rx = 3 * M_PI / 180, ry = -2 * M_PI / 180, rz = 4 * M_PI / 180, sp[3] = {0.9765625, 0.9765625, 2.5};
    const double Rx[9] = {1, 0, 0, 0, cos(rx), -sin(rx), 0, sin(rx), cos(rx)}, Ry[9] = {cos(ry), 0, sin(ry), 0, 1, 0, -sin(ry), 0, cos(ry)},
                 Rz[9] = {cos(rz), -sin(rz), 0, sin(rz), cos(rz), 0, 0, 0, 1};
    double T[9], R[9];
    for (int a = 0; a < 3; ++a)
        for (int b = 0; b < 3; ++b) {
            T[a * 3 + b] = 0;
            for (int c = 0; c < 3; ++c) T[a * 3 + b] += Ry[a * 3 + c] * Rx[c * 3 + b];
        }
    for (int a = 0; a < 3; ++a)
        for (int b = 0; b < 3; ++b) {
            R[a * 3 + b] = 0;
            for (int c = 0; c < 3; ++c) R[a * 3 + b] += Rz[a * 3 + c] * T[c * 3 + b];
        }
    Mat A;
    const double cen[3] = {(NX - 1) / 2.0, (NY - 1) / 2.0, (NZ - 1) / 2.0};
    for (int a = 0; a < 3; ++a) {
        double t = cen[a] + (a == 0 ? 5.0 : a == 1 ? -3.0 : 7.0) / sp[a];
        for (int b = 0; b < 3; ++b) {
            A.a[a * 4 + b] = (float)(R[a * 3 + b] * sp[b] / sp[a]);
            t -= R[a * 3 + b] * sp[b] / sp[a] * cen[b];
        }
        A.a[a * 4 + 3] = (float)t;
    }
    const dim3 grid((NX + 31) / 32, (NY + 7) / 8, (NZ + 7) / 8);
    for (int r = 0; r < 3; ++r) resample_tex_kernel<8><<<grid, 256>>>(tex, d_out, NX, NY, NZ, A, -3001.0f);
    CK(cudaGetLastError());
    CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(e0));
    for (int r = 0; r < 10; ++r) resample_tex_kernel<8><<<grid, 256>>>(tex, d_out, NX, NY, NZ, A, -3001.0f);
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    CK(cudaEventElapsedTime(&ms, e0, e1));
    const double us = 1e3 * ms / 10, nv = (double)NX * NY * NZ;
    printf("resample through 2 x tld4 per voxel, 512 x 512 x 300 int16: %.1f us, %.1f Gvoxel/s (L2-resident after the first pass: 157 MB array)\n",
           us, nv / us / 1e3);
    CK(cudaMemcpy(h.data(), d_out, bytes, cudaMemcpyDeviceToHost));
    long long sum = 0, nbg = 0;
    for (size_t q = 0; q < h.size(); ++q) {
        sum += h[q];
        nbg += h[q] == -3001;
    }
    printf("checksum %lld, background voxels %lld\n", sum, nbg);
    return 0;
}
