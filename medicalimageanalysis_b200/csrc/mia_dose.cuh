// dose -> image-lattice resample under an ROI mask, and the dose-volume-histogram statistics of the
// sample (structure/dose.py:738-816: sitk.ResampleImageFilter linear / default 0.0, dose_arr[np.where(mask > 0)],
// then min / max / mean / median / std / np.percentile / counts below dose levels).
//
// Design: the mask decides which voxels are worth a trilinear gather (an ROI is a few per cent of the lattice), so
// the path is (1) count the mask per 4096-voxel chunk, (2) one-block exclusive scan of the chunk counts, (3) a
// second pass over the mask that samples only the masked voxels and writes them COMPACTED in C order (the order of
// np.where) -- 2 B/voxel of mask traffic plus the gathers; (4) statistics on the compact sample: moments and
// level counts in one pass with a fixed-order two-stage reduction, and EXACT order statistics for the percentiles
// by a three-pass radix select on the order-preserving integer image of the float (12 + 10 + 10 bits, all
// requested ranks at once) -- no sort, no host round trip between the passes.
#pragma once
#include "mia_device.cuh"

namespace mia {

constexpr int kMaskChunk = 4096;            // voxels per block: 256 threads x 16 mask bytes
constexpr int kMaxLevels = 64, kMaxRanks = 64;

__device__ __forceinline__ unsigned nonzero_bytes(unsigned w)
{
    // one bit per non-zero byte, then popcount
    const unsigned t = (w | ((w & 0x7f7f7f7fu) + 0x7f7f7f7fu)) & 0x80808080u;
    return (unsigned)__popc(t);
}

__device__ __forceinline__ uint4 load_mask16(const unsigned char *mask, unsigned long long base, unsigned long long n)
{
    uint4 w = make_uint4(0u, 0u, 0u, 0u);
    if (base + 16 <= n && ((reinterpret_cast<uintptr_t>(mask + base) & 15u) == 0)) {
        w = __ldg(reinterpret_cast<const uint4 *>(mask + base));
    } else if (base < n) {
        unsigned b[4] = {0u, 0u, 0u, 0u};
        for (int e = 0; e < 16 && base + e < n; ++e) b[e >> 2] |= (unsigned)mask[base + e] << (8 * (e & 3));
        w = make_uint4(b[0], b[1], b[2], b[3]);
    }
    return w;
}

__global__ void __launch_bounds__(256) mask_count_kernel(const unsigned char *mask, unsigned long long n, unsigned *counts)
{
    __shared__ unsigned s[8];
    const unsigned long long base = (unsigned long long)blockIdx.x * kMaskChunk + threadIdx.x * 16u;
    const uint4 w = load_mask16(mask, base, n);
    unsigned c = nonzero_bytes(w.x) + nonzero_bytes(w.y) + nonzero_bytes(w.z) + nonzero_bytes(w.w);
    c = warp_sum_u(c);
    if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) counts[blockIdx.x] = s[0] + s[1] + s[2] + s[3] + s[4] + s[5] + s[6] + s[7];
}

// One block: exclusive scan of the chunk counts (64-bit offsets) and the total.
__global__ void __launch_bounds__(1024) mask_scan_kernel(const unsigned *counts, unsigned nchunks, unsigned long long *offsets,
                                                        unsigned long long *total)
{
    __shared__ unsigned long long s[1024];
    const unsigned per = (nchunks + 1023u) / 1024u;
    const unsigned b0 = threadIdx.x * per, b1 = min(nchunks, b0 + per);
    unsigned long long mine = 0;
    for (unsigned b = b0; b < b1; ++b) mine += counts[b];
    s[threadIdx.x] = mine;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {
        const unsigned long long add = threadIdx.x >= (unsigned)o ? s[threadIdx.x - o] : 0ull;
        __syncthreads();
        s[threadIdx.x] += add;
        __syncthreads();
    }
    unsigned long long run = s[threadIdx.x] - mine;
    for (unsigned b = b0; b < b1; ++b) {
        offsets[b] = run;
        run += counts[b];
    }
    if (threadIdx.x == 1023) *total = s[1023];
}

// Second pass: sample the masked voxels only, compacted in C order.
template <typename TM, bool CLOSED>
__global__ void __launch_bounds__(256) resample_masked_kernel(const VolView mov, const unsigned char *mask, int ox, int oy, int oz,
                                                              const Mat34 A, float background,
                                                              const unsigned long long *offsets, float *out)
{
    __shared__ unsigned s_w[8];
    const unsigned long long n = (unsigned long long)ox * oy * oz;
    const unsigned long long base = (unsigned long long)blockIdx.x * kMaskChunk + threadIdx.x * 16u;
    const uint4 w = load_mask16(mask, base, n);
    const unsigned words[4] = {w.x, w.y, w.z, w.w};
    const unsigned c = nonzero_bytes(w.x) + nonzero_bytes(w.y) + nonzero_bytes(w.z) + nonzero_bytes(w.w);
    // exclusive scan of c over the block: warp scan + 8 warp totals
    unsigned incl = c;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) s_w[warp] = incl;
    __syncthreads();
    unsigned before = 0;
    for (int q = 0; q < warp; ++q) before += s_w[q];
    if (c == 0) return;
    unsigned long long dst = offsets[blockIdx.x] + before + (incl - c);
    const TM *mv = static_cast<const TM *>(mov.p);
    const unsigned long long plane = (unsigned long long)ox * oy;
    for (int e = 0; e < 16; ++e) {
        if (((words[e >> 2] >> (8 * (e & 3))) & 0xffu) == 0u) continue;
        const unsigned long long idx = base + e;
        const int k = (int)(idx / plane);
        const unsigned r2 = (unsigned)(idx - (unsigned long long)k * plane);
        const int j = (int)(r2 / (unsigned)ox), i = (int)(r2 - (unsigned)j * (unsigned)ox);
        const RowCtx r = make_row_ctx(A.a, i, j, k, mov.nx, mov.ny, mov.nz);
        float val;
        const bool in = sample_voxel<TM, CLOSED>(mv, mov.nx, mov.ny, mov.nz, r, 0, val);
        out[dst++] = in ? val : background;
    }
}

// ---------------------------------------------------------------------------------------------------------------
// Statistics of a compact float32 sample.
__device__ __forceinline__ unsigned float_key(float f)
{
    const unsigned b = __float_as_uint(f);
    return b ^ ((b >> 31) ? 0xffffffffu : 0x80000000u);       // unsigned order == float order
}
__device__ __forceinline__ float key_float(unsigned k)
{
    return __uint_as_float(k ^ ((k >> 31) ? 0x80000000u : 0xffffffffu));
}

struct StatsWs {                         // device workspace layout (all offsets in bytes, see stats_ws_layout)
    double *part;                        // [grid][4]: sum, min, max, (unused)
    unsigned long long *part_below;      // [grid][kMaxLevels]
    unsigned *ticket;
    float *levels;                       // [kMaxLevels]
    unsigned long long *ranks;           // [kMaxRanks] remaining rank inside the current prefix
    unsigned *prefix;                    // [kMaxRanks]
    unsigned *hist1;                     // [4096]
    unsigned *histn;                     // [kMaxRanks][1024]
    double *part_ssq;                    // [grid]
};

// Pass A: sum / min / max / counts below each level; the last block reduces the block partials in block order.
// result: [0] min [1] max [2] sum [3] mean [4] sum (x - mean)^2 (pass B) [5] n [6 .. 6+L) below [6+L ..) order stats
__global__ void __launch_bounds__(256) stats_moments_kernel(const float *v, unsigned long long n, int n_levels, StatsWs ws, double *result)
{
    __shared__ unsigned long long s_below[kMaxLevels];
    __shared__ float s_levels[kMaxLevels];
    __shared__ double s_sum[8];
    __shared__ float s_min[8], s_max[8];
    __shared__ bool s_last;
    const int tid = threadIdx.x;
    if (tid < kMaxLevels) {
        s_below[tid] = 0ull;
        s_levels[tid] = tid < n_levels ? ws.levels[tid] : 0.0f;
    }
    __syncthreads();
    double sum = 0.0;
    float mn = INFINITY, mx = -INFINITY;
    // warp-uniform trip count: every lane stays in the loop so the ballots below see the whole warp
    for (unsigned long long b0 = (unsigned long long)blockIdx.x * 256u + (tid & ~31); b0 < n; b0 += (unsigned long long)gridDim.x * 256u) {
        const unsigned long long i = b0 + (tid & 31);
        const bool valid = i < n;
        const float x = valid ? __ldg(v + i) : 0.0f;
        if (valid) {
            sum += (double)x;
            mn = fminf(mn, x);
            mx = fmaxf(mx, x);
        }
        // count x < level for every level: one vote per level and warp, one shared atomic per vote
        for (int l = 0; l < n_levels; ++l) {
            const unsigned b = __ballot_sync(0xffffffffu, valid && x < s_levels[l]);
            if (b && (tid & 31) == 0) atomicAdd(&s_below[l], (unsigned long long)__popc(b));
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        sum += __shfl_xor_sync(0xffffffffu, sum, o);
        mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    }
    if ((tid & 31) == 0) { s_sum[tid >> 5] = sum; s_min[tid >> 5] = mn; s_max[tid >> 5] = mx; }
    __syncthreads();
    if (tid == 0) {
        double s = 0.0;
        float a = INFINITY, b = -INFINITY;
        for (int w = 0; w < 8; ++w) { s += s_sum[w]; a = fminf(a, s_min[w]); b = fmaxf(b, s_max[w]); }
        ws.part[4 * blockIdx.x + 0] = s;
        ws.part[4 * blockIdx.x + 1] = (double)a;
        ws.part[4 * blockIdx.x + 2] = (double)b;
    }
    if (tid < n_levels) ws.part_below[(size_t)blockIdx.x * kMaxLevels + tid] = s_below[tid];
    __threadfence();
    __syncthreads();
    if (tid == 0) s_last = atomicAdd(ws.ticket, 1u) == gridDim.x - 1;
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    // fixed-order second stage: thread t sums partials t, t + 256, ...; then a fixed tree over the 256 threads
    {
        __shared__ double f_sum[256], f_min[256], f_max[256];
        double s = 0.0, a = INFINITY, b = -INFINITY;
        for (unsigned g = tid; g < gridDim.x; g += 256) {
            s += __ldcg(ws.part + 4 * g);
            a = fmin(a, __ldcg(ws.part + 4 * g + 1));
            b = fmax(b, __ldcg(ws.part + 4 * g + 2));
        }
        f_sum[tid] = s; f_min[tid] = a; f_max[tid] = b;
        __syncthreads();
        for (int o = 128; o > 0; o >>= 1) {
            if (tid < o) {
                f_sum[tid] += f_sum[tid + o];
                f_min[tid] = fmin(f_min[tid], f_min[tid + o]);
                f_max[tid] = fmax(f_max[tid], f_max[tid + o]);
            }
            __syncthreads();
        }
        if (tid == 0) {
            result[0] = f_min[0]; result[1] = f_max[0]; result[2] = f_sum[0];
            result[3] = n ? f_sum[0] / (double)n : 0.0;
            result[5] = (double)n;
            *ws.ticket = 0u;
        }
    }
    // level counts: warp w takes levels w, w + 8, ...; integer sums, any order
    for (int l = tid >> 5; l < n_levels; l += 8) {
        unsigned long long c = 0;
        for (unsigned g = tid & 31; g < gridDim.x; g += 32) c += __ldcg(ws.part_below + (size_t)g * kMaxLevels + l);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
        if ((tid & 31) == 0) result[6 + l] = (double)c;
    }
}

// Pass B: sum of squared deviations from the mean (np.std is two-pass) and the 12-bit histogram of the keys.
__global__ void __launch_bounds__(256) stats_hist1_kernel(const float *v, unsigned long long n, int n_levels, StatsWs ws, double *result)
{
    __shared__ unsigned s_h[4096];
    __shared__ double s_sum[8];
    __shared__ bool s_last;
    const int tid = threadIdx.x;
    for (int b = tid; b < 4096; b += 256) s_h[b] = 0u;
    __syncthreads();
    const double mean = result[3];
    double ssq = 0.0;
    for (unsigned long long i = (unsigned long long)blockIdx.x * 256u + tid; i < n; i += (unsigned long long)gridDim.x * 256u) {
        const float x = __ldg(v + i);
        const double d = (double)x - mean;
        ssq += d * d;
        atomicAdd(&s_h[float_key(x) >> 20], 1u);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) ssq += __shfl_xor_sync(0xffffffffu, ssq, o);
    if ((tid & 31) == 0) s_sum[tid >> 5] = ssq;
    __syncthreads();
    for (int b = tid; b < 4096; b += 256)
        if (s_h[b]) atomicAdd(ws.hist1 + b, s_h[b]);
    if (tid == 0) {
        double s = 0.0;
        for (int w = 0; w < 8; ++w) s += s_sum[w];
        ws.part_ssq[blockIdx.x] = s;
    }
    __threadfence();
    __syncthreads();
    if (tid == 0) s_last = atomicAdd(ws.ticket, 1u) == gridDim.x - 1;
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    {
        __shared__ double f_sum[256];
        double s = 0.0;
        for (unsigned g = tid; g < gridDim.x; g += 256) s += __ldcg(ws.part_ssq + g);
        f_sum[tid] = s;
        __syncthreads();
        for (int o = 128; o > 0; o >>= 1) {
            if (tid < o) f_sum[tid] += f_sum[tid + o];
            __syncthreads();
        }
        if (tid == 0) {
            result[4] = f_sum[0];
            *ws.ticket = 0u;
        }
    }
    (void)n_levels;
}

// Resolve one radix level for every rank: find the bin of the rank's current histogram in which the cumulative
// count passes the remaining rank.  One block per rank: each thread owns nb / 256 consecutive bins, a block scan of
// the per-thread totals locates the owning thread, which then walks its own bins.
__global__ void __launch_bounds__(256) stats_resolve_kernel(StatsWs ws, int n_ranks, int level, int n_levels, double *result)
{
    __shared__ unsigned long long s_w[8];
    const int r = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (r >= n_ranks) return;
    const unsigned *h = level == 0 ? ws.hist1 : ws.histn + (size_t)r * 1024;
    const int per = level == 0 ? 16 : 4;
    unsigned c[16];
    unsigned long long mine = 0;
#pragma unroll
    for (int e = 0; e < 16; ++e) {
        c[e] = e < per ? __ldcg(h + tid * per + e) : 0u;
        mine += c[e];
    }
    unsigned long long incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned long long t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) s_w[warp] = incl;
    __syncthreads();
    unsigned long long before = incl - mine;
    for (int q = 0; q < warp; ++q) before += s_w[q];
    const unsigned long long rem = ws.ranks[r];
    __syncthreads();                                   // everyone has read ranks[r] before the owner rewrites it
    if (rem >= before && rem < before + mine) {        // exactly one thread (rank < total count)
        unsigned long long left = rem - before;
        int e = 0;
        for (; e < per - 1; ++e) {
            if (left < c[e]) break;
            left -= c[e];
        }
        const unsigned b = (unsigned)(tid * per + e);
        ws.ranks[r] = left;
        const unsigned pre = level == 0 ? b : ((ws.prefix[r] << 10) | b);
        ws.prefix[r] = pre;
        if (level == 2) result[6 + n_levels + r] = (double)key_float(pre);
    }
}

// Passes C / D: 10-bit histograms below each rank's prefix (pbits = 12 or 22 leading bits already fixed).
__global__ void __launch_bounds__(256) stats_histn_kernel(const float *v, unsigned long long n, StatsWs ws, int n_ranks, int pbits)
{
    __shared__ unsigned s_pre[kMaxRanks];
    const int tid = threadIdx.x;
    if (tid < n_ranks) s_pre[tid] = ws.prefix[tid];
    __syncthreads();
    const int shift = 32 - pbits, sub = shift - 10;
    for (unsigned long long i = (unsigned long long)blockIdx.x * 256u + tid; i < n; i += (unsigned long long)gridDim.x * 256u) {
        const unsigned k = float_key(__ldg(v + i));
        const unsigned top = k >> shift, bin = (k >> sub) & 1023u;
        for (int r = 0; r < n_ranks; ++r)
            if (s_pre[r] == top) atomicAdd(ws.histn + (size_t)r * 1024 + bin, 1u);
    }
}

}  // namespace mia
