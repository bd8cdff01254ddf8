// Kernels of the rigid resample + metric hot path (sm_100a).  See DESIGN.md for the roofline of each.
#pragma once
#include "mia_device.cuh"

namespace mia {

enum Metric { kNone = 0, kSsd = 1, kHist = 2, kMattes = 3 };

struct KParams {
    VolView mov;                 // moving volume
    const void *fixed;           // fixed volume on the output lattice (metrics) or nullptr
    const unsigned char *fmask;  // optional fixed-image mask on the output lattice (metrics): 0 = voxel ignored
    const unsigned char *mmask;  // optional moving-image mask on the moving lattice, looked up at the nearest voxel of
                                 // the mapped point (per-voxel kernel only)
    void *out;                   // resampled output or nullptr
    int ox, oy, oz;              // output lattice
    int ntx, nty;                // tiles along x / y
    unsigned magic_fold;         // tile path: -(kMagicBits * ((BY+1)*BX+1)) * sizeof(TM) mod 2^32, see tile_compute
    long long ntiles;            // ntx * nty * oz
    Mat34 A0;                    // matrix 0 by value (host-matrix path, n_mats == 1)
    const double *dA;            // [n_mats][12] device matrices, or nullptr -> A0
    float background;
    int round_mode;              // 0 half-up, 1 trunc
    // metric outputs
    double *ws_ssd;              // [n_mats][gridDim.x] block partials
    unsigned *ws_cnt;            // [n_mats][gridDim.x]
    unsigned *ws_ticket;         // [n_mats]
    double *d_ssd;               // [n_mats]
    unsigned long long *d_count; // [n_mats]
    unsigned *d_hist;            // [n_mats][bins][bins]
    unsigned long long *d_hist64; // kMattes: [n_mats][bins][bins] in units of 1 / kParzenOne sample
    float fmin, fscale, mmin, mscale;
    int bins;
};

template <typename T, int N>
struct alignas(sizeof(T) * N) Vec {
    T v[N];
};

__device__ __forceinline__ int bin_of(float v, float vmin, float scale, int bins)
{
    float b = floorf((v - vmin) * scale);
    b = fminf(fmaxf(b, 0.0f), (float)(bins - 1));   // NaN -> 0 via fmaxf
    return (int)b;
}

// ---- Mattes mutual information: Parzen-window joint histogram ------------------------------------------
// Structure of itk::MattesMutualInformationImageToImageMetricv4 (what "AdvancedMattesMutualInformation" +
// NumberOfHistogramBins ask of Elastix, utils/deformable/simpleitk.py:155-156): B bins with 2 padding bins on either
// side, a zero-order (box) window on the fixed image and a cubic B-spline window over 4 bins on the moving image.
// To keep the histogram an INTEGER one (order-independent, bit-identical for any grid, sharding or GPU count) the
// B-spline is evaluated on a 1/256 sub-bin lattice with 10-bit weights that sum to kParzenOne exactly; the table is
// built with integer arithmetic only, so host oracle and device agree bit for bit.
constexpr int kParzenOne = 1024, kParzenSteps = 256;

__host__ __device__ inline void parzen_weights(int k, unsigned short q[4])
{
    // t = k / 256;  w0 = (1-t)^3 / 6, w1 = (4 - 6 t^2 + 3 t^3) / 6, w3 = t^3 / 6, w2 = the rest; q = round(1024 w)
    const long long T = kParzenSteps, T3 = T * T * T, den = 6 * T3, one = kParzenOne;
    const long long a = T - k;
    const long long n0 = a * a * a, n1 = 4 * T3 - 6 * (long long)k * k * T + 3 * (long long)k * k * k, n3 = (long long)k * k * k;
    const long long q0 = (2 * one * n0 + den) / (2 * den), q1 = (2 * one * n1 + den) / (2 * den), q3 = (2 * one * n3 + den) / (2 * den);
    q[0] = (unsigned short)q0; q[1] = (unsigned short)q1; q[3] = (unsigned short)q3;
    q[2] = (unsigned short)(one - q0 - q1 - q3);
}

// lut: [kParzenSteps + 1][4] unsigned short in shared memory, right behind the bins * bins counters
__device__ __forceinline__ void parzen_lut_fill(unsigned short *lut, int tid, int nthreads)
{
    for (int k = tid; k <= kParzenSteps; k += nthreads) parzen_weights(k, lut + 4 * k);
}

// One sample into the CTA's histogram.  p.fscale / p.mscale are (bins - 4) / range here.
__device__ __forceinline__ void mattes_add(unsigned *s_hist, const unsigned short *lut, float f, float v, float fmin,
                                           float fscale, float mmin, float mscale, int bins)
{
    const int fb = bin_of(f, fmin, fscale, bins - 4) + 2;
    float b = fmaxf((v - mmin) * mscale, 0.0f);                       // NaN -> 0 like bin_of
    const float i0 = fminf(floorf(b), (float)(bins - 5));
    const float t = fminf(b - i0, 1.0f);
    const int tq = (int)floorf(fmaf(t, (float)kParzenSteps, 0.5f));
    const ushort4 w = *reinterpret_cast<const ushort4 *>(lut + 4 * tq);
    unsigned *cell = s_hist + fb * bins + (int)i0 + 1;                // first Parzen bin = (i0 + 2) - 1
    if (w.x) atomicAdd(cell, (unsigned)w.x);
    atomicAdd(cell + 1, (unsigned)w.y);
    atomicAdd(cell + 2, (unsigned)w.z);
    if (w.w) atomicAdd(cell + 3, (unsigned)w.w);
}

// One launch does everything: grid.x blocks stride over the (x-tile, y-tile, z) tiles of the output
// lattice, grid.y = candidate matrix.  METRIC selects the fused reduction.
template <typename TM, typename TF, typename TO, int METRIC, bool STORE, bool CLOSED>
__global__ void __launch_bounds__(kThreads) resample_kernel(const KParams p)
{
    extern __shared__ unsigned s_hist[];        // METRIC == kHist: bins*bins counters
    __shared__ double s_red[kThreads / 32];
    __shared__ unsigned s_redc[kThreads / 32];
    __shared__ bool s_last;

    const int m = blockIdx.y;
    const double *A = p.dA ? p.dA + 12 * m : p.A0.a;
    const TM *mov = static_cast<const TM *>(p.mov.p);
    const TF *fixed = static_cast<const TF *>(p.fixed);
    TO *out = static_cast<TO *>(p.out);
    const int nx = p.mov.nx, ny = p.mov.ny, nz = p.mov.nz;
    const int tx = threadIdx.x, ty = threadIdx.y;
    const int tid = ty * kBlockX + tx;

    unsigned short *s_lut = reinterpret_cast<unsigned short *>(s_hist + p.bins * p.bins);
    if (METRIC == kHist || METRIC == kMattes) {
        for (int b = tid; b < p.bins * p.bins; b += kThreads) s_hist[b] = 0u;
        if (METRIC == kMattes) parzen_lut_fill(s_lut, tid, kThreads);
        __syncthreads();
    }

    double acc = 0.0;
    unsigned cnt = 0;

    for (long long tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
        const int tix = (int)(tile % p.ntx);
        const long long t2 = tile / p.ntx;
        const int tiy = (int)(t2 % p.nty);
        const int k = (int)(t2 / p.nty);
        const int j = tiy * kBlockY + ty;
        if (j >= p.oy) continue;
        const size_t rowbase = ((size_t)k * p.oy + j) * p.ox;
        // Shift the 4-voxel chunks of this row so that every full chunk is a naturally aligned vector:
        // rows of an auto-cropped lattice have arbitrary length, so the misalignment differs per row.
        unsigned mis;
        if (STORE) mis = (unsigned)(reinterpret_cast<uintptr_t>(out + rowbase) / sizeof(TO)) & (kV - 1);
        else mis = (unsigned)(reinterpret_cast<uintptr_t>(fixed + rowbase) / sizeof(TF)) & (kV - 1);
        const int i0 = tix * kTileX + tx * kV - (int)mis;
        if (i0 >= p.ox || i0 + kV <= 0) continue;
        const bool full = (i0 >= 0) && (i0 + kV <= p.ox);
        const size_t row = rowbase + i0;          // may point before the row when i0 < 0: only used when full

        const RowCtx r = make_row_ctx(A, i0, j, k, nx, ny, nz);

        float fv[kV];
        if (METRIC != kNone) {
            const bool fvec = full && ((reinterpret_cast<uintptr_t>(fixed + row) & (sizeof(TF) * kV - 1)) == 0);
            if (fvec) {
                const Vec<TF, kV> f = *reinterpret_cast<const Vec<TF, kV> *>(fixed + row);
#pragma unroll
                for (int v = 0; v < kV; ++v) fv[v] = (float)f.v[v];
            } else {
#pragma unroll
                for (int v = 0; v < kV; ++v)
                    fv[v] = (i0 + v >= 0 && i0 + v < p.ox) ? (float)fixed[rowbase + i0 + v] : 0.0f;
            }
        }

        Vec<TO, kV> o;
        float part = 0.0f;
#pragma unroll
        for (int v = 0; v < kV; ++v) {
            float val;
            bool live = full || (i0 + v >= 0 && i0 + v < p.ox);
            if (METRIC != kNone && p.fmask && live) live = __ldg(p.fmask + rowbase + i0 + v) != 0;
            bool in = live && sample_voxel<TM, CLOSED>(mov, nx, ny, nz, r, v, val);
            if (METRIC != kNone && p.mmask && in) in = nearest_mask_voxel(p.mmask, nx, ny, nz, r, v);
            if (STORE) o.v[v] = in ? to_out<TO>(val, p.round_mode) : to_out<TO>(p.background, p.round_mode);
            if (METRIC == kSsd && in) {
                const float d = fv[v] - val;
                part = fmaf(d, d, part);
                ++cnt;
            }
            if (METRIC == kHist && in) {
                const int bf = bin_of(fv[v], p.fmin, p.fscale, p.bins);
                const int bm = bin_of(val, p.mmin, p.mscale, p.bins);
                atomicAdd(&s_hist[bf * p.bins + bm], 1u);
                ++cnt;
            }
            if (METRIC == kMattes && in) {
                mattes_add(s_hist, s_lut, fv[v], val, p.fmin, p.fscale, p.mmin, p.mscale, p.bins);
                ++cnt;
            }
        }
        if (METRIC == kSsd) acc += (double)part;
        if (STORE) {
            if (full) {
                *reinterpret_cast<Vec<TO, kV> *>(out + row) = o;
            } else {
#pragma unroll
                for (int v = 0; v < kV; ++v)
                    if (i0 + v >= 0 && i0 + v < p.ox) out[rowbase + i0 + v] = o.v[v];
            }
        }
    }

    if (METRIC == kNone) return;

    // ---- block reduction (fixed order => deterministic for a given grid) ----
    const int lane = tid & 31, warp = tid >> 5;
    if (METRIC == kSsd) {
        // double warp reduction through two 32-bit shuffles per step
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    }
    cnt = warp_sum_u(cnt);
    if (lane == 0) {
        s_red[warp] = acc;
        s_redc[warp] = cnt;
    }
    __syncthreads();

    if (METRIC == kHist) {
        // flush the privatised histogram; integer adds commute, so order does not matter
        unsigned *gh = p.d_hist + (size_t)m * p.bins * p.bins;
        for (int b = tid; b < p.bins * p.bins; b += kThreads) {
            const unsigned c = s_hist[b];
            if (c) atomicAdd(gh + b, c);
        }
    }
    if (METRIC == kMattes) {
        unsigned long long *gh = p.d_hist64 + (size_t)m * p.bins * p.bins;
        for (int b = tid; b < p.bins * p.bins; b += kThreads) {
            const unsigned c = s_hist[b];
            if (c) atomicAdd(gh + b, (unsigned long long)c);
        }
    }

    if (tid == 0) {
        double s = 0.0;
        unsigned c = 0;
        for (int w = 0; w < kThreads / 32; ++w) {
            s += s_red[w];
            c += s_redc[w];
        }
        p.ws_ssd[(size_t)m * gridDim.x + blockIdx.x] = s;
        p.ws_cnt[(size_t)m * gridDim.x + blockIdx.x] = c;
        __threadfence();
        const unsigned t = atomicAdd(p.ws_ticket + m, 1u);
        s_last = (t == gridDim.x - 1);
    }
    __syncthreads();
    if (!s_last) return;

    // last block of this matrix: fixed-order second stage over all block partials
    __threadfence();
    double s = 0.0;
    unsigned long long c = 0;
    for (unsigned b = tid; b < gridDim.x; b += kThreads) {
        s += __ldcg(p.ws_ssd + (size_t)m * gridDim.x + b);
        c += __ldcg(p.ws_cnt + (size_t)m * gridDim.x + b);
    }
    __shared__ double s_fin[kThreads];
    __shared__ unsigned long long s_finc[kThreads];
    s_fin[tid] = s;
    s_finc[tid] = c;
    __syncthreads();
    for (int o = kThreads / 2; o > 0; o >>= 1) {
        if (tid < o) {
            s_fin[tid] += s_fin[tid + o];
            s_finc[tid] += s_finc[tid + o];
        }
        __syncthreads();
    }
    if (tid == 0) {
        if (METRIC == kSsd) p.d_ssd[m] = s_fin[0];
        p.d_count[m] = s_finc[0];
        p.ws_ticket[m] = 0u;     // self-reset: the workspace is reusable without a memset
    }
}

// ---------------------------------------------------------------------------------------------
// MI from an integer joint histogram: one block per matrix, float64, fixed-order tree reductions.
template <typename CT>
__global__ void __launch_bounds__(256) mi_kernel(const CT *__restrict__ hist, int bins, double *mi)
{
    extern __shared__ double s_buf[];        // [bins] row sums, [bins] col sums, [256] scratch
    double *rowsum = s_buf, *colsum = s_buf + bins, *red = s_buf + 2 * bins;
    const CT *h = hist + (size_t)blockIdx.x * bins * bins;
    const int tid = threadIdx.x;
    for (int b = tid; b < bins; b += 256) {
        double rs = 0.0, cs = 0.0;
        for (int q = 0; q < bins; ++q) {
            rs += (double)h[b * bins + q];
            cs += (double)h[q * bins + b];
        }
        rowsum[b] = rs;
        colsum[b] = cs;
    }
    __syncthreads();
    double part = 0.0;
    for (int b = tid; b < bins; b += 256) part += rowsum[b];
    red[tid] = part;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (tid < o) red[tid] += red[tid + o];
        __syncthreads();
    }
    const double n = red[0];
    __syncthreads();
    part = 0.0;
    if (n > 0.0) {
        for (int e = tid; e < bins * bins; e += 256) {
            const CT c = h[e];
            if (c) {
                const double hc = (double)c;
                part += (hc / n) * log(hc * n / (rowsum[e / bins] * colsum[e % bins]));
            }
        }
    }
    red[tid] = part;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (tid < o) red[tid] += red[tid + o];
        __syncthreads();
    }
    if (tid == 0) mi[blockIdx.x] = red[0];
}

// ---------------------------------------------------------------------------------------------
// Exact axis permutation / flip (Feet-First-Supine normalisation).  Axes in [z,y,x] order.
struct PermParams {
    const void *in;
    void *out;
    int idim[3];     // input dims  [z,y,x]
    int odim[3];     // output dims [z,y,x]
    int perm[3];     // output axis a reads input axis perm[a]
    int flip[3];     // output axis a reversed
};

// Path A: input x stays the fastest output axis (possibly reversed): row copies, coalesced both sides.
template <typename T>
__global__ void __launch_bounds__(256) permute_rows_kernel(const PermParams p)
{
    const T *in = static_cast<const T *>(p.in);
    T *out = static_cast<T *>(p.out);
    const int ox = p.odim[2];
    const long long rows = (long long)p.odim[0] * p.odim[1];
    const int rows_per_block = blockDim.y;
    for (long long r = (long long)blockIdx.x * rows_per_block + threadIdx.y; r < rows;
         r += (long long)gridDim.x * rows_per_block) {
        const int o0 = (int)(r / p.odim[1]), o1 = (int)(r % p.odim[1]);
        int idx[3];
        idx[p.perm[0]] = p.flip[0] ? p.odim[0] - 1 - o0 : o0;
        idx[p.perm[1]] = p.flip[1] ? p.odim[1] - 1 - o1 : o1;
        const T *src = in + ((size_t)idx[0] * p.idim[1] + idx[1]) * p.idim[2];
        T *dst = out + (size_t)r * ox;
        if (p.flip[2]) {
            for (int x = threadIdx.x; x < ox; x += blockDim.x) dst[x] = __ldg(src + (ox - 1 - x));
        } else {
            for (int x = threadIdx.x; x < ox; x += blockDim.x) dst[x] = __ldg(src + x);
        }
    }
}

// Path B: input x becomes output axis pa != 2 -> 32x32 (x4 rows per thread) transpose through shared memory.
// The tile plane is (input x) x (input axis q = perm[2] that becomes output x).
template <typename T>
__global__ void __launch_bounds__(256) permute_transpose_kernel(const PermParams p, int pa, int pc)
{
    // pa: output axis fed by input x; pc: the remaining output axis (batch)
    __shared__ T tile[32][33];
    const T *in = static_cast<const T *>(p.in);
    T *out = static_cast<T *>(p.out);
    const int q = p.perm[2];                 // input axis that becomes output x
    const int c_in = p.perm[pc];             // input axis of the batch dimension
    const int nxin = p.idim[2];              // length along input x   (= odim[pa])
    const int nq = p.idim[q];                // length along input q   (= odim[2])
    const int tiles_x = (nxin + 31) / 32, tiles_q = (nq + 31) / 32;
    const long long ntiles = (long long)tiles_x * tiles_q * p.odim[pc];
    for (long long t = blockIdx.x; t < ntiles; t += gridDim.x) {
        const int tx0 = (int)(t % tiles_x) * 32;
        const long long t2 = t / tiles_x;
        const int tq0 = (int)(t2 % tiles_q) * 32;
        const int oc = (int)(t2 / tiles_q);  // output index along pc
        const int ic = p.flip[pc] ? p.odim[pc] - 1 - oc : oc;
        // load: threads sweep input x (fast), rows along input q
        for (int rr = threadIdx.y; rr < 32; rr += 8) {
            const int xi = tx0 + threadIdx.x, qi = tq0 + rr;
            if (xi < nxin && qi < nq) {
                int idx[3];
                idx[2] = xi;
                idx[q] = qi;
                idx[c_in] = ic;
                tile[rr][threadIdx.x] = __ldg(in + ((size_t)idx[0] * p.idim[1] + idx[1]) * p.idim[2] + idx[2]);
            }
        }
        __syncthreads();
        // store: threads sweep output x (input q), rows along output pa (input x)
        for (int rr = threadIdx.y; rr < 32; rr += 8) {
            const int qi = tq0 + threadIdx.x, xi = tx0 + rr;
            if (xi < nxin && qi < nq) {
                int o[3];
                o[2] = p.flip[2] ? nq - 1 - qi : qi;
                o[pa] = p.flip[pa] ? nxin - 1 - xi : xi;
                o[pc] = oc;
                out[((size_t)o[0] * p.odim[1] + o[1]) * p.odim[2] + o[2]] = tile[threadIdx.x][rr];
            }
        }
        __syncthreads();
    }
}


// ---------------------------------------------------------------------------------------------
// int16 fast paths of the FFS permutation: 16-byte / 4-byte accesses instead of 2-byte ones.
// HBM-bound copies (4 B per voxel); the scalar kernels above remain for odd sizes.

__device__ __forceinline__ unsigned swap_halves(unsigned w) { return __byte_perm(w, 0u, 0x1032); }

// Path A (x stays x): one thread per 8-voxel chunk.  Needs odim[2] % 8 == 0 and 16-byte aligned bases.
__global__ void __launch_bounds__(256) permute_rows_vec_i16_kernel(const PermParams p, unsigned chunks_per_row, unsigned long long nchunks)
{
    const uint4 *in = static_cast<const uint4 *>(p.in);
    uint4 *out = static_cast<uint4 *>(p.out);
    for (unsigned long long c = (unsigned long long)blockIdx.x * 256u + threadIdx.x; c < nchunks;
         c += (unsigned long long)gridDim.x * 256u) {
        const unsigned r = (unsigned)(c / chunks_per_row), cx = (unsigned)(c % chunks_per_row);
        const int o0 = (int)(r / (unsigned)p.odim[1]), o1 = (int)(r % (unsigned)p.odim[1]);
        int idx[3];
        idx[p.perm[0]] = p.flip[0] ? p.odim[0] - 1 - o0 : o0;
        idx[p.perm[1]] = p.flip[1] ? p.odim[1] - 1 - o1 : o1;
        const size_t srow = ((size_t)idx[0] * p.idim[1] + idx[1]) * chunks_per_row;
        uint4 v;
        if (p.flip[2]) {
            const uint4 s = __ldg(in + srow + (chunks_per_row - 1 - cx));
            v.x = swap_halves(s.w); v.y = swap_halves(s.z); v.z = swap_halves(s.y); v.w = swap_halves(s.x);
        } else {
            v = __ldg(in + srow + cx);
        }
        out[c] = v;
    }
}

// Path B (input x becomes output axis pa): 64 x 64 tile, 2x2 register transposes so that both the global
// loads (along input x) and the global stores (along input q = output x) are 4-byte, 128 B per warp.
// Needs even idim[2] and even idim[q].
//
// Persistent and software-pipelined: a block walks tiles blockIdx.x + n * gridDim.x and keeps DEPTH tiles of
// loads in flight in registers (8 words per thread and tile) while it transposes and stores the oldest one.
// The one-tile-per-block first cut was latency-bound at 42 % of the HBM roofline: a block held 8 KB of loads
// in flight only until they landed and then spent the rest of its life on shared memory, stores and the block
// relaunch, so an SM averaged ~30 KB of reads in flight where Little's law asks for ~45 KB at 6.5 TB/s.
template <int DEPTH>
__global__ void __launch_bounds__(256, (DEPTH <= 1 ? 8 : (DEPTH <= 3 ? 4 : 3))) permute_transpose_vec_i16_kernel(const PermParams p, int pa, int pc)
{
    __shared__ unsigned te[32][33], to[32][33];      // even / odd input x; [x / 2][q / 2]
    const int q = p.perm[2], c_in = p.perm[pc];
    const int nxin = p.idim[2], nq = p.idim[q];
    const int tiles_x = (nxin + 63) / 64, tiles_q = (nq + 63) / 64;
    const long long ntiles = (long long)tiles_x * tiles_q * p.odim[pc];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // element strides: input axes [z,y,x] and output axes [z,y,x]
    const size_t istr[3] = {(size_t)p.idim[1] * p.idim[2], (size_t)p.idim[2], 1};
    const size_t ostr[3] = {(size_t)p.odim[1] * p.odim[2], (size_t)p.odim[2], 1};
    const unsigned sq = (unsigned)istr[q];             // input stride of the axis that becomes output x
    const unsigned sa = (unsigned)ostr[pa];            // output stride of the axis fed by input x
    const size_t sc_in = istr[c_in], sc_out = ostr[pc];
    const bool fa = p.flip[pa] != 0, f2 = p.flip[2] != 0, fc = p.flip[pc] != 0;
    const int nc = p.odim[pc];
    const long long G = gridDim.x;

    struct Loc { int tx0, tq0, oc; };
    auto locate = [&](long long t) {
        Loc l;
        l.tx0 = (int)(t % tiles_x) * 64;
        const long long t2 = t / tiles_x;
        l.tq0 = (int)(t2 % tiles_q) * 64;
        l.oc = (int)(t2 / tiles_q);
        return l;
    };
    auto load = [&](long long t, unsigned (&w0)[4], unsigned (&w1)[4]) {
        if (t >= ntiles) return;
        const Loc l = locate(t);
        const int ic = fc ? nc - 1 - l.oc : l.oc;
        // tile base (64-bit once per tile); everything inside the tile is a 32-bit offset
        const int16_t *ibase = static_cast<const int16_t *>(p.in) + (size_t)ic * sc_in + (size_t)l.tq0 * sq + l.tx0;
        const int nxt = min(64, nxin - l.tx0), nqt = min(64, nq - l.tq0);
#pragma unroll
        for (int it = 0; it < 4; ++it) {
            const int rr = warp + 8 * it;
            w0[it] = w1[it] = 0u;
            if (2 * lane < nxt && 2 * rr < nqt) {
                const unsigned e = (unsigned)(2 * rr) * sq + 2u * lane;
                w0[it] = __ldg(reinterpret_cast<const unsigned *>(ibase + e));
                w1[it] = __ldg(reinterpret_cast<const unsigned *>(ibase + e + sq));
            }
        }
    };
    unsigned w0[DEPTH][4], w1[DEPTH][4];
#pragma unroll
    for (int d = 0; d < DEPTH; ++d) load((long long)blockIdx.x + d * G, w0[d], w1[d]);

    for (long long tb = blockIdx.x; tb < ntiles; tb += DEPTH * G) {
#pragma unroll
        for (int d = 0; d < DEPTH; ++d) {
            const long long t = tb + d * G;
            if (t >= ntiles) break;
#pragma unroll
            for (int it = 0; it < 4; ++it) {
                const int rr = warp + 8 * it;
                te[lane][rr] = __byte_perm(w0[d][it], w1[d][it], 0x5410);      // (a[q][x],   a[q+1][x])
                to[lane][rr] = __byte_perm(w0[d][it], w1[d][it], 0x7632);      // (a[q][x+1], a[q+1][x+1])
            }
            load(t + DEPTH * G, w0[d], w1[d]);        // the registers are free again: refill them DEPTH tiles ahead
            __syncthreads();
            const Loc l = locate(t);
            const int xo0 = fa ? nxin - 1 - l.tx0 : l.tx0;          // output index along pa of tile row 0
            const int qo0 = f2 ? nq - 2 - l.tq0 : l.tq0;             // output x of the tile's first pair
            int16_t *obase = static_cast<int16_t *>(p.out) + (size_t)l.oc * sc_out + (size_t)xo0 * sa + qo0;
            const int nxt = min(64, nxin - l.tx0), nqt = min(64, nq - l.tq0);
#pragma unroll
            for (int it = 0; it < 8; ++it) {
                const int xr = warp + 8 * it;
                if (xr < nxt && 2 * lane < nqt) {
                    unsigned w = (xr & 1) ? to[xr >> 1][lane] : te[xr >> 1][lane];
                    const long long row = fa ? -(long long)xr * sa : (long long)xr * sa;
                    long long col = 2 * lane;
                    if (f2) {
                        col = -col;
                        w = swap_halves(w);
                    }
                    *reinterpret_cast<unsigned *>(obase + row + col) = w;
                }
            }
            __syncthreads();
        }
    }
}

// Path B, 16-byte version (needs idim[2] % 8 == 0, idim[q] % 8 == 0, 16-byte aligned bases): the 4-byte kernel above
// is ISSUE-bound, not memory-bound (ncu: 83 % issue slots, 37 % DRAM, 36 lane-instructions per voxel -- 64-bit tile
// arithmetic and one memory instruction per 2 voxels).  Here a thread moves 16 voxels per tile with 2 LDG.128,
// 8 PRMT (2x2 transposes), 8 STS.32, 2 LDS.128 and 2 STG.128, and all tile arithmetic is 32-bit.
// Shared tile T[x][q / 2] (64 rows of 32 words, no padding) with the 4-word column group XOR-swizzled by x / 8:
// the writes of a warp (8 x-chunks x 4 row pairs) and the LDS.128 of a quarter warp (8 groups of one row) both
// touch 32 different banks.
// W16: output rows are 16-byte aligned (idim[q] % 8 == 0); otherwise idim[q] % 4 == 0 and the 8 voxels of a thread
// leave as two 8-byte stores (the last group of a row may hold only 4).
template <int DEPTH, bool W16>
__global__ void __launch_bounds__(256) permute_transpose_v16_i16_kernel(const PermParams p, int pa, int pc)
{
    __shared__ __align__(16) unsigned T[64 * 32];
    const int q = p.perm[2], c_in = p.perm[pc];
    const int nxin = p.idim[2], nq = p.idim[q];
    const unsigned tiles_x = (unsigned)(nxin + 63) / 64u, tiles_q = (unsigned)(nq + 63) / 64u;
    const unsigned ntiles = tiles_x * tiles_q * (unsigned)p.odim[pc];
    const size_t istr[3] = {(size_t)p.idim[1] * p.idim[2], (size_t)p.idim[2], 1};
    const size_t ostr[3] = {(size_t)p.odim[1] * p.odim[2], (size_t)p.odim[2], 1};
    const size_t sq = istr[q], sa = ostr[pa], sc_in = istr[c_in], sc_out = ostr[pc];
    const bool fa = p.flip[pa] != 0, f2 = p.flip[2] != 0, fc = p.flip[pc] != 0;
    const int nc = p.odim[pc];
    const int tid = threadIdx.x;
    const int lc = tid & 7, lr = tid >> 3;            // load role: x chunk (8 voxels), row pair
    const unsigned G = gridDim.x;
    const int16_t *in = static_cast<const int16_t *>(p.in);
    int16_t *out = static_cast<int16_t *>(p.out);

    auto load = [&](unsigned t, uint4 &a, uint4 &b) {
        a = make_uint4(0u, 0u, 0u, 0u);
        b = a;
        if (t >= ntiles) return;
        const unsigned t2 = t / tiles_x;
        const int tx0 = (int)(t - t2 * tiles_x) * 64;
        const unsigned oc = t2 / tiles_q;
        const int tq0 = (int)(t2 - oc * tiles_q) * 64;
        const int ic = fc ? nc - 1 - (int)oc : (int)oc;
        if (tx0 + 8 * lc < nxin && tq0 + 2 * lr < nq) {
            const int16_t *src = in + (size_t)ic * sc_in + (size_t)(tq0 + 2 * lr) * sq + (size_t)(tx0 + 8 * lc);
            a = __ldg(reinterpret_cast<const uint4 *>(src));
            b = __ldg(reinterpret_cast<const uint4 *>(src + sq));
        }
    };
    uint4 ra[DEPTH], rb[DEPTH];
#pragma unroll
    for (int d = 0; d < DEPTH; ++d) load(blockIdx.x + (unsigned)d * G, ra[d], rb[d]);

    const unsigned wcol = (unsigned)(lr ^ (lc << 2));
    for (unsigned tb = blockIdx.x; tb < ntiles; tb += (unsigned)DEPTH * G) {
#pragma unroll
        for (int d = 0; d < DEPTH; ++d) {
            const unsigned t = tb + (unsigned)d * G;
            if (t >= ntiles) break;
            {
                const unsigned aw[4] = {ra[d].x, ra[d].y, ra[d].z, ra[d].w}, bw[4] = {rb[d].x, rb[d].y, rb[d].z, rb[d].w};
                unsigned *row = T + (unsigned)(8 * lc) * 32u + wcol;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    row[(2 * i) * 32] = __byte_perm(aw[i], bw[i], 0x5410);        // (a[q][x],   a[q+1][x])
                    row[(2 * i + 1) * 32] = __byte_perm(aw[i], bw[i], 0x7632);    // (a[q][x+1], a[q+1][x+1])
                }
            }
            load(t + (unsigned)DEPTH * G, ra[d], rb[d]);
            __syncthreads();
            const unsigned t2 = t / tiles_x;
            const int tx0 = (int)(t - t2 * tiles_x) * 64;
            const unsigned oc = t2 / tiles_q;
            const int tq0 = (int)(t2 - oc * tiles_q) * 64;
            const int g = tid & 7;
            const int qcol = tq0 + 8 * g;
            if (qcol < nq) {
                int16_t *ob = out + (size_t)oc * sc_out;
                const bool hi_ok = qcol + 4 < nq;            // always true when W16
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const int x = (tid >> 3) + 32 * h;
                    if (tx0 + x < nxin) {
                        uint4 w = *reinterpret_cast<const uint4 *>(T + x * 32 + 4 * (g ^ (x >> 3)));
                        if (f2) w = make_uint4(swap_halves(w.w), swap_halves(w.z), swap_halves(w.y), swap_halves(w.x));
                        const int xo = fa ? nxin - 1 - (tx0 + x) : tx0 + x;
                        int16_t *orow = ob + (size_t)xo * sa;
                        if (W16) {
                            *reinterpret_cast<uint4 *>(orow + (f2 ? nq - 8 - qcol : qcol)) = w;
                        } else if (!f2) {
                            *reinterpret_cast<uint2 *>(orow + qcol) = make_uint2(w.x, w.y);
                            if (hi_ok) *reinterpret_cast<uint2 *>(orow + qcol + 4) = make_uint2(w.z, w.w);
                        } else {                              // reversed: voxels q..q+3 are the upper half of w
                            *reinterpret_cast<uint2 *>(orow + (nq - 4 - qcol)) = make_uint2(w.z, w.w);
                            if (hi_ok) *reinterpret_cast<uint2 *>(orow + (nq - 8 - qcol)) = make_uint2(w.x, w.y);
                        }
                    }
                }
            }
            __syncthreads();
        }
    }
}

}  // namespace mia
