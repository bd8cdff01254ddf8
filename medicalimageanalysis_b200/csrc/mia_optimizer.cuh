// Device-resident optimiser step of the rigid registration loop.
//
// The reference has no intensity-based rigid optimiser (rigid = ICP or manual nudges,
// structure/rigid.py:950-989); the parameter-update semantics restated here are the reference's:
// 'xyz' Euler degrees about a pivot, composed as T(+c) @ R @ T(-c) (rigid.py:957-965), translations in
// mm (rigid.py:977-980).  The algorithm is a regular-step gradient descent with central differences
// (13 probes per iteration) and step halving when the cost rises -- the same one the host optimiser in
// medicalimageanalysis_b200/registration.py runs, so CPU-oracle and GPU trajectories can be compared.
//
// Running it as a one-thread kernel between two metric launches removes the host round trip
// (D2H of 13 scalars + H2D of 13 matrices + two syncs, ~30 us) from every iteration, and lets the whole
// N-iteration loop be captured in one CUDA graph.
#pragma once
#include <cuda_runtime.h>

namespace mia {

constexpr int kOptProbes = 13;
constexpr int kOptMaxIters = 4096;

struct OptState {
    double params[6], prev_params[6], delta[6];
    double center[3];
    double P2W[16], W2P[16];
    double scale, min_scale, prev_cost;
    int iter, converged, maximize, has_prev, n_reject, pad;
    double hist_cost[kOptMaxIters];
    double hist_scale[kOptMaxIters];
    double hist_params[kOptMaxIters][6];
};

__host__ __device__ inline void opt_mat4_mul(const double *a, const double *b, double *c)
{
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) {
            double s = 0.0;
            for (int k = 0; k < 4; ++k) s += a[i * 4 + k] * b[k * 4 + j];
            c[i * 4 + j] = s;
        }
}

// params -> A (3x4): A = (W2P_mov @ T(t) @ T(c) @ Rz Ry Rx @ T(-c) @ P2W_fixed)[:3]
__host__ __device__ inline void opt_params_to_A(const double *prm, const double *center, const double *P2W,
                                                const double *W2P, double *A)
{
    const double d2r = 0.017453292519943295;
    const double cx = cos(prm[0] * d2r), sx = sin(prm[0] * d2r);
    const double cy = cos(prm[1] * d2r), sy = sin(prm[1] * d2r);
    const double cz = cos(prm[2] * d2r), sz = sin(prm[2] * d2r);
    // R = Rz @ Ry @ Rx  (scipy 'xyz' extrinsic)
    double R[9] = {cz * cy, cz * sy * sx - sz * cx, cz * sy * cx + sz * sx,
                   sz * cy, sz * sy * sx + cz * cx, sz * sy * cx - cz * sx,
                   -sy,     cy * sx,                cy * cx};
    double M[16];
    for (int i = 0; i < 3; ++i) {
        for (int j = 0; j < 3; ++j) M[i * 4 + j] = R[i * 3 + j];
        M[i * 4 + 3] = center[i] - (R[i * 3 + 0] * center[0] + R[i * 3 + 1] * center[1] + R[i * 3 + 2] * center[2]) +
                       prm[3 + i];
    }
    M[12] = 0.0; M[13] = 0.0; M[14] = 0.0; M[15] = 1.0;
    double T1[16], T2[16];
    opt_mat4_mul(M, P2W, T1);
    opt_mat4_mul(W2P, T1, T2);
    for (int i = 0; i < 12; ++i) A[i] = T2[i];
}

__device__ inline void opt_emit_probes(const OptState *st, double *dA)
{
    double prm[6];
    for (int q = 0; q < kOptProbes; ++q) {
        for (int i = 0; i < 6; ++i) prm[i] = st->params[i];
        if (q > 0) {
            const int i = (q - 1) >> 1;
            const double h = st->scale * st->delta[i];
            prm[i] += ((q - 1) & 1) ? -h : h;
        }
        opt_params_to_A(prm, st->center, st->P2W, st->W2P, dA + 12 * q);
    }
}

__device__ inline void opt_shift_probes(const double *dA, double *dA_slab, int k0)
{
    for (int q = 0; q < kOptProbes; ++q)
        for (int r = 0; r < 3; ++r) {
            const double *a = dA + 12 * q + 4 * r;
            double *b = dA_slab + 12 * q + 4 * r;
            b[0] = a[0]; b[1] = a[1]; b[2] = a[2];
            b[3] = a[3] + a[2] * (double)k0;
        }
}

__global__ void optimizer_init_kernel(OptState *st, double *dA)
{
    if (threadIdx.x == 0 && blockIdx.x == 0) opt_emit_probes(st, dA);
}

// CNT = unsigned long long (the metric kernels' own counters) or double (the packed, all-reduced buffer of the sharded
// loops).  dA_slab, when given, receives the same probes shifted to a z-slab that starts at fixed plane k0
// (A[:, 3] += A[:, 2] * k0): the sharded loop then needs no kernel of its own for the shift.
template <typename CNT>
__device__ inline void opt_step(OptState *st, const double *metric, const CNT *count, int metric_is_ssd, double *dA,
                                double *dA_slab, int k0)
{
    double c[kOptProbes];
    for (int q = 0; q < kOptProbes; ++q) {
        if (metric_is_ssd) {
            const double n = (double)count[q];
            c[q] = n > 0.0 ? metric[q] / n : 1.0e300;
        } else {
            c[q] = st->maximize ? -metric[q] : metric[q];
        }
    }
    const int it = st->iter;
    if (it < kOptMaxIters) {
        st->hist_cost[it] = c[0];
        st->hist_scale[it] = st->scale;
        for (int i = 0; i < 6; ++i) st->hist_params[it][i] = st->params[i];
    }
    st->iter = it + 1;
    if (!st->converged) {
        if (st->has_prev && c[0] > st->prev_cost) {
            for (int i = 0; i < 6; ++i) st->params[i] = st->prev_params[i];
            st->scale *= 0.5;
            st->n_reject += 1;
            if (st->scale < st->min_scale) st->converged = 1;
        } else {
            for (int i = 0; i < 6; ++i) st->prev_params[i] = st->params[i];
            st->prev_cost = c[0];
            st->has_prev = 1;
            double sg[6], norm2 = 0.0;
            for (int i = 0; i < 6; ++i) {
                const double h = st->scale * st->delta[i];
                const double g = (c[2 * i + 1] - c[2 * i + 2]) / (2.0 * h);
                sg[i] = g * st->delta[i];
                norm2 += sg[i] * sg[i];
            }
            const double norm = sqrt(norm2);
            if (norm > 0.0 && norm < 1.0e300) {
                for (int i = 0; i < 6; ++i) st->params[i] -= st->scale * st->delta[i] * (sg[i] / norm);
            } else {
                st->converged = 1;
            }
        }
    }
    opt_emit_probes(st, dA);
    if (dA_slab) opt_shift_probes(dA, dA_slab, k0);
}

template <typename CNT>
__global__ void optimizer_step_kernel(OptState *st, const double *metric, const CNT *count, int metric_is_ssd, double *dA,
                                      double *dA_slab = nullptr, int k0 = 0)
{
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    opt_step<CNT>(st, metric, count, metric_is_ssd, dA, dA_slab, k0);
}

// ---- fused exchange over NVLink peer memory ------------------------------------------------------
// Every rank owns one FusedBuf in symmetric memory (same layout on every GPU, peer-mapped).  Per iteration, rank r
// stores its 26 packed doubles into slot [parity][r] of EVERY rank's buffer (plain stores through NVLink), fences, and
// then writes the iteration number into flag [parity][r] of every rank.  The optimiser step on each rank waits until
// all `world` flags of the parity carry the iteration number, adds the slots in rank order (deterministic, identical on
// every rank) and steps.  No NCCL call, no host involvement: two small kernels per iteration, capturable.
// Parity = iteration & 1 is enough: a rank can only write iteration n + 1 after its own step n, which needed every
// peer's flag n, which a peer writes only after it finished reading iteration n - 1 (the previous use of that parity).
constexpr int kFusedMaxWorld = 16;

struct FusedBuf {
    double data[2][kFusedMaxWorld][2 * kOptProbes];
    unsigned long long flag[2][kFusedMaxWorld];
    unsigned long long timeouts;          // number of waits that gave up (a peer never arrived): results are then invalid
};

__global__ void fused_push_kernel(const OptState *st, const double *packed_local, FusedBuf *const *peers, int rank, int world)
{
    const unsigned long long seq = (unsigned long long)st->iter + 1ull;
    const int par = (int)(seq & 1ull);
    constexpr int n = 2 * kOptProbes;
    for (int t = threadIdx.x; t < n * world; t += blockDim.x) peers[t / n]->data[par][rank][t % n] = packed_local[t % n];
    __threadfence_system();               // this thread's payload stores are visible system-wide before any flag is
    __syncthreads();
    if ((int)threadIdx.x < world) {
        volatile unsigned long long *f = &peers[threadIdx.x]->flag[par][rank];
        asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(f), "l"(seq) : "memory");
    }
}

__global__ void optimizer_step_fused_kernel(OptState *st, FusedBuf *mine, int world, int metric_is_ssd, double *dA,
                                            double *dA_slab, int k0)
{
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    const unsigned long long seq = (unsigned long long)st->iter + 1ull;
    const int par = (int)(seq & 1ull);
    const long long t0 = clock64();
    for (int r = 0; r < world; ++r) {
        const unsigned long long *f = &mine->flag[par][r];
        unsigned long long v;
        do {
            asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(f) : "memory");
            if (v != seq && clock64() - t0 > 4000000000ll) {       // ~2 s at 2 GHz: a peer is gone; do not hang the GPU
                mine->timeouts += 1ull;
                break;
            }
        } while (v != seq);
    }
    double packed[2 * kOptProbes];
    for (int i = 0; i < 2 * kOptProbes; ++i) {
        double s = 0.0;
        for (int r = 0; r < world; ++r) s += *(volatile const double *)&mine->data[par][r][i];
        packed[i] = s;
    }
    opt_step<double>(st, packed, packed + kOptProbes, metric_is_ssd, dA, dA_slab, k0);
}

// packed[0 .. n_total) = metric, packed[n_total .. 2 n_total) = (double)count, both zero outside this rank's share
// [q0, q0 + n_local): ONE all-reduce(SUM) of 2 * n_total doubles then combines z-slabs (every rank holds all probes,
// q0 = 0) or probe shares (adding zeros is exact).  Counts stay exact in float64 below 2^53 voxels.
__global__ void metric_pack_kernel(const double *metric, const unsigned long long *count, int n_local, int q0, int n_total,
                                   double *packed)
{
    const int q = threadIdx.x;
    if (q >= n_total) return;
    const bool mine = q >= q0 && q < q0 + n_local;
    packed[q] = mine ? metric[q - q0] : 0.0;
    packed[n_total + q] = (mine && count) ? (double)count[q - q0] : 0.0;
}

}  // namespace mia
