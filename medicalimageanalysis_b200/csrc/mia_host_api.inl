// Second half of the C ABI: permutation, auto-crop helper, optimiser entry points, host-buffer pipeline.
// Included at the end of mia_b200.cu.

namespace {

void mat4_mul_host(const double *a, const double *b, double *c) { opt_mat4_mul(a, b, c); }

bool mat4_inv_host(const double *m, double *inv)
{
    // Gauss-Jordan with partial pivoting on a 4x8 augmented matrix
    double a[4][8];
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) {
            a[i][j] = m[i * 4 + j];
            a[i][j + 4] = i == j ? 1.0 : 0.0;
        }
    for (int c = 0; c < 4; ++c) {
        int piv = c;
        for (int r = c + 1; r < 4; ++r)
            if (std::fabs(a[r][c]) > std::fabs(a[piv][c])) piv = r;
        if (std::fabs(a[piv][c]) < 1e-300) return false;
        if (piv != c)
            for (int j = 0; j < 8; ++j) std::swap(a[c][j], a[piv][j]);
        const double d = a[c][c];
        for (int j = 0; j < 8; ++j) a[c][j] /= d;
        for (int r = 0; r < 4; ++r)
            if (r != c) {
                const double f = a[r][c];
                if (f != 0.0)
                    for (int j = 0; j < 8; ++j) a[r][j] -= f * a[c][j];
            }
    }
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) inv[i * 4 + j] = a[i][j + 4];
    return true;
}

}  // namespace

extern "C" {

int mia_autocrop_extent(const double M[16], const double P2W_mov[16], const int mdims[3],
                        const double out_spacing[3], int extent_rule, int odims[3], double oorigin[3])
{
    if (!M || !P2W_mov || !out_spacing || !odims || !oorigin) return fail(MIA_ERR_ARG, "mia_autocrop_extent: null pointer");
    if (!dims_ok(mdims)) return fail(MIA_ERR_ARG, "mia_autocrop_extent: dims must be positive");
    double Minv[16];
    if (!mat4_inv_host(M, Minv)) return fail(MIA_ERR_ARG, "mia_autocrop_extent: singular matrix");
    double T[16];
    mat4_mul_host(Minv, P2W_mov, T);
    double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
    for (int c = 0; c < 8; ++c) {
        const double idx[4] = {(double)((mdims[0] - 1) * (c & 1)), (double)((mdims[1] - 1) * ((c >> 1) & 1)),
                               (double)((mdims[2] - 1) * ((c >> 2) & 1)), 1.0};
        double q[4];
        for (int i = 0; i < 4; ++i) q[i] = T[i * 4] * idx[0] + T[i * 4 + 1] * idx[1] + T[i * 4 + 2] * idx[2] + T[i * 4 + 3];
        for (int a = 0; a < 3; ++a) {
            const double v = q[a] / q[3];
            if (v < lo[a]) lo[a] = v;
            if (v > hi[a]) hi[a] = v;
        }
    }
    for (int a = 0; a < 3; ++a) {
        if (!(out_spacing[a] > 0.0)) return fail(MIA_ERR_ARG, "mia_autocrop_extent: spacing must be positive");
        const double e = std::fabs((hi[a] - lo[a]) / out_spacing[a]);
        const double last = extent_rule == 0 ? std::floor(e + 0.5) : std::ceil(e);
        if (last > 2.0e9) return fail(MIA_ERR_ARG, "mia_autocrop_extent: extent overflow");
        odims[a] = (int)last + 1;
        oorigin[a] = lo[a];
    }
    return MIA_OK;
}

int mia_permute_axes(const void *d_in, int dtype, const int idims[3], void *d_out, const int perm[3],
                     const int flip[3], void *stream)
{
    if (!d_in || !d_out || !perm || !flip) return fail(MIA_ERR_ARG, "mia_permute_axes: null pointer");
    if (!dims_ok(idims)) return fail(MIA_ERR_ARG, "mia_permute_axes: dims must be positive");
    if (!dtype_ok(dtype)) return fail(MIA_ERR_DTYPE, "mia_permute_axes: bad dtype");
    int seen[3] = {0, 0, 0};
    for (int a = 0; a < 3; ++a) {
        if (perm[a] < 0 || perm[a] > 2 || seen[perm[a]]) return fail(MIA_ERR_ARG, "mia_permute_axes: perm is not a permutation");
        seen[perm[a]] = 1;
    }
    PermParams p;
    p.in = d_in;
    p.out = d_out;
    p.idim[0] = idims[2]; p.idim[1] = idims[1]; p.idim[2] = idims[0];
    for (int a = 0; a < 3; ++a) {
        p.perm[a] = perm[a];
        p.flip[a] = flip[a] ? 1 : 0;
        p.odim[a] = p.idim[perm[a]];
    }
    cudaStream_t s = (cudaStream_t)stream;
    const int sms = sm_count();
    if (perm[2] == 2) {
        const long long rows = (long long)p.odim[0] * p.odim[1];
        int bx = p.odim[2] >= 256 ? 256 : (p.odim[2] >= 128 ? 128 : (p.odim[2] >= 64 ? 64 : 32));
        dim3 block(bx, 256 / bx);
        long long g = (rows + block.y - 1) / block.y;
        if (g > (long long)sms * 32) g = (long long)sms * 32;
        const bool vec = dtype == MIA_I16 && p.odim[2] % 8 == 0 && (((uintptr_t)d_in | (uintptr_t)d_out) & 15u) == 0;
        if (vec) {
            const unsigned cpr = (unsigned)p.odim[2] / 8u;
            const unsigned long long nchunks = (unsigned long long)rows * cpr;
            unsigned long long gb = (nchunks + 255u) / 256u;
            if (gb > (unsigned long long)sms * 16) gb = (unsigned long long)sms * 16;
            permute_rows_vec_i16_kernel<<<(unsigned)gb, 256, 0, s>>>(p, cpr, nchunks);
        } else if (dtype == MIA_I16) permute_rows_kernel<int16_t><<<(unsigned)g, block, 0, s>>>(p);
        else permute_rows_kernel<float><<<(unsigned)g, block, 0, s>>>(p);
    } else {
        const int pa = perm[0] == 2 ? 0 : 1;   // output axis fed by input x
        const int pc = 1 - pa;                 // remaining (batch) output axis
        const long long tiles = (long long)((p.idim[2] + 31) / 32) * ((p.idim[perm[2]] + 31) / 32) * p.odim[pc];
        long long g = tiles < (long long)sms * 32 ? tiles : (long long)sms * 32;
        dim3 block(32, 8);
        const bool vec = dtype == MIA_I16 && p.idim[2] % 2 == 0 && p.idim[perm[2]] % 2 == 0 &&
                         (((uintptr_t)d_in | (uintptr_t)d_out) & 3u) == 0;
        if (vec) {
            // persistent blocks, DEPTH tiles of loads in flight per block (measured choice; env overrides for tuning)
            const long long vt = (long long)((p.idim[2] + 63) / 64) * ((p.idim[perm[2]] + 63) / 64) * p.odim[pc];
            static const int depth = [] { const char *e = getenv("MIA_TR_DEPTH"); return e ? atoi(e) : 2; }();
            static const int bps = [] { const char *e = getenv("MIA_TR_BLOCKS"); return e ? atoi(e) : 4; }();
            long long vg = (long long)sms * (bps > 0 ? bps : 4);
            if (vg > vt) vg = vt;
            const int nqd = p.idim[perm[2]];
            const bool v16 = p.idim[2] % 8 == 0 && nqd % 4 == 0 && vt < 0x7fffffffLL &&
                             (((uintptr_t)d_in | (uintptr_t)d_out) & 15u) == 0;
            const bool w16 = nqd % 8 == 0;
            if (!v16) permute_transpose_vec_i16_kernel<1><<<(unsigned)vg, 256, 0, s>>>(p, pa, pc);
            else if (!w16) permute_transpose_v16_i16_kernel<2, false><<<(unsigned)vg, 256, 0, s>>>(p, pa, pc);
            else if (depth == 1) permute_transpose_v16_i16_kernel<1, true><<<(unsigned)vg, 256, 0, s>>>(p, pa, pc);
            else if (depth == 3) permute_transpose_v16_i16_kernel<3, true><<<(unsigned)vg, 256, 0, s>>>(p, pa, pc);
            else permute_transpose_v16_i16_kernel<2, true><<<(unsigned)vg, 256, 0, s>>>(p, pa, pc);
        } else if (dtype == MIA_I16) permute_transpose_kernel<int16_t><<<(unsigned)g, block, 0, s>>>(p, pa, pc);
        else permute_transpose_kernel<float><<<(unsigned)g, block, 0, s>>>(p, pa, pc);
    }
    CUDA_TRY(cudaGetLastError());
    return MIA_OK;
}

size_t mia_optimizer_state_bytes(void) { return (sizeof(OptState) + 255) & ~(size_t)255; }

int mia_optimizer_init(void *d_state, const double params0[6], const double center[3],
                       const double P2W_fixed[16], const double W2P_mov[16], double rot_step_deg,
                       double trans_step_mm, double min_step_scale, int maximize, double *d_A, void *stream)
{
    if (!d_state || !params0 || !center || !P2W_fixed || !W2P_mov || !d_A)
        return fail(MIA_ERR_ARG, "mia_optimizer_init: null pointer");
    if (!(rot_step_deg > 0.0) || !(trans_step_mm > 0.0) || !(min_step_scale > 0.0))
        return fail(MIA_ERR_ARG, "mia_optimizer_init: steps must be positive");
    cudaStream_t s = (cudaStream_t)stream;
    // build the fixed-size head of the state on the host; the history tail is zeroed on the device
    std::vector<char> head(offsetof(OptState, hist_cost), 0);
    OptState *h = reinterpret_cast<OptState *>(head.data());
    for (int i = 0; i < 6; ++i) {
        h->params[i] = params0[i];
        h->prev_params[i] = params0[i];
        h->delta[i] = i < 3 ? rot_step_deg : trans_step_mm;
    }
    for (int i = 0; i < 3; ++i) h->center[i] = center[i];
    for (int i = 0; i < 16; ++i) {
        h->P2W[i] = P2W_fixed[i];
        h->W2P[i] = W2P_mov[i];
    }
    h->scale = 1.0;
    h->min_scale = min_step_scale;
    h->prev_cost = 0.0;
    h->maximize = maximize ? 1 : 0;
    CUDA_TRY(cudaMemsetAsync(d_state, 0, sizeof(OptState), s));
    CUDA_TRY(cudaMemcpyAsync(d_state, head.data(), head.size(), cudaMemcpyHostToDevice, s));
    CUDA_TRY(cudaStreamSynchronize(s));   // `head` is a stack-lifetime host buffer
    optimizer_init_kernel<<<1, 32, 0, s>>>(static_cast<OptState *>(d_state), d_A);
    CUDA_TRY(cudaGetLastError());
    return MIA_OK;
}

int mia_optimizer_step(void *d_state, const double *d_metric, const unsigned long long *d_count,
                       int metric_is_ssd, double *d_A, void *stream)
{
    if (!d_state || !d_metric || !d_A) return fail(MIA_ERR_ARG, "mia_optimizer_step: null pointer");
    if (metric_is_ssd && !d_count) return fail(MIA_ERR_ARG, "mia_optimizer_step: SSD needs d_count");
    optimizer_step_kernel<unsigned long long><<<1, 32, 0, (cudaStream_t)stream>>>(static_cast<OptState *>(d_state), d_metric,
                                                                                  d_count, metric_is_ssd, d_A);
    CUDA_TRY(cudaGetLastError());
    return MIA_OK;
}

int mia_metric_pack(const double *d_metric, const unsigned long long *d_count, int n_local, int q0, int n_total,
                    double *d_packed, void *stream)
{
    if (!d_packed || (n_local > 0 && !d_metric)) return fail(MIA_ERR_ARG, "mia_metric_pack: null pointer");
    if (n_total < 1 || n_total > MIA_MAX_MATS || n_local < 0 || q0 < 0 || q0 + n_local > n_total)
        return fail(MIA_ERR_ARG, "mia_metric_pack: need 0 <= q0, q0 + n_local <= n_total <= %d", MIA_MAX_MATS);
    metric_pack_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(d_metric, d_count, n_local, q0, n_total, d_packed);
    CUDA_TRY(cudaGetLastError());
    return MIA_OK;
}

int mia_optimizer_step_packed(void *d_state, const double *d_packed, int metric_is_ssd, double *d_A, double *d_A_slab,
                              int k0, void *stream)
{
    if (!d_state || !d_packed || !d_A) return fail(MIA_ERR_ARG, "mia_optimizer_step_packed: null pointer");
    optimizer_step_kernel<double><<<1, 32, 0, (cudaStream_t)stream>>>(static_cast<OptState *>(d_state), d_packed,
                                                                      d_packed + kOptProbes, metric_is_ssd, d_A, d_A_slab, k0);
    CUDA_TRY(cudaGetLastError());
    return MIA_OK;
}

size_t mia_fused_exchange_bytes(void) { return (sizeof(FusedBuf) + 255) & ~(size_t)255; }

int mia_fused_push(const void *d_state, const double *d_packed, void *const *d_peer_bufs, int rank, int world, void *stream)
{
    if (!d_state || !d_packed || !d_peer_bufs) return fail(MIA_ERR_ARG, "mia_fused_push: null pointer");
    if (world < 1 || world > kFusedMaxWorld || rank < 0 || rank >= world)
        return fail(MIA_ERR_ARG, "mia_fused_push: need 0 <= rank < world <= %d", kFusedMaxWorld);
    fused_push_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(static_cast<const OptState *>(d_state), d_packed,
                                                          reinterpret_cast<FusedBuf *const *>(d_peer_bufs), rank, world);
    CUDA_TRY(cudaGetLastError());
    return MIA_OK;
}

int mia_optimizer_step_fused(void *d_state, void *d_my_buf, int world, int metric_is_ssd, double *d_A, double *d_A_slab,
                             int k0, void *stream)
{
    if (!d_state || !d_my_buf || !d_A) return fail(MIA_ERR_ARG, "mia_optimizer_step_fused: null pointer");
    if (world < 1 || world > kFusedMaxWorld) return fail(MIA_ERR_ARG, "mia_optimizer_step_fused: world must be 1..%d", kFusedMaxWorld);
    optimizer_step_fused_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(static_cast<OptState *>(d_state), static_cast<FusedBuf *>(d_my_buf),
                                                                   world, metric_is_ssd, d_A, d_A_slab, k0);
    CUDA_TRY(cudaGetLastError());
    return MIA_OK;
}

int mia_fused_timeouts(const void *d_my_buf, unsigned long long *timeouts, void *stream)
{
    if (!d_my_buf || !timeouts) return fail(MIA_ERR_ARG, "mia_fused_timeouts: null pointer");
    CUDA_TRY(cudaMemcpyAsync(timeouts, &static_cast<const FusedBuf *>(d_my_buf)->timeouts, sizeof(unsigned long long),
                             cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    CUDA_TRY(cudaStreamSynchronize((cudaStream_t)stream));
    return MIA_OK;
}

int mia_optimizer_read(const void *d_state, double params[6], double *scale, int *iters, int *converged,
                       int *n_reject, double *hist_cost, double *hist_params, double *hist_scale,
                       int max_hist, void *stream)
{
    if (!d_state) return fail(MIA_ERR_ARG, "mia_optimizer_read: null pointer");
    cudaStream_t s = (cudaStream_t)stream;
    std::vector<char> head(offsetof(OptState, hist_cost));
    CUDA_TRY(cudaMemcpyAsync(head.data(), d_state, head.size(), cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaStreamSynchronize(s));
    const OptState *h = reinterpret_cast<const OptState *>(head.data());
    if (params) for (int i = 0; i < 6; ++i) params[i] = h->params[i];
    if (scale) *scale = h->scale;
    if (iters) *iters = h->iter;
    if (converged) *converged = h->converged;
    if (n_reject) *n_reject = h->n_reject;
    int n = h->iter < kOptMaxIters ? h->iter : kOptMaxIters;
    if (n > max_hist) n = max_hist;
    const char *base = static_cast<const char *>(d_state);
    if (n > 0 && hist_cost)
        CUDA_TRY(cudaMemcpyAsync(hist_cost, base + offsetof(OptState, hist_cost), sizeof(double) * n, cudaMemcpyDeviceToHost, s));
    if (n > 0 && hist_scale)
        CUDA_TRY(cudaMemcpyAsync(hist_scale, base + offsetof(OptState, hist_scale), sizeof(double) * n, cudaMemcpyDeviceToHost, s));
    if (n > 0 && hist_params)
        CUDA_TRY(cudaMemcpyAsync(hist_params, base + offsetof(OptState, hist_params), sizeof(double) * 6 * n, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaStreamSynchronize(s));
    return MIA_OK;
}

}  // extern "C"

#include "mia_host_pipeline.inl"
