// C-ABI of the masked dose sample and its statistics; see mia_dose.cuh and include/mia_b200.h.

namespace {

struct MaskWsLayout { size_t off_counts, off_offsets, total; unsigned nchunks; };

bool mask_ws_layout(const int odims[3], MaskWsLayout &w)
{
    const unsigned long long n = (unsigned long long)odims[0] * odims[1] * odims[2];
    const unsigned long long nchunks = (n + mia::kMaskChunk - 1) / mia::kMaskChunk;
    if (nchunks > 0x7fffffffull) return false;
    w.nchunks = (unsigned)nchunks;
    w.off_counts = 0;
    w.off_offsets = (sizeof(unsigned) * nchunks + 255) & ~(size_t)255;
    w.total = w.off_offsets + ((sizeof(unsigned long long) * nchunks + 255) & ~(size_t)255);
    return true;
}

constexpr unsigned kStatsGridMax = 148 * 8;

struct StatsWsLayout { size_t part, below, ticket, levels, ranks, prefix, hist1, histn, ssq, total; };

StatsWsLayout stats_ws_layout()
{
    StatsWsLayout w;
    size_t o = 0;
    auto take = [&](size_t bytes) { const size_t at = o; o += (bytes + 255) & ~(size_t)255; return at; };
    w.part = take(sizeof(double) * 4 * kStatsGridMax);
    w.below = take(sizeof(unsigned long long) * mia::kMaxLevels * kStatsGridMax);
    w.ssq = take(sizeof(double) * kStatsGridMax);
    w.levels = take(sizeof(float) * mia::kMaxLevels);
    w.ranks = take(sizeof(unsigned long long) * mia::kMaxRanks);
    // everything from here on is zeroed at the start of every call
    w.ticket = take(sizeof(unsigned));
    w.prefix = take(sizeof(unsigned) * mia::kMaxRanks);
    w.hist1 = take(sizeof(unsigned) * 4096);
    w.histn = take(sizeof(unsigned) * 1024 * mia::kMaxRanks);
    w.total = o;
    return w;
}

}  // namespace

extern "C" {

size_t mia_mask_workspace_bytes(const int odims[3])
{
    MaskWsLayout w;
    if (!dims_ok(odims) || !mask_ws_layout(odims, w)) return 0;
    return w.total;
}

int mia_mask_scan(const unsigned char *d_mask, const int odims[3], unsigned long long *d_total, void *d_ws,
                  size_t ws_bytes, void *stream)
{
    if (!d_mask || !d_total || !d_ws) return fail(MIA_ERR_ARG, "mia_mask_scan: null pointer");
    if (!dims_ok(odims)) return fail(MIA_ERR_ARG, "mia_mask_scan: dims must be positive");
    MaskWsLayout w;
    if (!mask_ws_layout(odims, w)) return fail(MIA_ERR_UNSUPPORTED, "mia_mask_scan: lattice too large");
    if (ws_bytes < w.total) return fail(MIA_ERR_WORKSPACE, "mia_mask_scan: workspace %zu < %zu bytes", ws_bytes, w.total);
    if (((uintptr_t)d_ws & 15) != 0) return fail(MIA_ERR_WORKSPACE, "mia_mask_scan: workspace must be 16-byte aligned");
    cudaStream_t s = (cudaStream_t)stream;
    char *ws = static_cast<char *>(d_ws);
    unsigned *counts = reinterpret_cast<unsigned *>(ws + w.off_counts);
    unsigned long long *offsets = reinterpret_cast<unsigned long long *>(ws + w.off_offsets);
    const unsigned long long n = (unsigned long long)odims[0] * odims[1] * odims[2];
    mia::mask_count_kernel<<<w.nchunks, 256, 0, s>>>(d_mask, n, counts);
    mia::mask_scan_kernel<<<1, 1024, 0, s>>>(counts, w.nchunks, offsets, d_total);
    CUDA_TRY(cudaGetLastError());
    return MIA_OK;
}

int mia_resample_masked(const void *d_mov, int mov_dtype, const int mdims[3], const unsigned char *d_mask,
                        const int odims[3], const double A[12], double background, int mode, const void *d_ws,
                        size_t ws_bytes, float *d_out, void *stream)
{
    if (!d_mov || !d_mask || !A || !d_ws) return fail(MIA_ERR_ARG, "mia_resample_masked: null pointer");
    if (!dims_ok(mdims) || !dims_ok(odims)) return fail(MIA_ERR_ARG, "mia_resample_masked: dims must be positive");
    if (!dtype_ok(mov_dtype)) return fail(MIA_ERR_DTYPE, "mia_resample_masked: bad dtype");
    if (mode != MIA_MODE_VTK && mode != MIA_MODE_ITK) return fail(MIA_ERR_ARG, "mia_resample_masked: bad mode");
    for (int i = 0; i < 12; ++i)
        if (!std::isfinite(A[i])) return fail(MIA_ERR_ARG, "mia_resample_masked: non-finite matrix");
    MaskWsLayout w;
    if (!mask_ws_layout(odims, w)) return fail(MIA_ERR_UNSUPPORTED, "mia_resample_masked: lattice too large");
    if (ws_bytes < w.total) return fail(MIA_ERR_WORKSPACE, "mia_resample_masked: workspace %zu < %zu bytes", ws_bytes, w.total);
    mia::VolView mv;
    mv.p = d_mov; mv.nx = mdims[0]; mv.ny = mdims[1]; mv.nz = mdims[2];
    mia::Mat34 M;
    memcpy(M.a, A, sizeof(double) * 12);
    const unsigned long long *offsets =
        reinterpret_cast<const unsigned long long *>(static_cast<const char *>(d_ws) + w.off_offsets);
    cudaStream_t s = (cudaStream_t)stream;
    const bool closed = mode == MIA_MODE_VTK;
    if (mov_dtype == MIA_I16) {
        if (closed) mia::resample_masked_kernel<int16_t, true><<<w.nchunks, 256, 0, s>>>(mv, d_mask, odims[0], odims[1], odims[2], M, (float)background, offsets, d_out);
        else mia::resample_masked_kernel<int16_t, false><<<w.nchunks, 256, 0, s>>>(mv, d_mask, odims[0], odims[1], odims[2], M, (float)background, offsets, d_out);
    } else {
        if (closed) mia::resample_masked_kernel<float, true><<<w.nchunks, 256, 0, s>>>(mv, d_mask, odims[0], odims[1], odims[2], M, (float)background, offsets, d_out);
        else mia::resample_masked_kernel<float, false><<<w.nchunks, 256, 0, s>>>(mv, d_mask, odims[0], odims[1], odims[2], M, (float)background, offsets, d_out);
    }
    CUDA_TRY(cudaGetLastError());
    return MIA_OK;
}

size_t mia_sample_stats_workspace_bytes(void) { return stats_ws_layout().total; }

int mia_sample_stats(const float *d_values, unsigned long long n, const float *levels, int n_levels,
                     const unsigned long long *ranks, int n_ranks, double *d_result, void *d_ws, size_t ws_bytes,
                     void *stream)
{
    if (!d_result || !d_ws) return fail(MIA_ERR_ARG, "mia_sample_stats: null pointer");
    if (n_levels < 0 || n_levels > mia::kMaxLevels || n_ranks < 0 || n_ranks > mia::kMaxRanks)
        return fail(MIA_ERR_ARG, "mia_sample_stats: at most %d levels and %d ranks", mia::kMaxLevels, mia::kMaxRanks);
    if ((n_levels && !levels) || (n_ranks && !ranks)) return fail(MIA_ERR_ARG, "mia_sample_stats: null levels / ranks");
    if (n == 0) return fail(MIA_ERR_ARG, "mia_sample_stats: empty sample");
    if (!d_values) return fail(MIA_ERR_ARG, "mia_sample_stats: null pointer");
    for (int r = 0; r < n_ranks; ++r)
        if (ranks[r] >= n) return fail(MIA_ERR_ARG, "mia_sample_stats: rank %d out of range", r);
    const StatsWsLayout w = stats_ws_layout();
    if (ws_bytes < w.total) return fail(MIA_ERR_WORKSPACE, "mia_sample_stats: workspace %zu < %zu bytes", ws_bytes, w.total);
    if (((uintptr_t)d_ws & 15) != 0) return fail(MIA_ERR_WORKSPACE, "mia_sample_stats: workspace must be 16-byte aligned");
    cudaStream_t s = (cudaStream_t)stream;
    char *ws = static_cast<char *>(d_ws);
    mia::StatsWs k;
    k.part = reinterpret_cast<double *>(ws + w.part);
    k.part_below = reinterpret_cast<unsigned long long *>(ws + w.below);
    k.part_ssq = reinterpret_cast<double *>(ws + w.ssq);
    k.levels = reinterpret_cast<float *>(ws + w.levels);
    k.ranks = reinterpret_cast<unsigned long long *>(ws + w.ranks);
    k.ticket = reinterpret_cast<unsigned *>(ws + w.ticket);
    k.prefix = reinterpret_cast<unsigned *>(ws + w.prefix);
    k.hist1 = reinterpret_cast<unsigned *>(ws + w.hist1);
    k.histn = reinterpret_cast<unsigned *>(ws + w.histn);
    CUDA_TRY(cudaMemsetAsync(ws + w.ticket, 0, w.total - w.ticket, s));
    if (n_levels) CUDA_TRY(cudaMemcpyAsync(k.levels, levels, sizeof(float) * n_levels, cudaMemcpyHostToDevice, s));
    if (n_ranks) CUDA_TRY(cudaMemcpyAsync(k.ranks, ranks, sizeof(unsigned long long) * n_ranks, cudaMemcpyHostToDevice, s));
    unsigned long long gb = (n + 256 * 8 - 1) / (256 * 8);            // ~8 values per thread before adding blocks
    const unsigned grid = (unsigned)(gb < kStatsGridMax ? gb : kStatsGridMax);
    mia::stats_moments_kernel<<<grid, 256, 0, s>>>(d_values, n, n_levels, k, d_result);
    mia::stats_hist1_kernel<<<grid, 256, 0, s>>>(d_values, n, n_levels, k, d_result);
    if (n_ranks) {
        mia::stats_resolve_kernel<<<n_ranks, 256, 0, s>>>(k, n_ranks, 0, n_levels, d_result);
        mia::stats_histn_kernel<<<grid, 256, 0, s>>>(d_values, n, k, n_ranks, 12);
        mia::stats_resolve_kernel<<<n_ranks, 256, 0, s>>>(k, n_ranks, 1, n_levels, d_result);
        CUDA_TRY(cudaMemsetAsync(k.histn, 0, sizeof(unsigned) * 1024 * mia::kMaxRanks, s));
        mia::stats_histn_kernel<<<grid, 256, 0, s>>>(d_values, n, k, n_ranks, 22);
        mia::stats_resolve_kernel<<<n_ranks, 256, 0, s>>>(k, n_ranks, 2, n_levels, d_result);
    }
    CUDA_TRY(cudaGetLastError());
    return MIA_OK;
}

}  // extern "C"
