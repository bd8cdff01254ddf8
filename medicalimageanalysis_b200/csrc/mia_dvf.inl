// C-ABI of the displacement-field kernels; see mia_dvf.cuh and include/mia_b200.h.

namespace {

template <typename TM, typename TO>
void launch_warp_dvf(const mia::DvfParams &p, int mode, unsigned grid, cudaStream_t s)
{
    if (mode == MIA_MODE_VTK) mia::warp_dvf_kernel<TM, TO, true><<<grid, mia::kDvfThreads, 0, s>>>(p);
    else mia::warp_dvf_kernel<TM, TO, false><<<grid, mia::kDvfThreads, 0, s>>>(p);
}

}  // namespace

extern "C" {

int mia_warp_dvf(const void *d_mov, int mov_dtype, const int mdims[3], const float *d_dvf, void *d_out, int out_dtype,
                 const int odims[3], const double A[12], const double B[9], double background, int mode, void *stream)
{
    if (!d_mov || !d_dvf || !d_out || !A || !B) return fail(MIA_ERR_ARG, "mia_warp_dvf: null pointer");
    if (!dims_ok(mdims) || !dims_ok(odims)) return fail(MIA_ERR_ARG, "mia_warp_dvf: dims must be positive");
    if (!dtype_ok(mov_dtype) || !dtype_ok(out_dtype)) return fail(MIA_ERR_DTYPE, "mia_warp_dvf: bad dtype");
    if (mov_dtype != out_dtype)
        return fail(MIA_ERR_UNSUPPORTED, "mia_warp_dvf: output must have the image dtype (sitk.Resample keeps the pixel type)");
    if (mode != MIA_MODE_VTK && mode != MIA_MODE_ITK) return fail(MIA_ERR_ARG, "mia_warp_dvf: bad mode");
    for (int i = 0; i < 12; ++i)
        if (!std::isfinite(A[i])) return fail(MIA_ERR_ARG, "mia_warp_dvf: non-finite matrix");
    for (int i = 0; i < 9; ++i)
        if (!std::isfinite(B[i])) return fail(MIA_ERR_ARG, "mia_warp_dvf: non-finite matrix");
    mia::DvfParams p;
    p.mov.p = d_mov;
    p.mov.nx = mdims[0]; p.mov.ny = mdims[1]; p.mov.nz = mdims[2];
    p.dvf = d_dvf;
    p.out = d_out;
    p.ox = odims[0]; p.oy = odims[1]; p.oz = odims[2];
    memcpy(p.A.a, A, sizeof(double) * 12);
    memcpy(p.B, B, sizeof(double) * 9);
    p.background = (float)background;
    p.round_mode = mode == MIA_MODE_VTK ? 0 : 1;
    const long long nwork = (long long)((odims[0] + 31) / 32) * ((odims[1] + 7) / 8) * ((odims[2] + mia::kDvfTZ - 1) / mia::kDvfTZ);
    long long gx = (long long)sm_count() * 8;
    if (gx > nwork) gx = nwork;
    cudaStream_t s = (cudaStream_t)stream;
    if (mov_dtype == MIA_I16) launch_warp_dvf<int16_t, int16_t>(p, mode, (unsigned)gx, s);
    else launch_warp_dvf<float, float>(p, mode, (unsigned)gx, s);
    CUDA_TRY(cudaGetLastError());
    return MIA_OK;
}

int mia_sample_field_points(const float *d_field, const int dims[3], int channels, const double *d_pts, long long n,
                            double scale, double *d_out, void *stream)
{
    if (n < 0) return fail(MIA_ERR_ARG, "mia_sample_field_points: negative point count");
    if (!dims_ok(dims)) return fail(MIA_ERR_ARG, "mia_sample_field_points: dims must be positive");
    if (channels < 1 || channels > 16) return fail(MIA_ERR_ARG, "mia_sample_field_points: channels must be 1..16");
    if (n == 0) return MIA_OK;                 // an empty point set has no buffers to check
    if (!d_field || !d_pts || !d_out) return fail(MIA_ERR_ARG, "mia_sample_field_points: null pointer");
    long long gx = (n + mia::kDvfThreads - 1) / mia::kDvfThreads;
    const long long cap = (long long)sm_count() * 16;
    if (gx > cap) gx = cap;
    mia::sample_field_points_kernel<<<(unsigned)gx, mia::kDvfThreads, 0, (cudaStream_t)stream>>>(
        d_field, dims[0], dims[1], dims[2], channels, d_pts, n, scale, d_out);
    CUDA_TRY(cudaGetLastError());
    return MIA_OK;
}

}  // extern "C"
