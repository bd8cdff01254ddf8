// C-ABI of the pyramid level; see mia_pyramid.cuh and include/mia_b200.h.

extern "C" {

int mia_shrink_mean(const void *d_in, int dtype, const int idims[3], void *d_out, const int factors[3], void *stream)
{
    if (!d_in || !d_out || !factors) return fail(MIA_ERR_ARG, "mia_shrink_mean: null pointer");
    if (!dims_ok(idims)) return fail(MIA_ERR_ARG, "mia_shrink_mean: dims must be positive");
    if (!dtype_ok(dtype)) return fail(MIA_ERR_DTYPE, "mia_shrink_mean: bad dtype");
    for (int a = 0; a < 3; ++a)
        if (factors[a] < 1 || factors[a] > 8) return fail(MIA_ERR_ARG, "mia_shrink_mean: factors must be 1..8");
    const int ox = (idims[0] + factors[0] - 1) / factors[0], oy = (idims[1] + factors[1] - 1) / factors[1],
              oz = (idims[2] + factors[2] - 1) / factors[2];
    const unsigned long long n = (unsigned long long)ox * oy * oz;
    unsigned long long gb = (n + 255) / 256;
    const unsigned long long cap = (unsigned long long)sm_count() * 16;
    const unsigned grid = (unsigned)(gb < cap ? gb : cap);
    cudaStream_t s = (cudaStream_t)stream;
    const bool x2 = factors[0] == 2 && idims[0] % 2 == 0;
    if (dtype == MIA_I16) {
        if (x2 && ((uintptr_t)d_in & 3u) == 0)
            mia::shrink_mean_x2_kernel<int16_t, short2><<<grid, 256, 0, s>>>(static_cast<const int16_t *>(d_in), idims[0], idims[1], idims[2],
                                                                            static_cast<int16_t *>(d_out), ox, oy, oz, factors[1], factors[2]);
        else
            mia::shrink_mean_kernel<int16_t><<<grid, 256, 0, s>>>(static_cast<const int16_t *>(d_in), idims[0], idims[1], idims[2],
                                                                  static_cast<int16_t *>(d_out), ox, oy, oz, factors[0], factors[1], factors[2]);
    } else {
        if (x2 && ((uintptr_t)d_in & 7u) == 0)
            mia::shrink_mean_x2_kernel<float, float2><<<grid, 256, 0, s>>>(static_cast<const float *>(d_in), idims[0], idims[1], idims[2],
                                                                          static_cast<float *>(d_out), ox, oy, oz, factors[1], factors[2]);
        else
            mia::shrink_mean_kernel<float><<<grid, 256, 0, s>>>(static_cast<const float *>(d_in), idims[0], idims[1], idims[2],
                                                                static_cast<float *>(d_out), ox, oy, oz, factors[0], factors[1], factors[2]);
    }
    CUDA_TRY(cudaGetLastError());
    return MIA_OK;
}

}  // extern "C"
