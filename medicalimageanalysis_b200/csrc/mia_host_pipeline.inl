// Host-buffer entry point: H2D / kernel / D2H pipelined by z-slab over three streams, so the two PCIe
// directions and the kernel overlap.  This is the call the reference boundary maps to (NumPy in,
// NumPy out: structure/rigid.py:672 and :194-199).

namespace {

struct HostCtx {
    bool init = false;
    void *d_mov = nullptr, *d_out = nullptr;
    size_t mov_cap = 0, out_cap = 0;
    cudaStream_t s_in = nullptr, s_k = nullptr, s_out = nullptr;
    std::vector<cudaEvent_t> ev_in, ev_k;
};

constexpr int kMaxDevices = 16;
HostCtx g_host[kMaxDevices];
std::mutex g_host_mu[kMaxDevices];       // one pipeline per device; calls on different devices do not serialise

int ensure_events(std::vector<cudaEvent_t> &v, size_t n)
{
    while (v.size() < n) {
        cudaEvent_t e;
        if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) return -1;
        v.push_back(e);
    }
    return 0;
}

int grow(void **p, size_t *cap, size_t need)
{
    if (*cap >= need) return 0;
    if (*p) cudaFree(*p);
    *p = nullptr;
    *cap = 0;
    // round up so that a sequence of slightly different volumes does not reallocate every call
    size_t want = (need + ((size_t)32 << 20) - 1) & ~(((size_t)32 << 20) - 1);
    if (cudaMalloc(p, want) != cudaSuccess) return -1;
    *cap = want;
    return 0;
}

}  // namespace

extern "C" {

void *mia_host_alloc(size_t bytes)
{
    void *p = nullptr;
    if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocPortable) != cudaSuccess) {
        fail(MIA_ERR_CUDA, "mia_host_alloc: cudaHostAlloc(%zu) failed", bytes);
        return nullptr;
    }
    return p;
}

int mia_host_free(void *p)
{
    if (p) CUDA_TRY(cudaFreeHost(p));
    return MIA_OK;
}

int mia_host_register(void *p, size_t bytes)
{
    if (!p || !bytes) return fail(MIA_ERR_ARG, "mia_host_register: null pointer or empty range");
    CUDA_TRY(cudaHostRegister(p, bytes, cudaHostRegisterPortable));
    return MIA_OK;
}

int mia_host_unregister(void *p)
{
    if (p) CUDA_TRY(cudaHostUnregister(p));
    return MIA_OK;
}

int mia_resample_rigid_host(const void *h_mov, int mov_dtype, const int mdims[3], void *h_out,
                            int out_dtype, const int odims[3], const double A[12], double background,
                            int mode, int device)
{
    if (!h_mov || !h_out || !A) return fail(MIA_ERR_ARG, "mia_resample_rigid_host: null pointer");
    if (!dims_ok(mdims) || !dims_ok(odims)) return fail(MIA_ERR_ARG, "mia_resample_rigid_host: dims must be positive");
    if (!dtype_ok(mov_dtype) || !dtype_ok(out_dtype)) return fail(MIA_ERR_DTYPE, "mia_resample_rigid_host: bad dtype");
    if (device < 0 || device >= kMaxDevices) return fail(MIA_ERR_ARG, "mia_resample_rigid_host: bad device");
    std::lock_guard<std::mutex> lock(g_host_mu[device]);
    int prev = 0;
    CUDA_TRY(cudaGetDevice(&prev));
    CUDA_TRY(cudaSetDevice(device));
    HostCtx &c = g_host[device];
    if (!c.init) {
        CUDA_TRY(cudaStreamCreateWithFlags(&c.s_in, cudaStreamNonBlocking));
        CUDA_TRY(cudaStreamCreateWithFlags(&c.s_k, cudaStreamNonBlocking));
        CUDA_TRY(cudaStreamCreateWithFlags(&c.s_out, cudaStreamNonBlocking));
        c.init = true;
    }
    const size_t mplane = (size_t)mdims[0] * mdims[1] * esize(mov_dtype);
    const size_t oplane = (size_t)odims[0] * odims[1] * esize(out_dtype);
    const size_t mbytes = mplane * mdims[2], obytes = oplane * odims[2];
    if (grow(&c.d_mov, &c.mov_cap, mbytes) || grow(&c.d_out, &c.out_cap, obytes)) {
        cudaSetDevice(prev);
        return fail(MIA_ERR_CUDA, "mia_resample_rigid_host: device allocation failed");
    }
    // slabs of about 8 MiB, at most 64 per side
    auto nslabs = [](size_t bytes, int nz) {
        size_t n = (bytes + ((size_t)8 << 20) - 1) / ((size_t)8 << 20);
        if (n < 1) n = 1;
        if (n > 64) n = 64;
        if (n > (size_t)nz) n = (size_t)nz;
        return (int)n;
    };
    // output slabs start on multiples of the kernels' 8-plane tile depth: every slab then runs the very tiles a
    // one-piece launch runs, and the host path returns the bytes of the device path
    const int otiles = (odims[2] + 7) / 8;
    const int nin = nslabs(mbytes, mdims[2]), nout = nslabs(obytes, otiles);
    if (ensure_events(c.ev_in, nin) || ensure_events(c.ev_k, nout)) {
        cudaSetDevice(prev);
        return fail(MIA_ERR_CUDA, "mia_resample_rigid_host: event creation failed");
    }
    int rc = MIA_OK;
    auto zsplit = [](int nz, int n, int i) { return (int)(((long long)nz * i) / n); };
    for (int i = 0; i < nin && rc == MIA_OK; ++i) {
        const int z0 = zsplit(mdims[2], nin, i), z1 = zsplit(mdims[2], nin, i + 1);
        if (cudaMemcpyAsync((char *)c.d_mov + mplane * z0, (const char *)h_mov + mplane * z0, mplane * (z1 - z0),
                            cudaMemcpyHostToDevice, c.s_in) != cudaSuccess ||
            cudaEventRecord(c.ev_in[i], c.s_in) != cudaSuccess)
            rc = fail(MIA_ERR_CUDA, "mia_resample_rigid_host: H2D failed");
    }
    for (int o = 0; o < nout && rc == MIA_OK; ++o) {
        const int k0 = 8 * zsplit(otiles, nout, o), k1 = o + 1 == nout ? odims[2] : 8 * zsplit(otiles, nout, o + 1);
        // highest moving z plane this output slab can touch (8 corners, +1 for the upper tap)
        double zmax = -1e300;
        for (int cc = 0; cc < 8; ++cc) {
            const double i = (cc & 1) ? odims[0] - 1 : 0, j = (cc & 2) ? odims[1] - 1 : 0, k = (cc & 4) ? k1 - 1 : k0;
            const double z = A[8] * i + A[9] * j + A[10] * k + A[11];
            if (z > zmax) zmax = z;
        }
        long long zneed = (long long)std::floor(zmax) + 2;
        if (zneed < 0) zneed = 0;
        if (zneed > mdims[2] - 1) zneed = mdims[2] - 1;
        int slab = 0;
        while (slab + 1 < nin && zsplit(mdims[2], nin, slab + 1) <= zneed) ++slab;
        double As[12];
        memcpy(As, A, sizeof As);
        for (int a = 0; a < 3; ++a) As[a * 4 + 3] += As[a * 4 + 2] * k0;
        const int od[3] = {odims[0], odims[1], k1 - k0};
        if (cudaStreamWaitEvent(c.s_k, c.ev_in[slab], 0) != cudaSuccess) { rc = fail(MIA_ERR_CUDA, "wait failed"); break; }
        rc = mia_resample_rigid(c.d_mov, mov_dtype, mdims, (char *)c.d_out + oplane * k0, out_dtype, od, As, background,
                                mode, c.s_k);
        if (rc != MIA_OK) break;
        if (cudaEventRecord(c.ev_k[o], c.s_k) != cudaSuccess || cudaStreamWaitEvent(c.s_out, c.ev_k[o], 0) != cudaSuccess ||
            cudaMemcpyAsync((char *)h_out + oplane * k0, (const char *)c.d_out + oplane * k0, oplane * (k1 - k0),
                            cudaMemcpyDeviceToHost, c.s_out) != cudaSuccess)
            rc = fail(MIA_ERR_CUDA, "mia_resample_rigid_host: D2H failed");
    }
    cudaError_t e1 = cudaStreamSynchronize(c.s_in), e2 = cudaStreamSynchronize(c.s_k), e3 = cudaStreamSynchronize(c.s_out);
    cudaSetDevice(prev);
    if (rc != MIA_OK) return rc;
    if (e1 != cudaSuccess || e2 != cudaSuccess || e3 != cudaSuccess)
        return fail(MIA_ERR_CUDA, "mia_resample_rigid_host: %s", cudaGetErrorString(e1 != cudaSuccess ? e1 : (e2 != cudaSuccess ? e2 : e3)));
    return MIA_OK;
}

}  // extern "C"
