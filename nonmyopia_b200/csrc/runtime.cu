// Error plumbing, device queries and pipe-peak micro-benchmarks.
#include <stdarg.h>
#include "common.cuh"

namespace nmb {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int check_cuda(cudaError_t e, const char* what) {
    if (e == cudaSuccess) return 0;
    set_error("CUDA error %d (%s) at %s", (int)e, cudaGetErrorString(e), what);
    return (int)e;
}

int sm_count() {
    static int cached = 0;
    if (cached > 0) return cached;
    int dev = 0, n = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 148;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) return 148;
    cached = n;
    return n;
}

// ---- pipe peaks: 8 independent chains per thread, 4096 iterations
template <typename T, int WHICH>
__global__ void __launch_bounds__(256) peak_kernel(T* out, T seed, int iters) {
    T a[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = seed + T(i) * T(0.001) + T(threadIdx.x) * T(1e-6);
    const T m = T(0.999999), c = T(1e-7);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (WHICH == 0) {
                a[i] = fma(a[i], m, c);
            } else if (WHICH == 1) {
                T s, co;
                Num<T>::sincospi(a[i], &s, &co);
                a[i] = s + co * T(0.5) + T(3.0);
            } else {
                a[i] = Num<T>::exp(a[i] * T(-0.25)) + T(0.5);
            }
        }
    }
    T r = T(0);
#pragma unroll
    for (int i = 0; i < 8; ++i) r += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

template <typename T, int WHICH> static int run_peak(double* host_result) {
    const int sms = sm_count();
    const int blocks = sms * 8, threads = 256;
    const int iters = WHICH == 0 ? 8192 : 512;
    T* buf = nullptr;
    NMB_CUDA(cudaMalloc(&buf, sizeof(T) * blocks * threads));
    cudaEvent_t e0, e1;
    NMB_CUDA(cudaEventCreate(&e0));
    NMB_CUDA(cudaEventCreate(&e1));
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        NMB_CUDA(cudaEventRecord(e0));
        peak_kernel<T, WHICH><<<blocks, threads>>>(buf, T(0.25), iters);
        NMB_CUDA(cudaEventRecord(e1));
        NMB_CUDA(cudaEventSynchronize(e1));
        float ms = 0;
        NMB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        if (rep > 0 && ms < best) best = ms;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(buf);
    const double ops = (double)blocks * threads * 8.0 * iters * (WHICH == 0 ? 2.0 : 1.0);
    *host_result = ops / (best * 1e-3);
    return 0;
}

// ---- device-side evaluation of the fastmath.cuh kernels (accuracy tests: rcp.approx only exists on the device)
__global__ void fastmath_probe_kernel(int which, const double* __restrict__ x, double* __restrict__ y0,
                                      double* __restrict__ y1, long long n) {
    __shared__ __align__(16) double tab[NMB_FAST_TAB_LEN];
    load_fast_tab(tab, threadIdx.x, blockDim.x);
    __syncthreads();
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double v = x[i];
    double a = 0.0, b = 0.0;
    switch (which) {
        case 0: a = nmb_exp_tab(v, tab); b = nmb_exp(v); break;
        case 1: a = nmb_expm1_tab(v, tab); b = nmb_expm1(v); break;
        case 2: nmb_sincospi_tab(v, tab, &a, &b); break;
        case 3: a = nmb_sigmoid_tab(v, tab); b = nmb_sigmoid(v); break;
        case 4: a = nmb_tanh_tab(v, tab); b = nmb_tanh(v); break;
        case 5: a = nmb_rcp2(v); b = nmb_rcp(v); break;
        case 6: a = nmb_elu_tab(v, tab); b = elu_<double>(v); break;
        default: nmb_sincospi(v, &a, &b); break;
    }
    y0[i] = a;
    y1[i] = b;
}

}  // namespace nmb

extern "C" int nmb_fastmath_probe(int which, const void* x, void* y0, void* y1, int64_t n, void* stream) {
    using namespace nmb;
    NMB_REQUIRE(x && y0 && y1 && n >= 0, NMB_EINVAL, "nmb_fastmath_probe: null argument");
    NMB_REQUIRE(which >= 0 && which <= 7, NMB_EINVAL, "nmb_fastmath_probe: which=%d", which);
    if (n == 0) return 0;
    fastmath_probe_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        which, (const double*)x, (double*)y0, (double*)y1, (long long)n);
    NMB_LAUNCH_CHECK();
    return 0;
}

extern "C" const char* nmb_last_error(void) { return nmb::g_err; }
extern "C" int nmb_version(void) { return 100; }
extern "C" int nmb_device_sm_count(void) {
    int dev = 0, n = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) { nmb::set_error("no CUDA device"); return NMB_ENODEV; }
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return NMB_ENODEV;
    return n;
}

extern "C" int nmb_pipe_peak(int dtype, int which, double* host_result) {
    using namespace nmb;
    NMB_REQUIRE(host_result, NMB_EINVAL, "nmb_pipe_peak: null result");
    NMB_REQUIRE(which >= 0 && which <= 2, NMB_EINVAL, "nmb_pipe_peak: which=%d", which);
    if (dtype == NMB_F64) {
        if (which == 0) return run_peak<double, 0>(host_result);
        if (which == 1) return run_peak<double, 1>(host_result);
        return run_peak<double, 2>(host_result);
    } else if (dtype == NMB_F32) {
        if (which == 0) return run_peak<float, 0>(host_result);
        if (which == 1) return run_peak<float, 1>(host_result);
        return run_peak<float, 2>(host_result);
    }
    set_error("nmb_pipe_peak: dtype %d", dtype);
    return NMB_EINVAL;
}
