// Shared device/host helpers for the nonmyopia_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include "../../include/nonmyopia_b200.h"
#include "fastmath.cuh"

namespace nmb {

// ---------------------------------------------------------------- errors
void set_error(const char* fmt, ...);
int  check_cuda(cudaError_t e, const char* what);

#define NMB_REQUIRE(cond, code, ...)            \
    do {                                        \
        if (!(cond)) {                          \
            nmb::set_error(__VA_ARGS__);        \
            return (code);                      \
        }                                       \
    } while (0)

#define NMB_CUDA(expr)                                              \
    do {                                                            \
        int _rc = nmb::check_cuda((expr), #expr);                   \
        if (_rc) return _rc;                                        \
    } while (0)

#define NMB_LAUNCH_CHECK() NMB_CUDA(cudaGetLastError())

int sm_count();

// ---------------------------------------------------------------- math
template <typename T> struct Num;
template <> struct Num<double> {
    // straight-line DFMA kernels (fastmath.cuh): libm's versions cost 3-6x the instructions
    static __device__ __forceinline__ void sincospi(double x, double* s, double* c) { nmb_sincospi(x, s, c); }
    static __device__ __forceinline__ double exp(double x) { return nmb_exp(x); }
    static __device__ __forceinline__ double sqrt(double x) { return ::sqrt(x); }
    static __device__ __forceinline__ double tanh(double x) { return nmb_tanh(x); }
    static __device__ __forceinline__ double expm1(double x) { return nmb_expm1(x); }
    static __device__ __forceinline__ double fabs(double x) { return ::fabs(x); }
    static __device__ __forceinline__ double rsqrt(double x) { return 1.0 / ::sqrt(x); }
    static constexpr double kPi = 3.14159265358979323846;
    static constexpr double kChJitter = 1e-8;   // psd_safe_cholesky jitter (GPyTorch, double)
};
template <> struct Num<float> {
    static __device__ __forceinline__ void sincospi(float x, float* s, float* c) { ::sincospif(x, s, c); }
    static __device__ __forceinline__ float exp(float x) { return ::expf(x); }
    static __device__ __forceinline__ float sqrt(float x) { return ::sqrtf(x); }
    static __device__ __forceinline__ float tanh(float x) { return ::tanhf(x); }
    static __device__ __forceinline__ float expm1(float x) { return ::expm1f(x); }
    static __device__ __forceinline__ float fabs(float x) { return ::fabsf(x); }
    static __device__ __forceinline__ float rsqrt(float x) { return ::rsqrtf(x); }
    static constexpr float kPi = 3.14159265358979323846f;
    static constexpr float kChJitter = 1e-6f;
};

template <typename T> __device__ __forceinline__ T sigmoid_(T x) {
    return T(1) / (T(1) + Num<T>::exp(-x));
}
template <> __device__ __forceinline__ double sigmoid_<double>(double x) { return nmb_sigmoid(x); }
template <typename T> __device__ __forceinline__ T elu_(T x) {
    return x > T(0) ? x : Num<T>::expm1(x);
}
// branch-free: expm1 of min(x, 0) is evaluated for every lane (no divergence), selected afterwards
template <> __device__ __forceinline__ double elu_<double>(double x) {
    const double e = nmb_expm1(fmin(x, 0.0));
    return x > 0.0 ? x : e;
}

// Stationary kernel profile.  r2 = sum_k ((x_k - x'_k)/ls_k)^2.
// Returns kappa(r2) (unit amplitude) and gcoef with
//   d kappa / d x_k = gcoef * ((x_k - x'_k)/ls_k) / ls_k .
// GPyTorch forms the distance as sqrt(clamp_min(r2, 1e-30)); below the clamp the
// reference's autograd gradient is zero - reproduced for Matern-1/2, whose
// gradient is otherwise singular at r = 0.
template <typename T>
__device__ __forceinline__ void kernel_profile(int kernel_id, T r2, T& kappa, T& gcoef) {
    if (kernel_id == NMB_KERNEL_RBF) {
        kappa = Num<T>::exp(T(-0.5) * r2);
        gcoef = -kappa;
        return;
    }
    const bool clamped = r2 < T(1e-30);
    const T r = Num<T>::sqrt(clamped ? T(1e-30) : r2);
    if (kernel_id == NMB_KERNEL_MATERN52) {
        const T s5 = T(2.23606797749978969641);
        const T e = Num<T>::exp(-s5 * r);
        const T lin = T(1) + s5 * r;
        kappa = (lin + T(5.0 / 3.0) * r * r) * e;
        gcoef = T(-5.0 / 3.0) * lin * e;
    } else if (kernel_id == NMB_KERNEL_MATERN32) {
        const T s3 = T(1.73205080756887729353);
        const T e = Num<T>::exp(-s3 * r);
        kappa = (T(1) + s3 * r) * e;
        gcoef = T(-3) * e;
    } else {  // Matern-1/2
        kappa = Num<T>::exp(-r);
        gcoef = clamped ? T(0) : -kappa / r;
    }
}

// table-driven variant of kernel_profile (FP64, `tab` = shared-memory copy of g_fast_tab)
__device__ __forceinline__ void kernel_profile_tab(int kernel_id, double r2, const double* __restrict__ tab, double& kappa,
                                                   double& gcoef) {
    if (kernel_id == NMB_KERNEL_RBF) {
        kappa = nmb_exp_tab(-0.5 * r2, tab);
        gcoef = -kappa;
        return;
    }
    const bool clamped = r2 < 1e-30;
    const double r = ::sqrt(clamped ? 1e-30 : r2);
    if (kernel_id == NMB_KERNEL_MATERN52) {
        const double s5 = 2.23606797749978969641;
        const double e = nmb_exp_tab(-s5 * r, tab);
        const double lin = 1.0 + s5 * r;
        kappa = (lin + (5.0 / 3.0) * r * r) * e;
        gcoef = (-5.0 / 3.0) * lin * e;
    } else if (kernel_id == NMB_KERNEL_MATERN32) {
        const double s3 = 1.73205080756887729353;
        const double e = nmb_exp_tab(-s3 * r, tab);
        kappa = (1.0 + s3 * r) * e;
        gcoef = -3.0 * e;
    } else {
        kappa = nmb_exp_tab(-r, tab);
        gcoef = clamped ? 0.0 : -kappa / r;
    }
}

// lookup tables of the table-driven transcendentals (fastmath.cuh); kernels copy them into shared memory
static __device__ const double g_fast_tab[NMB_FAST_TAB_LEN] = NMB_FAST_TAB_INIT;
__device__ __forceinline__ void load_fast_tab(double* __restrict__ dst, int tid, int nthr) {
    for (int e = tid; e < NMB_FAST_TAB_LEN; e += nthr) dst[e] = g_fast_tab[e];
}

// ---------------------------------------------------------------- warp helpers
template <typename T> __device__ __forceinline__ T shfl_xor_(T v, int m) {
    return __shfl_xor_sync(0xffffffffu, v, m);
}
// sum over groups of `width` consecutive lanes (width power of two <= 32)
template <typename T> __device__ __forceinline__ T group_sum(T v, int width) {
    for (int m = width >> 1; m > 0; m >>= 1) v += shfl_xor_(v, m);
    return v;
}
template <typename T> __device__ __forceinline__ T warp_sum(T v) { return group_sum(v, 32); }

}  // namespace nmb
