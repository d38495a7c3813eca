// Thompson-sample path evaluation, device side (reference acqfs.py:362,409-411).
//   f(x) = c + feat_scale * sum_m [ sin(z_m) w_m + cos(z_m) w_{M/2+m} ] + amp * sum_j kappa(x, X_j) v_j
//   z_m  = pi * sum_k w_phase[m,k] x_k          (w_phase = W / (lengthscale * pi))
// A group of `nl` threads shares one point; every thread returns PARTIAL sums of the
// value and of the d gradient components over its slice of the frequencies and of the
// training points; the caller reduces over the group.  The constants are read
// through the read-only path (L1/L2 resident: <= 32 KB + 8 KB + n*(d+1)*8 B).
#pragma once
#include "common.cuh"

namespace nmb {

// sin(pi x), cos(pi x) in FP32: x = q/2 + r with |r| <= 1/4 exactly (magic-number rounding), then the MUFU sine /
// cosine on pi r in [-pi/4, pi/4] (absolute error ~2^-21) and the quadrant swap - the FP32 / TF32 variant's path
// features (tolerance 1e-3); libm's sincospif costs ~3x the instructions.  Overloads keep the template code generic.
__device__ __forceinline__ void nmb_sincospi_fast(float x, float* sn, float* cs) {
    const float kd = fmaf(x, 2.0f, 12582912.0f);                 // 1.5 * 2^23
    const int q = __float_as_int(kd);
    const float r = fmaf(kd - 12582912.0f, -0.5f, x);
    const float a = 3.14159265358979f * r;
    const float s = __sinf(a), c = __cosf(a);
    const float u = (q & 1) ? c : s, v = (q & 1) ? s : c;
    *sn = (q & 2) ? -u : u;
    *cs = ((q + 1) & 2) ? -v : v;
}
__device__ __forceinline__ void nmb_sincospi_fast(double x, double* sn, double* cs) { nmb_sincospi(x, sn, cs); }

template <typename T> struct PathDev {
    int d, n_freq, n_train, n_paths, kernel_id;
    T feat_scale, amplitude, mean_const;
    const T* __restrict__ w_phase;
    const T* __restrict__ w_feat;
    const T* __restrict__ inv_ls;
    const T* __restrict__ train_x;
    const T* __restrict__ v_update;
};

template <typename T> inline PathDev<T> to_dev(const nmb_path_t& p) {
    PathDev<T> o;
    o.d = p.d; o.n_freq = p.n_freq; o.n_train = p.n_train; o.n_paths = p.n_paths;
    o.kernel_id = p.kernel_id;
    o.feat_scale = (T)p.feat_scale; o.amplitude = (T)p.amplitude; o.mean_const = (T)p.mean_const;
    o.w_phase = (const T*)p.w_phase; o.w_feat = (const T*)p.w_feat; o.inv_ls = (const T*)p.inv_ls;
    o.train_x = (const T*)p.train_x; o.v_update = (const T*)p.v_update;
    return o;
}

// x: the point (registers, DMAX padded), s: path index, (lane, nl): this thread's slot
// in its group.  Outputs are partial: y_part excludes mean_const.
// TAB (FP64 only): sincospi / exp through the shared-memory lookup tables at `tab` (fastmath.cuh).
template <typename T, int DMAX, bool GRAD = true, bool TAB = false>
__device__ __forceinline__ void path_partial(const PathDev<T>& P, const T (&x)[DMAX], int s, int lane,
                                             int nl, T& y_part, T (&g_part)[DMAX], const double* __restrict__ tab = nullptr) {
    const int d = P.d;
    T acc = T(0);
    T gacc[DMAX];
#pragma unroll
    for (int k = 0; k < DMAX; ++k) gacc[k] = T(0);

    // ---- random-Fourier-feature prior
    const T* wsin = P.w_feat + (size_t)s * 2 * P.n_freq;
    const T* wcos = wsin + P.n_freq;
#pragma unroll 4
    for (int m = lane; m < P.n_freq; m += nl) {
        const T* wr = P.w_phase + (size_t)m * d;
        T wk[DMAX];
        T u = T(0);
#pragma unroll
        for (int k = 0; k < DMAX; ++k) {
            wk[k] = (k < d) ? __ldg(wr + k) : T(0);
            u = fma(wk[k], x[k], u);
        }
        T sn, cs;
        if constexpr (TAB && sizeof(T) == 8) nmb_sincospi_tab(u, tab, &sn, &cs);
        else if constexpr (TAB) nmb_sincospi_fast(u, &sn, &cs);          // float: exact reduction + MUFU
        else Num<T>::sincospi(u, &sn, &cs);
        const T ws = __ldg(wsin + m), wc = __ldg(wcos + m);
        acc = fma(sn, ws, fma(cs, wc, acc));
        if (GRAD) {
            const T dphi = cs * ws - sn * wc;   // d/dz of (sin z ws + cos z wc)
#pragma unroll
            for (int k = 0; k < DMAX; ++k) gacc[k] = fma(dphi, wk[k], gacc[k]);
        }
    }
    acc *= P.feat_scale;
    const T gs = P.feat_scale * Num<T>::kPi;
#pragma unroll
    for (int k = 0; k < DMAX; ++k) gacc[k] *= gs;

    // ---- pathwise update  amp * k(x, X_tr) . v
    T ils[DMAX];
#pragma unroll
    for (int k = 0; k < DMAX; ++k) ils[k] = (k < d) ? __ldg(P.inv_ls + k) : T(0);
    const T* vv = P.v_update + (size_t)s * P.n_train;
    T kacc = T(0);
    T kg[DMAX];
#pragma unroll
    for (int k = 0; k < DMAX; ++k) kg[k] = T(0);
    for (int j = lane; j < P.n_train; j += nl) {
        const T* xr = P.train_x + (size_t)j * d;
        T sd[DMAX];
        T r2 = T(0);
#pragma unroll
        for (int k = 0; k < DMAX; ++k) {
            sd[k] = (k < d) ? (x[k] - __ldg(xr + k)) * ils[k] : T(0);
            r2 = fma(sd[k], sd[k], r2);
        }
        T kap, gc;
        if constexpr (TAB && sizeof(T) == 8) kernel_profile_tab(P.kernel_id, r2, tab, kap, gc);
        else kernel_profile<T>(P.kernel_id, r2, kap, gc);
        const T vj = __ldg(vv + j);
        kacc = fma(kap, vj, kacc);
        if (GRAD) {
            const T gv = gc * vj;
#pragma unroll
            for (int k = 0; k < DMAX; ++k) kg[k] = fma(gv, sd[k], kg[k]);
        }
    }
    y_part = fma(P.amplitude, kacc, acc);
#pragma unroll
    for (int k = 0; k < DMAX; ++k) g_part[k] = fma(P.amplitude * ils[k], kg[k], gacc[k]);
}

}  // namespace nmb
