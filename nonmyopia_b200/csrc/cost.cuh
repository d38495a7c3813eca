// Spotlight cost of one move (reference acqfs.py:604-633): shared device helpers.
#pragma once
#include "common.cuh"

namespace nmb {

template <typename T> struct CostDev {
    T radius, k, max_noise, p_norm, discount, discount_threshold, prev_cost, disc_mult;
};

template <typename T> inline CostDev<T> to_dev(const nmb_cost_t& c) {
    CostDev<T> o;
    o.radius = (T)c.radius; o.k = (T)c.k; o.max_noise = (T)c.max_noise; o.p_norm = (T)c.p_norm;
    o.discount = (T)c.discount; o.discount_threshold = (T)c.discount_threshold; o.prev_cost = (T)c.prev_cost;
    // the reference forms 1 - discount * mask in float32 (acqfs.py:626-631)
    o.disc_mult = (T)(1.0f - (float)c.discount);
    return o;
}

// distance of the move and its derivative direction (torch.cdist p-norm on 1x1 blocks)
template <typename T, int DMAX>
__device__ __forceinline__ T move_dist(const CostDev<T>& c, int d, const T (&cur)[DMAX], const T (&prev)[DMAX],
                                       T (&dir)[DMAX]) {
    T dist = T(0);
    if (c.p_norm == T(2)) {
#pragma unroll
        for (int k = 0; k < DMAX; ++k) if (k < d) { const T df = cur[k] - prev[k]; dist = fma(df, df, dist); }
        dist = Num<T>::sqrt(dist);
#pragma unroll
        for (int k = 0; k < DMAX; ++k) dir[k] = (k < d && dist > T(0)) ? (cur[k] - prev[k]) / dist : T(0);
    } else if (c.p_norm == T(1)) {
#pragma unroll
        for (int k = 0; k < DMAX; ++k) {
            const T df = k < d ? cur[k] - prev[k] : T(0);
            dist += Num<T>::fabs(df);
            dir[k] = df > T(0) ? T(1) : (df < T(0) ? T(-1) : T(0));
        }
    } else {
        const double p = (double)c.p_norm;
        double acc = 0.0;
#pragma unroll
        for (int k = 0; k < DMAX; ++k) if (k < d) acc += ::pow(::fabs((double)(cur[k] - prev[k])), p);
        const double dd = ::pow(acc, 1.0 / p);
        dist = (T)dd;
#pragma unroll
        for (int k = 0; k < DMAX; ++k) {
            const double df = k < d ? (double)(cur[k] - prev[k]) : 0.0;
            dir[k] = (k < d && dd > 0.0) ? (T)((df > 0 ? 1.0 : -1.0) * ::pow(::fabs(df) / dd, p - 1.0)) : T(0);
        }
    }
    return dist;
}

// hinge + noise, before the discount multiplier
template <typename T> __device__ __forceinline__ T cost_hinge(const CostDev<T>& c, T dist, T noise) {
    const T h = c.k * (dist - c.radius);
    return (h > T(0) ? h : T(0)) + noise * c.max_noise;
}

}  // namespace nmb
