// Branch-free FP64 transcendentals for the rollout kernels.
//
// ncu (profiles/ncu_lines_r01b_fwd.txt) showed the forward rollout spending 56 % of its warp
// instructions in CUDA libm's sincospi (74 instr/call), expm1 (~130), exp and tanh - all with slow
// paths and special-case handling the policy network / Thompson path never need.  These versions are
// straight-line DFMA sequences (magic-number rounding, Cody-Waite reduction, Taylor/Horner kernels):
//   nmb_exp      ~20 FP64 ops   |rel err| < 3e-16 on [-700, 700]
//   nmb_expm1    ~22 FP64 ops   |err| < 3e-16 * max(|result|, ...) incl. near 0 (k = 0 keeps the polynomial)
//   nmb_sincospi ~30 FP64 ops   |abs err| < 3e-16 for |x| < 2^30
//   nmb_rcp      rcp.approx + 2 Newton steps (positive, normal arguments only)
// Accuracy is checked on the host against libm by tests/test_fastmath.py (nvcc builds this header
// for the CPU as well: everything is __host__ __device__).
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <string.h>

namespace nmb {

#define NMB_HD __host__ __device__ __forceinline__

NMB_HD int64_t f64_bits(double x) {
#ifdef __CUDA_ARCH__
    return __double_as_longlong(x);
#else
    int64_t b; memcpy(&b, &x, 8); return b;
#endif
}
NMB_HD double bits_f64(int64_t b) {
#ifdef __CUDA_ARCH__
    return __longlong_as_double(b);
#else
    double x; memcpy(&x, &b, 8); return x;
#endif
}

// 1 / a for a in [2^-500, 2^500], a > 0
NMB_HD double nmb_rcp(double a) {
#ifdef __CUDA_ARCH__
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(a));
    double e = fma(-a, r, 1.0);
    r = fma(r, e, r);
    e = fma(-a, r, 1.0);
    r = fma(r, e, r);
    e = fma(-a, r, 1.0);
    return fma(r, e, r);
#else
    return 1.0 / a;
#endif
}

// shared reduction: x = k ln2 + r, |r| <= ln2/2; returns p = expm1(r) and k
NMB_HD double exp_reduce(double x, int& k) {
    const double magic = 6755399441055744.0;     // 1.5 * 2^52
    const double kd = fma(x, 1.44269504088896338700e+00, magic);
    k = (int)(uint32_t)f64_bits(kd);
    const double kf = kd - magic;
    double r = fma(kf, -6.93147180369123816490e-01, x);
    r = fma(kf, -1.90821492927058770002e-10, r);
    // expm1(r) = r + r^2 q(r), q = sum_j r^j / (j+2)!  (j <= 11) as even(z) + r odd(z), z = r^2: two independent
    // Horner chains of depth 5 instead of one of depth 11
    const double z = r * r;
    double qe = 2.08767569878681e-09;              // 1/12!
    double qo = 1.6059043836821613e-10;            // 1/13!
    qe = fma(qe, z, 2.755731922398589e-07);        // 1/10!
    qo = fma(qo, z, 2.505210838544172e-08);        // 1/11!
    qe = fma(qe, z, 2.48015873015873e-05);         // 1/8!
    qo = fma(qo, z, 2.7557319223985893e-06);       // 1/9!
    qe = fma(qe, z, 1.388888888888889e-03);        // 1/6!
    qo = fma(qo, z, 1.984126984126984e-04);        // 1/7!
    qe = fma(qe, z, 4.1666666666666664e-02);       // 1/4!
    qo = fma(qo, z, 8.333333333333333e-03);        // 1/5!
    qe = fma(qe, z, 0.5);                          // 1/2!
    qo = fma(qo, z, 1.6666666666666666e-01);       // 1/3!
    const double q = fma(qo, r, qe);
    return fma(q, z, r);
}

NMB_HD double pow2i(int k) { return bits_f64((int64_t)(k + 1023) << 52); }

NMB_HD double nmb_exp(double x) {
    x = fmin(fmax(x, -700.0), 700.0);
    int k;
    const double p = exp_reduce(x, k);
    const double s = pow2i(k);
    return fma(s, p, s);
}

NMB_HD double nmb_expm1(double x) {
    x = fmin(fmax(x, -60.0), 700.0);
    int k;
    const double p = exp_reduce(x, k);
    const double s = pow2i(k);
    return fma(s, p, s - 1.0);
}

// sin(pi x), cos(pi x): x = q/2 + r, |r| <= 1/4 (exact); Taylor kernels in z = r^2; quadrant swap
NMB_HD void nmb_sincospi(double x, double* sn, double* cs) {
    const double magic = 6755399441055744.0;
    const double kd = fma(x, 2.0, magic);
    const int q = (int)(uint32_t)f64_bits(kd);
    const double r = fma(kd - magic, -0.5, x);
    const double z = r * r;
    // sin(pi r) = r * sum (-1)^n pi^(2n+1)/(2n+1)! z^n
    double s = 7.952054001475513e-07;
    s = fma(s, z, -2.1915353447830217e-05);
    s = fma(s, z, 4.6630280576761255e-04);
    s = fma(s, z, -7.3704309457143504e-03);
    s = fma(s, z, 8.214588661112823e-02);
    s = fma(s, z, -5.992645293207921e-01);
    s = fma(s, z, 2.5501640398773455);
    s = fma(s, z, -5.16771278004997);
    s = fma(s, z, 3.141592653589793);
    s *= r;
    // cos(pi r) = sum (-1)^n pi^(2n)/(2n)! z^n
    double c = -1.3878952462213771e-07;
    c = fma(c, z, 4.303069587032947e-06);
    c = fma(c, z, -1.046381049248457e-04);
    c = fma(c, z, 1.9295743094039231e-03);
    c = fma(c, z, -2.580689139001406e-02);
    c = fma(c, z, 2.353306303588932e-01);
    c = fma(c, z, -1.3352627688545895);
    c = fma(c, z, 4.0587121264167685);
    c = fma(c, z, -4.934802200544679);
    c = fma(c, z, 1.0);
    // quadrant q mod 4: 0 (s, c)  1 (c, -s)  2 (-s, -c)  3 (-c, s)
    const double a = (q & 1) ? c : s, b = (q & 1) ? s : c;
    *sn = (q & 2) ? -a : a;
    *cs = ((q + 1) & 2) ? -b : b;
}

// logistic function and tanh on top of the exp kernels (arguments clamped where the result saturates)
NMB_HD double nmb_sigmoid(double x) {
    x = fmin(fmax(x, -40.0), 40.0);
    return nmb_rcp(1.0 + nmb_exp(-x));
}
NMB_HD double nmb_tanh(double x) {
    x = fmin(fmax(x, -20.0), 20.0);
    const double em = nmb_expm1(2.0 * x);
    return em * nmb_rcp(em + 2.0);
}

}  // namespace nmb
