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
#include "fast_tables.h"

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


// ---------------------------------------------------------------- table-driven variants
// ncu (profiles/ncu_lines_r01b_fwd.txt, round 1 final): the Taylor kernels above still made the forward rollout
// issue-bound - 57 warp instructions per sincospi and 67 per activation, 47 % of all instructions of the launch.
// With a small table in shared memory (fast_tables.h: 2^(j/64), j < 64 and (sin, cos)(pi j/128), j < 256) the
// reduced argument is 64 x / 128 x smaller and the polynomials shrink from degree 12 / 9+9 to 5 / 2+3:
//   nmb_exp_tab / nmb_expm1_tab   12 FP64 ops + 1 LDS.64     |rel err| < 3e-16
//   nmb_sincospi_tab              14 FP64 ops + 1 LDS.128    |abs err| < 3e-16
// `tab` points at the NMB_FAST_TAB_LEN doubles of fast_tables.h (16-byte aligned).
NMB_HD double exp_reduce_tab(double x, const double* __restrict__ tab, double& s) {
    const double magic = 6755399441055744.0;     // 1.5 * 2^52
    const double kd = fma(x, 92.33248261689366, magic);          // 64 / ln 2
    const int ki = (int)(uint32_t)f64_bits(kd);
    const double kf = kd - magic;
    double r = fma(kf, -0.010830424696249145, x);                // ln 2 / 64, high part
    r = fma(kf, -3.623510646634843e-19, r);                      // low part
    // expm1(r), |r| <= ln2/128: r + r^2 (1/2 + r/6 + r^2/24 + r^3/120), next term r^6/720 < 4e-17
    const double z = r * r;
    double q = fma(8.333333333333333e-03, r, 4.1666666666666664e-02);
    q = fma(q, r, 1.6666666666666666e-01);
    q = fma(q, r, 0.5);
    const double p = fma(q, z, r);
    // s = 2^(ki/64) = tab[ki & 63] * 2^(ki >> 6): the exponent is added to the table entry's bits
#ifdef __CUDA_ARCH__
    // the exponent goes into the HIGH word only (the 64-bit form costs an add-with-carry pair and two wide shifts)
    const double t = tab[ki & 63];
    s = __hiloint2double(__double2hiint(t) + ((ki >> 6) << 20), __double2loint(t));
#else
    s = bits_f64(f64_bits(tab[ki & 63]) + ((int64_t)(ki >> 6) << 52));
#endif
    return p;
}

NMB_HD double nmb_exp_tab(double x, const double* __restrict__ tab) {
    x = fmin(fmax(x, -700.0), 700.0);
    double s;
    const double p = exp_reduce_tab(x, tab, s);
    return fma(s, p, s);
}

NMB_HD double nmb_expm1_tab(double x, const double* __restrict__ tab) {
    x = fmin(fmax(x, -60.0), 700.0);
    double s;
    const double p = exp_reduce_tab(x, tab, s);
    return fma(s, p, s - 1.0);
}
// the same for arguments the caller has already confined to [-700, 700] (a 64-bit min/max costs three instructions:
// the activations below clamp once, not twice)
NMB_HD double nmb_exp_tab_nc(double x, const double* __restrict__ tab) {
    double s;
    const double p = exp_reduce_tab(x, tab, s);
    return fma(s, p, s);
}
NMB_HD double nmb_expm1_tab_nc(double x, const double* __restrict__ tab) {
    double s;
    const double p = exp_reduce_tab(x, tab, s);
    return fma(s, p, s - 1.0);
}

// sin(pi x), cos(pi x): x = j/128 + r, |r| <= 1/256 (exact); angle addition with the table entry of j mod 256
NMB_HD void nmb_sincospi_tab(double x, const double* __restrict__ tab, double* sn, double* cs) {
    const double magic = 6755399441055744.0;
    const double kd = fma(x, 128.0, magic);
    const int j = (int)(uint32_t)f64_bits(kd) & 255;
    const double r = fma(kd - magic, -0.0078125, x);
    const double z = r * r;
#ifdef __CUDA_ARCH__
    const double2 t = *reinterpret_cast<const double2*>(tab + 64 + 2 * j);
    const double ts = t.x, tc = t.y;
#else
    const double ts = tab[64 + 2 * j], tc = tab[64 + 2 * j + 1];
#endif
    // s = sin(pi r) = r (pi - pi^3/6 z + pi^5/120 z^2)      (next term: 7e-16 relative to s, |s| < 0.0123)
    double s = fma(2.5501640398773455, z, -5.16771278004997);
    s = fma(s, z, 3.141592653589793);
    s *= r;
    // dc = cos(pi r) - 1 = z (-pi^2/2 + pi^4/24 z - pi^6/720 z^2)   (next term < 2e-20)
    double dc = fma(-1.3352627688545895, z, 4.0587121264167685);
    dc = fma(dc, z, -4.934802200544679);
    dc *= z;
    *sn = fma(ts, dc, fma(tc, s, ts));
    *cs = fma(tc, dc, fma(-ts, s, tc));
}


NMB_HD double nmb_rcp2(double a) {      // 1 / a, a in [2^-500, 2^500]: rcp.approx (2^-23) + 2 Newton steps
#ifdef __CUDA_ARCH__
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(a));
    double e = fma(-a, r, 1.0);
    r = fma(r, e, r);
    e = fma(-a, r, 1.0);
    return fma(r, e, r);
#else
    return 1.0 / a;
#endif
}

// |x| <= lim (+ at most 2^-20 relative: the low word is kept) for a limit whose low word is zero (40, 20, ...),
// decided on the high word alone: three integer instructions (and, unsigned min, or) instead of the fourteen of
// fmin(fmax()) with its NaN handling (a NaN comes out as +-lim).  The callers only need "no overflow past lim".
NMB_HD double nmb_clamp_abs(double x, double lim) {
#ifdef __CUDA_ARCH__
    const int hi = __double2hiint(x);
    const unsigned a = min((unsigned)(hi & 0x7fffffff), (unsigned)__double2hiint(lim));
    return __hiloint2double((int)((unsigned)(hi & 0x80000000) | a), __double2loint(x));
#else
    return fmin(fmax(x, -lim), lim);
#endif
}
// The ELU's exponential argument: negative x with the magnitude clamped at 60 (-60.00..: low word kept).  Negative
// doubles order like their high words read as unsigned, so ONE unsigned min does it; x >= 0 passes through
// unchanged - the ELU selects x itself there and discards the exponential (whatever it evaluates to).
NMB_HD double nmb_neg_part_clamped(double x) {
#ifdef __CUDA_ARCH__
    const unsigned hi = (unsigned)__double2hiint(x);
    return __hiloint2double((int)min(hi, 0xC04E0000u), __double2loint(x));
#else
    return fmax(fmin(x, 0.0), -60.0);
#endif
}

// rcp.approx.f64 works on the upper 32 bits (20 mantissa bits): two Newton steps reach 2^-80, i.e. full double
// precision (measured on the device: identical worst-case error to three steps, tests/test_fastmath.py)
NMB_HD double nmb_sigmoid_tab(double x, const double* __restrict__ tab) {
    x = nmb_clamp_abs(x, 40.0);
    return nmb_rcp2(1.0 + nmb_exp_tab_nc(-x, tab));
}
NMB_HD double nmb_tanh_tab(double x, const double* __restrict__ tab) {
    x = nmb_clamp_abs(x, 20.0);
    const double em = nmb_expm1_tab_nc(2.0 * x, tab);
    return em * nmb_rcp2(em + 2.0);
}
NMB_HD double nmb_elu_tab(double x, const double* __restrict__ tab) {
    const double e = nmb_expm1_tab_nc(nmb_neg_part_clamped(x), tab);
    return x > 0.0 ? x : e;
}

}  // namespace nmb
