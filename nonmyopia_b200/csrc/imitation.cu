// Amortized pre-training of Actor.reset_parameters (reference actor.py:164-220) on the rollout kernels.
//
// One epoch = for every level l: one policy step of the fused rollout kernel in imitation mode (rollout.cu), then a
// draw of the surrogate's MARGINAL posterior at the step's output as the next observation
//     y = mu(x) + sigma(x) z,   z ~ N(0, 1)          (model(x.unsqueeze(-2)).sample(), actor.py:186-191)
// from the shared-root whitening kernel (nmb_gp_whiten_fwd: m0 = k.alpha, q = |L^-1 k|^2) and the sample kernel below;
// then the L1 imitation loss with its adjoint seeds and the early-stop bookkeeping (objective kernel below), the
// reverse rollout, the split-K weight gradients and the fused AdamW step of the query loop, unchanged.
#include "common.cuh"

namespace nmb {
namespace {

__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
    const unsigned M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int i = 0; i < 10; ++i) {
        const unsigned hi0 = __umulhi(M0, c.x), lo0 = M0 * c.x;
        const unsigned hi1 = __umulhi(M1, c.z), lo1 = M1 * c.z;
        c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
        k.x += W0; k.y += W1;
    }
    return c;
}

// sqrt of a 1x1 covariance with GPyTorch's psd_safe_cholesky jitter rule, as in mc_tree.cu
template <typename T> __device__ __forceinline__ T safe_sigma1(T var) {
    if (var > T(0)) return Num<T>::sqrt(var);
    T j = Num<T>::kChJitter;
    for (int i = 0; i < 3; ++i, j *= T(10))
        if (var + j > T(0)) return Num<T>::sqrt(var + j);
    return T(0);
}

template <typename T>
__global__ void marginal_sample_kernel(T amp, T mean_const, const T* __restrict__ m0, const T* __restrict__ q,
                                       long long n, const T* __restrict__ z, const T* __restrict__ z_ring, int ring_len,
                                       int levels, int level, unsigned long long seed, long long chain_offset,
                                       const long long* __restrict__ step_ptr, const int* __restrict__ done_flag,
                                       T* __restrict__ y) {
    if (done_flag && *done_flag) return;
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const long long step = step_ptr ? *step_ptr : 0;
    T zz;
    if (z_ring) {
        zz = z_ring[((size_t)(step % ring_len) * levels + level) * n + i];
    } else if (z) {
        zz = z[i];
    } else {
        const long long gc = i + chain_offset;
        const uint4 r = philox4x32_10(make_uint4((unsigned)gc, (unsigned)(gc >> 32), (unsigned)step, 0x40000000u | (unsigned)level),
                                      make_uint2((unsigned)seed, (unsigned)(seed >> 32)));
        const double u1 = ((double)r.x * 4294967296.0 + (double)r.y + 1.0) * (1.0 / 18446744073709551616.0);
        const double u2 = ((double)r.z * 4294967296.0 + (double)r.w) * (1.0 / 18446744073709551616.0);
        double sn, cs;
        ::sincospi(2.0 * u2, &sn, &cs);
        zz = (T)(::sqrt(-2.0 * ::log(u1)) * cs);
    }
    y[i] = mean_const + m0[i] + safe_sigma1<T>(amp - q[i]) * zz;
}

// L1 imitation loss  mean |target - out|  over the (L+1) N d outputs of the rollout, its adjoint seeds
// sign(out - target) / numel (torch: d|x|/dx = sign(x), 0 at 0) and the early-stop bookkeeping of actor.py:200-219:
// a new best loss resets the counter, otherwise it grows; the step still runs; the loop stops after it once the
// counter has reached `patience`.  One CTA; fixed reduction order.
template <typename T>
__global__ void __launch_bounds__(256) imitation_objective_kernel(int d, int L, long long N, const T* __restrict__ X,
                                                                  const T* __restrict__ A, const T* __restrict__ target,
                                                                  T* __restrict__ g_X, T* __restrict__ g_A,
                                                                  T* __restrict__ g_loss, T* __restrict__ g_cost,
                                                                  T* __restrict__ hist, long long max_epochs, int patience,
                                                                  nmb_loop_state_t* state) {
    if (state->done) return;
    const long long per = N * d, numel = (long long)(L + 1) * per;
    const T inv = T(1) / (T)numel;
    T acc = T(0);
    for (long long e = threadIdx.x; e < numel; e += blockDim.x) {
        const int l = (int)(e / per);
        const long long r = e - (long long)l * per;
        const T out = l < L ? X[e] : A[r];
        const T df = out - target[e];
        acc += Num<T>::fabs(df);
        const T sg = df > T(0) ? inv : (df < T(0) ? -inv : T(0));
        if (l < L) g_X[e] = sg; else g_A[r] = sg;
    }
    for (long long c = threadIdx.x; c < N; c += blockDim.x) { g_loss[c] = T(0); g_cost[c] = T(0); }
    __shared__ T red[256];
    red[threadIdx.x] = acc;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
        if ((int)threadIdx.x < s) red[threadIdx.x] += red[threadIdx.x + s];
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        const double loss = (double)red[0] * (double)inv;
        if (hist && state->step < max_epochs) hist[state->step] = (T)loss;
        state->loss_mean = loss;
        const bool better = state->step == 0 || loss < state->best_min;
        state->improved = better ? 1 : 0;
        if (better) { state->best_min = loss; state->stall = 0; }
        else state->stall += 1;
        state->done_next = (patience >= 0 && state->stall >= patience) ? 1 : 0;
    }
}

}  // namespace
}  // namespace nmb

using namespace nmb;

extern "C" int nmb_gp_marginal_sample(int dtype, double amplitude, double mean_const, const void* m0, const void* q,
                                      int64_t n, const void* z, const void* z_ring, int ring_len, int n_levels, int level,
                                      uint64_t seed, int64_t chain_offset, const int64_t* step_ptr,
                                      const int32_t* done_flag, void* y, void* stream) {
    NMB_REQUIRE(m0 && q && y && n >= 1, NMB_EINVAL, "nmb_gp_marginal_sample: null argument");
    NMB_REQUIRE(!z_ring || (ring_len > 0 && n_levels > 0 && level >= 0 && level < n_levels), NMB_EINVAL,
                "nmb_gp_marginal_sample: ring of %d x %d levels, level %d", ring_len, n_levels, level);
    cudaStream_t st = (cudaStream_t)stream;
    const unsigned grid = (unsigned)((n + 127) / 128);
    if (dtype == NMB_F64)
        marginal_sample_kernel<double><<<grid, 128, 0, st>>>(amplitude, mean_const, (const double*)m0, (const double*)q, n,
                                                              (const double*)z, (const double*)z_ring, ring_len, n_levels,
                                                              level, seed, chain_offset, (const long long*)step_ptr,
                                                              done_flag, (double*)y);
    else if (dtype == NMB_F32)
        marginal_sample_kernel<float><<<grid, 128, 0, st>>>((float)amplitude, (float)mean_const, (const float*)m0,
                                                             (const float*)q, n, (const float*)z, (const float*)z_ring,
                                                             ring_len, n_levels, level, seed, chain_offset,
                                                             (const long long*)step_ptr, done_flag, (float*)y);
    else { set_error("nmb_gp_marginal_sample: dtype %d", dtype); return NMB_EINVAL; }
    NMB_LAUNCH_CHECK();
    return 0;
}

extern "C" int nmb_imitation_objective(int dtype, int d, int L, int64_t n_chains, const void* X, const void* A,
                                       const void* target, void* g_X, void* g_A, void* g_loss, void* g_cost, void* hist,
                                       int64_t max_epochs, int patience, nmb_loop_state_t* state, void* stream) {
    NMB_REQUIRE(A && target && g_A && g_loss && g_cost && state && (L == 0 || (X && g_X)), NMB_EINVAL,
                "nmb_imitation_objective: null argument");
    NMB_REQUIRE(d >= 1 && L >= 0 && n_chains >= 1, NMB_EINVAL, "nmb_imitation_objective: d=%d L=%d N=%lld", d, L,
                (long long)n_chains);
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == NMB_F64)
        imitation_objective_kernel<double><<<1, 256, 0, st>>>(d, L, n_chains, (const double*)X, (const double*)A,
                                                               (const double*)target, (double*)g_X, (double*)g_A,
                                                               (double*)g_loss, (double*)g_cost, (double*)hist, max_epochs,
                                                               patience, state);
    else if (dtype == NMB_F32)
        imitation_objective_kernel<float><<<1, 256, 0, st>>>(d, L, n_chains, (const float*)X, (const float*)A,
                                                              (const float*)target, (float*)g_X, (float*)g_A, (float*)g_loss,
                                                              (float*)g_cost, (float*)hist, max_epochs, patience, state);
    else { set_error("nmb_imitation_objective: dtype %d", dtype); return NMB_EINVAL; }
    NMB_LAUNCH_CHECK();
    return 0;
}
