// nmb_path_eval: standalone Thompson-path evaluation (value + input gradient).
// Replaces `self.f(X)` of the reference (acqfs.py:362, 409-411) when it is called
// outside the fused rollout: free-parameter HES-TS, posterior-surface plots
// (draw_trajectory_gif.py:123-125 evaluates 40,000 points) and tests.
#include "path.cuh"

namespace nmb {

// One group of TPP threads per point; block = 256 threads.
template <typename T, int DMAX>
__global__ void __launch_bounds__(256) path_eval_kernel(PathDev<T> P, const T* __restrict__ x, long long n_points,
                                                        long long points_per_path, int tpp, T* __restrict__ y,
                                                        T* __restrict__ dydx) {
    __shared__ T red[8][DMAX + 1];
    const int groups_per_block = 256 / tpp;
    const int g = threadIdx.x / tpp, lane = threadIdx.x % tpp;
    const long long stride = (long long)gridDim.x * groups_per_block;
    for (long long base = (long long)blockIdx.x * groups_per_block; base < n_points; base += stride) {
        const long long i = base + g;
        const bool live = i < n_points;
        T xv[DMAX];
        T yp = T(0), gp[DMAX];
#pragma unroll
        for (int k = 0; k < DMAX; ++k) { xv[k] = T(0); gp[k] = T(0); }
        if (live) {
#pragma unroll
            for (int k = 0; k < DMAX; ++k) if (k < P.d) xv[k] = x[i * P.d + k];
            const int s = (int)(i / points_per_path);
            path_partial<T, DMAX>(P, xv, s, lane, tpp, yp, gp);
        }
        const int w = tpp < 32 ? tpp : 32;
        yp = group_sum(yp, w);
#pragma unroll
        for (int k = 0; k < DMAX; ++k) gp[k] = group_sum(gp[k], w);
        if (tpp > 32) {   // whole warps per point: finish across warps in shared memory
            const int warp = threadIdx.x >> 5;
            if ((threadIdx.x & 31) == 0) {
                red[warp][0] = yp;
#pragma unroll
                for (int k = 0; k < DMAX; ++k) red[warp][k + 1] = gp[k];
            }
            __syncthreads();
            if (lane == 0) {
                const int w0 = threadIdx.x >> 5, nw = tpp >> 5;
                yp = T(0);
#pragma unroll
                for (int k = 0; k < DMAX; ++k) gp[k] = T(0);
                for (int q = 0; q < nw; ++q) {
                    yp += red[w0 + q][0];
#pragma unroll
                    for (int k = 0; k < DMAX; ++k) gp[k] += red[w0 + q][k + 1];
                }
            }
            __syncthreads();
        }
        if (live && lane == 0) {
            if (y) y[i] = yp + P.mean_const;
            if (dydx) {
#pragma unroll
                for (int k = 0; k < DMAX; ++k) if (k < P.d) dydx[i * P.d + k] = gp[k];
            }
        }
    }
}

template <typename T>
static int launch_path_eval(const nmb_path_t* p, const void* x, int64_t n, int64_t ppp, void* y, void* dydx,
                            cudaStream_t st) {
    PathDev<T> P = to_dev<T>(*p);
    const int sms = sm_count();
    // enough threads per point to fill the machine at small n, a warp per point at large n
    int tpp = 32;
    while (tpp < 256 && n * tpp < (int64_t)sms * 256 * 4) tpp <<= 1;
    const int gpb = 256 / tpp;
    int64_t blocks = (n + gpb - 1) / gpb;
    const int64_t cap = (int64_t)sms * 8;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
#define NMB_PE(DM) path_eval_kernel<T, DM><<<(unsigned)blocks, 256, 0, st>>>(P, (const T*)x, n, ppp, tpp, (T*)y, (T*)dydx)
    if (p->d <= 2) NMB_PE(2);
    else if (p->d <= 4) NMB_PE(4);
    else if (p->d <= 8) NMB_PE(8);
    else NMB_PE(16);
#undef NMB_PE
    NMB_LAUNCH_CHECK();
    return 0;
}

}  // namespace nmb

extern "C" int nmb_path_eval(const nmb_path_t* path, int dtype, const void* x, int64_t n_points,
                             int64_t points_per_path, void* y, void* dydx, void* stream) {
    using namespace nmb;
    NMB_REQUIRE(path && x, NMB_EINVAL, "nmb_path_eval: null path or x");
    NMB_REQUIRE(path->d >= 1 && path->d <= NMB_MAX_DIM, NMB_ELIMIT, "nmb_path_eval: d=%d outside [1,%d]", path->d,
                NMB_MAX_DIM);
    NMB_REQUIRE(path->kernel_id >= 0 && path->kernel_id <= 3, NMB_EINVAL, "nmb_path_eval: kernel_id %d",
                path->kernel_id);
    NMB_REQUIRE(n_points >= 0 && points_per_path >= 1, NMB_EINVAL, "nmb_path_eval: bad sizes");
    NMB_REQUIRE((n_points + points_per_path - 1) / points_per_path <= path->n_paths || n_points == 0, NMB_EINVAL,
                "nmb_path_eval: %lld points / %lld per path exceed %d paths", (long long)n_points,
                (long long)points_per_path, path->n_paths);
    NMB_REQUIRE(dtype == NMB_F64 || dtype == NMB_F32, NMB_EINVAL, "nmb_path_eval: dtype %d", dtype);
    if (n_points == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    return dtype == NMB_F64 ? launch_path_eval<double>(path, x, n_points, points_per_path, y, dydx, st)
                            : launch_path_eval<float>(path, x, n_points, points_per_path, y, dydx, st);
}
