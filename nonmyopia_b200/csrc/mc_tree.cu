// Exact-GP posterior over a fantasy tree: shared-root whitening (GEMM against the explicit
// inverse Cholesky root, kernel rows generated on the fly), per-level tail recursion, the
// hand-derived reverse pass, and the per-restart reductions (reference acqfs.py:363-373,
// 413-415, 443-447).
//
//   whiten_fwd : CTA = 32 query points.  K tile [n][32] (k-major) is generated in shared memory
//                from the points and X_tr (one coalesced pass, fused m0 = k.alpha), then
//                U[32 x 64 cols] = K^T L^-T column block by column block with the L^-T operand
//                streamed from L2 through the double-buffered stage of policy_dev.cuh
//                (only the j <= i part is touched: half the flops of a dense product).
//   whiten_bwd : same tiling, g_K = gU L^-1 fused with the kernel-gradient contraction
//                g_x = sum_j (g_K[j] + g_m0 alpha_j) d k(x, X_j) / dx  (K rows are recomputed).
//   level_fwd  : one warp per node: the s dots u(z_t).u(x) against its ancestors (contiguous
//                n-vectors), then the O(s^2) scalar recursion.
//   level_bwd  : one thread per node: reverse recursion; the node's contributions to its ancestors'
//                adjoint slots go to a scratch record per (target level, field, node) and
//                level_bwd_reduce sums every ancestor's records over its descendants in a fixed
//                order (deterministic: no atomics; levels are issued leaf -> root, so a node's
//                slots are complete when its own level runs).
//   gather_gu  : one warp per node: gU = 2 g_q u + sum_t g_dot_t u(z_t) + sum_desc g_dot u(desc),
//                descendants visited in a fixed order.
#include <algorithm>
#include "policy_mma.cuh"

namespace nmb {

template <typename T> struct GpDev {
    int d, n, kernel_id, ldu, ldl;
    T amp, noise, mean_const;
    const T *train_x, *inv_ls, *linv, *linv_t, *alpha, *linv_tiles, *linv_t_tiles;
};

template <typename T> static GpDev<T> to_dev(const nmb_gp_t& g) {
    GpDev<T> o;
    o.d = g.d; o.n = g.n; o.kernel_id = g.kernel_id; o.ldu = g.ldu; o.ldl = g.ldl;
    o.amp = (T)g.amplitude; o.noise = (T)g.noise; o.mean_const = (T)g.mean_const;
    o.train_x = (const T*)g.train_x; o.inv_ls = (const T*)g.inv_ls; o.linv = (const T*)g.linv;
    o.linv_t = (const T*)g.linv_t; o.alpha = (const T*)g.alpha;
    o.linv_tiles = (const T*)g.linv_tiles; o.linv_t_tiles = (const T*)g.linv_t_tiles;
    return o;
}

}  // namespace nmb
#include "whiten_mma.cuh"
namespace nmb {

extern __shared__ __align__(16) unsigned char mc_smem_raw[];

// ------------------------------------------------------------------ whiten forward
template <typename T, int DMAX>
__global__ void __launch_bounds__(kNT, 1) whiten_fwd_kernel(GpDev<T> g, const T* __restrict__ pts, long long n_pts,
                                                            T* __restrict__ U, T* __restrict__ m0,
                                                            T* __restrict__ q) {
    const int n = g.n, d = g.d, tid = threadIdx.x;
    const int npad = (n + 3) & ~3;
    T* Kt = reinterpret_cast<T*>(mc_smem_raw);        // [npad][kRS]
    T* stage = Kt + (size_t)npad * kRS;               // [2][kKC][kOB]
    T* xs = stage + 2 * kKC * kOB;                    // [kR][DMAX] scaled points
    T* red = xs + kR * DMAX;                          // [8][kR]
    const long long p0 = (long long)blockIdx.x * kR;
    const int nvalid = (int)min((long long)kR, n_pts - p0);
    for (int e = tid; e < kR * DMAX; e += kNT) {
        const int r = e / DMAX, k = e - r * DMAX;
        xs[e] = (r < nvalid && k < d) ? pts[(p0 + r) * d + k] * __ldg(g.inv_ls + k) : T(0);
    }
    __syncthreads();
    // ---- kernel rows: warp w handles training points j = w, w+8, ...; lane = query point
    {
        const int r = tid & 31, w = tid >> 5;
        T xr[DMAX];
#pragma unroll
        for (int k = 0; k < DMAX; ++k) xr[k] = xs[r * DMAX + k];
        T macc = T(0);
        for (int j = w; j < npad; j += 8) {
            T kv = T(0);
            if (j < n) {
                T r2 = T(0);
#pragma unroll
                for (int k = 0; k < DMAX; ++k) {
                    if (k < d) {
                        const T df = xr[k] - __ldg(g.train_x + (size_t)j * d + k) * __ldg(g.inv_ls + k);
                        r2 = fma(df, df, r2);
                    }
                }
                T kap, gc;
                kernel_profile<T>(g.kernel_id, r2, kap, gc);
                kv = g.amp * kap;
                macc = fma(kv, __ldg(g.alpha + j), macc);
            }
            Kt[(size_t)j * kRS + r] = (r < nvalid) ? kv : T(0);
        }
        red[w * kR + r] = macc;
    }
    __syncthreads();
    if (tid < nvalid) {
        T s = T(0);
#pragma unroll
        for (int w = 0; w < 8; ++w) s += red[w * kR + tid];
        m0[p0 + tid] = s;
    }
    // ---- U tile = K^T L^-T, 64 output columns at a time; only k < cb + 64 contributes
    const TileCoord tc;
    T qacc[2] = {T(0), T(0)};
    for (int cb = 0; cb < n; cb += kOB) {
        T acc[2][4];
        zero_acc(acc);
        const int nc = min(kOB, n - cb);
        gemm_block<T>(Kt, min(n, cb + kOB), g.linv_t, g.ldl, cb, nc, stage, tc, acc);
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const int r = tc.r0 + i;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int c = cb + tc.c0 + j;
                if (c < n) {
                    qacc[i] = fma(acc[i][j], acc[i][j], qacc[i]);
                    if (r < nvalid) U[(size_t)(p0 + r) * g.ldu + c] = acc[i][j];
                }
            }
        }
    }
    // |u|^2: 4 lanes (lane & 3) x 4 warps (warp & 3) share a row
#pragma unroll
    for (int i = 0; i < 2; ++i) { qacc[i] += shfl_xor_(qacc[i], 1); qacc[i] += shfl_xor_(qacc[i], 2); }
    __syncthreads();
    if ((tid & 3) == 0) {
        const int cw = (tid >> 5) & 3;
        red[cw * kR + tc.r0] = qacc[0];
        red[cw * kR + tc.r0 + 1] = qacc[1];
    }
    __syncthreads();
    if (tid < nvalid) q[p0 + tid] = red[tid] + red[kR + tid] + red[2 * kR + tid] + red[3 * kR + tid];
}

// ------------------------------------------------------------------ whiten backward
template <typename T, int DMAX>
__global__ void __launch_bounds__(kNT, 1) whiten_bwd_kernel(GpDev<T> g, const T* __restrict__ pts, long long n_pts,
                                                            const T* __restrict__ gU, const T* __restrict__ g_m0,
                                                            T* __restrict__ g_pts) {
    const int n = g.n, d = g.d, tid = threadIdx.x;
    const int npad = (n + 3) & ~3;
    T* Gt = reinterpret_cast<T*>(mc_smem_raw);        // [npad][kRS]  gU tile, k-major
    T* stage = Gt + (size_t)npad * kRS;
    T* xs = stage + 2 * kKC * kOB;                    // [kR][DMAX]
    T* red = xs + kR * DMAX;                          // [4][kR][DMAX]
    const long long p0 = (long long)blockIdx.x * kR;
    const int nvalid = (int)min((long long)kR, n_pts - p0);
    for (int e = tid; e < kR * DMAX; e += kNT) {
        const int r = e / DMAX, k = e - r * DMAX;
        xs[e] = (r < nvalid && k < d) ? pts[(p0 + r) * d + k] * __ldg(g.inv_ls + k) : T(0);
    }
    // gU tile: coalesced global reads along i, transposed into the k-major tile
    for (int e = tid; e < kR * npad; e += kNT) {
        const int r = e / npad, i = e - r * npad;
        Gt[(size_t)i * kRS + r] = (r < nvalid && i < n) ? gU[(size_t)(p0 + r) * g.ldu + i] : T(0);
    }
    __syncthreads();
    const TileCoord tc;
    T gx[2][DMAX];
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int k = 0; k < DMAX; ++k) gx[i][k] = T(0);
    T gm[2], xr[2][DMAX];
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        gm[i] = (tc.r0 + i < nvalid) ? g_m0[p0 + tc.r0 + i] : T(0);
#pragma unroll
        for (int k = 0; k < DMAX; ++k) xr[i][k] = xs[(tc.r0 + i) * DMAX + k];
    }
    for (int cb = 0; cb < n; cb += kOB) {
        T acc[2][4];
        zero_acc(acc);
        const int nc = min(kOB, n - cb);
        // g_K[r][j] = sum_{i >= j} gU[r][i] Linv[i][j]: rows i >= cb of Linv, columns cb..cb+63
        gemm_block<T>(Gt + (size_t)cb * kRS, n - cb, g.linv + (size_t)cb * g.ldl, g.ldl, cb, nc, stage, tc, acc);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int c = cb + tc.c0 + j;
            if (c < n) {
                const T al = __ldg(g.alpha + c);
                T tj[DMAX];
#pragma unroll
                for (int k = 0; k < DMAX; ++k) tj[k] = k < d ? __ldg(g.train_x + (size_t)c * d + k) * __ldg(g.inv_ls + k) : T(0);
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    T sd[DMAX], r2 = T(0);
#pragma unroll
                    for (int k = 0; k < DMAX; ++k) { sd[k] = xr[i][k] - tj[k]; r2 = fma(sd[k], sd[k], r2); }
                    T kap, gc;
                    kernel_profile<T>(g.kernel_id, r2, kap, gc);
                    const T coef = (acc[i][j] + gm[i] * al) * g.amp * gc;
#pragma unroll
                    for (int k = 0; k < DMAX; ++k) gx[i][k] = fma(coef, sd[k], gx[i][k]);
                }
            }
        }
    }
    // reduce over the 4 lanes (lane & 3) and the 4 column warps (warp & 3) that share a row
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int k = 0; k < DMAX; ++k) { gx[i][k] += shfl_xor_(gx[i][k], 1); gx[i][k] += shfl_xor_(gx[i][k], 2); }
    __syncthreads();
    if ((tid & 3) == 0) {
        const int cw = (tid >> 5) & 3;
#pragma unroll
        for (int i = 0; i < 2; ++i)
#pragma unroll
            for (int k = 0; k < DMAX; ++k) red[(cw * kR + tc.r0 + i) * DMAX + k] = gx[i][k];
    }
    __syncthreads();
    for (int e = tid; e < kR * DMAX; e += kNT) {
        const int r = e / DMAX, k = e - r * DMAX;
        if (r < nvalid && k < d) {
            const T s = red[e] + red[kR * DMAX + e] + red[2 * kR * DMAX + e] + red[3 * kR * DMAX + e];
            g_pts[(p0 + r) * d + k] += s * __ldg(g.inv_ls + k);
        }
    }
}

template <typename T, int DMAX> static size_t whiten_smem(int n, bool bwd) {
    const size_t npad = (n + 3) & ~3;
    return (npad * kRS + 2 * kKC * kOB + kR * DMAX + (bwd ? 4 * kR * DMAX : 8 * kR)) * sizeof(T);
}


// ------------------------------------------------------------------ tree bookkeeping
struct TreeDev {
    int n_levels, obs_noise;
    long long cnt[NMB_MAX_LEVELS], off[NMB_MAX_LEVELS + 1];
    const void* base[NMB_MAX_LEVELS];
    double action_base_mean;
};

static int make_tree(const nmb_tree_t* t, TreeDev& o) {
    NMB_REQUIRE(t && t->n_levels >= 1 && t->n_levels <= NMB_MAX_LEVELS, NMB_EINVAL, "mc tree: n_levels");
    o.n_levels = t->n_levels; o.obs_noise = t->obs_noise_fantasy; o.action_base_mean = t->action_base_mean;
    o.off[0] = 0;
    for (int s = 0; s < t->n_levels; ++s) {
        NMB_REQUIRE(t->counts[s] >= 1 && (s == 0 || t->counts[s] % t->counts[s - 1] == 0), NMB_EINVAL,
                    "mc tree: level %d has %lld nodes", s, (long long)t->counts[s]);
        NMB_REQUIRE(s == t->n_levels - 1 || t->base[s], NMB_EINVAL, "mc tree: level %d has no base samples", s);
        o.cnt[s] = t->counts[s]; o.off[s + 1] = o.off[s] + t->counts[s]; o.base[s] = t->base[s];
    }
    return 0;
}

template <typename T> struct StateDev {
    const T* pts;
    T *U, *m0, *q, *W, *var, *mean, *cst, *sig, *bv, *leaf_loss;
    T *Gw, *Gdot, *Gc, *Gb, *Gq, *Gm0, *gU, *g_pts, *Gscr;
};
template <typename T> static StateDev<T> to_dev(const nmb_tree_state_t& s) {
    StateDev<T> o;
    o.pts = (const T*)s.pts; o.U = (T*)s.U; o.m0 = (T*)s.m0; o.q = (T*)s.q; o.W = (T*)s.W; o.var = (T*)s.var;
    o.mean = (T*)s.mean; o.cst = (T*)s.cst; o.sig = (T*)s.sig; o.bv = (T*)s.bv; o.leaf_loss = (T*)s.leaf_loss;
    o.Gw = (T*)s.Gw; o.Gdot = (T*)s.Gdot; o.Gc = (T*)s.Gc; o.Gb = (T*)s.Gb; o.Gq = (T*)s.Gq; o.Gm0 = (T*)s.Gm0;
    o.gU = (T*)s.gU; o.g_pts = (T*)s.g_pts; o.Gscr = (T*)s.Gscr;
    return o;
}

// sqrt of a 1x1 covariance with GPyTorch's psd_safe_cholesky jitter rule:
// var <= 0 -> retry with var + 1e-8 (1e-6 in FP32), x10, x100.
template <typename T> __device__ __forceinline__ T safe_sigma(T var, T& jit) {
    jit = T(0);
    if (var > T(0)) return Num<T>::sqrt(var);
    T j = Num<T>::kChJitter;
    for (int i = 0; i < 3; ++i, j *= T(10)) {
        if (var + j > T(0)) { jit = j; return Num<T>::sqrt(var + j); }
    }
    jit = j;
    return T(0);
}

// amp * kappa(a, b) and its gradient w.r.t. a (scaled points are formed here)
template <typename T, int DMAX>
__device__ __forceinline__ T kern_pair(const GpDev<T>& g, const T* __restrict__ a, const T* __restrict__ b, T (&ga)[DMAX]) {
    T sd[DMAX], r2 = T(0);
#pragma unroll
    for (int k = 0; k < DMAX; ++k) {
        sd[k] = k < g.d ? (a[k] - b[k]) * __ldg(g.inv_ls + k) : T(0);
        r2 = fma(sd[k], sd[k], r2);
    }
    T kap, gc;
    kernel_profile<T>(g.kernel_id, r2, kap, gc);
#pragma unroll
    for (int k = 0; k < DMAX; ++k) ga[k] = k < g.d ? g.amp * gc * sd[k] * __ldg(g.inv_ls + k) : T(0);
    return g.amp * kap;
}

// ------------------------------------------------------------------ level forward: warp per node
template <typename T, int DMAX>
__global__ void __launch_bounds__(256) level_fwd_kernel(GpDev<T> g, const __grid_constant__ TreeDev tr, StateDev<T> st, int s) {
    const int lane = threadIdx.x & 31;
    const long long warp0 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    const int NL = tr.n_levels;
    const bool leaf = (s == NL - 1);
    // restart-major order inside the level: the nodes of one restart's subtree run on neighbouring warps, so the
    // ancestors' n-vectors (shared by up to hundreds of nodes) are L2 hits; the s dots are accumulated together
    // from one pass over the node's own vector (independent load streams instead of s dependent reductions).
    const long long n_r = tr.cnt[0], per = tr.cnt[s] / n_r;
    for (long long item = warp0; item < tr.cnt[s]; item += nwarps) {
        const long long i = (item % per) * n_r + item / per;
        const long long node = tr.off[s] + i;
        const T* ux = st.U + (size_t)node * g.ldu;
        T w[NMB_MAX_LEVELS];
        T sumw2 = T(0), mb = T(0);
        // b of this node's branch under its parent's model: sigma_parent * z_base[f] / c_parent
        T bown = T(0);
        if (s > 0) {
            const long long par = tr.off[s - 1] + i % tr.cnt[s - 1];
            const T zb = ((const T*)tr.base[s - 1])[i / tr.cnt[s - 1]];
            bown = tr.obs_noise ? zb : st.sig[par] * zb / st.cst[par];
        }
        T dots[NMB_MAX_LEVELS];
        for (int t0 = 0; t0 < s; t0 += 4) {
            T d4[4] = {T(0), T(0), T(0), T(0)};
            const T* uz[4];
#pragma unroll
            for (int b = 0; b < 4; ++b) uz[b] = st.U + (size_t)(tr.off[min(t0 + b, s - 1)] + i % tr.cnt[min(t0 + b, s - 1)]) * g.ldu;
            // (independent load streams: unrolled so that ~20 loads per lane are in flight - the kernel ran at 1.9 TB/s)
#pragma unroll 4
            for (int k = lane; k < g.n; k += 32) {
                const T x = ux[k];
#pragma unroll
                for (int b = 0; b < 4; ++b) d4[b] = fma(x, uz[b][k], d4[b]);
            }
#pragma unroll
            for (int b = 0; b < 4; ++b) if (t0 + b < s) dots[t0 + b] = warp_sum(d4[b]);
        }
        for (int t = 0; t < s; ++t) {
            const long long anc = tr.off[t] + i % tr.cnt[t];
            const T dot = dots[t];
            T ga[DMAX];
            const T kx = kern_pair<T, DMAX>(g, st.pts + node * g.d, st.pts + anc * g.d, ga);
            T acc = kx - dot;
            const T* wz = st.W + (size_t)anc * NL;
            for (int t2 = 0; t2 < t; ++t2) acc = fma(-wz[t2], w[t2], acc);
            w[t] = acc / st.cst[anc];
            sumw2 = fma(w[t], w[t], sumw2);
            mb = fma(w[t], t == s - 1 ? bown : st.bv[tr.off[t + 1] + i % tr.cnt[t + 1]], mb);
        }
        const T var = g.amp - st.q[node] - sumw2;
        const T mean = g.mean_const + st.m0[node] + mb;
        T jit;
        const T sig = safe_sigma<T>(var, jit);
        if (lane == 0) {
            for (int t = 0; t < s; ++t) st.W[(size_t)node * NL + t] = w[t];
            st.var[node] = var; st.mean[node] = mean; st.sig[node] = sig;
            st.cst[node] = Num<T>::sqrt(var + g.noise);
            if (leaf) st.leaf_loss[i] = -(mean + sig * (T)tr.action_base_mean);
        }
        if (s > 0 && lane == 0) st.bv[node] = bown;
    }
}

// ------------------------------------------------------------------ level backward: thread per node
template <typename T, int DMAX>
__global__ void __launch_bounds__(128) level_bwd_kernel(GpDev<T> g, const __grid_constant__ TreeDev tr, StateDev<T> st, int s,
                                                        const T* __restrict__ g_leaf, const T* __restrict__ g_leaf_sig) {
    const int NL = tr.n_levels;
    const bool leaf = (s == NL - 1);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < tr.cnt[s];
         i += (long long)gridDim.x * blockDim.x) {
        const long long node = tr.off[s] + i;
        const T sig = st.sig[node], c = st.cst[node];
        T g_var = T(0), g_mean = T(0);
        if (leaf) {
            // HES: g_leaf is the adjoint of leaf_loss = -(mean + sigma zbar); otherwise (g_leaf, g_leaf_sig)
            // are the adjoints of the leaf's (mean, sigma)
            const T gl = g_leaf[i];
            g_mean = g_leaf_sig ? gl : -gl;
            const T g_sig = g_leaf_sig ? g_leaf_sig[i] : -gl * (T)tr.action_base_mean;
            g_var = sig > T(0) ? g_sig / (T(2) * sig) : T(0);
        } else {
            // c = sqrt(var + noise) is used by every descendant's division; sigma only through b = sigma z / c
            T g_c = st.Gc[node], g_sig = T(0);
            if (!tr.obs_noise) {
                const long long nf = tr.cnt[s + 1] / tr.cnt[s];
                const T* zb = (const T*)tr.base[s];
                for (long long f = 0; f < nf; ++f) {
                    const T gb = st.Gb[tr.off[s + 1] + i + f * tr.cnt[s]];
                    g_sig = fma(gb, zb[f] / c, g_sig);
                    g_c = fma(-gb, sig * zb[f] / (c * c), g_c);
                }
            }
            g_var = g_c / (T(2) * c) + (sig > T(0) ? g_sig / (T(2) * sig) : T(0));
        }
        T w[NMB_MAX_LEVELS], gw[NMB_MAX_LEVELS];
        // scratch record of (target level t, field f) for this node: Gscr[(t R + f) cnt_s + i], R = 2 + d + NL;
        // fields: 0 -> Gc, 1 -> Gb, 2.. -> g_pts[d], 2 + d.. -> Gw[t2 < t]
        const int R = 2 + g.d + NL;
        const size_t cs = (size_t)tr.cnt[s];
        T* scr = st.Gscr + i;
        for (int t = 0; t < s; ++t) {
            w[t] = st.W[(size_t)node * NL + t];
            const long long bn = tr.off[t + 1] + i % tr.cnt[t + 1];
            gw[t] = st.Gw[(size_t)node * NL + t] - T(2) * w[t] * g_var + st.bv[bn] * g_mean;
            // b of the branch under ancestor level t + 1 (the node itself when t + 1 == s)
            if (t + 1 == s) st.Gb[node] += w[t] * g_mean;
            else scr[((size_t)(t + 1) * R + 1) * cs] = w[t] * g_mean;
        }
        if (s > 0) scr[(size_t)1 * cs] = T(0);          // level-0 nodes have no b
        T gx[DMAX];
#pragma unroll
        for (int k = 0; k < DMAX; ++k) gx[k] = T(0);
        for (int t = s - 1; t >= 0; --t) {
            const long long anc = tr.off[t] + i % tr.cnt[t];
            const T ct = st.cst[anc];
            const T g_acc = gw[t] / ct;
            T* rec = scr + (size_t)t * R * cs;
            rec[0] = -gw[t] * w[t] / ct;
            st.Gdot[(size_t)node * NL + t] = -g_acc;
            T ga[DMAX];
            kern_pair<T, DMAX>(g, st.pts + node * g.d, st.pts + anc * g.d, ga);
#pragma unroll
            for (int k = 0; k < DMAX; ++k) {
                if (k < g.d) {
                    gx[k] = fma(g_acc, ga[k], gx[k]);
                    rec[(size_t)(2 + k) * cs] = -g_acc * ga[k];
                }
            }
            const T* wz = st.W + (size_t)anc * NL;
            for (int t2 = 0; t2 < t; ++t2) {
                rec[(size_t)(2 + g.d + t2) * cs] = -g_acc * w[t2];
                gw[t2] = fma(-g_acc, wz[t2], gw[t2]);
            }
        }
        st.Gq[node] = -g_var;
        st.Gm0[node] = g_mean;
#pragma unroll
        for (int k = 0; k < DMAX; ++k) if (k < g.d) st.g_pts[node * g.d + k] += gx[k];
    }
}

// Sum the level-s nodes' scratch records into their level-t ancestors' adjoint slots (blockIdx.y = t < s).
// Ancestor a of level t owns the nodes i = a + j cnt_t, j < cnt_s / cnt_t; a group of G lanes (G a power of two,
// <= 32, warp-uniform) walks j = lane, lane + G, ... and the partial sums meet in a fixed xor-shuffle order.
template <typename T>
__global__ void __launch_bounds__(256) level_bwd_reduce_kernel(const __grid_constant__ TreeDev tr, StateDev<T> st, int s, int d) {
    const int t = blockIdx.y, NL = tr.n_levels, R = 2 + d + NL;
    const int nf = 2 + d + t;                          // fields that level t receives
    const long long ct = tr.cnt[t], cs = tr.cnt[s], desc = cs / ct;
    int G = 1;
    while (G < 32 && G < desc) G <<= 1;
    const long long items = ct * nf;
    const int lane = threadIdx.x & (G - 1);
    for (long long item = ((long long)blockIdx.x * blockDim.x + threadIdx.x) / G; item < ((items + 31) / 32) * 32;
         item += (long long)gridDim.x * blockDim.x / G) {
        const bool live = item < items;
        const int f = live ? (int)(item / ct) : 0;
        const long long a = live ? item % ct : 0;
        T acc = T(0);
        if (live && !(f == 1 && t == 0)) {
            const T* src = st.Gscr + ((size_t)t * R + f) * cs + a;
            for (long long j = lane; j < desc; j += G) acc += src[(size_t)j * ct];
        }
        for (int o = G >> 1; o > 0; o >>= 1) acc += shfl_xor_(acc, o);
        if (live && lane == 0) {
            const long long anc = tr.off[t] + a;
            if (f == 0) st.Gc[anc] += acc;
            else if (f == 1) st.Gb[anc] += acc;
            else if (f < 2 + d) st.g_pts[anc * d + (f - 2)] += acc;
            else st.Gw[(size_t)anc * NL + (f - 2 - d)] += acc;
        }
    }
}

// ------------------------------------------------------------------ gU gather
// gU[x] = 2 Gq[x] u(x) + sum_{t < s} Gdot[x][t] u(z_t) + sum_{y below x} Gdot[y][s] u(y), descendants in a fixed order.
// Pure streaming of n-vectors: (1 + s + #descendants) reads per node.  Two things matter:
//  * work is issued restart-major (all nodes of one restart's tree, ~MBs, by neighbouring warps at the same time), so
//    a vector read by its s ancestors and by itself is served from L2 after the first touch instead of HBM;
//  * a node with hundreds of descendants (the root of a [64,4,2] tree has 832) gets a whole CTA: the warps split the
//    descendant list, partial sums meet in shared memory in warp order (deterministic).
// Descendants are visited four at a time so that 16 independent loads per lane are in flight.
template <typename T> struct DescIter {
    const TreeDev& tr;
    int s;
    long long i;
    __device__ DescIter(const TreeDev& t, int s_, long long i_) : tr(t), s(s_), i(i_) {}
    __device__ long long count() const {
        long long c = 0;
        for (int s2 = s + 1; s2 < tr.n_levels; ++s2) c += tr.cnt[s2] / tr.cnt[s];
        return c;
    }
    // e-th descendant in (level, then sibling) order -> global node index
    __device__ long long node(long long e) const {
        for (int s2 = s + 1; s2 < tr.n_levels; ++s2) {
            const long long c = tr.cnt[s2] / tr.cnt[s];
            if (e < c) return tr.off[s2] + i + e * tr.cnt[s];
            e -= c;
        }
        return -1;
    }
};

template <typename T>
__device__ __forceinline__ void gather_desc(const GpDev<T>& g, const TreeDev& tr, const StateDev<T>& st, int s, long long i,
                                            long long e0, long long e1, long long estep, int k0, int lane, T (&a)[4]) {
    const DescIter<T> it(tr, s, i);
    const int NL = tr.n_levels;
    for (long long e = e0; e < e1; e += 4 * estep) {
        T gd[4], v[4][4];
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            const long long eb = e + b * estep;
            const long long dn = eb < e1 ? it.node(eb) : -1;
            gd[b] = dn >= 0 ? st.Gdot[(size_t)dn * NL + s] : T(0);
            const T* uy = st.U + (size_t)(dn >= 0 ? dn : 0) * g.ldu;
#pragma unroll
            for (int u = 0; u < 4; ++u) { const int k = k0 + lane + 32 * u; v[b][u] = (dn >= 0 && k < g.n) ? uy[k] : T(0); }
        }
#pragma unroll
        for (int b = 0; b < 4; ++b)
#pragma unroll
            for (int u = 0; u < 4; ++u) a[u] = fma(gd[b], v[b][u], a[u]);
    }
}

// restart-major work order over the levels [s_lo, s_hi): item -> (level, index inside the level)
__device__ __forceinline__ void restart_major(const TreeDev& tr, int s_lo, int s_hi, long long item, int& s, long long& i) {
    const long long n_r = tr.cnt[0];
    long long per = 0;
    for (int l = s_lo; l < s_hi; ++l) per += tr.cnt[l] / n_r;
    const long long r = item / per;
    long long m = item - r * per;
    s = s_lo;
    while (m >= tr.cnt[s] / n_r) { m -= tr.cnt[s] / n_r; ++s; }
    i = m * n_r + r;
}

template <typename T, bool CTA_PER_NODE>
__global__ void __launch_bounds__(256) gather_gu_kernel(GpDev<T> g, const __grid_constant__ TreeDev tr, StateDev<T> st, int s_lo,
                                                        int s_hi, long long n_items) {
    __shared__ T part[8][128];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int NL = tr.n_levels;
    const long long unit0 = CTA_PER_NODE ? blockIdx.x : (((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    const long long nunits = CTA_PER_NODE ? gridDim.x : (((long long)gridDim.x * blockDim.x) >> 5);
    // CTA per node: a work unit is (node, block of 128 columns) - the columns are independent, and with few restarts
    // the upper-level nodes alone (one per restart, 832 descendant rows each at nf = [64, 4, 2]) left most SMs idle
    const int nkb = CTA_PER_NODE ? (g.n + 127) / 128 : 1;
    for (long long unit = unit0; unit < n_items * nkb; unit += nunits) {
        const long long item = unit / nkb;
        int s;
        long long i;
        restart_major(tr, s_lo, s_hi, item, s, i);
        const long long node = tr.off[s] + i;
        const T* ux = st.U + (size_t)node * g.ldu;
        T* out = st.gU + (size_t)node * g.ldu;
        const T gq2 = T(2) * st.Gq[node];
        const long long D = DescIter<T>(tr, s, i).count();
        const int k_lo = CTA_PER_NODE ? (int)(unit - item * nkb) * 128 : 0;
        const int k_hi = CTA_PER_NODE ? min(g.n, k_lo + 128) : g.n;
        for (int k0 = k_lo; k0 < k_hi; k0 += 128) {
            T a[4] = {T(0), T(0), T(0), T(0)};
            if (CTA_PER_NODE) {
                gather_desc<T>(g, tr, st, s, i, warp, D, 8, k0, lane, a);
#pragma unroll
                for (int u = 0; u < 4; ++u) part[warp][lane + 32 * u] = a[u];
                __syncthreads();
                if (threadIdx.x < 128 && k0 + threadIdx.x < g.n) {
                    const int k = k0 + threadIdx.x;
                    T v = gq2 * ux[k];
                    for (int t = 0; t < s; ++t)
                        v = fma(st.Gdot[(size_t)node * NL + t], st.U[(size_t)(tr.off[t] + i % tr.cnt[t]) * g.ldu + k], v);
#pragma unroll
                    for (int w = 0; w < 8; ++w) v += part[w][threadIdx.x];
                    out[k] = v;
                }
                __syncthreads();
            } else {
#pragma unroll
                for (int u = 0; u < 4; ++u) { const int k = k0 + lane + 32 * u; a[u] = k < g.n ? gq2 * ux[k] : T(0); }
                for (int t = 0; t < s; ++t) {
                    const T gd = st.Gdot[(size_t)node * NL + t];
                    const T* uz = st.U + (size_t)(tr.off[t] + i % tr.cnt[t]) * g.ldu;
#pragma unroll
                    for (int u = 0; u < 4; ++u) { const int k = k0 + lane + 32 * u; if (k < g.n) a[u] = fma(gd, uz[k], a[u]); }
                }
                gather_desc<T>(g, tr, st, s, i, 0, D, 1, k0, lane, a);
#pragma unroll
                for (int u = 0; u < 4; ++u) { const int k = k0 + lane + 32 * u; if (k < g.n) out[k] = a[u]; }
            }
        }
    }
}

// ------------------------------------------------------------------ per-restart mean over leaves
template <typename T>
__global__ void __launch_bounds__(256) restart_mean_kernel(const T* __restrict__ leaf, long long n_leaf, long long n_r,
                                                           T* __restrict__ out) {
    // one warp per restart: lanes stride over its leaves, shuffle tree reduction
    const int lane = threadIdx.x & 31;
    const long long r = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (r >= n_r) return;
    const long long per = n_leaf / n_r;
    T acc = T(0);
    for (long long f = lane; f < per; f += 32) acc += leaf[r + f * n_r];
    acc = warp_sum(acc);
    if (lane == 0) out[r] = acc / (T)per;
}

// ------------------------------------------------------------------ baseline utilities (q = 1)
template <typename T>
__global__ void __launch_bounds__(256) utility_kernel(int which, T best_f, T beta, const T* __restrict__ mean,
                                                      const T* __restrict__ sig, long long n, const T* __restrict__ base,
                                                      int S, T* __restrict__ value, T* __restrict__ d_mean,
                                                      T* __restrict__ d_sig) {
    // one warp per point, lanes stride over the S base samples
    const int lane = threadIdx.x & 31;
    const long long p = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (p >= n) return;
    const T m = mean[p], sg = sig[p];
    T v = T(0), dm = T(0), ds = T(0);
    if (which == 3) {
        // mu^ = mean_s y = m + sg * zbar;  u_s = mu^ + k |y_s - mu^| = mu^ + k sg |z_s - zbar|
        T zs = T(0);
        for (int a = lane; a < S; a += 32) zs += base[a];
        const T zbar = warp_sum(zs) / (T)S;
        const T kq = Num<T>::sqrt(beta * Num<T>::kPi / T(2));
        T ab = T(0);
        for (int a = lane; a < S; a += 32) ab += Num<T>::fabs(sg * (base[a] - zbar));
        ab = warp_sum(ab) / (T)S;
        v = m + sg * zbar + kq * ab;
        dm = T(1);
        ds = zbar + (sg != T(0) ? kq * ab / sg : T(0));
    } else {
        for (int a = lane; a < S; a += 32) {
            const T z = base[a];
            const T y = fma(sg, z, m);
            if (which == 0) {
                if (y - best_f > T(0)) { v += y - best_f; dm += T(1); ds += z; }
            } else if (which == 1) {
                const T sgm = sigmoid_((y - best_f) / T(1e-3));
                const T dd = sgm * (T(1) - sgm) / T(1e-3);
                v += sgm; dm += dd; ds += dd * z;
            } else {
                v += y; dm += T(1); ds += z;
            }
        }
        v = warp_sum(v) / (T)S; dm = warp_sum(dm) / (T)S; ds = warp_sum(ds) / (T)S;
    }
    if (lane == 0) { value[p] = v; d_mean[p] = dm; d_sig[p] = ds; }
}
}  // namespace nmb

using namespace nmb;

static int validate_gp(const nmb_gp_t* gp, int dtype, const char* who) {
    NMB_REQUIRE(gp, NMB_EINVAL, "%s: null gp", who);
    NMB_REQUIRE(dtype == NMB_F64 || dtype == NMB_F32, NMB_EINVAL, "%s: dtype %d", who, dtype);
    NMB_REQUIRE(gp->d >= 1 && gp->d <= NMB_MAX_DIM, NMB_ELIMIT, "%s: d=%d outside [1,%d]", who, gp->d, NMB_MAX_DIM);
    NMB_REQUIRE(gp->n >= 1 && gp->ldu >= gp->n, NMB_EINVAL, "%s: n=%d ldu=%d", who, gp->n, gp->ldu);
    NMB_REQUIRE(gp->kernel_id >= 0 && gp->kernel_id <= 3, NMB_EINVAL, "%s: kernel_id %d", who, gp->kernel_id);
    NMB_REQUIRE(gp->ldl >= gp->n && gp->ldl % 8 == 0, NMB_EINVAL, "%s: ldl=%d must be a multiple of 8 and >= n=%d", who,
                gp->ldl, gp->n);
    NMB_REQUIRE(gp->train_x && gp->inv_ls && gp->linv && gp->linv_t && gp->alpha, NMB_EINVAL, "%s: null gp array", who);
    return 0;
}

#define NMB_MC_D(d, CALL)                 \
    do {                                  \
        if ((d) <= 2) { CALL(2); }        \
        else if ((d) <= 4) { CALL(4); }   \
        else if ((d) <= 8) { CALL(8); }   \
        else { CALL(16); }                \
    } while (0)

// FP64: tensor-pipe kernels (whiten_mma.cuh); rows per CTA / ring depth chosen so the tile fits shared memory
template <int DM>
static int launch_whiten_mma_d(const nmb_gp_t* gp, bool bwd, const double* pts, int64_t n_pts, double* U, double* m0,
                               double* q, const double* gU, const double* g_m0, double* g_pts, cudaStream_t st) {
    NMB_REQUIRE(gp->linv_tiles && gp->linv_t_tiles, NMB_EINVAL, "gp whiten (FP64): linv_tiles / linv_t_tiles are null");
    GpDev<double> g = to_dev<double>(*gp);
    static const int cand[5][2] = {{32, 4}, {32, 3}, {24, 4}, {24, 3}, {16, 3}};
    int R = 0, nst = 0;
    size_t smem = 0;
    for (int i = 0; i < 5 && !R; ++i) {
        smem = wh_mma_bytes<DM>(gp->ldl, cand[i][0], cand[i][1], bwd);
        if (smem <= 227 * 1024) { R = cand[i][0]; nst = cand[i][1]; }
    }
    NMB_REQUIRE(R > 0, NMB_ELIMIT, "gp whiten: n=%d training points exceed shared memory", gp->n);
    const unsigned grid = (unsigned)((n_pts + R - 1) / R), nthr = (unsigned)((R / 8) * 2 + 1) * 32;
#define NMB_WHM(NST)                                                                                                  \
    if (!bwd) {                                                                                                       \
        NMB_CUDA(cudaFuncSetAttribute(whiten_fwd_mma_kernel<DM, NST>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
        whiten_fwd_mma_kernel<DM, NST><<<grid, (unsigned)((R / 8) * (8 / kWhNT) + 1) * 32, smem, st>>>(g, gp->ldl, R, pts, n_pts, U, m0, q); \
    } else {                                                                                                          \
        NMB_CUDA(cudaFuncSetAttribute(whiten_bwd_mma_kernel<DM, NST>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
        whiten_bwd_mma_kernel<DM, NST><<<grid, (unsigned)((R / 8) * (8 / kWhNT) + 1) * 32, smem, st>>>(g, gp->ldl, R, pts, n_pts, gU, g_m0, g_pts); \
    }
    if (nst == 4) { NMB_WHM(4) } else { NMB_WHM(3) }
#undef NMB_WHM
    NMB_LAUNCH_CHECK();
    return 0;
}

static int launch_whiten_mma(const nmb_gp_t* gp, bool bwd, const void* pts, int64_t n_pts, void* U, void* m0, void* q,
                             const void* gU, const void* g_m0, void* g_pts, cudaStream_t st) {
#define NMB_WD(DM) return launch_whiten_mma_d<DM>(gp, bwd, (const double*)pts, n_pts, (double*)U, (double*)m0, (double*)q, \
                                                  (const double*)gU, (const double*)g_m0, (double*)g_pts, st)
    NMB_MC_D(gp->d, NMB_WD);
#undef NMB_WD
    return 0;
}

template <typename T>
static int launch_whiten(const nmb_gp_t* gp, bool bwd, const void* pts, int64_t n_pts, void* U, void* m0, void* q,
                         const void* gU, const void* g_m0, void* g_pts, cudaStream_t st) {
    GpDev<T> g = to_dev<T>(*gp);
    const unsigned grid = (unsigned)((n_pts + kR - 1) / kR);
#define NMB_WH(DM)                                                                                              \
    {                                                                                                           \
        const size_t smem = whiten_smem<T, DM>(gp->n, bwd);                                                     \
        NMB_REQUIRE(smem <= 227 * 1024, NMB_ELIMIT, "gp whiten: n=%d training points exceed shared memory", gp->n); \
        if (!bwd) {                                                                                             \
            NMB_CUDA(cudaFuncSetAttribute(whiten_fwd_kernel<T, DM>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
            whiten_fwd_kernel<T, DM><<<grid, kNT, smem, st>>>(g, (const T*)pts, n_pts, (T*)U, (T*)m0, (T*)q);   \
        } else {                                                                                                \
            NMB_CUDA(cudaFuncSetAttribute(whiten_bwd_kernel<T, DM>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
            whiten_bwd_kernel<T, DM><<<grid, kNT, smem, st>>>(g, (const T*)pts, n_pts, (const T*)gU, (const T*)g_m0, (T*)g_pts); \
        }                                                                                                       \
    }
    NMB_MC_D(gp->d, NMB_WH);
#undef NMB_WH
    NMB_LAUNCH_CHECK();
    return 0;
}

extern "C" int nmb_gp_whiten_fwd(const nmb_gp_t* gp, int dtype, const void* pts, int64_t n_pts, void* U, void* m0,
                                 void* q, void* stream) {
    int rc = validate_gp(gp, dtype, "nmb_gp_whiten_fwd");
    if (rc) return rc;
    NMB_REQUIRE(pts && U && m0 && q && n_pts >= 0, NMB_EINVAL, "nmb_gp_whiten_fwd: null argument");
    if (n_pts == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    return dtype == NMB_F64 ? launch_whiten_mma(gp, false, pts, n_pts, U, m0, q, nullptr, nullptr, nullptr, st)
                            : launch_whiten<float>(gp, false, pts, n_pts, U, m0, q, nullptr, nullptr, nullptr, st);
}

extern "C" int nmb_gp_whiten_bwd(const nmb_gp_t* gp, int dtype, const void* pts, int64_t n_pts, const void* gU,
                                 const void* g_m0, void* g_pts, void* stream) {
    int rc = validate_gp(gp, dtype, "nmb_gp_whiten_bwd");
    if (rc) return rc;
    NMB_REQUIRE(pts && gU && g_m0 && g_pts && n_pts >= 0, NMB_EINVAL, "nmb_gp_whiten_bwd: null argument");
    if (n_pts == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    return dtype == NMB_F64 ? launch_whiten_mma(gp, true, pts, n_pts, nullptr, nullptr, nullptr, gU, g_m0, g_pts, st)
                            : launch_whiten<float>(gp, true, pts, n_pts, nullptr, nullptr, nullptr, gU, g_m0, g_pts, st);
}

static unsigned warp_grid(long long n_warps) {
    long long b = (n_warps + 7) / 8;
    const long long cap = (long long)sm_count() * 16;
    return (unsigned)(b < 1 ? 1 : (b > cap ? cap : b));
}

template <typename T>
static int launch_level(const nmb_gp_t* gp, const TreeDev& tr, const nmb_tree_state_t* state, int level, bool bwd,
                        const void* g_leaf, const void* g_leaf_sig, cudaStream_t st) {
    GpDev<T> g = to_dev<T>(*gp);
    StateDev<T> sd = to_dev<T>(*state);
    if (!bwd) {
#define NMB_LF(DM) level_fwd_kernel<T, DM><<<warp_grid(tr.cnt[level]), 256, 0, st>>>(g, tr, sd, level)
        NMB_MC_D(gp->d, NMB_LF);
#undef NMB_LF
    } else {
        long long b = (tr.cnt[level] + 127) / 128;
        const long long cap = (long long)sm_count() * 16;
        const unsigned grid = (unsigned)(b > cap ? cap : b);
#define NMB_LB(DM) level_bwd_kernel<T, DM><<<grid, 128, 0, st>>>(g, tr, sd, level, (const T*)g_leaf, (const T*)g_leaf_sig)
        NMB_MC_D(gp->d, NMB_LB);
#undef NMB_LB
        if (level > 0) {
            long long mx = 0;
            for (int t = 0; t < level; ++t) {
                long long G = 1;
                const long long desc = tr.cnt[level] / tr.cnt[t];
                while (G < 32 && G < desc) G <<= 1;
                mx = std::max(mx, tr.cnt[t] * (2 + gp->d + t) * G);
            }
            long long rb = (mx + 255) / 256;
            rb = rb < 1 ? 1 : (rb > cap ? cap : rb);
            level_bwd_reduce_kernel<T><<<dim3((unsigned)rb, (unsigned)level), 256, 0, st>>>(tr, sd, level, gp->d);
        }
    }
    NMB_LAUNCH_CHECK();
    return 0;
}

extern "C" int nmb_mc_level_fwd(const nmb_gp_t* gp, const nmb_tree_t* tree, const nmb_tree_state_t* st, int dtype,
                                int level, void* stream) {
    int rc = validate_gp(gp, dtype, "nmb_mc_level_fwd");
    if (rc) return rc;
    TreeDev tr;
    rc = make_tree(tree, tr);
    if (rc) return rc;
    NMB_REQUIRE(st && level >= 0 && level < tr.n_levels, NMB_EINVAL, "nmb_mc_level_fwd: level %d", level);
    NMB_REQUIRE(st->pts && st->U && st->m0 && st->q && st->W && st->var && st->mean && st->cst && st->sig && st->bv,
                NMB_EINVAL, "nmb_mc_level_fwd: null state array");
    NMB_REQUIRE(level < tr.n_levels - 1 || st->leaf_loss, NMB_EINVAL, "nmb_mc_level_fwd: null leaf_loss");
    cudaStream_t s = (cudaStream_t)stream;
    return dtype == NMB_F64 ? launch_level<double>(gp, tr, st, level, false, nullptr, nullptr, s)
                            : launch_level<float>(gp, tr, st, level, false, nullptr, nullptr, s);
}

extern "C" int64_t nmb_mc_level_bwd_scratch_elems(const nmb_tree_t* tree, int d) {
    if (!tree || tree->n_levels < 1 || tree->n_levels > NMB_MAX_LEVELS) return -1;
    int64_t mx = 1;
    for (int s = 1; s < tree->n_levels; ++s)
        mx = std::max<int64_t>(mx, tree->counts[s] * (int64_t)s * (2 + d + tree->n_levels));
    return mx;
}

extern "C" int nmb_mc_level_bwd(const nmb_gp_t* gp, const nmb_tree_t* tree, const nmb_tree_state_t* st, int dtype,
                                int level, const void* g_leaf, const void* g_leaf_sig, void* stream) {
    int rc = validate_gp(gp, dtype, "nmb_mc_level_bwd");
    if (rc) return rc;
    TreeDev tr;
    rc = make_tree(tree, tr);
    if (rc) return rc;
    NMB_REQUIRE(st && level >= 0 && level < tr.n_levels, NMB_EINVAL, "nmb_mc_level_bwd: level %d", level);
    NMB_REQUIRE(st->Gw && st->Gdot && st->Gc && st->Gb && st->Gq && st->Gm0 && st->g_pts, NMB_EINVAL,
                "nmb_mc_level_bwd: null adjoint array");
    NMB_REQUIRE(level == 0 || st->Gscr, NMB_EINVAL, "nmb_mc_level_bwd: null Gscr (nmb_mc_level_bwd_scratch_elems)");
    NMB_REQUIRE(level < tr.n_levels - 1 || g_leaf, NMB_EINVAL, "nmb_mc_level_bwd: null g_leaf");
    cudaStream_t s = (cudaStream_t)stream;
    return dtype == NMB_F64 ? launch_level<double>(gp, tr, st, level, true, g_leaf, g_leaf_sig, s)
                            : launch_level<float>(gp, tr, st, level, true, g_leaf, g_leaf_sig, s);
}

extern "C" int nmb_mc_gather_gu(const nmb_gp_t* gp, const nmb_tree_t* tree, const nmb_tree_state_t* st, int dtype,
                                void* stream) {
    int rc = validate_gp(gp, dtype, "nmb_mc_gather_gu");
    if (rc) return rc;
    TreeDev tr;
    rc = make_tree(tree, tr);
    if (rc) return rc;
    NMB_REQUIRE(st && st->U && st->gU && st->Gq && st->Gdot, NMB_EINVAL, "nmb_mc_gather_gu: null array");
    cudaStream_t s = (cudaStream_t)stream;
    // levels whose nodes have >= 64 descendants get a CTA per node, the rest a warp per node
    int s_heavy = 0;
    while (s_heavy < tr.n_levels) {
        long long desc = 0;
        for (int s2 = s_heavy + 1; s2 < tr.n_levels; ++s2) desc += tr.cnt[s2] / tr.cnt[s_heavy];
        if (desc < 64) break;
        ++s_heavy;
    }
    const long long n_heavy = tr.off[s_heavy], n_light = tr.off[tr.n_levels] - n_heavy;
#define NMB_GG(T)                                                                                                       \
    do {                                                                                                                \
        if (n_heavy > 0) {                                                                                              \
            const unsigned grid = (unsigned)std::min<long long>(n_heavy * ((gp->n + 127) / 128), 16LL * sm_count());    \
            gather_gu_kernel<T, true><<<grid, 256, 0, s>>>(to_dev<T>(*gp), tr, to_dev<T>(*st), 0, s_heavy, n_heavy);    \
        }                                                                                                               \
        if (n_light > 0)                                                                                                \
            gather_gu_kernel<T, false><<<warp_grid(n_light), 256, 0, s>>>(to_dev<T>(*gp), tr, to_dev<T>(*st), s_heavy,  \
                                                                         tr.n_levels, n_light);                         \
    } while (0)
    if (dtype == NMB_F64) NMB_GG(double);
    else NMB_GG(float);
#undef NMB_GG
    NMB_LAUNCH_CHECK();
    return 0;
}

extern "C" int nmb_tree_restart_mean(int dtype, const void* leaf, int64_t n_leaf, int64_t n_r, void* out, void* stream) {
    NMB_REQUIRE(leaf && out && n_r >= 1 && n_leaf >= n_r && n_leaf % n_r == 0, NMB_EINVAL,
                "nmb_tree_restart_mean: bad arguments");
    cudaStream_t s = (cudaStream_t)stream;
    const unsigned grid = (unsigned)((n_r * 32 + 255) / 256);
    if (dtype == NMB_F64) restart_mean_kernel<double><<<grid, 256, 0, s>>>((const double*)leaf, n_leaf, n_r, (double*)out);
    else if (dtype == NMB_F32) restart_mean_kernel<float><<<grid, 256, 0, s>>>((const float*)leaf, n_leaf, n_r, (float*)out);
    else { set_error("nmb_tree_restart_mean: dtype %d", dtype); return NMB_EINVAL; }
    NMB_LAUNCH_CHECK();
    return 0;
}

extern "C" int nmb_mc_utility(int dtype, int which, double best_f, double beta, const void* mean, const void* sig,
                              int64_t n, const void* base, int n_samples, void* value, void* d_mean, void* d_sig,
                              void* stream) {
    NMB_REQUIRE(mean && sig && base && value && d_mean && d_sig && n >= 0 && n_samples >= 1, NMB_EINVAL,
                "nmb_mc_utility: bad arguments");
    NMB_REQUIRE(which >= 0 && which <= 3, NMB_EINVAL, "nmb_mc_utility: which=%d", which);
    if (n == 0) return 0;
    cudaStream_t s = (cudaStream_t)stream;
    const unsigned grid = (unsigned)((n * 32 + 255) / 256);
    if (dtype == NMB_F64)
        utility_kernel<double><<<grid, 256, 0, s>>>(which, best_f, beta, (const double*)mean, (const double*)sig, n,
                                                    (const double*)base, n_samples, (double*)value, (double*)d_mean,
                                                    (double*)d_sig);
    else if (dtype == NMB_F32)
        utility_kernel<float><<<grid, 256, 0, s>>>(which, (float)best_f, (float)beta, (const float*)mean,
                                                   (const float*)sig, n, (const float*)base, n_samples, (float*)value,
                                                   (float*)d_mean, (float*)d_sig);
    else { set_error("nmb_mc_utility: dtype %d", dtype); return NMB_EINVAL; }
    NMB_LAUNCH_CHECK();
    return 0;
}
