// Shared-root whitening on the FP64 tensor pipe (DMMA m8n8k4), FP64 only.
//
//   forward : U[p][:]  = k(x_p, X_tr) L^-T        (A = kernel rows generated in shared memory, row-major;
//                                                   B = L^-T streamed from L2, only rows k <= column block end)
//   backward: g_K[p][:] = gU[p][:] L^-1            (A = the gU tile; B = L^-1, only rows i >= column block start),
//             fused with g_x = sum_j (g_K[j] + g_m0 alpha_j) d k(x, X_j) / dx  in the accumulator epilogue.
//
// Same machinery as the policy GEMMs (policy_mma.cuh): a producer warp streams [32 k-rows][64 cols] chunks of the
// triangular operand with cp.async.bulk into an mbarrier ring (row stride kSW = 68 doubles: conflict-free B fragments),
// consumer warps own 8-row x 32-col accumulator tiles, A fragments come from a row-major tile with stride
// KS = 4 mod 16 doubles (conflict free).  Replaces the DFMA 2x4 register-tile product (18 % of the FP64 peak at
// n = 500: bound by the shared-memory data pipe, see the header of policy_mma.cuh).
// The triangular operands are passed zero-padded to ldl = n rounded up to 8 (k extents are multiples of 4) and
// pre-tiled by 64-column blocks, [ceil(ldl/64)][ldl][68], so that a chunk is one contiguous, 16-byte aligned bulk copy.
// Reference: GPyTorch's exact prediction strategy behind model.posterior (acqfs.py:366, 414).
#pragma once
#include "policy_mma.cuh"

namespace nmb {

__host__ __device__ inline int wh_ks(int ldl) { return (ldl + 15) / 16 * 16 + 4; }

// per consumer warp (backward only): its 32 training points of the current column block, scaled, + their alpha
// (row stride DMAX + 1: the 4 k-lanes of a fragment read points 2 apart - an even stride of 2 * DMAX doubles put all
// four in one bank, 160 M conflicts per launch at cfg 3)
template <int DMAX> __host__ __device__ constexpr int wh_tp_elems() { return 32 * (DMAX + 2); }

template <int DMAX>
inline size_t wh_mma_bytes(int ldl, int R, int nst, bool bwd) {
    const size_t tp = bwd ? (size_t)(R / 8) * 2 * wh_tp_elems<DMAX>() : 0;
    // + 64: the 2^(j/64) table of the table-driven exp (fastmath.cuh) behind the kernel profile
    return ((size_t)nst * kStageElems + (size_t)R * wh_ks(ldl) + (size_t)R * DMAX + (size_t)2 * R * DMAX + tp + 2 * nst + 64) * 8;
}

struct WhSmem {
    double *stage, *A, *xs, *red, *tp, *tab;
    unsigned full0, empty0;
    int KS;
    template <int DMAX>
    __device__ void carve(double* base, int ldl, int R, int nst, bool bwd = false) {
        KS = wh_ks(ldl);
        double* p = base;
        stage = p; p += (size_t)nst * kStageElems;
        A = p; p += (size_t)R * KS;
        xs = p; p += R * DMAX;
        red = p; p += 2 * R * DMAX;
        tp = p; p += bwd ? (R / 8) * 2 * wh_tp_elems<DMAX>() : 0;
        full0 = smem_u32(p);
        empty0 = full0 + 8 * nst;
        tab = p + 2 * nst;
    }
    // the exponential of the kernel profile only indexes the first 64 entries of the fast-math table
    __device__ void load_tab(int tid, int nthr) const {
        for (int e = tid; e < 64; e += nthr) tab[e] = g_fast_tab[e];
    }
};

// Producer warp: chunks of the triangular operand in consumption order.  The operand is pre-tiled on the host,
// Bt[column block][k][kSW] (64 columns + 4 pad = the stage row stride), so a [kn k-rows][kSW] chunk is ONE contiguous
// bulk copy that lands in the stage layout as is (32 per-row copies of 512 B left the consumers waiting on the
// copy engine: 18 % of the samples on the full-barrier wait).
// forward (LOWER_ROWS = false): column block cb needs rows [0, cb + 64); backward: rows [cb, ldl).
template <int NST, bool LOWER_ROWS>
__device__ __forceinline__ void wh_producer(const double* __restrict__ Bt, int n, int ldl, const MmaRing& rg) {
    if ((threadIdx.x & 31) != 0) return;
    unsigned c = 0;
    for (int cb = 0; cb < n; cb += kOB) {
        const double* tile = Bt + (size_t)(cb / kOB) * ldl * kSW;
        const int k_lo = LOWER_ROWS ? cb : 0, k_hi = LOWER_ROWS ? ldl : min(cb + kOB, ldl);
        for (int k0 = k_lo; k0 < k_hi; k0 += kKC, ++c) {
            const unsigned slot = c % NST, fill = c / NST;
            if (fill > 0) mbar_wait(rg.empty0 + 8 * slot, (fill - 1) & 1);
            const unsigned bytes = (unsigned)min(kKC, k_hi - k0) * kSW * 8u;
            const unsigned full = rg.full0 + 8 * slot;
            mbar_arrive_expect_tx(full, bytes);
            bulk_g2s(smem_u32(rg.stage + (size_t)slot * kStageElems), tile + (size_t)k0 * kSW, bytes, full);
        }
    }
}

__device__ __forceinline__ void wh_init_barriers(const WhSmem& sm, int nst, int n_consumer_warps) {
    if (threadIdx.x == 0) {
        for (int s = 0; s < nst; ++s) {
            mbar_init(sm.full0 + 8 * s, 1);
            mbar_init(sm.empty0 + 8 * s, n_consumer_warps);
        }
        fence_barrier_init();
    }
    __syncthreads();
}

// ------------------------------------------------------------------ forward
// NT = 8-column tiles per warp: 2 -> four warps per 8-row block (16 consumer warps at 32 rows: the 8-warp form ran at
// 2.2 warps per scheduler with `wait` as its first stall reason), 4 -> two warps per block.
constexpr int kWhNT = 2;
template <int DMAX, int NST, int NT = kWhNT>
__global__ void __launch_bounds__(32 * (32 / 8) * (8 / NT) + 32, 1) whiten_fwd_mma_kernel(GpDev<double> g, int ldl, int R,
                                                                const double* __restrict__ pts, long long n_pts,
                                                                double* __restrict__ U, double* __restrict__ m0,
                                                                double* __restrict__ q) {
    extern __shared__ __align__(16) unsigned char wh_smem_raw[];
    const int n = g.n, d = g.d, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int ncw = (R >> 3) * (8 / NT), nthr = ncw * 32;      // every warp generates NT of the R kernel rows
    WhSmem sm;
    sm.carve<DMAX>(reinterpret_cast<double*>(wh_smem_raw), ldl, R, NST);
    wh_init_barriers(sm, NST, ncw);
    const MmaRing ring{sm.stage, sm.full0, sm.empty0, sm.KS};
    if (warp == ncw) {
        wh_producer<NST, false>(g.linv_t_tiles, n, ldl, ring);
        return;
    }
    const int KS = sm.KS;
    const long long p0 = (long long)blockIdx.x * R;
    const int nvalid = (int)min((long long)R, n_pts - p0);
    sm.load_tab(tid, nthr);
    for (int e = tid; e < R * DMAX; e += nthr) {
        const int r = e / DMAX, k = e - r * DMAX;
        sm.xs[e] = (r < nvalid && k < d) ? pts[(p0 + r) * d + k] * __ldg(g.inv_ls + k) : 0.0;
    }
    consumer_sync(nthr);
    // ---- kernel rows: lane = training point (its scaled coordinates loaded once, reused for the warp's 4 query
    //      rows = 4 independent sqrt/exp chains), conflict-free row-major stores
    {
        double ils[DMAX], macc[NT];
#pragma unroll
        for (int rr = 0; rr < NT; ++rr) macc[rr] = 0.0;
#pragma unroll
        for (int k = 0; k < DMAX; ++k) ils[k] = k < d ? __ldg(g.inv_ls + k) : 0.0;
        for (int j = lane; j < KS; j += 32) {
            double tj[DMAX], al = 0.0;
            const bool live = j < n;
#pragma unroll
            for (int k = 0; k < DMAX; ++k) tj[k] = (live && k < d) ? __ldg(g.train_x + (size_t)j * d + k) * ils[k] : 0.0;
            if (live) al = __ldg(g.alpha + j);
#pragma unroll
            for (int rr = 0; rr < NT; ++rr) {
                const int r = warp + rr * ncw;
                double r2 = 0.0;
#pragma unroll
                for (int k = 0; k < DMAX; ++k) {
                    const double df = sm.xs[r * DMAX + k] - tj[k];
                    r2 = fma(df, df, r2);
                }
                double kap, gc;
                kernel_profile_tab(g.kernel_id, r2, sm.tab, kap, gc);
                const double kv = (live && r < nvalid) ? g.amp * kap : 0.0;
                macc[rr] = fma(kv, al, macc[rr]);
                sm.A[(size_t)r * KS + j] = kv;
            }
        }
#pragma unroll
        for (int rr = 0; rr < NT; ++rr) {
            const int r = warp + rr * ncw;
            const double s = warp_sum(macc[rr]);
            if (lane == 0 && r < nvalid) m0[p0 + r] = s;
        }
    }
    consumer_sync(nthr);
    // ---- U tile = K L^-T, 64 output columns at a time; only k < cb + 64 contributes
    MmaWarp w;
    w.init(warp, nthr, n, NT);
    const int row = w.row0 + w.fr;
    double qacc = 0.0;
    for (int cb = 0; cb < n; cb += kOB) {
        double acc[NT][2];
        zero_acc4(acc);
        w.active = cb + w.col0 < n;
        mma_gemm<NST, NT>(ring, w, sm.A, min(cb + kOB, ldl), acc);
#pragma unroll
        for (int t = 0; t < NT; ++t) {
            const int c = cb + w.col0 + 8 * t + 2 * w.fk;
            if (c < n) {
                const bool two = c + 1 < n;
                qacc = fma(acc[t][0], acc[t][0], qacc);
                if (two) qacc = fma(acc[t][1], acc[t][1], qacc);
                if (row < nvalid) {
                    double* up = U + (size_t)(p0 + row) * g.ldu + c;
                    if (two) *reinterpret_cast<double2*>(up) = make_double2(acc[t][0], acc[t][1]);
                    else *up = acc[t][0];
                }
            }
        }
    }
    qacc += shfl_xor_(qacc, 1);
    qacc += shfl_xor_(qacc, 2);
    constexpr int wpr = 8 / NT;                                   // column warps that share a row block
    if (w.fk == 0) sm.red[(warp % wpr) * R + row] = qacc;
    consumer_sync(nthr);
    if (tid < nvalid) {
        double qs = 0.0;
#pragma unroll
        for (int c = 0; c < wpr; ++c) qs += sm.red[c * R + tid];
        q[p0 + tid] = qs;
    }
}

// ------------------------------------------------------------------ backward
template <int DMAX, int NST, int NT = kWhNT>
__global__ void __launch_bounds__(32 * (32 / 8) * (8 / NT) + 32, 1) whiten_bwd_mma_kernel(GpDev<double> g, int ldl, int R,
                                                                const double* __restrict__ pts, long long n_pts,
                                                                const double* __restrict__ gU,
                                                                const double* __restrict__ g_m0,
                                                                double* __restrict__ g_pts) {
    extern __shared__ __align__(16) unsigned char wh_smem_raw[];
    const int n = g.n, d = g.d, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int ncw = (R >> 3) * (8 / NT), nthr = ncw * 32;
    WhSmem sm;
    sm.carve<DMAX>(reinterpret_cast<double*>(wh_smem_raw), ldl, R, NST, true);
    wh_init_barriers(sm, NST, ncw);
    const MmaRing ring{sm.stage, sm.full0, sm.empty0, sm.KS};
    if (warp == ncw) {
        wh_producer<NST, true>(g.linv_tiles, n, ldl, ring);
        return;
    }
    const int KS = sm.KS;
    const long long p0 = (long long)blockIdx.x * R;
    const int nvalid = (int)min((long long)R, n_pts - p0);
    sm.load_tab(tid, nthr);
    for (int e = tid; e < R * DMAX; e += nthr) {
        const int r = e / DMAX, k = e - r * DMAX;
        sm.xs[e] = (r < nvalid && k < d) ? pts[(p0 + r) * d + k] * __ldg(g.inv_ls + k) : 0.0;
    }
    // gU tile: rows are contiguous in HBM and in the row-major A tile
    for (int r = warp; r < R; r += ncw) {
        const double* src = gU + (size_t)(p0 + r) * g.ldu;
        for (int i = lane; i < KS; i += 32) sm.A[(size_t)r * KS + i] = (r < nvalid && i < n) ? src[i] : 0.0;
    }
    consumer_sync(nthr);
    MmaWarp w;
    w.init(warp, nthr, n, NT);
    const int row = w.row0 + w.fr;
    const double gm = row < nvalid ? g_m0[p0 + row] : 0.0;
    double xr[DMAX], gx[DMAX];
#pragma unroll
    for (int k = 0; k < DMAX; ++k) { xr[k] = sm.xs[row * DMAX + k]; gx[k] = 0.0; }
    double ils[DMAX];
#pragma unroll
    for (int k = 0; k < DMAX; ++k) ils[k] = k < d ? __ldg(g.inv_ls + k) : 0.0;
    // this warp's 8 NT training points of the column block and their alpha: [8 NT][DMAX + 1] then [8 NT]
    constexpr int WC = 8 * NT;
    double* tp = sm.tp + (size_t)warp * (WC * (DMAX + 2));
    for (int cb = 0; cb < n; cb += kOB) {
        double acc[NT][2];
        zero_acc4(acc);
        w.active = cb + w.col0 < n;
        // this warp's 32 training points of the block (one per lane), fetched under the GEMM and parked in shared
        // memory for the epilogue (per-column global loads inside the epilogue were 19 % of the stall samples)
        const int cpt = cb + w.col0 + lane;
        const bool mine = lane < WC && cpt < n;
        double tjr[DMAX], alr = 0.0;
#pragma unroll
        for (int k = 0; k < DMAX; ++k) tjr[k] = (mine && k < d) ? __ldg(g.train_x + (size_t)cpt * d + k) * ils[k] : 0.0;
        if (mine) alr = __ldg(g.alpha + cpt);
        // g_K[r][j] = sum_{i >= j} gU[r][i] Linv[i][j]: rows i >= cb of Linv, columns cb..cb+63
        mma_gemm<NST, NT>(ring, w, sm.A + cb, ldl - cb, acc);
        __syncwarp();
        if (lane < WC) {
#pragma unroll
            for (int k = 0; k < DMAX; ++k) tp[lane * (DMAX + 1) + k] = tjr[k];
            tp[WC * (DMAX + 1) + lane] = alr;
        }
        __syncwarp();
#pragma unroll
        for (int t = 0; t < NT; ++t) {
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int cl = 8 * t + 2 * w.fk + e;
                if (cb + w.col0 + cl < n) {
                    const double al = tp[WC * (DMAX + 1) + cl];
                    double sd[DMAX], r2 = 0.0;
#pragma unroll
                    for (int k = 0; k < DMAX; ++k) {
                        sd[k] = xr[k] - tp[cl * (DMAX + 1) + k];
                        r2 = fma(sd[k], sd[k], r2);
                    }
                    double kap, gc;
                    kernel_profile_tab(g.kernel_id, r2, sm.tab, kap, gc);
                    const double coef = (acc[t][e] + gm * al) * g.amp * gc;
#pragma unroll
                    for (int k = 0; k < DMAX; ++k) gx[k] = fma(coef, sd[k], gx[k]);
                }
            }
        }
    }
    // reduce over the 4 lanes of a fragment row and the 2 column warps that share the rows
#pragma unroll
    for (int k = 0; k < DMAX; ++k) { gx[k] += shfl_xor_(gx[k], 1); gx[k] += shfl_xor_(gx[k], 2); }
    // (the gU tile is dead once every warp has left its last GEMM: it holds the 8 / NT partial sums per row)
    constexpr int wpr = 8 / NT;
    consumer_sync(nthr);
    double* red = sm.A;
    if (w.fk == 0) {
#pragma unroll
        for (int k = 0; k < DMAX; ++k) red[((warp % wpr) * R + row) * DMAX + k] = gx[k];
    }
    consumer_sync(nthr);
    for (int e = tid; e < R * DMAX; e += nthr) {
        const int r = e / DMAX, k = e - r * DMAX;
        if (r < nvalid && k < d) {
            double v = 0.0;
#pragma unroll
            for (int c = 0; c < wpr; ++c) v += red[c * R * DMAX + e];
            g_pts[(p0 + r) * d + k] += v * __ldg(g.inv_ls + k);
        }
    }
}

}  // namespace nmb
