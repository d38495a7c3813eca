// FP32 / TF32 rollout kernels (included by rollout.cu): the float twins of ts_rollout_fwd_mma_kernel /
// ts_rollout_bwd_mma_kernel on mma.sync.m16n8k8.tf32 (policy_tf32.cuh).  One row group per CTA, rows a multiple of 16,
// 8 consumer threads per chain (16 x 16 warp tiles), which share the chain in the Thompson-path phase.
#pragma once
#include "policy_tf32.cuh"

namespace nmb {

template <int DMAX>
__global__ void __launch_bounds__(kMaxRowsTf32 * 8 + 32, 1)
    ts_rollout_fwd_tf32_kernel(ProblemDev<float> pb, RolloutDev<float> ro0, RolloutDev<float> ro1, int rows) {
    using T = float;
    if (pb.done_flag && *pb.done_flag) return;
    const RolloutDev<T>& ro = blockIdx.y == 0 ? ro0 : ro1;
    const int d = pb.d, H = pb.H, L = pb.L;
    const long long N = pb.N;
    const int ncw = (rows >> 4) * 4, nthr = ncw * 32;
    const long long step = pb.step_ptr ? *pb.step_ptr : 0;
    const int tid = threadIdx.x;
    const long long chain0 = (long long)blockIdx.x * rows;
    const int nvalid = (int)max(0LL, min((long long)rows, N - chain0));

    Tf32Smem sm;
    SmallParF sp;
    sm.carve(reinterpret_cast<T*>(nmb_smem_raw), d, H, rows, SmallPar::elems(pb.lay));
    sp.init(pb.lay, sm.par);
    if (tid == 0) {
        for (int s = 0; s < kStagesF32; ++s) {
            mbar_init(sm.full0 + 8 * s, 1);
            mbar_init(sm.empty0 + 8 * s, ncw);
        }
        fence_barrier_init();
    }
    sp.load(pb.lay, ro.theta_t, sm.par, tid, (int)blockDim.x);
    __syncthreads();
    if (tid >= nthr) {
        tf32_producer<false>(pb.lay, ro.theta_t, L, sm.ring());
        return;
    }
    Tf32Warp w;
    w.init(tid >> 5, nthr, H);
    const int KS = sm.KS;
    int shiftX[kGemmsPerStep], shiftA[kGemmsPerStep];
#pragma unroll
    for (int g = 0; g < kGemmsPerStep; ++g) {
        int off, ldw;
        gemm_src(pb.lay, false, g, off, ldw);
        shiftX[g] = off & 3;
        gemm_src(pb.lay, true, g, off, ldw);
        shiftA[g] = off & 3;
    }

    for (int e = tid; e < (d + 1) * rows; e += nthr) {
        const int k = e / rows, r = e - k * rows;
        T v = T(0);
        if (r < nvalid) v = k < d ? pb.prev_x[(chain0 + r) * d + k] : pb.prev_y[chain0 + r];
        sm.IN[k * rows + r] = v;
    }
    for (int e = tid; e < H * rows; e += nthr) {
        const int r = e / H, k = e - r * H;
        const T v = r < nvalid ? pb.prev_h[(chain0 + r) * H + k] : T(0);
        sm.HA[r * KS + k] = v;
        if (ro.has_save && r < nvalid) ro.sv.HS[(chain0 + r) * H + k] = v;
    }
    if (tid < rows) sm.cost_acc[tid] = T(0);
    consumer_sync(nthr);

    const SaveAreas<T>* sv = ro.has_save ? &ro.sv : nullptr;
    for (int l = 0; l <= L; ++l) {
        const bool head_a = (l == L);
        const long long node0 = (long long)l * N + chain0;
        T* hs_out = ro.has_save ? ro.sv.HS + ((size_t)(l + 1) * N + chain0) * H : nullptr;
        T* hprev = (l & 1) ? sm.HB : sm.HA;
        T* hnew = (l & 1) ? sm.HA : sm.HB;
        tf32_policy_forward(pb.lay, head_a, sm, w, hprev, hnew, sv, node0, nvalid, hs_out, sp, head_a ? shiftA : shiftX);

        // ---- output point, Thompson path, cost: 8 threads per chain
        const int r = tid >> 3, lane = tid & 7;
        const long long chain = chain0 + r;
        const bool live = r < nvalid;
        T xin[DMAX], pt[DMAX];
#pragma unroll
        for (int k = 0; k < DMAX; ++k) {
            xin[k] = k < d ? sm.IN[k * rows + r] : T(0);
            pt[k] = k < d ? sm.OUT[k * rows + r] : T(0);
        }
        const bool noisy = head_a ? ro.noisy_action : ro.noisy_design;
        if (noisy && live) {
            T q[4]; int di, dj;
            get_rotation<T>(ro, pb, step, l, chain, q, di, dj);
            T vi = T(0), vj = T(0);
#pragma unroll
            for (int k = 0; k < DMAX; ++k) { if (k == di) vi = pt[k] - xin[k]; if (k == dj) vj = pt[k] - xin[k]; }
            const T ri = q[0] * vi + q[1] * vj, rj = q[2] * vi + q[3] * vj;
#pragma unroll
            for (int k = 0; k < DMAX; ++k) {
                T v = pt[k] - xin[k];
                if (k == di) v = ri;
                if (k == dj) v = rj;
                if (k < d) pt[k] = v + xin[k];
            }
        }
        T yp = T(0), gp[DMAX];
#pragma unroll
        for (int k = 0; k < DMAX; ++k) gp[k] = T(0);
        if (live) {
            if (ro.has_save) path_partial<T, DMAX, true, true>(pb.path, pt, (int)(chain / pb.cpp), lane, 8, yp, gp);
            else path_partial<T, DMAX, false, true>(pb.path, pt, (int)(chain / pb.cpp), lane, 8, yp, gp);
        }
        yp = group_sum(yp, 8);
        if (ro.has_save) {
#pragma unroll
            for (int k = 0; k < DMAX; ++k) gp[k] = group_sum(gp[k], 8);
        }
        const T yv = yp + pb.path.mean_const;
        consumer_sync(nthr);
        if (live && lane == 0) {
            T dir[DMAX];
            const T dist = move_dist<T, DMAX>(pb.cost, d, pt, xin, dir);
            const T nz = get_cost_noise<T>(ro, pb, step, l, chain);
            T h = pb.cost.k * (dist - pb.cost.radius);
            h = (h > T(0) ? h : T(0)) + nz * pb.cost.max_noise;
            T mult = T(1);
            if (pb.cost.discount > T(0)) {
                const T acc = sm.cost_acc[r] + pb.cost.prev_cost;
                if (acc + h > pb.cost.discount_threshold) mult = pb.cost.disc_mult;
                h *= mult;
            }
            sm.cost_acc[r] += h;
            if (ro.has_save) {
                ro.sv.CM[node0 + r] = mult;
#pragma unroll
                for (int k = 0; k < DMAX; ++k) if (k < d) ro.sv.DYDX[(node0 + r) * d + k] = gp[k];
            }
            T* dst = head_a ? ro.A + chain * d : ro.X + ((size_t)l * N + chain) * d;
#pragma unroll
            for (int k = 0; k < DMAX; ++k) if (k < d) dst[k] = pt[k];
            if (ro.Y) ro.Y[(size_t)l * N + chain] = yv;
            if (head_a) {
                ro.loss[chain] = -yv;
                ro.cost[chain] = sm.cost_acc[r];
            }
#pragma unroll
            for (int k = 0; k < DMAX; ++k) if (k < d) sm.IN[k * rows + r] = pt[k];
            sm.IN[d * rows + r] = yv;
        }
        consumer_sync(nthr);
    }
}

// ================================================================== backward
template <int DMAX>
__global__ void __launch_bounds__(kMaxRowsTf32 * 8 + 32, 1)
    ts_rollout_bwd_tf32_kernel(ProblemDev<float> pb, RolloutDev<float> ro, const float* __restrict__ g_loss,
                               const float* __restrict__ g_cost, const float* __restrict__ g_X,
                               const float* __restrict__ g_A, float* __restrict__ g_prev, int rows) {
    using T = float;
    if (pb.done_flag && *pb.done_flag) return;
    const int d = pb.d, H = pb.H, L = pb.L;
    const long long N = pb.N;
    const long long chain0 = (long long)blockIdx.x * rows;
    const int nvalid = (int)min((long long)rows, N - chain0);
    const int tid = threadIdx.x;
    const int ncw = (rows >> 4) * 4, nthr = ncw * 32;
    const long long step = pb.step_ptr ? *pb.step_ptr : 0;

    Tf32BwdSmem sm;
    sm.carve(reinterpret_cast<T*>(nmb_smem_raw), d, DMAX, H, rows);
    if (tid == 0) {
        for (int s = 0; s < kStagesF32; ++s) {
            mbar_init(sm.full0 + 8 * s, 1);
            mbar_init(sm.empty0 + 8 * s, ncw);
        }
        fence_barrier_init();
    }
    __syncthreads();
    if (tid >= nthr) {
        tf32_producer<true>(pb.lay, ro.theta, L, sm.ring());
        return;
    }
    Tf32Warp w;
    w.init(tid >> 5, nthr, H);
    int shiftX[kGemmsPerStep], shiftA[kGemmsPerStep];
#pragma unroll
    for (int g = 0; g < kGemmsPerStep; ++g) {
        int off, ldw;
        gemm_src_bwd(pb.lay, false, g, off, ldw);
        shiftX[g] = off & 3;
        gemm_src_bwd(pb.lay, true, g, off, ldw);
        shiftA[g] = off & 3;
    }
    const int KS = sm.KS;
    T* carry = sm.carry;      // [(DMAX+1)][rows]: adjoint of (x_{l+1}, y_{l+1}) from the level above

    for (int e = tid; e < rows * KS; e += nthr) sm.GH[e] = T(0);
    for (int e = tid; e < (DMAX + 1) * rows; e += nthr) carry[e] = T(0);
    consumer_sync(nthr);

    const SaveAreas<T>& sv = ro.sv;
    for (int l = L; l >= 0; --l) {
        const bool head_a = (l == L);
        const long long node0 = (long long)l * N + chain0;
        if (tid < rows) {
            const int r = tid;
            if (r < nvalid) {
                const long long chain = chain0 + r;
                const size_t node = (size_t)(node0 + r);
                T pt[DMAX], xin[DMAX], gout[DMAX];
                const T* psrc = head_a ? ro.A + chain * d : ro.X + ((size_t)l * N + chain) * d;
#pragma unroll
                for (int k = 0; k < DMAX; ++k) {
                    pt[k] = k < d ? psrc[k] : T(0);
                    xin[k] = k < d ? sv.XY[node * (d + 1) + k] : T(0);
                }
                const T gy = head_a ? -g_loss[chain] : carry[DMAX * rows + r];
#pragma unroll
                for (int k = 0; k < DMAX; ++k) {
                    T g = T(0);
                    if (k < d) {
                        g = carry[k * rows + r] + gy * sv.DYDX[node * d + k];
                        if (head_a) { if (g_A) g += g_A[chain * d + k]; }
                        else if (g_X) g += g_X[((size_t)l * N + chain) * d + k];
                    }
                    gout[k] = g;
                }
                T dir[DMAX];
                const T dist = move_dist<T, DMAX>(pb.cost, d, pt, xin, dir);
                const T hinge = pb.cost.k * (dist - pb.cost.radius);
                const T gc = hinge > T(0) ? g_cost[chain] * pb.cost.k * sv.CM[node] : T(0);
                T gprev[DMAX];
#pragma unroll
                for (int k = 0; k < DMAX; ++k) { gout[k] += gc * dir[k]; gprev[k] = -gc * dir[k]; }
                const bool noisy = head_a ? ro.noisy_action : ro.noisy_design;
                T go[DMAX];
#pragma unroll
                for (int k = 0; k < DMAX; ++k) go[k] = gout[k];
                if (noisy) {
                    T q[4]; int di, dj;
                    get_rotation<T>(ro, pb, step, l, chain, q, di, dj);
                    T gi = T(0), gj = T(0);
#pragma unroll
                    for (int k = 0; k < DMAX; ++k) { if (k == di) gi = gout[k]; if (k == dj) gj = gout[k]; }
                    const T ti = q[0] * gi + q[2] * gj, tj = q[1] * gi + q[3] * gj;
#pragma unroll
                    for (int k = 0; k < DMAX; ++k) {
                        if (k == di) go[k] = ti;
                        if (k == dj) go[k] = tj;
                        gprev[k] += gout[k] - go[k];
                    }
                }
#pragma unroll
                for (int k = 0; k < DMAX; ++k) {
                    if (k < d) sm.GO[k * rows + r] = go[k];
                    carry[k * rows + r] = gprev[k];
                }
            } else {
                for (int k = 0; k < d; ++k) sm.GO[k * rows + r] = T(0);
            }
        }
        consumer_sync(nthr);
        tf32_policy_backward(pb.lay, ro.theta, head_a, sm, w, sv, node0, nvalid, head_a ? shiftA : shiftX);
        if (tid < rows) {
            const int r = tid;
#pragma unroll
            for (int k = 0; k < DMAX; ++k) if (k < d) carry[k * rows + r] += sm.GXY[k * rows + r];
            carry[DMAX * rows + r] = sm.GXY[d * rows + r];
        }
        consumer_sync(nthr);
    }
    if (g_prev) {
        const int wd = d + 1 + H;
        for (int e = tid; e < wd * rows; e += nthr) {
            const int r = e / wd, k = e - r * wd;
            if (r < nvalid) {
                T v;
                if (k < d) v = carry[k * rows + r];
                else if (k == d) v = carry[DMAX * rows + r];
                else v = sm.GH[r * KS + (k - d - 1)];
                g_prev[(chain0 + r) * wd + k] = v;
            }
        }
    }
}


// float rows are TMA-able when every segment offset is even (d even) and H is a multiple of 8 (k-steps of 8)
static bool tf32_fwd_ok(const nmb_ts_problem_t* p, const nmb_ts_rollout_t* r, int n) {
    if (p->dtype != NMB_F32 || (p->d & 1) || p->H < 16 || (p->H & 7) || getenv("NMB_FP32_NO_TF32")) return false;
    for (int i = 0; i < n; ++i)
        if (((uintptr_t)r[i].theta_t & 15) != 0 || r[i].y_obs) return false;
    return true;
}

template <int DMAX>
static int launch_fwd_tf32(const nmb_ts_problem_t* p, const nmb_ts_rollout_t* r, int n, cudaStream_t st) {
    ProblemDev<float> pb = make_problem<float>(*p);
    RolloutDev<float> r0 = make_rollout<float>(*p, r[0]);
    RolloutDev<float> r1 = n > 1 ? make_rollout<float>(*p, r[1]) : r0;
    const int rows = tf32_rows_per_cta(p->n_chains, n, sm_count(), [&](int r) { return Tf32Smem::bytes(p->d, p->H, r); });
    const size_t smem = Tf32Smem::bytes(p->d, p->H, rows);
    auto kern = ts_rollout_fwd_tf32_kernel<DMAX>;
    NMB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid((unsigned)((p->n_chains + rows - 1) / rows), (unsigned)n);
    kern<<<grid, rows * 8 + 32, smem, st>>>(pb, r0, r1, rows);
    NMB_LAUNCH_CHECK();
    return 0;
}


static bool tf32_bwd_ok(const nmb_ts_problem_t* p, const nmb_ts_rollout_t* r) {
    return p->dtype == NMB_F32 && !(p->d & 1) && p->H >= 16 && !(p->H & 7) && ((uintptr_t)r->theta & 15) == 0 &&
           !getenv("NMB_FP32_NO_TF32");
}

template <int DMAX>
static int launch_bwd_tf32(const nmb_ts_problem_t* p, const nmb_ts_rollout_t* r, const void* gl, const void* gc,
                           const void* gX, const void* gA, void* gprev, cudaStream_t st) {
    ProblemDev<float> pb = make_problem<float>(*p);
    RolloutDev<float> r0 = make_rollout<float>(*p, *r);
    const int rows = tf32_rows_per_cta(p->n_chains, 1, sm_count(), [&](int r) { return Tf32BwdSmem::bytes(p->d, DMAX, p->H, r); });
    const size_t smem = Tf32BwdSmem::bytes(p->d, DMAX, p->H, rows);
    auto kern = ts_rollout_bwd_tf32_kernel<DMAX>;
    NMB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid((unsigned)((p->n_chains + rows - 1) / rows));
    kern<<<grid, rows * 8 + 32, smem, st>>>(pb, r0, (const float*)gl, (const float*)gc, (const float*)gX, (const float*)gA,
                                            (float*)gprev, rows);
    NMB_LAUNCH_CHECK();
    return 0;
}

}  // namespace nmb
