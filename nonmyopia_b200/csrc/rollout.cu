// Fused HES-TS-AM rollout, forward and backward (reference acqfs.py:277-456 with an
// amortized map and enable_ts; SURVEY App. C "Fused TS-AM rollout dataflow").
//
// One CTA (256 threads) owns a tile of 32 chains and walks all L+1 policy steps:
//   policy step (policy_dev.cuh)  ->  [planar rotation about the input point]
//   -> Thompson-path value + gradient (path.cuh, 8 threads per chain)
//   -> spotlight cost of the move  -> next step's input.
// The backward kernel walks the same chain in reverse (BPTT through the GRU and through
// f's input gradient) and leaves every layer's pre-activation adjoint in the save
// workspace for the split-K weight-gradient kernel (wgrad.cu).
#include <stdlib.h>
#include "cost.cuh"
#include "path.cuh"
#include "policy_dev.cuh"
#include "policy_mma.cuh"

namespace nmb {

// ------------------------------------------------------------------ counter RNG (Philox4x32-10)
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

template <typename T> struct RolloutDev {
    const T* theta;
    const T* theta_t;
    int noisy_design, noisy_action;
    const T* rot_q;
    const int* rot_dims;
    const T* cost_noise;
    unsigned rng_stream;
    T *X, *A, *Y, *loss, *cost;
    SaveAreas<T> sv;
    int has_save;
    const float2* ring_cs;
    const int* ring_dims;
    const T* ring_noise;
    int ring_len;
    // imitation mode (Actor.reset_parameters, reference actor.py:164-220): observations come from outside
    const T* y_obs;
    int level_begin, level_end;
};

template <typename T> struct ProblemDev {
    int d, H, L;
    long long N, cpp;
    PathDev<T> path;
    CostDev<T> cost;
    PolicyLayout lay;
    const T *prev_x, *prev_y, *prev_h;
    unsigned long long seed;
    const long long* step_ptr;
    const int* done_flag;
    long long chain_offset;
};

// rotation block + coordinate pair of (call, chain): given tensors or in-kernel draws that
// mimic the reference's float32 angle / float32 cos, sin (amortized_network.py:136,144-153)
template <typename T>
__device__ __forceinline__ void get_rotation(const RolloutDev<T>& ro, const ProblemDev<T>& pb, long long step, int call,
                                             long long chain, T (&q)[4], int& di, int& dj) {
    if (ro.ring_len > 0) {
        // host-stream ring: slot of this Adam iteration, float32 (cos, sin) as the reference rounds them
        const size_t idx = ((size_t)(step % ro.ring_len) * (pb.L + 1) + call) * pb.N + chain;
        const float2 cs = ro.ring_cs[idx];
        q[0] = (T)cs.x; q[1] = (T)(-cs.y); q[2] = (T)cs.y; q[3] = (T)cs.x;
        const int2 dd = reinterpret_cast<const int2*>(ro.ring_dims)[idx];
        di = dd.x; dj = dd.y;
    } else if (ro.rot_q) {
        const T* p = ro.rot_q + ((size_t)call * pb.N + chain) * 4;
        q[0] = p[0]; q[1] = p[1]; q[2] = p[2]; q[3] = p[3];
        const int* dd = ro.rot_dims + ((size_t)call * pb.N + chain) * 2;
        di = dd[0]; dj = dd[1];
    } else {
        const long long gc = chain + pb.chain_offset;       // global chain index: shard-invariant stream
        const uint4 r = philox4x32_10(make_uint4((unsigned)gc, (unsigned)(gc >> 32), (unsigned)step,
                                                 (ro.rng_stream << 8) | (unsigned)call),
                                      make_uint2((unsigned)pb.seed, (unsigned)(pb.seed >> 32)));
        const float u = (float)(r.x >> 8) * (1.0f / 16777216.0f);
        const float th = u * 2.0f * 3.14159265358979323846f - 3.14159265358979323846f;
        const float c = cosf(th), s = sinf(th);
        q[0] = (T)c; q[1] = (T)(-s); q[2] = (T)s; q[3] = (T)c;
        di = (int)(r.y % (unsigned)pb.d);
        dj = pb.d > 1 ? (di + 1 + (int)(r.z % (unsigned)(pb.d - 1))) % pb.d : di;
    }
}

template <typename T>
__device__ __forceinline__ T get_cost_noise(const RolloutDev<T>& ro, const ProblemDev<T>& pb, long long step, int call,
                                            long long chain) {
    if (ro.ring_len > 0)
        return ro.ring_noise[((size_t)(step % ro.ring_len) * (pb.L + 1) + call) * pb.N + chain];
    if (ro.cost_noise) return ro.cost_noise[(size_t)call * pb.N + chain];
    const long long gc = chain + pb.chain_offset;
    const uint4 r = philox4x32_10(make_uint4((unsigned)gc, (unsigned)(gc >> 32), (unsigned)step,
                                             0x80000000u | (ro.rng_stream << 8) | (unsigned)call),
                                  make_uint2((unsigned)pb.seed, (unsigned)(pb.seed >> 32)));
    const double u1 = ((double)r.x * 4294967296.0 + (double)r.y + 1.0) * (1.0 / 18446744073709551616.0);
    const double u2 = ((double)r.z * 4294967296.0 + (double)r.w) * (1.0 / 18446744073709551616.0);
    double sn, cs;
    ::sincospi(2.0 * u2, &sn, &cs);
    return (T)(::sqrt(-2.0 * ::log(u1)) * cs);
}

extern __shared__ __align__(16) unsigned char nmb_smem_raw[];

// ================================================================== forward
template <typename T, int DMAX>
__global__ void __launch_bounds__(kNT, 1) ts_rollout_fwd_kernel(ProblemDev<T> pb, RolloutDev<T> ro0, RolloutDev<T> ro1) {
    if (pb.done_flag && *pb.done_flag) return;
    const RolloutDev<T>& ro = blockIdx.y == 0 ? ro0 : ro1;
    const int d = pb.d, H = pb.H, L = pb.L;
    const long long N = pb.N;
    const long long chain0 = (long long)blockIdx.x * kR;
    const int nvalid = (int)min((long long)kR, N - chain0);
    const int tid = threadIdx.x;
    const long long step = pb.step_ptr ? *pb.step_ptr : 0;

    T* smem = reinterpret_cast<T*>(nmb_smem_raw);
    FwdSmem<T> sm;
    sm.carve(smem, d, H);
    T* cost_acc = sm.stage + 2 * kKC * kOB;     // [kR] running cost per chain

    // ---- load the tile's inputs
    for (int e = tid; e < (d + 1) * kR; e += kNT) {
        const int k = e >> 5, r = e & 31;
        T v = T(0);
        if (r < nvalid) v = k < d ? pb.prev_x[(chain0 + r) * d + k] : pb.prev_y[chain0 + r];
        sm.IN[k * kRS + r] = v;
    }
    for (int e = tid; e < H * kR; e += kNT) {
        const int k = e >> 5, r = e & 31;
        const T v = r < nvalid ? pb.prev_h[(chain0 + r) * H + k] : T(0);
        sm.HB[k * kRS + r] = v;
        if (ro.has_save && r < nvalid) ro.sv.HS[(chain0 + r) * H + k] = v;
    }
    if (tid < kR) cost_acc[tid] = T(0);
    __syncthreads();

    const SaveAreas<T>* sv = ro.has_save ? &ro.sv : nullptr;
    for (int l = 0; l <= L; ++l) {
        const bool head_a = (l == L);
        const long long node0 = (long long)l * N + chain0;
        T* hs_out = ro.has_save ? ro.sv.HS + ((size_t)(l + 1) * N + chain0) * H : nullptr;
        policy_forward_tile<T>(pb.lay, ro.theta_t, head_a, sm, sv, node0, nvalid, hs_out);

        // ---- output point, Thompson path, cost: 8 threads per chain
        const int r = tid >> 3, lane = tid & 7;
        const long long chain = chain0 + r;
        const bool live = r < nvalid;
        T xin[DMAX], pt[DMAX];
#pragma unroll
        for (int k = 0; k < DMAX; ++k) {
            xin[k] = k < d ? sm.IN[k * kRS + r] : T(0);
            pt[k] = k < d ? sm.OUT[k * kRS + r] : T(0);
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
                if (k == dj) v = rj;     // randperm gives di != dj for d >= 2
                if (k < d) pt[k] = v + xin[k];
            }
        }
        T yp = T(0), gp[DMAX];
#pragma unroll
        for (int k = 0; k < DMAX; ++k) gp[k] = T(0);
        if (live) path_partial<T, DMAX>(pb.path, pt, (int)(chain / pb.cpp), lane, 8, yp, gp);
        yp = group_sum(yp, 8);
#pragma unroll
        for (int k = 0; k < DMAX; ++k) gp[k] = group_sum(gp[k], 8);
        const T yv = yp + pb.path.mean_const;
        __syncthreads();     // every thread has read sm.IN / sm.OUT of this step
        if (live && lane == 0) {
            // spotlight cost of the move x_l -> point (acqfs.py:426-441)
            T dir[DMAX];
            const T dist = move_dist<T, DMAX>(pb.cost, d, pt, xin, dir);
            const T nz = get_cost_noise<T>(ro, pb, step, l, chain);
            T h = pb.cost.k * (dist - pb.cost.radius);
            h = (h > T(0) ? h : T(0)) + nz * pb.cost.max_noise;
            T mult = T(1);
            if (pb.cost.discount > T(0)) {
                const T acc = cost_acc[r] + pb.cost.prev_cost;
                if (acc + h > pb.cost.discount_threshold) mult = pb.cost.disc_mult;
                h *= mult;
            }
            cost_acc[r] += h;
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
                ro.cost[chain] = cost_acc[r];
            }
            // next step's input
#pragma unroll
            for (int k = 0; k < DMAX; ++k) if (k < d) sm.IN[k * kRS + r] = pt[k];
            sm.IN[d * kRS + r] = yv;
        }
        __syncthreads();
    }
}

// ================================================================== forward, FP64 tensor pipe + TMA weight stream
// rows (multiple of 8, <= 56) chains per CTA; rows/4 consumer warps + 1 producer warp (policy_mma.cuh).
// NST: depth of the weight ring.  kStages for full CTAs; with <= 16 rows per CTA (few chains: 64 restarts on the whole
// GPU) the activation buffers are small and the kernel is bound by the LATENCY of the weight stream (phase clocks at 64
// chains: 2.7 k cycles per GEMM against ~0.6 k of tensor work), so the ring takes the freed shared memory.
template <int DMAX, int NST = kStages>
__global__ void __launch_bounds__(kMaxRowsMma * 8 + 32, 1)
    ts_rollout_fwd_mma_kernel(ProblemDev<double> pb, RolloutDev<double> ro0, RolloutDev<double> ro1, int rows_cta) {
    using T = double;
    if (pb.done_flag && *pb.done_flag) return;
    const RolloutDev<T>& ro = blockIdx.y == 0 ? ro0 : ro1;
    const int d = pb.d, H = pb.H, L = pb.L;
    const long long N = pb.N;
    const int ncw_all = rows_cta >> 2, nthr_all = ncw_all * 32;
    const long long step = pb.step_ptr ? *pb.step_ptr : 0;
    // Two independent row groups per CTA (e.g. 56 rows = 32 + 24) walk the same layer sequence on their own named
    // barrier and shared-memory buffers but consume ONE weight ring: they drift apart by up to the ring depth, so one
    // group's GEMM fills the FP64 pipe while the other sits in an activation epilogue, a barrier or the path phase.
    const int rowsA = ((rows_cta >> 3) + 1) / 2 * 8;
    const int grp = (int)threadIdx.x >= rowsA * 8 ? 1 : 0;
    const int rows = grp ? rows_cta - rowsA : rowsA;               // rows, threads, first chain of THIS group
    const int tid = (int)threadIdx.x - (grp ? rowsA * 8 : 0);
    const int nthr = rows * 8;
    const long long chain0 = (long long)blockIdx.x * rows_cta + (grp ? rowsA : 0);
    const int nvalid = (int)max(0LL, min((long long)rows, N - chain0));

    MmaSmem sm;
    SmallPar sp;
    sm.carve_group(reinterpret_cast<T*>(nmb_smem_raw), d, H, rows_cta, grp ? rowsA : 0, rows, SmallPar::elems(pb.lay), NST);
    sp.init(pb.lay, sm.par);
    if (threadIdx.x == 0) {
        for (int s = 0; s < NST; ++s) {
            mbar_init(sm.full0 + 8 * s, 1);
            mbar_init(sm.empty0 + 8 * s, ncw_all);
        }
        fence_barrier_init();
    }
    load_fast_tab(sm.tab, (int)threadIdx.x, (int)blockDim.x);
    sp.load(pb.lay, ro.theta_t, sm.par, (int)threadIdx.x, (int)blockDim.x);
    __syncthreads();
    if ((int)threadIdx.x >= nthr_all) {      // producer warp
        mma_producer<NST, false>(pb.lay, ro.theta_t, L, sm.ring(), ro.level_begin, ro.level_end);
        return;
    }
    MmaWarp w;
    w.init(tid >> 5, nthr, H);
    w.tid = tid;
    w.bar = 1 + grp;
    const int KS = sm.KS;
    PhaseClk pc;
    pc.start();
    // imitation mode: levels [l0, l1) of the rollout in this launch; the observation that feeds level l > 0 is
    // y_obs[l - 1] (a draw of the surrogate at the previous output, made between the launches), its input point is
    // the previous output X[l - 1], its hidden state the saved HS[l]
    const bool imit = ro.y_obs != nullptr;
    const int l0 = ro.level_begin, l1 = ro.level_end;

    // ---- load the tile's inputs
    for (int e = tid; e < (d + 1) * rows; e += nthr) {
        const int k = e / rows, r = e - k * rows;
        T v = T(0);
        if (r < nvalid) {
            if (l0 == 0) v = k < d ? pb.prev_x[(chain0 + r) * d + k] : pb.prev_y[chain0 + r];
            else v = k < d ? ro.X[((size_t)(l0 - 1) * N + chain0 + r) * d + k] : ro.y_obs[(size_t)(l0 - 1) * N + chain0 + r];
        }
        sm.IN[k * rows + r] = v;
    }
    T* h_first = (l0 & 1) ? sm.HB : sm.HA;
    for (int e = tid; e < H * rows; e += nthr) {
        const int r = e / H, k = e - r * H;
        T v = T(0);
        if (r < nvalid) v = l0 == 0 ? pb.prev_h[(chain0 + r) * H + k] : ro.sv.HS[((size_t)l0 * N + chain0 + r) * H + k];
        h_first[r * KS + k] = v;
        if (ro.has_save && r < nvalid && l0 == 0) ro.sv.HS[(chain0 + r) * H + k] = v;
    }
    if (tid < rows) sm.cost_acc[tid] = T(0);
    consumer_sync(nthr, w.bar);
    pc.mark(7);

    const SaveAreas<T>* sv = ro.has_save ? &ro.sv : nullptr;
    for (int l = l0; l < l1; ++l) {
        const bool head_a = (l == L);
        const long long node0 = (long long)l * N + chain0;
        T* hs_out = ro.has_save ? ro.sv.HS + ((size_t)(l + 1) * N + chain0) * H : nullptr;
        T* hprev = (l & 1) ? sm.HB : sm.HA;
        T* hnew = (l & 1) ? sm.HA : sm.HB;
        mma_policy_forward<NST>(pb.lay, ro.theta_t, head_a, sm, w, hprev, hnew, sv, node0, nvalid, hs_out, pc, sp);

        // ---- output point, Thompson path, cost: 8 threads per chain (nthr = 8 rows)
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
        if (imit) {
            // the next observation is drawn outside the kernel and carries no gradient: dy/dx = 0, no cost
            consumer_sync(nthr, w.bar);
            if (live && lane == 0) {
                if (ro.has_save) {
                    ro.sv.CM[node0 + r] = T(1);
#pragma unroll
                    for (int k = 0; k < DMAX; ++k) if (k < d) ro.sv.DYDX[(node0 + r) * d + k] = T(0);
                }
                T* dst = head_a ? ro.A + chain * d : ro.X + ((size_t)l * N + chain) * d;
#pragma unroll
                for (int k = 0; k < DMAX; ++k) if (k < d) dst[k] = pt[k];
                if (head_a) { ro.loss[chain] = T(0); ro.cost[chain] = T(0); }
            }
            consumer_sync(nthr, w.bar);
            continue;               // the host launches level by level in this mode (l1 == l0 + 1)
        }
        if (live) {
            if (ro.has_save) path_partial<T, DMAX, true, true>(pb.path, pt, (int)(chain / pb.cpp), lane, 8, yp, gp, sm.tab);
            else path_partial<T, DMAX, false, true>(pb.path, pt, (int)(chain / pb.cpp), lane, 8, yp, gp, sm.tab);
        }
        yp = group_sum(yp, 8);
        if (ro.has_save) {
#pragma unroll
            for (int k = 0; k < DMAX; ++k) gp[k] = group_sum(gp[k], 8);
        }
        const T yv = yp + pb.path.mean_const;
        pc.mark(5);
        consumer_sync(nthr, w.bar);     // every thread has read sm.IN / sm.OUT of this step
        pc.mark(3);
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
        pc.mark(6);
        consumer_sync(nthr, w.bar);
        pc.mark(3);
    }
    pc.flush(threadIdx.x == 0);
}

#ifdef NMB_PHASE_CLOCKS
extern "C" int nmb_debug_phase_clocks(unsigned long long* host_out, int reset) {
    cudaDeviceSynchronize();
    if (host_out) cudaMemcpyFromSymbol(host_out, g_phase_clk, sizeof(unsigned long long) * 16);
    if (reset) { unsigned long long z[16] = {0}; cudaMemcpyToSymbol(g_phase_clk, z, sizeof(z)); }
    return 0;
}
#endif

// ================================================================== backward
template <typename T, int DMAX>
__global__ void __launch_bounds__(kNT, 1) ts_rollout_bwd_kernel(ProblemDev<T> pb, RolloutDev<T> ro, const T* __restrict__ g_loss,
                                                                const T* __restrict__ g_cost, const T* __restrict__ g_X,
                                                                const T* __restrict__ g_A, T* __restrict__ g_prev) {
    if (pb.done_flag && *pb.done_flag) return;
    const int d = pb.d, H = pb.H, L = pb.L;
    const long long N = pb.N;
    const long long chain0 = (long long)blockIdx.x * kR;
    const int nvalid = (int)min((long long)kR, N - chain0);
    const int tid = threadIdx.x;
    const long long step = pb.step_ptr ? *pb.step_ptr : 0;

    T* smem = reinterpret_cast<T*>(nmb_smem_raw);
    BwdSmem<T> sm;
    sm.carve(smem, d, H);
    T* carry = sm.stage + 2 * kKC * kOB;        // [(DMAX+1)][kR]: adjoint of (x_{l+1}, y_{l+1}) from the level above

    for (int e = tid; e < H * kR; e += kNT) sm.GH[e] = T(0);
    for (int e = tid; e < (DMAX + 1) * kR; e += kNT) carry[e] = T(0);
    __syncthreads();

    const SaveAreas<T>& sv = ro.sv;
    for (int l = L; l >= 0; --l) {
        const bool head_a = (l == L);
        const long long node0 = (long long)l * N + chain0;
        // ---- per-chain adjoint of the step's output point -> GO
        if (tid < kR) {
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
                // adjoint of y at this point: -g_loss for the action, the next step's input adjoint otherwise
                const T gy = head_a ? -g_loss[chain] : carry[DMAX * kR + r];
#pragma unroll
                for (int k = 0; k < DMAX; ++k) {
                    T g = T(0);
                    if (k < d) {
                        g = carry[k * kR + r] + gy * sv.DYDX[node * d + k];
                        if (head_a) { if (g_A) g += g_A[chain * d + k]; }
                        else if (g_X) g += g_X[((size_t)l * N + chain) * d + k];
                    }
                    gout[k] = g;
                }
                // cost of the move x_l -> pt
                T dir[DMAX];
                const T dist = move_dist<T, DMAX>(pb.cost, d, pt, xin, dir);
                const T hinge = pb.cost.k * (dist - pb.cost.radius);
                const T gc = hinge > T(0) ? g_cost[chain] * pb.cost.k * sv.CM[node] : T(0);
                T gprev[DMAX];
#pragma unroll
                for (int k = 0; k < DMAX; ++k) { gout[k] += gc * dir[k]; gprev[k] = -gc * dir[k]; }
                // rotation: pt = x + M (o - x)  =>  go = M^T gout,  gx += gout - go
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
                    if (k < d) sm.GO[k * kRS + r] = go[k];
                    carry[k * kR + r] = gprev[k];     // completed below with the network's input adjoint
                }
            } else {
                for (int k = 0; k < d; ++k) sm.GO[k * kRS + r] = T(0);
            }
        }
        __syncthreads();
        policy_backward_tile<T>(pb.lay, ro.theta, head_a, sm, sv, node0, node0, nvalid);
        if (tid < kR) {
            const int r = tid;
#pragma unroll
            for (int k = 0; k < DMAX; ++k) if (k < d) carry[k * kR + r] += sm.GXY[k * kRS + r];
            carry[DMAX * kR + r] = sm.GXY[d * kRS + r];
        }
        __syncthreads();
    }
    if (g_prev) {
        // adjoints of (prev_x, prev_y, prev_h)
        const int w = d + 1 + H;
        for (int e = tid; e < w * kR; e += kNT) {
            const int k = e / kR, r = e % kR;
            if (r < nvalid) {
                T v;
                if (k < d) v = carry[k * kR + r];
                else if (k == d) v = carry[DMAX * kR + r];
                else v = sm.GH[(k - d - 1) * kRS + r];
                g_prev[(chain0 + r) * w + k] = v;
            }
        }
    }
}

// ================================================================== backward, FP64 tensor pipe + TMA weight stream
template <int DMAX, int NST = kStagesBwd, int KC = kKC, int MAXROWS = kMaxRowsMmaBwd, int MINB = 1>
__global__ void __launch_bounds__(MAXROWS * (32 / kNTB) + 32, MINB)
    ts_rollout_bwd_mma_kernel(ProblemDev<double> pb, RolloutDev<double> ro, const double* __restrict__ g_loss,
                              const double* __restrict__ g_cost, const double* __restrict__ g_X,
                              const double* __restrict__ g_A, double* __restrict__ g_prev, int rows) {
    using T = double;
    if (pb.done_flag && *pb.done_flag) return;
    const int d = pb.d, H = pb.H, L = pb.L;
    const long long N = pb.N;
    const long long chain0 = (long long)blockIdx.x * rows;
    const int nvalid = (int)min((long long)rows, N - chain0);
    const int tid = threadIdx.x;
    const int ncw = (rows >> 3) * (8 / kNTB), nthr = ncw * 32;
    const long long step = pb.step_ptr ? *pb.step_ptr : 0;

    MmaBwdSmem sm;
    sm.carve(reinterpret_cast<T*>(nmb_smem_raw), d, DMAX, H, rows, NST, KC);
    if (tid == 0) {
        for (int s = 0; s < NST; ++s) {
            mbar_init(sm.full0 + 8 * s, 1);
            mbar_init(sm.empty0 + 8 * s, ncw);
        }
        fence_barrier_init();
    }
    __syncthreads();
    if (tid >= nthr) {
        mma_producer<NST, true, KC>(pb.lay, ro.theta, L, sm.ring());
        return;
    }
    MmaWarp w;
    w.init(tid >> 5, nthr, H, kNTB);
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
        mma_policy_backward<NST, KC>(pb.lay, ro.theta, head_a, sm, w, sv, node0, nvalid);
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

// ================================================================== host side
template <typename T> static ProblemDev<T> make_problem(const nmb_ts_problem_t& p) {
    ProblemDev<T> o;
    o.d = p.d; o.H = p.H; o.L = p.L; o.N = p.n_chains; o.cpp = p.chains_per_path;
    o.path = to_dev<T>(*p.path);
    o.cost = to_dev<T>(p.cost);
    o.lay.init(p.d, p.H, 1);
    o.prev_x = (const T*)p.prev_x; o.prev_y = (const T*)p.prev_y; o.prev_h = (const T*)p.prev_h;
    o.seed = p.seed; o.step_ptr = (const long long*)p.step_ptr; o.done_flag = p.done_flag;
    o.chain_offset = p.chain_offset;
    return o;
}

template <typename T> static RolloutDev<T> make_rollout(const nmb_ts_problem_t& p, const nmb_ts_rollout_t& r) {
    RolloutDev<T> o;
    o.theta = (const T*)r.theta; o.theta_t = (const T*)r.theta_t;
    o.noisy_design = r.noisy_design; o.noisy_action = r.noisy_action;
    o.rot_q = (const T*)r.rot_q; o.rot_dims = r.rot_dims; o.cost_noise = (const T*)r.cost_noise;
    o.rng_stream = r.rng_stream;
    o.X = (T*)r.X; o.A = (T*)r.A; o.Y = (T*)r.Y; o.loss = (T*)r.loss; o.cost = (T*)r.cost;
    o.ring_cs = (const float2*)r.ring_cs; o.ring_dims = r.ring_dims; o.ring_noise = (const T*)r.ring_noise;
    o.ring_len = r.ring_cs ? r.ring_len : 0;
    o.y_obs = (const T*)r.y_obs;
    o.level_begin = r.y_obs ? r.level_begin : 0;
    o.level_end = r.y_obs ? r.level_end : p.L + 1;
    o.has_save = r.save != nullptr;
    if (r.save) o.sv = carve_save<T>(r.save, p.d, p.H, p.L, p.n_chains);
    else memset(&o.sv, 0, sizeof(o.sv));
    return o;
}

static int validate_problem(const nmb_ts_problem_t* p, const char* who) {
    NMB_REQUIRE(p && p->path, NMB_EINVAL, "%s: null problem/path", who);
    NMB_REQUIRE(p->dtype == NMB_F64 || p->dtype == NMB_F32, NMB_EINVAL, "%s: dtype %d", who, p->dtype);
    NMB_REQUIRE(p->d >= 1 && p->d <= NMB_MAX_DIM, NMB_ELIMIT, "%s: d=%d outside [1,%d]", who, p->d, NMB_MAX_DIM);
    NMB_REQUIRE(p->d == p->path->d, NMB_EINVAL, "%s: path dim %d != d %d", who, p->path->d, p->d);
    NMB_REQUIRE(p->H >= 4 && p->H <= kMaxH && p->H % 4 == 0, NMB_ELIMIT, "%s: hidden_dim=%d must be a multiple of 4 in [4,%d]",
                who, p->H, kMaxH);
    NMB_REQUIRE(p->L >= 0 && p->L <= 62, NMB_ELIMIT, "%s: lookahead=%d outside [0,62]", who, p->L);
    NMB_REQUIRE(p->n_chains >= 1 && p->chains_per_path >= 1, NMB_EINVAL, "%s: bad chain counts", who);
    NMB_REQUIRE((p->n_chains + p->chains_per_path - 1) / p->chains_per_path <= p->path->n_paths, NMB_EINVAL,
                "%s: %lld chains / %lld per path exceed %d paths", who, (long long)p->n_chains,
                (long long)p->chains_per_path, p->path->n_paths);
    NMB_REQUIRE(p->prev_x && p->prev_y && p->prev_h, NMB_EINVAL, "%s: null inputs", who);
    return 0;
}

template <typename T, int DMAX>
static int launch_fwd(const nmb_ts_problem_t* p, const nmb_ts_rollout_t* r, int n, cudaStream_t st) {
    ProblemDev<T> pb = make_problem<T>(*p);
    RolloutDev<T> r0 = make_rollout<T>(*p, r[0]);
    RolloutDev<T> r1 = n > 1 ? make_rollout<T>(*p, r[1]) : r0;
    const size_t smem = (FwdSmem<T>::elems(p->d, p->H) + kR) * sizeof(T);
    auto kern = ts_rollout_fwd_kernel<T, DMAX>;
    NMB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid((unsigned)((p->n_chains + kR - 1) / kR), (unsigned)n);
    kern<<<grid, kNT, smem, st>>>(pb, r0, r1);
    NMB_LAUNCH_CHECK();
    return 0;
}

// the FP64 tensor-pipe forward needs 16-byte aligned weight rows: every segment offset even (d even), aligned base
static bool mma_fwd_ok(const nmb_ts_problem_t* p, const nmb_ts_rollout_t* r, int n) {
    if (p->dtype != NMB_F64 || (p->d & 1) || p->H < 8) return false;
    for (int i = 0; i < n; ++i)
        if (((uintptr_t)r[i].theta_t & 15) != 0) return false;
    return true;
}

template <int DMAX>
static int launch_fwd_mma(const nmb_ts_problem_t* p, const nmb_ts_rollout_t* r, int n, cudaStream_t st) {
    ProblemDev<double> pb = make_problem<double>(*p);
    RolloutDev<double> r0 = make_rollout<double>(*p, r[0]);
    RolloutDev<double> r1 = n > 1 ? make_rollout<double>(*p, r[1]) : r0;
    const int rows = mma_rows_per_cta(p->n_chains, n, sm_count());
    dim3 grid((unsigned)((p->n_chains + rows - 1) / rows), (unsigned)n);
    constexpr int kStagesSmall = 10;
    static const bool no_deep = getenv("NMB_FWD_NO_DEEP_RING") != nullptr;         // A/B switch
    if (rows <= 16 && !no_deep && MmaSmem::bytes(p->d, p->H, rows, kStagesSmall) <= (size_t)227 * 1024) {
        const size_t smem = MmaSmem::bytes(p->d, p->H, rows, kStagesSmall);
        auto kern = ts_rollout_fwd_mma_kernel<DMAX, kStagesSmall>;
        NMB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        kern<<<grid, rows * 8 + 32, smem, st>>>(pb, r0, r1, rows);
    } else {
        const size_t smem = MmaSmem::bytes(p->d, p->H, rows);
        auto kern = ts_rollout_fwd_mma_kernel<DMAX>;
        NMB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        kern<<<grid, rows * 8 + 32, smem, st>>>(pb, r0, r1, rows);
    }
    NMB_LAUNCH_CHECK();
    return 0;
}

template <typename T, int DMAX>
static int launch_bwd(const nmb_ts_problem_t* p, const nmb_ts_rollout_t* r, const void* gl, const void* gc,
                      const void* gX, const void* gA, void* gprev, cudaStream_t st) {
    ProblemDev<T> pb = make_problem<T>(*p);
    RolloutDev<T> r0 = make_rollout<T>(*p, *r);
    const size_t smem = (BwdSmem<T>::elems(p->d, p->H) + (DMAX + 1) * kR) * sizeof(T);
    auto kern = ts_rollout_bwd_kernel<T, DMAX>;
    NMB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid((unsigned)((p->n_chains + kR - 1) / kR));
    kern<<<grid, kNT, smem, st>>>(pb, r0, (const T*)gl, (const T*)gc, (const T*)gX, (const T*)gA, (T*)gprev);
    NMB_LAUNCH_CHECK();
    return 0;
}

static bool mma_bwd_ok(const nmb_ts_problem_t* p, const nmb_ts_rollout_t* r) {
    return p->dtype == NMB_F64 && !(p->d & 1) && p->H >= 8 && ((uintptr_t)r->theta & 15) == 0;
}

template <int DMAX>
static int launch_bwd_mma(const nmb_ts_problem_t* p, const nmb_ts_rollout_t* r, const void* gl, const void* gc,
                          const void* gX, const void* gA, void* gprev, cudaStream_t st) {
    ProblemDev<double> pb = make_problem<double>(*p);
    RolloutDev<double> r0 = make_rollout<double>(*p, *r);
    const int rows = mma_rows_per_cta(p->n_chains, 1, sm_count(), kMaxRowsMmaBwd);
    dim3 grid((unsigned)((p->n_chains + rows - 1) / rows));
    constexpr int kStagesSmall = 10;         // few chains: small adjoint buffers, the weight stream's latency binds
    static const bool no_deep = getenv("NMB_FWD_NO_DEEP_RING") != nullptr;
    if (rows <= 16 && !no_deep && MmaBwdSmem::bytes(p->d, DMAX, p->H, rows, kStagesSmall) <= (size_t)227 * 1024) {
        const size_t smem = MmaBwdSmem::bytes(p->d, DMAX, p->H, rows, kStagesSmall);
        auto kern = ts_rollout_bwd_mma_kernel<DMAX, kStagesSmall>;
        NMB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        kern<<<grid, rows * (32 / kNTB) + 32, smem, st>>>(pb, r0, (const double*)gl, (const double*)gc, (const double*)gX,
                                                           (const double*)gA, (double*)gprev, rows);
    } else {
        const size_t smem = MmaBwdSmem::bytes(p->d, DMAX, p->H, rows);
        auto kern = ts_rollout_bwd_mma_kernel<DMAX>;
        NMB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        kern<<<grid, rows * (32 / kNTB) + 32, smem, st>>>(pb, r0, (const double*)gl, (const double*)gc, (const double*)gX,
                                                           (const double*)gA, (double*)gprev, rows);
    }
    NMB_LAUNCH_CHECK();
    return 0;
}

}  // namespace nmb
#include "rollout_tf32.cuh"
namespace nmb {

// the in-kernel draws of (step, call, chain) exactly as the rollout kernels make them (same device functions)
__global__ void ts_device_draws_kernel(ProblemDev<double> pb, RolloutDev<double> ro, long long step, int n_calls,
                                       double* cs, int* dims, double* noise) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= pb.N * n_calls) return;
    const int call = (int)(i / pb.N);
    const long long chain = i % pb.N;
    double q[4];
    int di, dj;
    get_rotation(ro, pb, step, call, chain, q, di, dj);
    cs[2 * i] = q[0];
    cs[2 * i + 1] = q[2];
    dims[2 * i] = di;
    dims[2 * i + 1] = dj;
    noise[i] = get_cost_noise(ro, pb, step, call, chain);
}

}  // namespace nmb

using namespace nmb;

extern "C" int nmb_ts_device_draws(uint64_t seed, int64_t step, unsigned rng_stream, int n_calls, int64_t n_chains,
                                   int64_t chain_offset, int d, double* cs, int32_t* dims, double* noise, void* stream) {
    NMB_REQUIRE(cs && dims && noise && n_calls >= 1 && n_calls <= 63 && n_chains >= 1 && d >= 1 && d <= NMB_MAX_DIM,
                NMB_EINVAL, "nmb_ts_device_draws: bad argument");
    ProblemDev<double> pb;
    memset(&pb, 0, sizeof(pb));
    pb.d = d; pb.L = n_calls - 1; pb.N = n_chains; pb.seed = seed; pb.chain_offset = chain_offset;
    RolloutDev<double> ro;
    memset(&ro, 0, sizeof(ro));
    ro.rng_stream = rng_stream;
    const long long total = (long long)n_chains * n_calls;
    ts_device_draws_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(pb, ro, step, n_calls, cs,
                                                                                            dims, noise);
    NMB_LAUNCH_CHECK();
    return 0;
}

extern "C" int64_t nmb_policy_param_count(int d, int H, int n_actions) {
    PolicyLayout lay;
    lay.init(d, H, n_actions);
    return lay.total();
}

extern "C" int64_t nmb_ts_save_bytes(int dtype, int d, int H, int L, int64_t n_chains) {
    return (int64_t)save_elems(d, H, L, n_chains) * (dtype == NMB_F64 ? 8 : 4);
}

extern "C" int64_t nmb_ts_save_hidden_offset(int d, int H, int L, int64_t n_chains) {
    const size_t nodes = (size_t)(L + 1) * n_chains;
    return (int64_t)(((nodes * (d + 1) + 1) & ~(size_t)1) + nodes * H * 2);
}

#define NMB_DISPATCH_D(FN, T, d, ...)                   \
    ((d) <= 2 ? FN<T, 2>(__VA_ARGS__)                   \
     : (d) <= 4 ? FN<T, 4>(__VA_ARGS__)                 \
     : (d) <= 8 ? FN<T, 8>(__VA_ARGS__)                 \
                : FN<T, 16>(__VA_ARGS__))

extern "C" int nmb_ts_rollout_fwd(const nmb_ts_problem_t* prob, const nmb_ts_rollout_t* rollouts, int n_rollouts,
                                  void* stream) {
    int rc = validate_problem(prob, "nmb_ts_rollout_fwd");
    if (rc) return rc;
    NMB_REQUIRE(rollouts && (n_rollouts == 1 || n_rollouts == 2), NMB_EINVAL, "nmb_ts_rollout_fwd: n_rollouts=%d",
                n_rollouts);
    for (int i = 0; i < n_rollouts; ++i) {
        const nmb_ts_rollout_t& r = rollouts[i];
        NMB_REQUIRE(r.theta_t && r.A && r.loss && r.cost && (prob->L == 0 || r.X), NMB_EINVAL,
                    "nmb_ts_rollout_fwd: rollout %d has null theta_t/outputs", i);
        NMB_REQUIRE((r.rot_q == nullptr) == (r.rot_dims == nullptr), NMB_EINVAL,
                    "nmb_ts_rollout_fwd: rot_q and rot_dims must be given together");
        NMB_REQUIRE(r.ring_cs == nullptr || (r.ring_dims && r.ring_noise && r.ring_len > 0 && prob->step_ptr), NMB_EINVAL,
                    "nmb_ts_rollout_fwd: the host-stream ring needs ring_dims, ring_noise, ring_len > 0 and step_ptr");
    }
    for (int i = 0; i < n_rollouts; ++i) {
        const nmb_ts_rollout_t& r = rollouts[i];
        if (!r.y_obs) continue;
        NMB_REQUIRE(n_rollouts == 1 && r.save && r.level_begin >= 0 && r.level_end == r.level_begin + 1 &&
                        r.level_end <= prob->L + 1 && !r.noisy_design && !r.noisy_action,
                    NMB_EINVAL, "nmb_ts_rollout_fwd: imitation mode takes one saved, noise-free rollout and one level per "
                                "launch (levels [%d, %d) of %d)", r.level_begin, r.level_end, prob->L + 1);
        NMB_REQUIRE(mma_fwd_ok(prob, rollouts, n_rollouts), NMB_ELIMIT,
                    "nmb_ts_rollout_fwd: imitation mode needs the FP64 tensor-pipe kernel (float64, even d, H >= 8)");
    }
    cudaStream_t st = (cudaStream_t)stream;
    if (mma_fwd_ok(prob, rollouts, n_rollouts)) {
        const int d = prob->d;
        return d <= 2 ? launch_fwd_mma<2>(prob, rollouts, n_rollouts, st)
               : d <= 4 ? launch_fwd_mma<4>(prob, rollouts, n_rollouts, st)
               : d <= 8 ? launch_fwd_mma<8>(prob, rollouts, n_rollouts, st)
                        : launch_fwd_mma<16>(prob, rollouts, n_rollouts, st);
    }
    if (prob->dtype == NMB_F64) return NMB_DISPATCH_D(launch_fwd, double, prob->d, prob, rollouts, n_rollouts, st);
    if (tf32_fwd_ok(prob, rollouts, n_rollouts)) {
        const int d = prob->d;
        return d <= 2 ? launch_fwd_tf32<2>(prob, rollouts, n_rollouts, st)
               : d <= 4 ? launch_fwd_tf32<4>(prob, rollouts, n_rollouts, st)
               : d <= 8 ? launch_fwd_tf32<8>(prob, rollouts, n_rollouts, st)
                        : launch_fwd_tf32<16>(prob, rollouts, n_rollouts, st);
    }
    return NMB_DISPATCH_D(launch_fwd, float, prob->d, prob, rollouts, n_rollouts, st);
}

extern "C" int nmb_ts_rollout_bwd(const nmb_ts_problem_t* prob, const nmb_ts_rollout_t* rollout, const void* g_loss,
                                  const void* g_cost, const void* g_X, const void* g_A, void* g_prev, void* stream) {
    int rc = validate_problem(prob, "nmb_ts_rollout_bwd");
    if (rc) return rc;
    NMB_REQUIRE(rollout && rollout->save && rollout->theta, NMB_EINVAL,
                "nmb_ts_rollout_bwd: needs the saved rollout and theta");
    NMB_REQUIRE(g_loss && g_cost, NMB_EINVAL, "nmb_ts_rollout_bwd: null seeds");
    cudaStream_t st = (cudaStream_t)stream;
    if (mma_bwd_ok(prob, rollout)) {
        const int d = prob->d;
        return d <= 2 ? launch_bwd_mma<2>(prob, rollout, g_loss, g_cost, g_X, g_A, g_prev, st)
               : d <= 4 ? launch_bwd_mma<4>(prob, rollout, g_loss, g_cost, g_X, g_A, g_prev, st)
               : d <= 8 ? launch_bwd_mma<8>(prob, rollout, g_loss, g_cost, g_X, g_A, g_prev, st)
                        : launch_bwd_mma<16>(prob, rollout, g_loss, g_cost, g_X, g_A, g_prev, st);
    }
    if (prob->dtype == NMB_F64)
        return NMB_DISPATCH_D(launch_bwd, double, prob->d, prob, rollout, g_loss, g_cost, g_X, g_A, g_prev, st);
    if (tf32_bwd_ok(prob, rollout)) {
        const int d = prob->d;
        return d <= 2 ? launch_bwd_tf32<2>(prob, rollout, g_loss, g_cost, g_X, g_A, g_prev, st)
               : d <= 4 ? launch_bwd_tf32<4>(prob, rollout, g_loss, g_cost, g_X, g_A, g_prev, st)
               : d <= 8 ? launch_bwd_tf32<8>(prob, rollout, g_loss, g_cost, g_X, g_A, g_prev, st)
                        : launch_bwd_tf32<16>(prob, rollout, g_loss, g_cost, g_X, g_A, g_prev, st);
    }
    return NMB_DISPATCH_D(launch_bwd, float, prob->d, prob, rollout, g_loss, g_cost, g_X, g_A, g_prev, st);
}
