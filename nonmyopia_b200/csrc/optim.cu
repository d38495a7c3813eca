// Optimiser-loop kernels: parameter transpose, per-restart objective + best-restart
// bookkeeping, conditional snapshot, fused AdamW (reference actor.py:489-662).
#include <algorithm>
#include <stdlib.h>
#include "policy_dev.cuh"

namespace nmb {

struct SegTab { int off[S_COUNT + 1]; int O[S_COUNT]; int K[S_COUNT]; int splits[S_COUNT]; };

static SegTab make_segtab(int d, int H, int na, const int32_t* splits) {
    PolicyLayout lay;
    lay.init(d, H, na);
    SegTab t;
    for (int s = 0; s <= S_COUNT; ++s) t.off[s] = lay.off[s];
    for (int s = 0; s < S_COUNT; ++s) { t.O[s] = lay.O[s]; t.K[s] = lay.K[s]; t.splits[s] = splits ? splits[s] : 1; }
    return t;
}

// index of parameter i inside the transposed buffer
__device__ __forceinline__ int transposed_index(const SegTab& tb, int i, int& seg) {
    int s = 0;
    while (s + 1 < S_COUNT && i >= tb.off[s + 1]) ++s;
    seg = s;
    if (tb.K[s] == 0) return i;
    const int loc = i - tb.off[s];
    const int o = loc / tb.K[s], k = loc - o * tb.K[s];
    return tb.off[s] + k * tb.O[s] + o;
}

template <typename T>
__global__ void transpose_kernel(SegTab tb, const T* __restrict__ theta, T* __restrict__ theta_t, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int seg;
    theta_t[transposed_index(tb, i, seg)] = theta[i];
}

// torch.optim.AdamW (single-tensor path): p *= 1 - lr*wd; m.lerp_(g, 1-b1); v = b2 v + (1-b2) g^2;
// denom = sqrt(v)/sqrt(1-b2^t) + eps; p -= (lr/(1-b1^t)) * m/denom
// bias corrections of step t, computed by one thread per CTA (two FP64 pow calls) and broadcast through shared memory
__device__ __forceinline__ void adamw_bias_corrections(const nmb_adamw_t& hp, long long t, double* s_bc) {
    if (threadIdx.x == 0) {
        s_bc[0] = 1.0 - ::pow(hp.beta1, (double)t);
        s_bc[1] = ::sqrt(1.0 - ::pow(hp.beta2, (double)t));
    }
    __syncthreads();
}

// Second CTA counter of the scratch area behind the loop state (the first belongs to objective_kernel): the last CTA
// of an optimiser launch advances the loop state, after every CTA has read step / done (each CTA takes its ticket
// after its last read of the state), so no separate one-thread kernel is needed.
__device__ __forceinline__ unsigned int* advance_ticket(nmb_loop_state_t* state) {
    return reinterpret_cast<unsigned int*>(reinterpret_cast<char*>(state) + 128) + 1;
}
__device__ __forceinline__ bool last_cta_of_launch(unsigned int* ticket) {
    __shared__ int s_last;
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        const unsigned prev = atomicAdd(ticket, 1u);
        s_last = prev + 1 == gridDim.x;
        if (s_last) *ticket = 0;                         // ready for the next launch
    }
    __syncthreads();
    return s_last != 0;
}

template <typename T>
__device__ __forceinline__ void adamw_apply(const nmb_adamw_t& hp, T g, double bc1, double bc2s, long long i, int ti,
                                            T* __restrict__ theta, T* __restrict__ theta_t, T* __restrict__ m,
                                            T* __restrict__ v) {
    T p = theta[i] * (T)(1.0 - hp.lr * hp.weight_decay);
    T mi = m[i], vi = v[i];
    mi = mi + (g - mi) * (T)(1.0 - hp.beta1);
    vi = vi * (T)hp.beta2 + (T)(1.0 - hp.beta2) * g * g;
    const T denom = Num<T>::sqrt(vi) / (T)bc2s + (T)hp.eps;
    p = p - (T)(hp.lr / bc1) * (mi / denom);
    theta[i] = p; m[i] = mi; v[i] = vi;
    if (theta_t) theta_t[ti] = p;
}

template <typename T>
__global__ void adamw_kernel(SegTab tb, int use_tab, nmb_adamw_t hp, T* __restrict__ theta, T* __restrict__ theta_t,
                             T* __restrict__ m, T* __restrict__ v, const T* __restrict__ partials, long long n,
                             nmb_loop_state_t* state, int advance) {
    const bool stopped = state && state->done;
    if (stopped) return;                                  // uniform over the grid: done only changes in the last CTA's tail
    __shared__ double s_bc[2];
    adamw_bias_corrections(hp, (state ? state->step : 0) + 1, s_bc);
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        int seg = 0, ti = (int)i, nsp = 1;
        if (use_tab) { ti = transposed_index(tb, (int)i, seg); nsp = tb.splits[seg]; }
        if (nsp > 0) {
            // split-K partials summed in split order (deterministic); up to 32 loads in flight per thread
            T g = T(0);
            for (int q0 = 0; q0 < nsp; q0 += 32) {
                T pv[32];
#pragma unroll
                for (int u = 0; u < 32; ++u) pv[u] = q0 + u < nsp ? partials[(size_t)(q0 + u) * n + i] : T(0);
#pragma unroll
                for (int u = 0; u < 32; ++u) g += pv[u];
            }
            adamw_apply<T>(hp, g, s_bc[0], s_bc[1], i, ti, theta, theta_t, m, v);
        }
    }
    if (advance && state && last_cta_of_launch(advance_ticket(state))) {
        if (threadIdx.x == 0) {
            state->step += 1;
            state->done = state->done_next;
        }
    }
}

// actor.py:622-639: compare this iteration's best total (with |budget - cost|) with the min of
// best_loss + best_cost + best_KL of the last snapshot (without) - reproduced literally (SURVEY
// App. D#5).  cand_total / cand_lck are this iteration's min total and min loss+cost+KL over ALL
// restarts (all-reduced MIN over ranks before this runs when restarts are sharded).
__device__ __forceinline__ void loop_decide(nmb_loop_state_t* state, int past_initial, int patience) {
    const bool improve = state->step == 0 || state->cand_total < state->best_min;
    state->improved = improve ? 1 : 0;
    if (improve) {
        state->best_min = state->cand_lck;
        state->stall = 0;
        state->n_snapshots += 1;
    } else if (past_initial || state->step >= 200) {
        state->stall += 1;
    }
    state->done_next = (patience >= 0 && state->stall > patience) ? 1 : 0;
}

// ---- objective + bookkeeping: up to kObjMaxCtas CTAs of 256 threads; per-CTA partial results go to the scratch
// area behind the loop state (NMB_LOOP_STATE_BYTES) and the last CTA to finish reduces them in CTA order
// (deterministic) and takes the snapshot / early-stop decision.
constexpr int kObjMaxCtas = 96;
struct ObjPartial { double best, lck, sl, sc; long long idx; };
struct ObjScratch { unsigned int count; unsigned int count_adv; ObjPartial part[kObjMaxCtas]; };   // count_adv: advance_ticket()
static_assert(128 + sizeof(ObjScratch) <= NMB_LOOP_STATE_BYTES, "scratch must fit behind the loop state");

template <typename T>
__global__ void __launch_bounds__(256) objective_kernel(int d, int L, long long N, long long n_global, T budget,
                                                         int past_initial, int patience, const T* __restrict__ loss,
                                                         const T* __restrict__ cost, const T* __restrict__ X,
                                                         const T* __restrict__ A, const T* __restrict__ Xo,
                                                         const T* __restrict__ Ao, T* __restrict__ KL,
                                                         T* __restrict__ total, T* __restrict__ g_loss,
                                                         T* __restrict__ g_cost, T* __restrict__ g_X,
                                                         T* __restrict__ g_A, T* __restrict__ hist_loss,
                                                         T* __restrict__ hist_cost, T* __restrict__ hist_X,
                                                         long long hist_cap, int defer_decision, int lpc,
                                                         nmb_loop_state_t* state) {
    if (state->done) return;
    __shared__ double s_val[32];
    __shared__ double s_lck[32];
    __shared__ long long s_idx[32];
    __shared__ double s_sum[2][32];
    const int tid = threadIdx.x;
    const T inv_n = T(1) / (T)n_global;
    const T eps = T(1e-6);
    double best = 1.0 / 0.0;
    long long best_i = 0x7fffffffffffffffLL;
    double sl = 0.0, sc = 0.0, lck_min = 1.0 / 0.0;
    ObjScratch* scr = reinterpret_cast<ObjScratch*>(reinterpret_cast<char*>(state) + 128);
    // `lpc` lanes per chain (a power of two <= 32): the L + 1 policy outputs of a chain are handled by neighbouring
    // lanes (the per-level sqrt / divide chains run side by side instead of one after the other), the chain's own
    // terms by sub-lane 0
    const int sub = tid & (lpc - 1);
    const long long cpb = blockDim.x / lpc;
    const bool hist_on = state->step < hist_cap;
    const long long hstep = state->step;
    for (long long base = (long long)blockIdx.x * cpb; base < N; base += (long long)gridDim.x * cpb) {
        const long long c = base + tid / lpc;
        const bool valid = c < N;
        T kl = T(0);
        if (valid) {
            for (int l = sub; l <= L; l += lpc) {
                const T* xm = l < L ? X + ((size_t)l * N + c) * d : A + (size_t)c * d;
                const T* xo = nullptr;
                if (Xo) xo = l < L ? Xo + ((size_t)l * N + c) * d : Ao + (size_t)c * d;
                T* gdst = l < L ? (g_X ? g_X + ((size_t)l * N + c) * d : nullptr) : (g_A ? g_A + (size_t)c * d : nullptr);
                if (hist_X && hist_on && l < L)
                    for (int k = 0; k < d; ++k) hist_X[(((size_t)hstep * L + l) * N + c) * d + k] = xm[k];
                if (!xo) {
                    if (gdst) for (int k = 0; k < d; ++k) gdst[k] = T(0);
                    continue;
                }
                // F.cosine_similarity(dim=-1, eps=1e-6): x.y / (max(|x|, eps) * max(|y|, eps))
                T dot = T(0), nm = T(0), no = T(0);
                for (int k = 0; k < d; ++k) { dot = fma(xm[k], xo[k], dot); nm = fma(xm[k], xm[k], nm); no = fma(xo[k], xo[k], no); }
                const T rm = Num<T>::sqrt(nm), rqo = Num<T>::sqrt(no);
                const T cm = rm > eps ? rm : eps, co = rqo > eps ? rqo : eps;
                const T cs = dot / (cm * co);
                kl += T(1) - cs;
                if (gdst) {
                    // d(1 - cos)/d xm = -( xo/(cm co) - cos * xm / cm^2 )   (norm clamp inactive: |xm| > eps)
                    const T icc = T(1) / (cm * co), w = rm > eps ? cs / (cm * cm) : T(0);
                    for (int k = 0; k < d; ++k) gdst[k] = -inv_n * (xo[k] * icc - w * xm[k]);
                }
            }
        }
        for (int m = lpc >> 1; m > 0; m >>= 1) kl += __shfl_xor_sync(0xffffffffu, kl, m);   // fixed order: deterministic
        if (!valid || sub != 0) continue;
        const T lc = loss[c], cc = cost[c];
        const T gap = budget - cc;
        const T tot = lc + Num<T>::fabs(gap) + kl;
        if (KL) KL[c] = kl;
        total[c] = tot;
        g_loss[c] = inv_n;
        g_cost[c] = gap > T(0) ? -inv_n : (gap < T(0) ? inv_n : T(0));
        if (hist_loss && hist_on) {
            hist_loss[(size_t)hstep * N + c] = lc;
            hist_cost[(size_t)hstep * N + c] = cc;
        }
        sl += (double)lc; sc += (double)cc;
        if ((double)tot < best) { best = (double)tot; best_i = c; }   // ascending chains per lane: first minimum kept
        const double lck = (double)((lc + cc) + kl);                  // best_loss + best_cost + best_KL of a snapshot
        lck_min = lck < lck_min ? lck : lck_min;
    }
    // block argmin (lowest index among equal minima, like torch.argmin)
    for (int m = 16; m > 0; m >>= 1) {
        const double ov = __shfl_xor_sync(0xffffffffu, best, m);
        const long long oi = __shfl_xor_sync(0xffffffffu, best_i, m);
        if (ov < best || (ov == best && oi < best_i)) { best = ov; best_i = oi; }
        sl += __shfl_xor_sync(0xffffffffu, sl, m);
        sc += __shfl_xor_sync(0xffffffffu, sc, m);
        const double ol = __shfl_xor_sync(0xffffffffu, lck_min, m);
        lck_min = ol < lck_min ? ol : lck_min;
    }
    if ((tid & 31) == 0) { s_lck[tid >> 5] = lck_min; s_val[tid >> 5] = best; s_idx[tid >> 5] = best_i; s_sum[0][tid >> 5] = sl; s_sum[1][tid >> 5] = sc; }
    __syncthreads();
    if (tid == 0) {
        const int nw = blockDim.x >> 5;
        for (int w = 1; w < nw; ++w) {
            if (s_val[w] < best || (s_val[w] == best && s_idx[w] < best_i)) { best = s_val[w]; best_i = s_idx[w]; }
            sl += s_sum[0][w]; sc += s_sum[1][w];
            lck_min = s_lck[w] < lck_min ? s_lck[w] : lck_min;
        }
        if (gridDim.x > 1) {
            ObjPartial& me = scr->part[blockIdx.x];
            me.best = best; me.lck = lck_min; me.sl = sl; me.sc = sc; me.idx = best_i;
            __threadfence();
            const unsigned prev = atomicAdd(&scr->count, 1u);
            if (prev + 1 != gridDim.x) return;          // not the last CTA
            __threadfence();
            scr->count = 0;                              // ready for the next launch
            best = 1.0 / 0.0; best_i = 0x7fffffffffffffffLL; sl = 0.0; sc = 0.0; lck_min = 1.0 / 0.0;
            for (unsigned b = 0; b < gridDim.x; ++b) {
                const volatile ObjPartial& o = scr->part[b];
                const double ob = o.best; const long long oi = o.idx;
                if (ob < best || (ob == best && oi < best_i)) { best = ob; best_i = oi; }
                sl += o.sl; sc += o.sc;
                const double ol = o.lck;
                lck_min = ol < lck_min ? ol : lck_min;
            }
        }
        state->chosen_idx = (int)best_i;
        state->loss_mean = sl / (double)N;
        state->cost_mean = sc / (double)N;
        state->cand_total = best;
        state->cand_lck = lck_min;
        if (!defer_decision) loop_decide(state, past_initial, patience);
    }
}

__global__ void decide_kernel(nmb_loop_state_t* state, int past_initial, int patience) {
    if (state->done) return;
    loop_decide(state, past_initial, patience);
}


// ------------------------------------------------------------------ restart shards over NVLink peer memory
// One exchange per Adam iteration when the restarts are sharded (SURVEY 8e, C2 + C3), written against the ranks'
// peer-mapped buffers instead of two NCCL all-reduces:
//   publish   : this rank's policy gradient (split-K partials already reduced) and its two decision candidates go
//               into slot[epoch & 1] of its own buffer; the last CTA to finish raises flag[rank] = epoch + 1 in
//               EVERY rank's buffer (system-scope release).
//   wait      : one warp spins (acquire) until all ranks' flags reached epoch + 1, MIN-reduces the candidates read
//               from the peers and takes the snapshot / early-stop decision (identical on every rank).
//   adamw     : every thread sums its gradient element over the ranks in rank order (bitwise identical everywhere,
//               so the replicated weights never drift) and applies AdamW: the all-reduce is fused into the update.
// Double-buffered slots: a rank can run at most one exchange ahead of the slowest one (it cannot pass `wait` of
// epoch e + 1 before every rank has published e + 1, which each does after its own reads of epoch e).
// Peer loads use ld.cv (never served from a stale L1 line; peer addresses bypass the local L2).
struct PeerHdr {
    unsigned long long flag[NMB_MAX_PEERS];   // flag[r]: last epoch + 1 published by rank r
    unsigned long long epoch;                 // exchanges completed by this rank
    unsigned int ticket;                      // CTA counter of publish
    unsigned int pad;
};
static_assert(sizeof(PeerHdr) <= NMB_PEER_HEADER_BYTES, "peer header");

struct PeerTab {
    unsigned char* base[NMB_MAX_PEERS];
    int rank, world;
    long long n;
    long long slot_bytes, cand_off;
    long long ll_off;         // LL lines of the fused exchange: [2 parities][NMB_MAX_PEERS][n + 2] x 16 bytes behind the two slots
};

template <typename T> static PeerTab make_peertab(const nmb_peer_exchange_t* px) {
    PeerTab t;
    for (int r = 0; r < NMB_MAX_PEERS; ++r) t.base[r] = (unsigned char*)px->base[r];
    t.rank = px->rank; t.world = px->world; t.n = px->n;
    t.cand_off = (px->n * (long long)sizeof(T) + 15) / 16 * 16;
    t.slot_bytes = t.cand_off + 16;
    t.ll_off = NMB_PEER_HEADER_BYTES + 2 * t.slot_bytes;
    return t;
}

__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

template <typename T>
__global__ void __launch_bounds__(256) peer_publish_kernel(PeerTab pt, SegTab tb, int use_tab, const T* __restrict__ grad,
                                                            const nmb_loop_state_t* state) {
    if (state->done) return;
    PeerHdr* hdr = reinterpret_cast<PeerHdr*>(pt.base[pt.rank]);
    const unsigned long long e = hdr->epoch;
    unsigned char* slot = pt.base[pt.rank] + NMB_PEER_HEADER_BYTES + (e & 1) * pt.slot_bytes;
    T* dst = reinterpret_cast<T*>(slot);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < pt.n; i += (long long)gridDim.x * blockDim.x) {
        int nsp = 1;
        if (use_tab) { int seg; transposed_index(tb, (int)i, seg); nsp = tb.splits[seg]; }
        // split-K partials of the weight-gradient kernel summed in split order (deterministic), 16 loads in flight
        T g = T(0);
        for (int q0 = 0; q0 < nsp; q0 += 16) {
            T pv[16];
#pragma unroll
            for (int u = 0; u < 16; ++u) pv[u] = q0 + u < nsp ? grad[(size_t)(q0 + u) * pt.n + i] : T(0);
#pragma unroll
            for (int u = 0; u < 16; ++u) g += pv[u];
        }
        dst[i] = g;
    }
    __shared__ int last;
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) last = atomicAdd(&hdr->ticket, 1u) == gridDim.x - 1;
    __syncthreads();
    if (!last) return;
    if (threadIdx.x == 0) {
        hdr->ticket = 0;
        double* cand = reinterpret_cast<double*>(slot + pt.cand_off);
        cand[0] = state->cand_total;
        cand[1] = state->cand_lck;
        __threadfence_system();
    }
    __syncthreads();
    if (threadIdx.x < pt.world) {
        __threadfence_system();
        st_release_sys(&reinterpret_cast<PeerHdr*>(pt.base[threadIdx.x])->flag[pt.rank], e + 1);
    }
}

__global__ void __launch_bounds__(32) peer_wait_decide_kernel(PeerTab pt, nmb_loop_state_t* state, int past_initial,
                                                              int patience, long long timeout_cycles) {
    if (state->done) return;
    const PeerHdr* hdr = reinterpret_cast<const PeerHdr*>(pt.base[pt.rank]);
    const unsigned long long e = hdr->epoch;
    const int lane = threadIdx.x;
    double ct = INFINITY, cl = INFINITY;
    bool timed_out = false;
    if (lane < pt.world) {
        // a rank that died must not hang this GPU: give up after the timeout (NMB_PEER_TIMEOUT_MS, default 10 s,
        // converted with the SM clock rate on the host) and stop the loop
        const long long t0 = clock64();
        while (ld_acquire_sys(&hdr->flag[lane]) < e + 1) {
            __nanosleep(40);
            if (clock64() - t0 > timeout_cycles) { timed_out = true; break; }
        }
    }
    if (__any_sync(0xffffffffu, timed_out)) {
        if (lane == 0) { state->stall = NMB_PEER_TIMEOUT_STALL; state->done_next = 1; state->done = 1; }
        return;
    }
    if (lane < pt.world) {
        const double* cand = reinterpret_cast<const double*>(pt.base[lane] + NMB_PEER_HEADER_BYTES + (e & 1) * pt.slot_bytes +
                                                             pt.cand_off);
        ct = __ldcv(cand);
        cl = __ldcv(cand + 1);
    }
    for (int m = 16; m > 0; m >>= 1) {
        ct = fmin(ct, __shfl_xor_sync(0xffffffffu, ct, m));
        cl = fmin(cl, __shfl_xor_sync(0xffffffffu, cl, m));
    }
    if (lane == 0) {
        state->cand_total = ct;
        state->cand_lck = cl;
        loop_decide(state, past_initial, patience);
    }
}

template <typename T>
__global__ void __launch_bounds__(256) peer_adamw_kernel(PeerTab pt, SegTab tb, nmb_adamw_t hp, T* __restrict__ theta,
                                                         T* __restrict__ theta_t, T* __restrict__ m, T* __restrict__ v,
                                                         nmb_loop_state_t* state) {
    if (state->done) return;
    PeerHdr* hdr = reinterpret_cast<PeerHdr*>(pt.base[pt.rank]);
    const unsigned long long e = hdr->epoch;
    __shared__ double s_bc[2];
    adamw_bias_corrections(hp, state->step + 1, s_bc);
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < pt.n) {
        const long long off = NMB_PEER_HEADER_BYTES + (long long)(e & 1) * pt.slot_bytes;
        T pv[NMB_MAX_PEERS];
#pragma unroll
        for (int r = 0; r < NMB_MAX_PEERS; ++r)
            pv[r] = r < pt.world ? __ldcv(reinterpret_cast<const T*>(pt.base[r] + off) + i) : T(0);
        T g = T(0);
#pragma unroll
        for (int r = 0; r < NMB_MAX_PEERS; ++r) g += pv[r];
        int seg;
        const int ti = transposed_index(tb, (int)i, seg);
        adamw_apply<T>(hp, g, s_bc[0], s_bc[1], i, ti, theta, theta_t, m, v);
    }
    // end of a sharded iteration: the exchange epoch and the loop state advance together, in the last CTA to finish
    // (every CTA has read epoch / step / done before it takes its ticket)
    if (last_cta_of_launch(advance_ticket(state)) && threadIdx.x == 0) {
        hdr->epoch = e + 1;
        state->step += 1;
        state->done = state->done_next;
    }
}


// ---- the whole exchange in ONE kernel (the three-launch form above stays for callers that need the phases apart).
// Measured on 2 GPUs (tools/peer_clocks.py, -DNMB_PEER_CLOCKS): a flag-per-slice version of this kernel spent 7 us of
// every CTA's 16 us in `fence.sys` + the release store (213 CTAs fencing at once; whole kernel 31 us against 8.6 us
// for the single-GPU AdamW kernel).  This version needs no fence and no flag: values travel as self-validating
// 16-byte lines {lo, seq, hi, seq} (8-byte lines {bits, seq} in FP32) - the LL scheme of NCCL: an 8-byte store is
// atomic on NVLink, so a line whose sequence words both equal epoch + 1 is complete.
//   1. thread i reduces element i of the split-K partials and PUSHES it into recv[epoch & 1][rank][i] of every peer
//      (one NVLink trip instead of flag trip + read round trip); CTA 0 pushes the rank's two decision candidates;
//   2. it polls its own recv lines of the other ranks (local memory, L1 bypassed) until they carry epoch + 1, sums in
//      rank order with its own value from the register, applies AdamW: bitwise identical weights on every rank;
//   3. the last CTA to finish takes the MIN of the candidates, the snapshot / early-stop decision (loop_decide) and
//      advances the exchange epoch and the loop state.
// Two parity buffers: a rank can start exchange e + 2 (which overwrites parity e & 1) only after it has received every
// peer's lines of e + 1, which a peer sends after it finished reading e.  The snapshot copy (rollout outputs only,
// never the weights) runs AFTER this kernel and reads state->improved.
__device__ __forceinline__ void ll_store(void* p, double v, unsigned seq) {
    const unsigned lo = (unsigned)__double2loint(v), hi = (unsigned)__double2hiint(v);
    asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(lo), "r"(seq), "r"(hi), "r"(seq) : "memory");
}
__device__ __forceinline__ void ll_store(void* p, float v, unsigned seq) {
    asm volatile("st.volatile.global.v2.u32 [%0], {%1, %2};" ::"l"(p), "r"(__float_as_uint(v)), "r"(seq) : "memory");
}
__device__ __forceinline__ bool ll_try_load(const void* p, unsigned seq, double& v) {
    unsigned lo, f0, hi, f1;
    asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(lo), "=r"(f0), "=r"(hi), "=r"(f1) : "l"(p) : "memory");
    v = __hiloint2double((int)hi, (int)lo);
    return f0 == seq && f1 == seq;
}
__device__ __forceinline__ bool ll_try_load(const void* p, unsigned seq, float& v) {
    unsigned b, f;
    asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];" : "=r"(b), "=r"(f) : "l"(p) : "memory");
    v = __uint_as_float(b);
    return f == seq;
}
// line (of_rank, i) of parity `par` in rank `in_rank`'s buffer; lines are 16 bytes apart in both precisions;
// i = n, n + 1: the two decision candidates (always doubles)
__device__ __forceinline__ unsigned char* ll_line(const PeerTab& pt, int in_rank, int par, int of_rank, long long i) {
    return pt.base[in_rank] + pt.ll_off + (((size_t)par * NMB_MAX_PEERS + of_rank) * (size_t)(pt.n + 2) + (size_t)i) * 16;
}

#ifdef NMB_PEER_CLOCKS
static __device__ unsigned long long g_peer_clk[12];
static __device__ unsigned long long g_peer_span[2] = {~0ULL, 0ULL};
__device__ __forceinline__ unsigned long long gtime() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
#define NMB_PCLK(expr) do { if (threadIdx.x == 0) { expr; } } while (0)
extern "C" int nmb_debug_peer_clocks(unsigned long long* out12) {
    return cudaMemcpyFromSymbol(out12, g_peer_clk, sizeof(g_peer_clk)) == cudaSuccess ? 0 : -1;
}
#else
#define NMB_PCLK(expr) do { } while (0)
#endif

template <typename T>
__global__ void __launch_bounds__(256) peer_exchange_adamw_kernel(PeerTab pt, SegTab tb, const T* __restrict__ partials,
                                                                  nmb_adamw_t hp, T* __restrict__ theta,
                                                                  T* __restrict__ theta_t, T* __restrict__ m,
                                                                  T* __restrict__ v, nmb_loop_state_t* state,
                                                                  int past_initial, int patience, long long timeout_cycles) {
    if (state->done) return;
    PeerHdr* hdr = reinterpret_cast<PeerHdr*>(pt.base[pt.rank]);
    const unsigned long long e = hdr->epoch;
    const unsigned seq = (unsigned)(e + 1);
    const int par = (int)(e & 1);
    const int tid = threadIdx.x;
    const long long i = (long long)blockIdx.x * blockDim.x + tid;
    __shared__ double s_bc[2];
#ifdef NMB_PEER_CLOCKS
    unsigned long long tk0 = 0, tk1 = 0, tk2 = 0, tk3 = 0;
#endif
    NMB_PCLK(tk0 = gtime());
    // ---- 1. reduce + push
    T g = T(0);
    int ti = 0, nsp = 0;
    if (i < pt.n) {
        int seg;
        ti = transposed_index(tb, (int)i, seg);
        nsp = tb.splits[seg];
        for (int q0 = 0; q0 < nsp; q0 += 32) {
            T pv[32];
#pragma unroll
            for (int u = 0; u < 32; ++u) pv[u] = q0 + u < nsp ? partials[(size_t)(q0 + u) * pt.n + i] : T(0);
#pragma unroll
            for (int u = 0; u < 32; ++u) g += pv[u];
        }
        for (int r = 0; r < pt.world; ++r)
            if (r != pt.rank) ll_store(ll_line(pt, r, par, pt.rank, i), g, seq);
    }
    if (blockIdx.x == 0 && tid < 2) {
        const double c = tid == 0 ? state->cand_total : state->cand_lck;
        for (int r = 0; r < pt.world; ++r) ll_store(ll_line(pt, r, par, pt.rank, pt.n + tid), c, seq);
    }
    NMB_PCLK(tk1 = gtime());
    adamw_bias_corrections(hp, state->step + 1, s_bc);
    // ---- 2. receive, sum in rank order, update
    bool timed_out = false;
    if (i < pt.n && nsp > 0) {
        T pv[NMB_MAX_PEERS];
        unsigned pending = 0;
#pragma unroll
        for (int r = 0; r < NMB_MAX_PEERS; ++r) {
            pv[r] = T(0);
            if (r < pt.world) {
                if (r == pt.rank) pv[r] = g;
                else if (!ll_try_load(ll_line(pt, pt.rank, par, r, i), seq, pv[r])) pending |= 1u << r;
            }
        }
        if (pending) {
            const long long t0 = clock64();
            while (pending) {
#pragma unroll
                for (int r = 0; r < NMB_MAX_PEERS; ++r)
                    if ((pending >> r) & 1u)
                        if (ll_try_load(ll_line(pt, pt.rank, par, r, i), seq, pv[r])) pending &= ~(1u << r);
                if (pending && clock64() - t0 > timeout_cycles) { timed_out = true; break; }
            }
        }
        NMB_PCLK(tk2 = gtime());
        if (!timed_out) {
            T gs = T(0);
#pragma unroll
            for (int r = 0; r < NMB_MAX_PEERS; ++r) gs += pv[r];          // rank order; absent ranks hold exact zeros
            adamw_apply<T>(hp, gs, s_bc[0], s_bc[1], i, ti, theta, theta_t, m, v);
        }
    }
    if (__syncthreads_or(timed_out)) {
        // a dead peer: stop the loop and do NOT take a ticket (no CTA becomes the last one, the state is not advanced)
        if (tid == 0) { state->stall = NMB_PEER_TIMEOUT_STALL; state->done_next = 1; state->done = 1; }
        return;
    }
#ifdef NMB_PEER_CLOCKS
    if (tid == 0) {
        tk3 = gtime();
        atomicMin(&g_peer_span[0], tk0);
        atomicMax(&g_peer_span[1], tk3);
        if (blockIdx.x == 100) {
            atomicAdd(&g_peer_clk[0], tk1 - tk0); atomicAdd(&g_peer_clk[1], tk2 - tk1); atomicAdd(&g_peer_clk[2], tk3 - tk2);
            atomicAdd(&g_peer_clk[3], 1ULL);
        }
    }
#endif
    // ---- 3. decision + advance, once per launch
    if (!last_cta_of_launch(advance_ticket(state))) return;
    if (tid < 32) {
        double c0 = INFINITY, c1 = INFINITY;
        bool to = false;
        if (tid < pt.world) {
            const long long t0 = clock64();
            bool ok0 = false, ok1 = false;
            while (!(ok0 && ok1)) {
                if (!ok0) ok0 = ll_try_load(ll_line(pt, pt.rank, par, tid, pt.n), seq, c0);
                if (!ok1) ok1 = ll_try_load(ll_line(pt, pt.rank, par, tid, pt.n + 1), seq, c1);
                if (!(ok0 && ok1) && clock64() - t0 > timeout_cycles) { to = true; break; }
            }
        }
        to = __any_sync(0xffffffffu, to);
        double ct = to ? INFINITY : c0, cl = to ? INFINITY : c1;
        for (int q = 16; q > 0; q >>= 1) {
            ct = fmin(ct, __shfl_xor_sync(0xffffffffu, ct, q));
            cl = fmin(cl, __shfl_xor_sync(0xffffffffu, cl, q));
        }
        if (tid == 0) {
            if (to) {
                state->stall = NMB_PEER_TIMEOUT_STALL; state->done_next = 1; state->done = 1;
            } else {
                state->cand_total = ct;
                state->cand_lck = cl;
                loop_decide(state, past_initial, patience);
                hdr->epoch = e + 1;
                state->step += 1;
                state->done = state->done_next;
            }
#ifdef NMB_PEER_CLOCKS
            const unsigned long long cend = gtime();
            g_peer_clk[7] += cend - g_peer_span[0];
            g_peer_clk[8] += g_peer_span[1] - g_peer_span[0];
            g_peer_clk[10] += 1;
            g_peer_span[0] = ~0ULL; g_peer_span[1] = 0;
#endif
        }
    }
}

struct CopyList { const void* src[8]; void* dst[8]; long long n[8]; int count; };

template <typename T>
__global__ void snapshot_kernel(CopyList cl, const nmb_loop_state_t* state) {
    if (state->done || !state->improved) return;
    for (int q = 0; q < cl.count; ++q) {
        const T* s = (const T*)cl.src[q];
        T* dptr = (T*)cl.dst[q];
        for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < cl.n[q];
             i += (long long)gridDim.x * blockDim.x)
            dptr[i] = s[i];
    }
}

}  // namespace nmb

using namespace nmb;

extern "C" int nmb_policy_transpose(int dtype, int d, int H, int n_actions, const void* theta, void* theta_t,
                                    void* stream) {
    NMB_REQUIRE(theta && theta_t, NMB_EINVAL, "nmb_policy_transpose: null pointer");
    SegTab tb = make_segtab(d, H, n_actions, nullptr);
    const int n = tb.off[S_COUNT];
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == NMB_F64) transpose_kernel<double><<<(n + 255) / 256, 256, 0, st>>>(tb, (const double*)theta, (double*)theta_t, n);
    else if (dtype == NMB_F32) transpose_kernel<float><<<(n + 255) / 256, 256, 0, st>>>(tb, (const float*)theta, (float*)theta_t, n);
    else { set_error("nmb_policy_transpose: dtype %d", dtype); return NMB_EINVAL; }
    NMB_LAUNCH_CHECK();
    return 0;
}

extern "C" int nmb_ts_objective(int dtype, int d, int L, int64_t n_chains, int64_t n_global, double budget,
                                int past_initial, int patience, const void* loss, const void* cost, const void* X, const void* A,
                                const void* X_orig, const void* A_orig, void* KL, void* total, void* g_loss,
                                void* g_cost, void* g_X, void* g_A, void* hist_loss, void* hist_cost, void* hist_X,
                                int64_t hist_capacity, int defer_decision, nmb_loop_state_t* state, void* stream) {
    NMB_REQUIRE(loss && cost && A && total && g_loss && g_cost && state, NMB_EINVAL, "nmb_ts_objective: null argument");
    NMB_REQUIRE((X_orig == nullptr) == (A_orig == nullptr), NMB_EINVAL, "nmb_ts_objective: X_orig/A_orig mismatch");
    NMB_REQUIRE(L == 0 || X, NMB_EINVAL, "nmb_ts_objective: X is null");
    cudaStream_t st = (cudaStream_t)stream;
    int lpc = 1;
    while (lpc < 32 && lpc < L + 1) lpc <<= 1;
    const int obj_ctas = (int)std::min<long long>(kObjMaxCtas, (n_chains * lpc + 255) / 256);
    if (dtype == NMB_F64) {
        objective_kernel<double><<<obj_ctas, 256, 0, st>>>(d, L, n_chains, n_global, budget, past_initial, patience, (const double*)loss,
            (const double*)cost, (const double*)X, (const double*)A, (const double*)X_orig, (const double*)A_orig,
            (double*)KL, (double*)total, (double*)g_loss, (double*)g_cost, (double*)g_X, (double*)g_A,
            (double*)hist_loss, (double*)hist_cost, (double*)hist_X, hist_capacity, defer_decision, lpc, state);
    } else if (dtype == NMB_F32) {
        objective_kernel<float><<<obj_ctas, 256, 0, st>>>(d, L, n_chains, n_global, (float)budget, past_initial, patience, (const float*)loss,
            (const float*)cost, (const float*)X, (const float*)A, (const float*)X_orig, (const float*)A_orig,
            (float*)KL, (float*)total, (float*)g_loss, (float*)g_cost, (float*)g_X, (float*)g_A,
            (float*)hist_loss, (float*)hist_cost, (float*)hist_X, hist_capacity, defer_decision, lpc, state);
    } else { set_error("nmb_ts_objective: dtype %d", dtype); return NMB_EINVAL; }
    NMB_LAUNCH_CHECK();
    return 0;
}

extern "C" int nmb_snapshot_if_improved(int dtype, int n_copies, const void* const* src_host, void* const* dst_host,
                                        const int64_t* n_elems_host, const nmb_loop_state_t* state, void* stream) {
    NMB_REQUIRE(n_copies >= 1 && n_copies <= 8 && src_host && dst_host && n_elems_host && state, NMB_EINVAL,
                "nmb_snapshot_if_improved: bad arguments");
    CopyList cl;
    cl.count = n_copies;
    long long mx = 0;
    for (int i = 0; i < n_copies; ++i) {
        cl.src[i] = src_host[i]; cl.dst[i] = dst_host[i]; cl.n[i] = n_elems_host[i];
        mx = cl.n[i] > mx ? cl.n[i] : mx;
    }
    long long blocks = (mx + 255) / 256;
    const long long cap = (long long)sm_count() * 8;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == NMB_F64) snapshot_kernel<double><<<(unsigned)blocks, 256, 0, st>>>(cl, state);
    else if (dtype == NMB_F32) snapshot_kernel<float><<<(unsigned)blocks, 256, 0, st>>>(cl, state);
    else { set_error("nmb_snapshot_if_improved: dtype %d", dtype); return NMB_EINVAL; }
    NMB_LAUNCH_CHECK();
    return 0;
}

template <typename T>
static int launch_adamw(SegTab tb, int use_tab, const nmb_adamw_t* hp, void* theta, void* theta_t, void* m, void* v,
                        const void* partials, long long n, nmb_loop_state_t* state, int advance, cudaStream_t st) {
    adamw_kernel<T><<<(unsigned)((n + 255) / 256), 256, 0, st>>>(tb, use_tab, *hp, (T*)theta, (T*)theta_t, (T*)m, (T*)v,
                                                                 (const T*)partials, n, state, advance);
    NMB_LAUNCH_CHECK();
    return 0;
}

extern "C" int nmb_policy_adamw_step(int dtype, int d, int H, int n_actions, const nmb_adamw_t* hp, void* theta,
                                     void* theta_t, void* exp_avg, void* exp_avg_sq, const void* partials,
                                     const int32_t* seg_splits_host, nmb_loop_state_t* state, int advance_state,
                                     void* stream) {
    NMB_REQUIRE(hp && theta && exp_avg && exp_avg_sq && partials && seg_splits_host, NMB_EINVAL,
                "nmb_policy_adamw_step: null argument");
    SegTab tb = make_segtab(d, H, n_actions, seg_splits_host);
    const long long n = tb.off[S_COUNT];
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == NMB_F64) return launch_adamw<double>(tb, 1, hp, theta, theta_t, exp_avg, exp_avg_sq, partials, n, state, advance_state, st);
    if (dtype == NMB_F32) return launch_adamw<float>(tb, 1, hp, theta, theta_t, exp_avg, exp_avg_sq, partials, n, state, advance_state, st);
    set_error("nmb_policy_adamw_step: dtype %d", dtype);
    return NMB_EINVAL;
}

extern "C" int nmb_adamw_flat(int dtype, int64_t n, const nmb_adamw_t* hp, void* param, void* exp_avg, void* exp_avg_sq,
                              const void* grad, nmb_loop_state_t* state, int advance_state, void* stream) {
    NMB_REQUIRE(hp && param && exp_avg && exp_avg_sq && grad && n >= 0, NMB_EINVAL, "nmb_adamw_flat: bad argument");
    if (n == 0) return 0;
    SegTab tb;
    memset(&tb, 0, sizeof(tb));
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == NMB_F64) return launch_adamw<double>(tb, 0, hp, param, nullptr, exp_avg, exp_avg_sq, grad, n, state, advance_state, st);
    if (dtype == NMB_F32) return launch_adamw<float>(tb, 0, hp, param, nullptr, exp_avg, exp_avg_sq, grad, n, state, advance_state, st);
    set_error("nmb_adamw_flat: dtype %d", dtype);
    return NMB_EINVAL;
}

extern "C" int nmb_loop_decide(nmb_loop_state_t* state, int past_initial, int patience, void* stream) {
    NMB_REQUIRE(state, NMB_EINVAL, "nmb_loop_decide: null state");
    decide_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(state, past_initial, patience);
    NMB_LAUNCH_CHECK();
    return 0;
}

static int validate_peers(const nmb_peer_exchange_t* px, const char* who) {
    NMB_REQUIRE(px, NMB_EINVAL, "%s: null exchange", who);
    NMB_REQUIRE(px->world >= 1 && px->world <= NMB_MAX_PEERS && px->rank >= 0 && px->rank < px->world, NMB_EINVAL,
                "%s: rank %d / world %d (at most %d peers)", who, px->rank, px->world, NMB_MAX_PEERS);
    NMB_REQUIRE(px->n >= 1, NMB_EINVAL, "%s: n=%lld", who, (long long)px->n);
    for (int r = 0; r < px->world; ++r) NMB_REQUIRE(px->base[r], NMB_EINVAL, "%s: peer %d buffer is null", who, r);
    return 0;
}

extern "C" int64_t nmb_peer_exchange_bytes(int dtype, int64_t n) {
    const int64_t esz = dtype == NMB_F32 ? 4 : 8;
    // header + two gradient slots (each followed by the two decision candidates) + the LL receive lines of the fused
    // exchange: [2 parities][NMB_MAX_PEERS][n + 2] x 16 bytes
    return NMB_PEER_HEADER_BYTES + 2 * ((n * esz + 15) / 16 * 16 + 16) + (int64_t)2 * NMB_MAX_PEERS * (n + 2) * 16;
}

extern "C" int nmb_peer_publish(const nmb_peer_exchange_t* px, int dtype, const void* grad, int d, int H, int n_actions,
                                const int32_t* seg_splits_host, const nmb_loop_state_t* state, void* stream) {
    int rc = validate_peers(px, "nmb_peer_publish");
    if (rc) return rc;
    NMB_REQUIRE(grad && state, NMB_EINVAL, "nmb_peer_publish: null argument");
    SegTab tb;
    memset(&tb, 0, sizeof(tb));
    const int use_tab = seg_splits_host != nullptr;
    if (use_tab) {
        tb = make_segtab(d, H, n_actions, seg_splits_host);
        NMB_REQUIRE(tb.off[S_COUNT] == px->n, NMB_EINVAL, "nmb_peer_publish: exchange holds %lld elements, policy has %d",
                    (long long)px->n, tb.off[S_COUNT]);
    }
    const unsigned grid = (unsigned)std::min<long long>((px->n + 255) / 256, 2LL * sm_count());
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == NMB_F64) peer_publish_kernel<double><<<grid, 256, 0, st>>>(make_peertab<double>(px), tb, use_tab, (const double*)grad, state);
    else if (dtype == NMB_F32) peer_publish_kernel<float><<<grid, 256, 0, st>>>(make_peertab<float>(px), tb, use_tab, (const float*)grad, state);
    else { set_error("nmb_peer_publish: dtype %d", dtype); return NMB_EINVAL; }
    NMB_LAUNCH_CHECK();
    return 0;
}

static long long peer_timeout_cycles();

extern "C" int nmb_peer_wait_decide(const nmb_peer_exchange_t* px, int dtype, nmb_loop_state_t* state, int past_initial,
                                    int patience, void* stream) {
    int rc = validate_peers(px, "nmb_peer_wait_decide");
    if (rc) return rc;
    NMB_REQUIRE(state && (dtype == NMB_F64 || dtype == NMB_F32), NMB_EINVAL, "nmb_peer_wait_decide: bad argument");
    const PeerTab pt = dtype == NMB_F64 ? make_peertab<double>(px) : make_peertab<float>(px);
    const long long timeout_cycles = peer_timeout_cycles();
    peer_wait_decide_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(pt, state, past_initial, patience, timeout_cycles);
    NMB_LAUNCH_CHECK();
    return 0;
}

extern "C" int nmb_peer_adamw_step(const nmb_peer_exchange_t* px, int dtype, int d, int H, int n_actions,
                                   const nmb_adamw_t* hp, void* theta, void* theta_t, void* exp_avg, void* exp_avg_sq,
                                   nmb_loop_state_t* state, void* stream) {
    int rc = validate_peers(px, "nmb_peer_adamw_step");
    if (rc) return rc;
    NMB_REQUIRE(hp && theta && exp_avg && exp_avg_sq && state, NMB_EINVAL, "nmb_peer_adamw_step: null argument");
    SegTab tb = make_segtab(d, H, n_actions, nullptr);
    NMB_REQUIRE(tb.off[S_COUNT] == px->n, NMB_EINVAL, "nmb_peer_adamw_step: exchange holds %lld elements, policy has %d",
                (long long)px->n, tb.off[S_COUNT]);
    cudaStream_t st = (cudaStream_t)stream;
    const unsigned grid = (unsigned)((px->n + 255) / 256);
    PeerTab pt;
    if (dtype == NMB_F64) {
        pt = make_peertab<double>(px);
        peer_adamw_kernel<double><<<grid, 256, 0, st>>>(pt, tb, *hp, (double*)theta, (double*)theta_t, (double*)exp_avg,
                                                         (double*)exp_avg_sq, state);
    } else if (dtype == NMB_F32) {
        pt = make_peertab<float>(px);
        peer_adamw_kernel<float><<<grid, 256, 0, st>>>(pt, tb, *hp, (float*)theta, (float*)theta_t, (float*)exp_avg,
                                                        (float*)exp_avg_sq, state);
    } else { set_error("nmb_peer_adamw_step: dtype %d", dtype); return NMB_EINVAL; }
    NMB_LAUNCH_CHECK();
    return 0;
}

static long long peer_timeout_cycles() {
    static long long timeout_cycles = 0;
    if (timeout_cycles == 0) {
        int dev = 0, khz = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev);
        const char* env = getenv("NMB_PEER_TIMEOUT_MS");
        const double ms = env && atof(env) > 0 ? atof(env) : 10000.0;
        timeout_cycles = (long long)(ms * (khz > 0 ? (double)khz : 2.0e6));
    }
    return timeout_cycles;
}

extern "C" int nmb_peer_exchange_adamw_step(const nmb_peer_exchange_t* px, int dtype, const void* partials, int d, int H,
                                            int n_actions, const int32_t* seg_splits_host, const nmb_adamw_t* hp,
                                            void* theta, void* theta_t, void* exp_avg, void* exp_avg_sq,
                                            nmb_loop_state_t* state, int past_initial, int patience, void* stream) {
    int rc = validate_peers(px, "nmb_peer_exchange_adamw_step");
    if (rc) return rc;
    NMB_REQUIRE(partials && seg_splits_host && hp && theta && exp_avg && exp_avg_sq && state, NMB_EINVAL,
                "nmb_peer_exchange_adamw_step: null argument");
    SegTab tb = make_segtab(d, H, n_actions, seg_splits_host);
    NMB_REQUIRE(tb.off[S_COUNT] == px->n, NMB_EINVAL, "nmb_peer_exchange_adamw_step: exchange holds %lld elements, policy has %d",
                (long long)px->n, tb.off[S_COUNT]);
    const long long grid = (px->n + 255) / 256;
    // every CTA spins on its peers' flags: the whole grid has to be resident (8 CTAs of 256 threads per SM)
    NMB_REQUIRE(grid <= 8LL * sm_count(), NMB_ELIMIT,
                "nmb_peer_exchange_adamw_step: %lld slices do not fit the GPU at once (use publish / wait_decide / adamw_step)",
                grid);
    cudaStream_t st = (cudaStream_t)stream;
    const long long tmo = peer_timeout_cycles();
    if (dtype == NMB_F64) {
        peer_exchange_adamw_kernel<double><<<(unsigned)grid, 256, 0, st>>>(
            make_peertab<double>(px), tb, (const double*)partials, *hp, (double*)theta, (double*)theta_t, (double*)exp_avg,
            (double*)exp_avg_sq, state, past_initial, patience, tmo);
    } else if (dtype == NMB_F32) {
        peer_exchange_adamw_kernel<float><<<(unsigned)grid, 256, 0, st>>>(
            make_peertab<float>(px), tb, (const float*)partials, *hp, (float*)theta, (float*)theta_t, (float*)exp_avg,
            (float*)exp_avg_sq, state, past_initial, patience, tmo);
    } else { set_error("nmb_peer_exchange_adamw_step: dtype %d", dtype); return NMB_EINVAL; }
    NMB_LAUNCH_CHECK();
    return 0;
}
