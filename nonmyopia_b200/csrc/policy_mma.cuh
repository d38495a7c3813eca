// FP64 policy network on the FP64 tensor pipe (DMMA m8n8k4) with a TMA-fed weight stream.
//
// Why (measured on B200, tools/microbench.cu -> profiles/microbench_r01.jsonl):
//   * DFMA and mma.sync.m8n8k4.f64 both peak at 36.9 TFLOP/s, but a DMMA moves 256 FMAs per warp
//     instruction with 2 operand doubles per lane, whereas a 2x4 DFMA register tile needs three
//     LDS.128 (2-4 data-pipe cycles each, broadcast does not help) per 8 DFMAs: the DFMA form of these
//     32..64-row GEMMs is bound by the shared-memory data pipe at ~1/3 of the FP64 peak.
//   * the weight chunks were staged through registers (LDG -> STS -> __syncthreads) with the L2
//     latency exposed at the start of each of the 10 GEMMs of a policy step.
// Here a producer warp streams the step's weight chunks ([32 k-rows][H cols], row stride kSW = 68 doubles so
// the B fragments are bank-conflict free) with cp.async.bulk into a 4-stage mbarrier ring, running
// ahead across layers and steps; consumer warps own 8-row x 32-col tiles of the CTA's rows x 64 block.
// Activations are row-major [row][KS = H + 4] so A fragments are conflict free as well.
// Reference semantics: amortized_network.py:97-133.
#pragma once
#include "policy_dev.cuh"

namespace nmb {

#ifndef NMB_FWD_STAGES
#define NMB_FWD_STAGES 4
#endif
constexpr int kStages = NMB_FWD_STAGES;        // weight ring depth (forward rollout)
// Stage row stride (doubles).  A 64-bit shared load is served as two half-warp wavefronts (lanes 0-15, 16-31); the
// B fragment of lane l sits at (l & 3) * kSW + (l >> 2), so a half-warp touches k = 0..3 x n = 0..3 and needs
// kSW = 4 (mod 16) for 16 distinct 8-byte banks.  (72 = 8 mod 16 looked conflict-free for an arbitrary lane split but
// measured 2-way conflicts on every B load: ncu l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_ld, 1.7x the
// ideal wavefronts in all five tensor-pipe kernels.)
constexpr int kSW = 68;
constexpr int kStageElems = kKC * kSW;
constexpr int kMaxRowsMma = 56;   // 14 consumer warps + 1 producer warp = 480 threads
constexpr int kGemmsPerStep = 10;

// ---------------------------------------------------------------- PTX helpers
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(unsigned bar) {
    asm volatile("{\n .reg .b64 st;\n mbarrier.arrive.shared::cta.b64 st, [%0];\n}" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(unsigned bar, unsigned bytes) {
    asm volatile("{\n .reg .b64 st;\n mbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n}" ::"r"(bar), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity) {
    asm volatile(
        "{\n .reg .pred p;\n"
        "WAIT_%=:\n"
        " mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        " @p bra DONE_%=;\n"
        " bra WAIT_%=;\n"
        "DONE_%=:\n}" ::"r"(bar), "r"(parity)
        : "memory");
}
// global -> shared bulk copy (TMA, 1-D), completion counted in bytes on an mbarrier
__device__ __forceinline__ void bulk_g2s(unsigned dst, const void* src, unsigned bytes, unsigned bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
// barrier over the consumer warps only (id 1); the producer warp never joins it
// (id 1 by default; the forward rollout runs two row groups per CTA on ids 1 and 2)
__device__ __forceinline__ void consumer_sync(int nthr, int bar = 1) {
    asm volatile("bar.sync %1, %0;" ::"r"(nthr), "r"(bar) : "memory");
}

__device__ __forceinline__ void dmma884(double (&c)[2], double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c[0]), "+d"(c[1])
                 : "d"(a), "d"(b));
}

// ---------------------------------------------------------------- optional phase clocks (-DNMB_PHASE_CLOCKS)
// Cycle accounting of one consumer warp's timeline inside the forward rollout (tools/phase_clocks.py): 0 embedding,
// 1 GEMM loops (incl. waiting for weight chunks), 2 activation epilogues, 3 named barriers, 4 output layer, 5 rotation +
// Thompson path, 6 cost / stores / step tail, 7 input load.  Compiles to nothing in the product build.
#ifdef NMB_PHASE_CLOCKS
static __device__ unsigned long long g_phase_clk[16];
struct PhaseClk {
    long long t;
    unsigned long long acc[8];
    __device__ void start() { t = clock64(); for (int i = 0; i < 8; ++i) acc[i] = 0; }
    __device__ __forceinline__ void mark(int i) { const long long n = clock64(); acc[i] += (unsigned long long)(n - t); t = n; }
    __device__ void flush(bool who) {
        if (!who) return;
        for (int i = 0; i < 8; ++i) atomicAdd(&g_phase_clk[i], acc[i]);
        atomicAdd(&g_phase_clk[8], 1ULL);
    }
};
#else
struct PhaseClk {
    __device__ __forceinline__ void start() {}
    __device__ __forceinline__ void mark(int) {}
    __device__ __forceinline__ void flush(bool) {}
};
#endif

// ---------------------------------------------------------------- geometry / shared memory
// Shared-memory copy of the policy's vectors and thin matrices (embedding [d+1][H], output layers [H][d], every bias):
// the activation epilogues read them right after a GEMM, where a global load - even an L1 hit - sat on the critical
// path (ncu r02a: 17 % of the forward rollout's stall samples were long-scoreboard waits on these __ldg).
struct SmallPar {
    const double* base;
    int off[S_COUNT];
    __host__ __device__ static bool small(int seg) {
        return seg == S_WE || seg == S_BE || seg == S_BP || seg == S_BIH || seg == S_BHH || seg == S_AB0 || seg == S_AB3 ||
               seg == S_AW6 || seg == S_AB6 || seg == S_XB0 || seg == S_XB3 || seg == S_XW6 || seg == S_XB6;
    }
    __host__ __device__ static int elems(const PolicyLayout& lay) {
        int n = 0;
        for (int sg = 0; sg < S_COUNT; ++sg)
            if (small(sg)) n += (lay.off[sg + 1] - lay.off[sg] + 1) & ~1;      // even: double2 loads stay aligned
        return n;
    }
    __device__ void init(const PolicyLayout& lay, const double* smem) {
        base = smem;
        int n = 0;
        for (int sg = 0; sg < S_COUNT; ++sg) {
            off[sg] = n;
            if (small(sg)) n += (lay.off[sg + 1] - lay.off[sg] + 1) & ~1;
        }
    }
    __device__ void load(const PolicyLayout& lay, const double* __restrict__ theta_t, double* smem, int tid, int nthr) const {
        for (int sg = 0; sg < S_COUNT; ++sg) {
            if (!small(sg)) continue;
            const int n = lay.off[sg + 1] - lay.off[sg];
            for (int e = tid; e < n; e += nthr) smem[off[sg] + e] = __ldg(theta_t + lay.off[sg] + e);
        }
    }
    __device__ __forceinline__ const double* at(int seg) const { return base + off[seg]; }
};

// weight ring as seen by producer and consumers
struct MmaRing {
    double* stage;
    unsigned full0, empty0;     // shared addresses of the mbarrier arrays
    int KS;                     // activation row stride
};

struct MmaSmem {
    double *IN, *OUT, *B0, *B1, *HA, *HB, *stage, *cost_acc, *tab, *par;
    unsigned full0, empty0;     // shared addresses of the mbarrier arrays
    int rows, KS;
    __device__ void carve(double* base, int d, int H, int rows_, int nst = kStages) {
        rows = rows_;
        KS = H + 4;
        double* p = base;
        stage = p; p += nst * kStageElems;                 // 16-byte aligned bulk-copy destinations first
        tab = p; p += NMB_FAST_TAB_LEN;
        par = nullptr;
        B0 = p; p += rows * KS;
        B1 = p; p += rows * KS;
        HA = p; p += rows * KS;
        HB = p; p += rows * KS;
        IN = p; p += (d + 1) * rows;
        OUT = p; p += d * rows;
        cost_acc = p; p += rows;
        full0 = smem_u32(p);
        empty0 = full0 + 8 * nst;
    }
    // Row group [row_off, row_off + rows_g) of a CTA of rows_total rows: the groups' buffers lie one after the other
    // behind the shared weight ring, the mbarriers behind the last group (same footprint as carve(rows_total)).
    __device__ void carve_group(double* base, int d, int H, int rows_total, int row_off, int rows_g, int n_par,
                                int nst = kStages) {
        rows = rows_g;
        KS = H + 4;
        const int per_row = 4 * KS + 2 * d + 2;
        stage = base;
        tab = base + nst * kStageElems;
        par = tab + NMB_FAST_TAB_LEN;
        double* p = par + n_par + (size_t)row_off * per_row;
        B0 = p; p += rows * KS;
        B1 = p; p += rows * KS;
        HA = p; p += rows * KS;
        HB = p; p += rows * KS;
        IN = p; p += (d + 1) * rows;
        OUT = p; p += d * rows;
        cost_acc = p;
        full0 = smem_u32(par + n_par + (size_t)rows_total * per_row);
        empty0 = full0 + 8 * nst;
    }
    static size_t bytes(int d, int H, int rows, int nst = kStages) {
        PolicyLayout lay;
        lay.init(d, H, 1);
        return ((size_t)nst * kStageElems + NMB_FAST_TAB_LEN + SmallPar::elems(lay) + (size_t)4 * rows * (H + 4) +
                (size_t)(2 * d + 2) * rows + 2 * nst) * 8;
    }
    __device__ MmaRing ring() const { return MmaRing{stage, full0, empty0, KS}; }
};

// rows per CTA (multiple of 8, <= 56) minimising waves x rows for `n_rollouts` x N chains on `sms` SMs
inline int mma_rows_per_cta(long long N, int n_rollouts, int sms, int max_rows = kMaxRowsMma) {
    int best = 8;
    double best_cost = 1e300;
    for (int r = 8; r <= max_rows; r += 8) {
        const long long ctas = ((N + r - 1) / r) * n_rollouts;
        const long long waves = (ctas + sms - 1) / sms;
        // a CTA of r rows costs about (r + 8): fixed per-step latency of the serial layer chain
        const double cost = (double)waves * (r + 8);
        if (cost < best_cost - 1e-9) { best_cost = cost; best = r; }
    }
    return best;
}

// weight chunk source of GEMM g (0..9) of a policy step: element offset in theta_t and leading dimension
__device__ __forceinline__ void gemm_src(const PolicyLayout& lay, bool head_a, int g, int& off, int& ldw) {
    const int H = lay.H;
    const int sW0 = head_a ? S_AW0 : S_XW0, sW3 = head_a ? S_AW3 : S_XW3;
    switch (g) {
        case 0: off = lay.off[S_WP]; ldw = H; break;
        case 1: off = lay.off[S_WIH]; ldw = 3 * H; break;
        case 2: off = lay.off[S_WHH]; ldw = 3 * H; break;
        case 3: off = lay.off[S_WIH] + H; ldw = 3 * H; break;
        case 4: off = lay.off[S_WHH] + H; ldw = 3 * H; break;
        case 5: off = lay.off[S_WIH] + 2 * H; ldw = 3 * H; break;
        case 6: off = lay.off[S_WHH] + 2 * H; ldw = 3 * H; break;
        case 7: off = lay.off[sW0]; ldw = H; break;
        case 8: off = lay.off[sW0] + H * H; ldw = H; break;
        default: off = lay.off[sW3]; ldw = H; break;
    }
}

// weight chunk source of GEMM g (0..9) of a BACKWARD policy step: element offset in theta ([O][K] torch layout,
// contraction over O) and leading dimension
__device__ __forceinline__ void gemm_src_bwd(const PolicyLayout& lay, bool head_a, int g, int& off, int& ldw) {
    const int H = lay.H;
    const int sW0 = head_a ? S_AW0 : S_XW0, sW3 = head_a ? S_AW3 : S_XW3;
    switch (g) {
        case 0: off = lay.off[sW3]; ldw = H; break;
        case 1: off = lay.off[sW0]; ldw = 2 * H; break;
        case 2: off = lay.off[sW0] + H; ldw = 2 * H; break;
        case 3: off = lay.off[S_WIH]; ldw = H; break;
        case 4: off = lay.off[S_WIH] + H * H; ldw = H; break;
        case 5: off = lay.off[S_WIH] + 2 * H * H; ldw = H; break;
        case 6: off = lay.off[S_WHH]; ldw = H; break;
        case 7: off = lay.off[S_WHH] + H * H; ldw = H; break;
        case 8: off = lay.off[S_WHH] + 2 * H * H; ldw = H; break;
        default: off = lay.off[S_WP]; ldw = H; break;
    }
}

// Producer warp: stream every weight chunk of all L+1 steps in consumption order (forward: steps 0..L from
// theta_t; backward: steps L..0 from theta).
template <int NST, bool BWD, int KC = kKC>
__device__ __forceinline__ void mma_producer(const PolicyLayout& lay, const double* __restrict__ weights, int L,
                                             const MmaRing& rg, int it_begin = 0, int it_end = -1) {
    const int lane = threadIdx.x & 31;
    const int H = lay.H;
    const unsigned row_bytes = (unsigned)H * 8u;
    unsigned c = 0;
    if (it_end < 0) it_end = L + 1;
    for (int it = it_begin; it < it_end; ++it) {
        const int l = BWD ? L - it : it;
        for (int g = 0; g < kGemmsPerStep; ++g) {
            int off, ldw;
            if (BWD) gemm_src_bwd(lay, l == L, g, off, ldw);
            else gemm_src(lay, l == L, g, off, ldw);
            for (int k0 = 0; k0 < H; k0 += KC, ++c) {
                const unsigned slot = c % NST, fill = c / NST;
                if (fill > 0) mbar_wait(rg.empty0 + 8 * slot, (fill - 1) & 1);
                const int kn = min(KC, H - k0);
                const unsigned full = rg.full0 + 8 * slot;
                if (lane == 0) mbar_arrive_expect_tx(full, row_bytes * (unsigned)kn);
                __syncwarp();
                if (lane < kn)
                    bulk_g2s(smem_u32(rg.stage + (size_t)slot * (KC * kSW) + lane * kSW),
                             weights + off + (size_t)(k0 + lane) * ldw, row_bytes, full);
            }
        }
    }
}

// Consumer-side state of one warp
struct MmaWarp {
    int lane, row0, col0, nthr;
    int tid, bar;          // thread index inside the consumer group and the group's named barrier
    bool active;           // this warp's 32 columns intersect [0, H)
    unsigned chunk;        // running chunk counter (same order as the producer)
    int fr, fk;            // fragment coordinates: lane >> 2, lane & 3
    // nt = 8-column tiles per warp: 4 (two warps per 8-row block) or 2 (four warps per block)
    __device__ void init(int warp, int nthr_, int H, int nt = 4) {
        lane = threadIdx.x & 31;
        const int wpr = 8 / nt;
        row0 = (warp / wpr) * 8;
        col0 = (warp % wpr) * (8 * nt);
        nthr = nthr_;
        active = col0 < H;
        chunk = 0;
        tid = threadIdx.x;
        bar = 1;
        fr = lane >> 2;
        fk = lane & 3;
    }
};

// acc[t] (t = 0..3: 8-column tiles of the warp's 32 columns) += A[row0.., 0..H) . W chunk stream
// A: shared, row-major with stride KS, pointing at the first contraction column.
template <int NST, int NT = 4, int KC = kKC>
__device__ __forceinline__ void mma_gemm(const MmaRing& sm, MmaWarp& w, const double* __restrict__ A, int H,
                                         double (&acc)[NT][2]) {
    const double* arow = A + (size_t)(w.row0 + w.fr) * sm.KS + w.fk;
    for (int k0 = 0; k0 < H; k0 += KC) {
        const unsigned slot = w.chunk % NST, fill = w.chunk / NST;
        mbar_wait(sm.full0 + 8 * slot, fill & 1);
        if (w.active) {
            const double* st = sm.stage + (size_t)slot * (KC * kSW) + w.fk * kSW + w.col0 + w.fr;
            const int kn = min(KC, H - k0);
#pragma unroll 4
            for (int kk = 0; kk < kn; kk += 4) {
                const double a = arow[k0 + kk];
                double b[NT];
#pragma unroll
                for (int t = 0; t < NT; ++t) b[t] = st[kk * kSW + 8 * t];
#pragma unroll
                for (int t = 0; t < NT; ++t) dmma884(acc[t], a, b[t]);
            }
        }
        __syncwarp();
        if (w.lane == 0) mbar_arrive(sm.empty0 + 8 * slot);
        ++w.chunk;
    }
}

template <int NT> __device__ __forceinline__ void zero_acc4(double (&acc)[NT][2]) {
#pragma unroll
    for (int t = 0; t < NT; ++t) { acc[t][0] = 0.0; acc[t][1] = 0.0; }
}

// One policy step, forward, for the CTA's `rows` chains (consumer warps only).
//   in : sm.IN [k][rows] = [x; y];  hprev (sm.HA or sm.HB) = previous hidden state, row-major
//   out: sm.OUT [k][rows] = sigmoid output;  hnew (the other H buffer) = new hidden state
// B0 holds the embedding then u1, B1 holds p, u2 overwrites hprev (dead after the GRU).
template <int NST = kStages>
__device__ __forceinline__ void mma_policy_forward(const PolicyLayout& lay, const double* __restrict__ theta_t, bool head_a,
                                                   const MmaSmem& sm, MmaWarp& w, double* hprev, double* hnew,
                                                   const SaveAreas<double>* save, long long node0, int nvalid,
                                                   double* __restrict__ hs_out, PhaseClk& pc, const SmallPar& sp) {
    const int d = lay.d, H = lay.H, rows = sm.rows, KS = sm.KS;
    const int tid = w.tid, nthr = w.nthr;
    const MmaRing ring = sm.ring();
    const double* tab = sm.tab;
    auto nmb_elu_tab_ = [tab](double x) { return nmb_elu_tab(x, tab); };
    auto nmb_sigmoid_tab_ = [tab](double x) { return nmb_sigmoid_tab(x, tab); };
    auto nmb_tanh_tab_ = [tab](double x) { return nmb_tanh_tab(x, tab); };
    // ---- embedding (K = d+1): B0 = elu(We [x;y] + be)
    {
        const double* Wt = sp.at(S_WE);                  // [d+1][H]
        const double* b = sp.at(S_BE);
#pragma unroll 4
        for (int e = tid; e < rows * H; e += nthr) {
            const int r = e / H, o = e - r * H;
            double a = b[o];
            for (int k = 0; k <= d; ++k) a = fma(sm.IN[k * rows + r], Wt[k * H + o], a);
            const double v = nmb_elu_tab_(a);
            sm.B0[r * KS + o] = v;
            if (save && r < nvalid) save->EA[(node0 + r) * H + o] = v;
        }
        if (save) {
            for (int e = tid; e < rows * (d + 1); e += nthr) {
                const int r = e / (d + 1), k = e - r * (d + 1);
                if (r < nvalid) save->XY[(node0 + r) * (d + 1) + k] = sm.IN[k * rows + r];
            }
        }
    }
    pc.mark(0);
    consumer_sync(nthr, w.bar);
    pc.mark(3);
    const int row = w.row0 + w.fr;
    const bool rv = row < nvalid;
    const size_t nrow = (size_t)(node0 + row) * H;
    double acc[4][2];
    // ---- prepro: B1 = p = elu(Wp ea + bp)
    pc.mark(2);
    zero_acc4(acc);
    mma_gemm<NST>(ring, w, sm.B0, H, acc);
    pc.mark(1);
#pragma unroll
    for (int t = 0; t < 4; ++t) {
        const int o = w.col0 + 8 * t + 2 * w.fk;
        if (o < H) {
            const double2 bb = *reinterpret_cast<const double2*>(sp.at(S_BP) + o);
            const double2 v = make_double2(nmb_elu_tab_(acc[t][0] + bb.x), nmb_elu_tab_(acc[t][1] + bb.y));
            *reinterpret_cast<double2*>(sm.B1 + row * KS + o) = v;
            if (save && rv) *reinterpret_cast<double2*>(save->P + nrow + o) = v;
        }
    }
    pc.mark(2);
    consumer_sync(nthr, w.bar);
    pc.mark(3);
    // ---- GRU cell (gates r, z, n of the same (row, col) stay in registers)
    {
        double rg[4][2], zg[4][2], gh[4][2];
        const double* bih = sp.at(S_BIH);
        const double* bhh = sp.at(S_BHH);
        pc.mark(2);
        zero_acc4(acc);
        mma_gemm<NST>(ring, w, sm.B1, H, acc);
        mma_gemm<NST>(ring, w, hprev, H, acc);
        pc.mark(1);
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            const int o = min(w.col0 + 8 * t + 2 * w.fk, H - 2);
#pragma unroll
            for (int j = 0; j < 2; ++j) rg[t][j] = nmb_sigmoid_tab_(acc[t][j] + bih[o + j] + bhh[o + j]);
        }
        pc.mark(2);
        zero_acc4(acc);
        mma_gemm<NST>(ring, w, sm.B1, H, acc);
        mma_gemm<NST>(ring, w, hprev, H, acc);
        pc.mark(1);
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            const int o = min(w.col0 + 8 * t + 2 * w.fk, H - 2);
#pragma unroll
            for (int j = 0; j < 2; ++j) zg[t][j] = nmb_sigmoid_tab_(acc[t][j] + bih[H + o + j] + bhh[H + o + j]);
        }
        pc.mark(2);
        zero_acc4(acc);
        zero_acc4(gh);
        mma_gemm<NST>(ring, w, sm.B1, H, acc);
        mma_gemm<NST>(ring, w, hprev, H, gh);
        pc.mark(1);
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            const int o = w.col0 + 8 * t + 2 * w.fk;
            if (o < H) {
                double hn[2], ng[2], hv[2];
#pragma unroll
                for (int j = 0; j < 2; ++j) {
                    hn[j] = gh[t][j] + bhh[2 * H + o + j];
                    ng[j] = nmb_tanh_tab_(acc[t][j] + bih[2 * H + o + j] + rg[t][j] * hn[j]);
                    const double hp = hprev[row * KS + o + j];
                    hv[j] = (1.0 - zg[t][j]) * ng[j] + zg[t][j] * hp;
                }
                *reinterpret_cast<double2*>(hnew + row * KS + o) = make_double2(hv[0], hv[1]);
                if (rv) {
                    if (save) {
                        *reinterpret_cast<double2*>(save->RG + nrow + o) = make_double2(rg[t][0], rg[t][1]);
                        *reinterpret_cast<double2*>(save->ZG + nrow + o) = make_double2(zg[t][0], zg[t][1]);
                        *reinterpret_cast<double2*>(save->NG + nrow + o) = make_double2(ng[0], ng[1]);
                        *reinterpret_cast<double2*>(save->HN + nrow + o) = make_double2(hn[0], hn[1]);
                    }
                    if (hs_out) *reinterpret_cast<double2*>(hs_out + (size_t)row * H + o) = make_double2(hv[0], hv[1]);
                }
            }
        }
    }
    pc.mark(2);
    consumer_sync(nthr, w.bar);
    pc.mark(3);
    // ---- post MLP layer 0: B0 = u1 = elu(W0 [p; h'] + b0)
    const int sB0 = head_a ? S_AB0 : S_XB0, sB3 = head_a ? S_AB3 : S_XB3;
    const int sW6 = head_a ? S_AW6 : S_XW6, sB6 = head_a ? S_AB6 : S_XB6;
    pc.mark(2);
    zero_acc4(acc);
    mma_gemm<NST>(ring, w, sm.B1, H, acc);
    mma_gemm<NST>(ring, w, hnew, H, acc);
    pc.mark(1);
#pragma unroll
    for (int t = 0; t < 4; ++t) {
        const int o = w.col0 + 8 * t + 2 * w.fk;
        if (o < H) {
            const double2 bb = *reinterpret_cast<const double2*>(sp.at(sB0) + o);
            const double2 v = make_double2(nmb_elu_tab_(acc[t][0] + bb.x), nmb_elu_tab_(acc[t][1] + bb.y));
            *reinterpret_cast<double2*>(sm.B0 + row * KS + o) = v;
            if (save && rv) *reinterpret_cast<double2*>(save->U1 + nrow + o) = v;
        }
    }
    pc.mark(2);
    consumer_sync(nthr, w.bar);
    pc.mark(3);
    // ---- layer 3: u2 = elu(W3 u1 + b3) -> hprev buffer (dead)
    pc.mark(2);
    zero_acc4(acc);
    mma_gemm<NST>(ring, w, sm.B0, H, acc);
    pc.mark(1);
#pragma unroll
    for (int t = 0; t < 4; ++t) {
        const int o = w.col0 + 8 * t + 2 * w.fk;
        if (o < H) {
            const double2 bb = *reinterpret_cast<const double2*>(sp.at(sB3) + o);
            const double2 v = make_double2(nmb_elu_tab_(acc[t][0] + bb.x), nmb_elu_tab_(acc[t][1] + bb.y));
            *reinterpret_cast<double2*>(hprev + row * KS + o) = v;
            if (save && rv) *reinterpret_cast<double2*>(save->U2 + nrow + o) = v;
        }
    }
    pc.mark(2);
    consumer_sync(nthr, w.bar);
    pc.mark(3);
    // ---- output layer (O = d): out = sigmoid(W6 u2 + b6); 4 threads per output, 2 interleaved chains each
    {
        const double* Wt = sp.at(sW6);                   // [H][d]
        const double* b = sp.at(sB6);
        const int sub = tid & 3;
        for (int e = tid >> 2; e < ((rows * d + 7) & ~7); e += nthr >> 2) {
            const bool ok = e < rows * d;
            const int r = ok ? e / d : 0, o = ok ? e - r * d : 0;
            double a0 = 0.0, a1 = 0.0;
            const double* u = hprev + r * KS;
            for (int k = sub; k < H; k += 8) {
                a0 = fma(u[k], Wt[k * d + o], a0);
                if (k + 4 < H) a1 = fma(u[k + 4], Wt[(k + 4) * d + o], a1);
            }
            double a = a0 + a1;
            a += __shfl_xor_sync(0xffffffffu, a, 1);
            a += __shfl_xor_sync(0xffffffffu, a, 2);
            if (ok && sub == 0) {
                const double v = nmb_sigmoid_tab_(a + b[o]);
                sm.OUT[o * rows + r] = v;
                if (save && r < nvalid) save->OUT[(node0 + r) * d + o] = v;
            }
        }
    }
    pc.mark(4);
    consumer_sync(nthr, w.bar);
    pc.mark(3);
}

// ================================================================== backward step
constexpr int kStagesBwd = 5;          // 8 consumer warps drain a chunk in ~0.5 us: the ring must cover the L2 latency
// 8-row x 16-col warp tiles in the reverse rollout: 16 consumer warps for 32 rows instead of 8.  Two independent DMMA
// chains per warp already saturate the tensor pipe (dmma_chain micro-benchmark: 16.4 cycles per DMMA with 2 tiles); the
// reverse kernel was starved at 2.25 warps per scheduler (ncu r02a: DMMA 37 % active, `wait` + scoreboard stalls
// around the prefetched activations and the epilogues), not short of ILP inside a warp.
constexpr int kNTB = 2;
constexpr int kMaxRowsMmaBwd = 32;    // 7 row-major adjoint buffers (122 KB) + 5 weight stages (92 KB) in 227 KB

struct MmaBwdSmem {
    double *X0, *X1, *X2, *X3, *X4, *X5, *GH, *GO, *GXY, *carry, *stage;
    unsigned full0, empty0;
    int rows, KS;
    __device__ void carve(double* base, int d, int dmax, int H, int rows_, int nst = kStagesBwd, int kc = kKC) {
        rows = rows_;
        KS = H + 4;
        double* p = base;
        stage = p; p += nst * kc * kSW;
        X0 = p; p += rows * KS;
        X1 = p; p += rows * KS;
        X2 = p; p += rows * KS;
        X3 = p; p += rows * KS;
        X4 = p; p += rows * KS;
        X5 = p; p += rows * KS;
        GH = p; p += rows * KS;
        GO = p; p += d * rows;
        GXY = p; p += (d + 1) * rows;
        carry = p; p += (dmax + 1) * rows;
        full0 = smem_u32(p);
        empty0 = full0 + 8 * nst;
    }
    static size_t bytes(int d, int dmax, int H, int rows, int nst = kStagesBwd, int kc = kKC) {
        return ((size_t)nst * kc * kSW + (size_t)7 * rows * (H + 4) + (size_t)(2 * d + dmax + 2) * rows +
                2 * nst) * 8;
    }
    __device__ MmaRing ring() const { return MmaRing{stage, full0, empty0, KS}; }
};

// the lane's 4 x double2 slice of a node-major [node][H] save area (issued BEFORE the GEMM whose epilogue uses it,
// so the HBM/L2 latency overlaps the tensor-pipe work)
template <int NT>
__device__ __forceinline__ void prefetch_tile(double2 (&dst)[NT], const double* __restrict__ area, size_t nrow,
                                              const MmaWarp& w, int H, bool rv) {
#pragma unroll
    for (int t = 0; t < NT; ++t) {
        const int k = w.col0 + 8 * t + 2 * w.fk;
        dst[t] = (rv && k < H) ? *reinterpret_cast<const double2*>(area + nrow + k) : make_double2(0.0, 0.0);
    }
}

// One policy step backward for the CTA's rows (consumer warps only).
//   in : sm.GO [k][rows] adjoint of the sigmoid output;  sm.GH adjoint of the new hidden state (row-major)
//   out: sm.GXY [(d+1)][rows] adjoint of [x; y];  sm.GH adjoint of the previous hidden state
// Leaves the pre-activation adjoints (DO, DU2, DU1, DR, DZ, DN, DNH, DP, DE) of the node in the save
// workspace for the weight-gradient kernel.  theta: torch layout [O][K].
template <int NST = kStagesBwd, int KC = kKC>
__device__ __forceinline__ void mma_policy_backward(const PolicyLayout& lay, const double* __restrict__ theta, bool head_a,
                                                    const MmaBwdSmem& sm, MmaWarp& w, const SaveAreas<double>& sv,
                                                    long long node0, int nvalid) {
    const int d = lay.d, H = lay.H, rows = sm.rows, KS = sm.KS;
    const int tid = threadIdx.x, nthr = w.nthr;
    const MmaRing ring = sm.ring();
    const int sW6 = head_a ? S_AW6 : S_XW6;
    // ---- output layer: do_pre = go * o (1 - o)
    for (int e = tid; e < d * rows; e += nthr) {
        const int o = e / rows, r = e - o * rows;
        double v = 0.0;
        if (r < nvalid) {
            const double out = sv.OUT[(node0 + r) * d + o];
            v = sm.GO[o * rows + r] * out * (1.0 - out);
            sv.DO[(node0 + r) * d + o] = v;
        }
        sm.GO[o * rows + r] = v;
    }
    consumer_sync(nthr);
    // ---- du2_pre = (do_pre . W6) * elu'(u2)  (K = d) -> X0
    {
        const double* W = theta + lay.off[sW6];      // [d][H]
        for (int e0 = tid; e0 < rows * H; e0 += 4 * nthr) {
            double u2[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {      // 4 independent loads in flight
                const int e = e0 + i * nthr;
                const int r = e / H, k = e - r * H;
                u2[i] = (e < rows * H && r < nvalid) ? sv.U2[(node0 + r) * H + k] : 0.0;
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int e = e0 + i * nthr;
                if (e < rows * H) {
                    const int r = e / H, k = e - r * H;
                    double v = 0.0;
                    if (r < nvalid) {
                        double a = 0.0;
                        for (int o = 0; o < d; ++o) a = fma(sm.GO[o * rows + r], __ldg(W + o * H + k), a);
                        v = a * elu_grad_from_out(u2[i]);
                        sv.DU2[(node0 + r) * H + k] = v;
                    }
                    sm.X0[r * KS + k] = v;
                }
            }
        }
    }
    consumer_sync(nthr);
    const int row = w.row0 + w.fr;
    const bool rv = row < nvalid;
    const size_t nrow = (size_t)(node0 + row) * H;
    double acc[kNTB][2];
    // ---- du1_pre = (du2_pre . W3) * elu'(u1) -> X1
    double2 pf[kNTB];
    prefetch_tile(pf, sv.U1, nrow, w, H, rv);
    zero_acc4(acc);
    mma_gemm<NST, kNTB, KC>(ring, w, sm.X0, H, acc);
#pragma unroll
    for (int t = 0; t < kNTB; ++t) {
        const int k = w.col0 + 8 * t + 2 * w.fk;
        if (k < H) {
            double2 v = make_double2(0.0, 0.0);
            if (rv) {
                const double2 u = pf[t];
                v = make_double2(acc[t][0] * elu_grad_from_out(u.x), acc[t][1] * elu_grad_from_out(u.y));
                *reinterpret_cast<double2*>(sv.DU1 + nrow + k) = v;
            }
            *reinterpret_cast<double2*>(sm.X1 + row * KS + k) = v;
        }
    }
    consumer_sync(nthr);
    // ---- d[p; h'] = du1_pre . W0 : first H columns -> raw dp (X2), second H columns -> GRU backward
    double2 prg[kNTB], pzg[kNTB], png[kNTB], phn[kNTB], php[kNTB];
    prefetch_tile(prg, sv.RG, nrow, w, H, rv);
    prefetch_tile(pzg, sv.ZG, nrow, w, H, rv);
    prefetch_tile(png, sv.NG, nrow, w, H, rv);
    prefetch_tile(phn, sv.HN, nrow, w, H, rv);
    prefetch_tile(php, sv.HS, nrow, w, H, rv);      // HS[level] = h before the step
    zero_acc4(acc);
    mma_gemm<NST, kNTB, KC>(ring, w, sm.X1, H, acc);
#pragma unroll
    for (int t = 0; t < kNTB; ++t) {
        const int k = w.col0 + 8 * t + 2 * w.fk;
        if (k < H) *reinterpret_cast<double2*>(sm.X2 + row * KS + k) = make_double2(acc[t][0], acc[t][1]);
    }
    zero_acc4(acc);
    mma_gemm<NST, kNTB, KC>(ring, w, sm.X1, H, acc);
    // X0 (du2_pre), X3..X5 are dead for every warp here (barriers above / end of the previous step)
#pragma unroll
    for (int t = 0; t < kNTB; ++t) {
        const int k = w.col0 + 8 * t + 2 * w.fk;
        if (k < H) {
            double dr[2] = {0, 0}, dz[2] = {0, 0}, dn[2] = {0, 0}, dnh[2] = {0, 0}, dhp[2] = {0, 0};
            if (rv) {
                const double2 rg = prg[t], zg = pzg[t], ng = png[t], hn = phn[t], hp = php[t];
                const double2 gh = *reinterpret_cast<const double2*>(sm.GH + row * KS + k);
                const double rgv[2] = {rg.x, rg.y}, zgv[2] = {zg.x, zg.y}, ngv[2] = {ng.x, ng.y};
                const double hnv[2] = {hn.x, hn.y}, hpv[2] = {hp.x, hp.y}, ghv[2] = {gh.x, gh.y};
#pragma unroll
                for (int j = 0; j < 2; ++j) {
                    const double dh = acc[t][j] + ghv[j];
                    const double dng = dh * (1.0 - zgv[j]);
                    dn[j] = dng * (1.0 - ngv[j] * ngv[j]);
                    dz[j] = dh * (hpv[j] - ngv[j]) * zgv[j] * (1.0 - zgv[j]);
                    dr[j] = dn[j] * hnv[j] * rgv[j] * (1.0 - rgv[j]);
                    dnh[j] = dn[j] * rgv[j];
                    dhp[j] = dh * zgv[j];
                }
                *reinterpret_cast<double2*>(sv.DR + nrow + k) = make_double2(dr[0], dr[1]);
                *reinterpret_cast<double2*>(sv.DZ + nrow + k) = make_double2(dz[0], dz[1]);
                *reinterpret_cast<double2*>(sv.DN + nrow + k) = make_double2(dn[0], dn[1]);
                *reinterpret_cast<double2*>(sv.DNH + nrow + k) = make_double2(dnh[0], dnh[1]);
            }
            *reinterpret_cast<double2*>(sm.X0 + row * KS + k) = make_double2(dr[0], dr[1]);
            *reinterpret_cast<double2*>(sm.X3 + row * KS + k) = make_double2(dz[0], dz[1]);
            *reinterpret_cast<double2*>(sm.X4 + row * KS + k) = make_double2(dn[0], dn[1]);
            *reinterpret_cast<double2*>(sm.X5 + row * KS + k) = make_double2(dnh[0], dnh[1]);
            *reinterpret_cast<double2*>(sm.GH + row * KS + k) = make_double2(dhp[0], dhp[1]);   // own element only
        }
    }
    consumer_sync(nthr);
    // ---- dp_pre = (raw dp + dgi . Wih) * elu'(p) -> X2
    prefetch_tile(pf, sv.P, nrow, w, H, rv);
    double2 pea[kNTB];
    prefetch_tile(pea, sv.EA, nrow, w, H, rv);
    zero_acc4(acc);
    mma_gemm<NST, kNTB, KC>(ring, w, sm.X0, H, acc);
    mma_gemm<NST, kNTB, KC>(ring, w, sm.X3, H, acc);
    mma_gemm<NST, kNTB, KC>(ring, w, sm.X4, H, acc);
#pragma unroll
    for (int t = 0; t < kNTB; ++t) {
        const int k = w.col0 + 8 * t + 2 * w.fk;
        if (k < H) {
            double2 v = make_double2(0.0, 0.0);
            if (rv) {
                const double2 pp = pf[t];
                const double2 raw = *reinterpret_cast<const double2*>(sm.X2 + row * KS + k);
                v = make_double2((acc[t][0] + raw.x) * elu_grad_from_out(pp.x), (acc[t][1] + raw.y) * elu_grad_from_out(pp.y));
                *reinterpret_cast<double2*>(sv.DP + nrow + k) = v;
            }
            *reinterpret_cast<double2*>(sm.X2 + row * KS + k) = v;     // own element only
        }
    }
    // ---- dh_prev += dr . Whh_r + dz . Whh_z + dnh . Whh_n
    zero_acc4(acc);
    mma_gemm<NST, kNTB, KC>(ring, w, sm.X0, H, acc);
    mma_gemm<NST, kNTB, KC>(ring, w, sm.X3, H, acc);
    mma_gemm<NST, kNTB, KC>(ring, w, sm.X5, H, acc);
#pragma unroll
    for (int t = 0; t < kNTB; ++t) {
        const int k = w.col0 + 8 * t + 2 * w.fk;
        if (k < H) {
            double2* g = reinterpret_cast<double2*>(sm.GH + row * KS + k);
            const double2 o = *g;
            *g = make_double2(o.x + acc[t][0], o.y + acc[t][1]);
        }
    }
    consumer_sync(nthr);
    // ---- de_pre = (dp_pre . Wp) * elu'(ea) -> X1
    zero_acc4(acc);
    mma_gemm<NST, kNTB, KC>(ring, w, sm.X2, H, acc);
#pragma unroll
    for (int t = 0; t < kNTB; ++t) {
        const int k = w.col0 + 8 * t + 2 * w.fk;
        if (k < H) {
            double2 v = make_double2(0.0, 0.0);
            if (rv) {
                const double2 ea = pea[t];
                v = make_double2(acc[t][0] * elu_grad_from_out(ea.x), acc[t][1] * elu_grad_from_out(ea.y));
                *reinterpret_cast<double2*>(sv.DE + nrow + k) = v;
            }
            *reinterpret_cast<double2*>(sm.X1 + row * KS + k) = v;
        }
    }
    consumer_sync(nthr);
    // ---- d[x; y] = de_pre . We  (We is [H][d+1]); 4 threads per output
    {
        const double* W = theta + lay.off[S_WE];
        const int sub = tid & 3, nout = rows * (d + 1);
        for (int e = tid >> 2; e < ((nout + 7) & ~7); e += nthr >> 2) {
            const bool ok = e < nout;
            const int r = ok ? e / (d + 1) : 0, k = ok ? e - r * (d + 1) : 0;
            const double* g = sm.X1 + r * KS;
            double a0 = 0.0, a1 = 0.0;
            for (int o = sub; o < H; o += 8) {
                a0 = fma(g[o], __ldg(W + o * (d + 1) + k), a0);
                if (o + 4 < H) a1 = fma(g[o + 4], __ldg(W + (o + 4) * (d + 1) + k), a1);
            }
            double a = a0 + a1;
            a += __shfl_xor_sync(0xffffffffu, a, 1);
            a += __shfl_xor_sync(0xffffffffu, a, 2);
            if (ok && sub == 0) sm.GXY[k * rows + r] = a;
        }
    }
    consumer_sync(nthr);
}

}  // namespace nmb
