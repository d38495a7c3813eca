// FP32 variant of the policy network on the tensor cores: TF32 mma.sync.m16n8k8 with FP32 accumulation, weights
// streamed by the same TMA (cp.async.bulk) + mbarrier ring as the FP64 path (policy_mma.cuh).
//
// Why (round-2 evidence, north_star (b)/(d) "TF32/FP32 variant ... only if it pays off"):
//   * tools/microbench_tf32.cu on B200: FFMA 70 TFLOP/s, legacy warp-level mma.sync TF32 276 TFLOP/s, DMMA 37 TFLOP/s -
//     the warp-level TF32 MMA needs no 128-row tile (tcgen05 does) and fits the 16..64-row CTAs of the rollout;
//   * tools/tf32_tolerance.py: rounding every policy GEMM operand to TF32 moves acqf_loss by 1e-4 at BASELINE
//     configs[3] (tolerance 1e-3), same arg-min restart;
//   * the FP32 register-tile kernels of round 1 were SLOWER than the FP64 tensor-pipe path (0.74 vs 0.66 ms per Adam
//     iteration at 4096 chains on Cosine8).
// Only the K = H contractions run as TF32; the K = d + 1 embedding, the O = d output layer, every activation, the
// Thompson path, the rotation and the cost stay FP32 FMA / MUFU (phases, kernel distances and pivots must
// not lose mantissa).  Reference semantics: amortized_network.py:97-133.
#pragma once
#include <algorithm>
#include "policy_mma.cuh"

namespace nmb {

// Stage row stride in floats: the B fragment of lane l is (k = l & 3 [+4], n = l >> 2), so a warp touches 4 rows x 8
// columns per load: 72 = 8 (mod 32) puts them on 32 distinct banks.  The activation row stride H + 4 = 4 (mod 32)
// does the same for the A fragment (8 rows x 4 columns).
constexpr int kSWf = 72;
constexpr int kStageElemsF32 = kKC * kSWf;
constexpr int kStagesF32 = 4;
constexpr int kMaxRowsTf32 = 64;          // 16 consumer warps (16-row x 16-col tiles) + 1 producer warp
constexpr int kNTf = 2;                   // 8-column MMA tiles per warp: small tiles = more warps to hide the FP32 epilogues

__device__ __forceinline__ void mma_tf32_16x8x8(float (&c)[4], unsigned a0, unsigned a1, unsigned a2, unsigned a3,
                                                unsigned b0, unsigned b1) {
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
// round-to-nearest TF32 (the tolerance experiment rounds; raw FP32 bits would truncate)
__device__ __forceinline__ unsigned to_tf32(float x) {
    unsigned r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return r;
}

// FP32 activations on the MUFU unit (ex2 / rcp): ~2 ulp, far inside the 1e-3 tolerance of the FP32 variant; libm's
// expm1f / tanhf / sincospif cost 30-60 instructions with slow paths the network never needs
__device__ __forceinline__ float fast_sigmoid(float x) { return __fdividef(1.0f, 1.0f + __expf(-x)); }
__device__ __forceinline__ float fast_tanh(float x) {
    const float e = __expf(-2.0f * fabsf(x));
    return copysignf(__fdividef(1.0f - e, 1.0f + e), x);
}
__device__ __forceinline__ float fast_elu(float x) {
    // expm1 near zero: exp(x) - 1 loses digits in FP32, use the series for |x| < 1/8
    const float xm = fminf(x, 0.0f);
    const float em = xm > -0.125f ? xm * (1.0f + xm * (0.5f + xm * (0.16666667f + xm * 0.041666668f))) : __expf(xm) - 1.0f;
    return x > 0.0f ? x : em;
}

struct Tf32Ring {
    float* stage;
    unsigned full0, empty0;
    int KS;
};

struct Tf32Smem {
    float *IN, *OUT, *B0, *B1, *HA, *HB, *stage, *cost_acc, *par;
    unsigned full0, empty0;
    int rows, KS;
    __device__ void carve(float* base, int d, int H, int rows_, int n_par) {
        rows = rows_;
        KS = H + 4;
        float* p = base;
        stage = p; p += kStagesF32 * kStageElemsF32;            // 16-byte aligned bulk-copy destinations first
        par = p; p += (n_par + 3) & ~3;
        B0 = p; p += rows * KS;
        B1 = p; p += rows * KS;
        HA = p; p += rows * KS;
        HB = p; p += rows * KS;
        IN = p; p += (d + 1) * rows;
        OUT = p; p += d * rows;
        cost_acc = p; p += rows;
        p += ((size_t)(p - base) & 1);                           // mbarriers are 8-byte objects
        full0 = smem_u32(p);
        empty0 = full0 + 8 * kStagesF32;
    }
    static size_t bytes(int d, int H, int rows) {
        PolicyLayout lay;
        lay.init(d, H, 1);
        const size_t n_par = ((size_t)SmallPar::elems(lay) + 3) & ~(size_t)3;
        return ((size_t)kStagesF32 * kStageElemsF32 + n_par + (size_t)4 * rows * (H + 4) + (size_t)(2 * d + 2) * rows + 2) * 4 +
               16 * kStagesF32;
    }
    __device__ Tf32Ring ring() const { return Tf32Ring{stage, full0, empty0, KS}; }
};

// shared-memory copy of the small parameters (float twin of SmallPar)
struct SmallParF {
    const float* base;
    int off[S_COUNT];
    __device__ void init(const PolicyLayout& lay, const float* smem) {
        base = smem;
        int n = 0;
        for (int sg = 0; sg < S_COUNT; ++sg) {
            off[sg] = n;
            if (SmallPar::small(sg)) n += (lay.off[sg + 1] - lay.off[sg] + 1) & ~1;
        }
    }
    __device__ void load(const PolicyLayout& lay, const float* __restrict__ theta, float* smem, int tid, int nthr) const {
        for (int sg = 0; sg < S_COUNT; ++sg) {
            if (!SmallPar::small(sg)) continue;
            const int n = lay.off[sg + 1] - lay.off[sg];
            for (int e = tid; e < n; e += nthr) smem[off[sg] + e] = __ldg(theta + lay.off[sg] + e);
        }
    }
    __device__ __forceinline__ const float* at(int seg) const { return base + off[seg]; }
};

// rows per CTA (multiple of 16, <= 64) for `n_rollouts` x N chains: CTAs per SM follow from the shared-memory
// footprint and the thread count, cost = waves x (rows + fixed per-step latency)
template <typename F>
inline int tf32_rows_per_cta(long long N, int n_rollouts, int sms, F smem_bytes) {
    int best = 16;
    double best_cost = 1e300;
    for (int r = 16; r <= kMaxRowsTf32; r += 16) {
        const int per_sm = (int)std::max<size_t>(1, std::min<size_t>({(size_t)(220 * 1024) / (smem_bytes(r) + 1024),
                                                                       (size_t)2048 / (size_t)(r * 8 + 32), (size_t)4}));
        const long long ctas = ((N + r - 1) / r) * n_rollouts;
        const long long slots = (long long)sms * per_sm;
        const long long waves = (ctas + slots - 1) / slots;
        // CTAs sharing an SM share its pipes: a wave of per_sm CTAs costs about per_sm x rows
        const double cost = (double)waves * (per_sm * r + 16);
        if (cost < best_cost - 1e-9) { best_cost = cost; best = r; }
    }
    return best;
}

// Producer warp.  A float weight row starts at an 8-byte (not always 16-byte) boundary - the segments behind the
// d-element output bias are offset by d floats - so the bulk copy starts `shift` floats early (shift = off & 3, the
// same for every row of a GEMM because every leading dimension is a multiple of 4) and the consumers read their
// columns `shift` floats further right.
template <bool BWD>
__device__ __forceinline__ void tf32_producer(const PolicyLayout& lay, const float* __restrict__ weights, int L,
                                              const Tf32Ring& rg) {
    const int lane = threadIdx.x & 31;
    const int H = lay.H;
    unsigned c = 0;
    for (int it = 0; it <= L; ++it) {
        const int l = BWD ? L - it : it;
        for (int g = 0; g < kGemmsPerStep; ++g) {
            int off, ldw;
            if (BWD) gemm_src_bwd(lay, l == L, g, off, ldw);
            else gemm_src(lay, l == L, g, off, ldw);
            const int shift = off & 3;
            const unsigned row_bytes = (unsigned)(H + (shift ? 4 : 0)) * 4u;
            for (int k0 = 0; k0 < H; k0 += kKC, ++c) {
                const unsigned slot = c % kStagesF32, fill = c / kStagesF32;
                if (fill > 0) mbar_wait(rg.empty0 + 8 * slot, (fill - 1) & 1);
                const int kn = min(kKC, H - k0);
                const unsigned full = rg.full0 + 8 * slot;
                if (lane == 0) mbar_arrive_expect_tx(full, row_bytes * (unsigned)kn);
                __syncwarp();
                if (lane < kn)
                    bulk_g2s(smem_u32(rg.stage + (size_t)slot * kStageElemsF32 + lane * kSWf),
                             weights + off - shift + (size_t)(k0 + lane) * ldw, row_bytes, full);
            }
        }
    }
}

struct Tf32Warp {
    int lane, row0, col0, nthr, tid;
    bool active;
    unsigned chunk;
    int g, t4;              // fragment coordinates: lane >> 2, lane & 3
    __device__ void init(int warp, int nthr_, int H) {
        lane = threadIdx.x & 31;
        row0 = (warp >> 2) * 16;
        col0 = (warp & 3) * 16;
        nthr = nthr_;
        active = col0 < H;
        chunk = 0;
        tid = threadIdx.x;
        g = lane >> 2;
        t4 = lane & 3;
    }
};

// acc[t] (t = 0..3: 8-column tiles of the warp's 32 columns; c0,c1 = row g, c2,c3 = row g + 8) += A . W chunk stream.
// A: shared, row-major FP32 with stride KS, pointing at the first contraction column; `shift`: see tf32_producer.
__device__ __forceinline__ void tf32_gemm(const Tf32Ring& sm, Tf32Warp& w, const float* __restrict__ A, int H, int shift,
                                          float (&acc)[kNTf][4]) {
    const float* ar0 = A + (size_t)(w.row0 + w.g) * sm.KS + w.t4;
    const float* ar1 = ar0 + 8 * sm.KS;
    for (int k0 = 0; k0 < H; k0 += kKC) {
        const unsigned slot = w.chunk % kStagesF32, fill = w.chunk / kStagesF32;
        mbar_wait(sm.full0 + 8 * slot, fill & 1);
        if (w.active) {
            const float* st = sm.stage + (size_t)slot * kStageElemsF32 + w.t4 * kSWf + w.col0 + w.g + shift;
            const int kn = min(kKC, H - k0);
#pragma unroll 4
            for (int kk = 0; kk < kn; kk += 8) {
                const unsigned a0 = to_tf32(ar0[k0 + kk]), a1 = to_tf32(ar1[k0 + kk]);
                const unsigned a2 = to_tf32(ar0[k0 + kk + 4]), a3 = to_tf32(ar1[k0 + kk + 4]);
#pragma unroll
                for (int t = 0; t < kNTf; ++t) {
                    const unsigned b0 = to_tf32(st[kk * kSWf + 8 * t]), b1 = to_tf32(st[(kk + 4) * kSWf + 8 * t]);
                    mma_tf32_16x8x8(acc[t], a0, a1, a2, a3, b0, b1);
                }
            }
        }
        __syncwarp();
        if (w.lane == 0) mbar_arrive(sm.empty0 + 8 * slot);
        ++w.chunk;
    }
}

__device__ __forceinline__ void zero_acc_f(float (&acc)[kNTf][4]) {
#pragma unroll
    for (int t = 0; t < kNTf; ++t)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[t][j] = 0.0f;
}


// One policy step, forward, FP32 / TF32, for the CTA's `rows` chains (consumer warps only); the float twin of
// mma_policy_forward (same buffers, same save areas).  shift[g]: column shift of the g-th GEMM's staged rows.
__device__ __forceinline__ void tf32_policy_forward(const PolicyLayout& lay, bool head_a, const Tf32Smem& sm, Tf32Warp& w,
                                                    float* hprev, float* hnew, const SaveAreas<float>* save,
                                                    long long node0, int nvalid, float* __restrict__ hs_out,
                                                    const SmallParF& sp, const int (&shift)[kGemmsPerStep]) {
    const int d = lay.d, H = lay.H, rows = sm.rows, KS = sm.KS;
    const int tid = w.tid, nthr = w.nthr;
    const Tf32Ring ring = sm.ring();
    // ---- embedding (K = d + 1): B0 = elu(We [x; y] + be)
    {
        const float* Wt = sp.at(S_WE);
        const float* b = sp.at(S_BE);
#pragma unroll 4
        for (int e = tid; e < rows * H; e += nthr) {
            const int r = e / H, o = e - r * H;
            float a = b[o];
            for (int k = 0; k <= d; ++k) a = fmaf(sm.IN[k * rows + r], Wt[k * H + o], a);
            const float v = fast_elu(a);
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
    consumer_sync(nthr);
    float acc[kNTf][4];
    // epilogue helper: element (tile t, row half hf, column j) -> (row, col)
#define NMB_TF32_FOREACH(...)                                               \
    _Pragma("unroll") for (int t = 0; t < kNTf; ++t) {                         \
        const int o = w.col0 + 8 * t + 2 * w.t4;                            \
        if (o < H) {                                                        \
            _Pragma("unroll") for (int hf = 0; hf < 2; ++hf) {              \
                const int row = w.row0 + w.g + 8 * hf;                      \
                const bool rv = row < nvalid;                               \
                const size_t nrow = (size_t)(node0 + row) * H;              \
                __VA_ARGS__                                                 \
            }                                                               \
        }                                                                   \
    }
    // ---- prepro: B1 = p = elu(Wp ea + bp)
    zero_acc_f(acc);
    tf32_gemm(ring, w, sm.B0, H, shift[0], acc);
    NMB_TF32_FOREACH({
        const float2 bb = *reinterpret_cast<const float2*>(sp.at(S_BP) + o);
        const float2 v = make_float2(fast_elu(acc[t][2 * hf] + bb.x), fast_elu(acc[t][2 * hf + 1] + bb.y));
        *reinterpret_cast<float2*>(sm.B1 + row * KS + o) = v;
        if (save && rv) *reinterpret_cast<float2*>(save->P + nrow + o) = v;
    })
    consumer_sync(nthr);
    // ---- GRU cell
    {
        float rg[kNTf][4], zg[kNTf][4], gh[kNTf][4];
        const float* bih = sp.at(S_BIH);
        const float* bhh = sp.at(S_BHH);
        zero_acc_f(acc);
        tf32_gemm(ring, w, sm.B1, H, shift[1], acc);
        tf32_gemm(ring, w, hprev, H, shift[2], acc);
#pragma unroll
        for (int t = 0; t < kNTf; ++t) {
            const int o = min(w.col0 + 8 * t + 2 * w.t4, H - 2);
#pragma unroll
            for (int q = 0; q < 4; ++q) rg[t][q] = fast_sigmoid(acc[t][q] + bih[o + (q & 1)] + bhh[o + (q & 1)]);
        }
        zero_acc_f(acc);
        tf32_gemm(ring, w, sm.B1, H, shift[3], acc);
        tf32_gemm(ring, w, hprev, H, shift[4], acc);
#pragma unroll
        for (int t = 0; t < kNTf; ++t) {
            const int o = min(w.col0 + 8 * t + 2 * w.t4, H - 2);
#pragma unroll
            for (int q = 0; q < 4; ++q) zg[t][q] = fast_sigmoid(acc[t][q] + bih[H + o + (q & 1)] + bhh[H + o + (q & 1)]);
        }
        zero_acc_f(acc);
        zero_acc_f(gh);
        tf32_gemm(ring, w, sm.B1, H, shift[5], acc);
        tf32_gemm(ring, w, hprev, H, shift[6], gh);
        NMB_TF32_FOREACH({
            float hn[2];
            float ng[2];
            float hv[2];
            const float2 hp = *reinterpret_cast<const float2*>(hprev + row * KS + o);
            const float hpv[2] = {hp.x, hp.y};
            _Pragma("unroll") for (int j = 0; j < 2; ++j) {
                const int q = 2 * hf + j;
                hn[j] = gh[t][q] + bhh[2 * H + o + j];
                ng[j] = fast_tanh(acc[t][q] + bih[2 * H + o + j] + rg[t][q] * hn[j]);
                hv[j] = (1.0f - zg[t][q]) * ng[j] + zg[t][q] * hpv[j];
            }
            *reinterpret_cast<float2*>(hnew + row * KS + o) = make_float2(hv[0], hv[1]);
            if (rv) {
                if (save) {
                    *reinterpret_cast<float2*>(save->RG + nrow + o) = make_float2(rg[t][2 * hf], rg[t][2 * hf + 1]);
                    *reinterpret_cast<float2*>(save->ZG + nrow + o) = make_float2(zg[t][2 * hf], zg[t][2 * hf + 1]);
                    *reinterpret_cast<float2*>(save->NG + nrow + o) = make_float2(ng[0], ng[1]);
                    *reinterpret_cast<float2*>(save->HN + nrow + o) = make_float2(hn[0], hn[1]);
                }
                if (hs_out) *reinterpret_cast<float2*>(hs_out + (size_t)row * H + o) = make_float2(hv[0], hv[1]);
            }
        })
    }
    consumer_sync(nthr);
    // ---- post MLP layer 0: B0 = u1 = elu(W0 [p; h'] + b0)
    const int sB0 = head_a ? S_AB0 : S_XB0, sB3 = head_a ? S_AB3 : S_XB3;
    const int sW6 = head_a ? S_AW6 : S_XW6, sB6 = head_a ? S_AB6 : S_XB6;
    zero_acc_f(acc);
    tf32_gemm(ring, w, sm.B1, H, shift[7], acc);
    tf32_gemm(ring, w, hnew, H, shift[8], acc);
    NMB_TF32_FOREACH({
        const float2 bb = *reinterpret_cast<const float2*>(sp.at(sB0) + o);
        const float2 v = make_float2(fast_elu(acc[t][2 * hf] + bb.x), fast_elu(acc[t][2 * hf + 1] + bb.y));
        *reinterpret_cast<float2*>(sm.B0 + row * KS + o) = v;
        if (save && rv) *reinterpret_cast<float2*>(save->U1 + nrow + o) = v;
    })
    consumer_sync(nthr);
    // ---- layer 3: u2 = elu(W3 u1 + b3) -> hprev buffer (dead)
    zero_acc_f(acc);
    tf32_gemm(ring, w, sm.B0, H, shift[9], acc);
    NMB_TF32_FOREACH({
        const float2 bb = *reinterpret_cast<const float2*>(sp.at(sB3) + o);
        const float2 v = make_float2(fast_elu(acc[t][2 * hf] + bb.x), fast_elu(acc[t][2 * hf + 1] + bb.y));
        *reinterpret_cast<float2*>(hprev + row * KS + o) = v;
        if (save && rv) *reinterpret_cast<float2*>(save->U2 + nrow + o) = v;
    })
#undef NMB_TF32_FOREACH
    consumer_sync(nthr);
    // ---- output layer (O = d): out = sigmoid(W6 u2 + b6); 4 threads per output
    {
        const float* Wt = sp.at(sW6);
        const float* b = sp.at(sB6);
        const int sub = tid & 3;
        for (int e = tid >> 2; e < ((rows * d + 7) & ~7); e += nthr >> 2) {
            const bool ok = e < rows * d;
            const int r = ok ? e / d : 0, o = ok ? e - r * d : 0;
            float a0 = 0.0f, a1 = 0.0f;
            const float* u = hprev + r * KS;
            for (int k = sub; k < H; k += 8) {
                a0 = fmaf(u[k], Wt[k * d + o], a0);
                if (k + 4 < H) a1 = fmaf(u[k + 4], Wt[(k + 4) * d + o], a1);
            }
            float a = a0 + a1;
            a += __shfl_xor_sync(0xffffffffu, a, 1);
            a += __shfl_xor_sync(0xffffffffu, a, 2);
            if (ok && sub == 0) {
                const float v = fast_sigmoid(a + b[o]);
                sm.OUT[o * rows + r] = v;
                if (save && r < nvalid) save->OUT[(node0 + r) * d + o] = v;
            }
        }
    }
    consumer_sync(nthr);
}


// ================================================================== backward step (float twin of mma_policy_backward)
struct Tf32BwdSmem {
    float *X0, *X1, *X2, *X3, *X4, *X5, *GH, *GO, *GXY, *carry, *stage;
    unsigned full0, empty0;
    int rows, KS;
    __device__ void carve(float* base, int d, int dmax, int H, int rows_) {
        rows = rows_;
        KS = H + 4;
        float* p = base;
        stage = p; p += kStagesF32 * kStageElemsF32;
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
        p += ((size_t)(p - base) & 1);
        full0 = smem_u32(p);
        empty0 = full0 + 8 * kStagesF32;
    }
    static size_t bytes(int d, int dmax, int H, int rows) {
        return ((size_t)kStagesF32 * kStageElemsF32 + (size_t)7 * rows * (H + 4) + (size_t)(2 * d + dmax + 2) * rows + 2) * 4 +
               16 * kStagesF32;
    }
    __device__ Tf32Ring ring() const { return Tf32Ring{stage, full0, empty0, KS}; }
};

//   in : sm.GO [k][rows] adjoint of the sigmoid output;  sm.GH adjoint of the new hidden state (row-major)
//   out: sm.GXY [(d+1)][rows] adjoint of [x; y];  sm.GH adjoint of the previous hidden state
// Leaves the pre-activation adjoints of the node in the save workspace for the weight-gradient kernel.
__device__ __forceinline__ void tf32_policy_backward(const PolicyLayout& lay, const float* __restrict__ theta, bool head_a,
                                                     const Tf32BwdSmem& sm, Tf32Warp& w, const SaveAreas<float>& sv,
                                                     long long node0, int nvalid, const int (&shift)[kGemmsPerStep]) {
    const int d = lay.d, H = lay.H, rows = sm.rows, KS = sm.KS;
    const int tid = w.tid, nthr = w.nthr;
    const Tf32Ring ring = sm.ring();
    const int sW6 = head_a ? S_AW6 : S_XW6;
    // ---- output layer: do_pre = go * o (1 - o)
    for (int e = tid; e < d * rows; e += nthr) {
        const int o = e / rows, r = e - o * rows;
        float v = 0.0f;
        if (r < nvalid) {
            const float out = sv.OUT[(node0 + r) * d + o];
            v = sm.GO[o * rows + r] * out * (1.0f - out);
            sv.DO[(node0 + r) * d + o] = v;
        }
        sm.GO[o * rows + r] = v;
    }
    consumer_sync(nthr);
    // ---- du2_pre = (do_pre . W6) * elu'(u2)  (K = d) -> X0
    {
        const float* W = theta + lay.off[sW6];      // [d][H]
#pragma unroll 4
        for (int e = tid; e < rows * H; e += nthr) {
            const int r = e / H, k = e - r * H;
            float v = 0.0f;
            if (r < nvalid) {
                float a = 0.0f;
                for (int o = 0; o < d; ++o) a = fmaf(sm.GO[o * rows + r], __ldg(W + o * H + k), a);
                v = a * elu_grad_from_out(sv.U2[(node0 + r) * H + k]);
                sv.DU2[(node0 + r) * H + k] = v;
            }
            sm.X0[r * KS + k] = v;
        }
    }
    consumer_sync(nthr);
    float acc[kNTf][4];
#define NMB_TF32_FOREACH(...)                                               \
    _Pragma("unroll") for (int t = 0; t < kNTf; ++t) {                         \
        const int k = w.col0 + 8 * t + 2 * w.t4;                            \
        if (k < H) {                                                        \
            _Pragma("unroll") for (int hf = 0; hf < 2; ++hf) {              \
                const int row = w.row0 + w.g + 8 * hf;                      \
                const bool rv = row < nvalid;                               \
                const size_t nrow = (size_t)(node0 + row) * H;              \
                const float a0 = acc[t][2 * hf], a1 = acc[t][2 * hf + 1];   \
                __VA_ARGS__                                                 \
            }                                                               \
        }                                                                   \
    }
#define NMB_LD2(area) (rv ? *reinterpret_cast<const float2*>((area) + nrow + k) : make_float2(0.0f, 0.0f))
    // ---- du1_pre = (du2_pre . W3) * elu'(u1) -> X1
    zero_acc_f(acc);
    tf32_gemm(ring, w, sm.X0, H, shift[0], acc);
    NMB_TF32_FOREACH({
        float2 v = make_float2(0.0f, 0.0f);
        if (rv) {
            const float2 u = NMB_LD2(sv.U1);
            v = make_float2(a0 * elu_grad_from_out(u.x), a1 * elu_grad_from_out(u.y));
            *reinterpret_cast<float2*>(sv.DU1 + nrow + k) = v;
        }
        *reinterpret_cast<float2*>(sm.X1 + row * KS + k) = v;
    })
    consumer_sync(nthr);
    // ---- d[p; h'] = du1_pre . W0 : first H columns -> raw dp (X2), second H columns -> GRU backward
    zero_acc_f(acc);
    tf32_gemm(ring, w, sm.X1, H, shift[1], acc);
    NMB_TF32_FOREACH({ *reinterpret_cast<float2*>(sm.X2 + row * KS + k) = make_float2(a0, a1); (void)rv; (void)nrow; })
    zero_acc_f(acc);
    tf32_gemm(ring, w, sm.X1, H, shift[2], acc);
    NMB_TF32_FOREACH({
        float dr[2] = {0, 0};
        float dz[2] = {0, 0};
        float dn[2] = {0, 0};
        float dnh[2] = {0, 0};
        float dhp[2] = {0, 0};
        if (rv) {
            const float2 rg = NMB_LD2(sv.RG);
            const float2 zg = NMB_LD2(sv.ZG);
            const float2 ng = NMB_LD2(sv.NG);
            const float2 hn = NMB_LD2(sv.HN);
            const float2 hp = NMB_LD2(sv.HS);          // HS[level] = h before the step
            const float2 gh = *reinterpret_cast<const float2*>(sm.GH + row * KS + k);
            const float rgv[2] = {rg.x, rg.y};
            const float zgv[2] = {zg.x, zg.y};
            const float ngv[2] = {ng.x, ng.y};
            const float hnv[2] = {hn.x, hn.y};
            const float hpv[2] = {hp.x, hp.y};
            const float ghv[2] = {gh.x, gh.y};
            const float av[2] = {a0, a1};
            _Pragma("unroll") for (int j = 0; j < 2; ++j) {
                const float dh = av[j] + ghv[j];
                const float dng = dh * (1.0f - zgv[j]);
                dn[j] = dng * (1.0f - ngv[j] * ngv[j]);
                dz[j] = dh * (hpv[j] - ngv[j]) * zgv[j] * (1.0f - zgv[j]);
                dr[j] = dn[j] * hnv[j] * rgv[j] * (1.0f - rgv[j]);
                dnh[j] = dn[j] * rgv[j];
                dhp[j] = dh * zgv[j];
            }
            *reinterpret_cast<float2*>(sv.DR + nrow + k) = make_float2(dr[0], dr[1]);
            *reinterpret_cast<float2*>(sv.DZ + nrow + k) = make_float2(dz[0], dz[1]);
            *reinterpret_cast<float2*>(sv.DN + nrow + k) = make_float2(dn[0], dn[1]);
            *reinterpret_cast<float2*>(sv.DNH + nrow + k) = make_float2(dnh[0], dnh[1]);
        }
        *reinterpret_cast<float2*>(sm.X0 + row * KS + k) = make_float2(dr[0], dr[1]);
        *reinterpret_cast<float2*>(sm.X3 + row * KS + k) = make_float2(dz[0], dz[1]);
        *reinterpret_cast<float2*>(sm.X4 + row * KS + k) = make_float2(dn[0], dn[1]);
        *reinterpret_cast<float2*>(sm.X5 + row * KS + k) = make_float2(dnh[0], dnh[1]);
        *reinterpret_cast<float2*>(sm.GH + row * KS + k) = make_float2(dhp[0], dhp[1]);   // own element only
    })
    consumer_sync(nthr);
    // ---- dp_pre = (raw dp + dgi . Wih) * elu'(p) -> X2
    zero_acc_f(acc);
    tf32_gemm(ring, w, sm.X0, H, shift[3], acc);
    tf32_gemm(ring, w, sm.X3, H, shift[4], acc);
    tf32_gemm(ring, w, sm.X4, H, shift[5], acc);
    NMB_TF32_FOREACH({
        float2 v = make_float2(0.0f, 0.0f);
        if (rv) {
            const float2 pp = NMB_LD2(sv.P);
            const float2 raw = *reinterpret_cast<const float2*>(sm.X2 + row * KS + k);
            v = make_float2((a0 + raw.x) * elu_grad_from_out(pp.x), (a1 + raw.y) * elu_grad_from_out(pp.y));
            *reinterpret_cast<float2*>(sv.DP + nrow + k) = v;
        }
        *reinterpret_cast<float2*>(sm.X2 + row * KS + k) = v;     // own element only
    })
    // ---- dh_prev += dr . Whh_r + dz . Whh_z + dnh . Whh_n
    zero_acc_f(acc);
    tf32_gemm(ring, w, sm.X0, H, shift[6], acc);
    tf32_gemm(ring, w, sm.X3, H, shift[7], acc);
    tf32_gemm(ring, w, sm.X5, H, shift[8], acc);
    NMB_TF32_FOREACH({
        float2* gp = reinterpret_cast<float2*>(sm.GH + row * KS + k);
        const float2 o = *gp;
        *gp = make_float2(o.x + a0, o.y + a1);
        (void)rv; (void)nrow;
    })
    consumer_sync(nthr);
    // ---- de_pre = (dp_pre . Wp) * elu'(ea) -> X1
    zero_acc_f(acc);
    tf32_gemm(ring, w, sm.X2, H, shift[9], acc);
    NMB_TF32_FOREACH({
        float2 v = make_float2(0.0f, 0.0f);
        if (rv) {
            const float2 ea = NMB_LD2(sv.EA);
            v = make_float2(a0 * elu_grad_from_out(ea.x), a1 * elu_grad_from_out(ea.y));
            *reinterpret_cast<float2*>(sv.DE + nrow + k) = v;
        }
        *reinterpret_cast<float2*>(sm.X1 + row * KS + k) = v;
    })
#undef NMB_LD2
#undef NMB_TF32_FOREACH
    consumer_sync(nthr);
    // ---- d[x; y] = de_pre . We  (We is [H][d+1]); 4 threads per output
    {
        const float* W = theta + lay.off[S_WE];
        const int sub = tid & 3, nout = rows * (d + 1);
        for (int e = tid >> 2; e < ((nout + 7) & ~7); e += nthr >> 2) {
            const bool ok = e < nout;
            const int r = ok ? e / (d + 1) : 0, k = ok ? e - r * (d + 1) : 0;
            const float* g = sm.X1 + r * KS;
            float a0 = 0.0f, a1 = 0.0f;
            for (int o = sub; o < H; o += 8) {
                a0 = fmaf(g[o], __ldg(W + o * (d + 1) + k), a0);
                if (o + 4 < H) a1 = fmaf(g[o + 4], __ldg(W + (o + 4) * (d + 1) + k), a1);
            }
            float a = a0 + a1;
            a += __shfl_xor_sync(0xffffffffu, a, 1);
            a += __shfl_xor_sync(0xffffffffu, a, 2);
            if (ok && sub == 0) sm.GXY[k * rows + r] = a;
        }
    }
    consumer_sync(nthr);
}

}  // namespace nmb
