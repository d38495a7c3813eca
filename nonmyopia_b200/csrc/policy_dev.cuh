// Amortized policy network (reference amortized_network.py:97-133) as CTA-level device
// code: one CTA of 256 threads carries a tile of R = 32 chains through one policy step,
// forward or backward.  Activations live in shared memory in k-major layout
// (buf[k * RS + r], r = chain within the tile) so that every layer is a small GEMM
//     Y[r][o] = sum_k A[r][k] * Wt[k][o]
// with the weight operand streamed from global/L2 in k-chunks through a double-buffered
// shared-memory stage.  Thread tile 2 rows x 4 cols, warp tile 16 x 16, CTA tile 32 x 64:
// per k-step a warp issues 8 FMAs per lane against three 16-byte shared loads (A pair
// broadcast over 4 lanes, W quad broadcast over 8 lanes) -> FP64-pipe bound, not LDS bound.
//
// FP64 has no tcgen05 kind (f16/tf32/f8f6f4/i8/mx only), so the FP64 product path is DFMA.
#pragma once
#include "common.cuh"

namespace nmb {

constexpr int kR = 32;        // chains per CTA tile
constexpr int kRS = 32;       // row stride of k-major smem buffers
constexpr int kNT = 256;      // threads per CTA
constexpr int kKC = 32;       // k rows per staged weight chunk
constexpr int kOB = 64;       // output columns per GEMM block
constexpr int kMaxH = 64;

// ---- parameter layout (reference state_dict order; torch [O][K] row-major)
enum Seg {
    S_WE = 0, S_BE, S_WP, S_BP, S_WIH, S_WHH, S_BIH, S_BHH,
    S_AW0, S_AB0, S_AW3, S_AB3, S_AW6, S_AB6,
    S_XW0, S_XB0, S_XW3, S_XB3, S_XW6, S_XB6, S_COUNT
};

struct PolicyLayout {
    int d, H, na;
    int off[S_COUNT + 1];
    int O[S_COUNT], K[S_COUNT];   // K == 0 for bias vectors
    __host__ __device__ void init(int d_, int H_, int na_) {
        d = d_; H = H_; na = na_;
        const int o_[S_COUNT] = {H, H, H, H, 3 * H, 3 * H, 3 * H, 3 * H, H, H, H, H, d * na, d * na, H, H, H, H, d, d};
        const int k_[S_COUNT] = {d + 1, 0, H, 0, H, H, 0, 0, 2 * H, 0, H, 0, H, 0, 2 * H, 0, H, 0, H, 0};
        int acc = 0;
        for (int s = 0; s < S_COUNT; ++s) {
            O[s] = o_[s]; K[s] = k_[s];
            off[s] = acc;
            acc += o_[s] * (k_[s] ? k_[s] : 1);
        }
        off[S_COUNT] = acc;
    }
    __host__ __device__ int total() const { return off[S_COUNT]; }
};

// ---- thread coordinates inside the 32 x 64 CTA tile
struct TileCoord {
    int r0;   // first of 2 rows
    int c0;   // first of 4 cols inside the 64-col block
    __device__ TileCoord() {
        const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
        r0 = ((warp >> 2) << 4) + ((lane >> 2) << 1);
        c0 = ((warp & 3) << 4) + ((lane & 3) << 2);
    }
};

__device__ __forceinline__ void ld2(const double* p, double& a, double& b) {
    const double2 v = *reinterpret_cast<const double2*>(p); a = v.x; b = v.y;
}
__device__ __forceinline__ void ld2(const float* p, float& a, float& b) {
    const float2 v = *reinterpret_cast<const float2*>(p); a = v.x; b = v.y;
}
__device__ __forceinline__ void ld4(const double* p, double& a, double& b, double& c, double& d) {
    const double2 v0 = reinterpret_cast<const double2*>(p)[0], v1 = reinterpret_cast<const double2*>(p)[1];
    a = v0.x; b = v0.y; c = v1.x; d = v1.y;
}
__device__ __forceinline__ void ld4(const float* p, float& a, float& b, float& c, float& d) {
    const float4 v = *reinterpret_cast<const float4*>(p); a = v.x; b = v.y; c = v.z; d = v.w;
}

// acc[2][4] += sum_{k < K} A[k*kRS + r] * Wg[k*ldw + cb + c],  c in [0,64) masked by ncols.
// A: shared, k-major.  Wg: global, k-major with leading dim ldw.  stage: 2*kKC*kOB elements.
// All 256 threads must call it (contains __syncthreads); A must be visible (synced) before.
template <typename T>
__device__ __forceinline__ void gemm_block(const T* __restrict__ A, int K, const T* __restrict__ Wg, int ldw, int cb,
                                           int ncols, T* __restrict__ stage, const TileCoord& tc, T (&acc)[2][4]) {
    constexpr int PER = kKC * kOB / kNT;   // 8 elements per thread per chunk
    const int tid = threadIdx.x;
    T reg[PER];
    auto fetch = [&](int k0) {
#pragma unroll
        for (int i = 0; i < PER; ++i) {
            const int e = tid + i * kNT;
            const int kk = e >> 6, c = e & 63;
            const int k = k0 + kk;
            reg[i] = (k < K && c < ncols) ? __ldg(Wg + (size_t)k * ldw + cb + c) : T(0);
        }
    };
    auto put = [&](T* buf) {
#pragma unroll
        for (int i = 0; i < PER; ++i) buf[tid + i * kNT] = reg[i];
    };
    const int nch = (K + kKC - 1) / kKC;
    fetch(0);
    put(stage);
    __syncthreads();
    for (int ch = 0; ch < nch; ++ch) {
        const T* cur = stage + (ch & 1) * (kKC * kOB);
        if (ch + 1 < nch) fetch((ch + 1) * kKC);
        const int k0 = ch * kKC;
        const int kn = min(kKC, K - k0);
        const T* a = A + (size_t)k0 * kRS + tc.r0;
        const T* w = cur + tc.c0;
#pragma unroll 8
        for (int kk = 0; kk < kn; ++kk) {
            T a0, a1, w0, w1, w2, w3;
            ld2(a + kk * kRS, a0, a1);
            ld4(w + kk * kOB, w0, w1, w2, w3);
            acc[0][0] = fma(a0, w0, acc[0][0]); acc[0][1] = fma(a0, w1, acc[0][1]);
            acc[0][2] = fma(a0, w2, acc[0][2]); acc[0][3] = fma(a0, w3, acc[0][3]);
            acc[1][0] = fma(a1, w0, acc[1][0]); acc[1][1] = fma(a1, w1, acc[1][1]);
            acc[1][2] = fma(a1, w2, acc[1][2]); acc[1][3] = fma(a1, w3, acc[1][3]);
        }
        if (ch + 1 < nch) put(stage + ((ch + 1) & 1) * (kKC * kOB));
        __syncthreads();
    }
}

template <typename T> __device__ __forceinline__ void zero_acc(T (&acc)[2][4]) {
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = T(0);
}

// elu'(x) recovered from y = elu(x):  x > 0 -> 1, else exp(x) = y + 1
template <typename T> __device__ __forceinline__ T elu_grad_from_out(T y) { return y > T(0) ? T(1) : y + T(1); }

// ---- global save areas for the backward pass (one per objective evaluation)
// node = level * N + chain, level in [0, L]; HS has L+2 levels (HS[0] = incoming hidden).
template <typename T> struct SaveAreas {
    T *XY, *EA, *P, *HS, *RG, *ZG, *NG, *HN, *U1, *U2, *OUT, *DYDX, *CM;
    T *DO, *DU2, *DU1, *DR, *DZ, *DN, *DNH, *DP, *DE;
};

inline size_t save_elems(int d, int H, int L, long long N) {
    // every area starts on a 16-byte boundary (even element count) so tiles can be stored as double2 / float4 pairs
    const size_t nodes = (size_t)(L + 1) * N;
    auto ev = [](size_t n) { return (n + 1) & ~(size_t)1; };
    return ev(nodes * (d + 1)) + nodes * H * 2 + (size_t)(L + 2) * N * H + nodes * H * 4 + nodes * H * 2 +
           ev(nodes * d) * 2 + ev(nodes) + ev(nodes * d) + nodes * H * 8;
}

template <typename T> inline SaveAreas<T> carve_save(void* base, int d, int H, int L, long long N) {
    SaveAreas<T> s;
    T* p = (T*)base;
    const size_t nodes = (size_t)(L + 1) * N;
    auto take = [&](size_t n) { T* q = p; p += (n + 1) & ~(size_t)1; return q; };
    s.XY = take(nodes * (d + 1));
    s.EA = take(nodes * H); s.P = take(nodes * H);
    s.HS = take((size_t)(L + 2) * N * H);
    s.RG = take(nodes * H); s.ZG = take(nodes * H); s.NG = take(nodes * H); s.HN = take(nodes * H);
    s.U1 = take(nodes * H); s.U2 = take(nodes * H);
    s.OUT = take(nodes * d); s.DYDX = take(nodes * d); s.CM = take(nodes);
    s.DO = take(nodes * d);
    s.DU2 = take(nodes * H); s.DU1 = take(nodes * H);
    s.DR = take(nodes * H); s.DZ = take(nodes * H); s.DN = take(nodes * H); s.DNH = take(nodes * H);
    s.DP = take(nodes * H); s.DE = take(nodes * H);
    return s;
}

// ---- shared-memory carve for the forward step
template <typename T> struct FwdSmem {
    T *IN, *EA, *PH, *HB, *U1, *U2, *OUT, *stage;
    __device__ void carve(T* base, int d, int H) {
        T* p = base;
        IN = p; p += (d + 1) * kRS;
        EA = p; p += H * kRS;
        PH = p; p += 2 * H * kRS;
        HB = p; p += H * kRS;
        U1 = p; p += H * kRS;
        U2 = p; p += H * kRS;
        OUT = p; p += ((d + 3) & ~3) * kRS;
        stage = p; p += 2 * kKC * kOB;
    }
    static size_t elems(int d, int H) { return (size_t)(d + 1) * kRS + 6 * H * kRS + ((d + 3) & ~3) * kRS + 2 * kKC * kOB; }
};

// One policy step for the CTA's tile, forward.
//   in : sm.IN  = [x; y] (k-major), sm.HB = previous hidden state
//   out: sm.OUT = sigmoid output [d][kRS], sm.PH second half = new hidden state (also copied to HB)
// theta_t: parameters with weight matrices transposed ([K][O]); head_a selects postpro_A.
// save != nullptr stores the activations of node (level, chain0 + r) for the backward pass.
template <typename T>
__device__ void policy_forward_tile(const PolicyLayout& lay, const T* __restrict__ theta_t, bool head_a, FwdSmem<T>& sm,
                                    const SaveAreas<T>* save, long long node0, int nvalid, T* __restrict__ hs_out) {
    const int d = lay.d, H = lay.H;
    const int tid = threadIdx.x;
    const TileCoord tc;
    // ---- embedding (K = d+1, tiny): e = We [x;y] + be ; EA = elu(e)
    {
        const T* Wt = theta_t + lay.off[S_WE];    // [d+1][H]
        const T* b = theta_t + lay.off[S_BE];
        for (int e = tid; e < H * kR; e += kNT) {
            const int o = e >> 5, r = e & 31;
            T acc = __ldg(b + o);
            for (int k = 0; k <= d; ++k) acc = fma(sm.IN[k * kRS + r], __ldg(Wt + k * H + o), acc);
            const T v = elu_(acc);
            sm.EA[o * kRS + r] = v;
            if (save && r < nvalid) save->EA[(node0 + r) * H + o] = v;
        }
        if (save) {
            for (int e = tid; e < (d + 1) * kR; e += kNT) {
                const int k = e >> 5, r = e & 31;
                if (r < nvalid) save->XY[(node0 + r) * (d + 1) + k] = sm.IN[k * kRS + r];
            }
        }
    }
    __syncthreads();
    T acc[2][4];
    // ---- prepro: p = elu(Wp ea + bp) -> PH[0:H]
    for (int cb = 0; cb < H; cb += kOB) {
        zero_acc(acc);
        gemm_block<T>(sm.EA, H, theta_t + lay.off[S_WP], H, cb, min(kOB, H - cb), sm.stage, tc, acc);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int o = cb + tc.c0 + j;
            if (o < H) {
                const T bb = __ldg(theta_t + lay.off[S_BP] + o);
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    const T v = elu_(acc[i][j] + bb);
                    sm.PH[o * kRS + tc.r0 + i] = v;
                    if (save && tc.r0 + i < nvalid) save->P[(node0 + tc.r0 + i) * H + o] = v;
                }
            }
        }
    }
    __syncthreads();
    // ---- GRU cell: gates r, z, n for the same (row, col) positions stay in registers
    for (int cb = 0; cb < H; cb += kOB) {
        const int nc = min(kOB, H - cb);
        T gi[2][4], gh[2][4], rg[2][4], zg[2][4];
        // r gate
        zero_acc(gi);
        gemm_block<T>(sm.PH, H, theta_t + lay.off[S_WIH], 3 * H, cb, nc, sm.stage, tc, gi);
        gemm_block<T>(sm.HB, H, theta_t + lay.off[S_WHH], 3 * H, cb, nc, sm.stage, tc, gi);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int o = cb + tc.c0 + j;
            const T bb = o < H ? __ldg(theta_t + lay.off[S_BIH] + o) + __ldg(theta_t + lay.off[S_BHH] + o) : T(0);
#pragma unroll
            for (int i = 0; i < 2; ++i) rg[i][j] = sigmoid_(gi[i][j] + bb);
        }
        // z gate
        zero_acc(gi);
        gemm_block<T>(sm.PH, H, theta_t + lay.off[S_WIH], 3 * H, H + cb, nc, sm.stage, tc, gi);
        gemm_block<T>(sm.HB, H, theta_t + lay.off[S_WHH], 3 * H, H + cb, nc, sm.stage, tc, gi);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int o = cb + tc.c0 + j;
            const T bb = o < H ? __ldg(theta_t + lay.off[S_BIH] + H + o) + __ldg(theta_t + lay.off[S_BHH] + H + o) : T(0);
#pragma unroll
            for (int i = 0; i < 2; ++i) zg[i][j] = sigmoid_(gi[i][j] + bb);
        }
        // n gate: tanh(gi_n + b_in + r * (gh_n + b_hn))
        zero_acc(gi);
        zero_acc(gh);
        gemm_block<T>(sm.PH, H, theta_t + lay.off[S_WIH], 3 * H, 2 * H + cb, nc, sm.stage, tc, gi);
        gemm_block<T>(sm.HB, H, theta_t + lay.off[S_WHH], 3 * H, 2 * H + cb, nc, sm.stage, tc, gh);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int o = cb + tc.c0 + j;
            if (o < H) {
                const T bi = __ldg(theta_t + lay.off[S_BIH] + 2 * H + o);
                const T bh = __ldg(theta_t + lay.off[S_BHH] + 2 * H + o);
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    const int r = tc.r0 + i;
                    const T hn = gh[i][j] + bh;
                    const T n = Num<T>::tanh(gi[i][j] + bi + rg[i][j] * hn);
                    const T hp = sm.HB[o * kRS + r];
                    const T hnew = (T(1) - zg[i][j]) * n + zg[i][j] * hp;
                    sm.PH[(H + o) * kRS + r] = hnew;
                    if (r < nvalid) {
                        if (save) {
                            const size_t q = (size_t)(node0 + r) * H + o;
                            save->RG[q] = rg[i][j]; save->ZG[q] = zg[i][j]; save->NG[q] = n; save->HN[q] = hn;
                        }
                        if (hs_out) hs_out[(size_t)r * H + o] = hnew;
                    }
                }
            }
        }
    }
    __syncthreads();
    // hidden state for the next step (HB is read by no one until then)
    for (int e = tid; e < H * kR; e += kNT) sm.HB[e] = sm.PH[H * kRS + e];
    // ---- post MLP layer 0: u1 = elu(W0 [p; h'] + b0)
    const int sW0 = head_a ? S_AW0 : S_XW0, sB0 = head_a ? S_AB0 : S_XB0;
    const int sW3 = head_a ? S_AW3 : S_XW3, sB3 = head_a ? S_AB3 : S_XB3;
    const int sW6 = head_a ? S_AW6 : S_XW6, sB6 = head_a ? S_AB6 : S_XB6;
    for (int cb = 0; cb < H; cb += kOB) {
        zero_acc(acc);
        gemm_block<T>(sm.PH, 2 * H, theta_t + lay.off[sW0], H, cb, min(kOB, H - cb), sm.stage, tc, acc);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int o = cb + tc.c0 + j;
            if (o < H) {
                const T bb = __ldg(theta_t + lay.off[sB0] + o);
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    const T v = elu_(acc[i][j] + bb);
                    sm.U1[o * kRS + tc.r0 + i] = v;
                    if (save && tc.r0 + i < nvalid) save->U1[(node0 + tc.r0 + i) * H + o] = v;
                }
            }
        }
    }
    __syncthreads();
    for (int cb = 0; cb < H; cb += kOB) {
        zero_acc(acc);
        gemm_block<T>(sm.U1, H, theta_t + lay.off[sW3], H, cb, min(kOB, H - cb), sm.stage, tc, acc);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int o = cb + tc.c0 + j;
            if (o < H) {
                const T bb = __ldg(theta_t + lay.off[sB3] + o);
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    const T v = elu_(acc[i][j] + bb);
                    sm.U2[o * kRS + tc.r0 + i] = v;
                    if (save && tc.r0 + i < nvalid) save->U2[(node0 + tc.r0 + i) * H + o] = v;
                }
            }
        }
    }
    __syncthreads();
    // ---- output layer (O = d, tiny): out = sigmoid(W6 u2 + b6)
    {
        const T* Wt = theta_t + lay.off[sW6];     // [H][d]
        const T* b = theta_t + lay.off[sB6];
        for (int e = tid; e < d * kR; e += kNT) {
            const int o = e >> 5, r = e & 31;
            T a = __ldg(b + o);
            for (int k = 0; k < H; ++k) a = fma(sm.U2[k * kRS + r], __ldg(Wt + k * d + o), a);
            const T v = sigmoid_(a);
            sm.OUT[o * kRS + r] = v;
            if (save && r < nvalid) save->OUT[(node0 + r) * d + o] = v;
        }
    }
    __syncthreads();
}

// ---- shared-memory carve for the backward step
template <typename T> struct BwdSmem {
    T *GO, *GU2, *GU1, *G4, *GP, *GE, *GH, *GXY, *stage;
    __device__ void carve(T* base, int d, int H) {
        T* p = base;
        GO = p; p += ((d + 3) & ~3) * kRS;
        GU2 = p; p += H * kRS;
        GU1 = p; p += H * kRS;
        G4 = p; p += 4 * H * kRS;
        GP = p; p += H * kRS;
        GE = p; p += H * kRS;
        GH = p; p += H * kRS;
        GXY = p; p += ((d + 4) & ~3) * kRS;
        stage = p; p += 2 * kKC * kOB;
    }
    static size_t elems(int d, int H) {
        return (size_t)((d + 3) & ~3) * kRS + 9 * H * kRS + ((d + 4) & ~3) * kRS + 2 * kKC * kOB;
    }
};

// One policy step backward for the CTA's tile.
//   in : sm.GO = adjoint of the sigmoid output [d][kRS];  sm.GH = adjoint of the new hidden state
//   out: sm.GXY = adjoint of [x; y] [(d+1)][kRS];  sm.GH = adjoint of the previous hidden state
// theta: parameters in torch layout ([O][K]).  Writes the pre-activation adjoints
// (DO, DU2, DU1, DR, DZ, DN, DNH, DP, DE) of the node for the weight-gradient kernel.
template <typename T>
__device__ void policy_backward_tile(const PolicyLayout& lay, const T* __restrict__ theta, bool head_a, BwdSmem<T>& sm,
                                     const SaveAreas<T>& sv, long long node0, long long hs_prev0, int nvalid) {
    const int d = lay.d, H = lay.H;
    const int tid = threadIdx.x;
    const TileCoord tc;
    const int sW0 = head_a ? S_AW0 : S_XW0, sW3 = head_a ? S_AW3 : S_XW3, sW6 = head_a ? S_AW6 : S_XW6;
    // ---- output layer: do_pre = go * o (1 - o)
    for (int e = tid; e < d * kR; e += kNT) {
        const int o = e >> 5, r = e & 31;
        T v = T(0);
        if (r < nvalid) {
            const T out = sv.OUT[(node0 + r) * d + o];
            v = sm.GO[o * kRS + r] * out * (T(1) - out);
            sv.DO[(node0 + r) * d + o] = v;
        }
        sm.GO[o * kRS + r] = v;
    }
    __syncthreads();
    // du2 = do_pre . W6 (K = d), du2_pre = du2 * elu'(u2)
    {
        const T* W = theta + lay.off[sW6];     // [d][H]
        for (int e = tid; e < H * kR; e += kNT) {
            const int k = e >> 5, r = e & 31;
            T v = T(0);
            if (r < nvalid) {
                T a = T(0);
                for (int o = 0; o < d; ++o) a = fma(sm.GO[o * kRS + r], __ldg(W + o * H + k), a);
                v = a * elu_grad_from_out(sv.U2[(node0 + r) * H + k]);
                sv.DU2[(node0 + r) * H + k] = v;
            }
            sm.GU2[k * kRS + r] = v;
        }
    }
    __syncthreads();
    T acc[2][4];
    // du1 = du2_pre . W3 ; du1_pre = du1 * elu'(u1)
    for (int cb = 0; cb < H; cb += kOB) {
        zero_acc(acc);
        gemm_block<T>(sm.GU2, H, theta + lay.off[sW3], H, cb, min(kOB, H - cb), sm.stage, tc, acc);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int k = cb + tc.c0 + j;
            if (k < H) {
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    const int r = tc.r0 + i;
                    T v = T(0);
                    if (r < nvalid) {
                        v = acc[i][j] * elu_grad_from_out(sv.U1[(node0 + r) * H + k]);
                        sv.DU1[(node0 + r) * H + k] = v;
                    }
                    sm.GU1[k * kRS + r] = v;
                }
            }
        }
    }
    __syncthreads();
    // d[p; h'] = du1_pre . W0   (W0 is [H][2H]); first H cols -> GP (raw dp), second H cols -> GRU backward
    for (int cb = 0; cb < H; cb += kOB) {
        zero_acc(acc);
        gemm_block<T>(sm.GU1, H, theta + lay.off[sW0], 2 * H, cb, min(kOB, H - cb), sm.stage, tc, acc);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int k = cb + tc.c0 + j;
            if (k < H) {
#pragma unroll
                for (int i = 0; i < 2; ++i) sm.GP[k * kRS + tc.r0 + i] = acc[i][j];
            }
        }
    }
    for (int cb = 0; cb < H; cb += kOB) {
        zero_acc(acc);
        gemm_block<T>(sm.GU1, H, theta + lay.off[sW0], 2 * H, H + cb, min(kOB, H - cb), sm.stage, tc, acc);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int k = cb + tc.c0 + j;
            if (k < H) {
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    const int r = tc.r0 + i;
                    T dr = T(0), dz = T(0), dn = T(0), dnh = T(0), dhp = T(0);
                    if (r < nvalid) {
                        const size_t q = (size_t)(node0 + r) * H + k;
                        const T dh = acc[i][j] + sm.GH[k * kRS + r];
                        const T rg = sv.RG[q], zg = sv.ZG[q], ng = sv.NG[q], hn = sv.HN[q];
                        const T hp = sv.HS[(size_t)(hs_prev0 + r) * H + k];
                        const T dng = dh * (T(1) - zg);
                        dn = dng * (T(1) - ng * ng);
                        dz = dh * (hp - ng) * zg * (T(1) - zg);
                        dr = dn * hn * rg * (T(1) - rg);
                        dnh = dn * rg;
                        dhp = dh * zg;
                        sv.DR[q] = dr; sv.DZ[q] = dz; sv.DN[q] = dn; sv.DNH[q] = dnh;
                    }
                    sm.G4[k * kRS + r] = dr;
                    sm.G4[(H + k) * kRS + r] = dz;
                    sm.G4[(2 * H + k) * kRS + r] = dn;
                    sm.G4[(3 * H + k) * kRS + r] = dnh;
                    sm.GH[k * kRS + r] = dhp;     // own element only: no hazard
                }
            }
        }
    }
    __syncthreads();
    // dp += dgi . Wih (K = 3H) ; dp_pre = dp * elu'(p)
    for (int cb = 0; cb < H; cb += kOB) {
        zero_acc(acc);
        gemm_block<T>(sm.G4, 3 * H, theta + lay.off[S_WIH], H, cb, min(kOB, H - cb), sm.stage, tc, acc);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int k = cb + tc.c0 + j;
            if (k < H) {
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    const int r = tc.r0 + i;
                    T v = T(0);
                    if (r < nvalid) {
                        v = (acc[i][j] + sm.GP[k * kRS + r]) * elu_grad_from_out(sv.P[(node0 + r) * H + k]);
                        sv.DP[(node0 + r) * H + k] = v;
                    }
                    sm.GP[k * kRS + r] = v;
                }
            }
        }
    }
    // dh_prev += dgh . Whh : rows [0,2H) of G4 with Whh[0:2H], rows [3H,4H) with Whh[2H:3H]
    for (int cb = 0; cb < H; cb += kOB) {
        zero_acc(acc);
        gemm_block<T>(sm.G4, 2 * H, theta + lay.off[S_WHH], H, cb, min(kOB, H - cb), sm.stage, tc, acc);
        gemm_block<T>(sm.G4 + 3 * H * kRS, H, theta + lay.off[S_WHH] + (size_t)2 * H * H, H, cb, min(kOB, H - cb),
                      sm.stage, tc, acc);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int k = cb + tc.c0 + j;
            if (k < H) {
#pragma unroll
                for (int i = 0; i < 2; ++i) sm.GH[k * kRS + tc.r0 + i] += acc[i][j];
            }
        }
    }
    __syncthreads();
    // dea = dp_pre . Wp ; de_pre = dea * elu'(ea)
    for (int cb = 0; cb < H; cb += kOB) {
        zero_acc(acc);
        gemm_block<T>(sm.GP, H, theta + lay.off[S_WP], H, cb, min(kOB, H - cb), sm.stage, tc, acc);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int k = cb + tc.c0 + j;
            if (k < H) {
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    const int r = tc.r0 + i;
                    T v = T(0);
                    if (r < nvalid) {
                        v = acc[i][j] * elu_grad_from_out(sv.EA[(node0 + r) * H + k]);
                        sv.DE[(node0 + r) * H + k] = v;
                    }
                    sm.GE[k * kRS + r] = v;
                }
            }
        }
    }
    __syncthreads();
    // d[x; y] = de_pre . We (We is [H][d+1])
    {
        const T* W = theta + lay.off[S_WE];
        for (int e = tid; e < (d + 1) * kR; e += kNT) {
            const int k = e >> 5, r = e & 31;
            T a = T(0);
            for (int o = 0; o < H; ++o) a = fma(sm.GE[o * kRS + r], __ldg(W + o * (d + 1) + k), a);
            sm.GXY[k * kRS + r] = a;
        }
    }
    __syncthreads();
}

}  // namespace nmb
