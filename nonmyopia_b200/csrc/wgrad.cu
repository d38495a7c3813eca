// Weight gradients of the amortized policy: dW[o][k] = sum_nodes dY[node][o] * A[node][k]
// for every layer, over every (level, chain) node of a back-propagated rollout
// (the autograd.grad(loss, network parameters) of reference actor.py:641-644).
//
// Split-K GEMMs: a work item is (layer tile <= 64 x 64, chunk of nodes).  256 threads, 4 x 4
// outputs per thread; dY and A chunks (32 nodes) are staged node-major in shared memory,
// register-prefetched one chunk ahead.  Each item writes its partial tile to
// partials[split][param]; the reduction over splits is done in a fixed order by the
// AdamW kernel (optim.cu) or nmb_policy_reduce_partials -> bitwise deterministic.
#include "policy_dev.cuh"
#include "policy_mma.cuh"
#include "policy_tf32.cuh"

namespace nmb {

constexpr int kNC = 32;          // nodes per staged chunk
constexpr int kMaxTasks = 20;

template <typename T> struct WgTask {
    const T* dY; int ldY, o0, On;        // dY[node*ldY + o0 + o], o < On (<= 64)
    const T* A;  int ldA, k0, Kn;        // A[node*ldA + k0 + k],  k < Kn (<= 64)
    long long n_begin, n_end;            // node range
    int dst, ldK;                        // partial[dst + o*ldK + k]
    int bias_dst;                        // partial[bias_dst + o] or -1
    int item0, n_items;                  // work items of this task
    int ybulk, abulk;                    // operand rows are 16-byte multiples at 16-byte aligned addresses (TMA-able)
    int pair;                            // both operands are dense [node][64] areas: two nodes travel per bulk copy
};

template <typename T> struct WgArgs {
    WgTask<T> t[kMaxTasks];
    int n_tasks;
    int chunk;                           // nodes per work item
    int n_params;
    T* partials;
    const int* done_flag;
};

template <typename T>
__global__ void __launch_bounds__(256, 2) wgrad_kernel(const __grid_constant__ WgArgs<T> args) {
    if (args.done_flag && *args.done_flag) return;
    extern __shared__ __align__(16) unsigned char wg_smem_raw[];
    T (*Ys)[kNC * 64] = reinterpret_cast<T (*)[kNC * 64]>(wg_smem_raw);
    T (*As)[kNC * 64] = reinterpret_cast<T (*)[kNC * 64]>(wg_smem_raw + 2 * kNC * 64 * sizeof(T));
    const int item = blockIdx.x;
    int ti = 0;
    while (ti + 1 < args.n_tasks && item >= args.t[ti + 1].item0) ++ti;
    const WgTask<T>& tk = args.t[ti];
    const int split = item - tk.item0;
    const long long nb = tk.n_begin + (long long)split * args.chunk;
    const long long ne = min(tk.n_end, nb + args.chunk);
    const int tid = threadIdx.x;
    const int ty = tid >> 4, tx = tid & 15;

    T acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = T(0);
    T bacc = T(0);

    T ry[8], ra[8];
    auto fetch = [&](long long n0) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int e = tid + i * 256;
            const int nn = e >> 6, c = e & 63;
            const long long node = n0 + nn;
            const bool ok = node < ne;
            ry[i] = (ok && c < tk.On) ? tk.dY[(size_t)node * tk.ldY + tk.o0 + c] : T(0);
            ra[i] = (ok && c < tk.Kn) ? tk.A[(size_t)node * tk.ldA + tk.k0 + c] : T(0);
        }
    };
    auto put = [&](int b) {
#pragma unroll
        for (int i = 0; i < 8; ++i) { Ys[b][tid + i * 256] = ry[i]; As[b][tid + i * 256] = ra[i]; }
    };
    const int nch = (int)((ne - nb + kNC - 1) / kNC);
    if (nch > 0) {
        fetch(nb);
        put(0);
    }
    __syncthreads();
    for (int ch = 0; ch < nch; ++ch) {
        const int b = ch & 1;
        if (ch + 1 < nch) fetch(nb + (long long)(ch + 1) * kNC);
        const T* ys = Ys[b] + ty * 4;
        const T* as = As[b] + tx * 4;
#pragma unroll 8
        for (int nn = 0; nn < kNC; ++nn) {
            T y0, y1, y2, y3, a0, a1, a2, a3;
            ld4(ys + nn * 64, y0, y1, y2, y3);
            ld4(as + nn * 64, a0, a1, a2, a3);
            acc[0][0] = fma(y0, a0, acc[0][0]); acc[0][1] = fma(y0, a1, acc[0][1]);
            acc[0][2] = fma(y0, a2, acc[0][2]); acc[0][3] = fma(y0, a3, acc[0][3]);
            acc[1][0] = fma(y1, a0, acc[1][0]); acc[1][1] = fma(y1, a1, acc[1][1]);
            acc[1][2] = fma(y1, a2, acc[1][2]); acc[1][3] = fma(y1, a3, acc[1][3]);
            acc[2][0] = fma(y2, a0, acc[2][0]); acc[2][1] = fma(y2, a1, acc[2][1]);
            acc[2][2] = fma(y2, a2, acc[2][2]); acc[2][3] = fma(y2, a3, acc[2][3]);
            acc[3][0] = fma(y3, a0, acc[3][0]); acc[3][1] = fma(y3, a1, acc[3][1]);
            acc[3][2] = fma(y3, a2, acc[3][2]); acc[3][3] = fma(y3, a3, acc[3][3]);
        }
        if (tk.bias_dst >= 0 && tid < 64) {
#pragma unroll 8
            for (int nn = 0; nn < kNC; ++nn) bacc += Ys[b][nn * 64 + tid];
        }
        if (ch + 1 < nch) put(b ^ 1);
        __syncthreads();
    }
    T* out = args.partials + (size_t)split * args.n_params;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int o = ty * 4 + i;
        if (o < tk.On) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int k = tx * 4 + j;
                if (k < tk.Kn) out[tk.dst + (size_t)o * tk.ldK + k] = acc[i][j];
            }
        }
    }
    if (tk.bias_dst >= 0 && tid < tk.On) out[tk.bias_dst + tid] = bacc;
}

// ------------------------------------------------------------------ FP64 tensor-pipe variant (H x H tasks)
// 8 consumer warps own 16 (o) x 32 (k) tiles of the 64 x 64 weight tile; the contraction runs over nodes,
// so BOTH operands (dY and A, node-major in the save workspace) are k-major: a producer warp streams 32-node
// chunks of both with cp.async.bulk row copies into a 3-stage ring (row stride 72 doubles: conflict-free
// fragments), 2 CTAs per SM.
constexpr int kWgStages = 3;
constexpr int kWgStageElems = 2 * kNC * kSW;      // dY rows then A rows


// ---- producer warp of the tensor-core weight-gradient kernels (FP64 and TF32).
// Single-row mode: lane = node row of the 32-node chunk, one bulk copy per row into a padded row (stride SW).
// Pair mode (both operands dense [node][64]: every H x H task at H = 64): lane p < 16 moves nodes 2p and 2p + 1 as ONE
// copy of 128 elements into pair row p (stride SWP = 128 + pad) - half as many copy operations for the same bytes;
// the kernels were bound by the issue rate of the per-row copies (same time for 256-byte FP32 rows and 512-byte FP64
// rows).  The consumers then walk the nodes of a chunk in the order (pair rows t = 0..3: first halves, second halves),
// which keeps every fragment load on distinct banks (see the kernels).
template <typename T> struct WgLay {
    static constexpr int SW = sizeof(T) == 8 ? kSW : kSWf;                  // single-row stride
    static constexpr int SWP = 128 + (sizeof(T) == 8 ? 4 : 8);              // pair-row stride: 4 (mod 16) doubles / 8 (mod 32) floats
    static constexpr int stage_elems = 2 * kNC * SW;                        // >= 2 * 16 * SWP
};

template <typename T>
__device__ __forceinline__ void wg_producer(const WgTask<T>& tk, T* stage, unsigned full0, unsigned empty0, long long nb,
                                            long long ne, int nch, int lane) {
    constexpr int SW = WgLay<T>::SW, SWP = WgLay<T>::SWP, SE = WgLay<T>::stage_elems;
    const unsigned ybytes = (unsigned)tk.On * (unsigned)sizeof(T), abytes = (unsigned)tk.Kn * (unsigned)sizeof(T);
    for (int ch = 0; ch < nch; ++ch) {
        const unsigned slot = ch % kWgStages, fill = ch / kWgStages;
        if (fill > 0) mbar_wait(empty0 + 8 * slot, (fill - 1) & 1);
        const long long n0 = nb + (long long)ch * kNC;
        const int valid = (int)min((long long)kNC, ne - n0);
        const unsigned full = full0 + 8 * slot;
        T* base = stage + (size_t)slot * SE;
        if (tk.pair) {
            T* ys = base + lane * SWP;                       // lanes 0..15: pair rows of dY, then of A
            T* as = base + 16 * SWP + lane * SWP;
            const int have = lane < 16 ? max(0, min(2, valid - 2 * lane)) : 0;     // nodes of this pair inside the range
            if (lane < 16 && have < 2) {                      // ragged tail: zero what no copy will write
                for (int c = have * 64; c < 128; ++c) { ys[c] = T(0); as[c] = T(0); }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive_expect_tx(full, (ybytes + abytes) * (unsigned)valid);
            __syncwarp();
            if (have > 0) {
                const long long node = n0 + 2 * lane;
                bulk_g2s(smem_u32(ys), tk.dY + (size_t)node * 64, ybytes * (unsigned)have, full);
                bulk_g2s(smem_u32(as), tk.A + (size_t)node * 64, abytes * (unsigned)have, full);
            }
            continue;
        }
        T* ys = base + lane * SW;
        T* as = ys + kNC * SW;
        if (lane >= valid) {      // ragged tail: zero rows
            for (int c = 0; c < 64; ++c) { ys[c] = T(0); as[c] = T(0); }
        }
        // rows that are not 16-byte multiples (K = d+1) are copied by the lane itself
        if (lane < valid) {
            const long long node = n0 + lane;
            if (!tk.ybulk) for (int c = 0; c < tk.On; ++c) ys[c] = tk.dY[(size_t)node * tk.ldY + tk.o0 + c];
            if (!tk.abulk) for (int c = 0; c < tk.Kn; ++c) as[c] = tk.A[(size_t)node * tk.ldA + tk.k0 + c];
        }
        __syncwarp();
        if (lane == 0) mbar_arrive_expect_tx(full, ((tk.ybulk ? ybytes : 0u) + (tk.abulk ? abytes : 0u)) * (unsigned)valid);
        __syncwarp();
        if (lane < valid) {
            const long long node = n0 + lane;
            if (tk.ybulk) bulk_g2s(smem_u32(ys), tk.dY + (size_t)node * tk.ldY + tk.o0, ybytes, full);
            if (tk.abulk) bulk_g2s(smem_u32(as), tk.A + (size_t)node * tk.ldA + tk.k0, abytes, full);
        }
    }
}

__global__ void __launch_bounds__(288, 2) wgrad_mma_kernel(const __grid_constant__ WgArgs<double> args) {
    if (args.done_flag && *args.done_flag) return;
    extern __shared__ __align__(16) unsigned char wg_smem_raw[];
    double* stage = reinterpret_cast<double*>(wg_smem_raw);
    const unsigned full0 = smem_u32(stage + kWgStages * kWgStageElems), empty0 = full0 + 8 * kWgStages;
    const int item = blockIdx.x;
    int ti = 0;
    while (ti + 1 < args.n_tasks && item >= args.t[ti + 1].item0) ++ti;
    const WgTask<double>& tk = args.t[ti];
    const int split = item - tk.item0;
    const long long nb = tk.n_begin + (long long)split * args.chunk;
    const long long ne = min(tk.n_end, nb + args.chunk);
    const int nch = (int)((ne - nb + kNC - 1) / kNC);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) {
        for (int s = 0; s < kWgStages; ++s) {
            mbar_init(full0 + 8 * s, 1);
            mbar_init(empty0 + 8 * s, 8);
        }
        fence_barrier_init();
    }
    // thin operands (K = d+1 or O = d) fill only the first columns of their rows: the rest must read as zero
    if (tk.On < 64 || tk.Kn < 64) {
        for (int e = tid; e < kWgStages * kWgStageElems; e += blockDim.x) stage[e] = 0.0;
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();
    if (warp == 8) {
        wg_producer<double>(tk, stage, full0, empty0, nb, ne, nch, lane);
        return;
    }
    // ---- consumers
    const int fr = lane >> 2, fk = lane & 3;
    const int o0 = (warp >> 1) * 16, k0 = (warp & 1) * 32;
    double acc[2][4][2];
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int t = 0; t < 4; ++t) { acc[i][t][0] = 0.0; acc[i][t][1] = 0.0; }
    double bacc = 0.0;
    for (int ch = 0; ch < nch; ++ch) {
        const unsigned slot = ch % kWgStages, fill = ch / kWgStages;
        mbar_wait(full0 + 8 * slot, fill & 1);
        if (tk.pair) {
            // pair rows: k index fk of a DMMA = pair row 4 q + fk, first halves (even nodes) then second halves (odd nodes);
            // a half-warp touches pair rows fk = 0..3 x columns 0..3: (4 fk + c) mod 16 distinct 8-byte banks
            constexpr int SWP = WgLay<double>::SWP;
            const double* yp = stage + (size_t)slot * kWgStageElems + fk * SWP + o0 + fr;
            const double* ap = stage + (size_t)slot * kWgStageElems + 16 * SWP + fk * SWP + k0 + fr;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
#pragma unroll
                for (int half = 0; half < 2; ++half) {
                    const int o = q * 4 * SWP + half * 64;
                    const double a0 = yp[o], a1 = yp[o + 8];
                    const double b0 = ap[o], b1 = ap[o + 8], b2 = ap[o + 16], b3 = ap[o + 24];
                    dmma884(acc[0][0], a0, b0); dmma884(acc[0][1], a0, b1); dmma884(acc[0][2], a0, b2); dmma884(acc[0][3], a0, b3);
                    dmma884(acc[1][0], a1, b0); dmma884(acc[1][1], a1, b1); dmma884(acc[1][2], a1, b2); dmma884(acc[1][3], a1, b3);
                }
            }
            if (tk.bias_dst >= 0 && tid < 64) {
                const double* yb = stage + (size_t)slot * kWgStageElems + tid;
#pragma unroll 8
                for (int pr = 0; pr < 16; ++pr) bacc += yb[pr * SWP] + yb[pr * SWP + 64];
            }
        } else {
        const double* ys = stage + (size_t)slot * kWgStageElems + fk * kSW + o0 + fr;
        const double* as = stage + (size_t)slot * kWgStageElems + kNC * kSW + fk * kSW + k0 + fr;
#pragma unroll
        for (int kk = 0; kk < kNC; kk += 4) {
            const double a0 = ys[kk * kSW], a1 = ys[kk * kSW + 8];
            const double b0 = as[kk * kSW], b1 = as[kk * kSW + 8], b2 = as[kk * kSW + 16], b3 = as[kk * kSW + 24];
            dmma884(acc[0][0], a0, b0); dmma884(acc[0][1], a0, b1); dmma884(acc[0][2], a0, b2); dmma884(acc[0][3], a0, b3);
            dmma884(acc[1][0], a1, b0); dmma884(acc[1][1], a1, b1); dmma884(acc[1][2], a1, b2); dmma884(acc[1][3], a1, b3);
        }
        if (tk.bias_dst >= 0 && tid < 64) {
            const double* yb = stage + (size_t)slot * kWgStageElems + tid;
#pragma unroll 8
            for (int nn = 0; nn < kNC; ++nn) bacc += yb[nn * kSW];
        }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(empty0 + 8 * slot);
    }
    double* out = args.partials + (size_t)split * args.n_params;
    const bool vec_ok = !(tk.ldK & 1) && !(tk.dst & 1) && !(args.n_params & 1);
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const int o = o0 + 8 * i + fr;
        if (o < tk.On) {
#pragma unroll
            for (int t = 0; t < 4; ++t) {
                const int k = k0 + 8 * t + 2 * fk;
                double* pp = out + tk.dst + (size_t)o * tk.ldK + k;
                if (vec_ok && k + 1 < tk.Kn) {
                    *reinterpret_cast<double2*>(pp) = make_double2(acc[i][t][0], acc[i][t][1]);
                } else {
                    if (k < tk.Kn) pp[0] = acc[i][t][0];
                    if (k + 1 < tk.Kn) pp[1] = acc[i][t][1];
                }
            }
        }
    }
    if (tk.bias_dst >= 0 && tid < tk.On) out[tk.bias_dst + tid] = bacc;
}

// ------------------------------------------------------------------ FP32 variant on the TF32 tensor cores
// Same work split, ring and producer as wgrad_mma_kernel; the 8 consumer warps own 16 (o) x 32 (k) tiles and run
// mma.sync.m16n8k8.tf32 over the 32 nodes of a chunk (operands rounded to TF32 with cvt.rna, FP32 accumulation).
// Stage row stride 72 floats = 8 (mod 32): the A fragment (4 node rows x 8 o columns per load) and the B fragment
// (4 node rows x 8 k columns) each touch 32 distinct banks.  Replaces the FFMA register tiles for the H x H tasks of
// the FP32 variant (0.070 ms per launch at 4096 chains, slower than the FP64 tensor-pipe kernel).
constexpr int kWgStageElemsF = 2 * kNC * kSWf;

__global__ void __launch_bounds__(288, 2) wgrad_tf32_kernel(const __grid_constant__ WgArgs<float> args) {
    if (args.done_flag && *args.done_flag) return;
    extern __shared__ __align__(16) unsigned char wg_smem_raw[];
    float* stage = reinterpret_cast<float*>(wg_smem_raw);
    const unsigned full0 = smem_u32(stage + kWgStages * kWgStageElemsF), empty0 = full0 + 8 * kWgStages;
    const int item = blockIdx.x;
    int ti = 0;
    while (ti + 1 < args.n_tasks && item >= args.t[ti + 1].item0) ++ti;
    const WgTask<float>& tk = args.t[ti];
    const int split = item - tk.item0;
    const long long nb = tk.n_begin + (long long)split * args.chunk;
    const long long ne = min(tk.n_end, nb + args.chunk);
    const int nch = (int)((ne - nb + kNC - 1) / kNC);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) {
        for (int s = 0; s < kWgStages; ++s) {
            mbar_init(full0 + 8 * s, 1);
            mbar_init(empty0 + 8 * s, 8);
        }
        fence_barrier_init();
    }
    if (tk.On < 64 || tk.Kn < 64) {
        for (int e = tid; e < kWgStages * kWgStageElemsF; e += blockDim.x) stage[e] = 0.0f;
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();
    if (warp == 8) {
        wg_producer<float>(tk, stage, full0, empty0, nb, ne, nch, lane);
        return;
    }
    // ---- consumers
    const int g = lane >> 2, t = lane & 3;
    const int o0 = (warp >> 1) * 16, k0 = (warp & 1) * 32;
    float acc[4][4];
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int q = 0; q < 4; ++q) acc[j][q] = 0.0f;
    float bacc = 0.0f;
    for (int ch = 0; ch < nch; ++ch) {
        const unsigned slot = ch % kWgStages, fill = ch / kWgStages;
        mbar_wait(full0 + 8 * slot, fill & 1);
        if (tk.pair) {
            // pair rows: k index t of an MMA = even node of pair row 4 q + t, k index t + 4 = its odd node (64 floats on):
            // a warp touches pair rows t = 0..3 x 8 columns: (8 t + g) mod 32 distinct banks
            constexpr int SWP = WgLay<float>::SWP;
            const float* yp = stage + (size_t)slot * kWgStageElemsF + t * SWP + o0 + g;
            const float* ap = stage + (size_t)slot * kWgStageElemsF + 16 * SWP + t * SWP + k0 + g;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int o = q * 4 * SWP;
                const unsigned a0 = to_tf32(yp[o]), a1 = to_tf32(yp[o + 8]);
                const unsigned a2 = to_tf32(yp[o + 64]), a3 = to_tf32(yp[o + 72]);
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const unsigned b0 = to_tf32(ap[o + 8 * j]), b1 = to_tf32(ap[o + 64 + 8 * j]);
                    mma_tf32_16x8x8(acc[j], a0, a1, a2, a3, b0, b1);
                }
            }
            if (tk.bias_dst >= 0 && tid < 64) {
                const float* yb = stage + (size_t)slot * kWgStageElemsF + tid;
#pragma unroll 8
                for (int pr = 0; pr < 16; ++pr) bacc += yb[pr * SWP] + yb[pr * SWP + 64];
            }
        } else {
        const float* ys = stage + (size_t)slot * kWgStageElemsF + t * kSWf + o0 + g;
        const float* as = stage + (size_t)slot * kWgStageElemsF + kNC * kSWf + t * kSWf + k0 + g;
#pragma unroll
        for (int kk = 0; kk < kNC; kk += 8) {
            const unsigned a0 = to_tf32(ys[kk * kSWf]), a1 = to_tf32(ys[kk * kSWf + 8]);
            const unsigned a2 = to_tf32(ys[(kk + 4) * kSWf]), a3 = to_tf32(ys[(kk + 4) * kSWf + 8]);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const unsigned b0 = to_tf32(as[kk * kSWf + 8 * j]), b1 = to_tf32(as[(kk + 4) * kSWf + 8 * j]);
                mma_tf32_16x8x8(acc[j], a0, a1, a2, a3, b0, b1);
            }
        }
        if (tk.bias_dst >= 0 && tid < 64) {
            const float* yb = stage + (size_t)slot * kWgStageElemsF + tid;
#pragma unroll 8
            for (int nn = 0; nn < kNC; ++nn) bacc += yb[nn * kSWf];
        }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(empty0 + 8 * slot);
    }
    float* out = args.partials + (size_t)split * args.n_params;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const int o = o0 + 8 * i + g;
        if (o < tk.On) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int k = k0 + 8 * j + 2 * t;
                float* pp = out + tk.dst + (size_t)o * tk.ldK + k;
                if (k < tk.Kn) pp[0] = acc[j][2 * i];
                if (k + 1 < tk.Kn) pp[1] = acc[j][2 * i + 1];
            }
        }
    }
    if (tk.bias_dst >= 0 && tid < tk.On) out[tk.bias_dst + tid] = bacc;
}

constexpr int kThinMax = NMB_MAX_DIM + 1;     // a "thin" operand side: K = d + 1 or O = d

static void launch_wgrad_mma(const WgArgs<double>& am, int items, cudaStream_t st) {
    const size_t smem = (size_t)kWgStages * kWgStageElems * 8 + 2 * kWgStages * 8;
    cudaFuncSetAttribute(wgrad_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    wgrad_mma_kernel<<<items, 288, smem, st>>>(am);
}
static void launch_wgrad_mma(const WgArgs<float>& am, int items, cudaStream_t st) {
    const size_t smem = (size_t)kWgStages * kWgStageElemsF * 4 + 2 * kWgStages * 8;
    cudaFuncSetAttribute(wgrad_tf32_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    wgrad_tf32_kernel<<<items, 288, smem, st>>>(am);
}

// nodes per work item: a multiple of 32, at most 32 splits per segment, and - for the H x H tasks (8 over all
// nodes, 4 over the design levels, 4 over the action level, thin ones included) - no more items than the 2 CTAs/SM that are resident
static int wgrad_chunk(int L, long long N) {
    const long long nodes = (long long)(L + 1) * N;
    long long ch = (nodes + 31) / 32;
    ch = ((ch + kNC - 1) / kNC) * kNC;
    if (ch < 64) ch = 64;
    const long long slots = 2LL * sm_count();
    auto items = [&](long long c) {
        return 8 * ((nodes + c - 1) / c) + (L > 0 ? 4 * (((long long)L * N + c - 1) / c) : 0) + 4 * ((N + c - 1) / c);
    };
    if (items(ch) > slots) {
        // the first chunk size that fits in one wave, unless it is a poor fit anyway (then keep 2 balanced waves)
        long long c = ch;
        while (items(c) > slots && c < 2 * ch) c += kNC;
        if (items(c) <= slots) ch = c;
    }
    return (int)ch;
}

template <typename T>
static int launch_wgrad(const nmb_ts_problem_t* p, const void* save, void* partials, int32_t* seg_splits,
                        cudaStream_t st) {
    const int d = p->d, H = p->H, L = p->L;
    const long long N = p->n_chains;
    PolicyLayout lay;
    lay.init(d, H, 1);
    SaveAreas<T> sv = carve_save<T>(const_cast<void*>(save), d, H, L, N);
    // H x H tasks go to the FP64 tensor-pipe kernel (FP64, 16-byte aligned rows: d even), the rest (K = d+1,
    // O = d, FP32) to the FMA kernel
    // FP64: DMMA kernel; FP32: TF32 kernel (env NMB_FP32_NO_TF32 keeps the FFMA tiles, like the rollout kernels)
    static const bool no_tf32 = getenv("NMB_FP32_NO_TF32") != nullptr;
    const bool use_mma = (sizeof(T) == 8 ? !(d & 1) : !no_tf32) && H >= 8 && ((uintptr_t)save & 15) == 0 &&
                         ((uintptr_t)partials & 15) == 0;
    WgArgs<T> a, am;
    memset(&a, 0, sizeof(a));
    a.chunk = wgrad_chunk(L, N);
    a.n_params = lay.total();
    a.partials = (T*)partials;
    a.done_flag = p->done_flag;
    am = a;
    const long long all0 = 0, all1 = (long long)(L + 1) * N;
    const long long x0 = 0, x1 = (long long)L * N, a0 = (long long)L * N, a1 = all1;
    int nt = 0, item_small = 0, ntm = 0, itemm = 0;
    for (int s = 0; s < S_COUNT; ++s) seg_splits[s] = 0;
    auto add = [&](const T* dY, int ldY, int o0, int On, const T* A, int ldA, int k0, int Kn, long long nb,
                   long long ne, int wseg, int orow0, int kcol0, int bseg) {
        if (ne <= nb) return;
        const bool big = use_mma && (On == H || On <= kThinMax) && (Kn == H || Kn <= kThinMax);
        WgTask<T>& t = big ? am.t[ntm++] : a.t[nt++];
        int& item = big ? itemm : item_small;
        t.dY = dY; t.ldY = ldY; t.o0 = o0; t.On = On;
        t.A = A; t.ldA = ldA; t.k0 = k0; t.Kn = Kn;
        t.n_begin = nb; t.n_end = ne;
        t.ldK = lay.K[wseg];
        t.dst = lay.off[wseg] + orow0 * t.ldK + kcol0;
        t.bias_dst = bseg >= 0 ? lay.off[bseg] + orow0 : -1;
        // rows that a bulk copy can move: 16-byte multiples at 16-byte aligned addresses
        constexpr int al = 16 / (int)sizeof(T) - 1;
        t.ybulk = !(ldY & al) && !(On & al) && ((uintptr_t)(dY + o0) & 15) == 0;
        t.abulk = !(ldA & al) && !(Kn & al) && ((uintptr_t)(A + k0) & 15) == 0;
        static const bool no_pair = getenv("NMB_WGRAD_NO_PAIR") != nullptr;      // A/B switch
        t.pair = sizeof(T) == 4 && big && !no_pair && t.ybulk && t.abulk && On == 64 && Kn == 64 && ldY == 64 && ldA == 64 && o0 == 0 && k0 == 0;
        t.item0 = item;
        t.n_items = (int)((ne - nb + a.chunk - 1) / a.chunk);
        item += t.n_items;
        seg_splits[wseg] = t.n_items;
        if (bseg >= 0) seg_splits[bseg] = t.n_items;
    };
    add(sv.DE, H, 0, H, sv.XY, d + 1, 0, d + 1, all0, all1, S_WE, 0, 0, S_BE);
    add(sv.DP, H, 0, H, sv.EA, H, 0, H, all0, all1, S_WP, 0, 0, S_BP);
    add(sv.DR, H, 0, H, sv.P, H, 0, H, all0, all1, S_WIH, 0, 0, S_BIH);
    add(sv.DZ, H, 0, H, sv.P, H, 0, H, all0, all1, S_WIH, H, 0, S_BIH);
    add(sv.DN, H, 0, H, sv.P, H, 0, H, all0, all1, S_WIH, 2 * H, 0, S_BIH);
    add(sv.DR, H, 0, H, sv.HS, H, 0, H, all0, all1, S_WHH, 0, 0, S_BHH);
    add(sv.DZ, H, 0, H, sv.HS, H, 0, H, all0, all1, S_WHH, H, 0, S_BHH);
    add(sv.DNH, H, 0, H, sv.HS, H, 0, H, all0, all1, S_WHH, 2 * H, 0, S_BHH);
    const T* hnew = sv.HS + (size_t)N * H;      // hidden state after the step = HS[level + 1]
    add(sv.DU1, H, 0, H, sv.P, H, 0, H, x0, x1, S_XW0, 0, 0, S_XB0);
    add(sv.DU1, H, 0, H, hnew, H, 0, H, x0, x1, S_XW0, 0, H, -1);
    add(sv.DU2, H, 0, H, sv.U1, H, 0, H, x0, x1, S_XW3, 0, 0, S_XB3);
    add(sv.DO, d, 0, d, sv.U2, H, 0, H, x0, x1, S_XW6, 0, 0, S_XB6);
    add(sv.DU1, H, 0, H, sv.P, H, 0, H, a0, a1, S_AW0, 0, 0, S_AB0);
    add(sv.DU1, H, 0, H, hnew, H, 0, H, a0, a1, S_AW0, 0, H, -1);
    add(sv.DU2, H, 0, H, sv.U1, H, 0, H, a0, a1, S_AW3, 0, 0, S_AB3);
    add(sv.DO, d, 0, d, sv.U2, H, 0, H, a0, a1, S_AW6, 0, 0, S_AB6);
    a.n_tasks = nt;
    am.n_tasks = ntm;
    // bias rows of the gate blocks: offsets inside bih/bhh follow orow0 (handled by bias_dst above)
    if (itemm > 0) launch_wgrad_mma(am, itemm, st);
    if (item_small > 0) {
        const size_t smem = 4 * kNC * 64 * sizeof(T);
        NMB_CUDA(cudaFuncSetAttribute(wgrad_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        wgrad_kernel<T><<<item_small, 256, smem, st>>>(a);
    }
    NMB_LAUNCH_CHECK();
    return 0;
}

// grad[i] = sum_{s < splits(seg(i))} partials[s][i]
template <typename T> struct SegTable { int off[S_COUNT + 1]; int splits[S_COUNT]; };

template <typename T>
__global__ void reduce_partials_kernel(SegTable<T> tb, const T* __restrict__ partials, int n_params, T* __restrict__ grad) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_params) return;
    int s = 0;
    while (s + 1 < S_COUNT && i >= tb.off[s + 1]) ++s;
    T g = T(0);
    for (int q = 0; q < tb.splits[s]; ++q) g += partials[(size_t)q * n_params + i];
    grad[i] = g;
}

}  // namespace nmb

using namespace nmb;

extern "C" int64_t nmb_policy_wgrad_max_splits(int L, int64_t n_chains) {
    const int ch = wgrad_chunk(L, n_chains);
    return ((int64_t)(L + 1) * n_chains + ch - 1) / ch;
}

extern "C" int nmb_policy_wgrad(const nmb_ts_problem_t* prob, const void* save, void* partials,
                                int32_t* seg_splits_host, void* stream) {
    NMB_REQUIRE(prob && save && partials && seg_splits_host, NMB_EINVAL, "nmb_policy_wgrad: null argument");
    NMB_REQUIRE(prob->H >= 4 && prob->H <= kMaxH && prob->H % 4 == 0, NMB_ELIMIT, "nmb_policy_wgrad: hidden_dim=%d",
                prob->H);
    NMB_REQUIRE(prob->d >= 1 && prob->d <= NMB_MAX_DIM, NMB_ELIMIT, "nmb_policy_wgrad: d=%d", prob->d);
    cudaStream_t st = (cudaStream_t)stream;
    if (prob->dtype == NMB_F64) return launch_wgrad<double>(prob, save, partials, seg_splits_host, st);
    if (prob->dtype == NMB_F32) return launch_wgrad<float>(prob, save, partials, seg_splits_host, st);
    set_error("nmb_policy_wgrad: dtype %d", prob->dtype);
    return NMB_EINVAL;
}

extern "C" int nmb_policy_reduce_partials(int dtype, int d, int H, int n_actions, const void* partials,
                                          const int32_t* seg_splits_host, void* grad, void* stream) {
    NMB_REQUIRE(partials && seg_splits_host && grad, NMB_EINVAL, "nmb_policy_reduce_partials: null argument");
    PolicyLayout lay;
    lay.init(d, H, n_actions);
    cudaStream_t st = (cudaStream_t)stream;
    const int n = lay.total();
    if (dtype == NMB_F64) {
        SegTable<double> tb;
        for (int s = 0; s <= S_COUNT; ++s) tb.off[s] = lay.off[s];
        for (int s = 0; s < S_COUNT; ++s) tb.splits[s] = seg_splits_host[s];
        reduce_partials_kernel<double><<<(n + 255) / 256, 256, 0, st>>>(tb, (const double*)partials, n, (double*)grad);
    } else if (dtype == NMB_F32) {
        SegTable<float> tb;
        for (int s = 0; s <= S_COUNT; ++s) tb.off[s] = lay.off[s];
        for (int s = 0; s < S_COUNT; ++s) tb.splits[s] = seg_splits_host[s];
        reduce_partials_kernel<float><<<(n + 255) / 256, 256, 0, st>>>(tb, (const float*)partials, n, (float*)grad);
    } else {
        set_error("nmb_policy_reduce_partials: dtype %d", dtype);
        return NMB_EINVAL;
    }
    NMB_LAUNCH_CHECK();
    return 0;
}
