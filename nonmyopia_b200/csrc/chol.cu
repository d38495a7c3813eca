// nmb_cholesky_root: lower Cholesky factor L of K + s2 I and its explicit inverse L^-1, one CTA per matrix
// (one warp per matrix for n <= 32), batched over grid.x.  Replaces the cuSOLVER potrf + cuBLAS trsm pair behind
// SingleTaskGP's prediction-strategy caches that the reference builds once per BO step (GPyTorch psd_safe_cholesky +
// the inverse root it multiplies by; reached from acqfs.py:258-266 / 364-373; SURVEY A.3, north_star item (b)).
//
// Blocked right-looking factorisation, block size 32, matrix in place in global memory (n <= 800: <= 5 MB, L2 resident):
//   1. diagonal block -> shared memory; one warp factors it (lane = row) and inverts it (lane = column);
//   2. panel  P = A[below, block] L_jj^-T: one thread per row against the shared inverse (broadcast reads), result kept
//      in shared memory (row stride 34: 16-byte aligned rows for the paired broadcast loads of step 3);
//   3. trailing update A -= P P^T on the lower triangle: a warp per 32 x 32 tile, lane = column holds its panel row in
//      registers, the other operand is broadcast from shared memory, stores are coalesced.
// Inverse, block row by block row: X_ij = -L_ii^-1 sum_{j <= k < i} L_ik X_kj with the L row panel staged in shared
// memory, a warp per 32 x 32 tile.  Fixed summation order: results do not depend on the launch.
// info[b] follows LAPACK potrf: 0, or k > 0 if the leading minor of order k is not positive definite.
#include <stdlib.h>
#include <algorithm>
#include "common.cuh"

namespace nmb {

constexpr int kCB = 32;        // block size
constexpr int kCD = 33;        // row stride of the diagonal-block buffers
constexpr int kCP = 34;        // row stride of the panel buffer
constexpr int kCholMaxN = 800; // GPyTorch's max_cholesky_size
constexpr int kCholMultiMinN = 128;   // one matrix of at least this order runs on the multi-CTA path

template <typename T>
__global__ void __launch_bounds__(256) chol_root_kernel(int n, T* __restrict__ A_all, long long lda, long long strideA,
                                                        T* __restrict__ X_all, long long ldx, long long strideX,
                                                        int* __restrict__ info_all) {
    extern __shared__ __align__(16) unsigned char chol_smem_raw[];
    T* D = reinterpret_cast<T*>(chol_smem_raw);      // [32][33] diagonal block / its factor
    T* Di = D + kCB * kCD;                           // [32][33] inverse of the diagonal factor
    T* P = Di + kCB * kCD + 2;                       // panel [rows][34] (16-byte aligned: 2 * 32 * 33 + 2 elements)
    __shared__ int fail;
    T* A = A_all + (size_t)blockIdx.x * strideA;
    T* X = X_all ? X_all + (size_t)blockIdx.x * strideX : nullptr;
    const int tid = threadIdx.x, nthr = blockDim.x, lane = tid & 31, warp = tid >> 5, nw = nthr >> 5;
    if (tid == 0) fail = 0;
    __syncthreads();

    for (int j0 = 0; j0 < n; j0 += kCB) {
        const int nb = min(kCB, n - j0);
        // ---- 1. diagonal block
        for (int e = tid; e < kCB * kCB; e += nthr) {
            const int r = e >> 5, c = e & 31;
            D[r * kCD + c] = (r < nb && c <= r) ? A[(size_t)(j0 + r) * lda + j0 + c] : T(0);
            Di[r * kCD + c] = T(0);
        }
        __syncthreads();
        if (warp == 0) {
            for (int k = 0; k < nb; ++k) {
                const T dkk = D[k * kCD + k];
                if (!(dkk > T(0)) || !(dkk < T(INFINITY))) {      // also catches NaN
                    if (lane == 0) fail = j0 + k + 1;
                    break;
                }
                const T s = Num<T>::sqrt(dkk);
                T lrk = T(0);
                if (lane == k) D[k * kCD + k] = s;
                if (lane > k && lane < nb) { lrk = D[lane * kCD + k] / s; D[lane * kCD + k] = lrk; }
                __syncwarp();
                for (int c = k + 1; c < nb; ++c)
                    if (c <= lane && lane < nb) D[lane * kCD + c] -= lrk * D[c * kCD + k];
                __syncwarp();
            }
        }
        __syncthreads();
        if (fail) break;
        if (warp == 0 && lane < nb) {
            // column `lane` of the inverse: L x = e_lane by forward substitution
            const int c = lane;
            Di[c * kCD + c] = T(1) / D[c * kCD + c];
            for (int r = c + 1; r < nb; ++r) {
                T acc = T(0);
                for (int t = c; t < r; ++t) acc = fma(D[r * kCD + t], Di[t * kCD + c], acc);
                Di[r * kCD + c] = -acc / D[r * kCD + r];
            }
        }
        __syncthreads();
        for (int e = tid; e < nb * nb; e += nthr) {
            const int r = e / nb, c = e - r * nb;
            A[(size_t)(j0 + r) * lda + j0 + c] = c <= r ? D[r * kCD + c] : T(0);
            if (X) X[(size_t)(j0 + r) * ldx + j0 + c] = c <= r ? Di[r * kCD + c] : T(0);
        }
        const int m = n - j0 - nb;                   // rows below the block
        if (m <= 0) break;
        // ---- 2. panel: P[r][:] = A[r][block] L_jj^-T   (P[r][c] = sum_{k <= c} Di[c][k] a[k])
        for (int r = tid; r < m; r += nthr) {
            T* arow = A + (size_t)(j0 + nb + r) * lda + j0;
            T a[kCB];
#pragma unroll
            for (int k = 0; k < kCB; ++k) a[k] = k < nb ? arow[k] : T(0);
#pragma unroll
            for (int c = 0; c < kCB; ++c) {
                T p = T(0);
#pragma unroll
                for (int k = 0; k <= c; ++k) p = fma(Di[c * kCD + k], a[k], p);
                if (c < nb) arow[c] = p;
                P[(size_t)r * kCP + c] = c < nb ? p : T(0);
            }
        }
        // zero the tail of the last partial tile so the tile loops need no row guards on reads
        const int mpad = (m + kCB - 1) / kCB * kCB;
        for (int e = m * kCP + tid; e < mpad * kCP; e += nthr) P[e] = T(0);
        __syncthreads();
        // ---- 3. trailing update on the lower triangle, warp per 32 x 32 tile
        const int nt = mpad / kCB, ntiles = nt * (nt + 1) / 2;
        for (int tile = warp; tile < ntiles; tile += nw) {
            int ti = 0;
            while ((ti + 1) * (ti + 2) / 2 <= tile) ++ti;
            const int tk = tile - ti * (ti + 1) / 2;
            T pk[kCB];
            const T* prow = P + (size_t)(tk * kCB + lane) * kCP;
#pragma unroll
            for (int c = 0; c < kCB; ++c) pk[c] = prow[c];
            const int col = tk * kCB + lane;
            for (int i = 0; i < kCB; ++i) {
                const int row = ti * kCB + i;
                if (row >= m) break;
                const T* pi = P + (size_t)row * kCP;
                T dot = T(0);
#pragma unroll
                for (int c = 0; c < kCB; ++c) dot = fma(pi[c], pk[c], dot);
                if (col <= row) A[(size_t)(j0 + nb + row) * lda + j0 + nb + col] -= dot;
            }
        }
        __syncthreads();
    }
    if (tid == 0 && info_all) info_all[blockIdx.x] = fail;
    if (fail) return;
    // strict upper triangle of L (still the input there) -> 0, like torch.linalg.cholesky
    for (long long e = tid; e < (long long)n * n; e += nthr) {
        const int r = (int)(e / n), c = (int)(e - (long long)r * n);
        if (c > r && (c / kCB) > (r / kCB)) A[(size_t)r * lda + c] = T(0);
    }
    if (!X) return;
    __syncthreads();
    // ---- inverse: off-diagonal blocks, block row by block row
    const int nblk = (n + kCB - 1) / kCB;
    for (int bi = 1; bi < nblk; ++bi) {
        const int i0 = bi * kCB, nbi = min(kCB, n - i0);
        const int ldp = i0 + 2;                      // L row panel [32][i0 (+2)] in the panel buffer
        for (int e = tid; e < kCB * i0; e += nthr) {
            const int r = e / i0, k = e - r * i0;
            P[(size_t)r * ldp + k] = r < nbi ? A[(size_t)(i0 + r) * lda + k] : T(0);
        }
        for (int e = tid; e < kCB * kCB; e += nthr) {
            const int r = e >> 5, c = e & 31;
            Di[r * kCD + c] = (r < nbi && c <= r) ? X[(size_t)(i0 + r) * ldx + i0 + c] : T(0);
        }
        __syncthreads();
        for (int bj = warp; bj < bi; bj += nw) {
            T acc[kCB];
#pragma unroll
            for (int r = 0; r < kCB; ++r) acc[r] = T(0);
            const int col = bj * kCB + lane;          // < i0 <= n always
            for (int k = bj * kCB; k < i0; ++k) {
                const T x = X[(size_t)k * ldx + col];   // rows k < i0 are final (earlier block rows / diagonal blocks)
                const T* lp = P + k;
#pragma unroll
                for (int r = 0; r < kCB; ++r) acc[r] = fma(lp[(size_t)r * ldp], x, acc[r]);
            }
#pragma unroll
            for (int r = 0; r < kCB; ++r) {
                T o = T(0);
#pragma unroll
                for (int t = 0; t <= r; ++t) o = fma(Di[r * kCD + t], acc[t], o);
                if (r < nbi) X[(size_t)(i0 + r) * ldx + col] = -o;
            }
        }
        __syncthreads();
    }
    for (long long e = tid; e < (long long)n * n; e += nthr) {
        const int r = (int)(e / n), c = (int)(e - (long long)r * n);
        if (c > r && (c / kCB) > (r / kCB)) X[(size_t)r * ldx + c] = T(0);
    }
}

// ------------------------------------------------------------------ one matrix over many CTAs
// The same blocked algorithm with the three phases of a block step as separate launches, so that the panel rows and the
// trailing tiles of ONE matrix spread over the SMs (the single-CTA kernel above takes 5 ms at n = 500; it stays the
// batched / small-n path).  Every element goes through the same FMA sequence as above: bit-identical results.
//   chol_diag_kernel   (1 warp)            factor + invert the diagonal block: L_jj -> A, L_jj^-1 -> X
//   chol_panel_kernel  (CTA per 128 rows)  P = A[below, block] L_jj^-T in place
//   chol_update_kernel (CTA per tile)      A[tile] -= P_i P_k^T on the lower triangle
//   chol_inverse_kernel (CTA per block column) off-diagonal blocks of L^-1, block row by block row
template <typename T>
__global__ void __launch_bounds__(32) chol_diag_kernel(int n, int j0, T* __restrict__ A, long long lda, T* __restrict__ X,
                                                       long long ldx, int* __restrict__ info) {
    // One warp, the block and its inverse in shared memory: literally step 1 of the single-CTA kernel above (same
    // loops, same expressions: bit-identical).  A register-resident variant of this kernel (rows in registers, columns
    // travelling by shuffle, fully unrolled) took 110 us per block - 1.8 of the 2.7 ms of a factorisation at n = 500;
    // this one is bound by ~500 dependent shared-memory FMAs of the forward substitution.
    __shared__ T D[kCB * kCD];
    __shared__ T Di[kCB * kCD];
    if (*info) return;
    const int lane = threadIdx.x, nb = min(kCB, n - j0);
    for (int e = lane; e < kCB * kCB; e += 32) {
        const int r = e >> 5, c = e & 31;
        D[r * kCD + c] = (r < nb && c <= r) ? A[(size_t)(j0 + r) * lda + j0 + c] : T(0);
        Di[r * kCD + c] = T(0);
    }
    __syncwarp();
    int fail = 0;
    for (int k = 0; k < nb; ++k) {
        const T dkk = D[k * kCD + k];
        if (!(dkk > T(0)) || !(dkk < T(INFINITY))) {      // also catches NaN (uniform over the warp)
            fail = j0 + k + 1;
            break;
        }
        const T s = Num<T>::sqrt(dkk);
        T lrk = T(0);
        __syncwarp();
        if (lane == k) D[k * kCD + k] = s;
        if (lane > k && lane < nb) { lrk = D[lane * kCD + k] / s; D[lane * kCD + k] = lrk; }
        __syncwarp();
        for (int c = k + 1; c < nb; ++c)
            if (c <= lane && lane < nb) D[lane * kCD + c] -= lrk * D[c * kCD + k];
        __syncwarp();
    }
    if (fail) { if (lane == 0) *info = fail; return; }
    if (lane < nb) {
        // column `lane` of the inverse: L x = e_lane by forward substitution
        const int c = lane;
        Di[c * kCD + c] = T(1) / D[c * kCD + c];
        for (int r = c + 1; r < nb; ++r) {
            T acc = T(0);
            for (int t = c; t < r; ++t) acc = fma(D[r * kCD + t], Di[t * kCD + c], acc);
            Di[r * kCD + c] = -acc / D[r * kCD + r];
        }
    }
    __syncwarp();
    for (int e = lane; e < nb * nb; e += 32) {
        const int r = e / nb, c = e - r * nb;
        A[(size_t)(j0 + r) * lda + j0 + c] = c <= r ? D[r * kCD + c] : T(0);
        X[(size_t)(j0 + r) * ldx + j0 + c] = c <= r ? Di[r * kCD + c] : T(0);
    }
}

template <typename T>
__global__ void __launch_bounds__(128) chol_panel_kernel(int n, int j0, T* __restrict__ A, long long lda,
                                                         const T* __restrict__ X, long long ldx, const int* __restrict__ info) {
    __shared__ T Di[kCB * kCD];
    if (*info) return;
    const int nb = min(kCB, n - j0), m = n - j0 - nb;
    for (int e = threadIdx.x; e < kCB * kCB; e += 128) {
        const int r = e >> 5, c = e & 31;
        Di[r * kCD + c] = (r < nb && c <= r) ? X[(size_t)(j0 + r) * ldx + j0 + c] : T(0);
    }
    __syncthreads();
    const int r = blockIdx.x * 128 + threadIdx.x;
    if (r >= m) return;
    T* arow = A + (size_t)(j0 + nb + r) * lda + j0;
    T a[kCB];
#pragma unroll
    for (int k = 0; k < kCB; ++k) a[k] = k < nb ? arow[k] : T(0);
#pragma unroll
    for (int c = 0; c < kCB; ++c) {
        T p = T(0);
#pragma unroll
        for (int k = 0; k <= c; ++k) p = fma(Di[c * kCD + k], a[k], p);
        if (c < nb) arow[c] = p;
    }
}

template <typename T>
__global__ void __launch_bounds__(256) chol_update_kernel(int n, int j0, T* __restrict__ A, long long lda,
                                                          const int* __restrict__ info) {
    __shared__ T Pi[kCB * kCD], Pk[kCB * kCD];
    if (*info) return;
    const int nb = min(kCB, n - j0), m = n - j0 - nb;
    int ti = 0;
    const int tile = blockIdx.x;
    while ((ti + 1) * (ti + 2) / 2 <= tile) ++ti;
    const int tk = tile - ti * (ti + 1) / 2;
    const T* base = A + (size_t)(j0 + nb) * lda + j0;
    for (int e = threadIdx.x; e < kCB * kCB; e += 256) {
        const int r = e >> 5, c = e & 31;
        const int ri = ti * kCB + r, rk = tk * kCB + r;
        Pi[r * kCD + c] = (ri < m && c < nb) ? base[(size_t)ri * lda + c] : T(0);
        Pk[r * kCD + c] = (rk < m && c < nb) ? base[(size_t)rk * lda + c] : T(0);
    }
    __syncthreads();
    const int lane = threadIdx.x & 31, wy = threadIdx.x >> 5;
    const int col = tk * kCB + lane;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const int i = wy * 4 + q, row = ti * kCB + i;
        if (row >= m) break;
        T dot = T(0);
#pragma unroll
        for (int c = 0; c < kCB; ++c) dot = fma(Pi[i * kCD + c], Pk[lane * kCD + c], dot);
        if (col <= row) A[(size_t)(j0 + nb + row) * lda + j0 + nb + col] -= dot;
    }
}

template <typename T>
__global__ void __launch_bounds__(256) chol_inverse_kernel(int n, const T* __restrict__ A, long long lda, T* __restrict__ X,
                                                           long long ldx, const int* __restrict__ info) {
    extern __shared__ __align__(16) unsigned char chol_inv_smem[];
    T* Di = reinterpret_cast<T*>(chol_inv_smem);     // [32][33]
    T* ACC = Di + kCB * kCD;                         // [32][33] column sums of the tile before the diagonal solve
    T* P = ACC + kCB * kCD + 2;                      // L row panel [32][ldp]
    if (*info) return;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int bj = blockIdx.x, k_lo = bj * kCB, nblk = (n + kCB - 1) / kCB;
    const int col = k_lo + lane;
    for (int bi = bj + 1; bi < nblk; ++bi) {
        const int i0 = bi * kCB, nbi = min(kCB, n - i0), w = i0 - k_lo, ldp = w + 2;
        for (int e = tid; e < kCB * w; e += 256) {
            const int r = e / w, k = e - r * w;
            P[(size_t)r * ldp + k] = r < nbi ? A[(size_t)(i0 + r) * lda + k_lo + k] : T(0);
        }
        for (int e = tid; e < kCB * kCB; e += 256) {
            const int r = e >> 5, c = e & 31;
            Di[r * kCD + c] = (r < nbi && c <= r) ? X[(size_t)(i0 + r) * ldx + i0 + c] : T(0);
        }
        __syncthreads();
        // warp w owns rows 4w..4w+3 of the tile; every element sums over k in the single-CTA kernel's order
        T acc[4] = {T(0), T(0), T(0), T(0)};
        for (int k8 = 0; k8 < w; k8 += 8) {          // w is a multiple of 32; 8 independent L2 loads in flight
            T x[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) x[u] = (col < n) ? X[(size_t)(k_lo + k8 + u) * ldx + col] : T(0);
#pragma unroll
            for (int u = 0; u < 8; ++u)
#pragma unroll
                for (int q = 0; q < 4; ++q) acc[q] = fma(P[(size_t)(warp * 4 + q) * ldp + k8 + u], x[u], acc[q]);
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) ACC[(warp * 4 + q) * kCD + lane] = acc[q];
        __syncthreads();
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int r = warp * 4 + q;
            T o = T(0);
            for (int t = 0; t <= r; ++t) o = fma(Di[r * kCD + t], ACC[t * kCD + lane], o);
            if (r < nbi && col < n) X[(size_t)(i0 + r) * ldx + col] = -o;
        }
        __syncthreads();          // the block row just written is read (from global) by the next bi of this CTA
    }
}

template <typename T>
__global__ void chol_zero_upper_kernel(int n, T* __restrict__ A, long long lda, T* __restrict__ X, long long ldx,
                                       const int* __restrict__ info) {
    if (*info) return;
    for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < (long long)n * n;
         e += (long long)gridDim.x * blockDim.x) {
        const int r = (int)(e / n), c = (int)(e - (long long)r * n);
        if (c > r && (c / kCB) > (r / kCB)) { A[(size_t)r * lda + c] = T(0); X[(size_t)r * ldx + c] = T(0); }
    }
}

template <typename T>
static int launch_chol_multi(int n, void* A_, int64_t lda, void* X_, int64_t ldx, int32_t* info, cudaStream_t st) {
    T* A = (T*)A_;
    T* X = (T*)X_;
    NMB_CUDA(cudaMemsetAsync(info, 0, sizeof(int32_t), st));
    for (int j0 = 0; j0 < n; j0 += kCB) {
        const int nb = std::min(kCB, n - j0), m = n - j0 - nb;
        chol_diag_kernel<T><<<1, 32, 0, st>>>(n, j0, A, lda, X, ldx, info);
        if (m <= 0) break;
        chol_panel_kernel<T><<<(m + 127) / 128, 128, 0, st>>>(n, j0, A, lda, X, ldx, info);
        const int nt = (m + kCB - 1) / kCB;
        chol_update_kernel<T><<<nt * (nt + 1) / 2, 256, 0, st>>>(n, j0, A, lda, info);
    }
    const int nblk = (n + kCB - 1) / kCB;
    if (nblk > 1) {
        const size_t smem = ((size_t)2 * kCB * kCD + 2 + (size_t)kCB * (n + 2)) * sizeof(T);
        NMB_CUDA(cudaFuncSetAttribute(chol_inverse_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        chol_inverse_kernel<T><<<nblk - 1, 256, smem, st>>>(n, A, lda, X, ldx, info);
    }
    chol_zero_upper_kernel<T><<<64, 256, 0, st>>>(n, A, lda, X, ldx, info);
    NMB_LAUNCH_CHECK();
    return 0;
}

template <typename T>
static int launch_chol(int n, int batch, void* A, int64_t lda, int64_t strideA, void* X, int64_t ldx, int64_t strideX,
                       int32_t* info, cudaStream_t st) {
    // one large matrix with its inverse: spread over the SMs (bit-identical to the single-CTA kernel)
    if (batch == 1 && n >= kCholMultiMinN && X && info && !getenv("NMB_CHOL_SINGLE_CTA"))
        return launch_chol_multi<T>(n, A, lda, X, ldx, info, st);
    const int mpad = n > kCB ? (n - 1) / kCB * kCB : 0;                      // panel rows, rounded up to tiles
    const size_t panel = std::max<size_t>((size_t)mpad * kCP, (size_t)kCB * (n + 2));
    const size_t smem = ((size_t)2 * kCB * kCD + 2 + panel) * sizeof(T);
    NMB_REQUIRE(smem <= 227 * 1024, NMB_ELIMIT, "nmb_cholesky_root: n=%d does not fit shared memory", n);
    NMB_CUDA(cudaFuncSetAttribute(chol_root_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int threads = n <= kCB ? 32 : 256;
    chol_root_kernel<T><<<batch, threads, smem, st>>>(n, (T*)A, lda, strideA, (T*)X, ldx, strideX, info);
    NMB_LAUNCH_CHECK();
    return 0;
}

}  // namespace nmb

using namespace nmb;

extern "C" int nmb_cholesky_root(int dtype, int n, int batch, void* A, int64_t lda, int64_t stride_a, void* Linv,
                                 int64_t ldi, int64_t stride_i, int32_t* info, void* stream) {
    NMB_REQUIRE(A && n >= 1 && batch >= 0 && lda >= n, NMB_EINVAL, "nmb_cholesky_root: bad arguments");
    NMB_REQUIRE(n <= kCholMaxN, NMB_ELIMIT, "nmb_cholesky_root: n=%d above %d (GPyTorch switches to CG there)", n,
                kCholMaxN);
    NMB_REQUIRE(!Linv || ldi >= n, NMB_EINVAL, "nmb_cholesky_root: ldi=%lld < n", (long long)ldi);
    NMB_REQUIRE(batch <= 1 || (stride_a >= (int64_t)n * lda && (!Linv || stride_i >= (int64_t)n * ldi)), NMB_EINVAL,
                "nmb_cholesky_root: batch strides overlap");
    if (batch == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == NMB_F64) return launch_chol<double>(n, batch, A, lda, stride_a, Linv, ldi, stride_i, info, st);
    if (dtype == NMB_F32) return launch_chol<float>(n, batch, A, lda, stride_a, Linv, ldi, stride_i, info, st);
    set_error("nmb_cholesky_root: dtype %d", dtype);
    return NMB_EINVAL;
}
