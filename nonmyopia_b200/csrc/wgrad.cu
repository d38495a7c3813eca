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

static int wgrad_chunk(int L, long long N) {
    const long long nodes = (long long)(L + 1) * N;
    long long ch = (nodes + 31) / 32;          // at most 32 splits per segment
    ch = ((ch + kNC - 1) / kNC) * kNC;
    if (ch < 64) ch = 64;
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
    WgArgs<T> a;
    memset(&a, 0, sizeof(a));
    a.chunk = wgrad_chunk(L, N);
    a.n_params = lay.total();
    a.partials = (T*)partials;
    a.done_flag = p->done_flag;
    const long long all0 = 0, all1 = (long long)(L + 1) * N;
    const long long x0 = 0, x1 = (long long)L * N, a0 = (long long)L * N, a1 = all1;
    int nt = 0, item = 0;
    for (int s = 0; s < S_COUNT; ++s) seg_splits[s] = 0;
    auto add = [&](const T* dY, int ldY, int o0, int On, const T* A, int ldA, int k0, int Kn, long long nb,
                   long long ne, int wseg, int orow0, int kcol0, int bseg) {
        if (ne <= nb) return;
        WgTask<T>& t = a.t[nt++];
        t.dY = dY; t.ldY = ldY; t.o0 = o0; t.On = On;
        t.A = A; t.ldA = ldA; t.k0 = k0; t.Kn = Kn;
        t.n_begin = nb; t.n_end = ne;
        t.ldK = lay.K[wseg];
        t.dst = lay.off[wseg] + orow0 * t.ldK + kcol0;
        t.bias_dst = bseg >= 0 ? lay.off[bseg] + orow0 : -1;
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
    // bias rows of the gate blocks: offsets inside bih/bhh follow orow0 (handled by bias_dst above)
    const size_t smem = 4 * kNC * 64 * sizeof(T);
    NMB_CUDA(cudaFuncSetAttribute(wgrad_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    wgrad_kernel<T><<<item, 256, smem, st>>>(a);
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
