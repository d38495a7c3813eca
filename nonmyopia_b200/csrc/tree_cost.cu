// nmb_tree_cost_fwd / nmb_tree_cost_bwd: accumulated spotlight cost along every
// root-to-leaf chain of a lookahead tree (reference acqfs.py:421-441 calling
// qCostFunction.forward acqfs.py:604-633 once per level).
//
// Tree layout (reference tensors [*nf_rev[:s], n_r, 1, d] flattened): level s holds
// cnt_s = n_r * prod(nf[:s]) points with the restart index fastest, so the level-s
// ancestor of leaf i is i % cnt_s and the parent of level-0 node i is root[i % n_root].
// The reference adds the level costs with broadcasting, i.e. every leaf carries the sum of
// the hinges on its own chain; the non-Markovian discount looks at that running sum, so the
// multiplier of a level-s move is a function of the level-s node alone.
//
// Forward : one thread per leaf walks its chain (<= L+1 moves of d-dim distance; the work
//           of the shared upper levels is recomputed instead of communicated - it is O(d)).
// Backward: one thread per node sums the leaf adjoints below it in a fixed order (no
//           atomics => bitwise deterministic) and emits d acc / d point for the move that
//           ends in the node and for the moves that start from it.
#include "cost.cuh"

namespace nmb {

constexpr int kMaxLevels = 64;

struct TreeShape {
    int n_levels;
    long long cnt[kMaxLevels];
    long long off[kMaxLevels + 1];
    long long n_root;
};

template <typename T, int DMAX>
__device__ __forceinline__ void load_pt(const T* __restrict__ p, int d, T (&v)[DMAX]) {
#pragma unroll
    for (int k = 0; k < DMAX; ++k) v[k] = k < d ? p[k] : T(0);
}

template <typename T, int DMAX>
__global__ void __launch_bounds__(256) tree_cost_fwd_kernel(const __grid_constant__ TreeShape ts, int d, CostDev<T> c,
                                                            const T* __restrict__ root, const T* __restrict__ pts,
                                                            const T* __restrict__ noise, T* __restrict__ acc_out,
                                                            T* __restrict__ mult_out) {
    const int nl = ts.n_levels;
    const long long n_leaf = ts.cnt[nl - 1];
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_leaf;
         i += (long long)gridDim.x * blockDim.x) {
        T prev[DMAX], cur[DMAX], dir[DMAX];
        load_pt<T, DMAX>(root + (i % ts.n_root) * d, d, prev);
        T acc = T(0);
        for (int s = 0; s < nl; ++s) {
            const long long j = i % ts.cnt[s];
            load_pt<T, DMAX>(pts + (ts.off[s] + j) * d, d, cur);
            const T dist = move_dist<T, DMAX>(c, d, cur, prev, dir);
            T h = cost_hinge<T>(c, dist, noise ? noise[ts.off[s] + j] : T(0));
            T mult = T(1);
            if (c.discount > T(0)) {
                if (acc + c.prev_cost + h > c.discount_threshold) mult = c.disc_mult;
                h *= mult;
            }
            // the reference starts the sum with the first level's cost (no 0 + h rounding)
            acc = s == 0 ? h : acc + h;
            if (mult_out && j == i) mult_out[ts.off[s] + j] = mult;   // written by the first descendant only
#pragma unroll
            for (int k = 0; k < DMAX; ++k) prev[k] = cur[k];
        }
        acc_out[i] = acc;
    }
}

// sum of the leaf adjoints below node (s, j): leaves l == j (mod cnt_s), ascending
template <typename T>
__device__ __forceinline__ T leaf_sum(const TreeShape& ts, int s, long long j, const T* __restrict__ g) {
    const long long n_leaf = ts.cnt[ts.n_levels - 1];
    T v = T(0);
    for (long long l = j; l < n_leaf; l += ts.cnt[s]) v += g[l];
    return v;
}

template <typename T, int DMAX>
__global__ void __launch_bounds__(256) tree_cost_bwd_kernel(const __grid_constant__ TreeShape ts, int d, CostDev<T> c,
                                                            const T* __restrict__ root, const T* __restrict__ pts,
                                                            const T* __restrict__ mult, const T* __restrict__ g_acc,
                                                            T* __restrict__ g_pts, T* __restrict__ g_root,
                                                            long long first_node, int skip_root) {
    const int nl = ts.n_levels;
    const long long n_nodes = ts.off[nl];
    // nodes [first_node, n_nodes) = tree nodes of the lower levels; [n_nodes, n_nodes + n_root) = root points (the
    // upper levels and - when they have many leaves below them - the root points go to tree_cost_bwd_heavy_kernel)
    const long long t_end = n_nodes + (skip_root ? 0 : ts.n_root);
    for (long long t = first_node + (long long)blockIdx.x * blockDim.x + threadIdx.x; t < t_end;
         t += (long long)gridDim.x * blockDim.x) {
        int s = -1;
        long long j = t - n_nodes;
        if (t < n_nodes) {
            s = 0;
            while (t >= ts.off[s + 1]) ++s;
            j = t - ts.off[s];
        }
        T me[DMAX], other[DMAX], dir[DMAX], g[DMAX];
#pragma unroll
        for (int k = 0; k < DMAX; ++k) g[k] = T(0);
        load_pt<T, DMAX>(s < 0 ? root + j * d : pts + (ts.off[s] + j) * d, d, me);
        // the move that ends in this node
        if (s >= 0) {
            const T G = leaf_sum<T>(ts, s, j, g_acc);
            const T* pp = s == 0 ? root + (j % ts.n_root) * d : pts + (ts.off[s - 1] + j % ts.cnt[s - 1]) * d;
            load_pt<T, DMAX>(pp, d, other);
            const T dist = move_dist<T, DMAX>(c, d, me, other, dir);
            if (c.k * (dist - c.radius) > T(0)) {
                const T w = G * c.k * mult[ts.off[s] + j];
#pragma unroll
                for (int k = 0; k < DMAX; ++k) g[k] = fma(w, dir[k], g[k]);
            }
        }
        // the moves that start from this node
        if (s + 1 < nl) {
            const long long step = s < 0 ? ts.n_root : ts.cnt[s];
            for (long long ch = j; ch < ts.cnt[s + 1]; ch += step) {
                const T G = leaf_sum<T>(ts, s + 1, ch, g_acc);
                load_pt<T, DMAX>(pts + (ts.off[s + 1] + ch) * d, d, other);
                const T dist = move_dist<T, DMAX>(c, d, other, me, dir);
                if (c.k * (dist - c.radius) > T(0)) {
                    const T w = G * c.k * mult[ts.off[s + 1] + ch];
#pragma unroll
                    for (int k = 0; k < DMAX; ++k) g[k] = fma(-w, dir[k], g[k]);
                }
            }
        }
        T* dst = s < 0 ? (g_root ? g_root + j * d : nullptr) : g_pts + (ts.off[s] + j) * d;
        if (dst) {
#pragma unroll
            for (int k = 0; k < DMAX; ++k) if (k < d) dst[k] = g[k];
        }
    }
}


// Upper levels of the tree (and the root points): a node there has hundreds of leaves below it, and the thread-per-node
// kernel above walks them one after the other (level-0 node of the [64, 4, 2] tree: 512 + 64 x 8 dependent-latency
// loads: 0.2 - 0.3 ms for a trivial amount of arithmetic, and no faster with fewer restarts).  Here a WARP owns a node:
// the leaf adjoints below it are summed lane-strided + shuffle tree (fixed order: deterministic), the moves that start
// from it are spread over the lanes when there are >= 32 children, otherwise walked one by one with lane-parallel leaf
// sums.  Items: nodes [0, off[h_levels]) then (root_heavy) the n_root root points.
template <typename T, int DMAX>
__global__ void __launch_bounds__(256) tree_cost_bwd_heavy_kernel(const __grid_constant__ TreeShape ts, int d, CostDev<T> c,
                                                                  const T* __restrict__ root, const T* __restrict__ pts,
                                                                  const T* __restrict__ mult, const T* __restrict__ g_acc,
                                                                  T* __restrict__ g_pts, T* __restrict__ g_root,
                                                                  int h_levels, int root_heavy) {
    const int nl = ts.n_levels;
    const int lane = threadIdx.x & 31;
    const long long warp0 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    const long long n_heavy = ts.off[h_levels];
    const long long n_items = n_heavy + (root_heavy ? ts.n_root : 0);
    const long long n_leaf = ts.cnt[nl - 1];
    for (long long item = warp0; item < n_items; item += nwarps) {
        int s = -1;
        long long j = item - n_heavy;
        if (item < n_heavy) {
            s = 0;
            while (item >= ts.off[s + 1]) ++s;
            j = item - ts.off[s];
        }
        T me[DMAX], other[DMAX], dir[DMAX], g[DMAX];
#pragma unroll
        for (int k = 0; k < DMAX; ++k) g[k] = T(0);
        load_pt<T, DMAX>(s < 0 ? root + j * d : pts + (ts.off[s] + j) * d, d, me);
        // the move that ends in this node
        if (s >= 0) {
            T G = T(0);
            for (long long l = j + (long long)lane * ts.cnt[s]; l < n_leaf; l += 32 * ts.cnt[s]) G += g_acc[l];
            G = warp_sum(G);
            if (lane == 0) {
                const T* pp = s == 0 ? root + (j % ts.n_root) * d : pts + (ts.off[s - 1] + j % ts.cnt[s - 1]) * d;
                load_pt<T, DMAX>(pp, d, other);
                const T dist = move_dist<T, DMAX>(c, d, me, other, dir);
                if (c.k * (dist - c.radius) > T(0)) {
                    const T w = G * c.k * mult[ts.off[s] + j];
#pragma unroll
                    for (int k = 0; k < DMAX; ++k) g[k] = fma(w, dir[k], g[k]);
                }
            }
        }
        // the moves that start from this node
        if (s + 1 < nl) {
            const long long step = s < 0 ? ts.n_root : ts.cnt[s];
            const long long nch = ts.cnt[s + 1] / step;
            if (nch >= 32) {
                for (long long q = lane; q < nch; q += 32) {
                    const long long ch = j + q * step;
                    const T G = leaf_sum<T>(ts, s + 1, ch, g_acc);
                    load_pt<T, DMAX>(pts + (ts.off[s + 1] + ch) * d, d, other);
                    const T dist = move_dist<T, DMAX>(c, d, other, me, dir);
                    if (c.k * (dist - c.radius) > T(0)) {
                        const T w = G * c.k * mult[ts.off[s + 1] + ch];
#pragma unroll
                        for (int k = 0; k < DMAX; ++k) g[k] = fma(-w, dir[k], g[k]);
                    }
                }
            } else {
                for (long long q = 0; q < nch; ++q) {
                    const long long ch = j + q * step;
                    T G = T(0);
                    for (long long l = ch + (long long)lane * ts.cnt[s + 1]; l < n_leaf; l += 32 * ts.cnt[s + 1]) G += g_acc[l];
                    G = warp_sum(G);
                    if (lane == 0) {
                        load_pt<T, DMAX>(pts + (ts.off[s + 1] + ch) * d, d, other);
                        const T dist = move_dist<T, DMAX>(c, d, other, me, dir);
                        if (c.k * (dist - c.radius) > T(0)) {
                            const T w = G * c.k * mult[ts.off[s + 1] + ch];
#pragma unroll
                            for (int k = 0; k < DMAX; ++k) g[k] = fma(-w, dir[k], g[k]);
                        }
                    }
                }
            }
        }
#pragma unroll
        for (int k = 0; k < DMAX; ++k) g[k] = warp_sum(g[k]);
        T* dst = s < 0 ? (g_root ? g_root + j * d : nullptr) : g_pts + (ts.off[s] + j) * d;
        if (dst && lane == 0) {
#pragma unroll
            for (int k = 0; k < DMAX; ++k) if (k < d) dst[k] = g[k];
        }
    }
}

static int make_shape(int n_levels, const int64_t* counts, int64_t n_root, TreeShape& ts) {
    NMB_REQUIRE(n_levels >= 1 && n_levels <= kMaxLevels && counts, NMB_EINVAL, "tree cost: n_levels=%d", n_levels);
    NMB_REQUIRE(n_root >= 1, NMB_EINVAL, "tree cost: n_root=%lld", (long long)n_root);
    ts.n_levels = n_levels;
    ts.n_root = n_root;
    ts.off[0] = 0;
    for (int s = 0; s < n_levels; ++s) {
        const long long below = s == 0 ? n_root : counts[s - 1];
        NMB_REQUIRE(counts[s] >= below && counts[s] % below == 0, NMB_EINVAL,
                    "tree cost: level %d has %lld nodes, not a multiple of %lld", s, (long long)counts[s], below);
        ts.cnt[s] = counts[s];
        ts.off[s + 1] = ts.off[s] + counts[s];
    }
    return 0;
}

static unsigned grid_for(long long n) {
    long long b = (n + 255) / 256;
    const long long cap = (long long)sm_count() * 8;
    return (unsigned)(b < 1 ? 1 : (b > cap ? cap : b));
}

}  // namespace nmb

using namespace nmb;

#define NMB_TC_DISPATCH(KERN, T, d, grid, st, ...)                    \
    do {                                                              \
        if ((d) <= 2) KERN<T, 2><<<grid, 256, 0, st>>>(__VA_ARGS__);  \
        else if ((d) <= 4) KERN<T, 4><<<grid, 256, 0, st>>>(__VA_ARGS__); \
        else if ((d) <= 8) KERN<T, 8><<<grid, 256, 0, st>>>(__VA_ARGS__); \
        else KERN<T, 16><<<grid, 256, 0, st>>>(__VA_ARGS__);          \
    } while (0)

extern "C" int nmb_tree_cost_fwd(int dtype, int d, int n_levels, const int64_t* counts_host, const nmb_cost_t* cost,
                                 const void* root, int64_t n_root, const void* pts, const void* noise, void* acc,
                                 void* mult, void* stream) {
    NMB_REQUIRE(cost && root && pts && acc, NMB_EINVAL, "nmb_tree_cost_fwd: null pointer");
    NMB_REQUIRE(d >= 1 && d <= NMB_MAX_DIM, NMB_ELIMIT, "nmb_tree_cost_fwd: d=%d", d);
    TreeShape ts;
    int rc = make_shape(n_levels, counts_host, n_root, ts);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    const unsigned grid = grid_for(ts.cnt[n_levels - 1]);
    if (dtype == NMB_F64)
        NMB_TC_DISPATCH(tree_cost_fwd_kernel, double, d, grid, st, ts, d, to_dev<double>(*cost), (const double*)root,
                        (const double*)pts, (const double*)noise, (double*)acc, (double*)mult);
    else if (dtype == NMB_F32)
        NMB_TC_DISPATCH(tree_cost_fwd_kernel, float, d, grid, st, ts, d, to_dev<float>(*cost), (const float*)root,
                        (const float*)pts, (const float*)noise, (float*)acc, (float*)mult);
    else { set_error("nmb_tree_cost_fwd: dtype %d", dtype); return NMB_EINVAL; }
    NMB_LAUNCH_CHECK();
    return 0;
}

extern "C" int nmb_tree_cost_bwd(int dtype, int d, int n_levels, const int64_t* counts_host, const nmb_cost_t* cost,
                                 const void* root, int64_t n_root, const void* pts, const void* mult,
                                 const void* g_acc, void* g_pts, void* g_root, void* stream) {
    NMB_REQUIRE(cost && root && pts && mult && g_acc && g_pts, NMB_EINVAL, "nmb_tree_cost_bwd: null pointer");
    NMB_REQUIRE(d >= 1 && d <= NMB_MAX_DIM, NMB_ELIMIT, "nmb_tree_cost_bwd: d=%d", d);
    TreeShape ts;
    int rc = make_shape(n_levels, counts_host, n_root, ts);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    // upper levels (>= 64 leaves below a node) and, with as many leaves per root point, the roots: warp per node
    const long long n_leaf = ts.cnt[n_levels - 1];
    int h_levels = 0;
    while (h_levels < n_levels - 1 && n_leaf / ts.cnt[h_levels] >= 64) ++h_levels;
    const int root_heavy = n_leaf / n_root >= 64 ? 1 : 0;
    const long long first_light = ts.off[h_levels];
    const long long n_heavy = first_light + (root_heavy ? n_root : 0);
    const long long n_light = ts.off[n_levels] - first_light + (root_heavy ? 0 : n_root);
    const unsigned grid = grid_for(n_light);
    const unsigned grid_h = grid_for(n_heavy * 32);
    if (dtype == NMB_F64) {
        if (n_heavy > 0)
            NMB_TC_DISPATCH(tree_cost_bwd_heavy_kernel, double, d, grid_h, st, ts, d, to_dev<double>(*cost), (const double*)root,
                            (const double*)pts, (const double*)mult, (const double*)g_acc, (double*)g_pts,
                            (double*)g_root, h_levels, root_heavy);
        if (n_light > 0)
            NMB_TC_DISPATCH(tree_cost_bwd_kernel, double, d, grid, st, ts, d, to_dev<double>(*cost), (const double*)root,
                            (const double*)pts, (const double*)mult, (const double*)g_acc, (double*)g_pts,
                            (double*)g_root, first_light, root_heavy);
    } else if (dtype == NMB_F32) {
        if (n_heavy > 0)
            NMB_TC_DISPATCH(tree_cost_bwd_heavy_kernel, float, d, grid_h, st, ts, d, to_dev<float>(*cost), (const float*)root,
                            (const float*)pts, (const float*)mult, (const float*)g_acc, (float*)g_pts, (float*)g_root,
                            h_levels, root_heavy);
        if (n_light > 0)
            NMB_TC_DISPATCH(tree_cost_bwd_kernel, float, d, grid, st, ts, d, to_dev<float>(*cost), (const float*)root,
                            (const float*)pts, (const float*)mult, (const float*)g_acc, (float*)g_pts, (float*)g_root,
                            first_light, root_heavy);
    }
    else { set_error("nmb_tree_cost_bwd: dtype %d", dtype); return NMB_EINVAL; }
    NMB_LAUNCH_CHECK();
    return 0;
}
