/*
 * nonmyopia_b200 - C ABI of the B200-native lookahead-acquisition hot path.
 *
 * The reference (sangttruong/nonmyopia, synthetic_experiment/) is pure Python and
 * has no FFI of its own; its de-facto boundary is the Python class API of
 * acqfs.py / actor.py / amortized_network.py.  The host-side mirror
 * of those classes lives in nonmyopia_b200/{acqfs,actor,amortized_network}.py and
 * reaches the GPU only through the entry points declared here (ctypes binding in
 * nonmyopia_b200/_lib.py; the C++ torch.library layer in csrc/torch_ops.cpp calls
 * the same symbols).  Each entry point names the reference code it replaces.
 *
 * Conventions
 *  - every pointer is a DEVICE pointer unless the name says host; tensors are dense
 *    row-major; `dtype` is NMB_F64 or NMB_F32 and applies to every floating tensor
 *    of that call;
 *  - `stream` is a cudaStream_t passed as void* (0 = legacy default stream);
 *  - functions are asynchronous on `stream`, never allocate or free device memory
 *    (workspaces are caller-provided; nmb_*_workspace_bytes tell the size) and
 *    return 0 on success, a negative NMB_E* code for bad arguments, or a positive
 *    cudaError_t; nmb_last_error() gives the message for the calling thread.
 *  - no CPU fallback exists: without a CUDA device every compute entry point fails.
 */
#ifndef NONMYOPIA_B200_H
#define NONMYOPIA_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NMB_F64 0
#define NMB_F32 1

/* stationary kernels of the GP surrogate (reference main.py:86-92) */
#define NMB_KERNEL_RBF      0
#define NMB_KERNEL_MATERN12 1
#define NMB_KERNEL_MATERN32 2
#define NMB_KERNEL_MATERN52 3

#define NMB_EINVAL  (-1)   /* bad argument (shape, dtype, null pointer) */
#define NMB_ELIMIT  (-2)   /* a size exceeds what the kernels were built for */
#define NMB_ENODEV  (-3)   /* no CUDA device */

#define NMB_MAX_DIM 16     /* input dimension d <= 16 (reference envs: 2,4,6,8) */

const char* nmb_last_error(void);
int         nmb_version(void);
/* number of SMs of the current device (grid sizing), <0 on error */
int         nmb_device_sm_count(void);

/* ------------------------------------------------------------------------
 * Thompson-sample path  f(x) = c + feat_scale * [sin(z), cos(z)] . w_s + k(x, X_tr) . v_s
 * with z = W (x / lengthscale).  Replaces the BoTorch MatheronPath call
 * `self.f(X)` at reference acqfs.py:362 and :409-411.
 * The host prepares the constants once per BO step (dump_model, acqfs.py:258-266).
 * ---------------------------------------------------------------------- */
typedef struct nmb_path {
    int32_t d;            /* input dim */
    int32_t n_freq;       /* M/2 spectral frequencies */
    int32_t n_train;      /* n GP training points */
    int32_t n_paths;      /* S jointly drawn paths */
    int32_t kernel_id;    /* NMB_KERNEL_* */
    int32_t _pad;
    double  feat_scale;   /* sqrt(outputscale * 2 / M) */
    double  amplitude;    /* outputscale (1 for a bare kernel) */
    double  mean_const;   /* constant mean c */
    const void* w_phase;  /* [n_freq, d]  W[m,k] / (lengthscale[k] * pi): phase in half turns */
    const void* w_feat;   /* [n_paths, 2*n_freq] feature weights (sin block, then cos block) */
    const void* inv_ls;   /* [d] 1 / lengthscale (isotropic kernels repeat the value) */
    const void* train_x;  /* [n_train, d] */
    const void* v_update; /* [n_paths, n_train] pathwise update weights K^-1 (y - prior - eps) */
} nmb_path_t;

/* y[i] = f_{s(i)}(x[i]) and dydx[i,:] = grad_x f (either output may be NULL).
 * Point i uses path s(i) = i / points_per_path (points_per_path = N for one path).
 * The backward of the op is gx = gy * dydx (elementwise), so one launch serves
 * forward and backward. */
int nmb_path_eval(const nmb_path_t* path, int dtype, const void* x, int64_t n_points,
                  int64_t points_per_path, void* y, void* dydx, void* stream);

/* ------------------------------------------------------------------------
 * FP64 / FP32 pipe micro-benchmarks used for the roofline denominators
 * (MEASURED_PEAKS.json has no FP64 entry).
 * which: 0 = dependent-chain-free FMA throughput, 1 = sincospi, 2 = exp.
 * Writes the achieved operations/second (FMA counted as 2 flops for which=0,
 * one op per call otherwise) to *host_result. Synchronous.
 * ---------------------------------------------------------------------- */
int nmb_pipe_peak(int dtype, int which, double* host_result);

#ifdef __cplusplus
}
#endif
#endif /* NONMYOPIA_B200_H */
