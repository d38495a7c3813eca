// Pipe micro-benchmarks that decide the register / shared-memory tiling of the FP64 policy GEMMs:
//   1. DFMA vs FP64 tensor-core mma.sync (m8n8k4, m16n8k8) throughput per SM
//   2. cost of LDS.128 / LDS.64 / LDS.32 with broadcast patterns (shared data-pipe cycles / instr)
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/microbench.bin tools/microbench.cu
#include <cstdio>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); return 1; } } while (0)

__device__ __forceinline__ void dmma884(double (&c)[2], double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(c[0]), "+d"(c[1]) : "d"(a), "d"(b));
}
__device__ __forceinline__ void dmma1688(double (&c)[4], const double (&a)[4], const double (&b)[2]) {
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
                 : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3])
                 : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(b[0]), "d"(b[1]));
}
__device__ __forceinline__ void dmma16816(double (&c)[4], const double (&a)[8], const double (&b)[4]) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7,%8,%9,%10,%11}, {%12,%13,%14,%15}, {%0,%1,%2,%3};\n"
                 : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3])
                 : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(a[4]), "d"(a[5]), "d"(a[6]), "d"(a[7]),
                   "d"(b[0]), "d"(b[1]), "d"(b[2]), "d"(b[3]));
}

// WHICH: 0 DFMA (8 chains), 1 m8n8k4 (8 acc tiles), 2 m16n8k8 (4 acc tiles), 3 m16n8k16 (4 acc tiles)
template <int WHICH> __global__ void __launch_bounds__(256) math_kernel(double* out, int iters) {
    const double a0 = 1.0 + threadIdx.x * 1e-9, b0 = 0.999999;
    double r = 0;
    if (WHICH == 0) {
        double c[8];
        for (int i = 0; i < 8; ++i) c[i] = i * 0.001;
        for (int it = 0; it < iters; ++it)
#pragma unroll
            for (int i = 0; i < 8; ++i) c[i] = fma(c[i], b0, a0);
        for (int i = 0; i < 8; ++i) r += c[i];
    } else if (WHICH == 1) {
        double c[8][2];
        for (int i = 0; i < 8; ++i) c[i][0] = c[i][1] = i * 0.001;
        for (int it = 0; it < iters; ++it)
#pragma unroll
            for (int i = 0; i < 8; ++i) dmma884(c[i], a0, b0);
        for (int i = 0; i < 8; ++i) r += c[i][0] + c[i][1];
    } else if (WHICH == 2) {
        double c[4][4], a[4] = {a0, a0, a0, a0}, b[2] = {b0, b0};
        for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) c[i][j] = i * 0.001;
        for (int it = 0; it < iters; ++it)
#pragma unroll
            for (int i = 0; i < 4; ++i) dmma1688(c[i], a, b);
        for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) r += c[i][j];
    } else {
        double c[4][4], a[8] = {a0, a0, a0, a0, a0, a0, a0, a0}, b[4] = {b0, b0, b0, b0};
        for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) c[i][j] = i * 0.001;
        for (int it = 0; it < iters; ++it)
#pragma unroll
            for (int i = 0; i < 4; ++i) dmma16816(c[i], a, b);
        for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) r += c[i][j];
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

// DFMA and DMMA issued by different warps of the same SM sub-partitions: do they share one pipe?
__global__ void __launch_bounds__(256) mixed_kernel(double* out, int iters) {
    const double a0 = 1.0 + threadIdx.x * 1e-9, b0 = 0.999999;
    double r = 0;
    if ((threadIdx.x >> 5) & 4) {       // warps 4-7: one per sub-partition
        double c[8];
        for (int i = 0; i < 8; ++i) c[i] = i * 0.001;
        for (int it = 0; it < iters * 8; ++it)
#pragma unroll
            for (int i = 0; i < 8; ++i) c[i] = fma(c[i], b0, a0);
        for (int i = 0; i < 8; ++i) r += c[i];
    } else {                            // warps 0-3
        double c[8][2];
        for (int i = 0; i < 8; ++i) c[i][0] = c[i][1] = i * 0.001;
        for (int it = 0; it < iters; ++it)
#pragma unroll
            for (int i = 0; i < 8; ++i) dmma884(c[i], a0, b0);
        for (int i = 0; i < 8; ++i) r += c[i][0] + c[i][1];
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

static int run_mixed() {
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    const int blocks = sms * 4, threads = 256, iters = 4096;
    double* buf;
    CK(cudaMalloc(&buf, sizeof(double) * blocks * threads));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep) {
        CK(cudaEventRecord(e0));
        mixed_kernel<<<blocks, threads>>>(buf, iters);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        if (rep && ms < best) best = ms;
    }
    // per block: 4 warps x iters x 8 DMMA x 256 FMA  +  4 warps x iters*8 x 8 DFMA x 32 FMA  (equal FMA counts)
    const double fma = (double)blocks * 4.0 * iters * 8.0 * (256.0 + 8.0 * 32.0);
    printf("{\"bench\": \"dfma_and_dmma_concurrent\", \"ms\": %.4f, \"tflops\": %.3f}\n", best, 2.0 * fma / (best * 1e-3) / 1e12);
    cudaFree(buf);
    return 0;
}

// DMMA dependent-issue latency: 1 warp per SM sub-partition, NACC independent accumulator tiles
template <int NACC> __global__ void __launch_bounds__(128) dmma_chain_kernel(double* out, long long* cyc, int iters) {
    const double a0 = 1.0 + threadIdx.x * 1e-9, b0 = 0.999999;
    double c[NACC][2];
    for (int i = 0; i < NACC; ++i) c[i][0] = c[i][1] = i * 0.001;
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it)
#pragma unroll
        for (int i = 0; i < NACC; ++i) dmma884(c[i], a0, b0);
    const long long t1 = clock64();
    double r = 0;
    for (int i = 0; i < NACC; ++i) r += c[i][0] + c[i][1];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
template <int NACC> static int run_chain() {
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    double* buf; long long* cyc;
    CK(cudaMalloc(&buf, sizeof(double) * sms * 128));
    CK(cudaMalloc(&cyc, sizeof(long long) * sms));
    const int iters = 4096;
    for (int rep = 0; rep < 2; ++rep) dmma_chain_kernel<NACC><<<sms, 128>>>(buf, cyc, iters);
    CK(cudaDeviceSynchronize());
    long long h[4];
    CK(cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost));
    printf("{\"bench\": \"dmma_chain\", \"independent_tiles\": %d, \"cycles_per_dmma\": %.2f}\n", NACC,
           (double)h[0] / ((double)iters * NACC));
    cudaFree(buf); cudaFree(cyc);
    return 0;
}

// shared-load cost: PAT 0 all lanes one address, 1 = 8 distinct (lane>>2), 2 = 4 distinct (lane&3) stride 32B,
// 3 = all lanes distinct consecutive.  WIDTH 4/8/16 bytes.  Reports SM cycles per warp-instruction with 8 warps/SM.
template <int WIDTH, int PAT> __global__ void __launch_bounds__(256) lds_kernel(double* out, long long* cyc, int iters) {
    __shared__ __align__(16) double sm[4096];
    for (int i = threadIdx.x; i < 4096; i += blockDim.x) sm[i] = i * 1e-3;
    __syncthreads();
    const int lane = threadIdx.x & 31;
    int idx;   // in units of WIDTH bytes
    if (PAT == 0) idx = 0;
    else if (PAT == 1) idx = lane >> 2;
    else if (PAT == 2) idx = (lane & 3) * (32 / WIDTH);
    else idx = lane;
    const char* base = reinterpret_cast<const char*>(sm) + idx * WIDTH + (threadIdx.x >> 5) * 1024;
    double acc0 = 0, acc1 = 0;
    float facc = 0;
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const char* p = base + u * 512 + (it & 1) * 4096;
            const unsigned sa = (unsigned)__cvta_generic_to_shared(p);
            if (WIDTH == 16) { double x, y; asm volatile("ld.volatile.shared.v2.f64 {%0,%1}, [%2];" : "=d"(x), "=d"(y) : "r"(sa)); acc0 += x; acc1 += y; }
            else if (WIDTH == 8) { double x; asm volatile("ld.volatile.shared.f64 %0, [%1];" : "=d"(x) : "r"(sa)); acc0 += x; }
            else { float x; asm volatile("ld.volatile.shared.f32 %0, [%1];" : "=f"(x) : "r"(sa)); facc += x; }
        }
    }
    const long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc0 + acc1 + facc;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int WHICH> static int run_math(const char* name, double fma_per_inst_warp) {
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    const int blocks = sms * 4, threads = 256, iters = 4096;
    double* buf;
    CK(cudaMalloc(&buf, sizeof(double) * blocks * threads));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep) {
        CK(cudaEventRecord(e0));
        math_kernel<WHICH><<<blocks, threads>>>(buf, iters);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        if (rep && ms < best) best = ms;
    }
    const int n_inst = WHICH == 0 ? 8 : (WHICH == 1 ? 8 : 4);
    const double fma = (double)blocks * (threads / 32) * iters * n_inst * fma_per_inst_warp;
    printf("{\"bench\": \"%s\", \"ms\": %.4f, \"tflops\": %.3f}\n", name, best, 2.0 * fma / (best * 1e-3) / 1e12);
    cudaFree(buf);
    return 0;
}

template <int WIDTH, int PAT> static int run_lds() {
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    const int blocks = sms, threads = 256, iters = 2048;
    double* buf; long long* cyc;
    CK(cudaMalloc(&buf, sizeof(double) * blocks * threads));
    CK(cudaMalloc(&cyc, sizeof(long long) * blocks));
    for (int rep = 0; rep < 2; ++rep) lds_kernel<WIDTH, PAT><<<blocks, threads>>>(buf, cyc, iters);
    CK(cudaDeviceSynchronize());
    long long h[8];
    CK(cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost));
    const double inst = (double)iters * 8 * (threads / 32);     // warp-instructions per SM
    printf("{\"bench\": \"lds\", \"width\": %d, \"pattern\": %d, \"sm_cycles_per_warp_inst\": %.3f}\n", WIDTH, PAT,
           (double)h[0] / inst);
    cudaFree(buf); cudaFree(cyc);
    return 0;
}

int main() {
    if (run_math<0>("dfma", 32.0)) return 1;
    if (run_math<1>("dmma_m8n8k4", 256.0)) return 1;
    if (run_math<2>("dmma_m16n8k8", 1024.0)) return 1;
    if (run_math<3>("dmma_m16n8k16", 2048.0)) return 1;
    if (run_mixed()) return 1;
    run_chain<1>(); run_chain<2>(); run_chain<4>(); run_chain<8>();
    run_lds<16, 0>(); run_lds<16, 1>(); run_lds<16, 2>(); run_lds<16, 3>();
    run_lds<8, 0>(); run_lds<8, 1>(); run_lds<8, 2>(); run_lds<8, 3>();
    run_lds<4, 0>(); run_lds<4, 3>();
    return 0;
}
