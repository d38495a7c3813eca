// N1 evidence: what do the FP32 options of the policy GEMMs cost on B200?
//   FFMA (the FP32 variant's register tiles today), legacy warp-level mma.sync.m16n8k8 TF32 (no 128-row tile needed,
//   fits the 8..56-row CTAs of the rollout), and - for scale - the FP64 DMMA the FP64 path uses.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/microbench_tf32.bin tools/microbench_tf32.cu
#include <cstdio>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); return 1; } } while (0)

__device__ __forceinline__ void mma_tf32(float (&c)[4], const unsigned (&a)[4], const unsigned (&b)[2]) {
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}
__device__ __forceinline__ void dmma884(double (&c)[2], double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(c[0]), "+d"(c[1]) : "d"(a), "d"(b));
}

// WHICH: 0 FFMA (16 chains), 1 TF32 m16n8k8 (NACC accumulator tiles), 2 DMMA m8n8k4 (8 tiles)
template <int WHICH, int NACC> __global__ void __launch_bounds__(256) k(float* out, int iters) {
    float r = 0;
    if (WHICH == 0) {
        float c[16];
        const float a0 = 1.0f + threadIdx.x * 1e-7f, b0 = 0.9999f;
        for (int i = 0; i < 16; ++i) c[i] = i * 0.001f;
        for (int it = 0; it < iters; ++it)
#pragma unroll
            for (int i = 0; i < 16; ++i) c[i] = fmaf(c[i], b0, a0);
        for (int i = 0; i < 16; ++i) r += c[i];
    } else if (WHICH == 1) {
        float c[NACC][4];
        unsigned a[4] = {0x3f800000u, 0x3f800000u, 0x3f800000u, 0x3f800000u}, b[2] = {0x3f7fe000u, 0x3f7fe000u};
        for (int i = 0; i < NACC; ++i) for (int j = 0; j < 4; ++j) c[i][j] = i * 0.001f;
        for (int it = 0; it < iters; ++it)
#pragma unroll
            for (int i = 0; i < NACC; ++i) mma_tf32(c[i], a, b);
        for (int i = 0; i < NACC; ++i) for (int j = 0; j < 4; ++j) r += c[i][j];
    } else {
        double c[8][2];
        for (int i = 0; i < 8; ++i) c[i][0] = c[i][1] = i * 0.001;
        for (int it = 0; it < iters; ++it)
#pragma unroll
            for (int i = 0; i < 8; ++i) dmma884(c[i], 1.0 + threadIdx.x * 1e-9, 0.999999);
        for (int i = 0; i < 8; ++i) r += (float)(c[i][0] + c[i][1]);
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

template <int WHICH, int NACC> static int run(const char* name, double fma_per_inst_warp, int n_inst) {
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    const int blocks = sms * 4, threads = 256, iters = 4096;
    float* buf;
    CK(cudaMalloc(&buf, sizeof(float) * blocks * threads));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep) {
        CK(cudaEventRecord(e0));
        k<WHICH, NACC><<<blocks, threads>>>(buf, iters);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        if (rep && ms < best) best = ms;
    }
    const double fma = (double)blocks * (threads / 32) * iters * n_inst * fma_per_inst_warp;
    printf("{\"bench\": \"%s\", \"ms\": %.4f, \"tflops\": %.3f}\n", name, best, 2.0 * fma / (best * 1e-3) / 1e12);
    cudaFree(buf);
    return 0;
}

int main() {
    if (run<0, 1>("ffma_fp32", 32.0, 16)) return 1;
    if (run<1, 2>("mma_sync_m16n8k8_tf32_2tiles", 1024.0, 2)) return 1;
    if (run<1, 4>("mma_sync_m16n8k8_tf32_4tiles", 1024.0, 4)) return 1;
    if (run<1, 8>("mma_sync_m16n8k8_tf32_8tiles", 1024.0, 8)) return 1;
    if (run<2, 8>("dmma_m8n8k4_fp64", 256.0, 8)) return 1;
    return 0;
}
