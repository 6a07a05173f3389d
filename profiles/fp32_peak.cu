// fp32_peak.cu -- FFMA throughput microbenchmark: the FP32 roofline denominator of bench.py.
//
// MEASURED_PEAKS.json holds HBM and bf16 tensor peaks only; the blend kernels are bound by the FP32 /
// issue pipe, so their peak is measured here instead of derived from 148 SM x 128 lanes x 2 x clock.
// Every thread runs 16 independent FFMA chains (enough ILP to cover the 4-cycle FMA latency at 8 warps per
// scheduler); the grid is 8 CTAs x 256 threads per SM.  Built by __graft_entry__.build() into
// profiles/libfp32peak.so (measurement infrastructure, not part of the product library).
#include <cuda_runtime.h>

__global__ void __launch_bounds__(256) ffma_kernel(float* out, int iters, float a, float b) {
    float v[16];
#pragma unroll
    for (int k = 0; k < 16; k++) v[k] = (float)(threadIdx.x + k);
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int k = 0; k < 16; k++) v[k] = fmaf(v[k], a, b);
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 16; k++) s += v[k];
    if (s == 123.456f) out[0] = s;   // never true: keeps the chains alive
}

// returns TFLOP/s (2 flops per FFMA) of the best of `reps` launches, or a negative cudaError_t
extern "C" double eags_fp32_peak_tflops(int iters, int reps) {
    int dev = 0, sms = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return -1.0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    float* out = nullptr;
    if (cudaMalloc(&out, 4) != cudaSuccess) return -2.0;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int grid = sms * 8, block = 256;
    double best = 0.0;
    for (int r = 0; r < reps + 2; r++) {
        cudaEventRecord(e0);
        ffma_kernel<<<grid, block>>>(out, iters, 1.0000001f, 1e-7f);
        cudaEventRecord(e1);
        if (cudaEventSynchronize(e1) != cudaSuccess) { best = -3.0; break; }
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        const double tf = 2.0 * 16.0 * (double)iters * grid * block / (ms * 1e-3) / 1e12;
        if (r >= 2 && tf > best) best = tf;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(out);
    return best;
}
