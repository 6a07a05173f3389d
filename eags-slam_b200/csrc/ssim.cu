// ssim.cu -- fused SSIM (11x11 Gaussian window, sigma 1.5, zero padding) forward + backward for the
// mapping colour loss (reference src/entities/losses.py:28-113, used at
// src/entities/mapper.py:162-163 as  (1 - lambda) L1 + lambda (1 - ssim(image, gt_image))).
//
// The reference runs five grouped conv2d's + ~15 elementwise kernels forward and their autograd
// backward (five transposed convolutions).  Here:
//   ssim_fwd_kernel : one CTA per 32x32 tile and channel stages the (32+10)^2 halo of both images in
//                     shared memory, runs the separable window once for the five moments (x, y, x^2,
//                     y^2, xy), forms the SSIM map in registers, block-reduces it into one fp64
//                     accumulator and stores the three per-pixel partials the backward needs
//                     (dS/dmu1, dS/ds11, dS/ds12).
//   ssim_bwd_kernel : dL/dimg1 = scale * (w * a + 2 img1 (w * b) + img2 (w * c)) -- the same separable
//                     window over the three partial maps; `scale` = upstream gradient / (C H W), read
//                     from device memory (no host sync).
// HBM-bound streaming with a 10-pixel halo; no tensor cores (an 11-tap separable stencil).
#include "kernels.h"

namespace {

constexpr int TS = 32;            // output tile
constexpr int R = 5;              // window radius
constexpr int HS = TS + 2 * R;    // halo tile
constexpr int SS_THREADS = 256;

struct SsimWindow { float g[2 * R + 1]; };

// separable pass over `NQ` staged quantities: horizontal into s_h, then vertical from s_h
template <int NQ>
__device__ __forceinline__ void horizontal_pass(const float (*s_in)[HS][HS + 1], float (*s_h)[HS][TS],
                                                const SsimWindow& w) {
    for (int o = threadIdx.x; o < HS * TS; o += SS_THREADS) {
        const int r = o / TS, c = o - r * TS;
        float acc[NQ];
#pragma unroll
        for (int q = 0; q < NQ; q++) acc[q] = 0.f;
#pragma unroll
        for (int k = 0; k <= 2 * R; k++) {
#pragma unroll
            for (int q = 0; q < NQ; q++) acc[q] = fmaf(w.g[k], s_in[q][r][c + k], acc[q]);
        }
#pragma unroll
        for (int q = 0; q < NQ; q++) s_h[q][r][c] = acc[q];
    }
}

__global__ void __launch_bounds__(SS_THREADS)
ssim_fwd_kernel(int C, int H, int W, const float* __restrict__ img1, const float* __restrict__ img2,
                SsimWindow w, float* __restrict__ part_a, float* __restrict__ part_b,
                float* __restrict__ part_c, double* __restrict__ sum) {
    __shared__ float s_in[2][HS][HS + 1];     // the two images; the three products are formed on the fly
    __shared__ float s_h[5][HS][TS];
    __shared__ double s_red[SS_THREADS / 32];
    const int ch = blockIdx.z, x0 = blockIdx.x * TS, y0 = blockIdx.y * TS;
    const float* p1 = img1 + (size_t)ch * H * W;
    const float* p2 = img2 + (size_t)ch * H * W;
    for (int o = threadIdx.x; o < HS * HS; o += SS_THREADS) {
        const int r = o / HS, c = o - r * HS;
        const int y = y0 + r - R, x = x0 + c - R;
        float a = 0.f, b = 0.f;   // zero padding (F.conv2d padding = window_size // 2)
        if (y >= 0 && y < H && x >= 0 && x < W) { a = p1[(size_t)y * W + x]; b = p2[(size_t)y * W + x]; }
        s_in[0][r][c] = a; s_in[1][r][c] = b;
    }
    __syncthreads();
    for (int o = threadIdx.x; o < HS * TS; o += SS_THREADS) {
        const int r = o / TS, c = o - r * TS;
        float acc[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int k = 0; k <= 2 * R; k++) {
            const float a = s_in[0][r][c + k], b = s_in[1][r][c + k], g = w.g[k];
            acc[0] = fmaf(g, a, acc[0]); acc[1] = fmaf(g, b, acc[1]);
            acc[2] = fmaf(g, a * a, acc[2]); acc[3] = fmaf(g, b * b, acc[3]); acc[4] = fmaf(g, a * b, acc[4]);
        }
#pragma unroll
        for (int q = 0; q < 5; q++) s_h[q][r][c] = acc[q];
    }
    __syncthreads();
    double local = 0.0;
    const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;
    for (int o = threadIdx.x; o < TS * TS; o += SS_THREADS) {
        const int r = o / TS, c = o - r * TS;
        const int y = y0 + r, x = x0 + c;
        float m[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int k = 0; k <= 2 * R; k++) {
#pragma unroll
            for (int q = 0; q < 5; q++) m[q] = fmaf(w.g[k], s_h[q][r + k][c], m[q]);
        }
        if (y < H && x < W) {
            const float mu1 = m[0], mu2 = m[1];
            const float mu1mu2 = mu1 * mu2, mu1sq = mu1 * mu1, mu2sq = mu2 * mu2;
            const float sig1 = m[2] - mu1sq, sig2 = m[3] - mu2sq, sig12 = m[4] - mu1mu2;
            const float A1 = 2.f * mu1mu2 + C1, A2 = 2.f * sig12 + C2;
            const float B1 = mu1sq + mu2sq + C1, B2 = sig1 + sig2 + C2;
            const float inv = 1.f / (B1 * B2);
            const float S = A1 * A2 * inv;
            local += (double)S;
            const size_t pix = (size_t)ch * H * W + (size_t)y * W + x;
            // partials of S w.r.t. the three moments img1 enters (mu1, s11 = w*x^2, s12 = w*xy)
            part_a[pix] = 2.f * mu2 * (A2 - A1) * inv - S * 2.f * mu1 * (1.f / B1 - 1.f / B2);
            part_b[pix] = -S / B2;
            part_c[pix] = 2.f * A1 * inv;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = local;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int k = 0; k < SS_THREADS / 32; k++) t += s_red[k];
        atomicAdd(sum, t);
    }
}

__global__ void __launch_bounds__(SS_THREADS)
ssim_bwd_kernel(int C, int H, int W, const float* __restrict__ img1, const float* __restrict__ img2,
                SsimWindow w, const float* __restrict__ part_a, const float* __restrict__ part_b,
                const float* __restrict__ part_c, const float* __restrict__ upstream, float inv_count,
                float* __restrict__ grad1) {
    __shared__ float s_in[3][HS][HS + 1];
    __shared__ float s_h[3][HS][TS];
    const int ch = blockIdx.z, x0 = blockIdx.x * TS, y0 = blockIdx.y * TS;
    const size_t plane = (size_t)ch * H * W;
    for (int o = threadIdx.x; o < HS * HS; o += SS_THREADS) {
        const int r = o / HS, c = o - r * HS;
        const int y = y0 + r - R, x = x0 + c - R;
        float a = 0.f, b = 0.f, cc = 0.f;
        if (y >= 0 && y < H && x >= 0 && x < W) {
            const size_t pix = plane + (size_t)y * W + x;
            a = part_a[pix]; b = part_b[pix]; cc = part_c[pix];
        }
        s_in[0][r][c] = a; s_in[1][r][c] = b; s_in[2][r][c] = cc;
    }
    __syncthreads();
    horizontal_pass<3>(s_in, s_h, w);
    __syncthreads();
    const float scale = __ldg(upstream) * inv_count;
    for (int o = threadIdx.x; o < TS * TS; o += SS_THREADS) {
        const int r = o / TS, c = o - r * TS;
        const int y = y0 + r, x = x0 + c;
        if (y >= H || x >= W) continue;
        float m[3] = {0.f, 0.f, 0.f};
#pragma unroll
        for (int k = 0; k <= 2 * R; k++) {
#pragma unroll
            for (int q = 0; q < 3; q++) m[q] = fmaf(w.g[k], s_h[q][r + k][c], m[q]);
        }
        const size_t pix = plane + (size_t)y * W + x;
        grad1[pix] = scale * (m[0] + 2.f * img1[pix] * m[1] + img2[pix] * m[2]);
    }
}

__global__ void ssim_finish_kernel(const double* __restrict__ sum, float inv_count, float* __restrict__ out) {
    out[0] = (float)(sum[0] * (double)inv_count);
}

SsimWindow make_window() {   // losses.py:28-41: float32 exp values divided by their float32 sum
    SsimWindow w;
    float s = 0.f;
    for (int k = 0; k <= 2 * R; k++) {
        w.g[k] = (float)exp(-(double)((k - R) * (k - R)) / (2.0 * 1.5 * 1.5));
        s += w.g[k];
    }
    for (int k = 0; k <= 2 * R; k++) w.g[k] /= s;
    return w;
}

}  // namespace

int launch_ssim_fwd(int C, int H, int W, const float* img1, const float* img2, float* part_abc, double* sum,
                    float* out, cudaStream_t s) {
    const size_t n = (size_t)C * H * W;
    if (cudaMemsetAsync(sum, 0, sizeof(double), s) != cudaSuccess) return -1;
    dim3 grid((W + TS - 1) / TS, (H + TS - 1) / TS, C);
    ssim_fwd_kernel<<<grid, SS_THREADS, 0, s>>>(C, H, W, img1, img2, make_window(), part_abc, part_abc + n,
                                                part_abc + 2 * n, sum);
    ssim_finish_kernel<<<1, 1, 0, s>>>(sum, 1.f / (float)n, out);
    return cudaPeekAtLastError() == cudaSuccess ? 2 : -1;
}

int launch_ssim_bwd(int C, int H, int W, const float* img1, const float* img2, const float* part_abc,
                    const float* upstream, float* grad1, cudaStream_t s) {
    const size_t n = (size_t)C * H * W;
    dim3 grid((W + TS - 1) / TS, (H + TS - 1) / TS, C);
    ssim_bwd_kernel<<<grid, SS_THREADS, 0, s>>>(C, H, W, img1, img2, make_window(), part_abc, part_abc + n,
                                                part_abc + 2 * n, upstream, 1.f / (float)n, grad1);
    return cudaPeekAtLastError() == cudaSuccess ? 1 : -1;
}
