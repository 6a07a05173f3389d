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

// Register-blocked separable window (round 2).  Round 1 computed every output with its own 11 shared-memory
// loads per quantity: 84 LDS per pixel, LDS-bound at 109 us for 1752x1168x3.  Now a thread slides the window
// along the filter direction in registers:
//   horizontal pass: one thread = one halo row x 8 consecutive output columns: 18 loads per input feed
//                    8 outputs (lanes run over rows: row stride 43 floats -> conflict-free; the results go to
//                    s_h with row stride 33 -> conflict-free stores);
//   vertical pass:   one thread = one column x 4 consecutive output rows: 14 loads per quantity feed 4 outputs
//                    (lanes run over columns).
// ~5x fewer shared-memory instructions; the kernels end up FMA-bound.
constexpr int HB = 8;                 // outputs per thread, horizontal pass
constexpr int VB = 4;                 // outputs per thread, vertical pass
constexpr int SH_STRIDE = TS + 1;     // row stride of the horizontally filtered tile

// s_h[q][r][c0 + j] = sum_k g[k] * f_q(in[r][c0 + j + k]) for the NQ quantities produced by `emit`
template <int NIN, int NQ, typename Emit>
__device__ __forceinline__ void horizontal_pass(const float (*s_in)[HS][HS + 1], float (*s_h)[HS][SH_STRIDE],
                                                const SsimWindow& w, Emit emit) {
    for (int o = threadIdx.x; o < HS * (TS / HB); o += SS_THREADS) {
        const int seg = o / HS, r = o - seg * HS, c0 = seg * HB;
        float v[NIN][HB + 2 * R];
#pragma unroll
        for (int i = 0; i < NIN; i++)
#pragma unroll
            for (int k = 0; k < HB + 2 * R; k++) v[i][k] = s_in[i][r][c0 + k];
#pragma unroll
        for (int jj = 0; jj < HB; jj++) {
            float acc[NQ];
#pragma unroll
            for (int q = 0; q < NQ; q++) acc[q] = 0.f;
#pragma unroll
            for (int k = 0; k <= 2 * R; k++) {
                float in[NIN], f[NQ];
#pragma unroll
                for (int i = 0; i < NIN; i++) in[i] = v[i][jj + k];
                emit(in, f);
#pragma unroll
                for (int q = 0; q < NQ; q++) acc[q] = fmaf(w.g[k], f[q], acc[q]);
            }
#pragma unroll
            for (int q = 0; q < NQ; q++) s_h[q][r][c0 + jj] = acc[q];
        }
    }
}

// m[q][j] = sum_k g[k] * s_h[q][r0 + j + k][c] for the thread's VB rows
template <int NQ>
__device__ __forceinline__ void vertical_window(const float (*s_h)[HS][SH_STRIDE], const SsimWindow& w, int r0, int c,
                                                float m[NQ][VB]) {
#pragma unroll
    for (int q = 0; q < NQ; q++) {
        float v[VB + 2 * R];
#pragma unroll
        for (int k = 0; k < VB + 2 * R; k++) v[k] = s_h[q][r0 + k][c];
#pragma unroll
        for (int jj = 0; jj < VB; jj++) {
            float acc = 0.f;
#pragma unroll
            for (int k = 0; k <= 2 * R; k++) acc = fmaf(w.g[k], v[jj + k], acc);
            m[q][jj] = acc;
        }
    }
}

__global__ void __launch_bounds__(SS_THREADS)
ssim_fwd_kernel(int C, int H, int W, const float* __restrict__ img1, const float* __restrict__ img2,
                SsimWindow w, float* __restrict__ part_a, float* __restrict__ part_b,
                float* __restrict__ part_c, double* __restrict__ sum) {
    __shared__ float s_in[2][HS][HS + 1];     // the two images; the three products are formed on the fly
    __shared__ float s_h[5][HS][SH_STRIDE];
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
    horizontal_pass<2, 5>(s_in, s_h, w, [](const float in[2], float f[5]) {
        f[0] = in[0]; f[1] = in[1]; f[2] = in[0] * in[0]; f[3] = in[1] * in[1]; f[4] = in[0] * in[1];
    });
    __syncthreads();
    double local = 0.0;
    const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;
    {
        static_assert(TS * (TS / VB) == SS_THREADS, "one vertical item per thread");
        const int c = threadIdx.x % TS, r0 = (threadIdx.x / TS) * VB;
        float m[5][VB];
        vertical_window<5>(s_h, w, r0, c, m);
        const int x = x0 + c;
#pragma unroll
        for (int jj = 0; jj < VB; jj++) {
            const int y = y0 + r0 + jj;
            if (y < H && x < W) {
                const float mu1 = m[0][jj], mu2 = m[1][jj];
                const float mu1mu2 = mu1 * mu2, mu1sq = mu1 * mu1, mu2sq = mu2 * mu2;
                const float sig1 = m[2][jj] - mu1sq, sig2 = m[3][jj] - mu2sq, sig12 = m[4][jj] - mu1mu2;
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
    __shared__ float s_h[3][HS][SH_STRIDE];
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
    horizontal_pass<3, 3>(s_in, s_h, w, [](const float in[3], float f[3]) { f[0] = in[0]; f[1] = in[1]; f[2] = in[2]; });
    __syncthreads();
    const float scale = __ldg(upstream) * inv_count;
    const int c = threadIdx.x % TS, r0 = (threadIdx.x / TS) * VB;
    float m[3][VB];
    vertical_window<3>(s_h, w, r0, c, m);
    const int x = x0 + c;
#pragma unroll
    for (int jj = 0; jj < VB; jj++) {
        const int y = y0 + r0 + jj;
        if (y >= H || x >= W) continue;
        const size_t pix = plane + (size_t)y * W + x;
        grad1[pix] = scale * (m[0][jj] + 2.f * img1[pix] * m[1][jj] + img2[pix] * m[2][jj]);
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
