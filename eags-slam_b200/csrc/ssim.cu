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
//
// FUSE (the whole per-view mapping loss in two kernels, mapper.py:161-165):
//   forward : the tile's centre pixels are already in shared memory, so the masked colour L1 and (channel-0
//             CTAs) the masked depth L1 are summed in the same pass; per-CTA partials, combined in fixed order
//             by a one-CTA finish kernel, which also forms  (1 - lambda) L1 + lambda (1 - SSIM)  and the depth loss.
//   backward: dL/dimage = -lambda up_c dSSIM/dimage + (1 - lambda) up_c sign(image - gt) m / n_c  and
//             dL/ddepth = up_d sign(depth - gt_depth) m / n_d  (the signs are recomputed from the pixels the
//             kernel reads anyway: no raw-gradient images, no scaling pass, no autograd adds).
#include "kernels.h"

namespace {

constexpr int TS = 32;            // output tile
constexpr int R = 5;              // window radius
constexpr int HS = TS + 2 * R;    // halo tile
constexpr int SS_THREADS = 256;

struct SsimWindow { float g[2 * R + 1]; };

// the L1 terms of the mapping loss riding on the SSIM kernels (FUSE = true)
struct MapFwd {
    const float* depth;        // (H,W) rendered
    const float* gt_depth;     // (H,W)
    const uint8_t* mask;       // (H,W) or NULL (all valid)
    int valid_depth;           // AND the mask with (gt_depth > 0) & !isnan(depth)  (mapper.py:159)
    double* partial;           // [n_cta][4] = {SSIM sum, colour |d| sum, depth |d| sum, masked pixels}
    float lambda;
    float* out;                // [8] = {colour sum, n_c, depth sum, n_d, colour L1, depth L1, SSIM, colour loss}
};
struct MapBwd {
    const float* depth;
    const float* gt_depth;
    const uint8_t* mask;
    int valid_depth;
    const float* out;          // the forward's out[8]
    const float* up_c;         // dL/d(colour loss), device scalar
    const float* up_d;         // dL/d(depth loss), device scalar
    float lambda;
    float* grad_depth;         // (H,W)
};

__device__ __forceinline__ float sgnf(float v) { return (v > 0.f) ? 1.f : ((v < 0.f) ? -1.f : 0.f); }

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

// The (32+10)^2 halo tile of NMAPS planes, global -> shared with cp.async (LDGSTS): no register staging, all of a
// thread's ~7 x NMAPS loads in flight at once (ncu: the staged loop left both kernels waiting on long_scoreboard),
// zero fill outside the image through the src-size operand (F.conv2d's zero padding).  (r, c) advance
// incrementally: 256 elements further = 6 rows and 4 columns.
template <int NMAPS>
__device__ __forceinline__ void halo_load_async(float (*s_in)[HS][HS + 1], const float* const (&plane)[NMAPS], int H, int W,
                                                int x0, int y0) {
    static_assert(SS_THREADS == 6 * HS + 4, "incremental (r, c) step");
    int r = threadIdx.x / HS, c = threadIdx.x - r * HS;
    for (int o = threadIdx.x; o < HS * HS; o += SS_THREADS) {
        const int y = y0 + r - R, x = x0 + c - R;
        const bool in = y >= 0 && y < H && x >= 0 && x < W;
        const size_t off = in ? (size_t)y * W + x : 0;
        const unsigned nbytes = in ? 4u : 0u;
#pragma unroll
        for (int m = 0; m < NMAPS; m++) {
            const unsigned dst = (unsigned)__cvta_generic_to_shared(&s_in[m][r][c]);
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(dst), "l"(plane[m] + off), "r"(nbytes) : "memory");
        }
        r += 6; c += 4;
        if (c >= HS) { c -= HS; r += 1; }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 0;" ::: "memory");
}

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

template <bool FUSE>
__global__ void __launch_bounds__(SS_THREADS)
ssim_fwd_kernel(int C, int H, int W, const float* __restrict__ img1, const float* __restrict__ img2,
                SsimWindow w, float* __restrict__ part_a, float* __restrict__ part_b,
                float* __restrict__ part_c, double* __restrict__ sum, MapFwd mf) {
    __shared__ float s_in[2][HS][HS + 1];     // the two images; the three products are formed on the fly
    __shared__ float s_h[5][HS][SH_STRIDE];
    __shared__ double s_red[FUSE ? 4 : 1][SS_THREADS / 32];
    const int ch = blockIdx.z, x0 = blockIdx.x * TS, y0 = blockIdx.y * TS;
    const float* p1 = img1 + (size_t)ch * H * W;
    const float* p2 = img2 + (size_t)ch * H * W;
    {
        const float* const planes[2] = {p1, p2};
        halo_load_async<2>(s_in, planes, H, W, x0, y0);
    }
    __syncthreads();
    horizontal_pass<2, 5>(s_in, s_h, w, [](const float in[2], float f[5]) {
        f[0] = in[0]; f[1] = in[1]; f[2] = in[0] * in[0]; f[3] = in[1] * in[1]; f[4] = in[0] * in[1];
    });
    __syncthreads();
    double local = 0.0;
    float l1_c = 0.f, l1_d = 0.f, cnt = 0.f;     // FUSE: at most VB pixels per thread, exact in fp32 order
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
                if constexpr (FUSE) {
                    const size_t p2d = (size_t)y * W + x;
                    bool on = mf.mask ? (mf.mask[p2d] != 0) : true;
                    float dep = 0.f, gtd = 0.f;
                    if (mf.valid_depth || ch == 0) { dep = __ldg(mf.depth + p2d); gtd = __ldg(mf.gt_depth + p2d); }
                    if (mf.valid_depth) on = on && gtd > 0.f && !isnan(dep);
                    // selects, not products: a masked-out pixel may hold nan (that is what the mask is for)
                    l1_c += on ? fabsf(s_in[0][r0 + jj + R][c + R] - s_in[1][r0 + jj + R][c + R]) : 0.f;
                    cnt += on ? 1.f : 0.f;
                    if (ch == 0) l1_d += on ? fabsf(dep - gtd) : 0.f;
                }
            }
        }
    }
    if constexpr (!FUSE) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
        if ((threadIdx.x & 31) == 0) s_red[0][threadIdx.x >> 5] = local;
        __syncthreads();
        if (threadIdx.x == 0) {
            double t = 0.0;
            for (int k = 0; k < SS_THREADS / 32; k++) t += s_red[0][k];
            atomicAdd(sum, t);
        }
    } else {
        // per-CTA partials in fixed order (the gradient only depends on the exact counts anyway)
        // the L1 sums of a warp (<= 128 terms of magnitude <= a few units) stay in fp32, the count is an integer;
        // fp64 from the warp totals on
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            local += __shfl_xor_sync(0xffffffffu, local, o);
            l1_c += __shfl_xor_sync(0xffffffffu, l1_c, o);
            if (ch == 0) l1_d += __shfl_xor_sync(0xffffffffu, l1_d, o);
        }
        const unsigned cnt_w = __reduce_add_sync(0xffffffffu, (unsigned)cnt);
        if ((threadIdx.x & 31) == 0) {
            const int wi = threadIdx.x >> 5;
            s_red[0][wi] = local; s_red[1][wi] = (double)l1_c; s_red[2][wi] = (double)l1_d; s_red[3][wi] = (double)cnt_w;
        }
        __syncthreads();
        const unsigned me = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
        if (threadIdx.x < 4) {
            double t = 0.0;
#pragma unroll
            for (int k = 0; k < SS_THREADS / 32; k++) t += s_red[threadIdx.x][k];
            mf.partial[(size_t)me * 4 + threadIdx.x] = t;     // combined by mapping_loss_finish_kernel
        }
    }
}

// One CTA sums the per-CTA partials in a fixed order (thread t takes CTAs t, t + 1024, ...; then a fixed tree):
// the loss values are reproducible run to run.  A separate launch instead of a last-CTA epilogue: the epilogue
// needed a __threadfence in every CTA (which exposed the latency of its 12 partial-map stores per thread) and
// left ~10 dependent rounds of L2 loads running alone at the end of the kernel.
constexpr int FIN_THREADS = 1024;
__global__ void __launch_bounds__(FIN_THREADS)
mapping_loss_finish_kernel(MapFwd mf, unsigned n_cta, unsigned per_ch, double inv_n) {
    __shared__ double s_fin[5][FIN_THREADS / 32];
    double t[5] = {0.0, 0.0, 0.0, 0.0, 0.0};   // SSIM, colour, depth, masked elements of all channels, of channel 0
    for (unsigned b = threadIdx.x; b < n_cta; b += FIN_THREADS) {
        const double2 lo = __ldcg(reinterpret_cast<const double2*>(mf.partial + (size_t)b * 4));
        const double2 hi = __ldcg(reinterpret_cast<const double2*>(mf.partial + (size_t)b * 4 + 2));
        t[0] += lo.x; t[1] += lo.y; t[2] += hi.x; t[3] += hi.y;
        if (b < per_ch) t[4] += hi.y;
    }
#pragma unroll
    for (int k = 0; k < 5; k++) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) t[k] += __shfl_xor_sync(0xffffffffu, t[k], o);
        if ((threadIdx.x & 31) == 0) s_fin[k][threadIdx.x >> 5] = t[k];
    }
    __syncthreads();
    if (threadIdx.x < 32) {
        double f[5];
#pragma unroll
        for (int k = 0; k < 5; k++) {
            f[k] = s_fin[k][threadIdx.x];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) f[k] += __shfl_xor_sync(0xffffffffu, f[k], o);
        }
        if (threadIdx.x == 0) {
            const float sc = (float)f[1], nc = (float)f[3], sd = (float)f[2], nd = (float)f[4];
            const float ssim = (float)(f[0] * inv_n);
            const float l1c = sc / nc, l1d = sd / nd;       // 0/0 -> nan, as the reference's mean of nothing
            mf.out[0] = sc; mf.out[1] = nc; mf.out[2] = sd; mf.out[3] = nd;
            mf.out[4] = l1c; mf.out[5] = l1d; mf.out[6] = ssim;
            mf.out[7] = (1.0f - mf.lambda) * l1c + mf.lambda * (1.0f - ssim);   // mapper.py:162-163
        }
    }
}

template <bool FUSE>
__global__ void __launch_bounds__(SS_THREADS)
ssim_bwd_kernel(int C, int H, int W, const float* __restrict__ img1, const float* __restrict__ img2,
                SsimWindow w, const float* __restrict__ part_a, const float* __restrict__ part_b,
                const float* __restrict__ part_c, const float* __restrict__ upstream, float inv_count,
                float* __restrict__ grad1, MapBwd mb) {
    __shared__ float s_in[3][HS][HS + 1];
    __shared__ float s_h[3][HS][SH_STRIDE];
    const int ch = blockIdx.z, x0 = blockIdx.x * TS, y0 = blockIdx.y * TS;
    const size_t plane = (size_t)ch * H * W;
    {
        const float* const planes[3] = {part_a + plane, part_b + plane, part_c + plane};
        halo_load_async<3>(s_in, planes, H, W, x0, y0);
    }
    __syncthreads();
    horizontal_pass<3, 3>(s_in, s_h, w, [](const float in[3], float f[3]) { f[0] = in[0]; f[1] = in[1]; f[2] = in[2]; });
    __syncthreads();
    float scale, k_c = 0.f, k_d = 0.f;
    if constexpr (FUSE) {
        const float up_c = __ldg(mb.up_c), n_c = __ldg(mb.out + 1), n_d = __ldg(mb.out + 3);
        scale = -mb.lambda * up_c * inv_count;                   // d(lambda (1 - ssim))
        k_c = n_c > 0.f ? (1.0f - mb.lambda) * up_c / n_c : 0.f;
        k_d = n_d > 0.f ? __ldg(mb.up_d) / n_d : 0.f;
    } else {
        scale = __ldg(upstream) * inv_count;
    }
    const int c = threadIdx.x % TS, r0 = (threadIdx.x / TS) * VB;
    float m[3][VB];
    vertical_window<3>(s_h, w, r0, c, m);
    const int x = x0 + c;
#pragma unroll
    for (int jj = 0; jj < VB; jj++) {
        const int y = y0 + r0 + jj;
        if (y >= H || x >= W) continue;
        const size_t pix = plane + (size_t)y * W + x;
        const float a = img1[pix], b = img2[pix];
        float g = scale * (m[0][jj] + 2.f * a * m[1][jj] + b * m[2][jj]);
        if constexpr (FUSE) {
            const size_t p2d = (size_t)y * W + x;
            bool on = mb.mask ? (mb.mask[p2d] != 0) : true;
            float dep = 0.f, gtd = 0.f;
            if (mb.valid_depth || ch == 0) { dep = __ldg(mb.depth + p2d); gtd = __ldg(mb.gt_depth + p2d); }
            if (mb.valid_depth) on = on && gtd > 0.f && !isnan(dep);
            g += on ? k_c * sgnf(a - b) : 0.f;
            if (ch == 0) mb.grad_depth[p2d] = on ? k_d * sgnf(dep - gtd) : 0.f;
        }
        grad1[pix] = g;
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
    ssim_fwd_kernel<false><<<grid, SS_THREADS, 0, s>>>(C, H, W, img1, img2, make_window(), part_abc, part_abc + n,
                                                       part_abc + 2 * n, sum, MapFwd{});
    ssim_finish_kernel<<<1, 1, 0, s>>>(sum, 1.f / (float)n, out);
    return cudaPeekAtLastError() == cudaSuccess ? 2 : -1;
}

int launch_ssim_bwd(int C, int H, int W, const float* img1, const float* img2, const float* part_abc,
                    const float* upstream, float* grad1, cudaStream_t s) {
    const size_t n = (size_t)C * H * W;
    dim3 grid((W + TS - 1) / TS, (H + TS - 1) / TS, C);
    ssim_bwd_kernel<false><<<grid, SS_THREADS, 0, s>>>(C, H, W, img1, img2, make_window(), part_abc, part_abc + n,
                                                       part_abc + 2 * n, upstream, 1.f / (float)n, grad1, MapBwd{});
    return cudaPeekAtLastError() == cudaSuccess ? 1 : -1;
}

// ---- the fused per-view mapping loss (three-channel image + depth) ------------------------------------
// scratch: [3 partial maps of 3 H W floats, 256-aligned][n_cta x 4 doubles]
static size_t map_parts_bytes(int H, int W) { return eags_align256((size_t)3 * 3 * H * W * sizeof(float)); }
static size_t map_n_cta(int H, int W) { return (size_t)((W + TS - 1) / TS) * ((H + TS - 1) / TS) * 3; }

size_t mapping_loss_scratch_bytes(int W, int H) {
    return map_parts_bytes(H, W) + map_n_cta(H, W) * 4 * sizeof(double);
}

int launch_mapping_loss_fwd(int W, int H, float lambda, const float* image, const float* depth, const float* gt_image,
                            const float* gt_depth, const uint8_t* mask, int valid_depth, void* scratch, float* out8,
                            cudaStream_t s) {
    const size_t n = (size_t)3 * H * W;
    float* part = (float*)scratch;
    MapFwd mf;
    mf.depth = depth; mf.gt_depth = gt_depth; mf.mask = mask; mf.valid_depth = valid_depth; mf.lambda = lambda; mf.out = out8;
    mf.partial = (double*)((char*)scratch + map_parts_bytes(H, W));
    dim3 grid((W + TS - 1) / TS, (H + TS - 1) / TS, 3);
    ssim_fwd_kernel<true><<<grid, SS_THREADS, 0, s>>>(3, H, W, image, gt_image, make_window(), part, part + n,
                                                      part + 2 * n, nullptr, mf);
    mapping_loss_finish_kernel<<<1, FIN_THREADS, 0, s>>>(mf, grid.x * grid.y * grid.z, grid.x * grid.y,
                                                         1.0 / (double)n);
    return cudaPeekAtLastError() == cudaSuccess ? 2 : -1;
}

int launch_mapping_loss_bwd(int W, int H, float lambda, const float* image, const float* depth, const float* gt_image,
                            const float* gt_depth, const uint8_t* mask, int valid_depth, const void* scratch,
                            const float* out8, const float* up_color, const float* up_depth, float* grad_image,
                            float* grad_depth, cudaStream_t s) {
    const size_t n = (size_t)3 * H * W;
    const float* part = (const float*)scratch;
    MapBwd mb;
    mb.depth = depth; mb.gt_depth = gt_depth; mb.mask = mask; mb.valid_depth = valid_depth; mb.out = out8; mb.up_c = up_color; mb.up_d = up_depth;
    mb.lambda = lambda; mb.grad_depth = grad_depth;
    dim3 grid((W + TS - 1) / TS, (H + TS - 1) / TS, 3);
    ssim_bwd_kernel<true><<<grid, SS_THREADS, 0, s>>>(3, H, W, image, gt_image, make_window(), part, part + n,
                                                      part + 2 * n, nullptr, 1.f / (float)n, grad_image, mb);
    return cudaPeekAtLastError() == cudaSuccess ? 1 : -1;
}
