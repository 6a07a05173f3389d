// loss.cu -- fused image-space losses that bracket the rasterizer in every tracking / mapping
// iteration (reference src/entities/tracker.py:99-128,
// src/entities/mapper.py:161-165, src/entities/losses.py:8-27).
//
// The reference builds these from ~25 elementwise / reduction torch kernels per iteration.  Here:
//   K9  image_l1_fwd   : ONE pass over the pixels -> the four reductions (colour / depth loss sums and
//                        their normalisers) and the raw (un-normalised) gradient images
//                        dL/dcolor, dL/ddepth, dL/dalpha.  Deterministic: per-CTA partials are
//                        combined in fixed order by the last CTA to finish.
//   K10 image_l1_scale : scales the raw gradients by upstream_grad / normaliser (device scalars).
// HBM-bound streaming: 9 floats + 1 mask byte read, 5 floats written per pixel.
#include <initializer_list>
#include <mutex>

#include "kernels.h"

namespace {

constexpr int LOSS_THREADS = 256;

struct LossArgs {
    int N;                 // pixels
    int soft_alpha;        // weight both losses by alpha^3 (tracker.py:118-121)
    int mask_color;        // apply the mask to the colour loss too (tracker.py:122-125)
    int norm_mode;         // 0: normalise by #(loss element > 0) (tracker.py:127-128)
                           // 1: normalise by #masked elements (mean over the mask; mapper.py:161-165, l1_loss)
    float eps;             // added to |x - y| before weighting (1e-8 in the tracker, 0 for a plain L1)
    const float* color;    // (3,N)
    const float* depth;    // (N)
    const float* alpha;    // (N) or NULL
    const float* gt_color; // (3,N)
    const float* gt_depth; // (N)
    const uint8_t* mask;   // (N) or NULL (all valid)
    const uint8_t* mask_depth;   // (N) or NULL (= mask)
    float alpha_min;       // soft_alpha == 2: opacity threshold of the depth term
    float* g_color;        // (3,N) raw gradient
    float* g_depth;        // (N)
    float* g_alpha;        // (N) or NULL
    double* partial;       // [grid][4]
    unsigned* ticket;      // zero-initialised by the caller's memset
    float* out;            // [6] = {colour sum, colour normaliser, depth sum, depth normaliser, colour loss, depth loss}
};

__device__ __forceinline__ float sgn(float v) { return (v > 0.f) ? 1.f : ((v < 0.f) ? -1.f : 0.f); }

// the loss terms and raw gradients of ONE pixel (the arithmetic of the header comment, in one place for the
// scalar and the 4-pixels-per-thread paths)
struct PixelOut { float g[3], gd, ga_c, ga_d; };
__device__ __forceinline__ PixelOut pixel_terms(const LossArgs& a, const float col[3], const float gt[3], float dep,
                                                float gtd, float al, float m, float md, float& s_c, float& n_c,
                                                float& s_d, float& n_d) {
    PixelOut o;
    const float w = a.soft_alpha == 1 ? al * al * al : 1.f;
    float wc = w * ((a.mask_color || !a.soft_alpha) ? m : 1.f);   // tracker.py:118-125
    float wd = w * md;
    if (a.soft_alpha == 2) {   // registration loss, src/gsr/loss.py:66-95
        wc = al * m;
        wd = md * (al > a.alpha_min ? 1.f : 0.f);
    }
    float lc_sum = 0.f;
#pragma unroll
    for (int c = 0; c < 3; c++) {
        const float d = col[c] - gt[c];
        const float l = (fabsf(d) + a.eps) * wc;
        lc_sum += fabsf(d) + a.eps;
        s_c += l;
        n_c += (a.norm_mode == 0) ? (l > 0.f ? 1.f : 0.f) : (a.norm_mode == 1 ? m : 1.f);
        o.g[c] = sgn(d) * wc;
    }
    const float dd = dep - gtd;
    const float ld0 = fabsf(dd) + a.eps;
    const float ld = ld0 * wd;
    s_d += ld;
    n_d += (a.norm_mode == 0) ? (ld > 0.f ? 1.f : 0.f) : (a.norm_mode == 1 ? md : 1.f);
    o.gd = sgn(dd) * wd;
    // d(alpha^3)/dalpha = 3 alpha^2, applied to the unweighted loss terms it multiplies; the colour and depth
    // parts are scaled by different factors later, so they are kept apart: colour part in slot p, depth
    // part in slot N + p of g_alpha
    const float d3 = a.soft_alpha == 1 ? 3.f * al * al : 0.f;
    o.ga_c = a.soft_alpha == 2 ? lc_sum * m : d3 * lc_sum * ((a.mask_color) ? m : 1.f);
    o.ga_d = a.soft_alpha == 2 ? 0.f : d3 * ld0 * md;   // (alpha > min) has no gradient
    return o;
}

// VEC: four pixels per thread and iteration through 16-byte loads / stores (N % 4 == 0, 16-byte aligned
// planes -- every torch image of the reference's sizes); all loads of an iteration are issued before its
// stores.  The scalar path handles any other size.
template <bool VEC>
__global__ void __launch_bounds__(LOSS_THREADS)
image_l1_fwd_kernel(LossArgs a) {
    float s_c = 0.f, n_c = 0.f, s_d = 0.f, n_d = 0.f;
    if constexpr (VEC) {
        const int N4 = a.N >> 2;
        const size_t N = (size_t)a.N;
        for (int q = blockIdx.x * LOSS_THREADS + threadIdx.x; q < N4; q += gridDim.x * LOSS_THREADS) {
            float4 c4[3], t4[3];
#pragma unroll
            for (int c = 0; c < 3; c++) {
                c4[c] = __ldg(reinterpret_cast<const float4*>(a.color + c * N) + q);
                t4[c] = __ldg(reinterpret_cast<const float4*>(a.gt_color + c * N) + q);
            }
            const float4 d4 = __ldg(reinterpret_cast<const float4*>(a.depth) + q);
            const float4 e4 = __ldg(reinterpret_cast<const float4*>(a.gt_depth) + q);
            const float4 al4 = a.alpha ? __ldg(reinterpret_cast<const float4*>(a.alpha) + q) : make_float4(1.f, 1.f, 1.f, 1.f);
            const uchar4 m4 = a.mask ? __ldg(reinterpret_cast<const uchar4*>(a.mask) + q) : make_uchar4(1, 1, 1, 1);
            const uchar4 k4 = a.mask_depth ? __ldg(reinterpret_cast<const uchar4*>(a.mask_depth) + q) : m4;
            const float cc[3][4] = {{c4[0].x, c4[0].y, c4[0].z, c4[0].w}, {c4[1].x, c4[1].y, c4[1].z, c4[1].w},
                                    {c4[2].x, c4[2].y, c4[2].z, c4[2].w}};
            const float tt[3][4] = {{t4[0].x, t4[0].y, t4[0].z, t4[0].w}, {t4[1].x, t4[1].y, t4[1].z, t4[1].w},
                                    {t4[2].x, t4[2].y, t4[2].z, t4[2].w}};
            const float dv[4] = {d4.x, d4.y, d4.z, d4.w}, ev[4] = {e4.x, e4.y, e4.z, e4.w};
            const float av[4] = {al4.x, al4.y, al4.z, al4.w};
            const unsigned char mv[4] = {m4.x, m4.y, m4.z, m4.w}, kv[4] = {k4.x, k4.y, k4.z, k4.w};
            float og[3][4], ogd[4], oac[4], oad[4];
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const float col[3] = {cc[0][u], cc[1][u], cc[2][u]}, gt[3] = {tt[0][u], tt[1][u], tt[2][u]};
                const PixelOut o = pixel_terms(a, col, gt, dv[u], ev[u], av[u], mv[u] ? 1.f : 0.f, kv[u] ? 1.f : 0.f,
                                               s_c, n_c, s_d, n_d);
                og[0][u] = o.g[0]; og[1][u] = o.g[1]; og[2][u] = o.g[2];
                ogd[u] = o.gd; oac[u] = o.ga_c; oad[u] = o.ga_d;
            }
#pragma unroll
            for (int c = 0; c < 3; c++)
                reinterpret_cast<float4*>(a.g_color + c * N)[q] = make_float4(og[c][0], og[c][1], og[c][2], og[c][3]);
            reinterpret_cast<float4*>(a.g_depth)[q] = make_float4(ogd[0], ogd[1], ogd[2], ogd[3]);
            if (a.g_alpha) {
                reinterpret_cast<float4*>(a.g_alpha)[q] = make_float4(oac[0], oac[1], oac[2], oac[3]);
                reinterpret_cast<float4*>(a.g_alpha + N)[q] = make_float4(oad[0], oad[1], oad[2], oad[3]);
            }
        }
    } else {
        for (int p = blockIdx.x * LOSS_THREADS + threadIdx.x; p < a.N; p += gridDim.x * LOSS_THREADS) {
            const float m = a.mask ? (a.mask[p] ? 1.f : 0.f) : 1.f;
            const float md = a.mask_depth ? (a.mask_depth[p] ? 1.f : 0.f) : m;
            const float al = a.alpha ? a.alpha[p] : 1.f;
            const float col[3] = {a.color[p], a.color[(size_t)a.N + p], a.color[(size_t)2 * a.N + p]};
            const float gt[3] = {a.gt_color[p], a.gt_color[(size_t)a.N + p], a.gt_color[(size_t)2 * a.N + p]};
            const PixelOut o = pixel_terms(a, col, gt, a.depth[p], a.gt_depth[p], al, m, md, s_c, n_c, s_d, n_d);
#pragma unroll
            for (int c = 0; c < 3; c++) a.g_color[(size_t)c * a.N + p] = o.g[c];
            a.g_depth[p] = o.gd;
            if (a.g_alpha) {
                a.g_alpha[p] = o.ga_c;
                a.g_alpha[(size_t)a.N + p] = o.ga_d;
            }
        }
    }
    // block reduction (fixed order inside the block), then a fixed-order pass over the partials
    __shared__ double sh[4][LOSS_THREADS / 32];
    double v[4] = {(double)s_c, (double)n_c, (double)s_d, (double)n_d};
#pragma unroll
    for (int k = 0; k < 4; k++) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v[k] += __shfl_xor_sync(0xffffffffu, v[k], o);
        if ((threadIdx.x & 31) == 0) sh[k][threadIdx.x >> 5] = v[k];
    }
    __syncthreads();
    __shared__ bool last;
    if (threadIdx.x < 4) {
        double t = 0.0;
#pragma unroll
        for (int w = 0; w < LOSS_THREADS / 32; w++) t += sh[threadIdx.x][w];
        a.partial[(size_t)blockIdx.x * 4 + threadIdx.x] = t;
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) last = (atomicAdd(a.ticket, 1u) == gridDim.x - 1);
    __syncthreads();
    if (last && threadIdx.x < 32) {
        __threadfence();
        double t[4] = {0.0, 0.0, 0.0, 0.0};
        for (unsigned b = threadIdx.x; b < gridDim.x; b += 32)
#pragma unroll
            for (int k = 0; k < 4; k++) t[k] += __ldcg(a.partial + (size_t)b * 4 + k);
#pragma unroll
        for (int k = 0; k < 4; k++)
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) t[k] += __shfl_xor_sync(0xffffffffu, t[k], o);
        if (threadIdx.x == 0) {
            const float sc = (float)t[0], nc = (float)t[1], sd = (float)t[2], nd = (float)t[3];
            a.out[0] = sc; a.out[1] = nc; a.out[2] = sd; a.out[3] = nd;
            a.out[4] = sc / nc;     // 0/0 -> nan, as the reference's division
            a.out[5] = sd / nd;
        }
    }
}

// g *= up / normaliser; `sums` = the forward's out[4]; `up` = {dL/dcolour_loss, dL/ddepth_loss} (device)
template <bool VEC>
__global__ void __launch_bounds__(LOSS_THREADS)
image_l1_scale_kernel(int N, const float* __restrict__ sums, const float* __restrict__ up_c,
                      const float* __restrict__ up_d, float* __restrict__ g_color,
                      float* __restrict__ g_depth, float* __restrict__ g_alpha) {
    const float kc = (sums[1] > 0.f) ? __ldg(up_c) / sums[1] : 0.f;
    const float kd = (sums[3] > 0.f) ? __ldg(up_d) / sums[3] : 0.f;
    if constexpr (VEC) {
        const int N4 = N >> 2;
        float4* gc = reinterpret_cast<float4*>(g_color);
        float4* gd = reinterpret_cast<float4*>(g_depth);
        float4* ga = reinterpret_cast<float4*>(g_alpha);
        for (int q = blockIdx.x * LOSS_THREADS + threadIdx.x; q < N4; q += gridDim.x * LOSS_THREADS) {
            float4 c0 = gc[q], c1 = gc[N4 + q], c2 = gc[2 * N4 + q], d = gd[q];
            float4 a0 = make_float4(0.f, 0.f, 0.f, 0.f), a1 = a0;
            if (g_alpha) { a0 = ga[q]; a1 = ga[N4 + q]; }
            c0.x *= kc; c0.y *= kc; c0.z *= kc; c0.w *= kc;
            c1.x *= kc; c1.y *= kc; c1.z *= kc; c1.w *= kc;
            c2.x *= kc; c2.y *= kc; c2.z *= kc; c2.w *= kc;
            d.x *= kd; d.y *= kd; d.z *= kd; d.w *= kd;
            gc[q] = c0; gc[N4 + q] = c1; gc[2 * N4 + q] = c2; gd[q] = d;
            if (g_alpha)
                ga[q] = make_float4(a0.x * kc + a1.x * kd, a0.y * kc + a1.y * kd, a0.z * kc + a1.z * kd, a0.w * kc + a1.w * kd);
        }
    } else {
        for (int p = blockIdx.x * LOSS_THREADS + threadIdx.x; p < N; p += gridDim.x * LOSS_THREADS) {
            g_color[p] *= kc;
            g_color[(size_t)N + p] *= kc;
            g_color[(size_t)2 * N + p] *= kc;
            g_depth[p] *= kd;
            if (g_alpha) g_alpha[p] = g_alpha[p] * kc + g_alpha[(size_t)N + p] * kd;
        }
    }
}

// 16-byte vector path: N divisible by 4 (so every plane offset c * N keeps the alignment) and aligned bases
bool vec_ok(int N, std::initializer_list<const void*> ptrs) {
    if (N & 3) return false;
    for (const void* p : ptrs)
        if (p && ((uintptr_t)p & 15)) return false;
    return true;
}

int loss_grid(int N) {
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int want = (N + LOSS_THREADS - 1) / LOSS_THREADS;
    const int cap = sms * 4;
    return want < cap ? (want < 1 ? 1 : want) : cap;
}

}  // namespace

size_t image_l1_scratch_bytes(int W, int H) {
    (void)W; (void)H;
    return 256 + (size_t)148 * 8 * 4 * sizeof(double);   // ticket line + partials for up to 1184 CTAs
}

int launch_image_l1_fwd(int W, int H, int soft_alpha, int mask_color, int norm_mode, float eps,
                        const float* color, const float* depth, const float* alpha, const float* gt_color,
                        const float* gt_depth, const uint8_t* mask, const uint8_t* mask_depth, float alpha_min,
                        float* g_color, float* g_depth,
                        float* g_alpha, void* scratch, float* out4, cudaStream_t s) {
    LossArgs a;
    a.N = W * H; a.soft_alpha = soft_alpha; a.mask_color = mask_color; a.norm_mode = norm_mode; a.eps = eps;
    a.color = color; a.depth = depth; a.alpha = alpha; a.gt_color = gt_color; a.gt_depth = gt_depth;
    a.mask = mask; a.mask_depth = mask_depth; a.alpha_min = alpha_min; a.g_color = g_color; a.g_depth = g_depth; a.g_alpha = g_alpha;
    a.ticket = (unsigned*)scratch;
    a.partial = (double*)((char*)scratch + 256);
    a.out = out4;
    if (cudaMemsetAsync(scratch, 0, 256, s) != cudaSuccess) return -1;
    const bool vec = vec_ok(a.N, {color, depth, alpha, gt_color, gt_depth, g_color, g_depth, g_alpha}) &&
                     (!mask || ((uintptr_t)mask & 3) == 0) && (!mask_depth || ((uintptr_t)mask_depth & 3) == 0);
    int grid = loss_grid(vec ? a.N / 4 : a.N);
    if (grid > 148 * 8) grid = 148 * 8;
    if (vec) image_l1_fwd_kernel<true><<<grid, LOSS_THREADS, 0, s>>>(a);
    else image_l1_fwd_kernel<false><<<grid, LOSS_THREADS, 0, s>>>(a);
    return cudaPeekAtLastError() == cudaSuccess ? 1 : -1;
}

// ---- image_l1_fwd with library-owned scratch (eags_image_l1_forward_static) ------------------------------
// For callers that may not allocate (CUDA-graph capture, serving loops, harnesses that forbid cudaMalloc): the
// kernel's scratch -- the 256-byte ticket line and the per-CTA partial sums -- lives in this file-scope device
// array instead of a caller buffer.  38 144 bytes, held by every GPU that loads the library until the process
// exits.  ALL calls share it: two calls that overlap on the GPU (different streams) race on the ticket and the
// partials; calls on one stream are ordered and each one clears the ticket first (the partials are overwritten
// before they are read).
constexpr size_t L1_STATIC_BYTES = 256 + (size_t)148 * 8 * 4 * sizeof(double);   // == image_l1_scratch_bytes()
static __device__ __align__(256) unsigned char g_l1_static_scratch[L1_STATIC_BYTES];

namespace {
constexpr int MAX_DEVICES = 64;
std::mutex g_l1_static_mutex;
void* g_l1_static_ptr[MAX_DEVICES];      // device address of the array on each device, resolved once per device
}  // namespace

// In host code `g_l1_static_scratch` is only a handle: the device address comes from cudaGetSymbolAddress, asked
// once per device with that device current.  Returns nullptr (and leaves the CUDA error set) on failure.
static void* l1_static_scratch_for_current_device() {
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= MAX_DEVICES) return nullptr;
    std::lock_guard<std::mutex> lock(g_l1_static_mutex);
    if (!g_l1_static_ptr[dev]) {
        void* p = nullptr;
        if (cudaGetSymbolAddress(&p, g_l1_static_scratch) != cudaSuccess || !p) return nullptr;
        g_l1_static_ptr[dev] = p;
    }
    return g_l1_static_ptr[dev];
}

// -2: the scratch address could not be resolved; otherwise as launch_image_l1_fwd.  No allocation of any kind.
int launch_image_l1_fwd_static(int W, int H, int soft_alpha, int mask_color, int norm_mode, float eps,
                               const float* color, const float* depth, const float* alpha, const float* gt_color,
                               const float* gt_depth, const uint8_t* mask, const uint8_t* mask_depth, float alpha_min,
                               float* g_color, float* g_depth, float* g_alpha, float* out4, cudaStream_t s) {
    static_assert(L1_STATIC_BYTES % 256 == 0, "whole 256-byte lines");
    void* scratch = l1_static_scratch_for_current_device();
    if (!scratch || image_l1_scratch_bytes(W, H) > L1_STATIC_BYTES) return -2;
    return launch_image_l1_fwd(W, H, soft_alpha, mask_color, norm_mode, eps, color, depth, alpha, gt_color, gt_depth,
                               mask, mask_depth, alpha_min, g_color, g_depth, g_alpha, scratch, out4, s);
}

int launch_image_l1_scale(int W, int H, const float* sums, const float* up_c, const float* up_d,
                          float* g_color, float* g_depth, float* g_alpha, cudaStream_t s) {
    const int N = W * H;
    if (vec_ok(N, {g_color, g_depth, g_alpha}))
        image_l1_scale_kernel<true><<<loss_grid(N / 4), LOSS_THREADS, 0, s>>>(N, sums, up_c, up_d, g_color, g_depth, g_alpha);
    else
        image_l1_scale_kernel<false><<<loss_grid(N), LOSS_THREADS, 0, s>>>(N, sums, up_c, up_d, g_color, g_depth, g_alpha);
    return cudaPeekAtLastError() == cudaSuccess ? 1 : -1;
}

// ---- isotropic regulariser of the mapping loss (reference src/entities/losses.py:115-126) -------------
//   loss = mean_{i,j} | s_ij - mean_j s_ij |   over scaling (P,3);  one pass: value (fp64 accumulator)
//   and the raw gradient  d(sum)/ds_ij = sign(d_ij) - (1/3) sum_k sign(d_ik)   (sign(0) = 0 as torch.abs)
namespace {
__global__ void __launch_bounds__(256)
isotropic_kernel(int P, const float* __restrict__ s, float* __restrict__ grad, double* __restrict__ sum) {
    __shared__ double s_red[8];
    const int i = blockIdx.x * 256 + threadIdx.x;
    double local = 0.0;
    if (i < P) {
        const float a = s[(size_t)3 * i], b = s[(size_t)3 * i + 1], c = s[(size_t)3 * i + 2];
        const float m = (a + b + c) / 3.0f;
        const float d0 = a - m * 1.0f, d1 = b - m * 1.0f, d2 = c - m * 1.0f;
        local = (double)(fabsf(d0) + fabsf(d1) + fabsf(d2));
        const float g0 = (d0 > 0.f) - (d0 < 0.f), g1 = (d1 > 0.f) - (d1 < 0.f), g2 = (d2 > 0.f) - (d2 < 0.f);
        const float gm = (g0 + g1 + g2) / 3.0f;
        grad[(size_t)3 * i] = g0 - gm; grad[(size_t)3 * i + 1] = g1 - gm; grad[(size_t)3 * i + 2] = g2 - gm;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = local;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int k = 0; k < 8; k++) t += s_red[k];
        atomicAdd(sum, t);
    }
}
__global__ void isotropic_finish_kernel(const double* __restrict__ sum, float inv_count, float* __restrict__ out) {
    out[0] = (float)(sum[0] * (double)inv_count);
}
}  // namespace

int launch_isotropic(int P, const float* scaling, float* grad, double* sum, float* out, cudaStream_t s) {
    if (cudaMemsetAsync(sum, 0, sizeof(double), s) != cudaSuccess) return -1;
    isotropic_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, scaling, grad, sum);
    isotropic_finish_kernel<<<1, 1, 0, s>>>(sum, 1.f / (3.f * (float)P), out);
    return cudaPeekAtLastError() == cudaSuccess ? 2 : -1;
}
