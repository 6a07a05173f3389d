// preprocess.cu -- per-Gaussian kernels of the rasterizer hot path (sm_100a).
//
//   K1 preprocess_fwd : cull, Sigma3D, EWA projection (+0.3 blur), conic, radius, pixel mean, tile
//                       rectangle, SH->RGB, depth; per-tile key counts (RED atomics) and K.
//                       Replaces upstream forward.cu::preprocessCUDA + the CUB InclusiveSum
//                       (SURVEY.md 2.3 / A.1); the per-Gaussian prefix sum disappears because keys
//                       are bucketed per tile (binning.cu).
//   K3 duplicate      : one (depth|index) key per covered tile, scattered into the tile's bucket
//                       (upstream rasterizer_impl.cu::duplicateWithKeys).
//   (K8 preprocess_bwd lives in preprocess_bwd.cu, a separate translation unit built WITH FMA.)
//
// THIS FILE IS COMPILED WITH -fmad=false: every expression is evaluated in IEEE binary32 in the
// order written, so depth keys, radii and tile rectangles are bit-identical to the CPU oracle
// (oracle/eags_oracle.c, built with -ffp-contract=off).  The kernels are HBM/latency bound, the
// missing FMA contraction costs nothing measurable.
#include "preprocess_common.cuh"

namespace {

// ------------------------------------------------------------------------------------------------
// K1
// ------------------------------------------------------------------------------------------------
// Geometry of one Gaussian up to its tile rectangle; false = culled.  CamPtr is
// `const float*` (registers) or `const volatile float*` (shared memory, re-read at every use): the
// arithmetic and its order are the same, so the results are bit-identical either way.
struct Splat {
    float p[3];
    float px, py, depth;
    float4 con;
    int r, x0, y0, x1, y1;
};

// Inputs of one Gaussian requested up front (the aggregating kernel): the scale / rotation / opacity / SH-DC
// loads do not depend on the culls, so issuing them together with the mean costs one DRAM latency per
// Gaussian instead of three.
struct RawIn {
    float sc[3];
    float4 q;
    float opacity;
    float sh[3];
};

template <typename CamPtr>
__device__ __forceinline__ bool project_one(const PreArgs& a, CamPtr V, CamPtr F, int i, Splat& s,
                                            bool pre = false, const RawIn& in = RawIn()) {
    float pv[3];
    pv[0] = V[0] * s.p[0] + V[4] * s.p[1] + V[8] * s.p[2] + V[12];
    pv[1] = V[1] * s.p[0] + V[5] * s.p[1] + V[9] * s.p[2] + V[13];
    pv[2] = V[2] * s.p[0] + V[6] * s.p[1] + V[10] * s.p[2] + V[14];
    if (pv[2] <= 0.2f) return false;
    float hx = F[0] * s.p[0] + F[4] * s.p[1] + F[8] * s.p[2] + F[12];
    float hy = F[1] * s.p[0] + F[5] * s.p[1] + F[9] * s.p[2] + F[13];
    float hw = F[3] * s.p[0] + F[7] * s.p[1] + F[11] * s.p[2] + F[15];
    float pw = 1.0f / (hw + 0.0000001f);
    float ppx = hx * pw, ppy = hy * pw;
    float c6[6];
    if (pre) cov3d_from_scale_rot(in.sc, a.scale_modifier, in.q, c6);
    else load_cov6(a, i, c6);
    float fx = (float)a.W / (2.f * a.tanfovx);
    float fy = (float)a.H / (2.f * a.tanfovy);
    Ewa e;
    ewa_project_lean(V, fx, fy, a.tanfovx, a.tanfovy, pv, c6, e);
    float ca = e.cov[0], cb = e.cov[1], cc = e.cov[2];
    float det = ca * cc - cb * cb;
    if (det == 0.0f) return false;
    float det_inv = 1.f / det;
    float mid = 0.5f * (ca + cc);
    float sq = sqrtf(fmaxf(0.1f, mid * mid - det));
    float lambda1 = mid + sq, lambda2 = mid - sq;
    float my_radius = ceilf(3.f * sqrtf(fmaxf(lambda1, lambda2)));
    s.px = ((ppx + 1.0f) * (float)a.W - 1.0f) * 0.5f;
    s.py = ((ppy + 1.0f) * (float)a.H - 1.0f) * 0.5f;
    s.r = (int)my_radius;
    tile_rect(s.px, s.py, s.r, a.gx, a.gy, s.x0, s.y0, s.x1, s.y1);
    if ((s.x1 - s.x0) * (s.y1 - s.y0) == 0) return false;
    s.con = make_float4(cc * det_inv, -cb * det_inv, ca * det_inv, 0.f);
    s.depth = pv[2];
    return true;
}

// colour + the per-Gaussian geom records of a surviving Gaussian
__device__ __forceinline__ void emit_one(const PreArgs& a, const GeomPtrs& geom, int i, Splat& s,
                                         const float* sh_row, bool pre = false, const RawIn& in = RawIn()) {
    float rgb[3];
    unsigned clamped = 0;
    if (pre) {
        float cp[3] = {0.f, 0.f, 0.f};                 // degree 0: the view direction is not used
        sh_to_rgb(0, s.p, cp, in.sh, rgb, clamped);
    } else if (a.colors_precomp) {
        rgb[0] = a.colors_precomp[(size_t)3 * i];
        rgb[1] = a.colors_precomp[(size_t)3 * i + 1];
        rgb[2] = a.colors_precomp[(size_t)3 * i + 2];
    } else {
        float cp[3] = {__ldg(a.campos), __ldg(a.campos + 1), __ldg(a.campos + 2)};
        sh_to_rgb(a.degree, s.p, cp, sh_row, rgb, clamped);
    }
    // the record in the form the blend kernels read (blend.cu): log2-scaled conic, alpha >= 1/255 reach
    const float opacity = pre ? in.opacity : a.opacities[i];
    const float L = 1.4426950408889634f;
    float4* rec = geom.splat_rec + (size_t)3 * i;
    rec[0] = make_float4(s.px, s.py, __uint_as_float((unsigned)i), s.depth);
    rec[1] = make_float4(-0.5f * L * s.con.x, -L * s.con.y, -0.5f * L * s.con.z, opacity);
    rec[2] = make_float4(rgb[0], rgb[1], rgb[2], __log2f(255.f * opacity));
    geom.clamped[i] = (uint8_t)clamped;
}

// One thread per Gaussian, one global atomic per (Gaussian, tile): the general kernel (wide SH rows,
// images with more tiles than the aggregating kernel's shared-memory histogram holds).
template <bool WIDE>
__global__ void __launch_bounds__(PRE_THREADS)
preprocess_fwd_kernel(PreArgs a, GeomPtrs geom, int32_t* __restrict__ radii,
                      uint32_t* __restrict__ tile_count, EagsCtrl* __restrict__ ctrl,
                      unsigned long long* __restrict__ bucket, unsigned tile_cap) {
    extern __shared__ float s_sh[];   // WIDE only: SH rows of the CTA, PRE_THREADS x S floats
    const int i = blockIdx.x * PRE_THREADS + threadIdx.x;
    unsigned my_tiles = 0;
    // phase A: geometry up to the tile rectangle; `alive` = survives every cull
    bool alive = false;
    Splat sp;
    sp.r = 0; sp.x0 = sp.y0 = sp.x1 = sp.y1 = 0; sp.depth = 0.f;
    if (i < a.P) {
        float V[16], F[16];
        load_cam(a, V, F);
        sp.p[0] = a.means3D[(size_t)3 * i]; sp.p[1] = a.means3D[(size_t)3 * i + 1]; sp.p[2] = a.means3D[(size_t)3 * i + 2];
        alive = project_one(a, (const float*)V, (const float*)F, i, sp);
    }
    // phase B (wide SH rows): the warp fetches the rows of its surviving Gaussians, coalesced
    const float* sh_row = nullptr;
    if constexpr (WIDE) {
        const int L = 3 * a.M, S = L | 1;
        float* s_w = s_sh + (size_t)(threadIdx.x >> 5) * 32 * S;
        const long long i0w = (long long)i - (threadIdx.x & 31);
        warp_rows_in(a.shs, i0w, L, S, (int)max(0ll, min(32ll, (long long)a.P - i0w)),
                     __ballot_sync(0xffffffffu, alive), s_w);
        sh_row = s_w + (threadIdx.x & 31) * S;
    } else if (alive && !a.colors_precomp) {
        sh_row = a.shs + (size_t)i * a.M * 3;
    }
    // phase C: colour, per-Gaussian records, tile counters / buckets
    if (i < a.P) {
        int r_out = 0;
        if (alive) {
            emit_one(a, geom, i, sp, sh_row);
            r_out = sp.r;
            const int x0 = sp.x0, y0 = sp.y0, x1 = sp.x1, y1 = sp.y1;
            my_tiles = (unsigned)((x1 - x0) * (y1 - y0));
            if (bucket == nullptr) {
                // exact path: per-tile key counts only (no-return RED atomics spread over G addresses);
                // duplicate_kernel scatters the keys after the tile scan
                for (int y = y0; y < y1; y++)
                    for (int x = x0; x < x1; x++) atomicAdd(&tile_count[(size_t)(y * a.gx + x) * EAGS_CTR_STRIDE], 1u);
            } else {
                // bucketed path: the returning atomic IS the slot inside the tile's fixed-stride bucket,
                // so the key is scattered here and there is no second pass over the Gaussians.  Four
                // atomics are kept in flight before the dependent stores.
                const unsigned long long key = ((unsigned long long)__float_as_uint(sp.depth) << 32) | (unsigned)i;
                const int n = (int)my_tiles;
                int x = x0, y = y0;
                bool over = false;
                for (int b = 0; b < n; b += 4) {
                    unsigned slot[4], tile[4];
#pragma unroll
                    for (int u = 0; u < 4; u++) {
                        if (b + u < n) {
                            tile[u] = (unsigned)(y * a.gx + x);
                            slot[u] = atomicAdd(&tile_count[(size_t)tile[u] * EAGS_CTR_STRIDE], 1u);
                            if (++x == x1) { x = x0; y++; }
                        }
                    }
#pragma unroll
                    for (int u = 0; u < 4; u++) {
                        if (b + u < n) {
                            if (slot[u] < tile_cap) bucket[(size_t)tile[u] * tile_cap + slot[u]] = key;
                            else over = true;
                        }
                    }
                }
                if (over) ctrl->bucket_overflow = 1u;   // the caller re-issues on the exact path
            }
        }
        radii[i] = r_out;
        geom.tiles_touched[i] = my_tiles;
    }
    // K = sum tiles_touched: warp shuffle reduce, one atomic per warp
    unsigned s = my_tiles;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0 && s) atomicAdd(&ctrl->num_rendered, (unsigned long long)s);
}

// CTA-aggregated K1 (the default for narrow SH rows).  The one-atomic-per-key kernel above is bound by
// the L2 atomic transaction rate (~47 G atomics/s: 886 K keys = 19 us at c2, whatever else the kernel
// does).  Here one CTA per SM owns a contiguous span of P / #SM Gaussians and
//   phase 1  runs the per-Gaussian preprocess and counts the span's keys per tile in a SHARED-memory
//            histogram (shared atomics),
//   phase 2  reserves, with ONE global atomic per (CTA, touched tile), a run of that many slots in the
//            tile's bucket (or, on the exact path, adds the count with a RED) and turns the histogram
//            entry into the run's cursor,
//   phase 3  re-derives each Gaussian's tile rectangle from the records it wrote in phase 1 (L2-resident,
//            same thread) and scatters its keys at cursor++ (shared atomics again).
// Global atomics per view drop from K to (#CTAs x tiles a span touches): 5x fewer at c2 for Gaussians in
// random order, ~#tiles-per-span fewer when the submap is stored in image order (SLAM seeds Gaussians
// from keyframe pixels, so neighbours in memory share tiles).  The order inside a bucket is arbitrary
// either way: sort_pack orders by (depth, index).
#ifndef EAGS_AGG_THREADS
#define EAGS_AGG_THREADS 1024
#endif
constexpr int AGG_THREADS = EAGS_AGG_THREADS;
constexpr int AGG_CTAS_PER_SM = 1024 / AGG_THREADS;
constexpr int AGG_MAX_TILES = 24 * 1024 / AGG_CTAS_PER_SM;   // <= 96 KB of histogram per SM

__global__ void __launch_bounds__(AGG_THREADS, AGG_CTAS_PER_SM)
preprocess_fwd_agg_kernel(PreArgs a, GeomPtrs geom, int32_t* __restrict__ radii,
                          uint32_t* __restrict__ tile_count, EagsCtrl* __restrict__ ctrl,
                          unsigned long long* __restrict__ bucket, unsigned tile_cap, int span) {
    extern __shared__ unsigned s_hist[];   // G counters, then the two camera matrices (32 floats)
    __shared__ unsigned s_sum[AGG_THREADS / 32];
    const int G = a.gx * a.gy, tid = threadIdx.x;
    float* s_cam = reinterpret_cast<float*>(s_hist + G);
    for (int t = tid; t < G; t += AGG_THREADS) s_hist[t] = 0u;
    if (tid < 32) s_cam[tid] = tid < 16 ? __ldg(a.view + tid) : __ldg(a.proj + tid - 16);
    __syncthreads();
    const volatile float* V = s_cam;
    const volatile float* F = s_cam + 16;
    const int begin = blockIdx.x * span, end = min(a.P, begin + span);
    unsigned my_total = 0;
    // ---- phase 1
    // the call shape of every reference call site (scale + rotation, SH degree 0): all inputs up front
    const bool fast = !a.cov3D_precomp && !a.colors_precomp && a.degree == 0;   // kernel-uniform
    for (int i = begin + tid; i < end; i += AGG_THREADS) {
        Splat sp;
        RawIn in;
        unsigned my_tiles = 0;
        int r_out = 0;
        sp.p[0] = a.means3D[(size_t)3 * i]; sp.p[1] = a.means3D[(size_t)3 * i + 1]; sp.p[2] = a.means3D[(size_t)3 * i + 2];
        if (fast) {
            in.sc[0] = a.scales[(size_t)3 * i]; in.sc[1] = a.scales[(size_t)3 * i + 1]; in.sc[2] = a.scales[(size_t)3 * i + 2];
            in.q = *reinterpret_cast<const float4*>(a.rotations + (size_t)4 * i);
            in.opacity = a.opacities[i];
            const float* shr = a.shs + (size_t)i * a.M * 3;
            in.sh[0] = shr[0]; in.sh[1] = shr[1]; in.sh[2] = shr[2];
        }
        if (project_one(a, V, F, i, sp, fast, in)) {
            emit_one(a, geom, i, sp, a.colors_precomp ? nullptr : a.shs + (size_t)i * a.M * 3, fast, in);
            r_out = sp.r;
            my_tiles = (unsigned)((sp.x1 - sp.x0) * (sp.y1 - sp.y0));
            for (int y = sp.y0; y < sp.y1; y++)
                for (int x = sp.x0; x < sp.x1; x++) {
                    EAGS_CHECK(y >= 0 && x >= 0 && y * a.gx + x < G);
                    atomicAdd(&s_hist[y * a.gx + x], 1u);
                }
        }
        radii[i] = r_out;
        geom.tiles_touched[i] = my_tiles;
        my_total += my_tiles;
    }
    __syncthreads();
    // ---- phase 2
    for (int t = tid; t < G; t += AGG_THREADS) {
        const unsigned c = s_hist[t];
        if (c) {
            if (bucket) s_hist[t] = atomicAdd(&tile_count[(size_t)t * EAGS_CTR_STRIDE], c);
            else atomicAdd(&tile_count[(size_t)t * EAGS_CTR_STRIDE], c);
        }
    }
    // ---- phase 3
    if (bucket) {
        __syncthreads();
        bool over = false;
        for (int i = begin + tid; i < end; i += AGG_THREADS) {
            const int r = radii[i];
            if (r <= 0) continue;
            const float4 r0 = geom.splat_rec[(size_t)3 * i];
            const float depth = r0.w;
            int x0, y0, x1, y1;
            tile_rect(r0.x, r0.y, r, a.gx, a.gy, x0, y0, x1, y1);
            const unsigned long long key = ((unsigned long long)__float_as_uint(depth) << 32) | (unsigned)i;
            for (int y = y0; y < y1; y++)
                for (int x = x0; x < x1; x++) {
                    const unsigned t = (unsigned)(y * a.gx + x);
                    EAGS_CHECK(t < (unsigned)G);
                    const unsigned slot = atomicAdd(&s_hist[t], 1u);
                    if (slot < tile_cap) bucket[(size_t)t * tile_cap + slot] = key;
                    else over = true;
                }
        }
        if (over) ctrl->bucket_overflow = 1u;   // the caller re-issues on the exact path
    }
    // K = sum tiles_touched: one atomic per CTA
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) my_total += __shfl_xor_sync(0xffffffffu, my_total, o);
    if ((tid & 31) == 0) s_sum[tid >> 5] = my_total;
    __syncthreads();
    if (tid < 32) {
        unsigned v = s_sum[tid];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (tid == 0 && v) atomicAdd(&ctrl->num_rendered, (unsigned long long)v);
    }
}

// ------------------------------------------------------------------------------------------------
// K3
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(PRE_THREADS)
duplicate_kernel(int P, int gx, int gy, const int32_t* __restrict__ radii, GeomPtrs geom,
                 uint32_t* __restrict__ tile_cursor, const EagsCtrl* __restrict__ ctrl,
                 unsigned long long* __restrict__ unsorted, long long capacity) {
    if (ctrl->num_rendered > (unsigned long long)capacity) return;   // caller re-issues with more room
    const int i = blockIdx.x * PRE_THREADS + threadIdx.x;
    if (i >= P) return;
    const int r = radii[i];
    if (r <= 0) return;
    const float4 r0 = geom.splat_rec[(size_t)3 * i];
    const float depth = r0.w;
    int x0, y0, x1, y1;
    tile_rect(r0.x, r0.y, r, gx, gy, x0, y0, x1, y1);
    // depth > 0.2 is a positive float: its bit pattern orders like the value;
    // ties are broken by the Gaussian index == upstream's emission order inside one tile.
    const unsigned long long key = ((unsigned long long)__float_as_uint(depth) << 32) | (unsigned)i;
    for (int y = y0; y < y1; y++)
        for (int x = x0; x < x1; x++) {
            unsigned slot = atomicAdd(&tile_cursor[(size_t)(y * gx + x) * EAGS_CTR_STRIDE + EAGS_CTR_CURSOR], 1u);
            unsorted[slot] = key;
        }
}

}  // namespace

// "preprocess_agg" option (c_api.cu): 1 = CTA-aggregated K1 where it applies (default), 0 = always the
// one-atomic-per-key kernel (A/B measurements, parity tests of both)
int launch_preprocess_fwd(const EagsView& v, const EagsGaussians& g, GeomPtrs geom, ImagePtrs img,
                          int32_t* radii, unsigned long long* bucket, unsigned tile_cap, int aggregate,
                          cudaStream_t s) {
    if (g.P <= 0) return 0;
    PreArgs a = make_args(v, g);
    if (wide_sh(a)) {
        const int grid = (g.P + PRE_THREADS - 1) / PRE_THREADS;
        if (!allow_wide_smem(preprocess_fwd_kernel<true>)) return -1;
        preprocess_fwd_kernel<true><<<grid, PRE_THREADS, wide_sh_smem(a), s>>>(
            a, geom, radii, img.tile_count, img.ctrl, tile_cap ? bucket : nullptr, tile_cap);
    } else if (aggregate && img.G <= AGG_MAX_TILES) {
        int dev = 0, sms = 148;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        // one CTA per SM; spans are multiples of the CTA width so every pass of a CTA is fully coalesced
        const int ctas = sms * AGG_CTAS_PER_SM;
        int span = (g.P + ctas - 1) / ctas;
        span = ((span + AGG_THREADS - 1) / AGG_THREADS) * AGG_THREADS;
        const int grid = (g.P + span - 1) / span;
        const size_t smem = ((size_t)img.G + 32) * sizeof(unsigned);
        static bool opted[64] = {};
        if (smem > 48 * 1024 && dev >= 0 && dev < 64 && !opted[dev]) {
            if (cudaFuncSetAttribute(preprocess_fwd_agg_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)(((size_t)AGG_MAX_TILES + 32) * sizeof(unsigned))) != cudaSuccess)
                return -1;
            opted[dev] = true;
        }
        preprocess_fwd_agg_kernel<<<grid, AGG_THREADS, smem, s>>>(
            a, geom, radii, img.tile_count, img.ctrl, tile_cap ? bucket : nullptr, tile_cap, span);
    } else {
        const int grid = (g.P + PRE_THREADS - 1) / PRE_THREADS;
        preprocess_fwd_kernel<false><<<grid, PRE_THREADS, 0, s>>>(
            a, geom, radii, img.tile_count, img.ctrl, tile_cap ? bucket : nullptr, tile_cap);
    }
    return cudaPeekAtLastError() == cudaSuccess ? 1 : -1;
}

int launch_duplicate(const EagsView& v, int P, const int32_t* radii, GeomPtrs geom, ImagePtrs img,
                     BinningPtrs bin, cudaStream_t s) {
    (void)v;
    if (P <= 0) return 0;
    duplicate_kernel<<<(P + PRE_THREADS - 1) / PRE_THREADS, PRE_THREADS, 0, s>>>(
        P, img.gx, img.gy, radii, geom, img.tile_count, img.ctrl, bin.unsorted, bin.capacity);
    return cudaPeekAtLastError() == cudaSuccess ? 1 : -1;
}

