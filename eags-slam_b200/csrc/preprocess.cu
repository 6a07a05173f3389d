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
    float p[3] = {0.f, 0.f, 0.f};
    float px = 0.f, py = 0.f, depth = 0.f;
    float4 con = make_float4(0.f, 0.f, 0.f, 0.f);
    int r = 0, x0 = 0, y0 = 0, x1 = 0, y1 = 0;
    if (i < a.P) {
        float V[16], F[16];
        load_cam(a, V, F);
        do {
            p[0] = a.means3D[(size_t)3 * i]; p[1] = a.means3D[(size_t)3 * i + 1]; p[2] = a.means3D[(size_t)3 * i + 2];
            float pv[3];
            pv[0] = V[0] * p[0] + V[4] * p[1] + V[8] * p[2] + V[12];
            pv[1] = V[1] * p[0] + V[5] * p[1] + V[9] * p[2] + V[13];
            pv[2] = V[2] * p[0] + V[6] * p[1] + V[10] * p[2] + V[14];
            if (pv[2] <= 0.2f) break;
            float hx = F[0] * p[0] + F[4] * p[1] + F[8] * p[2] + F[12];
            float hy = F[1] * p[0] + F[5] * p[1] + F[9] * p[2] + F[13];
            float hw = F[3] * p[0] + F[7] * p[1] + F[11] * p[2] + F[15];
            float pw = 1.0f / (hw + 0.0000001f);
            float ppx = hx * pw, ppy = hy * pw;
            float c6[6];
            load_cov6(a, i, c6);
            float fx = (float)a.W / (2.f * a.tanfovx);
            float fy = (float)a.H / (2.f * a.tanfovy);
            Ewa e;
            ewa_project(V, fx, fy, a.tanfovx, a.tanfovy, pv, c6, e);
            float ca = e.cov[0], cb = e.cov[1], cc = e.cov[2];
            float det = ca * cc - cb * cb;
            if (det == 0.0f) break;
            float det_inv = 1.f / det;
            float mid = 0.5f * (ca + cc);
            float sq = sqrtf(fmaxf(0.1f, mid * mid - det));
            float lambda1 = mid + sq, lambda2 = mid - sq;
            float my_radius = ceilf(3.f * sqrtf(fmaxf(lambda1, lambda2)));
            px = ((ppx + 1.0f) * (float)a.W - 1.0f) * 0.5f;
            py = ((ppy + 1.0f) * (float)a.H - 1.0f) * 0.5f;
            r = (int)my_radius;
            tile_rect(px, py, r, a.gx, a.gy, x0, y0, x1, y1);
            if ((x1 - x0) * (y1 - y0) == 0) break;
            con = make_float4(cc * det_inv, -cb * det_inv, ca * det_inv, 0.f);
            depth = pv[2];
            alive = true;
        } while (0);
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
            float rgb[3];
            unsigned clamped = 0;
            if (a.colors_precomp) {
                rgb[0] = a.colors_precomp[(size_t)3 * i];
                rgb[1] = a.colors_precomp[(size_t)3 * i + 1];
                rgb[2] = a.colors_precomp[(size_t)3 * i + 2];
            } else {
                float cp[3] = {__ldg(a.campos), __ldg(a.campos + 1), __ldg(a.campos + 2)};
                sh_to_rgb(a.degree, p, cp, sh_row, rgb, clamped);
            }
            geom.xy[i] = make_float2(px, py);
            con.w = a.opacities[i];
            geom.conic_opacity[i] = con;
            geom.rgbd[i] = make_float4(rgb[0], rgb[1], rgb[2], depth);
            geom.clamped[i] = (uint8_t)clamped;
            r_out = r;
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
                const unsigned long long key = ((unsigned long long)__float_as_uint(depth) << 32) | (unsigned)i;
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
    const float2 xy = geom.xy[i];
    const float depth = geom.rgbd[i].w;
    int x0, y0, x1, y1;
    tile_rect(xy.x, xy.y, r, gx, gy, x0, y0, x1, y1);
    // depth > 0.2 is a positive float: its bit pattern orders like the value;
    // ties are broken by the Gaussian index == upstream's emission order inside one tile.
    const unsigned long long key = ((unsigned long long)__float_as_uint(depth) << 32) | (unsigned)i;
    for (int y = y0; y < y1; y++)
        for (int x = x0; x < x1; x++) {
            unsigned slot = atomicAdd(&tile_cursor[(size_t)(y * gx + x) * EAGS_CTR_STRIDE], 1u);
            unsorted[slot] = key;
        }
}

}  // namespace

int launch_preprocess_fwd(const EagsView& v, const EagsGaussians& g, GeomPtrs geom, ImagePtrs img,
                          int32_t* radii, unsigned long long* bucket, unsigned tile_cap, cudaStream_t s) {
    if (g.P <= 0) return 0;
    PreArgs a = make_args(v, g);
    const int grid = (g.P + PRE_THREADS - 1) / PRE_THREADS;
    if (wide_sh(a)) {
        if (!allow_wide_smem(preprocess_fwd_kernel<true>)) return -1;
        preprocess_fwd_kernel<true><<<grid, PRE_THREADS, wide_sh_smem(a), s>>>(
            a, geom, radii, img.tile_count, img.ctrl, tile_cap ? bucket : nullptr, tile_cap);
    } else {
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

