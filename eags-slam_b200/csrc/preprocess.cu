// preprocess.cu -- per-Gaussian kernels of the rasterizer hot path (sm_100a).
//
//   K1 preprocess_fwd : cull, Sigma3D, EWA projection (+0.3 blur), conic, radius, pixel mean, tile
//                       rectangle, SH->RGB, depth; per-tile key counts (RED atomics) and K.
//                       Replaces upstream forward.cu::preprocessCUDA + the CUB InclusiveSum
//                       (SURVEY.md 2.3 / A.1); the per-Gaussian prefix sum disappears because keys
//                       are bucketed per tile (binning.cu).
//   K3 duplicate      : one (depth|index) key per covered tile, scattered into the tile's bucket
//                       (upstream rasterizer_impl.cu::duplicateWithKeys).
//   K8 preprocess_bwd : conic/mean2D/depth/colour gradients -> means3D, scales, rotations, SH,
//                       cov3D and the closed-form camera-pose gradient (upstream
//                       backward.cu::computeCov2DCUDA + preprocessCUDA).
//
// THIS FILE IS COMPILED WITH -fmad=false: every expression is evaluated in IEEE binary32 in the
// order written, so depth keys, radii and tile rectangles are bit-identical to the CPU oracle
// (oracle/eags_oracle.c, built with -ffp-contract=off).  The kernels are HBM/latency bound, the
// missing FMA contraction costs nothing measurable.
#include "kernels.h"

#define SH_C0 0.28209479177387814f
#define SH_C1 0.4886025119029199f
__device__ __constant__ float SH_C2[5] = {1.0925484305920792f, -1.0925484305920792f,
                                          0.31539156525252005f, -1.0925484305920792f,
                                          0.5462742152960396f};
__device__ __constant__ float SH_C3[7] = {-0.5900435899266435f, 2.890611442640554f,
                                          -0.4570457994644658f, 0.3731763325901154f,
                                          -0.4570457994644658f, 1.445305721320277f,
                                          -0.5900435899266435f};

namespace {

constexpr int PRE_THREADS = 256;

struct PreArgs {
    int P, M, degree, W, H, gx, gy;
    float tanfovx, tanfovy, scale_modifier;
    const float* view;
    const float* proj;
    const float* campos;
    const float* means3D;
    const float* opacities;
    const float* shs;
    const float* colors_precomp;
    const float* scales;
    const float* rotations;
    const float* cov3D_precomp;
};

__device__ __forceinline__ void quat_to_R(const float4 q, float R[3][3]) {
    // (w,x,y,z), reference src/utils/gaussian_model_utils.py:182-211
    float r = q.x, x = q.y, y = q.z, z = q.w;
    R[0][0] = 1.f - 2.f * (y * y + z * z);
    R[0][1] = 2.f * (x * y - r * z);
    R[0][2] = 2.f * (x * z + r * y);
    R[1][0] = 2.f * (x * y + r * z);
    R[1][1] = 1.f - 2.f * (x * x + z * z);
    R[1][2] = 2.f * (y * z - r * x);
    R[2][0] = 2.f * (x * z - r * y);
    R[2][1] = 2.f * (y * z + r * x);
    R[2][2] = 1.f - 2.f * (x * x + y * y);
}

// Sigma = (R S)(R S)^T, 6-vector (xx,xy,xz,yy,yz,zz): reference src/entities/gaussian_model.py:70-74
__device__ __forceinline__ void cov3d_from_scale_rot(const float s[3], float mod, const float4 q,
                                                     float c6[6]) {
    float R[3][3], Mx[3][3];
    quat_to_R(q, R);
    float sx = mod * s[0], sy = mod * s[1], sz = mod * s[2];
#pragma unroll
    for (int i = 0; i < 3; i++) {
        Mx[i][0] = R[i][0] * sx;
        Mx[i][1] = R[i][1] * sy;
        Mx[i][2] = R[i][2] * sz;
    }
    c6[0] = Mx[0][0] * Mx[0][0] + Mx[0][1] * Mx[0][1] + Mx[0][2] * Mx[0][2];
    c6[1] = Mx[0][0] * Mx[1][0] + Mx[0][1] * Mx[1][1] + Mx[0][2] * Mx[1][2];
    c6[2] = Mx[0][0] * Mx[2][0] + Mx[0][1] * Mx[2][1] + Mx[0][2] * Mx[2][2];
    c6[3] = Mx[1][0] * Mx[1][0] + Mx[1][1] * Mx[1][1] + Mx[1][2] * Mx[1][2];
    c6[4] = Mx[1][0] * Mx[2][0] + Mx[1][1] * Mx[2][1] + Mx[1][2] * Mx[2][2];
    c6[5] = Mx[2][0] * Mx[2][0] + Mx[2][1] * Mx[2][1] + Mx[2][2] * Mx[2][2];
}

struct Ewa {
    float t[3];
    float xmul, ymul;
    float J00, J02, J11, J12;
    float Wm[3][3];   // R_w2c
    float Mt[2][3];   // J * R_w2c
    float cov[3];     // a, b, c (+0.3 on the diagonal)
};

__device__ __forceinline__ void ewa_project(const float* V, float fx, float fy, float tanfovx,
                                            float tanfovy, const float pview[3], const float c6[6],
                                            Ewa& e) {
    float limx = 1.3f * tanfovx, limy = 1.3f * tanfovy;
    float tz = pview[2];
    float txtz = pview[0] / tz, tytz = pview[1] / tz;
    float cx = fminf(limx, fmaxf(-limx, txtz));
    float cy = fminf(limy, fmaxf(-limy, tytz));
    e.xmul = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
    e.ymul = (tytz < -limy || tytz > limy) ? 0.f : 1.f;
    e.t[0] = cx * tz;
    e.t[1] = cy * tz;
    e.t[2] = tz;
    e.J00 = fx / tz;
    e.J02 = -(fx * e.t[0]) / (tz * tz);
    e.J11 = fy / tz;
    e.J12 = -(fy * e.t[1]) / (tz * tz);
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
        for (int k = 0; k < 3; k++) e.Wm[i][k] = V[k * 4 + i];
#pragma unroll
    for (int k = 0; k < 3; k++) {
        e.Mt[0][k] = e.J00 * e.Wm[0][k] + e.J02 * e.Wm[2][k];
        e.Mt[1][k] = e.J11 * e.Wm[1][k] + e.J12 * e.Wm[2][k];
    }
    float Sg[3][3] = {{c6[0], c6[1], c6[2]}, {c6[1], c6[3], c6[4]}, {c6[2], c6[4], c6[5]}};
    float N[2][3];
#pragma unroll
    for (int i = 0; i < 2; i++)
#pragma unroll
        for (int k = 0; k < 3; k++)
            N[i][k] = e.Mt[i][0] * Sg[0][k] + e.Mt[i][1] * Sg[1][k] + e.Mt[i][2] * Sg[2][k];
    float c00 = N[0][0] * e.Mt[0][0] + N[0][1] * e.Mt[0][1] + N[0][2] * e.Mt[0][2];
    float c01 = N[0][0] * e.Mt[1][0] + N[0][1] * e.Mt[1][1] + N[0][2] * e.Mt[1][2];
    float c11 = N[1][0] * e.Mt[1][0] + N[1][1] * e.Mt[1][1] + N[1][2] * e.Mt[1][2];
    e.cov[0] = c00 + 0.3f;
    e.cov[1] = c01;
    e.cov[2] = c11 + 0.3f;
}

// SH basis of reference src/utils/gaussian_model_utils.py:59-114; shs layout (P, M, 3).
__device__ __forceinline__ void sh_to_rgb(int degree, const float* p, const float* campos,
                                          const float* sh, float rgb[3], unsigned& clamped) {
    float res[3];
#pragma unroll
    for (int c = 0; c < 3; c++) res[c] = SH_C0 * sh[0 * 3 + c];
    if (degree > 0) {
        float dx = p[0] - campos[0], dy = p[1] - campos[1], dz = p[2] - campos[2];
        float len = sqrtf(dx * dx + dy * dy + dz * dz);
        float x = dx / len, y = dy / len, z = dz / len;
#pragma unroll
        for (int c = 0; c < 3; c++)
            res[c] = res[c] - SH_C1 * y * sh[1 * 3 + c] + SH_C1 * z * sh[2 * 3 + c] -
                     SH_C1 * x * sh[3 * 3 + c];
        if (degree > 1) {
            float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
#pragma unroll
            for (int c = 0; c < 3; c++)
                res[c] = res[c] + SH_C2[0] * xy * sh[4 * 3 + c] + SH_C2[1] * yz * sh[5 * 3 + c] +
                         SH_C2[2] * (2.f * zz - xx - yy) * sh[6 * 3 + c] +
                         SH_C2[3] * xz * sh[7 * 3 + c] + SH_C2[4] * (xx - yy) * sh[8 * 3 + c];
            if (degree > 2) {
#pragma unroll
                for (int c = 0; c < 3; c++)
                    res[c] = res[c] + SH_C3[0] * y * (3.f * xx - yy) * sh[9 * 3 + c] +
                             SH_C3[1] * xy * z * sh[10 * 3 + c] +
                             SH_C3[2] * y * (4.f * zz - xx - yy) * sh[11 * 3 + c] +
                             SH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy) * sh[12 * 3 + c] +
                             SH_C3[4] * x * (4.f * zz - xx - yy) * sh[13 * 3 + c] +
                             SH_C3[5] * z * (xx - yy) * sh[14 * 3 + c] +
                             SH_C3[6] * x * (xx - 3.f * yy) * sh[15 * 3 + c];
            }
        }
    }
    clamped = 0;
#pragma unroll
    for (int c = 0; c < 3; c++) {
        res[c] = res[c] + 0.5f;
        if (res[c] < 0.f) clamped |= 1u << c;
        rgb[c] = fmaxf(res[c], 0.f);
    }
}

// tile rectangle of a splat (upstream auxiliary.h::getRect; C (int) truncation)
__device__ __forceinline__ void tile_rect(float px, float py, int r, int gx, int gy, int& x0,
                                          int& y0, int& x1, int& y1) {
    float rf = (float)r;
    x0 = min(gx, max(0, (int)((px - rf) / 16.0f)));
    y0 = min(gy, max(0, (int)((py - rf) / 16.0f)));
    x1 = min(gx, max(0, (int)((px + rf + 16.0f - 1.0f) / 16.0f)));
    y1 = min(gy, max(0, (int)((py + rf + 16.0f - 1.0f) / 16.0f)));
}

__device__ __forceinline__ void load_cam(const PreArgs& a, float V[16], float F[16]) {
#pragma unroll
    for (int k = 0; k < 16; k++) {
        V[k] = __ldg(a.view + k);
        F[k] = __ldg(a.proj + k);
    }
}

__device__ __forceinline__ void load_cov6(const PreArgs& a, int i, float c6[6]) {
    if (a.cov3D_precomp) {
#pragma unroll
        for (int k = 0; k < 6; k++) c6[k] = a.cov3D_precomp[(size_t)6 * i + k];
    } else {
        float s[3] = {a.scales[(size_t)3 * i], a.scales[(size_t)3 * i + 1], a.scales[(size_t)3 * i + 2]};
        float4 q = *reinterpret_cast<const float4*>(a.rotations + (size_t)4 * i);
        cov3d_from_scale_rot(s, a.scale_modifier, q, c6);
    }
}

// ------------------------------------------------------------------------------------------------
// Wide SH rows (M > 1: global-map refinement, evaluate_merged_map.py:67 runs sh_degree 3 -> 48 floats per
// Gaussian).  A thread walking its own 192-byte row makes every load / store instruction of the warp
// touch 32 different lines (32 LSU wavefronts for 128 useful bytes).  Instead each warp moves the rows it
// needs through shared memory with fully coalesced accesses: row r of the warp sits at s_w[r * S .. ],
// S = 3M | 1 (odd -> the per-thread walk over its own row is bank-conflict free).
// ------------------------------------------------------------------------------------------------
constexpr int SH_WIDE_MAX_M = 16;
constexpr int SH_WIDE_STRIDE_MAX = (3 * SH_WIDE_MAX_M) | 1;
constexpr size_t SH_WIDE_SMEM = (size_t)PRE_THREADS * SH_WIDE_STRIDE_MAX * sizeof(float);

__device__ __forceinline__ void cp_async4(float* smem_dst, const float* gmem_src) {
    const unsigned dst = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(gmem_src) : "memory");
}

// rows (bit r of `rows` set; n = rows of the warp inside the cloud) of the warp whose first Gaussian is
// i0w: global -> shared.  Asynchronous 4-byte copies (the odd row stride rules out wider ones): nothing
// waits until every row of the warp is in flight.
__device__ __forceinline__ void warp_rows_in(const float* __restrict__ g, long long i0w, int L, int S, int n,
                                             unsigned rows, float* s_w) {
    const int lane = threadIdx.x & 31;
    if (4 * __popc(rows) >= 3 * n) {
        // most rows are wanted: the warp's rows are one contiguous run of n * L floats
        const float* src = g + (size_t)i0w * L;
        const int total = n * L;
        int off = lane;                      // row * S + col, advanced without a division
        int col = lane;
        while (col >= L) { col -= L; off += S - L; }
#pragma unroll 4
        for (int idx = lane; idx < total; idx += 32) {
            cp_async4(s_w + off, src + idx);
            col += 32; off += 32;
            while (col >= L) { col -= L; off += S - L; }
        }
    } else {
        while (rows) {
            const int r = __ffs(rows) - 1;
            rows &= rows - 1;
            const float* src = g + (size_t)(i0w + r) * L;
            for (int j = lane; j < L; j += 32) cp_async4(s_w + r * S + j, src + j);
        }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncwarp();
}

// all `n` rows of the warp: shared (rows in `rows`) or zeros -> global, as one contiguous run
__device__ __forceinline__ void warp_rows_out(float* __restrict__ g, long long i0w, int L, int S, int n,
                                              unsigned rows, const float* s_w) {
    const int lane = threadIdx.x & 31;
    __syncwarp();
    float* dst = g + (size_t)i0w * L;
    const int total = n * L;
    int off = lane, col = lane, row = 0;
    while (col >= L) { col -= L; off += S - L; row++; }
#pragma unroll 4
    for (int idx = lane; idx < total; idx += 32) {
        dst[idx] = ((rows >> row) & 1u) ? s_w[off] : 0.f;
        col += 32; off += 32;
        while (col >= L) { col -= L; off += S - L; row++; }
    }
}

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

// ------------------------------------------------------------------------------------------------
// K8
// ------------------------------------------------------------------------------------------------
struct BwdArgs {
    const int32_t* radii;
    const uint8_t* clamped;
    const float4* grad_acc;   // [P][3] float4: {dmean2D.x, dmean2D.y, dconic.x, dconic.y}
                              //                {dconic.z, dopacity, dcolor.r, dcolor.g}
                              //                {dcolor.b, ddepth, -, -}
    float* dmeans3D;
    float* dmeans2D;
    float* dopacities;
    float* dshs;
    float* dcolors_precomp;
    float* dscales;
    float* drotations;
    float* dcov3D;
    double* pose_acc;   // 15 doubles or NULL
    // sharded mapping (collective.cu): when route_world > 0 the xyz / scale / rotation / opacity rows
    // of Gaussian i go to route_base[i / route_chunk] (the owner's receive slot of this rank, a peer
    // mapping) at offsets 0 / 3 stride / 6 stride / 10 stride instead of the four pointers above
    float* route_base[EAGS_MAX_PEERS];
    int route_world;
    int route_chunk;
    long long route_stride;
    // optional local partial sums (xyz, scale, rotation, opacity rows of the views this rank already
    // rendered this step) added to this pass's rows before they leave: several views per rank per step
    const float* route_acc[4];
};

#ifndef PRE_BWD_MIN_CTAS
#define PRE_BWD_MIN_CTAS 2
#endif
template <bool WIDE>
__global__ void __launch_bounds__(PRE_THREADS, PRE_BWD_MIN_CTAS)
preprocess_bwd_kernel(PreArgs a, BwdArgs b) {
    __shared__ __align__(16) float s_stage[7 * PRE_THREADS];   // routed stores only (collective.cu)
    extern __shared__ float s_sh[];   // WIDE only: SH rows in, dL/dSH rows out (same place)
    const int i = blockIdx.x * PRE_THREADS + threadIdx.x;
    const bool vis = i < a.P && b.radii[i] > 0;
    float* sh_row = nullptr;
    unsigned sh_rows = 0;
    if constexpr (WIDE) {
        const int L = 3 * a.M, S = L | 1;
        float* s_w = s_sh + (size_t)(threadIdx.x >> 5) * 32 * S;
        sh_rows = __ballot_sync(0xffffffffu, vis);
        const long long i0w = (long long)i - (threadIdx.x & 31);
        warp_rows_in(a.shs, i0w, L, S, (int)max(0ll, min(32ll, (long long)a.P - i0w)), sh_rows, s_w);
        sh_row = s_w + (threadIdx.x & 31) * S;
    }
    float pose[15];
#pragma unroll
    for (int k = 0; k < 15; k++) pose[k] = 0.f;
    if (i < a.P) {
        float dm[3] = {0.f, 0.f, 0.f}, dscale[3] = {0.f, 0.f, 0.f}, drot[4] = {0.f, 0.f, 0.f, 0.f};
        float dS[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        float g2x = 0.f, g2y = 0.f, dop = 0.f;
        float dcol[3] = {0.f, 0.f, 0.f};
        float dmean_sh[3] = {0.f, 0.f, 0.f};
        float p[3] = {0.f, 0.f, 0.f};
        unsigned clamped = 0;
        if (vis) {
            float V[16], F[16];
            load_cam(a, V, F);
            const float4 ga = b.grad_acc[(size_t)3 * i], gb = b.grad_acc[(size_t)3 * i + 1],
                         gc = b.grad_acc[(size_t)3 * i + 2];
            g2x = ga.x; g2y = ga.y;
            const float gxx = ga.z, gxy = ga.w, gyy = gb.x;
            dop = gb.y;
            dcol[0] = gb.z; dcol[1] = gb.w; dcol[2] = gc.x;
            const float dd = gc.y;
            clamped = b.clamped[i];
            p[0] = a.means3D[(size_t)3 * i]; p[1] = a.means3D[(size_t)3 * i + 1]; p[2] = a.means3D[(size_t)3 * i + 2];
            float pv[3];
            pv[0] = V[0] * p[0] + V[4] * p[1] + V[8] * p[2] + V[12];
            pv[1] = V[1] * p[0] + V[5] * p[1] + V[9] * p[2] + V[13];
            pv[2] = V[2] * p[0] + V[6] * p[1] + V[10] * p[2] + V[14];
            float c6[6];
            load_cov6(a, i, c6);
            const float fx = (float)a.W / (2.f * a.tanfovx);
            const float fy = (float)a.H / (2.f * a.tanfovy);
            Ewa e;
            ewa_project(V, fx, fy, a.tanfovx, a.tanfovy, pv, c6, e);
            const float ca = e.cov[0], cb = e.cov[1], cc = e.cov[2];
            const float denom = ca * cc - cb * cb;
            const float denom2inv = 1.0f / (denom * denom + 0.0000001f);   // Q2
            float dLa = 0.f, dLb = 0.f, dLc = 0.f;
            float dMt[2][3] = {{0.f, 0.f, 0.f}, {0.f, 0.f, 0.f}};
            if (denom2inv != 0.f) {
                dLa = denom2inv * (-cc * cc * gxx + 2.f * cb * cc * gxy + (denom - ca * cc) * gyy);
                dLc = denom2inv * (-ca * ca * gyy + 2.f * ca * cb * gxy + (denom - ca * cc) * gxx);
                dLb = denom2inv * 2.f * (cb * cc * gxx - (denom + 2.f * cb * cb) * gxy + ca * cb * gyy);
                const float(*Mt)[3] = e.Mt;
                dS[0] = Mt[0][0] * Mt[0][0] * dLa + Mt[0][0] * Mt[1][0] * dLb + Mt[1][0] * Mt[1][0] * dLc;
                dS[3] = Mt[0][1] * Mt[0][1] * dLa + Mt[0][1] * Mt[1][1] * dLb + Mt[1][1] * Mt[1][1] * dLc;
                dS[5] = Mt[0][2] * Mt[0][2] * dLa + Mt[0][2] * Mt[1][2] * dLb + Mt[1][2] * Mt[1][2] * dLc;
                dS[1] = 2.f * Mt[0][0] * Mt[0][1] * dLa + (Mt[0][0] * Mt[1][1] + Mt[0][1] * Mt[1][0]) * dLb +
                        2.f * Mt[1][0] * Mt[1][1] * dLc;
                dS[2] = 2.f * Mt[0][0] * Mt[0][2] * dLa + (Mt[0][0] * Mt[1][2] + Mt[0][2] * Mt[1][0]) * dLb +
                        2.f * Mt[1][0] * Mt[1][2] * dLc;
                dS[4] = 2.f * Mt[0][2] * Mt[0][1] * dLa + (Mt[0][1] * Mt[1][2] + Mt[0][2] * Mt[1][1]) * dLb +
                        2.f * Mt[1][1] * Mt[1][2] * dLc;
                const float Sg[3][3] = {{c6[0], c6[1], c6[2]}, {c6[1], c6[3], c6[4]}, {c6[2], c6[4], c6[5]}};
#pragma unroll
                for (int k = 0; k < 3; k++) {
                    float m0 = Mt[0][0] * Sg[0][k] + Mt[0][1] * Sg[1][k] + Mt[0][2] * Sg[2][k];
                    float m1 = Mt[1][0] * Sg[0][k] + Mt[1][1] * Sg[1][k] + Mt[1][2] * Sg[2][k];
                    dMt[0][k] = 2.f * m0 * dLa + m1 * dLb;
                    dMt[1][k] = 2.f * m1 * dLc + m0 * dLb;
                }
            }
            // dL/dJ = dL/dMt * W^T
            const float dJ00 = dMt[0][0] * e.Wm[0][0] + dMt[0][1] * e.Wm[0][1] + dMt[0][2] * e.Wm[0][2];
            const float dJ02 = dMt[0][0] * e.Wm[2][0] + dMt[0][1] * e.Wm[2][1] + dMt[0][2] * e.Wm[2][2];
            const float dJ11 = dMt[1][0] * e.Wm[1][0] + dMt[1][1] * e.Wm[1][1] + dMt[1][2] * e.Wm[1][2];
            const float dJ12 = dMt[1][0] * e.Wm[2][0] + dMt[1][1] * e.Wm[2][1] + dMt[1][2] * e.Wm[2][2];
            const float tz = 1.f / e.t[2], tz2 = tz * tz, tz3 = tz2 * tz;
            float dt[3];
            dt[0] = e.xmul * (-fx * tz2 * dJ02);
            dt[1] = e.ymul * (-fy * tz2 * dJ12);
            dt[2] = -fx * tz2 * dJ00 - fy * tz2 * dJ11 + (2.f * fx * e.t[0]) * tz3 * dJ02 +
                    (2.f * fy * e.t[1]) * tz3 * dJ12;
            float dmean_cov[3];
#pragma unroll
            for (int k = 0; k < 3; k++)
                dmean_cov[k] = e.Wm[0][k] * dt[0] + e.Wm[1][k] * dt[1] + e.Wm[2][k] * dt[2];
            // projection path (Q3)
            const float hx = F[0] * p[0] + F[4] * p[1] + F[8] * p[2] + F[12];
            const float hy = F[1] * p[0] + F[5] * p[1] + F[9] * p[2] + F[13];
            const float hw = F[3] * p[0] + F[7] * p[1] + F[11] * p[2] + F[15];
            const float mw = 1.0f / (hw + 0.0000001f);
            const float mul1 = hx * mw * mw, mul2 = hy * mw * mw;
            float dmean_proj[3];
            dmean_proj[0] = (F[0] * mw - F[3] * mul1) * g2x + (F[1] * mw - F[3] * mul2) * g2y;
            dmean_proj[1] = (F[4] * mw - F[7] * mul1) * g2x + (F[5] * mw - F[7] * mul2) * g2y;
            dmean_proj[2] = (F[8] * mw - F[11] * mul1) * g2x + (F[9] * mw - F[11] * mul2) * g2y;
            // depth path: depth = row 2 of the view transform
            const float dmean_dep[3] = {V[2] * dd, V[6] * dd, V[10] * dd};
            // colour path: SH backward (incl. the view-direction term for degree > 0)
            // (wide rows: dsh overwrites sh in shared memory, so within each band every sh[m] is read
            // before dsh[m] is written; the view-direction term is also due when the SH stay frozen)
            if (!a.colors_precomp && (WIDE || b.dshs)) {
                float dRGB[3];
#pragma unroll
                for (int k = 0; k < 3; k++) dRGB[k] = (clamped >> k & 1u) ? 0.f : dcol[k];
                float* dsh = WIDE ? sh_row : b.dshs + (size_t)i * a.M * 3;
                const float* sh = WIDE ? sh_row : a.shs + (size_t)i * a.M * 3;
#pragma unroll
                for (int k = 0; k < 3; k++) dsh[k] = SH_C0 * dRGB[k];
                const int used = (a.degree + 1) * (a.degree + 1);
                for (int m = used; m < a.M; m++) {
                    dsh[m * 3] = 0.f; dsh[m * 3 + 1] = 0.f; dsh[m * 3 + 2] = 0.f;
                }
                if (a.degree > 0) {
                    const float cpx = __ldg(a.campos), cpy = __ldg(a.campos + 1), cpz = __ldg(a.campos + 2);
                    const float ox = p[0] - cpx, oy = p[1] - cpy, oz = p[2] - cpz;
                    const float len = sqrtf(ox * ox + oy * oy + oz * oz);
                    const float x = ox / len, y = oy / len, z = oz / len;
                    float dRGBdx[3], dRGBdy[3], dRGBdz[3];
#pragma unroll
                    for (int k = 0; k < 3; k++) {
                        dRGBdx[k] = -SH_C1 * sh[3 * 3 + k];
                        dRGBdy[k] = -SH_C1 * sh[1 * 3 + k];
                        dRGBdz[k] = SH_C1 * sh[2 * 3 + k];
                        dsh[1 * 3 + k] = -SH_C1 * y * dRGB[k];
                        dsh[2 * 3 + k] = SH_C1 * z * dRGB[k];
                        dsh[3 * 3 + k] = -SH_C1 * x * dRGB[k];
                    }
                    if (a.degree > 1) {
                        const float xx = x * x, yy = y * y, zz = z * z, xy_ = x * y, yz = y * z, xz = x * z;
#pragma unroll
                        for (int k = 0; k < 3; k++) {
                            dRGBdx[k] += SH_C2[0] * y * sh[4 * 3 + k] + SH_C2[2] * 2.f * -x * sh[6 * 3 + k] +
                                         SH_C2[3] * z * sh[7 * 3 + k] + SH_C2[4] * 2.f * x * sh[8 * 3 + k];
                            dRGBdy[k] += SH_C2[0] * x * sh[4 * 3 + k] + SH_C2[1] * z * sh[5 * 3 + k] +
                                         SH_C2[2] * 2.f * -y * sh[6 * 3 + k] + SH_C2[4] * 2.f * -y * sh[8 * 3 + k];
                            dRGBdz[k] += SH_C2[1] * y * sh[5 * 3 + k] + SH_C2[2] * 2.f * 2.f * z * sh[6 * 3 + k] +
                                         SH_C2[3] * x * sh[7 * 3 + k];
                            dsh[4 * 3 + k] = SH_C2[0] * xy_ * dRGB[k];
                            dsh[5 * 3 + k] = SH_C2[1] * yz * dRGB[k];
                            dsh[6 * 3 + k] = SH_C2[2] * (2.f * zz - xx - yy) * dRGB[k];
                            dsh[7 * 3 + k] = SH_C2[3] * xz * dRGB[k];
                            dsh[8 * 3 + k] = SH_C2[4] * (xx - yy) * dRGB[k];
                        }
                        if (a.degree > 2) {
#pragma unroll
                            for (int k = 0; k < 3; k++) {
                                dRGBdx[k] += SH_C3[0] * sh[9 * 3 + k] * 3.f * 2.f * xy_ +
                                             SH_C3[1] * sh[10 * 3 + k] * yz +
                                             SH_C3[2] * sh[11 * 3 + k] * -2.f * xy_ +
                                             SH_C3[3] * sh[12 * 3 + k] * -3.f * 2.f * xz +
                                             SH_C3[4] * sh[13 * 3 + k] * (-3.f * xx + 4.f * zz - yy) +
                                             SH_C3[5] * sh[14 * 3 + k] * 2.f * xz +
                                             SH_C3[6] * sh[15 * 3 + k] * 3.f * (xx - yy);
                                dRGBdy[k] += SH_C3[0] * sh[9 * 3 + k] * 3.f * (xx - yy) +
                                             SH_C3[1] * sh[10 * 3 + k] * xz +
                                             SH_C3[2] * sh[11 * 3 + k] * (-3.f * yy + 4.f * zz - xx) +
                                             SH_C3[3] * sh[12 * 3 + k] * -3.f * 2.f * yz +
                                             SH_C3[4] * sh[13 * 3 + k] * -2.f * xy_ +
                                             SH_C3[5] * sh[14 * 3 + k] * -2.f * yz +
                                             SH_C3[6] * sh[15 * 3 + k] * -3.f * 2.f * xy_;
                                dRGBdz[k] += SH_C3[1] * sh[10 * 3 + k] * xy_ +
                                             SH_C3[2] * sh[11 * 3 + k] * 4.f * 2.f * yz +
                                             SH_C3[3] * sh[12 * 3 + k] * 3.f * (2.f * zz - xx - yy) +
                                             SH_C3[4] * sh[13 * 3 + k] * 4.f * 2.f * xz +
                                             SH_C3[5] * sh[14 * 3 + k] * (xx - yy);
                                dsh[9 * 3 + k] = SH_C3[0] * y * (3.f * xx - yy) * dRGB[k];
                                dsh[10 * 3 + k] = SH_C3[1] * xy_ * z * dRGB[k];
                                dsh[11 * 3 + k] = SH_C3[2] * y * (4.f * zz - xx - yy) * dRGB[k];
                                dsh[12 * 3 + k] = SH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy) * dRGB[k];
                                dsh[13 * 3 + k] = SH_C3[4] * x * (4.f * zz - xx - yy) * dRGB[k];
                                dsh[14 * 3 + k] = SH_C3[5] * z * (xx - yy) * dRGB[k];
                                dsh[15 * 3 + k] = SH_C3[6] * x * (xx - 3.f * yy) * dRGB[k];
                            }
                        }
                    }
                    float ddir[3] = {0.f, 0.f, 0.f};
#pragma unroll
                    for (int k = 0; k < 3; k++) {
                        ddir[0] += dRGBdx[k] * dRGB[k];
                        ddir[1] += dRGBdy[k] * dRGB[k];
                        ddir[2] += dRGBdz[k] * dRGB[k];
                    }
                    const float inv = 1.f / len;
                    const float dot = (x * ddir[0] + y * ddir[1] + z * ddir[2]);
                    dmean_sh[0] = (ddir[0] - x * dot) * inv;
                    dmean_sh[1] = (ddir[1] - y * dot) * inv;
                    dmean_sh[2] = (ddir[2] - z * dot) * inv;
                }
            }
#pragma unroll
            for (int k = 0; k < 3; k++) dm[k] = dmean_cov[k] + dmean_proj[k] + dmean_dep[k] + dmean_sh[k];
            // Sigma -> scale, rotation
            if (!a.cov3D_precomp && (b.dscales || b.drotations)) {
                const float4 q = *reinterpret_cast<const float4*>(a.rotations + (size_t)4 * i);
                float R[3][3];
                quat_to_R(q, R);
                const float s[3] = {a.scale_modifier * a.scales[(size_t)3 * i],
                                    a.scale_modifier * a.scales[(size_t)3 * i + 1],
                                    a.scale_modifier * a.scales[(size_t)3 * i + 2]};
                const float Dm[3][3] = {{dS[0], 0.5f * dS[1], 0.5f * dS[2]},
                                        {0.5f * dS[1], dS[3], 0.5f * dS[4]},
                                        {0.5f * dS[2], 0.5f * dS[4], dS[5]}};
                float Mx[3][3], dM[3][3], dR[3][3];
#pragma unroll
                for (int r = 0; r < 3; r++)
#pragma unroll
                    for (int k = 0; k < 3; k++) Mx[r][k] = R[r][k] * s[k];
#pragma unroll
                for (int r = 0; r < 3; r++)
#pragma unroll
                    for (int k = 0; k < 3; k++)
                        dM[r][k] = 2.f * (Dm[r][0] * Mx[0][k] + Dm[r][1] * Mx[1][k] + Dm[r][2] * Mx[2][k]);
#pragma unroll
                for (int k = 0; k < 3; k++) {
                    dscale[k] = a.scale_modifier * (dM[0][k] * R[0][k] + dM[1][k] * R[1][k] + dM[2][k] * R[2][k]);
#pragma unroll
                    for (int r = 0; r < 3; r++) dR[r][k] = dM[r][k] * s[k];
                }
                const float qr = q.x, qx = q.y, qy = q.z, qz = q.w;
                drot[0] = 2.f * (-qz * dR[0][1] + qy * dR[0][2] + qz * dR[1][0] - qx * dR[1][2] -
                                 qy * dR[2][0] + qx * dR[2][1]);
                drot[1] = 2.f * (qy * dR[0][1] + qz * dR[0][2] + qy * dR[1][0] - 2.f * qx * dR[1][1] -
                                 qr * dR[1][2] + qz * dR[2][0] + qr * dR[2][1] - 2.f * qx * dR[2][2]);
                drot[2] = 2.f * (-2.f * qy * dR[0][0] + qx * dR[0][1] + qr * dR[0][2] + qx * dR[1][0] +
                                 qz * dR[1][2] - qr * dR[2][0] + qz * dR[2][1] - 2.f * qy * dR[2][2]);
                drot[3] = 2.f * (-2.f * qz * dR[0][0] - qr * dR[0][1] + qx * dR[0][2] + qr * dR[1][0] -
                                 2.f * qz * dR[1][1] + qy * dR[1][2] + qx * dR[2][0] + qy * dR[2][1]);
            }
            // camera-pose pieces (SURVEY A.5 closed form): g = dL/dp_C with R fixed, G = dL/dR_w2c
            if (b.pose_acc) {
                const float dm_pc[3] = {dmean_cov[0] + dmean_proj[0] + dmean_dep[0],
                                        dmean_cov[1] + dmean_proj[1] + dmean_dep[1],
                                        dmean_cov[2] + dmean_proj[2] + dmean_dep[2]};
                float g[3];
#pragma unroll
                for (int r = 0; r < 3; r++)
                    g[r] = e.Wm[r][0] * dm_pc[0] + e.Wm[r][1] * dm_pc[1] + e.Wm[r][2] * dm_pc[2];
                pose[0] = g[0]; pose[1] = g[1]; pose[2] = g[2];
                pose[3] = pv[1] * g[2] - pv[2] * g[1];
                pose[4] = pv[2] * g[0] - pv[0] * g[2];
                pose[5] = pv[0] * g[1] - pv[1] * g[0];
                const float Jm[2][3] = {{e.J00, 0.f, e.J02}, {0.f, e.J11, e.J12}};
#pragma unroll
                for (int m = 0; m < 3; m++)
#pragma unroll
                    for (int k = 0; k < 3; k++)
                        pose[6 + 3 * m + k] = Jm[0][m] * dMt[0][k] + Jm[1][m] * dMt[1][k];
            }
        } else if (!WIDE && !a.colors_precomp && b.dshs) {   // (wide rows: zeros come from warp_rows_out)
            float* dsh = b.dshs + (size_t)i * a.M * 3;
            for (int m = 0; m < a.M * 3; m++) dsh[m] = 0.f;
        }
        // stores (all P rows are written: the binding hands in torch.empty buffers)
        if (b.route_world > 0) {
            // sharded mapping: rows go to their owner rank over NVLink.  xyz / scale / opacity are staged
            // in shared memory and leave as 16-byte vectors below (4-byte peer stores waste the link);
            // the rotation row already is one vector.
            if (b.route_acc[0]) {   // kernel-uniform
#pragma unroll
                for (int k = 0; k < 3; k++) {
                    dm[k] += b.route_acc[0][(size_t)3 * i + k];
                    dscale[k] += b.route_acc[1][(size_t)3 * i + k];
                }
                const float4 ar = *reinterpret_cast<const float4*>(b.route_acc[2] + (size_t)4 * i);
                drot[0] += ar.x; drot[1] += ar.y; drot[2] += ar.z; drot[3] += ar.w;
                dop += b.route_acc[3][i];
            }
            s_stage[3 * threadIdx.x] = dm[0]; s_stage[3 * threadIdx.x + 1] = dm[1]; s_stage[3 * threadIdx.x + 2] = dm[2];
            s_stage[3 * PRE_THREADS + 3 * threadIdx.x] = dscale[0];
            s_stage[3 * PRE_THREADS + 3 * threadIdx.x + 1] = dscale[1];
            s_stage[3 * PRE_THREADS + 3 * threadIdx.x + 2] = dscale[2];
            s_stage[6 * PRE_THREADS + threadIdx.x] = dop;
            float* base = b.route_base[min(i / b.route_chunk, b.route_world - 1)];
            *reinterpret_cast<float4*>(base + 6 * b.route_stride + (size_t)4 * i) = make_float4(drot[0], drot[1], drot[2], drot[3]);
        } else {
            if (b.dmeans3D) {
                b.dmeans3D[(size_t)3 * i] = dm[0]; b.dmeans3D[(size_t)3 * i + 1] = dm[1]; b.dmeans3D[(size_t)3 * i + 2] = dm[2];
            }
            if (b.dopacities) b.dopacities[i] = dop;
            if (b.dscales) {
                b.dscales[(size_t)3 * i] = dscale[0]; b.dscales[(size_t)3 * i + 1] = dscale[1]; b.dscales[(size_t)3 * i + 2] = dscale[2];
            }
            if (b.drotations)
                *reinterpret_cast<float4*>(b.drotations + (size_t)4 * i) = make_float4(drot[0], drot[1], drot[2], drot[3]);
        }
        if (b.dmeans2D) {
            b.dmeans2D[(size_t)3 * i] = g2x; b.dmeans2D[(size_t)3 * i + 1] = g2y; b.dmeans2D[(size_t)3 * i + 2] = 0.f;
        }
        if (b.dcolors_precomp) {
            b.dcolors_precomp[(size_t)3 * i] = dcol[0]; b.dcolors_precomp[(size_t)3 * i + 1] = dcol[1];
            b.dcolors_precomp[(size_t)3 * i + 2] = dcol[2];
        }
        if (b.dcov3D) {
#pragma unroll
            for (int k = 0; k < 6; k++) b.dcov3D[(size_t)6 * i + k] = dS[k];
        }
    }
    if constexpr (WIDE) {
        if (b.dshs) {
            const long long i0w = (long long)i - (threadIdx.x & 31);
            const int L = 3 * a.M, n = (int)max(0ll, min(32ll, (long long)a.P - i0w));
            warp_rows_out(b.dshs, i0w, L, L | 1, n, sh_rows, sh_row - (threadIdx.x & 31) * (L | 1));
        }
    }
    if (b.route_world > 0) {   // kernel-uniform
        if (i >= a.P) {        // rows past the cloud pad the last vectors with zeros
#pragma unroll
            for (int k = 0; k < 3; k++) {
                s_stage[3 * threadIdx.x + k] = 0.f;
                s_stage[3 * PRE_THREADS + 3 * threadIdx.x + k] = 0.f;
            }
            s_stage[6 * PRE_THREADS + threadIdx.x] = 0.f;
        }
        __syncthreads();
        // the CTA's 256 rows are 192 + 192 + 64 vectors; owner boundaries are multiples of 4 Gaussians, so
        // a vector never straddles two owners
        const long long i0 = (long long)blockIdx.x * PRE_THREADS;
        constexpr int V3 = 3 * PRE_THREADS / 4, V1 = PRE_THREADS / 4;
        for (int t = threadIdx.x; t < 2 * V3 + V1; t += PRE_THREADS) {
            const int seg = t < V3 ? 0 : (t < 2 * V3 ? 1 : 2);
            const int tt = t - seg * V3, k = seg < 2 ? 3 : 1;
            const long long gidx = i0 + (4 * tt) / k;
            if (gidx >= a.P) continue;
            float* base = b.route_base[min((int)(gidx / b.route_chunk), b.route_world - 1)];
            const long long off = (seg == 0 ? 0 : (seg == 1 ? 3 : 10)) * b.route_stride + k * i0 + 4 * tt;
            *reinterpret_cast<float4*>(base + off) =
                *reinterpret_cast<const float4*>(&s_stage[(seg == 0 ? 0 : (seg == 1 ? 3 : 6)) * PRE_THREADS + 4 * tt]);
        }
    }
    if (b.pose_acc) {
        // block reduction of the 15 pose pieces: warp shuffles, then one fp64 atomic per block
        __shared__ float s_pose[PRE_THREADS / 32][16];
#pragma unroll
        for (int k = 0; k < 15; k++) {
            float v = pose[k];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            if ((threadIdx.x & 31) == 0) s_pose[threadIdx.x >> 5][k] = v;
        }
        __syncthreads();
        if (threadIdx.x < 15) {
            double acc = 0.0;
#pragma unroll
            for (int w = 0; w < PRE_THREADS / 32; w++) acc += (double)s_pose[w][threadIdx.x];
            if (acc != 0.0) atomicAdd(&b.pose_acc[threadIdx.x], acc);
        }
    }
}

// dL/dtau = [sum g | sum p_C x g + vee(A - A^T)],  A = (sum G) R_w2c^T   
__global__ void pose_finalize_kernel(const double* __restrict__ pose, const float* __restrict__ view,
                                     float* __restrict__ dL_dtau) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    double Wm[3][3], A[3][3];
    for (int i = 0; i < 3; i++)
        for (int k = 0; k < 3; k++) Wm[i][k] = (double)view[k * 4 + i];
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) {
            A[i][j] = 0.0;
            for (int k = 0; k < 3; k++) A[i][j] += pose[6 + 3 * i + k] * Wm[j][k];
        }
    dL_dtau[0] = (float)pose[0];
    dL_dtau[1] = (float)pose[1];
    dL_dtau[2] = (float)pose[2];
    dL_dtau[3] = (float)(pose[3] + (A[2][1] - A[1][2]));
    dL_dtau[4] = (float)(pose[4] + (A[0][2] - A[2][0]));
    dL_dtau[5] = (float)(pose[5] + (A[1][0] - A[0][1]));
}

PreArgs make_args(const EagsView& v, const EagsGaussians& g) {
    PreArgs a;
    a.P = g.P; a.M = g.M; a.degree = v.sh_degree; a.W = v.image_width; a.H = v.image_height;
    a.gx = (a.W + EAGS_TILE - 1) / EAGS_TILE; a.gy = (a.H + EAGS_TILE - 1) / EAGS_TILE;
    a.tanfovx = v.tanfovx; a.tanfovy = v.tanfovy; a.scale_modifier = v.scale_modifier;
    a.view = v.viewmatrix; a.proj = v.projmatrix; a.campos = v.campos;
    a.means3D = g.means3D; a.opacities = g.opacities; a.shs = g.shs;
    a.colors_precomp = g.colors_precomp; a.scales = g.scales; a.rotations = g.rotations;
    a.cov3D_precomp = g.cov3D_precomp;
    return a;
}

// SH rows wider than one coefficient go through the shared-memory staging of the <true> kernels
inline bool wide_sh(const PreArgs& a) {
    return a.shs && !a.colors_precomp && a.M > 1 && a.M <= SH_WIDE_MAX_M;
}
inline size_t wide_sh_smem(const PreArgs& a) { return (size_t)PRE_THREADS * ((3 * a.M) | 1) * sizeof(float); }

// > 48 KB of dynamic shared memory is an opt-in per kernel and device
template <typename K>
bool allow_wide_smem(K kernel) {
    static bool done[64] = {};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return false;
    if (!done[dev]) {
        if (cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SH_WIDE_SMEM) != cudaSuccess)
            return false;
        done[dev] = true;
    }
    return true;
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

int launch_preprocess_bwd(const EagsView& v, const EagsGaussians& g, const int32_t* radii,
                          GeomPtrs geom, const float* grad_acc, const EagsInputGrads& gout,
                          double* pose_acc, const EagsPeerRoute* route, cudaStream_t s) {
    if (g.P <= 0) return 0;
    PreArgs a = make_args(v, g);
    BwdArgs b;
    b.route_world = 0; b.route_chunk = 1; b.route_stride = 0;
    for (int k = 0; k < EAGS_MAX_PEERS; k++) b.route_base[k] = nullptr;
    for (int k = 0; k < 4; k++) b.route_acc[k] = nullptr;
    b.radii = radii; b.clamped = geom.clamped; b.grad_acc = (const float4*)grad_acc;
    b.dmeans3D = gout.dL_dmeans3D; b.dmeans2D = gout.dL_dmeans2D; b.dopacities = gout.dL_dopacities;
    b.dshs = gout.dL_dshs; b.dcolors_precomp = gout.dL_dcolors_precomp; b.dscales = gout.dL_dscales;
    b.drotations = gout.dL_drotations; b.dcov3D = gout.dL_dcov3D_precomp;
    b.pose_acc = gout.dL_dtau ? pose_acc : nullptr;
    if (route) {
        const long long stride = ((long long)g.P + 3) & ~3ll;
        b.route_world = route->world;
        b.route_chunk = (int)(((((long long)g.P + route->world - 1) / route->world) + 3) & ~3ll);
        b.route_stride = stride;
        for (int o = 0; o < route->world; o++) b.route_base[o] = route->recv[o] + (long long)route->rank * 11 * stride;
        // non-null so that the scale / rotation chain is evaluated; the stores follow the route
        b.dmeans3D = b.dscales = b.drotations = b.dopacities = b.route_base[route->rank];
        if (route->accum[0])
            for (int k = 0; k < 4; k++) b.route_acc[k] = route->accum[k];
    }
    const int grid = (g.P + PRE_THREADS - 1) / PRE_THREADS;
    if (wide_sh(a)) {
        if (!allow_wide_smem(preprocess_bwd_kernel<true>)) return -1;
        preprocess_bwd_kernel<true><<<grid, PRE_THREADS, wide_sh_smem(a), s>>>(a, b);
    } else {
        preprocess_bwd_kernel<false><<<grid, PRE_THREADS, 0, s>>>(a, b);
    }
    if (cudaPeekAtLastError() != cudaSuccess) return -1;
    int n = 1;
    if (gout.dL_dtau) {
        pose_finalize_kernel<<<1, 32, 0, s>>>(pose_acc, v.viewmatrix, gout.dL_dtau);
        if (cudaPeekAtLastError() != cudaSuccess) return -1;
        n++;
    }
    return n;
}
