// preprocess_common.cuh -- device helpers shared by the per-Gaussian kernels (sm_100a).
//
// Included by preprocess.cu (K1 preprocess_fwd + K3 duplicate; compiled with -fmad=false so that depth
// keys, radii and tile rectangles are bit-identical to the CPU oracle) and by preprocess_bwd.cu (K8;
// compiled WITH FMA contraction: gradients are compared at 1e-3, not bit-exact).  Everything here is
// `static` / in an anonymous namespace: each translation unit gets its own copy, built with its own
// floating-point flags.
#pragma once
#include "kernels.h"

#define SH_C0 0.28209479177387814f
#define SH_C1 0.4886025119029199f
static __device__ __constant__ float SH_C2[5] = {1.0925484305920792f, -1.0925484305920792f,
                                          0.31539156525252005f, -1.0925484305920792f,
                                          0.5462742152960396f};
static __device__ __constant__ float SH_C3[7] = {-0.5900435899266435f, 2.890611442640554f,
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

// K8 variant: R_w2c is NOT copied into e.Wm (the caller re-reads it from shared memory at every use)
template <typename CamPtr>
__device__ __forceinline__ void ewa_project_lean(CamPtr V, float fx, float fy, float tanfovx,
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
    for (int k = 0; k < 3; k++) {
        e.Mt[0][k] = e.J00 * V[k * 4 + 0] + e.J02 * V[k * 4 + 2];
        e.Mt[1][k] = e.J11 * V[k * 4 + 1] + e.J12 * V[k * 4 + 2];
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

// > 48 KB of dynamic shared memory is an opt-in per kernel and device (idempotent and cheap, so it is
// simply repeated: a per-type cache would be shared by kernels with the same signature)
template <typename K>
bool allow_wide_smem(K kernel) {
    return cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SH_WIDE_SMEM) == cudaSuccess;
}

}  // namespace
