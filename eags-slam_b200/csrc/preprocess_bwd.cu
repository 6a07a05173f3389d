// preprocess_bwd.cu -- K8 preprocess_bwd (+ pose_finalize): conic / mean2D / depth / colour gradients of
// every Gaussian -> means3D, scales, rotations, SH, cov3D and the closed-form camera-pose gradient
// (upstream backward.cu::computeCov2DCUDA + preprocessCUDA).
//
// Its own translation unit so that it is compiled WITH FMA contraction: only the forward's depth keys,
// radii and tile rectangles need IEEE evaluation order (preprocess.cu, -fmad=false); the gradients are
// compared with the oracle at 1e-3 relative.  The EWA projection is recomputed here instead of stored.
#include "preprocess_common.cuh"

namespace {

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
#define PRE_BWD_MIN_CTAS 3
#endif
// POSE: the 15 camera-pose pieces are formed and block-reduced (variant B / the tracking fast path);
// the mapping iteration (no pose gradient) runs the <.., false> instantiation, which keeps none of
// them in registers.
template <bool WIDE, bool POSE>
__global__ void __launch_bounds__(PRE_THREADS, PRE_BWD_MIN_CTAS)
preprocess_bwd_kernel(PreArgs a, BwdArgs b) {
    __shared__ __align__(16) float s_stage[7 * PRE_THREADS];   // routed stores only (collective.cu)
    extern __shared__ float s_sh[];   // WIDE only: SH rows in, dL/dSH rows out (same place)
    // the two 4x4 camera matrices live in shared memory and are re-read (warp-uniform broadcast loads)
    // where they are used: held in registers they cost 32 of the kernel's registers for its whole length
    __shared__ float s_cam_store[32];
    if (threadIdx.x < 32)
        s_cam_store[threadIdx.x] = threadIdx.x < 16 ? __ldg(a.view + threadIdx.x) : __ldg(a.proj + threadIdx.x - 16);
    __syncthreads();
    const volatile float* V = s_cam_store;
    const volatile float* F = s_cam_store + 16;
#define WM(i, k) V[(k) * 4 + (i)]
    const int i = blockIdx.x * PRE_THREADS + threadIdx.x;
    // every input of the Gaussian is requested before its radius is known: one DRAM latency per thread
    // instead of two (the ~10 % culled Gaussians cost a few wasted sectors)
    const bool have_sr = !a.cov3D_precomp;      // kernel-uniform
    float4 ga = make_float4(0.f, 0.f, 0.f, 0.f), gb = ga, gc = ga, q_in = ga;
    float p_in[3] = {0.f, 0.f, 0.f}, sc_in[3] = {0.f, 0.f, 0.f};
    int radius = 0;
    if (i < a.P) {
        radius = b.radii[i];
        ga = b.grad_acc[(size_t)3 * i]; gb = b.grad_acc[(size_t)3 * i + 1]; gc = b.grad_acc[(size_t)3 * i + 2];
        p_in[0] = a.means3D[(size_t)3 * i]; p_in[1] = a.means3D[(size_t)3 * i + 1]; p_in[2] = a.means3D[(size_t)3 * i + 2];
        if (have_sr) {
            sc_in[0] = a.scales[(size_t)3 * i]; sc_in[1] = a.scales[(size_t)3 * i + 1]; sc_in[2] = a.scales[(size_t)3 * i + 2];
            q_in = *reinterpret_cast<const float4*>(a.rotations + (size_t)4 * i);
        }
    }
    const bool vis = radius > 0;
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
    float pose[POSE ? 15 : 1];
#pragma unroll
    for (int k = 0; k < (POSE ? 15 : 1); k++) pose[k] = 0.f;
    if (i < a.P) {
        float dm[3] = {0.f, 0.f, 0.f}, dscale[3] = {0.f, 0.f, 0.f}, drot[4] = {0.f, 0.f, 0.f, 0.f};
        float dS[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        float g2x = 0.f, g2y = 0.f, dop = 0.f;
        float dcol[3] = {0.f, 0.f, 0.f};
        float dmean_sh[3] = {0.f, 0.f, 0.f};
        float p[3] = {0.f, 0.f, 0.f};
        unsigned clamped = 0;
        if (vis) {
            g2x = ga.x; g2y = ga.y;
            const float gxx = ga.z, gxy = ga.w, gyy = gb.x;
            dop = gb.y;
            dcol[0] = gb.z; dcol[1] = gb.w; dcol[2] = gc.x;
            const float dd = gc.y;
            clamped = b.clamped[i];
            p[0] = p_in[0]; p[1] = p_in[1]; p[2] = p_in[2];
            float pv[3];
            pv[0] = V[0] * p[0] + V[4] * p[1] + V[8] * p[2] + V[12];
            pv[1] = V[1] * p[0] + V[5] * p[1] + V[9] * p[2] + V[13];
            pv[2] = V[2] * p[0] + V[6] * p[1] + V[10] * p[2] + V[14];
            float c6[6];
            if (have_sr) cov3d_from_scale_rot(sc_in, a.scale_modifier, q_in, c6);
            else load_cov6(a, i, c6);
            const float fx = (float)a.W / (2.f * a.tanfovx);
            const float fy = (float)a.H / (2.f * a.tanfovy);
            Ewa e;
            ewa_project_lean(V, fx, fy, a.tanfovx, a.tanfovy, pv, c6, e);
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
            const float dJ00 = dMt[0][0] * WM(0, 0) + dMt[0][1] * WM(0, 1) + dMt[0][2] * WM(0, 2);
            const float dJ02 = dMt[0][0] * WM(2, 0) + dMt[0][1] * WM(2, 1) + dMt[0][2] * WM(2, 2);
            const float dJ11 = dMt[1][0] * WM(1, 0) + dMt[1][1] * WM(1, 1) + dMt[1][2] * WM(1, 2);
            const float dJ12 = dMt[1][0] * WM(2, 0) + dMt[1][1] * WM(2, 1) + dMt[1][2] * WM(2, 2);
            const float tz = 1.f / e.t[2], tz2 = tz * tz, tz3 = tz2 * tz;
            float dt[3];
            dt[0] = e.xmul * (-fx * tz2 * dJ02);
            dt[1] = e.ymul * (-fy * tz2 * dJ12);
            dt[2] = -fx * tz2 * dJ00 - fy * tz2 * dJ11 + (2.f * fx * e.t[0]) * tz3 * dJ02 +
                    (2.f * fy * e.t[1]) * tz3 * dJ12;
            float dmean_cov[3];
#pragma unroll
            for (int k = 0; k < 3; k++)
                dmean_cov[k] = WM(0, k) * dt[0] + WM(1, k) * dt[1] + WM(2, k) * dt[2];
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
                const float4 q = q_in;
                float R[3][3];
                quat_to_R(q, R);
                const float s[3] = {a.scale_modifier * sc_in[0], a.scale_modifier * sc_in[1],
                                    a.scale_modifier * sc_in[2]};
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
            if constexpr (POSE) {
                const float dm_pc[3] = {dmean_cov[0] + dmean_proj[0] + dmean_dep[0],
                                        dmean_cov[1] + dmean_proj[1] + dmean_dep[1],
                                        dmean_cov[2] + dmean_proj[2] + dmean_dep[2]};
                float g[3];
#pragma unroll
                for (int r = 0; r < 3; r++)
                    g[r] = WM(r, 0) * dm_pc[0] + WM(r, 1) * dm_pc[1] + WM(r, 2) * dm_pc[2];
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
    if constexpr (POSE) {
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

}  // namespace

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
    const bool pose = b.pose_acc != nullptr;
    if (wide_sh(a)) {
        if (pose) {
            if (!allow_wide_smem(preprocess_bwd_kernel<true, true>)) return -1;
            preprocess_bwd_kernel<true, true><<<grid, PRE_THREADS, wide_sh_smem(a), s>>>(a, b);
        } else {
            if (!allow_wide_smem(preprocess_bwd_kernel<true, false>)) return -1;
            preprocess_bwd_kernel<true, false><<<grid, PRE_THREADS, wide_sh_smem(a), s>>>(a, b);
        }
    } else if (pose) {
        preprocess_bwd_kernel<false, true><<<grid, PRE_THREADS, 0, s>>>(a, b);
    } else {
        preprocess_bwd_kernel<false, false><<<grid, PRE_THREADS, 0, s>>>(a, b);
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
