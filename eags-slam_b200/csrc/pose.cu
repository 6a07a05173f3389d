// pose.cu -- the tracking fast path.
//
// The reference's tracker (src/entities/tracker.py:85-97) moves ALL P Gaussians by an optimisable
// relative pose in PyTorch every iteration -- means <- R(q) p + t, rotations <- q (x) r -- renders them
// from the fixed previous camera and back-propagates through dL/dmeans3D (P,3) and dL/drotations (P,4).
// Rendering the moved Gaussians from camera L is the same as rendering the ORIGINAL Gaussians from the
// camera L * Rel (p_view and J W Sigma W^T J^T are identical), so here the pose is folded into the view
// matrices by a one-thread kernel and the 7 pose gradients are recovered from the rasterizer's own
// camera-pose gradient dL/dtau (left perturbation of the full w2c) by a second one:
//   K11 pose_compose : (q, t, viewmatrix, projmatrix) -> viewmatrix', projmatrix', campos'
//   K12 pose_chain   : (q, t, viewmatrix, dL/dtau)   -> dL/dq (4), dL/dt (3)
// No P-sized tensor is created or differentiated.
#include "kernels.h"

namespace {

__device__ void quat_to_R_normalised(const float* q, float R[3][3], float qh[4], float& norm) {
    // F.normalize(q) then build_rotation (reference src/utils/gaussian_model_utils.py:182-211), wxyz
    norm = sqrtf(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
    const float inv = 1.f / fmaxf(norm, 1e-12f);
    for (int k = 0; k < 4; k++) qh[k] = q[k] * inv;
    const float r = qh[0], x = qh[1], y = qh[2], z = qh[3];
    R[0][0] = 1.f - 2.f * (y * y + z * z); R[0][1] = 2.f * (x * y - r * z); R[0][2] = 2.f * (x * z + r * y);
    R[1][0] = 2.f * (x * y + r * z); R[1][1] = 1.f - 2.f * (x * x + z * z); R[1][2] = 2.f * (y * z - r * x);
    R[2][0] = 2.f * (x * z - r * y); R[2][1] = 2.f * (y * z + r * x); R[2][2] = 1.f - 2.f * (x * x + y * y);
}

__global__ void pose_compose_kernel(const float* __restrict__ q, const float* __restrict__ t,
                                    const float* __restrict__ view, const float* __restrict__ proj,
                                    const float* __restrict__ campos, float* __restrict__ view_out,
                                    float* __restrict__ proj_out, float* __restrict__ campos_out) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    float R[3][3], qh[4], norm;
    quat_to_R_normalised(q, R, qh, norm);
    // row-vector convention: p' = p M with M = Rel^T, so view' = M view and proj' = M proj
    float M[4][4];
    for (int i = 0; i < 3; i++) {
        for (int k = 0; k < 3; k++) M[i][k] = R[k][i];
        M[i][3] = 0.f;
        M[3][i] = t[i];
    }
    M[3][3] = 1.f;
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 4; j++) {
            float a = 0.f, b = 0.f;
            for (int k = 0; k < 4; k++) {
                a += M[i][k] * view[k * 4 + j];
                b += M[i][k] * proj[k * 4 + j];
            }
            view_out[i * 4 + j] = a;
            proj_out[i * 4 + j] = b;
        }
    // camera centre expressed in the un-moved frame: Rel^-1 campos = R^T (campos - t)
    const float d[3] = {campos[0] - t[0], campos[1] - t[1], campos[2] - t[2]};
    for (int i = 0; i < 3; i++) campos_out[i] = R[0][i] * d[0] + R[1][i] * d[1] + R[2][i] * d[2];
}

__device__ void cross3(const float a[3], const float b[3], float o[3]) {
    o[0] = a[1] * b[2] - a[2] * b[1];
    o[1] = a[2] * b[0] - a[0] * b[2];
    o[2] = a[0] * b[1] - a[1] * b[0];
}

// dL/dtau = [g_rho, g_theta] is the gradient w.r.t. the left perturbation of T = L Rel.  A left
// perturbation xi' of Rel maps to tau = Ad_L xi', hence
//   g_rho'   = R_L^T g_rho,        g_theta' = R_L^T (g_theta + g_rho x t_L)
//   dL/dt    = g_rho',             w = g_theta' - t x g_rho'
//   dL/dq^   = 2 (0, w) (x) q^,    dL/dq = (I - q^ q^T) dL/dq^ / |q|
// (checked against float64 autograd through the reference's SE3_exp, tests/test_pose_chain.py).
__global__ void pose_chain_kernel(const float* __restrict__ q, const float* __restrict__ t,
                                  const float* __restrict__ view, const float* __restrict__ dtau,
                                  float* __restrict__ dq, float* __restrict__ dt) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    float RL[3][3], tL[3];
    for (int i = 0; i < 3; i++) {
        for (int j = 0; j < 3; j++) RL[i][j] = view[j * 4 + i];
        tL[i] = view[3 * 4 + i];
    }
    const float g_rho[3] = {dtau[0], dtau[1], dtau[2]}, g_th[3] = {dtau[3], dtau[4], dtau[5]};
    float c[3];
    cross3(g_rho, tL, c);
    float grp[3], gtp[3];
    for (int i = 0; i < 3; i++) {
        grp[i] = RL[0][i] * g_rho[0] + RL[1][i] * g_rho[1] + RL[2][i] * g_rho[2];
        gtp[i] = RL[0][i] * (g_th[0] + c[0]) + RL[1][i] * (g_th[1] + c[1]) + RL[2][i] * (g_th[2] + c[2]);
    }
    const float tt[3] = {t[0], t[1], t[2]};
    float txg[3];
    cross3(tt, grp, txg);
    const float w[3] = {gtp[0] - txg[0], gtp[1] - txg[1], gtp[2] - txg[2]};
    float R[3][3], qh[4], norm;
    quat_to_R_normalised(q, R, qh, norm);
    // (0, w) (x) q^   (Hamilton product, wxyz)
    float dqh[4];
    dqh[0] = 2.f * (-w[0] * qh[1] - w[1] * qh[2] - w[2] * qh[3]);
    dqh[1] = 2.f * (w[0] * qh[0] + w[1] * qh[3] - w[2] * qh[2]);
    dqh[2] = 2.f * (-w[0] * qh[3] + w[1] * qh[0] + w[2] * qh[1]);
    dqh[3] = 2.f * (w[0] * qh[2] - w[1] * qh[1] + w[2] * qh[0]);
    const float dot = qh[0] * dqh[0] + qh[1] * dqh[1] + qh[2] * dqh[2] + qh[3] * dqh[3];
    const float inv = 1.f / fmaxf(norm, 1e-12f);
    for (int k = 0; k < 4; k++) dq[k] = (dqh[k] - qh[k] * dot) * inv;
    for (int i = 0; i < 3; i++) dt[i] = grp[i];
}

}  // namespace

int launch_pose_compose(const float* q, const float* t, const float* view, const float* proj,
                        const float* campos, float* view_out, float* proj_out, float* campos_out,
                        cudaStream_t s) {
    pose_compose_kernel<<<1, 32, 0, s>>>(q, t, view, proj, campos, view_out, proj_out, campos_out);
    return cudaPeekAtLastError() == cudaSuccess ? 1 : -1;
}

int launch_pose_chain(const float* q, const float* t, const float* view, const float* dtau, float* dq,
                      float* dt, cudaStream_t s) {
    pose_chain_kernel<<<1, 32, 0, s>>>(q, t, view, dtau, dq, dt);
    return cudaPeekAtLastError() == cudaSuccess ? 1 : -1;
}
