// device_math.cuh -- per-edge math of the BA hot path (FP64), written for sm_100a.
// Behaviour follows the reference kernels' formulas (cited per function, paths
// relative to the reference repo); the code layout is ours: everything is
// computed in registers from SoA loads, nothing is staged through errors/Xcs
// buffers as the reference does.
#pragma once
#include <cuda_runtime.h>
#include <math.h>

namespace cugo_b200 {

enum { RK_NONE = 0, RK_CAUCHY = 1, RK_TUKEY = 2, RK_HUBER = 3 };
// edge kinds (the `dim` code of an edge set).  2 / 3 are the reference's MonoEdgeSet / StereoEdgeSet (their measurement
// dimension); 4 is its DepthEdgeSet (include/ba_types.h:171-234: three rows, the third the inverse depth); 5 / 6 its
// pose-only ICP sets PlaneEdgeSet / LineEdgeSet (include/icp_types.h:19-107, one row; kernels_poseonly.cuh).
enum { EDGE_MONO = 2, EDGE_STEREO = 3, EDGE_DEPTH = 4, EDGE_PLANE = 5, EDGE_LINE = 6 };

// Robust kernels rho(x), rho'(x), x = omega |e|^2.
// None/Tukey/Cauchy: src/cuda/cuda_block_solver.cu:972-1027.  Huber (absent from the
// reference) is g2o's RobustKernelHuber.  Plain arguments replace the reference's
// process-global device object (:1030-1053), so each edge set has its own kernel.
__device__ __forceinline__ double rk_rho(int type, double delta, double x) {
    const double d2 = delta * delta;
    if (type == RK_NONE) return x;
    if (type == RK_TUKEY) {
        const double maxv = (1.0 / 3) * d2;
        if (x <= d2) { const double t = 1 - x / d2; return maxv * (1 - t * t * t); }
        return maxv;
    }
    if (type == RK_CAUCHY) return d2 * log((1.0 / d2) * x + 1.0);
    return x <= d2 ? x : 2 * delta * sqrt(x) - d2;   // Huber
}
__device__ __forceinline__ double rk_drho(int type, double delta, double x) {
    const double d2 = delta * delta;
    if (type == RK_NONE) return 1.0;
    if (type == RK_TUKEY) {
        if (x <= d2) { const double t = 1 - x / d2; return t * t; }
        return 0.0;
    }
    if (type == RK_CAUCHY) return 1.0 / ((1.0 / d2) * x + 1.0);
    return x <= d2 ? 1.0 : delta / sqrt(x);           // Huber
}

// Xc = R(q) Xw + t with q = (x,y,z,w):  cuda_block_solver.cu:379-402
__device__ __forceinline__ void world_to_camera(const double q[4], const double t[3],
                                                const double X[3], double Xc[3]) {
    double a0 = q[1] * X[2] - q[2] * X[1];
    double a1 = q[2] * X[0] - q[0] * X[2];
    double a2 = q[0] * X[1] - q[1] * X[0];
    a0 += a0; a1 += a1; a2 += a2;
    const double b0 = q[1] * a2 - q[2] * a1;
    const double b1 = q[2] * a0 - q[0] * a2;
    const double b2 = q[0] * a1 - q[1] * a0;
    Xc[0] = X[0] + q[3] * a0 + b0 + t[0];
    Xc[1] = X[1] + q[3] * a1 + b1 + t[1];
    Xc[2] = X[2] + q[3] * a2 + b2 + t[2];
}

// residual e = proj(Xc) - meas, mono (dim 2) or stereo (dim 3):
// cuda_block_solver.cu:409-424, 1100-1105.  cam = fx fy cx cy bf.
// Depth (dim code 4): proj = (u, v, 1/Z), camProjectDepth :426-432.  The reference's depth kernel takes meas - proj and
// pairs it with the stereo Jacobian (:2341-2343, include/ba_types.h:217), which makes every step an ascent direction
// (SURVEY.md 9.12); here the residual keeps the sign of the other edges and has its own third Jacobian row.
__device__ __forceinline__ void residual(int dim, const double Xc[3], const double cam[5],
                                         double mu, double mv, double mr, double e[3]) {
    const double invZ = 1.0 / Xc[2];
    const double pu = cam[0] * invZ * Xc[0] + cam[2];
    const double pv = cam[1] * invZ * Xc[1] + cam[3];
    e[0] = pu - mu;
    e[1] = pv - mv;
    e[2] = (dim == 3) ? (pu - cam[4] * invZ) - mr : (dim == EDGE_DEPTH ? invZ - mr : 0.0);
}

// rotation matrix of q: cuda_block_solver.cu:449-478.  R[r][c].
__device__ __forceinline__ void quat_to_R(const double q[4], double R[3][3]) {
    const double x = q[0], y = q[1], z = q[2], w = q[3];
    const double tx = 2 * x, ty = 2 * y, tz = 2 * z;
    const double twx = tx * w, twy = ty * w, twz = tz * w;
    const double txx = tx * x, txy = ty * x, txz = tz * x;
    const double tyy = ty * y, tyz = tz * y, tzz = tz * z;
    R[0][0] = 1 - (tyy + tzz); R[0][1] = txy - twz;       R[0][2] = txz + twy;
    R[1][0] = txy + twz;       R[1][1] = 1 - (txx + tzz); R[1][2] = tyz - twx;
    R[2][0] = txz - twy;       R[2][1] = tyz + twx;       R[2][2] = 1 - (txx + tyy);
}

// Pose Jacobian JP[r][c] (dim x 6), J = -d proj / d x, g2o linearizeOplus sign
// convention: mono cuda_block_solver.cu:515-527, stereo :558-577.  Row 2 is zero
// for mono edges so callers can run fixed-trip loops.  Depth edges: rows 0-1 as stereo, row 2 of 1/Z.
__device__ __forceinline__ void jacobian_pose(int dim, const double Xc[3], const double cam[5],
                                              double JP[3][6]) {
    const double X = Xc[0], Y = Xc[1], Z = Xc[2];
    const double fu = cam[0], fv = cam[1];
    const double invZ = 1.0 / Z;
    if (dim == 2) {
        const double x = invZ * X, y = invZ * Y;
        const double fu_invZ = fu * invZ, fv_invZ = fv * invZ;
        JP[0][0] = fu * x * y;       JP[0][1] = -fu * (1 + x * x); JP[0][2] = fu * y;
        JP[0][3] = -fu_invZ;         JP[0][4] = 0;                 JP[0][5] = fu_invZ * x;
        JP[1][0] = fv * (1 + y * y); JP[1][1] = -fv * x * y;       JP[1][2] = -fv * x;
        JP[1][3] = 0;                JP[1][4] = -fv_invZ;          JP[1][5] = fv_invZ * y;
#pragma unroll
        for (int c = 0; c < 6; c++) JP[2][c] = 0;
    } else {
        const double invZZ = invZ * invZ, bf = cam[4];
        JP[0][0] = X * Y * invZZ * fu;       JP[0][1] = -(1 + (X * X * invZZ)) * fu;
        JP[0][2] = Y * invZ * fu;            JP[0][3] = -1 * invZ * fu;
        JP[0][4] = 0;                        JP[0][5] = X * invZZ * fu;
        JP[1][0] = (1 + Y * Y * invZZ) * fv; JP[1][1] = -X * Y * invZZ * fv;
        JP[1][2] = -X * invZ * fv;           JP[1][3] = 0;
        JP[1][4] = -1 * invZ * fv;           JP[1][5] = Y * invZZ * fv;
        if (dim == 3) {
            JP[2][0] = JP[0][0] - bf * Y * invZZ; JP[2][1] = JP[0][1] + bf * X * invZZ;
            JP[2][2] = JP[0][2];                  JP[2][3] = JP[0][3];
            JP[2][4] = 0;                         JP[2][5] = JP[0][5] - bf * invZZ;
        } else {    // depth: -d(1/Z)/d(delta) = (1/Z^2) dZ/d(delta), dZ/d(delta) = (Y, -X, 0, 0, 0, 1)
            JP[2][0] = Y * invZZ; JP[2][1] = -X * invZZ; JP[2][2] = 0;
            JP[2][3] = 0;         JP[2][4] = 0;          JP[2][5] = invZZ;
        }
    }
}

// Landmark Jacobian JL[r][c] (dim x 3): mono :508-513, stereo :546-556.
__device__ __forceinline__ void jacobian_landmark(int dim, const double Xc[3], const double q[4],
                                                  const double cam[5], double JL[3][3]) {
    double R[3][3];
    quat_to_R(q, R);
    const double X = Xc[0], Y = Xc[1], Z = Xc[2];
    const double fu = cam[0], fv = cam[1];
    const double invZ = 1.0 / Z;
    if (dim == 2) {
        const double x = invZ * X, y = invZ * Y;
        const double fu_invZ = fu * invZ, fv_invZ = fv * invZ;
#pragma unroll
        for (int c = 0; c < 3; c++) {
            JL[0][c] = -fu_invZ * (R[0][c] - x * R[2][c]);
            JL[1][c] = -fv_invZ * (R[1][c] - y * R[2][c]);
            JL[2][c] = 0;
        }
    } else {
        const double invZZ = invZ * invZ, bf = cam[4];
#pragma unroll
        for (int c = 0; c < 3; c++) {
            JL[0][c] = -fu * R[0][c] * invZ + fu * X * R[2][c] * invZZ;
            JL[1][c] = -fv * R[1][c] * invZ + fv * Y * R[2][c] * invZZ;
            JL[2][c] = (dim == 3) ? JL[0][c] - bf * R[2][c] * invZZ : R[2][c] * invZZ;   // depth: -d(1/Z)/dXw
        }
    }
}

// inverse of a symmetric 3x3 given its six unique entries (adjugate /
// determinant, same expression tree as cuda_block_solver.cu:639-669)
__device__ __forceinline__ void sym3_inverse(double A00, double A01, double A11, double A02,
                                             double A12, double A22, double B[6]) {
    const double det = A00 * A11 * A22 + A01 * A12 * A02 + A02 * A01 * A12 - A00 * A12 * A12 -
                       A02 * A11 * A02 - A01 * A01 * A22;
    const double invDet = 1 / det;
    B[0] = invDet * (A11 * A22 - A12 * A12);   // B00
    B[1] = invDet * (A02 * A12 - A01 * A22);   // B01
    B[2] = invDet * (A00 * A22 - A02 * A02);   // B11
    B[3] = invDet * (A01 * A12 - A02 * A11);   // B02
    B[4] = invDet * (A02 * A01 - A00 * A12);   // B12
    B[5] = invDet * (A00 * A11 - A01 * A01);   // B22
}

// T <- Exp(delta) * T, delta = (omega, upsilon); quaternion renormalised with w >= 0.
// cuda_block_solver.cu:671-719 (skew, addOmega), :721-754 (R -> q), :781-823.
__device__ __forceinline__ void se3_exp_update(const double d[6], double q[4], double t[3]) {
    const double wx = d[0], wy = d[1], wz = d[2];
    const double theta = sqrt(wx * wx + wy * wy + wz * wz);
    double a1, a2, v1, v2;    // R = I + a1 O1 + a2 O2 ; V = I + v1 O1 + v2 O2
    if (theta < 0.00001) {
        a1 = 1.0; a2 = 0.5; v1 = 0.5; v2 = 1.0 / 6;
    } else {
        const double s = sin(theta), c = cos(theta);
        a1 = s / theta;
        a2 = (1 - c) / (theta * theta);
        v1 = a2;
        v2 = (theta - s) / (theta * theta * theta);
    }
    const double xx = wx * wx, yy = wy * wy, zz = wz * wz, xy = wx * wy, yz = wy * wz, zx = wz * wx;
    double O1[3][3] = {{0, -wz, wy}, {wz, 0, -wx}, {-wy, wx, 0}};
    double O2[3][3] = {{-yy - zz, xy, zx}, {xy, -zz - xx, yz}, {zx, yz, -xx - yy}};
    double R[3][3], V[3][3];
#pragma unroll
    for (int r = 0; r < 3; r++)
#pragma unroll
        for (int c = 0; c < 3; c++) {
            const double id = (r == c) ? 1.0 : 0.0;
            R[r][c] = id + a1 * O1[r][c] + a2 * O2[r][c];
            V[r][c] = id + v1 * O1[r][c] + v2 * O2[r][c];
        }
    double dq[4];
    double tr = R[0][0] + R[1][1] + R[2][2];
    if (tr > 0) {
        double s = sqrt(tr + 1);
        dq[3] = 0.5 * s;
        s = 0.5 / s;
        dq[0] = (R[2][1] - R[1][2]) * s;
        dq[1] = (R[0][2] - R[2][0]) * s;
        dq[2] = (R[1][0] - R[0][1]) * s;
    } else {
        int i = 0;
        if (R[1][1] > R[0][0]) i = 1;
        if (R[2][2] > R[i][i]) i = 2;
        const int j = (i + 1) % 3, k = (j + 1) % 3;
        double s = sqrt(R[i][i] - R[j][j] - R[k][k] + 1);
        dq[i] = 0.5 * s;
        s = 0.5 / s;
        dq[3] = (R[k][j] - R[j][k]) * s;
        dq[j] = (R[j][i] + R[i][j]) * s;
        dq[k] = (R[k][i] + R[i][k]) * s;
    }
    double dt[3];
#pragma unroll
    for (int r = 0; r < 3; r++) dt[r] = V[r][0] * d[3] + V[r][1] * d[4] + V[r][2] * d[5];
    // t <- dt + dq (x) t
    const double zero[3] = {0, 0, 0};
    double u[3];
    world_to_camera(dq, zero, t, u);
    t[0] = dt[0] + u[0]; t[1] = dt[1] + u[1]; t[2] = dt[2] + u[2];
    // q <- normalize(dq * q)
    double r[4];
    r[3] = dq[3] * q[3] - dq[0] * q[0] - dq[1] * q[1] - dq[2] * q[2];
    r[0] = dq[3] * q[0] + dq[0] * q[3] + dq[1] * q[2] - dq[2] * q[1];
    r[1] = dq[3] * q[1] + dq[1] * q[3] + dq[2] * q[0] - dq[0] * q[2];
    r[2] = dq[3] * q[2] + dq[2] * q[3] + dq[0] * q[1] - dq[1] * q[0];
    double invn = 1 / sqrt(r[0] * r[0] + r[1] * r[1] + r[2] * r[2] + r[3] * r[3]);
    if (r[3] < 0) invn = -invn;
    q[0] = invn * r[0]; q[1] = invn * r[1]; q[2] = invn * r[2]; q[3] = invn * r[3];
}

// ---- reductions -------------------------------------------------------------
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_down_sync(0xffffffffu, v, o));
    return v;
}
// sum over a CTA in a fixed order (deterministic); result valid on thread 0
template <int TPB>
__device__ __forceinline__ double block_sum(double v, double* sh /* [TPB/32] */) {
    v = warp_sum(v);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    __syncthreads();
    if (lane == 0) sh[wid] = v;
    __syncthreads();
    double r = 0;
    if (threadIdx.x == 0)
        for (int i = 0; i < TPB / 32; i++) r += sh[i];
    return r;
}
template <int TPB>
__device__ __forceinline__ double block_max(double v, double* sh) {
    v = warp_max(v);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    __syncthreads();
    if (lane == 0) sh[wid] = v;
    __syncthreads();
    double r = 0;
    if (threadIdx.x == 0)
        for (int i = 0; i < TPB / 32; i++) r = fmax(r, sh[i]);
    return r;
}

}  // namespace cugo_b200
