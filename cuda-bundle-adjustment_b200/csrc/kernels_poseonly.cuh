// kernels_poseonly.cuh -- SURVEY.md 8(f) row 4: what the reference does OUTSIDE the Schur path.
//   * pose-only ICP edges, PlaneEdgeSet / LineEdgeSet (include/icp_types.h:19-107; kernels
//     computeActiveErrorsKernel_Plane/_Line src/cuda/cuda_block_solver.cu:2380-2451, constructQuadraticFormKernel_Plane/_Line
//     :2453-2536): one scalar residual per edge, contributions to Hpp / bp only;
//   * graphs without free landmarks (BlockSolver::doSchur_ == false, src/block_solver.cpp:80,226-234,344-354): Hpp is block
//     diagonal there (every edge touches one free pose), the reference expands it to a dense 6P x 6P matrix and calls
//     cusolverDn getrf/getrs (DenseLinearSolver, src/cholesky.hpp:314-516).  Here: one 6x6 damped Cholesky per pose.
// Residuals follow the reference: plane e = n . Pw - d (signedDistance :372-377), line e = |(Pw - a) x (Pw - b)| / length
// (distance :352-370), Pw = R(q) pointP + t (se3MulVec :334-342).  The reference's ICP Jacobians are unfinished
// (computeJacobians_Line multiplies by an uninitialised matrix, :2286-2288; computeJacobians_Plane interleaves the rotation and
// translation columns, :2208-2227; the Line error kernel sums e instead of omega e^2, :2407-2409), so the Jacobians here are
// derived from the residuals and pinned by finite differences, while the residuals themselves are pinned to the reference's
// own output (tests/test_poseonly.py, goldens ref_f4_*), as SURVEY.md 8(f) asks ("fix ... rather than copy").
// Like the BA edges: J = -de/d(delta) for T <- Exp(delta) T, delta = (omega, upsilon); b = J^T w e; H = J^T w J; H x = b.
#pragma once
#include "device_math.cuh"
#include "engine.h"
#include "kernels_edge.cuh"

namespace cugo_b200 {

// residual e of ICP edge k, the transformed point Pw and g = de/dPw
__device__ __forceinline__ double icp_eval(int kind, const double q[4], const double t[3], const double* __restrict__ pay,
                                           size_t nI, int k, double Pw[3], double g[3]) {
    const double* __restrict__ m = pay + k;
    if (kind == EDGE_PLANE) {
        const double n[3] = {__ldg(m), __ldg(m + nI), __ldg(m + 2 * nI)};
        const double dist = __ldg(m + 3 * nI);
        const double p[3] = {__ldg(m + 4 * nI), __ldg(m + 5 * nI), __ldg(m + 6 * nI)};
        world_to_camera(q, t, p, Pw);
        g[0] = n[0]; g[1] = n[1]; g[2] = n[2];
        return n[0] * Pw[0] + n[1] * Pw[1] + n[2] * Pw[2] - dist;
    }
    const double a[3] = {__ldg(m), __ldg(m + nI), __ldg(m + 2 * nI)};
    const double b[3] = {__ldg(m + 3 * nI), __ldg(m + 4 * nI), __ldg(m + 5 * nI)};
    const double invLen = 1.0 / __ldg(m + 6 * nI);
    const double p[3] = {__ldg(m + 7 * nI), __ldg(m + 8 * nI), __ldg(m + 9 * nI)};
    world_to_camera(q, t, p, Pw);
    const double u[3] = {Pw[0] - a[0], Pw[1] - a[1], Pw[2] - a[2]};
    const double v[3] = {Pw[0] - b[0], Pw[1] - b[1], Pw[2] - b[2]};
    const double c[3] = {u[1] * v[2] - u[2] * v[1], u[2] * v[0] - u[0] * v[2], u[0] * v[1] - u[1] * v[0]};
    const double nc = sqrt(c[0] * c[0] + c[1] * c[1] + c[2] * c[2]);
    // d|c| = c^ . ((b - a) x dPw) = dPw . (c^ x (b - a))
    if (nc > 1e-12) {
        const double s = invLen / nc;
        const double mx = b[0] - a[0], my = b[1] - a[1], mz = b[2] - a[2];
        g[0] = s * (c[1] * mz - c[2] * my);
        g[1] = s * (c[2] * mx - c[0] * mz);
        g[2] = s * (c[0] * my - c[1] * mx);
    } else {
        g[0] = g[1] = g[2] = 0.0;        // the point lies on the line: the norm has no derivative there
    }
    return nc * invLen;
}

// chi2 of the ICP edges: partials behind those of k_errors (d.chiPartial[d.errGrid + CTA]), fixed grid-stride order
template <int TPB>
__global__ void __launch_bounds__(TPB) k_icp_errors(Dev d, IcpDev ic, int which, int guarded) {
    __shared__ double sh[TPB / 32];
    pdl_trigger(); pdl_wait();   // engine.h: nothing mutable is touched before the predecessor has completed
    const LMState* st = d.st;
    if (guarded && st->done) return;
    const int buf = which ? 1 - st->cur : st->cur;
    const double* __restrict__ pose = d.pose[buf];
    double sum = 0;
    for (int k = blockIdx.x * TPB + threadIdx.x; k < ic.nI; k += gridDim.x * TPB) {
        const int set = ic.set[k];
        double q[4], t[3], Pw[3], g[3];
        load_pose(pose, __ldg(ic.iP + k), q, t);
        const double e = icp_eval(ic.kind[set], q, t, ic.pay, (size_t)ic.nI, k, Pw, g);
        sum += rk_rho(ic.rkType[set], ic.rkDelta[set], __ldg(ic.w + k) * e * e);
    }
    const double tot = block_sum<TPB>(sum, sh);
    if (threadIdx.x == 0) d.chiPartial[d.errGrid + blockIdx.x] = tot;
}

// J^T w J (21 unique entries) and J^T w e (6) of one slice of a pose's ICP edges -> ic.part[unit]
template <int TPB>
__global__ void __launch_bounds__(TPB, 2) k_icp_build(Dev d, IcpDev ic) {
    __shared__ double red[TPB / 32][27];
    pdl_trigger(); pdl_wait();
    const LMState* st = d.st;
    if (st->done || !st->needBuild) return;
    const int2 un = ic.units[blockIdx.x];
    const int p = un.x;
    double q[4], t[3];
    load_pose(d.pose[st->cur], p, q, t);
    double a[27];
#pragma unroll
    for (int i = 0; i < 27; i++) a[i] = 0.0;
    const int k0 = ic.posePtr[p] + un.y * ICP_SLICE, k1 = min(ic.posePtr[p + 1], k0 + ICP_SLICE);
#pragma unroll 2
    for (int k = k0 + threadIdx.x; k < k1; k += TPB) {
        const int set = ic.set[k];
        double Pw[3], g[3];
        const double e = icp_eval(ic.kind[set], q, t, ic.pay, (size_t)ic.nI, k, Pw, g);
        const double omega = __ldg(ic.w + k);
        const double w = omega * rk_drho(ic.rkType[set], ic.rkDelta[set], omega * e * e);
        const double J[6] = {-(Pw[1] * g[2] - Pw[2] * g[1]), -(Pw[2] * g[0] - Pw[0] * g[2]), -(Pw[0] * g[1] - Pw[1] * g[0]),
                             -g[0], -g[1], -g[2]};
        int n = 0;
#pragma unroll
        for (int c = 0; c < 6; c++)
#pragma unroll
            for (int r = c; r < 6; r++) a[n++] += w * (J[r] * J[c]);
#pragma unroll
        for (int r = 0; r < 6; r++) a[21 + r] += w * (J[r] * e);
    }
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
    for (int i = 0; i < 27; i++) {
        const double s = warp_sum(a[i]);
        if (lane == 0) red[wid][i] = s;
    }
    __syncthreads();
    if (threadIdx.x < 27) {
        double s = 0;
        for (int w2 = 0; w2 < TPB / 32; w2++) s += red[w2][threadIdx.x];
        ic.part[(size_t)blockIdx.x * 27 + threadIdx.x] = s;
    }
}

// Hpp / bp / max diagonal of every free pose += its ICP units (one CTA per pose: a scan matched against a map puts
// hundreds of units on ONE pose).  Warp g sums the units g, g + G, ... in that order, four loads in flight; warp 0 adds the
// G partial sums in warp order -- a fixed order, so runs repeat bit for bit.  init: the graph has no BA edges, so nothing
// was written before -- set instead of add.
template <int TPB>
__global__ void __launch_bounds__(TPB) k_icp_add(Dev d, IcpDev ic, int init) {
    constexpr int G = TPB / 32;
    __shared__ double sh[G][27];
    pdl_trigger(); pdl_wait();
    const LMState* st = d.st;
    if (st->done || !st->needBuild) return;
    const int lane = threadIdx.x & 31, grp = threadIdx.x >> 5;
    const int p = blockIdx.x;
    double s = 0.0;
    if (lane < 27) {
        const int u1 = ic.slicePtr[p + 1];
        int u = ic.slicePtr[p] + grp;
        const double* __restrict__ part = ic.part + lane;
        for (; u + 3 * G < u1; u += 4 * G) {
            const double a0 = __ldcg(part + (size_t)u * 27), a1 = __ldcg(part + (size_t)(u + G) * 27);
            const double a2 = __ldcg(part + (size_t)(u + 2 * G) * 27), a3 = __ldcg(part + (size_t)(u + 3 * G) * 27);
            s += a0; s += a1; s += a2; s += a3;
        }
        for (; u < u1; u += G) s += __ldcg(part + (size_t)u * 27);
        sh[grp][lane] = s;
    }
    __syncthreads();
    if (grp != 0) return;
    s = 0.0;
    if (lane < 27)
#pragma unroll
        for (int g = 0; g < G; g++) s += sh[g][lane];
    double mx = 0.0;
#pragma unroll
    for (int round = 0; round < 2; round++) {
        const int idx = lane + 32 * round;
        const int i = idx < 36 ? idx : 0;
        int r = i % 6, c = i / 6;
        if (r < c) { const int tmp = r; r = c; c = tmp; }
        const double v = __shfl_sync(0xffffffffu, s, c * 6 - c * (c - 1) / 2 + (r - c));
        if (idx < 36) {
            double* h = d.Hpp + (size_t)p * 36 + idx;
            const double hv = (init ? 0.0 : *h) + v;
            *h = hv;
            if (idx % 7 == 0) mx = fmax(mx, hv);
        }
    }
    const double vb = __shfl_sync(0xffffffffu, s, 21 + (lane < 6 ? lane : 0));
    if (lane < 6) {
        double* bb = d.b + (size_t)p * 6 + lane;
        *bb = (init ? 0.0 : *bb) + vb;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if (lane == 0) d.maxDiagP[p] = mx;
}

// Graph without free landmarks: (Hpp_p + lambda I) x_p = b_p for every free pose, one thread each (6x6 Cholesky in
// registers; a pivot <= 1e-14 fails the trial like the zero-pivot test of the sparse path, src/cholesky.hpp:79-92).
template <int TPB>
__global__ void __launch_bounds__(TPB) k_po_solve(Dev d) {
    pdl_trigger(); pdl_wait();
    LMState* st = d.st;
    if (st->done) return;
    const int p = blockIdx.x * TPB + threadIdx.x;
    if (p >= d.nP) return;
    const double lam = st->lambda;
    const double* __restrict__ H = d.Hpp + (size_t)p * 36;
    double L[6][6];
#pragma unroll
    for (int c = 0; c < 6; c++)
#pragma unroll
        for (int r = c; r < 6; r++) L[r][c] = H[c * 6 + r] + (r == c ? lam : 0.0);
    bool bad = false;
#pragma unroll
    for (int j = 0; j < 6; j++) {
        double dj = L[j][j];
#pragma unroll
        for (int k = 0; k < j; k++) dj -= L[j][k] * L[j][k];
        if (!(dj > 1e-14)) { bad = true; dj = 1.0; }
        const double lj = sqrt(dj);
        L[j][j] = lj;
#pragma unroll
        for (int i = j + 1; i < 6; i++) {
            double s = L[i][j];
#pragma unroll
            for (int k = 0; k < j; k++) s -= L[i][k] * L[j][k];
            L[i][j] = s / lj;
        }
    }
    double y[6];
#pragma unroll
    for (int i = 0; i < 6; i++) {
        double s = d.b[(size_t)p * 6 + i];
#pragma unroll
        for (int k = 0; k < i; k++) s -= L[i][k] * y[k];
        y[i] = s / L[i][i];
    }
#pragma unroll
    for (int i = 5; i >= 0; i--) {
        double s = y[i];
#pragma unroll
        for (int k = i + 1; k < 6; k++) s -= L[k][i] * y[k];
        y[i] = s / L[i][i];
    }
#pragma unroll
    for (int i = 0; i < 6; i++) d.x[(size_t)p * 6 + i] = bad ? 0.0 : y[i];
    if (bad) st->fail = 1;
}

// raw ICP arrays of one set (AoS as the caller holds them, `cols` doubles per edge) -> the unsorted column-major payload
__global__ void k_icp_scatter(const double* meas, const double* omega, int n, int cols, int nomega, size_t nI, size_t off,
                              double* pay, double* w) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    for (int c = 0; c < ICP_COLS; c++) pay[(size_t)c * nI + off + i] = c < cols ? meas[(size_t)i * cols + c] : 0.0;
    w[off + i] = nomega > 1 ? omega[i] : omega[0];
}
// sorted by pose: src[k] = raw index of the k-th edge
__global__ void k_icp_gather(const int* src, size_t nI, const uint8_t* rset, const double* rw, const double* rpay,
                             uint8_t* set, double* w, double* pay) {
    const size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= nI) return;
    const size_t j = (size_t)src[k];
    set[k] = rset[j]; w[k] = rw[j];
    for (int c = 0; c < ICP_COLS; c++) pay[(size_t)c * nI + k] = rpay[(size_t)c * nI + j];
}
__global__ void k_icp_slice_count(const int* posePtr, int nP, int* cnt) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p > nP) return;
    const int deg = p < nP ? posePtr[p + 1] - posePtr[p] : 0;
    cnt[p] = p < nP ? max(1, (deg + ICP_SLICE - 1) / ICP_SLICE) : 0;
}

}  // namespace cugo_b200
