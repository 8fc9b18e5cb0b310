// kernels_schur.cuh -- Schur complement and back-substitution.
//   k_schur_inv     invHll = (Hll + lambda I)^-1                (reference computeBschureKernel, Sym3x3Inv
//   k_schur_blocks  Hsc, bsc                                     cuda_block_solver.cu:639-669, 1286-1345)
//   k_backsub       xl = invHll (bl - Hpl^T xp), landmark update (schurComplementPostKernel :1419-1442,
//                                                                  updateLandmarksKernel :1457-1469)
//   k_update_poses  T <- Exp(xp) T                               (updatePosesKernel :1444-1455)
// Damping is applied on the fly (Hpp/Hll stay undamped), so the reference's
// addLambda / restoreDiagonal passes (:1255-1284) and its backup buffers vanish.
// The reference subtracts each 6x6 product into Hsc with 36 atomics; here every
// Hsc block is owned by one warp that walks the block's product list.
#pragma once
#include "device_math.cuh"
#include "engine.h"

namespace cugo_b200 {

template <int TPB>
__global__ void __launch_bounds__(TPB) k_schur_inv(Dev d) {
    const LMState* st = d.st;
    if (st->done) return;
    const int l = blockIdx.x * TPB + threadIdx.x;
    if (l >= d.nL) return;
    const double lam = st->lambda;
    const double* __restrict__ H = d.Hll + (size_t)l * 9;
    double B[6];
    sym3_inverse(H[0] + lam, H[3], H[4] + lam, H[2], H[7], H[8] + lam, B);
    double* __restrict__ o = d.invHll + (size_t)l * 9;
    o[0] = B[0]; o[1] = B[1]; o[2] = B[3];
    o[3] = B[1]; o[4] = B[2]; o[5] = B[4];
    o[6] = B[3]; o[7] = B[4]; o[8] = B[5];
}

// One warp per upper-triangular Hsc block (i,j).  Products (one per landmark seen
// by both poses) are dealt to lanes round-robin; each lane forms
// T = Hpl_i invHll (6x3) and accumulates T Hpl_j^T (6x6); diagonal blocks also
// accumulate T bl for bsc.  Lane partials are combined by a shuffle tree, so the
// summation order is fixed.  Hsc(i,i) = Hpp_i + lambda I - sum, bsc_i = bp_i - sum.
template <int WARPS>
__global__ void __launch_bounds__(WARPS * 32) k_schur_blocks(Dev d) {
    const LMState* st = d.st;
    if (st->done) return;
    const int blk = blockIdx.x * WARPS + (threadIdx.x >> 5);
    if (blk >= d.nblk) return;
    const int lane = threadIdx.x & 31;
    const int row = d.bRowOf[blk], col = d.bColInd[blk];
    const bool diag = (row == col);
    const double* __restrict__ bl = d.b + (size_t)d.nP * 6;
    double acc[36], accb[6];
#pragma unroll
    for (int i = 0; i < 36; i++) acc[i] = 0.0;
#pragma unroll
    for (int i = 0; i < 6; i++) accb[i] = 0.0;
    const int p0 = d.prodPtr[blk], p1 = d.prodPtr[blk + 1];
    for (int p = p0 + lane; p < p1; p += 32) {
        const int2 ij = d.prodIJ[p];
        if (ij.x < 0) continue;
        const int l = d.hplLm[ij.x];
        const double* __restrict__ A = d.Hpl + (size_t)ij.x * 18;
        const double* __restrict__ iH = d.invHll + (size_t)l * 9;
        double T[18];
        double ih[9];
#pragma unroll
        for (int i = 0; i < 9; i++) ih[i] = iH[i];
#pragma unroll
        for (int r = 0; r < 6; r++) {
            const double a0 = A[r], a1 = A[r + 6], a2 = A[r + 12];
#pragma unroll
            for (int c = 0; c < 3; c++) T[c * 6 + r] = a0 * ih[c * 3] + a1 * ih[c * 3 + 1] + a2 * ih[c * 3 + 2];
        }
        const double* __restrict__ B = d.Hpl + (size_t)ij.y * 18;
#pragma unroll
        for (int c = 0; c < 6; c++) {
            const double b0 = B[c], b1 = B[c + 6], b2 = B[c + 12];
#pragma unroll
            for (int r = 0; r < 6; r++) acc[c * 6 + r] += T[r] * b0 + T[r + 6] * b1 + T[r + 12] * b2;
        }
        if (diag) {
            const double g0 = bl[(size_t)l * 3], g1 = bl[(size_t)l * 3 + 1], g2 = bl[(size_t)l * 3 + 2];
#pragma unroll
            for (int r = 0; r < 6; r++) accb[r] += T[r] * g0 + T[r + 6] * g1 + T[r + 12] * g2;
        }
    }
#pragma unroll
    for (int i = 0; i < 36; i++) acc[i] = warp_sum(acc[i]);
    if (diag) {
#pragma unroll
        for (int i = 0; i < 6; i++) accb[i] = warp_sum(accb[i]);
    }
    if (lane == 0) {
        double* __restrict__ out = d.Hsc + (size_t)(d.hscSlot ? d.hscSlot[blk] : blk) * 36;
        if (diag) {
            // with several ranks Hpp / bp are partial sums over the rank's edges and lambda is added once
            const double lam = d.rank == 0 ? st->lambda : 0.0;
            const double* __restrict__ Hp = d.Hpp + (size_t)row * 36;
#pragma unroll
            for (int i = 0; i < 36; i++) out[i] = Hp[i] + ((i % 7 == 0) ? lam : 0.0) - acc[i];
#pragma unroll
            for (int i = 0; i < 6; i++) d.bsc[(size_t)row * 6 + i] = d.b[(size_t)row * 6 + i] - accb[i];
        } else {
#pragma unroll
            for (int i = 0; i < 36; i++) out[i] = -acc[i];
        }
    }
}

// Back-substitution, landmark-major with the same chunking as k_build_lm.  Work
// item = (Hpl block, landmark coordinate c): a 6-term dot of one contiguous
// 48-byte column of the block with xp of its pose (coalesced over Hpl).  Per
// landmark the segment head sums its items in CSC order, applies invHll, writes
// xl and the updated landmark into the trial buffer, and adds xl (lambda xl + bl)
// to the CTA's partial of the LM gain denominator (reference computeScaleKernel
// :1471-1490).
template <int TPB>
__global__ void __launch_bounds__(TPB) k_backsub(Dev d) {
    __shared__ double acc[LM_CHUNK * 3];
    __shared__ double part[TPB * 3];
    __shared__ double sh[TPB / 32];
    const LMState* st = d.st;
    if (st->done) return;
    const int l0 = d.chunkLm[blockIdx.x];
    const int l1 = min(d.chunkLm[blockIdx.x + 1], d.nL);
    const int tid = threadIdx.x;
    if (l0 >= l1) {
        if (tid == 0) d.scalePartial[blockIdx.x] = 0.0;
        return;
    }
    const int nl = l1 - l0;
    const double* __restrict__ xp = d.x;
    const double* __restrict__ bl = d.b + (size_t)d.nP * 6;
    double* __restrict__ xl = d.x + (size_t)d.nP * 6;
    for (int i = tid; i < nl * 3; i += TPB) acc[i] = 0.0;
    __syncthreads();
    const int s0 = d.colPtr[l0], s1 = d.colPtr[l1];
    for (int base = s0; base < s1; base += TPB) {
        const int nslot = min(TPB, s1 - base);
        for (int it = tid; it < nslot * 3; it += TPB) {
            const int slot = base + it / 3, c = it % 3;
            const double* __restrict__ h = d.Hpl + (size_t)slot * 18 + c * 6;
            const double* __restrict__ xr = xp + (size_t)__ldg(d.rowInd + slot) * 6;
            double s = h[0] * xr[0];
#pragma unroll
            for (int i = 1; i < 6; i++) s += h[i] * xr[i];
            part[it] = s;
        }
        __syncthreads();
        if (tid < nslot) {
            const int slot = base + tid;
            const int l = __ldg(d.hplLm + slot);
            if (tid == 0 || __ldg(d.hplLm + slot - 1) != l) {
                double s[3] = {part[tid * 3], part[tid * 3 + 1], part[tid * 3 + 2]};
                for (int j = tid + 1; j < nslot && __ldg(d.hplLm + base + j) == l; j++) {
                    s[0] += part[j * 3]; s[1] += part[j * 3 + 1]; s[2] += part[j * 3 + 2];
                }
                double* a = acc + (size_t)(l - l0) * 3;
                a[0] += s[0]; a[1] += s[1]; a[2] += s[2];
            }
        }
        __syncthreads();
    }
    const double lam = st->lambda;
    const double* __restrict__ lmc = d.lm[st->cur];
    double* __restrict__ lmn = d.lm[1 - st->cur];
    double sc = 0.0;
    for (int i = tid; i < nl; i += TPB) {
        const size_t l = (size_t)(l0 + i);
        const double g0 = bl[l * 3], g1 = bl[l * 3 + 1], g2 = bl[l * 3 + 2];
        const double c0 = g0 - acc[i * 3], c1 = g1 - acc[i * 3 + 1], c2 = g2 - acc[i * 3 + 2];
        const double* __restrict__ iH = d.invHll + l * 9;
        const double x0 = iH[0] * c0 + iH[3] * c1 + iH[6] * c2;
        const double x1 = iH[1] * c0 + iH[4] * c1 + iH[7] * c2;
        const double x2 = iH[2] * c0 + iH[5] * c1 + iH[8] * c2;
        xl[l * 3] = x0; xl[l * 3 + 1] = x1; xl[l * 3 + 2] = x2;
        lmn[l * 3] = lmc[l * 3] + x0; lmn[l * 3 + 1] = lmc[l * 3 + 1] + x1; lmn[l * 3 + 2] = lmc[l * 3 + 2] + x2;
        sc += x0 * (lam * x0 + g0) + x1 * (lam * x1 + g1) + x2 * (lam * x2 + g2);
    }
    const double tot = block_sum<TPB>(sc, sh);
    if (tid == 0) d.scalePartial[blockIdx.x] = tot;
}

template <int TPB>
__global__ void __launch_bounds__(TPB) k_update_poses(Dev d) {
    const LMState* st = d.st;
    if (st->done) return;
    const int p = blockIdx.x * TPB + threadIdx.x;
    if (p >= d.nP) return;
    const double* __restrict__ src = d.pose[st->cur] + (size_t)p * POSE_STRIDE;
    double* __restrict__ dst = d.pose[1 - st->cur] + (size_t)p * POSE_STRIDE;
    double q[4] = {src[0], src[1], src[2], src[3]};
    double t[3] = {src[4], src[5], src[6]};
    double dx[6];
#pragma unroll
    for (int i = 0; i < 6; i++) dx[i] = d.x[(size_t)p * 6 + i];
    se3_exp_update(dx, q, t);
    dst[0] = q[0]; dst[1] = q[1]; dst[2] = q[2]; dst[3] = q[3];
    dst[4] = t[0]; dst[5] = t[1]; dst[6] = t[2]; dst[7] = 0.0;
}

}  // namespace cugo_b200
