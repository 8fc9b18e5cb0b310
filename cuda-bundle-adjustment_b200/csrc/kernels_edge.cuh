// kernels_edge.cuh -- per-edge stages of the LM iteration:
//   k_errors      computeError  (reference computeActiveErrorsKernel + computeChiValueKernel,
//                 src/cuda/cuda_block_solver.cu:1059-1133)
//   k_build_lm    buildSystem, landmark side: Hll, bl, Hpl   (constructQuadraticFormKernel
//   k_build_pose  buildSystem, pose side:     Hpp, bp          :1151-1220)
// The reference issues 54 FP64 global atomics per edge; here edges are stored twice,
// sorted by landmark and by pose, and every block is reduced inside one CTA in a
// fixed order (no global atomics, bit-reproducible).
#pragma once
#include "device_math.cuh"
#include "engine.h"

namespace cugo_b200 {

__device__ __forceinline__ void load_pose(const double* __restrict__ pose, int iP, double q[4], double t[3]) {
    const double2* p = reinterpret_cast<const double2*>(pose + (size_t)iP * POSE_STRIDE);
    const double2 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2), e = __ldg(p + 3);
    q[0] = a.x; q[1] = a.y; q[2] = b.x; q[3] = b.y;
    t[0] = c.x; t[1] = c.y; t[2] = e.x;
}
__device__ __forceinline__ void load_cam(const Dev& d, const SetParams& sp, const int* camIdx, int k, int set,
                                         double cam[5]) {
    if (camIdx) {
        const double* c = d.cams + 5 * (size_t)__ldg(camIdx + k);
#pragma unroll
        for (int i = 0; i < 5; i++) cam[i] = __ldg(c + i);
    } else {
#pragma unroll
        for (int i = 0; i < 5; i++) cam[i] = sp.cam[set][i];
    }
}

// ---------------------------------------------------------------------------
// chi2 = sum over active, non-outlier edges of rho(omega |e|^2).
// which = 0: accepted estimates est[cur]; 1: trial estimates est[1-cur].
// One partial per CTA (fixed grid-stride assignment => deterministic).
// ---------------------------------------------------------------------------
template <int TPB>
__global__ void __launch_bounds__(TPB) k_errors(Dev d, SetParams sp, int which, int guarded) {
    __shared__ double sh[TPB / 32];
    pdl_trigger(); pdl_wait();   // engine.h: nothing mutable is touched before the predecessor has completed
    LMState* st = d.st;
    if (guarded && st->done) return;
    const int buf = which ? 1 - st->cur : st->cur;
    if (blockIdx.x == 0 && threadIdx.x == 0) st->lastEval = buf;
    const double* __restrict__ pose = d.pose[buf];
    const double* __restrict__ lm = d.lm[buf];
    double sum = 0;
    for (int k = blockIdx.x * TPB + threadIdx.x; k < d.E; k += gridDim.x * TPB) {
        if (d.s_out && d.s_out[k]) continue;
        const int iP = __ldg(d.s_iP + k), iL = __ldg(d.s_iL + k);
        const int set = d.s_set[k];
        const int dim = sp.dim[set];
        double q[4], t[3], cam[5], Xc[3], e[3];
        load_pose(pose, iP, q, t);
        const double X[3] = {__ldg(lm + 3 * (size_t)iL), __ldg(lm + 3 * (size_t)iL + 1), __ldg(lm + 3 * (size_t)iL + 2)};
        load_cam(d, sp, d.s_cam, k, set, cam);
        const double2 uv = __ldg(d.s_uv + k);
        const double mr = (dim != 2) ? __ldg(d.s_r + k) : 0.0;
        world_to_camera(q, t, X, Xc);
        residual(dim, Xc, cam, uv.x, uv.y, mr, e);
        const double chi = rk_rho(sp.rkType[set], sp.rkDelta[set],
                                  __ldg(d.s_w + k) * (e[0] * e[0] + e[1] * e[1] + e[2] * e[2]));
        sum += chi;
        if (d.chiEdge) d.chiEdge[d.s_gid[k]] = chi;
    }
    const double tot = block_sum<TPB>(sum, sh);
    if (threadIdx.x == 0) d.chiPartial[blockIdx.x] = tot;
}

// ---------------------------------------------------------------------------
// Landmark-major pass.  CTA c owns landmarks [chunkLm[c], chunkLm[c+1]) and all
// their edges (contiguous in the landmark-sorted arrays).  Per edge: residual,
// JP, JL, weight; Hpl block -> shared staging -> one coalesced store (the blocks
// of consecutive edges are consecutive in the CSC); JL^T JL w and JL^T e w are
// summed per landmark by the segment-head thread in edge order.
// ---------------------------------------------------------------------------
template <int TPB, typename HT>
__device__ __forceinline__ void build_lm_cta(const Dev& d, const SetParams& sp, const int blk) {
    __shared__ double acc[LM_CHUNK * 9];      // per landmark: H00 H10 H11 H20 H21 H22 b0 b1 b2
    __shared__ double val[9 * TPB];           // per edge of the tile, component-major
    __shared__ __align__(16) HT hst[TPB * 18];   // Hpl staging
    __shared__ double shm[TPB / 32];
    const LMState* st = d.st;
    if (st->done || !st->needBuild) return;
    const int l0 = d.chunkLm[blk], l1 = d.chunkLm[blk + 1];
    if (l0 >= l1) return;
    const int nl = l1 - l0;
    const int e0 = d.lmEdgePtr[l0], e1 = d.lmEdgePtr[l1];
    const double* __restrict__ pose = d.pose[st->cur];
    const double* __restrict__ lm = d.lm[st->cur];
    const int tid = threadIdx.x;
    for (int i = tid; i < nl * 9; i += TPB) acc[i] = 0.0;
    __syncthreads();
    for (int base = e0; base < e1; base += TPB) {
        const int k = base + tid;
        const bool in = k < e1;
        int iL = -1;
        double v[9];
#pragma unroll
        for (int i = 0; i < 9; i++) v[i] = 0.0;
        const int pre0 = d.s_hplPre[base];
        if (in) {
            iL = __ldg(d.s_iL + k);
            const bool live = iL < d.nL && !(d.s_out && d.s_out[k]);
            if (live) {
                const int iP = __ldg(d.s_iP + k);
                const int set = d.s_set[k];
                const int dim = sp.dim[set];
                double q[4], t[3], cam[5], Xc[3], e[3], JL[3][3];
                load_pose(pose, iP, q, t);
                const double X[3] = {__ldg(lm + 3 * (size_t)iL), __ldg(lm + 3 * (size_t)iL + 1),
                                     __ldg(lm + 3 * (size_t)iL + 2)};
                load_cam(d, sp, d.s_cam, k, set, cam);
                const double2 uv = __ldg(d.s_uv + k);
                const double mr = (dim != 2) ? __ldg(d.s_r + k) : 0.0;
                world_to_camera(q, t, X, Xc);
                residual(dim, Xc, cam, uv.x, uv.y, mr, e);
                const double omega = __ldg(d.s_w + k);
                const double w = omega * rk_drho(sp.rkType[set], sp.rkDelta[set],
                                                 omega * (e[0] * e[0] + e[1] * e[1] + e[2] * e[2]));
                jacobian_landmark(dim, Xc, q, cam, JL);
                v[0] = w * (JL[0][0] * JL[0][0] + JL[1][0] * JL[1][0] + JL[2][0] * JL[2][0]);
                v[1] = w * (JL[0][1] * JL[0][0] + JL[1][1] * JL[1][0] + JL[2][1] * JL[2][0]);
                v[2] = w * (JL[0][1] * JL[0][1] + JL[1][1] * JL[1][1] + JL[2][1] * JL[2][1]);
                v[3] = w * (JL[0][2] * JL[0][0] + JL[1][2] * JL[1][0] + JL[2][2] * JL[2][0]);
                v[4] = w * (JL[0][2] * JL[0][1] + JL[1][2] * JL[1][1] + JL[2][2] * JL[2][1]);
                v[5] = w * (JL[0][2] * JL[0][2] + JL[1][2] * JL[1][2] + JL[2][2] * JL[2][2]);
                v[6] = w * (JL[0][0] * e[0] + JL[1][0] * e[1] + JL[2][0] * e[2]);
                v[7] = w * (JL[0][1] * e[0] + JL[1][1] * e[1] + JL[2][1] * e[2]);
                v[8] = w * (JL[0][2] * e[0] + JL[1][2] * e[1] + JL[2][2] * e[2]);
                if (iP < d.nP) {
                    // Hpl = JP^T JL w, 6x3 column-major
                    double JP[3][6];
                    jacobian_pose(dim, Xc, cam, JP);
                    HT* h = hst + (size_t)(d.s_hplPre[k] - pre0) * 18;
#pragma unroll
                    for (int c = 0; c < 3; c++)
#pragma unroll
                        for (int r = 0; r < 6; r++)
                            h[c * 6 + r] = (HT)(w * (JP[0][r] * JL[0][c] + JP[1][r] * JL[1][c] + JP[2][r] * JL[2][c]));
                }
            } else if (d.s_out && d.s_hplPre[k + 1] != d.s_hplPre[k]) {
                // an edge flagged on the device after the structure was built keeps its Hpl slot: a zero block
                HT* h = hst + (size_t)(d.s_hplPre[k] - pre0) * 18;
#pragma unroll
                for (int i = 0; i < 18; i++) h[i] = (HT)0;
            }
        }
#pragma unroll
        for (int i = 0; i < 9; i++) val[i * TPB + tid] = v[i];
        __syncthreads();
        // segment heads add their run (edge order) into the landmark accumulator
        if (in && iL < d.nL && (tid == 0 || __ldg(d.s_iL + k - 1) != iL)) {
            double s[9];
#pragma unroll
            for (int i = 0; i < 9; i++) s[i] = val[i * TPB + tid];
            for (int j = tid + 1; j < TPB && base + j < e1 && __ldg(d.s_iL + base + j) == iL; j++)
#pragma unroll
                for (int i = 0; i < 9; i++) s[i] += val[i * TPB + j];
            double* a = acc + (size_t)(iL - l0) * 9;
#pragma unroll
            for (int i = 0; i < 9; i++) a[i] += s[i];
        }
        // store of this tile's Hpl blocks: they are one contiguous run of the CSC array
        const int last = (base + TPB < e1) ? base + TPB : e1;
        const int cnt = (d.s_hplPre[last] - pre0) * 18;
        HT* __restrict__ dst = HplOf<HT>::get(d) + (size_t)pre0 * 18;
        if constexpr (sizeof(HT) == 8) {
            // FP64 blocks (144 bytes each, 16-byte aligned): ONE bulk copy shared -> global by the TMA unit (SASS: UBLKCP)
            // instead of 18 load/store pairs per thread; the staging buffer is free again once the copy has read it
            if (tid == 0 && cnt > 0) {
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");     // the generic-proxy writes above -> async proxy
                asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst),
                             "r"((unsigned)__cvta_generic_to_shared(hst)), "r"((unsigned)cnt * 8u) : "memory");
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            }
        } else {
            for (int i = tid; i < cnt; i += TPB) dst[i] = hst[i];
        }
        __syncthreads();
    }
    if (sizeof(HT) == 8 && tid == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");   // the bulk stores have landed
    // write Hll (full 3x3, column-major) and bl for the free landmarks of the chunk
    const int nlf = (l1 < d.nL ? l1 : d.nL) - l0;
    const int map9[9] = {0, 1, 3, 1, 2, 4, 3, 4, 5};
    double* __restrict__ Hll = d.Hll + (size_t)l0 * 9;
    for (int i = tid; i < nlf * 9; i += TPB) Hll[i] = acc[(i / 9) * 9 + map9[i % 9]];
    double* __restrict__ bl = d.b + (size_t)d.nP * 6 + (size_t)l0 * 3;
    for (int i = tid; i < nlf * 3; i += TPB) bl[i] = acc[(i / 3) * 9 + 6 + i % 3];
    double mx = 0.0;
    for (int i = tid; i < nlf; i += TPB) mx = fmax(mx, fmax(acc[i * 9], fmax(acc[i * 9 + 2], acc[i * 9 + 5])));
    mx = block_max<TPB>(mx, shm);
    if (tid == 0) d.maxDiagL[blk] = mx;
}
template <int TPB, typename HT>
__global__ void __launch_bounds__(TPB) k_build_lm(Dev d, SetParams sp) {
    pdl_trigger(); pdl_wait();   // engine.h: nothing mutable is touched before the predecessor has completed
    build_lm_cta<TPB, HT>(d, sp, blockIdx.x);
}

// ---------------------------------------------------------------------------
// Pose-major pass.  One CTA per free pose; threads stride over the pose's edges
// (pose-sorted copy of the edge arrays), accumulate the 21 unique entries of
// JP^T JP w and the 6 of JP^T e w in registers, then a fixed-order tree.
// ---------------------------------------------------------------------------
template <int TPB>
__device__ __forceinline__ void build_pose_cta(const Dev& d, const SetParams& sp, const int unit) {
    __shared__ double red[TPB / 32][27];
    __shared__ int ticket;
    const LMState* st = d.st;
    if (st->done || !st->needBuild) return;
    const int2 un = d.poseUnits[unit];
    const int p = un.x;
    const int nsl = d.poseSlicePtr[p + 1] - d.poseSlicePtr[p];
    const double* __restrict__ pose = d.pose[st->cur];
    const double* __restrict__ lm = d.lm[st->cur];
    double q[4], t[3];
    load_pose(pose, p, q, t);
    double a[27];
#pragma unroll
    for (int i = 0; i < 27; i++) a[i] = 0.0;
    const int k0 = d.posePtr[p] + un.y * POSE_SLICE, k1 = min(d.posePtr[p + 1], k0 + POSE_SLICE);
    for (int k = k0 + threadIdx.x; k < k1; k += TPB) {
        if (d.p_out && d.p_out[k]) continue;
        const int iL = __ldg(d.p_iL + k);
        const int set = d.p_set[k];
        const int dim = sp.dim[set];
        double cam[5], Xc[3], e[3], JP[3][6];
        const double X[3] = {__ldg(lm + 3 * (size_t)iL), __ldg(lm + 3 * (size_t)iL + 1), __ldg(lm + 3 * (size_t)iL + 2)};
        load_cam(d, sp, d.p_cam, k, set, cam);
        const double2 uv = __ldg(d.p_uv + k);
        const double mr = (dim != 2) ? __ldg(d.p_r + k) : 0.0;
        world_to_camera(q, t, X, Xc);
        residual(dim, Xc, cam, uv.x, uv.y, mr, e);
        const double omega = __ldg(d.p_w + k);
        const double w = omega * rk_drho(sp.rkType[set], sp.rkDelta[set],
                                         omega * (e[0] * e[0] + e[1] * e[1] + e[2] * e[2]));
        jacobian_pose(dim, Xc, cam, JP);
        int n = 0;
#pragma unroll
        for (int c = 0; c < 6; c++)
#pragma unroll
            for (int r = c; r < 6; r++)
                a[n++] += w * (JP[0][r] * JP[0][c] + JP[1][r] * JP[1][c] + JP[2][r] * JP[2][c]);
#pragma unroll
        for (int r = 0; r < 6; r++) a[21 + r] += w * (JP[0][r] * e[0] + JP[1][r] * e[1] + JP[2][r] * e[2]);
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
        red[0][threadIdx.x] = s;
    }
    __syncthreads();
    if (nsl > 1) {       // the last slice of the pose to arrive adds the partial sums in slice order
        if (threadIdx.x < 27) d.posePart[(size_t)unit * 27 + threadIdx.x] = red[0][threadIdx.x];
        __threadfence();
        __syncthreads();
        if (threadIdx.x == 0) ticket = atomicAdd(d.poseTicket + p, 1);
        __syncthreads();
        if (ticket != nsl - 1) return;
        __threadfence();
        if (threadIdx.x < 27) {
            const double* part = d.posePart + (size_t)d.poseSlicePtr[p] * 27 + threadIdx.x;
            double s = 0;
            for (int j = 0; j < nsl; j++) s += __ldcg(part + (size_t)j * 27);
            red[0][threadIdx.x] = s;
        }
        if (threadIdx.x == 0) d.poseTicket[p] = 0;
        __syncthreads();
    }
    if (threadIdx.x < 36) {
        // full symmetric 6x6, column-major; packed index of (r>=c): c*6 - c(c-1)/2 + (r-c)
        int r = threadIdx.x % 6, c = threadIdx.x / 6;
        if (r < c) { const int tmp = r; r = c; c = tmp; }
        d.Hpp[(size_t)p * 36 + threadIdx.x] = red[0][c * 6 - c * (c - 1) / 2 + (r - c)];
    } else if (threadIdx.x < 42) {
        d.b[(size_t)p * 6 + threadIdx.x - 36] = red[0][21 + threadIdx.x - 36];
    } else if (threadIdx.x == 42) {
        double mx = 0;
        for (int c = 0; c < 6; c++) mx = fmax(mx, red[0][c * 6 - c * (c - 1) / 2]);
        d.maxDiagP[p] = mx;
    }
}
template <int TPB>
__global__ void __launch_bounds__(TPB) k_build_pose(Dev d, SetParams sp) {
    pdl_trigger(); pdl_wait();   // engine.h: nothing mutable is touched before the predecessor has completed
    build_pose_cta<TPB>(d, sp, blockIdx.x);
}
// Both passes in one launch (they share inputs and write disjoint outputs): CTAs [0, nPoseUnits) take the
// pose slices, the rest the landmark chunks.
template <int TPB, typename HT>
__global__ void __launch_bounds__(TPB, 5) k_build(Dev d, SetParams sp) {
    pdl_trigger(); pdl_wait();
    if ((int)blockIdx.x < d.nPoseUnits) build_pose_cta<TPB>(d, sp, blockIdx.x);
    else build_lm_cta<TPB, HT>(d, sp, blockIdx.x - d.nPoseUnits);
}

}  // namespace cugo_b200
