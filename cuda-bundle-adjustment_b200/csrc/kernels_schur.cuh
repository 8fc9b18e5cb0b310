// kernels_schur.cuh -- Schur complement and back-substitution.
//   k_schur_inv     invHll = (Hll + lambda I)^-1                (reference computeBschureKernel, Sym3x3Inv
//   k_schur_rows    Hsc, bsc (by block rows, below)              cuda_block_solver.cu:639-669, 1286-1345)
//   k_backsub       xl = invHll (bl - Hpl^T xp), landmark update (schurComplementPostKernel :1419-1442,
//                                                                  updateLandmarksKernel :1457-1469)
//   k_update_poses  T <- Exp(xp) T                               (updatePosesKernel :1444-1455)
// Damping is applied on the fly (Hpp/Hll stay undamped), so the reference's
// addLambda / restoreDiagonal passes (:1255-1284) and its backup buffers vanish.
// The reference subtracts each 6x6 product into Hsc with 36 atomics; here every
// Hsc block is summed by the warps of one CTA in a fixed order.
#pragma once
#include "device_math.cuh"
#include "engine.h"
#include "kernels_chol_common.cuh"

namespace cugo_b200 {

template <int TPB>
__global__ void __launch_bounds__(TPB) k_schur_inv(Dev d) {
    pdl_trigger(); pdl_wait();   // engine.h: nothing mutable is touched before the predecessor has completed
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

// Back-substitution, landmark-major with the same chunking as k_build_lm.  Work
// item = (Hpl block, landmark coordinate c): a 6-term dot of one contiguous
// 48-byte column of the block with xp of its pose (coalesced over Hpl).  Per
// landmark the segment head sums its items in CSC order, applies invHll, writes
// xl and the updated landmark into the trial buffer, and adds xl (lambda xl + bl)
// to the CTA's partial of the LM gain denominator (reference computeScaleKernel
// :1471-1490).
template <int TPB, typename HT>
__global__ void __launch_bounds__(TPB) k_backsub(Dev d) {
    __shared__ double acc[LM_CHUNK * 3];
    __shared__ double part[TPB * 3];
    __shared__ double sh[TPB / 32];
    pdl_trigger(); pdl_wait();   // engine.h: nothing mutable is touched before the predecessor has completed
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
            const HT* __restrict__ h = HplOf<HT>::get(d) + (size_t)slot * 18 + c * 6;
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
    pdl_trigger(); pdl_wait();   // engine.h: nothing mutable is touched before the predecessor has completed
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

namespace cugo_b200 {

// ---------------------------------------------------------------------------
// Second-generation Schur complement: CTAs work on block ROWS of Hsc.
//   unit     a block row i, or -- for rows with more than SROW_UNIT products -- a run of its blocks
//            (cut at block boundaries, about SROW_UNIT products each; built by k_row_unit_*).
//   phase 1  T_e = Hpl_e invHll for every Hpl block of pose i -> shared memory, so that the row's blocks
//            are read from L2 once per unit instead of once per product; bsc_i = bp_i - sum_e T_e bl.
//            Poses with more blocks than the stage holds are done in several passes.
//   phase 2  per product a lane PAIR forms T_i B_j^T (each lane 3 columns, 54 FMAs), 16 products of one
//            Hsc block per warp round.  Blocks with more than SROW_BIG products are shared by the CTA's
//            warps (partials combined in warp order); the others are handed to warps one at a time through
//            a shared counter.  The 16 pair sums are combined by a fixed reduce-scatter tree (19
//            exchanges), which leaves the 36 totals spread over the lanes for the store.
// All gathers (Hpl_e, invHll, bl in phase 1; Hpl_j in phase 2) are 16/8-byte cp.async copies issued by
// CONSECUTIVE lanes for consecutive pieces of a block, into shared memory: a thread-per-block gather costs
// one L1 tag lookup per lane and instruction (the tag stage, one line per clock, was the bound of both
// k_schur_blocks and the first row kernel), this one about two per 144-byte block.  Phase 2 keeps two
// rounds in flight per warp (metadata prefetched one more round ahead).
// No atomics on data, fixed summation order.
// ---------------------------------------------------------------------------
constexpr int SROW_TPB = 256;
constexpr int SROW_NW = SROW_TPB / 32;
constexpr int SROW_RING = 2 * 16 * 18;         // doubles per warp: two stages of 16 Hpl blocks
constexpr int SROW_SIDE_LD = 14;               // doubles per edge of the phase-1 side stage: invHll 9, bl 3, pad
constexpr int SROW_REGION = (SROW_NW * SROW_RING > SROW_TPB * SROW_SIDE_LD) ? SROW_NW * SROW_RING : SROW_TPB * SROW_SIDE_LD;
constexpr int SROW_MAX_CAP = 464;              // Hpl blocks of the T stage: two CTAs per SM (with the 2 KB of block descriptors)
constexpr int SROW_NDESC = 256;                // Hsc blocks of a unit whose (first product, count) are staged in shared memory
constexpr int SROW_UNIT = 2048;                // least products per work unit (engine: max(this, products / 1200))
constexpr int SROW_BIG = 256;                  // blocks with more products are shared by all warps

__host__ __device__ __forceinline__ int srow_units(int nprod, int unit) { return nprod > unit ? (nprod + unit - 1) / unit : 1; }
__host__ __device__ __forceinline__ size_t srow_smem_bytes(int cap) { return ((size_t)cap * 18 + SROW_REGION) * sizeof(double); }

__device__ __forceinline__ void cp_async16(double* smemDst, const double* gsrc) {
    const unsigned dst = (unsigned)__cvta_generic_to_shared(smemDst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(dst), "l"(gsrc));
}
// asynchronous copy of two consecutive elements of an Hpl block (16 bytes of doubles, 8 bytes of floats)
__device__ __forceinline__ void cp_async_pair(double* smemDst, const double* gsrc) { cp_async16(smemDst, gsrc); }
__device__ __forceinline__ void cp_async_pair(float* smemDst, const float* gsrc) {
    const unsigned dst = (unsigned)__cvta_generic_to_shared(smemDst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;\n" ::"r"(dst), "l"(gsrc));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait_group() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory"); }

// final value of element idx (column-major 6x6) of Hsc block b of row i
__device__ __forceinline__ void srow_store(const Dev& d, const LMState* st, int i, int b, int idx, double s, bool first) {
    double* __restrict__ out = d.HscPart + (size_t)(d.hscSlot ? d.hscSlot[b] : b) * 36;
    if (!first) {            // later pass of a pose whose blocks do not fit the shared-memory stage at once
        out[idx] -= s;
    } else if (d.bColInd[b] == i) {
        const double lam = d.rank == 0 ? st->lambda : 0.0;
        out[idx] = d.Hpp[(size_t)i * 36 + idx] + ((idx % 7 == 0) ? lam : 0.0) - s;
    } else {
        out[idx] = -s;
    }
}

// element (column-major 6x6) whose total srow_segment leaves in out[j] of this lane, -1 = none
__device__ __forceinline__ int srow_elem(int lane, int j) {
    const int k3 = j + ((lane & 16) ? 2 : 0);
    const int k2 = k3 + ((lane & 8) ? 3 : 0);
    const int k1 = k2 + ((lane & 4) ? 5 : 0);
    if (k3 >= 3 || k2 >= 5 || k1 >= 9) return -1;
    return 18 * (lane & 1) + ((lane & 2) ? 9 : 0) + k1;
}

// metadata of round r of the segment [pa, pb): sj[m] = Hpl slot whose piece (lane + 32 m) % 9 this lane
// copies (-1 none), rk = row rank of the product of this lane's pair (-1 none)
__device__ __forceinline__ void srow_meta(const Dev& d, int pa, int pb, int r, int lane, int sj[5], int& rk) {
#pragma unroll
    for (int m = 0; m < 5; m++) {
        const int x = lane + 32 * m, p = pa + 16 * r + x / 9;
        sj[m] = (x < 144 && p < pb) ? d.prodIJ[p].y : -1;
    }
    const int p = pa + 16 * r + (lane >> 1);
    rk = p < pb ? d.prodRank[p] : -1;
}
template <typename HT>
__device__ __forceinline__ void srow_issue(const Dev& d, double* buf, int lane, const int sj[5]) {
#pragma unroll
    for (int m = 0; m < 5; m++) {
        const int x = lane + 32 * m;
        if (sj[m] >= 0) cp_async_pair(reinterpret_cast<HT*>(buf) + 2 * x, HplOf<HT>::get(d) + (size_t)sj[m] * 18 + 2 * (x % 9));
    }
    cp_async_commit();
}

// Sum of T_i(e) B_j(e)^T over the products [pa, pb) whose row rank lies in the stage [c0, c0 + cap).
// Lane pair lane/2 takes every 16th product, lane parity picks columns 0-2 or 3-5 of the 6x6 result.
template <typename HT>
__device__ __forceinline__ void srow_segment(const Dev& d, const double* T, double* ring, int c0, int cap, int pa, int pb,
                                             int lane, double out[2]) {
    const int half = lane & 1, pr = lane >> 1;
    double acc[18];
#pragma unroll
    for (int j = 0; j < 18; j++) acc[j] = 0.0;
    const int nr = (pb - pa + 15) >> 4;
    if (nr > 0) {
        int sjn[5], rkc, rkn;
        srow_meta(d, pa, pb, 0, lane, sjn, rkc);
        srow_issue<HT>(d, ring, lane, sjn);
        srow_meta(d, pa, pb, 1, lane, sjn, rkn);
        for (int r = 0; r < nr; r++) {
            srow_issue<HT>(d, ring + ((r + 1) & 1) * (16 * 18), lane, sjn);      // round r + 1 (nothing past the end)
            int rk2;
            srow_meta(d, pa, pb, r + 2, lane, sjn, rk2);
            cp_async_wait_group<1>();                                         // round r has landed
            __syncwarp();
            const int e = rkc - c0;
            const bool ok = e >= 0 && e < cap;
            if (__any_sync(0xffffffffu, ok)) {
                if (ok) {
                    const HT* B = reinterpret_cast<const HT*>(ring + (r & 1) * (16 * 18)) + pr * 18 + 3 * half;   // (c, k) at 6 k + c
                    double bv[3][3];
#pragma unroll
                    for (int k = 0; k < 3; k++)
#pragma unroll
                        for (int cc = 0; cc < 3; cc++) bv[k][cc] = (double)B[6 * k + cc];
                    const double2* t2 = reinterpret_cast<const double2*>(T + (size_t)e * 18);
#pragma unroll
                    for (int rr = 0; rr < 3; rr++) {
                        const double2 u0 = t2[rr], u1 = t2[rr + 3], u2 = t2[rr + 6];
#pragma unroll
                        for (int cc = 0; cc < 3; cc++) {
                            acc[cc * 6 + 2 * rr] += u0.x * bv[0][cc] + u1.x * bv[1][cc] + u2.x * bv[2][cc];
                            acc[cc * 6 + 2 * rr + 1] += u0.y * bv[0][cc] + u1.y * bv[1][cc] + u2.y * bv[2][cc];
                        }
                    }
                }
            }
            __syncwarp();                     // the stage is overwritten by the copies of round r + 2
            rkc = rkn; rkn = rk2;
        }
    }
    // reduce-scatter over the 16 lane pairs: 18 -> 9 -> 5 -> 3 -> 2 values per lane
    const bool h1 = lane & 2, h2 = lane & 4, h3 = lane & 8, h4 = lane & 16;
    double a1[9], a2[5], a3[3];
#pragma unroll
    for (int j = 0; j < 9; j++) {
        const double keep = h1 ? acc[j + 9] : acc[j], send = h1 ? acc[j] : acc[j + 9];
        a1[j] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
    }
#pragma unroll
    for (int j = 0; j < 5; j++) {
        const double lo = a1[j], hi = (j + 5 < 9) ? a1[(j + 5 < 9) ? j + 5 : 0] : 0.0;
        const double keep = h2 ? hi : lo, send = h2 ? lo : hi;
        a2[j] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
    }
#pragma unroll
    for (int j = 0; j < 3; j++) {
        const double lo = a2[j], hi = (j + 3 < 5) ? a2[(j + 3 < 5) ? j + 3 : 0] : 0.0;
        const double keep = h3 ? hi : lo, send = h3 ? lo : hi;
        a3[j] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
    }
#pragma unroll
    for (int j = 0; j < 2; j++) {
        const double lo = a3[j], hi = (j + 2 < 3) ? a3[(j + 2 < 3) ? j + 2 : 0] : 0.0;
        const double keep = h4 ? hi : lo, send = h4 ? lo : hi;
        out[j] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
    }
}

// The same sum on the FP64 tensor pipe (round 2): the 16 products of a round are ONE product of a 6 x 48 by a 48 x 6 matrix --
// [T_i(e_0) | T_i(e_1) | ...] [B_j(e_0) | B_j(e_1) | ...]^T, K index 3 p + k for product p, landmark dimension k -- i.e. twelve
// DMMA m8n8k4 steps with both operands read straight from shared memory (T stage, copy ring) instead of 16 x 54 FMAs behind ~450
// instructions of operand shuffling.  The accumulator fragment IS the block: lane (g, q) ends with elements (g, 2q), (g, 2q + 1),
// so there is no reduction tree either.  out[] follows srow_elem_mma.
__device__ __forceinline__ int srow_elem_mma(int lane, int j) {
    const int g = lane >> 2, q = lane & 3;
    return (g < 6 && q < 3) ? (2 * q + j) * 6 + g : -1;
}
template <typename HT>
__device__ __forceinline__ void srow_segment_mma(const Dev& d, const double* T, double* ring, int c0, int cap, int pa, int pb,
                                                 int lane, double out[2]) {
    const int g = lane >> 2, q = lane & 3, g6 = g < 6 ? g : 0;
    double acc0 = 0.0, acc1 = 0.0, acd0 = 0.0, acd1 = 0.0;      // two accumulator chains (even / odd steps)
    const int nr = (pb - pa + 15) >> 4;
    if (nr > 0) {
        int sjn[5], rkc, rkn;
        srow_meta(d, pa, pb, 0, lane, sjn, rkc);
        srow_issue<HT>(d, ring, lane, sjn);
        srow_meta(d, pa, pb, 1, lane, sjn, rkn);
        for (int r = 0; r < nr; r++) {
            srow_issue<HT>(d, ring + ((r + 1) & 1) * (16 * 18), lane, sjn);      // round r + 1 (nothing past the end)
            int rk2;
            srow_meta(d, pa, pb, r + 2, lane, sjn, rk2);
            cp_async_wait_group<1>();                                         // round r has landed
            __syncwarp();
            // T offset of this lane pair's product (lanes 2p, 2p + 1 hold product p), -1 if it is not in the stage
            const int e = rkc - c0;
            const int toff = (e >= 0 && e < cap) ? 18 * e : -1;
            const HT* B = reinterpret_cast<const HT*>(ring + (r & 1) * (16 * 18));
            const int np = min(16, pb - pa - 16 * r);
            const int nst = (3 * np + 3) >> 2;
#pragma unroll 4
            for (int s4 = 0; s4 < nst; s4++) {
                const int kap = 4 * s4 + q, p = kap / 3, k = kap - 3 * p;
                const int to = __shfl_sync(0xffffffffu, toff, 2 * p);
                const bool ok = to >= 0 && p < np;
                double a = T[(ok ? to : 0) + 6 * k + g6];
                double b = (double)B[p * 18 + 6 * k + g6];
                if (!ok) { a = 0.0; b = 0.0; }
                if (g >= 6) { a = 0.0; b = 0.0; }
                if (s4 & 1) dmma884(acd0, acd1, a, b);
                else dmma884(acc0, acc1, a, b);
            }
            __syncwarp();                     // the stage is overwritten by the copies of round r + 2
            rkc = rkn; rkn = rk2;
        }
    }
    out[0] = acc0 + acd0;
    out[1] = acc1 + acd1;
}

// blkPass (several passes only): per (block, pass) the range of the block's products whose row rank lies in
// the pass's stage -- products of a block are ordered by landmark, as are the slots of a pose, so the range
// is tight; the per-product rank test stays as the guarantee.
template <typename HT, bool MMA>
__global__ void __launch_bounds__(SROW_TPB, 2) k_schur_rows(Dev d, int cap /* Hpl blocks the stage holds */, int dbg,
                                                            int unitProd, int npass, const int2* __restrict__ blkPass) {
    extern __shared__ __align__(16) double T[];     // [cap][18], block e at T + 18 e, column-major 6x3
    double* const region = T + (size_t)cap * 18;    // phase 1: side stage; phase 2: the warps' copy rings
    __shared__ double part[2][SROW_NW][36];
    __shared__ double red[SROW_NW][6];
    __shared__ int nextBlk;
    __shared__ int2 bdesc[SROW_NDESC];      // (first product, products) of the unit's blocks in this pass: one parallel load instead of a
                                            // dependent global round trip per block and warp
    pdl_trigger(); pdl_wait();   // engine.h: nothing mutable is touched before the predecessor has completed
#ifndef CUGO_B200_DIAG
    dbg = 0;                         // the cycle counters exist in diagnostic builds only (make DIAG=1)
#endif
    const long long tStart = dbg ? clock64() : 0;
    const LMState* st = d.st;
    if (st->done) return;
    const int2 un = d.rowUnits[blockIdx.x];
    const int i = un.x, u = un.y, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int s0 = d.poseSlotPtr[i], deg = d.poseSlotPtr[i + 1] - s0;
    const double* __restrict__ bl = d.b + (size_t)d.nP * 6;
    const int b0 = d.bRowPtr[i], b1 = d.bRowPtr[i + 1];
    const int pBeg = d.prodPtr[b0], pEnd = d.prodPtr[b1];
    // the unit's blocks: those whose first product lies in [pBeg + u q, pBeg + (u + 1) q)
    const int nun = srow_units(pEnd - pBeg, unitProd);
    int ub0 = b0, ub1 = b1;
    if (nun > 1) {
        const int q = (pEnd - pBeg + nun - 1) / nun;
        const int lower = pBeg + u * q;
        int lo = b0, hi = b1;
        while (lo < hi) { const int mid = (lo + hi) >> 1; if (d.prodPtr[mid] < lower) lo = mid + 1; else hi = mid; }
        ub0 = lo;
        if (u + 1 < nun) {
            hi = b1;
            while (lo < hi) { const int mid = (lo + hi) >> 1; if (d.prodPtr[mid] < lower + q) lo = mid + 1; else hi = mid; }
            ub1 = lo;
        }
    }
    if (ub0 >= ub1 && u > 0) return;                // a block larger than the unit size swallowed this unit
    const int e0 = MMA ? srow_elem_mma(lane, 0) : srow_elem(lane, 0), e1 = MMA ? srow_elem_mma(lane, 1) : srow_elem(lane, 1);
    double* const ring = region + wid * SROW_RING;
    double tb[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
    int par = 0;
    long long dbgRounds = 0, dbgBlocks = 0;      // diagnostics: rounds / blocks handled by this warp
    for (int c0 = 0; c0 == 0 || c0 < deg; c0 += cap) {     // one pass unless deg > cap (loop-closure poses)
        const bool first = (c0 == 0);
        // ---- phase 1: T_e = Hpl_e invHll for the pose's blocks [c0, c0 + cap)
        const int nE = min(deg - c0, cap);
        for (int x0 = tid; x0 < nE * 9; x0 += 4 * SROW_TPB) {       // index loads of four copies first
            int sl[4];
#pragma unroll
            for (int v = 0; v < 4; v++) {
                const int x = x0 + v * SROW_TPB;
                sl[v] = x < nE * 9 ? d.poseSlots[s0 + c0 + x / 9] : -1;
            }
#pragma unroll
            for (int v = 0; v < 4; v++) {
                const int x = x0 + v * SROW_TPB;
                if (sl[v] >= 0)      // block e lands at the start of its T slot (floats fill half of it)
                    cp_async_pair(reinterpret_cast<HT*>(T + 18 * (x / 9)) + 2 * (x % 9), HplOf<HT>::get(d) + (size_t)sl[v] * 18 + 2 * (x % 9));
            }
        }
        for (int w0 = 0; w0 < nE; w0 += SROW_TPB) {
            const int nW = min(SROW_TPB, nE - w0);
            int lv[12];                                             // piece x = tid + 256 m: edge x / 12, double x % 12
#pragma unroll
            for (int m = 0; m < 12; m++) {
                const int x = tid + m * SROW_TPB;
                lv[m] = x < nW * 12 ? d.poseLm[s0 + c0 + w0 + x / 12] : -1;
            }
#pragma unroll
            for (int m = 0; m < 12; m++) {
                const int x = tid + m * SROW_TPB, e = x / 12, c = x - 12 * e;
                if (lv[m] >= 0)
                    cp_async8(region + e * SROW_SIDE_LD + c,
                              c < 9 ? d.invHll + (size_t)lv[m] * 9 + c : bl + (size_t)lv[m] * 3 + (c - 9), true);
            }
            cp_async_wait_all();
            __syncthreads();
            if (tid < nW) {
                double2* t2 = reinterpret_cast<double2*>(T + (size_t)(w0 + tid) * 18);
                const double* sd = region + tid * SROW_SIDE_LD;
                double a[18], ih[9];
                if (sizeof(HT) == 8) {
#pragma unroll
                    for (int k = 0; k < 9; k++) { const double2 v = t2[k]; a[2 * k] = v.x; a[2 * k + 1] = v.y; }
                } else {
                    const float2* f2 = reinterpret_cast<const float2*>(t2);
#pragma unroll
                    for (int k = 0; k < 9; k++) { const float2 v = f2[k]; a[2 * k] = v.x; a[2 * k + 1] = v.y; }
                }
#pragma unroll
                for (int k = 0; k < 9; k++) ih[k] = sd[k];
                const double g0 = sd[9], g1 = sd[10], g2 = sd[11];
                double o[18];
#pragma unroll
                for (int r = 0; r < 6; r++) {
                    const double t0 = a[r] * ih[0] + a[r + 6] * ih[1] + a[r + 12] * ih[2];
                    const double t1 = a[r] * ih[3] + a[r + 6] * ih[4] + a[r + 12] * ih[5];
                    const double t2v = a[r] * ih[6] + a[r + 6] * ih[7] + a[r + 12] * ih[8];
                    o[r] = t0; o[r + 6] = t1; o[r + 12] = t2v;
                    tb[r] += t0 * g0 + t1 * g1 + t2v * g2;
                }
#pragma unroll
                for (int k = 0; k < 9; k++) t2[k] = make_double2(o[2 * k], o[2 * k + 1]);
            }
            __syncthreads();        // the side stage is reused; T is complete after the last wave
        }
        if (tid == 0) nextBlk = ub0;
        const int ps = c0 / cap;
        for (int k = tid; k < min(ub1 - ub0, SROW_NDESC); k += SROW_TPB) {
            const int b = ub0 + k;
            int2 dsc;
            if (blkPass) { const int2 rg = blkPass[(size_t)b * npass + ps]; dsc = make_int2(rg.x, max(rg.y - rg.x, 0)); }
            else { const int pb0 = d.prodPtr[b]; dsc = make_int2(pb0, d.prodPtr[b + 1] - pb0); }
            bdesc[k] = dsc;
        }
        __syncthreads();
        if (dbg && tid == 0 && blockIdx.x < 2048 && c0 == 0) g_schurDbg[2048 + blockIdx.x] = clock64() - tStart;   // phase 1 of the first pass
        // ---- phase 2a: the large blocks, all warps on each
        for (int b = ub0; b < ub1; b++) {
            int bBeg, n;
            if (b - ub0 < SROW_NDESC) { const int2 dsc = bdesc[b - ub0]; bBeg = dsc.x; n = dsc.y; }
            else if (blkPass) { const int2 rg = blkPass[(size_t)b * npass + ps]; bBeg = rg.x; n = max(rg.y - rg.x, 0); }
            else { bBeg = d.prodPtr[b]; n = d.prodPtr[b + 1] - bBeg; }
            if (n <= SROW_BIG) continue;
            const int qw = ((n + SROW_NW - 1) / SROW_NW + 15) & ~15;
            const int pa = bBeg + wid * qw;
            dbgBlocks += 1000; dbgRounds += (max(min(pa + qw, bBeg + n) - pa, 0) + 15) >> 4;
            double o[2];
            if (MMA) srow_segment_mma<HT>(d, T, ring, c0, cap, pa, min(pa + qw, bBeg + n), lane, o);
            else srow_segment<HT>(d, T, ring, c0, cap, pa, min(pa + qw, bBeg + n), lane, o);
            if (e0 >= 0) part[par][wid][e0] = o[0];
            if (e1 >= 0) part[par][wid][e1] = o[1];
            __syncthreads();
            if (tid < 36) {
                double v = 0.0;
#pragma unroll
                for (int w = 0; w < SROW_NW; w++) v += part[par][w][tid];
                srow_store(d, st, i, b, tid, v, first);
            }
            par ^= 1;           // the other buffer: its readers are past the barrier above
        }
        // ---- phase 2b: the other blocks, one warp each
        for (;;) {
            int b = 0;
            if (lane == 0) b = atomicAdd(&nextBlk, 1);
            b = __shfl_sync(0xffffffffu, b, 0);
            if (b >= ub1) break;
            int bBeg, n;
            if (b - ub0 < SROW_NDESC) { const int2 dsc = bdesc[b - ub0]; bBeg = dsc.x; n = dsc.y; }
            else if (blkPass) { const int2 rg = blkPass[(size_t)b * npass + ps]; bBeg = rg.x; n = max(rg.y - rg.x, 0); }
            else { bBeg = d.prodPtr[b]; n = d.prodPtr[b + 1] - bBeg; }
            if (n > SROW_BIG || (n == 0 && !first)) continue;
            dbgBlocks++; dbgRounds += (n + 15) >> 4;
            double o[2];
            if (MMA) srow_segment_mma<HT>(d, T, ring, c0, cap, bBeg, bBeg + n, lane, o);
            else srow_segment<HT>(d, T, ring, c0, cap, bBeg, bBeg + n, lane, o);
            if (e0 >= 0) srow_store(d, st, i, b, e0, o[0], first);
            if (e1 >= 0) srow_store(d, st, i, b, e1, o[1], first);
        }
        __syncthreads();            // the stage and the counter are reused by the next pass
    }
    if (u == 0) {
#pragma unroll
        for (int r = 0; r < 6; r++) {
            const double v = warp_sum(tb[r]);
            if (lane == 0) red[wid][r] = v;
        }
        __syncthreads();
        if (tid < 6) {
            double s = 0.0;
            for (int w = 0; w < SROW_NW; w++) s += red[w][tid];
            d.bscPart[(size_t)i * 6 + tid] = d.b[(size_t)i * 6 + tid] - s;
        }
    }
    if (dbg && tid == 0 && blockIdx.x < 2048) { g_schurDbg[blockIdx.x] = clock64() - tStart; g_schurDbg[4096 + blockIdx.x] = dbgRounds; g_schurDbg[6144 + blockIdx.x] = dbgBlocks; }
}

// work units of k_schur_rows: rows in launch order (heavy first), each cut into srow_units(products) units
__global__ void k_row_unit_count(const int* workSorted, int nP, int unitProd, int* cnt) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k > nP) return;
    cnt[k] = k < nP ? srow_units(workSorted[k], unitProd) : 0;
}
// per (block, pass) product range of k_schur_rows
__global__ void k_blk_pass_init(int n, int2* blkPass) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < n) blkPass[k] = make_int2(0x7fffffff, 0);
}
__global__ void k_blk_pass(const int* prodRank, const int* head, const int* blkOf, int n, int cap, int npass, int2* blkPass) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const int rk = prodRank[p];
    if (rk < 0) return;
    int2* rg = blkPass + (size_t)(blkOf[p] + head[p] - 1) * npass + rk / cap;
    atomicMin(&rg->x, p);
    atomicMax(&rg->y, p + 1);
}
__global__ void k_row_unit_fill(const int* rowOrder, const int* unitPtr, int nP, int2* units) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= nP) return;
    const int beg = unitPtr[k], end = unitPtr[k + 1], row = rowOrder[k];
    for (int u = beg; u < end; u++) units[u] = make_int2(row, u - beg);
}

// product count of every block row (sort key of the launch order: heavy rows first)
__global__ void k_row_work(const int* bRowPtr, const int* prodPtr, int nP, int* work, int* ids) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nP) return;
    work[i] = prodPtr[bRowPtr[i + 1]] - prodPtr[bRowPtr[i]];
    ids[i] = i;
}

// pose-major view of the Hpl blocks for k_schur_rows (structure build)
__global__ void k_pm_flags(const int* p_src, const int* flag, int E, int* pmFlag) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k > E) return;
    pmFlag[k] = k < E ? flag[p_src[k]] : 0;
}
__global__ void k_pm_slots(const int* p_src, const int* flag, const int* hplPre, const int* pmPre, const int* poseOfPm,
                           const int* posePtr, int E, int* poseSlots, int* slotRank) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= E) return;
    const int e = p_src[k];
    if (!flag[e]) return;
    const int slot = hplPre[e];
    poseSlots[pmPre[k]] = slot;
    slotRank[slot] = pmPre[k] - pmPre[posePtr[poseOfPm[k]]];
}
__global__ void k_pose_slot_ptr(const int* pmPre, const int* posePtr, int nP, int* poseSlotPtr, int* deg) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p > nP) return;
    poseSlotPtr[p] = pmPre[posePtr[p]];
    if (p < nP) deg[p] = pmPre[posePtr[p + 1]] - pmPre[posePtr[p]];
}
__global__ void k_pose_lm(const int* poseSlots, const int* hplLm, int n, int* poseLm) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < n) poseLm[k] = hplLm[poseSlots[k]];
}
__global__ void k_prod_rank(const int2* prodIJ, const int* slotRank, int n, int* prodRank) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const int s = prodIJ[p].x;
    prodRank[p] = s < 0 ? -1 : slotRank[s];
}

}  // namespace cugo_b200
