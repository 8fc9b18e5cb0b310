// kernels_chol.cuh -- numeric supernodal (multifrontal) Cholesky of the reduced
// camera system and the two triangular solves.  Replaces the reference's
// cusolverSp{D}csrcholFactor / csrcholZeroPivot / csrcholSolve calls
// (src/cholesky.hpp:109-156, 245-283) and the BSR->CSR expansion + value
// permutation that feed them (cuda_block_solver.cu:1492-1501, 2117-2123).
//
// Per level of the assembly tree (leaves first) three kernels run:
//   k_chol_diag    one CTA per supernode: assemble + factor the diagonal block L11
//   k_chol_trsm    one CTA per 64-row slab: assemble rows below, solve X L11^T = A21
//   k_chol_update  one CTA per 32x32 tile of the update matrix:
//                  U_t = sum_children ext(U_c) - L21 L21^T   (FP64 DMMA tiles)
// The right-hand side rides along as one extra row of every front, so the forward
// substitution is part of the factorisation; k_chol_bwd walks the levels back down.
// Children contributions are *pulled* through inverse index maps (no scatter, no
// atomics): results are bit-reproducible.
#pragma once
#include "engine.h"

namespace cugo_b200 {

constexpr int CHOL_MAXW = 16;               // block columns per supernode (panel width <= 96)
constexpr int CHOL_RS = 64;                 // rows per trsm slab
constexpr int CHOL_TS = 32;                 // update tile
constexpr double CHOL_PIVOT_TOL = 1e-14;    // reference zero-pivot tolerance, src/cholesky.hpp:79-92

struct SnInfo {
    int t, s, nt, rt, fb, m, nc, u;
};
__device__ __forceinline__ SnInfo sn_info(const Chol& c, int t) {
    SnInfo i;
    i.t = t;
    i.s = c.snStart[t];
    i.nt = c.snStart[t + 1] - i.s;
    i.rt = c.snRowPtr[t + 1] - c.snRowPtr[t];
    i.fb = i.nt + i.rt;
    i.m = 6 * i.fb + 1;
    i.nc = 6 * i.nt;
    i.u = 6 * i.rt + 1;
    return i;
}
// scalar row of child ch's update matrix landing on scalar row s of the parent's
// front (6*fb block rows, then the rhs row), or -1
__device__ __forceinline__ int child_row(const Chol& c, int ch, int s, int fb) {
    if (s == 6 * fb) return 6 * (c.snRowPtr[ch + 1] - c.snRowPtr[ch]);
    const int k = c.inv[c.invPtr[ch] + s / 6];
    return k < 0 ? -1 : 6 * k + s % 6;
}
// sum over the children of front entry (r, cc), r >= cc (scalar front coordinates)
__device__ __forceinline__ double pull_children(const Chol& c, const SnInfo& n, int r, int cc) {
    double acc = 0.0;
    for (int q = c.childPtr[n.t]; q < c.childPtr[n.t + 1]; q++) {
        const int ch = c.childIdx[q];
        const int ccol = child_row(c, ch, cc, n.fb);
        if (ccol < 0) continue;
        const int crow = child_row(c, ch, r, n.fb);
        if (crow < 0) continue;
        const int uc = 6 * (c.snRowPtr[ch + 1] - c.snRowPtr[ch]) + 1;
        acc += c.U[c.Uptr[ch] + (size_t)ccol * uc + crow];
    }
    return acc;
}

// ---------------------------------------------------------------------------
template <int TPB>
__global__ void __launch_bounds__(TPB) k_chol_diag(Dev d, Chol c, int lev) {
    extern __shared__ double D[];            // nc x nc, column-major
    LMState* st = d.st;
    if (st->done) return;
    const SnInfo n = sn_info(c, c.levelSn[c.levelPtr[lev] + blockIdx.x]);
    const int nc = n.nc, tid = threadIdx.x;
    for (int i = tid; i < nc * nc; i += TPB) D[i] = 0.0;
    __syncthreads();
    // original entries (blocks whose row lies inside the diagonal part)
    for (int k = c.snBlkPtr[n.t] + (tid / 36); k < c.snBlkPtr[n.t + 1]; k += TPB / 36) {
        if (tid >= (TPB / 36) * 36) break;
        const int fr = c.asmRow[k];
        if (fr >= n.nt) continue;
        const int fcRaw = c.asmCol[k];
        const int fc = fcRaw & 0x3fffffff, tr = fcRaw >> 30;
        const int e = tid % 36, rr = e % 6, cc = e / 6;
        const double* __restrict__ B = d.Hsc + (size_t)c.asmBlk[k] * 36;
        D[(6 * fc + cc) * nc + 6 * fr + rr] = tr ? B[rr * 6 + cc] : B[cc * 6 + rr];
    }
    __syncthreads();
    if (c.childPtr[n.t] < c.childPtr[n.t + 1])
        for (int i = tid; i < nc * nc; i += TPB) {
            const int r = i % nc, cc = i / nc;
            if (r >= cc) D[i] += pull_children(c, n, r, cc);
        }
    __syncthreads();
    // right-looking factorisation by 6-column blocks
    for (int jb = 0; jb < n.nt; jb++) {
        const int j0 = 6 * jb, j1 = j0 + 6;
        if (tid == 0) {
#pragma unroll
            for (int k = 0; k < 6; k++) {
                double x = D[(j0 + k) * nc + j0 + k];
                if (!(x > CHOL_PIVOT_TOL)) { st->fail = 1; x = 1.0; }
                const double dd = sqrt(x), inv = 1.0 / dd;
                D[(j0 + k) * nc + j0 + k] = dd;
                for (int r = k + 1; r < 6; r++) D[(j0 + k) * nc + j0 + r] *= inv;
                for (int cc = k + 1; cc < 6; cc++) {
                    const double f = D[(j0 + k) * nc + j0 + cc];
                    for (int r = cc; r < 6; r++) D[(j0 + cc) * nc + j0 + r] -= f * D[(j0 + k) * nc + j0 + r];
                }
            }
        }
        __syncthreads();
        // rows below the diagonal block: x L_D^T = a
        for (int r = j1 + tid; r < nc; r += TPB) {
            double x[6];
#pragma unroll
            for (int k = 0; k < 6; k++) {
                double s = D[(j0 + k) * nc + r];
#pragma unroll
                for (int p = 0; p < k; p++) s -= x[p] * D[(j0 + p) * nc + j0 + k];
                x[k] = s / D[(j0 + k) * nc + j0 + k];
            }
#pragma unroll
            for (int k = 0; k < 6; k++) D[(j0 + k) * nc + r] = x[k];
        }
        __syncthreads();
        // rank-6 update of the trailing lower triangle
        const int w = nc - j1;
        for (int i = tid; i < w * w; i += TPB) {
            const int r = j1 + i % w, cc = j1 + i / w;
            if (r < cc) continue;
            double s = D[cc * nc + r];
#pragma unroll
            for (int k = 0; k < 6; k++) s -= D[(j0 + k) * nc + r] * D[(j0 + k) * nc + cc];
            D[cc * nc + r] = s;
        }
        __syncthreads();
    }
    double* __restrict__ Pn = c.L + c.Lptr[n.t];
    for (int i = tid; i < nc * nc; i += TPB) {
        const int r = i % nc, cc = i / nc;
        if (r >= cc) Pn[(size_t)cc * n.m + r] = D[i];
    }
}

// ---------------------------------------------------------------------------
__device__ __forceinline__ void find_task(const int* ptr, int q0, int q1, int task, int* q, int* local) {
    // ptr is a prefix array over levelSn positions; find q in [q0,q1) with ptr[q] <= task < ptr[q+1]
    int lo = q0, hi = q1;
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (ptr[mid] <= task) lo = mid; else hi = mid;
    }
    *q = lo;
    *local = task - ptr[lo];
}

template <int TPB>
__global__ void __launch_bounds__(TPB) k_chol_trsm(Dev d, Chol c, int lev) {
    extern __shared__ double sm[];
    const LMState* st = d.st;
    if (st->done) return;
    const int q0 = c.levelPtr[lev], q1 = c.levelPtr[lev + 1];
    int q, slab;
    find_task(c.trsmPtr, q0, q1, c.trsmPtr[q0] + blockIdx.x, &q, &slab);
    const SnInfo n = sn_info(c, c.levelSn[q]);
    const int nc = n.nc, tid = threadIdx.x;
    double* L11 = sm;                        // nc x nc
    double* S = sm + nc * nc;                // CHOL_RS x nc, column-major, ld = CHOL_RS
    const int r0 = nc + slab * CHOL_RS;      // first front row of this slab
    const int nr = min(CHOL_RS, n.m - r0);
    double* __restrict__ Pn = c.L + c.Lptr[n.t];
    for (int i = tid; i < nc * nc; i += TPB) {
        const int r = i % nc, cc = i / nc;
        L11[i] = (r >= cc) ? Pn[(size_t)cc * n.m + r] : 0.0;
    }
    for (int i = tid; i < CHOL_RS * nc; i += TPB) S[i] = 0.0;
    __syncthreads();
    for (int k = c.snBlkPtr[n.t] + (tid / 36); k < c.snBlkPtr[n.t + 1]; k += TPB / 36) {
        if (tid >= (TPB / 36) * 36) break;
        const int fr = c.asmRow[k];
        const int rowBase = 6 * fr - r0;
        if (fr < n.nt || rowBase + 5 < 0 || rowBase >= nr) continue;
        const int fcRaw = c.asmCol[k];
        const int fc = fcRaw & 0x3fffffff, tr = fcRaw >> 30;
        const int e = tid % 36, rr = e % 6, cc = e / 6;
        const int lr = rowBase + rr;
        if (lr < 0 || lr >= nr) continue;
        const double* __restrict__ B = d.Hsc + (size_t)c.asmBlk[k] * 36;
        S[(6 * fc + cc) * CHOL_RS + lr] = tr ? B[rr * 6 + cc] : B[cc * 6 + rr];
    }
    __syncthreads();
    // rhs row (last front row): bsc in permuted order
    if (r0 + nr == n.m)
        for (int cc = tid; cc < nc; cc += TPB)
            S[cc * CHOL_RS + nr - 1] = d.bsc[(size_t)c.perm[n.s + cc / 6] * 6 + cc % 6];
    __syncthreads();
    if (c.childPtr[n.t] < c.childPtr[n.t + 1])
        for (int i = tid; i < nr * nc; i += TPB) {
            const int lr = i % nr, cc = i / nr;
            S[cc * CHOL_RS + lr] += pull_children(c, n, r0 + lr, cc);
        }
    __syncthreads();
    for (int jb = 0; jb < n.nt; jb++) {
        const int j0 = 6 * jb, j1 = j0 + 6;
        if (tid < nr) {
            double x[6];
#pragma unroll
            for (int k = 0; k < 6; k++) {
                double s = S[(j0 + k) * CHOL_RS + tid];
#pragma unroll
                for (int p = 0; p < k; p++) s -= x[p] * L11[(j0 + p) * nc + j0 + k];
                x[k] = s / L11[(j0 + k) * nc + j0 + k];
            }
#pragma unroll
            for (int k = 0; k < 6; k++) S[(j0 + k) * CHOL_RS + tid] = x[k];
        }
        __syncthreads();
        const int w = nc - j1;
        for (int i = tid; i < nr * w; i += TPB) {
            const int lr = i % nr, cc = j1 + i / nr;
            double s = S[cc * CHOL_RS + lr];
#pragma unroll
            for (int k = 0; k < 6; k++) s -= S[(j0 + k) * CHOL_RS + lr] * L11[(j0 + k) * nc + cc];
            S[cc * CHOL_RS + lr] = s;
        }
        __syncthreads();
    }
    for (int i = tid; i < nr * nc; i += TPB) {
        const int lr = i % nr, cc = i / nr;
        Pn[(size_t)cc * n.m + r0 + lr] = S[cc * CHOL_RS + lr];
    }
}

// ---------------------------------------------------------------------------
// D = A(8x4, row) * B(4x8, col) + C on the FP64 tensor pipe (DMMA)
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

// One CTA (4 warps) per 32x32 tile of the lower triangle of U_t; warp w owns the
// 16x16 quadrant (w&1, w>>1).  L21 rows of the tile's row range and column range are
// staged in shared memory (row stride 36 doubles => conflict-free fragment loads).
__global__ void __launch_bounds__(128) k_chol_update(Dev d, Chol c, int lev) {
    constexpr int LDS = CHOL_TS + 4;
    constexpr int KC = 48;                    // K chunk staged per pass
    __shared__ double As[KC * LDS];           // [k][row] rows of the tile's row range
    __shared__ double Bs[KC * LDS];           // [k][row] rows of the tile's column range
    const LMState* st = d.st;
    if (st->done) return;
    const int q0 = c.levelPtr[lev], q1 = c.levelPtr[lev + 1];
    int q, tile;
    find_task(c.updPtr, q0, q1, c.updPtr[q0] + blockIdx.x, &q, &tile);
    const SnInfo n = sn_info(c, c.levelSn[q]);
    // tile -> (ti >= tj) in the lower triangle
    int ti = (int)((sqrt(8.0 * tile + 1.0) - 1.0) * 0.5);
    while ((ti + 1) * (ti + 2) / 2 <= tile) ti++;
    while (ti * (ti + 1) / 2 > tile) ti--;
    const int tj = tile - ti * (ti + 1) / 2;
    const int i0 = ti * CHOL_TS, j0 = tj * CHOL_TS;   // in U coordinates
    const int tid = threadIdx.x, nc = n.nc;
    const double* __restrict__ Pn = c.L + c.Lptr[n.t];
    const int lane = tid & 31, wid = tid >> 5;
    const int wi = (wid & 1) * 16, wj = (wid >> 1) * 16;
    double acc[2][2][2];
#pragma unroll
    for (int a = 0; a < 2; a++)
#pragma unroll
        for (int b = 0; b < 2; b++) acc[a][b][0] = acc[a][b][1] = 0.0;
    const int fr = lane >> 2, fk = lane & 3;
    for (int kc = 0; kc < nc; kc += KC) {
        const int kn = min(KC, nc - kc);      // multiple of 6
        __syncthreads();
        for (int i = tid; i < kn * CHOL_TS; i += 128) {
            const int r = i % CHOL_TS, k = i / CHOL_TS;
            As[k * LDS + r] = (i0 + r < n.u) ? Pn[(size_t)(kc + k) * n.m + nc + i0 + r] : 0.0;
            Bs[k * LDS + r] = (j0 + r < n.u) ? Pn[(size_t)(kc + k) * n.m + nc + j0 + r] : 0.0;
        }
        __syncthreads();
        for (int k0 = 0; k0 < kn; k0 += 4) {  // the last step of a chunk may be half empty
            double af[2], bf[2];
            const bool kin = (k0 + fk) < kn;
#pragma unroll
            for (int a = 0; a < 2; a++) af[a] = kin ? As[(k0 + fk) * LDS + wi + 8 * a + fr] : 0.0;
#pragma unroll
            for (int b = 0; b < 2; b++) bf[b] = kin ? Bs[(k0 + fk) * LDS + wj + 8 * b + fr] : 0.0;
#pragma unroll
            for (int a = 0; a < 2; a++)
#pragma unroll
                for (int b = 0; b < 2; b++) dmma884(acc[a][b][0], acc[a][b][1], af[a], bf[b]);
        }
    }
    double* __restrict__ Ut = c.U + c.Uptr[n.t];
    const bool hasChildren = c.childPtr[n.t] < c.childPtr[n.t + 1];
#pragma unroll
    for (int a = 0; a < 2; a++)
#pragma unroll
        for (int b = 0; b < 2; b++)
#pragma unroll
            for (int e = 0; e < 2; e++) {
                const int r = i0 + wi + 8 * a + (lane >> 2);
                const int cc = j0 + wj + 8 * b + 2 * (lane & 3) + e;
                if (r < n.u && cc < n.u && r >= cc) {
                    double v = -acc[a][b][e];
                    if (hasChildren) v += pull_children(c, n, nc + r, nc + cc);
                    Ut[(size_t)cc * n.u + r] = v;
                }
            }
}

// ---------------------------------------------------------------------------
// Backward substitution, one CTA per supernode, levels walked top-down:
//   x_t = L11^-T (y_t - L21^T x_R),  y_t = the factorised rhs row of the panel.
template <int TPB>
__global__ void __launch_bounds__(TPB) k_chol_bwd(Dev d, Chol c, int lev) {
    extern __shared__ double sm[];
    const LMState* st = d.st;
    if (st->done) return;
    const SnInfo n = sn_info(c, c.levelSn[c.levelPtr[lev] + blockIdx.x]);
    const int nc = n.nc, tid = threadIdx.x, nr = 6 * n.rt;
    double* L11 = sm;                 // nc x nc
    double* w = sm + nc * nc;         // nc
    double* xR = w + nc;              // 6 rt
    const double* __restrict__ Pn = c.L + c.Lptr[n.t];
    const int* __restrict__ Rt = c.snRows + c.snRowPtr[n.t];
    for (int i = tid; i < nc * nc; i += TPB) {
        const int r = i % nc, cc = i / nc;
        L11[i] = (r >= cc) ? Pn[(size_t)cc * n.m + r] : 0.0;
    }
    for (int i = tid; i < nr; i += TPB) xR[i] = c.xperm[(size_t)Rt[i / 6] * 6 + i % 6];
    __syncthreads();
    const int lane = tid & 31, wid = tid >> 5;
    for (int cc = wid; cc < nc; cc += TPB / 32) {
        const double* __restrict__ col = Pn + (size_t)cc * n.m + nc;
        double s = 0.0;
        for (int r = lane; r < nr; r += 32) s += col[r] * xR[r];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
        if (lane == 0) w[cc] = Pn[(size_t)cc * n.m + n.m - 1] - s;
    }
    __syncthreads();
    if (wid == 0) {
        for (int j = nc - 1; j >= 0; j--) {
            double s = 0.0;
            for (int r = j + 1 + lane; r < nc; r += 32) s += L11[j * nc + r] * w[r];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
            if (lane == 0) w[j] = (w[j] - s) / L11[j * nc + j];
            __syncwarp();
        }
    }
    __syncthreads();
    for (int i = tid; i < nc; i += TPB) {
        const double v = w[i];
        c.xperm[(size_t)(n.s + i / 6) * 6 + i % 6] = v;
        d.x[(size_t)c.perm[n.s + i / 6] * 6 + i % 6] = v;
    }
}

}  // namespace cugo_b200
