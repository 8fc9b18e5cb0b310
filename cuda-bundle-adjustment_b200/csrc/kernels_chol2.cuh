// kernels_chol2.cuh -- second generation of the numeric multifrontal Cholesky
// (same storage and index maps as kernels_chol.cuh, new kernels).  Replaces the
// reference's cusolverSp csrcholFactor / csrcholSolve (src/cholesky.hpp:109-156,
// 245-283).  Per level of the assembly tree (leaves first) two kernels run:
//
//   k_chol_panel    one CTA per (supernode, slab of <= CHOL_RB block rows of L21).
//                   The CTA assembles the diagonal block D (all slabs, redundantly)
//                   and its slab S into shared memory -- Hsc blocks scattered, the
//                   children's update matrices *pulled* 6x6 block by block through
//                   the inverse index maps -- and factors the stacked matrix [D; S]
//                   left-looking by 6-column blocks:
//                     block column jb  -=  [D;S](:, <jb) * D(jb, <jb)^T    FP64 DMMA m8n8k4
//                     every thread factors the 6x6 diagonal block in registers
//                     (L_D and W = L_D^-1), then one thread per row: x = a W^T
//                   Two barriers per block column; no single-thread section.
//   k_chol_update2  one CTA per 48x48 tile of the update matrix
//                     U_t = sum_children ext(U_c) - L21 L21^T
//                   (DMMA over the whole K = 6 n_t staged in shared memory, children
//                   pulled block-wise in the epilogue).
//
// The right-hand side rides along as the last row of every front, so the forward
// substitution is part of the factorisation.  k_chol_bwd2 walks the levels back down:
//   x_t = L11^-T (y_t - L21^T x_R), blocked by 6 with the stored W blocks.
// No atomics, fixed summation order: bit-reproducible.
#pragma once
#include "device_math.cuh"
#include "kernels_chol.cuh"

namespace cugo_b200 {

constexpr int CHOL_RB = 10;            // block rows of L21 per panel slab (60 rows + rhs row)
constexpr int PANEL_LD = 164;          // smem leading dimension: >= 6 MAXW + 6 RB + 1, == 4 (mod 16)
constexpr int PANEL_TPB = 256;
constexpr int CHOL_TB = 8;             // block rows per update tile (48 rows)
constexpr int UPD_LD = 52;             // == 4 (mod 16): conflict-free DMMA fragment loads
constexpr int UPD_CLD = 50;
constexpr int CHOL_MAXCH = 32;         // children whose descriptors are cached in shared memory
static_assert(PANEL_LD >= 6 * CHOL_MAXW + 6 * CHOL_RB + 1 && PANEL_LD % 16 == 4, "panel leading dimension");
static_assert(PANEL_TPB >= 6 * CHOL_MAXW + 6 * CHOL_RB + 1, "one thread per panel row");

struct ChildRef {
    const int* inv;      // parent front block -> row block of the child's R (or -1)
    const double* U;     // child's update matrix, uc x uc column-major (lower triangle valid)
    int uc;
    int pad;
};
__device__ __forceinline__ ChildRef load_child(const Chol& c, int q) {
    const int ch = c.childIdx[q];
    ChildRef r;
    r.inv = c.inv + c.invPtr[ch];
    r.uc = 6 * (c.snRowPtr[ch + 1] - c.snRowPtr[ch]) + 1;
    r.U = c.U + c.Uptr[ch];
    r.pad = 0;
    return r;
}
// phase timestamps of CTA 0 of each panel launch (diagnostics, enabled by Chol::dbg)
__device__ long long g_cholDbg[64 * 24];
__device__ long long g_cholTrace[96];       // diagnostics: clock of CTA 0 / thread 0 at four points of every factor-loop step of level 2
__device__ long long g_cholTime[64 * 8];    // diagnostics: %globaltimer (ns) of CTA 0 per level: panel (after wait, end), update, backward solve
__device__ __forceinline__ long long global_ns() { long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
#define CHOL_GT(slot) do { if (c.dbg && blockIdx.x == 0 && threadIdx.x == 0 && lev < 64) g_cholTime[lev * 8 + (slot)] = global_ns(); } while (0)
#define CHOL_STAMP(ph) do { if (c.dbg && blockIdx.x == 0 && threadIdx.x == 0 && lev < 64) g_cholDbg[lev * 24 + (ph)] = clock64(); } while (0)
__device__ __forceinline__ int tri(int i, int k) { return i * (i + 1) / 2 + k; }   // packed lower, k <= i

// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(PANEL_TPB) k_chol_panel(Dev d, Chol c, int lev) {
    extern __shared__ double A[];                 // PANEL_LD x nc, column-major
    __shared__ ChildRef chs[CHOL_MAXCH];
    LMState* st = d.st;
    if (st->done) return;
    const int q0 = c.levelPtr[lev], q1 = c.levelPtr[lev + 1];
    int q, slab;
    find_task(c.trsmPtr, q0, q1, c.trsmPtr[q0] + blockIdx.x, &q, &slab);
    const SnInfo n = sn_info(c, c.levelSn[q]);
    const int nc = n.nc, nt = n.nt, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int rb0 = nt + slab * CHOL_RB;          // first front block row of the slab
    const int rb1 = min(rb0 + CHOL_RB, n.fb);
    const bool last = (rb1 == n.fb);              // the last slab carries the rhs row
    const int nS = 6 * (rb1 - rb0) + (last ? 1 : 0);
    const int R = nc + nS;
    const int ch0 = c.childPtr[n.t], nch = c.childPtr[n.t + 1] - ch0;

    CHOL_STAMP(0);
    for (int i = tid; i < nc * PANEL_LD; i += PANEL_TPB) A[i] = 0.0;
    if (tid < nch && tid < CHOL_MAXCH) chs[tid] = load_child(c, ch0 + tid);
    __syncthreads();
    CHOL_STAMP(1);
    // ---- original entries of Hsc
    if (tid < (PANEL_TPB / 36) * 36) {
        const int e = tid % 36, rr = e % 6, cc = e / 6;
        for (int k = c.snBlkPtr[n.t] + tid / 36; k < c.snBlkPtr[n.t + 1]; k += PANEL_TPB / 36) {
            const int fr = c.asmRow[k];
            int lr;
            if (fr < nt) lr = 6 * fr;
            else if (fr >= rb0 && fr < rb1) lr = nc + 6 * (fr - rb0);
            else continue;
            const int fcRaw = c.asmCol[k];
            const int fc = fcRaw & 0x3fffffff, tr = fcRaw >> 30;
            const double* __restrict__ B = d.Hsc + (size_t)c.asmBlk[k] * 36;
            A[(6 * fc + cc) * PANEL_LD + lr + rr] = tr ? B[rr * 6 + cc] : B[cc * 6 + rr];
        }
    }
    if (last)
        for (int cc = tid; cc < nc; cc += PANEL_TPB)
            A[cc * PANEL_LD + R - 1] = d.bsc[(size_t)c.perm[n.s + cc / 6] * 6 + cc % 6];
    __syncthreads();
    CHOL_STAMP(2);
    // ---- children: pull their update matrices block by block
    if (nch > 0) {
        const int nD = nt * nt, nSb = (rb1 - rb0) * nt;
        const int nBlkUnits = (nD + nSb) * 6;
        const int nUnits = nBlkUnits + (last ? nc : 0);
        for (int wu = tid; wu < nUnits; wu += PANEL_TPB) {
            int fr, fc, c6, lr, rmin = 0;
            if (wu < nBlkUnits) {
                const int item = wu / 6;
                c6 = wu % 6;
                if (item < nD) {
                    fr = item / nt; fc = item % nt;
                    if (fc > fr) continue;
                    lr = 6 * fr;
                    if (fr == fc) rmin = c6;
                } else {
                    const int i2 = item - nD;
                    fr = rb0 + i2 / nt; fc = i2 % nt;
                    lr = nc + 6 * (fr - rb0);
                }
            } else {
                const int col = wu - nBlkUnits;
                fc = col / 6; c6 = col % 6; fr = -1; lr = R - 1;
            }
            double acc[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
            for (int i = 0; i < nch; i++) {
                const ChildRef cr = (i < CHOL_MAXCH) ? chs[i] : load_child(c, ch0 + i);
                const int kc = __ldg(cr.inv + fc);
                if (kc < 0) continue;
                if (fr >= 0) {
                    const int kr = __ldg(cr.inv + fr);
                    if (kr < 0) continue;
                    const double* __restrict__ src = cr.U + (size_t)(6 * kc + c6) * cr.uc + 6 * kr;
#pragma unroll
                    for (int rr = 0; rr < 6; rr++)
                        if (rr >= rmin) acc[rr] += src[rr];
                } else {
                    acc[0] += cr.U[(size_t)(6 * kc + c6) * cr.uc + cr.uc - 1];
                }
            }
            double* dst = A + (6 * fc + c6) * PANEL_LD + lr;
            if (fr >= 0) {
#pragma unroll
                for (int rr = 0; rr < 6; rr++)
                    if (rr >= rmin) dst[rr] += acc[rr];
            } else {
                dst[0] += acc[0];
            }
        }
    }
    __syncthreads();
    CHOL_STAMP(3);
    // ---- factorisation, left-looking by 6-column blocks
    const int fr8 = lane >> 2, fk = lane & 3;
    for (int jb = 0; jb < nt; jb++) {
        const int j0 = 6 * jb;
        if (jb > 0) {
            const int nmt = (R - j0 + 7) >> 3;            // 8-row tiles from row j0 down
            const int nmine = (nmt - wid + 7) >> 3;       // tiles wid, wid+8, wid+16 (warp-uniform)
            if (nmine > 0) {
                double acc[3][2][2];
#pragma unroll
                for (int i = 0; i < 3; i++) acc[i][0][0] = acc[i][0][1] = acc[i][1][0] = acc[i][1][1] = 0.0;
                const double* Bp = A + j0 + fr8;
                for (int k0 = 0; k0 < j0; k0 += 8) {
                    const int ka = k0 + fk, kb = k0 + 4 + fk;
                    const bool va = ka < j0, vb = kb < j0;
                    const double b0 = va ? Bp[ka * PANEL_LD] : 0.0;
                    const double b1 = vb ? Bp[kb * PANEL_LD] : 0.0;
#pragma unroll
                    for (int i = 0; i < 3; i++)
                        if (i < nmine) {
                            const double* Ap = A + j0 + 8 * (wid + 8 * i) + fr8;
                            const double a0 = va ? Ap[ka * PANEL_LD] : 0.0;
                            const double a1 = vb ? Ap[kb * PANEL_LD] : 0.0;
                            dmma884(acc[i][0][0], acc[i][0][1], a0, b0);
                            dmma884(acc[i][1][0], acc[i][1][1], a1, b1);
                        }
                }
                const int cl = 2 * fk;
                if (cl < 6) {
#pragma unroll
                    for (int i = 0; i < 3; i++)
                        if (i < nmine) {
                            const int r = j0 + 8 * (wid + 8 * i) + fr8;
                            if (r < R) {
                                A[(j0 + cl) * PANEL_LD + r] -= acc[i][0][0] + acc[i][1][0];
                                A[(j0 + cl + 1) * PANEL_LD + r] -= acc[i][0][1] + acc[i][1][1];
                            }
                        }
                }
            }
            __syncthreads();
        }
        // every thread: L_D = chol(D_jj), W = L_D^-1 (registers, broadcast loads)
        double Ld[21], W[21], invd[6];
#pragma unroll
        for (int i = 0; i < 6; i++)
#pragma unroll
            for (int k = 0; k <= i; k++) Ld[tri(i, k)] = A[(j0 + k) * PANEL_LD + j0 + i];
        bool bad = false;
#pragma unroll
        for (int k = 0; k < 6; k++) {
            double x = Ld[tri(k, k)];
            if (!(x > CHOL_PIVOT_TOL)) { bad = true; x = 1.0; }
            const double inv = rsqrt(x);
            invd[k] = inv;
            Ld[tri(k, k)] = x * inv;
#pragma unroll
            for (int i = k + 1; i < 6; i++) Ld[tri(i, k)] *= inv;
#pragma unroll
            for (int cc = k + 1; cc < 6; cc++)
#pragma unroll
                for (int i = cc; i < 6; i++) Ld[tri(i, cc)] -= Ld[tri(i, k)] * Ld[tri(cc, k)];
        }
#pragma unroll
        for (int k = 0; k < 6; k++) {
            W[tri(k, k)] = invd[k];
#pragma unroll
            for (int i = k + 1; i < 6; i++) {
                double s = 0.0;
#pragma unroll
                for (int p = k; p < i; p++) s += Ld[tri(i, p)] * W[tri(p, k)];
                W[tri(i, k)] = -invd[i] * s;
            }
        }
        const int r = j0 + tid;
        if (tid < 6) {
            // rows of the diagonal block: store L_D
#pragma unroll
            for (int i = 0; i < 6; i++)
                if (tid == i) {
#pragma unroll
                    for (int k = 0; k <= i; k++) A[(j0 + k) * PANEL_LD + r] = Ld[tri(i, k)];
                }
        } else if (r < R) {
            double a[6], x[6];
#pragma unroll
            for (int k = 0; k < 6; k++) a[k] = A[(j0 + k) * PANEL_LD + r];
#pragma unroll
            for (int k = 0; k < 6; k++) {
                double s = a[0] * W[tri(k, 0)];
#pragma unroll
                for (int p = 1; p <= k; p++) s += a[p] * W[tri(k, p)];
                x[k] = s;
            }
#pragma unroll
            for (int k = 0; k < 6; k++) A[(j0 + k) * PANEL_LD + r] = x[k];
        }
        if (slab == 0 && tid == 0) {
            double* __restrict__ Wg = c.Winv + (size_t)(n.s + jb) * 21;
#pragma unroll
            for (int i = 0; i < 21; i++) Wg[i] = W[i];
            if (bad) st->fail = 1;
        }
        __syncthreads();
    }
    CHOL_STAMP(4);
    // ---- write the panel rows this CTA owns
    double* __restrict__ Pn = c.L + c.Lptr[n.t];
    if (slab == 0)
        for (int i = tid; i < nc * nc; i += PANEL_TPB) {
            const int r = i % nc, cc = i / nc;
            if (r >= cc) Pn[(size_t)cc * n.m + r] = A[cc * PANEL_LD + r];
        }
    for (int i = tid; i < nS * nc; i += PANEL_TPB) {
        const int lr = i % nS, cc = i / nS;
        Pn[(size_t)cc * n.m + 6 * rb0 + lr] = A[cc * PANEL_LD + nc + lr];
    }
    CHOL_STAMP(5);
}

// ---------------------------------------------------------------------------
// U_t tile (ti >= tj), 48x48: 4 warps, warp (wr, wc) owns the 24x24 quadrant.
__global__ void __launch_bounds__(128) k_chol_update2(Dev d, Chol c, int lev, int kpadMax) {
    extern __shared__ double sm[];
    __shared__ ChildRef chs[CHOL_MAXCH];
    const LMState* st = d.st;
    if (st->done) return;
    const int q0 = c.levelPtr[lev], q1 = c.levelPtr[lev + 1];
    int q, tile;
    find_task(c.updPtr, q0, q1, c.updPtr[q0] + blockIdx.x, &q, &tile);
    const SnInfo n = sn_info(c, c.levelSn[q]);
    int ti = (int)((sqrtf(8.0f * tile + 1.0f) - 1.0f) * 0.5f);
    while ((ti + 1) * (ti + 2) / 2 <= tile) ti++;
    while (ti * (ti + 1) / 2 > tile) ti--;
    const int tj = tile - ti * (ti + 1) / 2;
    constexpr int TS = 6 * CHOL_TB;
    const int i0 = ti * TS, j0 = tj * TS;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int nc = n.nc, kp = (nc + 7) & ~7;
    double* As = sm;
    double* Bs = (ti == tj) ? sm : sm + kpadMax * UPD_LD;
    double* Cs = sm + 2 * kpadMax * UPD_LD;
    const double* __restrict__ Pn = c.L + c.Lptr[n.t];
    const int ch0 = c.childPtr[n.t], nch = c.childPtr[n.t + 1] - ch0;
    CHOL_STAMP(8);
    if (tid < nch && tid < CHOL_MAXCH) chs[tid] = load_child(c, ch0 + tid);
    for (int i = tid; i < kp * TS; i += 128) {
        const int r = i % TS, k = i / TS;
        const bool kin = k < nc;
        As[k * UPD_LD + r] = (kin && i0 + r < n.u) ? Pn[(size_t)k * n.m + nc + i0 + r] : 0.0;
        if (ti != tj) Bs[k * UPD_LD + r] = (kin && j0 + r < n.u) ? Pn[(size_t)k * n.m + nc + j0 + r] : 0.0;
    }
    __syncthreads();
    CHOL_STAMP(9);
    const int wr = wid & 1, wc = wid >> 1;
    if (!(ti == tj && wc > wr)) {
        double acc[2][3][3][2];
#pragma unroll
        for (int h = 0; h < 2; h++)
#pragma unroll
            for (int x = 0; x < 3; x++)
#pragma unroll
                for (int y = 0; y < 3; y++) acc[h][x][y][0] = acc[h][x][y][1] = 0.0;
        const int fr8 = lane >> 2, fk = lane & 3;
        const double* Aw = As + wr * 24 + fr8;
        const double* Bw = Bs + wc * 24 + fr8;
        for (int k0 = 0; k0 < kp; k0 += 8) {
            double a[2][3], b[2][3];
#pragma unroll
            for (int h = 0; h < 2; h++)
#pragma unroll
                for (int x = 0; x < 3; x++) {
                    a[h][x] = Aw[(k0 + 4 * h + fk) * UPD_LD + 8 * x];
                    b[h][x] = Bw[(k0 + 4 * h + fk) * UPD_LD + 8 * x];
                }
#pragma unroll
            for (int h = 0; h < 2; h++)
#pragma unroll
                for (int x = 0; x < 3; x++)
#pragma unroll
                    for (int y = 0; y < 3; y++) dmma884(acc[h][x][y][0], acc[h][x][y][1], a[h][x], b[h][y]);
        }
#pragma unroll
        for (int x = 0; x < 3; x++)
#pragma unroll
            for (int y = 0; y < 3; y++)
#pragma unroll
                for (int e = 0; e < 2; e++) {
                    const int r = wr * 24 + 8 * x + fr8;
                    const int cc = wc * 24 + 8 * y + 2 * fk + e;
                    Cs[cc * UPD_CLD + r] = -(acc[0][x][y][e] + acc[1][x][y][e]);
                }
    }
    __syncthreads();
    CHOL_STAMP(10);
    // epilogue: add the children's blocks, store the lower triangle of U_t
    double* __restrict__ Ut = c.U + c.Uptr[n.t];
    const int rt = n.rt;
    for (int wu = tid; wu < CHOL_TB * CHOL_TB * 6; wu += 128) {
        const int c6 = wu % 6, item = wu / 6, br = item % CHOL_TB, bc = item / CHOL_TB;
        const int gbr = CHOL_TB * ti + br, gbc = CHOL_TB * tj + bc;
        if (gbr > rt || gbc > gbr || gbc >= rt) continue;   // block row rt = the rhs row; the rhs column is never used
        const bool isRhs = (gbr == rt);
        const int rmin = (gbr == gbc) ? c6 : 0;
        double v[6];
#pragma unroll
        for (int rr = 0; rr < 6; rr++) v[rr] = Cs[(6 * bc + c6) * UPD_CLD + 6 * br + rr];
        const int fc = n.nt + gbc, fr = n.nt + gbr;
        for (int i = 0; i < nch; i++) {
            const ChildRef cr = (i < CHOL_MAXCH) ? chs[i] : load_child(c, ch0 + i);
            const int kc = __ldg(cr.inv + fc);
            if (kc < 0) continue;
            if (!isRhs) {
                const int kr = __ldg(cr.inv + fr);
                if (kr < 0) continue;
                const double* __restrict__ src = cr.U + (size_t)(6 * kc + c6) * cr.uc + 6 * kr;
#pragma unroll
                for (int rr = 0; rr < 6; rr++)
                    if (rr >= rmin) v[rr] += src[rr];
            } else {
                v[0] += cr.U[(size_t)(6 * kc + c6) * cr.uc + cr.uc - 1];
            }
        }
        double* __restrict__ dst = Ut + (size_t)(6 * gbc + c6) * n.u + 6 * gbr;
        if (!isRhs) {
#pragma unroll
            for (int rr = 0; rr < 6; rr++)
                if (rr >= rmin) dst[rr] = v[rr];
        } else {
            dst[0] = v[0];
        }
    }
    CHOL_STAMP(11);
}

// ---------------------------------------------------------------------------
// Backward substitution, one CTA per supernode, levels walked top-down.
template <int TPB>
__global__ void __launch_bounds__(TPB) k_chol_bwd2(Dev d, Chol c, int lev) {
    extern __shared__ double sm[];
    const LMState* st = d.st;
    if (st->done) return;
    const SnInfo n = sn_info(c, c.levelSn[c.levelPtr[lev] + blockIdx.x]);
    const int nc = n.nc, nt = n.nt, tid = threadIdx.x, nr = 6 * n.rt, ldL = nc + 1;
    double* L11 = sm;                   // nc columns, leading dimension ldL (odd: conflict-free row reads)
    double* w = L11 + nc * ldL;         // nc
    double* Wsm = w + nc;               // 21 nt
    double* xR = Wsm + 21 * nt;         // 6 rt
    const double* __restrict__ Pn = c.L + c.Lptr[n.t];
    const int* __restrict__ Rt = c.snRows + c.snRowPtr[n.t];
    CHOL_STAMP(16);
    for (int i = tid; i < nr; i += TPB) xR[i] = c.xperm[(size_t)Rt[i / 6] * 6 + i % 6];
    for (int i = tid; i < nc * nc; i += TPB) {
        const int r = i % nc, cc = i / nc;
        if (r > cc) L11[cc * ldL + r] = Pn[(size_t)cc * n.m + r];
    }
    for (int i = tid; i < 21 * nt; i += TPB) Wsm[i] = c.Winv[(size_t)n.s * 21 + i];
    __syncthreads();
    CHOL_STAMP(17);
    constexpr int NW = TPB / 32;
    constexpr int CPW = (6 * CHOL_MAXW + NW - 1) / NW;      // columns per warp
    const int lane = tid & 31, wid = tid >> 5;
    {
        double acc[CPW];
#pragma unroll
        for (int j = 0; j < CPW; j++) acc[j] = 0.0;
        for (int r = lane; r < nr; r += 32) {
            const double xv = xR[r];
            const double* __restrict__ p = Pn + nc + r;
#pragma unroll
            for (int j = 0; j < CPW; j++) {
                const int cc = wid + NW * j;
                if (cc < nc) acc[j] += p[(size_t)cc * n.m] * xv;
            }
        }
#pragma unroll
        for (int j = 0; j < CPW; j++) {
            const int cc = wid + NW * j;
            const double s = warp_sum(acc[j]);
            if (lane == 0 && cc < nc) w[cc] = Pn[(size_t)cc * n.m + n.m - 1] - s;
        }
    }
    __syncthreads();
    CHOL_STAMP(18);
    if (wid == 0) {
        for (int jb = nt - 1; jb >= 0; jb--) {
            const int j0 = 6 * jb;
            double xk = 0.0;
            if (lane < 6) {
                const double* Wb = Wsm + 21 * jb;
#pragma unroll
                for (int p = 0; p < 6; p++)
                    if (p >= lane) xk += Wb[tri(p, lane)] * w[j0 + p];
            }
            __syncwarp();
            if (lane < 6) w[j0 + lane] = xk;
            __syncwarp();
            for (int cc = lane; cc < j0; cc += 32) {
                double s = w[cc];
#pragma unroll
                for (int k = 0; k < 6; k++) s -= L11[cc * ldL + j0 + k] * w[j0 + k];
                w[cc] = s;
            }
            __syncwarp();
        }
    }
    __syncthreads();
    CHOL_STAMP(19);
    for (int i = tid; i < nc; i += TPB) {
        const double v = w[i];
        c.xperm[(size_t)(n.s + i / 6) * 6 + i % 6] = v;
        d.x[(size_t)c.perm[n.s + i / 6] * 6 + i % 6] = v;
    }
}

}  // namespace cugo_b200
