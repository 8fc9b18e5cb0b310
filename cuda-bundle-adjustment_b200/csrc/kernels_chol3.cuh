// kernels_chol3.cuh -- third-generation panel kernel of the multifrontal Cholesky.
// Same task decomposition and storage as k_chol_panel (kernels_chol2.cuh): one CTA per
// (supernode, slab of <= CHOL_RB block rows of L21), [D; S] assembled and factored in
// shared memory.  What changed, driven by the phase timestamps of the second generation
// (profiles/r01b_*): per block column the 6x6 diagonal factor was computed redundantly by
// all 256 threads (FP64-issue bound, ~3000 cycles) after a DMMA update that sat on the
// critical path.  Here
//   * 512 threads; warp 0 alone factors the diagonal block (L_D, W = L_D^-1 -> shared
//     memory) while warps 1..15 already apply all *finished* columns (< j0) to the NEXT
//     block column (look-ahead DMMA update, off the critical path);
//   * one thread per row then applies x = a W^T, and a rank-6 DMMA update brings the next
//     block column up to date -- three barriers per block column;
//   * assembly: Hsc blocks scattered by (entry, column) work units, children pulled in
//     groups of four with all loads issued before the (fixed-order) accumulation.
#pragma once
#include "kernels_chol2.cuh"

namespace cugo_b200 {

constexpr int PANEL3_TPB = 512;
constexpr int PANEL3_NW = PANEL3_TPB / 32;

// pull the children's contributions to 6 consecutive rows (or the rhs element) of one front column
__device__ __forceinline__ void pull_unit(const Chol& c, const ChildRef* chs, int ch0, int nch, int fr, int fc, int c6,
                                          int rmin, double acc[6]) {
    for (int i0 = 0; i0 < nch; i0 += 4) {
        const double* src[4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
            src[j] = nullptr;
            const int i = i0 + j;
            if (i < nch) {
                const ChildRef cr = (i < CHOL_MAXCH) ? chs[i] : load_child(c, ch0 + i);
                const int kc = __ldg(cr.inv + fc);
                if (kc >= 0) {
                    if (fr >= 0) {
                        const int kr = __ldg(cr.inv + fr);
                        if (kr >= 0) src[j] = cr.U + (size_t)(6 * kc + c6) * cr.uc + 6 * kr;
                    } else {
                        src[j] = cr.U + (size_t)(6 * kc + c6) * cr.uc + cr.uc - 1;
                    }
                }
            }
        }
        double v[4][6];
#pragma unroll
        for (int j = 0; j < 4; j++)
#pragma unroll
            for (int rr = 0; rr < 6; rr++)
                v[j][rr] = (src[j] && rr >= rmin && (fr >= 0 || rr == 0)) ? src[j][rr] : 0.0;
#pragma unroll
        for (int j = 0; j < 4; j++)
#pragma unroll
            for (int rr = 0; rr < 6; rr++) acc[rr] += v[j][rr];
    }
}

__global__ void __launch_bounds__(PANEL3_TPB) k_chol_panel3(Dev d, Chol c, int lev) {
    extern __shared__ double A[];                 // PANEL_LD x nc, column-major
    __shared__ ChildRef chs[CHOL_MAXCH];
    __shared__ double Wall[21 * CHOL_MAXW];   // W = L_D^-1 of every block column (packed lower)
    __shared__ int badFlag;
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
    if (tid == 0) badFlag = 0;
    for (int i = tid; i < nc * PANEL_LD; i += PANEL3_TPB) A[i] = 0.0;
    if (tid < nch && tid < CHOL_MAXCH) chs[tid] = load_child(c, ch0 + tid);
    __syncthreads();
    CHOL_STAMP(1);
    // ---- original entries of Hsc: work unit = (entry, block column c6)
    {
        const int k0 = c.snBlkPtr[n.t], nent = c.snBlkPtr[n.t + 1] - k0;
        for (int u = tid; u < nent * 6; u += PANEL3_TPB) {
            const int k = k0 + u / 6, c6 = u % 6;
            const int fr = __ldg(c.asmRow + k);
            int lr;
            if (fr < nt) lr = 6 * fr;
            else if (fr >= rb0 && fr < rb1) lr = nc + 6 * (fr - rb0);
            else continue;
            const int fcRaw = __ldg(c.asmCol + k);
            const int fc = fcRaw & 0x3fffffff, tr = fcRaw >> 30;
            const double* __restrict__ B = d.Hsc + (size_t)__ldg(c.asmBlk + k) * 36;
            double v[6];
#pragma unroll
            for (int rr = 0; rr < 6; rr++) v[rr] = tr ? B[rr * 6 + c6] : B[c6 * 6 + rr];
            double* dst = A + (6 * fc + c6) * PANEL_LD + lr;
#pragma unroll
            for (int rr = 0; rr < 6; rr++) dst[rr] = v[rr];
        }
    }
    if (last)
        for (int cc = tid; cc < nc; cc += PANEL3_TPB)
            A[cc * PANEL_LD + R - 1] = d.bsc[(size_t)c.perm[n.s + cc / 6] * 6 + cc % 6];
    __syncthreads();
    CHOL_STAMP(2);
    // ---- children: pull their update matrices block by block
    if (nch > 0) {
        const int nD = nt * nt, nSb = (rb1 - rb0) * nt;
        const int nBlkUnits = (nD + nSb) * 6;
        const int nUnits = nBlkUnits + (last ? nc : 0);
        for (int wu = tid; wu < nUnits; wu += PANEL3_TPB) {
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
            pull_unit(c, chs, ch0, nch, fr, fc, c6, rmin, acc);
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
    // ---- factorisation by 6-column blocks with one block column of look-ahead
    const int fr8 = lane >> 2, fk = lane & 3;
    const bool dbgT = c.dbg && blockIdx.x == 0 && tid == 0 && lev < 64;
    long long tF = 0, tW = 0, tS = 0, tU = 0, t0 = 0, t1 = 0;
    for (int jb = 0; jb < nt; jb++) {
        const int j0 = 6 * jb, j1 = j0 + 6;
        const bool more = (jb + 1 < nt);
        if (dbgT) t0 = clock64();
        // phase F: warp 0 factors D_jj; the others apply the finished columns [0, j0) to block column jb+1
        if (wid == 0) {
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
            if (lane == 0) {
#pragma unroll
                for (int i = 0; i < 6; i++)
#pragma unroll
                    for (int k = 0; k <= i; k++) A[(j0 + k) * PANEL_LD + j0 + i] = Ld[tri(i, k)];
#pragma unroll
                for (int i = 0; i < 21; i++) Wall[21 * jb + i] = W[i];
                if (bad) badFlag = 1;
            }
        } else if (more && j0 > 0 && (wid & 3) != 0) {
            // warps 4, 8, 12 share warp 0's scheduler / FP64 pipe: they stay idle so that the DMMAs of the
            // look-ahead never delay the dependent DFMA chain of the diagonal factor
            const int nmt = (R - j1 + 7) >> 3;          // <= 19 tiles of 8 rows: at most two per worker warp
            constexpr int NLA = 3 * (PANEL3_NW / 4);    // 12 look-ahead warps
            const int ta = (wid >> 2) * 3 + (wid & 3) - 1, tb = ta + NLA;
            if (ta < nmt) {
                const bool two = tb < nmt;
                const double* Bp = A + j1 + fr8;
                const double* Ap = A + j1 + 8 * ta + fr8;
                const double* Aq = A + j1 + 8 * (two ? tb : ta) + fr8;
                double c00 = 0.0, c01 = 0.0, c10 = 0.0, c11 = 0.0, e00 = 0.0, e01 = 0.0, e10 = 0.0, e11 = 0.0;
#pragma unroll 2
                for (int k0 = 0; k0 < j0; k0 += 8) {
                    const int ka = k0 + fk, kb = k0 + 4 + fk;
                    const bool va = ka < j0, vb = kb < j0;
                    const double b0 = va ? Bp[ka * PANEL_LD] : 0.0;
                    const double b1 = vb ? Bp[kb * PANEL_LD] : 0.0;
                    const double a0 = va ? Ap[ka * PANEL_LD] : 0.0;
                    const double a1 = vb ? Ap[kb * PANEL_LD] : 0.0;
                    dmma884(c00, c01, a0, b0);
                    dmma884(c10, c11, a1, b1);
                    if (two) {
                        const double q0 = va ? Aq[ka * PANEL_LD] : 0.0;
                        const double q1 = vb ? Aq[kb * PANEL_LD] : 0.0;
                        dmma884(e00, e01, q0, b0);
                        dmma884(e10, e11, q1, b1);
                    }
                }
                const int cl = 2 * fk;
                if (cl < 6) {
                    const int r = j1 + 8 * ta + fr8;
                    if (r < R) {
                        A[(j1 + cl) * PANEL_LD + r] -= c00 + c10;
                        A[(j1 + cl + 1) * PANEL_LD + r] -= c01 + c11;
                    }
                    const int r2 = j1 + 8 * tb + fr8;
                    if (two && r2 < R) {
                        A[(j1 + cl) * PANEL_LD + r2] -= e00 + e10;
                        A[(j1 + cl + 1) * PANEL_LD + r2] -= e01 + e11;
                    }
                }
            }
        }
        if (dbgT) { t1 = clock64(); tF += t1 - t0; }
        __syncthreads();
        if (dbgT) { t0 = clock64(); tW += t0 - t1; }
        // phase S: one thread per row below the diagonal block, x = a W^T
        {
            const int r = j1 + tid;
            if (r < R) {
                double W[21], a[6], x[6];
#pragma unroll
                for (int i = 0; i < 21; i++) W[i] = Wall[21 * jb + i];
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
        }
        __syncthreads();
        if (dbgT) { t1 = clock64(); tS += t1 - t0; }
        // phase U: rank-6 update of block column jb+1 with the block column just finished
        if (more) {
            const int nmt = (R - j1 + 7) >> 3;
            const double* Bp = A + j1 + fr8;
            for (int t = wid; t < nmt; t += PANEL3_NW) {
                const double* Ap = A + j1 + 8 * t + fr8;
                const double a0 = Ap[(j0 + fk) * PANEL_LD], b0 = Bp[(j0 + fk) * PANEL_LD];
                const double a1 = fk < 2 ? Ap[(j0 + 4 + fk) * PANEL_LD] : 0.0;
                const double b1 = fk < 2 ? Bp[(j0 + 4 + fk) * PANEL_LD] : 0.0;
                double c0 = 0.0, c1 = 0.0;
                dmma884(c0, c1, a0, b0);
                dmma884(c0, c1, a1, b1);
                const int r = j1 + 8 * t + fr8, cl = 2 * fk;
                if (cl < 6 && r < R) {
                    A[(j1 + cl) * PANEL_LD + r] -= c0;
                    A[(j1 + cl + 1) * PANEL_LD + r] -= c1;
                }
            }
            __syncthreads();
        }
        if (dbgT) tU += clock64() - t1;
    }
    if (dbgT) {
        g_cholDbg[lev * 24 + 12] = tF; g_cholDbg[lev * 24 + 13] = tW;
        g_cholDbg[lev * 24 + 14] = tS; g_cholDbg[lev * 24 + 15] = tU;
    }
    CHOL_STAMP(4);
    // ---- write the panel rows this CTA owns (a warp per column, lanes along the rows)
    double* __restrict__ Pn = c.L + c.Lptr[n.t];
    if (slab == 0) {
        for (int i = tid; i < 21 * nt; i += PANEL3_TPB) c.Winv[(size_t)n.s * 21 + i] = Wall[i];
        if (tid == 0 && badFlag) st->fail = 1;
    }
    for (int cc = wid; cc < nc; cc += PANEL3_NW) {
        double* __restrict__ dst = Pn + (size_t)cc * n.m;
        const double* src = A + cc * PANEL_LD;
        if (slab == 0)
            for (int r = cc + lane; r < nc; r += 32) dst[r] = src[r];
        for (int lr = lane; lr < nS; lr += 32) dst[6 * rb0 + lr] = src[nc + lr];
    }
    CHOL_STAMP(5);
}


// ---------------------------------------------------------------------------
// 8-byte asynchronous global -> shared copy (LDGSTS); `valid == false` zero-fills.
__device__ __forceinline__ void cp_async8(double* smemDst, const double* gsrc, bool valid) {
    const unsigned dst = (unsigned)__cvta_generic_to_shared(smemDst);
    const int bytes = valid ? 8 : 0;
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;\n" ::"r"(dst), "l"(gsrc), "r"(bytes));
}
__device__ __forceinline__ void cp_async_wait_all() {
    asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;\n" ::: "memory");
}

// U_t tile (ti >= tj), 48x48, 8 warps: warp (wr, wc, kh) owns the 24x24 quadrant (wr, wc) over
// half kh of K.  Staging is asynchronous (no load -> store serialisation), the two K halves are
// combined through shared memory in a fixed order.
constexpr int UPD3_TPB = 256;
__global__ void __launch_bounds__(UPD3_TPB) k_chol_update3(Dev d, Chol c, int lev, int kpadMax) {
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
    const int nc = n.nc, kp = (nc + 15) & ~15;           // two K halves of whole 8-steps
    double* As = sm;
    double* Bs = (ti == tj) ? sm : sm + kpadMax * UPD_LD;
    double* Cs = sm + 2 * kpadMax * UPD_LD;               // 48 x UPD_CLD
    double* Cs2 = Cs + TS * UPD_CLD;                      // second K half
    const double* __restrict__ Pn = c.L + c.Lptr[n.t];
    const int ch0 = c.childPtr[n.t], nch = c.childPtr[n.t + 1] - ch0;
    CHOL_STAMP(8);
    if (tid < nch && tid < CHOL_MAXCH) chs[tid] = load_child(c, ch0 + tid);
    for (int i = tid; i < kp * TS; i += UPD3_TPB) {
        const int r = i % TS, k = i / TS;
        const bool kin = k < nc;
        cp_async8(As + k * UPD_LD + r, Pn + (size_t)(kin ? k : 0) * n.m + nc + i0 + r, kin && i0 + r < n.u);
        if (ti != tj)
            cp_async8(Bs + k * UPD_LD + r, Pn + (size_t)(kin ? k : 0) * n.m + nc + j0 + r, kin && j0 + r < n.u);
    }
    cp_async_wait_all();
    __syncthreads();
    CHOL_STAMP(9);
    const int wq = wid & 3, kh = wid >> 2, wr = wq & 1, wc = wq >> 1;
    const bool active = !(ti == tj && wc > wr);
    const int fr8 = lane >> 2, fk = lane & 3;
    double acc[2][3][3][2];
    if (active) {
#pragma unroll
        for (int h = 0; h < 2; h++)
#pragma unroll
            for (int x = 0; x < 3; x++)
#pragma unroll
                for (int y = 0; y < 3; y++) acc[h][x][y][0] = acc[h][x][y][1] = 0.0;
        const double* Aw = As + wr * 24 + fr8;
        const double* Bw = Bs + wc * 24 + fr8;
        const int kHalf = kp >> 1;
        for (int k0 = kh * kHalf; k0 < (kh + 1) * kHalf; k0 += 8) {
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
        double* Cd = kh ? Cs2 : Cs;
#pragma unroll
        for (int x = 0; x < 3; x++)
#pragma unroll
            for (int y = 0; y < 3; y++)
#pragma unroll
                for (int e = 0; e < 2; e++) {
                    const int r = wr * 24 + 8 * x + fr8;
                    const int cc = wc * 24 + 8 * y + 2 * fk + e;
                    Cd[cc * UPD_CLD + r] = acc[0][x][y][e] + acc[1][x][y][e];
                }
    }
    __syncthreads();
    CHOL_STAMP(10);
    // epilogue: add the children's blocks, store the lower triangle of U_t
    double* __restrict__ Ut = c.U + c.Uptr[n.t];
    const int rt = n.rt;
    for (int wu = tid; wu < CHOL_TB * CHOL_TB * 6; wu += UPD3_TPB) {
        const int c6 = wu % 6, item = wu / 6, br = item % CHOL_TB, bc = item / CHOL_TB;
        const int gbr = CHOL_TB * ti + br, gbc = CHOL_TB * tj + bc;
        if (gbr > rt || gbc > gbr || gbc >= rt) continue;   // block row rt = the rhs row; the rhs column is never used
        const bool isRhs = (gbr == rt);
        const int rmin = (gbr == gbc) ? c6 : 0;
        double v[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
        pull_unit(c, chs, ch0, nch, isRhs ? -1 : n.nt + gbr, n.nt + gbc, c6, rmin, v);
        const int o = (6 * bc + c6) * UPD_CLD + 6 * br;
        double* __restrict__ dst = Ut + (size_t)(6 * gbc + c6) * n.u + 6 * gbr;
        if (!isRhs) {
#pragma unroll
            for (int rr = 0; rr < 6; rr++)
                if (rr >= rmin) dst[rr] = v[rr] - (Cs[o + rr] + Cs2[o + rr]);
        } else {
            dst[0] = v[0] - (Cs[o] + Cs2[o]);
        }
    }
    CHOL_STAMP(11);
}

// ---------------------------------------------------------------------------
// Backward substitution, one CTA per supernode, levels walked top-down (see k_chol_bwd2).
// L11 / W are staged asynchronously; the L21^T x_R product issues 4 x CPW independent loads per
// lane before the first use.
template <int TPB>
__global__ void __launch_bounds__(TPB) k_chol_bwd3(Dev d, Chol c, int lev) {
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
    for (int i = tid; i < nc * nc; i += TPB) {
        const int r = i % nc, cc = i / nc;
        if (r > cc) cp_async8(L11 + cc * ldL + r, Pn + (size_t)cc * n.m + r, true);
    }
    for (int i = tid; i < 21 * nt; i += TPB) cp_async8(Wsm + i, c.Winv + (size_t)n.s * 21 + i, true);
    for (int i = tid; i < nr; i += TPB) xR[i] = c.xperm[(size_t)__ldg(Rt + i / 6) * 6 + i % 6];
    __syncthreads();                    // xR visible; L11 / W are only needed after the product
    CHOL_STAMP(17);
    constexpr int NW = TPB / 32;
    constexpr int CPW = (6 * CHOL_MAXW + NW - 1) / NW;      // columns per warp
    const int lane = tid & 31, wid = tid >> 5;
    {
        double acc[CPW];
#pragma unroll
        for (int j = 0; j < CPW; j++) acc[j] = 0.0;
        for (int r0 = 0; r0 < nr; r0 += 128) {
            double v[4][CPW], xv[4];
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const int r = r0 + 32 * u + lane;
                const bool ok = r < nr;
                xv[u] = ok ? xR[r] : 0.0;
#pragma unroll
                for (int j = 0; j < CPW; j++) {
                    const int cc = wid + NW * j;
                    v[u][j] = (ok && cc < nc) ? Pn[(size_t)cc * n.m + nc + r] : 0.0;
                }
            }
#pragma unroll
            for (int u = 0; u < 4; u++)
#pragma unroll
                for (int j = 0; j < CPW; j++) acc[j] += v[u][j] * xv[u];
        }
#pragma unroll
        for (int j = 0; j < CPW; j++) {
            const int cc = wid + NW * j;
            const double s = warp_sum(acc[j]);
            if (lane == 0 && cc < nc) w[cc] = Pn[(size_t)cc * n.m + n.m - 1] - s;
        }
    }
    cp_async_wait_all();
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
