// kernels_chol3.cuh -- numeric multifrontal Cholesky of the reduced camera system and its solves
// (replaces cusolverSpDcsrcholFactor / csrcholSolve, src/cholesky.hpp:109-156): k_chol_panel3 (one CTA per
// (supernode, slab of block rows of L21): [D; S] assembled and factored in shared memory by a data-flow loop -- a chain warp for
// the 6x6 diagonal blocks, a warp per block row of the diagonal part, slab warps behind published counters, FP64 DMMA
// throughout), k_chol_update3 / k_chol_update_small (48x48 / 24x24 tiles of the update matrices), k_chol_zinv (inverse of L11
// on a side branch) and k_chol_bwd3 (backward solves).  DESIGN.md section 4 has the data flow and the measurements behind
// each choice.
#pragma once
#include "kernels_chol_common.cuh"

namespace cugo_b200 {

constexpr int PANEL3_TPB = 768;
constexpr int BWD_MAXG = 8;            // CTAs that may share one supernode in the backward solve
constexpr int PANEL3_NW = PANEL3_TPB / 32;

// pull the children's contributions to 6 consecutive rows (or the rhs element) of one front column
__device__ __forceinline__ void pull_unit(const Chol& c, const ChildRef* chs, int ch0, int nch, int fr, int fc, int c6,
                                          int rmin, double acc[6], int iStart = 0) {
    for (int i0 = iStart; i0 < nch; i0 += 4) {
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

// ---- children pulled through inverse maps staged in shared memory -----------------------------------
// invS[i * W + k] = row block of child i's update matrix that lands on the k-th front block this CTA touches
// (-1: none).  One shared-memory lookup replaces the dependent global load of pull_unit, so a unit costs a
// single L2 round trip, and two units are kept in flight per thread.
constexpr int PULL_W = CHOL_MAXW + CHOL_RB;        // panel: the nt diagonal blocks + the slab's block rows
struct PullUnit {
    const double* src[2];
    int kr, kc, c6, rmin;
};
__device__ __forceinline__ void pull_src(const ChildRef* chs, const int* invS, int W, int i, int nchS, PullUnit& u, int j) {
    u.src[j] = nullptr;
    if (i < nchS) {
        const int kc = invS[i * W + u.kc];
        if (kc >= 0) {
            const ChildRef cr = chs[i];
            if (u.kr >= 0) {
                const int kr = invS[i * W + u.kr];
                if (kr >= 0) u.src[j] = cr.U + (size_t)(6 * kc + u.c6) * cr.uc + 6 * kr;
            } else {
                u.src[j] = cr.U + (size_t)(6 * kc + u.c6) * cr.uc + cr.uc - 1;
            }
        }
    }
}
__device__ __forceinline__ void pull_load(const PullUnit& u, int j, double v[6]) {
#pragma unroll
    for (int rr = 0; rr < 6; rr++) v[rr] = (u.src[j] && rr >= u.rmin && (u.kr >= 0 || rr == 0)) ? u.src[j][rr] : 0.0;
}
// sums of the children's contributions to two units (children in index order, as pull_unit)
__device__ __forceinline__ void pull_pair(const ChildRef* chs, const int* invS, int W, int nchS, PullUnit& a, PullUnit& b,
                                          bool hasB, double accA[6], double accB[6]) {
    for (int i0 = 0; i0 < nchS; i0 += 2) {
        pull_src(chs, invS, W, i0, nchS, a, 0);
        pull_src(chs, invS, W, i0 + 1, nchS, a, 1);
        if (hasB) {
            pull_src(chs, invS, W, i0, nchS, b, 0);
            pull_src(chs, invS, W, i0 + 1, nchS, b, 1);
        } else {
            b.src[0] = b.src[1] = nullptr;
        }
        double va[2][6], vb[2][6];
        pull_load(a, 0, va[0]); pull_load(a, 1, va[1]);
        pull_load(b, 0, vb[0]); pull_load(b, 1, vb[1]);
#pragma unroll
        for (int j = 0; j < 2; j++)
#pragma unroll
            for (int rr = 0; rr < 6; rr++) { accA[rr] += va[j][rr]; accB[rr] += vb[j][rr]; }
    }
}

// publish / observe a step counter in shared memory with release / acquire semantics at CTA scope
__device__ __forceinline__ void st_release_cta(int* p, int v) {
    asm volatile("st.release.cta.shared.s32 [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(p)), "r"(v) : "memory");
}
__device__ __forceinline__ int ld_acquire_cta(const int* p) {
    int v;
    asm volatile("ld.acquire.cta.shared.s32 %0, [%1];" : "=r"(v) : "r"((unsigned)__cvta_generic_to_shared(p)) : "memory");
    return v;
}
// Cholesky factor of the 6x6 diagonal block jb of the shared-memory panel by one warp (every lane the same
// arithmetic, lane 0 stores): L_D into the panel, the reciprocal diagonal into invAll.
__device__ __forceinline__ void panel_factor_diag(double* A, double* invAll, int* badFlag, int jb, int lane, double* Wsm = nullptr) {
    const int j0 = 6 * jb;
    double Ld[21], invd[6];
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
    if (lane == 0) {
#pragma unroll
        for (int i = 0; i < 6; i++)
#pragma unroll
            for (int k = 0; k <= i; k++) A[(j0 + k) * PANEL_LD + j0 + i] = Ld[tri(i, k)];
#pragma unroll
        for (int k = 0; k < 6; k++) invAll[6 * jb + k] = invd[k];
        if (bad) *badFlag = 1;
    }
    if (Wsm != nullptr) {
        // W = L_D^-1 (lower triangular), column k = lane by forward substitution on the unit vector e_k:
        // w_i = ([i == k] - sum_{p<i} l_ip w_p) / l_ii -- the same instruction stream in every lane (entries above the
        // diagonal come out as exact zeros), lanes 0..5 store.  Dense, row-major: the worker warps multiply their
        // rows by W^T on the FP64 tensor pipe instead of substituting (panel_tile_solve).
        double w[6];
#pragma unroll
        for (int i = 0; i < 6; i++) {
            double sv = (lane == i) ? 1.0 : 0.0;
#pragma unroll
            for (int p2 = 0; p2 < i; p2++) sv -= Ld[tri(i, p2)] * w[p2];
            w[i] = sv * invd[i];
        }
        if (lane < 6) {
#pragma unroll
            for (int i = 0; i < 6; i++) Wsm[i * 6 + lane] = w[i];
        }
    }
}

// C-fragment pair (x0, x1) = X[g][2q], X[g][2q+1] of an 8x6 block -> its A-fragment pair X[g][q], X[g][4+q] (the latter zero for
// q >= 2), which is also the B-fragment of X^T.  Four 64-bit shuffles inside the quad of row g.
__device__ __forceinline__ void frag_c_to_a(double x0, double x1, int lane, double& a0, double& a1) {
    const int g4 = lane & ~3, q = lane & 3;
    const double v0 = __shfl_sync(0xffffffffu, x0, g4 + (q >> 1)), v1 = __shfl_sync(0xffffffffu, x1, g4 + (q >> 1));
    const double w0 = __shfl_sync(0xffffffffu, x0, g4 + 2), w1 = __shfl_sync(0xffffffffu, x1, g4 + 2);
    a0 = (q & 1) ? v1 : v0;
    a1 = q >= 2 ? 0.0 : (q ? w1 : w0);
}

// The chain's factor of the data-flow loop: Cholesky of D_jb AND W = L_D^-1 in one pass (row k of W needs row k of L and
// the rows of W above it, so it fills the bubbles of the next pivot's rsqrt), W published through `pub` (value jb + 1)
// BEFORE L_D and the reciprocal diagonal are stored -- nothing inside the loop reads those two.
// Lane `lane` carries column min(lane, 5)... every lane runs the same instruction stream: column k = lane of W is
// w_i = ([i == k] - sum_{p<i} l_ip w_p) / l_ii, exact zeros above the diagonal; lanes 0..5 store.
// 1 / sqrt(x) for the chain: the hardware approximation (about 20 bits) and ONE third-order step
// y (1 + e/2 + 3/8 e^2), e = 1 - x y^2 -- error e^3 ~ 2^-60 before rounding; no special cases (x is a pivot > 1e-14, or the
// factorisation has failed anyway).  55 dependent cycles against 74 for rsqrt().
__device__ __forceinline__ double chain_rsqrt(double x) {
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    const double e = fma(-x * y, y, 1.0);
    const double p = fma(e, 0.375, 0.5);
    return fma(y * e, p, y);
}
__device__ __forceinline__ void panel_chain_core(double (&Ld)[21], double* A, double* invAll, int* badFlag, int jb, int lane, double* Wsm, int* pub) {
    const int j0 = 6 * jb;
    double invd[6], w[6];
    bool bad = false;
#pragma unroll
    for (int k = 0; k < 6; k++) {
        // a pivot <= tol fails the factorisation (the trial is rejected: st->fail); the test is kept OFF the dependent chain --
        // what follows a bad pivot is never used, it only has to terminate
        const double x = Ld[tri(k, k)];
        bad = bad || !(x > CHOL_PIVOT_TOL);
        double sv = (lane == k) ? 1.0 : 0.0;            // row k of W: independent of this pivot's rsqrt up to the last product
#pragma unroll
        for (int p2 = 0; p2 < k; p2++) sv -= Ld[tri(k, p2)] * w[p2];
        const double inv = chain_rsqrt(x);
        invd[k] = inv;
        w[k] = sv * inv;
        Ld[tri(k, k)] = x * inv;
#pragma unroll
        for (int i = k + 1; i < 6; i++) Ld[tri(i, k)] *= inv;
#pragma unroll
        for (int cc = k + 1; cc < 6; cc++)
#pragma unroll
            for (int i = cc; i < 6; i++) Ld[tri(i, cc)] -= Ld[tri(i, k)] * Ld[tri(cc, k)];
    }
    if (lane < 6) {
#pragma unroll
        for (int i = 0; i < 6; i++) Wsm[i * 6 + lane] = w[i];
    }
    __syncwarp();
    if (lane == 0) {
        st_release_cta(pub, jb + 1);
#pragma unroll
        for (int i = 0; i < 6; i++)
#pragma unroll
            for (int k = 0; k <= i; k++) A[(j0 + k) * PANEL_LD + j0 + i] = Ld[tri(i, k)];
#pragma unroll
        for (int k = 0; k < 6; k++) invAll[6 * jb + k] = invd[k];
        if (bad) *badFlag = 1;
    }
}
__device__ __forceinline__ void panel_chain_factor(double* A, double* invAll, int* badFlag, int jb, int lane, double* Wsm, int* pub) {
    const int j0 = 6 * jb;
    double Ld[21];
#pragma unroll
    for (int i = 0; i < 6; i++)
#pragma unroll
        for (int k = 0; k <= i; k++) Ld[tri(i, k)] = A[(j0 + k) * PANEL_LD + j0 + i];
    panel_chain_core(Ld, A, invAll, badFlag, jb, lane, Wsm, pub);
}
// One step of the chain for block row r >= 1 (see the loop in k_chol_panel3): X[r, r-1] = C W_(r-1)^T on the tensor pipe,
// published; D_r -= X X^T in fragment registers, handed to every lane by shuffles (entry (i, k) sits in lane 4 i + k / 2);
// factor + W_r (panel_chain_core).
__device__ __forceinline__ void panel_chain_step(double* A, double* invAll, int* badFlag, int r, int R, int lane, double (*Wall)[36],
                                                 int* pubW, int* pubL) {
    const int g = lane >> 2, q = lane & 3, cl = 2 * q, r0 = 6 * r, j0 = r0 - 6, row = r0 + g;
    const int clc = cl < 6 ? cl : 0, n6 = g < 6 ? g : 0;
    const double c0 = A[(j0 + clc) * PANEL_LD + row], c1 = A[(j0 + clc + 1) * PANEL_LD + row];     // C[r, r-1]
    double d0 = A[(r0 + clc) * PANEL_LD + row], d1 = A[(r0 + clc + 1) * PANEL_LD + row];           // D_r so far
    double wb0 = Wall[r - 1][n6 * 6 + q], wb1 = Wall[r - 1][n6 * 6 + 4 + (q & 1)];
    if (q >= 2) wb1 = 0.0;
    if (g >= 6) { wb0 = 0.0; wb1 = 0.0; }
    double ca0, ca1;
    frag_c_to_a(c0, c1, lane, ca0, ca1);
    double x0 = 0.0, x1 = 0.0;
    dmma884(x0, x1, ca0, wb0);
    dmma884(x0, x1, ca1, wb1);
    if (cl < 6 && g < 6) {
        A[(j0 + cl) * PANEL_LD + row] = x0;
        A[(j0 + cl + 1) * PANEL_LD + row] = x1;
    }
    __syncwarp();
    if (lane == 0) st_release_cta(pubL, r);
    double xa0, xa1;
    frag_c_to_a(x0, x1, lane, xa0, xa1);
    dmma884(d0, d1, -xa0, xa0);
    dmma884(d0, d1, -xa1, xa1);
    double Ld[21];
#pragma unroll
    for (int i = 0; i < 6; i++)
#pragma unroll
        for (int k = 0; k <= i; k++) {
            const double v = __shfl_sync(0xffffffffu, (k & 1) ? d1 : d0, 4 * i + (k >> 1));
            Ld[tri(i, k)] = v;
        }
    panel_chain_core(Ld, A, invAll, badFlag, r, lane, Wall[r], pubW);
}

// S-phase of one 8-row tile on the tensor pipe: X = A_blk W^T for rows [r0, r0 + 8) of block column jb (W = L_D(jb)^-1 from
// panel_factor_diag).  Fragment layout of m8n8k4: lane (g = lane >> 2, q = lane & 3) holds A[g][q] / B[q][g] and receives
// C[g][2q], C[g][2q + 1]; K = 6 takes two instructions, the second with k = 4, 5 only.
__device__ __forceinline__ void panel_tile_solve(double* A, const double* Wsm, int jb, int r0, int R, int lane, int nrows = 8) {
    const int j0 = 6 * jb, g = lane >> 2, q = lane & 3;
    const int r = r0 + g;
    double a0 = A[(j0 + q) * PANEL_LD + r], a1 = A[(j0 + 4 + (q & 1)) * PANEL_LD + r];
    const int n = g < 6 ? g : 0;
    double b0 = Wsm[n * 6 + q], b1 = Wsm[n * 6 + 4 + (q & 1)];
    if (q >= 2) { a1 = 0.0; b1 = 0.0; }
    if (g >= 6) { b0 = 0.0; b1 = 0.0; }
    double x0 = 0.0, x1 = 0.0;
    dmma884(x0, x1, a0, b0);
    dmma884(x0, x1, a1, b1);
    if (q < 3 && r < R && g < nrows) {
        A[(j0 + 2 * q) * PANEL_LD + r] = x0;
        A[(j0 + 2 * q + 1) * PANEL_LD + r] = x1;
    }
}


// two tile solves in one instruction stream (rows r0a.. and, if `two`, r0b..), see panel_tile_solve
__device__ __forceinline__ void panel_tile_solve2(double* A, const double* Wsm, int jb, int r0a, int r0b, bool two, int R, int lane, int nrows) {
    const int j0 = 6 * jb, g = lane >> 2, q = lane & 3;
    const int ra = r0a + g, rb = (two ? r0b : r0a) + g;
    double a0 = A[(j0 + q) * PANEL_LD + ra], a1 = A[(j0 + 4 + (q & 1)) * PANEL_LD + ra];
    double p0 = A[(j0 + q) * PANEL_LD + rb], p1 = A[(j0 + 4 + (q & 1)) * PANEL_LD + rb];
    const int n = g < 6 ? g : 0;
    double b0 = Wsm[n * 6 + q], b1 = Wsm[n * 6 + 4 + (q & 1)];
    if (q >= 2) { a1 = 0.0; p1 = 0.0; b1 = 0.0; }
    if (g >= 6) { b0 = 0.0; b1 = 0.0; }
    double x0 = 0.0, x1 = 0.0, y0 = 0.0, y1 = 0.0;
    dmma884(x0, x1, a0, b0);
    dmma884(y0, y1, p0, b0);
    dmma884(x0, x1, a1, b1);
    dmma884(y0, y1, p1, b1);
    if (q < 3 && g < nrows) {
        if (ra < R) {
            A[(j0 + 2 * q) * PANEL_LD + ra] = x0;
            A[(j0 + 2 * q + 1) * PANEL_LD + ra] = x1;
        }
        if (two && rb < R) {
            A[(j0 + 2 * q) * PANEL_LD + rb] = y0;
            A[(j0 + 2 * q + 1) * PANEL_LD + rb] = y1;
        }
    }
    __syncwarp();
}

// Rows below the diagonal part (the slab rows and the rhs row) are nobody's operand inside the panel: X = S L11^-T can lag
// behind the factorisation of D.  A slab warp owns one or two fixed 8-row tiles and runs its own left-looking loop over the
// block columns, tied to the chain only by the published step counter: for block column cb
//   C = A[tile, cb] - sum_{k < 6 cb} X[tile, k] L[6 cb .., k]^T        (FP64 DMMA, K = 6 cb, two accumulators per tile)
//   X[tile, cb] = C W_cb^T                                            (panel_tile_solve, through the panel itself)
// Reads: its own rows, and rows [6 cb, 6 cb + 8) of the finished columns -- final once the chain has factored D_cb.
__device__ __forceinline__ void panel_slab_tiles(double* A, const double* Wcb, int cb, int ra0, int rb0, bool two, int R, int lane, int phase = 3) {
    const int g = lane >> 2, q = lane & 3, cl = 2 * q;
    const bool st = cl < 6;
    const int c0 = 6 * cb, ra = ra0 + g, rb = rb0 + g;
    const int cu = st ? c0 + cl : c0;
    if (cb > 0 && (phase & 1)) {
        const double ta0 = A[cu * PANEL_LD + ra], ta1 = A[(cu + 1) * PANEL_LD + ra];
        const double tb0 = A[cu * PANEL_LD + rb], tb1 = A[(cu + 1) * PANEL_LD + rb];
        double c00 = 0.0, c01 = 0.0, c10 = 0.0, c11 = 0.0, e00 = 0.0, e01 = 0.0, e10 = 0.0, e11 = 0.0;
        const double* Bp = A + c0 + g + q * PANEL_LD;
        const double* Ap = A + ra + q * PANEL_LD;
        const double* Aq = A + rb + q * PANEL_LD;
        const int nfull = c0 >> 3, rem = c0 & 7;
        const bool va = q < rem, vb = q + 4 < rem;
        if (two) {
#pragma unroll 2
            for (int it = 0; it < nfull; it++) {
                const double b0 = Bp[0], b1 = Bp[4 * PANEL_LD];
                const double a0 = Ap[0], a1 = Ap[4 * PANEL_LD];
                const double q0 = Aq[0], q1 = Aq[4 * PANEL_LD];
                dmma884(c00, c01, a0, b0);
                dmma884(c10, c11, a1, b1);
                dmma884(e00, e01, q0, b0);
                dmma884(e10, e11, q1, b1);
                Bp += 8 * PANEL_LD; Ap += 8 * PANEL_LD; Aq += 8 * PANEL_LD;
            }
            if (rem) {                      // columns up to c0 + 7 exist only for cb + 1 < nt: clamp the second half
                const int o1 = (q + 4 < rem) ? 4 * PANEL_LD : 0;
                double b0 = Bp[0], b1 = Bp[o1], a0 = Ap[0], a1 = Ap[o1], q0 = Aq[0], q1 = Aq[o1];
                if (!va) { b0 = 0.0; a0 = 0.0; q0 = 0.0; }
                if (!vb) { b1 = 0.0; a1 = 0.0; q1 = 0.0; }
                dmma884(c00, c01, a0, b0);
                dmma884(c10, c11, a1, b1);
                dmma884(e00, e01, q0, b0);
                dmma884(e10, e11, q1, b1);
            }
        } else {
#pragma unroll 2
            for (int it = 0; it < nfull; it++) {
                const double b0 = Bp[0], b1 = Bp[4 * PANEL_LD];
                const double a0 = Ap[0], a1 = Ap[4 * PANEL_LD];
                dmma884(c00, c01, a0, b0);
                dmma884(c10, c11, a1, b1);
                Bp += 8 * PANEL_LD; Ap += 8 * PANEL_LD;
            }
            if (rem) {
                const int o1 = (q + 4 < rem) ? 4 * PANEL_LD : 0;
                double b0 = Bp[0], b1 = Bp[o1], a0 = Ap[0], a1 = Ap[o1];
                if (!va) { b0 = 0.0; a0 = 0.0; }
                if (!vb) { b1 = 0.0; a1 = 0.0; }
                dmma884(c00, c01, a0, b0);
                dmma884(c10, c11, a1, b1);
            }
        }
        if (st && ra < R) {
            A[(c0 + cl) * PANEL_LD + ra] = ta0 - (c00 + c10);
            A[(c0 + cl + 1) * PANEL_LD + ra] = ta1 - (c01 + c11);
        }
        if (two && st && rb < R) {
            A[(c0 + cl) * PANEL_LD + rb] = tb0 - (e00 + e10);
            A[(c0 + cl + 1) * PANEL_LD + rb] = tb1 - (e01 + e11);
        }
        __syncwarp();
    }
    if (phase & 2) panel_tile_solve2(A, Wcb, cb, ra0, rb0, two, R, lane, 8);
}

// spin until a published counter reaches `need` (every lane acquires)
__device__ __forceinline__ void wait_counter(const int* p, int need, bool nap) {
    int v;
    for (;;) {
        asm volatile("ld.acquire.cta.shared.s32 %0, [%1];" : "=r"(v) : "r"((unsigned)__cvta_generic_to_shared(p)) : "memory");
        if (v >= need) break;
        if (nap) __nanosleep(40);
    }
    __syncwarp();
}

// One block row r (rows 6r .. 6r+5; 8 are computed) of the diagonal part, all its column blocks t = 0 .. r-2, by one warp.
// The running target C[r, t] and the diagonal block D_r live in REGISTERS (C-fragment layout) from stage to stage; shared
// memory sees only the solved X[r, t] (which other warps read) and, at the end, C[r, r-1] and D_r for the chain.  Per stage:
//   cn  = A[r, t+1] - sum_{k < 6t} X[r, k] L[t+1, k]^T        long sum, before the chain is needed (rowProg[t+1] >= t)
//   X   = C W_t^T                                           after pubW > t; stored; A-fragment of X by shuffles
//   D  -= X X^T ;  cn -= X L[t+1, t]^T  (after pubL > t) ;  C = cn
__device__ __forceinline__ void panel_row_run(double* A, const double (*Wall)[36], int r, int R, int lane, const int* pubW,
                                              const int* pubL, int* rowProg, int* preReady) {
    const int g = lane >> 2, q = lane & 3, cl = 2 * q, r0 = 6 * r, row = r0 + g;
    const bool st = cl < 6 && g < 6 && row < R;
    const int clc = cl < 6 ? cl : 0;
    double c0 = A[clc * PANEL_LD + row], c1 = A[(clc + 1) * PANEL_LD + row];                       // C[r, 0]
    double d0 = A[(r0 + clc) * PANEL_LD + row], d1 = A[(r0 + clc + 1) * PANEL_LD + row];           // D_r
    const int n6 = g < 6 ? g : 0;
    for (int t = 0; t + 1 < r; t++) {
        const int j0 = 6 * t, j1 = j0 + 6;
        // ---- next target and its long sum
        double n0 = A[(j1 + clc) * PANEL_LD + row], n1 = A[(j1 + clc + 1) * PANEL_LD + row];
        if (t > 0) {
            wait_counter(rowProg + t + 1, t, true);
            double s00 = 0.0, s01 = 0.0, s10 = 0.0, s11 = 0.0;
            const double* Bp = A + j1 + g + q * PANEL_LD;
            const double* Ap = A + row + q * PANEL_LD;
            const int nfull = j0 >> 3, rem = j0 & 7;
#pragma unroll 4
            for (int it = 0; it < nfull; it++) {
                const double b0 = Bp[0], b1 = Bp[4 * PANEL_LD];
                const double a0 = Ap[0], a1 = Ap[4 * PANEL_LD];
                dmma884(s00, s01, a0, b0);
                dmma884(s10, s11, a1, b1);
                Bp += 8 * PANEL_LD; Ap += 8 * PANEL_LD;
            }
            if (rem) {
                const bool va = q < rem, vb = q + 4 < rem;
                const int o1 = vb ? 4 * PANEL_LD : 0;
                double b0 = Bp[0], b1 = Bp[o1], a0 = Ap[0], a1 = Ap[o1];
                if (!va) { b0 = 0.0; a0 = 0.0; }
                if (!vb) { b1 = 0.0; a1 = 0.0; }
                dmma884(s00, s01, a0, b0);
                dmma884(s10, s11, a1, b1);
            }
            n0 -= s00 + s10; n1 -= s01 + s11;
        }
        // ---- solve: X = C W_t^T
        double ca0, ca1;
        frag_c_to_a(c0, c1, lane, ca0, ca1);
        wait_counter(pubW, t + 1, true);
        double wb0 = Wall[t][n6 * 6 + q], wb1 = Wall[t][n6 * 6 + 4 + (q & 1)];
        if (q >= 2) wb1 = 0.0;
        if (g >= 6) { wb0 = 0.0; wb1 = 0.0; }
        double x0 = 0.0, x1 = 0.0;
        dmma884(x0, x1, ca0, wb0);
        dmma884(x0, x1, ca1, wb1);
        if (st) {
            A[(j0 + cl) * PANEL_LD + row] = x0;
            A[(j0 + cl + 1) * PANEL_LD + row] = x1;
        }
        __syncwarp();
        if (lane == 0) st_release_cta(rowProg + r, t + 1);
        double xa0, xa1;
        frag_c_to_a(x0, x1, lane, xa0, xa1);
        // ---- D_r -= X X^T
        dmma884(d0, d1, -xa0, xa0);
        dmma884(d0, d1, -xa1, xa1);
        // ---- cn -= X L[t+1, t]^T
        wait_counter(pubL, t + 1, true);
        double lb0 = A[(j0 + q) * PANEL_LD + j1 + g], lb1 = A[(j0 + 4 + (q & 1)) * PANEL_LD + j1 + g];
        if (q >= 2) lb1 = 0.0;
        dmma884(n0, n1, -xa0, lb0);
        dmma884(n0, n1, -xa1, lb1);
        c0 = n0; c1 = n1;
    }
    if (st) {
        if (r >= 2) {
            A[(r0 - 6 + cl) * PANEL_LD + row] = c0;            // C[r, r-1] for the chain
            A[(r0 - 6 + cl + 1) * PANEL_LD + row] = c1;
        }
        A[(r0 + cl) * PANEL_LD + row] = d0;
        A[(r0 + cl + 1) * PANEL_LD + row] = d1;
    }
    __syncwarp();
    if (lane == 0) st_release_cta(preReady + r, 1);
}

// NT 8-row slab tiles (rows rt[i] .. rt[i] + 7, below the diagonal part) through ALL block columns t = 0 .. nt-1 by one warp, in one
// instruction stream; the structure of panel_row_run without a diagonal block: the running targets stay in registers, shared
// memory sees the solved X only.  The long sum of the next target needs block row t + 1 of L up to column 6 t (rowProg), the solve
// W_t (pubW), the rank-6 rest L[t+1, t] (pubL) -- so a slab tile finishes one solve after the chain.
template <int NT>
__device__ __forceinline__ void panel_slab_run(double* A, const double (*Wall)[36], const int (&rt)[NT], int nt, int R, int lane,
                                               const int* pubW, const int* pubL, const int* rowProg) {
    const int g = lane >> 2, q = lane & 3, cl = 2 * q;
    const int clc = cl < 6 ? cl : 0, n6 = g < 6 ? g : 0;
    int row[NT];
    bool st[NT];
    double c0[NT], c1[NT];
#pragma unroll
    for (int i = 0; i < NT; i++) {
        row[i] = min(rt[i] + g, PANEL_LD - 1);           // rows past R are computed and dropped; keep their loads inside the panel
        st[i] = cl < 6 && rt[i] + g < R;
        c0[i] = A[clc * PANEL_LD + row[i]]; c1[i] = A[(clc + 1) * PANEL_LD + row[i]];
    }
    for (int t = 0; t < nt; t++) {
        const int j0 = 6 * t, j1 = j0 + 6;
        const bool more = t + 1 < nt;
        double n0[NT], n1[NT];
        if (more) {
#pragma unroll
            for (int i = 0; i < NT; i++) { n0[i] = A[(j1 + clc) * PANEL_LD + row[i]]; n1[i] = A[(j1 + clc + 1) * PANEL_LD + row[i]]; }
            if (t > 0) {
                wait_counter(rowProg + t + 1, t, true);
                double s0[NT], s1[NT], u0[NT], u1[NT];
#pragma unroll
                for (int i = 0; i < NT; i++) { s0[i] = s1[i] = u0[i] = u1[i] = 0.0; }
                const double* Bp = A + j1 + g + q * PANEL_LD;
                const int qo = q * PANEL_LD;
                const int nfull = j0 >> 3, rem = j0 & 7;
#pragma unroll 2
                for (int it = 0; it < nfull; it++) {
                    const double b0 = Bp[0], b1 = Bp[4 * PANEL_LD];
                    double a0[NT], a1[NT];
#pragma unroll
                    for (int i = 0; i < NT; i++) { a0[i] = A[qo + (8 * it) * PANEL_LD + row[i]]; a1[i] = A[qo + (8 * it + 4) * PANEL_LD + row[i]]; }
#pragma unroll
                    for (int i = 0; i < NT; i++) { dmma884(s0[i], s1[i], a0[i], b0); dmma884(u0[i], u1[i], a1[i], b1); }
                    Bp += 8 * PANEL_LD;
                }
                if (rem) {
                    const bool va = q < rem, vb = q + 4 < rem;
                    const int o1 = vb ? 4 * PANEL_LD : 0;
                    double b0 = Bp[0], b1 = Bp[o1];
                    if (!va) b0 = 0.0;
                    if (!vb) b1 = 0.0;
#pragma unroll
                    for (int i = 0; i < NT; i++) {
                        double a0 = A[qo + (8 * nfull) * PANEL_LD + row[i]], a1 = A[qo + (8 * nfull) * PANEL_LD + o1 + row[i]];
                        if (!va) a0 = 0.0;
                        if (!vb) a1 = 0.0;
                        dmma884(s0[i], s1[i], a0, b0);
                        dmma884(u0[i], u1[i], a1, b1);
                    }
                }
#pragma unroll
                for (int i = 0; i < NT; i++) { n0[i] -= s0[i] + u0[i]; n1[i] -= s1[i] + u1[i]; }
            }
        }
        // ---- solve
        double ca0[NT], ca1[NT];
#pragma unroll
        for (int i = 0; i < NT; i++) frag_c_to_a(c0[i], c1[i], lane, ca0[i], ca1[i]);
        wait_counter(pubW, t + 1, true);
        double wb0 = Wall[t][n6 * 6 + q], wb1 = Wall[t][n6 * 6 + 4 + (q & 1)];
        if (q >= 2) wb1 = 0.0;
        if (g >= 6) { wb0 = 0.0; wb1 = 0.0; }
        double x0[NT], x1[NT];
#pragma unroll
        for (int i = 0; i < NT; i++) { x0[i] = 0.0; x1[i] = 0.0; dmma884(x0[i], x1[i], ca0[i], wb0); }
#pragma unroll
        for (int i = 0; i < NT; i++) dmma884(x0[i], x1[i], ca1[i], wb1);
#pragma unroll
        for (int i = 0; i < NT; i++)
            if (st[i]) { A[(j0 + cl) * PANEL_LD + row[i]] = x0[i]; A[(j0 + cl + 1) * PANEL_LD + row[i]] = x1[i]; }
        __syncwarp();
        if (!more) break;
        double xa0[NT], xa1[NT];
#pragma unroll
        for (int i = 0; i < NT; i++) frag_c_to_a(x0[i], x1[i], lane, xa0[i], xa1[i]);
        wait_counter(pubL, t + 1, true);
        double lb0 = A[(j0 + q) * PANEL_LD + j1 + g], lb1 = A[(j0 + 4 + (q & 1)) * PANEL_LD + j1 + g];
        if (q >= 2) lb1 = 0.0;
#pragma unroll
        for (int i = 0; i < NT; i++) {
            dmma884(n0[i], n1[i], -xa0[i], lb0);
            dmma884(n0[i], n1[i], -xa1[i], lb1);
            c0[i] = n0[i]; c1[i] = n1[i];
        }
    }
}


// `early` (set by the host for lev > 0 under programmatic dependent launch): the predecessor kernel -- the
// update or panel kernel of the level below -- signals its dependents only after its own pdl_wait(), so when this
// CTA runs, everything up to the Schur kernels has completed and the Hsc / bsc scatter may run before pdl_wait(),
// overlapping the predecessor's tail.  Zeroing and the structure reads always do.
// DC (decoupled chain, the default): warp 0 runs the whole chain of diagonal-block factors and is tied to the
// worker warps by named barriers only -- see the loop.  DC = false keeps the CTA-wide two-barrier loop.
template <bool DC>
__global__ void __launch_bounds__(PANEL3_TPB) k_chol_panel3(Dev d, Chol c, int lev, int early) {
    extern __shared__ __align__(16) double A[];   // PANEL_LD x nc, column-major
    __shared__ ChildRef chs[CHOL_MAXCH];
    __shared__ __align__(16) double invAll[6 * CHOL_MAXW];  // reciprocal diagonal of every factored 6x6 block
    __shared__ int badFlag;
    __shared__ int invS[CHOL_MAXCH * PULL_W];
    __shared__ int rowTab[4][PANEL_LD];
    __shared__ long long tBaseS;
    __shared__ __align__(16) double Wall[CHOL_MAXW][36];   // L_D^-1 of every block column (panel_factor_diag)
    __shared__ int pubW, pubL, preReady[CHOL_MAXW], rowProg[CHOL_MAXW];   // counters / flags of the data-flow loop (see there)
    LMState* st = d.st;
    const int q0 = c.levelPtr[lev], q1 = c.levelPtr[lev + 1];
    int q, slab;
    find_task(c.trsmPtr, q0, q1, c.trsmPtr[q0] + blockIdx.x, &q, &slab);
    const SnInfo n = sn_info(c, c.levelSn[q]);
    const int nc = n.nc, nt = n.nt, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int rbL = c.rbLev[lev];
    const int rb0 = nt + slab * rbL;           // first front block row of the slab
    const int rb1 = min(rb0 + rbL, n.fb);
    const bool last = (rb1 == n.fb);              // the last slab carries the rhs row
    const int nS = 6 * (rb1 - rb0) + (last ? 1 : 0);
    const int R = nc + nS;
    const int ch0 = c.childPtr[n.t], nch = c.childPtr[n.t + 1] - ch0;

    CHOL_STAMP(0);
    if (tid == 0) { badFlag = 0; pubW = 0; pubL = 0; }
    if (tid < CHOL_MAXW) { preReady[tid] = 0; rowProg[tid] = 0; }
    for (int i = tid; i < nc * PANEL_LD; i += PANEL3_TPB) A[i] = 0.0;
    if (tid < nch && tid < CHOL_MAXCH) chs[tid] = load_child(c, ch0 + tid);
    __syncthreads();
    CHOL_STAMP(1);
    const int nchS = min(nch, CHOL_MAXCH), pullW = nt + (rb1 - rb0);
    for (int i = tid; i < nchS * pullW; i += PANEL3_TPB) {         // structure data: may run before pdl_wait
        const int k = i % pullW;
        invS[i] = __ldg(chs[i / pullW].inv + (k < nt ? k : rb0 + (k - nt)));
    }
    // ---- original entries of Hsc: work unit = (entry, block column c6)
    if (!early) {
        pdl_wait(); pdl_trigger();
        if (st->done) return;
        CHOL_GT(0);
    }
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
    if (early) {
        pdl_wait(); pdl_trigger();
        if (st->done) return;
        CHOL_GT(0);
    }
    __syncthreads();
    CHOL_STAMP(2);
    // ---- children: extend-add of their update matrices (lower triangle of D, the slab, the rhs row).
    // Four children at a time: a table rowTab[j][r] = row of child j's update matrix that lands on panel row r
    // (-1: none) is built first, so the pull itself is ~6 instructions per entry (the block-wise version before it
    // spent its time on index arithmetic and on the L1 tag stage: one lookup per lane and load).  A warp takes two
    // panel columns at a time with its lanes along the ROWS -- consecutive panel rows are mostly consecutive rows of
    // a child's column, so a warp load touches 2-3 lines -- and keeps 16 loads per lane in flight.
    if (nchS > 0) {
        const int rhsRow = last ? R - 1 : -1;
        for (int g0 = 0; g0 < nchS; g0 += 4) {
            const int ng = min(4, nchS - g0);
            if (g0 > 0) __syncthreads();
            for (int idx = tid; idx < ng * R; idx += PANEL3_TPB) {
                const int j = idx / R, r = idx - j * R;
                int row;
                if (r == rhsRow) {
                    row = chs[g0 + j].uc - 1;
                } else {
                    const int la = r < nc ? r : r - nc, b = la / 6;
                    const int k = invS[(g0 + j) * pullW + (r < nc ? b : nt + b)];
                    row = k < 0 ? -1 : 6 * k + (la - 6 * b);
                }
                rowTab[j][r] = row;
            }
            __syncthreads();
            for (int ca = wid; ca < nc; ca += 2 * PANEL3_NW) {
                const int cb = ca + PANEL3_NW;
                const bool hasB = cb < nc;
                const double* colA[4];
                const double* colB[4];
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    colA[j] = nullptr; colB[j] = nullptr;
                    if (j < ng) {
                        const ChildRef cr = chs[g0 + j];
                        const int ka = invS[(g0 + j) * pullW + ca / 6];
                        const int kb = hasB ? invS[(g0 + j) * pullW + cb / 6] : -1;
                        if (ka >= 0) colA[j] = cr.U + (size_t)(6 * ka + ca % 6) * cr.uc;
                        if (kb >= 0) colB[j] = cr.U + (size_t)(6 * kb + cb % 6) * cr.uc;
                    }
                }
                for (int off = lane; ca + off < R; off += 128) {
                    double sA[4] = {0.0, 0.0, 0.0, 0.0}, sB[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
                    for (int jj = 0; jj < 4; jj += 2) {
                        if (jj < ng) {
                            double v[2][2][4];
#pragma unroll
                            for (int j = 0; j < 2; j++)
#pragma unroll
                                for (int u = 0; u < 4; u++) {
                                    const int ra = ca + off + 32 * u, rb = cb + off + 32 * u;
                                    const int rowA = ra < R ? rowTab[jj + j][ra] : -1;
                                    const int rowB = rb < R ? rowTab[jj + j][rb] : -1;
                                    const bool va = rowA >= 0 && colA[jj + j] != nullptr;
                                    const bool vb = rowB >= 0 && colB[jj + j] != nullptr;
                                    v[j][0][u] = va ? __ldcg(colA[jj + j] + rowA) : 0.0;
                                    v[j][1][u] = vb ? __ldcg(colB[jj + j] + rowB) : 0.0;
                                }
#pragma unroll
                            for (int j = 0; j < 2; j++)
#pragma unroll
                                for (int u = 0; u < 4; u++) { sA[u] += v[j][0][u]; sB[u] += v[j][1][u]; }
                        }
                    }
#pragma unroll
                    for (int u = 0; u < 4; u++) {
                        const int ra = ca + off + 32 * u, rb = cb + off + 32 * u;
                        if (ra < R) A[ca * PANEL_LD + ra] += sA[u];
                        if (hasB && rb < R) A[cb * PANEL_LD + rb] += sB[u];
                    }
                }
            }
        }
    }
    if (nch > CHOL_MAXCH) {      // descriptors beyond the shared-memory cache: block-wise pull with global map lookups
        __syncthreads();
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
            pull_unit(c, chs, ch0, nch, fr, fc, c6, rmin, acc, CHOL_MAXCH);
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
    // ---- factorisation by 6-column blocks.  Per block column jb two phases:
    //   S   every row below the diagonal block solves x L_D^T = a (one thread per row);
    //   X   warp 0 applies block column jb to the next diagonal block and factors it, while the worker
    //       warps apply block column jb to the rest of block column jb+1 and then the finished columns
    //       [0, 6 jb + 6) to block column jb+2 (look-ahead) -- so the chain of dependent diagonal factors
    //       never waits for the bulk updates.
    // Warps 4, 8, 12 share warp 0's scheduler / FP64 pipe and sit phase X out.
    if constexpr (DC) {
        // Data-flow loop: no barrier inside it, three published counters in shared memory (release / acquire at CTA scope).
        // Roles (16 warps): warp 0 = chain; warps 4, 8, 12 share its scheduler and sit the loop out; of the other twelve,
        // row warps 0..7 own the block rows of the DIAGONAL part (row warp w: block rows w + 1 and w + 9), slab warps 0..3
        // own the rows below it (two fixed 8-row tiles each).  Everything but the 6x6 diagonal blocks is LEFT-looking
        // and lags behind the chain at its own pace:
        //   chain     step r (block row r = 1 .. nt-1), having factored D_(r-1) and published W_(r-1) (pubW = r):
        //               wait preReady[r]; X[r, r-1] = C[r, r-1] W_(r-1)^T; publish "block row r final" (pubL = r);
        //               D_r -= X X^T; factor D_r, W_r = L_D(r)^-1; publish (pubW = r + 1)
        //   row warp  block row r, column blocks t = 0 .. r-2 in turn: C[r, t+1] -= sum_{m < t} X[r, m] L[t+1, m]^T as soon as
        //               rowProg[t+1] >= t says those entries of block row t+1 are final (the long sum, two steps ahead of
        //               its need; rowProg[r] = columns of block row r solved so far); wait pubW > t, X[r, t] = C[r, t] W_t^T; wait pubL > t, the rank-6 rest X[r, t] L[t+1, t]^T.
        //               D_r -= sum_m X[r, m] X[r, m]^T likewise in a long and a rank-6 part; then preReady[r] = 1.
        //   slab warp column block cb: wait pubW > cb, C = A[tile, cb] - sum_{m < cb} X[tile, m] L[cb, m]^T, X = C W_cb^T.
        // W is kept for every block column because row and slab warps may lag several steps.
        // Warps (24): 0 = chain; 4, 8, .., 20 share its scheduler and sit the loop out (work there was measured to slow the chain
        // by half).  Of the other eighteen, fifteen take one block row of the diagonal part each (rows 1 .. 15), three the slab
        // tiles (a tile each, or pairs with stride NWS when a CTA has more: planPanelSlabs).
        static_assert(PANEL3_NW == 24, "warp roles of the data-flow loop");
        const int nTs = (nS + 7) >> 3;
        const int NWS = nTs > 6 ? 6 : 3, NWR = 18 - NWS;     // tall slabs (leaf levels): six slab warps, rows 13-15 double up
        const bool dbgT = CHOL_DIAG(c) && blockIdx.x == 0 && tid == 0 && lev < 64;
        const long long tLoop0 = dbgT ? clock64() : 0;
        // diagnostics (make DIAG=1): clocks of one level (CUGO_B200_DEBUG = 1 + 256 (level + 1); default level 2), CTA 0, per step r,
        // against the loop start: chain [0] pubW = r  [1] preReady[r] seen  [2] pubL = r  [3] D_r factored;
        // [4] the row warp's preReady[r] = 1.  [128..131] slab warps done, [136..143] row warps done.
        const bool trAny = CHOL_DIAG(c) && blockIdx.x == 0 && lev == ((CHOL_DIAG(c) >> 8) ? (CHOL_DIAG(c) >> 8) - 1 : 2);
        if (trAny && tid == 0) { g_cholTrace[126] = R; g_cholTrace[127] = nt; tBaseS = tLoop0; }
        if (trAny) __syncthreads();
        const long long tB = trAny ? tBaseS : 0;
        const int wk = (wid >> 2) * 3 + (wid & 3) - 1;   // 0..11 over the warps with (wid & 3) != 0
        if (wid == 0) {
            const bool trc = trAny && lane == 0;
            panel_chain_factor(A, invAll, &badFlag, 0, lane, Wall[0], &pubW);
            for (int r = 1; r < nt; r++) {
                if (trc && r < 16) g_cholTrace[8 * r] = clock64() - tB;
                wait_counter(&preReady[r], 1, false);
                if (trc && r < 16) g_cholTrace[8 * r + 1] = clock64() - tB;
                panel_chain_step(A, invAll, &badFlag, r, R, lane, Wall, &pubW, &pubL);
                if (trc && r < 16) g_cholTrace[8 * r + 3] = clock64() - tB;
            }
        } else if ((wid & 3) != 0 && wk < NWR) {
            for (int r = wk + 1; r < nt; r += NWR) {
                panel_row_run(A, Wall, r, R, lane, &pubW, &pubL, rowProg, preReady);
                if (trAny && lane == 0 && r < 16) g_cholTrace[8 * r + 4] = clock64() - tB;
            }
            if (trAny && lane == 0) g_cholTrace[136 + wk] = clock64() - tB;
        } else if ((wid & 3) != 0) {
            const int sw = wk - NWR;                     // slab warp: tiles sw, sw + NWS, sw + 2 NWS, ... of the rows from nc, in pairs
            if (sw < nTs) {
                if (nTs <= NWS) {
                    // a tile per slab warp (the upper levels): running target in registers, one solve behind the chain
                    const int rt1[1] = {nc + 8 * sw};
                    panel_slab_run<1>(A, Wall, rt1, nt, R, lane, &pubW, &pubL, rowProg);
                } else {
                    // tall slabs (leaf levels): pairs of tiles through shared memory, left-looking per block column (measured
                    // faster than two or three register-resident tiles per warp, which fall behind the chain)
                    for (int cb = 0; cb < nt; cb++) {
                        wait_counter(&pubL, cb, true);           // the long sum needs block row cb of L only
                        for (int tp = sw; tp < nTs; tp += 2 * NWS) {
                            const bool two = tp + NWS < nTs;
                            panel_slab_tiles(A, Wall[cb], cb, nc + 8 * tp, nc + 8 * (two ? tp + NWS : tp), two, R, lane, 1);
                        }
                        wait_counter(&pubW, cb + 1, true);
                        for (int tp = sw; tp < nTs; tp += 2 * NWS) {
                            const bool two = tp + NWS < nTs;
                            panel_slab_tiles(A, Wall[cb], cb, nc + 8 * tp, nc + 8 * (two ? tp + NWS : tp), two, R, lane, 2);
                        }
                    }
                }
                if (trAny && lane == 0) g_cholTrace[128 + sw] = clock64() - tB;      // slab warp done
            }
        }
        __syncthreads();
        if (dbgT) g_cholDbg[lev * 24 + 15] = clock64() - tLoop0;
    } else {
        const int fr8 = lane >> 2, fk = lane & 3;
        const bool dbgT = CHOL_DIAG(c) && blockIdx.x == 0 && tid == 0 && lev < 64;
        long long tF = 0, tW = 0, tS = 0, tU = 0, t0 = 0, t1 = 0;
        const bool dbgK = (CHOL_DIAG(c) & 2) && blockIdx.x == 0 && tid == 32 && lev < 64;   // a worker warp's view
        long long wU = 0, wL = 0, w0 = 0, w1 = 0;
        const long long tLoop0 = dbgT ? clock64() : 0;
        if (wid == 0) panel_factor_diag(A, invAll, &badFlag, 0, lane);
        __syncthreads();
        constexpr int NWK = 3 * (PANEL3_NW / 4);        // 12 worker warps
        const int wk = (wid >> 2) * 3 + (wid & 3) - 1;   // worker index of warps with (wid & 3) != 0
        for (int jb = 0; jb < nt; jb++) {
            const int j0 = 6 * jb, j1 = j0 + 6, j2 = j1 + 6;
            const bool more = (jb + 1 < nt);
            if (dbgT) { t0 = clock64(); if (lev == 2 && jb < 16) g_cholTrace[4 * jb] = t0 - tLoop0; }
            // phase S
            {
                const int r = j1 + tid + (dbgT ? (int)((unsigned long long)t0 >> 63) : 0);   // (timing: keeps the phase behind the clock read)
                if (r < R) {
                    double x[6], Lc[5][6], iv[6];
    #pragma unroll
                    for (int k = 0; k < 6; k++) x[k] = A[(j0 + k) * PANEL_LD + r];
    #pragma unroll
                    for (int p = 0; p < 5; p++) {
                        const double2* lp = reinterpret_cast<const double2*>(A + (j0 + p) * PANEL_LD + j0);
    #pragma unroll
                        for (int h = 0; h < 3; h++) { const double2 v = lp[h]; Lc[p][2 * h] = v.x; Lc[p][2 * h + 1] = v.y; }
                    }
                    {
                        const double2* ip = reinterpret_cast<const double2*>(invAll + 6 * jb);
    #pragma unroll
                        for (int h = 0; h < 3; h++) { const double2 v = ip[h]; iv[2 * h] = v.x; iv[2 * h + 1] = v.y; }
                    }
    #pragma unroll
                    for (int k = 0; k < 6; k++) {
                        double sv = x[k];
    #pragma unroll
                        for (int p = 0; p < k; p++) sv -= x[p] * Lc[p][k];
                        x[k] = sv * iv[k];
                    }
    #pragma unroll
                    for (int k = 0; k < 6; k++) A[(j0 + k) * PANEL_LD + r] = x[k];
                }
            }
            __syncthreads();
            if (dbgT) { t1 = clock64(); tS += t1 - t0; if (lev == 2 && jb < 16) g_cholTrace[4 * jb + 1] = t1 - tLoop0; }
            if (!more) break;
            // phase X
            if (wid == 0) {
                if (lane < 21) {                          // D_(jb+1) -= L_(jb+1,jb) L_(jb+1,jb)^T, one entry per lane
                    int i = 0;
    #pragma unroll
                    for (int q = 1; q < 6; q++) i += (lane >= tri(q, 0)) ? 1 : 0;
                    const int k = lane - tri(i, 0);
                    double sv = 0.0;
    #pragma unroll
                    for (int p = 0; p < 6; p++) sv += A[(j0 + p) * PANEL_LD + j1 + i] * A[(j0 + p) * PANEL_LD + j1 + k];
                    A[(j1 + k) * PANEL_LD + j1 + i] -= sv;
                }
                __syncwarp();
                panel_factor_diag(A, invAll, &badFlag, jb + 1, lane);
                if (dbgT) { t0 = clock64(); tF += t0 - t1; if (lev == 2 && jb < 16) g_cholTrace[4 * jb + 2] = t0 - tLoop0; }
            } else if ((wid & 3) != 0) {
                // Worker warp: 8-row tiles ta (and tb) of the rows from j2.  All operand and target loads are issued
                // before the first store (stores to the panel would otherwise serialise the next tile's loads):
                //   U: block column jb -> block column jb+1 (rank 6, two DMMAs per tile),
                //   L: finished columns [0, j1) -> block column jb+2 (look-ahead).
                if (dbgK) w0 = clock64();
                const int nmt = (R - j2 + 7) >> 3;          // <= 19 tiles: at most two per worker warp
                const int ta = wk, tb = ta + NWK;
                if (ta < nmt) {
                    const bool two = tb < nmt, la = jb + 2 < nt;
                    const int cl = 2 * fk;
                    const bool st = cl < 6;                 // lanes whose two output columns exist
                    const int ra = j2 + 8 * ta + fr8, rb = j2 + 8 * (two ? tb : ta) + fr8;
                    const int cu = st ? j1 + cl : j1, cv = (st && la) ? j2 + cl : j1;      // valid columns for every lane
                    // ---- U operands (columns j0 + fk and j0 + 4 + fk <= j1 + 1 exist) and targets
                    double bu0 = A[(j0 + fk) * PANEL_LD + j1 + fr8], bu1 = A[(j0 + 4 + fk) * PANEL_LD + j1 + fr8];
                    double au0 = A[(j0 + fk) * PANEL_LD + ra], au1 = A[(j0 + 4 + fk) * PANEL_LD + ra];
                    double qu0 = A[(j0 + fk) * PANEL_LD + rb], qu1 = A[(j0 + 4 + fk) * PANEL_LD + rb];
                    if (fk >= 2) { bu1 = 0.0; au1 = 0.0; qu1 = 0.0; }
                    const double tu0 = A[cu * PANEL_LD + ra], tu1 = A[(cu + 1) * PANEL_LD + ra];
                    const double tv0 = A[cu * PANEL_LD + rb], tv1 = A[(cu + 1) * PANEL_LD + rb];
                    const double tc0 = A[cv * PANEL_LD + ra], tc1 = A[(cv + 1) * PANEL_LD + ra];
                    const double te0 = A[cv * PANEL_LD + rb], te1 = A[(cv + 1) * PANEL_LD + rb];
                    double u0 = 0.0, u1 = 0.0, v0 = 0.0, v1 = 0.0;
                    dmma884(u0, u1, au0, bu0);
                    dmma884(u0, u1, au1, bu1);
                    if (two) {
                        dmma884(v0, v1, qu0, bu0);
                        dmma884(v0, v1, qu1, bu1);
                    }
                    if (dbgK) { w1 = clock64(); wU += w1 - w0; }
                    // ---- L: lane (fr8, fk) reads row (tile base + fr8) of column k0 + fk [+ 4]; pointers advance by
                    // eight columns per iteration, the partial last group is handled after the loop
                    double c00 = 0.0, c01 = 0.0, c10 = 0.0, c11 = 0.0, e00 = 0.0, e01 = 0.0, e10 = 0.0, e11 = 0.0;
                    if (la) {
                        const double* Bp = A + j2 + fr8 + fk * PANEL_LD;
                        const double* Ap = A + ra + fk * PANEL_LD;
                        const double* Aq = A + rb + fk * PANEL_LD;
                        const int nfull = j1 >> 3, rem = j1 & 7;
                        const bool va = fk < rem, vb = fk + 4 < rem;
                        if (two) {
    #pragma unroll 2
                            for (int it = 0; it < nfull; it++) {
                                const double b0 = Bp[0], b1 = Bp[4 * PANEL_LD];
                                const double a0 = Ap[0], a1 = Ap[4 * PANEL_LD];
                                const double q0 = Aq[0], q1 = Aq[4 * PANEL_LD];
                                dmma884(c00, c01, a0, b0);
                                dmma884(c10, c11, a1, b1);
                                dmma884(e00, e01, q0, b0);
                                dmma884(e10, e11, q1, b1);
                                Bp += 8 * PANEL_LD; Ap += 8 * PANEL_LD; Aq += 8 * PANEL_LD;
                            }
                            if (rem) {                  // columns up to j1 + 7 < nc exist: load, then select
                                double b0 = Bp[0], b1 = Bp[4 * PANEL_LD], a0 = Ap[0], a1 = Ap[4 * PANEL_LD];
                                double q0 = Aq[0], q1 = Aq[4 * PANEL_LD];
                                if (!va) { b0 = 0.0; a0 = 0.0; q0 = 0.0; }
                                if (!vb) { b1 = 0.0; a1 = 0.0; q1 = 0.0; }
                                dmma884(c00, c01, a0, b0);
                                dmma884(c10, c11, a1, b1);
                                dmma884(e00, e01, q0, b0);
                                dmma884(e10, e11, q1, b1);
                            }
                        } else {
    #pragma unroll 2
                            for (int it = 0; it < nfull; it++) {
                                const double b0 = Bp[0], b1 = Bp[4 * PANEL_LD];
                                const double a0 = Ap[0], a1 = Ap[4 * PANEL_LD];
                                dmma884(c00, c01, a0, b0);
                                dmma884(c10, c11, a1, b1);
                                Bp += 8 * PANEL_LD; Ap += 8 * PANEL_LD;
                            }
                            if (rem) {
                                double b0 = Bp[0], b1 = Bp[4 * PANEL_LD], a0 = Ap[0], a1 = Ap[4 * PANEL_LD];
                                if (!va) { b0 = 0.0; a0 = 0.0; }
                                if (!vb) { b1 = 0.0; a1 = 0.0; }
                                dmma884(c00, c01, a0, b0);
                                dmma884(c10, c11, a1, b1);
                            }
                        }
                    }
                    // ---- stores
                    if (st && ra < R) {
                        A[(j1 + cl) * PANEL_LD + ra] = tu0 - u0;
                        A[(j1 + cl + 1) * PANEL_LD + ra] = tu1 - u1;
                        if (la) {
                            A[(j2 + cl) * PANEL_LD + ra] = tc0 - (c00 + c10);
                            A[(j2 + cl + 1) * PANEL_LD + ra] = tc1 - (c01 + c11);
                        }
                    }
                    if (two && st && rb < R) {
                        A[(j1 + cl) * PANEL_LD + rb] = tv0 - v0;
                        A[(j1 + cl + 1) * PANEL_LD + rb] = tv1 - v1;
                        if (la) {
                            A[(j2 + cl) * PANEL_LD + rb] = te0 - (e00 + e10);
                            A[(j2 + cl + 1) * PANEL_LD + rb] = te1 - (e01 + e11);
                        }
                    }
                }
                if (dbgK) wL += clock64() - w1;
            }
            __syncthreads();
            if (dbgT) { t1 = clock64(); tW += t1 - t0; if (lev == 2 && jb < 16) g_cholTrace[4 * jb + 3] = t1 - tLoop0; }
        }
        if (dbgK) { g_cholDbg[lev * 24 + 20] = wU; g_cholDbg[lev * 24 + 21] = wL; }
        if (dbgT) {
            tU = clock64() - tLoop0;          // whole factor loop incl. the first diagonal factor
            g_cholDbg[lev * 24 + 12] = tF; g_cholDbg[lev * 24 + 13] = tW;
            g_cholDbg[lev * 24 + 14] = tS; g_cholDbg[lev * 24 + 15] = tU;
        }
    }
    CHOL_STAMP(4);
    // ---- write the panel rows this CTA owns (a warp per column, lanes along the rows)
    double* __restrict__ Pn = c.L + c.Lptr[n.t];
    if (slab == 0) {
        // W = L_D^-1 of every block column for the solves (k_chol_bwd3): one thread per (block, column)
        if (tid < 6 * nt) {
            const int jb = tid / 6, k = tid % 6, j0 = 6 * jb;
            double w[6];
#pragma unroll
            for (int i = 0; i < 6; i++) w[i] = (i == k) ? invAll[6 * jb + i] : 0.0;
#pragma unroll
            for (int i = 1; i < 6; i++) {
                if (i > k) {
                    double sv = 0.0;
#pragma unroll
                    for (int p = 0; p < i; p++) sv += A[(j0 + p) * PANEL_LD + j0 + i] * w[p];   // w[p] = 0 for p < k
                    w[i] = -invAll[6 * jb + i] * sv;
                }
            }
#pragma unroll
            for (int i = 0; i < 6; i++)
                if (i >= k) c.Winv[(size_t)(n.s + jb) * 21 + tri(i, k)] = w[i];
        }
        if (tid == 0 && badFlag) st->fail = 1;
    }
    for (int cc = wid; cc < nc; cc += PANEL3_NW) {
        double* __restrict__ dst = Pn + (size_t)cc * n.m;
        const double* src = A + cc * PANEL_LD;
        if (slab == 0)
            for (int r = cc + lane; r < nc; r += 32) dst[r] = src[r];
        for (int lr = lane; lr < nS; lr += 32) dst[6 * rb0 + lr] = src[nc + lr];
    }
    if (lev == c.nlev - 1 && blockIdx.x == 0 && tid == 0) c.xFlag[c.nsn] += 1;     // a new solve: epoch of the backward flags
    CHOL_STAMP(5);
    CHOL_GT(1);
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
    __shared__ int invS[CHOL_MAXCH * 2 * CHOL_TB];
    const LMState* st = d.st;
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
    double* Cs = sm;                                      // 48 x UPD_CLD, over the operand stage once the MMAs are done
    double* Cs2 = Cs + TS * UPD_CLD;                      // second K half
    const double* __restrict__ Pn = c.L + c.Lptr[n.t];
    const int ch0 = c.childPtr[n.t], nch = c.childPtr[n.t + 1] - ch0;
    CHOL_STAMP(8);
    if (tid < nch && tid < CHOL_MAXCH) chs[tid] = load_child(c, ch0 + tid);
    const int nchS = min(nch, CHOL_MAXCH);
    if (nch > 0) {                       // inverse maps of the tile's 8 row blocks and 8 column blocks (pull_pair)
        __syncthreads();
        for (int i = tid; i < nchS * 2 * CHOL_TB; i += UPD3_TPB) {
            const int k = i % (2 * CHOL_TB);
            const int fb = n.nt + (k < CHOL_TB ? CHOL_TB * ti + k : CHOL_TB * tj + (k - CHOL_TB));
            invS[i] = fb < n.fb ? __ldg(chs[i / (2 * CHOL_TB)].inv + fb) : -1;
        }
    }
    pdl_wait(); pdl_trigger();          // structure reads above; the panel rows below are the predecessor's output
    if (st->done) return;
    CHOL_GT(2);
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
    }
    __syncthreads();            // every warp has read its operands: the stage becomes the C tiles (two CTAs per SM fit)
    if (active) {
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
    constexpr int NU = CHOL_TB * CHOL_TB * 6;
    static_assert(NU <= 2 * UPD3_TPB, "two units per thread");
    {
        PullUnit u[2];
        int o[2];
        double* dst[2];
        bool ok[2];
#pragma unroll
        for (int j = 0; j < 2; j++) {
            const int wu = tid + j * UPD3_TPB;
            const int c6 = wu % 6, item = wu / 6, br = item % CHOL_TB, bc = item / CHOL_TB;
            const int gbr = CHOL_TB * ti + br, gbc = CHOL_TB * tj + bc;
            // block row rt = the rhs row; the rhs column is never used
            ok[j] = wu < NU && !(gbr > rt || gbc > gbr || gbc >= rt);
            u[j].kr = (gbr == rt) ? -1 : br;
            u[j].kc = CHOL_TB + bc;
            u[j].c6 = c6;
            u[j].rmin = (gbr == gbc) ? c6 : 0;
            o[j] = (6 * bc + c6) * UPD_CLD + 6 * br;
            dst[j] = Ut + (size_t)(6 * gbc + c6) * n.u + 6 * gbr;
        }
        if (ok[0] || ok[1]) {
            if (!ok[0]) { u[0] = u[1]; o[0] = o[1]; dst[0] = dst[1]; ok[0] = true; ok[1] = false; }
            double v0[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0}, v1[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
            pull_pair(chs, invS, 2 * CHOL_TB, nchS, u[0], u[1], ok[1], v0, v1);
            if (nch > CHOL_MAXCH) {
                const int g0 = n.nt + CHOL_TB * tj + (u[0].kc - CHOL_TB);
                pull_unit(c, chs, ch0, nch, u[0].kr < 0 ? -1 : n.nt + CHOL_TB * ti + u[0].kr, g0, u[0].c6, u[0].rmin, v0, CHOL_MAXCH);
                if (ok[1]) {
                    const int g1 = n.nt + CHOL_TB * tj + (u[1].kc - CHOL_TB);
                    pull_unit(c, chs, ch0, nch, u[1].kr < 0 ? -1 : n.nt + CHOL_TB * ti + u[1].kr, g1, u[1].c6, u[1].rmin, v1, CHOL_MAXCH);
                }
            }
            if (u[0].kr >= 0) {
#pragma unroll
                for (int rr = 0; rr < 6; rr++)
                    if (rr >= u[0].rmin) dst[0][rr] = v0[rr] - (Cs[o[0] + rr] + Cs2[o[0] + rr]);
            } else {
                dst[0][0] = v0[0] - (Cs[o[0]] + Cs2[o[0]]);
            }
            if (ok[1]) {
                if (u[1].kr >= 0) {
#pragma unroll
                    for (int rr = 0; rr < 6; rr++)
                        if (rr >= u[1].rmin) dst[1][rr] = v1[rr] - (Cs[o[1] + rr] + Cs2[o[1] + rr]);
                } else {
                    dst[1][0] = v1[0] - (Cs[o[1]] + Cs2[o[1]]);
                }
            }
        }
    }
    CHOL_STAMP(11);
    CHOL_GT(3);
}

// The same update for the UPPER tree levels, where a supernode has a few dozen tiles and most SMs idle: 24x24 tiles (four
// times as many CTAs), K dealt to the 8 warps by 4-column steps (warp w: steps w, w + 8, ...), the eight partial tiles summed in
// warp order -- deterministic.  Which levels use it: Solver::planPanelSlabs (tile height per level, Chol::tbLev).
constexpr int UPD_TBS = 4;                  // block rows per small update tile (24 rows)
constexpr int UPD_SCLD = 26;
__global__ void __launch_bounds__(UPD3_TPB) k_chol_update_small(Dev d, Chol c, int lev, int kpadMax) {
    extern __shared__ double sm[];
    __shared__ ChildRef chs[CHOL_MAXCH];
    __shared__ int invS[CHOL_MAXCH * 2 * UPD_TBS];
    const LMState* st = d.st;
    const int q0 = c.levelPtr[lev], q1 = c.levelPtr[lev + 1];
    int q, tile;
    find_task(c.updPtr, q0, q1, c.updPtr[q0] + blockIdx.x, &q, &tile);
    const SnInfo n = sn_info(c, c.levelSn[q]);
    int ti = (int)((sqrtf(8.0f * tile + 1.0f) - 1.0f) * 0.5f);
    while ((ti + 1) * (ti + 2) / 2 <= tile) ti++;
    while (ti * (ti + 1) / 2 > tile) ti--;
    const int tj = tile - ti * (ti + 1) / 2;
    constexpr int TS = 6 * UPD_TBS;
    const int i0 = ti * TS, j0 = tj * TS;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int nc = n.nc, kp = (nc + 3) & ~3;
    double* As = sm;
    double* Bs = (ti == tj) ? sm : sm + kpadMax * UPD_LD;
    double* Cs = sm;                                      // 8 partial tiles of 24 x UPD_SCLD over the operand stage
    const double* __restrict__ Pn = c.L + c.Lptr[n.t];
    const int ch0 = c.childPtr[n.t], nch = c.childPtr[n.t + 1] - ch0;
    if (tid < nch && tid < CHOL_MAXCH) chs[tid] = load_child(c, ch0 + tid);
    const int nchS = min(nch, CHOL_MAXCH);
    if (nch > 0) {
        __syncthreads();
        for (int i = tid; i < nchS * 2 * UPD_TBS; i += UPD3_TPB) {
            const int k = i % (2 * UPD_TBS);
            const int fb = n.nt + (k < UPD_TBS ? UPD_TBS * ti + k : UPD_TBS * tj + (k - UPD_TBS));
            invS[i] = fb < n.fb ? __ldg(chs[i / (2 * UPD_TBS)].inv + fb) : -1;
        }
    }
    pdl_wait(); pdl_trigger();
    if (st->done) return;
    CHOL_GT(2);
    for (int i = tid; i < kp * TS; i += UPD3_TPB) {
        const int r = i % TS, k = i / TS;
        const bool kin = k < nc;
        cp_async8(As + k * UPD_LD + r, Pn + (size_t)(kin ? k : 0) * n.m + nc + i0 + r, kin && i0 + r < n.u);
        if (ti != tj)
            cp_async8(Bs + k * UPD_LD + r, Pn + (size_t)(kin ? k : 0) * n.m + nc + j0 + r, kin && j0 + r < n.u);
    }
    cp_async_wait_all();
    __syncthreads();
    const int fr8 = lane >> 2, fk = lane & 3;
    double acc[3][3][2];
#pragma unroll
    for (int x = 0; x < 3; x++)
#pragma unroll
        for (int y = 0; y < 3; y++) acc[x][y][0] = acc[x][y][1] = 0.0;
    for (int s4 = wid; s4 < (kp >> 2); s4 += UPD3_TPB / 32) {
        double a[3], b[3];
#pragma unroll
        for (int x = 0; x < 3; x++) {
            a[x] = As[(4 * s4 + fk) * UPD_LD + 8 * x + fr8];
            b[x] = Bs[(4 * s4 + fk) * UPD_LD + 8 * x + fr8];
        }
#pragma unroll
        for (int x = 0; x < 3; x++)
#pragma unroll
            for (int y = 0; y < 3; y++) dmma884(acc[x][y][0], acc[x][y][1], a[x], b[y]);
    }
    __syncthreads();            // every warp has read its operands: the stage becomes the partial tiles
    {
        double* Cd = Cs + wid * (TS * UPD_SCLD);
#pragma unroll
        for (int x = 0; x < 3; x++)
#pragma unroll
            for (int y = 0; y < 3; y++)
#pragma unroll
                for (int e = 0; e < 2; e++) Cd[(8 * y + 2 * fk + e) * UPD_SCLD + 8 * x + fr8] = acc[x][y][e];
    }
    __syncthreads();
    // epilogue: add the children's blocks, store the lower triangle of U_t; one (block, column) unit per thread
    double* __restrict__ Ut = c.U + c.Uptr[n.t];
    const int rt = n.rt;
    constexpr int NU = UPD_TBS * UPD_TBS * 6;
    if (tid < NU) {
        const int c6 = tid % 6, item = tid / 6, br = item % UPD_TBS, bc = item / UPD_TBS;
        const int gbr = UPD_TBS * ti + br, gbc = UPD_TBS * tj + bc;
        if (!(gbr > rt || gbc > gbr || gbc >= rt)) {
            PullUnit u0, u1;
            u0.kr = (gbr == rt) ? -1 : br;
            u0.kc = UPD_TBS + bc;
            u0.c6 = c6;
            u0.rmin = (gbr == gbc) ? c6 : 0;
            u1 = u0;
            const int o = (6 * bc + c6) * UPD_SCLD + 6 * br;
            double* dst = Ut + (size_t)(6 * gbc + c6) * n.u + 6 * gbr;
            double v0[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0}, v1[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
            pull_pair(chs, invS, 2 * UPD_TBS, nchS, u0, u1, false, v0, v1);
            if (nch > CHOL_MAXCH)
                pull_unit(c, chs, ch0, nch, u0.kr < 0 ? -1 : n.nt + UPD_TBS * ti + u0.kr, n.nt + UPD_TBS * tj + bc, c6, u0.rmin, v0, CHOL_MAXCH);
            const int nrr = u0.kr >= 0 ? 6 : 1;
#pragma unroll
            for (int rr = 0; rr < 6; rr++) {
                if (rr < nrr && rr >= u0.rmin) {
                    double sv = 0.0;
#pragma unroll
                    for (int w8 = 0; w8 < UPD3_TPB / 32; w8++) sv += Cs[w8 * (TS * UPD_SCLD) + o + rr];
                    dst[rr] = v0[rr] - sv;
                }
            }
        }
    }
    CHOL_GT(3);
}

// ---------------------------------------------------------------------------
// ---------------------------------------------------------------------------
// Z = L11^-1 of every supernode of a level -> c.Zinv (block lower triangle, nc x nc column-major per supernode at
// offset 6 s * 6 MAXW).  Runs on a side branch of the trial graph, concurrently with the rest of the
// factorisation, so that the backward solve of the level is a product with Z^T instead of a 16-step triangular
// solve (k_chol_bwd3).  Z_jj = W_j (the 6x6 inverses the panel kernel left), Z_ij = -W_i sum_k L_ik Z_kj.
template <int TPB>
__global__ void __launch_bounds__(TPB) k_chol_zinv(Dev d, Chol c, int lev) {
    extern __shared__ double sm[];
    const LMState* st = d.st;
    if (st->done) return;
    const SnInfo n = sn_info(c, c.levelSn[c.levelPtr[lev] + blockIdx.x]);
    const int nc = n.nc, nt = n.nt, tid = threadIdx.x, ldL = nc + 1;
    double* L11 = sm;                   // nc columns, leading dimension ldL
    double* Wsm = L11 + nc * ldL;       // 21 nt
    double* T = Wsm + 21 * nt;          // per-warp 6x6 scratch
    double* Zs = T + 36 * (TPB / 32);   // nc columns, leading dimension ldL
    const double* __restrict__ Pn = c.L + c.Lptr[n.t];
    for (int i = tid; i < nc * nc; i += TPB) {
        const int r = i % nc, cc = i / nc;
        if (r > cc) cp_async8(L11 + cc * ldL + r, Pn + (size_t)cc * n.m + r, true);
    }
    for (int i = tid; i < 21 * nt; i += TPB) cp_async8(Wsm + i, c.Winv + (size_t)n.s * 21 + i, true);
    cp_async_wait_all();
    __syncthreads();
    // ---- Z = L11^-1 into its own buffer, block column by block column: Z_jj = W_j, Z_ij = -W_i sum_k L_ik Z_kj.
    // The block columns are independent, so a warp owns two of them (j and nt-1-j: equal work) and only needs
    // __syncwarp; 18 lanes hold a 2x1 piece of the 6x6 block each.
    {
        const int lane = tid & 31, wid = tid >> 5;
        double* Sw = T + 36 * wid;                      // per-warp 6x6 scratch (T holds 6 nc >= 36 * 8 only if nc >= 48: see host)
        const int half = (nt + 1) >> 1;
        for (int pass = 0; pass < 2; pass++) {
            const int j = pass == 0 ? wid : nt - 1 - wid;
            if (pass == 0 ? (wid >= half) : (j < half)) continue;
            if (lane < 21) {
                int p = 0;
#pragma unroll
                for (int q = 1; q < 6; q++) p += (lane >= tri(q, 0)) ? 1 : 0;
                const int q = lane - tri(p, 0);
                Zs[(6 * j + q) * ldL + 6 * j + p] = Wsm[21 * j + lane];
            }
            __syncwarp();
            const int cz = lane / 3, rp = lane - 3 * cz;        // lanes < 18: column cz, rows 2 rp, 2 rp + 1 of the block
            for (int i = j + 1; i < nt; i++) {
                if (lane < 18) {
                    const double* zc = Zs + (6 * j + cz) * ldL;
                    const double* lr = L11 + 6 * i + 2 * rp;
                    double s0a = 0.0, s0b = 0.0, s1a = 0.0, s1b = 0.0;
                    int s2 = 6 * j + cz;
                    for (; s2 + 1 < 6 * i; s2 += 2) {
                        const double z0 = zc[s2], z1 = zc[s2 + 1];
                        s0a += lr[s2 * ldL] * z0; s1a += lr[s2 * ldL + 1] * z0;
                        s0b += lr[(s2 + 1) * ldL] * z1; s1b += lr[(s2 + 1) * ldL + 1] * z1;
                    }
                    if (s2 < 6 * i) { const double z0 = zc[s2]; s0a += lr[s2 * ldL] * z0; s1a += lr[s2 * ldL + 1] * z0; }
                    Sw[cz * 6 + 2 * rp] = s0a + s0b;
                    Sw[cz * 6 + 2 * rp + 1] = s1a + s1b;
                }
                __syncwarp();
                if (lane < 18) {
                    const double* Wb = Wsm + 21 * i;
#pragma unroll
                    for (int e = 0; e < 2; e++) {
                        const int rr = 2 * rp + e;
                        double sv = 0.0;
#pragma unroll
                        for (int k = 0; k < 6; k++)
                            if (k <= rr) sv += Wb[tri(rr, k)] * Sw[cz * 6 + k];
                        Zs[(6 * j + cz) * ldL + 6 * i + rr] = -sv;
                    }
                }
                __syncwarp();
            }
        }
    }
    __syncthreads();
    double* __restrict__ Zg = c.Zinv + (size_t)(6 * n.s) * (6 * CHOL_MAXW);
    for (int i = tid; i < nc * nc; i += TPB) {
        const int r = i % nc, cc = i / nc;
        if (r / 6 > cc / 6 || (r / 6 == cc / 6 && r >= cc)) Zg[(size_t)cc * nc + r] = Zs[cc * ldL + r];
    }
}

// Backward substitution, one CTA per supernode, levels walked top-down: x_t = L11^-T (y_t - L21^T x_R).
// useZ: Z = L11^-1 is in c.Zinv (k_chol_zinv) and the solve is a product with Z^T; otherwise L11 / W are staged
// and warp 0 runs the blocked triangular solve.  The L21^T x_R product issues 4 x CPW independent loads per lane
// before the first use.
template <int TPB>
__global__ void __launch_bounds__(TPB) k_chol_bwd3(Dev d, Chol c, int lev, int early, int useZ, int G) {
    extern __shared__ double sm[];
    __shared__ int shTicket;
    const LMState* st = d.st;
    // `early` (host: every level but the first one solved, under programmatic dependent launch): the predecessor
    // is a backward kernel that signals after its own pdl_wait(), so the whole factorisation has completed and
    // L11 / W can be staged while the predecessor still solves.
    if (!early) pdl_wait();
    // G CTAs share a supernode: each forms the L21^T x_R product over its slice of the rows (a single SM pulls L21 out
    // of L2 at ~25 B/clk -- that product was 5-8 us of every level), the last one to arrive sums the partial vectors
    // in slice order and solves.
    const int q = blockIdx.x / G, g = blockIdx.x - q * G;
    const SnInfo n = sn_info(c, c.levelSn[c.levelPtr[lev] + q]);
    const int nc = n.nc, nt = n.nt, tid = threadIdx.x, nr = 6 * n.rt, ldL = nc + 1;
    const int Ge = max(1, min(G, nr / 64));        // slices worth their fence + ticket: at least 64 rows each
    if (g >= Ge) return;
    const int chunk = (((nr + Ge - 1) / Ge) + 31) & ~31;
    const int rLo = min(nr, g * chunk), rHi = min(nr, rLo + chunk);
    double* L11 = sm;                   // nc columns, leading dimension ldL (odd: conflict-free row reads)
    double* w = L11 + nc * ldL;         // nc
    double* Wsm = w + nc;               // 21 nt
    double* xR = Wsm + 21 * nt;         // 6 rt
    const double* __restrict__ Pn = c.L + c.Lptr[n.t];
    const int* __restrict__ Rt = c.snRows + c.snRowPtr[n.t];
    CHOL_STAMP(16);
    if (useZ) {          // L11 := Z (block lower triangle incl. the diagonal blocks)
        const double* __restrict__ Zg = c.Zinv + (size_t)(6 * n.s) * (6 * CHOL_MAXW);
        for (int i = tid; i < nc * nc; i += TPB) {
            const int r = i % nc, cc = i / nc;
            if (r >= cc) cp_async8(L11 + cc * ldL + r, Zg + (size_t)cc * nc + r, true);
        }
    } else {
        for (int i = tid; i < nc * nc; i += TPB) {
            const int r = i % nc, cc = i / nc;
            if (r > cc) cp_async8(L11 + cc * ldL + r, Pn + (size_t)cc * n.m + r, true);
        }
        for (int i = tid; i < 21 * nt; i += TPB) cp_async8(Wsm + i, c.Winv + (size_t)n.s * 21 + i, true);
    }
    // Under programmatic dependent launch (`early`) the levels below become resident one after the other right away (signal on
    // entry) and do NOT wait for the level above as a whole: the rows of x_R come from the ancestors of this supernode, the
    // far ones were solved levels ago.  Every supernode publishes "x_t stored" (c.xFlag[t] = the solve's epoch, set by the top
    // panel kernel); the product below takes its rows in chunks from the far end and waits per row for the flag of the
    // ancestor that owns it -- so only the parent's rows are left when the parent finishes.  (Safe before pdl_wait: the
    // factorisation completed before the first backward kernel signalled; x is ordered by the flags.)
    if (early) pdl_trigger();
    if (st->done) { cp_async_wait_all(); if (early) pdl_wait(); else pdl_trigger(); return; }
    const int epoch = __ldcg(c.xFlag + c.nsn);
    CHOL_GT(4);
    CHOL_STAMP(17);
    CHOL_GT(7);
    constexpr int NW = TPB / 32;
    constexpr int CPW = (6 * CHOL_MAXW + NW - 1) / NW;      // columns per warp
    const int lane = tid & 31, wid = tid >> 5;
    {
        double acc[CPW];
#pragma unroll
        for (int j = 0; j < CPW; j++) acc[j] = 0.0;
        const int nch = (rHi - rLo + 127) >> 7;
        for (int ch = nch - 1; ch >= 0; ch--) {
            const int r0 = rLo + 128 * ch;
            if (tid < 128 && r0 + tid < rHi) {          // this chunk of x_R, each row once its owner has published
                const int r = r0 + tid, col = __ldg(Rt + r / 6);
                const int* fl = c.xFlag + __ldg(c.colToSn + col);
                int fv;
                do { asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(fv) : "l"(fl) : "memory"); } while (fv != epoch);
                xR[r] = __ldcg(c.xperm + (size_t)col * 6 + r % 6);
            }
            __syncthreads();
            double v[4][CPW], xv[4];
            // unconditional loads (rows / columns beyond the block are clamped and their products vanish): guarded
            // loads compile to branches and would serialise the 4 x CPW loads a lane keeps in flight
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const int r = r0 + 32 * u + lane;
                const bool ok = r < rHi;
                xv[u] = ok ? xR[r] : 0.0;
                const int rc = ok ? r : 0;
#pragma unroll
                for (int j = 0; j < CPW; j++) {
                    const int cc = wid + NW * j;
                    v[u][j] = Pn[(size_t)(cc < nc ? cc : 0) * n.m + nc + rc];
                }
            }
#pragma unroll
            for (int u = 0; u < 4; u++)
#pragma unroll
                for (int j = 0; j < CPW; j++) acc[j] += v[u][j] * xv[u];
        }
        if (early) pdl_wait();          // before the first global write: kernels still complete in launch order
        else pdl_trigger();
#pragma unroll
        for (int j = 0; j < CPW; j++) {
            const int cc = wid + NW * j;
            const double s = warp_sum(acc[j]);
            if (lane == 0 && cc < nc) {
                if (Ge == 1) w[cc] = Pn[(size_t)cc * n.m + n.m - 1] - s;
                else c.bwdPart[((size_t)n.t * BWD_MAXG + g) * (6 * CHOL_MAXW) + cc] = s;
            }
        }
    }
    if (Ge > 1) {
        __threadfence();
        __syncthreads();
        if (tid == 0) shTicket = atomicAdd(c.bwdCnt + n.t, 1);
        __syncthreads();
        if (shTicket != Ge - 1) { cp_async_wait_all(); return; }
        __threadfence();
        for (int cc = tid; cc < nc; cc += TPB) {
            double sv = 0.0;
            for (int g2 = 0; g2 < Ge; g2++) sv += __ldcg(c.bwdPart + ((size_t)n.t * BWD_MAXG + g2) * (6 * CHOL_MAXW) + cc);
            w[cc] = Pn[(size_t)cc * n.m + n.m - 1] - sv;
        }
        if (tid == 0) c.bwdCnt[n.t] = 0;       // ready for the next trial
    }
    cp_async_wait_all();
    __syncthreads();
    CHOL_STAMP(18);
    CHOL_GT(6);
    if (useZ) {          // x = Z^T w: a warp per column, lanes along the rows r >= c; the result goes to xR's place
        double acc[CPW];
#pragma unroll
        for (int j = 0; j < CPW; j++) {
            const int cc = wid + NW * j;
            acc[j] = 0.0;
            if (cc < nc) {
                const double* zc = L11 + cc * ldL;
#pragma unroll
                for (int u = 0; u < (6 * CHOL_MAXW + 31) / 32; u++) {
                    const int r = 32 * u + lane;
                    if (r >= cc && r < nc) acc[j] += zc[r] * w[r];
                }
            }
        }
#pragma unroll
        for (int j = 0; j < CPW; j++) {
            const int cc = wid + NW * j;
            const double sv = warp_sum(acc[j]);
            if (lane == 0 && cc < nc) Wsm[cc] = sv;       // Wsm (21 nt >= nc doubles) is free in this mode
        }
        __syncthreads();
        for (int i = tid; i < nc; i += TPB) w[i] = Wsm[i];
    } else if (wid == 0) {
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
    __threadfence();
    __syncthreads();
    if (tid == 0) asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(c.xFlag + n.t), "r"(epoch) : "memory");   // x_t is out
    CHOL_GT(5);
}

}  // namespace cugo_b200
