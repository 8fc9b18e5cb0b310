// Host-side test harness for csrc/symbolic.h: runs the same symbolic analysis the
// GPU runs (it is host/device-shared code) and a plain CPU emulation of the
// multifrontal data flow the CUDA kernels implement (assemble -> panel factor ->
// update matrix -> backward solve), so the index maps can be validated on a CPU
// box against a dense solve.  TEST-ONLY: never linked into libcugo_b200.so.
#include "../../cuda-bundle-adjustment_b200/csrc/symbolic.h"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

using namespace cugo_b200;

struct Analysis {
    int P = 0, nblk = 0;
    std::vector<int> adjPtr, adj, verts, iperm, mark, level, queue, tmp, stack;
    std::vector<int> snStart, snParent, colToSn, snRowPtr, snLevel, childHead, childNext, snRows;
    std::vector<int> levelPtr, levelSn, childPtr, childIdx, invPtr, inv, snBlkPtr, asmBlk, asmRow, asmCol;
    std::vector<long long> Lptr, Uptr;
    SymWork w;
    SymCounts c;
    SymMaps m;
};

static void analyze(Analysis& A, int P, const int* bRowPtr, const int* bColInd, int leaf, int maxw) {
    A.P = P; A.nblk = bRowPtr[P];
    A.adjPtr.assign(P + 1, 0); A.adj.assign(2 * A.nblk + 1, 0);
    A.verts.assign(P, 0); A.iperm.assign(P, 0); A.mark.assign(P, 0); A.level.assign(P, 0);
    A.queue.assign(P, 0); A.tmp.assign(P + 1, 0); A.stack.assign(4 * P + 8, 0);
    A.snStart.assign(P + 1, 0); A.snParent.assign(P, 0); A.colToSn.assign(P, 0);
    A.snRowPtr.assign(P + 1, 0); A.snLevel.assign(P, 0); A.childHead.assign(P, 0); A.childNext.assign(P, 0);
    const long long cap = sym_rows_cap(P);
    A.snRows.assign(cap, 0);
    SymWork& w = A.w;
    w.P = P; w.bRowPtr = bRowPtr; w.bColInd = bColInd; w.leaf = leaf; w.maxw = maxw;
    w.adjPtr = A.adjPtr.data(); w.adj = A.adj.data(); w.verts = A.verts.data(); w.iperm = A.iperm.data();
    w.mark = A.mark.data(); w.level = A.level.data(); w.queue = A.queue.data(); w.tmp = A.tmp.data();
    w.stack = A.stack.data(); w.snStart = A.snStart.data(); w.snParent = A.snParent.data();
    w.colToSn = A.colToSn.data(); w.snRowPtr = A.snRowPtr.data(); w.snLevel = A.snLevel.data();
    w.childHead = A.childHead.data(); w.childNext = A.childNext.data(); w.snRows = A.snRows.data();
    w.rowsCap = cap;
    sym_analyze(w, &A.c);
    const SymCounts& c = A.c;
    A.levelPtr.assign(c.nlev + 1, 0); A.levelSn.assign(c.nsn + 1, 0);
    A.Lptr.assign(c.nsn + 1, 0); A.Uptr.assign(c.nsn + 1, 0);
    A.childPtr.assign(c.nsn + 1, 0); A.childIdx.assign(c.nsn + 1, 0);
    A.invPtr.assign(c.nsn + 1, 0); A.inv.assign(c.ninv + 1, 0);
    A.snBlkPtr.assign(c.nsn + 1, 0); A.asmBlk.assign(A.nblk + 1, 0);
    A.asmRow.assign(A.nblk + 1, 0); A.asmCol.assign(A.nblk + 1, 0);
    SymMaps& m = A.m;
    m.levelPtr = A.levelPtr.data(); m.levelSn = A.levelSn.data(); m.Lptr = A.Lptr.data(); m.Uptr = A.Uptr.data();
    m.childPtr = A.childPtr.data(); m.childIdx = A.childIdx.data(); m.invPtr = A.invPtr.data(); m.inv = A.inv.data();
    m.snBlkPtr = A.snBlkPtr.data(); m.asmBlk = A.asmBlk.data(); m.asmRow = A.asmRow.data(); m.asmCol = A.asmCol.data();
    sym_fill(w, c, m);
}

// scalar row of child `ch`'s update matrix that lands on scalar row `s` of the
// parent's front (front rows: 6 (n_p + r_p) block rows, then the rhs row), or -1
static inline int child_row(const Analysis& A, int ch, int s, int frontBlocks) {
    const int rc = A.snRowPtr[ch + 1] - A.snRowPtr[ch];
    if (s == 6 * frontBlocks) return 6 * rc;
    const int k = A.inv[A.invPtr[ch] + s / 6];
    return k < 0 ? -1 : 6 * k + s % 6;
}

// Multifrontal Cholesky + solve, emulating the GPU kernels' data flow.
static int mf_solve(const Analysis& A, const double* Hsc, const double* rhs, double* x) {
    const SymCounts& c = A.c;
    std::vector<double> L(c.lsize, 0.0), U(c.usize, 0.0), xp(6 * (size_t)A.P, 0.0);
    for (int lev = 0; lev < c.nlev; lev++)
        for (int q = A.levelPtr[lev]; q < A.levelPtr[lev + 1]; q++) {
            const int t = A.levelSn[q];
            const int s = A.snStart[t], nt = A.snStart[t + 1] - s;
            const int rt = A.snRowPtr[t + 1] - A.snRowPtr[t];
            const int fb = nt + rt, m = 6 * fb + 1, nc = 6 * nt, u = 6 * rt + 1;
            double* Pn = L.data() + A.Lptr[t];
            // --- panel kernel: assemble
            for (int k = A.snBlkPtr[t]; k < A.snBlkPtr[t + 1]; k++) {
                const double* B = Hsc + 36 * (size_t)A.asmBlk[k];
                const int fr = A.asmRow[k], fc = A.asmCol[k] & 0x3fffffff, tr = (A.asmCol[k] >> 30) & 1;
                for (int cc = 0; cc < 6; cc++)
                    for (int rr = 0; rr < 6; rr++)
                        Pn[(size_t)(6 * fc + cc) * m + 6 * fr + rr] = tr ? B[rr * 6 + cc] : B[cc * 6 + rr];
            }
            for (int cc = 0; cc < nc; cc++) Pn[(size_t)cc * m + m - 1] = rhs[6 * A.verts[s + cc / 6] + cc % 6];
            for (int q2 = A.childPtr[t]; q2 < A.childPtr[t + 1]; q2++) {
                const int ch = A.childIdx[q2];
                const int uc = 6 * (A.snRowPtr[ch + 1] - A.snRowPtr[ch]) + 1;
                const double* Uc = U.data() + A.Uptr[ch];
                for (int cc = 0; cc < nc; cc++) {
                    const int ccol = child_row(A, ch, cc, fb);
                    if (ccol < 0) continue;
                    for (int rr = cc; rr < m; rr++) {
                        const int crow = child_row(A, ch, rr, fb);
                        if (crow >= 0) Pn[(size_t)cc * m + rr] += Uc[(size_t)ccol * uc + crow];
                    }
                }
            }
            // --- panel kernel: factor (right-looking by 6-column blocks in the GPU
            // kernel; column by column here, same arithmetic up to summation order)
            for (int j = 0; j < nc; j++) {
                double d = Pn[(size_t)j * m + j];
                if (!(d > 1e-14)) return 0;
                d = std::sqrt(d);
                Pn[(size_t)j * m + j] = d;
                for (int r = j + 1; r < m; r++) Pn[(size_t)j * m + r] /= d;
                for (int k = j + 1; k < nc; k++) {
                    const double f = Pn[(size_t)j * m + k];
                    for (int r = k; r < m; r++) Pn[(size_t)k * m + r] -= f * Pn[(size_t)j * m + r];
                }
            }
            // --- update kernel: U_t = sum_children ext(U_c) - L21 L21^T  (lower, incl. rhs row)
            double* Ut = U.data() + A.Uptr[t];
            for (int cc = 0; cc < u; cc++)
                for (int rr = cc; rr < u; rr++) {
                    double acc = 0;
                    for (int q2 = A.childPtr[t]; q2 < A.childPtr[t + 1]; q2++) {
                        const int ch = A.childIdx[q2];
                        const int uc = 6 * (A.snRowPtr[ch + 1] - A.snRowPtr[ch]) + 1;
                        const int ccol = child_row(A, ch, nc + cc, fb), crow = child_row(A, ch, nc + rr, fb);
                        if (ccol >= 0 && crow >= 0) acc += U[A.Uptr[ch] + (size_t)ccol * uc + crow];
                    }
                    double dot = 0;
                    for (int k = 0; k < nc; k++) dot += Pn[(size_t)k * m + nc + rr] * Pn[(size_t)k * m + nc + cc];
                    Ut[(size_t)cc * u + rr] = acc - dot;
                }
        }
    // backward solve, top-down
    for (int lev = c.nlev - 1; lev >= 0; lev--)
        for (int q = A.levelPtr[lev]; q < A.levelPtr[lev + 1]; q++) {
            const int t = A.levelSn[q];
            const int s = A.snStart[t], nt = A.snStart[t + 1] - s;
            const int rt = A.snRowPtr[t + 1] - A.snRowPtr[t];
            const int m = 6 * (nt + rt) + 1, nc = 6 * nt;
            const double* Pn = L.data() + A.Lptr[t];
            const int* Rt = A.snRows.data() + A.snRowPtr[t];
            for (int j = nc - 1; j >= 0; j--) {
                double v = Pn[(size_t)j * m + m - 1];          // y_j (forward-solved rhs row)
                for (int r = j + 1; r < nc; r++) v -= Pn[(size_t)j * m + r] * xp[6 * (size_t)s + r];
                for (int r = 0; r < 6 * rt; r++) v -= Pn[(size_t)j * m + nc + r] * xp[6 * (size_t)Rt[r / 6] + r % 6];
                xp[6 * (size_t)s + j] = v / Pn[(size_t)j * m + j];
            }
        }
    for (int i = 0; i < A.P; i++)
        for (int k = 0; k < 6; k++) x[6 * A.verts[i] + k] = xp[6 * (size_t)i + k];
    return 1;
}

extern "C" int cugo_sym_test(int P, const int* bRowPtr, const int* bColInd, int leaf, int maxw,
                             const double* Hsc, const double* rhs, double* x, int* perm_out,
                             int* counts_out /* nsn nlev nrows maxFront maxCols err */, double* lsize_usize) {
    Analysis A;
    analyze(A, P, bRowPtr, bColInd, leaf, maxw);
    counts_out[0] = A.c.nsn; counts_out[1] = A.c.nlev; counts_out[2] = A.c.nrows;
    counts_out[3] = A.c.maxFront; counts_out[4] = A.c.maxCols; counts_out[5] = A.c.err;
    lsize_usize[0] = (double)A.c.lsize; lsize_usize[1] = (double)A.c.usize;
    {   // flop count (FMA = 2) and per-level critical path of the panel work
        double fl = 0, crit = 0;
        std::vector<double> lev(A.c.nlev, 0.0);
        for (int t = 0; t < A.c.nsn; t++) {
            const double nc = 6.0 * (A.snStart[t + 1] - A.snStart[t]), r = 6.0 * (A.snRowPtr[t + 1] - A.snRowPtr[t]) + 1;
            const double f = nc * nc * nc / 3 + r * nc * nc + r * r * nc;
            fl += f;
            const double pf = nc * nc * nc / 3 + r * nc * nc;
            if (pf > lev[A.snLevel[t]]) lev[A.snLevel[t]] = pf;
        }
        for (double v : lev) crit += v;
        lsize_usize[2] = fl; lsize_usize[3] = crit;
        if (getenv("CUGO_SYM_VERBOSE"))
            for (int t = 0; t < A.c.nsn; t++)
                fprintf(stderr, "sn %d level %d cols %d rows %d parent %d\n", t, A.snLevel[t],
                        A.snStart[t + 1] - A.snStart[t], A.snRowPtr[t + 1] - A.snRowPtr[t], A.snParent[t]);
    }
    if (A.c.err) return -1;
    memcpy(perm_out, A.verts.data(), sizeof(int) * P);
    return mf_solve(A, Hsc, rhs, x);
}

// The host twin of the device symbolic analysis: every array of symbolic.h for a given pattern, flattened into
// `out` as [count, values...] records in the order
//   verts snStart snRowPtr snRows snParent snLevel levelSn childIdx inv asmBlk asmRow asmCol
// Returns the number of ints written (or -1 if `cap` is too small / the analysis failed).
extern "C" long cugo_sym_twin(int P, const int* bRowPtr, const int* bColInd, int leaf, int maxw, int* out, long cap) {
    Analysis A;
    analyze(A, P, bRowPtr, bColInd, leaf, maxw);
    if (A.c.err) return -1;
    const int nsn = A.c.nsn, nblk = A.nblk;
    const int nchild = A.childPtr[nsn];
    long o = 0;
    auto put = [&](const int* p, long n) {
        if (o + 1 + n > cap) { o = cap + 1; return; }
        out[o++] = (int)n;
        for (long i = 0; i < n; i++) out[o++] = p[i];
    };
    put(A.verts.data(), P); put(A.snStart.data(), nsn + 1); put(A.snRowPtr.data(), nsn + 1); put(A.snRows.data(), A.c.nrows);
    put(A.snParent.data(), nsn); put(A.snLevel.data(), nsn); put(A.levelSn.data(), nsn); put(A.childIdx.data(), nchild);
    put(A.inv.data(), A.c.ninv); put(A.asmBlk.data(), nblk); put(A.asmRow.data(), nblk); put(A.asmCol.data(), nblk);
    return o > cap ? -1 : o;
}
