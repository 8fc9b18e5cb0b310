// symbolic.h -- symbolic analysis of the reduced camera system Hsc (block level).
//
// Replaces the reference's stage "Symbolic Decomposition":
// cusolverSpXcsrmetisndHost on the 6P x 6P scalar pattern (src/cholesky.hpp:292-297)
// + cusolverSpXcsrcholAnalysis (src/cholesky.hpp:94-101).  Here the analysis runs on
// the P-node *block* graph (36x smaller), produces a nested-dissection ordering,
// a supernode partition (ND separators / leaves, split into panels of <= maxw block
// columns), each supernode's row structure, the assembly tree with its level
// schedule and the index maps the numeric multifrontal kernels use.
//
// The code is plain sequential integer work written once and compiled twice:
// as device code (run by one thread of a kernel, so the pattern never leaves the
// GPU) and as host code (unit tests on CPU boxes, and the CPU twin the GPU result
// is compared against bit for bit).
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define CUGO_HD __host__ __device__
#else
#define CUGO_HD
#endif

namespace cugo_b200 {

struct SymCounts {
    int nsn;          // supernodes
    int nlev;         // levels of the assembly tree
    int nrows;        // sum of row-structure lengths (blocks)
    int ninv;         // sum over non-root supernodes of parent's front size (blocks)
    int maxFront;     // max (n_t + r_t) in blocks
    int maxCols;      // max n_t
    long long lsize;  // doubles in L panel storage
    long long usize;  // doubles in update-matrix storage
    int err;          // 0 ok
};

// Worst-case scratch sizes (ints) for a P-node graph with nadj adjacency entries.
CUGO_HD inline long long sym_rows_cap(int P) { return (long long)P * (P + 1) / 2 + P; }

struct SymWork {
    // inputs: upper BSR of Hsc (rows ascending, cols ascending, diagonal present)
    int P;
    const int* bRowPtr;
    const int* bColInd;
    int leaf;   // ND recursion stops at regions of <= leaf vertices
    int maxw;   // max block columns per supernode (panel width)
    // scratch (caller allocated)
    int* adjPtr;    // [P+1]
    int* adj;       // [2*(nblk-P)] (<= 2*nblk)
    int* verts;     // [P]   becomes perm (new -> old)
    int* iperm;     // [P]   old -> new
    int* mark;      // [P]
    int* level;     // [P]
    int* queue;     // [P]
    int* tmp;       // [P]
    int* stack;     // [4*P + 8]
    // outputs sized by P
    int* snStart;   // [P+1]
    int* snParent;  // [P]
    int* colToSn;   // [P]
    int* snRowPtr;  // [P+1]
    int* snLevel;   // [P]
    int* childHead; // [P]
    int* childNext; // [P]
    // output with worst-case capacity
    int* snRows;    // [rowsCap]
    long long rowsCap;
};

// ---- small helpers ---------------------------------------------------------
CUGO_HD inline void sym_sort_ints(int* a, int n) {
    // heap sort: O(n log n), no recursion, no extra memory
    if (n < 2) return;
    for (int start = n / 2 - 1; start >= 0; start--) {
        int root = start;
        for (;;) {
            int child = 2 * root + 1;
            if (child >= n) break;
            if (child + 1 < n && a[child] < a[child + 1]) child++;
            if (a[root] >= a[child]) break;
            int t = a[root]; a[root] = a[child]; a[child] = t;
            root = child;
        }
    }
    for (int end = n - 1; end > 0; end--) {
        int t = a[0]; a[0] = a[end]; a[end] = t;
        int root = 0;
        for (;;) {
            int child = 2 * root + 1;
            if (child >= end) break;
            if (child + 1 < end && a[child] < a[child + 1]) child++;
            if (a[root] >= a[child]) break;
            int t2 = a[root]; a[root] = a[child]; a[child] = t2;
            root = child;
        }
    }
}

CUGO_HD inline int sym_lower_bound(const int* a, int n, int v) {
    int lo = 0, hi = n;
    while (lo < hi) {
        int mid = (lo + hi) >> 1;
        if (a[mid] < v) lo = mid + 1; else hi = mid;
    }
    return lo;
}

// BFS inside the region tagged `tag` in mark[], from `src`.  Fills queue[0..n)
// in visit order and level[v]; returns the number of vertices reached and the
// number of levels through *nlev.  Visited vertices get mark = tag+1.
CUGO_HD inline int sym_bfs(const SymWork& w, int src, int tag, int* nlev) {
    int head = 0, tail = 0;
    w.queue[tail++] = src;
    w.mark[src] = tag + 1;
    w.level[src] = 0;
    int maxl = 0;
    while (head < tail) {
        const int v = w.queue[head++];
        const int lv = w.level[v];
        if (lv > maxl) maxl = lv;
        for (int k = w.adjPtr[v]; k < w.adjPtr[v + 1]; k++) {
            const int u = w.adj[k];
            if (w.mark[u] == tag) {
                w.mark[u] = tag + 1;
                w.level[u] = lv + 1;
                w.queue[tail++] = u;
            }
        }
    }
    *nlev = maxl + 1;
    return tail;
}

// ---- phase 1: ordering, supernodes, row structures, assembly tree ----------
CUGO_HD inline void sym_analyze(const SymWork& w, SymCounts* out) {
    const int P = w.P;
    SymCounts c;
    c.nsn = c.nlev = c.nrows = c.ninv = c.maxFront = c.maxCols = 0;
    c.lsize = c.usize = 0; c.err = 0;
    if (P == 0) { *out = c; return; }

    // symmetric adjacency (no self loops) from the upper BSR
    for (int i = 0; i <= P; i++) w.adjPtr[i] = 0;
    for (int r = 0; r < P; r++)
        for (int k = w.bRowPtr[r]; k < w.bRowPtr[r + 1]; k++) {
            const int cc = w.bColInd[k];
            if (cc != r) { w.adjPtr[r + 1]++; w.adjPtr[cc + 1]++; }
        }
    for (int i = 0; i < P; i++) w.adjPtr[i + 1] += w.adjPtr[i];
    for (int i = 0; i < P; i++) w.tmp[i] = w.adjPtr[i];
    for (int r = 0; r < P; r++)
        for (int k = w.bRowPtr[r]; k < w.bRowPtr[r + 1]; k++) {
            const int cc = w.bColInd[k];
            if (cc != r) { w.adj[w.tmp[r]++] = cc; w.adj[w.tmp[cc]++] = r; }
        }

    // nested dissection by BFS level structures; verts[] becomes the permutation
    int* isStart = w.colToSn;
    for (int i = 0; i < P; i++) { w.verts[i] = i; w.mark[i] = 0; isStart[i] = 0; }
    int tag = 2;
    int sp = 0;
    w.stack[sp++] = 0; w.stack[sp++] = P;
    while (sp > 0) {
        const int hi = w.stack[--sp];
        const int lo = w.stack[--sp];
        const int n = hi - lo;
        if (n <= 0) continue;
        bool emit = (n <= w.leaf);
        int sepLo = lo;   // start of the range emitted as a supernode
        if (!emit) {
            for (int i = lo; i < hi; i++) w.mark[w.verts[i]] = tag;
            int nlev = 0;
            int reached = sym_bfs(w, w.verts[lo], tag, &nlev);
            tag += 2;
            if (reached < n) {
                // disconnected: peel off the reached component, keep the rest
                int a = 0;
                for (int i = 0; i < reached; i++) w.tmp[a++] = w.queue[i];
                for (int i = lo; i < hi; i++)
                    if (w.mark[w.verts[i]] == tag - 2) w.tmp[a++] = w.verts[i];
                for (int i = 0; i < n; i++) w.verts[lo + i] = w.tmp[i];
                w.stack[sp++] = lo + reached; w.stack[sp++] = hi;
                w.stack[sp++] = lo; w.stack[sp++] = lo + reached;
                continue;
            }
            // connected: walk to a pseudo-peripheral vertex (two more sweeps)
            for (int it = 0; it < 2; it++) {
                const int far = w.queue[n - 1];
                for (int i = lo; i < hi; i++) w.mark[w.verts[i]] = tag;
                int nl2 = 0;
                sym_bfs(w, far, tag, &nl2);
                tag += 2;
                const bool better = nl2 > nlev;
                nlev = nl2;
                if (!better) break;
            }
            if (nlev <= 2) {
                emit = true;           // (nearly) a clique: one dense supernode
            } else {
                int* cnt = w.tmp;      // per-level counts
                for (int l = 0; l < nlev; l++) cnt[l] = 0;
                for (int i = 0; i < n; i++) cnt[w.level[w.queue[i]]]++;
                // separator = the smallest level that leaves >= 25 % of the rest on
                // each side; if no level balances that well, the best-balanced one
                int best = -1, bestSize = 0x7fffffff, fb = 1, fbBal = -1;
                int below = cnt[0];
                for (int l = 1; l <= nlev - 2; l++) {
                    const int above = n - below - cnt[l];
                    const int bal = below < above ? below : above;
                    if (4 * bal >= (n - cnt[l]) && cnt[l] < bestSize) { best = l; bestSize = cnt[l]; }
                    if (bal > fbBal) { fb = l; fbBal = bal; }
                    below += cnt[l];
                }
                if (best < 0) best = fb;
                // queue[] is in BFS order => already grouped by level
                int nA = 0;
                for (int l = 0; l < best; l++) nA += cnt[l];
                const int nS = cnt[best];
                int a = lo;
                for (int i = 0; i < nA; i++) w.verts[a++] = w.queue[i];
                for (int i = nA + nS; i < n; i++) w.verts[a++] = w.queue[i];
                for (int i = nA; i < nA + nS; i++) w.verts[a++] = w.queue[i];
                sepLo = hi - nS;
                w.stack[sp++] = lo + nA; w.stack[sp++] = hi - nS;   // B
                w.stack[sp++] = lo; w.stack[sp++] = lo + nA;        // A
                emit = true;
            }
        }
        if (emit)
            for (int s = sepLo; s < hi; s += w.maxw) isStart[s] = 1;
    }
    for (int i = 0; i < P; i++) w.iperm[w.verts[i]] = i;

    // supernode column ranges in ascending order
    int nsn = 0;
    for (int i = 0; i < P; i++) if (isStart[i]) w.snStart[nsn++] = i;
    w.snStart[nsn] = P;
    for (int t = 0; t < nsn; t++) {
        for (int j = w.snStart[t]; j < w.snStart[t + 1]; j++) w.colToSn[j] = t;
        w.childHead[t] = -1; w.childNext[t] = -1; w.snParent[t] = -1; w.snLevel[t] = 0;
        const int nt = w.snStart[t + 1] - w.snStart[t];
        if (nt > c.maxCols) c.maxCols = nt;
    }

    // row structures bottom-up (children always have smaller ids)
    for (int i = 0; i < P; i++) w.mark[i] = -1;
    long long nrows = 0;
    w.snRowPtr[0] = 0;
    for (int t = 0; t < nsn; t++) {
        const int s = w.snStart[t], e = w.snStart[t + 1];
        int* rows = w.snRows + nrows;
        int r = 0;
        for (int j = s; j < e; j++) {
            const int v = w.verts[j];
            for (int k = w.adjPtr[v]; k < w.adjPtr[v + 1]; k++) {
                const int u = w.iperm[w.adj[k]];
                if (u >= e && w.mark[u] != t) {
                    if (nrows + r >= w.rowsCap) { c.err = 1; *out = c; return; }
                    w.mark[u] = t; rows[r++] = u;
                }
            }
        }
        for (int ch = w.childHead[t]; ch >= 0; ch = w.childNext[ch])
            for (int k = w.snRowPtr[ch]; k < w.snRowPtr[ch + 1]; k++) {
                const int u = w.snRows[k];
                if (u >= e && w.mark[u] != t) {
                    if (nrows + r >= w.rowsCap) { c.err = 1; *out = c; return; }
                    w.mark[u] = t; rows[r++] = u;
                }
            }
        sym_sort_ints(rows, r);
        nrows += r;
        w.snRowPtr[t + 1] = (int)nrows;
        if (r > 0) {
            const int p = w.colToSn[rows[0]];
            w.snParent[t] = p;
            // append to the parent's child list keeping ascending child order
            if (w.childHead[p] < 0) w.childHead[p] = t;
            else { int q = w.childHead[p]; while (w.childNext[q] >= 0) q = w.childNext[q]; w.childNext[q] = t; }
            if (w.snLevel[t] + 1 > w.snLevel[p]) w.snLevel[p] = w.snLevel[t] + 1;
            c.ninv += (w.snStart[p + 1] - w.snStart[p]);   // + r_p added below
        }
        const int nt = e - s;
        if (nt + r > c.maxFront) c.maxFront = nt + r;
        const long long m = 6LL * (nt + r) + 1, u = 6LL * r + 1;
        c.lsize += m * 6 * nt;
        c.usize += u * u;
    }
    for (int t = 0; t < nsn; t++) {
        if (w.snLevel[t] + 1 > c.nlev) c.nlev = w.snLevel[t] + 1;
        const int p = w.snParent[t];
        if (p >= 0) c.ninv += w.snRowPtr[p + 1] - w.snRowPtr[p];
    }
    c.nsn = nsn;
    c.nrows = (int)nrows;
    *out = c;
}

// ---- phase 2: schedule, storage offsets and index maps (exact-size arrays) ---
struct SymMaps {
    int* levelPtr;        // [nlev+1]
    int* levelSn;         // [nsn]   supernodes grouped by level (ascending id inside)
    long long* Lptr;      // [nsn+1] panel offsets (doubles); panel t is m_t x 6 n_t,
                          //         column-major, ld = m_t = 6 (n_t + r_t) + 1
    long long* Uptr;      // [nsn+1] update-matrix offsets; u_t x u_t, u_t = 6 r_t + 1
    int* childPtr;        // [nsn+1]
    int* childIdx;        // [#non-root]
    int* invPtr;          // [nsn]   offset into inv[] for supernode c as a child (-1 for roots)
    int* inv;             // [ninv]  front block index of parent -> row index in R_c, or -1
    int* snBlkPtr;        // [nsn+1]
    int* asmBlk;          // [nblk]  BSR block ids grouped by destination supernode
    int* asmRow;          // [nblk]  destination block row in the front
    int* asmCol;          // [nblk]  destination block column; bit 30 set => transpose
};

CUGO_HD inline void sym_fill(const SymWork& w, const SymCounts& c, const SymMaps& m) {
    const int nsn = c.nsn, P = w.P;
    // level schedule
    for (int l = 0; l <= c.nlev; l++) m.levelPtr[l] = 0;
    for (int t = 0; t < nsn; t++) m.levelPtr[w.snLevel[t] + 1]++;
    for (int l = 0; l < c.nlev; l++) m.levelPtr[l + 1] += m.levelPtr[l];
    for (int l = 0; l < c.nlev; l++) w.tmp[l] = m.levelPtr[l];
    for (int t = 0; t < nsn; t++) m.levelSn[w.tmp[w.snLevel[t]]++] = t;
    // storage offsets, children
    long long lo = 0, uo = 0;
    int nch = 0, ninv = 0;
    m.childPtr[0] = 0;
    for (int t = 0; t < nsn; t++) {
        const int nt = w.snStart[t + 1] - w.snStart[t];
        const int rt = w.snRowPtr[t + 1] - w.snRowPtr[t];
        m.Lptr[t] = lo; m.Uptr[t] = uo;
        lo += (6LL * (nt + rt) + 1) * 6 * nt;
        uo += (6LL * rt + 1) * (6LL * rt + 1);
        for (int ch = w.childHead[t]; ch >= 0; ch = w.childNext[ch]) m.childIdx[nch++] = ch;
        m.childPtr[t + 1] = nch;
    }
    m.Lptr[nsn] = lo; m.Uptr[nsn] = uo;
    // inverse relative-index maps: for child ch of p, inv[f] = k  <=>  R_ch[k] is
    // front block f of p (f < n_p: column snStart[p]+f; else R_p[f - n_p])
    for (int t = 0; t < nsn; t++) {
        const int p = w.snParent[t];
        if (p < 0) { m.invPtr[t] = -1; continue; }
        const int sp = w.snStart[p], np = w.snStart[p + 1] - sp;
        const int* Rp = w.snRows + w.snRowPtr[p];
        const int rp = w.snRowPtr[p + 1] - w.snRowPtr[p];
        int* inv = m.inv + ninv;
        m.invPtr[t] = ninv;
        for (int f = 0; f < np + rp; f++) inv[f] = -1;
        const int* Rt = w.snRows + w.snRowPtr[t];
        const int rt = w.snRowPtr[t + 1] - w.snRowPtr[t];
        for (int k = 0; k < rt; k++) {
            const int u = Rt[k];
            const int f = (u < sp + np) ? (u - sp) : np + sym_lower_bound(Rp, rp, u);
            inv[f] = k;
        }
        ninv += np + rp;
    }
    // assembly of the BSR blocks of Hsc into the fronts
    for (int t = 0; t <= nsn; t++) m.snBlkPtr[t] = 0;
    for (int r = 0; r < P; r++)
        for (int k = w.bRowPtr[r]; k < w.bRowPtr[r + 1]; k++) {
            const int a = w.iperm[r], b = w.iperm[w.bColInd[k]];
            m.snBlkPtr[w.colToSn[a < b ? a : b] + 1]++;
        }
    for (int t = 0; t < nsn; t++) m.snBlkPtr[t + 1] += m.snBlkPtr[t];
    for (int t = 0; t < nsn; t++) w.tmp[t] = m.snBlkPtr[t];
    for (int r = 0; r < P; r++)
        for (int k = w.bRowPtr[r]; k < w.bRowPtr[r + 1]; k++) {
            const int a = w.iperm[r], b = w.iperm[w.bColInd[k]];
            const int lo2 = a < b ? a : b, hi2 = a < b ? b : a;
            const int t = w.colToSn[lo2];
            const int s = w.snStart[t], nt = w.snStart[t + 1] - s;
            const int* Rt = w.snRows + w.snRowPtr[t];
            const int rt = w.snRowPtr[t + 1] - w.snRowPtr[t];
            const int frow = (hi2 < s + nt) ? (hi2 - s) : nt + sym_lower_bound(Rt, rt, hi2);
            const int pos = w.tmp[t]++;
            m.asmBlk[pos] = k;
            m.asmRow[pos] = frow;
            // stored block is H(r, col) with r <= col (original ids).  The front holds
            // the lower triangle in permuted order: entry (hi2, lo2).  If the row
            // vertex r is the later one (a > b) the block is used as is, else transposed.
            m.asmCol[pos] = (lo2 - s) | ((a < b) ? (1 << 30) : 0);
        }
}

}  // namespace cugo_b200
