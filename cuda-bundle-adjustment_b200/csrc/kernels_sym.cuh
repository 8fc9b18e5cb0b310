// kernels_sym.cuh -- the symbolic analysis of the reduced camera system as PARALLEL device work.
//
// Replaces the reference's "Symbolic Decomposition" (METIS nested dissection on the host inside
// cusolverSpXcsrmetisndHost + cusolverSpXcsrcholAnalysis, src/cholesky.hpp:94-101,292-297) and the host loops of
// HschurSparseBlockMatrix (src/sparse_block_matrix.cpp:63-156).  symbolic.h states the algorithm once as plain
// sequential code (it stays the HOST TWIN the device result is compared against bit for bit, and the CPU-testable
// specification); the kernels here compute exactly the same arrays with the parallelism the algorithm has:
//
//   adjacency        one radix sort of the 2 (nblk - P) directed block pairs            (k_sym_pairs, k_sym_adj)
//   nested           the regions of one dissection depth are independent: one CTA, a warp per region, every
//   dissection       breadth-first sweep warp-cooperative (32 neighbours tested per step, appended in adjacency order
//                    by ballot + prefix count = the sequential visit order), adjacency and sweep scratch in shared
//                    memory when they fit                                                (k_sym_nd)
//   row structures   one bitmap of P bits per supernode; a supernode's structure is the OR of its columns' adjacency
//                    and of its children's bitmaps, supernodes of one dissection depth in parallel (a warp each),
//                    depths bottom-up; the sorted row lists fall out of the bitmaps      (k_sym_rows, k_sym_extract)
//   maps             level schedule / offsets / children (O(supernodes), one thread on staged arrays), inverse
//                    relative indices (a warp per supernode), Hsc -> front assembly (a stable sort of the blocks by
//                    destination supernode)                                              (k_sym_tables, k_sym_inv, k_sym_asm_*)
//
// No P^2 workspace: the row lists are sized exactly from the bitmap counts.
#pragma once
#include "engine.h"
#include "symbolic.h"

namespace cugo_b200 {

constexpr int SYM_TPB = 1024;
constexpr int SYM_NW = SYM_TPB / 32;

// ---- adjacency: both directions of every off-diagonal block, sorted by (vertex, neighbour) ----------------------
// Every block row starts with its diagonal block (columns ascending, diagonal present: k_gen_products), so the
// off-diagonal block k of row r is the (k - r - 1)-th off-diagonal block overall.
__global__ void k_sym_pairs(const int* rowOf, const int* colInd, int nblk, unsigned long long P, unsigned long long* keys) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= nblk) return;
    const int r = rowOf[k], c = colInd[k];
    if (r == c) return;
    const size_t j = (size_t)(k - r - 1);
    keys[2 * j] = (unsigned long long)r * P + (unsigned long long)c;
    keys[2 * j + 1] = (unsigned long long)c * P + (unsigned long long)r;
}
__global__ void k_sym_adj(const unsigned long long* keys, int n, unsigned long long P, int* adj, int* vOf) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    adj[i] = (int)(keys[i] % P);
    vOf[i] = (int)(keys[i] / P);
}

struct SymPar {
    int P, leaf, maxw, nadj, useSmem;
    const int* adjPtr;     // [P+1]
    const int* adj;        // [nadj] sorted neighbour lists
    int* verts;            // [P]   out: permutation new -> old
    int* iperm;            // [P]   out: old -> new
    int* isStart;          // [P]   scratch: 1 at the first column of every supernode
    int* snStart;          // [P+1] out
    int* colToSn;          // [P]   out
    int* mark; int* level; int* queue; int* tmp;     // [P] / [P+1] global scratch when shared memory is too small
    int2* regA; int2* regB;                          // [P] region lists of two consecutive depths
    int4* emits;           // [P]   (first column, end column, dissection depth, 0) of every emitted column range
    int* counters;         // [0] emitted ranges [1] supernodes [2] deepest dissection depth
};

// breadth-first sweep of the region tagged `tag`, by one warp; q[0..reached) in visit order
__device__ __forceinline__ int sym_bfs_warp(const int* adjPtr, const int* adj, int* mark, int* level, int* q, int src, int tag,
                                            int* nlev, int lane) {
    int head = 0, tail = 1, maxl = 0;
    if (lane == 0) { q[0] = src; mark[src] = tag + 1; level[src] = 0; }
    __syncwarp();
    while (head < tail) {
        const int v = q[head++];
        const int lv = level[v];
        if (lv > maxl) maxl = lv;
        const int k0 = adjPtr[v], k1 = adjPtr[v + 1];
        for (int kb = k0; kb < k1; kb += 32) {
            const int k = kb + lane;
            const int u = k < k1 ? adj[k] : -1;
            const bool hit = u >= 0 && mark[u] == tag;
            const unsigned bal = __ballot_sync(0xffffffffu, hit);
            if (hit) {
                const int pos = tail + __popc(bal & ((1u << lane) - 1u));
                mark[u] = tag + 1; level[u] = lv + 1; q[pos] = u;
            }
            tail += __popc(bal);
            __syncwarp();
        }
    }
    *nlev = maxl + 1;
    return tail;
}

// Nested dissection (symbolic.h: sym_analyze, first half).  One CTA; the regions of a depth are dealt to the warps.
__global__ void __launch_bounds__(SYM_TPB) k_sym_nd(SymPar p) {
    extern __shared__ int symSh[];
    __shared__ int nCur, nNext, nEmit;
    __shared__ int scanTmp[SYM_TPB];
    const int P = p.P, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int* adjPtr = p.adjPtr;
    const int* adj = p.adj;
    int *mark = p.mark, *level = p.level, *queue = p.queue, *tmp = p.tmp, *verts = p.verts;
    if (p.useSmem) {
        int* sAdj = symSh;
        int* sPtr = sAdj + p.nadj;
        mark = sPtr + (P + 1); level = mark + P; queue = level + P; tmp = queue + P; verts = tmp + (P + 1);
        for (int i = tid; i < p.nadj; i += SYM_TPB) sAdj[i] = p.adj[i];
        for (int i = tid; i <= P; i += SYM_TPB) sPtr[i] = p.adjPtr[i];
        adj = sAdj; adjPtr = sPtr;
    }
    for (int i = tid; i < P; i += SYM_TPB) { verts[i] = i; mark[i] = 0; p.isStart[i] = 0; }
    if (tid == 0) { p.regA[0] = make_int2(0, P); nCur = 1; nNext = 0; nEmit = 0; }
    __syncthreads();
    int2* cur = p.regA;
    int2* nxt = p.regB;
    int depth = 0;
    for (;;) {
        const int nreg = nCur;
        if (nreg == 0) break;
        for (int ri = wid; ri < nreg; ri += SYM_NW) {
            const int2 rg = cur[ri];
            const int lo = rg.x, hi = rg.y, n = hi - lo;
            if (n <= 0) continue;
            int sepLo = lo;
            bool emit = n <= p.leaf;
            if (!emit) {
                const int tag = 8 * depth + 2;
                int* q = queue + lo;
                int* t = tmp + lo;
                for (int i = lo + lane; i < hi; i += 32) mark[verts[i]] = tag;
                __syncwarp();
                int nlev = 0;
                const int reached = sym_bfs_warp(adjPtr, adj, mark, level, q, verts[lo], tag, &nlev, lane);
                if (reached < n) {
                    // disconnected: the reached component first (visit order), then the rest in its old order
                    for (int i = lane; i < reached; i += 32) t[i] = q[i];
                    int a = reached;
                    for (int base = lo; base < hi; base += 32) {
                        const int i = base + lane;
                        const int v = i < hi ? verts[i] : -1;
                        const bool keep = v >= 0 && mark[v] == tag;
                        const unsigned bal = __ballot_sync(0xffffffffu, keep);
                        if (keep) t[a + __popc(bal & ((1u << lane) - 1u))] = v;
                        a += __popc(bal);
                    }
                    __syncwarp();
                    for (int i = lane; i < n; i += 32) verts[lo + i] = t[i];
                    if (lane == 0) {
                        const int s = atomicAdd(&nNext, 2);
                        nxt[s] = make_int2(lo + reached, hi);
                        nxt[s + 1] = make_int2(lo, lo + reached);
                    }
                    __syncwarp();
                    continue;
                }
                // connected: walk to a pseudo-peripheral vertex (at most two more sweeps)
                for (int it = 0; it < 2; it++) {
                    const int far = q[n - 1];
                    const int tg = tag + 2 * (it + 1);
                    __syncwarp();
                    for (int i = lo + lane; i < hi; i += 32) mark[verts[i]] = tg;
                    __syncwarp();
                    int nl2 = 0;
                    sym_bfs_warp(adjPtr, adj, mark, level, q, far, tg, &nl2, lane);
                    const bool better = nl2 > nlev;
                    nlev = nl2;
                    if (!better) break;
                }
                if (nlev <= 2) {
                    emit = true;                   // (nearly) a clique: one dense run of supernodes
                } else {
                    // q is in visit order = grouped by level: t[l] = first position of level l
                    for (int i = lane; i < n; i += 32) {
                        const int l = level[q[i]];
                        if (i == 0 || level[q[i - 1]] != l) t[l] = i;
                    }
                    __syncwarp();
                    int best = -1, bestSize = 0x7fffffff, fb = 1, fbBal = -1;
                    int below = (nlev > 1 ? t[1] : n) - t[0];
                    for (int l = 1; l <= nlev - 2; l++) {
                        const int cl = t[l + 1] - t[l];
                        const int above = n - below - cl;
                        const int bal = below < above ? below : above;
                        if (4 * bal >= (n - cl) && cl < bestSize) { best = l; bestSize = cl; }
                        if (bal > fbBal) { fb = l; fbBal = bal; }
                        below += cl;
                    }
                    if (best < 0) best = fb;
                    const int nA = t[best];
                    const int nS = (best + 1 < nlev ? t[best + 1] : n) - t[best];
                    __syncwarp();
                    for (int i = lane; i < nA; i += 32) verts[lo + i] = q[i];
                    for (int i = nA + nS + lane; i < n; i += 32) verts[lo + i - nS] = q[i];
                    for (int i = nA + lane; i < nA + nS; i += 32) verts[hi - nS + (i - nA)] = q[i];
                    sepLo = hi - nS;
                    if (lane == 0) {
                        const int s = atomicAdd(&nNext, 2);
                        nxt[s] = make_int2(lo + nA, hi - nS);
                        nxt[s + 1] = make_int2(lo, lo + nA);
                    }
                    emit = true;
                }
            }
            if (emit && lane == 0) {
                for (int s2 = sepLo; s2 < hi; s2 += p.maxw) p.isStart[s2] = 1;
                p.emits[atomicAdd(&nEmit, 1)] = make_int4(sepLo, hi, depth, 0);
            }
            __syncwarp();
        }
        __syncthreads();
        if (tid == 0) { nCur = nNext; nNext = 0; }
        int2* sw = cur; cur = nxt; nxt = sw;
        depth++;
        __syncthreads();
    }
    // permutation out, inverse permutation, supernode column ranges (a scan of the start flags)
    for (int i = tid; i < P; i += SYM_TPB) { const int v = verts[i]; p.verts[i] = v; p.iperm[v] = i; }
    const int per = (P + SYM_TPB - 1) / SYM_TPB;
    const int i0 = min(P, tid * per), i1 = min(P, i0 + per);
    int c = 0;
    for (int i = i0; i < i1; i++) c += p.isStart[i];
    scanTmp[tid] = c;
    __syncthreads();
    if (tid == 0) {
        int run = 0;
        for (int i = 0; i < SYM_TPB; i++) { const int v = scanTmp[i]; scanTmp[i] = run; run += v; }
        p.counters[0] = nEmit; p.counters[1] = run; p.counters[2] = depth - 1;
        p.snStart[run] = P;
    }
    __syncthreads();
    int t = scanTmp[tid] - 1;
    for (int i = i0; i < i1; i++) {
        if (p.isStart[i]) { t++; p.snStart[t] = i; }
        p.colToSn[i] = t;
    }
}

struct SymRowsPar {
    int P, maxw, W;        // W = words per bitmap
    const int* adjPtr; const int* adj; const int* verts; const int* iperm; const int* snStart; const int* colToSn;
    const int4* emits; const int* counters;
    unsigned* bits;        // [nsn x W], zero on entry
    int* snParent; int* snLevel; int* childHead; int* childNext; int* rcount;   // [nsn]
    int* snRowPtr;         // [nsn + 1]
    SymCounts* out;
};

// Row structures, parents, levels (symbolic.h: sym_analyze, second half).  One CTA.
__global__ void __launch_bounds__(SYM_TPB) k_sym_rows(SymRowsPar p) {
    __shared__ int scanTmp[SYM_TPB];
    __shared__ unsigned long long sL, sU;
    __shared__ int sMaxFront, sMaxCols, sNlev, sNinv;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int nEmit = p.counters[0], nsn = p.counters[1], maxDepth = p.counters[2], W = p.W;
    for (int t = tid; t < nsn; t += SYM_TPB) { p.snParent[t] = -1; p.snLevel[t] = 0; p.childHead[t] = -1; p.childNext[t] = -1; p.rcount[t] = 0; }
    if (tid == 0) { sL = 0; sU = 0; sMaxFront = 0; sMaxCols = 0; sNlev = 0; sNinv = 0; }
    __syncthreads();
    for (int d = maxDepth; d >= 0; d--) {
        for (int ei = wid; ei < nEmit; ei += SYM_NW) {
            const int4 em = p.emits[ei];
            if (em.z != d) continue;
            for (int s = em.x; s < em.y; s += p.maxw) {      // the column range's supernodes depend on each other in order
                const int e = min(s + p.maxw, em.y);
                const int t = p.colToSn[s];
                unsigned* Bt = p.bits + (size_t)t * W;
                for (int j = s; j < e; j++) {
                    const int v = p.verts[j];
                    for (int k = p.adjPtr[v] + lane; k < p.adjPtr[v + 1]; k += 32) {
                        const int u = p.iperm[p.adj[k]];
                        if (u >= e) atomicOr(&Bt[u >> 5], 1u << (u & 31));
                    }
                }
                __syncwarp();
                const int we = e >> 5;
                const unsigned me = ~((1u << (e & 31)) - 1u);      // bits >= e inside word we
                // (bitmaps, child lists and levels are written by other warps / by atomics: read them past L1)
                for (int ch = __ldcg(p.childHead + t); ch >= 0; ch = __ldcg(p.childNext + ch)) {
                    const unsigned* Bc = p.bits + (size_t)ch * W;
                    for (int w = we + lane; w < W; w += 32) {
                        unsigned b = __ldcg(Bc + w);
                        if (w == we) b &= me;
                        if (b) atomicOr(&Bt[w], b);
                    }
                }
                __syncwarp();
                int cnt = 0, umin = 0x7fffffff;
                for (int w = we + lane; w < W; w += 32) {
                    const unsigned b = __ldcg(Bt + w);
                    if (b) { cnt += __popc(b); umin = min(umin, 32 * w + __ffs(b) - 1); }
                }
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) {
                    cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
                    umin = min(umin, __shfl_xor_sync(0xffffffffu, umin, o));
                }
                if (lane == 0) {
                    p.rcount[t] = cnt;
                    if (cnt > 0) {
                        const int par = p.colToSn[umin];
                        p.snParent[t] = par;
                        p.childNext[t] = atomicExch(&p.childHead[par], t);
                        atomicMax(&p.snLevel[par], __ldcg(p.snLevel + t) + 1);
                    }
                }
                __syncwarp();
            }
        }
        __syncthreads();
    }
    // row pointers (scan of the counts) and the size statistics
    const int per = (nsn + SYM_TPB - 1) / SYM_TPB;
    const int t0 = min(nsn, tid * per), t1 = min(nsn, t0 + per);
    int c = 0;
    for (int t = t0; t < t1; t++) c += p.rcount[t];
    scanTmp[tid] = c;
    __syncthreads();
    if (tid == 0) {
        int run = 0;
        for (int i = 0; i < SYM_TPB; i++) { const int v = scanTmp[i]; scanTmp[i] = run; run += v; }
        p.snRowPtr[nsn] = run;
    }
    __syncthreads();
    int run = scanTmp[tid];
    for (int t = t0; t < t1; t++) {
        p.snRowPtr[t] = run;
        const int r = p.rcount[t];
        run += r;
        const int nt = p.snStart[t + 1] - p.snStart[t];
        atomicMax(&sMaxFront, nt + r); atomicMax(&sMaxCols, nt); atomicMax(&sNlev, p.snLevel[t] + 1);
        const unsigned long long m = 6ULL * (nt + r) + 1, u = 6ULL * r + 1;
        atomicAdd(&sL, m * 6ULL * nt); atomicAdd(&sU, u * u);
        const int par = p.snParent[t];
        if (par >= 0) atomicAdd(&sNinv, (p.snStart[par + 1] - p.snStart[par]) + p.rcount[par]);
    }
    __syncthreads();
    if (tid == 0) {
        SymCounts o;
        o.nsn = nsn; o.nlev = sNlev; o.nrows = p.snRowPtr[nsn]; o.ninv = sNinv; o.maxFront = sMaxFront; o.maxCols = sMaxCols;
        o.lsize = (long long)sL; o.usize = (long long)sU; o.err = 0;
        *p.out = o;
    }
}

// sorted row lists out of the bitmaps: a warp per supernode
__global__ void k_sym_extract(const unsigned* bits, int W, int nsn, const int* snRowPtr, int* snRows) {
    const int t = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (t >= nsn) return;
    const unsigned* Bt = bits + (size_t)t * W;
    int base = snRowPtr[t];
    for (int w0 = 0; w0 < W; w0 += 32) {
        const int w = w0 + lane;
        unsigned b = w < W ? __ldcg(Bt + w) : 0u;
        const int c = __popc(b);
        int pre = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(0xffffffffu, pre, o); if (lane >= o) pre += v; }
        int pos = base + pre - c;
        while (b) { const int bit = __ffs(b) - 1; snRows[pos++] = 32 * w + bit; b &= b - 1; }
        base += __shfl_sync(0xffffffffu, pre, 31);
    }
}

// Level schedule, storage offsets, children (ascending), inverse-map offsets and the task tables of the numeric kernels
// (symbolic.h: sym_fill, first half, and the loop k_sym_fill ran): all O(supernodes); one thread on arrays staged in
// shared memory by the whole CTA.
__global__ void __launch_bounds__(SYM_TPB) k_sym_tables(int nsn, int nlev, const int* snStart, const int* snRowPtr, const int* snParent,
                                                        const int* snLevel, SymMaps m, int* trsmPtr, int* updPtr, int rb, int tb,
                                                        int* levMaxNt, double* factStats, long long lsize) {
    extern __shared__ int sh[];
    int* nt_ = sh;                 // [nsn] block columns
    int* rt_ = nt_ + nsn;          // [nsn] block rows below
    int* par_ = rt_ + nsn;         // [nsn]
    int* lev_ = par_ + nsn;        // [nsn]
    int* tmp_ = lev_ + nsn;        // [nsn + nlev + 2]
    for (int t = threadIdx.x; t < nsn; t += SYM_TPB) {
        nt_[t] = snStart[t + 1] - snStart[t]; rt_[t] = snRowPtr[t + 1] - snRowPtr[t]; par_[t] = snParent[t]; lev_[t] = snLevel[t];
    }
    __syncthreads();
    if (threadIdx.x != 0) return;
    // level schedule: supernodes grouped by level, ascending id inside
    for (int l = 0; l <= nlev; l++) m.levelPtr[l] = 0;
    for (int t = 0; t < nsn; t++) m.levelPtr[lev_[t] + 1]++;
    for (int l = 0; l < nlev; l++) m.levelPtr[l + 1] += m.levelPtr[l];
    for (int l = 0; l < nlev; l++) tmp_[l] = m.levelPtr[l];
    for (int t = 0; t < nsn; t++) m.levelSn[tmp_[lev_[t]]++] = t;
    // storage offsets, inverse-map offsets
    long long lo = 0, uo = 0;
    int ninv = 0;
    for (int t = 0; t < nsn; t++) {
        m.Lptr[t] = lo; m.Uptr[t] = uo;
        lo += (6LL * (nt_[t] + rt_[t]) + 1) * 6 * nt_[t];
        uo += (6LL * rt_[t] + 1) * (6LL * rt_[t] + 1);
        const int pp = par_[t];
        if (pp < 0) m.invPtr[t] = -1;
        else { m.invPtr[t] = ninv; ninv += nt_[pp] + rt_[pp]; }
    }
    m.Lptr[nsn] = lo; m.Uptr[nsn] = uo;
    // children of every supernode in ascending order
    for (int t = 0; t <= nsn; t++) m.childPtr[t] = 0;
    for (int t = 0; t < nsn; t++) if (par_[t] >= 0) m.childPtr[par_[t] + 1]++;
    for (int t = 0; t < nsn; t++) m.childPtr[t + 1] += m.childPtr[t];
    for (int t = 0; t < nsn; t++) tmp_[t] = m.childPtr[t];
    for (int t = 0; t < nsn; t++) if (par_[t] >= 0) m.childIdx[tmp_[par_[t]]++] = t;
    // task tables in levelSn order + algorithmic flops of one factorisation and solve
    int a = 0, b = 0;
    double fPanel = 0.0, fUpd = 0.0, fBwd = 0.0;
    for (int l = 0; l < nlev; l++) levMaxNt[l] = 0;
    for (int q = 0; q < nsn; q++) {
        const int t = m.levelSn[q];
        const int rt = rt_[t], nt = nt_[t];
        if (nt > levMaxNt[lev_[t]]) levMaxNt[lev_[t]] = nt;
        trsmPtr[q] = a; updPtr[q] = b;
        const double ncd = 6.0 * nt, rd = 6.0 * rt + 1.0;
        fPanel += ncd * ncd * ncd / 3.0 + rd * ncd * ncd;
        fUpd += rt > 0 ? rd * rd * ncd : 0.0;
        fBwd += ncd * ncd + 2.0 * 6.0 * rt * ncd;
        const int ns = (rt + rb - 1) / rb;
        a += ns > 0 ? ns : 1;
        if (rt > 0) { const int n2 = (rt + tb) / tb; b += n2 * (n2 + 1) / 2; }
    }
    trsmPtr[nsn] = a; updPtr[nsn] = b;
    factStats[0] = fPanel; factStats[1] = fUpd; factStats[2] = fBwd; factStats[3] = (double)lsize;
}

// inverse relative-index maps (symbolic.h: sym_fill): for child t of p, inv[f] = k  <=>  R_t[k] is front block f of p.
// A warp per supernode.
__global__ void k_sym_inv(int nsn, const int* snStart, const int* snRowPtr, const int* snRows, const int* snParent,
                          const int* invPtr, int* invAll) {
    const int t = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (t >= nsn) return;
    const int pp = snParent[t];
    if (pp < 0) return;
    const int sp = snStart[pp], np = snStart[pp + 1] - sp;
    const int* Rp = snRows + snRowPtr[pp];
    const int rp = snRowPtr[pp + 1] - snRowPtr[pp];
    int* inv = invAll + invPtr[t];
    for (int f = lane; f < np + rp; f += 32) inv[f] = -1;
    __syncwarp();
    const int* Rt = snRows + snRowPtr[t];
    const int rt = snRowPtr[t + 1] - snRowPtr[t];
    for (int k = lane; k < rt; k += 32) {
        const int u = Rt[k];
        inv[(u < sp + np) ? (u - sp) : np + sym_lower_bound(Rp, rp, u)] = k;
    }
}

// assembly of the Hsc blocks into the fronts: key = destination supernode of block k (stable sort keeps block order)
__global__ void k_sym_asm_keys(const int* rowOf, const int* colInd, int nblk, const int* iperm, const int* colToSn, int* keys,
                               int* vals) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= nblk) return;
    const int a = iperm[rowOf[k]], b = iperm[colInd[k]];
    keys[k] = colToSn[a < b ? a : b];
    vals[k] = k;
}
__global__ void k_sym_asm_fill(const int* sortedSn, const int* sortedBlk, int nblk, const int* rowOf, const int* colInd,
                               const int* iperm, const int* snStart, const int* snRowPtr, const int* snRows, int* asmBlk,
                               int* asmRow, int* asmCol) {
    const int pos = blockIdx.x * blockDim.x + threadIdx.x;
    if (pos >= nblk) return;
    const int k = sortedBlk[pos], t = sortedSn[pos];
    const int a = iperm[rowOf[k]], b = iperm[colInd[k]];
    const int lo2 = a < b ? a : b, hi2 = a < b ? b : a;
    const int s = snStart[t], nt = snStart[t + 1] - s;
    const int* Rt = snRows + snRowPtr[t];
    const int rt = snRowPtr[t + 1] - snRowPtr[t];
    asmBlk[pos] = k;
    asmRow[pos] = (hi2 < s + nt) ? (hi2 - s) : nt + sym_lower_bound(Rt, rt, hi2);
    // stored block is H(r, col), r <= col (original ids); the front holds the lower triangle in permuted order:
    // entry (hi2, lo2).  If the row vertex is the later one (a > b) the block is used as is, else transposed.
    asmCol[pos] = (lo2 - s) | ((a < b) ? (1 << 30) : 0);
}

}  // namespace cugo_b200
