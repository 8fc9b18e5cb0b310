// kernels_struct.cuh -- structure build ("Build Structure" + "Symbolic Decomposition"
// of the reference, src/block_solver.cpp:139-248) as GPU integer work:
// edge sorting into landmark-major / pose-major SoA, Hpl CSC
// (gpu::buildHplStructure, cuda_block_solver.cu:1571-1604), Hsc block pattern
// (HschurSparseBlockMatrix::constructFromVertices, src/sparse_block_matrix.cpp:63-156,
// a single-threaded host loop over a dense P x P byte map there; a key sort +
// unique here), product lists (findHschureMulBlockIndices :1606-1634) and the
// symbolic factorisation (symbolic.h).
#pragma once
#include "engine.h"
#include "symbolic.h"

namespace cugo_b200 {

// raw edge arrays in input order (sets concatenated): global edge id = position
struct RawEdges {
    int E;
    const int* iP;
    const int* iL;
    const double* m0;
    const double* m1;
    const double* m2;
    const double* w;
    const uint8_t* set;
    const int* cam;        // or nullptr
    const uint8_t* out;    // or nullptr
};

__global__ void k_make_lm_keys(RawEdges r, int nP_all, unsigned long long* keys, int* vals) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= r.E) return;
    keys[k] = (unsigned long long)r.iL[k] * (unsigned long long)nP_all + (unsigned long long)r.iP[k];
    vals[k] = k;
}

// gather into landmark-major SoA; flag = edge owns an Hpl block (both ends free, not an outlier)
__global__ void k_gather_lm(RawEdges r, const int* gid, int nP, int nL, int* s_iP, int* s_iL, double2* s_uv,
                            double* s_r, double* s_w, uint8_t* s_set, int* s_cam, uint8_t* s_out, int* flag) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k > r.E) return;
    if (k == r.E) { flag[k] = 0; return; }
    const int g = gid[k];
    const int iP = r.iP[g], iL = r.iL[g];
    s_iP[k] = iP; s_iL[k] = iL;
    s_uv[k] = make_double2(r.m0[g], r.m1[g]);
    s_r[k] = r.m2[g];
    s_w[k] = r.w[g];
    s_set[k] = r.set[g];
    if (s_cam) s_cam[k] = r.cam[g];
    const int o = r.out ? r.out[g] : 0;
    if (s_out) s_out[k] = (uint8_t)o;
    flag[k] = (iP < nP && iL < nL && !o) ? 1 : 0;
}

// ptr[v] = first position k with sortedKey[k] >= v, for v in [0, n]  (sortedKey ascending)
__global__ void k_lower_bounds(const int* sortedKey, int E, int n, int* ptr) {
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v > n) return;
    int lo = 0, hi = E;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (sortedKey[mid] < v) lo = mid + 1; else hi = mid;
    }
    ptr[v] = lo;
}

__global__ void k_fill_hpl(int E, const int* s_iP, const int* s_iL, const int* gid, const int* flag,
                           const int* pre, const int* lmEdgePtr, int nL, int* colPtr, int* rowInd, int* hplLm,
                           int* edge2Hpl) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k <= nL) colPtr[k] = pre[lmEdgePtr[k]];
    if (k >= E) return;
    if (flag[k]) {
        const int slot = pre[k];
        rowInd[slot] = s_iP[k];
        hplLm[slot] = s_iL[k];
        edge2Hpl[gid[k]] = slot;
    } else {
        edge2Hpl[gid[k]] = -1;
    }
}

// landmark weights for CTA chunking: max(degree, 1)
__global__ void k_lm_weights(const int* lmEdgePtr, int nL_all, int* wgt) {
    const int l = blockIdx.x * blockDim.x + threadIdx.x;
    if (l > nL_all) return;
    wgt[l] = (l < nL_all) ? max(lmEdgePtr[l + 1] - lmEdgePtr[l], 1) : 0;
}
// chunk(l) = Wexcl[l] / LM_CHUNK ; chunkLm[c] = first landmark of chunk c (empty chunks allowed)
__global__ void k_chunk_bounds(const int* wexcl, int nL_all, int nChunks, int* chunkLm) {
    const int l = blockIdx.x * blockDim.x + threadIdx.x;
    if (l > nL_all) return;
    const int cprev = (l == 0) ? -1 : wexcl[l - 1] / LM_CHUNK;
    const int ccur = (l == nL_all) ? nChunks : wexcl[l] / LM_CHUNK;
    for (int c = cprev + 1; c <= ccur; c++) chunkLm[c] = l;
}

// pose-major copy
__global__ void k_make_pose_keys(const int* s_iP, int E, int* keys, int* vals) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= E) return;
    keys[k] = s_iP[k];
    vals[k] = k;
}
__global__ void k_gather_pose(int E, const int* src, const int* s_iL, const double2* s_uv, const double* s_r,
                              const double* s_w, const uint8_t* s_set, const int* s_cam, const uint8_t* s_out,
                              int* p_iL, double2* p_uv, double* p_r, double* p_w, uint8_t* p_set, int* p_cam,
                              uint8_t* p_out) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= E) return;
    const int s = src[k];
    p_iL[k] = s_iL[s]; p_uv[k] = s_uv[s]; p_r[k] = s_r[s]; p_w[k] = s_w[s]; p_set[k] = s_set[s];
    if (p_cam) p_cam[k] = s_cam[s];
    if (p_out) p_out[k] = s_out[s];
}

// products: landmark l contributes d(d+1)/2 pairs (i <= j) of its Hpl column
__global__ void k_prod_counts(const int* colPtr, int nL, int* cnt) {
    const int l = blockIdx.x * blockDim.x + threadIdx.x;
    if (l > nL) return;
    if (l == nL) { cnt[l] = 0; return; }
    const int dd = colPtr[l + 1] - colPtr[l];
    cnt[l] = dd * (dd + 1) / 2;
}
__global__ void k_gen_products(const int* colPtr, const int* rowInd, const int* off, int nL, int nP, int nmul,
                               unsigned long long* keys, unsigned long long* vals) {
    const int l = blockIdx.x * blockDim.x + threadIdx.x;
    if (l < nL) {
        int o = off[l];
        const int c0 = colPtr[l], c1 = colPtr[l + 1];
        for (int i = c0; i < c1; i++) {
            const unsigned long long pi = (unsigned long long)rowInd[i];
            for (int j = i; j < c1; j++) {
                keys[o] = pi * (unsigned long long)nP + (unsigned long long)rowInd[j];
                vals[o] = ((unsigned long long)(unsigned)i << 32) | (unsigned)j;
                o++;
            }
        }
    }
    // explicit (empty) diagonal products so every pose owns a diagonal block
    if (l < nP) {
        keys[nmul + l] = (unsigned long long)l * (unsigned long long)nP + (unsigned long long)l;
        vals[nmul + l] = 0xffffffffffffffffULL;
    }
}
__global__ void k_prod_heads(const unsigned long long* keys, int n, int* head) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k > n) return;
    head[k] = (k < n && (k == 0 || keys[k] != keys[k - 1])) ? 1 : 0;
}
__global__ void k_fill_blocks(const unsigned long long* keys, const unsigned long long* vals, const int* head,
                              const int* blkOf, int n, int nP, int nblk, int* bRowOf, int* bColInd, int* prodPtr,
                              int2* prodIJ) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k == 0) prodPtr[nblk] = n;
    if (k >= n) return;
    const unsigned long long v = vals[k];
    prodIJ[k] = (v == 0xffffffffffffffffULL) ? make_int2(-1, -1) : make_int2((int)(v >> 32), (int)(v & 0xffffffffu));
    if (head[k]) {
        const int b = blkOf[k];
        bRowOf[b] = (int)(keys[k] / (unsigned long long)nP);
        bColInd[b] = (int)(keys[k] % (unsigned long long)nP);
        prodPtr[b] = k;
    }
}

// ---- multi-GPU: union of the ranks' Hsc block patterns --------------------------------------------
constexpr unsigned long long KEY_NONE = 0xffffffffffffffffULL;
// key of every local block, padded with KEY_NONE up to `padded` entries (equal-size all-gather)
__global__ void k_block_keys(const int* bRowOf, const int* bColInd, int nblk, int padded, unsigned long long nP,
                             unsigned long long* keys) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= padded) return;
    keys[k] = k < nblk ? (unsigned long long)bRowOf[k] * nP + (unsigned long long)bColInd[k] : KEY_NONE;
}
__global__ void k_key_heads(const unsigned long long* keys, int n, int* head) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k > n) return;
    head[k] = (k < n && keys[k] != KEY_NONE && (k == 0 || keys[k] != keys[k - 1])) ? 1 : 0;
}
__global__ void k_compact_keys(const unsigned long long* keys, const int* head, const int* idx, int n,
                               unsigned long long nP, unsigned long long* ukeys, int* rowOf, int* colInd) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n || !head[k]) return;
    const int g = idx[k];
    ukeys[g] = keys[k];
    rowOf[g] = (int)(keys[k] / nP);
    colInd[g] = (int)(keys[k] % nP);
}
__global__ void k_map_slots(const unsigned long long* lkeys, int nblk, const unsigned long long* ukeys, int nblkG,
                            int* slot) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= nblk) return;
    const unsigned long long key = lkeys[k];
    int lo = 0, hi = nblkG;
    while (lo < hi) { const int mid = (lo + hi) >> 1; if (ukeys[mid] < key) lo = mid + 1; else hi = mid; }
    slot[k] = lo;
}

// work units of the pose-major build pass: slices of POSE_SLICE edges per pose
__global__ void k_pose_slice_count(const int* posePtr, int nP, int* cnt) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p > nP) return;
    const int deg = p < nP ? posePtr[p + 1] - posePtr[p] : 0;
    cnt[p] = p < nP ? max(1, (deg + POSE_SLICE - 1) / POSE_SLICE) : 0;
}
__global__ void k_pose_units_fill(const int* slicePtr, int nP, int2* units) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= nP) return;
    for (int u = slicePtr[p]; u < slicePtr[p + 1]; u++) units[u] = make_int2(p, u - slicePtr[p]);
}

// ---- symbolic factorisation: the shared host/device code run by one thread ----
// adjCap > 0: the adjacency and the BFS scratch (the arrays every sweep of the nested dissection walks) live
// in shared memory -- a single thread pays ~30 cycles per dependent access there instead of an L2 round trip.
__global__ void k_sym_analyze(SymWork w, SymCounts* out, int adjCap) {
    extern __shared__ int symSh[];
    if (blockIdx.x != 0 || threadIdx.x != 0) return;
    if (adjCap > 0) {
        w.adj = symSh; w.adjPtr = symSh + adjCap; w.mark = w.adjPtr + (w.P + 1); w.level = w.mark + w.P;
        w.queue = w.level + w.P;
    }
    sym_analyze(w, out);
}
// rb > 0: second-generation task tables (kernels_chol2.cuh): slabs of rb block rows, tiles of tb
// block rows, and the widest supernode of every level (levMaxNt) for shared-memory sizing.
// rb == 0: first-generation tables in scalar rows (rs rows per slab, ts x ts tiles).
__global__ void k_sym_fill(SymWork w, const SymCounts* c, SymMaps m, int* trsmPtr, int* updPtr, int rs, int ts, int rb,
                           int tb, int* levMaxNt, double* factStats) {
    if (blockIdx.x != 0 || threadIdx.x != 0) return;
    sym_fill(w, *c, m);
    // task tables of the numeric kernels, in levelSn order
    int a = 0, b = 0;
    double fPanel = 0.0, fUpd = 0.0, fBwd = 0.0;      // algorithmic flops of one factorisation + solve
    for (int l = 0; l < c->nlev; l++) levMaxNt[l] = 0;
    for (int q = 0; q < c->nsn; q++) {
        const int t = m.levelSn[q];
        const int rt = w.snRowPtr[t + 1] - w.snRowPtr[t];
        const int nt = w.snStart[t + 1] - w.snStart[t];
        if (nt > levMaxNt[w.snLevel[t]]) levMaxNt[w.snLevel[t]] = nt;
        trsmPtr[q] = a; updPtr[q] = b;
        {
            const double ncd = 6.0 * nt, rd = 6.0 * rt + 1.0;
            fPanel += ncd * ncd * ncd / 3.0 + rd * ncd * ncd;   // Cholesky of L11 + triangular solves of the rows below (rhs row included)
            fUpd += rt > 0 ? rd * rd * ncd : 0.0;               // symmetric rank-nc update, lower triangle
            fBwd += ncd * ncd + 2.0 * 6.0 * rt * ncd;           // backward substitution
        }
        if (rb > 0) {
            const int ns = (rt + rb - 1) / rb;
            a += ns > 0 ? ns : 1;
            if (rt > 0) { const int n2 = (rt + tb) / tb; b += n2 * (n2 + 1) / 2; }
        } else {
            a += (6 * rt + 1 + rs - 1) / rs;
            if (rt > 0) { const int n2 = (6 * rt + 1 + ts - 1) / ts; b += n2 * (n2 + 1) / 2; }
        }
    }
    trsmPtr[c->nsn] = a; updPtr[c->nsn] = b;
    factStats[0] = fPanel; factStats[1] = fUpd; factStats[2] = fBwd; factStats[3] = (double)c->lsize;
}

}  // namespace cugo_b200
