// kernels_chol_common.cuh -- constants and helpers shared by the kernels of the multifrontal Cholesky
// (kernels_chol3.cuh): front geometry of a supernode, task lookup in the level tables, child descriptors, the FP64
// tensor-core primitive (DMMA m8n8k4) and the optional phase-clock diagnostics.
#pragma once
#include "device_math.cuh"
#include "engine.h"

namespace cugo_b200 {

constexpr int CHOL_MAXW = 16;               // block columns per supernode (panel width <= 96)
constexpr double CHOL_PIVOT_TOL = 1e-14;    // reference zero-pivot tolerance, src/cholesky.hpp:79-92
constexpr int CHOL_RB = 16;                 // most block rows of L21 per panel slab (96 rows + rhs row); per level: Solver::planPanelSlabs
constexpr int PANEL_LD = 196;               // smem leading dimension: >= 6 MAXW + 6 RB + 1, == 4 (mod 16)
constexpr int CHOL_TB = 8;                  // block rows per update tile (48 rows)
constexpr int UPD_LD = 52;                  // == 4 (mod 16): conflict-free DMMA fragment loads
constexpr int UPD_CLD = 50;
constexpr int CHOL_MAXCH = 32;              // children whose descriptors are cached in shared memory
static_assert(PANEL_LD >= 6 * CHOL_MAXW + 6 * CHOL_RB + 1 && PANEL_LD % 16 == 4, "panel leading dimension");

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

// D = A(8x4, row) * B(4x8, col) + C on the FP64 tensor pipe (DMMA)
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

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
// Diagnostics (phase clocks of CTA 0 per tree level), compiled in only with -DCUGO_B200_DIAG (make DIAG=1): the shipped
// library carries neither the device arrays nor the stamps.
#ifdef CUGO_B200_DIAG
__device__ long long g_cholDbg[64 * 24];
__device__ long long g_cholTrace[176];       // clock of CTA 0 / thread 0 at points of every factor-loop step of level 2
__device__ long long g_cholTime[64 * 8];    // %globaltimer (ns) of CTA 0 per level: panel (after wait, end), update, backward solve
__device__ long long g_schurDbg[8192];      // cycles per unit of k_schur_rows
__device__ __forceinline__ long long global_ns() { long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
#define CHOL_DIAG(c) ((c).dbg)
#define CHOL_GT(slot) do { if (c.dbg && blockIdx.x == 0 && threadIdx.x == 0 && lev < 64) g_cholTime[lev * 8 + (slot)] = global_ns(); } while (0)
#define CHOL_STAMP(ph) do { if (c.dbg && blockIdx.x == 0 && threadIdx.x == 0 && lev < 64) g_cholDbg[lev * 24 + (ph)] = clock64(); } while (0)
#else
// (the stamps below sit behind `if (CHOL_DIAG(c) && ...)` = if (0 && ...): dead code, never addressed)
#define g_cholDbg ((long long*)nullptr)
#define g_cholTrace ((long long*)nullptr)
#define g_cholTime ((long long*)nullptr)
#define g_schurDbg ((long long*)nullptr)
#define CHOL_DIAG(c) 0
#define CHOL_GT(slot) do { } while (0)
#define CHOL_STAMP(ph) do { } while (0)
#endif
__device__ __forceinline__ int tri(int i, int k) { return i * (i + 1) / 2 + k; }   // packed lower, k <= i

}  // namespace cugo_b200
