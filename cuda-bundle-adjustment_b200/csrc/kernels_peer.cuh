// kernels_peer.cuh -- the exchange step of a sharded trial over NVLink peer memory (several B200s of one NVSwitch box, one
// process per GPU): the partial reduced camera systems [Hsc | bsc] of the ranks are summed by ONE kernel per rank that reads the
// peers' partials and writes the peers' results directly (two-shot: every rank reduces its 1/N slice in rank order -- so all
// ranks hold bit-identical sums -- and pushes it to everybody), tied together by flags in peer memory.  No counterpart in the
// reference (single GPU); it replaces the ncclAllReduce of round 1 on this step (55 us for 6.3 MB on two GPUs, most of it
// latency).  The small reductions (chi2 / gain / pivot flag; the diagonal maxima of the first trial) use the one-shot variant.
// Memory: every rank cudaMalloc's one exchange block and opens the others' through CUDA IPC (Solver::peerSetup).
#pragma once
#include "engine.h"
#include "kernels_lm.cuh"

namespace cugo_b200 {

constexpr int PEER_MAXR = 8;            // ranks of one NVSwitch box
constexpr int PEER_CTAS = 48;           // CTAs of the all-reduce kernel (each with its own flag slots)
constexpr int PEER_TPB = 512;
constexpr int PEER_SMALL_MAX = 16384;   // doubles of a one-shot reduction (6 P + ranks for the diagonal maxima)

// layout of a rank's exchange block (all offsets in bytes from its base, identical on every rank)
struct PeerCtx {
    char* base[PEER_MAXR];      // base[r]: rank r's block as mapped into THIS process (base[rank] = the local allocation)
    int nranks, rank;
    size_t offPart;             // [nred] doubles: this rank's partial system (k_schur_rows writes it)
    size_t offDst;              // [nred] doubles: the summed system (every rank's slice is pushed here)
    size_t offSmall;            // [PEER_MAXR][PEER_SMALL_MAX] doubles: one-shot slots, slot s written by rank s
    size_t offFlag;             // int flags: [3 kinds][PEER_MAXR source ranks][PEER_CTAS]
    size_t offEpoch;            // int [2][PEER_CTAS]: local launch counters (big kernel per CTA, small kernel)
};

__device__ __forceinline__ void st_release_sys(int* p, int v) { asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ int ld_acquire_sys(const int* p) {
    int v;
    asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ int* peer_flag(const PeerCtx& x, int onRank, int kind, int src, int cta) {
    return reinterpret_cast<int*>(x.base[onRank] + x.offFlag) + ((size_t)kind * PEER_MAXR + src) * PEER_CTAS + cta;
}
// every rank's CTA `cta` has passed this point with epoch >= ep (thread 0 signals and waits; the others wait at the barrier)
__device__ __forceinline__ void peer_barrier(const PeerCtx& x, int kind, int cta, int ep) {
    __syncthreads();
    if (threadIdx.x < x.nranks) {
        const int p = threadIdx.x;
        __threadfence_system();
        st_release_sys(peer_flag(x, p, kind, x.rank, cta), ep);
        while (ld_acquire_sys(peer_flag(x, x.rank, kind, p, cta)) < ep) { }
    }
    __syncthreads();
}

// dst = sum over the ranks of part, n doubles (n even), on every rank.  Rank r owns the slice [r n / N, (r + 1) n / N) (in double2
// units); CTA c a contiguous share of it.
__global__ void __launch_bounds__(PEER_TPB) k_peer_allreduce(PeerCtx x, size_t n, const LMState* st) {
    if (st && st->done) return;             // identical on every rank (the decision is taken on reduced values)
    const int cta = blockIdx.x, N = x.nranks;
    int* epoch = reinterpret_cast<int*>(x.base[x.rank] + x.offEpoch) + cta;
    const int ep = *epoch + 1;
    peer_barrier(x, 0, cta, ep);            // every rank's partial is complete (its Schur kernel precedes this one in its stream)
    const size_t n2 = n >> 1;
    const size_t sBeg = n2 * x.rank / N, sEnd = n2 * (x.rank + 1) / N;
    const size_t per = (sEnd - sBeg + gridDim.x - 1) / gridDim.x;
    const size_t cBeg = sBeg + per * cta, cEnd = min(sEnd, cBeg + per);
    const double2* part[PEER_MAXR];
    double2* dst[PEER_MAXR];
#pragma unroll
    for (int p = 0; p < PEER_MAXR; p++) {
        part[p] = reinterpret_cast<const double2*>(x.base[p < N ? p : 0] + x.offPart);
        dst[p] = reinterpret_cast<double2*>(x.base[p < N ? p : 0] + x.offDst);
    }
    for (size_t i0 = cBeg + threadIdx.x; i0 < cEnd; i0 += 2 * PEER_TPB) {
        const size_t i1 = i0 + PEER_TPB;
        const bool two = i1 < cEnd;
        double2 v0[PEER_MAXR], v1[PEER_MAXR];
#pragma unroll
        for (int p = 0; p < PEER_MAXR; p++)
            if (p < N) {
                v0[p] = __ldcg(part[p] + i0);      // peer memory: past the L1
                if (two) v1[p] = __ldcg(part[p] + i1);
            }
        double2 a = v0[0], b = two ? v1[0] : make_double2(0.0, 0.0);
#pragma unroll
        for (int p = 1; p < PEER_MAXR; p++)
            if (p < N) {
                a.x += v0[p].x; a.y += v0[p].y;
                if (two) { b.x += v1[p].x; b.y += v1[p].y; }
            }
#pragma unroll
        for (int p = 0; p < PEER_MAXR; p++)
            if (p < N) {
                dst[p][i0] = a;
                if (two) dst[p][i1] = b;
            }
    }
    peer_barrier(x, 1, cta, ep);            // every rank's pushes of its CTA `cta` have landed
    if (threadIdx.x == 0) *epoch = ep;
}

// One-shot sum of a short vector (n <= PEER_SMALL_MAX doubles) over the ranks, in place in `buf` (local memory): every rank writes
// its vector into slot `rank` of everybody, then sums the slots in rank order.  One CTA.
__global__ void __launch_bounds__(PEER_TPB) k_peer_small(PeerCtx x, double* buf, int n, const LMState* st, int checkDone) {
    if (checkDone && st->done) return;
    const int N = x.nranks;
    int* epoch = reinterpret_cast<int*>(x.base[x.rank] + x.offEpoch) + PEER_CTAS;
    const int ep = *epoch + 1;
    for (int i = threadIdx.x; i < n; i += PEER_TPB) {
        const double v = buf[i];
        for (int p = 0; p < N; p++)
            reinterpret_cast<double*>(x.base[p] + x.offSmall)[(size_t)x.rank * PEER_SMALL_MAX + i] = v;
    }
    peer_barrier(x, 2, 0, ep);
    const double* slots = reinterpret_cast<const double*>(x.base[x.rank] + x.offSmall);
    for (int i = threadIdx.x; i < n; i += PEER_TPB) {
        double s = 0.0;
        for (int p = 0; p < N; p++) s += __ldcg(slots + (size_t)p * PEER_SMALL_MAX + i);
        buf[i] = s;
    }
    // the slots may be overwritten by the next call only after everybody has read them: the next call's barrier comes after its
    // writes, so close this one with a second barrier
    peer_barrier(x, 2, 1, ep);
    if (threadIdx.x == 0) *epoch = ep;
}

// End of a sharded trial in ONE kernel: the rank's chi2 / gain denominator / pivot flag (k_lm_sums), their sum over the ranks
// through peer memory (one-shot, rank order: every rank decides on identical numbers) and the decision (lm_decide_tail) --
// instead of k_lm_sums, a reduction kernel and k_lm_decide.  One CTA.  Slots: the tail of the one-shot region.
template <int TPB>
__global__ void __launch_bounds__(TPB) k_lm_decide_peer(Dev d, PeerCtx x) {
    __shared__ double sh[TPB / 32];
    __shared__ double mine[3], tot[3];
    pdl_trigger(); pdl_wait();
    LMState* st = d.st;
    if (st->done) return;                   // identical on every rank
    const double chi = sum_array<TPB>(d.chiPartial, d.nChi, sh);          // (block sums: valid on thread 0 only)
    const double lamTrue = st->lambda, lam = d.rank == 0 ? lamTrue : 0.0;
    double s = 0.0;
    for (int i = threadIdx.x; i < 6 * d.nP; i += TPB) s += d.x[i] * (lam * d.x[i] + d.b[i]);
    for (int i = threadIdx.x; i < d.nChunks; i += TPB) s += d.scalePartial[i];
    s = block_sum<TPB>(s, sh);
    if (threadIdx.x == 0) { mine[0] = chi; mine[1] = s; mine[2] = st->fail ? 1.0 : 0.0; }
    __syncthreads();
    const int N = x.nranks;
    int* epoch = reinterpret_cast<int*>(x.base[x.rank] + x.offEpoch) + PEER_CTAS + 1;
    const int ep = *epoch + 1;
    constexpr int SLOT = PEER_SMALL_MAX - 4;
    if (threadIdx.x < N) {
        double* slot = reinterpret_cast<double*>(x.base[threadIdx.x] + x.offSmall) + (size_t)x.rank * PEER_SMALL_MAX + SLOT;
        slot[0] = mine[0]; slot[1] = mine[1]; slot[2] = mine[2];
    }
    peer_barrier(x, 2, 2, ep);
    if (threadIdx.x < 3) {
        const double* slots = reinterpret_cast<const double*>(x.base[x.rank] + x.offSmall) + SLOT + threadIdx.x;
        double v = 0.0;
        for (int p = 0; p < N; p++) v += __ldcg(slots + (size_t)p * PEER_SMALL_MAX);
        tot[threadIdx.x] = v;
    }
    peer_barrier(x, 2, 3, ep);              // everybody has read the slots before anybody writes the next trial's
    if (threadIdx.x != 0) return;
    *epoch = ep;
    lm_decide_tail(d, st, tot[0], tot[1], tot[2] > 0.0, lamTrue);
}

}  // namespace cugo_b200
