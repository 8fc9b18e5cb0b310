// kernels_lm.cuh -- device-resident Levenberg-Marquardt control.
// The reference runs this logic on the host with 2-3 scalar read-backs and >= 15
// stream syncs per trial (src/cuda_graph_optimisation.cpp:48-154).
// Here the state lives in LMState; every kernel of a "trial slot" reads it, and
// k_lm_decide applies exactly the reference's accept / reject / termination rules.
// A CUDA graph of N slots runs N trials without the host; kernels of slots after
// termination exit on st->done.
#pragma once
#include "device_math.cuh"
#include "engine.h"

namespace cugo_b200 {

template <int TPB>
__device__ __forceinline__ double sum_array(const double* a, int n, double* sh) {
    double s = 0.0;
    for (int i = threadIdx.x; i < n; i += TPB) s += a[i];
    return block_sum<TPB>(s, sh);
}

// start of optimize(): F = chi2 of the current estimates (k_errors ran before)
template <int TPB>
__global__ void __launch_bounds__(TPB) k_lm_begin(Dev d, int niter) {
    __shared__ double sh[TPB / 32];
    double F = sum_array<TPB>(d.chiPartial, d.nChi, sh);
    if (threadIdx.x == 0) {
        LMState* st = d.st;
        if (d.nranks > 1) F = d.redC[0];
        st->F = F; st->Fhat = F; st->lambda = 0.0; st->nu = 2.0; st->rho = -1.0; st->scale = 0.0;
        st->iter = 0; st->q = 0; st->done = (niter <= 0) ? 1 : 0; st->needBuild = 1; st->fail = 0;
        st->niter = niter; st->nstats = 0; st->slots = 0;
    }
}

// after the (conditional) system build: lambda0 = tau * max diagonal on the first
// iteration (reference :69-72, maxDiagonalKernel cuda_block_solver.cu:1222-1253)
template <int TPB>
__global__ void __launch_bounds__(TPB) k_lm_prepare(Dev d) {
    __shared__ double sh[TPB / 32];
    pdl_trigger(); pdl_wait();   // engine.h: nothing mutable is touched before the predecessor has completed
    LMState* st = d.st;
    if (st->done) return;
    if (st->needBuild) {
        double m = 0.0;
        if (d.nranks > 1) {     // exchanged in the first trial of an optimize() only: that is when lambda0 is formed
            if (st->iter == 0)
                for (int i = threadIdx.x; i < 6 * d.nP + d.nranks; i += TPB) m = fmax(m, d.redA[i]);
        } else {
            for (int i = threadIdx.x; i < d.nChunks; i += TPB) m = fmax(m, d.maxDiagL[i]);
            for (int i = threadIdx.x; i < d.nP; i += TPB) m = fmax(m, d.maxDiagP[i]);
        }
        m = block_max<TPB>(m, sh);
        if (threadIdx.x == 0) {
            if (d.nranks == 1 || st->iter == 0) st->maxDiag = m;
            if (st->iter == 0) st->lambda = 1e-5 * m;
            st->needBuild = 0;
        }
    }
    if (threadIdx.x == 0) { st->fail = 0; st->slots++; }
}

// the decision itself (one thread): gain ratio, accept / reject, damping update, termination
__device__ __forceinline__ void lm_decide_tail(const Dev& d, LMState* st, double Fhat, double scaleRaw, bool fail, double lam) {
    const int maxq = 10;
    const double F = st->F;
    const double scale = scaleRaw + 1e-3;
    const double Fdiff = Fhat - F;
    const double rho = fail ? -1.0 : (F - Fhat) / scale;
    st->Fhat = Fhat; st->scale = scale; st->rho = rho;
    bool endIter = false;
    int q = st->q;
    if (rho > 0) {
        const double a = 1.0 - (2.0 * rho - 1.0) * (2.0 * rho - 1.0) * (2.0 * rho - 1.0);
        st->lambda = lam * fmax(1.0 / 3.0, fmin(a, 2.0 / 3.0));
        st->nu = 2.0;
        st->F = Fhat;
        st->cur = 1 - st->cur;       // accept: the trial buffer becomes current
        st->needBuild = 1;
        endIter = true;
    } else {
        st->lambda = lam * st->nu;
        st->nu *= 2.0;
        if (!isfinite(st->lambda) || Fdiff < 1e-4) {
            endIter = true;
        } else {
            q++;
            if (!(q < maxq && rho < 0)) endIter = true;
        }
    }
    st->q = q;
    if (endIter) {
        const int n = st->nstats;
        if (n < MAX_STATS) {
            st->statChi2[n] = st->F; st->statLambda[n] = st->lambda; st->statTrials[n] = q;
        }
        st->nstats = n + 1;
        if (q == maxq || rho < 1e-6 || !isfinite(st->lambda)) {
            st->done = 1;
        } else {
            st->iter++;
            st->q = 0;
            if (st->iter >= st->niter) st->done = 1;
        }
    }
}

// end of a trial: gain ratio, accept / reject, damping update, termination.
template <int TPB>
__global__ void __launch_bounds__(TPB) k_lm_decide(Dev d) {
    __shared__ double sh[TPB / 32];
    pdl_trigger(); pdl_wait();   // engine.h: nothing mutable is touched before the predecessor has completed
    LMState* st = d.st;
    if (st->done) return;
    double Fhat = sum_array<TPB>(d.chiPartial, d.nChi, sh);
    const double lam = st->lambda;
    double s = 0.0;
    for (int i = threadIdx.x; i < 6 * d.nP; i += TPB) s += d.x[i] * (lam * d.x[i] + d.b[i]);
    for (int i = threadIdx.x; i < d.nChunks; i += TPB) s += d.scalePartial[i];
    double scaleRaw = block_sum<TPB>(s, sh);
    if (threadIdx.x != 0) return;
    bool fail = st->fail != 0;
    if (d.nranks > 1) { Fhat = d.redC[0]; scaleRaw = d.redC[1]; fail = d.redC[2] > 0.0; }   // k_lm_sums + all-reduce ran before
    lm_decide_tail(d, st, Fhat, scaleRaw, fail, lam);
}

// ---- multi-GPU: local partial sums that are all-reduced before the LM kernels read them ------------
// redA = [diag(Hpp_local) (6 nP) | one-hot max diagonal of the local Hll (nranks)]: after a SUM all-reduce
// every rank can take the max over diag(Hpp_total) and all ranks' landmark maxima.
template <int TPB>
__global__ void __launch_bounds__(TPB) k_pack_A(Dev d) {
    __shared__ double sh[TPB / 32];
    pdl_trigger(); pdl_wait();   // engine.h: nothing mutable is touched before the predecessor has completed
    for (int i = blockIdx.x * TPB + threadIdx.x; i < 6 * d.nP; i += gridDim.x * TPB)
        d.redA[i] = d.Hpp[(size_t)(i / 6) * 36 + (i % 6) * 7];
    if (blockIdx.x == 0) {
        double m = 0.0;
        for (int i = threadIdx.x; i < d.nChunks; i += TPB) m = fmax(m, d.maxDiagL[i]);
        m = block_max<TPB>(m, sh);
        if (threadIdx.x < d.nranks) d.redA[6 * d.nP + threadIdx.x] = 0.0;
        __syncthreads();
        if (threadIdx.x == 0) d.redA[6 * d.nP + d.rank] = m;
    }
}
// redC = [chi2 of the rank's edges | gain denominator: pose part with the rank's partial bp (lambda term on
// rank 0 only) + the rank's landmarks | this rank's factorisation hit a non-positive pivot]; the first two are linear
// in the per-rank parts, so SUM gives the global values; the third is > 0 iff any rank failed (with the factorisation
// distributed by subtree a pivot fails on its owner only, and every rank must take the same decision).
template <int TPB>
__global__ void __launch_bounds__(TPB) k_lm_sums(Dev d) {
    __shared__ double sh[TPB / 32];
    pdl_trigger(); pdl_wait();   // engine.h: nothing mutable is touched before the predecessor has completed
    const double chi = sum_array<TPB>(d.chiPartial, d.nChi, sh);
    const double lam = d.rank == 0 ? d.st->lambda : 0.0;
    double s = 0.0;
    for (int i = threadIdx.x; i < 6 * d.nP; i += TPB) s += d.x[i] * (lam * d.x[i] + d.b[i]);
    for (int i = threadIdx.x; i < d.nChunks; i += TPB) s += d.scalePartial[i];
    s = block_sum<TPB>(s, sh);
    if (threadIdx.x == 0) { d.redC[0] = chi; d.redC[1] = s; d.redC[2] = d.st->fail ? 1.0 : 0.0; }
}

// ---- helpers of the host-driven (staged) API ---------------------------------
__global__ void k_state_set(LMState* st, double lambda, int needBuild) {
    st->done = 0; st->needBuild = needBuild; st->lambda = lambda; st->fail = 0;
}
__global__ void k_state_flip(LMState* st) { st->cur = 1 - st->cur; }
template <int TPB>
__global__ void __launch_bounds__(TPB) k_reduce(const double* a, int n, const double* a2, int n2, int isMax,
                                                double* out) {
    __shared__ double sh[TPB / 32];
    double v = 0.0;
    if (isMax) {
        for (int i = threadIdx.x; i < n; i += TPB) v = fmax(v, a[i]);
        for (int i = threadIdx.x; i < n2; i += TPB) v = fmax(v, a2[i]);
        v = block_max<TPB>(v, sh);
    } else {
        for (int i = threadIdx.x; i < n; i += TPB) v += a[i];
        for (int i = threadIdx.x; i < n2; i += TPB) v += a2[i];
        v = block_sum<TPB>(v, sh);
    }
    if (threadIdx.x == 0) *out = v;
}
template <int TPB>
__global__ void __launch_bounds__(TPB) k_scale_poses(Dev d, double lambda, double* out) {
    __shared__ double sh[TPB / 32];
    double s = 0.0;
    for (int i = threadIdx.x; i < 6 * d.nP; i += TPB) s += d.x[i] * (lambda * d.x[i] + d.b[i]);
    for (int i = threadIdx.x; i < d.nChunks; i += TPB) s += d.scalePartial[i];
    s = block_sum<TPB>(s, sh);
    if (threadIdx.x == 0) *out = s;
}

// computeScale with a lambda other than the one of the last solve(): the reference's sum over the whole
// [xp | xl] / [bp | bl] vectors (computeScaleKernel cuda_block_solver.cu:1471-1490), one CTA, fixed order
template <int TPB>
__global__ void __launch_bounds__(TPB) k_scale_full(Dev d, double lambda, double* out) {
    __shared__ double sh[TPB / 32];
    double s = 0.0;
    const size_t n = (size_t)6 * d.nP + (size_t)3 * d.nL;
    for (size_t i = threadIdx.x; i < n; i += TPB) s += d.x[i] * (lambda * d.x[i] + d.b[i]);
    s = block_sum<TPB>(s, sh);
    if (threadIdx.x == 0) *out = s;
}

}  // namespace cugo_b200
