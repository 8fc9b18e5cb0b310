// Microbenchmarks (B200): dependent-chain latency and per-SM throughput of the FP64 operations the
// factorisation kernels are built from.  nvcc -O3 -gencode arch=compute_100a,code=sm_100a
#include <cstdio>
#include <cuda_runtime.h>
#define N 2048
__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
template <int OP>
__global__ void lat(double* out, long long* cyc, double seed) {
    double x = seed + threadIdx.x * 1e-3, y = 1.0000001, z = 0.5;
    double c0 = 0, c1 = 0;
    long long t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N; i++) {
        if (OP == 0) x = fma(x, y, z);
        if (OP == 1) x = x * y;
        if (OP == 2) x = rsqrt(x) + 1.5;
        if (OP == 3) x = 1.0 / x + 0.5;
        if (OP == 4) x = sqrt(x) + 1.5;
        if (OP == 5) dmma(c0, c1, x, y);
        if (OP == 6) x = x + y;
        if (OP == 7) { float f = rsqrtf((float)x); double r = (double)f; r = r * (1.5 - 0.5 * x * r * r); r = r * (1.5 - 0.5 * x * r * r); x = r + 1.5; }
    }
    long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = x + c0 + c1;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
// throughput: ILP independent chains per thread, nthreads per CTA, one CTA per SM
template <int OP>
__global__ void thr(double* out, long long* cyc, double seed) {
    double x[8], c[8][2];
    for (int j = 0; j < 8; j++) { x[j] = seed + j + threadIdx.x * 1e-3; c[j][0] = c[j][1] = 0; }
    double y = 1.0000001, z = 0.5;
    __syncthreads();
    long long t0 = clock64();
    for (int i = 0; i < N / 8; i++) {
#pragma unroll
        for (int j = 0; j < 8; j++) {
            if (OP == 0) x[j] = fma(x[j], y, z);
            if (OP == 5) dmma(c[j][0], c[j][1], x[j], y);
        }
    }
    __syncthreads();
    long long t1 = clock64();
    double s = 0;
    for (int j = 0; j < 8; j++) s += x[j] + c[j][0] + c[j][1];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
// whole-GPU rate: every SM busy, timed with CUDA events (no clock64 involved)
__global__ void __launch_bounds__(1024) dmma_full(double* out, int iters, double seed) {
    double x[8], c[8][2];
    for (int j = 0; j < 8; j++) { x[j] = seed + j + threadIdx.x * 1e-3; c[j][0] = c[j][1] = 0; }
    double y = 1.0000001;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int j = 0; j < 8; j++) dmma(c[j][0], c[j][1], x[j], y);
    }
    double s = 0;
    for (int j = 0; j < 8; j++) s += c[j][0] + c[j][1];
    out[(blockIdx.x * blockDim.x + threadIdx.x) & 0xffff] = s;
}
// accuracy of the chain's reciprocal square root (kernels_chol3.cuh: chain_rsqrt) against rsqrt()
__global__ void rsq_acc(double* out) {
    double worst = 0.0, worst0 = 0.0;
    for (int i = threadIdx.x; i < (1 << 20); i += blockDim.x) {
        const double x = exp2(-46.0 + 110.0 * i / (1 << 20)) * (1.0 + 0.37 * (i % 97) / 97.0);
        double y;
        asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
        const double ref = rsqrt(x);
        worst0 = fmax(worst0, fabs(y - ref) / ref);
        const double e = fma(-x * y, y, 1.0);
        const double p = fma(e, 0.375, 0.5);
        y = fma(y * e, p, y);
        worst = fmax(worst, fabs(y - ref) / ref);
    }
    out[threadIdx.x] = worst; out[256 + threadIdx.x] = worst0;
}
__global__ void ldslat(double* out, long long* cyc, int stride) {
    __shared__ int idx[1024];
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) idx[i] = (i + stride) & 1023;
    __syncthreads();
    int p = threadIdx.x;
    long long t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N; i++) p = idx[p];
    long long t1 = clock64();
    out[threadIdx.x] = p;
    if (threadIdx.x == 0) *cyc = t1 - t0;
}
__global__ void synclat(long long* cyc) {
    long long t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N; i++) __syncthreads();
    long long t1 = clock64();
    if (threadIdx.x == 0) *cyc = t1 - t0;
}
int main() {
    double* out; long long* cyc; cudaMalloc(&out, 1 << 22); cudaMalloc(&cyc, 8);
    long long h;
    const char* names[] = {"dfma", "dmul", "rsqrt(double)+add", "1/x+add", "sqrt+add", "dmma m8n8k4", "dadd", "rsqrtf+2NR+add"};
#define LAT(OP) lat<OP><<<1, 32>>>(out, cyc, 1.25); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost); printf("latency %-20s %.1f cycles/op (1 warp)\n", names[OP], (double)h / N);
    LAT(0) LAT(1) LAT(2) LAT(3) LAT(4) LAT(5) LAT(6) LAT(7)
    for (int nt = 32; nt <= 1024; nt *= 2) {
        thr<0><<<1, nt>>>(out, cyc, 1.25); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
        printf("throughput dfma  %4d thr/SM: %.2f FMA/clk/SM\n", nt, (double)N * nt / h);
        thr<5><<<1, nt>>>(out, cyc, 1.25); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
        printf("throughput dmma  %4d thr/SM: %.2f FMA/clk/SM (%.1f cycles per warp-MMA per SM)\n", nt, (double)N * (nt / 32) * 256 / h, (double)h / (N * (nt / 32)));
    }
    ldslat<<<1, 32>>>(out, cyc, 33); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost); printf("LDS dependent latency %.1f cycles\n", (double)h / N);
    for (int nt = 64; nt <= 256; nt *= 2) { synclat<<<1, nt>>>(cyc); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost); printf("__syncthreads %d threads: %.1f cycles\n", nt, (double)h / N); }
    {
        int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
        cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
        for (int nt = 256; nt <= 1024; nt *= 2) {
            const int iters = 1 << 15;
            dmma_full<<<sms, nt>>>(out, 64, 1.25);
            cudaEventRecord(e0); dmma_full<<<sms, nt>>>(out, iters, 1.25); cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
            const double flop = 2.0 * 256 * 8.0 * iters * (nt / 32) * sms;
            printf("whole-GPU dmma %4d thr/SM x %d SMs: %.2f TFLOP/s (%.3f ms, CUDA events)\n", nt, sms, flop / (ms * 1e-3) / 1e12, ms);
        }
    }
    {
        rsq_acc<<<1, 256>>>(out);
        static double h2[512]; cudaMemcpy(h2, out, sizeof(h2), cudaMemcpyDeviceToHost);
        double w = 0, w0 = 0; for (int i = 0; i < 256; i++) { if (h2[i] > w) w = h2[i]; if (h2[256 + i] > w0) w0 = h2[256 + i]; }
        printf("rsqrt.approx.ftz.f64: max relative error %.3e; after one third-order step %.3e (against rsqrt(), x in [2^-46, 2^64])\n", w0, w);
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
