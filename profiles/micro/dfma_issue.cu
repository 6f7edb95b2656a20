// Microbenchmark (tuning aid, not product): how many SMSP issue cycles does one DFMA cost on sm_100a, and do
// integer instructions co-issue with it?  8 independent DFMA chains per thread, NI integer ops per DFMA.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o dfma_issue dfma_issue.cu && ./dfma_issue
#include <cstdio>
#include <cuda_runtime.h>

template <int NI, bool UCONST>
__global__ void k(double* out, const double* cin, int iters, unsigned seed)
{
    double a[8];
    unsigned x[8];
    const double c0 = UCONST ? 0.999 : cin[threadIdx.x & 1], c1 = UCONST ? 1e-3 : cin[2 + (threadIdx.x & 1)];
    for (int i = 0; i < 8; i++) { a[i] = threadIdx.x * 1e-3 + i; x[i] = seed + threadIdx.x * 8 + i; }
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int u = 0; u < 4; u++) {
#pragma unroll
            for (int i = 0; i < 8; i++) {
                a[i] = fma(a[i], c0, c1);
#pragma unroll
                for (int j = 0; j < NI; j++) x[i] = (x[i] & 0x3ffeu) ^ (x[(i + 1) & 7] >> 3);
            }
        }
    }
    double s = 0; unsigned t = 0;
    for (int i = 0; i < 8; i++) { s += a[i]; t ^= x[i]; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s + t;
}

template <int NI, bool UC>
void run(const char* name, int warps_per_sm, double* out, double* cin)
{
    const int iters = 4096, threads = 32 * warps_per_sm / 4 * 4;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<NI, UC><<<148, threads>>>(out, cin, 16, 1);
    cudaEventRecord(e0);
    k<NI, UC><<<148, threads>>>(out, cin, iters, 1);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    const double dfma_per_smsp = (double)iters * 32 * (warps_per_sm / 4.0);
    // cycles at the max clock (an upper bound of the cycle count if the clock was lower)
    printf("%-28s warps/SM %2d  %.3f ms  %.2f cycles(@%d MHz) per DFMA per SMSP\n", name, warps_per_sm, ms,
           ms * 1e-3 * clk * 1e3 / dfma_per_smsp, clk / 1000);
}

int main()
{
    double *out, *cin;
    cudaMalloc(&out, 148 * 1024 * 8); cudaMalloc(&cin, 64);
    double h[4] = {0.999, 0.999, 1e-3, 1e-3}; cudaMemcpy(cin, h, 32, cudaMemcpyHostToDevice);
    for (int w : {4, 8, 16, 32}) {
        run<0, true>("dfma only (imm const)", w, out, cin);
        run<0, false>("dfma only (reg const)", w, out, cin);
        run<1, false>("dfma + 2 int ops (LOP3,SHF)", w, out, cin);
        run<2, false>("dfma + 4 int ops", w, out, cin);
    }
    printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
