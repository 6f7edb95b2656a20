// Microbenchmark (tuning aid): throughput of the FP64<->FP32 conversions next to DFMA on sm_100a.
#include <cstdio>
#include <cuda_runtime.h>
template <int MODE>
__global__ void k(double* out, const double* cin, int iters)
{
    double a[8]; float f[8];
    const double c0 = cin[threadIdx.x & 1], c1 = cin[2 + (threadIdx.x & 1)];
    for (int i = 0; i < 8; i++) { a[i] = threadIdx.x * 1e-3 + i; f[i] = (float)a[i]; }
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int u = 0; u < 4; u++)
#pragma unroll
            for (int i = 0; i < 8; i++) {
                if (MODE == 0) a[i] = fma(a[i], c0, c1);
                if (MODE == 1) { f[i] = (float)a[i]; a[i] = a[i] + (double)f[i]; }        // F2F.F32.F64 + F2F.F64.F32 + DADD
                if (MODE == 2) { f[i] = (float)a[i]; a[i] = fma(a[i], c0, (double)__float_as_int(f[i])); } // one F2F.F32.F64 + I2F? avoid
                if (MODE == 3) { a[i] = fma(a[i], c0, c1); f[i] = (float)a[i] + f[i]; }   // DFMA + F2F.F32.F64 + FADD
            }
    }
    double s = 0;
    for (int i = 0; i < 8; i++) s += a[i] + f[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int MODE>
void run(const char* name, double* out, double* cin)
{
    const int iters = 2048, threads = 512;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE><<<148, threads>>>(out, cin, 8);
    cudaEventRecord(e0);
    k<MODE><<<148, threads>>>(out, cin, iters);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    const double per_smsp = (double)iters * 32 * (threads / 32 / 4.0);
    printf("%-44s %.3f ms  %.2f cycles(@%d MHz) per loop body per SMSP\n", name, ms, ms * 1e-3 * clk * 1e3 / per_smsp, clk / 1000);
}
int main()
{
    double *out, *cin;
    cudaMalloc(&out, 148 * 1024 * 8); cudaMalloc(&cin, 64);
    double h[4] = {0.999, 0.999, 1e-3, 1e-3}; cudaMemcpy(cin, h, 32, cudaMemcpyHostToDevice);
    run<0>("DFMA", out, cin);
    run<1>("F2F.F32.F64 + F2F.F64.F32 + DADD", out, cin);
    run<3>("DFMA + F2F.F32.F64 + FADD", out, cin);
    printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
