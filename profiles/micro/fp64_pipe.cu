// Microbenchmark (tuning aid, not product): what the FP64 pipe of one SMSP of sm_100a does with dependent DFMA chains.
// Measured in SM cycles with clock64() inside the kernel, so the numbers do not depend on the clock the GPU runs at.
//   cycles per DFMA per SMSP as a function of (warps per SMSP, independent chains per thread)
//   -> issue cost of a DFMA (the plateau) and its dependent-issue latency (chains x warps needed to reach it)
//   cost of F2F.F32.F64, F2I.S32.F32, I2F, DADD, 64-bit SHFL relative to it
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_pipe fp64_pipe.cu && ./fp64_pipe
#include <cstdio>
#include <cuda_runtime.h>

template <int CH>
__global__ void dfma_chains(double* out, long long* cyc, const double* cin, int iters)
{
    double a[CH];
    const double c0 = cin[0], c1 = cin[1];
#pragma unroll
    for (int i = 0; i < CH; i++) a[i] = threadIdx.x * 1e-3 + i;
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int u = 0; u < 8; u++)
#pragma unroll
            for (int i = 0; i < CH; i++) a[i] = fma(a[i], c0, c1);
    }
    const long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int i = 0; i < CH; i++) s += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if ((threadIdx.x & 31) == 0) cyc[blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32] = t1 - t0;
}

// MODE 0: F2F.F32.F64  1: F2I(F32)  2: I2F  3: DADD  4: 64-bit SHFL  5: F2F + DFMA mix (1 F2F per 4 DFMA)
template <int MODE>
__global__ void misc_ops(double* out, long long* cyc, const double* cin, int iters)
{
    double a[8]; float f[8]; int q[8];
    const double c0 = cin[0], c1 = cin[1];
#pragma unroll
    for (int i = 0; i < 8; i++) { a[i] = threadIdx.x * 1e-3 + i; f[i] = (float)i + threadIdx.x; q[i] = i + threadIdx.x; }
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int u = 0; u < 8; u++)
#pragma unroll
            for (int i = 0; i < 8; i++) {
                if (MODE == 0) { f[i] = (float)a[i]; a[i] = __hiloint2double(__double2hiint(a[i]) ^ __float_as_int(f[i]), __double2loint(a[i])); }
                if (MODE == 1) { q[i] = __float2int_rz(f[i]); f[i] = __int_as_float(__float_as_int(f[i]) ^ (q[i] & 1)); }
                if (MODE == 2) { f[i] = (float)q[i]; q[i] ^= __float_as_int(f[i]) & 3; }
                if (MODE == 3) { a[i] = a[i] + c1; }
                if (MODE == 4) { a[i] = __shfl_up_sync(0xffffffffu, a[i], 1); }
                if (MODE == 5) { a[i] = fma(a[i], c0, c1); if ((i & 3) == 0) { f[i] = (float)a[i]; q[i] ^= __float_as_int(f[i]); } }
            }
    }
    const long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s += a[i] + f[i] + q[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if ((threadIdx.x & 31) == 0) cyc[blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32] = t1 - t0;
}

static double mean_cycles(long long* d_cyc, int n)
{
    static long long h[148 * 64];
    cudaMemcpy(h, d_cyc, n * sizeof(long long), cudaMemcpyDeviceToHost);
    double s = 0;
    for (int i = 0; i < n; i++) s += (double)h[i];
    return s / n;
}

template <int CH>
void run_chains(double* out, long long* cyc, double* cin)
{
    const int iters = 512;
    for (int wps : {1, 2, 3, 4, 6, 8}) {
        const int threads = 32 * 4 * wps;
        dfma_chains<CH><<<148, threads>>>(out, cyc, cin, 8);
        dfma_chains<CH><<<148, threads>>>(out, cyc, cin, iters);
        cudaDeviceSynchronize();
        const double c = mean_cycles(cyc, 148 * 4 * wps);
        const double per = c / ((double)iters * 8 * CH * wps);      // cycles per warp-DFMA per SMSP
        printf("dfma chains/thread %d  warps/SMSP %d  in flight/SMSP %2d : %.2f cycles per DFMA per SMSP  (one chain step every %.1f cycles)\n",
               CH, wps, CH * wps, per, c / ((double)iters * 8));
    }
}

template <int MODE>
void run_misc(const char* name, int ops_per_iter, double* out, long long* cyc, double* cin)
{
    const int iters = 256;
    for (int wps : {1, 4}) {
        const int threads = 32 * 4 * wps;
        misc_ops<MODE><<<148, threads>>>(out, cyc, cin, 8);
        misc_ops<MODE><<<148, threads>>>(out, cyc, cin, iters);
        cudaDeviceSynchronize();
        const double c = mean_cycles(cyc, 148 * 4 * wps);
        printf("%-34s warps/SMSP %d : %.2f cycles per warp-op per SMSP\n", name, wps, c / ((double)iters * ops_per_iter * wps));
    }
}

int main()
{
    double *out, *cin; long long* cyc;
    cudaMalloc(&out, 148 * 1024 * 8); cudaMalloc(&cin, 64); cudaMalloc(&cyc, 148 * 64 * 8);
    double h[2] = {0.999, 1e-3}; cudaMemcpy(cin, h, 16, cudaMemcpyHostToDevice);
    run_chains<1>(out, cyc, cin);
    run_chains<2>(out, cyc, cin);
    run_chains<4>(out, cyc, cin);
    run_chains<6>(out, cyc, cin);
    run_chains<8>(out, cyc, cin);
    run_chains<12>(out, cyc, cin);
    run_misc<0>("F2F.F32.F64 (+2 int ops)", 64, out, cyc, cin);
    run_misc<1>("F2I.S32.F32 (+2 int ops)", 64, out, cyc, cin);
    run_misc<2>("I2F.F32.S32 (+2 int ops)", 64, out, cyc, cin);
    run_misc<3>("DADD", 64, out, cyc, cin);
    run_misc<4>("SHFL.UP 64-bit (2 SHFL)", 64, out, cyc, cin);
    run_misc<5>("DFMA + 1/4 F2F (+ 1/4 int)", 64, out, cyc, cin);
    printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
