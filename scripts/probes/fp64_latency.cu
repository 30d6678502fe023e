// FP64 pipe characterisation on one SM: dependent-issue latency of DFMA and sustained rate as a function of resident warps
// and independent chains per warp.   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_latency fp64_latency.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int ILP>
__global__ void chain(double* out, int iters, long long* cyc)
{
    double a[ILP];
#pragma unroll
    for (int k = 0; k < ILP; k++) a[k] = 1.0 + threadIdx.x * 1e-9 + k;
    const double m = 1.0000001, c = 1e-9;
    __syncthreads();
    long long t0 = clock64();
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int u = 0; u < 16; u++)
#pragma unroll
            for (int k = 0; k < ILP; k++) a[k] = fma(a[k], m, c);
    }
    long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int k = 0; k < ILP; k++) s += a[k];
    if (s == 123.456) out[0] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int ILP>
void run(int warps)
{
    double* out;
    long long* cyc;
    cudaMalloc(&out, 8);
    cudaMalloc(&cyc, 8);
    const int iters = 2000;
    chain<ILP><<<1, warps * 32>>>(out, 10, cyc);
    chain<ILP><<<1, warps * 32>>>(out, iters, cyc);
    long long h;
    cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    const double n_inst = (double)iters * 16 * ILP;   // per warp
    printf("warps/SM %2d  ILP %d : %.2f cycles per DFMA per warp, %.3f DFMA warp-inst/cycle/SM\n", warps, ILP, h / n_inst,
           n_inst * warps / h);
    cudaFree(out);
    cudaFree(cyc);
}

int main()
{
    for (int w : {1, 4, 8, 16, 24, 32}) { run<1>(w); run<2>(w); run<4>(w); run<8>(w); }
    return 0;
}
