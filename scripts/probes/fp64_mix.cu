// Can other instructions issue in the second cycle a DFMA occupies the FP64 pipe?  Per thread: NF independent DFMA chains
// and NI independent integer (IMAD) chains, interleaved; 16 warps per SM; cycles per loop body per SMSP.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_mix fp64_mix.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int NF, int NI, int NL>
__global__ void mix(double* out, int iters, long long* cyc, const double* tabg)
{
    __shared__ double tab[512];
    for (int i = threadIdx.x; i < 512; i += blockDim.x) tab[i] = tabg ? tabg[i] : 1.0 + i;
    double a[NF > 0 ? NF : 1];
    int b[NI > 0 ? NI : 1];
    double l[NL > 0 ? NL : 1];
    for (int k = 0; k < NF; k++) a[k] = 1.0 + threadIdx.x * 1e-9 + k;
    for (int k = 0; k < NI; k++) b[k] = threadIdx.x + k;
    for (int k = 0; k < NL; k++) l[k] = 0;
    const double m = 1.0000001, c = 1e-9;
    __syncthreads();
    long long t0 = clock64();
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int u = 0; u < 8; u++) {
#pragma unroll
            for (int k = 0; k < (NF > NI ? NF : NI); k++) {
                if (k < NF) a[k] = fma(a[k], m, c);
                if (k < NI) b[k] = b[k] * 3 + i;
                if (k < NL) l[k] += tab[(b[k % (NI > 0 ? NI : 1)] + k) & 511];
            }
        }
    }
    long long t1 = clock64();
    double s = 0;
    for (int k = 0; k < NF; k++) s += a[k];
    for (int k = 0; k < NI; k++) s += b[k];
    for (int k = 0; k < NL; k++) s += l[k];
    if (s == 123.456) out[0] = s;
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
}

template <int NF, int NI, int NL>
void run(int warps)
{
    double* out;
    long long* cyc;
    cudaMalloc(&out, 8);
    cudaMalloc(&cyc, 8);
    const int iters = 2000;
    mix<NF, NI, NL><<<1, warps * 32>>>(out, 10, cyc, nullptr);
    mix<NF, NI, NL><<<1, warps * 32>>>(out, iters, cyc, nullptr);
    long long h;
    cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    // per SMSP: warps/4 warps each issuing iters*8*(NF+NI+..)
    const double per_body = (double)h / ((double)iters * 8) / (warps / 4.0);
    printf("warps %2d  DFMA %d  IMAD %d  LDS %d : %.2f cycles per (body of one warp) per SMSP\n", warps, NF, NI, NL, per_body);
    cudaFree(out);
    cudaFree(cyc);
}

int main()
{
    run<4, 0, 0>(16);
    run<0, 4, 0>(16);
    run<4, 4, 0>(16);
    run<4, 8, 0>(16);
    run<4, 2, 0>(16);
    run<2, 4, 0>(16);
    run<4, 4, 2>(16);
    run<4, 4, 4>(16);
    run<4, 0, 4>(16);
    return 0;
}
