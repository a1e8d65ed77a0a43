// standalone debug harness for acq_fft.cuh (not part of the product)
#include "acq_fft.cuh"
#include <cstdio>
#include <cmath>
#include <vector>
using namespace gb200;
__global__ void k(FftPlan pl, const float2* in, const float2* gA, const float2* gB, float2* out)
{
    extern __shared__ __align__(16) float2 sm[];
    float2* x = sm; float2* sA = sm + pl.m; float2* sB = sA + pl.n_hi;
    fft_load_tables(pl, gA, gB, sA, sB);
    for (int i = threadIdx.x; i < pl.m; i += blockDim.x) x[i] = in[i];
    __syncthreads();
    int n_s = pl.m;
    for (int s = 0; s < pl.n_stages; s++)
        {
            if (threadIdx.x == 0) printf("stage %d R=%d n_s=%d inv=%u\n", s, pl.radix[s], n_s, pl.inv_ms[s]);
            fft_smem_stage_any<false>(pl.radix[s], x, pl.m, n_s, pl.inv_ms[s], pl.n, sA, sB);
            n_s /= pl.radix[s];
            if (threadIdx.x == 0) for (int i = 0; i < pl.m; i++) printf("  x[%d]=%f %f\n", i, x[i].x, x[i].y);
            __syncthreads();
        }
    for (int i = threadIdx.x; i < pl.m; i += blockDim.x) out[i] = x[i];
}
int main()
{
    FftPlan pl{}; pl.n = 8; pl.m = 8; pl.g_radix = 1; pl.n_stages = 2; pl.n_hi = 1; pl.radix[0] = 4; pl.radix[1] = 2;
    pl.inv_ms[0] = (unsigned)((0x100000000ull + 1) / 2); pl.inv_ms[1] = 0;
    std::vector<float2> A(1), B(256), in(8);
    A[0] = make_float2(1, 0);
    for (int i = 0; i < 256; i++) B[i] = make_float2((float)cos(-2 * M_PI * i / 8), (float)sin(-2 * M_PI * i / 8));
    for (int i = 0; i < 8; i++) in[i] = make_float2(i + 1, 0.5f * i);
    float2 *dA, *dB, *di, *dout;
    cudaMalloc(&dA, 8); cudaMalloc(&dB, 2048); cudaMalloc(&di, 64); cudaMalloc(&dout, 64);
    cudaMemcpy(dA, A.data(), 8, cudaMemcpyHostToDevice); cudaMemcpy(dB, B.data(), 2048, cudaMemcpyHostToDevice); cudaMemcpy(di, in.data(), 64, cudaMemcpyHostToDevice);
    k<<<1, 128, 64 + 257 * 8>>>(pl, di, dA, dB, dout);
    printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
