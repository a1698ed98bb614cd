// Micro-benchmark: issue cost of shuffles / shared-memory accesses next to a dependent FP64 chain (one warp).
#include <cstdio>
#include <cuda_runtime.h>
#define ITERS 512
template <int N, int MODE>
__global__ void bench(double* out, long long* cyc, double seed)
{
	__shared__ double sm[64];
	const int lane = threadIdx.x;
	double a = seed + lane, acc = 0;
	double v[12];
	for (int k = 0; k < 12; k++) v[k] = seed * k + lane;
	sm[lane] = a; sm[32 + lane] = a;
	__syncwarp();
	long long t0 = clock64();
#pragma unroll 4
	for (int it = 0; it < ITERS; it++)
	{
		a = fma(a, 1.0000001, 1e-9);                       // the dependent chain
		if (MODE == 0)
		{
#pragma unroll
			for (int k = 0; k < N; k++) v[k] = __shfl_sync(0xffffffffu, v[k], (lane + k + 1) & 31);      // N independent double shuffles
		}
		if (MODE == 1)
		{
#pragma unroll
			for (int k = 0; k < N; k++) v[k] += sm[(lane + k + it) & 63];                               // N independent LDS.64
		}
		if (MODE == 2)
		{
#pragma unroll
			for (int k = 0; k < N; k++) { float f = __shfl_sync(0xffffffffu, (float)v[k], (lane + k + 1) & 31); v[k] = f; }
		}
	}
	long long t1 = clock64();
	for (int k = 0; k < 12; k++) acc += v[k];
	out[lane] = a + acc;
	if (lane == 0) cyc[0] = t1 - t0;
}
template <int N, int MODE> void run(const char* name)
{
	double* out; long long* cyc;
	cudaMalloc(&out, 32 * sizeof(double)); cudaMalloc(&cyc, 8);
	bench<N, MODE><<<1, 32>>>(out, cyc, 1.37);
	bench<N, MODE><<<1, 32>>>(out, cyc, 1.37);
	long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
	printf("%-40s N=%2d: %.1f cycles per iteration\n", name, N, (double)h / ITERS);
	cudaFree(out); cudaFree(cyc);
}
int main()
{
	run<0, 0>("DFMA only"); run<1, 0>("double shuffles"); run<2, 0>("double shuffles"); run<4, 0>("double shuffles"); run<6, 0>("double shuffles"); run<10, 0>("double shuffles");
	run<2, 1>("LDS.64 + DADD"); run<4, 1>("LDS.64 + DADD"); run<6, 1>("LDS.64 + DADD"); run<10, 1>("LDS.64 + DADD");
	return 0;
}
