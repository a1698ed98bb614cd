// Micro-benchmark: FP64 / FP32 / conversion / select issue cost for ONE warp (cycles per instruction when independent).
#include <cstdio>
#include <cuda_runtime.h>
#define ITERS 256
template <int N, int MODE>
__global__ void bench(double* out, long long* cyc, double seed)
{
	const int lane = threadIdx.x;
	double v[32]; float f[32];
	for (int k = 0; k < 32; k++) { v[k] = seed * (k + 1) + lane; f[k] = (float)v[k]; }
	long long t0 = clock64();
#pragma unroll 2
	for (int it = 0; it < ITERS; it++)
	{
#pragma unroll
		for (int k = 0; k < N; k++)
		{
			if (MODE == 0) v[k] = fma(v[k], 1.0000001, 1e-9);
			if (MODE == 1) f[k] = fmaf(f[k], 1.0000001f, 1e-9f);
			if (MODE == 2) v[k] = (double)f[k] + v[k];             // F2F + DADD
			if (MODE == 3) v[k] = (lane & (1 << (it & 3))) ? v[k] : v[(k + 1) % N];   // double select
			if (MODE == 4) v[k] = v[k] * 1.0000001;
		}
	}
	long long t1 = clock64();
	double acc = 0;
	for (int k = 0; k < 32; k++) acc += v[k] + f[k];
	out[lane] = acc;
	if (lane == 0) cyc[0] = t1 - t0;
}
template <int N, int MODE> void run(const char* name)
{
	double* out; long long* cyc;
	cudaMalloc(&out, 32 * sizeof(double)); cudaMalloc(&cyc, 8);
	bench<N, MODE><<<1, 32>>>(out, cyc, 1.37);
	bench<N, MODE><<<1, 32>>>(out, cyc, 1.37);
	long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
	printf("%-28s N=%2d independent: %.1f cycles per iteration = %.2f per instruction\n", name, N, (double)h / ITERS, (double)h / ITERS / N);
	cudaFree(out); cudaFree(cyc);
}
int main()
{
	run<8, 0>("DFMA"); run<16, 0>("DFMA"); run<32, 0>("DFMA");
	run<16, 4>("DMUL");
	run<16, 1>("FFMA"); run<32, 1>("FFMA");
	run<16, 2>("F2F.F64.F32 + DADD");
	run<16, 3>("double select");
	return 0;
}
