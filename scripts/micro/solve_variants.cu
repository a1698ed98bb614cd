// Micro-benchmark: variants of the 8x9 Gauss-Jordan elimination used by the tracker's LM step (one warp).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 solve_variants.cu -o solve_variants
#include <cstdio>
#include <cuda_runtime.h>
#define ITERS 256
__device__ __forceinline__ double shfl_d(double v, int src) { return __shfl_sync(0xffffffffu, v, src); }

template <int V>
__device__ __forceinline__ void solve(double& m0, double& m1, double& m2, int lane)
{
	const int i = lane & 7, g = (lane >> 3) == 3 ? 0 : (lane >> 3);
	const int i3 = i / 3;
	const double cand = i3 == 0 ? m0 : (i3 == 1 ? m1 : m2);
	const int myexp = (__double2hiint(cand) >> 20) & 0x7ff;
	int ek[8];
	if (V & 1)
	{
#pragma unroll
		for (int k = 0; k < 8; k++) ek[k] = __shfl_sync(0xffffffffu, myexp, k + 8 * (k % 3));
	}
#pragma unroll
	for (int k = 0; k < 8; k++)
	{
		const int e = (V & 1) ? ek[k] : __shfl_sync(0xffffffffu, myexp, k + 8 * (k % 3));
		const double sc = __hiloint2double((2046 - e) << 20, 0), isc = __hiloint2double(e << 20, 0);
		const bool self = (i == k);
		const double mk = (k / 3 == 0) ? m0 : ((k / 3 == 1) ? m1 : m2);
		const double fr = shfl_d(mk, i + 8 * (k % 3));
		double f, piv;
		const double t0 = m0 * sc, t1 = m1 * sc, t2 = m2 * sc;
		const double pivr = shfl_d(mk, k + 8 * (k % 3));
		const double p0 = shfl_d(m0, k + 8 * g), p1 = shfl_d(m1, k + 8 * g), p2 = shfl_d(m2, k + 8 * g);
		if (V & 2)
		{
			// arithmetic instead of selects: multipliers prepared from the lane's own flags
			const double one_self = self ? 1.0 : 0.0;       // off the chain
			f = fr * (self ? 0.0 : sc);
			piv = fma(one_self, isc - pivr, pivr);          // self: isc, else pivr (needs pivr: on chain)
		}
		else
		{
			f = self ? 0.0 : fr * sc;
			piv = self ? isc : pivr;
		}
		if (V & 4) { f = fr * sc; piv = pivr; }
		m0 = fma(-f, p0, piv * t0);
		m1 = fma(-f, p1, piv * t1);
		m2 = fma(-f, p2, piv * t2);
	}
}

// the same elimination with the pivot row / pivot column published through shared memory (double-buffered) instead of
// shuffles: one LDS.64 moves a double, a shuffle needs two SHFL
template <int V>
__device__ __forceinline__ void solve_smem(double& m0, double& m1, double& m2, int lane, double (*Ms)[8][10])
{
	const int i = lane & 7, g = (lane >> 3) == 3 ? 0 : (lane >> 3);
	const int i3 = i / 3;
	const double cand = i3 == 0 ? m0 : (i3 == 1 ? m1 : m2);
	const int myexp = (__double2hiint(cand) >> 20) & 0x7ff;
	int ek[8];
#pragma unroll
	for (int k = 0; k < 8; k++) ek[k] = __shfl_sync(0xffffffffu, myexp, k + 8 * (k % 3));
	Ms[0][i][g] = m0; Ms[0][i][g + 3] = m1; Ms[0][i][g + 6] = m2;
	__syncwarp();
#pragma unroll
	for (int k = 0; k < 8; k++)
	{
		const int e = ek[k];
		const double sc = __hiloint2double((2046 - e) << 20, 0), isc = __hiloint2double(e << 20, 0);
		const bool self = (i == k);
		const double t0 = m0 * sc, t1 = m1 * sc, t2 = m2 * sc;
		double (*B)[10] = Ms[k & 1];
		const double pivr = B[k][k], fr = B[i][k];
		const double p0 = B[k][g], p1 = B[k][g + 3], p2 = B[k][g + 6];
		const double f = self ? 0.0 : fr * sc;
		const double piv = self ? isc : pivr;
		m0 = fma(-f, p0, piv * t0);
		m1 = fma(-f, p1, piv * t1);
		m2 = fma(-f, p2, piv * t2);
		if (k < 7)
		{
			double (*N)[10] = Ms[(k + 1) & 1];
			if (V == 0)
			{
				// publish only what step k+1 reads: row k+1 and column k+1
				if (i == k + 1) { N[i][g] = m0; N[i][g + 3] = m1; N[i][g + 6] = m2; }
				if (g == (k + 1) % 3) N[i][k + 1] = ((k + 1) / 3 == 0) ? m0 : (((k + 1) / 3 == 1) ? m1 : m2);
			}
			else { N[i][g] = m0; N[i][g + 3] = m1; N[i][g + 6] = m2; }
			__syncwarp();
		}
	}
}

template <int V>
__global__ void bench_smem(double* out, long long* cyc, double seed)
{
	__shared__ double Ms[2][8][10];
	const int lane = threadIdx.x;
	double m0 = seed + lane, m1 = seed * 2 + lane, m2 = seed * 3 - lane;
	long long t0 = clock64();
	for (int it = 0; it < ITERS; it++)
	{
		solve_smem<V>(m0, m1, m2, lane, Ms);
		m0 = m0 * 1e-3 + 2.0; m1 = m1 * 1e-3 + 3.0; m2 = m2 * 1e-3 + 5.0;
	}
	long long t1 = clock64();
	out[lane] = m0 + m1 + m2;
	if (lane == 0) cyc[0] = t1 - t0;
}
template <int V> void run_smem(const char* name)
{
	double* out; long long* cyc;
	cudaMalloc(&out, 32 * sizeof(double)); cudaMalloc(&cyc, 8);
	bench_smem<V><<<1, 32>>>(out, cyc, 1.37);
	bench_smem<V><<<1, 32>>>(out, cyc, 1.37);
	long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
	printf("%-50s %.0f cycles per 8-step elimination\n", name, (double)h / ITERS);
	cudaFree(out); cudaFree(cyc);
}

template <int V>
__global__ void bench(double* out, long long* cyc, double seed)
{
	const int lane = threadIdx.x;
	double m0 = seed + lane, m1 = seed * 2 + lane, m2 = seed * 3 - lane;
	long long t0 = clock64();
	for (int it = 0; it < ITERS; it++)
	{
		solve<V>(m0, m1, m2, lane);
		m0 = m0 * 1e-3 + 2.0; m1 = m1 * 1e-3 + 3.0; m2 = m2 * 1e-3 + 5.0;
	}
	long long t1 = clock64();
	out[lane] = m0 + m1 + m2;
	if (lane == 0) cyc[0] = t1 - t0;
}
template <int V> void run(const char* name)
{
	double* out; long long* cyc;
	cudaMalloc(&out, 32 * sizeof(double)); cudaMalloc(&cyc, 8);
	bench<V><<<1, 32>>>(out, cyc, 1.37);
	bench<V><<<1, 32>>>(out, cyc, 1.37);
	long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
	printf("%-50s %.0f cycles per 8-step elimination\n", name, (double)h / ITERS);
	cudaFree(out); cudaFree(cyc);
}
int main()
{
	run<0>("selects, exponent shuffles in the loop");
	run<1>("selects, exponent shuffles hoisted");
	run<2>("arithmetic self handling");
	run<3>("arithmetic self handling, hoisted");
	run<4>("no self handling (wrong result; floor)");
	run<5>("no self handling, hoisted");
	run_smem<0>("shared memory, publish row/column k+1 only");
	run_smem<1>("shared memory, publish everything");
	return 0;
}
