// Micro-benchmark: dependent-issue latency (cycles) of the operations on the tracker's LM chain, one warp on one SM.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 op_latency.cu -o op_latency
#include <cstdio>
#include <cuda_runtime.h>
#define ITERS 512
template <int OP>
__global__ void lat(double* out, long long* cyc, double seed)
{
	__shared__ double sm[4];
	double a = seed + threadIdx.x * 1e-3, b = 1.0000001, c = 1e-9, a2 = a + 1, a3 = a + 2, a4 = a + 3;
	float f = (float)seed + threadIdx.x;
	int lane = threadIdx.x & 31;
	long long t0 = clock64();
#pragma unroll 16
	for (int i = 0; i < ITERS; i++)
	{
		if (OP == 0) a = fma(a, b, c);
		if (OP == 1) a = a * b;
		if (OP == 2) a = a + c;
		if (OP == 3) a = __shfl_sync(0xffffffffu, a, (lane + 1) & 31);
		if (OP == 4) f = __shfl_sync(0xffffffffu, f, (lane + 1) & 31);
		if (OP == 5) f = fmaf(f, 1.0000001f, 1e-9f);
		if (OP == 6) { a = (double)f; f = (float)a + 1.f; }
		if (OP == 7) { double r = (double)__frcp_rn((float)a); r = r * fma(-a, r, 2.0); r = r * fma(-a, r, 2.0); a = r + 1.5; }
		if (OP == 8) a = 1.0 / a + 1.5;
		if (OP == 9) f = __fdiv_rn(1.0f, f) + 1.5f;
		if (OP == 10) f = sqrtf(f) + 1.5f;
		if (OP == 11) a = __hiloint2double((2046 - ((__double2hiint(a) >> 20) & 0x7ff)) << 20, 0) + a;
		if (OP == 12) f = __fdividef(1.0f, f) + 1.5f;
		if (OP == 13) a = fma(__shfl_sync(0xffffffffu, a, (lane + 1) & 31), b, c);
		if (OP == 14)
		{
			// one elimination step of the tracker's 8x9 solve: three entries per lane, pivot row by shuffle
			const double sc = 0.5;
			const double t0 = a * sc, t1 = a2 * sc, t2 = a3 * sc, fm = a * sc;
			const double piv = __shfl_sync(0xffffffffu, a, i & 7);
			const double p0 = __shfl_sync(0xffffffffu, a, (i & 7) + 8), p1 = __shfl_sync(0xffffffffu, a2, (i & 7) + 8), p2 = __shfl_sync(0xffffffffu, a3, (i & 7) + 8);
			a = fma(-fm, p0, piv * t0); a2 = fma(-fm, p1, piv * t1); a3 = fma(-fm, p2, piv * t2);
		}
		if (OP == 15)
		{
			// the same through shared memory: the pivot lane stores its entries, everybody loads them
			const double sc = 0.5;
			const double t0 = a * sc, t1 = a2 * sc, t2 = a3 * sc, fm = a * sc;
			if (lane == (i & 7)) { sm[0] = a; sm[1] = a2; sm[2] = a3; }
			__syncwarp();
			const double piv = sm[0], p0 = sm[0], p1 = sm[1], p2 = sm[2];
			__syncwarp();
			a = fma(-fm, p0, piv * t0); a2 = fma(-fm, p1, piv * t1); a3 = fma(-fm, p2, piv * t2);
		}
		if (OP == 16) { a = fma(a, b, c); a2 = fma(a2, b, c); a3 = fma(a3, b, c); a4 = fma(a4, b, c); }
		if (OP == 17) { float t = __shfl_sync(0xffffffffu, f, (lane + 1) & 31); f = fmaf(t, 1.0000001f, 1e-9f); }
	}
	long long t1 = clock64();
	out[threadIdx.x] = a + f + a2 + a3 + a4;
	if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
template <int OP> void run(const char* name)
{
	double* out; long long* cyc;
	cudaMalloc(&out, 1024 * sizeof(double)); cudaMalloc(&cyc, 8);
	lat<OP><<<1, 32>>>(out, cyc, 1.37);
	lat<OP><<<1, 32>>>(out, cyc, 1.37);
	long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
	printf("%-34s %.1f cycles per dependent step\n", name, (double)h / ITERS);
	cudaFree(out); cudaFree(cyc);
}
int main()
{
	run<0>("DFMA"); run<1>("DMUL"); run<2>("DADD"); run<3>("shfl double (2 SHFL)"); run<4>("shfl float"); run<5>("FFMA");
	run<6>("F2F f32->f64->f32 (+FADD)"); run<7>("fast_rcp (frcp + 2 Newton) + DADD"); run<8>("IEEE double division + DADD");
	run<13>("shfl double -> DFMA"); run<14>("elimination step via shuffles"); run<15>("elimination step via smem"); run<16>("4 independent DFMA"); run<17>("shfl float -> FFMA");
	run<9>("IEEE float division + FADD"); run<10>("sqrtf + FADD"); run<11>("exponent extract + DADD"); run<12>("__fdividef + FADD");
	return 0;
}
