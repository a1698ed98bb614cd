// Micro-benchmark: FP64 / FP32 issue rate and dependent latency on one SM (build: nvcc -arch=sm_100a -O3 fp64_rate.cu -o fp64_rate)
#include <cstdio>
#include <cuda_runtime.h>
template <typename T, int ILP>
__global__ void rate(T* out, int iters, long long* cyc)
{
	T a[ILP];
	for (int i = 0; i < ILP; i++) a[i] = (T)(threadIdx.x + i) * (T)1e-3;
	const T m = (T)1.0000001, c = (T)1e-9;
	__syncthreads();
	long long t0 = clock64();
	for (int it = 0; it < iters; it++)
	{
#pragma unroll
		for (int i = 0; i < ILP; i++) a[i] = a[i] * m + c;
	}
	long long t1 = clock64();
	T s = 0;
	for (int i = 0; i < ILP; i++) s += a[i];
	out[blockIdx.x * blockDim.x + threadIdx.x] = s;
	if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
template <typename T, int ILP> void run(const char* name, int threads)
{
	T* out; long long* cyc;
	cudaMalloc(&out, 1024 * sizeof(T)); cudaMalloc(&cyc, 8);
	const int iters = 2000;
	rate<T, ILP><<<1, threads>>>(out, iters, cyc);
	rate<T, ILP><<<1, threads>>>(out, iters, cyc);
	long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
	printf("%s ILP %d threads %4d: %.2f cycles per dependent step, %.2f FMA/clk/SM\n", name, ILP, threads, (double)h / iters, (double)threads * ILP * iters / h);
	cudaFree(out); cudaFree(cyc);
}
int main()
{
	run<double, 1>("f64", 32); run<double, 8>("f64", 32); run<double, 8>("f64", 128); run<double, 8>("f64", 256); run<double, 8>("f64", 1024);
	run<float, 1>("f32", 32); run<float, 8>("f32", 128); run<float, 8>("f32", 1024);
	return 0;
}
