// Micro-benchmark: L2 and HBM read bandwidth of the whole GPU (build: nvcc -arch=sm_100a -O3 l2_bw.cu -o l2_bw).
// Also the store-only HBM bandwidth (16-byte stores over 2 GB).
// A buffer that fits in L2 (32 MB) resp. one far larger than L2 (2 GB) is summed repeatedly with 16-byte loads.
#include <cstdio>
#include <cuda_runtime.h>
__global__ void __launch_bounds__(256) sum_kernel(const float4* __restrict__ p, size_t n, int reps, float* out)
{
	float4 acc = make_float4(0, 0, 0, 0);
	const size_t stride = (size_t)gridDim.x * blockDim.x;
	for (int r = 0; r < reps; r++)
		for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
		{
			const float4 v = __ldcg(p + i);
			acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
		}
	if (acc.x + acc.y + acc.z + acc.w == 12345.678f) out[0] = acc.x;
}
__global__ void __launch_bounds__(256) fill_kernel(float4* __restrict__ p, size_t n, int reps)
{
	const size_t stride = (size_t)gridDim.x * blockDim.x;
	for (int r = 0; r < reps; r++)
		for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) p[i] = make_float4((float)r, 1.f, 2.f, 3.f);
}
static double run_fill(size_t bytes, int reps)
{
	float4* p;
	cudaMalloc(&p, bytes);
	int sm = 0; cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, 0);
	cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
	fill_kernel<<<sm * 8, 256>>>(p, bytes / 16, 1);
	cudaEventRecord(e0);
	fill_kernel<<<sm * 8, 256>>>(p, bytes / 16, reps);
	cudaEventRecord(e1); cudaEventSynchronize(e1);
	float ms; cudaEventElapsedTime(&ms, e0, e1);
	cudaFree(p);
	return (double)bytes * reps / (ms * 1e-3) / 1e9;
}
static double run(size_t bytes, int reps)
{
	float4* p; float* o;
	cudaMalloc(&p, bytes); cudaMalloc(&o, 4);
	cudaMemset(p, 0, bytes);
	int sm = 0; cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, 0);
	cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
	sum_kernel<<<sm * 8, 256>>>(p, bytes / 16, 2, o);
	cudaEventRecord(e0);
	sum_kernel<<<sm * 8, 256>>>(p, bytes / 16, reps, o);
	cudaEventRecord(e1); cudaEventSynchronize(e1);
	float ms; cudaEventElapsedTime(&ms, e0, e1);
	cudaFree(p); cudaFree(o);
	return (double)bytes * reps / (ms * 1e-3) / 1e9;
}
int main()
{
	printf("{\"l2_read_GBps_32MB\": %.1f, \"l2_read_GBps_64MB\": %.1f, \"hbm_read_GBps_2GB\": %.1f, \"hbm_write_GBps_2GB\": %.1f}\n", run(32ull << 20, 200), run(64ull << 20, 100), run(2048ull << 20, 5), run_fill(2048ull << 20, 5));
	return 0;
}
