// Micro-benchmark: cost of barrier.cluster arrive / wait for a 16-CTA cluster (first use and repeated), seen by thread 0 of
// every CTA.  build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 cluster_barrier.cu -o cluster_barrier
#include <cstdio>
#include <cuda_runtime.h>
#include <cooperative_groups.h>
namespace cg = cooperative_groups;

__global__ void __launch_bounds__(256, 1) k(long long* out, int variant)
{
	extern __shared__ float4 dyn[];
	cg::cluster_group cluster = cg::this_cluster();
	const int rank = cluster.block_rank();
	long long t[8];
	unsigned long long g0; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g0));
	if (variant >= 2) { long long c0 = clock64(); while (clock64() - c0 < 8000) { } }   // arrive late: is the first arrive's cost a start-up cost?
	t[0] = clock64();
	if (variant == 0) asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
	else asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
	t[1] = clock64();
	asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
	t[2] = clock64();
	asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
	t[3] = clock64();
	asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
	t[4] = clock64();
	cluster.sync();
	t[5] = clock64();
	__syncthreads();
	t[6] = clock64();
	unsigned long long g; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g));
	t[7] = (long long)g;
	if (threadIdx.x == 0) { for (int i = 0; i < 8; i++) out[rank * 8 + i] = t[i]; out[128 + rank] = (long long)g0; }
	if (dyn[threadIdx.x].x == 123.f) out[0] = 0;
}

int main()
{
	long long* d; cudaMalloc(&d, 18 * 8 * 8);
	cudaFuncSetAttribute(k, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
	cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
	for (int variant = 0; variant < 3; variant++)
		for (int rep = 0; rep < 3; rep++)
		{
			cudaLaunchConfig_t cfg = {};
			cfg.gridDim = dim3(16); cfg.blockDim = dim3(256); cfg.dynamicSmemBytes = 16 * 1024;
			cudaLaunchAttribute at[1];
			at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = 16; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
			cfg.attrs = at; cfg.numAttrs = 1;
			cudaError_t e = cudaLaunchKernelEx(&cfg, k, d, variant);
			cudaDeviceSynchronize();
			long long h[144]; cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
			long long s0 = h[128], s1 = h[128];
			for (int r = 0; r < 16; r++) { if (h[128 + r] < s0) s0 = h[128 + r]; if (h[128 + r] > s1) s1 = h[128 + r]; }
			long long gmin = h[7], gmax = h[7];
			for (int r = 0; r < 16; r++) { if (h[r * 8 + 7] < gmin) gmin = h[r * 8 + 7]; if (h[r * 8 + 7] > gmax) gmax = h[r * 8 + 7]; }
			printf("%s rep %d (%s): CTA0 arrive %lld wait %lld | 2nd arrive %lld wait %lld | cluster.sync %lld | syncthreads %lld | end-time spread over CTAs %lld ns, start-time spread %lld ns\n",
			       variant == 0 ? "relaxed" : (variant == 1 ? "release" : "relaxed after 8000 cycles"), rep, cudaGetErrorString(e), h[1] - h[0], h[2] - h[1], h[3] - h[2], h[4] - h[3], h[5] - h[4], h[6] - h[5], gmax - gmin, s1 - s0);
			printf("   CTA15: arrive %lld wait %lld | 2nd arrive %lld wait %lld\n", h[121] - h[120], h[122] - h[121], h[123] - h[122], h[124] - h[123]);
		}
	return 0;
}
