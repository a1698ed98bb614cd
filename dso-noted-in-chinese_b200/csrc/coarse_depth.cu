// Next row N3 (first half): CoarseTracker::makeCoarseDepthL0 (FullSystem/CoarseTracker.cpp:135-292) on the device --
// the producer of the tracker's reference cloud.  Splat of the active points' projections into the newest key frame,
// 2x2 sum pyramid, one dilation pass per level, normalisation and compaction of the valid pixels of every level into
// (u, v, idepth, color) records in raster order (= the order of the reference's nested loops, so the tracker's
// point -> thread assignment is the one the uploaded-cloud path gets).
//
//   cd_hits / cd_splat / cd_splat_serial  :141-158   one thread per point, float atomics where at most two points share a
//                                  pixel (exact: addition commutes), the rest added in point order, one thread per pixel
//   cd_down_kernel      :161-186   ((a + b) + c) + d per output pixel, level by level
//   cd_dilate_kernel    :190-242   diagonal neighbours on levels 0-1, 4-neighbours above; reads only pixels that had a
//                                  weight before the pass, writes only pixels that had none: no ordering hazard
//   cd_count / cd_scan / cd_write  :246-290  deterministic stream compaction (block counts, scan, ranked writes)
// All of it is a few MB of streaming FP32 traffic once per key frame: HBM/L2-bound, far off the frame path.
#include "common.cuh"
#include <vector>

static size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

struct CDLevels
{
	int levels;
	int w[MAXL], h[MAXL];
	float* idepth[MAXL]; float* wsum[MAXL]; float* wsum2[MAXL];
	const float4* ref[MAXL];
	float4* pc[MAXL];
	int block_start[MAXL + 1];         // blocks of 1024 pixels, all levels back to back
};

struct CoarseDepthWS
{
	size_t cap_px, cap_pts; int cap_blocks;
	float* base; float* d_pts; int* d_block_count; int* d_block_off; int* d_pc_n; int* h_pc_n;
};

// :141-158.  Float addition commutes, so one or two points on a pixel can be added with atomics and still give the
// reference's sequential result bit for bit; pixels hit by three or more points (counted first) are left to
// cd_splat_serial_kernel, which adds them in point order.
__device__ __forceinline__ bool cd_point(int k, const float* __restrict__ cu, const float* __restrict__ cv, const float* __restrict__ cid,
                                         const float* __restrict__ hdi, int w, int h, int& pix, float& idw, float& weight)
{
	const int u = (int)fadd(cu[k], 0.5f), v = (int)fadd(cv[k], 0.5f);                 // :149-150
	if (u < 0 || v < 0 || u >= w || v >= h) return false;                              // the reference trusts its projections
	weight = __fsqrt_rn((float)(1e-3 / ((double)hdi[k] + 1e-12)));                     // :152 (double quotient, sqrtf)
	idw = fmul(cid[k], weight);
	pix = u + w * v;
	return true;
}

__global__ void cd_hits_kernel(int n, const float* __restrict__ cu, const float* __restrict__ cv, int w, int h, int* hits)
{
	const int k = blockIdx.x * blockDim.x + threadIdx.x;
	if (k >= n) return;
	const int u = (int)fadd(cu[k], 0.5f), v = (int)fadd(cv[k], 0.5f);
	if (u < 0 || v < 0 || u >= w || v >= h) return;
	atomicAdd(hits + u + w * v, 1);
}

__global__ void cd_splat_kernel(int n, const float* __restrict__ cu, const float* __restrict__ cv, const float* __restrict__ cid,
                                const float* __restrict__ hdi, int w, int h, const int* __restrict__ hits, float* idepth, float* wsum)
{
	const int k = blockIdx.x * blockDim.x + threadIdx.x;
	if (k >= n) return;
	int pix; float idw, weight;
	if (!cd_point(k, cu, cv, cid, hdi, w, h, pix, idw, weight) || hits[pix] > 2) return;
	atomicAdd(idepth + pix, idw);
	atomicAdd(wsum + pix, weight);
}

// the (rare) pixels with three or more points: chunks of 1024 points, flagged ones compacted in point order, then added
// per pixel in that order, one thread per pixel
__global__ void __launch_bounds__(1024) cd_splat_serial_kernel(int n, const float* __restrict__ cu, const float* __restrict__ cv,
                                                               const float* __restrict__ cid, const float* __restrict__ hdi, int w, int h,
                                                               const int* __restrict__ hits, float* idepth, float* wsum)
{
	__shared__ int s_pix[1024];
	__shared__ float s_idw[1024], s_w[1024];
	__shared__ int s_warp[32];
	__shared__ unsigned long long s_key[1024];
	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	for (int base = 0; base < n; base += 1024)
	{
		const int k = base + tid;
		int pix = 0; float idw = 0.f, weight = 0.f;
		const bool f = k < n && cd_point(k, cu, cv, cid, hdi, w, h, pix, idw, weight) && hits[pix] > 2;
		const unsigned m = __ballot_sync(0xffffffffu, f);
		if (lane == 0) s_warp[warp] = __popc(m);
		__syncthreads();
		int pre = 0, tot = 0;
		for (int q = 0; q < 32; q++) { const int c = s_warp[q]; if (q < warp) pre += c; tot += c; }
		if (f) { const int pos = pre + __popc(m & ((1u << lane) - 1u)); s_pix[pos] = pix; s_idw[pos] = idw; s_w[pos] = weight; }
		__syncthreads();
		// different pixels are independent; the entries of one pixel must be added in point order.  Sort (pixel, position
		// in the compacted list) -- a bitonic network over the next power of two, in shared memory -- so that the entries of
		// a pixel sit next to each other in point order; the thread holding the first entry of a run adds the whole run.
		int P2 = 1;
		while (P2 < tot) P2 <<= 1;
		if (tid < P2) s_key[tid] = tid < tot ? (((unsigned long long)s_pix[tid] << 10) | (unsigned long long)tid) : ~0ull;
		__syncthreads();
		for (int kk = 2; kk <= P2; kk <<= 1)
			for (int j = kk >> 1; j > 0; j >>= 1)
			{
				const int partner = tid ^ j;
				if (tid < P2 && partner > tid)
				{
					const unsigned long long x = s_key[tid], y = s_key[partner];
					const bool up = (tid & kk) == 0;
					if ((x > y) == up) { s_key[tid] = y; s_key[partner] = x; }
				}
				__syncthreads();
			}
		if (tid < tot)
		{
			const unsigned long long me = s_key[tid];
			const int p = (int)(me >> 10);
			if (tid == 0 || (int)(s_key[tid - 1] >> 10) != p)
			{
				float a = idepth[p], b = wsum[p];
				for (int q = tid; q < tot && (int)(s_key[q] >> 10) == p; q++)
				{
					const int pos = (int)(s_key[q] & 1023ull);
					a = fadd(a, s_idw[pos]); b = fadd(b, s_w[pos]);
				}
				idepth[p] = a; wsum[p] = b;
			}
		}
		__syncthreads();
	}
}

__global__ void cd_down_kernel(int wl, int hl, int wlm1, const float* __restrict__ id_lm, const float* __restrict__ ws_lm,
                               float* id_l, float* ws_l)
{
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= wl * hl) return;
	const int y = i / wl, x = i - y * wl;
	const int b = 2 * x + 2 * y * wlm1;
	id_l[i] = fadd(fadd(fadd(id_lm[b], id_lm[b + 1]), id_lm[b + wlm1]), id_lm[b + wlm1 + 1]);
	ws_l[i] = fadd(fadd(fadd(ws_lm[b], ws_lm[b + 1]), ws_lm[b + wlm1]), ws_lm[b + wlm1 + 1]);
}

__global__ void cd_dilate_kernel(int wl, int hl, int diagonal, float* idepth, const float* __restrict__ wsum, float* wsum2)
{
	const int i = blockIdx.x * blockDim.x + threadIdx.x, n = wl * hl;
	if (i >= n) return;
	const float wi = wsum[i];
	float wo = wi;
	if (i >= wl && i < n - wl && wi <= 0)
	{
		const int o0 = diagonal ? 1 + wl : 1, o1 = diagonal ? -1 - wl : -1, o2 = diagonal ? wl - 1 : wl, o3 = diagonal ? -wl + 1 : -wl;
		float sum = 0.f, num = 0.f, numn = 0.f;
		const int offs[4] = { o0, o1, o2, o3 };
#pragma unroll
		for (int q = 0; q < 4; q++)
		{
			const int j = i + offs[q];
			if (j < 0 || j >= n) continue;                 // the reference reads one element outside at the first / last row
			const float wj = wsum[j];
			if (wj > 0) { sum = fadd(sum, idepth[j]); num = fadd(num, wj); numn += 1.f; }
		}
		if (numn > 0) { idepth[i] = fdiv(sum, numn); wo = fdiv(num, numn); }
	}
	wsum2[i] = wo;
}

// validity of raster pixel i of level l and its record (:261-283)
__device__ __forceinline__ bool cd_record(const CDLevels& L, int l, int i, float4& rec)
{
	const int wl = L.w[l], hl = L.h[l];
	if (i >= wl * hl) return false;
	const int y = i / wl, x = i - y * wl;
	if (x < 2 || x >= wl - 2 || y < 2 || y >= hl - 2) return false;
	const float ws = L.wsum2[l][i];
	if (!(ws > 0)) return false;
	const float id = fdiv(L.idepth[l][i], ws);
	const float color = L.ref[l][i].x;
	if (!isfinite(color) || !(id > 0)) return false;
	rec = make_float4((float)x, (float)y, id, color);
	return true;
}

__device__ __forceinline__ int cd_level_of_block(const CDLevels& L, int b)
{
	int l = 0;
	while (l + 1 < L.levels && b >= L.block_start[l + 1]) l++;
	return l;
}

__global__ void __launch_bounds__(256) cd_count_kernel(CDLevels L, int* block_count)
{
	__shared__ int s[8];
	const int l = cd_level_of_block(L, blockIdx.x);
	const int base = (blockIdx.x - L.block_start[l]) * 1024 + threadIdx.x * 4;
	int c = 0;
	float4 r;
#pragma unroll
	for (int q = 0; q < 4; q++) c += cd_record(L, l, base + q, r) ? 1 : 0;
#pragma unroll
	for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
	if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = c;
	__syncthreads();
	if (threadIdx.x == 0) { int t = 0; for (int k = 0; k < 8; k++) t += s[k]; block_count[blockIdx.x] = t; }
}

// one block per level: exclusive scan of that level's block counts (sequential chunks of 1024 blocks)
__global__ void __launch_bounds__(1024) cd_scan_kernel(CDLevels L, const int* __restrict__ block_count, int* block_off, int* pc_n)
{
	__shared__ int s_warp[32];
	__shared__ int s_carry;
	const int l = blockIdx.x, b0 = L.block_start[l], nb = L.block_start[l + 1] - b0;
	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	if (tid == 0) s_carry = 0;
	__syncthreads();
	for (int base = 0; base < nb; base += 1024)
	{
		const int k = base + tid;
		const int v = k < nb ? block_count[b0 + k] : 0;
		int x = v;
#pragma unroll
		for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
		if (lane == 31) s_warp[warp] = x;
		__syncthreads();
		if (warp == 0)
		{
			int w = s_warp[lane];
#pragma unroll
			for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, w, o); if (lane >= o) w += y; }
			s_warp[lane] = w;
		}
		__syncthreads();
		const int incl = x + (warp ? s_warp[warp - 1] : 0);
		const int carry = s_carry;
		if (k < nb) block_off[b0 + k] = carry + incl - v;
		__syncthreads();
		if (tid == 1023) s_carry = carry + incl;
		__syncthreads();
	}
	if (tid == 0) pc_n[l] = s_carry;
}

__global__ void __launch_bounds__(256) cd_write_kernel(CDLevels L, const int* __restrict__ block_off)
{
	__shared__ int s[8];
	const int l = cd_level_of_block(L, blockIdx.x);
	const int base = (blockIdx.x - L.block_start[l]) * 1024 + threadIdx.x * 4;
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	float4 r[4];
	bool ok[4];
	int c = 0;
#pragma unroll
	for (int q = 0; q < 4; q++) { ok[q] = cd_record(L, l, base + q, r[q]); c += ok[q] ? 1 : 0; }
	int x = c;
#pragma unroll
	for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
	if (lane == 31) s[warp] = x;
	__syncthreads();
	int pre = 0;
	for (int k = 0; k < warp; k++) pre += s[k];
	int pos = block_off[blockIdx.x] + pre + x - c;
	float4* out = L.pc[l];
#pragma unroll
	for (int q = 0; q < 4; q++) if (ok[q]) out[pos++] = r[q];
}

// =========================================================================================================
static int cd_workspace(dso_b200_ctx* ctx, CoarseDepthWS** out, size_t npts)
{
	CoarseDepthWS* ws = (CoarseDepthWS*)ctx->coarse_depth;
	if (!ws)
	{
		ws = new CoarseDepthWS();
		memset(ws, 0, sizeof(*ws));
		size_t px = 0; int blocks = 0;
		for (int l = 0; l < MAXL; l++) { const size_t n = (size_t)(ctx->max_w >> l) * (ctx->max_h >> l); px += align_up(n * sizeof(float), 256) / sizeof(float); blocks += (int)((n + 1023) / 1024); }
		ws->cap_px = px; ws->cap_blocks = blocks;
		CK(cudaMalloc(&ws->base, 3 * px * sizeof(float)));
		CK(cudaMalloc(&ws->d_block_count, blocks * sizeof(int)));
		CK(cudaMalloc(&ws->d_block_off, blocks * sizeof(int)));
		CK(cudaMalloc(&ws->d_pc_n, MAXL * sizeof(int)));
		CK(cudaHostAlloc((void**)&ws->h_pc_n, MAXL * sizeof(int), cudaHostAllocDefault));
		ctx->coarse_depth = ws;
	}
	if (npts > ws->cap_pts)
	{
		if (ws->cap_pts) cudaFree(ws->d_pts);
		CK(cudaMalloc(&ws->d_pts, 4 * npts * sizeof(float)));
		ws->cap_pts = npts;
	}
	*out = ws;
	return 0;
}

void coarse_depth_destroy(dso_b200_ctx* ctx)
{
	CoarseDepthWS* ws = (CoarseDepthWS*)ctx->coarse_depth;
	if (!ws) return;
	cudaFree(ws->base); cudaFree(ws->d_block_count); cudaFree(ws->d_block_off); cudaFree(ws->d_pc_n); cudaFreeHost(ws->h_pc_n);
	if (ws->cap_pts) cudaFree(ws->d_pts);
	delete ws;
	ctx->coarse_depth = nullptr;
}

extern "C" int dso_b200_tracker_set_ref_points(dso_b200_ctx* ctx, int trk, int ref_frame_slot, int levels,
                                               float fx, float fy, float cx, float cy, int n_points,
                                               const float* cpt_u, const float* cpt_v, const float* cpt_idepth, const float* HdiF,
                                               double ref_aff_a, double ref_aff_b, float ref_ab_exposure, int* pc_n_out)
{
	ARG_CHECK(ctx && (trk == 0 || trk == 1), "bad tracker slot");
	ARG_CHECK(ref_frame_slot >= 0 && ref_frame_slot < ctx->n_slots && ctx->slots[ref_frame_slot].valid, "reference frame slot empty");
	ARG_CHECK(n_points >= 0 && (n_points == 0 || (cpt_u && cpt_v && cpt_idepth && HdiF)), "null point arrays");
	CK(cudaSetDevice(ctx->device));
	FrameSlot& F = ctx->slots[ref_frame_slot];
	ARG_CHECK(levels >= 1 && levels <= MAXL && levels <= F.levels, "levels exceed the reference frame's pyramid");
	TrackerSlot& T = ctx->trk[trk];
	int rc = frame_slot_wait(ctx, ref_frame_slot, T.stream, T.consumer);
	if (rc) return rc;
	const int w = F.w, h = F.h;
	tracker_make_K(T, w, h, levels, fx, fy, cx, cy);                                   // makeK, CoarseTracker.cpp:102-131
	CoarseDepthWS* ws;
	if ((rc = cd_workspace(ctx, &ws, (size_t)(n_points > 0 ? n_points : 1)))) return rc;
	cudaStream_t s = T.stream;      // the tracker slot being prepared (coarseTracker_forNewKF), not the one tracking
	CDLevels L;
	memset(&L, 0, sizeof(L));
	L.levels = levels;
	float* p = ws->base;
	int blocks = 0;
	for (int l = 0; l < levels; l++)
	{
		const size_t n = (size_t)(w >> l) * (h >> l), st = align_up(n * sizeof(float), 256) / sizeof(float);
		L.w[l] = w >> l; L.h[l] = h >> l;
		L.idepth[l] = p; p += st; L.wsum[l] = p; p += st; L.wsum2[l] = p; p += st;
		L.ref[l] = F.pyr[l]; L.pc[l] = T.pc[l];
		L.block_start[l] = blocks; blocks += (int)((n + 1023) / 1024);
	}
	for (int l = levels; l <= MAXL; l++) L.block_start[l] = blocks;
	const size_t n0 = (size_t)w * h;
	CK(cudaMemsetAsync(L.idepth[0], 0, n0 * sizeof(float), s));                        // :137-138
	CK(cudaMemsetAsync(L.wsum[0], 0, n0 * sizeof(float), s));
	if (n_points > 0)
	{
		const size_t nb = (size_t)n_points * sizeof(float);
		CK(cudaMemcpyAsync(ws->d_pts, cpt_u, nb, cudaMemcpyHostToDevice, s));
		CK(cudaMemcpyAsync(ws->d_pts + n_points, cpt_v, nb, cudaMemcpyHostToDevice, s));
		CK(cudaMemcpyAsync(ws->d_pts + 2 * (size_t)n_points, cpt_idepth, nb, cudaMemcpyHostToDevice, s));
		CK(cudaMemcpyAsync(ws->d_pts + 3 * (size_t)n_points, HdiF, nb, cudaMemcpyHostToDevice, s));
		const float* du = ws->d_pts; const float* dv = ws->d_pts + n_points;
		const float* di = ws->d_pts + 2 * (size_t)n_points; const float* dh = ws->d_pts + 3 * (size_t)n_points;
		int* hits = reinterpret_cast<int*>(L.wsum2[0]);            // free until the dilation pass writes it
		CK(cudaMemsetAsync(hits, 0, n0 * sizeof(int), s));
		cd_hits_kernel<<<(n_points + 255) / 256, 256, 0, s>>>(n_points, du, dv, w, h, hits);
		cd_splat_kernel<<<(n_points + 255) / 256, 256, 0, s>>>(n_points, du, dv, di, dh, w, h, hits, L.idepth[0], L.wsum[0]);
		cd_splat_serial_kernel<<<1, 1024, 0, s>>>(n_points, du, dv, di, dh, w, h, hits, L.idepth[0], L.wsum[0]);
		LAUNCHED(ctx, 3);
	}
	for (int l = 1; l < levels; l++)
	{
		const int n = L.w[l] * L.h[l];
		cd_down_kernel<<<(n + 255) / 256, 256, 0, s>>>(L.w[l], L.h[l], L.w[l - 1], L.idepth[l - 1], L.wsum[l - 1], L.idepth[l], L.wsum[l]);
		LAUNCHED(ctx, 1);
	}
	for (int l = 0; l < levels; l++)
	{
		const int n = L.w[l] * L.h[l];
		cd_dilate_kernel<<<(n + 255) / 256, 256, 0, s>>>(L.w[l], L.h[l], l < 2 ? 1 : 0, L.idepth[l], L.wsum[l], L.wsum2[l]);
		LAUNCHED(ctx, 1);
	}
	cd_count_kernel<<<blocks, 256, 0, s>>>(L, ws->d_block_count);
	cd_scan_kernel<<<levels, 1024, 0, s>>>(L, ws->d_block_count, ws->d_block_off, ws->d_pc_n);
	cd_write_kernel<<<blocks, 256, 0, s>>>(L, ws->d_block_off);
	LAUNCHED(ctx, 3);
	CK(cudaGetLastError());
	CK(cudaMemcpyAsync(ws->h_pc_n, ws->d_pc_n, levels * sizeof(int), cudaMemcpyDeviceToHost, s));
	CK(cudaStreamSynchronize(s));
	for (int l = 0; l < levels; l++)
	{
		T.pc_n[l] = ws->h_pc_n[l];
		if (pc_n_out) pc_n_out[l] = ws->h_pc_n[l];
	}
	T.ref_a = ref_aff_a; T.ref_b = ref_aff_b; T.ref_exposure = ref_ab_exposure;
	T.valid = true; T.last_lvl = -1; T.last_warped_n = 0;
	return 0;
}

// debug / test access to a tracker slot's cloud: (u, v, idepth, color) records of one level
extern "C" int dso_b200_tracker_get_cloud(dso_b200_ctx* ctx, int trk, int lvl, int max_points, float* out4, int* n_out)
{
	ARG_CHECK(ctx && (trk == 0 || trk == 1) && ctx->trk[trk].valid && n_out, "bad arguments");
	TrackerSlot& T = ctx->trk[trk];
	ARG_CHECK(lvl >= 0 && lvl < T.levels, "bad level");
	CK(cudaSetDevice(ctx->device));
	*n_out = T.pc_n[lvl];
	if (out4)
	{
		const int n = T.pc_n[lvl] < max_points ? T.pc_n[lvl] : max_points;
		CK(cudaMemcpyAsync(out4, T.pc[lvl], (size_t)n * sizeof(float4), cudaMemcpyDeviceToHost, T.stream));
		CK(cudaStreamSynchronize(T.stream));
	}
	return 0;
}
