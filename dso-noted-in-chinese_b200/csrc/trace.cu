// Next row N3 (second half): ImmaturePoint::traceOn (FullSystem/ImmaturePoint.cpp:77-438) for every immature point of
// every key frame on a new frame = FullSystem::traceNewCoarse (FullSystem.cpp:486-530).
//
//   trace_on_kernel   eight lanes per point (one per pattern pixel, four points per warp).  The per-point control flow
//                     (projection of the inverse-depth interval, the status exits, step direction, Gauss-Newton along the
//                     line, the new interval) is computed redundantly by the eight lanes; the eight pattern samples of a
//                     step are taken in parallel and summed in pattern order through shuffles, so energies, the chosen
//                     minimum, the quality ratio and the resulting interval are bit-identical to the reference's loops.
// Work per point: up to 99 search steps x 8 bilinear samples (4 texels of 16 B each) + 3 GN iterations; the texels of
// successive steps are neighbours, so this is L1/L2-latency-bound gather work, ~10-15 k points per frame.
#include "common.cuh"
#include <vector>

#define TR_PTS_PER_BLOCK 32
#define IPS_GOOD 0
#define IPS_OOB 1
#define IPS_OUTLIER 2
#define IPS_SKIPPED 3
#define IPS_BADCONDITION 4

struct TraceHosts
{
	float KRKi[MAXF][9], Kt[MAXF][3], aff[MAXF][2];
};

struct TraceWS
{
	int n, cap;
	float* d_u; float* d_v; float* d_color; float* d_weights; float* d_gradH; float* d_energyTH; int* d_host;
	float* d_idmin; float* d_idmax; float* d_quality; int* d_status; float* d_uv; float* d_interval;
};

__constant__ int c_trace_pattern[8][2] = { {0,-2}, {-1,-1}, {1,-1}, {-2,0}, {0,0}, {2,0}, {-1,1}, {0,2} };   // settings.cpp:274

// sum of the 8 lanes of a point's group in pattern order, starting from `init` (the reference's scalar "+=" loop)
__device__ __forceinline__ float tr_seq8(float init, float v, int base)
{
	float s = init;
#pragma unroll
	for (int k = 0; k < 8; k++) s = fadd(s, __shfl_sync(0xffffffffu, v, base + k));
	return s;
}

// getInterpolatedElement31 / 33 (util/globalFuncs.h:126-140, 75-89) on the float4 (I, dx, dy, |grad|^2) level-0 image
__device__ __forceinline__ void tr_interp(const float4* __restrict__ img, float x, float y, int width, float& h0, float& h1, float& h2, bool grad)
{
	const int ix = (int)x, iy = (int)y;
	const float dx = fsub(x, (float)ix), dy = fsub(y, (float)iy);
	const float dxdy = fmul(dx, dy);
	const float w11 = dxdy, w01 = fsub(dy, dxdy), w10 = fsub(dx, dxdy), w00 = fadd(fsub(fsub(1.f, dx), dy), dxdy);
	const float4* bp = img + ix + iy * width;
	const float4 p00 = __ldg(bp), p10 = __ldg(bp + 1), p01 = __ldg(bp + width), p11 = __ldg(bp + 1 + width);
	h0 = fadd(fadd(fadd(fmul(w11, p11.x), fmul(w01, p01.x)), fmul(w10, p10.x)), fmul(w00, p00.x));
	if (grad)
	{
		h1 = fadd(fadd(fadd(fmul(w11, p11.y), fmul(w01, p01.y)), fmul(w10, p10.y)), fmul(w00, p00.y));
		h2 = fadd(fadd(fadd(fmul(w11, p11.z), fmul(w01, p01.z)), fmul(w10, p10.z)), fmul(w00, p00.z));
	}
}

struct TracePoint
{
	float u, v, idmin, idmax, quality, energyTH, G[4];
	int status; float uv0, uv1, interval;
};

__device__ __forceinline__ int tr_fail(TracePoint& p, int st) { p.uv0 = -1.f; p.uv1 = -1.f; p.interval = 0.f; return p.status = st; }

// one point, executed by the 8 lanes of its group (idx = pattern pixel of this lane); control flow is uniform in the group
__device__ int trace_on_group(const dso_b200_settings& ST, TracePoint& p, float color, float weight, const float4* __restrict__ img, int wG0, int hG0,
                              const float* KRKi, const float* Kt, const float* aff, float* errors, int idx, int base, unsigned gmask)
{
	if (p.status == IPS_OOB) return p.status;                                          // :79
	const float maxPixSearch = fmul((float)(wG0 + hG0), ST.maxPixSearch);               // :85 setting_maxPixSearch
	float pr[3];
#pragma unroll
	for (int r = 0; r < 3; r++) pr[r] = dot3e(KRKi[3 * r], p.u, KRKi[3 * r + 1], p.v, KRKi[3 * r + 2], 1.f);   // :96
	float ptpMin[3];
#pragma unroll
	for (int r = 0; r < 3; r++) ptpMin[r] = fadd(pr[r], fmul(Kt[r], p.idmin));         // :98
	const float uMin = fdiv(ptpMin[0], ptpMin[2]), vMin = fdiv(ptpMin[1], ptpMin[2]);
	const float wlim = (float)(wG0 - 5), hlim = (float)(hG0 - 5);
	if (!(uMin > 4.f && vMin > 4.f && uMin < wlim && vMin < hlim)) return tr_fail(p, IPS_OOB);   // :103-110
	float dist, uMax, vMax;
	const bool finiteMax = isfinite(p.idmax);
	if (finiteMax)
	{
		float q[3];
#pragma unroll
		for (int r = 0; r < 3; r++) q[r] = fadd(pr[r], fmul(Kt[r], p.idmax));            // :119
		uMax = fdiv(q[0], q[2]); vMax = fdiv(q[1], q[2]);
		if (!(uMax > 4.f && vMax > 4.f && uMax < wlim && vMax < hlim)) return tr_fail(p, IPS_OOB);   // :124-130
		const float du = fsub(uMin, uMax), dv = fsub(vMin, vMax);
		dist = __fsqrt_rn(fadd(fmul(du, du), fmul(dv, dv)));                             // :135-136
		if (dist < ST.trace_slackInterval)                                               // :138-145 setting_trace_slackInterval
		{
			p.uv0 = fmul(fadd(uMax, uMin), 0.5f); p.uv1 = fmul(fadd(vMax, vMin), 0.5f);
			p.interval = dist;
			return p.status = IPS_SKIPPED;
		}
	}
	else
	{
		dist = maxPixSearch;                                                             // :150
		float q[3];
#pragma unroll
		for (int r = 0; r < 3; r++) q[r] = fadd(pr[r], fmul(Kt[r], 0.01f));              // :153
		uMax = fdiv(q[0], q[2]); vMax = fdiv(q[1], q[2]);
		const float ddx = fsub(uMax, uMin), ddy = fsub(vMax, vMin);
		const float d = fdiv(1.0f, __fsqrt_rn(fadd(fmul(ddx, ddx), fmul(ddy, ddy))));
		uMax = fadd(uMin, fmul(fmul(dist, ddx), d));                                     // :162
		vMax = fadd(vMin, fmul(fmul(dist, ddy), d));
		if (!(uMax > 4.f && vMax > 4.f && uMax < wlim && vMax < hlim)) return tr_fail(p, IPS_OOB);   // :166-172
	}
	if (!(p.idmin < 0.f || (ptpMin[2] > 0.75f && ptpMin[2] < 1.5f))) return tr_fail(p, IPS_OOB);     // :179-185
	float dx = fmul(ST.trace_stepsize, fsub(uMax, uMin));                               // :196 setting_trace_stepsize
	float dy = fmul(ST.trace_stepsize, fsub(vMax, vMin));
	float a, b;
	{
		const float t0 = fadd(fmul(dx, p.G[0]), fmul(dy, p.G[2])), t1 = fadd(fmul(dx, p.G[1]), fmul(dy, p.G[3]));   // :199
		a = fadd(fmul(t0, dx), fmul(t1, dy));
		const float s0 = fadd(fmul(dy, p.G[0]), fmul(-dx, p.G[2])), s1 = fadd(fmul(dy, p.G[1]), fmul(-dx, p.G[3]));   // :200
		b = fadd(fmul(s0, dy), fmul(s1, -dx));
	}
	float errorInPixel = fadd(0.2f, fdiv(fmul(0.2f, fadd(a, b)), a));                   // :201
	if (fmul(errorInPixel, ST.trace_minImprovementFactor) > dist && finiteMax)          // :203-209 setting_trace_minImprovementFactor
	{
		p.uv0 = fmul(fadd(uMax, uMin), 0.5f); p.uv1 = fmul(fadd(vMax, vMin), 0.5f);
		p.interval = dist;
		return p.status = IPS_BADCONDITION;
	}
	if (errorInPixel > 10.f) errorInPixel = 10.f;
	dx = fdiv(dx, dist);                                                                // :215
	dy = fdiv(dy, dist);
	if (dist > maxPixSearch)                                                            // :227-232
	{
		uMax = fadd(uMin, fmul(maxPixSearch, dx));
		vMax = fadd(vMin, fmul(maxPixSearch, dy));
		dist = maxPixSearch;
	}
	int numSteps = (int)fadd(1.9999f, fdiv(dist, ST.trace_stepsize));                   // :234
	const float randShift = fsub(fmul(uMin, 1000.f), floorf(fmul(uMin, 1000.f)));       // :238
	float ptx = fsub(uMin, fmul(randShift, dx));
	float pty = fsub(vMin, fmul(randShift, dy));
	const float px = (float)c_trace_pattern[idx][0], py = (float)c_trace_pattern[idx][1];
	const float rot0 = fadd(fmul(KRKi[0], px), fmul(KRKi[1], py));                      // :245 Rplane * pattern
	const float rot1 = fadd(fmul(KRKi[3], px), fmul(KRKi[4], py));
	if (!isfinite(dx) || !isfinite(dy)) return tr_fail(p, IPS_OOB);                     // :250-256
	float bestU = 0.f, bestV = 0.f, bestEnergy = 1e10f;
	int bestIdx = -1;
	if (numSteps >= 100) numSteps = 99;                                                 // :264
	for (int i = 0; i < numSteps; i++)                                                  // :267-300 discrete search along the line
	{
		float hit, g1, g2;
		tr_interp(img, fadd(ptx, rot0), fadd(pty, rot1), wG0, hit, g1, g2, false);
		float term;
		if (!isfinite(hit)) term = 1e5f;
		else
		{
			const float residual = fsub(hit, fadd(fmul(aff[0], color), aff[1]));
			const float ar = fabsf(residual);
			const float hw = ar < ST.huberTH ? 1.f : fdiv(ST.huberTH, ar);                 // setting_huberTH
			term = fmul(fmul(fmul(hw, residual), residual), fsub(2.f, hw));
		}
		// energy += term in pattern order; the reference adds the double literal 1e5 for a non-finite sample (:277)
		float energy = 0.f;
#pragma unroll
		for (int k = 0; k < 8; k++)
		{
			const float tk = __shfl_sync(gmask, term, base + k);
			const bool bad = __shfl_sync(gmask, !isfinite(hit) ? 1 : 0, base + k) != 0;
			energy = bad ? (float)((double)energy + 1e5) : fadd(energy, tk);
		}
		if (idx == 0) errors[i] = energy;
		if (energy < bestEnergy) { bestU = ptx; bestV = pty; bestEnergy = energy; bestIdx = i; }
		ptx = fadd(ptx, dx);
		pty = fadd(pty, dy);
	}
	__syncwarp(gmask);
	float secondBest = 1e10f;                                                           // :303-309 (a minimum: order-free)
	for (int i = idx; i < numSteps; i += 8)
		if ((i < bestIdx - ST.minTraceTestRadius || i > bestIdx + ST.minTraceTestRadius) && errors[i] < secondBest) secondBest = errors[i];   // setting_minTraceTestRadius
#pragma unroll
	for (int o = 4; o > 0; o >>= 1)
	{
		const float other = __shfl_xor_sync(gmask, secondBest, o);
		if (other < secondBest) secondBest = other;
	}
	const float newQuality = fdiv(secondBest, bestEnergy);
	if (newQuality < p.quality || numSteps > 10) p.quality = newQuality;                // :311
	float uBak = bestU, vBak = bestV, stepBack = 0.f;                                   // :315
	if (ST.trace_GNIterations > 0) bestEnergy = 1e5f;                                   // :316
	for (int it = 0; it < ST.trace_GNIterations; it++)                                  // :318-375 Gauss-Newton along the line
	{
		float h0, h1, h2;
		tr_interp(img, fadd(bestU, rot0), fadd(bestV, rot1), wG0, h0, h1, h2, true);
		const bool bad = !isfinite(h0);
		float tH = 0.f, tb = 0.f, tE = 0.f;
		if (!bad)
		{
			const float residual = fsub(h0, fadd(fmul(aff[0], color), aff[1]));
			const float dRes = fadd(fmul(dx, h1), fmul(dy, h2));
			const float ar = fabsf(residual);
			const float hw = ar < ST.huberTH ? 1.f : fdiv(ST.huberTH, ar);
			tH = fmul(fmul(hw, dRes), dRes);
			tb = fmul(fmul(hw, residual), dRes);
			tE = fmul(fmul(fmul(fmul(fmul(weight, weight), hw), residual), residual), fsub(2.f, hw));
		}
		float H = 1.f, bb = 0.f, energy = 0.f;
#pragma unroll
		for (int k = 0; k < 8; k++)
		{
			const bool bk = __shfl_sync(gmask, bad ? 1 : 0, base + k) != 0;
			const float hk = __shfl_sync(gmask, tH, base + k), bk2 = __shfl_sync(gmask, tb, base + k), ek = __shfl_sync(gmask, tE, base + k);
			if (bk) energy = (float)((double)energy + 1e5);                              // :328 (H, b untouched)
			else { H = fadd(H, hk); bb = fadd(bb, bk2); energy = fadd(energy, ek); }
		}
		if (energy > bestEnergy)                                                        // :341-351
		{
			stepBack = fmul(stepBack, 0.5f);
			bestU = fadd(uBak, fmul(stepBack, dx));
			bestV = fadd(vBak, fmul(stepBack, dy));
		}
		else                                                                            // :352-372
		{
			float step = fdiv(fmul(-1.f, bb), H);
			if (step < -0.5f) step = -0.5f;
			else if (step > 0.5f) step = 0.5f;
			if (!isfinite(step)) step = 0.f;
			uBak = bestU; vBak = bestV;
			stepBack = step;
			bestU = fadd(bestU, fmul(step, dx));
			bestV = fadd(bestV, fmul(step, dy));
			bestEnergy = energy;
		}
		if (fabsf(stepBack) < ST.trace_GNThreshold) break;                              // :374 setting_trace_GNThreshold
	}
	if (!(bestEnergy < fmul(p.energyTH, ST.trace_extraSlackOnTH)))                      // :380-391 setting_trace_extraSlackOnTH
	{
		p.interval = 0.f; p.uv0 = -1.f; p.uv1 = -1.f;
		if (p.status == IPS_OUTLIER) return p.status = IPS_OOB;
		return p.status = IPS_OUTLIER;
	}
	float imin, imax;
	if (fmul(dx, dx) > fmul(dy, dy))                                                    // :395-404
	{
		const float lo = fsub(bestU, fmul(errorInPixel, dx)), hi = fadd(bestU, fmul(errorInPixel, dx));
		imin = fdiv(fsub(fmul(pr[2], lo), pr[0]), fsub(Kt[0], fmul(Kt[2], lo)));
		imax = fdiv(fsub(fmul(pr[2], hi), pr[0]), fsub(Kt[0], fmul(Kt[2], hi)));
	}
	else
	{
		const float lo = fsub(bestV, fmul(errorInPixel, dy)), hi = fadd(bestV, fmul(errorInPixel, dy));
		imin = fdiv(fsub(fmul(pr[2], lo), pr[1]), fsub(Kt[1], fmul(Kt[2], lo)));
		imax = fdiv(fsub(fmul(pr[2], hi), pr[1]), fsub(Kt[1], fmul(Kt[2], hi)));
	}
	if (imin > imax) { const float t = imin; imin = imax; imax = t; }                   // :405
	p.idmin = imin; p.idmax = imax;
	if (!isfinite(imin) || !isfinite(imax) || (imax < 0.f)) return tr_fail(p, IPS_OUTLIER);   // :408-415
	p.interval = fmul(2.f, errorInPixel);                                               // :417
	p.uv0 = bestU; p.uv1 = bestV;
	return p.status = IPS_GOOD;
}

__global__ void __launch_bounds__(256) trace_on_kernel(dso_b200_settings ST, int n, const float4* __restrict__ img, int w, int h, TraceHosts H,
                                                       const float* __restrict__ pu, const float* __restrict__ pv,
                                                       const float* __restrict__ color8, const float* __restrict__ weights8,
                                                       const float* __restrict__ gradH4, const float* __restrict__ energyTH,
                                                       const int* __restrict__ host, float* idmin, float* idmax, float* quality,
                                                       int* status, float* uv2, float* interval)
{
	__shared__ float s_err[TR_PTS_PER_BLOCK][100];
	const int tid = threadIdx.x, lane = tid & 31, g = tid >> 3, idx = tid & 7, base = lane & ~7;
	const int k = blockIdx.x * TR_PTS_PER_BLOCK + g;
	if (k >= n) return;                                        // whole groups leave together
	const unsigned gmask = 0xffu << base;
	TracePoint p;
	p.u = pu[k]; p.v = pv[k]; p.idmin = idmin[k]; p.idmax = idmax[k]; p.quality = quality[k]; p.energyTH = energyTH[k];
#pragma unroll
	for (int q = 0; q < 4; q++) p.G[q] = gradH4[4 * (size_t)k + q];
	p.status = status[k]; p.uv0 = uv2[2 * (size_t)k]; p.uv1 = uv2[2 * (size_t)k + 1]; p.interval = interval[k];
	const int hst = host[k];
	trace_on_group(ST, p, color8[8 * (size_t)k + idx], weights8[8 * (size_t)k + idx], img, w, h, H.KRKi[hst], H.Kt[hst], H.aff[hst],
	               s_err[g], idx, base, gmask);
	if (idx == 0)
	{
		idmin[k] = p.idmin; idmax[k] = p.idmax; quality[k] = p.quality; status[k] = p.status;
		uv2[2 * (size_t)k] = p.uv0; uv2[2 * (size_t)k + 1] = p.uv1; interval[k] = p.interval;
	}
}

// =========================================================================================================
static int trace_ws(dso_b200_ctx* ctx, TraceWS** out, int n)
{
	TraceWS* T = (TraceWS*)ctx->trace;
	if (!T) { T = new TraceWS(); memset(T, 0, sizeof(*T)); ctx->trace = T; }
	if (n > T->cap)
	{
		if (T->cap)
		{
			cudaFree(T->d_u); cudaFree(T->d_v); cudaFree(T->d_color); cudaFree(T->d_weights); cudaFree(T->d_gradH); cudaFree(T->d_energyTH);
			cudaFree(T->d_host); cudaFree(T->d_idmin); cudaFree(T->d_idmax); cudaFree(T->d_quality); cudaFree(T->d_status); cudaFree(T->d_uv); cudaFree(T->d_interval);
		}
		const size_t m = (size_t)n;
		CK(cudaMalloc(&T->d_u, m * 4)); CK(cudaMalloc(&T->d_v, m * 4)); CK(cudaMalloc(&T->d_color, m * 32)); CK(cudaMalloc(&T->d_weights, m * 32));
		CK(cudaMalloc(&T->d_gradH, m * 16)); CK(cudaMalloc(&T->d_energyTH, m * 4)); CK(cudaMalloc(&T->d_host, m * 4));
		CK(cudaMalloc(&T->d_idmin, m * 4)); CK(cudaMalloc(&T->d_idmax, m * 4)); CK(cudaMalloc(&T->d_quality, m * 4)); CK(cudaMalloc(&T->d_status, m * 4));
		CK(cudaMalloc(&T->d_uv, m * 8)); CK(cudaMalloc(&T->d_interval, m * 4));
		T->cap = n;
	}
	*out = T;
	return 0;
}

void trace_destroy(dso_b200_ctx* ctx)
{
	TraceWS* T = (TraceWS*)ctx->trace;
	if (!T) return;
	if (T->cap)
	{
		cudaFree(T->d_u); cudaFree(T->d_v); cudaFree(T->d_color); cudaFree(T->d_weights); cudaFree(T->d_gradH); cudaFree(T->d_energyTH);
		cudaFree(T->d_host); cudaFree(T->d_idmin); cudaFree(T->d_idmax); cudaFree(T->d_quality); cudaFree(T->d_status); cudaFree(T->d_uv); cudaFree(T->d_interval);
	}
	delete T;
	ctx->trace = nullptr;
}

extern "C" int dso_b200_trace_set_points(dso_b200_ctx* ctx, int n, const float* u, const float* v, const float* color8, const float* weights8,
                                         const float* gradH4, const float* energyTH, const int* host_idx,
                                         const float* idepth_min, const float* idepth_max, const float* quality, const int* status,
                                         const float* lastTraceUV2, const float* lastTracePixelInterval)
{
	ARG_CHECK(ctx && n >= 1 && u && v && color8 && weights8 && gradH4 && energyTH && host_idx && idepth_min && idepth_max && quality && status,
	          "null argument");
	for (int i = 0; i < n; i++) ARG_CHECK(host_idx[i] >= 0 && host_idx[i] < MAXF, "host index out of range");
	CK(cudaSetDevice(ctx->device));
	TraceWS* T;
	int rc = trace_ws(ctx, &T, n);
	if (rc) return rc;
	cudaStream_t s = ctx->ba_stream;
	const size_t m = (size_t)n;
	CK(cudaMemcpyAsync(T->d_u, u, m * 4, cudaMemcpyHostToDevice, s)); CK(cudaMemcpyAsync(T->d_v, v, m * 4, cudaMemcpyHostToDevice, s));
	CK(cudaMemcpyAsync(T->d_color, color8, m * 32, cudaMemcpyHostToDevice, s)); CK(cudaMemcpyAsync(T->d_weights, weights8, m * 32, cudaMemcpyHostToDevice, s));
	CK(cudaMemcpyAsync(T->d_gradH, gradH4, m * 16, cudaMemcpyHostToDevice, s)); CK(cudaMemcpyAsync(T->d_energyTH, energyTH, m * 4, cudaMemcpyHostToDevice, s));
	CK(cudaMemcpyAsync(T->d_host, host_idx, m * 4, cudaMemcpyHostToDevice, s));
	CK(cudaMemcpyAsync(T->d_idmin, idepth_min, m * 4, cudaMemcpyHostToDevice, s)); CK(cudaMemcpyAsync(T->d_idmax, idepth_max, m * 4, cudaMemcpyHostToDevice, s));
	CK(cudaMemcpyAsync(T->d_quality, quality, m * 4, cudaMemcpyHostToDevice, s)); CK(cudaMemcpyAsync(T->d_status, status, m * 4, cudaMemcpyHostToDevice, s));
	if (lastTraceUV2) CK(cudaMemcpyAsync(T->d_uv, lastTraceUV2, m * 8, cudaMemcpyHostToDevice, s)); else CK(cudaMemsetAsync(T->d_uv, 0, m * 8, s));
	if (lastTracePixelInterval) CK(cudaMemcpyAsync(T->d_interval, lastTracePixelInterval, m * 4, cudaMemcpyHostToDevice, s)); else CK(cudaMemsetAsync(T->d_interval, 0, m * 4, s));
	CK(cudaStreamSynchronize(s));
	T->n = n;
	return 0;
}

extern "C" int dso_b200_trace_new_coarse(dso_b200_ctx* ctx, int frame_slot, int n_hosts, const float* KRKi9, const float* Kt3, const float* aff2,
                                         float* idepth_min_out, float* idepth_max_out, float* quality_out, int* status_out,
                                         float* lastTraceUV2_out, float* lastTracePixelInterval_out)
{
	ARG_CHECK(ctx && ctx->trace && ((TraceWS*)ctx->trace)->n > 0, "call dso_b200_trace_set_points first");
	ARG_CHECK(frame_slot >= 0 && frame_slot < ctx->n_slots && ctx->slots[frame_slot].valid, "frame slot empty");
	ARG_CHECK(n_hosts >= 1 && n_hosts <= MAXF && KRKi9 && Kt3 && aff2, "bad host arguments");
	CK(cudaSetDevice(ctx->device));
	int rc = frame_slot_wait(ctx, frame_slot, ctx->ba_stream, DSO_B200_STREAM_MAPPING);
	if (rc) return rc;
	TraceWS* T = (TraceWS*)ctx->trace;
	const FrameSlot& F = ctx->slots[frame_slot];
	TraceHosts H;
	memset(&H, 0, sizeof(H));
	for (int i = 0; i < n_hosts; i++)
	{
		memcpy(H.KRKi[i], KRKi9 + 9 * i, 36); memcpy(H.Kt[i], Kt3 + 3 * i, 12); memcpy(H.aff[i], aff2 + 2 * i, 8);
	}
	cudaStream_t s = ctx->ba_stream;
	const int n = T->n;
	trace_on_kernel<<<(n + TR_PTS_PER_BLOCK - 1) / TR_PTS_PER_BLOCK, 256, 0, s>>>(ctx->settings, n, F.pyr[0], F.w, F.h, H, T->d_u, T->d_v, T->d_color, T->d_weights, T->d_gradH,
	                                                                             T->d_energyTH, T->d_host, T->d_idmin, T->d_idmax, T->d_quality, T->d_status,
	                                                                             T->d_uv, T->d_interval);
	LAUNCHED(ctx, 1);
	CK(cudaGetLastError());
	const size_t m = (size_t)n;
	if (idepth_min_out) CK(cudaMemcpyAsync(idepth_min_out, T->d_idmin, m * 4, cudaMemcpyDeviceToHost, s));
	if (idepth_max_out) CK(cudaMemcpyAsync(idepth_max_out, T->d_idmax, m * 4, cudaMemcpyDeviceToHost, s));
	if (quality_out) CK(cudaMemcpyAsync(quality_out, T->d_quality, m * 4, cudaMemcpyDeviceToHost, s));
	if (status_out) CK(cudaMemcpyAsync(status_out, T->d_status, m * 4, cudaMemcpyDeviceToHost, s));
	if (lastTraceUV2_out) CK(cudaMemcpyAsync(lastTraceUV2_out, T->d_uv, m * 8, cudaMemcpyDeviceToHost, s));
	if (lastTracePixelInterval_out) CK(cudaMemcpyAsync(lastTracePixelInterval_out, T->d_interval, m * 4, cudaMemcpyDeviceToHost, s));
	if (idepth_min_out || idepth_max_out || quality_out || status_out || lastTraceUV2_out || lastTracePixelInterval_out) CK(cudaStreamSynchronize(s));
	return 0;
}

// =========================================================================================================
// FullSystem::optimizeImmaturePoint (FullSystemOptPoint.cpp:51-209) + ImmaturePoint::linearizeResidual
// (ImmaturePoint.cpp:499-570): the per-point Gauss-Newton on the inverse depth that decides about activation.
// Eight lanes per point again; the residuals to the <= 7 target frames are visited one after the other (the reference's
// running sums Hdd, bd, energy cross residual boundaries and keep the partial contribution of a residual that leaves
// early), the eight pattern samples of a residual in parallel with the sums taken in pattern order.
// =========================================================================================================
#define RS_IN 0
#define RS_OOB 1
#define RS_OUTLIER 2
#define ACT_MAXT (MAXF - 1)

struct ActTargets
{
	int nT, w;
	float fxl, fyl, cxl, cyl, fxli, fyli, wM3G, hM3G;
	const float4* img[ACT_MAXT];
	float RTll[ACT_MAXT][9], tTll[ACT_MAXT][3], aff[ACT_MAXT][2];
};

struct ActRes { int state, newState; float energy, newEnergy; };

// linearizeResidual for target t; returns the value the reference returns (as double), updates Hdd / bd
__device__ __forceinline__ double act_linearize(const dso_b200_settings& ST, const ActTargets& A, int t, float pu, float pv, float color, float weight, float energyTH,
                                                float slack, ActRes& r, float& Hdd, float& bd, float idepth, int idx, int base, unsigned gmask)
{
	if (r.state == RS_OOB) { r.newState = RS_OOB; return (double)r.energy; }            // :505-506
	const float dxp = (float)c_trace_pattern[idx][0], dyp = (float)c_trace_pattern[idx][1];
	// projectPoint, ResidualProjections.h:69-101
	const float k0 = fmul(fsub(fadd(pu, dxp), A.cxl), A.fxli), k1 = fmul(fsub(fadd(pv, dyp), A.cyl), A.fyli);
	const float* R = A.RTll[t]; const float* tt = A.tTll[t];
	float ptp[3];
#pragma unroll
	for (int i = 0; i < 3; i++) ptp[i] = fadd(dot3e(R[3 * i], k0, R[3 * i + 1], k1, R[3 * i + 2], 1.f), fmul(tt[i], idepth));
	const float drescale = fdiv(1.0f, ptp[2]);
	bool ok = drescale > 0.f;
	float u = 0.f, v = 0.f, Ku = 0.f, Kv = 0.f;
	if (ok)
	{
		u = fmul(ptp[0], drescale); v = fmul(ptp[1], drescale);
		Ku = fadd(fmul(u, A.fxl), A.cxl); Kv = fadd(fmul(v, A.fyl), A.cyl);
		ok = Ku > 1.1f && Kv > 1.1f && Ku < A.wM3G && Kv < A.hM3G;
	}
	float tE = 0.f, tH = 0.f, tb = 0.f;
	if (ok)
	{
		float h0, h1, h2;
		tr_interp(A.img[t], Ku, Kv, A.w, h0, h1, h2, true);
		if (!isfinite(h0)) ok = false;                                                    // :531
		else
		{
			const float residual = fsub(h0, fadd(fmul(A.aff[t][0], color), A.aff[t][1]));
			const float ar = fabsf(residual);
			float hw = ar < ST.huberTH ? 1.f : fdiv(ST.huberTH, ar);
			tE = fmul(fmul(fmul(fmul(fmul(weight, weight), hw), residual), residual), fsub(2.f, hw));
			const float dxI = fmul(h1, A.fxl), dyI = fmul(h2, A.fyl);
			const float d_id = fmul(fadd(fmul(fmul(dxI, drescale), fsub(tt[0], fmul(tt[2], u))), fmul(fmul(dyI, drescale), fsub(tt[1], fmul(tt[2], v)))), 1.0f);
			hw = fmul(hw, fmul(weight, weight));
			tH = fmul(fmul(hw, d_id), d_id);
			tb = fmul(fmul(hw, residual), d_id);
		}
	}
	// the first failing pattern pixel ends the residual; the pixels before it have already been added (:524-531)
	const unsigned failmask = (__ballot_sync(gmask, !ok) >> base) & 0xffu;
	const int firstFail = failmask ? (__ffs(failmask) - 1) : 8;
	float energyLeft = 0.f;
#pragma unroll
	for (int k = 0; k < 8; k++)
	{
		const float ek = __shfl_sync(gmask, tE, base + k), hk = __shfl_sync(gmask, tH, base + k), bk = __shfl_sync(gmask, tb, base + k);
		if (k < firstFail) { energyLeft = fadd(energyLeft, ek); Hdd = fadd(Hdd, hk); bd = fadd(bd, bk); }
	}
	if (firstFail < 8) { r.newState = RS_OOB; return (double)r.energy; }
	const float th = fmul(energyTH, slack);
	if (energyLeft > th) { energyLeft = th; r.newState = RS_OUTLIER; }                   // :555-563
	else r.newState = RS_IN;
	r.newEnergy = energyLeft;
	return (double)energyLeft;
}

__global__ void __launch_bounds__(256) optimize_immature_kernel(dso_b200_settings ST, int n, ActTargets A, int minObs, const float* __restrict__ pu, const float* __restrict__ pv,
                                                                const float* __restrict__ color8, const float* __restrict__ weights8,
                                                                const float* __restrict__ energyTH, const float* __restrict__ idmin,
                                                                const float* __restrict__ idmax, int* result, float* idepth_out, int* res_state)
{
	const int tid = threadIdx.x, lane = tid & 31, g = tid >> 3, idx = tid & 7, base = lane & ~7;
	const int k = blockIdx.x * TR_PTS_PER_BLOCK + g;
	if (k >= n) return;
	const unsigned gmask = 0xffu << base;
	const float u = pu[k], v = pv[k], color = color8[8 * (size_t)k + idx], weight = weights8[8 * (size_t)k + idx], eTH = energyTH[k];
	const int nres = A.nT;
	ActRes res[ACT_MAXT];
#pragma unroll
	for (int t = 0; t < ACT_MAXT; t++) { res[t].state = RS_IN; res[t].newState = RS_OUTLIER; res[t].energy = 0.f; res[t].newEnergy = 0.f; }   // :57-66
	float lastEnergy = 0.f, lastHdd = 0.f, lastbd = 0.f;
	float currentIdepth = fmul(fadd(idmax[k], idmin[k]), 0.5f);                           // :76
#pragma unroll
	for (int t = 0; t < ACT_MAXT; t++)
		if (t < nres)
		{
			lastEnergy = (float)((double)lastEnergy + act_linearize(ST, A, t, u, v, color, weight, eTH, 1000.f, res[t], lastHdd, lastbd, currentIdepth, idx, base, gmask));
			res[t].state = res[t].newState; res[t].energy = res[t].newEnergy;
		}
	int out = 1;
	if (!isfinite(lastEnergy) || lastHdd < ST.minIdepthH_act) out = 0;                     // :90-97 setting_minIdepthH_act
	if (out)
	{
		float lambda = 0.1f;
		for (int iteration = 0; iteration < ST.GNItsOnPointActivation; iteration++)       // setting_GNItsOnPointActivation
		{
			float H = lastHdd;
			H = fmul(H, fadd(1.f, lambda));
			const float step = (float)((1.0 / (double)H) * (double)lastbd);                 // :106
			const float newIdepth = fsub(currentIdepth, step);
			float newHdd = 0.f, newbd = 0.f, newEnergy = 0.f;
#pragma unroll
			for (int t = 0; t < ACT_MAXT; t++)
				if (t < nres)
					newEnergy = (float)((double)newEnergy + act_linearize(ST, A, t, u, v, color, weight, eTH, 1.f, res[t], newHdd, newbd, newIdepth, idx, base, gmask));
			if (!isfinite(lastEnergy) || newHdd < ST.minIdepthH_act) { out = 0; break; }                // :117-124
			if (newEnergy < lastEnergy)
			{
				currentIdepth = newIdepth; lastHdd = newHdd; lastbd = newbd; lastEnergy = newEnergy;
#pragma unroll
				for (int t = 0; t < ACT_MAXT; t++) { res[t].state = res[t].newState; res[t].energy = res[t].newEnergy; }
				lambda = fmul(lambda, 0.5f);
			}
			else lambda = fmul(lambda, 5.f);
			if ((double)fabsf(step) < 0.0001 * (double)currentIdepth) break;                // :151
		}
	}
	if (out)
	{
		if (!isfinite(currentIdepth)) out = -1;                                           // :155-159
		else
		{
			int good = 0;
#pragma unroll
			for (int t = 0; t < ACT_MAXT; t++) if (t < nres && res[t].state == RS_IN) good++;
			if (good < minObs || !isfinite(eTH)) out = -1;                                  // :162-174
		}
	}
	if (idx == 0)
	{
		result[k] = out;
		idepth_out[k] = out == 1 ? currentIdepth : __int_as_float(0x7fc00000);
#pragma unroll
		for (int t = 0; t < ACT_MAXT; t++) if (t < nres) res_state[(size_t)k * nres + t] = res[t].state;
	}
}

extern "C" int dso_b200_optimize_immature(dso_b200_ctx* ctx, int n_targets, const int* target_slots, float fx, float fy, float cx, float cy,
                                          const float* RTll9, const float* tTll3, const float* aff2,
                                          int n, const float* u, const float* v, const float* color8, const float* weights8, const float* energyTH,
                                          const float* idepth_min, const float* idepth_max, int minObs,
                                          int* result, float* idepth_out, int* res_state_out)
{
	ARG_CHECK(ctx && target_slots && RTll9 && tTll3 && aff2 && u && v && color8 && weights8 && energyTH && idepth_min && idepth_max && result && idepth_out,
	          "null argument");
	ARG_CHECK(n_targets >= 1 && n_targets <= ACT_MAXT && n >= 1, "1..7 targets, n >= 1");
	CK(cudaSetDevice(ctx->device));
	ActTargets A;
	memset(&A, 0, sizeof(A));
	A.nT = n_targets;
	for (int t = 0; t < n_targets; t++)
	{
		const int sl = target_slots[t];
		ARG_CHECK(sl >= 0 && sl < ctx->n_slots && ctx->slots[sl].valid, "target frame slot empty");
		ARG_CHECK(t == 0 || (ctx->slots[sl].w == A.w), "target frames differ in size");
		int rc = frame_slot_wait(ctx, sl, ctx->ba_stream, DSO_B200_STREAM_MAPPING);
		if (rc) return rc;
		A.w = ctx->slots[sl].w;
		A.img[t] = ctx->slots[sl].pyr[0];
		memcpy(A.RTll[t], RTll9 + 9 * t, 36); memcpy(A.tTll[t], tTll3 + 3 * t, 12); memcpy(A.aff[t], aff2 + 2 * t, 8);
	}
	const int h = ctx->slots[target_slots[0]].h;
	// CalibHessian::setValueScaled (HessianBlocks.h:383-397) and wM3G / hM3G (globalCalib.cpp:70-71)
	A.fxl = fx; A.fyl = fy; A.cxl = cx; A.cyl = cy; A.fxli = 1.0f / fx; A.fyli = 1.0f / fy;
	A.wM3G = (float)(A.w - 3); A.hM3G = (float)(h - 3);
	// scratch: reuse the trace workspace's arrays of the same shape (separate point set: allocate a private block)
	const size_t m = (size_t)n;
	// grow-only scratch kept in the context (cudaMalloc / cudaFree per call would stall the whole device, including the
	// tracking thread's stream)
	const size_t need = m * (4 + 4 + 32 + 32 + 4 + 4 + 4 + 4 + 4) + m * n_targets * 4;
	if (need > ctx->act_scratch_bytes)
	{
		CK(cudaStreamSynchronize(ctx->ba_stream));
		if (ctx->act_scratch) CK(cudaFree(ctx->act_scratch));
		ctx->act_scratch = nullptr; ctx->act_scratch_bytes = 0;
		const size_t cap = need + need / 2;
		CK(cudaMalloc(&ctx->act_scratch, cap));
		ctx->act_scratch_bytes = cap;
	}
	float* d = (float*)ctx->act_scratch;
	float* d_u = d; float* d_v = d_u + m; float* d_c = d_v + m; float* d_w = d_c + 8 * m; float* d_e = d_w + 8 * m;
	float* d_lo = d_e + m; float* d_hi = d_lo + m; float* d_id = d_hi + m; int* d_res = (int*)(d_id + m); int* d_rs = d_res + m;
	cudaStream_t s = ctx->ba_stream;
	cudaError_t e = cudaSuccess;
#define UP(dst, src, bytes) if (e == cudaSuccess) e = cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, s)
	UP(d_u, u, m * 4); UP(d_v, v, m * 4); UP(d_c, color8, m * 32); UP(d_w, weights8, m * 32); UP(d_e, energyTH, m * 4);
	UP(d_lo, idepth_min, m * 4); UP(d_hi, idepth_max, m * 4);
#undef UP
	if (e == cudaSuccess)
	{
		optimize_immature_kernel<<<(n + TR_PTS_PER_BLOCK - 1) / TR_PTS_PER_BLOCK, 256, 0, s>>>(ctx->settings, n, A, minObs, d_u, d_v, d_c, d_w, d_e, d_lo, d_hi, d_res, d_id, d_rs);
		LAUNCHED(ctx, 1);
		e = cudaGetLastError();
	}
	if (e == cudaSuccess) e = cudaMemcpyAsync(result, d_res, m * 4, cudaMemcpyDeviceToHost, s);
	if (e == cudaSuccess) e = cudaMemcpyAsync(idepth_out, d_id, m * 4, cudaMemcpyDeviceToHost, s);
	if (e == cudaSuccess && res_state_out) e = cudaMemcpyAsync(res_state_out, d_rs, m * n_targets * 4, cudaMemcpyDeviceToHost, s);
	if (e == cudaSuccess) e = cudaStreamSynchronize(s);
	CK(e);
	return 0;
}
