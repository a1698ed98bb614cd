// K2 -- frame-to-keyframe tracker, replaces CoarseTracker::calcRes / calcGSSSE / trackNewestCoarse
// (reference: FullSystem/CoarseTracker.cpp:296-354, 359-520, 551-747).
//
// One evaluation = for every reference point: warp, bounds test, bilinear sample of the new frame's
// float4 (I,dx,dy,.) texels, affine-corrected residual, Huber weight, cut-off, and the weighted outer
// product of the 8-DoF Jacobian [6 pose, a, b | r] -- calcRes and calcGSSSE fused, so the reference's
// buf_warped_* arrays never exist.  Everything that decides a mask or a count is computed with
// individually rounded float ops in the reference's evaluation order (bit-exact vs. its SSE2 build);
// the 44 sums of J^T W J / J^T W r are FP32 per thread, then a fixed-order warp-shuffle / shared-memory
// / cluster tree (deterministic, no atomics on data).
//
//   track_eval_kernel        one evaluation, grid-wide, last-CTA-done final sum (drop-in for the two
//                            private methods; host drives the LM loop)
//   track_persistent_kernel  the whole coarse-to-fine LM loop of trackNewestCoarse for up to 32 pose
//                            hypotheses, one thread-block cluster per hypothesis; partial sums are
//                            exchanged through distributed shared memory, the 8x8 LDL^T solve and the
//                            SE3 update run in FP64 on the device; no host round trip until all
//                            levels are done.
//
// Roofline: gather-bound (L2 at VGA, HBM for the large sweeps): per point 16 B of cloud + 4 x 16 B
// texels, ~170 flop.
#include "common.cuh"
#include <atomic>
#include "lm_math.cuh"
#include <cooperative_groups.h>
namespace cg = cooperative_groups;

#define SCALE_XI_ROT 1.0f
#define SCALE_XI_TRANS 0.5f
#define SCALE_A 10.0f
#define SCALE_B 1000.0f

#define EVAL_THREADS 256
#define PERS_THREADS 256
#define MAX_CLUSTER 16
#define DSO_B200_TRACK_MAX_NC 16      // clusters one hypothesis may be spread over

struct EvalParams
{
	float RKi[9], t[3], Ki[9];
	float affLL0, affLL1;
	float fx, fy, cx, cy;
	float wl_m3, hl_m3;
	int wl, n, lvl0;
	float cutoffTH, maxEnergy, huberTH;
	float b0;
};

struct TrackLevels
{
	int levels;
	int w[MAXL], h[MAXL];
	float fx[MAXL], fy[MAXL], cx[MAXL], cy[MAXL];
	float Ki[MAXL][9];
	const float4* pc[MAXL];
	int pc_n[MAXL];
	const float4* img[MAXL];
	double ref_a, ref_b;
	float ref_exposure, new_exposure;
};

// CoarseTracker.cpp:375-378, 385 : pose/affine (double) -> the float quantities of one calcRes call
HD void make_eval_params(const TrackLevels& TL, int lvl, const Pose& refToNew, double aff_a, double aff_b,
                         float cutoffTH, float huberTH, EvalParams& P)
{
	float Rf[9];
	for (int i = 0; i < 9; i++) Rf[i] = (float)refToNew.R[i];
	mat33f_mul(Rf, TL.Ki[lvl], P.RKi);
	for (int i = 0; i < 3; i++) P.t[i] = (float)refToNew.t[i];
	for (int i = 0; i < 9; i++) P.Ki[i] = TL.Ki[lvl][i];
	double a, b;
	aff_from_to(TL.ref_exposure, TL.new_exposure, TL.ref_a, TL.ref_b, aff_a, aff_b, a, b);
	P.affLL0 = (float)a; P.affLL1 = (float)b;
	P.fx = TL.fx[lvl]; P.fy = TL.fy[lvl]; P.cx = TL.cx[lvl]; P.cy = TL.cy[lvl];
	P.wl = TL.w[lvl];
	P.wl_m3 = (float)(TL.w[lvl] - 3); P.hl_m3 = (float)(TL.h[lvl] - 3);
	P.n = TL.pc_n[lvl];
	P.lvl0 = (lvl == 0);
	P.cutoffTH = cutoffTH;
	P.huberTH = huberTH;
#ifdef __CUDA_ARCH__
	P.maxEnergy = __fsub_rn(__fmul_rn(__fmul_rn(2.0f, huberTH), cutoffTH), __fmul_rn(huberTH, huberTH));
#else
	{ volatile float t0 = 2 * huberTH; volatile float t1 = t0 * cutoffTH; volatile float t2 = huberTH * huberTH; P.maxEnergy = t1 - t2; }
#endif
	P.b0 = (float)TL.ref_b;
}

__device__ __forceinline__ void proj3(const float* M, float x, float y, const float* t, float id, float sgn, float* pt)
{
#pragma unroll
	for (int r = 0; r < 3; r++)
	{
		float p = dot3e(M[r * 3], x, M[r * 3 + 1], y, M[r * 3 + 2], 1.0f);
		float tid = fmul(t[r], id);
		pt[r] = sgn > 0 ? fadd(p, tid) : fsub(p, tid);
	}
}

// One reference point of calcRes (CoarseTracker.cpp:402-486) + its calcGSSSE contribution (:313-334), in two stages so that
// a thread can have the texel gathers of several points in flight before it consumes the first one:
//   eval_issue   warp, flow indicators, bounds test, interpolation weights, the four 16-byte texel loads
//   eval_finish  bilinear sample, residual, Huber weight, cut-off, Jacobian and the 44 weighted products
struct PtIssue
{
	float4 p00, p10, p01, p11;
	float w00, w01, w10, w11;
	float u, v, new_idepth, refColor;
	int i;
	bool live;
};

template <bool DEBUG>
__device__ __forceinline__ void eval_issue(const float4 pc, int i, const EvalParams& P, const float4* __restrict__ img,
                                           float* acc, unsigned char* status, PtIssue& q)
{
	q.live = false; q.i = i;
	const float id = pc.z, x = pc.x, y = pc.y;
	float pt[3];
	proj3(P.RKi, x, y, P.t, id, 1.f, pt);                // :408
	const float u = fdiv(pt[0], pt[2]);
	const float v = fdiv(pt[1], pt[2]);
	const float Ku = fadd(fmul(P.fx, u), P.cx);
	const float Kv = fadd(fmul(P.fy, v), P.cy);
	const float new_idepth = fdiv(id, pt[2]);

	if (P.lvl0 && (i & 31) == 0)                         // :416-447 flow indicators
	{
		float t[3];
		proj3(P.Ki, x, y, P.t, id, 1.f, t);
		float KuT = fadd(fmul(P.fx, fdiv(t[0], t[2])), P.cx), KvT = fadd(fmul(P.fy, fdiv(t[1], t[2])), P.cy);
		proj3(P.Ki, x, y, P.t, id, -1.f, t);
		float KuT2 = fadd(fmul(P.fx, fdiv(t[0], t[2])), P.cx), KvT2 = fadd(fmul(P.fy, fdiv(t[1], t[2])), P.cy);
		proj3(P.RKi, x, y, P.t, id, -1.f, t);
		float Ku3 = fadd(fmul(P.fx, fdiv(t[0], t[2])), P.cx), Kv3 = fadd(fmul(P.fy, fdiv(t[1], t[2])), P.cy);
		float a0 = fsub(KuT, x), a1 = fsub(KvT, y), b0 = fsub(KuT2, x), b1 = fsub(KvT2, y);
		float c0 = fsub(Ku, x), c1 = fsub(Kv, y), d0 = fsub(Ku3, x), d1 = fsub(Kv3, y);
		acc[S_FLOW_T] += fadd(fmul(a0, a0), fmul(a1, a1));
		acc[S_FLOW_T] += fadd(fmul(b0, b0), fmul(b1, b1));
		acc[S_FLOW_RT] += fadd(fmul(c0, c0), fmul(c1, c1));
		acc[S_FLOW_RT] += fadd(fmul(d0, d0), fmul(d1, d1));
		acc[S_FLOW_N] += 2.f;
	}

	if (DEBUG) status[i] = 0;
	if (!(Ku > 2.f && Kv > 2.f && Ku < P.wl_m3 && Kv < P.hl_m3 && new_idepth > 0.f)) return;   // :449

	// getInterpolatedElement33, util/globalFuncs.h:75-89, on the float4 texel layout
	const int ix = (int)Ku, iy = (int)Kv;
	const float dx = fsub(Ku, (float)ix), dy = fsub(Kv, (float)iy);
	const float dxdy = fmul(dx, dy);
	q.w11 = dxdy; q.w01 = fsub(dy, dxdy); q.w10 = fsub(dx, dxdy);
	q.w00 = fadd(fsub(fsub(1.f, dx), dy), dxdy);
	const float4* bp = img + ix + iy * P.wl;
	q.p00 = __ldg(bp); q.p10 = __ldg(bp + 1); q.p01 = __ldg(bp + P.wl); q.p11 = __ldg(bp + 1 + P.wl);
	q.u = u; q.v = v; q.new_idepth = new_idepth; q.refColor = pc.w;
	q.live = true;
}

template <bool DEBUG>
__device__ __forceinline__ void eval_finish(const PtIssue& q, const EvalParams& P, float* acc, unsigned char* status, float* warped)
{
	if (!q.live) return;
	const float h0 = fadd(fadd(fadd(fmul(q.w11, q.p11.x), fmul(q.w01, q.p01.x)), fmul(q.w10, q.p10.x)), fmul(q.w00, q.p00.x));
	const float h1 = fadd(fadd(fadd(fmul(q.w11, q.p11.y), fmul(q.w01, q.p01.y)), fmul(q.w10, q.p10.y)), fmul(q.w00, q.p00.y));
	const float h2 = fadd(fadd(fadd(fmul(q.w11, q.p11.z), fmul(q.w01, q.p01.z)), fmul(q.w10, q.p10.z)), fmul(q.w00, q.p00.z));
	if (!isfinite(h0)) return;                           // :456
	const float u = q.u, v = q.v, new_idepth = q.new_idepth, refColor = q.refColor;
	const float residual = fsub(h0, fadd(fmul(P.affLL0, refColor), P.affLL1));   // :458
	const float ar = fabsf(residual);
	const float hw = ar < P.huberTH ? 1.f : fdiv(P.huberTH, ar);                  // :459

	if (ar > P.cutoffTH)                                 // :462-468
	{
		acc[S_E] += P.maxEnergy;
		acc[S_NTERMS] += 1.f;
		acc[S_NSAT] += 1.f;
		if (DEBUG) status[q.i] = 1;
		return;
	}
	acc[S_E] += fmul(fmul(fmul(hw, residual), residual), fsub(2.f, hw));   // :473
	acc[S_NTERMS] += 1.f;
	acc[S_NWARP] += 1.f;
	if (DEBUG)
	{
		status[q.i] = 2;
		float* wp = warped + 8 * (size_t)q.i;
		wp[0] = new_idepth; wp[1] = u; wp[2] = v; wp[3] = h1; wp[4] = h2; wp[5] = residual; wp[6] = hw; wp[7] = refColor;
	}
	// calcGSSSE :313-334
	const float gx = fmul(h1, P.fx), gy = fmul(h2, P.fy);
	float J[9];
	J[0] = fmul(new_idepth, gx);
	J[1] = fmul(new_idepth, gy);
	J[2] = -fmul(new_idepth, fadd(fmul(u, gx), fmul(v, gy)));
	J[3] = -fadd(fmul(fmul(u, v), gx), fmul(gy, fadd(1.f, fmul(v, v))));
	J[4] = fadd(fmul(fmul(u, v), gy), fmul(gx, fadd(1.f, fmul(u, u))));
	J[5] = fsub(fmul(u, gy), fmul(v, gx));
	J[6] = fmul(P.affLL0, fsub(P.b0, refColor));
	J[7] = -1.f;
	J[8] = residual;
	int k = S_ACC;
#pragma unroll
	for (int r = 0; r < 8; r++)
	{
		const float Jw = fmul(J[r], hw);
#pragma unroll
		for (int c = r; c < 8; c++) { acc[k] = fmaf(Jw, J[c], acc[k]); k++; }
	}
#pragma unroll
	for (int r = 0; r < 8; r++) { acc[k] = fmaf(fmul(J[r], hw), J[8], acc[k]); k++; }
}

template <bool DEBUG>
__device__ __forceinline__ void eval_point(const float4 pc, int i, const EvalParams& P, const float4* __restrict__ img,
                                           float* acc, unsigned char* status, float* warped)
{
	PtIssue q;
	eval_issue<DEBUG>(pc, i, P, img, acc, status, q);
	eval_finish<DEBUG>(q, P, acc, status, warped);
}

// Fixed-order CTA reduction of S_COUNT per-thread sums: xor-butterfly inside each warp, then the
// warps in index order.  Result in s_out[0..S_COUNT) (valid after the trailing __syncthreads).
template <int NWARPS>
__device__ __forceinline__ void block_reduce(float* acc, float (*s_w)[S_COUNT], float* s_out)
{
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
	for (int k = 0; k < S_COUNT; k++)
	{
		float v = acc[k];
#pragma unroll
		for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
		acc[k] = v;
	}
	if (lane == 0)
	{
#pragma unroll
		for (int k = 0; k < S_COUNT; k++) s_w[warp][k] = acc[k];
	}
	__syncthreads();
	if (threadIdx.x < S_COUNT)
	{
		float s = 0.f;
#pragma unroll
		for (int w = 0; w < NWARPS; w++) s += s_w[w][threadIdx.x];
		s_out[threadIdx.x] = s;
	}
	__syncthreads();
}

// ---- K2a ------------------------------------------------------------------------------------------
template <bool DEBUG>
__global__ void __launch_bounds__(EVAL_THREADS) track_eval_kernel(EvalParams P, const float4* __restrict__ pts,
                                                                  const float4* __restrict__ img, float* partials,
                                                                  unsigned int* ticket, float* sums_out,
                                                                  unsigned char* status, float* warped)
{
	__shared__ float s_w[EVAL_THREADS / 32][S_COUNT];
	__shared__ float s_out[S_COUNT];
	__shared__ bool s_last;
	float acc[S_COUNT];
#pragma unroll
	for (int k = 0; k < S_COUNT; k++) acc[k] = 0.f;
	for (int i = blockIdx.x * EVAL_THREADS + threadIdx.x; i < P.n; i += gridDim.x * EVAL_THREADS)
		eval_point<DEBUG>(__ldg(pts + i), i, P, img, acc, status, warped);
	block_reduce<EVAL_THREADS / 32>(acc, s_w, s_out);
	if (threadIdx.x < S_COUNT) partials[blockIdx.x * S_COUNT + threadIdx.x] = s_out[threadIdx.x];
	__threadfence();
	__syncthreads();
	if (threadIdx.x == 0)
	{
		unsigned int t = atomicAdd(ticket, 1u);
		s_last = (t == gridDim.x - 1);
	}
	__syncthreads();
	if (s_last)
	{
		__threadfence();
		if (threadIdx.x < S_COUNT)
		{
			float s = 0.f;
			for (unsigned b = 0; b < gridDim.x; b++) s += __ldcg(partials + b * S_COUNT + threadIdx.x);
			sums_out[threadIdx.x] = s;
		}
		if (threadIdx.x == 0) *ticket = 0;
	}
}

// ---- LM bookkeeping shared by the host loop and the persistent kernel --------------------------------
struct EvalResult
{
	double E, n, flowT, flowRT, sat;   // Vec6 rs of calcRes (:511-519); rs[3] is always 0
	double H[64], b[8];                // calcGSSSE output (:338-353)
};

// sums of one evaluation -> the reference's Vec6 and scaled H, b
HD void sums_to_result(const float* s, EvalResult& r)
{
	r.E = (double)s[S_E];
	r.n = (double)s[S_NTERMS];
	r.flowT = (double)s[S_FLOW_T] / ((double)s[S_FLOW_N] + 0.1);    // float / (float + 0.1) -> double
	r.flowRT = (double)s[S_FLOW_RT] / ((double)s[S_FLOW_N] + 0.1);
	r.sat = (double)(s[S_NSAT] / s[S_NTERMS]);                       // numSaturated / (float)numTermsInE
	const int nw = (int)s[S_NWARP];
	const int npad = (nw + 3) & ~3;                                  // :488-499 padding to a multiple of 4
	const float invn = 1.0f / (float)npad;
	const float sc[8] = { SCALE_XI_ROT, SCALE_XI_ROT, SCALE_XI_ROT, SCALE_XI_TRANS, SCALE_XI_TRANS, SCALE_XI_TRANS, SCALE_A, SCALE_B };
	int k = S_ACC;
	for (int i = 0; i < 8; i++)
		for (int j = i; j < 8; j++)
		{
			double v = (double)s[k++] * (double)invn;
			v *= (double)sc[j];
			v *= (double)sc[i];
			r.H[i * 8 + j] = v;
			r.H[j * 8 + i] = v;
		}
	for (int i = 0; i < 8; i++) r.b[i] = (double)s[k++] * (double)invn * (double)sc[i];
}

// State machine of trackNewestCoarse (CoarseTracker.cpp:551-747): feed it the result of each
// evaluation, it answers with the next evaluation to run (pose, affine, level, cut-off) or "done".
struct LM
{
	dso_b200_track_opts opt;
	int lvl, iteration, phase, haveRepeated, done, aborted;
	float levelCutoffRepeat, lambda;
	Pose cur, nw;
	double a_cur, b_cur, a_new, b_new;
	double resOldE, resOldN, resOldFlowT, resOldFlowRT;
	double H[64], b[8];
	double incnorm;
	double minRes[5];
	double ref_a, ref_b;            // lastRef_aff_g2l
	float ref_exposure, new_exposure;
	dso_b200_track_result* out;

	// request produced for the caller
	int req_lvl; float req_cutoff; Pose req_pose; double req_a, req_b;

	HD void request(int l, const Pose& p, double a, double b_)
	{
		req_lvl = l; req_pose = p; req_a = a; req_b = b_;
		req_cutoff = opt.coarseCutoffTH * levelCutoffRepeat;
	}
	HD void init(const dso_b200_track_opts& o, const TrackLevels& TL, const double* pose12, double a0, double b0,
	             int coarsestLvl, const double* minResForAbort, dso_b200_track_result* res)
	{
		opt = o; out = res;
		ref_a = TL.ref_a; ref_b = TL.ref_b; ref_exposure = TL.ref_exposure; new_exposure = TL.new_exposure;
		pose_from12(cur, pose12); a_cur = a0; b_cur = b0;
		haveRepeated = 0; done = 0; aborted = 0;
		for (int i = 0; i < 5; i++) minRes[i] = minResForAbort ? minResForAbort[i] : NAN;
		out->ok = 0; out->reserved = -1;
		for (int i = 0; i < 12; i++) out->pose[i] = pose12[i];
		out->aff_a = a0; out->aff_b = b0;
		for (int i = 0; i < 5; i++) out->lastResiduals[i] = NAN;        // :563
		for (int i = 0; i < 3; i++) out->lastFlowIndicators[i] = 1000;  // :564
		out->n_passes = 0; out->n_iterations = 0; out->n_evals = 0; out->trace_len = 0; out->point_evals = 0;
		start_level(coarsestLvl);
	}
	HD void start_level(int l)
	{
		lvl = l; phase = 0; levelCutoffRepeat = 1; iteration = 0;
		request(lvl, cur, a_cur, b_cur);   // :584
	}
	// :613-665 solve for the increment and propose the new state
	HD void propose()
	{
		double Hl[64], nb[8], inc[8];
		for (int i = 0; i < 64; i++) Hl[i] = H[i];
		for (int i = 0; i < 8; i++) Hl[i * 8 + i] *= (1 + lambda);
		for (int i = 0; i < 8; i++) nb[i] = -b[i];
		ldlt_solve(Hl, nb, inc, 8);
		const bool fixA = opt.affineOptModeA < 0, fixB = opt.affineOptModeB < 0;
		if (fixA && fixB) { ldlt_solve(Hl, nb, inc, 6); inc[6] = inc[7] = 0; }
		if (!fixA && fixB) { ldlt_solve(Hl, nb, inc, 7); inc[7] = 0; }
		if (fixA && !fixB)
		{
			double Hs[64], bs[8], is[8];
			for (int i = 0; i < 64; i++) Hs[i] = Hl[i];
			for (int i = 0; i < 8; i++) bs[i] = nb[i];
			for (int i = 0; i < 8; i++) Hs[i * 8 + 6] = Hs[i * 8 + 7];
			for (int i = 0; i < 8; i++) Hs[6 * 8 + i] = Hs[7 * 8 + i];
			bs[6] = bs[7];
			ldlt_solve(Hs, bs, is, 7);
			for (int i = 0; i < 6; i++) inc[i] = is[i];
			inc[6] = 0; inc[7] = is[6];
		}
		float extrapFac = 1;
		const float lambdaExtrapolationLimit = 0.001f;
		if (lambda < lambdaExtrapolationLimit) extrapFac = sqrtf(sqrtf(lambdaExtrapolationLimit / lambda));   // :650
		double nrm = 0;
		for (int i = 0; i < 8; i++) { inc[i] *= extrapFac; nrm += inc[i] * inc[i]; }
		incnorm = sqrt(nrm);
		double is[8];
		for (int i = 0; i < 8; i++) is[i] = inc[i];
		for (int i = 0; i < 3; i++) is[i] *= SCALE_XI_ROT;
		for (int i = 3; i < 6; i++) is[i] *= SCALE_XI_TRANS;
		is[6] *= SCALE_A; is[7] *= SCALE_B;
		double sum = 0;
		for (int i = 0; i < 8; i++) sum += is[i];
		if (!isfinite(sum)) for (int i = 0; i < 8; i++) is[i] = 0;   // :659
		Pose e;
		se3_exp(is, e);
		pose_mul(e, cur, nw);                                          // :661
		a_new = a_cur + is[6]; b_new = b_cur + is[7];
		request(lvl, nw, a_new, b_new);                                // :667
	}
	HD void finish_level()
	{
		const float rmse = sqrtf((float)(resOldE / resOldN));          // :710
		out->lastResiduals[lvl] = rmse;
		out->lastFlowIndicators[0] = resOldFlowT; out->lastFlowIndicators[1] = 0; out->lastFlowIndicators[2] = resOldFlowRT;
		if (out->n_passes < DSO_B200_MAX_PASSES) { out->pass_level[out->n_passes] = lvl; out->pass_rmse[out->n_passes] = rmse; out->n_passes++; }
		if (out->lastResiduals[lvl] > 1.5 * minRes[lvl]) { done = 1; aborted = 1; out->ok = 0; out->reserved = lvl; return; }   // :713
		int next = lvl - 1;
		if (levelCutoffRepeat > 1 && !haveRepeated) { next = lvl; haveRepeated = 1; }   // :717-722
		if (next < 0) { finalize(); return; }
		start_level(next);
	}
	HD void finalize()
	{
		done = 1;
		pose_to12(cur, out->pose);                                     // :727-728
		out->aff_a = a_cur; out->aff_b = b_cur;
		out->ok = 1;
		if ((opt.affineOptModeA != 0 && (fabsf((float)out->aff_a) > 1.2f)) || (opt.affineOptModeB != 0 && (fabsf((float)out->aff_b) > 200.f)))
		{ out->ok = 0; return; }
		double ra, rb;
		aff_from_to(ref_exposure, new_exposure, ref_a, ref_b, out->aff_a, out->aff_b, ra, rb);   // :735-739
		if ((opt.affineOptModeA == 0 && (fabsf(logf((float)ra)) > 1.5f)) || (opt.affineOptModeB == 0 && (fabsf((float)rb) > 200.f)))
		{ out->ok = 0; return; }
		if (opt.affineOptModeA < 0) out->aff_a = 0;
		if (opt.affineOptModeB < 0) out->aff_b = 0;
	}
	// feed one evaluation result; afterwards either done!=0 or req_* holds the next evaluation
	HD void step(const EvalResult& r, int n_points)
	{
		out->n_evals++; out->point_evals += n_points;
		if (phase == 0)
		{
			if (r.sat > 0.6 && levelCutoffRepeat < 50)                 // :586-593
			{
				levelCutoffRepeat *= 2;
				request(lvl, cur, a_cur, b_cur);
				return;
			}
			resOldE = r.E; resOldN = r.n; resOldFlowT = r.flowT; resOldFlowRT = r.flowRT;
			for (int i = 0; i < 64; i++) H[i] = r.H[i];
			for (int i = 0; i < 8; i++) b[i] = r.b[i];
			lambda = 0.01f;
			iteration = 0;
			phase = 1;
			if (iteration >= opt.maxIterations[lvl]) { finish_level(); return; }
			propose();
			return;
		}
		const bool accept = (r.E / r.n) < (resOldE / resOldN);         // :669
		if (out->trace_len < DSO_B200_TRACE_MARGINS)
		{
			out->trace_dE[out->trace_len] = (float)(r.E * resOldN - resOldE * r.n);
			out->trace_E[out->trace_len] = (float)(resOldE * r.n);
		}
		if (out->trace_len < 256) out->trace[out->trace_len++] = accept ? 1 : 0;
		out->n_iterations++;
		if (accept)
		{
			for (int i = 0; i < 64; i++) H[i] = r.H[i];
			for (int i = 0; i < 8; i++) b[i] = r.b[i];
			resOldE = r.E; resOldN = r.n; resOldFlowT = r.flowT; resOldFlowRT = r.flowRT;
			a_cur = a_new; b_cur = b_new; cur = nw;
			lambda *= 0.5f;
		}
		else
		{
			lambda *= 4;
			if (lambda < 0.001f) lambda = 0.001f;
		}
		iteration++;
		if (!(incnorm > 1e-3) || iteration >= opt.maxIterations[lvl]) { finish_level(); return; }   // :700, :613
		propose();
	}
};

// ---- K2b ------------------------------------------------------------------------------------------
struct TrackIn
{
	double pose[12];
	double aff_a, aff_b;
	double minRes[5];
};

// LM state of one hypothesis.  The serial dependency chain of trackNewestCoarse (solve -> exp -> warp -> reduce ->
// accept?) is the latency floor of this kernel, so the step between two evaluations is executed by one warp and its state
// lives in that warp's registers across iterations.  One warp issues in order, so the instruction count matters as much as
// the dependent chains: the 8x9 system [H | -b] is spread over 24 lanes -- lane i + 8g (row i = lane & 7, group g = 0..2)
// holds the three entries of row i in columns g, g+3, g+6 (column 8 is the right-hand side) -- lanes 0..11 hold the
// entries of the current and the proposed pose, every scalar is replicated in all lanes.  What the other warps need (the
// next EvalParams, the level, the done flag) and the two small vectors every lane reads with its own indices (the
// increment, the current pose) go through shared memory.
struct LMR
{
	double m[3];                                  // accepted system (scaled as CoarseTracker.cpp:338-353, without the common 1/n)
	double pe_cur, pe_new;                        // pose entry `lane` (R row-major 0..8, t 9..11) of refToNew_current / _new
	double a_cur, b_cur, a_new, b_new;
	double incnorm2;                              // squared norm of the last increment
	float oldE, oldN, oldFT, oldFRT, oldFN;       // raw float sums of the accepted state (resOld)
	float lambda, rep;                            // rep = levelCutoffRepeat
	int lvl, iteration, phase, haveRepeated, done;
	int n_evals, n_iter, trace_len, n_passes;
	double point_evals;
};
// per-lane constants of the LM warp
struct LMC
{
	int i, g;                                     // row, column group (lanes 24..31 mirror group 0)
	double si;                                    // SCALE of parameter i
	int off[3];                                   // offsets into the evaluation sums of this lane's three entries
	double scl[3];                                // their scale factors: SCALE_i * SCALE_c, -SCALE_i for the right-hand side
	bool diag[3];                                 // entry s is H_ii
	bool fixable[3];                              // entry lies in row / column 6 or 7 (fixed-affine modes)
	int xa, xb, xc;                               // pose lanes: rows r, r+1, r+2 (cyclic) -> indices into the increment's rotation part
	int pa, pb, pc;                               // and into the current pose (column c of those rows)
	bool tl;                                      // translation lane (9..11)
};
struct LMShared
{
	double x[8];                                  // scaled increment of the last solve
	double cur[12];                               // refToNew_current
};

__device__ __forceinline__ void lm_constants(int lane, LMC& K)
{
	const float sc[8] = { SCALE_XI_ROT, SCALE_XI_ROT, SCALE_XI_ROT, SCALE_XI_TRANS, SCALE_XI_TRANS, SCALE_XI_TRANS, SCALE_A, SCALE_B };
	K.i = lane & 7;
	K.g = (lane >> 3) == 3 ? 0 : (lane >> 3);
	float si = 0.f;
#pragma unroll
	for (int q = 0; q < 8; q++) if (q == K.i) si = sc[q];
	K.si = (double)si;
#pragma unroll
	for (int t = 0; t < 3; t++)
	{
		const int c = K.g + 3 * t;
		if (c < 8)
		{
			const int a = K.i < c ? K.i : c, b = K.i < c ? c : K.i;
			K.off[t] = S_ACC + a * 8 - ((a * (a - 1)) >> 1) + (b - a);
			float scc = 0.f;
#pragma unroll
			for (int q = 0; q < 8; q++) if (q == c) scc = sc[q];
			K.scl[t] = (double)si * (double)scc;
		}
		else { K.off[t] = S_ACC + 36 + K.i; K.scl[t] = -(double)si; }
		K.diag[t] = (c == K.i);
		K.fixable[t] = (K.i >= 6) || (c == 6 || c == 7);
	}
	const int r = (lane < 9) ? lane / 3 : (lane < 12 ? lane - 9 : 0);
	const int c = (lane < 9) ? lane - 3 * r : 3;
	const int r1 = (r + 1) % 3, r2 = (r + 2) % 3;
	K.xa = 3 + r; K.xb = 3 + r1; K.xc = 3 + r2;
	K.pa = c < 3 ? 3 * r + c : 9 + r; K.pb = c < 3 ? 3 * r1 + c : 9 + r1; K.pc = c < 3 ? 3 * r2 + c : 9 + r2;
	K.tl = (lane >= 9 && lane < 12);
}

__device__ __forceinline__ double shfl_d(double v, int src) { return __shfl_sync(0xffffffffu, v, src); }
// reciprocal to ~1e-15 relative: hardware seed (MUFU.RCP64H, ~20 bits) + two Newton steps; the operand is a pivot of an
// SPD system (normal range), non-finite results are caught by the isfinite test of the increment
__device__ __forceinline__ double fast_rcp(double x)
{
	double r;
	asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
	r = fma(fma(-x, r, 1.0), r, r);
	r = fma(fma(-x, r, 1.0), r, r);
	return r;
}
__device__ __forceinline__ float scale_of(int i)
{
	return i < 3 ? SCALE_XI_ROT : (i < 6 ? SCALE_XI_TRANS : (i == 6 ? SCALE_A : SCALE_B));
}
// c0 + c1 t + ... + c7 t^7 by Estrin's scheme: dependent depth 3 FMAs after t^2, t^4 instead of 7
__device__ __forceinline__ double estrin8(double t, double t2, double t4, double c0, double c1, double c2, double c3,
                                          double c4, double c5, double c6, double c7)
{
	const double p01 = fma(c1, t, c0), p23 = fma(c3, t, c2), p45 = fma(c5, t, c4), p67 = fma(c7, t, c6);
	const double p03 = fma(p23, t2, p01), p47 = fma(p67, t2, p45);
	return fma(p47, t4, p03);
}
// SE3 exponential coefficients A = sin(t)/t, B = (1-cos t)/t^2, C = (t - sin t)/t^3 as series in t^2 = th2, valid for
// th2 < 0.25 (truncation error < 1e-16; any LM increment is far below); branch-free so that the three chains interleave
__device__ __forceinline__ void exp_coeffs_series(double th2, double& A, double& B, double& C)
{
	const double t2 = th2 * th2, t4 = t2 * t2;
	A = estrin8(th2, t2, t4, 1.0, -1.0 / 6, 1.0 / 120, -1.0 / 5040, 1.0 / 362880, -1.0 / 39916800, 1.0 / 6227020800.0, -1.0 / 1307674368000.0);
	B = estrin8(th2, t2, t4, 0.5, -1.0 / 24, 1.0 / 720, -1.0 / 40320, 1.0 / 3628800, -1.0 / 479001600, 1.0 / 87178291200.0, -1.0 / 20922789888000.0);
	C = estrin8(th2, t2, t4, 1.0 / 6, -1.0 / 120, 1.0 / 5040, -1.0 / 362880, 1.0 / 39916800, -1.0 / 6227020800.0, 1.0 / 1307674368000.0, -1.0 / 355687428096000.0);
}
// exp(x) for the relative affine gain, |x| < 0.125 (it is ~1e-2 in tracking): degree-11 Taylor series (error < 1e-19)
__device__ __forceinline__ double exp_series(double x)
{
	const double x2 = x * x, x4 = x2 * x2, x8 = x4 * x4;
	const double lo = estrin8(x, x2, x4, 1.0, 1.0, 0.5, 1.0 / 6, 1.0 / 24, 1.0 / 120, 1.0 / 720, 1.0 / 5040);
	const double p89 = fma(1.0 / 362880, x, 1.0 / 40320), p1011 = fma(1.0 / 39916800, x, 1.0 / 3628800);
	return fma(fma(p1011, x2, p89), x8, lo);
}
// AffLight::fromToVecExposure (NumType.h:176-189) for the tracker's (reference, new frame) pair, general (slow) form
__device__ __forceinline__ void aff_ref_to_new(const TrackLevels& TL, double a_req, double b_req, double& a, double& b)
{
	double bb;
	aff_from_to(TL.ref_exposure, TL.new_exposure, TL.ref_a, TL.ref_b, a_req, b_req, a, bb);
	b = b_req - a * TL.ref_b;
}

// CoarseTracker.cpp:375-378 + :385 for the state to evaluate next, written into the shared EvalParams by warp 0 with one
// store per lane: lanes 0..8 one entry of R*K^-1 each, lanes 9..11 the translation, lanes 12/13 the affine pair; the level-
// and cut-off-dependent scalars only when they change.  pe = this lane's entry (R row-major 0..8, t 9..11) of the pose.
__device__ __forceinline__ void lm_write_params(EvalParams& P, const TrackLevels& TL, const dso_b200_track_opts& opt, int lane, int lvl,
                                                double pe, double affA, double affB, float rep, bool level_changed, bool cutoff_changed)
{
	static_assert(offsetof(EvalParams, t) == 9 * sizeof(float) && offsetof(EvalParams, affLL0) == 21 * sizeof(float) &&
	              offsetof(EvalParams, affLL1) == 22 * sizeof(float), "EvalParams layout");
	const float pef = (float)pe;
	const int r = (lane < 9) ? lane / 3 : 0, c = (lane < 9) ? lane - 3 * r : 0;
	const float k0 = TL.Ki[lvl][c], k1 = TL.Ki[lvl][3 + c], k2 = TL.Ki[lvl][6 + c];
	const float R0 = __shfl_sync(0xffffffffu, pef, 3 * r), R1 = __shfl_sync(0xffffffffu, pef, 3 * r + 1), R2 = __shfl_sync(0xffffffffu, pef, 3 * r + 2);
	const float rk = dot3e(R0, k0, R1, k1, R2, k2);
	const float val = lane < 9 ? rk : (lane < 12 ? pef : (lane == 12 ? (float)affA : (float)affB));
	if (lane < 14) reinterpret_cast<float*>(&P)[lane < 12 ? lane : lane + 9] = val;
	if (level_changed || cutoff_changed)
	{
		if (lane < 9 && level_changed) P.Ki[lane] = TL.Ki[lvl][lane];
		if (lane == 13)
		{
			const float cutoff = opt.coarseCutoffTH * rep;
			P.cutoffTH = cutoff;
			P.maxEnergy = fsub(fmul(fmul(2.0f, opt.huberTH), cutoff), fmul(opt.huberTH, opt.huberTH));
			if (level_changed)
			{
				P.fx = TL.fx[lvl]; P.fy = TL.fy[lvl]; P.cx = TL.cx[lvl]; P.cy = TL.cy[lvl];
				P.wl = TL.w[lvl];
				P.wl_m3 = (float)(TL.w[lvl] - 3); P.hl_m3 = (float)(TL.h[lvl] - 3);
				P.n = TL.pc_n[lvl];
				P.lvl0 = (lvl == 0);
				P.huberTH = opt.huberTH;
				P.b0 = (float)TL.ref_b;
			}
		}
	}
}

// Rare general cases of the LM step, redone with the plain formulas and kept out of line (the persistent kernel's
// instruction footprint matters): rotation beyond the range of the series, exposure ratio != 1, affine gain far from 1.
struct LMGeneral { double pe, affA, affB; };
__device__ __noinline__ LMGeneral lm_general_case(const LMShared* SH, float ref_exposure, float new_exposure, double ref_a, double ref_b,
                                                  bool propose, int lane, double a_req, double b_req, double pe_in)
{
	LMGeneral o;
	o.pe = pe_in;
	if (propose)
	{
		double x[8];
#pragma unroll
		for (int k = 0; k < 8; k++) x[k] = SH->x[k];
		Pose e, cur, nw;
		se3_exp(x, e);
#pragma unroll
		for (int k = 0; k < 9; k++) cur.R[k] = SH->cur[k];
#pragma unroll
		for (int k = 0; k < 3; k++) cur.t[k] = SH->cur[9 + k];
		pose_mul(e, cur, nw);
#pragma unroll
		for (int k = 0; k < 9; k++) if (lane == k) o.pe = nw.R[k];
#pragma unroll
		for (int k = 0; k < 3; k++) if (lane == 9 + k) o.pe = nw.t[k];
	}
	double bb;
	aff_from_to(ref_exposure, new_exposure, ref_a, ref_b, a_req, b_req, o.affA, bb);
	o.affB = b_req - o.affA * ref_b;
	return o;
}

// One step of the trackNewestCoarse state machine (CoarseTracker.cpp:584-722), executed by warp 0 of every CTA of the
// cluster on identical inputs (so all CTAs take identical decisions, no broadcast).  Consumes the sums of the evaluation
// just finished (s_tot, shared memory), leaves either R.done or the next EvalParams in P.
//
// This is the serial section of the kernel (one warp, in-order issue), so it is written as straight-line code with few
// instructions and short dependent chains: the new system entries are formed speculatively while the accept test is
// evaluated; the elimination works on 24 lanes x 3 entries with the pivots rescaled by powers of two estimated beforehand
// (nothing but shuffle -> multiply -> FMA per pivot on the chain); the series of the exponentials are evaluated by
// Estrin's scheme; every lane picks "its" rows of the rotation by index instead of by selects; and the rare general cases
// (large rotation, exposure ratio != 1, non-finite increment) are redone in a branch at the end.  Everything that is not
// needed to start the next evaluation is either a fire-and-forget store (trace, margins) or postponed to the end of the
// kernel (the per-pass RMSE with its FP64 division and square root, the flow indicators, the counters).
template <int V = 0>
__device__ __forceinline__ void lm_step_warp(LMR& R, const LMC& K, LMShared& SH, const float* s_tot, const TrackLevels& TL,
                                             const dso_b200_track_opts& opt, dso_b200_track_result* out, float* s_passE, float* s_passN,
                                             EvalParams& P, int* sCtl, int n_points, int lane, long long* lmt)
{
	long long q0 = lmt ? clock64() : 0;
	const int i = K.i;
	// ---- all loads of this step up front ----
	const float newE = s_tot[S_E], newN = s_tot[S_NTERMS], newSatN = s_tot[S_NSAT];
	const float nFT = s_tot[S_FLOW_T], nFRT = s_tot[S_FLOW_RT], nFN = s_tot[S_FLOW_N];
	const float h0 = s_tot[K.off[0]], h1 = s_tot[K.off[1]], h2 = s_tot[K.off[2]];
	R.n_evals++; R.point_evals += (double)n_points;
	const float rep_in = R.rep;
	const int lvl_in = R.lvl;
	// speculative entries of the new system (:338-353).  The common factor 1/n of H and b is left out: the increment
	// (H + lambda diag H)^-1 b does not depend on it and nothing else reads H or b.
	const double n0 = (double)h0 * K.scl[0], n1 = (double)h1 * K.scl[1], n2 = (double)h2 * K.scl[2];

	// ---- control ----
	bool take, level_done, propose;
	if (R.phase == 0)
	{
		const double newSat = (double)fdiv(newSatN, newN);                         // :517
		const bool repeat = newSat > 0.6 && R.rep < 50;                           // :586-593: same state, larger cut-off
		if (repeat) R.rep *= 2;
		take = !repeat;
		if (take) { R.lambda = 0.01f; R.iteration = 0; R.phase = 1; }
		level_done = take && (0 >= opt.maxIterations[R.lvl]);
		propose = take && !level_done;
	}
	else
	{
		// :669  (E'/n' < E/n) evaluated as E'*n < E*n' (n, n' >= 0; equal up to a 1-ulp tie of the two quotients,
		// and false like the reference's NaN comparison when a count is 0) -- keeps two FP64 divisions off the chain
		const double acc_l = (double)newE * (double)R.oldN, acc_r = (double)R.oldE * (double)newN;
		const bool accept = acc_l < acc_r;
		if (lane == 0)
		{
			if (R.trace_len < DSO_B200_TRACE_MARGINS) { out->trace_dE[R.trace_len] = (float)(acc_l - acc_r); out->trace_E[R.trace_len] = (float)acc_r; }
			if (R.trace_len < 256) out->trace[R.trace_len] = accept ? 1 : 0;
		}
		if (R.trace_len < 256) R.trace_len++;
		R.n_iter++;
		take = accept;
		if (accept)
		{
			R.a_cur = R.a_new; R.b_cur = R.b_new; R.pe_cur = R.pe_new;
			if (lane < 12) SH.cur[lane] = R.pe_new;
		}
		const float l4 = fmaxf(R.lambda * 4, 0.001f);
		R.lambda = accept ? R.lambda * 0.5f : l4;
		R.iteration++;
		level_done = (!(R.incnorm2 > 1e-6) || R.iteration >= opt.maxIterations[R.lvl]);   // :700 (|inc| > 1e-3 on the squared norm), :613
		propose = !level_done;
	}
	if (lmt) { long long q = clock64(); lmt[5] += q - q0; q0 = q; }
	R.m[0] = take ? n0 : R.m[0]; R.m[1] = take ? n1 : R.m[1]; R.m[2] = take ? n2 : R.m[2];
	R.oldE = take ? newE : R.oldE; R.oldN = take ? newN : R.oldN;
	R.oldFT = take ? nFT : R.oldFT; R.oldFRT = take ? nFRT : R.oldFRT; R.oldFN = take ? nFN : R.oldFN;   // divided when read (:514-516)
	if (lmt) { long long q = clock64(); lmt[4] += q - q0; q0 = q; }

	if (level_done)
	{
		// :708-722; the RMSE of this pass is computed from (E, n) at the end of the kernel
		if (lane == 0 && R.n_passes < DSO_B200_MAX_PASSES) { out->pass_level[R.n_passes] = R.lvl; s_passE[R.n_passes] = R.oldE; s_passN[R.n_passes] = R.oldN; }
		if (R.n_passes < DSO_B200_MAX_PASSES) R.n_passes++;
		int next = R.lvl - 1;
		if (R.rep > 1 && !R.haveRepeated) { next = R.lvl; R.haveRepeated = 1; }
		R.done = next < 0 ? 1 : 0;
		R.lvl = next; R.phase = 0; R.rep = 1; R.iteration = 0;                    // :584 of the next level pass
	}
	if (lmt) { long long q = clock64(); lmt[0] += q - q0; q0 = q; }

	const bool unit_exposure = (TL.ref_exposure == 1.f && TL.new_exposure == 1.f);
	double a_req = R.a_cur, b_req = R.b_cur;
	double pe = R.pe_cur;                          // this lane's pose entry of the state to evaluate next
	bool redo = !unit_exposure;                    // the general (slow) forms are needed
	bool bad_inc = false;
	if (propose)
	{
		// ---- :613-665: damped solve.  H + lambda*diag(H) is SPD, so no pivoting is needed (Eigen's pivoted LDLT gives the
		// same solution up to rounding).  Division-free Gauss-Jordan on [H | -b]: step k is, for every row i != k and every
		// column, M_i <- (piv*sc_k)*M_i - (M_ik*sc_k)*M_k, with sc_k a power of two (exact) close to 1/H_kk taken from
		// the diagonal BEFORE the elimination (pivots of an SPD matrix only shrink, by less than the conditioning of the
		// damped system), so that the magnitudes stay within a few hundred binades of where they started.  The pivot row
		// itself takes the same code path with the multipliers (1/sc_k, 0), which leaves it unchanged exactly.  Row
		// scaling does not change the solution; the reciprocal of the diagonal is taken once at the end. ----
		const float extrapFac = (R.lambda < 0.001f) ? sqrtf(sqrtf(0.001f / R.lambda)) : 1.f;      // :650
		const double damp = (double)(1.f + R.lambda);
		double m0 = K.diag[0] ? R.m[0] * damp : R.m[0];
		double m1 = K.diag[1] ? R.m[1] * damp : R.m[1];
		double m2 = K.diag[2] ? R.m[2] * damp : R.m[2];
		// :621-643 fixed-affine variants: the reference solves the 6x6 / 7x7 system without the fixed parameter's row and
		// column; replacing that row and column by the identity's (and its right-hand side by 0) yields the same
		// increments for the free parameters and exactly 0 for the fixed one.
		const bool fixA = opt.affineOptModeA < 0, fixB = opt.affineOptModeB < 0;
		if (fixA || fixB)
		{
			const bool fi = (i == 6 && fixA) || (i == 7 && fixB);
			const int c0 = K.g, c1 = K.g + 3, c2 = K.g + 6;
			if (fi || (c0 == 6 && fixA) || (c0 == 7 && fixB)) m0 = K.diag[0] ? 1.0 : 0.0;
			if (fi || (c1 == 6 && fixA) || (c1 == 7 && fixB)) m1 = K.diag[1] ? 1.0 : 0.0;
			if (fi || (c2 == 6 && fixA) || (c2 == 7 && fixB)) m2 = K.diag[2] ? 1.0 : 0.0;
		}
		// power-of-two scales from the diagonal: H_ii sits in lane i + 8*(i%3), entry i/3
		const int i3 = i / 3, im3 = i - 3 * i3;
		const double cand = i3 == 0 ? m0 : (i3 == 1 ? m1 : m2);
		const int myexp = (__double2hiint(cand) >> 20) & 0x7ff;
#pragma unroll
		for (int k = 0; k < ((V & 32) ? 0 : 8); k++)
		{
			const int ek = __shfl_sync(0xffffffffu, myexp, k + 8 * (k % 3));
			const double sc = __hiloint2double((2046 - ek) << 20, 0), isc = __hiloint2double(ek << 20, 0);
			const bool self = (i == k);
			// off the chain: everything that does not need the pivot row
			const double mk = (k / 3 == 0) ? m0 : ((k / 3 == 1) ? m1 : m2);       // compile-time choice
			const double fr = shfl_d(mk, i + 8 * (k % 3));                        // M_ik
			const double f = self ? 0.0 : fr * sc;
			const double t0 = m0 * sc, t1 = m1 * sc, t2 = m2 * sc;
			// on the chain: the pivot row arrives, one multiply, one FMA
			const double pivr = shfl_d(mk, k + 8 * (k % 3));
			const double p0 = shfl_d(m0, k + 8 * K.g), p1 = shfl_d(m1, k + 8 * K.g), p2 = shfl_d(m2, k + 8 * K.g);
			const double piv = self ? isc : pivr;
			m0 = fma(-f, p0, piv * t0);
			m1 = fma(-f, p1, piv * t1);
			m2 = fma(-f, p2, piv * t2);
		}
		// row i is now (0,..,d_i,..,0 | rhs_i): x_i = rhs_i / d_i
		const double dcand = i3 == 0 ? m0 : (i3 == 1 ? m1 : m2);
		const double di = shfl_d(dcand, i + 8 * im3);
		const double ri = shfl_d(m2, i + 16);
		const double xi = ri * fast_rcp(di) * (double)extrapFac;             // entry i of the (extrapolated) increment, in every group
		bad_inc = __any_sync(0xffffffffu, !isfinite(xi));                     // :659 (a non-finite entry makes the sum non-finite)
		const double isi = bad_inc ? 0.0 : xi * K.si;                         // :652-659: scaled, zero if not finite
		if (lane < 8) SH.x[lane] = isi;
		double nn = xi * xi;                                                  // |inc|^2 for :700 (NaN ends the level, as there)
		nn += shfl_d(nn, lane ^ 1); nn += shfl_d(nn, lane ^ 2); nn += shfl_d(nn, lane ^ 4);
		R.incnorm2 = nn;
		__syncwarp();
		if (lmt) { long long q = clock64(); lmt[1] += q - q0; q0 = q; }
		// every lane reads the entries it needs by index: the canonical rotation vector (for theta^2), "its" cyclic
		// permutation of it and of the translation part, column c of the matching rows of the current pose
		if (V & 16) { pe = xi; R.pe_new = xi; } else {
		const double w3 = SH.x[3], w4 = SH.x[4], w5 = SH.x[5], x6 = SH.x[6], x7 = SH.x[7];
		const double wa = SH.x[K.xa], wb = SH.x[K.xb], wc = SH.x[K.xc];
		const double ua = SH.x[K.xa - 3], ub = SH.x[K.xb - 3], uc = SH.x[K.xc - 3];
		const double pa = SH.cur[K.pa], pb = SH.cur[K.pb], pc = SH.cur[K.pc];
		a_req = R.a_cur + x6; b_req = R.b_cur + x7;
		// exp(inc) (Sophus SE3::exp): R = I + A W + B W^2, t = (I + B W + C W^2) u.  In the cyclic order (r, r+1, r+2) row r
		// of W is (0, -w_{r+2}, w_{r+1}) and row r of W^2 is w_r w - theta^2 e_r
		const double th2 = fma(w3, w3, fma(w4, w4, w5 * w5));
		redo = redo || !(th2 < 0.25);
		double A, B, C;
		exp_coeffs_series(th2, A, B, C);
		const double waa = fma(wa, wa, -th2), wab = wa * wb, wac = wa * wc;
		const double ea = fma(B, waa, 1.0), eb = fma(B, wab, -(A * wc)), ec = fma(B, wac, A * wb);
		const double va = fma(C, waa, 1.0), vb = fma(C, wab, -(B * wc)), vc = fma(C, wac, B * wb);
		const double et = fma(va, ua, fma(vb, ub, vc * uc));
		// refToNew_new = exp(inc) * refToNew_current (:661): entry (r,c) = sum_k e[r][k] cur[k][c]; t_r = sum_k e[r][k] t_k + et_r
		const double ne = fma(ea, pa, fma(eb, pb, fma(ec, pc, K.tl ? et : 0.0)));
		pe = ne;
		R.pe_new = ne; R.a_new = a_req; R.b_new = b_req;
		}
	}
	// affine pair of the state to evaluate (NumType.h:176-189 with unit exposures: a = exp(a_req - ref_a) exactly)
	const double da = a_req - TL.ref_a;
	redo = redo || !(fabs(da) < 0.125);
	double affA = exp_series(da);
	double affB = fma(-affA, TL.ref_b, b_req);
	if (lmt) { long long q = clock64(); lmt[2] += q - q0; q0 = q; }
	if (redo)
	{
		const LMGeneral gc = lm_general_case(&SH, TL.ref_exposure, TL.new_exposure, TL.ref_a, TL.ref_b, propose, lane, a_req, b_req, pe);
		pe = gc.pe; affA = gc.affA; affB = gc.affB;
		if (propose) R.pe_new = gc.pe;
	}
	if (V & 8) { if (lane < 12) reinterpret_cast<float*>(&P)[lane] = (float)(pe + affA + affB); }
	else if (!R.done) lm_write_params(P, TL, opt, lane, R.lvl, pe, affA, affB, R.rep, level_done, R.rep != rep_in);
	if (lane == 0) { sCtl[0] = R.done ? lvl_in : R.lvl; sCtl[1] = R.done; }
	__syncwarp();
	if (lmt) { long long q = clock64(); lmt[3] += q - q0; }
}

// ---- cluster exchange through remote shared memory + mbarrier (no cluster-wide barrier, no L1 invalidation) ----
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(void* bar, unsigned count)
{
	asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(void* bar, unsigned bytes)
{
	asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(void* bar, unsigned parity)
{
	asm volatile(
		"{\n"
		".reg .pred p;\n"
		"WAIT_LOOP:\n"
		"mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
		"@p bra WAIT_DONE;\n"
		"bra WAIT_LOOP;\n"
		"WAIT_DONE:\n"
		"}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// store one float into CTA `rank`'s copy of `local_addr` and count 4 bytes on that CTA's mbarrier
__device__ __forceinline__ void st_async_remote(const void* local_addr, const void* local_bar, unsigned rank, float v)
{
	unsigned raddr, rbar;
	asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(raddr) : "r"(smem_u32(local_addr)), "r"(rank));
	asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(rbar) : "r"(smem_u32(local_bar)), "r"(rank));
	asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b32 [%0], %1, [%2];" ::"r"(raddr), "r"(__float_as_uint(v)), "r"(rbar) : "memory");
}

// Fixed-tree warp reduction of S_COUNT (<= 64) values with recursive halving: 62 shuffles instead of
// 5 per value.  On return lane l holds the warp sums of entries 2l and 2l+1 in v[0], v[1].
__device__ __forceinline__ void warp_reduce_halving(float* v, int lane)
{
#pragma unroll
	for (int m = 64, o = 16; o >= 1; m >>= 1, o >>= 1)
	{
		const bool hi = (lane & o) != 0;
#pragma unroll
		for (int j = 0; j < m / 2; j++)
		{
			const float send = hi ? v[j] : v[j + m / 2];
			const float keep = hi ? v[j + m / 2] : v[j];
			v[j] = keep + __shfl_xor_sync(0xffffffffu, send, o);
		}
	}
}

// Totals of one evaluation when a hypothesis is spread over NC clusters (large clouds): cluster totals go through global
// memory and a monotonic arrival counter.  Executed by the LM warp of every CTA; kept out of line (rare path, and the
// persistent kernel's instruction footprint matters).  s0 / s1: this lane's two cluster totals in, grid totals out.
__device__ __noinline__ float2 mc_exchange(float s0, float s1, float* mc_part, unsigned int* mc_count, int hyp, int NC, int cidx,
                                           unsigned rank, unsigned int mc_round, int lane)
{
	const bool second = lane < S_COUNT - 32;
	float* mine = mc_part + ((size_t)(hyp * 2 + ((mc_round - 1) & 1)) * NC + cidx) * S_COUNT;
	if (rank == 0) { __stcg(mine + lane, s0); if (second) __stcg(mine + 32 + lane, s1); }
	__syncwarp();
	if (lane == 0)
	{
		if (rank == 0) { __threadfence(); atomicAdd(mc_count + hyp, 1u); }
		const unsigned int want = mc_round * (unsigned)NC;
		// the grid is launched cooperatively (all clusters co-resident or the launch is refused), so this wait always
		// ends; the time limit only turns a violated assumption into a launch failure instead of a hang
		unsigned long long t0 = 0, spins = 0;
		while (*((volatile unsigned int*)(mc_count + hyp)) < want)
		{
			if ((++spins & 0xfffu) == 0)
			{
				unsigned long long now;
				asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
				if (!t0) t0 = now;
				else if (now - t0 > 4000000000ull) __trap();
			}
		}
		__threadfence();
	}
	__syncwarp();
	const float* src = mc_part + (size_t)(hyp * 2 + ((mc_round - 1) & 1)) * NC * S_COUNT;
	s0 = 0.f; s1 = 0.f;
	for (int c0 = 0; c0 < NC; c0 += 8)                // clusters in index order, eight loads in flight
	{
		float t0v[8], t1v[8];
#pragma unroll
		for (int q = 0; q < 8; q++)
		{
			t0v[q] = (c0 + q < NC) ? __ldcg(src + (size_t)(c0 + q) * S_COUNT + lane) : 0.f;
			t1v[q] = (second && c0 + q < NC) ? __ldcg(src + (size_t)(c0 + q) * S_COUNT + 32 + lane) : 0.f;
		}
#pragma unroll
		for (int q = 0; q < 8; q++) if (c0 + q < NC) { s0 += t0v[q]; s1 += t1v[q]; }
	}
	return make_float2(s0, s1);
}

// What the LM loop postponed, at the end of the kernel: per-pass RMSE (:710), flow indicators (:711), :727-746
__device__ __noinline__ void lm_finalize(dso_b200_track_result* res, const float* s_passE, const float* s_passN, int lane, int n_passes,
                                         int n_iter, int n_evals, int trace_len, double point_evals, double pe_cur, double a_cur, double b_cur,
                                         float oldFT, float oldFRT, float oldFN, float modeA, float modeB,
                                         float ref_exposure, float new_exposure, double ref_a, double ref_b)
{
	// spread over the lanes (this runs once per frame, in front of the result's trip to the host): the passes' RMSEs by
	// one lane each, the flow indicators by lanes 1 and 2, the counters by lane 3, the acceptance test by lane 0
	static_assert(DSO_B200_MAX_PASSES <= 16, "one lane per pass below lane 16");
	const bool isPass = lane < n_passes;
	const double rmse = isPass ? (double)sqrtf((float)((double)s_passE[lane] / (double)s_passN[lane])) : 0.0;
	const int plv = isPass ? res->pass_level[lane] : -1 - lane;
	bool lastOfLevel = isPass;            // the reference's loop lets the last pass of a level win (:710)
	for (int q = 1; q < DSO_B200_MAX_PASSES; q++)
	{
		const int lq = __shfl_sync(0xffffffffu, plv, q);
		if (q > lane && q < n_passes && lq == plv) lastOfLevel = false;
	}
	if (isPass) res->pass_rmse[lane] = rmse;
	if (lane < 12) res->pose[lane] = pe_cur;
	if (lane >= 16 && lane < 21) res->lastResiduals[lane - 16] = NAN;      // :563
	__syncwarp();
	if (lastOfLevel) res->lastResiduals[plv] = rmse;
	if (lane == 1) { res->lastFlowIndicators[0] = (double)oldFT / ((double)oldFN + 0.1); res->lastFlowIndicators[1] = 0; }
	if (lane == 2) res->lastFlowIndicators[2] = (double)oldFRT / ((double)oldFN + 0.1);
	if (lane == 3)
	{
		res->reserved = -1;
		res->n_passes = n_passes; res->n_iterations = n_iter; res->n_evals = n_evals; res->trace_len = trace_len;
		res->point_evals = point_evals;
	}
	if (lane == 0)
	{
		res->aff_a = a_cur; res->aff_b = b_cur;
		int ok = 1;
		if ((modeA != 0 && (fabsf((float)a_cur) > 1.2f)) || (modeB != 0 && (fabsf((float)b_cur) > 200.f))) ok = 0;
		else
		{
			double ra, rb;
			aff_from_to(ref_exposure, new_exposure, ref_a, ref_b, a_cur, b_cur, ra, rb);
			if ((modeA == 0 && (fabsf(logf((float)ra)) > 1.5f)) || (modeB == 0 && (fabsf((float)rb) > 200.f))) ok = 0;
		}
		res->ok = ok;
		if (ok)                                                            // :743-744
		{
			if (modeA < 0) res->aff_a = 0;
			if (modeB < 0) res->aff_b = 0;
		}
	}
}

// K2b.  PROF = true compiles in the per-phase cycle counters (dso_b200_track_phase_cycles): SM cycles spent by thread 0 of
// CTA 0 in [point evaluation, CTA reduction, cluster exchange + barrier, LM step], number of evaluations, whole kernel
// cycles and ns, and the split of the LM step.  The production instantiation carries none of it.
template <int NT, bool PROF>
__global__ void __launch_bounds__(NT, 1)
track_persistent_kernel(TrackLevels TL, dso_b200_track_opts opt, const TrackIn* __restrict__ in,
                        dso_b200_track_result* out, int coarsestLvl, int smem_pts_cap, long long* phase_cycles,
                        dso_b200_track_result* host_out, unsigned int* host_flags, unsigned int seq,
                        int NC, float* mc_part, unsigned int* mc_count, const __grid_constant__ TrackIn in0)
{
	// in == nullptr: the single hypothesis travels in the launch parameters (in0), which saves the kernel's PCIe read
	// of the mapped input block.
	// NC > 1: one hypothesis is spread over NC clusters (large clouds); every CTA of the grid must be resident
	// (cooperative launch, see enqueue_track)
	cg::cluster_group cluster = cg::this_cluster();
	const unsigned CS = cluster.num_blocks();
	const unsigned rank = cluster.block_rank();
	const int hyp = blockIdx.x / (CS * NC);
	const int cidx = (blockIdx.x / CS) % NC;
	unsigned int mc_round = 0;
	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	long long tph[5] = { 0, 0, 0, 0, 0 };
	long long tc0 = 0, tk0 = 0, tg0 = 0;
	long long lmt[6] = { 0, 0, 0, 0, 0, 0 };   // LM step split: level, solve, exp+pose, params, sums->H, control
	long long t_pro = 0, t_stage = 0, t_loop_end = 0, t_x[5] = { 0, 0, 0, 0, 0 };   // outside the evaluations: prologue, cloud staging, epilogue
	if (PROF) { tk0 = clock64(); asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tg0)); }

	__shared__ EvalParams sP;
	__shared__ int sCtl[2];                               // level of the next evaluation, done flag
	__shared__ float s_w[NT / 32][64];
	__shared__ float xchg[2][MAX_CLUSTER][S_COUNT];
	__shared__ __align__(8) unsigned long long xbar[2];   // one mbarrier per exchange buffer
	__shared__ float s_tot[64];
	__shared__ float s_passE[DSO_B200_MAX_PASSES], s_passN[DSO_B200_MAX_PASSES];
	__shared__ LMShared s_lm;
	__shared__ int s_lvl_per[5], s_lvl_base[5];
	__shared__ dso_b200_track_result s_res;
	extern __shared__ float4 s_pts[];

	for (int i = tid; i < (int)(sizeof(dso_b200_track_result) / 4); i += NT) reinterpret_cast<unsigned int*>(&s_res)[i] = 0u;
	LMR R;
	LMC K;
	if (warp == 0)
	{
		lm_constants(lane, K);
		// the input block may live in mapped host memory: one coalesced read by 19 lanes; the pose entries stay in the
		// lanes that read them
		static_assert(sizeof(TrackIn) == 19 * sizeof(double), "TrackIn layout");
		const double vin = (lane < 19) ? (in ? reinterpret_cast<const double*>(in + hyp)[lane] : reinterpret_cast<const double*>(&in0)[lane]) : 0.0;
		R.pe_cur = R.pe_new = vin;
		if (lane < 12) s_lm.cur[lane] = vin;
		if (lane < 8) s_lm.x[lane] = 0.0;
		R.a_cur = R.a_new = shfl_d(vin, 12); R.b_cur = R.b_new = shfl_d(vin, 13);
		R.m[0] = R.m[1] = R.m[2] = 0.0;
		R.incnorm2 = 0.0;
		R.oldE = 0.f; R.oldN = 1.f; R.oldFT = 0.f; R.oldFRT = 0.f; R.oldFN = 0.f;
		R.lambda = 0.01f; R.rep = 1.f;
		R.lvl = coarsestLvl; R.iteration = 0; R.phase = 0; R.haveRepeated = 0; R.done = 0;
		R.n_evals = 0; R.n_iter = 0; R.trace_len = 0; R.n_passes = 0; R.point_evals = 0.0;
		// CoarseTracker.cpp:375-385 for the very first evaluation (:584)
		double affA, affB;
		aff_ref_to_new(TL, R.a_cur, R.b_cur, affA, affB);
		lm_write_params(sP, TL, opt, lane, coarsestLvl, R.pe_cur, affA, affB, 1.f, true, true);
		if (lane == 0) { sCtl[0] = coarsestLvl; sCtl[1] = 0; }
		if (PROF) t_x[0] = clock64() - tk0;
	}
	if (tid == 0) { mbar_init(&xbar[0], 1); mbar_init(&xbar[1], 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
	// this CTA's slice of the cloud of EVERY level goes into shared memory now, by the warps that have nothing to do with
	// the LM set-up: one round of (possibly cold) memory latency for all levels instead of one per level change
	int tot_slots = 0;
	int perL[5];                                           // one run-time division per level, not one per use
#pragma unroll
	for (int l = 0; l < 5; l++)
	{
		perL[l] = 0;
		if (l <= coarsestLvl)
		{
			perL[l] = (TL.pc_n[l] + CS * NC - 1) / (CS * NC);
			if (tid == 32) { s_lvl_per[l] = perL[l]; s_lvl_base[l] = tot_slots; }
			tot_slots += perL[l];
		}
	}
	const bool all_staged = tot_slots <= smem_pts_cap;
	if (all_staged && NT > 32 && warp > 0)
	{
		const int NS = NT - 32, t = tid - 32;
		for (int j0 = t; j0 < tot_slots; j0 += 4 * NS)
		{
			float4 v[4];
			bool have[4];
#pragma unroll
			for (int q = 0; q < 4; q++)
			{
				const int j = j0 + q * NS;
				have[q] = false;
				int off = 0;
#pragma unroll
				for (int l = 0; l < 5; l++)
					if (l <= coarsestLvl)
					{
						const int nl = TL.pc_n[l], per = perL[l];
						const int b = min((int)(cidx * CS + rank) * per, nl), e = min(b + per, nl);
						if (j >= off && j - off < e - b) { v[q] = __ldg(TL.pc[l] + b + (j - off)); have[q] = true; }
						off += per;
					}
			}
#pragma unroll
			for (int q = 0; q < 4; q++)
				if (have[q]) s_pts[j0 + q * NS] = v[q];
		}
	}
	// the new frame's pyramid may be cold (it has just come from the host, or other work has gone through the L2 since
	// makeImages wrote it): ask for it now, coarsest level first, behind the cloud loads above.  Fire and forget; the
	// levels are walked minutes (in GPU terms) later.  Only the levels this call uses.
	// (cp.async.bulk.prefetch.L2 in 16 KB pieces was tried instead: same first-evaluation times, kernel 3 us longer)
	// Only the two coarsest levels are requested here; the finer ones follow in slices, issued by the warps that idle
	// during the LM steps of the first evaluations (below), so that the requests do not queue in front of the first
	// evaluation's own loads (3 200 line requests per CTA are ~3 k cycles of load/store-unit time).
	if (warp > 0)
	{
		const int G = (int)(CS * NC) * (NT - 32), g = (int)(cidx * CS + rank) * (NT - 32) + (tid - 32);
		for (int l = coarsestLvl; l >= 0 && l >= coarsestLvl - 1; l--)
		{
			const char* base = reinterpret_cast<const char*>(TL.img[l]);
			const int bytes = TL.w[l] * TL.h[l] * (int)sizeof(float4);
			if (bytes > (8 << 20)) continue;                      // a large level is sampled sparsely: fetching all of it would cost more than it saves
			const int lines = (bytes + 127) >> 7;
			for (int i = g; i < lines; i += G) asm volatile("prefetch.global.L2 [%0];" :: "l"(base + ((size_t)i << 7)));
		}
	}
	__syncthreads();
	if (PROF) t_x[1] = clock64() - tk0;
	// every CTA's barriers must exist before anybody signals them: arrive here, wait in front of the first push, so the
	// first evaluation hides the cluster's launch skew and the barrier's own latency
	asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
	bool peers_ready = false;
	if (PROF) t_pro = clock64() - tk0;
	int buf = 0;
	unsigned xphase[2] = { 0u, 0u };
	int staged_lvl = -1;
	constexpr int PF_PER_EVAL = 4;
	int pf_lvl = coarsestLvl - 2, pf_round = 0;            // prefetch cursor of the idle warps: level, round of G lines
	while (true)
	{
		const int lvl = sCtl[0];
		const int n = TL.pc_n[lvl];
		// slice of this CTA and its offset in the staged cloud: tabulated in the prologue (run-time integer divisions are
		// ~100 cycles each, and this is the top of every evaluation)
		const int per = s_lvl_per[lvl];
		const int beg = min((int)(cidx * CS + rank) * per, n), end = min(beg + per, n);
		const float4* __restrict__ gp = TL.pc[lvl];
		const bool use_smem = all_staged || (end - beg) <= smem_pts_cap;
		const int s_base = all_staged ? s_lvl_base[lvl] : 0;
		if (!all_staged && use_smem && staged_lvl != lvl)
		{
			long long ts0 = 0;
			if (PROF) ts0 = clock64();
			for (int i = beg + tid; i < end; i += NT) s_pts[i - beg] = __ldg(gp + i);
			staged_lvl = lvl;
			__syncthreads();
			if (PROF) t_stage += clock64() - ts0;
		}
		const EvalParams P = sP;
		const float4* __restrict__ img = TL.img[lvl];
		if (PROF) { __syncthreads(); tc0 = clock64(); }
		float acc[64];
#pragma unroll
		for (int k = 0; k < 64; k++) acc[k] = 0.f;
		for (int i = beg + tid; i < end; i += NT)      // one copy of the evaluation code (instruction footprint)
		{
			float4 pc;
			if (use_smem) pc = s_pts[s_base + i - beg]; else pc = __ldg(gp + i);
			eval_point<false>(pc, i, P, img, acc, nullptr, nullptr);
		}
		if (PROF) { __syncthreads(); long long t = clock64(); tph[0] += t - tc0; if (tid == 0 && blockIdx.x == 0 && tph[4] < 24) phase_cycles[24 + 4 * tph[4]] = t - tc0; tc0 = t; }
		warp_reduce_halving(acc, lane);
		s_w[warp][2 * lane] = acc[0];
		s_w[warp][2 * lane + 1] = acc[1];
		__syncthreads();
		if (!peers_ready) { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); peers_ready = true; }
		// CTA partial (warps in index order), pushed into every CTA of the cluster through DSMEM
		if (tid < S_COUNT)
		{
			float s = 0.f;
#pragma unroll
			for (int w = 0; w < NT / 32; w++) s += s_w[w][tid];
			for (unsigned r = 0; r < CS; r++) st_async_remote(&xchg[buf][rank][tid], &xbar[buf], r, s);
		}
		if (tid == 0) mbar_arrive_expect_tx(&xbar[buf], CS * S_COUNT * 4);   // this CTA expects CS partials of S_COUNT floats
		if (PROF) { long long t = clock64(); tph[1] += t - tc0; if (tid == 0 && blockIdx.x == 0 && tph[4] < 24) phase_cycles[25 + 4 * tph[4]] = t - tc0; tc0 = t; }
		// only the LM warp needs the totals: the first 52 threads (two warps) add the ranks in index order and hand them
		// over through a two-warp named barrier; everybody else goes straight to the barrier at the end of the iteration
		if (warp < 2)
		{
			mbar_wait(&xbar[buf], xphase[buf]);
			if (tid < S_COUNT)
			{
				float v[MAX_CLUSTER];                  // loads issued together (fixed trip count), adds in order
#pragma unroll
				for (int r = 0; r < MAX_CLUSTER; r++) v[r] = (r < (int)CS) ? xchg[buf][r][tid] : 0.f;
				float sacc = 0.f;
#pragma unroll
				for (int r = 0; r < MAX_CLUSTER; r++) sacc += v[r];
				s_tot[tid] = sacc;
			}
			asm volatile("bar.sync 1, 64;" ::: "memory");
		}
		if (warp == 0)
		{
			if (NC > 1)
			{
				mc_round++;
				const float2 g = mc_exchange(s_tot[lane], s_tot[32 + lane], mc_part, mc_count, hyp, NC, cidx, rank, mc_round, lane);
				__syncwarp();
				s_tot[lane] = g.x; s_tot[32 + lane] = g.y;
				__syncwarp();
			}
			if (PROF) { long long t = clock64(); tph[2] += t - tc0; if (tid == 0 && blockIdx.x == 0 && tph[4] < 24) phase_cycles[26 + 4 * tph[4]] = t - tc0; tc0 = t; }
			lm_step_warp(R, K, s_lm, s_tot, TL, opt, &s_res, s_passE, s_passN, sP, sCtl, n, lane, PROF ? lmt : nullptr);
			if (PROF) { long long t = clock64(); tph[3] += t - tc0; if (tid == 0 && blockIdx.x == 0 && tph[4] < 24) phase_cycles[27 + 4 * tph[4]] = t - tc0; tph[4]++; }
		}
		if (NT > 64 && warp >= 2 && pf_lvl >= 0)
		{
			// idle until the LM warp is done: the next slice of the finer levels' prefetch (PF_PER_EVAL lines per thread)
			const int G = (int)(CS * NC) * (NT - 64), g = (int)(cidx * CS + rank) * (NT - 64) + (tid - 64);
#pragma unroll
			for (int q = 0; q < PF_PER_EVAL; q++)
			{
				if (pf_lvl < 0) break;
				const int bytes = TL.w[pf_lvl] * TL.h[pf_lvl] * (int)sizeof(float4);
				const int lines = bytes > (8 << 20) ? 0 : (bytes + 127) >> 7;
				const int i = pf_round * G + g;
				if (i < lines) asm volatile("prefetch.global.L2 [%0];" :: "l"(reinterpret_cast<const char*>(TL.img[pf_lvl]) + ((size_t)i << 7)));
				pf_round++;
				if (pf_round * G >= lines) { pf_lvl--; pf_round = 0; }        // uniform over the cluster: G and lines are
			}
		}
		xphase[buf] ^= 1u;
		__syncthreads();           // the next EvalParams / level / done flag are in shared memory
		if (sCtl[1]) break;
		buf ^= 1;
	}
	if (PROF) t_loop_end = clock64();
	// nobody leaves while a peer may still be writing into it: arrive now, wait after the result has been written (the
	// peers only wait for this CTA's arrival, and this CTA finds them all arrived when it gets there)
	asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
	if (rank == 0 && cidx == 0)
	{
		if (warp == 0)
			lm_finalize(&s_res, s_passE, s_passN, lane, R.n_passes, R.n_iter, R.n_evals, R.trace_len, R.point_evals, R.pe_cur, R.a_cur, R.b_cur,
			            R.oldFT, R.oldFRT, R.oldFN, opt.affineOptModeA, opt.affineOptModeB, TL.ref_exposure, TL.new_exposure, TL.ref_a, TL.ref_b);
		__syncthreads();
		if (PROF) t_x[2] = clock64() - t_loop_end;
		static_assert(sizeof(dso_b200_track_result) % 8 == 0, "result copied in 8-byte words");
		const int nwords = sizeof(dso_b200_track_result) / 8;
		const uint2* src = reinterpret_cast<const uint2*>(&s_res);
		if (host_out && warp == 0)
		{
			// results straight into mapped host memory by ONE warp, then the per-hypothesis flag: the host polls it instead
			// of waiting for a copy node and a stream synchronisation.  One warp = one system-scope fence to wait for.
			uint2* hdst = reinterpret_cast<uint2*>(host_out + hyp);
			for (int i = lane; i < nwords; i += 32) hdst[i] = src[i];
			if (PROF) t_x[3] = clock64() - t_loop_end;
			__threadfence_system();
			__syncwarp();
			if (PROF) t_x[4] = clock64() - t_loop_end;
			// no fence behind the flag: the end of the kernel publishes it at the latest, a fence would not make it earlier
			if (lane == 0) *reinterpret_cast<volatile unsigned int*>(host_flags + hyp) = seq;
		}
		else if (!host_out || warp > 0)
		{
			const int t0 = host_out ? tid - 32 : tid, nthr = host_out ? NT - 32 : NT;
			uint2* dst = reinterpret_cast<uint2*>(out + hyp);
			for (int i = t0; i < nwords; i += nthr) dst[i] = src[i];
		}
	}
	asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
	if (PROF && blockIdx.x == 0 && tid == 0)
	{
		for (int k = 0; k < 5; k++) phase_cycles[k] = tph[k];
		long long tg1; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tg1));
		phase_cycles[5] = clock64() - tk0;
		phase_cycles[6] = tg1 - tg0;
		for (int k = 0; k < 6; k++) phase_cycles[8 + k] = lmt[k];
		for (int k = 0; k < 5; k++) phase_cycles[16 + k] = t_x[k];
		phase_cycles[7] = t_pro; phase_cycles[14] = t_stage; phase_cycles[15] = clock64() - t_loop_end;
	}
}

static void mat33f_inverse(const float* m, float* out);
// Timing harness for the LM step alone (debug): one warp runs lm_step_warp `iters` times on a fixed, well-conditioned
// synthetic evaluation whose energy shrinks a little every time, so that every step is an accepted one with a new proposal.
template <int V>
__global__ void lm_step_bench_kernel(TrackLevels TL, dso_b200_track_opts opt, int iters, long long* cycles, double* sink)
{
	__shared__ EvalParams sP;
	__shared__ int sCtl[2];
	__shared__ float s_tot[64];
	__shared__ float s_passE[DSO_B200_MAX_PASSES], s_passN[DSO_B200_MAX_PASSES];
	__shared__ dso_b200_track_result s_res;
	__shared__ LMShared s_lm;
	const int lane = threadIdx.x;
	LMC K;
	lm_constants(lane, K);
	for (int k = lane; k < 64; k += 32) s_tot[k] = 0.f;
	__syncwarp();
	if (lane == 0)
	{
		s_tot[S_E] = 5.0e5f; s_tot[S_NTERMS] = 4000.f; s_tot[S_NSAT] = 100.f; s_tot[S_NWARP] = 3900.f;
		int k = S_ACC;
		for (int r = 0; r < 8; r++) for (int c = r; c < 8; c++) s_tot[k++] = (r == c) ? 3.0e6f * (1 + r) : 1.0e4f * (r + c);
		for (int r = 0; r < 8; r++) s_tot[k++] = 50.f * (r + 1);
	}
	__syncwarp();
	LMR R;
	R.pe_cur = R.pe_new = (lane == 0 || lane == 4 || lane == 8) ? 1.0 : 0.0;
	if (lane < 12) s_lm.cur[lane] = R.pe_cur;
	if (lane < 8) s_lm.x[lane] = 0.0;
	R.a_cur = R.a_new = 0.0; R.b_cur = R.b_new = 0.0;
	R.m[0] = R.m[1] = R.m[2] = 0.0;
	R.incnorm2 = 1.0;
	R.oldE = 6.0e5f; R.oldN = 4000.f; R.oldFT = 0.f; R.oldFRT = 0.f; R.oldFN = 0.f;
	R.lambda = 0.01f; R.rep = 1.f;
	R.lvl = 0; R.iteration = 0; R.phase = 1; R.haveRepeated = 0; R.done = 0;
	R.n_evals = 0; R.n_iter = 0; R.trace_len = 0; R.n_passes = 0; R.point_evals = 0.0;
	opt.maxIterations[0] = 1 << 30;
	long long lmt[6] = { 0, 0, 0, 0, 0, 0 };
	long long t0 = 0;
	for (int it = 0; it < iters + 8; it++)
	{
		if (it == 8) { t0 = clock64(); for (int k = 0; k < 6; k++) lmt[k] = 0; }
		lm_step_warp<V>(R, K, s_lm, s_tot, TL, opt, &s_res, s_passE, s_passN, sP, sCtl, 1000, lane, (V & 64) ? nullptr : lmt);
		R.trace_len = 0; R.lambda = 0.01f; R.incnorm2 = 1.0; R.iteration = 0;
		if (lane == 0) s_tot[S_E] *= 0.999f;
		__syncwarp();
	}
	const long long t1 = clock64();
	if (lane == 0) { cycles[0] = t1 - t0; for (int k = 0; k < 6; k++) cycles[1 + k] = lmt[k]; }
	sink[lane] = R.pe_new + R.a_new + sP.RKi[lane % 9];
}

extern "C" int dso_b200_debug_lm_bench(dso_b200_ctx* ctx, int iters, long long* cycles7)
{
	ARG_CHECK(ctx && cycles7 && iters > 0, "bad arguments");
	CK(cudaSetDevice(ctx->device));
	TrackLevels TL;
	memset(&TL, 0, sizeof(TL));
	TL.levels = 1; TL.w[0] = 640; TL.h[0] = 480; TL.fx[0] = 500; TL.fy[0] = 500; TL.cx[0] = 320; TL.cy[0] = 240;
	const float K[9] = { 500, 0, 320, 0, 500, 240, 0, 0, 1 };
	mat33f_inverse(K, TL.Ki[0]);
	TL.ref_exposure = 1.f; TL.new_exposure = 1.f;
	dso_b200_track_opts o; dso_b200_track_opts_default(&o);
	long long* d_c = nullptr; double* d_s = nullptr;
	CK(cudaMalloc(&d_c, 64)); CK(cudaMalloc(&d_s, 32 * sizeof(double)));
	const int variant = getenv("DSO_B200_LM_VARIANT") ? atoi(getenv("DSO_B200_LM_VARIANT")) : 0;
	if (variant == 64) lm_step_bench_kernel<64><<<1, 32, 0, ctx->stream>>>(TL, o, iters, d_c, d_s);
	else if (variant == 72) lm_step_bench_kernel<72><<<1, 32, 0, ctx->stream>>>(TL, o, iters, d_c, d_s);
	else if (variant == 80) lm_step_bench_kernel<80><<<1, 32, 0, ctx->stream>>>(TL, o, iters, d_c, d_s);
	else if (variant == 96) lm_step_bench_kernel<96><<<1, 32, 0, ctx->stream>>>(TL, o, iters, d_c, d_s);
	else if (variant == 120) lm_step_bench_kernel<120><<<1, 32, 0, ctx->stream>>>(TL, o, iters, d_c, d_s);
	else lm_step_bench_kernel<0><<<1, 32, 0, ctx->stream>>>(TL, o, iters, d_c, d_s);
	CK(cudaGetLastError());
	CK(cudaStreamSynchronize(ctx->stream));
	CK(cudaMemcpy(cycles7, d_c, 56, cudaMemcpyDeviceToHost));
	cudaFree(d_c); cudaFree(d_s);
	return 0;
}

// ---- host side ------------------------------------------------------------------------------------
static void fill_levels(dso_b200_ctx* ctx, TrackerSlot& T, const FrameSlot* fs, float new_exposure, TrackLevels& TL)
{
	memset(&TL, 0, sizeof(TL));
	TL.levels = T.levels;
	for (int l = 0; l < T.levels; l++)
	{
		TL.w[l] = T.w[l]; TL.h[l] = T.h[l];
		TL.fx[l] = T.fx[l]; TL.fy[l] = T.fy[l]; TL.cx[l] = T.cx[l]; TL.cy[l] = T.cy[l];
		memcpy(TL.Ki[l], T.Ki[l], 36);
		TL.pc[l] = T.pc[l]; TL.pc_n[l] = T.pc_n[l];
		TL.img[l] = fs ? fs->pyr[l] : nullptr;
	}
	TL.ref_a = T.ref_a; TL.ref_b = T.ref_b; TL.ref_exposure = T.ref_exposure; TL.new_exposure = new_exposure;
}

// Eigen's compute_inverse<Matrix3f> (cofactors of column 0, determinant by expansion along it), as used by
// CoarseTracker::makeK (CoarseTracker.cpp:125)
static float cof33(const float* m, int i, int j)
{
	int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
	volatile float a = m[i1 * 3 + j1] * m[i2 * 3 + j2];
	volatile float b = m[i1 * 3 + j2] * m[i2 * 3 + j1];
	return a - b;
}
static void mat33f_inverse(const float* m, float* out)
{
	float c0 = cof33(m, 0, 0), c1 = cof33(m, 1, 0), c2 = cof33(m, 2, 0);
	float det = dot3h(c0, m[0], c1, m[3], c2, m[6]);
	float invdet = 1.0f / det;
	out[0] = c0 * invdet; out[1] = c1 * invdet; out[2] = c2 * invdet;
	out[3] = cof33(m, 0, 1) * invdet; out[4] = cof33(m, 1, 1) * invdet; out[5] = cof33(m, 2, 1) * invdet;
	out[6] = cof33(m, 0, 2) * invdet; out[7] = cof33(m, 1, 2) * invdet; out[8] = cof33(m, 2, 2) * invdet;
}

extern "C" void dso_b200_track_opts_default(dso_b200_track_opts* o)
{
	o->huberTH = 9; o->coarseCutoffTH = 20; o->affineOptModeA = 1e12f; o->affineOptModeB = 1e8f;
	const int mi[5] = { 10, 20, 50, 50, 50 };
	for (int i = 0; i < 5; i++) o->maxIterations[i] = mi[i];
}

// CoarseTracker::makeK, CoarseTracker.cpp:102-131
void tracker_make_K(TrackerSlot& T, int w, int h, int levels, float fx, float fy, float cx, float cy)
{
	T.levels = levels;
	T.w[0] = w; T.h[0] = h; T.fx[0] = fx; T.fy[0] = fy; T.cx[0] = cx; T.cy[0] = cy;
	for (int l = 1; l < levels; l++)
	{
		T.w[l] = w >> l; T.h[l] = h >> l;
		T.fx[l] = T.fx[l - 1] * 0.5;
		T.fy[l] = T.fy[l - 1] * 0.5;
		T.cx[l] = (T.cx[0] + 0.5) / ((int)1 << l) - 0.5;
		T.cy[l] = (T.cy[0] + 0.5) / ((int)1 << l) - 0.5;
	}
	for (int l = 0; l < levels; l++)
	{
		float K[9] = { T.fx[l], 0, T.cx[l], 0, T.fy[l], T.cy[l], 0, 0, 1 };
		mat33f_inverse(K, T.Ki[l]);
	}
}

extern "C" int dso_b200_tracker_set_ref(dso_b200_ctx* ctx, int trk, int w, int h, int levels,
                                        float fx, float fy, float cx, float cy,
                                        const int* pc_n, const float* const* pc_u, const float* const* pc_v,
                                        const float* const* pc_idepth, const float* const* pc_color,
                                        double ref_aff_a, double ref_aff_b, float ref_ab_exposure)
{
	ARG_CHECK(ctx && (trk == 0 || trk == 1), "bad tracker slot");
	ARG_CHECK(levels >= 1 && levels <= MAXL && w <= ctx->max_w && h <= ctx->max_h, "bad size / levels");
	CK(cudaSetDevice(ctx->device));
	TrackerSlot& T = ctx->trk[trk];
	tracker_make_K(T, w, h, levels, fx, fy, cx, cy);
	// clouds: SoA host arrays -> one float4 per point, packed on the host, one copy per level
	std::vector<float4> tmp;
	for (int l = 0; l < levels; l++)
	{
		int n = pc_n[l];
		ARG_CHECK(n >= 0 && (size_t)n <= T.cap_points, "pc_n exceeds capacity");
		T.pc_n[l] = n;
		tmp.resize(n);
		for (int i = 0; i < n; i++) tmp[i] = make_float4(pc_u[l][i], pc_v[l][i], pc_idepth[l][i], pc_color[l][i]);
		if (n)
		{
			// pageable source: the call returns once the data sits in the driver's staging area, so tmp can be reused
			CK(cudaMemcpyAsync(T.pc[l], tmp.data(), sizeof(float4) * n, cudaMemcpyHostToDevice, T.stream));
		}
	}
	CK(cudaStreamSynchronize(T.stream));
	T.ref_a = ref_aff_a; T.ref_b = ref_aff_b; T.ref_exposure = ref_ab_exposure;
	T.valid = true; T.last_lvl = -1; T.last_warped_n = 0;
	return 0;
}

static int check_eval_args(dso_b200_ctx* ctx, int trk, int frame_slot, int lvl)
{
	ARG_CHECK(ctx && (trk == 0 || trk == 1) && ctx->trk[trk].valid, "tracker reference not set");
	ARG_CHECK(frame_slot >= 0 && frame_slot < ctx->n_slots && ctx->slots[frame_slot].valid, "frame slot empty");
	const TrackerSlot& T = ctx->trk[trk];
	const FrameSlot& F = ctx->slots[frame_slot];
	ARG_CHECK(lvl >= 0 && lvl < T.levels && lvl < F.levels, "bad level");
	ARG_CHECK(F.w == T.w[0] && F.h == T.h[0], "frame / tracker size mismatch");
	return frame_slot_wait(ctx, frame_slot, ctx->trk[trk].stream, ctx->trk[trk].consumer);   // an asynchronous build of this slot on another stream
}

static int grid_for(const dso_b200_ctx* ctx, int n)
{
	int g = (n + EVAL_THREADS - 1) / EVAL_THREADS;
	int cap = ctx->sm_count * 4;
	if (g > cap) g = cap;
	if (g < 1) g = 1;
	return g;
}

static int run_eval(dso_b200_ctx* ctx, int trk, int frame_slot, int lvl, const Pose& pose, double a, double b,
                    float new_exposure, float cutoffTH, float huberTH, bool debug, unsigned char* d_status, float* d_warped)
{
	TrackerSlot& T = ctx->trk[trk];
	TrackLevels TL;
	fill_levels(ctx, T, &ctx->slots[frame_slot], new_exposure, TL);
	EvalParams P;
	make_eval_params(TL, lvl, pose, a, b, cutoffTH, huberTH, P);
	int g = grid_for(ctx, P.n);
	if (debug)
		track_eval_kernel<true><<<g, EVAL_THREADS, 0, T.stream>>>(P, T.pc[lvl], TL.img[lvl], T.d_partials, T.d_ticket, T.d_sums, d_status, d_warped);
	else
		track_eval_kernel<false><<<g, EVAL_THREADS, 0, T.stream>>>(P, T.pc[lvl], TL.img[lvl], T.d_partials, T.d_ticket, T.d_sums, nullptr, nullptr);
	LAUNCHED(ctx, 1);
	CK(cudaGetLastError());
	CK(cudaMemcpyAsync(T.h_sums, T.d_sums, sizeof(float) * S_COUNT, cudaMemcpyDeviceToHost, T.stream));
	CK(cudaStreamSynchronize(T.stream));
	T.last_lvl = lvl;
	return 0;
}

extern "C" int dso_b200_calc_res(dso_b200_ctx* ctx, int trk, int frame_slot, int lvl, const double* refToNew12,
                                 double aff_a, double aff_b, float new_ab_exposure, float cutoffTH, double* out6)
{
	int rc = check_eval_args(ctx, trk, frame_slot, lvl);
	if (rc) return rc;
	CK(cudaSetDevice(ctx->device));
	Pose p; pose_from12(p, refToNew12);
	rc = run_eval(ctx, trk, frame_slot, lvl, p, aff_a, aff_b, new_ab_exposure, cutoffTH, ctx->settings.huberTH, false, nullptr, nullptr);
	if (rc) return rc;
	EvalResult er;
	sums_to_result(ctx->trk[trk].h_sums, er);
	out6[0] = er.E; out6[1] = er.n; out6[2] = er.flowT; out6[3] = 0; out6[4] = er.flowRT; out6[5] = er.sat;
	return 0;
}

extern "C" int dso_b200_calc_gs(dso_b200_ctx* ctx, int trk, double* H64, double* b8)
{
	ARG_CHECK(ctx && (trk == 0 || trk == 1) && ctx->trk[trk].valid && ctx->trk[trk].last_lvl >= 0, "no preceding calc_res");
	EvalResult er;
	sums_to_result(ctx->trk[trk].h_sums, er);
	memcpy(H64, er.H, sizeof(er.H));
	memcpy(b8, er.b, sizeof(er.b));
	return 0;
}

extern "C" int dso_b200_calc_res_debug(dso_b200_ctx* ctx, int trk, int frame_slot, int lvl, const double* refToNew12,
                                       double aff_a, double aff_b, float new_ab_exposure, float cutoffTH,
                                       unsigned char* status_out, float* warped8_out)
{
	int rc = check_eval_args(ctx, trk, frame_slot, lvl);
	if (rc) return rc;
	CK(cudaSetDevice(ctx->device));
	TrackerSlot& T = ctx->trk[trk];
	int n = T.pc_n[lvl];
	unsigned char* d_status = nullptr; float* d_warped = nullptr;
	CK(cudaMalloc(&d_status, n + 16));
	CK(cudaMalloc(&d_warped, sizeof(float) * 8 * (size_t)(n + 1)));
	CK(cudaMemsetAsync(d_warped, 0, sizeof(float) * 8 * (size_t)(n + 1), T.stream));
	Pose p; pose_from12(p, refToNew12);
	rc = run_eval(ctx, trk, frame_slot, lvl, p, aff_a, aff_b, new_ab_exposure, cutoffTH, ctx->settings.huberTH, true, d_status, d_warped);
	if (!rc)
	{
		cudaMemcpy(status_out, d_status, n, cudaMemcpyDeviceToHost);
		cudaMemcpy(warped8_out, d_warped, sizeof(float) * 8 * (size_t)n, cudaMemcpyDeviceToHost);
	}
	cudaFree(d_status); cudaFree(d_warped);
	return rc;
}

extern "C" int dso_b200_track_hostloop(dso_b200_ctx* ctx, int trk, int frame_slot, const double* pose12_init,
                                       double aff_a_init, double aff_b_init, float new_ab_exposure, int coarsestLvl,
                                       const double* minResForAbort5, const dso_b200_track_opts* opts,
                                       dso_b200_track_result* out)
{
	int rc = check_eval_args(ctx, trk, frame_slot, coarsestLvl);
	if (rc) return rc;
	ARG_CHECK(coarsestLvl < 5, "coarsestLvl must be < 5 (CoarseTracker.cpp:561)");
	CK(cudaSetDevice(ctx->device));
	dso_b200_track_opts o;
	if (opts) o = *opts; else { dso_b200_track_opts_default(&o); o.huberTH = ctx->settings.huberTH; }
	LM lm;
	TrackLevels TL0;
	fill_levels(ctx, ctx->trk[trk], &ctx->slots[frame_slot], new_ab_exposure, TL0);
	lm.init(o, TL0, pose12_init, aff_a_init, aff_b_init, coarsestLvl, minResForAbort5, out);
	while (!lm.done)
	{
		rc = run_eval(ctx, trk, frame_slot, lm.req_lvl, lm.req_pose, lm.req_a, lm.req_b, new_ab_exposure, lm.req_cutoff, o.huberTH, false, nullptr, nullptr);
		if (rc) return rc;
		EvalResult er;
		sums_to_result(ctx->trk[trk].h_sums, er);
		lm.step(er, ctx->trk[trk].pc_n[lm.req_lvl]);
	}
	return 0;
}

extern "C" int dso_b200_track_replay_abort(const dso_b200_track_result* r, const double* minRes, double* lastRes)
{
	double lr[5];
	for (int i = 0; i < 5; i++) lr[i] = NAN;
	int ok = 1;
	for (int p = 0; p < r->n_passes; p++)
	{
		int l = r->pass_level[p];
		lr[l] = r->pass_rmse[p];
		if (lr[l] > 1.5 * minRes[l]) { ok = 0; break; }
	}
	if (ok) ok = r->ok;
	if (lastRes) for (int i = 0; i < 5; i++) lastRes[i] = lr[i];
	return ok;
}

// sequence numbers of tracking calls: 1, 2, ..., 2^32-1, 1, ... (0 = "no host-visible results")
static inline unsigned int next_seq(unsigned int s) { return s + 1u == 0u ? 1u : s + 1u; }
static inline unsigned int seq_distance(unsigned int from, unsigned int to)
{
	unsigned int d = 0;
	while (from != to && d <= TRACK_IN_RING) { from = next_seq(from); d++; }
	return d;
}

// Enqueue (no synchronisation): the persistent kernel reads the initial poses from and writes its results to mapped
// pinned memory.
static int enqueue_track(dso_b200_ctx* ctx, int trk, int frame_slot, int nHyp, const double* pose12_init,
                         double aff_a_init, double aff_b_init, float new_ab_exposure, int coarsestLvl,
                         const dso_b200_track_opts* opts)
{
	int rc = check_eval_args(ctx, trk, frame_slot, coarsestLvl);
	if (rc) return rc;
	ARG_CHECK(nHyp >= 1 && nHyp <= DSO_B200_MAX_HYP, "nHyp out of range");
	ARG_CHECK(coarsestLvl < 5, "coarsestLvl must be < 5 (CoarseTracker.cpp:561)");
	CK(cudaSetDevice(ctx->device));
	TrackerSlot& T = ctx->trk[trk];
	dso_b200_track_opts o;
	if (opts) o = *opts; else { dso_b200_track_opts_default(&o); o.huberTH = ctx->settings.huberTH; }
	// pageable staging: cudaMemcpyAsync returns once the source has been copied to the driver's staging
	// area, so back-to-back asynchronous calls may reuse it
	static_assert(sizeof(TrackIn) * DSO_B200_MAX_HYP <= TRACK_IN_BLOCK, "input block too small");
	const TrackIn* kin = (const TrackIn*)T.last_in;
	unsigned int seq = 0;
	if (pose12_init)
	{
		// the kernel reads its inputs from a ring of mapped pinned blocks (one PCIe read by one warp) and writes its
		// results into the matching block of the result ring: no copy nodes in front of or behind the kernel
		if (seq_distance(T.fetched, T.seq) >= TRACK_IN_RING)
		{
			dso_set_error("dso_b200_track_batched_async: %d tracking calls on this slot are waiting for dso_b200_track_fetch; "
			              "fetch before enqueueing more", TRACK_IN_RING);
			return DSO_B200_E_STATE;
		}
		seq = T.seq = next_seq(T.seq);
		TrackIn* hin = (TrackIn*)((char*)T.h_track_in + (size_t)(seq % TRACK_IN_RING) * TRACK_IN_BLOCK);
		for (int i = 0; i < nHyp; i++)
		{
			memcpy(hin[i].pose, pose12_init + 12 * i, 96);
			hin[i].aff_a = aff_a_init; hin[i].aff_b = aff_b_init;
			for (int k = 0; k < 5; k++) hin[i].minRes[k] = NAN;
		}
		kin = hin;
		T.last_in = hin;
	}
	TrackLevels TL;
	fill_levels(ctx, T, &ctx->slots[frame_slot], new_ab_exposure, TL);

	// cluster size: as many CTAs per hypothesis as fit on the device, power of two, <= 16 (non-portable size,
	// allowed explicitly below; measured best at 256 threads x 16 CTAs for one VGA hypothesis)
	int cs = 16;
	while (cs > 1 && nHyp * cs > ctx->sm_count) cs >>= 1;
	int max_n = 0;
	for (int l = 0; l <= coarsestLvl; l++) max_n = T.pc_n[l] > max_n ? T.pc_n[l] : max_n;
	const size_t smem_cap_bytes = 160 * 1024;
	// shared-memory slots (cloud records) per CTA when the levels are split over `div` CTAs: the slices of all levels if
	// they fit together (the kernel then stages them once, up front), else the largest level's slice
	auto slots_for = [&](int div) -> int {
		size_t all = 0;
		for (int l = 0; l <= coarsestLvl; l++) all += (size_t)(T.pc_n[l] + div - 1) / div;
		if (all * sizeof(float4) <= smem_cap_bytes) return (int)all;
		return (int)std::min<size_t>((size_t)((max_n + div - 1) / div), smem_cap_bytes / sizeof(float4));
	};
	int cap = slots_for(cs);
	size_t dyn = (size_t)cap * sizeof(float4);
	// tuning knobs (dso_b200_track_set_tuning; the DSO_B200_TRACK_* environment variables give the defaults): threads per
	// CTA and cluster size
	static int env_threads = -1, env_cluster = -1;
	if (env_threads < 0)
	{
		const char* e = getenv("DSO_B200_TRACK_THREADS"); env_threads = e ? atoi(e) : 0;
		e = getenv("DSO_B200_TRACK_CLUSTER"); env_cluster = e ? atoi(e) : 0;
	}
	const int want_threads = ctx->tune_threads ? ctx->tune_threads : env_threads;
	const int want_cluster = ctx->tune_cluster ? ctx->tune_cluster : env_cluster;
	int nt = want_threads ? want_threads : PERS_THREADS;
	if (want_cluster) cs = want_cluster;
	while (cs > 1 && nHyp * cs > ctx->sm_count) cs >>= 1;
	cap = slots_for(cs);
	dyn = (size_t)cap * sizeof(float4);
	if (nt != 128 && nt != 384) nt = 256;
	const bool prof = ctx->d_phase_cycles != nullptr;
	typedef void (*kern_t)(TrackLevels, dso_b200_track_opts, const TrackIn*, dso_b200_track_result*, int, int, long long*,
	                       dso_b200_track_result*, unsigned int*, unsigned int, int, float*, unsigned int*, TrackIn);
	kern_t kern = nt == 128 ? (prof ? (kern_t)track_persistent_kernel<128, true> : (kern_t)track_persistent_kernel<128, false>)
	            : nt == 384 ? (prof ? (kern_t)track_persistent_kernel<384, true> : (kern_t)track_persistent_kernel<384, false>)
	                        : (prof ? (kern_t)track_persistent_kernel<256, true> : (kern_t)track_persistent_kernel<256, false>);
	static bool attr_set[64][6] = { { false } };
	const int ki = (nt == 128 ? 0 : (nt == 384 ? 4 : 2)) + (prof ? 1 : 0);
	if (!attr_set[ctx->device & 63][ki])
	{
		CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_cap_bytes));
		CK(cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
		attr_set[ctx->device & 63][ki] = true;
	}
	// large clouds, one hypothesis: spread it over several clusters (DSO_B200_TRACK_NCLUSTERS overrides; 1 = off).  All
	// CTAs exchange through a spin barrier, so the count is capped by what can be resident at once.
	int nc = 1;
	static int env_nc_s = -1;
	if (env_nc_s < 0) { const char* e = getenv("DSO_B200_TRACK_NCLUSTERS"); env_nc_s = e ? atoi(e) : 0; }
	const int env_nc = ctx->tune_nclusters ? ctx->tune_nclusters : env_nc_s;
	if (nHyp == 1 && (env_nc > 1 || (env_nc == 0 && max_n >= 32768)))
	{
		cudaLaunchConfig_t q = {};
		q.gridDim = dim3(cs, 1, 1); q.blockDim = dim3(nt, 1, 1);
		cudaLaunchAttribute qa[1];
		qa[0].id = cudaLaunchAttributeClusterDimension; qa[0].val.clusterDim.x = cs; qa[0].val.clusterDim.y = 1; qa[0].val.clusterDim.z = 1;
		q.attrs = qa; q.numAttrs = 1;
		q.dynamicSmemBytes = 0;
		int maxc = 0;
		nc = env_nc > 1 ? env_nc : (max_n + 16383) / 16384;
		// the per-CTA slice shrinks with nc, and so does the dynamic shared memory the query must assume
		for (int it = 0; it < 4; it++)
		{
			q.dynamicSmemBytes = (size_t)slots_for(cs * nc) * sizeof(float4);
			CK(cudaOccupancyMaxActiveClusters(&maxc, kern, &q));
			if (maxc < 1) maxc = 1;
			if (nc <= maxc) break;
			nc = maxc;
		}
		if (nc > DSO_B200_TRACK_MAX_NC) nc = DSO_B200_TRACK_MAX_NC;
		if (nc < 1) nc = 1;
		cap = slots_for(cs * nc);
		dyn = (size_t)cap * sizeof(float4);
		if (nc > 1) CK(cudaMemsetAsync(T.d_mc_count, 0, sizeof(unsigned int) * DSO_B200_MAX_HYP, T.stream));
	}
	// one hypothesis: its input block rides in the launch parameters (the ring block, host memory, stays the record the
	// pose-less replay reads); the device-resident default block of a fresh slot is passed by pointer
	const bool inline_in = nHyp == 1 && (const void*)kin != (const void*)T.d_track_in;
	TrackIn in0;
	memset(&in0, 0, sizeof(in0));
	if (inline_in) in0 = kin[0];
	cudaLaunchConfig_t cfg = {};
	cfg.blockDim = dim3(nt, 1, 1);
	cfg.stream = T.stream;
	cudaLaunchAttribute at[2];
	at[0].id = cudaLaunchAttributeClusterDimension;
	at[0].val.clusterDim.x = cs; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
	// nc > 1: the clusters of one hypothesis wait for each other in global memory, so they must all be resident at once.
	// A cooperative launch makes the driver guarantee that (or refuse the launch, e.g. when another context holds SMs);
	// a refused launch falls back to one cluster, which needs no co-residency.
	at[1].id = cudaLaunchAttributeCooperative; at[1].val.cooperative = 1;
	cfg.attrs = at;
	for (int attempt = 0; attempt < 2; attempt++)
	{
		cfg.gridDim = dim3(nHyp * cs * nc, 1, 1);
		cfg.dynamicSmemBytes = dyn;
		cfg.numAttrs = nc > 1 ? 2 : 1;
		cudaError_t le = cudaLaunchKernelEx(&cfg, kern, TL, o, inline_in ? (const TrackIn*)nullptr : kin, (dso_b200_track_result*)T.d_track_out, coarsestLvl, cap, (long long*)ctx->d_phase_cycles,
		                                    seq ? (dso_b200_track_result*)T.h_track_out + (size_t)(seq % TRACK_IN_RING) * DSO_B200_MAX_HYP : (dso_b200_track_result*)nullptr,
		                                    T.h_flags + (size_t)(seq % TRACK_IN_RING) * DSO_B200_MAX_HYP, seq,
		                                    nc, T.d_mc_part, T.d_mc_count, in0);
		if (le == cudaSuccess) break;
		if (nc > 1 && attempt == 0)
		{
			(void)cudaGetLastError();
			nc = 1;
			cap = slots_for(cs);
			dyn = (size_t)cap * sizeof(float4);
			continue;
		}
		CK(le);
	}
	ctx->last_shape[0] = nt; ctx->last_shape[1] = cs; ctx->last_shape[2] = nc;
	LAUNCHED(ctx, 1);
	return 0;
}

extern "C" int dso_b200_track_batched_async(dso_b200_ctx* ctx, int trk, int frame_slot, int nHyp, const double* pose12_init,
                                            double aff_a_init, double aff_b_init, float new_ab_exposure, int coarsestLvl,
                                            const dso_b200_track_opts* opts)
{
	return enqueue_track(ctx, trk, frame_slot, nHyp, pose12_init, aff_a_init, aff_b_init, new_ab_exposure, coarsestLvl, opts);
}

extern "C" int dso_b200_track_set_tuning(dso_b200_ctx* ctx, int threads, int cluster_size, int n_clusters)
{
	ARG_CHECK(ctx, "null context");
	ARG_CHECK(threads == 0 || threads == 128 || threads == 256, "threads must be 0, 128 or 256");
	ARG_CHECK(cluster_size == 0 || cluster_size == 1 || cluster_size == 2 || cluster_size == 4 || cluster_size == 8 || cluster_size == 16, "cluster_size must be 0 or a power of two <= 16");
	ARG_CHECK(n_clusters >= 0 && n_clusters <= DSO_B200_TRACK_MAX_NC, "n_clusters out of range");
	ctx->tune_threads = threads; ctx->tune_cluster = cluster_size; ctx->tune_nclusters = n_clusters;
	return 0;
}

extern "C" int dso_b200_track_last_shape(dso_b200_ctx* ctx, int* shape3)
{
	ARG_CHECK(ctx && shape3, "null argument");
	for (int i = 0; i < 3; i++) shape3[i] = ctx->last_shape[i];
	return 0;
}

extern "C" int dso_b200_track_phase_cycles(dso_b200_ctx* ctx, int enable, long long* cycles7)
{
	ARG_CHECK(ctx, "null context");
	CK(cudaSetDevice(ctx->device));
	if (enable && !ctx->d_phase_cycles) { CK(cudaMalloc(&ctx->d_phase_cycles, 960)); CK(cudaMemset(ctx->d_phase_cycles, 0, 960)); }
	if (cycles7 && ctx->d_phase_cycles)
	{
		CK(cudaStreamSynchronize(ctx->trk[0].stream));
		CK(cudaStreamSynchronize(ctx->trk[1].stream));
		CK(cudaMemcpy(cycles7, ctx->d_phase_cycles, 960, cudaMemcpyDeviceToHost));
	}
	if (!enable && ctx->d_phase_cycles) { cudaFree(ctx->d_phase_cycles); ctx->d_phase_cycles = nullptr; }
	return 0;
}

extern "C" int dso_b200_track_fetch(dso_b200_ctx* ctx, int trk, int nHyp, dso_b200_track_result* out)
{
	ARG_CHECK(ctx && (trk == 0 || trk == 1) && out && nHyp >= 1 && nHyp <= DSO_B200_MAX_HYP, "bad arguments");
	TrackerSlot& T = ctx->trk[trk];
	if (T.fetched != T.seq)
	{
		// oldest call whose results have not been handed out: poll the flags the kernel writes after its results (mapped
		// pinned memory); look at the stream now and then so that a failed launch or a device fault surfaces as an error
		// instead of a hang
		const unsigned int want = next_seq(T.fetched);
		const size_t blk = (size_t)(want % TRACK_IN_RING) * DSO_B200_MAX_HYP;
		volatile unsigned int* f = T.h_flags + blk;
		unsigned long long spins = 0;
		for (int i = 0; i < nHyp; i++)
			while (f[i] != want)
			{
				if ((++spins & 0x3fff) == 0)
				{
					cudaError_t e = cudaStreamQuery(T.stream);
					if (e != cudaSuccess && e != cudaErrorNotReady) CK(e);
					if (e == cudaSuccess && f[i] != want) { dso_set_error("dso_b200_track_fetch: stream idle but results missing"); return DSO_B200_E_STATE; }
				}
			}
		std::atomic_thread_fence(std::memory_order_acquire);
		memcpy(out, (const dso_b200_track_result*)T.h_track_out + blk, sizeof(dso_b200_track_result) * nHyp);
		T.fetched = want;
		return 0;
	}
	// nothing pending with host-visible results (kernel-only launches): the device-side results of the last launch
	CK(cudaSetDevice(ctx->device));
	CK(cudaStreamSynchronize(T.stream));
	CK(cudaMemcpy(out, T.d_track_out, sizeof(dso_b200_track_result) * nHyp, cudaMemcpyDeviceToHost));
	return 0;
}

extern "C" int dso_b200_track_batched(dso_b200_ctx* ctx, int trk, int frame_slot, int nHyp, const double* pose12_init,
                                      double aff_a_init, double aff_b_init, float new_ab_exposure, int coarsestLvl,
                                      const dso_b200_track_opts* opts, dso_b200_track_result* out)
{
	int rc = enqueue_track(ctx, trk, frame_slot, nHyp, pose12_init, aff_a_init, aff_b_init, new_ab_exposure, coarsestLvl, opts);
	if (rc) return rc;
	return dso_b200_track_fetch(ctx, trk, nHyp, out);
}

// The pipelined front end in one call: track the (already built) frame_slot, enqueue the upload + pyramid of the NEXT
// image into next_slot on the mapping stream while the tracker runs, then wait for the tracking result.
extern "C" int dso_b200_track_and_prefetch(dso_b200_ctx* ctx, int trk, int frame_slot, int nHyp, const double* pose12_init,
                                           double aff_a_init, double aff_b_init, float new_ab_exposure, int coarsestLvl,
                                           const dso_b200_track_opts* opts, int next_slot, const float* next_color_host,
                                           int w, int h, int levels, dso_b200_track_result* out)
{
	ARG_CHECK(pose12_init && out, "null argument");
	int rc = enqueue_track(ctx, trk, frame_slot, nHyp, pose12_init, aff_a_init, aff_b_init, new_ab_exposure, coarsestLvl, opts);
	if (rc) return rc;
	if (next_slot >= 0 && next_color_host)
	{
		rc = dso_b200_make_images_async(ctx, next_slot, next_color_host, w, h, levels);
		if (rc) return rc;                      // the tracking call stays in flight: dso_b200_track_fetch still returns it
	}
	return dso_b200_track_fetch(ctx, trk, nHyp, out);
}

extern "C" int dso_b200_track_frame(dso_b200_ctx* ctx, int trk, int frame_slot, const float* color_host, int w, int h,
                                    int levels, int nHyp, const double* pose12_init, double aff_a_init, double aff_b_init,
                                    float new_ab_exposure, int coarsestLvl, const dso_b200_track_opts* opts,
                                    dso_b200_track_result* out)
{
	ARG_CHECK(ctx && color_host && frame_slot >= 0 && frame_slot < ctx->n_slots, "bad arguments");
	ARG_CHECK(w >= 16 && h >= 16 && w <= ctx->max_w && h <= ctx->max_h && levels >= 1 && levels <= MAXL, "bad image size / levels");
	CK(cudaSetDevice(ctx->device));
	// the whole frame on the tracker slot's stream: upload, pyramid, tracking (no cross-stream hand-over)
	ARG_CHECK(trk == 0 || trk == 1, "bad tracker slot");
	cudaStream_t st = ctx->trk[trk].stream;
	CK(cudaMemcpyAsync(ctx->d_color, color_host, (size_t)w * h * sizeof(float), cudaMemcpyHostToDevice, st));
	int rc = pyramid_build_on(ctx, st, ctx->slots[frame_slot], ctx->d_color, w, h, levels, nullptr, false);
	if (rc) return rc;
	rc = frame_slot_produced(ctx, ctx->slots[frame_slot], st);
	if (rc) return rc;
	rc = enqueue_track(ctx, trk, frame_slot, nHyp, pose12_init, aff_a_init, aff_b_init, new_ab_exposure, coarsestLvl, opts);
	if (rc) return rc;
	return dso_b200_track_fetch(ctx, trk, nHyp, out);
}

// ---- evaluation without any allocation -----------------------------------------------------------------
// dso_b200_calc_res_noalloc: one calcRes + calcGSSSE pass for callers that must not allocate (see the header).
// The per-CTA partial sums and the "last CTA done" ticket that track_eval_kernel needs live in this fixed
// array instead of cudaMalloc'ed context memory.  It exists once per GPU that loads the library and is
// shared by every call, so calls that overlap on a GPU race on it.
#define NOALLOC_MAX_CTAS (DSO_B200_NOALLOC_MAX_POINTS / EVAL_THREADS)
#define NOALLOC_TICKET_FLOATS 64                       /* 256 bytes: the ticket word, padded */
static __device__ __align__(256) float g_noalloc_scratch[NOALLOC_TICKET_FLOATS + NOALLOC_MAX_CTAS * S_COUNT];

// In host code `g_noalloc_scratch` is only a handle for the runtime, not a device address; the address the
// kernel must be given comes from cudaGetSymbolAddress, looked up once per device with that device current.
static float* noalloc_scratch_ptr(int device)
{
	static float* ptrs[64] = { nullptr };
	if (device < 0 || device >= 64) { dso_set_error("dso_b200_calc_res_noalloc: device index %d not supported", device); return nullptr; }
	if (!ptrs[device])
	{
		float* p = nullptr;
		cudaError_t e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);
		if (e != cudaSuccess || !p)
		{
			dso_set_error("dso_b200_calc_res_noalloc: cudaGetSymbolAddress failed on device %d: %s", device, cudaGetErrorString(e));
			return nullptr;
		}
		ptrs[device] = p;
	}
	return ptrs[device];
}

extern "C" int dso_b200_calc_res_noalloc(const float* cloud_dev, int n_points, const float* frame_dev, int w, int h,
                                         int is_level0, float fx, float fy, float cx, float cy,
                                         const double* refToNew12, double ref_aff_a, double ref_aff_b, float ref_ab_exposure,
                                         double aff_a, double aff_b, float new_ab_exposure, float cutoffTH,
                                         void* stream, float* sums52_dev)
{
	ARG_CHECK(cloud_dev && frame_dev && refToNew12 && sums52_dev, "null pointer argument");
	ARG_CHECK(n_points >= 0 && w >= 8 && h >= 8, "bad sizes");
	if (n_points > DSO_B200_NOALLOC_MAX_POINTS)
	{
		dso_set_error("dso_b200_calc_res_noalloc: n_points = %d exceeds DSO_B200_NOALLOC_MAX_POINTS = %d (the scratch array is "
		              "fixed-size); use dso_b200_calc_res with a context for larger clouds", n_points, DSO_B200_NOALLOC_MAX_POINTS);
		return DSO_B200_E_LIMIT;
	}
	ARG_CHECK(((reinterpret_cast<size_t>(cloud_dev) | reinterpret_cast<size_t>(frame_dev)) & 15) == 0, "cloud_dev / frame_dev must be 16-byte aligned");
	int device = -1;
	CK(cudaGetDevice(&device));
	float* scratch = noalloc_scratch_ptr(device);
	if (!scratch) return DSO_B200_E_CUDA;
	cudaStream_t st = (cudaStream_t)stream;

	// the same float quantities as CoarseTracker.cpp:375-385, for this single level
	TrackLevels TL;
	memset(&TL, 0, sizeof(TL));
	TL.levels = 1;
	TL.w[0] = w; TL.h[0] = h; TL.fx[0] = fx; TL.fy[0] = fy; TL.cx[0] = cx; TL.cy[0] = cy;
	const float K[9] = { fx, 0, cx, 0, fy, cy, 0, 0, 1 };
	mat33f_inverse(K, TL.Ki[0]);
	TL.pc_n[0] = n_points;
	TL.ref_a = ref_aff_a; TL.ref_b = ref_aff_b; TL.ref_exposure = ref_ab_exposure; TL.new_exposure = new_ab_exposure;
	Pose pose; pose_from12(pose, refToNew12);
	EvalParams P;
	make_eval_params(TL, 0, pose, aff_a, aff_b, cutoffTH, 9.0f, P);
	P.lvl0 = is_level0 ? 1 : 0;

	// the array keeps whatever the previous call left: the ticket must start at zero (the partials are
	// fully rewritten by this launch before anything reads them)
	unsigned int* ticket = reinterpret_cast<unsigned int*>(scratch);
	float* partials = scratch + NOALLOC_TICKET_FLOATS;
	CK(cudaMemsetAsync(ticket, 0, NOALLOC_TICKET_FLOATS * sizeof(float), st));
	int g = (n_points + EVAL_THREADS - 1) / EVAL_THREADS;
	if (g < 1) g = 1;
	track_eval_kernel<false><<<g, EVAL_THREADS, 0, st>>>(P, reinterpret_cast<const float4*>(cloud_dev),
	                                                     reinterpret_cast<const float4*>(frame_dev), partials, ticket, sums52_dev,
	                                                     nullptr, nullptr);
	CK(cudaGetLastError());
	return 0;
}
