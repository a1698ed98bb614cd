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

// One reference point of calcRes (CoarseTracker.cpp:402-486) + its calcGSSSE contribution (:313-334).
template <bool DEBUG>
__device__ __forceinline__ void eval_point(const float4 pc, int i, const EvalParams& P, const float4* __restrict__ img,
                                           float* acc, unsigned char* status, float* warped)
{
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
		float q[3];
		proj3(P.Ki, x, y, P.t, id, 1.f, q);
		float KuT = fadd(fmul(P.fx, fdiv(q[0], q[2])), P.cx), KvT = fadd(fmul(P.fy, fdiv(q[1], q[2])), P.cy);
		proj3(P.Ki, x, y, P.t, id, -1.f, q);
		float KuT2 = fadd(fmul(P.fx, fdiv(q[0], q[2])), P.cx), KvT2 = fadd(fmul(P.fy, fdiv(q[1], q[2])), P.cy);
		proj3(P.RKi, x, y, P.t, id, -1.f, q);
		float Ku3 = fadd(fmul(P.fx, fdiv(q[0], q[2])), P.cx), Kv3 = fadd(fmul(P.fy, fdiv(q[1], q[2])), P.cy);
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
	const float w11 = dxdy, w01 = fsub(dy, dxdy), w10 = fsub(dx, dxdy);
	const float w00 = fadd(fsub(fsub(1.f, dx), dy), dxdy);
	const float4* bp = img + ix + iy * P.wl;
	const float4 p00 = __ldg(bp), p10 = __ldg(bp + 1), p01 = __ldg(bp + P.wl), p11 = __ldg(bp + 1 + P.wl);
	const float h0 = fadd(fadd(fadd(fmul(w11, p11.x), fmul(w01, p01.x)), fmul(w10, p10.x)), fmul(w00, p00.x));
	const float h1 = fadd(fadd(fadd(fmul(w11, p11.y), fmul(w01, p01.y)), fmul(w10, p10.y)), fmul(w00, p00.y));
	const float h2 = fadd(fadd(fadd(fmul(w11, p11.z), fmul(w01, p01.z)), fmul(w10, p10.z)), fmul(w00, p00.z));
	if (!isfinite(h0)) return;                           // :456
	const float refColor = pc.w;
	const float residual = fsub(h0, fadd(fmul(P.affLL0, refColor), P.affLL1));   // :458
	const float ar = fabsf(residual);
	const float hw = ar < P.huberTH ? 1.f : fdiv(P.huberTH, ar);                  // :459

	if (ar > P.cutoffTH)                                 // :462-468
	{
		acc[S_E] += P.maxEnergy;
		acc[S_NTERMS] += 1.f;
		acc[S_NSAT] += 1.f;
		if (DEBUG) status[i] = 1;
		return;
	}
	acc[S_E] += fmul(fmul(fmul(hw, residual), residual), fsub(2.f, hw));   // :473
	acc[S_NTERMS] += 1.f;
	acc[S_NWARP] += 1.f;
	if (DEBUG)
	{
		status[i] = 2;
		float* wp = warped + 8 * (size_t)i;
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

// LM state of one hypothesis, in shared memory.  The serial dependency chain of trackNewestCoarse
// (solve -> exp -> warp -> reduce -> accept?) is the latency floor of this kernel, so the step between
// two evaluations is executed by one warp: per-lane rows / pose entries in registers, uniform scalar
// work replicated in every lane (no divergent sections), shuffles instead of shared-memory round trips.
struct LMS
{
	double H[64], b[8];
	double cur[12], nw[12];                       // R (row-major 9) then t
	double a_cur, b_cur, a_new, b_new;
	double resOldE, resOldN, resOldFlowT, resOldFlowRT, resOldFlowN;   // flow sums kept raw
	double incnorm;                 // squared norm of the last increment
	double minRes[5];
	float lambda, levelCutoffRepeat;
	int lvl, iteration, phase, haveRepeated, done;
};

__device__ __forceinline__ double shfl_d(double v, int src) { return __shfl_sync(0xffffffffu, v, src); }
// reciprocal to ~1e-15 relative: float seed + two Newton steps (no IEEE division on the critical path)
__device__ __forceinline__ double fast_rcp(double x)
{
	double r = (double)__frcp_rn((float)x);
	r = r * fma(-x, r, 2.0);
	r = r * fma(-x, r, 2.0);
	return r;
}
__device__ __forceinline__ float scale_of(int i)
{
	return i < 3 ? SCALE_XI_ROT : (i < 6 ? SCALE_XI_TRANS : (i == 6 ? SCALE_A : SCALE_B));
}

// SE3 exponential coefficients A = sin(t)/t, B = (1-cos t)/t^2, C = (t - sin t)/t^3 as series in t^2
// (|t| < 0.5 for any LM increment; truncation error < 1e-16), falling back to sincos otherwise.
__device__ __forceinline__ void exp_coeffs(double th2, double& A, double& B, double& C)
{
	if (th2 < 0.25)
	{
		A = 1.0 + th2 * (-1.0 / 6 + th2 * (1.0 / 120 + th2 * (-1.0 / 5040 + th2 * (1.0 / 362880 + th2 * (-1.0 / 39916800 + th2 * (1.0 / 6227020800.0 + th2 * (-1.0 / 1307674368000.0)))))));
		B = 0.5 + th2 * (-1.0 / 24 + th2 * (1.0 / 720 + th2 * (-1.0 / 40320 + th2 * (1.0 / 3628800 + th2 * (-1.0 / 479001600 + th2 * (1.0 / 87178291200.0 + th2 * (-1.0 / 20922789888000.0)))))));
		C = 1.0 / 6 + th2 * (-1.0 / 120 + th2 * (1.0 / 5040 + th2 * (-1.0 / 362880 + th2 * (1.0 / 39916800 + th2 * (-1.0 / 6227020800.0 + th2 * (1.0 / 1307674368000.0 + th2 * (-1.0 / 355687428096000.0)))))));
	}
	else
	{
		const double th = sqrt(th2);
		double s, c;
		sincos(th, &s, &c);
		A = s / th; B = (1.0 - c) / th2; C = (th - s) / (th2 * th);
	}
}

// exp(x) for the relative affine gain: |x| is ~1e-2 in tracking, where a degree-11 Taylor series is accurate to
// < 1e-19 and much shorter than the general routine; anything larger goes through exp().
__device__ __forceinline__ double exp_small(double x)
{
	if (fabs(x) < 0.125)
	{
		double p = 1.0 / 39916800.0;
		p = fma(p, x, 1.0 / 3628800.0); p = fma(p, x, 1.0 / 362880.0); p = fma(p, x, 1.0 / 40320.0); p = fma(p, x, 1.0 / 5040.0);
		p = fma(p, x, 1.0 / 720.0); p = fma(p, x, 1.0 / 120.0); p = fma(p, x, 1.0 / 24.0); p = fma(p, x, 1.0 / 6.0);
		p = fma(p, x, 0.5); p = fma(p, x, 1.0); p = fma(p, x, 1.0);
		return p;
	}
	return exp(x);
}
// AffLight::fromToVecExposure (NumType.h:176-189) for the tracker's (reference, new frame) pair
__device__ __forceinline__ void aff_ref_to_new(const TrackLevels& TL, double a_req, double b_req, double& a, double& b)
{
	if (TL.ref_exposure == 1.f && TL.new_exposure == 1.f) a = exp_small(a_req - TL.ref_a);    // e*1/1 == e exactly
	else { double bb; aff_from_to(TL.ref_exposure, TL.new_exposure, TL.ref_a, TL.ref_b, a_req, b_req, a, bb); }
	b = b_req - a * TL.ref_b;
}

// One step of the trackNewestCoarse state machine (CoarseTracker.cpp:584-722), executed by warp 0 of
// every CTA of the cluster on identical inputs (so all CTAs take identical decisions, no broadcast).
// Consumes the sums of the evaluation just finished, leaves either S.done or the next EvalParams in P.
__device__ __forceinline__ void lm_step_warp(LMS& S, const float* s_tot, const TrackLevels& TL, const dso_b200_track_opts& opt,
                                             dso_b200_track_result* out, EvalParams& P, int n_points, int lane, long long* lmt)
{
	long long q0 = lmt ? clock64() : 0;
	const int i = lane & 7;
	// ---- sums -> this lane's row of the scaled H, b (:338-353) and the Vec6 (:511-519) ----
	const int nwarp = (int)s_tot[S_NWARP];
	const float invn = fdiv(1.0f, (float)((nwarp + 3) & ~3));
	const double si = (double)invn * (double)scale_of(i);
	double hn[8];
#pragma unroll
	for (int j = 0; j < 8; j++)
	{
		const int a = i < j ? i : j, b = i < j ? j : i;
		const int idx = a * 8 - ((a * (a - 1)) >> 1) + (b - a);
		hn[j] = (double)s_tot[S_ACC + idx] * si * (double)scale_of(j);
	}
	const double bn = (double)s_tot[S_ACC + 36 + i] * si;
	const double newE = (double)s_tot[S_E], newN = (double)s_tot[S_NTERMS];
	// flow indicators are only read when a level finishes: keep the raw sums, divide there (:514-516)
	const double newFlowT = (double)s_tot[S_FLOW_T], newFlowRT = (double)s_tot[S_FLOW_RT];
	const double newFlowN = (double)s_tot[S_FLOW_N];

	if (lmt) { long long q = clock64(); lmt[4] += q - q0; q0 = q; }
	// uniform copies of the scalar state
	int lvl = S.lvl, phase = S.phase, iteration = S.iteration;
	float lambda = S.lambda, rep = S.levelCutoffRepeat;
	const float rep_in = rep;
	double a_cur = S.a_cur, b_cur = S.b_cur;
	bool take = false, use_new = false, level_done = false, propose = false;

	if (phase == 0)
	{
		const double newSat = (double)fdiv(s_tot[S_NSAT], s_tot[S_NTERMS]);     // :517
		if (newSat > 0.6 && rep < 50) rep *= 2;                                // :586-593: same state, larger cut-off
		else { take = true; lambda = 0.01f; iteration = 0; phase = 1; level_done = (0 >= opt.maxIterations[lvl]); propose = !level_done; }
	}
	else
	{
		// :669  (E'/n' < E/n) evaluated as E'*n < E*n' (n, n' >= 0; equal up to a 1-ulp tie of the two quotients,
		// and false like the reference's NaN comparison when a count is 0) -- keeps two FP64 divisions off the chain
		const double acc_l = newE * S.resOldN, acc_r = S.resOldE * newN;
		const bool accept = acc_l < acc_r;
		if (accept)
		{
			take = true;
			a_cur = S.a_new; b_cur = S.b_new;
			if (lane < 12) S.cur[lane] = S.nw[lane];
			lambda *= 0.5f;
		}
		else { lambda *= 4; if (lambda < 0.001f) lambda = 0.001f; }
		iteration++;
		level_done = (!(S.incnorm > 1e-6) || iteration >= opt.maxIterations[lvl]);   // :700 (|inc| > 1e-3 on the squared norm), :613
		propose = !level_done;
		if (lane == 0)
		{
			if (out->trace_len < DSO_B200_TRACE_MARGINS) { out->trace_dE[out->trace_len] = (float)(acc_l - acc_r); out->trace_E[out->trace_len] = (float)acc_r; }
			if (out->trace_len < 256) out->trace[out->trace_len++] = accept ? 1 : 0;
			out->n_iterations++;
		}
	}
	if (lmt) { long long q = clock64(); lmt[5] += q - q0; q0 = q; }
	double resOldE = S.resOldE, resOldN = S.resOldN, resOldFlowT = S.resOldFlowT, resOldFlowRT = S.resOldFlowRT, resOldFlowN = S.resOldFlowN;
	if (take)
	{
		resOldE = newE; resOldN = newN; resOldFlowT = newFlowT; resOldFlowRT = newFlowRT; resOldFlowN = newFlowN;
		if (lane < 8)
		{
#pragma unroll
			for (int j = 0; j < 8; j++) S.H[i * 8 + j] = hn[j];
			S.b[i] = bn;
		}
	}
	__syncwarp();

	bool done = false;
	if (level_done)
	{
		// :708-722 (+ :727-746 when the finest level is finished)
		const float rmse = sqrtf((float)(resOldE / resOldN));
		const bool abort = (double)rmse > 1.5 * S.minRes[lvl];
		int next = lvl - 1;
		int haveRepeated = S.haveRepeated;
		if (rep > 1 && !haveRepeated) { next = lvl; haveRepeated = 1; }
		if (lane == 0)
		{
			out->lastResiduals[lvl] = rmse;
			out->lastFlowIndicators[0] = resOldFlowT / (resOldFlowN + 0.1); out->lastFlowIndicators[1] = 0;
			out->lastFlowIndicators[2] = resOldFlowRT / (resOldFlowN + 0.1);
			if (out->n_passes < DSO_B200_MAX_PASSES) { out->pass_level[out->n_passes] = lvl; out->pass_rmse[out->n_passes] = rmse; out->n_passes++; }
			S.haveRepeated = haveRepeated;
			if (abort) { out->ok = 0; out->reserved = lvl; }
			else if (next < 0)
			{
				for (int k = 0; k < 12; k++) out->pose[k] = S.cur[k];
				out->aff_a = a_cur; out->aff_b = b_cur;
				int ok = 1;
				if ((opt.affineOptModeA != 0 && (fabsf((float)a_cur) > 1.2f)) || (opt.affineOptModeB != 0 && (fabsf((float)b_cur) > 200.f))) ok = 0;
				else
				{
					double ra, rb;
					aff_from_to(TL.ref_exposure, TL.new_exposure, TL.ref_a, TL.ref_b, a_cur, b_cur, ra, rb);
					if ((opt.affineOptModeA == 0 && (fabsf(logf((float)ra)) > 1.5f)) || (opt.affineOptModeB == 0 && (fabsf((float)rb) > 200.f))) ok = 0;
				}
				out->ok = ok;
				if (ok)                                                            // :743-744
				{
					if (opt.affineOptModeA < 0) out->aff_a = 0;
					if (opt.affineOptModeB < 0) out->aff_b = 0;
				}
			}
		}
		done = abort || next < 0;
		lvl = next; phase = 0; rep = 1; iteration = 0;                         // :584 of the next level pass
	}

	if (lmt) { long long q = clock64(); lmt[0] += q - q0; q0 = q; }
	double a_req = a_cur, b_req = b_cur, affA = 1.0, affB = 0.0;
	double pe = (lane < 12) ? S.cur[lane] : 0.0;       // this lane's pose entry of the state to evaluate next
	double incnorm = S.incnorm;
	if (propose)
	{
		// ---- :613-665: damped solve, one lane per row.  H + lambda*diag(H) is SPD, so no pivoting is needed
		// (Eigen's pivoted LDLT gives the same solution up to rounding).  Division-free Gauss-Jordan: every step
		// is row_i <- g*row_i - f*row_k for all i != k with (g, f) = (piv, row_i[k]) * 2^-exponent(piv): a power-of-
		// two rescale (exact) replaces the reciprocal chain on the critical path and bounds the growth to 2x per
		// step; the eight reciprocals are taken once at the end, in parallel.  Row scaling does not change the
		// solution. ----
		const double damp = (double)(1.f + lambda);
		double row[8], x[8];
#pragma unroll
		for (int j = 0; j < 8; j++) row[j] = take ? hn[j] : S.H[i * 8 + j];
		double rhs = -(take ? bn : S.b[i]);
#pragma unroll
		for (int j = 0; j < 8; j++) if (j == i) row[j] *= damp;
		// :621-643 fixed-affine variants: the reference solves the 6x6 / 7x7 system without the fixed parameter's
		// row and column; replacing that row and column by the identity's (and its right-hand side by 0) yields
		// the same increments for the free parameters and exactly 0 for the fixed one.
		const bool fixA = opt.affineOptModeA < 0, fixB = opt.affineOptModeB < 0;
		if (fixA || fixB)
		{
			const bool fi = (i == 6 && fixA) || (i == 7 && fixB);
#pragma unroll
			for (int j = 0; j < 8; j++)
				if (fi || (j == 6 && fixA) || (j == 7 && fixB)) row[j] = (i == j) ? 1.0 : 0.0;
			if (fi) rhs = 0.0;
		}
#pragma unroll
		for (int k = 0; k < 8; k++)
		{
			const double piv = shfl_d(row[k], k);
			const double prhs = shfl_d(rhs, k);
			const bool self = (i == k);
			const double sc = __hiloint2double((2046 - ((__double2hiint(piv) >> 20) & 0x7ff)) << 20, 0);   // 2^-exponent(piv)
			const double f = self ? 0.0 : row[k] * sc;
			const double g = self ? 1.0 : piv * sc;
#pragma unroll
			for (int j = 0; j < 8; j++)
			{
				if (j <= k) continue;                       // columns <= k of the other rows become 0 / are never read again
				const double pj = shfl_d(row[j], k);
				row[j] = fma(-f, pj, g * row[j]);
			}
			rhs = fma(-f, prhs, g * rhs);
#pragma unroll
			for (int j = 0; j < 8; j++) if (j < k) row[j] = g * row[j];   // keep the already-diagonalised part consistent (only row[i], j == i, matters)
			if (!self) row[k] = 0.0;
		}
		// row i is now (0,..,d_i,..,0 | rhs_i): x_i = rhs_i / d_i
		double di = 0.0;
#pragma unroll
		for (int j = 0; j < 8; j++) if (j == i) di = row[j];
		const double xi = rhs * fast_rcp(di);
#pragma unroll
		for (int k = 0; k < 8; k++) x[k] = shfl_d(xi, k);
		if (lmt) { long long q = clock64(); lmt[1] += q - q0; q0 = q; }
		float extrapFac = 1.f;
		if (lambda < 0.001f) extrapFac = sqrtf(sqrtf(0.001f / lambda));       // :650
		double nrm = 0, sum = 0, is[8];
#pragma unroll
		for (int k = 0; k < 8; k++) { x[k] *= (double)extrapFac; nrm = fma(x[k], x[k], nrm); is[k] = x[k] * (double)scale_of(k); sum += is[k]; }
		if (!isfinite(sum))                                                    // :659
		{
#pragma unroll
			for (int k = 0; k < 8; k++) is[k] = 0;
		}
		incnorm = nrm;
		a_req = a_cur + is[6]; b_req = b_cur + is[7];
		aff_ref_to_new(TL, a_req, b_req, affA, affB);   // independent of the pose chain below: overlaps with it
		// exp(inc) (Sophus SE3::exp): R = I + A W + B W^2, t = (I + B W + C W^2) u
		const double wx = is[3], wy = is[4], wz = is[5];
		const double th2 = wx * wx + wy * wy + wz * wz;
		double A, B, C;
		exp_coeffs(th2, A, B, C);
		// this lane's row r of exp(inc) (R lanes: r = lane/3, t lanes: r = lane-9)
		const int r = (lane < 9) ? lane / 3 : (lane < 12 ? lane - 9 : 0);
		const int c = (lane < 9) ? lane - 3 * r : 0;
		const double wr = r == 0 ? wx : (r == 1 ? wy : wz);
		// W row r = (0,-wz,wy),(wz,0,-wx),(-wy,wx,0); W^2 row r = wr*w - th2*e_r
		const double W0 = r == 0 ? 0.0 : (r == 1 ? wz : -wy);
		const double W1 = r == 0 ? -wz : (r == 1 ? 0.0 : wx);
		const double W2 = r == 0 ? wy : (r == 1 ? -wx : 0.0);
		const double Q0 = wr * wx - (r == 0 ? th2 : 0.0), Q1 = wr * wy - (r == 1 ? th2 : 0.0), Q2 = wr * wz - (r == 2 ? th2 : 0.0);
		const double e0 = (r == 0 ? 1.0 : 0.0) + A * W0 + B * Q0;
		const double e1 = (r == 1 ? 1.0 : 0.0) + A * W1 + B * Q1;
		const double e2 = (r == 2 ? 1.0 : 0.0) + A * W2 + B * Q2;
		const double v0 = (r == 0 ? 1.0 : 0.0) + B * W0 + C * Q0;
		const double v1 = (r == 1 ? 1.0 : 0.0) + B * W1 + C * Q1;
		const double v2 = (r == 2 ? 1.0 : 0.0) + B * W2 + C * Q2;
		const double et = v0 * is[0] + v1 * is[1] + v2 * is[2];
		// refToNew_new = exp(inc) * refToNew_current (:661): entry (r,c) = sum_k e[r][k] cur[k][c]; t_r = sum_k e[r][k] t_k + et_r
		const int src0 = (lane < 9) ? c : 9, src1 = (lane < 9) ? 3 + c : 10, src2 = (lane < 9) ? 6 + c : 11;
		const double c0 = shfl_d(pe, src0), c1 = shfl_d(pe, src1), c2 = shfl_d(pe, src2);
		double ne = e0 * c0 + e1 * c1 + e2 * c2;
		if (lane >= 9) ne += et;
		pe = ne;
		if (lane < 12) S.nw[lane] = ne;
		use_new = true;
	}

	if (lmt) { long long q = clock64(); lmt[2] += q - q0; q0 = q; }
	// ---- CoarseTracker.cpp:375-378 + :385 for the state to evaluate next (skipped when done) ----
	if (!done)
	{
		const int r = (lane < 9) ? lane / 3 : 0, c = (lane < 9) ? lane - 3 * r : 0;
		const float R0 = (float)shfl_d(pe, 3 * r), R1 = (float)shfl_d(pe, 3 * r + 1), R2 = (float)shfl_d(pe, 3 * r + 2);
		const float rk = dot3e(R0, TL.Ki[lvl][c], R1, TL.Ki[lvl][3 + c], R2, TL.Ki[lvl][6 + c]);
		if (!use_new) aff_ref_to_new(TL, a_req, b_req, affA, affB);
		const double a = affA, b = affB;
		if (lane < 9) { P.RKi[lane] = rk; if (level_done) P.Ki[lane] = TL.Ki[lvl][lane]; }
		else if (lane < 12) P.t[lane - 9] = (float)pe;
		if (lane == 12) { P.affLL0 = (float)a; P.affLL1 = (float)b; }
		if (lane == 13 && (level_done || rep != rep_in))
		{
			// level- and cut-off-dependent parameters only change at a level switch / cut-off doubling
			const float cutoff = opt.coarseCutoffTH * rep;
			P.cutoffTH = cutoff;
			P.maxEnergy = fsub(fmul(fmul(2.0f, opt.huberTH), cutoff), fmul(opt.huberTH, opt.huberTH));
			if (level_done)
			{
				P.fx = TL.fx[lvl]; P.fy = TL.fy[lvl]; P.cx = TL.cx[lvl]; P.cy = TL.cy[lvl];
				P.wl = TL.w[lvl];
				P.wl_m3 = (float)(TL.w[lvl] - 3); P.hl_m3 = (float)(TL.h[lvl] - 3);
				P.n = TL.pc_n[lvl];
				P.lvl0 = (lvl == 0);
			}
		}
	}
	if (lane == 0)
	{
		S.lvl = lvl; S.phase = phase; S.iteration = iteration; S.lambda = lambda; S.levelCutoffRepeat = rep;
		S.a_cur = a_cur; S.b_cur = b_cur;
		if (use_new) { S.a_new = a_req; S.b_new = b_req; S.incnorm = incnorm; }
		S.resOldE = resOldE; S.resOldN = resOldN; S.resOldFlowT = resOldFlowT; S.resOldFlowRT = resOldFlowRT; S.resOldFlowN = resOldFlowN;
		S.done = done ? 1 : 0;
		out->n_evals++; out->point_evals += n_points;
	}
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

// phase_cycles (optional, debug): SM cycles spent by thread 0 of CTA 0 in [point evaluation, CTA
// reduction, cluster exchange + barrier, LM step], number of evaluations, whole kernel cycles and ns.
template <int NT>
__global__ void __launch_bounds__(NT, 1)
track_persistent_kernel(TrackLevels TL, dso_b200_track_opts opt, const TrackIn* __restrict__ in,
                        dso_b200_track_result* out, int coarsestLvl, int smem_pts_cap, long long* phase_cycles,
                        dso_b200_track_result* host_out, unsigned int* host_flags, unsigned int seq,
                        int NC, float* mc_part, unsigned int* mc_count)
{
	// NC > 1: one hypothesis is spread over NC clusters (large clouds): cluster totals go through global memory and a
	// monotonic arrival counter; every CTA of the grid must be resident (the host caps NC by cudaOccupancyMaxActiveClusters)
	cg::cluster_group cluster = cg::this_cluster();
	const unsigned CS = cluster.num_blocks();
	const unsigned rank = cluster.block_rank();
	const int hyp = blockIdx.x / (CS * NC);
	const int cidx = (blockIdx.x / CS) % NC;
	unsigned int mc_round = 0;
	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	long long tph[5] = { 0, 0, 0, 0, 0 };
	long long tc0 = 0, tk0 = 0, tg0 = 0;
	long long lmt[6] = { 0, 0, 0, 0, 0, 0 };   // LM step split: sums+control, solve, exp+pose, params+stores
	if (phase_cycles) { tk0 = clock64(); asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tg0)); }

	__shared__ EvalParams sP;
	__shared__ LMS S;
	__shared__ float s_w[NT / 32][64];
	__shared__ float xchg[2][MAX_CLUSTER][S_COUNT];
	__shared__ __align__(8) unsigned long long xbar[2];   // one mbarrier per exchange buffer
	__shared__ float s_tot[S_COUNT];
	__shared__ dso_b200_track_result s_res;
	extern __shared__ float4 s_pts[];

	for (int i = tid; i < (int)(sizeof(dso_b200_track_result) / 4); i += NT) reinterpret_cast<unsigned int*>(&s_res)[i] = 0u;
	__syncthreads();
	if (warp == 0)
	{
		// the input block may live in mapped host memory: one coalesced read by 19 lanes, then shared memory
		__shared__ TrackIn s_in;
		static_assert(sizeof(TrackIn) == 19 * sizeof(double), "TrackIn layout");
		if (lane < 19) reinterpret_cast<double*>(&s_in)[lane] = reinterpret_cast<const double*>(in + hyp)[lane];
		__syncwarp();
		if (lane == 0)
		{
			const TrackIn& I = s_in;
			for (int i = 0; i < 12; i++) S.cur[i] = I.pose[i];
			S.a_cur = I.aff_a; S.b_cur = I.aff_b; S.a_new = I.aff_a; S.b_new = I.aff_b;
			for (int i = 0; i < 5; i++) S.minRes[i] = I.minRes[i];
			S.haveRepeated = 0; S.done = 0; S.incnorm = 0; S.lambda = 0.01f;
			S.lvl = coarsestLvl; S.phase = 0; S.levelCutoffRepeat = 1; S.iteration = 0;
			S.resOldE = 0; S.resOldN = 1; S.resOldFlowT = 0; S.resOldFlowRT = 0; S.resOldFlowN = 0;
			s_res.ok = 0; s_res.reserved = -1;
			for (int i = 0; i < 12; i++) s_res.pose[i] = I.pose[i];
			s_res.aff_a = I.aff_a; s_res.aff_b = I.aff_b;
			for (int i = 0; i < 5; i++) s_res.lastResiduals[i] = NAN;        // :563
			for (int i = 0; i < 3; i++) s_res.lastFlowIndicators[i] = 1000;  // :564
			s_res.n_passes = 0; s_res.n_iterations = 0; s_res.n_evals = 0; s_res.trace_len = 0; s_res.point_evals = 0;
			// CoarseTracker.cpp:375-385 for the very first evaluation (:584)
			Pose p0; pose_from12(p0, I.pose);
			make_eval_params(TL, coarsestLvl, p0, I.aff_a, I.aff_b, opt.coarseCutoffTH, opt.huberTH, sP);
		}
		__syncwarp();
	}
	if (tid == 0) { mbar_init(&xbar[0], 1); mbar_init(&xbar[1], 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
	__syncthreads();
	cluster.sync();            // every CTA's barriers exist before anybody signals them
	int buf = 0;
	unsigned xphase[2] = { 0u, 0u };
	int staged_lvl = -1;
	while (true)
	{
		const int lvl = S.lvl;
		const int n = TL.pc_n[lvl];
		const int per = (n + CS * NC - 1) / (CS * NC);
		const int beg = min((int)(cidx * CS + rank) * per, n), end = min(beg + per, n);
		const float4* __restrict__ gp = TL.pc[lvl];
		const bool use_smem = (end - beg) <= smem_pts_cap;
		if (use_smem && staged_lvl != lvl)
		{
			for (int i = beg + tid; i < end; i += NT) s_pts[i - beg] = __ldg(gp + i);
			staged_lvl = lvl;
			__syncthreads();
		}
		const EvalParams P = sP;
		const float4* __restrict__ img = TL.img[lvl];
		if (phase_cycles) { __syncthreads(); tc0 = clock64(); }
		float acc[64];
#pragma unroll
		for (int k = 0; k < 64; k++) acc[k] = 0.f;
		if (use_smem)
			for (int i = beg + tid; i < end; i += NT) eval_point<false>(s_pts[i - beg], i, P, img, acc, nullptr, nullptr);
		else
			for (int i = beg + tid; i < end; i += NT) eval_point<false>(__ldg(gp + i), i, P, img, acc, nullptr, nullptr);
		if (phase_cycles) { __syncthreads(); long long t = clock64(); tph[0] += t - tc0; tc0 = t; }
		warp_reduce_halving(acc, lane);
		s_w[warp][2 * lane] = acc[0];
		s_w[warp][2 * lane + 1] = acc[1];
		__syncthreads();
		// CTA partial (warps in index order), pushed into every CTA of the cluster through DSMEM
		if (tid < S_COUNT)
		{
			float s = 0.f;
#pragma unroll
			for (int w = 0; w < NT / 32; w++) s += s_w[w][tid];
			for (unsigned r = 0; r < CS; r++) st_async_remote(&xchg[buf][rank][tid], &xbar[buf], r, s);
		}
		if (tid == 0) mbar_arrive_expect_tx(&xbar[buf], CS * S_COUNT * 4);   // this CTA expects CS partials of S_COUNT floats
		if (phase_cycles) { long long t = clock64(); tph[1] += t - tc0; tc0 = t; }
		mbar_wait(&xbar[buf], xphase[buf]);
		xphase[buf] ^= 1u;
		if (tid < S_COUNT)
		{
			// ranks in index order; loads issued together (fixed trip count), adds in order
			float v[MAX_CLUSTER];
#pragma unroll
			for (int r = 0; r < MAX_CLUSTER; r++) v[r] = (r < (int)CS) ? xchg[buf][r][tid] : 0.f;
			float s = 0.f;
#pragma unroll
			for (int r = 0; r < MAX_CLUSTER; r++) s += v[r];
			s_tot[tid] = s;
			if (NC > 1 && rank == 0) __stcg(mc_part + ((size_t)(hyp * 2 + (mc_round & 1)) * NC + cidx) * S_COUNT + tid, s);
		}
		__syncthreads();
		if (NC > 1)
		{
			mc_round++;
			if (tid == 0)
			{
				if (rank == 0) { __threadfence(); atomicAdd(mc_count + hyp, 1u); }
				const unsigned int want = mc_round * (unsigned)NC;
				// the grid is launched cooperatively (all clusters co-resident or the launch is refused), so this wait
				// always ends; the time limit only turns a violated assumption into a launch failure instead of a hang
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
			__syncthreads();
			if (tid < S_COUNT)
			{
				const float* src = mc_part + (size_t)(hyp * 2 + ((mc_round - 1) & 1)) * NC * S_COUNT + tid;
				float sum = 0.f;
				for (int c0 = 0; c0 < NC; c0 += 8)                // clusters in index order, eight loads in flight
				{
					float t[8];
#pragma unroll
					for (int q = 0; q < 8; q++) t[q] = (c0 + q < NC) ? __ldcg(src + (size_t)(c0 + q) * S_COUNT) : 0.f;
#pragma unroll
					for (int q = 0; q < 8; q++) if (c0 + q < NC) sum += t[q];
				}
				s_tot[tid] = sum;
			}
			__syncthreads();
		}
		if (phase_cycles) { long long t = clock64(); tph[2] += t - tc0; tc0 = t; }
		if (warp == 0) lm_step_warp(S, s_tot, TL, opt, &s_res, sP, n, lane, phase_cycles ? lmt : nullptr);
		__syncthreads();
		if (phase_cycles) { long long t = clock64(); tph[3] += t - tc0; tph[4]++; }
		if (S.done) break;
		buf ^= 1;
	}
	cluster.sync();                // nobody leaves while a peer may still be writing into it
	if (rank == 0 && cidx == 0)
	{
		const int nwords = sizeof(dso_b200_track_result) / 4;
		const unsigned int* src = reinterpret_cast<const unsigned int*>(&s_res);
		unsigned int* dst = reinterpret_cast<unsigned int*>(out + hyp);
		for (int i = tid; i < nwords; i += NT) dst[i] = src[i];
		if (host_out)
		{
			// results straight into mapped host memory, then the per-hypothesis flag: the host polls it instead of waiting
			// for a copy node and a stream synchronisation
			unsigned int* hdst = reinterpret_cast<unsigned int*>(host_out + hyp);
			for (int i = tid; i < nwords; i += NT) hdst[i] = src[i];
			__threadfence_system();
			__syncthreads();
			if (tid == 0) { *reinterpret_cast<volatile unsigned int*>(host_flags + hyp) = seq; __threadfence_system(); }
		}
	}
	if (phase_cycles && blockIdx.x == 0 && tid == 0)
	{
		for (int k = 0; k < 5; k++) phase_cycles[k] = tph[k];
		long long tg1; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tg1));
		phase_cycles[5] = clock64() - tk0;
		phase_cycles[6] = tg1 - tg0;
		for (int k = 0; k < 6; k++) phase_cycles[8 + k] = lmt[k];
	}
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
	rc = run_eval(ctx, trk, frame_slot, lvl, p, aff_a, aff_b, new_ab_exposure, cutoffTH, 9.0f, false, nullptr, nullptr);
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
	rc = run_eval(ctx, trk, frame_slot, lvl, p, aff_a, aff_b, new_ab_exposure, cutoffTH, 9.0f, true, d_status, d_warped);
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
	if (opts) o = *opts; else dso_b200_track_opts_default(&o);
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
	if (opts) o = *opts; else dso_b200_track_opts_default(&o);
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
	int per = (max_n + cs - 1) / cs;
	size_t smem_cap_bytes = 160 * 1024;
	int cap = (int)std::min<size_t>((size_t)per, smem_cap_bytes / sizeof(float4));
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
	per = (max_n + cs - 1) / cs;
	cap = (int)std::min<size_t>((size_t)per, smem_cap_bytes / sizeof(float4));
	dyn = (size_t)cap * sizeof(float4);
	void (*kern)(TrackLevels, dso_b200_track_opts, const TrackIn*, dso_b200_track_result*, int, int, long long*,
	             dso_b200_track_result*, unsigned int*, unsigned int, int, float*, unsigned int*) =
		nt == 128 ? track_persistent_kernel<128> : (nt == 512 ? track_persistent_kernel<512> : (nt == 384 ? track_persistent_kernel<384> : track_persistent_kernel<256>));
	if (nt != 128 && nt != 512 && nt != 384) nt = 256;
	static bool attr_set[64][4] = { { false } };
	const int ki = nt == 128 ? 0 : (nt == 256 ? 1 : (nt == 512 ? 2 : 3));
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
			q.dynamicSmemBytes = std::min<size_t>((size_t)((max_n + cs * nc - 1) / (cs * nc)), smem_cap_bytes / sizeof(float4)) * sizeof(float4);
			CK(cudaOccupancyMaxActiveClusters(&maxc, kern, &q));
			if (maxc < 1) maxc = 1;
			if (nc <= maxc) break;
			nc = maxc;
		}
		if (nc > DSO_B200_TRACK_MAX_NC) nc = DSO_B200_TRACK_MAX_NC;
		if (nc < 1) nc = 1;
		per = (max_n + cs * nc - 1) / (cs * nc);
		cap = (int)std::min<size_t>((size_t)per, smem_cap_bytes / sizeof(float4));
		dyn = (size_t)cap * sizeof(float4);
		if (nc > 1) CK(cudaMemsetAsync(T.d_mc_count, 0, sizeof(unsigned int) * DSO_B200_MAX_HYP, T.stream));
	}
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
		cudaError_t le = cudaLaunchKernelEx(&cfg, kern, TL, o, kin, (dso_b200_track_result*)T.d_track_out, coarsestLvl, cap, (long long*)ctx->d_phase_cycles,
		                                    seq ? (dso_b200_track_result*)T.h_track_out + (size_t)(seq % TRACK_IN_RING) * DSO_B200_MAX_HYP : (dso_b200_track_result*)nullptr,
		                                    T.h_flags + (size_t)(seq % TRACK_IN_RING) * DSO_B200_MAX_HYP, seq,
		                                    nc, T.d_mc_part, T.d_mc_count);
		if (le == cudaSuccess) break;
		if (nc > 1 && attempt == 0)
		{
			(void)cudaGetLastError();
			nc = 1;
			per = (max_n + cs - 1) / cs;
			cap = (int)std::min<size_t>((size_t)per, smem_cap_bytes / sizeof(float4));
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
	ARG_CHECK(threads == 0 || threads == 128 || threads == 256 || threads == 384 || threads == 512, "threads must be 0, 128, 256, 384 or 512");
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
	if (enable && !ctx->d_phase_cycles) { CK(cudaMalloc(&ctx->d_phase_cycles, 128)); CK(cudaMemset(ctx->d_phase_cycles, 0, 128)); }
	if (cycles7 && ctx->d_phase_cycles)
	{
		CK(cudaStreamSynchronize(ctx->trk[0].stream));
		CK(cudaStreamSynchronize(ctx->trk[1].stream));
		CK(cudaMemcpy(cycles7, ctx->d_phase_cycles, 112, cudaMemcpyDeviceToHost));
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
