// Next row N4 -- the two-frame initialiser's linearisation, replaces CoarseInitializer::calcResAndGS
// (reference: FullSystem/CoarseInitializer.cpp:370-732; makeK :1126-1155).
//
// Same family as the tracker's evaluation (warp, bilinear sample, Huber weight, 8-DoF Jacobian) with two additions: every
// point carries its own inverse depth, whose 1x1 block is eliminated per point (JbBuffer, the Schur complement acc9SC), and
// each point is observed through the 8-pixel residual pattern.  Eight lanes work on one point, one per pattern pixel; the
// 8-term sums that decide a mask or are handed back per point (energy, JbBuffer, maxstep) are taken in pattern order with
// individually rounded float operations, so isGood_new, energy_new, maxstep, lastHessian_new and JbBuffer_new are
// bit-identical to the reference; the 45 + 45 entries of the two 9x9 accumulators and the energy sum are FP32 per thread,
// then a fixed-order shuffle / shared-memory / last-block tree (deterministic, no atomics on data), within the 1e-4 bar.
//
// Roofline: gather-bound like K2a (per point 8 x (4 texels of the new frame + 4 of the first frame) of 16 B), launched a
// handful of times per bootstrap frame on <= ~10 k points: latency-bound at any realistic size.
#include "common.cuh"
#include "lm_math.cuh"

#define INIT_PTS_PER_BLOCK 32
#define INIT_NSUM 96            // 45 (acc9) + 45 (acc9SC) + energy sum + padding

struct InitParams
{
	float RKi[9], t[3];
	float aff0, aff1;
	float fxl, fyl, cxl, cyl;
	int wl, hl, npts;
	float huberTH;
	float alphaOpt, couplingWeight;
};

__constant__ int c_init_pattern[8][2] = { {0,-2}, {-1,-1}, {1,-1}, {-2,0}, {0,0}, {2,0}, {-1,1}, {0,2} };   // settings.cpp:274
__constant__ unsigned char c_tri9[45][2] = {
	{0,0},{0,1},{0,2},{0,3},{0,4},{0,5},{0,6},{0,7},{0,8}, {1,1},{1,2},{1,3},{1,4},{1,5},{1,6},{1,7},{1,8},
	{2,2},{2,3},{2,4},{2,5},{2,6},{2,7},{2,8}, {3,3},{3,4},{3,5},{3,6},{3,7},{3,8}, {4,4},{4,5},{4,6},{4,7},{4,8},
	{5,5},{5,6},{5,7},{5,8}, {6,6},{6,7},{6,8}, {7,7},{7,8}, {8,8} };

// sum of the 8 lanes of a point's group in pattern order (the reference's scalar "+=" loop starting from 0)
__device__ __forceinline__ float init_seq8(float v, int base)
{
	float s = __shfl_sync(0xffffffffu, v, base);
#pragma unroll
	for (int k = 1; k < 8; k++) s = fadd(s, __shfl_sync(0xffffffffu, v, base + k));
	return s;
}

// bilinear sample of the float4 (I, dx, dy, .) texel image: getInterpolatedElement33 / 31, util/globalFuncs.h:75-89, 126-140
__device__ __forceinline__ void init_interp(const float4* __restrict__ img, float x, float y, int width, float& h0, float& h1, float& h2, bool grad)
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

// One point per 8 lanes (lane = pattern pixel), 32 points per CTA.  partials: [gridDim.x][INIT_NSUM]; the last CTA to
// finish adds them in block order into sums_out: [0..44] acc9 upper triangle (row-major r <= c over dp0..dp7, r),
// [45..89] acc9SC, [90] the energy sum E.A.
__global__ void __launch_bounds__(256) init_res_gs_kernel(InitParams P, const float4* __restrict__ imgRef, const float4* __restrict__ imgNew,
                                                          const float* __restrict__ pu, const float* __restrict__ pv, const float* __restrict__ pid,
                                                          const unsigned char* __restrict__ isGood, const float* __restrict__ energy2,
                                                          const float* __restrict__ outlierTH, const float* __restrict__ iR,
                                                          unsigned char* isGood_new, float* energy_new2, float* maxstep_out, float* lastH_new,
                                                          float* Jb10, float* partials, unsigned int* ticket, float* sums_out)
{
	__shared__ float sJb[INIT_PTS_PER_BLOCK][12];
	__shared__ float s_w[8][INIT_NSUM];
	__shared__ bool s_last;
	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = tid >> 3, idx = tid & 7, base = lane & ~7;
	const int i = blockIdx.x * INIT_PTS_PER_BLOCK + g;
	const bool valid = i < P.npts;
	const int ii = valid ? i : 0;
	const float u0 = pu[ii], v0 = pv[ii], id = pid[ii];
	const bool good_in = valid && isGood[ii] != 0;

	// ---- projection (:463-470) ----
	const float x = fadd(u0, (float)c_init_pattern[idx][0]), y = fadd(v0, (float)c_init_pattern[idx][1]);
	float pt[3];
#pragma unroll
	for (int q = 0; q < 3; q++) pt[q] = fadd(dot3e(P.RKi[q * 3], x, P.RKi[q * 3 + 1], y, P.RKi[q * 3 + 2], 1.f), fmul(P.t[q], id));
	const float u = fdiv(pt[0], pt[2]), v = fdiv(pt[1], pt[2]);
	const float Ku = fadd(fmul(P.fxl, u), P.cxl), Kv = fadd(fmul(P.fyl, v), P.cyl);
	const float new_idepth = fdiv(id, pt[2]);
	bool ok = good_in && (Ku > 1.f && Kv > 1.f && Ku < (float)(P.wl - 2) && Kv < (float)(P.hl - 2) && new_idepth > 0.f);   // :473
	// the reference reads the first frame at (x, y) unchecked (setFirst keeps a pattern-sized border); a point closer to
	// the border than that is dropped here instead of read out of bounds
	ok = ok && (x >= 0.f && y >= 0.f && x < (float)(P.wl - 1) && y < (float)(P.hl - 1));
	float h0 = 0.f, h1 = 0.f, h2 = 0.f, rlR = 0.f, dmy;
	if (ok)
	{
		init_interp(imgNew, Ku, Kv, P.wl, h0, h1, h2, true);                       // :483
		init_interp(imgRef, x, y, P.wl, rlR, dmy, dmy, false);                     // :488
		if (!isfinite(rlR) || !isfinite(h0)) ok = false;                         // :491
	}
	const bool all_ok = ((__ballot_sync(0xffffffffu, ok) >> base) & 0xffu) == 0xffu;

	// ---- photometric error and Jacobians (:499-547) ----
	const float residual = fsub(fsub(h0, fmul(P.aff0, rlR)), P.aff1);
	const float ar = fabsf(residual);
	float hw = ar < P.huberTH ? 1.f : fdiv(P.huberTH, ar);
	const float e_i = fmul(fmul(fmul(hw, residual), residual), fsub(2.f, hw));
	const float dxdd = fdiv(fsub(P.t[0], fmul(P.t[2], u)), pt[2]);
	const float dydd = fdiv(fsub(P.t[1], fmul(P.t[2], v)), pt[2]);
	if (hw < 1.f) hw = __fsqrt_rn(hw);
	const float dxI = fmul(fmul(hw, h1), P.fxl), dyI = fmul(fmul(hw, h2), P.fyl);
	float J[9];
	J[0] = fmul(new_idepth, dxI);
	J[1] = fmul(new_idepth, dyI);
	J[2] = fmul(-new_idepth, fadd(fmul(u, dxI), fmul(v, dyI)));
	J[3] = fsub(fmul(fmul(-u, v), dxI), fmul(fadd(1.f, fmul(v, v)), dyI));
	J[4] = fadd(fmul(fadd(1.f, fmul(u, u)), dxI), fmul(fmul(u, v), dyI));
	J[5] = fadd(fmul(-v, dxI), fmul(u, dyI));
	J[6] = fmul(fmul(-hw, P.aff0), rlR);
	J[7] = -hw;
	J[8] = fmul(hw, residual);
	const float dd = fadd(fmul(dxI, dxdd), fmul(dyI, dydd));
	const float ma = fmul(dxdd, P.fxl), mb = fmul(dydd, P.fyl);
	const float ms = fdiv(1.0f, __fsqrt_rn(fadd(fmul(ma, ma), fmul(mb, mb))));   // :551

	// ---- per-point sums in pattern order (:504, :562-571) ----
	const float energy = init_seq8(e_i, base);
	float Jb[10];
#pragma unroll
	for (int k = 0; k < 8; k++) Jb[k] = init_seq8(fmul(J[k], dd), base);
	Jb[8] = init_seq8(fmul(J[8], dd), base);
	Jb[9] = init_seq8(fmul(dd, dd), base);
	float mstep = 1e10f;
#pragma unroll
	for (int k = 0; k < 8; k++)
	{
		const float mk = __shfl_sync(0xffffffffu, ms, base + k);
		const bool okk = __shfl_sync(0xffffffffu, ok ? 1 : 0, base + k) != 0;
		if (okk && mk < mstep) mstep = mk;
	}

	// ---- :574-584, :617-633, :652-675 ----
	const float oth = outlierTH[ii], e_old0 = energy2[2 * ii], e_old1 = energy2[2 * ii + 1];
	const bool good = good_in && all_ok && !(energy > fmul(oth, 20.f));
	float e_pt = valid ? e_old0 : 0.f;
	float wsc = 0.f;
	if (good)
	{
		e_pt = energy;
		const float lastH = Jb[9];
		const float idm1 = fsub(id, 1.f);
		Jb[8] = fadd(Jb[8], fmul(P.alphaOpt, idm1));
		Jb[9] = fadd(Jb[9], P.alphaOpt);
		if (P.alphaOpt == 0.f)
		{
			Jb[8] = fadd(Jb[8], fmul(P.couplingWeight, fsub(id, iR[ii])));
			Jb[9] = fadd(Jb[9], P.couplingWeight);
		}
		Jb[9] = fdiv(1.f, fadd(1.f, Jb[9]));
		wsc = Jb[9];
		if (idx == 0)
		{
			isGood_new[i] = 1;
			energy_new2[2 * i] = energy; energy_new2[2 * i + 1] = fmul(idm1, idm1);
			lastH_new[i] = lastH;
		}
	}
	else if (valid && idx == 0)
	{
		isGood_new[i] = 0;
		energy_new2[2 * i] = e_old0; energy_new2[2 * i + 1] = e_old1;
		lastH_new[i] = 0.f;
	}
	if (valid && idx == 0) maxstep_out[i] = mstep;
	// JbBuffer_new of the point (lane k stores entry k, lanes 0 and 1 also 8 and 9); also staged for the Schur products
#pragma unroll
	for (int k = 0; k < 10; k++)
		if ((k & 7) == idx && (k < 8 || idx < 2))
		{
			sJb[g][k] = Jb[k];
			if (valid) Jb10[10 * (size_t)i + k] = good ? Jb[k] : 0.f;
		}
	__syncwarp();

	// ---- acc9 (:587-597): this lane's pattern pixel; acc9SC (:670): entries idx, idx+8, ... of the point's outer product ----
	float a9[45];
	{
		int k = 0;
#pragma unroll
		for (int r = 0; r < 9; r++)
#pragma unroll
			for (int c = r; c < 9; c++) a9[k++] = good ? fmul(J[r], J[c]) : 0.f;
	}
	float asc[6];
#pragma unroll
	for (int m = 0; m < 6; m++)
	{
		const int e = idx + 8 * m;
		float val = 0.f;
		if (e < 45 && good) val = fmul(fmul(sJb[g][c_tri9[e][0]], wsc), sJb[g][c_tri9[e][1]]);
		asc[m] = val;
	}
	float esum = (idx == 0) ? e_pt : 0.f;

	// ---- fixed-order reduction: warp butterflies, warps in order, blocks in order ----
#pragma unroll
	for (int k = 0; k < 45; k++)
	{
		float s = a9[k];
#pragma unroll
		for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
		a9[k] = s;
	}
#pragma unroll
	for (int m = 0; m < 6; m++)
	{
		float s = asc[m];
		s += __shfl_xor_sync(0xffffffffu, s, 8);
		s += __shfl_xor_sync(0xffffffffu, s, 16);
		asc[m] = s;
	}
#pragma unroll
	for (int o = 16; o > 0; o >>= 1) esum += __shfl_xor_sync(0xffffffffu, esum, o);
	if (lane == 0)
	{
#pragma unroll
		for (int k = 0; k < 45; k++) s_w[warp][k] = a9[k];
		s_w[warp][90] = esum;
		for (int k = 91; k < INIT_NSUM; k++) s_w[warp][k] = 0.f;
	}
	if (lane < 8)
	{
#pragma unroll
		for (int m = 0; m < 6; m++) if (lane + 8 * m < 45) s_w[warp][45 + lane + 8 * m] = asc[m];
	}
	__syncthreads();
	if (tid < INIT_NSUM)
	{
		float s = 0.f;
#pragma unroll
		for (int w = 0; w < 8; w++) s += s_w[w][tid];
		partials[(size_t)blockIdx.x * INIT_NSUM + tid] = s;
	}
	__threadfence();
	__syncthreads();
	if (tid == 0)
	{
		const unsigned int t = atomicAdd(ticket, 1u);
		s_last = (t == gridDim.x - 1);
	}
	__syncthreads();
	if (s_last)
	{
		__threadfence();
		if (tid < INIT_NSUM)
		{
			float s = 0.f;
			for (unsigned b = 0; b < gridDim.x; b++) s += __ldcg(partials + (size_t)b * INIT_NSUM + tid);
			sums_out[tid] = s;
		}
		if (tid == 0) *ticket = 0;
	}
}

// ---- host side ------------------------------------------------------------------------------------
struct InitWS
{
	int cap;                    // points
	float* d_f;                 // u, v, idepth_new, energy2 (2), outlierTH, iR | energy_new2 (2), maxstep, lastH, Jb (10)  = 7 + 14 floats per point
	unsigned char* d_b;         // isGood, isGood_new
	float* d_partials; int cap_blocks;
	unsigned int* d_ticket; float* d_sums; float* h_sums;
};

void init_destroy(dso_b200_ctx* ctx)
{
	InitWS* ws = (InitWS*)ctx->init_ws;
	if (!ws) return;
	cudaFree(ws->d_f); cudaFree(ws->d_b); cudaFree(ws->d_partials); cudaFree(ws->d_ticket); cudaFree(ws->d_sums);
	if (ws->h_sums) cudaFreeHost(ws->h_sums);
	delete ws;
	ctx->init_ws = nullptr;
}

static int init_workspace(dso_b200_ctx* ctx, InitWS** out, int npts)
{
	InitWS* ws = (InitWS*)ctx->init_ws;
	if (!ws)
	{
		ws = new InitWS();
		memset(ws, 0, sizeof(*ws));
		ctx->init_ws = ws;
		CK(cudaMalloc(&ws->d_ticket, sizeof(unsigned int)));
		CK(cudaMemset(ws->d_ticket, 0, sizeof(unsigned int)));
		CK(cudaMalloc(&ws->d_sums, INIT_NSUM * sizeof(float)));
		CK(cudaHostAlloc((void**)&ws->h_sums, INIT_NSUM * sizeof(float), cudaHostAllocDefault));
	}
	if (npts > ws->cap)
	{
		CK(cudaStreamSynchronize(ctx->stream));
		cudaFree(ws->d_f); cudaFree(ws->d_b); cudaFree(ws->d_partials);
		ws->d_f = nullptr; ws->d_b = nullptr; ws->d_partials = nullptr; ws->cap = 0;
		const int cap = npts + npts / 2 + 1024;
		CK(cudaMalloc(&ws->d_f, (size_t)cap * 21 * sizeof(float)));
		CK(cudaMalloc(&ws->d_b, (size_t)cap * 2));
		ws->cap_blocks = (cap + INIT_PTS_PER_BLOCK - 1) / INIT_PTS_PER_BLOCK;
		CK(cudaMalloc(&ws->d_partials, (size_t)ws->cap_blocks * INIT_NSUM * sizeof(float)));
		ws->cap = cap;
	}
	*out = ws;
	return 0;
}

// Eigen's compute_inverse<Matrix3d> (cofactors of column 0, determinant by expansion along it): CoarseInitializer::makeK
// keeps K and K^-1 in double (CoarseInitializer.cpp:1146-1154)
static void mat33d_inverse_h(const double* m, double* out)
{
	auto cof = [&](int i, int j) { int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
	                               return dadd_r(dmul_r(m[i1 * 3 + j1], m[i2 * 3 + j2]), -dmul_r(m[i1 * 3 + j2], m[i2 * 3 + j1])); };
	const double c0 = cof(0, 0), c1 = cof(1, 0), c2 = cof(2, 0);
	const double det = dot3dd(c0, m[0], c1, m[3], c2, m[6]);
	const double invdet = 1.0 / det;
	out[0] = c0 * invdet; out[1] = c1 * invdet; out[2] = c2 * invdet;
	out[3] = cof(0, 1) * invdet; out[4] = cof(1, 1) * invdet; out[5] = cof(2, 1) * invdet;
	out[6] = cof(0, 2) * invdet; out[7] = cof(1, 2) * invdet; out[8] = cof(2, 2) * invdet;
}

extern "C" int dso_b200_init_calc_res_and_gs(dso_b200_ctx* ctx, int first_slot, int new_slot, int lvl,
                                             float fx, float fy, float cx, float cy,
                                             const double* refToNew12, double aff_a, double aff_b,
                                             int npts, const float* u, const float* v, const float* idepth_new, const unsigned char* isGood,
                                             const float* energy2, const float* outlierTH, const float* iR,
                                             float alphaW, float alphaK, float couplingWeight,
                                             float* H64, float* b8, float* Hsc64, float* bsc8, float* res3,
                                             unsigned char* isGood_new, float* energy_new2, float* maxstep, float* lastHessian_new,
                                             float* JbBuffer_new10)
{
	ARG_CHECK(ctx && refToNew12 && H64 && b8 && Hsc64 && bsc8 && res3, "null argument");
	ARG_CHECK(first_slot >= 0 && first_slot < ctx->n_slots && ctx->slots[first_slot].valid, "first frame slot empty");
	ARG_CHECK(new_slot >= 0 && new_slot < ctx->n_slots && ctx->slots[new_slot].valid, "new frame slot empty");
	const FrameSlot& F0 = ctx->slots[first_slot];
	const FrameSlot& F1 = ctx->slots[new_slot];
	ARG_CHECK(F0.w == F1.w && F0.h == F1.h, "frame size mismatch");
	ARG_CHECK(lvl >= 0 && lvl < F0.levels && lvl < F1.levels, "level not in the frames' pyramids");
	ARG_CHECK(npts >= 1 && u && v && idepth_new && isGood && energy2 && outlierTH && iR, "null point arrays");
	CK(cudaSetDevice(ctx->device));
	int rc;
	if ((rc = frame_slot_wait(ctx, first_slot, ctx->stream, DSO_B200_STREAM_MAIN))) return rc;
	if ((rc = frame_slot_wait(ctx, new_slot, ctx->stream, DSO_B200_STREAM_MAIN))) return rc;
	InitWS* ws;
	if ((rc = init_workspace(ctx, &ws, npts))) return rc;

	// CoarseInitializer::makeK for this level, in double (:1126-1155)
	double dfx = fx, dfy = fy, dcx = cx, dcy = cy;
	const double cx0 = dcx, cy0 = dcy;
	for (int level = 1; level <= lvl; level++)
	{
		dfx = dfx * 0.5; dfy = dfy * 0.5;
		dcx = (cx0 + 0.5) / ((int)1 << level) - 0.5;
		dcy = (cy0 + 0.5) / ((int)1 << level) - 0.5;
	}
	const double K[9] = { dfx, 0.0, dcx, 0.0, dfy, dcy, 0.0, 0.0, 1.0 };
	double Ki[9], RK[9];
	mat33d_inverse_h(K, Ki);
	Pose T; pose_from12(T, refToNew12);
	mat33d_mul(T.R, Ki, RK);
	InitParams P;
	for (int i = 0; i < 9; i++) P.RKi[i] = (float)RK[i];                      // :385
	for (int i = 0; i < 3; i++) P.t[i] = (float)T.t[i];                       // :387
	P.aff0 = (float)exp(aff_a); P.aff1 = (float)aff_b;                         // :390
	P.fxl = (float)dfx; P.fyl = (float)dfy; P.cxl = (float)dcx; P.cyl = (float)dcy;
	P.wl = F0.w >> lvl; P.hl = F0.h >> lvl; P.npts = npts;
	P.huberTH = ctx->settings.huberTH;
	// :635-649.  The reference feeds the regulariser energies into E instead of EAlpha (:617-633), so EAlpha.A is 0 and the
	// alpha energy depends on the translation only: it is known before the kernel runs
	const double tsq = dot3dd(T.t[0], T.t[0], T.t[1], T.t[1], T.t[2], T.t[2]);
	volatile float alphaEnergy = (float)((double)alphaW * (0.0 + tsq * npts));
	volatile float aK = alphaK * (float)npts;
	float alphaOpt;
	if (alphaEnergy > aK) { alphaOpt = 0; alphaEnergy = aK; }
	else alphaOpt = alphaW;
	P.alphaOpt = alphaOpt; P.couplingWeight = couplingWeight;

	cudaStream_t s = ctx->stream;
	const size_t n = (size_t)npts, c = (size_t)ws->cap;
	float* d_u = ws->d_f; float* d_v = d_u + c; float* d_id = d_v + c; float* d_e = d_id + c; float* d_oth = d_e + 2 * c; float* d_iR = d_oth + c;
	float* d_en = d_iR + c; float* d_ms = d_en + 2 * c; float* d_lh = d_ms + c; float* d_jb = d_lh + c;
	unsigned char* d_g = ws->d_b; unsigned char* d_gn = d_g + c;
	CK(cudaMemcpyAsync(d_u, u, n * 4, cudaMemcpyHostToDevice, s));
	CK(cudaMemcpyAsync(d_v, v, n * 4, cudaMemcpyHostToDevice, s));
	CK(cudaMemcpyAsync(d_id, idepth_new, n * 4, cudaMemcpyHostToDevice, s));
	CK(cudaMemcpyAsync(d_e, energy2, n * 8, cudaMemcpyHostToDevice, s));
	CK(cudaMemcpyAsync(d_oth, outlierTH, n * 4, cudaMemcpyHostToDevice, s));
	CK(cudaMemcpyAsync(d_iR, iR, n * 4, cudaMemcpyHostToDevice, s));
	CK(cudaMemcpyAsync(d_g, isGood, n, cudaMemcpyHostToDevice, s));
	const int grid = (npts + INIT_PTS_PER_BLOCK - 1) / INIT_PTS_PER_BLOCK;
	init_res_gs_kernel<<<grid, 256, 0, s>>>(P, F0.pyr[lvl], F1.pyr[lvl], d_u, d_v, d_id, d_g, d_e, d_oth, d_iR, d_gn, d_en, d_ms, d_lh, d_jb,
	                                        ws->d_partials, ws->d_ticket, ws->d_sums);
	LAUNCHED(ctx, 1);
	CK(cudaGetLastError());
	CK(cudaMemcpyAsync(ws->h_sums, ws->d_sums, INIT_NSUM * sizeof(float), cudaMemcpyDeviceToHost, s));
	if (isGood_new) CK(cudaMemcpyAsync(isGood_new, d_gn, n, cudaMemcpyDeviceToHost, s));
	if (energy_new2) CK(cudaMemcpyAsync(energy_new2, d_en, n * 8, cudaMemcpyDeviceToHost, s));
	if (maxstep) CK(cudaMemcpyAsync(maxstep, d_ms, n * 4, cudaMemcpyDeviceToHost, s));
	if (lastHessian_new) CK(cudaMemcpyAsync(lastHessian_new, d_lh, n * 4, cudaMemcpyDeviceToHost, s));
	if (JbBuffer_new10) CK(cudaMemcpyAsync(JbBuffer_new10, d_jb, n * 40, cudaMemcpyDeviceToHost, s));
	CK(cudaStreamSynchronize(s));

	// :680-694
	const float* S = ws->h_sums;
	float A9[9][9], ASC[9][9];
	int k = 0;
	for (int r = 0; r < 9; r++) for (int cc = r; cc < 9; cc++) { A9[r][cc] = A9[cc][r] = S[k]; ASC[r][cc] = ASC[cc][r] = S[45 + k]; k++; }
	for (int r = 0; r < 8; r++)
	{
		for (int cc = 0; cc < 8; cc++) { H64[r * 8 + cc] = A9[r][cc]; Hsc64[r * 8 + cc] = ASC[r][cc]; }
		b8[r] = A9[r][8]; bsc8[r] = ASC[r][8];
	}
	volatile float an = alphaOpt * (float)npts;
	H64[0] += an; H64[9] += an; H64[18] += an;
	double lg[6];
	pose_log(T, lg);
	for (int q = 0; q < 3; q++) { volatile float t0 = (float)lg[q] * alphaOpt; volatile float t1 = t0 * (float)npts; b8[q] += t1; }
	res3[0] = S[90]; res3[1] = alphaEnergy; res3[2] = (float)(2 * (size_t)npts);   // E.num counts every point twice (:617-633)
	return 0;
}
