// K3-K5 -- windowed photometric BA: linearisation fused with the top-Hessian and Schur-complement
// accumulation (reference: FullSystem/Residuals.cpp:78-347, OptimizationBackend/AccumulatedTopHessian.{h,cpp},
// AccumulatedSCHessian.{h,cpp}, EnergyFunctional.cpp:209-274, EnergyFunctionalStructs.cpp:39-49).
//
//   ba_linearize_kernel     PointFrameResidual::linearize: 8 lanes per residual (one per pattern pixel), the
//                           8-term sums taken in index order through shuffles, so energies, state decisions
//                           and the RawResidualJacobian are bit-identical to the reference's scalar loop.
//   ba_apply_res_kernel     applyRes(true) + EFResidual::takeDataF (the J "swap" is a per-residual buffer flip).
//   ba_accum_top_kernel     addPoint<0>: residuals pre-sorted by (host,target); one CTA per chunk of 64
//                           residuals of one pair stages them in shared memory, then thread e owns entry e of the
//                           13x13 block and sums it over the chunk in order -- no atomics, no cross-thread reduce.
//   ba_point_sums_kernel    per-point Hdd / bd / Hcd in residualsAll order.
//   ba_reduce_partials / ba_top_pair / ba_top_gather
//                           stitchDoubleInternal + symmetrisation: chunk partials summed in index order in FP64, one CTA
//                           per (host,target) pair does its adjoint sandwiches in FP64, every output block then gathers
//                           its contributions in the reference's pair order.
//   ba_schur_*              AccumulatedSCHessianSSE::addPoint as per-host "sum_p w_p a_p a_p^T" with thread-owned
//                           output entries over chunks of 16 points, then two FP64 stitch stages that write every
//                           contribution to scratch and gather them per output block in a fixed order.
// All of it is FP32 gather/outer-product work bound by L2 latency at the window sizes of BASELINE configs[1]
// (12 k residuals x ~850 B), HBM-bound only for the 100 k-point sweep.
#include "common.cuh"
#include "lm_math.cuh"
#include <algorithm>

#include "ba_internal.cuh"

__constant__ int c_pattern[8][2] = { {0,-2}, {-1,-1}, {1,-1}, {-2,0}, {0,0}, {2,0}, {-1,1}, {0,2} };   // settings.cpp:274


void ba_settings_changed(dso_b200_ctx* ctx);
int ba_create(dso_b200_ctx* ctx)
{
	BAWorkspace* W = new BAWorkspace();
	memset(&W->P, 0, sizeof(W->P));
	W->window_set = W->glue_set = W->points_set = W->res_set = W->lin_done = W->top_done = false;
	W->capP = W->capR = W->capChunks = W->capPC = 0;
	W->nChunks = W->nPC = 0;
	W->contrib_from_lin = false;
	const int N = 8 * MAXF + CPARS;
	CK(cudaMalloc(&W->d_precalc, MAXF * MAXF * 28 * sizeof(float)));
	CK(cudaMalloc(&W->d_adHost, MAXF * MAXF * 64 * sizeof(double)));
	CK(cudaMalloc(&W->d_adTarget, MAXF * MAXF * 64 * sizeof(double)));
	CK(cudaMalloc(&W->d_pair_chunk_begin, (MAXF * MAXF + 1) * sizeof(int)));
	CK(cudaMalloc(&W->d_host_pc_begin, (MAXF + 1) * sizeof(int)));
	// H_A | b_A | H_sc | b_sc | scalars in one block (fixed strides for the largest window): the step ends with one copy
	CK(cudaMalloc(&W->d_outblk, BA_OUT_DOUBLES * sizeof(double)));
	CK(cudaMemset(W->d_outblk, 0, BA_OUT_DOUBLES * sizeof(double)));
	W->d_HA = W->d_outblk; W->d_bA = W->d_HA + (size_t)N * N; W->d_HS = W->d_bA + N; W->d_bS = W->d_HS + (size_t)N * N; W->d_scalars = W->d_bS + N;
	// stage-1 contributions: per (i,j): C1,C2 (64 each), C3,C4 (MAXF*64 each), CEh,CEt (32 each), Cbh,Cbt (8 each)
	CK(cudaMalloc(&W->d_scContrib, (size_t)MAXF * MAXF * (128 + 2 * MAXF * 64 + 64 + 16) * sizeof(double)));
	CK(cudaMalloc(&W->d_scRed, (size_t)MAXF * (MAXF * MAXF * 64 + MAXF * 40 + 24) * sizeof(double)));
	CK(cudaMalloc(&W->d_topRed, (size_t)MAXF * MAXF * NTOP_PAD * sizeof(double)));
	CK(cudaMalloc(&W->d_topContrib, (size_t)MAXF * MAXF * 292 * sizeof(double)));
	CK(cudaHostAlloc((void**)&W->h_out, BA_OUT_DOUBLES * sizeof(double), cudaHostAllocDefault));
	W->d_energy_blocks = nullptr;
	CK(cudaMalloc(&W->d_adHTdeltaF, MAXF * MAXF * 8 * sizeof(float)));
	CK(cudaMalloc(&W->d_cDeltaF, 4 * sizeof(float)));
	CK(cudaMemset(W->d_adHTdeltaF, 0, MAXF * MAXF * 8 * sizeof(float)));
	CK(cudaMemset(W->d_cDeltaF, 0, 4 * sizeof(float)));
	for (int i = 0; i < 4; i++) W->h_cDeltaF[i] = 0.f;
	CK(cudaMalloc(&W->d_HL, (size_t)N * N * sizeof(double)));
	CK(cudaMalloc(&W->d_bL, N * sizeof(double)));
	CK(cudaMalloc(&W->d_priorH, N * sizeof(double)));
	CK(cudaMalloc(&W->d_priorD, N * sizeof(double)));
	CK(cudaMalloc(&W->d_delta, N * sizeof(double)));
	CK(cudaMemset(W->d_priorH, 0, N * sizeof(double)));
	CK(cudaMemset(W->d_priorD, 0, N * sizeof(double)));
	CK(cudaMemset(W->d_delta, 0, N * sizeof(double)));
	W->nLin = 0; W->S = nullptr;
	memset(&W->g_step, 0, sizeof(W->g_step)); memset(W->g_iter, 0, sizeof(W->g_iter));
	W->use_graphs = !getenv("DSO_B200_NO_GRAPH");
	CK(cudaStreamCreateWithFlags(&W->side_stream, cudaStreamNonBlocking));
	CK(cudaEventCreateWithFlags(&W->ev_fork, cudaEventDisableTiming));
	CK(cudaEventCreateWithFlags(&W->ev_join, cudaEventDisableTiming));
	W->side_busy = false;
	ctx->ba = W;
	ba_settings_changed(ctx);
	return 0;
}

// dso_b200_set_settings: the two thresholds PointFrameResidual::linearize reads travel in the kernel parameter block
void ba_settings_changed(dso_b200_ctx* ctx)
{
	BAWorkspace* W = ctx->ba;
	if (!W) return;
	W->P.huberTH = ctx->settings.huberTH;
	W->P.outlierTHSum = ctx->settings.outlierTHSumComponent;
	W->lin_done = false; W->top_done = false;
	ba_graphs_invalidate(W);
}

static void free_points(BAWorkspace* W)
{
	if (!W->capP) return;
	cudaFree(W->d_pt); cudaFree(W->d_color); cudaFree(W->d_weights); cudaFree(W->d_phost); cudaFree(W->d_priorF); cudaFree(W->d_deltaF);
	cudaFree(W->d_res_begin); cudaFree(W->d_Hdd); cudaFree(W->d_bd); cudaFree(W->d_Hcd); cudaFree(W->d_HdiF); cudaFree(W->d_bdSumF); cudaFree(W->d_idh);
	cudaFree(W->d_Hdd_L); cudaFree(W->d_bd_L); cudaFree(W->d_Hcd_L); cudaFree(W->d_HcdSum); cudaFree(W->d_HdiF_m); cudaFree(W->d_bdSumF_m);
	if (W->capPC) { cudaFree(W->d_pc_start); cudaFree(W->d_pc_cnt); cudaFree(W->d_pc_host); cudaFree(W->d_scPartials); }
	W->capP = 0; W->capPC = 0;
}
static void free_res(BAWorkspace* W)
{
	if (!W->capR) return;
	cudaFree(W->d_rhost); cudaFree(W->d_rpoint); cudaFree(W->d_rtarget); cudaFree(W->d_state); cudaFree(W->d_newState); cudaFree(W->d_energy); cudaFree(W->d_newEnergy);
	cudaFree(W->d_newEnergyWO); cudaFree(W->d_isLin); cudaFree(W->d_isAct); cudaFree(W->d_efsel); cudaFree(W->d_cpt); cudaFree(W->d_proj);
	cudaFree(W->d_J[0]); cudaFree(W->d_J[1]); cudaFree(W->d_JpJdF); cudaFree(W->d_contrib); cudaFree(W->d_resEnergy); cudaFree(W->d_perm); cudaFree(W->d_subset);
	cudaFree(W->d_energy_blocks); cudaFree(W->d_resToZero);
	if (W->capChunks) { cudaFree(W->d_chunk_start); cudaFree(W->d_chunk_cnt); cudaFree(W->d_topPartials); cudaFree(W->d_chunk_nres); }
	W->capR = 0; W->capChunks = 0;
}

void ba_graphs_invalidate(BAWorkspace* W)
{
	if (!W) return;
	BAGraph* gs[3] = { &W->g_step, &W->g_iter[0], &W->g_iter[1] };
	for (BAGraph* g : gs)
	{
		if (g->exec) cudaGraphExecDestroy(g->exec);
		g->exec = nullptr; g->valid = false;
	}
}

void ba_destroy(dso_b200_ctx* ctx)
{
	BAWorkspace* W = ctx->ba;
	if (!W) return;
	ba_solve_destroy(W);
	ba_graphs_invalidate(W);
	free_points(W); free_res(W);
	cudaFree(W->d_adHTdeltaF); cudaFree(W->d_cDeltaF); cudaFree(W->d_HL); cudaFree(W->d_bL); cudaFree(W->d_priorH); cudaFree(W->d_priorD); cudaFree(W->d_delta);
	cudaFree(W->d_precalc); cudaFree(W->d_adHost); cudaFree(W->d_adTarget); cudaFree(W->d_pair_chunk_begin); cudaFree(W->d_host_pc_begin);
	cudaFree(W->d_outblk); cudaFree(W->d_scContrib); cudaFree(W->d_scRed); cudaFree(W->d_topRed); cudaFree(W->d_topContrib);
	cudaFreeHost(W->h_out);
	if (W->side_stream) { cudaStreamSynchronize(W->side_stream); cudaStreamDestroy(W->side_stream); }
	if (W->ev_fork) cudaEventDestroy(W->ev_fork);
	if (W->ev_join) cudaEventDestroy(W->ev_join);
	delete W;
	ctx->ba = nullptr;
}

// ---------------------------------------------------------------------------------------------------------
// getInterpolatedElement33 (util/globalFuncs.h:75-89) on float4 texels
__device__ __forceinline__ void interp33(const float4* __restrict__ img, float x, float y, int width, float& h0, float& h1, float& h2)
{
	const int ix = (int)x, iy = (int)y;
	const float dx = fsub(x, (float)ix), dy = fsub(y, (float)iy);
	const float dxdy = fmul(dx, dy);
	const float w11 = dxdy, w01 = fsub(dy, dxdy), w10 = fsub(dx, dxdy), w00 = fadd(fsub(fsub(1.f, dx), dy), dxdy);
	const float4* bp = img + ix + iy * width;
	const float4 p00 = __ldg(bp), p10 = __ldg(bp + 1), p01 = __ldg(bp + width), p11 = __ldg(bp + 1 + width);
	h0 = fadd(fadd(fadd(fmul(w11, p11.x), fmul(w01, p01.x)), fmul(w10, p10.x)), fmul(w00, p00.x));
	h1 = fadd(fadd(fadd(fmul(w11, p11.y), fmul(w01, p01.y)), fmul(w10, p10.y)), fmul(w00, p00.y));
	h2 = fadd(fadd(fadd(fmul(w11, p11.z), fmul(w01, p01.z)), fmul(w10, p10.z)), fmul(w00, p00.z));
}

// sum of the 8 lanes of a residual group in index order 0..7 (the reference's scalar "+=" loop)
__device__ __forceinline__ float seq8(float v, int base)
{
	float s = __shfl_sync(0xffffffffu, v, base);
#pragma unroll
	for (int k = 1; k < 8; k++) s = fadd(s, __shfl_sync(0xffffffffu, v, base + k));
	return s;
}

// PointFrameResidual::linearize, Residuals.cpp:78-291.  APPLY = true fuses applyRes(true) + EFResidual::takeDataF
// (Residuals.cpp:323-347, EnergyFunctionalStructs.cpp:39-49) for the same residual into the tail of the kernel (the window
// step always applies what it has just linearised; saves a launch and the re-read of the Jacobian).
template <bool APPLY>
__global__ void __launch_bounds__(256) ba_linearize_kernel(BAParams B, const float* __restrict__ precalc, const float4* __restrict__ pt,
                                                           const float* __restrict__ color, const float* __restrict__ weights,
                                                           const int* __restrict__ phost, const int* __restrict__ rpoint,
                                                           const int* __restrict__ rtarget, int* state,
                                                           double* state_energy, unsigned char* efsel,
                                                           float* J0, float* J1, int* newState, float* newEnergy, float* newEnergyWO,
                                                           float* cpt, float* proj, double* resEnergy,
                                                           const int* __restrict__ subset, int nsub, int reset_oob,
                                                           unsigned char* isAct, float* JpJdF, float* contrib = nullptr,
                                                           const int* __restrict__ rhost = nullptr)
{
	__shared__ float sJ[32][JSZ + 2];
	const int tid = threadIdx.x, lane = tid & 31, g = tid >> 3, idx = tid & 7, base = lane & ~7;
	// subset != NULL: the per-object form (a batch of k residuals, e.g. FullSystem::flagPointsForRemoval, FullSystem.cpp:781);
	// reset_oob: PointFrameResidual::resetOOB first (Residuals.h:101-108: energies 0, state IN)
	const int slot = blockIdx.x * 32 + g;
	const bool valid = slot < (subset ? nsub : B.nR);
	const int r = valid ? (subset ? subset[slot] : slot) : 0;
	const int rr = r;
	const int p = rpoint[rr], t = rtarget[rr], h = rhost ? rhost[rr] : phost[p];
	const int st = reset_oob ? ST_IN : state[rr];
	const float4 P4 = __ldg(pt + p);            // u, v, idepth_scaled, idepth_zero_scaled
	const float* pre = precalc + (size_t)(h * B.nF + t) * 28;
	const float* R0 = pre;        // PRE_RTll_0
	const float* t0 = pre + 9;    // PRE_tTll_0
	const float* KRKi = pre + 12; // PRE_KRKiTll
	const float* Kt = pre + 21;   // PRE_KtTll
	const float affLL0 = pre[24], affLL1 = pre[25], b0 = pre[26];

	// ---- centre projection at the FEJ point, projectPoint (ResidualProjections.h:69-101), dx = dy = 0 ----
	const float Kl0 = fmul(fsub(fadd(P4.x, 0.f), B.cx), B.fxi);
	const float Kl1 = fmul(fsub(fadd(P4.y, 0.f), B.cy), B.fyi);
	float ptp[3];
#pragma unroll
	for (int k = 0; k < 3; k++) ptp[k] = fadd(dot3e(R0[k * 3], Kl0, R0[k * 3 + 1], Kl1, R0[k * 3 + 2], 1.f), fmul(t0[k], P4.w));
	const float drescale = fdiv(1.0f, ptp[2]);
	const float new_idepth = fmul(P4.w, drescale);
	const float u = fmul(ptp[0], drescale), v = fmul(ptp[1], drescale);
	const float Kuc = fadd(fmul(u, B.fx), B.cx), Kvc = fadd(fmul(v, B.fy), B.cy);
	const bool centre_ok = (drescale > 0.f) && (Kuc > 1.1f && Kvc > 1.1f && Kuc < B.wM3G && Kvc < B.hM3G);

	// ---- this lane's pattern pixel, projectPoint (ResidualProjections.h:55-65) ----
	const float up = fadd(P4.x, (float)c_pattern[idx][0]), vp = fadd(P4.y, (float)c_pattern[idx][1]);
	float q[3];
#pragma unroll
	for (int k = 0; k < 3; k++) q[k] = fadd(dot3e(KRKi[k * 3], up, KRKi[k * 3 + 1], vp, KRKi[k * 3 + 2], 1.f), fmul(Kt[k], P4.z));
	const float Ku = fdiv(q[0], q[2]), Kv = fdiv(q[1], q[2]);
	bool bad = !(Ku > 1.1f && Kv > 1.1f && Ku < B.wM3G && Kv < B.hM3G);
	float h0 = 0.f, h1 = 0.f, h2 = 0.f;
	const bool sample = valid && st != ST_OOB && centre_ok && !bad;
	if (sample)
	{
		interp33(B.img[t], Ku, Kv, B.w, h0, h1, h2);
		if (!isfinite(h0)) bad = true;
	}
	const unsigned badmask = (__ballot_sync(0xffffffffu, bad) >> base) & 0xffu;
	const bool oob = (st == ST_OOB) || !centre_ok || badmask != 0;

	const float col = __ldg(color + 8 * (size_t)p + idx), wgt = __ldg(weights + 8 * (size_t)p + idx);
	const float residual = fsub(h0, fadd(fmul(affLL0, col), affLL1));
	const float drdA = fsub(col, b0);
	float wv = __fsqrt_rn(fdiv(B.outlierTHSum, fadd(B.outlierTHSum, fadd(fmul(h1, h1), fmul(h2, h2)))));
	wv = fmul(0.5f, fadd(wv, wgt));
	const float ar = fabsf(residual);
	float hw = ar < B.huberTH ? 1.f : fdiv(B.huberTH, ar);
	const float e_i = fmul(fmul(fmul(fmul(fmul(wv, wv), hw), residual), residual), fsub(2.f, hw));   // :219
	if (hw < 1.f) hw = __fsqrt_rn(hw);
	hw = fmul(hw, wv);
	const float g1 = fmul(h1, hw), g2 = fmul(h2, hw);
	const float resF = fmul(residual, hw);
	const float ja = fmul(drdA, hw);

	const float energy = seq8(e_i, base);
	const float JI00 = seq8(fmul(g1, g1), base), JI11 = seq8(fmul(g2, g2), base), JI10 = seq8(fmul(g1, g2), base);
	const float JA00 = seq8(fmul(ja, g1), base), JA01 = seq8(fmul(ja, g2), base), JA10 = seq8(fmul(hw, g1), base), JA11 = seq8(fmul(hw, g2), base);
	const float AA00 = seq8(fmul(fmul(fmul(drdA, drdA), hw), hw), base), AA01 = seq8(fmul(fmul(drdA, hw), hw), base), AA11 = seq8(fmul(hw, hw), base);
	const float wJI2 = seq8(fmul(fmul(hw, hw), fadd(fmul(g1, g1), fmul(g2, g2))), base);
	// APPLY: JI_r = sum_i resF_i * JIdx_i of AccumulatedTopHessianSSE::addPoint<0> (AccumulatedTopHessian.cpp:107-118: a scalar
	// "+=" loop that starts from 0.f, hence the 0 + p of lane 0), so that the per-point terms of this residual (:142-150) can
	// leave with the Jacobian and the per-point sums need not wait for the top accumulation
	float JIr0 = 0.f, JIr1 = 0.f;
	if (APPLY)
	{
		const float p0 = fmul(resF, g1), p1 = fmul(resF, g2);
		JIr0 = seq8(idx == 0 ? fadd(0.f, p0) : p0, base);
		JIr1 = seq8(idx == 0 ? fadd(0.f, p1) : p1, base);
	}

	bool outlier_f = false;
	float eLeft_f = 0.f;
	if (valid)
	{
		if (oob)
		{
			// :82-83, :119-121, :197-198, :209-210: state OOB, energy of the previous state, nothing else touched
			if (idx == 0)
			{
				newState[r] = ST_OOB;
				newEnergyWO[r] = -1.f;
				resEnergy[r] = reset_oob ? 0.0 : state_energy[r];
				if (st != ST_OOB && centre_ok) { cpt[3 * r] = Kuc; cpt[3 * r + 1] = Kvc; cpt[3 * r + 2] = new_idepth; }   // :124 precedes the pattern loop
			}
		}
		else
		{
			const float th = fmaxf(B.frameEnergyTH[h], B.frameEnergyTH[t]);
			const bool outlier = (energy > th) || (wJI2 < 2.f);                 // :278
			const float eLeft = outlier ? th : energy;
			outlier_f = outlier; eLeft_f = eLeft;
			proj[16 * (size_t)r + 2 * idx] = Ku; proj[16 * (size_t)r + 2 * idx + 1] = Kv;
			float* s = sJ[g];
			s[J_RESF + idx] = resF; s[J_IDX0 + idx] = g1; s[J_IDX1 + idx] = g2; s[J_AB0 + idx] = B.fixA ? 0.f : ja; s[J_AB1 + idx] = B.fixB ? 0.f : hw;   // :257-258
			if (idx == 0)
			{
				newState[r] = outlier ? ST_OUTLIER : ST_IN;
				newEnergy[r] = eLeft; newEnergyWO[r] = energy; resEnergy[r] = (double)eLeft;
				cpt[3 * r] = Kuc; cpt[3 * r + 1] = Kvc; cpt[3 * r + 2] = new_idepth;
				// d_xi_x, :159-164
				s[J_PDXI0 + 0] = fmul(new_idepth, B.fx); s[J_PDXI0 + 1] = 0.f; s[J_PDXI0 + 2] = fmul(fmul(-new_idepth, u), B.fx);
				s[J_PDXI0 + 3] = fmul(fmul(-u, v), B.fx); s[J_PDXI0 + 4] = fmul(fadd(1.f, fmul(u, u)), B.fx); s[J_PDXI0 + 5] = fmul(-v, B.fx);
			}
			else if (idx == 1)
			{
				// d_xi_y, :166-171
				s[J_PDXI1 + 0] = 0.f; s[J_PDXI1 + 1] = fmul(new_idepth, B.fy); s[J_PDXI1 + 2] = fmul(fmul(-new_idepth, v), B.fy);
				s[J_PDXI1 + 3] = fmul(-fadd(1.f, fmul(v, v)), B.fy); s[J_PDXI1 + 4] = fmul(fmul(u, v), B.fy); s[J_PDXI1 + 5] = fmul(u, B.fy);
			}
			else if (idx == 2)
			{
				// d_C_x, :138-151
				float c2 = fmul(drescale, fsub(fmul(R0[6], u), R0[0]));
				float c3 = fmul(fmul(fmul(B.fx, drescale), fsub(fmul(R0[7], u), R0[1])), B.fyi);
				float c0 = fmul(Kl0, c2), c1 = fmul(Kl1, c3);
				s[J_PDC0 + 0] = fmul(fadd(c0, u), SCALE_F); s[J_PDC0 + 1] = fmul(c1, SCALE_F);
				s[J_PDC0 + 2] = fmul(fadd(c2, 1.f), SCALE_C); s[J_PDC0 + 3] = fmul(c3, SCALE_C);
			}
			else if (idx == 3)
			{
				// d_C_y, :143-156
				float c2 = fmul(fmul(fmul(B.fy, drescale), fsub(fmul(R0[6], v), R0[3])), B.fxi);
				float c3 = fmul(drescale, fsub(fmul(R0[7], v), R0[4]));
				float c0 = fmul(Kl0, c2), c1 = fmul(Kl1, c3);
				s[J_PDC1 + 0] = fmul(c0, SCALE_F); s[J_PDC1 + 1] = fmul(fadd(c1, v), SCALE_F);
				s[J_PDC1 + 2] = fmul(c2, SCALE_C); s[J_PDC1 + 3] = fmul(fadd(c3, 1.f), SCALE_C);
			}
			else if (idx == 4)
			{
				// d_d, :131-132
				s[J_PDD + 0] = fmul(fmul(fmul(drescale, fsub(t0[0], fmul(t0[2], u))), SCALE_IDEPTH), B.fx);
				s[J_PDD + 1] = fmul(fmul(fmul(drescale, fsub(t0[1], fmul(t0[2], v))), SCALE_IDEPTH), B.fy);
				s[J_IDX2 + 0] = JI00; s[J_IDX2 + 1] = JI10; s[J_IDX2 + 2] = JI10; s[J_IDX2 + 3] = JI11;
			}
			else if (idx == 5) { s[J_ABIDX + 0] = JA00; s[J_ABIDX + 1] = JA01; s[J_ABIDX + 2] = JA10; s[J_ABIDX + 3] = JA11; }
			else if (idx == 6) { s[J_AB2 + 0] = AA00; s[J_AB2 + 1] = AA01; s[J_AB2 + 2] = AA01; s[J_AB2 + 3] = AA11; }
		}
	}
	__syncwarp();
	if (valid && !oob)
	{
		// PointFrameResidual::J is the buffer EFResidual is not holding
		float* dst = (efsel[r] ? J0 : J1) + (size_t)r * JSZ;
		for (int k = idx; k < JSZ; k += 8) dst[k] = sJ[g][k];
	}
	if (APPLY && valid && st != ST_OOB)                   // applyRes: can never go back from OOB
	{
		const bool in = !oob && !outlier_f;
		if (in)
		{
			// takeDataF: JpJdF from the Jacobian just written (std::swap(J, data->J) = flipping efsel below)
			const float* Jl = sJ[g];
			const float d0 = Jl[J_PDD], d1 = Jl[J_PDD + 1];
			const float a0 = fadd(fmul(Jl[J_IDX2 + 0], d0), fmul(Jl[J_IDX2 + 1], d1));
			const float a1 = fadd(fmul(Jl[J_IDX2 + 2], d0), fmul(Jl[J_IDX2 + 3], d1));
			float val;
			if (idx < 6) val = fadd(fmul(Jl[J_PDXI0 + idx], a0), fmul(Jl[J_PDXI1 + idx], a1));
			else val = fadd(fmul(Jl[J_ABIDX + 2 * (idx - 6)], d0), fmul(Jl[J_ABIDX + 2 * (idx - 6) + 1], d1));
			JpJdF[8 * (size_t)r + idx] = val;
			if (contrib && idx < 6)
			{
				float c;
				if (idx == 0) c = fadd(fmul(JIr0, d0), fmul(JIr1, d1));
				else if (idx == 1) c = fadd(fmul(a0, d0), fmul(a1, d1));
				else c = fadd(fmul(Jl[J_PDC0 + idx - 2], a0), fmul(Jl[J_PDC1 + idx - 2], a1));
				contrib[6 * (size_t)r + idx] = c;
			}
		}
		if (idx == 0)
		{
			isAct[r] = in ? 1 : 0;
			if (in) efsel[r] = efsel[r] ^ 1;
			state[r] = oob ? ST_OOB : (outlier_f ? ST_OUTLIER : ST_IN);
			state_energy[r] = oob ? (double)newEnergy[r] : (double)eLeft_f;      // the OOB exits leave state_NewEnergy as it was
		}
	}
}

// applyRes(true), Residuals.cpp:323-347 + EFResidual::takeDataF, EnergyFunctionalStructs.cpp:39-49
__global__ void ba_apply_res_kernel(int nR, int* state, const int* __restrict__ newState, double* state_energy, const float* __restrict__ newEnergy,
                                    unsigned char* isAct, unsigned char* efsel, const float* J0, const float* J1, float* JpJdF,
                                    const int* __restrict__ subset, int reset_oob)
{
	const int slot = blockIdx.x * blockDim.x + threadIdx.x;
	if (slot >= nR) return;
	const int r = subset ? subset[slot] : slot;
	if (reset_oob) { state[r] = ST_IN; state_energy[r] = 0.0; }   // PointFrameResidual::resetOOB preceded the linearisation
	if (state[r] == ST_OOB) return;                       // can never go back from OOB
	const int ns = newState[r];
	if (ns == ST_IN)
	{
		isAct[r] = 1;
		const unsigned char sel = efsel[r] ^ 1;            // std::swap(J, data->J)
		efsel[r] = sel;
		const float* J = (sel ? J1 : J0) + (size_t)r * JSZ;
		const float d0 = J[J_PDD], d1 = J[J_PDD + 1];
		const float a0 = fadd(fmul(J[J_IDX2 + 0], d0), fmul(J[J_IDX2 + 1], d1));
		const float a1 = fadd(fmul(J[J_IDX2 + 2], d0), fmul(J[J_IDX2 + 3], d1));
		float* o = JpJdF + 8 * (size_t)r;
#pragma unroll
		for (int k = 0; k < 6; k++) o[k] = fadd(fmul(J[J_PDXI0 + k], a0), fmul(J[J_PDXI1 + k], a1));
		o[6] = fadd(fmul(J[J_ABIDX + 0], d0), fmul(J[J_ABIDX + 1], d1));
		o[7] = fadd(fmul(J[J_ABIDX + 2], d0), fmul(J[J_ABIDX + 3], d1));
	}
	else
		isAct[r] = 0;
	state[r] = ns;
	state_energy[r] = (double)newEnergy[r];
}

// deterministic double sum of per-residual energies: fixed tree per block, blocks summed in order by one thread
__global__ void ba_energy_blocks_kernel(int n, const double* __restrict__ e, double* blocks)
{
	__shared__ double s[256];
	const int i = blockIdx.x * 256 + threadIdx.x;
	s[threadIdx.x] = i < n ? e[i] : 0.0;
	__syncthreads();
	for (int o = 128; o > 0; o >>= 1) { if ((int)threadIdx.x < o) s[threadIdx.x] += s[threadIdx.x + o]; __syncthreads(); }
	if (threadIdx.x == 0) blocks[blockIdx.x] = s[0];
}
// both in one launch: the last block to finish adds the block sums in block order and re-arms the counter
__global__ void __launch_bounds__(256) ba_energy_kernel(int n, const double* __restrict__ e, double* blocks, unsigned int* counter, double* out)
{
	__shared__ double s[256];
	__shared__ bool last;
	const int i = blockIdx.x * 256 + threadIdx.x;
	s[threadIdx.x] = i < n ? e[i] : 0.0;
	__syncthreads();
	for (int o = 128; o > 0; o >>= 1) { if ((int)threadIdx.x < o) s[threadIdx.x] += s[threadIdx.x + o]; __syncthreads(); }
	if (threadIdx.x == 0)
	{
		blocks[blockIdx.x] = s[0];
		__threadfence();
		last = atomicAdd(counter, 1u) == gridDim.x - 1;
	}
	__syncthreads();
	if (last && threadIdx.x == 0)
	{
		__threadfence();
		const int nb = gridDim.x;
		double t = 0;
		for (int i0 = 0; i0 < nb; i0 += 16)                   // sixteen loads in flight, adds in block order
		{
			double v[16];
#pragma unroll
			for (int u = 0; u < 16; u++) v[u] = (i0 + u < nb) ? __ldcg(blocks + i0 + u) : 0.0;
#pragma unroll
			for (int u = 0; u < 16; u++) if (i0 + u < nb) t += v[u];
		}
		out[0] = t;
		*counter = 0u;
	}
}
__global__ void ba_energy_final_kernel(int nb, const double* __restrict__ blocks, double* out)
{
	if (threadIdx.x == 0 && blockIdx.x == 0)
	{
		double s = 0;
		for (int i0 = 0; i0 < nb; i0 += 16)                   // sixteen loads in flight, adds in block order
		{
			double t[16];
#pragma unroll
			for (int u = 0; u < 16; u++) t[u] = (i0 + u < nb) ? blocks[i0 + u] : 0.0;
#pragma unroll
			for (int u = 0; u < 16; u++) if (i0 + u < nb) s += t[u];
		}
		out[0] = s;
	}
}

// Jp*delta of EFResidual::fixLinearizationF / addPoint<1> (EnergyFunctionalStructs.cpp:92-100, AccumulatedTopHessian.cpp:82-87):
// d[0..1] = Jpdxi[k].dot(dp.head<6>()) + Jpdc[k].dot(dc) + Jpdd[k]*dd (fixed-size reductions in Eigen's unrolled order:
// 6 terms (p0+(p1+p2))+(p3+(p4+p5)), 4 terms (p0+p1)+(p2+p3)), d[2..3] = dp[6], dp[7]
__device__ __forceinline__ void ba_jp_delta(const float* __restrict__ J, const float* __restrict__ dp, const float* __restrict__ dc, float dd, float* d)
{
#pragma unroll
	for (int k = 0; k < 2; k++)
	{
		const float* xi = J + (k ? J_PDXI1 : J_PDXI0);
		const float* c = J + (k ? J_PDC1 : J_PDC0);
		const float s6 = fadd(fadd(fmul(xi[0], dp[0]), fadd(fmul(xi[1], dp[1]), fmul(xi[2], dp[2]))),
		                      fadd(fmul(xi[3], dp[3]), fadd(fmul(xi[4], dp[4]), fmul(xi[5], dp[5]))));
		const float s4 = fadd(fadd(fmul(c[0], dc[0]), fmul(c[1], dc[1])), fadd(fmul(c[2], dc[2]), fmul(c[3], dc[3])));
		d[k] = fadd(fadd(s6, s4), fmul(J[J_PDD + k], dd));
	}
	d[2] = dp[6]; d[3] = dp[7];
}

// AccumulatedTopHessianSSE::addPoint<MODE>, AccumulatedTopHessian.cpp:39-160, one chunk of one (host,target) pair.
// MODE 0: active, not linearised, r = resF.  MODE 1: active and linearised, r = res_toZeroF + J*delta (:73-101).
template <int MODE>
__global__ void __launch_bounds__(128) ba_accum_top_kernel(const int* __restrict__ perm, const int* __restrict__ chunk_start,
                                                           const int* __restrict__ chunk_cnt, const unsigned char* __restrict__ isAct,
                                                           const unsigned char* __restrict__ isLin, const unsigned char* __restrict__ efsel,
                                                           const float* __restrict__ J0, const float* __restrict__ J1, float* contrib,
                                                           float* partials, int* chunk_nres, BALinArgs LA)
{
	// contrib == NULL: the fused linearisation has already written the per-point terms of every active residual
	__shared__ float sx[CHUNK_R][10], sy[CHUNK_R][10], sw[CHUNK_R][16];   // sw: a,b,c | TR00,TR10,TR01,TR11,TR02,TR12 | BR0..5
	__shared__ int s_n;
	const int chunk = blockIdx.x, tid = threadIdx.x;
	const int start = chunk_start[chunk], cnt = chunk_cnt[chunk];
	if (tid == 0) s_n = 0;
	__syncthreads();
	if (tid < cnt)
	{
		const int r = perm[start + tid];
		const bool act = isAct[r] && (MODE == 0 ? !isLin[r] : (MODE == 1 ? isLin[r] != 0 : LA.pmask[LA.rpoint[r]] != 0));   // :55-67
		float x[10], y[10], w[15];
		if (act)
		{
			const float* J = (efsel[r] ? J1 : J0) + (size_t)r * JSZ;
#pragma unroll
			for (int k = 0; k < 4; k++) { x[k] = J[J_PDC0 + k]; y[k] = J[J_PDC1 + k]; }
#pragma unroll
			for (int k = 0; k < 6; k++) { x[4 + k] = J[J_PDXI0 + k]; y[4 + k] = J[J_PDXI1 + k]; }
			float res[8];
			if (MODE == 0)
			{
#pragma unroll
				for (int i = 0; i < 8; i++) res[i] = J[J_RESF + i];
			}
			else if (MODE == 2)
			{
				const float* tz = LA.resToZero + 8 * (size_t)r;           // marginalise: resApprox = res_toZeroF (:71-72)
#pragma unroll
				for (int i = 0; i < 8; i++) res[i] = tz[i];
			}
			else
			{
				const int p = LA.rpoint[r];
				float d[4];
				ba_jp_delta(J, LA.adHTdeltaF + 8 * (LA.phost[p] + LA.nF * LA.rtarget[r]), LA.cDeltaF, LA.deltaF[p], d);
				const float* tz = LA.resToZero + 8 * (size_t)r;
#pragma unroll
				for (int i = 0; i < 8; i++)
				{
					float rtz = tz[i];                                              // :93-99
					rtz = fadd(rtz, fmul(J[J_IDX0 + i], d[0]));
					rtz = fadd(rtz, fmul(J[J_IDX1 + i], d[1]));
					rtz = fadd(rtz, fmul(J[J_AB0 + i], d[2]));
					rtz = fadd(rtz, fmul(J[J_AB1 + i], d[3]));
					res[i] = rtz;
				}
			}
			float JI_r0 = 0.f, JI_r1 = 0.f, Jab_r0 = 0.f, Jab_r1 = 0.f, rr = 0.f;   // :107-118
#pragma unroll
			for (int i = 0; i < 8; i++)
			{
				const float ri = res[i];
				JI_r0 = fadd(JI_r0, fmul(ri, J[J_IDX0 + i]));
				JI_r1 = fadd(JI_r1, fmul(ri, J[J_IDX1 + i]));
				Jab_r0 = fadd(Jab_r0, fmul(ri, J[J_AB0 + i]));
				Jab_r1 = fadd(Jab_r1, fmul(ri, J[J_AB1 + i]));
				rr = fadd(rr, fmul(ri, ri));
			}
			w[0] = J[J_IDX2 + 0]; w[1] = J[J_IDX2 + 1]; w[2] = J[J_IDX2 + 3];
			w[3] = J[J_ABIDX + 0]; w[4] = J[J_ABIDX + 1]; w[5] = J[J_ABIDX + 2]; w[6] = J[J_ABIDX + 3]; w[7] = JI_r0; w[8] = JI_r1;
			w[9] = J[J_AB2 + 0]; w[10] = J[J_AB2 + 1]; w[11] = Jab_r0; w[12] = J[J_AB2 + 3]; w[13] = Jab_r1; w[14] = rr;
			// :142-150 per-point terms of this residual
			const float d0 = J[J_PDD], d1 = J[J_PDD + 1];
			const float q0 = fadd(fmul(J[J_IDX2 + 0], d0), fmul(J[J_IDX2 + 1], d1));
			const float q1 = fadd(fmul(J[J_IDX2 + 2], d0), fmul(J[J_IDX2 + 3], d1));
			if (contrib)
			{
				float* c = contrib + 6 * (size_t)r;
				c[0] = fadd(fmul(JI_r0, d0), fmul(JI_r1, d1));
				c[1] = fadd(fmul(q0, d0), fmul(q1, d1));
#pragma unroll
				for (int k = 0; k < 4; k++) c[2 + k] = fadd(fmul(J[J_PDC0 + k], q0), fmul(J[J_PDC1 + k], q1));
			}
			atomicAdd(&s_n, 1);
		}
		else
		{
#pragma unroll
			for (int k = 0; k < 10; k++) { x[k] = 0.f; y[k] = 0.f; }
#pragma unroll
			for (int k = 0; k < 15; k++) w[k] = 0.f;
		}
#pragma unroll
		for (int k = 0; k < 10; k++) { sx[tid][k] = x[k]; sy[tid][k] = y[k]; }
#pragma unroll
		for (int k = 0; k < 15; k++) sw[tid][k] = w[k];
	}
	__syncthreads();
	if (tid < NTOP)
	{
		float acc = 0.f;
		if (tid < 55)
		{
			int r = 0, rem = tid;
			while (rem >= 10 - r) { rem -= 10 - r; r++; }
			const int c = r + rem;
			for (int i = 0; i < cnt; i++)
			{
				const float a = sw[i][0], b = sw[i][1], cc = sw[i][2];
				const float xc = sx[i][c], xr = sx[i][r], yc = sy[i][c], yr = sy[i][r];
				// MatrixAccumulators.h:738-842: a*x[c]*x[r] + c*y[c]*y[r] + b*(x[c]*y[r] + y[c]*x[r])
				const float term = fadd(fadd(fmul(fmul(a, xc), xr), fmul(fmul(cc, yc), yr)), fmul(b, fadd(fmul(xc, yr), fmul(yc, xr))));
				acc = fadd(acc, term);
			}
		}
		else if (tid < 85)
		{
			const int i10 = (tid - 55) / 3, k = (tid - 55) - 3 * i10;
			for (int i = 0; i < cnt; i++)
				acc = fadd(acc, fadd(fmul(sx[i][i10], sw[i][3 + 2 * k]), fmul(sy[i][i10], sw[i][4 + 2 * k])));   // :845-897: x*TRk0 + y*TRk1
		}
		else
		{
			const int k = tid - 85;
			for (int i = 0; i < cnt; i++) acc = fadd(acc, sw[i][9 + k]);        // :899-913
		}
		partials[(size_t)chunk * NTOP_PAD + tid] = acc;
	}
	if (tid == 0) chunk_nres[chunk] = s_n;
}

// TopRight argument order of updateTopRight (AccumulatedTopHessian.cpp:134-139): TR00=JabJIdx(0,0), TR10=JabJIdx(0,1),
// TR01=JabJIdx(1,0), TR11=JabJIdx(1,1), TR02=JI_r[0], TR12=JI_r[1]  -> sw[3..8] above is already in (k0,k1) pairs.

// AccumulatedSCHessianSSE::addPoint (AccumulatedSCHessian.cpp:35-60) for one point without marginalisation mask: inverse
// Hessian and b from the AF and LF sums (shiftPriorToZero = true, EnergyFunctional.cpp:263, 271)
struct SchurPrepOut { float* HdiF; float* bdSumF; float* idh; float* HcdSum; const float* priorF; const float* deltaF; };
__device__ __forceinline__ void schur_prep_point(int p, const int* __restrict__ res_begin, const unsigned char* __restrict__ isAct,
                                                 float HddA, float bdA, const float* hcA, float HddL, float bdL, const float* hcL,
                                                 const SchurPrepOut& o, int ngood_known = -1, bool have_prior = false, float priorF_v = 0.f, float deltaF_v = 0.f)
{
	const float prF = have_prior ? priorF_v : o.priorF[p], dlF = have_prior ? deltaF_v : o.deltaF[p];
#pragma unroll
	for (int k = 0; k < 4; k++) o.HcdSum[4 * p + k] = fadd(hcA[k], hcL[k]);   // Hcd_accAF + Hcd_accLF (:62)
	int ngood = ngood_known;
	if (ngood < 0)
	{
		ngood = 0;
		for (int r = res_begin[p]; r < res_begin[p + 1]; r++) ngood += isAct[r] ? 1 : 0;
	}
	if (ngood == 0) { o.HdiF[p] = 0.f; o.bdSumF[p] = 0.f; o.idh[p] = 0.f; return; }
	float H = fadd(fadd(HddA, HddL), prF);           // Hdd_accAF + Hdd_accLF + priorF
	if (H < 1e-10) H = 1e-10;
	o.idh[p] = H;
	o.HdiF[p] = (float)(1.0 / (double)H);
	o.bdSumF[p] = fadd(fadd(bdA, bdL), fmul(prF, dlF));
}

// per-point sums in EFPoint::residualsAll order, AccumulatedTopHessian.cpp:142-160.  prep (modes 0 and 1, no mask): the
// Schur preparation of the same point follows in the same thread (saves a launch on the critical path); the sums of the
// other mode are read from where its own pass left them.
__global__ void ba_point_sums_kernel(int nP, const int* __restrict__ res_begin, const unsigned char* __restrict__ isAct,
                                     const unsigned char* __restrict__ isLin, const float* __restrict__ contrib,
                                     float* Hdd, float* bd, float* Hcd, int mode, const unsigned char* __restrict__ pmask,
                                     float* Hdd_A, float* bd_A, float* Hcd_A,
                                     bool prep = false, const float* Hdd_L = nullptr, const float* bd_L = nullptr, const float* Hcd_L = nullptr,
                                     SchurPrepOut po = SchurPrepOut())
{
	const int p = blockIdx.x * blockDim.x + threadIdx.x;
	if (p >= nP) return;
	if (mode == 2 && !pmask[p]) return;                        // addPoint<2> only visits the points being marginalised
	float bd_acc = 0.f, Hdd_acc = 0.f, h0 = 0.f, h1 = 0.f, h2 = 0.f, h3 = 0.f;
	int nact = 0;                                              // active residuals of the point (the Schur preparation's "ngood")
	const int r1 = res_begin[p + 1];
	// what the Schur preparation needs from elsewhere does not depend on the residual loop: asked for now
	float oHdd = 0.f, obd = 0.f, oth[4] = { 0.f, 0.f, 0.f, 0.f }, prF = 0.f, dlF = 0.f;
	if (prep)
	{
		prF = po.priorF[p]; dlF = po.deltaF[p];
		const float* src = mode == 0 ? Hcd_L : Hcd_A;
#pragma unroll
		for (int k = 0; k < 4; k++) oth[k] = src[4 * p + k];
		oHdd = mode == 0 ? Hdd_L[p] : Hdd_A[p];
		obd = mode == 0 ? bd_L[p] : bd_A[p];
	}
	for (int r0 = res_begin[p]; r0 < r1; r0 += 8)              // eight residuals' loads in flight (a point has at most nF-1), sums in residualsAll order
	{
		float c[8][6]; bool use[8];
#pragma unroll
		for (int u = 0; u < 8; u++)
		{
			const int r = r0 + u;
			const bool act = r < r1 && isAct[r];
			nact += act ? 1 : 0;
			use[u] = act && (mode == 0 ? !isLin[r] : (mode == 1 ? isLin[r] != 0 : true));
			// unconditional (clamped) loads: they go out together with the flag loads instead of one round behind them, and
			// predicated loads would be batched by the seven predicate registers; unused values are never touched
			const size_t rc = (size_t)min(r, r1 - 1);
#pragma unroll
			for (int k = 0; k < 6; k++) c[u][k] = contrib[6 * rc + k];
		}
#pragma unroll
		for (int u = 0; u < 8; u++)
			if (use[u])
			{
				bd_acc = fadd(bd_acc, c[u][0]); Hdd_acc = fadd(Hdd_acc, c[u][1]);
				h0 = fadd(h0, c[u][2]); h1 = fadd(h1, c[u][3]); h2 = fadd(h2, c[u][4]); h3 = fadd(h3, c[u][5]);
			}
	}
	Hdd[p] = Hdd_acc; bd[p] = bd_acc;
	Hcd[4 * p] = h0; Hcd[4 * p + 1] = h1; Hcd[4 * p + 2] = h2; Hcd[4 * p + 3] = h3;
	if (mode == 2)                                              // :168-175: the AF sums of a marginalised point are zero
	{
		Hdd_A[p] = 0.f; bd_A[p] = 0.f;
		Hcd_A[4 * p] = 0.f; Hcd_A[4 * p + 1] = 0.f; Hcd_A[4 * p + 2] = 0.f; Hcd_A[4 * p + 3] = 0.f;
	}
	if (prep)
	{
		const float own[4] = { h0, h1, h2, h3 };
		if (mode == 0) schur_prep_point(p, res_begin, isAct, Hdd_acc, bd_acc, own, oHdd, obd, oth, po, nact, true, prF, dlF);
		else schur_prep_point(p, res_begin, isAct, oHdd, obd, oth, Hdd_acc, bd_acc, own, po, nact, true, prF, dlF);
	}
}

// index of entry (i,j) of the symmetric 13x13 block in the 91-vector written by ba_accum_top_kernel
__device__ __forceinline__ int top_index(int i, int j)
{
	if (i > j) { int t = i; i = j; j = t; }
	if (j < 10) return i * 10 - (i * (i - 1)) / 2 + (j - i);
	if (i < 10) return 55 + 3 * i + (j - 10);
	// bottom-right 3x3: (10,10)->0 (10,11)->1 (10,12)->2 (11,11)->3 (11,12)->4 (12,12)->5
	const int a = i - 10, b = j - 10;
	return 85 + (a == 0 ? b : (a == 1 ? 2 + b : 5));
}

// v += sum_{k < n} EXPR(k) for n <= MAXF: every load is issued before the first add, the adds stay in index order
// (a plain loop over a run-time n makes each add wait for its own load: n dependent DRAM/L2 round trips)
#define ORDERED_SUM_MAXF(v, n, k, EXPR) do { double t_[MAXF]; _Pragma("unroll") for (int k = 0; k < MAXF; k++) t_[k] = (k < (n)) ? (EXPR) : 0.0; \
	_Pragma("unroll") for (int k = 0; k < MAXF; k++) if (k < (n)) v += t_[k]; } while (0)

// sum of the per-chunk partials of one entry over the chunks [cb, c1) in chunk order, FP64.  The loop is a chain of
// memory round trips (large windows have hundreds of chunks per group), so 32 loads are in flight per round.
#define CHUNK_SUM_U 32
__device__ __forceinline__ double ordered_chunk_sum(const float* __restrict__ partials, int stride, int e, int cb, int c1)
{
	double s = 0.0;
	for (int c0 = cb; c0 < c1; c0 += CHUNK_SUM_U)                // adds in chunk order
	{
		float t[CHUNK_SUM_U];
#pragma unroll
		for (int u = 0; u < CHUNK_SUM_U; u++) t[u] = partials[(size_t)min(c0 + u, c1 - 1) * stride + e];   // clamped, not predicated:
		// predicated loads would be batched by the number of predicate registers (7), not by CHUNK_SUM_U
		asm volatile("" ::: "memory");                             // every load of the round is issued before the first add
#pragma unroll
		for (int u = 0; u < CHUNK_SUM_U; u++) if (c0 + u < c1) s += (double)t[u];
	}
	return s;
}

// two entries at once (either may be -1): one loop, the same additions per entry as above
__device__ __forceinline__ void ordered_chunk_sum2(const float* __restrict__ partials, int stride, int eA, int eB, int cb, int c1,
                                                   double& sA, double& sB)
{
	sA = 0.0; sB = 0.0;
	const bool hasB = __any_sync(0xffffffffu, eB >= 0);          // warp-uniform: only slice k = 1 has second entries
	for (int c0 = cb; c0 < c1; c0 += CHUNK_SUM_U)
	{
		float ta[CHUNK_SUM_U], tb[CHUNK_SUM_U];
#pragma unroll
		for (int u = 0; u < CHUNK_SUM_U; u++)
		{
			const size_t row = (size_t)min(c0 + u, c1 - 1) * stride;
			ta[u] = partials[row + max(eA, 0)];
			tb[u] = hasB ? partials[row + max(eB, 0)] : 0.f;
		}
		asm volatile("" ::: "memory");
#pragma unroll
		for (int u = 0; u < CHUNK_SUM_U; u++) if (c0 + u < c1) { sA += (double)ta[u]; sB += (double)tb[u]; }
	}
	if (eA < 0) sA = 0.0;
	if (eB < 0) sB = 0.0;
}

// the same for every group (pair or host): one thread per entry
__global__ void ba_reduce_partials_kernel(const float* __restrict__ partials, int stride, int nEntries,
                                          const int* __restrict__ group_begin, double* out)
{
	const int e = blockIdx.x * blockDim.x + threadIdx.x, g = blockIdx.y;
	if (e >= nEntries) return;
	out[(size_t)g * nEntries + e] = ordered_chunk_sum(partials, stride, e, group_begin[g], group_begin[g + 1]);
}

// contribution layout per (host,target) pair: [HH 64][TT 64][HT 64][CH 32][CT 32][BH 8][BT 8][CC 16][BC 4]
#define TC_STRIDE 292
#define TC_HH 0
#define TC_TT 64
#define TC_HT 128
#define TC_CH 192
#define TC_CT 224
#define TC_BH 256
#define TC_BT 264
#define TC_CC 272
#define TC_BC 288

// stitchDoubleInternal (AccumulatedTopHessian.cpp:259-310) for one pair: the adjoint sandwiches of its 13x13 block
// partials != NULL (few chunks per pair): the block sums the chunk partials of its pair itself, in chunk order, instead of
// reading the output of a separate reduction launch
__global__ void __launch_bounds__(256) ba_top_pair_kernel(int nF, const double* __restrict__ topRed, const double* __restrict__ adHost,
                                                          const double* __restrict__ adTarget, double* contrib,
                                                          const float* __restrict__ partials = nullptr, const int* __restrict__ pair_chunk_begin = nullptr)
{
	__shared__ double accH[176], AH[64], AT[64], T1[192];
	const int k = blockIdx.x, tid = threadIdx.x;
	double* out = contrib + (size_t)k * TC_STRIDE;
	if (tid < 169)
	{
		const int i = tid / 13, j = tid - 13 * i;
		accH[tid] = partials ? ordered_chunk_sum(partials, NTOP_PAD, top_index(i, j), pair_chunk_begin[k], pair_chunk_begin[k + 1])
		                     : topRed[(size_t)k * NTOP_PAD + top_index(i, j)];
	}
	else if (tid < 233) AH[tid - 169] = adHost[(size_t)k * 64 + tid - 169];
	if (tid >= 192) AT[tid - 192] = adTarget[(size_t)k * 64 + tid - 192];
	__syncthreads();
	if (tid < 192)
	{
		const int s = tid >> 6, o = tid & 63, i = o >> 3, j = o & 7;
		const double* L = (s == 1) ? AT : AH;
		double v = 0.0;
#pragma unroll
		for (int q = 0; q < 8; q++) v += L[i * 8 + q] * accH[(CPARS + q) * 13 + CPARS + j];
		T1[s * 64 + o] = v;
	}
	else
	{
		const int o = tid - 192, s = o >> 5, i = (o & 31) >> 2, j = o & 3;   // calibration columns
		const double* L = s ? AT : AH;
		double v = 0.0;
#pragma unroll
		for (int q = 0; q < 8; q++) v += L[i * 8 + q] * accH[(CPARS + q) * 13 + j];
		out[(s ? TC_CT : TC_CH) + (o & 31)] = v;
	}
	__syncthreads();
	if (tid < 192)
	{
		const int s = tid >> 6, o = tid & 63, i = o >> 3, j = o & 7;
		const double* R = (s == 0) ? AH : AT;          // s=0: AH A AH^T, s=1: AT A AT^T, s=2: AH A AT^T
		double v = 0.0;
#pragma unroll
		for (int q = 0; q < 8; q++) v += T1[s * 64 + i * 8 + q] * R[j * 8 + q];
		out[(s == 0 ? TC_HH : (s == 1 ? TC_TT : TC_HT)) + o] = v;
	}
	else if (tid < 208)
	{
		const int o = tid - 192, s = o >> 3, i = o & 7;
		const double* L = s ? AT : AH;
		double v = 0.0;
#pragma unroll
		for (int q = 0; q < 8; q++) v += L[i * 8 + q] * accH[(CPARS + q) * 13 + CPARS + 8];
		out[(s ? TC_BT : TC_BH) + i] = v;
	}
	else if (tid < 224) { const int o = tid - 208; out[TC_CC + o] = accH[(o >> 2) * 13 + (o & 3)]; }
	else if (tid < 228) out[TC_BC + tid - 224] = accH[(tid - 224) * 13 + CPARS + 8];
}

// gather per output block in a fixed order, including "make diagonal by copying over parts" (AccumulatedTopHessian.h:126-138)
__global__ void __launch_bounds__(64) ba_top_gather_kernel(int nF, int nChunks, const int* __restrict__ chunk_nres,
                                                           const double* __restrict__ contrib, double* Hout, double* bout, double* scalars,
                                                           const double* __restrict__ priorH, const double* __restrict__ priorD,
                                                           const float* __restrict__ cDeltaF)
{
	const int N = 8 * nF + CPARS, o = threadIdx.x, i = o >> 3, j = o & 7;
	if ((int)blockIdx.x < nF * nF)
	{
		const int r = blockIdx.x / nF, c = blockIdx.x % nF;
		double v = 0.0;
		if (r == c)
		{
			// pairs with host r, then pairs with target r, alternating (fixed order)
			double th[MAXF], tt[MAXF];
#pragma unroll
			for (int q = 0; q < MAXF; q++)
			{
				th[q] = (q < nF) ? contrib[(size_t)(r + nF * q) * TC_STRIDE + TC_HH + o] : 0.0;
				tt[q] = (q < nF) ? contrib[(size_t)(q + nF * r) * TC_STRIDE + TC_TT + o] : 0.0;
			}
#pragma unroll
			for (int q = 0; q < MAXF; q++) if (q < nF) { v += th[q]; v += tt[q]; }
			if (priorH && i == j) v += priorH[CPARS + r * 8 + i];     // usePrior, AccumulatedTopHessian.cpp:314-324
		}
		else
			v = contrib[(size_t)(r + nF * c) * TC_STRIDE + TC_HT + o] + contrib[(size_t)(c + nF * r) * TC_STRIDE + TC_HT + j * 8 + i];
		Hout[(CPARS + r * 8 + i) * N + CPARS + c * 8 + j] = v;
	}
	else if ((int)blockIdx.x < nF * nF + nF)
	{
		// calibration rows / columns and b of frame r: one block per frame, one entry per thread (a single round of loads;
		// one block doing this for all frames, then the calibration block, then resInA was a dozen dependent rounds)
		const int r = blockIdx.x - nF * nF;
		if (o < 32)
		{
			const int q = o;
			double v = 0.0;
			{
				double th[MAXF], tt[MAXF];
#pragma unroll
				for (int m = 0; m < MAXF; m++)
				{
					th[m] = (m < nF) ? contrib[(size_t)(r + nF * m) * TC_STRIDE + TC_CH + q] : 0.0;
					tt[m] = (m < nF) ? contrib[(size_t)(m + nF * r) * TC_STRIDE + TC_CT + q] : 0.0;
				}
#pragma unroll
				for (int m = 0; m < MAXF; m++) if (m < nF) { v += th[m]; v += tt[m]; }
			}
			const int row = CPARS + r * 8 + (q >> 2), col = q & 3;
			Hout[row * N + col] = v; Hout[col * N + row] = v;
		}
		else if (o < 40)
		{
			const int q = o - 32, e = r * 8 + q;
			double v = 0.0;
			{
				double th[MAXF], tt[MAXF];
#pragma unroll
				for (int m = 0; m < MAXF; m++)
				{
					th[m] = (m < nF) ? contrib[(size_t)(r + nF * m) * TC_STRIDE + TC_BH + q] : 0.0;
					tt[m] = (m < nF) ? contrib[(size_t)(m + nF * r) * TC_STRIDE + TC_BT + q] : 0.0;
				}
#pragma unroll
				for (int m = 0; m < MAXF; m++) if (m < nF) { v += th[m]; v += tt[m]; }
			}
			if (priorH) v += __dmul_rn(priorH[CPARS + e], priorD[CPARS + e]);   // prior.cwiseProduct(delta_prior)
			bout[CPARS + e] = v;
		}
	}
	else if ((int)blockIdx.x == nF * nF + nF)
	{
		// calibration block / vector: sum over all pairs, three threads per entry (each a third of the pairs, in order), then
		// the three parts in order
		__shared__ double s_cc[3][20];
		if (o < 60)
		{
			const int e = o % 20, part = o / 20, np = nF * nF, per = (np + 2) / 3;
			const int k1 = min(np, (part + 1) * per);
			double v = 0.0;
			for (int k0 = part * per; k0 < k1; k0 += 8)
			{
				double t[8];
#pragma unroll
				for (int u = 0; u < 8; u++) t[u] = (k0 + u < k1) ? contrib[(size_t)(k0 + u) * TC_STRIDE + TC_CC + e] : 0.0;
#pragma unroll
				for (int u = 0; u < 8; u++) if (k0 + u < k1) v += t[u];
			}
			s_cc[part][e] = v;
		}
		__syncthreads();
		if (o < 20)
		{
			double v = (s_cc[0][o] + s_cc[1][o]) + s_cc[2][o];
			if (priorH && o < 16 && (o >> 2) == (o & 3)) v += priorH[o >> 2];
			if (priorH && o >= 16) v += __dmul_rn(priorH[o - 16], (double)cDeltaF[o - 16]);   // cPrior.cwiseProduct(cDeltaF.cast<double>())
			if (o < 16) Hout[(o >> 2) * N + (o & 3)] = v; else bout[o - 16] = v;
		}
	}
	else
	{
		// resInA: integer sum over the chunks, own block, eight loads per thread in flight
		__shared__ int s_n[2];
		int n = 0;
		for (int c0 = o; c0 < nChunks; c0 += 64 * 8)
		{
			int t[8];
#pragma unroll
			for (int u = 0; u < 8; u++) t[u] = chunk_nres[min(c0 + 64 * u, nChunks - 1)];
#pragma unroll
			for (int u = 0; u < 8; u++) if (c0 + 64 * u < nChunks) n += t[u];
		}
#pragma unroll
		for (int d = 16; d > 0; d >>= 1) n += __shfl_xor_sync(0xffffffffu, n, d);
		if ((o & 31) == 0) s_n[o >> 5] = n;
		__syncthreads();
		if (o == 0) scalars[1] = (double)(s_n[0] + s_n[1]);
	}
}

// AccumulatedSCHessianSSE::addPoint (AccumulatedSCHessian.cpp:35-60): per-point inverse Hessian and b
__global__ void ba_schur_prep_kernel(int nP, const int* __restrict__ res_begin, const unsigned char* __restrict__ isAct,
                                     const float* __restrict__ Hdd, const float* __restrict__ bd, const float* __restrict__ Hcd,
                                     const float* __restrict__ Hdd_L, const float* __restrict__ bd_L, const float* __restrict__ Hcd_L,
                                     const float* __restrict__ priorF, const float* __restrict__ deltaF,
                                     float* HdiF, float* bdSumF, float* idh, float* HcdSum, const unsigned char* __restrict__ pmask)
{
	const int p = blockIdx.x * blockDim.x + threadIdx.x;
	if (p >= nP) return;
	// marginalisation pass (pmask != NULL): only the flagged points contribute (weight HdiF = 0 for the others, their
	// idepth_hessian is left alone) and the prior is not shifted (addPoint(p, false), EnergyFunctional.cpp:670)
	if (pmask && !pmask[p])
	{
#pragma unroll
		for (int k = 0; k < 4; k++) HcdSum[4 * p + k] = 0.f;
		HdiF[p] = 0.f; bdSumF[p] = 0.f;
		return;
	}
#pragma unroll
	for (int k = 0; k < 4; k++) HcdSum[4 * p + k] = fadd(Hcd[4 * p + k], Hcd_L[4 * p + k]);   // Hcd_accAF + Hcd_accLF (:62)
	int ngood = 0;
	for (int r = res_begin[p]; r < res_begin[p + 1]; r++) ngood += isAct[r] ? 1 : 0;
	if (ngood == 0) { HdiF[p] = 0.f; bdSumF[p] = 0.f; idh[p] = 0.f; return; }
	float H = fadd(fadd(Hdd[p], Hdd_L[p]), priorF[p]);   // Hdd_accAF + Hdd_accLF + priorF
	if (H < 1e-10) H = 1e-10;
	idh[p] = H;
	HdiF[p] = (float)(1.0 / (double)H);
	float bs = fadd(bd[p], bd_L[p]);
	if (!pmask) bs = fadd(bs, fmul(priorF[p], deltaF[p]));   // shiftPriorToZero = true (EnergyFunctional.cpp:263, 271)
	bdSumF[p] = bs;
}

// optional fusion: the per-point sums of addPoint<0> and the Schur preparation of the chunk's own points (contrib == NULL: off)
struct SchurFusedSums
{
	const float* contrib; const unsigned char* isLin;
	float* Hdd; float* bd; float* Hcd;                       // out: the AF sums (what ba_point_sums_kernel writes in mode 0)
	const float* Hdd_L; const float* bd_L; const float* Hcd_L; const float* priorF; const float* deltaF;
	float* HdiF; float* bdSumF; float* idh; float* HcdSum;   // out: what the preparation writes
};

// accD / accE / accEB / accHcc / accbc contributions of one chunk of points of one host frame
// (AccumulatedSCHessian.cpp:64-84): thread-owned output entries, points visited in order.
__global__ void __launch_bounds__(256) ba_schur_accum_kernel(int nF, const int* __restrict__ pc_start, const int* __restrict__ pc_cnt,
                                                             const int* __restrict__ res_begin, const int* __restrict__ rtarget,
                                                             const unsigned char* __restrict__ isAct, const float* __restrict__ JpJdF,
                                                             const float* __restrict__ HdiF, const float* __restrict__ bdSumF,
                                                             const float* __restrict__ Hcd, float* partials, int scN,
                                                             SchurFusedSums fs = SchurFusedSums())
{
	__shared__ float sa[CHUNK_P][MAXF * 8];     // JpJdF by target slot (0 when inactive)
	__shared__ float swa[CHUNK_P][MAXF * 8];    // HdiF * JpJdF
	__shared__ float sh[CHUNK_P][8];            // Hcd[4], w, w*bdSum, w*Hcd handled on the fly
	__shared__ unsigned char s_t1[MAXF * MAXF], s_t2[MAXF * MAXF];   // pair index -> (t1, t2): no run-time divisions per entry
	static_assert(CHUNK_P * 16 == 256, "16 threads per point below");
	const int chunk = blockIdx.x, tid = threadIdx.x;
	const int start = pc_start[chunk], cnt = pc_cnt[chunk];
	for (int i = tid; i < CHUNK_P * MAXF * 8; i += 256) { (&sa[0][0])[i] = 0.f; (&swa[0][0])[i] = 0.f; }
	if (tid < nF * nF) { s_t1[tid] = (unsigned char)(tid / nF); s_t2[tid] = (unsigned char)(tid % nF); }
	__syncthreads();
	{
		// 16 threads per point: two residuals at a time, one JpJdF entry each (the serial per-point loop was a chain of
		// dependent global loads: begin -> flag -> target -> eight values, residual after residual)
		const int pl = tid >> 4, sub = tid & 15, k = sub & 7;
		float w_f = 0.f, hc_f = 0.f, wb_f = 0.f;
		if (fs.contrib)
		{
			// the per-point sums of addPoint<0> (AccumulatedTopHessian.cpp:142-160, residualsAll order) and the Schur
			// preparation (AccumulatedSCHessian.cpp:35-60) of this chunk's own points, here instead of in a launch of their own:
			// lanes 0-5 of a point's 16 sum one component each (bd, Hdd, Hcd[0..3]), lane 0 then prepares the point
			const int p = start + min(pl, cnt - 1);
			const int r0 = res_begin[p], r1 = res_begin[p + 1];
			const float prF = fs.priorF[p], dlF = fs.deltaF[p];
			const float oth = sub < 6 ? (sub == 0 ? fs.bd_L[p] : (sub == 1 ? fs.Hdd_L[p] : fs.Hcd_L[4 * p + sub - 2])) : 0.f;
			float acc = 0.f;
			int nact = 0;
			const int comp = min(sub, 5);
			for (int rb = r0; rb < r1; rb += 8)
			{
				float c[8]; bool use[8];
#pragma unroll
				for (int u = 0; u < 8; u++)
				{
					const int r = rb + u, rc = min(r, r1 - 1);
					const bool act = r < r1 && isAct[rc];
					nact += act ? 1 : 0;
					use[u] = act && !fs.isLin[rc];
					c[u] = fs.contrib[6 * (size_t)rc + comp];
				}
#pragma unroll
				for (int u = 0; u < 8; u++) if (use[u]) acc = fadd(acc, c[u]);
			}
			const int b16 = (tid & 31) & ~15;
			const float s_bd = __shfl_sync(0xffffffffu, acc, b16 + 0), s_Hdd = __shfl_sync(0xffffffffu, acc, b16 + 1);
			const float o_bd = __shfl_sync(0xffffffffu, oth, b16 + 0), o_Hdd = __shfl_sync(0xffffffffu, oth, b16 + 1);
			const float hsum = fadd(acc, oth);                          // lanes 2-5: Hcd_accAF + Hcd_accLF (:62)
			float H = fadd(fadd(s_Hdd, o_Hdd), prF);                   // Hdd_accAF + Hdd_accLF + priorF
			if (H < 1e-10) H = 1e-10;
			const float wv = nact == 0 ? 0.f : (float)(1.0 / (double)H);
			const float bs = nact == 0 ? 0.f : fadd(fadd(s_bd, o_bd), fmul(prF, dlF));
			if (pl < cnt)
			{
				if (sub == 0) { fs.bd[p] = acc; fs.HdiF[p] = wv; fs.bdSumF[p] = bs; fs.idh[p] = nact == 0 ? 0.f : H; }
				else if (sub == 1) fs.Hdd[p] = acc;
				else if (sub < 6) { fs.Hcd[4 * p + sub - 2] = acc; fs.HcdSum[4 * p + sub - 2] = hsum; }
			}
			w_f = wv; wb_f = fmul(wv, bs);
			hc_f = __shfl_sync(0xffffffffu, hsum, b16 + min(sub + 2, 5));   // lanes 0-3 fetch HcdSum[0..3] from lanes 2-5
		}
		if (pl < cnt)
		{
			const int p = start + pl;
			const float w = fs.contrib ? w_f : HdiF[p];
			if (sub < 4) sh[pl][sub] = fs.contrib ? hc_f : Hcd[4 * p + sub];
			else if (sub == 4) sh[pl][4] = w;
			else if (sub == 5) sh[pl][5] = fs.contrib ? wb_f : fmul(w, bdSumF[p]);
			const int r1 = res_begin[p + 1];
			for (int r = res_begin[p] + (sub >> 3); r < r1; r += 2)
			{
				if (!isAct[r]) continue;
				const int t = rtarget[r];
				const float a = JpJdF[8 * (size_t)r + k];
				sa[pl][t * 8 + k] = a; swa[pl][t * 8 + k] = fmul(w, a);
			}
		}
	}
	__syncthreads();
	const int nD = nF * nF * 64, nE = nF * 32, nEB = nF * 8;
	float* out = partials + (size_t)chunk * scN;
	for (int e = tid; e < nD + nE + nEB + 20; e += 256)
	{
		float acc = 0.f;
		if (e < nD)
		{
			const int t1 = s_t1[e >> 6], t2 = s_t2[e >> 6], i = (e >> 3) & 7, j = e & 7;
			for (int p = 0; p < cnt; p++) acc = fadd(acc, fmul(sa[p][t2 * 8 + j], swa[p][t1 * 8 + i]));       // A += (w L) R^T
		}
		else if (e < nD + nE)
		{
			const int o = e - nD, t = o >> 5, i = (o >> 2) & 7, c = o & 3;
			for (int p = 0; p < cnt; p++) acc = fadd(acc, fmul(sh[p][c], swa[p][t * 8 + i]));                   // accE: L = JpJdF, R = Hcd
		}
		else if (e < nD + nE + nEB)
		{
			const int o = e - nD - nE;
			for (int p = 0; p < cnt; p++) acc = fadd(acc, fmul(sh[p][5], sa[p][o]));                            // accEB: (HdiF*bdSumF) * JpJdF
		}
		else if (e < nD + nE + nEB + 16)
		{
			const int o = e - nD - nE - nEB, i = o >> 2, j = o & 3;
			for (int p = 0; p < cnt; p++) acc = fadd(acc, fmul(sh[p][j], fmul(sh[p][4], sh[p][i])));            // accHcc: (w Hcd) Hcd^T
		}
		else
		{
			const int i = e - nD - nE - nEB - 16;
			for (int p = 0; p < cnt; p++) acc = fadd(acc, fmul(sh[p][5], sh[p][i]));                            // accbc: (bdSumF*HdiF) * Hcd
		}
		out[e] = acc;
	}
}

// contribution buffer layout per (i,j): [C1 64][C2 64][C3 MAXF*64][C4 MAXF*64][CEh 32][CEt 32][Cbh 8][Cbt 8]
#define SCC_STRIDE (128 + 2 * MAXF * 64 + 64 + 16)
#define SCC_C1 0
#define SCC_C2 64
#define SCC_C3 128
#define SCC_C4 (128 + MAXF * 64)
#define SCC_CEH (128 + 2 * MAXF * 64)
#define SCC_CET (SCC_CEH + 32)
#define SCC_CBH (SCC_CEH + 64)
#define SCC_CBT (SCC_CEH + 72)

__device__ __forceinline__ void mm8(const double* A, const double* Bm, double* C, bool transB, int o)
{
	const int i = o >> 3, j = o & 7;
	double v = 0.0;
#pragma unroll
	for (int q = 0; q < 8; q++) v += A[i * 8 + q] * (transB ? Bm[j * 8 + q] : Bm[q * 8 + j]);
	C[o] = v;
}

// stage 1 of stitchDoubleInternal (AccumulatedSCHessian.cpp:86-165) for the pair (i = host, j = target):
// every adjoint sandwich that involves accD[i,j,*], accE[i,j], accEB[i,j], written out as separate contributions.
// The chunk partials of host i (accD[i,j,*], accE[i,j], accEB[i,j]) are summed here, in chunk order, by the threads that
// use them -- each entry has exactly one consumer -- and the pairs (i, 0) also sum accHcc / accbc of their host into scRed
// for the gather kernel: no separate reduction launch on the critical path.
__global__ void __launch_bounds__(64 * MAXF) ba_schur_stage1_kernel(int nF, const float* __restrict__ partials, int scN,
                                                                    const int* __restrict__ host_pc_begin, double* scRed, int nRed,
                                                                    const double* __restrict__ adHost, const double* __restrict__ adTarget,
                                                                    double* contrib)
{
	// one CTA per pair, one 64-thread slice per third frame k (the k-dependent products are independent of each other)
	__shared__ double Dk[MAXF][64], AHk[MAXF][64], ATk[MAXF][64], Vk[MAXF][64], Qk[MAXF][64];
	__shared__ double S1[64], AHij[64], ATij[64], E[32], EB[8];
	const int i = blockIdx.x % nF, j = blockIdx.x / nF, o = threadIdx.x & 63, k = threadIdx.x >> 6;
	const int ij = i + nF * j;
	const int nD = nF * nF * 64, nE = nF * 32;
	const int cb = host_pc_begin[i], ce = host_pc_begin[i + 1];      // chunks of host i
	double* out = contrib + (size_t)ij * SCC_STRIDE;
	const bool act = k < nF;
	if (k == 0) { AHij[o] = adHost[(size_t)ij * 64 + o]; ATij[o] = adTarget[(size_t)ij * 64 + o]; }
	// one pass over the chunks per thread: its accD entry and, for slice k = 1, one of accE / accEB / accHcc+accbc (no
	// divergent second and third pass)
	int eB = -1;
	if (k == 1)
	{
		if (o < 32) eB = nD + j * 32 + o;
		else if (o < 40) eB = nD + nE + j * 8 + (o - 32);
		else if (o < 60 && j == 0) eB = nD + nE + nF * 8 + (o - 40);
	}
	double sA = 0.0, sB = 0.0;
	const int eA = act ? (j * nF + k) * 64 + o : -1;
	if (partials) ordered_chunk_sum2(partials, scN, eA, eB, cb, ce, sA, sB);
	else
	{
		// large windows (hundreds of chunks per host): ba_reduce_partials_kernel has summed them with far more threads
		const double* red = scRed + (size_t)i * nRed;
		if (eA >= 0) sA = red[eA];
		if (eB >= 0) sB = red[eB];
	}
	if (k == 1 && o < 32) E[o] = sB;
	if (k == 1 && o >= 32 && o < 40) EB[o - 32] = sB;
	if (partials && k == 1 && o >= 40 && o < 60 && j == 0) scRed[(size_t)i * nRed + nD + nE + nF * 8 + (o - 40)] = sB;
	if (act)
	{
		const int ik = i + nF * k;
		Dk[k][o] = sA;
		AHk[k][o] = adHost[(size_t)ik * 64 + o]; ATk[k][o] = adTarget[(size_t)ik * 64 + o];
	}
	__syncthreads();
	if (act)
	{
		const double* D = Dk[k]; const double* AHik = AHk[k]; const double* ATik = ATk[k];
		const int r = o >> 3, c = o & 7;
		double v = 0.0, q = 0.0;
#pragma unroll
		for (int m = 0; m < 8; m++) { v += D[r * 8 + m] * AHik[c * 8 + m]; q += D[r * 8 + m] * ATik[c * 8 + m]; }
		Vk[k][o] = v;          // D[i,j,k] adHost[i,k]^T
		Qk[k][o] = q;          // D[i,j,k] adTarget[i,k]^T
	}
	__syncthreads();
	if (act)
	{
		mm8(ATij, Qk[k], out + SCC_C3 + k * 64, false, o);   // -> block (j,k)
		mm8(AHij, Qk[k], out + SCC_C4 + k * 64, false, o);   // -> block (i,k)
	}
	if (k == 0)
	{
		double sacc = 0.0;                                   // sum_k D[i,j,k] adHost[i,k]^T, k in order
		for (int q = 0; q < nF; q++) sacc += Vk[q][o];
		S1[o] = sacc;
	}
	if (k == 1 && o < 32)
	{
		const int r = o >> 2, c = o & 3;
		double vh = 0.0, vt = 0.0;
#pragma unroll
		for (int m = 0; m < 8; m++) { vh += AHij[r * 8 + m] * E[m * 4 + c]; vt += ATij[r * 8 + m] * E[m * 4 + c]; }
		out[SCC_CEH + o] = vh; out[SCC_CET + o] = vt;
	}
	if (k == 1 && o >= 32 && o < 40)
	{
		const int e = o - 32;
		double vh = 0.0, vt = 0.0;
#pragma unroll
		for (int m = 0; m < 8; m++) { vh += AHij[e * 8 + m] * EB[m]; vt += ATij[e * 8 + m] * EB[m]; }
		out[SCC_CBH + e] = vh; out[SCC_CBT + e] = vt;
	}
	__syncthreads();
	if (k == 0) mm8(AHij, S1, out + SCC_C1, false, o);       // -> block (i,i)
	if (k == 1) mm8(ATij, S1, out + SCC_C2, false, o);       // -> block (j,i)
}

// stage 2: gather the contributions of every output block in a fixed order
__global__ void __launch_bounds__(64) ba_schur_gather_kernel(int nF, const double* __restrict__ scRed,
                                                             const double* __restrict__ contrib, double* Hout, double* bout)
{
	const int N = 8 * nF + CPARS, o = threadIdx.x;
	if ((int)blockIdx.x < nF * nF)
	{
		const int r = blockIdx.x / nF, c = blockIdx.x % nF;
		double v = 0.0;
		const double c2 = contrib[(size_t)(c + nF * r) * SCC_STRIDE + SCC_C2 + o];                         // i = c, j = r
		if (r == c) ORDERED_SUM_MAXF(v, nF, j, contrib[(size_t)(r + nF * j) * SCC_STRIDE + SCC_C1 + o]);
		v += c2;
		ORDERED_SUM_MAXF(v, nF, i, contrib[(size_t)(i + nF * r) * SCC_STRIDE + SCC_C3 + c * 64 + o]);      // j = r, k = c
		ORDERED_SUM_MAXF(v, nF, j, contrib[(size_t)(r + nF * j) * SCC_STRIDE + SCC_C4 + c * 64 + o]);      // i = r, k = c
		Hout[(CPARS + r * 8 + (o >> 3)) * N + CPARS + c * 8 + (o & 7)] = v;
	}
	else if ((int)blockIdx.x < nF * nF + nF)
	{
		// calibration rows / columns and b of frame r: one block per frame, one entry per thread (a single round of loads)
		const int r = blockIdx.x - nF * nF;
		if (o < 32)
		{
			const int q = o;
			double v = 0.0;
			ORDERED_SUM_MAXF(v, nF, j, contrib[(size_t)(r + nF * j) * SCC_STRIDE + SCC_CEH + q]);
			ORDERED_SUM_MAXF(v, nF, i, contrib[(size_t)(i + nF * r) * SCC_STRIDE + SCC_CET + q]);
			const int row = CPARS + r * 8 + (q >> 2), col = q & 3;
			Hout[row * N + col] = v;
			Hout[col * N + row] = v;                                  // AccumulatedSCHessian.h:130-134
		}
		else if (o < 40)
		{
			const int q = o - 32;
			double v = 0.0;
			ORDERED_SUM_MAXF(v, nF, j, contrib[(size_t)(r + nF * j) * SCC_STRIDE + SCC_CBH + q]);
			ORDERED_SUM_MAXF(v, nF, i, contrib[(size_t)(i + nF * r) * SCC_STRIDE + SCC_CBT + q]);
			bout[CPARS + r * 8 + q] = v;
		}
	}
	else if (o < 20)
	{
		// accHcc / accbc: chunk partials already summed per host by ba_reduce_partials_kernel, hosts in order here
		const int nD = nF * nF * 64, nE = nF * 32, nEB = nF * 8;
		double v = 0.0;
		ORDERED_SUM_MAXF(v, nF, i, scRed[(size_t)i * (nD + nE + nEB + 20) + nD + nE + nEB + o]);
		if (o < 16) Hout[(o >> 2) * N + (o & 3)] = v; else bout[o - 16] = v;
	}
}

// ---- host side --------------------------------------------------------------------------------------------

extern "C" int dso_b200_ba_set_window(dso_b200_ctx* ctx, int nF, const int* frame_slots, int w, int h,
                                      float fx, float fy, float cx, float cy, const float* frameEnergyTH)
{
	ARG_CHECK(ctx && ctx->ba && frame_slots && frameEnergyTH, "null argument");
	ARG_CHECK(nF >= 2 && nF <= MAXF, "nF must be 2..8");
	BAWorkspace* W = ctx->ba;
	BAParams& P = W->P;
	for (int i = 0; i < nF; i++)
	{
		ARG_CHECK(frame_slots[i] >= 0 && frame_slots[i] < ctx->n_slots && ctx->slots[frame_slots[i]].valid, "frame slot empty");
		ARG_CHECK(ctx->slots[frame_slots[i]].w == w && ctx->slots[frame_slots[i]].h == h, "frame size mismatch");
		P.img[i] = ctx->slots[frame_slots[i]].pyr[0];
		if (frame_slot_wait(ctx, frame_slots[i], ctx->ba_stream, DSO_B200_STREAM_MAPPING)) return DSO_B200_E_CUDA;
		P.frameEnergyTH[i] = frameEnergyTH[i];
	}
	P.nF = nF; P.w = w; P.h = h;
	// CalibHessian::setValueScaled, HessianBlocks.h:383-397; globalCalib.cpp:70-71
	P.fx = fx; P.fy = fy; P.cx = cx; P.cy = cy;
	P.fxi = 1.0f / fx; P.fyi = 1.0f / fy;
	P.wM3G = (float)(w - 3); P.hM3G = (float)(h - 3);
	W->window_set = true; W->glue_set = false; W->res_set = false; W->lin_done = false; W->top_done = false;
	ba_graphs_invalidate(W);
	return 0;
}

extern "C" int dso_b200_ba_set_affine_modes(dso_b200_ctx* ctx, float modeA, float modeB)
{
	ARG_CHECK(ctx && ctx->ba, "null argument");
	ctx->ba->P.fixA = modeA < 0; ctx->ba->P.fixB = modeB < 0;
	ctx->ba->lin_done = false; ctx->ba->top_done = false;
	ba_graphs_invalidate(ctx->ba);
	return 0;
}

extern "C" int dso_b200_ba_set_glue(dso_b200_ctx* ctx, const dso_b200_precalc* precalc, const double* adHost,
                                    const double* adTarget, const float* adHTdeltaF, const float* cDeltaF4)
{
	ARG_CHECK(ctx && ctx->ba && precalc && adHost && adTarget, "null argument");
	BAWorkspace* W = ctx->ba;
	BA_CHECK(W->window_set, "call dso_b200_ba_set_window first");
	CK(cudaSetDevice(ctx->device));
	const int n2 = W->P.nF * W->P.nF;
	CK(cudaMemcpyAsync(W->d_precalc, precalc, (size_t)n2 * sizeof(dso_b200_precalc), cudaMemcpyHostToDevice, ctx->ba_stream));
	CK(cudaMemcpyAsync(W->d_adHost, adHost, (size_t)n2 * 64 * sizeof(double), cudaMemcpyHostToDevice, ctx->ba_stream));
	CK(cudaMemcpyAsync(W->d_adTarget, adTarget, (size_t)n2 * 64 * sizeof(double), cudaMemcpyHostToDevice, ctx->ba_stream));
	// setDeltaF products, read by the linearised-residual mode (addPoint<1>, fixLinearizationF); NULL = zero deltas
	if (adHTdeltaF) CK(cudaMemcpyAsync(W->d_adHTdeltaF, adHTdeltaF, (size_t)n2 * 8 * sizeof(float), cudaMemcpyHostToDevice, ctx->ba_stream));
	else CK(cudaMemsetAsync(W->d_adHTdeltaF, 0, (size_t)n2 * 8 * sizeof(float), ctx->ba_stream));
	if (cDeltaF4) memcpy(W->h_cDeltaF, cDeltaF4, 4 * sizeof(float));
	CK(cudaMemcpyAsync(W->d_cDeltaF, W->h_cDeltaF, 4 * sizeof(float), cudaMemcpyHostToDevice, ctx->ba_stream));
	CK(cudaStreamSynchronize(ctx->ba_stream));
	W->glue_set = true;
	return 0;
}

extern "C" int dso_b200_ba_set_points(dso_b200_ctx* ctx, int nP, const float* u, const float* v, const float* idepth_scaled,
                                      const float* idepth_zero_scaled, const float* color8, const float* weights8,
                                      const int* host, const float* priorF, const float* deltaF)
{
	ARG_CHECK(ctx && ctx->ba && u && v && idepth_scaled && idepth_zero_scaled && color8 && weights8 && host && priorF && deltaF, "null argument");
	ARG_CHECK(nP >= 1, "nP must be >= 1");
	BAWorkspace* W = ctx->ba;
	BA_CHECK(W->window_set, "call dso_b200_ba_set_window first");
	CK(cudaSetDevice(ctx->device));
	const int nF = W->P.nF;
	for (int i = 0; i < nP; i++)
	{
		ARG_CHECK(host[i] >= 0 && host[i] < nF, "point host out of range");
		ARG_CHECK(i == 0 || host[i] >= host[i - 1], "points must be grouped by host frame (EnergyFunctional::allPoints order)");
	}
	if (nP > W->capP)
	{
		free_points(W);
		const size_t n = (size_t)nP;
		CK(cudaMalloc(&W->d_pt, n * sizeof(float4))); CK(cudaMalloc(&W->d_color, n * 8 * sizeof(float))); CK(cudaMalloc(&W->d_weights, n * 8 * sizeof(float)));
		CK(cudaMalloc(&W->d_phost, n * sizeof(int))); CK(cudaMalloc(&W->d_priorF, n * sizeof(float))); CK(cudaMalloc(&W->d_deltaF, n * sizeof(float)));
		CK(cudaMalloc(&W->d_res_begin, (n + 1) * sizeof(int)));
		CK(cudaMalloc(&W->d_Hdd, n * sizeof(float))); CK(cudaMalloc(&W->d_bd, n * sizeof(float))); CK(cudaMalloc(&W->d_Hcd, n * 4 * sizeof(float)));
		CK(cudaMalloc(&W->d_HdiF, n * sizeof(float))); CK(cudaMalloc(&W->d_bdSumF, n * sizeof(float))); CK(cudaMalloc(&W->d_idh, n * sizeof(float)));
		CK(cudaMalloc(&W->d_Hdd_L, n * sizeof(float))); CK(cudaMalloc(&W->d_bd_L, n * sizeof(float))); CK(cudaMalloc(&W->d_Hcd_L, n * 4 * sizeof(float)));
		CK(cudaMalloc(&W->d_HcdSum, n * 4 * sizeof(float)));
		CK(cudaMalloc(&W->d_HdiF_m, n * sizeof(float))); CK(cudaMalloc(&W->d_bdSumF_m, n * sizeof(float)));
		W->capP = nP;
	}
	// *_accLF start at zero: accumulateLF_MT overwrites them when there are linearised residuals
	CK(cudaMemsetAsync(W->d_Hdd_L, 0, (size_t)nP * sizeof(float), ctx->ba_stream));
	CK(cudaMemsetAsync(W->d_bd_L, 0, (size_t)nP * sizeof(float), ctx->ba_stream));
	CK(cudaMemsetAsync(W->d_Hcd_L, 0, (size_t)nP * 4 * sizeof(float), ctx->ba_stream));
	std::vector<float4> pt(nP);
	for (int i = 0; i < nP; i++) pt[i] = make_float4(u[i], v[i], idepth_scaled[i], idepth_zero_scaled[i]);
	CK(cudaMemcpyAsync(W->d_pt, pt.data(), (size_t)nP * sizeof(float4), cudaMemcpyHostToDevice, ctx->ba_stream));
	CK(cudaMemcpyAsync(W->d_color, color8, (size_t)nP * 8 * sizeof(float), cudaMemcpyHostToDevice, ctx->ba_stream));
	CK(cudaMemcpyAsync(W->d_weights, weights8, (size_t)nP * 8 * sizeof(float), cudaMemcpyHostToDevice, ctx->ba_stream));
	CK(cudaMemcpyAsync(W->d_phost, host, (size_t)nP * sizeof(int), cudaMemcpyHostToDevice, ctx->ba_stream));
	CK(cudaMemcpyAsync(W->d_priorF, priorF, (size_t)nP * sizeof(float), cudaMemcpyHostToDevice, ctx->ba_stream));
	CK(cudaMemcpyAsync(W->d_deltaF, deltaF, (size_t)nP * sizeof(float), cudaMemcpyHostToDevice, ctx->ba_stream));
	// Schur chunks: CHUNK_P consecutive points of one host
	std::vector<int> pcs, pcc, pch, hb(nF + 1, 0);
	int i = 0;
	for (int hst = 0; hst < nF; hst++)
	{
		hb[hst] = (int)pcs.size();
		while (i < nP && host[i] == hst)
		{
			int j = i;
			while (j < nP && host[j] == hst && j - i < CHUNK_P) j++;
			pcs.push_back(i); pcc.push_back(j - i); pch.push_back(hst);
			i = j;
		}
	}
	hb[nF] = (int)pcs.size();
	W->nPC = (int)pcs.size();
	W->scN = nF * nF * 64 + nF * 32 + nF * 8 + 20;
	W->scN = (W->scN + 3) & ~3;
	if (W->nPC > W->capPC)
	{
		if (W->capPC) { cudaFree(W->d_pc_start); cudaFree(W->d_pc_cnt); cudaFree(W->d_pc_host); cudaFree(W->d_scPartials); }
		CK(cudaMalloc(&W->d_pc_start, W->nPC * sizeof(int))); CK(cudaMalloc(&W->d_pc_cnt, W->nPC * sizeof(int))); CK(cudaMalloc(&W->d_pc_host, W->nPC * sizeof(int)));
		CK(cudaMalloc(&W->d_scPartials, (size_t)W->nPC * (MAXF * MAXF * 64 + MAXF * 40 + 24) * sizeof(float)));
		W->capPC = W->nPC;
	}
	CK(cudaMemcpyAsync(W->d_pc_start, pcs.data(), pcs.size() * sizeof(int), cudaMemcpyHostToDevice, ctx->ba_stream));
	CK(cudaMemcpyAsync(W->d_pc_cnt, pcc.data(), pcc.size() * sizeof(int), cudaMemcpyHostToDevice, ctx->ba_stream));
	CK(cudaMemcpyAsync(W->d_pc_host, pch.data(), pch.size() * sizeof(int), cudaMemcpyHostToDevice, ctx->ba_stream));
	CK(cudaMemcpyAsync(W->d_host_pc_begin, hb.data(), (nF + 1) * sizeof(int), cudaMemcpyHostToDevice, ctx->ba_stream));
	CK(cudaStreamSynchronize(ctx->ba_stream));
	W->h_phost.assign(host, host + nP);
	W->P.nP = nP;
	W->points_set = true; W->res_set = false; W->lin_done = false; W->top_done = false;
	ba_graphs_invalidate(W);
	return 0;
}

extern "C" int dso_b200_ba_set_residuals(dso_b200_ctx* ctx, int nR, const int* point, const int* target, const int* state_state,
                                         const double* state_energy, const unsigned char* isLinearized, const unsigned char* isActive)
{
	ARG_CHECK(ctx && ctx->ba && point && target && state_state, "null argument");
	ARG_CHECK(nR >= 1, "nR must be >= 1");
	BAWorkspace* W = ctx->ba;
	BA_CHECK(W->points_set, "call dso_b200_ba_set_points first");
	CK(cudaSetDevice(ctx->device));
	const int nF = W->P.nF, nP = W->P.nP;
	std::vector<int> rb(nP + 1, 0);
	for (int i = 0; i < nR; i++)
	{
		ARG_CHECK(point[i] >= 0 && point[i] < nP && target[i] >= 0 && target[i] < nF, "residual point / target out of range");
		ARG_CHECK(i == 0 || point[i] >= point[i - 1], "residuals must be grouped by point (EFPoint::residualsAll order)");
		ARG_CHECK(target[i] != W->h_phost[point[i]], "residual target equals its host");
		if (isLinearized && isLinearized[i] && isActive && isActive[i])
		{
			dso_set_error("dso_b200_ba_set_residuals: active linearised residuals cannot be imported (their res_toZeroF and Jacobians live on the device): linearise them with dso_b200_ba_fix_linearization");
			return DSO_B200_E_STATE;
		}
		rb[point[i] + 1]++;
	}
	for (int p = 0; p < nP; p++) rb[p + 1] += rb[p];
	if (nR > W->capR)
	{
		free_res(W);
		const size_t n = (size_t)nR;
		CK(cudaMalloc(&W->d_rpoint, n * sizeof(int))); CK(cudaMalloc(&W->d_rtarget, n * sizeof(int))); CK(cudaMalloc(&W->d_state, n * sizeof(int)));
		CK(cudaMalloc(&W->d_rhost, n * sizeof(int)));
		CK(cudaMalloc(&W->d_newState, n * sizeof(int))); CK(cudaMalloc(&W->d_energy, n * sizeof(double))); CK(cudaMalloc(&W->d_newEnergy, n * sizeof(float)));
		CK(cudaMalloc(&W->d_newEnergyWO, n * sizeof(float))); CK(cudaMalloc(&W->d_isLin, n)); CK(cudaMalloc(&W->d_isAct, n)); CK(cudaMalloc(&W->d_efsel, n));
		CK(cudaMalloc(&W->d_cpt, n * 3 * sizeof(float))); CK(cudaMalloc(&W->d_proj, n * 16 * sizeof(float)));
		CK(cudaMalloc(&W->d_J[0], n * JSZ * sizeof(float))); CK(cudaMalloc(&W->d_J[1], n * JSZ * sizeof(float)));
		CK(cudaMalloc(&W->d_JpJdF, n * 8 * sizeof(float))); CK(cudaMalloc(&W->d_contrib, n * 6 * sizeof(float))); CK(cudaMalloc(&W->d_resEnergy, n * sizeof(double)));
		CK(cudaMalloc(&W->d_perm, n * sizeof(int)));
		CK(cudaMalloc(&W->d_subset, n * sizeof(int)));
		CK(cudaMalloc(&W->d_energy_blocks, ((n + 255) / 256 + 1) * sizeof(double)));      // + the fused kernel's block counter
		CK(cudaMemset(W->d_energy_blocks, 0, ((n + 255) / 256 + 1) * sizeof(double)));
		CK(cudaMalloc(&W->d_resToZero, n * 8 * sizeof(float)));
		W->capR = nR;
	}
	// pair-sorted permutation (stable: keeps point order inside a pair, the reference's accumulation order)
	std::vector<int> perm(nR), key(nR);
	for (int i = 0; i < nR; i++) { perm[i] = i; key[i] = W->h_phost[point[i]] + nF * target[i]; }
	std::stable_sort(perm.begin(), perm.end(), [&](int a, int b) { return key[a] < key[b]; });
	std::vector<int> cs, cc, pcb(nF * nF + 1, 0);
	int i = 0;
	for (int k = 0; k < nF * nF; k++)
	{
		pcb[k] = (int)cs.size();
		while (i < nR && key[perm[i]] == k)
		{
			int j = i;
			while (j < nR && key[perm[j]] == k && j - i < CHUNK_R) j++;
			cs.push_back(i); cc.push_back(j - i);
			i = j;
		}
	}
	pcb[nF * nF] = (int)cs.size();
	W->nChunks = (int)cs.size();
	if (W->nChunks > W->capChunks)
	{
		if (W->capChunks) { cudaFree(W->d_chunk_start); cudaFree(W->d_chunk_cnt); cudaFree(W->d_topPartials); cudaFree(W->d_chunk_nres); }
		CK(cudaMalloc(&W->d_chunk_start, W->nChunks * sizeof(int))); CK(cudaMalloc(&W->d_chunk_cnt, W->nChunks * sizeof(int)));
		CK(cudaMalloc(&W->d_topPartials, (size_t)W->nChunks * NTOP_PAD * sizeof(float))); CK(cudaMalloc(&W->d_chunk_nres, W->nChunks * sizeof(int)));
		CK(cudaMemset(W->d_topPartials, 0, (size_t)W->nChunks * NTOP_PAD * sizeof(float)));
		W->capChunks = W->nChunks;
	}
	std::vector<double> en(nR, 0.0);
	std::vector<float> nen(nR, 0.f);
	std::vector<unsigned char> lin(nR, 0), act(nR, 0);
	for (int k = 0; k < nR; k++)
	{
		if (state_energy) en[k] = state_energy[k];
		nen[k] = (float)en[k];
		if (isLinearized) lin[k] = isLinearized[k];
		if (isActive) act[k] = isActive[k];
	}
	cudaStream_t s = ctx->ba_stream;
	CK(cudaMemcpyAsync(W->d_rpoint, point, (size_t)nR * sizeof(int), cudaMemcpyHostToDevice, s));
	CK(cudaMemcpyAsync(W->d_rtarget, target, (size_t)nR * sizeof(int), cudaMemcpyHostToDevice, s));
	{
		// host frame per residual: the linearisation reaches the (host, target) precalc without the point -> host hop
		std::vector<int> rh((size_t)nR);
		for (int i = 0; i < nR; i++) rh[i] = W->h_phost[point[i]];
		CK(cudaMemcpyAsync(W->d_rhost, rh.data(), (size_t)nR * sizeof(int), cudaMemcpyHostToDevice, s));
		CK(cudaStreamSynchronize(s));          // rh leaves scope
	}
	CK(cudaMemcpyAsync(W->d_state, state_state, (size_t)nR * sizeof(int), cudaMemcpyHostToDevice, s));
	CK(cudaMemcpyAsync(W->d_energy, en.data(), (size_t)nR * sizeof(double), cudaMemcpyHostToDevice, s));
	CK(cudaMemcpyAsync(W->d_newEnergy, nen.data(), (size_t)nR * sizeof(float), cudaMemcpyHostToDevice, s));
	CK(cudaMemcpyAsync(W->d_isLin, lin.data(), (size_t)nR, cudaMemcpyHostToDevice, s));
	CK(cudaMemcpyAsync(W->d_isAct, act.data(), (size_t)nR, cudaMemcpyHostToDevice, s));
	CK(cudaMemcpyAsync(W->d_perm, perm.data(), (size_t)nR * sizeof(int), cudaMemcpyHostToDevice, s));
	CK(cudaMemcpyAsync(W->d_res_begin, rb.data(), (size_t)(nP + 1) * sizeof(int), cudaMemcpyHostToDevice, s));
	CK(cudaMemcpyAsync(W->d_chunk_start, cs.data(), cs.size() * sizeof(int), cudaMemcpyHostToDevice, s));
	CK(cudaMemcpyAsync(W->d_chunk_cnt, cc.data(), cc.size() * sizeof(int), cudaMemcpyHostToDevice, s));
	CK(cudaMemcpyAsync(W->d_pair_chunk_begin, pcb.data(), pcb.size() * sizeof(int), cudaMemcpyHostToDevice, s));
	CK(cudaMemsetAsync(W->d_efsel, 0, nR, s));
	CK(cudaMemsetAsync(W->d_resToZero, 0, (size_t)nR * 8 * sizeof(float), s));
	W->nLin = 0;
	{ std::vector<int> ns(nR, ST_OUTLIER); CK(cudaMemcpyAsync(W->d_newState, ns.data(), (size_t)nR * sizeof(int), cudaMemcpyHostToDevice, s)); CK(cudaStreamSynchronize(s)); }   // resetOOB, Residuals.h:97-105
	CK(cudaMemsetAsync(W->d_newEnergyWO, 0, (size_t)nR * sizeof(float), s));
	CK(cudaMemsetAsync(W->d_cpt, 0, (size_t)nR * 3 * sizeof(float), s));
	CK(cudaMemsetAsync(W->d_proj, 0, (size_t)nR * 16 * sizeof(float), s));
	CK(cudaMemsetAsync(W->d_J[0], 0, (size_t)nR * JSZ * sizeof(float), s));
	CK(cudaMemsetAsync(W->d_J[1], 0, (size_t)nR * JSZ * sizeof(float), s));
	CK(cudaMemsetAsync(W->d_JpJdF, 0, (size_t)nR * 8 * sizeof(float), s));
	CK(cudaMemsetAsync(W->d_contrib, 0, (size_t)nR * 6 * sizeof(float), s));
	CK(cudaStreamSynchronize(s));
	W->P.nR = nR;
	W->res_set = true; W->lin_done = false; W->top_done = false;
	ba_graphs_invalidate(W);
	return 0;
}


// ---- host-side glue: FrameFramePrecalc::set + EnergyFunctional::setAdjointsF ------------------
// Eigen's compute_inverse<Matrix3f> (cofactors of column 0, determinant by expansion along it)
static float cof33h(const float* m, int i, int j)
{
	int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
	volatile float a = m[i1 * 3 + j1] * m[i2 * 3 + j2];
	volatile float b = m[i1 * 3 + j2] * m[i2 * 3 + j1];
	return a - b;
}
static void mat33f_inverse_h(const float* m, float* out)
{
	float c0 = cof33h(m, 0, 0), c1 = cof33h(m, 1, 0), c2 = cof33h(m, 2, 0);
	float det = dot3h(c0, m[0], c1, m[3], c2, m[6]);
	float invdet = 1.0f / det;
	out[0] = c0 * invdet; out[1] = c1 * invdet; out[2] = c2 * invdet;
	out[3] = cof33h(m, 0, 1) * invdet; out[4] = cof33h(m, 1, 1) * invdet; out[5] = cof33h(m, 2, 1) * invdet;
	out[6] = cof33h(m, 0, 2) * invdet; out[7] = cof33h(m, 1, 2) * invdet; out[8] = cof33h(m, 2, 2) * invdet;
}

extern "C" int dso_b200_ba_set_frames(dso_b200_ctx* ctx, const double* evalPT12, const double* state10, const double* state_zero10,
                                      const float* ab_exposure, dso_b200_precalc* precalc_out, double* adHost_out, double* adTarget_out)
{
	ARG_CHECK(ctx && ctx->ba && evalPT12 && state10 && state_zero10 && ab_exposure, "null argument");
	BAWorkspace* W = ctx->ba;
	BA_CHECK(W->window_set, "call dso_b200_ba_set_window first");
	const int nF = W->P.nF;
	const float SXT = 0.5f, SXR = 1.0f, SA = 10.0f, SB = 1000.0f;   // SCALE_XI_TRANS, SCALE_XI_ROT, SCALE_A, SCALE_B (HessianBlocks.h:54-61)
	std::vector<Pose> evalPT(nF), w2c(nF), c2w(nF);
	std::vector<double> scaled(nF * 10);
	for (int f = 0; f < nF; f++)
	{
		pose_from12(evalPT[f], evalPT12 + 12 * f);
		const double* st = state10 + 10 * f;
		double* sc = &scaled[10 * f];
		// FrameHessian::setState, HessianBlocks.h:183-199
		for (int i = 0; i < 3; i++) sc[i] = SXT * st[i];
		for (int i = 3; i < 6; i++) sc[i] = SXR * st[i];
		sc[6] = SA * st[6]; sc[7] = SB * st[7]; sc[8] = SA * st[8]; sc[9] = SB * st[9];
		Pose e;
		se3_exp(sc, e);
		pose_mul(e, evalPT[f], w2c[f]);
		pose_inverse(w2c[f], c2w[f]);
	}
	std::vector<dso_b200_precalc> pre(nF * nF);
	std::vector<double> adH((size_t)nF * nF * 64), adT((size_t)nF * nF * 64);
	const float K[9] = { W->P.fx, 0, W->P.cx, 0, W->P.fy, W->P.cy, 0, 0, 1 };
	float Kinv[9];
	mat33f_inverse_h(K, Kinv);
	for (int h = 0; h < nF; h++)
		for (int t = 0; t < nF; t++)
		{
			// FrameFramePrecalc::set, HessianBlocks.cpp:213-250
			dso_b200_precalc& P = pre[h * nF + t];
			Pose inv, l0, l;
			pose_inverse(evalPT[h], inv);
			pose_mul(evalPT[t], inv, l0);
			pose_mul(w2c[t], c2w[h], l);
			float RTll[9], tTll[3], KR[9];
			for (int i = 0; i < 9; i++) { P.PRE_RTll_0[i] = (float)l0.R[i]; RTll[i] = (float)l.R[i]; }
			for (int i = 0; i < 3; i++) { P.PRE_tTll_0[i] = (float)l0.t[i]; tTll[i] = (float)l.t[i]; }
			mat33f_mul(K, RTll, KR);
			mat33f_mul(KR, Kinv, P.PRE_KRKiTll);
			for (int r = 0; r < 3; r++) P.PRE_KtTll[r] = dot3h(K[r * 3], tTll[0], K[r * 3 + 1], tTll[1], K[r * 3 + 2], tTll[2]);
			double a, b;
			aff_from_to(ab_exposure[h], ab_exposure[t], scaled[10 * h + 6], scaled[10 * h + 7], scaled[10 * t + 6], scaled[10 * t + 7], a, b);
			P.PRE_aff_mode[0] = (float)a; P.PRE_aff_mode[1] = (float)b;
			P.PRE_b0_mode = (float)(state_zero10[10 * h + 7] * SB);
			P.pad = 0;
			// setAdjointsF, EnergyFunctional.cpp:57-94: AH = -Adj(hostToTarget)^T, AT = I (6x6), affine rows, SCALE_* on rows
			double* AH = &adH[(size_t)(h + t * nF) * 64];
			double* AT = &adT[(size_t)(h + t * nF) * 64];
			for (int i = 0; i < 64; i++) { AH[i] = (i % 9 == 0) ? 1.0 : 0.0; AT[i] = (i % 9 == 0) ? 1.0 : 0.0; }
			double Ad[36];
			pose_adjoint(l0, Ad);
			for (int r = 0; r < 6; r++) for (int c = 0; c < 6; c++) AH[r * 8 + c] = -Ad[c * 6 + r];
			double a0, b0;
			aff_from_to(ab_exposure[h], ab_exposure[t], state_zero10[10 * h + 6] * SA, state_zero10[10 * h + 7] * SB,
			            state_zero10[10 * t + 6] * SA, state_zero10[10 * t + 7] * SB, a0, b0);
			const float affLL0 = (float)a0;
			AT[6 * 8 + 6] = -affLL0; AH[6 * 8 + 6] = affLL0; AT[7 * 8 + 7] = -1; AH[7 * 8 + 7] = affLL0;
			for (int c = 0; c < 8; c++)
			{
				for (int r = 0; r < 3; r++) { AH[r * 8 + c] *= SXT; AT[r * 8 + c] *= SXT; }
				for (int r = 3; r < 6; r++) { AH[r * 8 + c] *= SXR; AT[r * 8 + c] *= SXR; }
				AH[6 * 8 + c] *= SA; AT[6 * 8 + c] *= SA;
				AH[7 * 8 + c] *= SB; AT[7 * 8 + c] *= SB;
			}
		}
	if (precalc_out) memcpy(precalc_out, pre.data(), pre.size() * sizeof(dso_b200_precalc));
	if (adHost_out) memcpy(adHost_out, adH.data(), adH.size() * sizeof(double));
	if (adTarget_out) memcpy(adTarget_out, adT.data(), adT.size() * sizeof(double));
	// setDeltaF, EnergyFunctional.cpp:178-193: adHTdeltaF[h+t*nF] = delta_h^T adHostF + delta_t^T adTargetF in float, every
	// coefficient an unrolled 8-term reduction ((p0+p1)+(p2+p3))+((p4+p5)+(p6+p7))
	std::vector<float> adHTd((size_t)nF * nF * 8);
	for (int h = 0; h < nF; h++)
		for (int t = 0; t < nF; t++)
		{
			const int idx = h + t * nF;
			for (int c = 0; c < 8; c++)
			{
				volatile float p0[8], p1[8];
				for (int r = 0; r < 8; r++)
				{
					p0[r] = (float)(state10[10 * h + r] - state_zero10[10 * h + r]) * (float)adH[(size_t)idx * 64 + r * 8 + c];
					p1[r] = (float)(state10[10 * t + r] - state_zero10[10 * t + r]) * (float)adT[(size_t)idx * 64 + r * 8 + c];
				}
				const float s0 = ((p0[0] + p0[1]) + (p0[2] + p0[3])) + ((p0[4] + p0[5]) + (p0[6] + p0[7]));
				const float s1 = ((p1[0] + p1[1]) + (p1[2] + p1[3])) + ((p1[4] + p1[5]) + (p1[6] + p1[7]));
				adHTd[(size_t)idx * 8 + c] = s0 + s1;
			}
		}
	// EFFrame::takeData (EnergyFunctionalStructs.cpp:55-63): delta = state - state_zero, delta_prior = state - getPriorZero() (= 0)
	{
		double dl[8 * MAXF + CPARS], dp[8 * MAXF + CPARS];
		for (int i = 0; i < CPARS; i++) dl[i] = dp[i] = 0.0;
		for (int f = 0; f < nF; f++)
			for (int i = 0; i < 8; i++)
			{
				dl[CPARS + 8 * f + i] = state10[10 * f + i] - state_zero10[10 * f + i];
				dp[CPARS + 8 * f + i] = state10[10 * f + i];
			}
		CK(cudaSetDevice(ctx->device));
		CK(cudaMemcpyAsync(W->d_delta, dl, (8 * nF + CPARS) * sizeof(double), cudaMemcpyHostToDevice, ctx->ba_stream));
		CK(cudaMemcpyAsync(W->d_priorD, dp, (8 * nF + CPARS) * sizeof(double), cudaMemcpyHostToDevice, ctx->ba_stream));
		CK(cudaStreamSynchronize(ctx->ba_stream));
	}
	return dso_b200_ba_set_glue(ctx, pre.data(), adH.data(), adT.data(), adHTd.data(), nullptr);
}

// The step's launch sequence is not one chain: the energy sum only has to be ready at the end, and the top-Hessian stitch
// (three kernels) is independent of the Schur-complement chain (five).  With fork = true those side branches go to a second
// stream (captured into the same CUDA graph as parallel branches); ba_join_side makes the main stream wait for them.
static int ba_fork(dso_b200_ctx* ctx)
{
	BAWorkspace* W = ctx->ba;
	CK(cudaEventRecord(W->ev_fork, ctx->ba_stream));
	CK(cudaStreamWaitEvent(W->side_stream, W->ev_fork, 0));
	W->side_busy = true;
	return 0;
}
int ba_join_side(dso_b200_ctx* ctx)
{
	BAWorkspace* W = ctx->ba;
	if (!W->side_busy) return 0;
	CK(cudaEventRecord(W->ev_join, W->side_stream));
	CK(cudaStreamWaitEvent(ctx->ba_stream, W->ev_join, 0));
	W->side_busy = false;
	return 0;
}

int ba_enqueue_linearize(dso_b200_ctx* ctx, bool fork, bool apply)
{
	BAWorkspace* W = ctx->ba;
	const int nR = W->P.nR;
	if (apply)
		ba_linearize_kernel<true><<<(nR + 31) / 32, 256, 0, ctx->ba_stream>>>(W->P, W->d_precalc, W->d_pt, W->d_color, W->d_weights, W->d_phost, W->d_rpoint,
		                                                                 W->d_rtarget, W->d_state, W->d_energy, W->d_efsel, W->d_J[0], W->d_J[1],
		                                                                 W->d_newState, W->d_newEnergy, W->d_newEnergyWO, W->d_cpt, W->d_proj, W->d_resEnergy,
		                                                                 nullptr, 0, 0, W->d_isAct, W->d_JpJdF, W->d_contrib, W->d_rhost);
	else
		ba_linearize_kernel<false><<<(nR + 31) / 32, 256, 0, ctx->ba_stream>>>(W->P, W->d_precalc, W->d_pt, W->d_color, W->d_weights, W->d_phost, W->d_rpoint,
		                                                                  W->d_rtarget, W->d_state, W->d_energy, W->d_efsel, W->d_J[0], W->d_J[1],
		                                                                  W->d_newState, W->d_newEnergy, W->d_newEnergyWO, W->d_cpt, W->d_proj, W->d_resEnergy,
		                                                                  nullptr, 0, 0, nullptr, nullptr, nullptr, W->d_rhost);
	if (apply) W->top_done = false;
	const int nb = (nR + 255) / 256;
	cudaStream_t se = ctx->ba_stream;
	if (fork) { int rc = ba_fork(ctx); if (rc) return rc; se = W->side_stream; }
	ba_energy_kernel<<<nb, 256, 0, se>>>(nR, W->d_resEnergy, W->d_energy_blocks, reinterpret_cast<unsigned int*>(W->d_energy_blocks + nb), W->d_scalars);
	LAUNCHED(ctx, 2);
	CK(cudaGetLastError());
	W->lin_done = true;
	W->contrib_from_lin = apply;       // the fused form has written the per-point terms of every active residual
	return 0;
}

int ba_enqueue_apply(dso_b200_ctx* ctx)
{
	BAWorkspace* W = ctx->ba;
	const int nR = W->P.nR;
	ba_apply_res_kernel<<<(nR + 255) / 256, 256, 0, ctx->ba_stream>>>(nR, W->d_state, W->d_newState, W->d_energy, W->d_newEnergy, W->d_isAct, W->d_efsel,
	                                                             W->d_J[0], W->d_J[1], W->d_JpJdF, nullptr, 0);
	LAUNCHED(ctx, 1);
	CK(cudaGetLastError());
	W->top_done = false;
	return 0;
}

// mode 0: accumulateAF_MT -> d_HA, d_bA, per-point AF sums; mode 1: accumulateLF_MT -> d_HL, d_bL (+ priors), per-point LF sums
int ba_enqueue_top(dso_b200_ctx* ctx, int mode, const unsigned char* d_pmask, bool fork, bool prep)
{
	BAWorkspace* W = ctx->ba;
	const int nF = W->P.nF;
	cudaStream_t st = ctx->ba_stream;
	BALinArgs LA;
	LA.rpoint = W->d_rpoint; LA.rtarget = W->d_rtarget; LA.phost = W->d_phost; LA.deltaF = W->d_deltaF;
	LA.adHTdeltaF = W->d_adHTdeltaF; LA.cDeltaF = W->d_cDeltaF; LA.resToZero = W->d_resToZero; LA.nF = nF; LA.pmask = d_pmask;
	float* Hdd = mode ? W->d_Hdd_L : W->d_Hdd; float* bd = mode ? W->d_bd_L : W->d_bd; float* Hcd = mode ? W->d_Hcd_L : W->d_Hcd;
	double* Hout = mode ? W->d_HL : W->d_HA; double* bout = mode ? W->d_bL : W->d_bA;
	// mode 0 right after the fused linearisation: the per-point terms are already there, so the per-point sums (and with
	// them the whole Schur chain) do not wait for the top accumulation, which moves to the side branch with its stitch
	const bool early_sums = mode == 0 && fork && W->contrib_from_lin && !d_pmask;
	if (mode == 2) W->contrib_from_lin = false;               // addPoint<2> rewrites the terms of the marginalised points
	if (early_sums) { int rc = ba_fork(ctx); if (rc) return rc; st = W->side_stream; }
	if (mode == 2)
		ba_accum_top_kernel<2><<<W->nChunks, 128, 0, st>>>(W->d_perm, W->d_chunk_start, W->d_chunk_cnt, W->d_isAct, W->d_isLin, W->d_efsel,
		                                                 W->d_J[0], W->d_J[1], W->d_contrib, W->d_topPartials, W->d_chunk_nres, LA);
	else if (mode == 0)
		ba_accum_top_kernel<0><<<W->nChunks, 128, 0, st>>>(W->d_perm, W->d_chunk_start, W->d_chunk_cnt, W->d_isAct, W->d_isLin, W->d_efsel,
		                                                 W->d_J[0], W->d_J[1], early_sums ? nullptr : W->d_contrib, W->d_topPartials, W->d_chunk_nres, LA);
	else
		ba_accum_top_kernel<1><<<W->nChunks, 128, 0, st>>>(W->d_perm, W->d_chunk_start, W->d_chunk_cnt, W->d_isAct, W->d_isLin, W->d_efsel,
		                                                 W->d_J[0], W->d_J[1], W->d_contrib, W->d_topPartials, W->d_chunk_nres, LA);
	SchurPrepOut po;
	po.HdiF = W->d_HdiF; po.bdSumF = W->d_bdSumF; po.idh = W->d_idh; po.HcdSum = W->d_HcdSum; po.priorF = W->d_priorF; po.deltaF = W->d_deltaF;
	if (prep && (mode == 2 || d_pmask)) { dso_set_error("ba_enqueue_top: the fused Schur preparation has no marginalisation form"); return DSO_B200_E_ARG; }
	// no linearised residuals to add afterwards (prep with early sums): the Schur accumulation, which is next on the main
	// stream, computes the per-point sums and the preparation of its own points itself -- one launch less on the critical path
	W->sums_in_schur = early_sums && prep;
	if (!W->sums_in_schur)
	ba_point_sums_kernel<<<(W->P.nP + 127) / 128, 128, 0, early_sums ? ctx->ba_stream : st>>>(W->P.nP, W->d_res_begin, W->d_isAct, W->d_isLin, W->d_contrib, Hdd, bd, Hcd, mode, d_pmask, W->d_Hdd, W->d_bd, W->d_Hcd,
	                                                            prep, W->d_Hdd_L, W->d_bd_L, W->d_Hcd_L, po);
	if (fork && !early_sums) { int rc = ba_fork(ctx); if (rc) return rc; st = W->side_stream; }       // the stitch: independent of what follows on the main stream
	const bool fused_reduce = W->nChunks <= 48 * nF * nF;            // few chunks per pair: the pair kernel sums them itself
	if (!fused_reduce)
		ba_reduce_partials_kernel<<<dim3(1, nF * nF), NTOP_PAD, 0, st>>>(W->d_topPartials, NTOP_PAD, NTOP_PAD, W->d_pair_chunk_begin, W->d_topRed);
	ba_top_pair_kernel<<<nF * nF, 256, 0, st>>>(nF, W->d_topRed, W->d_adHost, W->d_adTarget, W->d_topContrib,
	                                          fused_reduce ? W->d_topPartials : nullptr, W->d_pair_chunk_begin);
	ba_top_gather_kernel<<<nF * nF + nF + 2, 64, 0, st>>>(nF, W->nChunks, W->d_chunk_nres, W->d_topContrib, Hout, bout, W->d_scalars + (mode ? 2 : 0),
	                                               mode == 1 ? W->d_priorH : nullptr, W->d_priorD, W->d_cDeltaF);
	LAUNCHED(ctx, (fused_reduce ? 4 : 5) - (W->sums_in_schur ? 1 : 0));
	CK(cudaGetLastError());
	if (mode == 0) W->top_done = true;
	return 0;
}

int ba_enqueue_schur(dso_b200_ctx* ctx, const unsigned char* d_pmask, bool prep_done)
{
	BAWorkspace* W = ctx->ba;
	const int nF = W->P.nF, nP = W->P.nP;
	float* HdiF = d_pmask ? W->d_HdiF_m : W->d_HdiF; float* bdSumF = d_pmask ? W->d_bdSumF_m : W->d_bdSumF;
	if (!prep_done)
		ba_schur_prep_kernel<<<(nP + 127) / 128, 128, 0, ctx->ba_stream>>>(nP, W->d_res_begin, W->d_isAct, W->d_Hdd, W->d_bd, W->d_Hcd,
		                                                              W->d_Hdd_L, W->d_bd_L, W->d_Hcd_L, W->d_priorF, W->d_deltaF,
		                                                              HdiF, bdSumF, W->d_idh, W->d_HcdSum, d_pmask);
	SchurFusedSums fs = SchurFusedSums();
	if (W->sums_in_schur)
	{
		BA_CHECK(prep_done && !d_pmask, "internal: fused per-point sums without the matching top accumulation");
		fs.contrib = W->d_contrib; fs.isLin = W->d_isLin; fs.Hdd = W->d_Hdd; fs.bd = W->d_bd; fs.Hcd = W->d_Hcd;
		fs.Hdd_L = W->d_Hdd_L; fs.bd_L = W->d_bd_L; fs.Hcd_L = W->d_Hcd_L; fs.priorF = W->d_priorF; fs.deltaF = W->d_deltaF;
		fs.HdiF = W->d_HdiF; fs.bdSumF = W->d_bdSumF; fs.idh = W->d_idh; fs.HcdSum = W->d_HcdSum;
		W->sums_in_schur = false;
	}
	ba_schur_accum_kernel<<<W->nPC, 256, 0, ctx->ba_stream>>>(nF, W->d_pc_start, W->d_pc_cnt, W->d_res_begin, W->d_rtarget, W->d_isAct, W->d_JpJdF,
	                                                     HdiF, bdSumF, W->d_HcdSum, W->d_scPartials, W->scN, fs);
	const int nRed = nF * nF * 64 + nF * 32 + nF * 8 + 20;          // accD, accE, accEB, accHcc + accbc per host
	// few chunks per host (a VGA window has ~18): stage 1 sums them itself, one launch less on the critical path; many
	// chunks: the reduction runs as its own, much wider launch (stage 1 has only nF^2 blocks)
	const bool fused_reduce = W->nPC <= 48 * nF;
	if (!fused_reduce)
		ba_reduce_partials_kernel<<<dim3((nRed + 255) / 256, nF), 256, 0, ctx->ba_stream>>>(W->d_scPartials, W->scN, nRed, W->d_host_pc_begin, W->d_scRed);
	ba_schur_stage1_kernel<<<nF * nF, 64 * MAXF, 0, ctx->ba_stream>>>(nF, fused_reduce ? W->d_scPartials : nullptr, W->scN, W->d_host_pc_begin, W->d_scRed, nRed,
	                                                               W->d_adHost, W->d_adTarget, W->d_scContrib);
	ba_schur_gather_kernel<<<nF * nF + nF + 1, 64, 0, ctx->ba_stream>>>(nF, W->d_scRed, W->d_scContrib, W->d_HS, W->d_bS);
	LAUNCHED(ctx, (prep_done ? 3 : 4) + (fused_reduce ? 0 : 1));
	CK(cudaGetLastError());
	return 0;
}

int ba_check_ready(dso_b200_ctx* ctx, const char* fn)
{
	if (!ctx || !ctx->ba) { dso_set_error("%s: null context", fn); return DSO_B200_E_ARG; }
	BAWorkspace* W = ctx->ba;
	if (!(W->window_set && W->glue_set && W->points_set && W->res_set))
	{
		dso_set_error("%s: window, glue, points and residuals must be set first", fn);
		return DSO_B200_E_STATE;
	}
	return 0;
}

extern "C" int dso_b200_ba_linearize(dso_b200_ctx* ctx, int* newState, float* newEnergy, float* newEnergyWithOutlier,
                                     float* centerProjectedTo3, float* projectedTo16, float* J74, double* energy_sum)
{
	int rc = ba_check_ready(ctx, __func__);
	if (rc) return rc;
	CK(cudaSetDevice(ctx->device));
	BAWorkspace* W = ctx->ba;
	const size_t nR = W->P.nR;
	rc = ba_enqueue_linearize(ctx);
	if (rc) return rc;
	cudaStream_t s = ctx->ba_stream;
	if (newState) CK(cudaMemcpyAsync(newState, W->d_newState, nR * sizeof(int), cudaMemcpyDeviceToHost, s));
	if (newEnergy) CK(cudaMemcpyAsync(newEnergy, W->d_newEnergy, nR * sizeof(float), cudaMemcpyDeviceToHost, s));
	if (newEnergyWithOutlier) CK(cudaMemcpyAsync(newEnergyWithOutlier, W->d_newEnergyWO, nR * sizeof(float), cudaMemcpyDeviceToHost, s));
	if (centerProjectedTo3) CK(cudaMemcpyAsync(centerProjectedTo3, W->d_cpt, nR * 3 * sizeof(float), cudaMemcpyDeviceToHost, s));
	if (projectedTo16) CK(cudaMemcpyAsync(projectedTo16, W->d_proj, nR * 16 * sizeof(float), cudaMemcpyDeviceToHost, s));
	if (energy_sum) CK(cudaMemcpyAsync(energy_sum, W->d_scalars, sizeof(double), cudaMemcpyDeviceToHost, s));
	CK(cudaStreamSynchronize(s));
	if (J74)
	{
		// PointFrameResidual::J of residual r is buffer (1 - efsel[r])
		std::vector<unsigned char> sel(nR);
		std::vector<float> j0(nR * JSZ), j1(nR * JSZ);
		CK(cudaMemcpy(sel.data(), W->d_efsel, nR, cudaMemcpyDeviceToHost));
		CK(cudaMemcpy(j0.data(), W->d_J[0], nR * JSZ * sizeof(float), cudaMemcpyDeviceToHost));
		CK(cudaMemcpy(j1.data(), W->d_J[1], nR * JSZ * sizeof(float), cudaMemcpyDeviceToHost));
		for (size_t r = 0; r < nR; r++) memcpy(J74 + r * JSZ, (sel[r] ? j0.data() : j1.data()) + r * JSZ, JSZ * sizeof(float));
	}
	return 0;
}

// PointFrameResidual::linearize / applyRes per object, as a batch of k residuals: FullSystem::flagPointsForRemoval
// (FullSystem.cpp:776-787) does r->resetOOB(); r->linearize(&Hcalib); r->applyRes(true) for the residuals of every point it
// marginalises, one by one; the host mirror collects them and forwards one batch.
extern "C" int dso_b200_ba_linearize_residuals(dso_b200_ctx* ctx, int k, const int* residual_idx, int reset_oob, int apply,
                                               int* newState_out, float* newEnergy_out, unsigned char* isActive_out)
{
	int rc = ba_check_ready(ctx, __func__);
	if (rc) return rc;
	BAWorkspace* W = ctx->ba;
	ARG_CHECK(k >= 0 && (k == 0 || residual_idx), "bad arguments");
	if (k == 0) return 0;
	const int nR = W->P.nR;
	for (int q = 0; q < k; q++) ARG_CHECK(residual_idx[q] >= 0 && residual_idx[q] < nR, "residual index out of range");
	CK(cudaSetDevice(ctx->device));
	cudaStream_t s = ctx->ba_stream;
	BA_CHECK(k <= nR, "more indices than residuals");
	int* d_sub = W->d_subset;
	CK(cudaMemcpyAsync(d_sub, residual_idx, (size_t)k * sizeof(int), cudaMemcpyHostToDevice, s));
	W->contrib_from_lin = false;            // the subset's Jacobians change under the per-point terms
	ba_linearize_kernel<false><<<(k + 31) / 32, 256, 0, s>>>(W->P, W->d_precalc, W->d_pt, W->d_color, W->d_weights, W->d_phost, W->d_rpoint,
	                                                  W->d_rtarget, W->d_state, W->d_energy, W->d_efsel, W->d_J[0], W->d_J[1],
	                                                  W->d_newState, W->d_newEnergy, W->d_newEnergyWO, W->d_cpt, W->d_proj, W->d_resEnergy,
	                                                  d_sub, k, reset_oob, nullptr, nullptr);
	LAUNCHED(ctx, 1);
	if (apply)
	{
		ba_apply_res_kernel<<<(k + 255) / 256, 256, 0, s>>>(k, W->d_state, W->d_newState, W->d_energy, W->d_newEnergy, W->d_isAct, W->d_efsel,
		                                                    W->d_J[0], W->d_J[1], W->d_JpJdF, d_sub, reset_oob);
		LAUNCHED(ctx, 1);
		W->top_done = false;
	}
	CK(cudaGetLastError());
	if (newState_out || newEnergy_out || isActive_out)
	{
		std::vector<int> ns(nR); std::vector<float> ne(nR); std::vector<unsigned char> ia(nR);
		CK(cudaMemcpyAsync(ns.data(), W->d_newState, (size_t)nR * sizeof(int), cudaMemcpyDeviceToHost, s));
		CK(cudaMemcpyAsync(ne.data(), W->d_newEnergy, (size_t)nR * sizeof(float), cudaMemcpyDeviceToHost, s));
		CK(cudaMemcpyAsync(ia.data(), W->d_isAct, (size_t)nR, cudaMemcpyDeviceToHost, s));
		CK(cudaStreamSynchronize(s));
		for (int q = 0; q < k; q++)
		{
			if (newState_out) newState_out[q] = ns[residual_idx[q]];
			if (newEnergy_out) newEnergy_out[q] = ne[residual_idx[q]];
			if (isActive_out) isActive_out[q] = ia[residual_idx[q]];
		}
	}
	else
		CK(cudaStreamSynchronize(s));
	ba_graphs_invalidate(W);      // nothing captured depends on it, but the residual states changed outside the captured sequence
	return 0;
}

extern "C" int dso_b200_ba_apply_res(dso_b200_ctx* ctx, unsigned char* isActive_out, float* JpJdF8_out)
{
	int rc = ba_check_ready(ctx, __func__);
	if (rc) return rc;
	BAWorkspace* W = ctx->ba;
	BA_CHECK(W->lin_done, "call dso_b200_ba_linearize first");
	CK(cudaSetDevice(ctx->device));
	rc = ba_enqueue_apply(ctx);
	if (rc) return rc;
	const size_t nR = W->P.nR;
	if (isActive_out) CK(cudaMemcpyAsync(isActive_out, W->d_isAct, nR, cudaMemcpyDeviceToHost, ctx->ba_stream));
	if (JpJdF8_out) CK(cudaMemcpyAsync(JpJdF8_out, W->d_JpJdF, nR * 8 * sizeof(float), cudaMemcpyDeviceToHost, ctx->ba_stream));
	CK(cudaStreamSynchronize(ctx->ba_stream));
	W->lin_done = false;
	return 0;
}

extern "C" int dso_b200_ba_accumulate_A(dso_b200_ctx* ctx, double* H, double* b, int* resInA,
                                        float* Hdd_accAF, float* bd_accAF, float* Hcd_accAF4)
{
	int rc = ba_check_ready(ctx, __func__);
	if (rc) return rc;
	ARG_CHECK(H && b, "null output");
	CK(cudaSetDevice(ctx->device));
	BAWorkspace* W = ctx->ba;
	rc = ba_enqueue_top(ctx, 0);
	if (rc) return rc;
	const int N = 8 * W->P.nF + CPARS;
	const size_t nP = W->P.nP;
	cudaStream_t s = ctx->ba_stream;
	CK(cudaMemcpyAsync(W->h_out, W->d_HA, (size_t)N * N * sizeof(double), cudaMemcpyDeviceToHost, s));
	CK(cudaMemcpyAsync(W->h_out + N * N, W->d_bA, N * sizeof(double), cudaMemcpyDeviceToHost, s));
	CK(cudaMemcpyAsync(W->h_out + N * N + N, W->d_scalars, 2 * sizeof(double), cudaMemcpyDeviceToHost, s));
	if (Hdd_accAF) CK(cudaMemcpyAsync(Hdd_accAF, W->d_Hdd, nP * sizeof(float), cudaMemcpyDeviceToHost, s));
	if (bd_accAF) CK(cudaMemcpyAsync(bd_accAF, W->d_bd, nP * sizeof(float), cudaMemcpyDeviceToHost, s));
	if (Hcd_accAF4) CK(cudaMemcpyAsync(Hcd_accAF4, W->d_Hcd, nP * 4 * sizeof(float), cudaMemcpyDeviceToHost, s));
	CK(cudaStreamSynchronize(s));
	memcpy(H, W->h_out, (size_t)N * N * sizeof(double));
	memcpy(b, W->h_out + N * N, N * sizeof(double));
	if (resInA) *resInA = (int)W->h_out[N * N + N + 1];
	return 0;
}

// The per-point results of the last accumulation as they stand on the device (after dso_b200_ba_step / _iteration /
// _solve or the piecewise calls): EFPoint::Hdd_accAF, bd_accAF, Hcd_accAF, HdiF, bdSumF and PointHessian::idepth_hessian
// (AccumulatedTopHessian.cpp:162-166, AccumulatedSCHessian.cpp:50-60).  Every pointer may be NULL.
extern "C" int dso_b200_ba_get_point_terms(dso_b200_ctx* ctx, float* Hdd_accAF, float* bd_accAF, float* Hcd_accAF4,
                                           float* HdiF, float* bdSumF, float* idepth_hessian)
{
	int rc = ba_check_ready(ctx, __func__);
	if (rc) return rc;
	CK(cudaSetDevice(ctx->device));
	BAWorkspace* W = ctx->ba;
	const size_t nP = (size_t)W->P.nP;
	cudaStream_t s = ctx->ba_stream;
	if (Hdd_accAF) CK(cudaMemcpyAsync(Hdd_accAF, W->d_Hdd, nP * sizeof(float), cudaMemcpyDeviceToHost, s));
	if (bd_accAF) CK(cudaMemcpyAsync(bd_accAF, W->d_bd, nP * sizeof(float), cudaMemcpyDeviceToHost, s));
	if (Hcd_accAF4) CK(cudaMemcpyAsync(Hcd_accAF4, W->d_Hcd, nP * 4 * sizeof(float), cudaMemcpyDeviceToHost, s));
	if (HdiF) CK(cudaMemcpyAsync(HdiF, W->d_HdiF, nP * sizeof(float), cudaMemcpyDeviceToHost, s));
	if (bdSumF) CK(cudaMemcpyAsync(bdSumF, W->d_bdSumF, nP * sizeof(float), cudaMemcpyDeviceToHost, s));
	if (idepth_hessian) CK(cudaMemcpyAsync(idepth_hessian, W->d_idh, nP * sizeof(float), cudaMemcpyDeviceToHost, s));
	CK(cudaStreamSynchronize(s));
	return 0;
}

extern "C" int dso_b200_ba_accumulate_SC(dso_b200_ctx* ctx, double* H_sc, double* b_sc,
                                         float* HdiF, float* bdSumF, float* idepth_hessian)
{
	int rc = ba_check_ready(ctx, __func__);
	if (rc) return rc;
	ARG_CHECK(H_sc && b_sc, "null output");
	BAWorkspace* W = ctx->ba;
	BA_CHECK(W->top_done, "call dso_b200_ba_accumulate_A first (it produces Hdd/bd/Hcd, as in EnergyFunctional::solveSystemF)");
	CK(cudaSetDevice(ctx->device));
	rc = ba_enqueue_schur(ctx);
	if (rc) return rc;
	const int N = 8 * W->P.nF + CPARS;
	const size_t nP = W->P.nP;
	cudaStream_t s = ctx->ba_stream;
	CK(cudaMemcpyAsync(W->h_out, W->d_HS, (size_t)N * N * sizeof(double), cudaMemcpyDeviceToHost, s));
	CK(cudaMemcpyAsync(W->h_out + N * N, W->d_bS, N * sizeof(double), cudaMemcpyDeviceToHost, s));
	if (HdiF) CK(cudaMemcpyAsync(HdiF, W->d_HdiF, nP * sizeof(float), cudaMemcpyDeviceToHost, s));
	if (bdSumF) CK(cudaMemcpyAsync(bdSumF, W->d_bdSumF, nP * sizeof(float), cudaMemcpyDeviceToHost, s));
	if (idepth_hessian) CK(cudaMemcpyAsync(idepth_hessian, W->d_idh, nP * sizeof(float), cudaMemcpyDeviceToHost, s));
	CK(cudaStreamSynchronize(s));
	memcpy(H_sc, W->h_out, (size_t)N * N * sizeof(double));
	memcpy(b_sc, W->h_out + N * N, N * sizeof(double));
	return 0;
}

extern "C" int dso_b200_ba_step(dso_b200_ctx* ctx, double* H_A, double* b_A, double* H_sc, double* b_sc,
                                double* energy_sum, int* resInA)
{
	int rc = ba_check_ready(ctx, __func__);
	if (rc) return rc;
	ARG_CHECK(H_A && b_A && H_sc && b_sc, "null output");
	CK(cudaSetDevice(ctx->device));
	BAWorkspace* W = ctx->ba;
	const int N = 8 * W->P.nF + CPARS;
	const size_t NN = (size_t)N * N;
	cudaStream_t s = ctx->ba_stream;
	rc = ba_run_graphed(ctx, W->g_step, BAGraphKey{ 0.0, 0.f, 0, 0, 0, 0 }, [&]() -> int {
		int r;
		if ((r = ba_enqueue_linearize(ctx, true, true))) return r;        // linearize + applyRes in one kernel
		if ((r = ba_enqueue_top(ctx, 0, nullptr, true, true))) return r;      // + the Schur preparation per point
		if ((r = ba_enqueue_schur(ctx, nullptr, true))) return r;
		if ((r = ba_join_side(ctx))) return r;
		CK(cudaMemcpyAsync(W->h_out, W->d_outblk, BA_OUT_DOUBLES * sizeof(double), cudaMemcpyDeviceToHost, s));
		return 0;
	});
	if (rc) return rc;
	W->top_done = true;
	CK(cudaStreamSynchronize(s));
	const size_t NM = 8 * MAXF + CPARS, oB = NM * NM, oS = oB + NM, oBS = oS + NM * NM, oSc = oBS + NM;   // strides of the block
	memcpy(H_A, W->h_out, NN * sizeof(double));
	memcpy(b_A, W->h_out + oB, N * sizeof(double));
	memcpy(H_sc, W->h_out + oS, NN * sizeof(double));
	memcpy(b_sc, W->h_out + oBS, N * sizeof(double));
	if (energy_sum) *energy_sum = W->h_out[oSc];
	if (resInA) *resInA = (int)W->h_out[oSc + 1];
	W->lin_done = false;
	return 0;
}
