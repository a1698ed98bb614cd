// Next rows of the windowed-BA path (SURVEY.md section 8f, N1 and the accumulate half of N2): everything between the
// Schur accumulation and the next linearisation stays on the device.
//
//   ba_fix_lin_kernel        EFResidual::fixLinearizationF (EnergyFunctionalStructs.cpp:89-115)
//   ba_accum_top_kernel<1>   accumulateLF_MT = addPoint<1> + stitch with priors (ba.cu, shared with the active mode)
//   ba_solve_kernel          EnergyFunctional::solveSystemF tail (EnergyFunctional.cpp:822-918): bM_top, HFinal_top /
//                            bFinal_top, damping, the diagonally scaled LDL^T solve (Eigen::LDLT pivot rule: largest
//                            remaining ORIGINAL diagonal entry), orthogonalize(&x, 0), and resubstituteF_MT's xAd (:276-293)
//   ba_resubstitute_kernel   resubstituteFPt (:304-330): per-point inverse-depth step
//   ba_points_step_kernel    the per-point half of FullSystem::doStepFromBackup (FullSystemOptimize.cpp:272-286)
//
// One CTA solves the (8nF+4)^2 <= 68^2 FP64 system out of shared memory: a latency-bound chain of 68 pivots, not a
// bandwidth problem; the point kernels are a few kB of coalesced FP32 traffic.
#include "ba_internal.cuh"
#include <cmath>
#include <cstring>
#include <cstdio>
#include <cstdlib>

#define NMAX (8 * MAXF + CPARS)
#define LDH (NMAX + 1)

#define SOLVER_SVD 1
#define SOLVER_ORTHOGONALIZE_SYSTEM 2
#define SOLVER_ORTHOGONALIZE_POINTMARG 4
#define SOLVER_ORTHOGONALIZE_FULL 8
#define SOLVER_SVD_CUT7 16
#define SOLVER_USE_GN 64
#define SOLVER_FIX_LAMBDA 128
#define SOLVER_ORTHOGONALIZE_X 256
#define SOLVER_ORTHOGONALIZE_X_LATER 2048

struct BASolve
{
	double* d_HM; double* d_bM;                      // marginalisation prior
	double* d_P; bool haveP;                         // 0.5 (N Npi^T + Npi N^T) of EnergyFunctional::orthogonalize
	double* d_x; float* d_xAd; double* d_HF; double* d_bF;
	int capP, capR;
	float* d_step; float* d_idepth_backup; double* d_stepBlocks; unsigned char* d_which;
	double* h_x;                                     // pinned: x (NMAX), step sums (4)
	unsigned char* d_pmask; int capPm; double* d_margH; double* d_eblocks;   // marginalisation / energy scratch
	long long* d_dbg;                                // DSO_B200_SOLVE_TIMING=1: clock64 stamps of ba_solve_kernel's phases
	bool priors_set;
	// non-default solver branches: "a frame with frameID 0 is in the window" (EnergyFunctional.cpp:688, :835), the solver
	// mode / delta marginalizePointsF sees, scratch for HT_act, bT_act and the products of orthogonalize()
	int haveFirstFrame, margSolverMode;
	double solverModeDelta;                          // setting_solverModeDelta (given with the null spaces or the solver state)
	double* d_work;
};

static int solve_state(dso_b200_ctx* ctx, BASolve** out)
{
	BAWorkspace* W = ctx->ba;
	if (!W->S)
	{
		BASolve* S = new BASolve();
		memset(S, 0, sizeof(*S));
		S->haveFirstFrame = 1; S->margSolverMode = SOLVER_FIX_LAMBDA | SOLVER_ORTHOGONALIZE_X_LATER; S->solverModeDelta = 1e-5;   // util/settings.cpp:60-61
		CK(cudaMalloc(&S->d_HM, (size_t)NMAX * NMAX * sizeof(double)));
		CK(cudaMalloc(&S->d_bM, NMAX * sizeof(double)));
		CK(cudaMalloc(&S->d_P, (size_t)NMAX * NMAX * sizeof(double)));
		CK(cudaMalloc(&S->d_x, NMAX * sizeof(double)));
		CK(cudaMalloc(&S->d_xAd, MAXF * MAXF * 8 * sizeof(float)));
		CK(cudaMalloc(&S->d_HF, (size_t)NMAX * NMAX * sizeof(double)));
		CK(cudaMalloc(&S->d_bF, NMAX * sizeof(double)));
		CK(cudaMemset(S->d_HM, 0, (size_t)NMAX * NMAX * sizeof(double)));
		CK(cudaMemset(S->d_bM, 0, NMAX * sizeof(double)));
		CK(cudaHostAlloc((void**)&S->h_x, (NMAX + 8) * sizeof(double), cudaHostAllocDefault));
		if (getenv("DSO_B200_SOLVE_TIMING")) { CK(cudaMalloc(&S->d_dbg, 16 * sizeof(long long))); CK(cudaMemset(S->d_dbg, 0, 16 * sizeof(long long))); }
		W->S = S;
	}
	*out = W->S;
	return 0;
}

static int nondefault_buffers(BASolve* S)
{
	if (!S->d_work) CK(cudaMalloc(&S->d_work, ((size_t)3 * NMAX * NMAX + 4 * NMAX) * sizeof(double)));   // Hact | bact | T (+ y) | spare
	return 0;
}

static int ensure_point_buffers(dso_b200_ctx* ctx, BASolve* S)
{
	BAWorkspace* W = ctx->ba;
	if (W->P.nP > S->capP)
	{
		if (S->capP) { cudaFree(S->d_step); cudaFree(S->d_idepth_backup); cudaFree(S->d_stepBlocks); }
		const size_t n = (size_t)W->P.nP;
		CK(cudaMalloc(&S->d_step, n * sizeof(float)));
		CK(cudaMalloc(&S->d_idepth_backup, n * sizeof(float)));
		CK(cudaMalloc(&S->d_stepBlocks, (((n + 255) / 256) * 2 + 1) * sizeof(double)));     // + the fused kernel's block counter
		CK(cudaMemset(S->d_stepBlocks, 0, (((n + 255) / 256) * 2 + 1) * sizeof(double)));
		CK(cudaMemset(S->d_step, 0, n * sizeof(float)));
		S->capP = W->P.nP;
	}
	return 0;
}

void ba_solve_destroy(BAWorkspace* W)
{
	BASolve* S = W->S;
	if (!S) return;
	cudaFree(S->d_HM); cudaFree(S->d_bM); cudaFree(S->d_P); cudaFree(S->d_x);
	cudaFree(S->d_xAd); cudaFree(S->d_HF); cudaFree(S->d_bF);
	if (S->capP) { cudaFree(S->d_step); cudaFree(S->d_idepth_backup); cudaFree(S->d_stepBlocks); }
	if (S->capR) cudaFree(S->d_which);
	cudaFreeHost(S->h_x);
	if (S->d_dbg) cudaFree(S->d_dbg);
	if (S->d_work) cudaFree(S->d_work);
	if (S->capPm) { cudaFree(S->d_pmask); cudaFree(S->d_eblocks); }
	if (S->d_margH) cudaFree(S->d_margH);
	delete S;
	W->S = nullptr;
}

// ---------------------------------------------------------------------------------------------------------
// EFResidual::fixLinearizationF, EnergyFunctionalStructs.cpp:89-115: res_toZeroF = resF - J * delta for the flagged
// active residuals, which become linearised
__device__ __forceinline__ void jp_delta(const float* __restrict__ J, const float* __restrict__ dp, const float* __restrict__ dc, float dd, float* d)
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

__global__ void ba_fix_lin_kernel(int nR, int nF, const unsigned char* __restrict__ which, const unsigned char* __restrict__ isAct,
                                  unsigned char* isLin, const unsigned char* __restrict__ efsel, const float* __restrict__ J0,
                                  const float* __restrict__ J1, const int* __restrict__ rpoint, const int* __restrict__ rtarget,
                                  const int* __restrict__ phost, const float* __restrict__ deltaF, const float* __restrict__ adHTdeltaF,
                                  const float* __restrict__ cDeltaF, float* resToZero)
{
	const int r = blockIdx.x * blockDim.x + threadIdx.x;
	if (r >= nR) return;
	if (!isAct[r] || (which && !which[r])) return;
	const float* J = (efsel[r] ? J1 : J0) + (size_t)r * JSZ;
	const int p = rpoint[r];
	float d[4];
	jp_delta(J, adHTdeltaF + 8 * (phost[p] + nF * rtarget[r]), cDeltaF, deltaF[p], d);
#pragma unroll
	for (int i = 0; i < 8; i++)
	{
		float rtz = J[J_RESF + i];
		rtz = fsub(rtz, fmul(J[J_IDX0 + i], d[0]));
		rtz = fsub(rtz, fmul(J[J_IDX1 + i], d[1]));
		rtz = fsub(rtz, fmul(J[J_AB0 + i], d[2]));
		rtz = fsub(rtz, fmul(J[J_AB1 + i], d[3]));
		resToZero[8 * (size_t)r + i] = rtz;
	}
	isLin[r] = 1;
}

// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ double shfl_dd(double v, int src) { return __shfl_sync(0xffffffffu, v, src); }

struct SolveArgs
{
	int nF, orth, haveL;
	double lambda;
	const double* HA; const double* bA; const double* HL; const double* bL; const double* HS; const double* bS;
	const double* HM; const double* bM; const double* delta; const double* priorH; const double* priorD; const float* cDeltaF;
	const double* P; const double* adHost; const double* adTarget;
	double* x; float* xAd; double* HF; double* bF;
	const double* Hpre; const double* bpre;          // SOLVER_ORTHOGONALIZE_SYSTEM: HT_act / bT_act, already projected (:832-851)
	int assemble_only;                               // SOLVER_SVD: stop after HFinal / bFinal (written to HF / bF)
	long long* dbg;
};

// solveSystemF tail (EnergyFunctional.cpp:822-918, default solver branch) + xAd of resubstituteF_MT (:283-293).
// All products and sums are separate IEEE operations in the reference's order (no contraction), so the only
// differences to the CPU path are the summation order of the back substitution and of Eigen's GEMV.
#define SOLVE_NT 1024
__global__ void __launch_bounds__(SOLVE_NT) ba_solve_kernel(SolveArgs a)
{
	__shared__ double A[NMAX * LDH];
	__shared__ double sv[NMAX], rhs[NMAX], yv[NMAX], xv[NMAX], dg[NMAX], dl[NMAX];
	__shared__ int ord[NMAX];
	const int tid = threadIdx.x, N = 8 * a.nF + CPARS;
#define STAMP(i) do { if (a.dbg && tid == 0) a.dbg[i] = clock64(); } while (0)
	STAMP(0);
	// getStitchedDeltaF (:952-958)
	if (tid < N) dl[tid] = tid < CPARS ? (double)a.cDeltaF[tid] : a.delta[tid];
	__syncthreads();
	// HM through shared memory first (coalesced, loads batched): bM_top = bM + HM * delta (:824) sums each row in order
	constexpr int NQ = (NMAX * NMAX + SOLVE_NT - 1) / SOLVE_NT;
	const float invN = 1.0f / (float)N;
#pragma unroll
	for (int q = 0; q < NQ; q++)
	{
		const int e = tid + SOLVE_NT * q;
		if (e < N * N)
		{
			int i = (int)((float)e * invN);
			i += ((i + 1) * N <= e) ? 1 : 0; i -= (i * N > e) ? 1 : 0;
			A[i * LDH + (e - i * N)] = a.HM[e];
		}
	}
	__syncthreads();
	if (tid < N)
	{
		// bFinal_top = bL_top + bM_top + bA_top - b_sc (:859)
		double s = 0.0;
		for (int j = 0; j < N; j++) s = __dadd_rn(s, __dmul_rn(A[tid * LDH + j], dl[j]));
		const double bMtop = __dadd_rn(a.bM[tid], s);
		const double bl = a.haveL ? a.bL[tid] : __dmul_rn(a.priorH[tid], tid < CPARS ? (double)a.cDeltaF[tid] : a.priorD[tid]);
		rhs[tid] = a.bpre ? __dadd_rn(a.bpre[tid], bMtop)                                  // :844
		                  : __dsub_rn(__dadd_rn(__dadd_rn(bl, bMtop), a.bA[tid]), a.bS[tid]);
	}
	STAMP(1);
	// HFinal_top = HL_top + HM + HA_top; diagonal *= (1+lambda); -= H_sc * (1.0f/(1+lambda))  (:857-866); every thread
	// rewrites the entries it staged itself
	const double damp = __dadd_rn(1.0, a.lambda), isc = __ddiv_rn(1.0, __dadd_rn(1.0, a.lambda));
	__syncthreads();
#pragma unroll
	for (int q = 0; q < NQ; q++)
	{
		const int e = tid + SOLVE_NT * q;
		if (e < N * N)
		{
			int i = (int)((float)e * invN);
			i += ((i + 1) * N <= e) ? 1 : 0; i -= (i * N > e) ? 1 : 0;
			const int j = e - i * N;
			double v;
			if (a.Hpre)
			{
				v = __dadd_rn(a.Hpre[e], A[i * LDH + j]);                                   // HT_act + HM (:843), damping (:849)
				if (i == j) v = __dmul_rn(v, damp);
			}
			else
			{
				const double hl = a.haveL ? a.HL[e] : (i == j ? a.priorH[i] : 0.0);
				v = __dadd_rn(__dadd_rn(hl, A[i * LDH + j]), a.HA[e]);
				if (i == j) v = __dmul_rn(v, damp);
				v = __dsub_rn(v, __dmul_rn(a.HS[e], isc));
			}
			A[i * LDH + j] = v;
			if (a.HF) a.HF[e] = v;
		}
	}
	__syncthreads();
	STAMP(2);
	if (tid < N && a.bF) a.bF[tid] = rhs[tid];
	if (a.assemble_only) return;                     // uniform: the SVD branch continues in ba_svd_solve_kernel
	if (tid < N) sv[tid] = __ddiv_rn(1.0, sqrt(__dadd_rn(A[tid * LDH + tid], 10.0)));    // :905
	__syncthreads();
	for (int e = tid; e < N * N; e += SOLVE_NT)
	{
		int i = (int)((float)e * invN);
		i += ((i + 1) * N <= e) ? 1 : 0; i -= (i * N > e) ? 1 : 0;
		const int j = e - i * N;
		A[i * LDH + j] = __dmul_rn(__dmul_rn(sv[i], A[i * LDH + j]), sv[j]);     // SVecI.asDiagonal() * H * SVecI.asDiagonal()
	}
	if (tid < N) { rhs[tid] = __dmul_rn(sv[tid], rhs[tid]); ord[tid] = tid; }
	__syncthreads();
	STAMP(3);
	if (tid < N) dg[tid] = fabs(A[tid * LDH + tid]);
	__syncthreads();
	// Eigen::LDLT pivot order: at step k the largest |diagonal| among positions k..N-1 (first one on ties) is swapped to
	// position k; the diagonal entries it looks at are the not-yet-updated ones, so the whole order follows from the
	// initial diagonal.  Without exact ties that is a descending sort (rank = number of larger entries); with ties the
	// swaps decide, and warp 0 replays the N selection steps.
	int tie = 0;
	if (tid < N)
	{
		const double v = dg[tid];
		int rank = 0;
		for (int j = 0; j < N; j++) { const double o = dg[j]; rank += (o > v) ? 1 : 0; tie |= (o == v && j != tid) ? 1 : 0; }
		if (!tie) ord[rank] = tid;
	}
	if (__syncthreads_or(tie))
	{
		if (tid < N) ord[tid] = tid;
		__syncthreads();
		if (tid < 32)
		{
			for (int k = 0; k < N; k++)
			{
				double best = -1.0; int bi = N;
				for (int i = k + tid; i < N; i += 32) { const double v = dg[i]; if (v > best) { best = v; bi = i; } }
#pragma unroll
				for (int o = 16; o > 0; o >>= 1)
				{
					const double ob = __shfl_xor_sync(0xffffffffu, best, o);
					const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
					if (ob > best || (ob == best && oi < bi)) { best = ob; bi = oi; }
				}
				if (tid == 0 && bi != k && bi < N)
				{
					const double t = dg[k]; dg[k] = dg[bi]; dg[bi] = t;
					const int u = ord[k]; ord[k] = ord[bi]; ord[bi] = u;
				}
				__syncwarp();
			}
		}
		__syncthreads();
	}
	// LDL^T of the permuted matrix, lower triangle, right-looking and blocked by panels of PANEL columns:
	// A(i,j) -= (L(i,k) * L(j,k)) * D(k) for k ascending, L(i,j) = A(i,j) / D(j) -- the same operations in the same order
	// per entry as the bordered form Eigen uses.  The matrix is first permuted physically (through registers), then per
	// panel warp 0 factors the PANEL columns in registers (three rows per lane, pivots and pivot rows by shuffle) and all
	// threads apply the rank-PANEL update to their entries of the trailing matrix: 2 block barriers per PANEL pivots.
	STAMP(4);
	constexpr int PANEL = 4;                             // 8 nF + 4 is a multiple of 4
	constexpr int MAXQ = (NMAX * (NMAX + 1) / 2 + SOLVE_NT - 1) / SOLVE_NT;
	const int T = N * (N + 1) / 2;
	int ei[MAXQ], ej[MAXQ];
	{
		double av[MAXQ];
#pragma unroll
		for (int q = 0; q < MAXQ; q++)
		{
			const int e = tid + SOLVE_NT * q;
			int i = (int)((sqrtf(8.f * (float)e + 1.f) - 1.f) * 0.5f);
			while (i * (i + 1) / 2 > e) i--;
			while ((i + 1) * (i + 2) / 2 <= e) i++;
			const int j = e - i * (i + 1) / 2;
			const bool ok = e < T;
			ei[q] = ok ? i : -1; ej[q] = ok ? j : -1;
			av[q] = ok ? A[ord[i] * LDH + ord[j]] : 0.0;
		}
		__syncthreads();
#pragma unroll
		for (int q = 0; q < MAXQ; q++) if (ei[q] >= 0) A[ei[q] * LDH + ej[q]] = av[q];
		if (tid < N) yv[tid] = rhs[ord[tid]];
	}
	__syncthreads();
	STAMP(5);
	const int lane = tid & 31;
	for (int k0 = 0; k0 < N; k0 += PANEL)
	{
		if (tid < 32)
		{
			double p[3][PANEL];
#pragma unroll
			for (int c = 0; c < 3; c++)
			{
				const int r = k0 + lane + 32 * c;
#pragma unroll
				for (int b = 0; b < PANEL; b++) p[c][b] = (r < N && k0 + b <= r) ? A[r * LDH + k0 + b] : 0.0;
			}
#pragma unroll
			for (int b = 0; b < PANEL; b++)
			{
				const double d = shfl_dd(p[0][b], b);                      // D(k0+b): row k0+b is lane b, c = 0
				// column / D: one IEEE reciprocal per pivot, then q0 = p*rd, e = p - q0*d (exact, FMA), q = q0 + e*rd,
				// which is the correctly rounded quotient (Markstein) -- three FMAs per row instead of a division each
				const double rd = (fabs(d) > 0.0) ? __ddiv_rn(1.0, d) : 0.0;
#pragma unroll
				for (int c = 0; c < 3; c++)
				{
					const int r = k0 + lane + 32 * c;
					if (r > k0 + b && r < N)
					{
						const double q0 = __dmul_rn(p[c][b], rd);
						const double e = __fma_rn(-q0, d, p[c][b]);
						p[c][b] = (fabs(d) > 0.0) ? __fma_rn(e, rd, q0) : 0.0;
					}
				}
#pragma unroll
				for (int b2 = b + 1; b2 < PANEL; b2++)
				{
					const double lrow = shfl_dd(p[0][b], b2);              // L(k0+b2, k0+b)
#pragma unroll
					for (int c = 0; c < 3; c++)
					{
						const int r = k0 + lane + 32 * c;
						if (r >= k0 + b2 && r < N) p[c][b2] = __dsub_rn(p[c][b2], __dmul_rn(__dmul_rn(p[c][b], lrow), d));
					}
				}
			}
#pragma unroll
			for (int c = 0; c < 3; c++)
			{
				const int r = k0 + lane + 32 * c;
#pragma unroll
				for (int b = 0; b < PANEL; b++) if (r < N && k0 + b <= r) A[r * LDH + k0 + b] = p[c][b];
			}
		}
		__syncthreads();
		double dk[PANEL];
#pragma unroll
		for (int b = 0; b < PANEL; b++) dk[b] = A[(k0 + b) * LDH + k0 + b];
#pragma unroll
		for (int q = 0; q < MAXQ; q++)
		{
			if (ej[q] < k0 + PANEL) continue;                            // also skips the unused slots (ej = -1)
			const int i = ei[q], j = ej[q];
			double li[PANEL], lj[PANEL];
#pragma unroll
			for (int b = 0; b < PANEL; b++) { li[b] = A[i * LDH + k0 + b]; lj[b] = A[j * LDH + k0 + b]; }
			double v = A[i * LDH + j];
#pragma unroll
			for (int b = 0; b < PANEL; b++) v = __dsub_rn(v, __dmul_rn(__dmul_rn(li[b], lj[b]), dk[b]));
			A[i * LDH + j] = v;
		}
		__syncthreads();
	}
	STAMP(6);
	// solve: permute (done above), L y = b, D, L^T, permute back (:908) -- warp 0 alone, three rows per lane in registers,
	// pivot entries broadcast by shuffle: no block barrier per column
	if (tid < 32)
	{
		double y[3];
#pragma unroll
		for (int c = 0; c < 3; c++) y[c] = (lane + 32 * c < N) ? yv[lane + 32 * c] : 0.0;
		for (int k = 0; k < N; k++)
		{
			const double cand = k < 32 ? y[0] : (k < 64 ? y[1] : y[2]);
			const double yk = shfl_dd(cand, k & 31);
#pragma unroll
			for (int c = 0; c < 3; c++)
			{
				const int r = lane + 32 * c;
				if (r > k && r < N) y[c] = __dsub_rn(y[c], __dmul_rn(A[r * LDH + k], yk));
			}
		}
#pragma unroll
		for (int c = 0; c < 3; c++)
		{
			const int r = lane + 32 * c;
			if (r < N)
			{
				const double d = A[r * LDH + r];
				y[c] = (fabs(d) > 2.2250738585072014e-308) ? __ddiv_rn(y[c], d) : 0.0;
			}
		}
		for (int k = N - 1; k > 0; k--)
		{
			const double cand = k < 32 ? y[0] : (k < 64 ? y[1] : y[2]);
			const double yk = shfl_dd(cand, k & 31);
#pragma unroll
			for (int c = 0; c < 3; c++)
			{
				const int r = lane + 32 * c;
				if (r < k) y[c] = __dsub_rn(y[c], __dmul_rn(A[k * LDH + r], yk));
			}
		}
#pragma unroll
		for (int c = 0; c < 3; c++) if (lane + 32 * c < N) yv[lane + 32 * c] = y[c];
	}
	__syncthreads();
	STAMP(7);
	if (tid < N) xv[ord[tid]] = __dmul_rn(sv[ord[tid]], yv[tid]);
	__syncthreads();
	if (a.orth)
	{
		// orthogonalize(&x, 0) (:741-792): x -= NNpiTS * x with the projector prepared by dso_b200_ba_set_nullspaces
		for (int e = tid; e < N * N; e += SOLVE_NT)               // the factors in A are no longer needed
		{
			int i = (int)((float)e * invN);
			i += ((i + 1) * N <= e) ? 1 : 0; i -= (i * N > e) ? 1 : 0;
			A[i * LDH + (e - i * N)] = a.P[e];
		}
		__syncthreads();
		double s = 0.0;
		if (tid < N) for (int j = 0; j < N; j++) s = __dadd_rn(s, __dmul_rn(A[tid * LDH + j], xv[j]));
		__syncthreads();
		if (tid < N) xv[tid] = __dsub_rn(xv[tid], s);
		__syncthreads();
	}
	STAMP(8);
	if (tid < N) a.x[tid] = xv[tid];
	// xAd[nF*h + t] = xF_h^T adHostF[h + nF*t] + xF_t^T adTargetF[h + nF*t]  (:289-292), 8-term unrolled reductions
	for (int e = tid; e < a.nF * a.nF * 8; e += SOLVE_NT)
	{
		const int j = e & 7, ht = e >> 3, h = ht / a.nF, t = ht - h * a.nF;
		const double* AH = a.adHost + (size_t)(h + a.nF * t) * 64;
		const double* AT_ = a.adTarget + (size_t)(h + a.nF * t) * 64;
		float ph[8], pt[8];
#pragma unroll
		for (int i = 0; i < 8; i++)
		{
			ph[i] = fmul((float)xv[CPARS + 8 * h + i], (float)AH[i * 8 + j]);
			pt[i] = fmul((float)xv[CPARS + 8 * t + i], (float)AT_[i * 8 + j]);
		}
		const float sh = fadd(fadd(fadd(ph[0], ph[1]), fadd(ph[2], ph[3])), fadd(fadd(ph[4], ph[5]), fadd(ph[6], ph[7])));
		const float st = fadd(fadd(fadd(pt[0], pt[1]), fadd(pt[2], pt[3])), fadd(fadd(pt[4], pt[5]), fadd(pt[6], pt[7])));
		a.xAd[e] = fadd(sh, st);
	}
	STAMP(9);
#undef STAMP
}

// resubstituteFPt, EnergyFunctional.cpp:304-330
__device__ __forceinline__ float resubstitute_point(int p, int nF, const int* __restrict__ res_begin, const int* __restrict__ rtarget,
                                                    const int* __restrict__ phost, const unsigned char* __restrict__ isAct,
                                                    const float* __restrict__ JpJdF, const float* __restrict__ HcdSum, const float* __restrict__ bdSumF,
                                                    const float* __restrict__ HdiF, const double* __restrict__ x, const float* __restrict__ xAd)
{
	// (the reference's "ngood == 0 -> step 0" is decided at the end: the loop below sees the same flags)
	int ngood = 0;
	float b = bdSumF[p];
	const float* hc = HcdSum + 4 * (size_t)p;
	const float x0 = (float)x[0], x1 = (float)x[1], x2 = (float)x[2], x3 = (float)x[3];
	b = fsub(b, fadd(fadd(fmul(x0, hc[0]), fmul(x1, hc[1])), fadd(fmul(x2, hc[2]), fmul(x3, hc[3]))));
	const int h = phost[p];
	const int r1 = res_begin[p + 1];
	for (int r0 = res_begin[p]; r0 < r1; r0 += 8)              // eight residuals per round (a point has at most nF-1), subtractions in residualsAll order
	{
		// round 1: flags, targets and the Jacobian terms of all eight (clamped, unconditional: nothing waits for a flag);
		// round 2: the xAd rows the targets point at
		bool use[8]; int tg[8]; float4 j0[8], j1[8];
#pragma unroll
		for (int u = 0; u < 8; u++)
		{
			const int r = r0 + u, rc = min(r, r1 - 1);
			use[u] = r < r1 && isAct[rc];
			tg[u] = rtarget[rc];
			const float4* jp = reinterpret_cast<const float4*>(JpJdF + 8 * (size_t)rc);
			j0[u] = jp[0]; j1[u] = jp[1];
			ngood += use[u] ? 1 : 0;
		}
		float d[8];
#pragma unroll
		for (int u = 0; u < 8; u++)
		{
			const float4* xa = reinterpret_cast<const float4*>(xAd + 8 * (h * nF + tg[u]));
			const float4 x0 = xa[0], x1 = xa[1];
			d[u] = fadd(fadd(fadd(fmul(x0.x, j0[u].x), fmul(x0.y, j0[u].y)), fadd(fmul(x0.z, j0[u].z), fmul(x0.w, j0[u].w))),
			            fadd(fadd(fmul(x1.x, j1[u].x), fmul(x1.y, j1[u].y)), fadd(fmul(x1.z, j1[u].z), fmul(x1.w, j1[u].w))));
		}
#pragma unroll
		for (int u = 0; u < 8; u++) if (use[u]) b = fsub(b, d[u]);
	}
	return ngood == 0 ? 0.f : fmul(-b, HdiF[p]);
}
__global__ void ba_resubstitute_kernel(int nP, int nF, const int* __restrict__ res_begin, const int* __restrict__ rtarget,
                                       const int* __restrict__ phost, const unsigned char* __restrict__ isAct,
                                       const float* __restrict__ JpJdF, const float* __restrict__ HcdSum, const float* __restrict__ bdSumF,
                                       const float* __restrict__ HdiF, const double* __restrict__ x, const float* __restrict__ xAd, float* step)
{
	const int p = blockIdx.x * blockDim.x + threadIdx.x;
	if (p < nP) step[p] = resubstitute_point(p, nF, res_begin, rtarget, phost, isAct, JpJdF, HcdSum, bdSumF, HdiF, x, xAd);
}

// per-point half of FullSystem::doStepFromBackup (FullSystemOptimize.cpp:272-286): idepth = idepth_backup + stepfacD * step,
// setIdepthZero of the same value (HessianBlocks.h:470-490; SCALE_IDEPTH = 1), EFPoint::deltaF = idepth - idepth_zero = 0;
// block sums of step^2 and |idepth_backup| in FP64 (fixed tree)
__global__ void __launch_bounds__(256) ba_points_step_kernel(int nP, float stepfacD, const float* __restrict__ step,
                                                             const float* __restrict__ backup, float4* pt, float* deltaF, double* blocks)
{
	__shared__ double s0[256], s1[256];
	const int p = blockIdx.x * 256 + threadIdx.x;
	double a = 0.0, b = 0.0;
	if (p < nP)
	{
		const float st = step[p], bk = backup[p];
		const float v = fadd(bk, fmul(stepfacD, st));
		float4 q = pt[p];
		q.z = v; q.w = v;
		pt[p] = q;
		deltaF[p] = 0.f;
		a = (double)fmul(st, st); b = (double)fabsf(bk);
	}
	s0[threadIdx.x] = a; s1[threadIdx.x] = b;
	__syncthreads();
	for (int o = 128; o > 0; o >>= 1)
	{
		if (threadIdx.x < o) { s0[threadIdx.x] += s0[threadIdx.x + o]; s1[threadIdx.x] += s1[threadIdx.x + o]; }
		__syncthreads();
	}
	if (threadIdx.x == 0) { blocks[2 * blockIdx.x] = s0[0]; blocks[2 * blockIdx.x + 1] = s1[0]; }
}
__global__ void ba_points_step_final_kernel(int nb, const double* __restrict__ blocks, double* out)
{
	if (threadIdx.x == 0 && blockIdx.x == 0)
	{
		double a = 0, b = 0;
		for (int i0 = 0; i0 < nb; i0 += 8)
		{
			double ta[8], tb[8];
#pragma unroll
			for (int u = 0; u < 8; u++) { ta[u] = (i0 + u < nb) ? blocks[2 * (i0 + u)] : 0.0; tb[u] = (i0 + u < nb) ? blocks[2 * (i0 + u) + 1] : 0.0; }
#pragma unroll
			for (int u = 0; u < 8; u++) if (i0 + u < nb) { a += ta[u]; b += tb[u]; }
		}
		out[0] = a; out[1] = b;
	}
}
__global__ void ba_points_backup_kernel(int nP, const float4* __restrict__ pt, float* backup, int restore, float4* pt_rw, float* deltaF)
{
	const int p = blockIdx.x * blockDim.x + threadIdx.x;
	if (p >= nP) return;
	if (!restore) backup[p] = pt[p].z;                      // FullSystem::backupState: ph->idepth_backup = ph->idepth
	else { float4 q = pt_rw[p]; q.z = backup[p]; q.w = backup[p]; pt_rw[p] = q; deltaF[p] = 0.f; }   // loadSateBackup
}

// resubstituteFPt + backupState + the point half of doStepFromBackup + the two step sums in ONE launch (the Gauss-Newton
// iteration's tail; each piece computes exactly what its stand-alone kernel above computes).  The last block to finish adds
// the block sums in index order, as ba_points_step_final_kernel does, and re-arms the counter for the next launch.
__global__ void __launch_bounds__(256) ba_resub_step_kernel(int nP, int nF, const int* __restrict__ res_begin, const int* __restrict__ rtarget,
                                                            const int* __restrict__ phost, const unsigned char* __restrict__ isAct,
                                                            const float* __restrict__ JpJdF, const float* __restrict__ HcdSum,
                                                            const float* __restrict__ bdSumF, const float* __restrict__ HdiF,
                                                            const double* __restrict__ x, const float* __restrict__ xAd, float* step,
                                                            float stepfacD, float* backup, float4* pt, float* deltaF, double* blocks,
                                                            unsigned int* counter, double* out)
{
	__shared__ double s0[256], s1[256];
	__shared__ bool last;
	const int p = blockIdx.x * 256 + threadIdx.x;
	double a = 0.0, b = 0.0;
	if (p < nP)
	{
		const float st = resubstitute_point(p, nF, res_begin, rtarget, phost, isAct, JpJdF, HcdSum, bdSumF, HdiF, x, xAd);
		step[p] = st;
		float4 q = pt[p];
		const float bk = q.z;                                   // backupState: idepth_backup = idepth
		backup[p] = bk;
		const float v = fadd(bk, fmul(stepfacD, st));
		q.z = v; q.w = v;
		pt[p] = q;
		deltaF[p] = 0.f;
		a = (double)fmul(st, st); b = (double)fabsf(bk);
	}
	s0[threadIdx.x] = a; s1[threadIdx.x] = b;
	__syncthreads();
	for (int o = 128; o > 0; o >>= 1)
	{
		if (threadIdx.x < o) { s0[threadIdx.x] += s0[threadIdx.x + o]; s1[threadIdx.x] += s1[threadIdx.x + o]; }
		__syncthreads();
	}
	if (threadIdx.x == 0)
	{
		blocks[2 * blockIdx.x] = s0[0]; blocks[2 * blockIdx.x + 1] = s1[0];
		__threadfence();
		last = atomicAdd(counter, 1u) == gridDim.x - 1;
	}
	__syncthreads();
	if (last && threadIdx.x == 0)
	{
		__threadfence();
		const int nb = gridDim.x;
		double sa = 0, sb = 0;
		for (int i0 = 0; i0 < nb; i0 += 8)
		{
			double ta[8], tb[8];
#pragma unroll
			for (int u = 0; u < 8; u++) { ta[u] = (i0 + u < nb) ? __ldcg(blocks + 2 * (i0 + u)) : 0.0; tb[u] = (i0 + u < nb) ? __ldcg(blocks + 2 * (i0 + u) + 1) : 0.0; }
#pragma unroll
			for (int u = 0; u < 8; u++) if (i0 + u < nb) { sa += ta[u]; sb += tb[u]; }
		}
		out[0] = sa; out[1] = sb;
		*counter = 0u;
	}
}

// ---------------------------------------------------------------------------------------------------------
// Non-default solver branches (setting_solverMode bits 1, 2, 4, 8, 16; the default is 128 | 2048)
// ---------------------------------------------------------------------------------------------------------
// SOLVER_ORTHOGONALIZE_SYSTEM, EnergyFunctional.cpp:837-838: HT_act = HL_top + HA_top - H_sc, bT_act = bL_top + bA_top - b_sc
__global__ void ba_assemble_act_kernel(int N, int haveL, const double* __restrict__ HL, const double* __restrict__ bL,
                                       const double* __restrict__ HA, const double* __restrict__ bA,
                                       const double* __restrict__ HS, const double* __restrict__ bS,
                                       const double* __restrict__ priorH, const double* __restrict__ priorD, const float* __restrict__ cDeltaF,
                                       double* Hact, double* bact)
{
	const int e = blockIdx.x * blockDim.x + threadIdx.x;
	if (e < N * N)
	{
		const int i = e / N, j = e - i * N;
		const double hl = haveL ? HL[e] : (i == j ? priorH[i] : 0.0);        // without linearised residuals H_L is the prior alone
		Hact[e] = __dsub_rn(__dadd_rn(hl, HA[e]), HS[e]);
	}
	if (e < N)
	{
		const double bl = haveL ? bL[e] : __dmul_rn(priorH[e], e < CPARS ? (double)cDeltaF[e] : priorD[e]);
		bact[e] = __dsub_rn(__dadd_rn(bl, bA[e]), bS[e]);
	}
}

// EnergyFunctional::orthogonalize(VecX* b, MatXX* H), :741-792, with NNpiTS = P prepared by dso_b200_ba_set_nullspaces:
// b -= P b; H -= (P H) P.  Stage 1: T = P H (row i per block) and y = P b (block N); stage 2: H -= T P, b -= y.
__global__ void __launch_bounds__(128) ba_orth_stage1_kernel(int N, const double* __restrict__ P, const double* __restrict__ H,
                                                             const double* __restrict__ b, double* T)
{
	const int i = blockIdx.x, j = threadIdx.x;
	if (j >= N) return;
	double s = 0.0;
	if (i < N)
	{
		for (int q = 0; q < N; q++) s = __dadd_rn(s, __dmul_rn(P[(size_t)i * N + q], H[(size_t)q * N + j]));
		T[(size_t)i * N + j] = s;
	}
	else
	{
		for (int q = 0; q < N; q++) s = __dadd_rn(s, __dmul_rn(P[(size_t)j * N + q], b[q]));
		T[(size_t)N * N + j] = s;
	}
}
__global__ void __launch_bounds__(128) ba_orth_stage2_kernel(int N, const double* __restrict__ P, double* H, double* b, const double* __restrict__ T)
{
	const int i = blockIdx.x, j = threadIdx.x;
	if (j >= N) return;
	if (i < N)
	{
		double s = 0.0;
		for (int q = 0; q < N; q++) s = __dadd_rn(s, __dmul_rn(T[(size_t)i * N + q], P[(size_t)q * N + j]));
		H[(size_t)i * N + j] = __dsub_rn(H[(size_t)i * N + j], s);
	}
	else if (b)
		b[j] = __dsub_rn(b[j], T[(size_t)N * N + j]);
}

// marginalizePointsF with the orthogonalisation options: stage 1 = H, b alone (:681-682), stage 2 = the accumulation (:694-695)
__global__ void ba_marg_accumulate_kernel(int N, double w, const double* __restrict__ H, const double* __restrict__ b, double* HM, double* bM)
{
	const int e = blockIdx.x * blockDim.x + threadIdx.x;
	if (e < N * N) HM[e] = __dadd_rn(HM[e], __dmul_rn(w, H[e]));
	if (e < N) bM[e] = __dadd_rn(bM[e], __dmul_rn(w, b[e]));
}
__global__ void ba_marg_diff_kernel(int N, const double* __restrict__ M, const double* __restrict__ Mb,
                                    const double* __restrict__ Msc, const double* __restrict__ Mbsc, double* H, double* b)
{
	const int e = blockIdx.x * blockDim.x + threadIdx.x;
	if (e < N * N) H[e] = __dsub_rn(M[e], Msc[e]);
	if (e < N) b[e] = __dsub_rn(Mb[e], Mbsc[e]);
}

// SOLVER_SVD (+ SOLVER_SVD_CUT7), EnergyFunctional.cpp:870-900, on HFinal / bFinal left in global memory by ba_solve_kernel's
// assemble-only mode: SVecI = 1/sqrt(diag), thin SVD of the scaled matrix, small singular values (and with CUT7 the last
// seven of the decreasing order Eigen returns) dropped, x = SVecI V (U^T b ./ S).  The SVD is a one-sided Jacobi (Hestenes)
// iteration like the oracle's, but with the N/2 disjoint column pairs of a round-robin schedule rotated at once, 16 lanes
// per pair; the singular triplets agree with any other SVD up to rounding.  One CTA; A and V column-major in dynamic smem.
#define SVD_NT 576
__global__ void __launch_bounds__(SVD_NT) ba_svd_solve_kernel(int N, const double* __restrict__ HF, const double* __restrict__ bF,
                                                             double delta, int cut7, double* x)
{
	extern __shared__ double sm[];
	const int LD = N + 1;
	double* A = sm;                       // [col * LD + row]
	double* V = sm + (size_t)N * LD;
	double* sv = V + (size_t)N * LD;      // N
	double* bs = sv + N;                  // N
	double* S = bs + N;                   // N
	double* Ub = S + N;                   // N
	int* ord = reinterpret_cast<int*>(Ub + N);
	__shared__ unsigned long long s_off;
	const int tid = threadIdx.x, grp = tid >> 4, l = tid & 15, npairs = N >> 1;
	if (tid < N) sv[tid] = __ddiv_rn(1.0, sqrt(HF[(size_t)tid * N + tid]));
	__syncthreads();
	for (int e = tid; e < N * N; e += SVD_NT)
	{
		const int i = e / N, j = e - i * N;
		A[j * LD + i] = __dmul_rn(__dmul_rn(sv[i], HF[e]), sv[j]);
		V[j * LD + i] = (i == j) ? 1.0 : 0.0;
	}
	if (tid < N) bs[tid] = __dmul_rn(sv[tid], bF[tid]);
	__syncthreads();
	for (int sweep = 0; sweep < 60; sweep++)
	{
		if (tid == 0) s_off = 0ull;
		__syncthreads();
		for (int step = 0; step < N - 1; step++)
		{
			// circle method: player N-1 stays, the others rotate
			int p = 0, q = 0;
			const bool active = grp < npairs;
			if (active)
			{
				const int m = N - 1;
				const int a = grp == 0 ? N - 1 : (step + grp) % m;
				const int b = grp == 0 ? step % m : (step - grp + m) % m;
				p = min(a, b); q = max(a, b);
			}
			double alpha = 0.0, beta = 0.0, gamma = 0.0;
			if (active)
				for (int r = l; r < N; r += 16)
				{
					const double ap = A[p * LD + r], aq = A[q * LD + r];
					alpha = fma(ap, ap, alpha); beta = fma(aq, aq, beta); gamma = fma(ap, aq, gamma);
				}
#pragma unroll
			for (int o = 8; o > 0; o >>= 1)
			{
				alpha += __shfl_xor_sync(0xffffffffu, alpha, o);
				beta += __shfl_xor_sync(0xffffffffu, beta, o);
				gamma += __shfl_xor_sync(0xffffffffu, gamma, o);
			}
			if (active && fabs(gamma) > 1e-300)
			{
				const double nrm = sqrt(alpha * beta);
				if (l == 0) atomicMax(&s_off, (unsigned long long)__double_as_longlong(fabs(gamma) / nrm));
				if (fabs(gamma) > 1e-17 * nrm)
				{
					const double zeta = (beta - alpha) / (2.0 * gamma);
					const double t = (zeta >= 0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
					const double c = 1.0 / sqrt(1.0 + t * t), sn = c * t;
					for (int r = l; r < N; r += 16)
					{
						const double ap = A[p * LD + r], aq = A[q * LD + r];
						A[p * LD + r] = c * ap - sn * aq; A[q * LD + r] = sn * ap + c * aq;
						const double vp = V[p * LD + r], vq = V[q * LD + r];
						V[p * LD + r] = c * vp - sn * vq; V[q * LD + r] = sn * vp + c * vq;
					}
				}
			}
			__syncthreads();
		}
		if (__longlong_as_double((long long)s_off) < 1e-15) break;      // uniform: read after the step's barrier
		__syncthreads();
	}
	// singular values = column norms, U = normalised columns (kept in A)
	if (tid < N)
	{
		double n2 = 0.0;
		for (int r = 0; r < N; r++) n2 = fma(A[tid * LD + r], A[tid * LD + r], n2);
		S[tid] = sqrt(n2);
	}
	__syncthreads();
	if (tid < N)
	{
		// position of column tid in the decreasing order (ties by index, like a stable sort)
		const double v = S[tid];
		int rank = 0;
		for (int j = 0; j < N; j++) rank += (S[j] > v || (S[j] == v && j < tid)) ? 1 : 0;
		ord[rank] = tid;
	}
	__syncthreads();
	if (tid < N)
	{
		const int c = ord[tid];
		const double Sc = S[c];
		double maxSv = 0.0;
		for (int j = 0; j < N; j++) maxSv = fmax(maxSv, S[j]);
		double t = 0.0;
		for (int r = 0; r < N; r++) t += (Sc > 0 ? A[c * LD + r] / Sc : 0.0) * bs[r];      // U^T b
		if (Sc < delta * maxSv) t = 0.0;                                                   // :886-887
		if (cut7 && tid >= N - 7) t = 0.0;                                                 // :889-890
		else t /= Sc;
		Ub[tid] = t;
	}
	__syncthreads();
	if (tid < N)
	{
		double t = 0.0;
		for (int j = 0; j < N; j++) t += V[ord[j] * LD + tid] * Ub[j];
		x[tid] = sv[tid] * t;
	}
}

// what follows the solve in solveSystemF when the solve itself ran elsewhere: orthogonalize(&x, 0) (:910-914) and the xAd of
// resubstituteF_MT (:283-293), exactly as at the end of ba_solve_kernel
__global__ void __launch_bounds__(512) ba_solve_tail_kernel(SolveArgs a)
{
	__shared__ double xv[NMAX];
	const int tid = threadIdx.x, N = 8 * a.nF + CPARS;
	if (tid < N) xv[tid] = a.x[tid];
	__syncthreads();
	if (a.orth)
	{
		double s = 0.0;
		if (tid < N) for (int j = 0; j < N; j++) s = __dadd_rn(s, __dmul_rn(a.P[(size_t)tid * N + j], xv[j]));
		__syncthreads();
		if (tid < N) { xv[tid] = __dsub_rn(xv[tid], s); a.x[tid] = xv[tid]; }
		__syncthreads();
	}
	for (int e = tid; e < a.nF * a.nF * 8; e += 512)
	{
		const int j = e & 7, ht = e >> 3, h = ht / a.nF, t = ht - h * a.nF;
		const double* AH = a.adHost + (size_t)(h + a.nF * t) * 64;
		const double* AT_ = a.adTarget + (size_t)(h + a.nF * t) * 64;
		float ph[8], pt[8];
#pragma unroll
		for (int i = 0; i < 8; i++)
		{
			ph[i] = fmul((float)xv[CPARS + 8 * h + i], (float)AH[i * 8 + j]);
			pt[i] = fmul((float)xv[CPARS + 8 * t + i], (float)AT_[i * 8 + j]);
		}
		const float sh = fadd(fadd(fadd(ph[0], ph[1]), fadd(ph[2], ph[3])), fadd(fadd(ph[4], ph[5]), fadd(ph[6], ph[7])));
		const float st = fadd(fadd(fadd(pt[0], pt[1]), fadd(pt[2], pt[3])), fadd(fadd(pt[4], pt[5]), fadd(pt[6], pt[7])));
		a.xAd[e] = fadd(sh, st);
	}
}

// =========================================================================================================
// host API
// =========================================================================================================
extern "C" int dso_b200_ba_set_calib_delta(dso_b200_ctx* ctx, const double* value_minus_value_zero4)
{
	ARG_CHECK(ctx && ctx->ba && value_minus_value_zero4, "null argument");
	BAWorkspace* W = ctx->ba;
	CK(cudaSetDevice(ctx->device));
	for (int i = 0; i < 4; i++) W->h_cDeltaF[i] = (float)value_minus_value_zero4[i];   // cDeltaF = value_minus_value_zero.cast<float>() (:193)
	CK(cudaMemcpyAsync(W->d_cDeltaF, W->h_cDeltaF, 4 * sizeof(float), cudaMemcpyHostToDevice, ctx->ba_stream));
	CK(cudaStreamSynchronize(ctx->ba_stream));
	return 0;
}

extern "C" int dso_b200_ba_set_priors(dso_b200_ctx* ctx, const double* cPrior4, const double* framePrior, const double* frameDeltaPrior,
                                      const double* frameDelta, const double* HM, const double* bM)
{
	ARG_CHECK(ctx && ctx->ba, "null argument");
	BAWorkspace* W = ctx->ba;
	BA_CHECK(W->window_set, "call dso_b200_ba_set_window first");
	CK(cudaSetDevice(ctx->device));
	BASolve* S;
	int rc = solve_state(ctx, &S);
	if (rc) return rc;
	const int nF = W->P.nF, N = 8 * nF + CPARS;
	double ph[NMAX];
	for (int i = 0; i < N; i++) ph[i] = 0.0;
	if (cPrior4) for (int i = 0; i < 4; i++) ph[i] = cPrior4[i];
	if (framePrior) for (int i = 0; i < 8 * nF; i++) ph[CPARS + i] = framePrior[i];
	cudaStream_t s = ctx->ba_stream;
	CK(cudaMemcpyAsync(W->d_priorH, ph, N * sizeof(double), cudaMemcpyHostToDevice, s));
	// delta_prior / delta are also derived by dso_b200_ba_set_frames from the frame states; explicit arrays override them
	if (frameDeltaPrior) CK(cudaMemcpyAsync(W->d_priorD + CPARS, frameDeltaPrior, 8 * nF * sizeof(double), cudaMemcpyHostToDevice, s));
	if (frameDelta) CK(cudaMemcpyAsync(W->d_delta + CPARS, frameDelta, 8 * nF * sizeof(double), cudaMemcpyHostToDevice, s));
	if (HM) CK(cudaMemcpyAsync(S->d_HM, HM, (size_t)N * N * sizeof(double), cudaMemcpyHostToDevice, s));
	else CK(cudaMemsetAsync(S->d_HM, 0, (size_t)N * N * sizeof(double), s));
	if (bM) CK(cudaMemcpyAsync(S->d_bM, bM, N * sizeof(double), cudaMemcpyHostToDevice, s));
	else CK(cudaMemsetAsync(S->d_bM, 0, N * sizeof(double), s));
	CK(cudaStreamSynchronize(s));
	S->priors_set = true;
	return 0;
}

// thin SVD by one-sided Jacobi rotations: A (m x k, row-major) is overwritten by U*diag(S); V (k x k) accumulates the rotations
static void svd_hestenes(std::vector<double>& A, int m, int k, std::vector<double>& S, std::vector<double>& V)
{
	V.assign((size_t)k * k, 0.0);
	for (int i = 0; i < k; i++) V[i * k + i] = 1.0;
	for (int sweep = 0; sweep < 60; sweep++)
	{
		double off = 0.0;
		for (int p = 0; p < k - 1; p++)
			for (int q = p + 1; q < k; q++)
			{
				double al = 0, be = 0, ga = 0;
				for (int i = 0; i < m; i++) { const double x = A[i * k + p], y = A[i * k + q]; al += x * x; be += y * y; ga += x * y; }
				if (ga == 0.0 || al == 0.0 || be == 0.0) continue;
				const double rel = std::fabs(ga) / std::sqrt(al * be);
				if (rel > off) off = rel;
				if (rel <= 1e-17) continue;
				const double zeta = (be - al) / (2.0 * ga);
				const double t = (zeta >= 0 ? 1.0 : -1.0) / (std::fabs(zeta) + std::sqrt(1.0 + zeta * zeta));
				const double c = 1.0 / std::sqrt(1.0 + t * t), sn = c * t;
				for (int i = 0; i < m; i++) { const double x = A[i * k + p], y = A[i * k + q]; A[i * k + p] = c * x - sn * y; A[i * k + q] = sn * x + c * y; }
				for (int i = 0; i < k; i++) { const double x = V[i * k + p], y = V[i * k + q]; V[i * k + p] = c * x - sn * y; V[i * k + q] = sn * x + c * y; }
			}
		if (off < 1e-15) break;
	}
	S.assign(k, 0.0);
	for (int j = 0; j < k; j++)
	{
		double n2 = 0;
		for (int i = 0; i < m; i++) n2 += A[i * k + j] * A[i * k + j];
		S[j] = std::sqrt(n2);
	}
}

extern "C" int dso_b200_ba_set_nullspaces(dso_b200_ctx* ctx, const double* ns, int n, double solverModeDelta)
{
	ARG_CHECK(ctx && ctx->ba && (ns || n == 0) && n >= 0 && n <= 16, "bad argument");
	BAWorkspace* W = ctx->ba;
	BA_CHECK(W->window_set, "call dso_b200_ba_set_window first");
	CK(cudaSetDevice(ctx->device));
	BASolve* S;
	int rc = solve_state(ctx, &S);
	if (rc) return rc;
	if (n == 0) { S->haveP = false; return 0; }
	// EnergyFunctional::orthogonalize, EnergyFunctional.cpp:745-785: N = normalised columns, Npi = U diag(1/s) V^T (singular
	// values below solverModeDelta * max dropped), NNpiTS = 0.5 (N Npi^T + Npi N^T).  Host FP64, once per optimisation.
	const int N = 8 * W->P.nF + CPARS, k = n;
	std::vector<double> Nn((size_t)N * k), US, Sv, V;
	for (int j = 0; j < k; j++)
	{
		double n2 = 0;
		for (int i = 0; i < N; i++) n2 += ns[(size_t)j * N + i] * ns[(size_t)j * N + i];
		const double nrm = std::sqrt(n2);
		for (int i = 0; i < N; i++) Nn[(size_t)i * k + j] = ns[(size_t)j * N + i] / nrm;
	}
	US = Nn;
	svd_hestenes(US, N, k, Sv, V);
	double maxSv = 0;
	for (int j = 0; j < k; j++) if (Sv[j] > maxSv) maxSv = Sv[j];
	std::vector<double> Npi((size_t)N * k, 0.0), P((size_t)N * N), PS((size_t)N * N);
	for (int i = 0; i < N; i++)
		for (int j = 0; j < k; j++)
		{
			double s = 0;
			for (int q = 0; q < k; q++)
				if (Sv[q] > solverModeDelta * maxSv) s += (US[(size_t)i * k + q] / Sv[q]) * (1.0 / Sv[q]) * V[(size_t)j * k + q];   // U = US / s, then * (1/s)
			Npi[(size_t)i * k + j] = s;
		}
	for (int i = 0; i < N; i++)
		for (int j = 0; j < N; j++)
		{
			double s = 0;
			for (int q = 0; q < k; q++) s += Nn[(size_t)i * k + q] * Npi[(size_t)j * k + q];
			P[(size_t)i * N + j] = s;
		}
	for (int i = 0; i < N; i++) for (int j = 0; j < N; j++) PS[(size_t)i * N + j] = 0.5 * (P[(size_t)i * N + j] + P[(size_t)j * N + i]);
	CK(cudaMemcpyAsync(S->d_P, PS.data(), (size_t)N * N * sizeof(double), cudaMemcpyHostToDevice, ctx->ba_stream));
	CK(cudaStreamSynchronize(ctx->ba_stream));
	S->haveP = true;
	S->solverModeDelta = solverModeDelta;
	return 0;
}

// The global state the non-default solver branches read besides setting_solverMode: whether a frame with frameID 0 is
// in the window ("haveFirstFrame", EnergyFunctional.cpp:686-688, :833-835) and the solver mode / delta marginalizePointsF
// sees (util/settings.cpp:60-61).  Defaults: first frame present, 128 | 2048, 1e-5.
extern "C" int dso_b200_ba_set_solver_state(dso_b200_ctx* ctx, int haveFirstFrame, int margSolverMode, double solverModeDelta)
{
	ARG_CHECK(ctx && ctx->ba, "null context");
	BASolve* S;
	int rc = solve_state(ctx, &S);
	if (rc) return rc;
	S->haveFirstFrame = haveFirstFrame ? 1 : 0;
	S->margSolverMode = margSolverMode;
	S->solverModeDelta = solverModeDelta;
	ba_graphs_invalidate(ctx->ba);
	return 0;
}

extern "C" int dso_b200_ba_fix_linearization(dso_b200_ctx* ctx, const unsigned char* which, float* res_toZeroF_out)
{
	int rc = ba_check_ready(ctx, __func__);
	if (rc) return rc;
	CK(cudaSetDevice(ctx->device));
	BAWorkspace* W = ctx->ba;
	BASolve* S;
	if ((rc = solve_state(ctx, &S))) return rc;
	const int nR = W->P.nR;
	cudaStream_t s = ctx->ba_stream;
	if (which)
	{
		if (nR > S->capR)
		{
			if (S->capR) cudaFree(S->d_which);
			CK(cudaMalloc(&S->d_which, (size_t)nR));
			S->capR = nR;
		}
		CK(cudaMemcpyAsync(S->d_which, which, (size_t)nR, cudaMemcpyHostToDevice, s));
	}
	ba_fix_lin_kernel<<<(nR + 255) / 256, 256, 0, s>>>(nR, W->P.nF, which ? S->d_which : nullptr, W->d_isAct, W->d_isLin, W->d_efsel, W->d_J[0], W->d_J[1],
	                                                 W->d_rpoint, W->d_rtarget, W->d_phost, W->d_deltaF, W->d_adHTdeltaF, W->d_cDeltaF, W->d_resToZero);
	LAUNCHED(ctx, 1);
	CK(cudaGetLastError());
	if (res_toZeroF_out) CK(cudaMemcpyAsync(res_toZeroF_out, W->d_resToZero, (size_t)nR * 8 * sizeof(float), cudaMemcpyDeviceToHost, s));
	CK(cudaStreamSynchronize(s));
	W->nLin = -1;                  // some residuals may now be linearised: accumulateLF has work
	ba_graphs_invalidate(W);
	W->top_done = false;
	return 0;
}

extern "C" int dso_b200_ba_accumulate_L(dso_b200_ctx* ctx, double* H, double* b, int* resInL,
                                        float* Hdd_accLF, float* bd_accLF, float* Hcd_accLF4)
{
	int rc = ba_check_ready(ctx, __func__);
	if (rc) return rc;
	ARG_CHECK(H && b, "null output");
	CK(cudaSetDevice(ctx->device));
	BAWorkspace* W = ctx->ba;
	if ((rc = ba_enqueue_top(ctx, 1))) return rc;
	const int N = 8 * W->P.nF + CPARS;
	const size_t nP = W->P.nP;
	cudaStream_t s = ctx->ba_stream;
	CK(cudaMemcpyAsync(W->h_out, W->d_HL, (size_t)N * N * sizeof(double), cudaMemcpyDeviceToHost, s));
	CK(cudaMemcpyAsync(W->h_out + N * N, W->d_bL, N * sizeof(double), cudaMemcpyDeviceToHost, s));
	CK(cudaMemcpyAsync(W->h_out + N * N + N, W->d_scalars + 2, 2 * sizeof(double), cudaMemcpyDeviceToHost, s));
	if (Hdd_accLF) CK(cudaMemcpyAsync(Hdd_accLF, W->d_Hdd_L, nP * sizeof(float), cudaMemcpyDeviceToHost, s));
	if (bd_accLF) CK(cudaMemcpyAsync(bd_accLF, W->d_bd_L, nP * sizeof(float), cudaMemcpyDeviceToHost, s));
	if (Hcd_accLF4) CK(cudaMemcpyAsync(Hcd_accLF4, W->d_Hcd_L, nP * 4 * sizeof(float), cudaMemcpyDeviceToHost, s));
	CK(cudaStreamSynchronize(s));
	memcpy(H, W->h_out, (size_t)N * N * sizeof(double));
	memcpy(b, W->h_out + N * N, N * sizeof(double));
	if (resInL) *resInL = (int)W->h_out[N * N + N + 1];
	return 0;
}

// EnergyFunctional::solveSystemF (EnergyFunctional.cpp:802-924): accumulateAF_MT, accumulateLF_MT, accumulateSCF_MT, the
// solve and resubstituteF_MT, all enqueued on the context stream; one synchronisation at the end.
static int enqueue_solve(dso_b200_ctx* ctx, BASolve* S, int iteration, double lambda, int solverMode, bool debug, bool withStep = false, float stepfacD = 0.f)
{
	BAWorkspace* W = ctx->ba;
	int rc;
	if (solverMode & SOLVER_USE_GN) lambda = 0;                      // :804
	if (solverMode & SOLVER_FIX_LAMBDA) lambda = 1e-5;              // :805
	// the stitch of the last top accumulation runs beside the Schur chain (the two top modes share their stitch buffers, so
	// only the last one may be forked)
	const bool haveL = W->nLin != 0;
	// the last top accumulation also prepares the Schur step per point (one launch less on the critical path)
	if ((rc = ba_enqueue_top(ctx, 0, nullptr, !haveL, !haveL))) return rc;  // :816
	if (haveL && (rc = ba_enqueue_top(ctx, 1, nullptr, true, true))) return rc;   // :817 (without linearised residuals H_L, b_L are the priors alone)
	if ((rc = ba_enqueue_schur(ctx, nullptr, true))) return rc;     // :820
	if ((rc = ba_join_side(ctx))) return rc;
	if ((rc = ensure_point_buffers(ctx, S))) return rc;
	SolveArgs a;
	a.nF = W->P.nF; a.lambda = lambda; a.haveL = haveL ? 1 : 0;
	a.orth = ((solverMode & SOLVER_ORTHOGONALIZE_X) || (iteration >= 2 && (solverMode & SOLVER_ORTHOGONALIZE_X_LATER))) && S->haveP;   // :913
	a.HA = W->d_HA; a.bA = W->d_bA; a.HL = W->d_HL; a.bL = W->d_bL; a.HS = W->d_HS; a.bS = W->d_bS;
	a.HM = S->d_HM; a.bM = S->d_bM; a.delta = W->d_delta; a.priorH = W->d_priorH; a.priorD = W->d_priorD; a.cDeltaF = W->d_cDeltaF;
	a.P = S->d_P; a.adHost = W->d_adHost; a.adTarget = W->d_adTarget;
	a.x = S->d_x; a.xAd = S->d_xAd; a.HF = debug ? S->d_HF : nullptr; a.bF = debug ? S->d_bF : nullptr;
	a.dbg = S->d_dbg;
	a.Hpre = nullptr; a.bpre = nullptr; a.assemble_only = 0;
	const int N = 8 * W->P.nF + CPARS;
	cudaStream_t s = ctx->ba_stream;
	if (solverMode & (SOLVER_SVD | SOLVER_ORTHOGONALIZE_SYSTEM))
	{
		if ((rc = nondefault_buffers(S))) return rc;
		BA_CHECK(!(solverMode & SOLVER_ORTHOGONALIZE_SYSTEM) || S->haveFirstFrame || S->haveP,
		         "SOLVER_ORTHOGONALIZE_SYSTEM without the first frame needs dso_b200_ba_set_nullspaces");
	}
	if (solverMode & SOLVER_ORTHOGONALIZE_SYSTEM)
	{
		// :832-851: HT_act / bT_act, projected off the null spaces unless the gauge-fixing first frame is in the window
		double* Hact = S->d_work; double* bact = Hact + (size_t)NMAX * NMAX; double* T = bact + NMAX;
		ba_assemble_act_kernel<<<(N * N + 255) / 256, 256, 0, s>>>(N, a.haveL, W->d_HL, W->d_bL, W->d_HA, W->d_bA, W->d_HS, W->d_bS,
		                                                          W->d_priorH, W->d_priorD, W->d_cDeltaF, Hact, bact);
		LAUNCHED(ctx, 1);
		if (!S->haveFirstFrame)
		{
			ba_orth_stage1_kernel<<<N + 1, 128, 0, s>>>(N, S->d_P, Hact, bact, T);
			ba_orth_stage2_kernel<<<N + 1, 128, 0, s>>>(N, S->d_P, Hact, bact, T);
			LAUNCHED(ctx, 2);
		}
		a.Hpre = Hact; a.bpre = bact;
	}
	if (S->d_dbg) CK(cudaMemsetAsync(S->d_dbg, 0, 16 * sizeof(long long), s));
	if (solverMode & SOLVER_SVD)
	{
		// :870-900: the assembled system goes through global memory to the SVD kernel; orthogonalize(x) and xAd follow
		a.assemble_only = 1; a.HF = S->d_HF; a.bF = S->d_bF;
		ba_solve_kernel<<<1, SOLVE_NT, 0, s>>>(a);
		const size_t smem = ((size_t)2 * N * (N + 1) + 4 * N) * sizeof(double) + (size_t)N * sizeof(int);
		static bool attr[64] = { false };
		if (!attr[ctx->device & 63]) { CK(cudaFuncSetAttribute(ba_svd_solve_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024)); attr[ctx->device & 63] = true; }
		ba_svd_solve_kernel<<<1, SVD_NT, smem, s>>>(N, S->d_HF, S->d_bF, S->solverModeDelta, (solverMode & SOLVER_SVD_CUT7) ? 1 : 0, S->d_x);
		ba_solve_tail_kernel<<<1, 512, 0, s>>>(a);
		LAUNCHED(ctx, 2);
	}
	else
		ba_solve_kernel<<<1, SOLVE_NT, 0, s>>>(a);
	if (withStep)
	{
		const int nb = (W->P.nP + 255) / 256;
		ba_resub_step_kernel<<<nb, 256, 0, ctx->ba_stream>>>(W->P.nP, W->P.nF, W->d_res_begin, W->d_rtarget, W->d_phost, W->d_isAct,
		                                                     W->d_JpJdF, W->d_HcdSum, W->d_bdSumF, W->d_HdiF, S->d_x, S->d_xAd, S->d_step,
		                                                     stepfacD, S->d_idepth_backup, W->d_pt, W->d_deltaF, S->d_stepBlocks,
		                                                     reinterpret_cast<unsigned int*>(S->d_stepBlocks + 2 * nb), W->d_scalars + 4);
	}
	else
		ba_resubstitute_kernel<<<(W->P.nP + 127) / 128, 128, 0, ctx->ba_stream>>>(W->P.nP, W->P.nF, W->d_res_begin, W->d_rtarget, W->d_phost, W->d_isAct,
		                                                                    W->d_JpJdF, W->d_HcdSum, W->d_bdSumF, W->d_HdiF, S->d_x, S->d_xAd, S->d_step);
	LAUNCHED(ctx, 2);
	CK(cudaGetLastError());
	return 0;
}

extern "C" int dso_b200_ba_solve(dso_b200_ctx* ctx, int iteration, double lambda, int solverMode, double* x_out, float* pointStep_out,
                                 double* HFinal_out, double* bFinal_out)
{
	int rc = ba_check_ready(ctx, __func__);
	if (rc) return rc;
	CK(cudaSetDevice(ctx->device));
	BAWorkspace* W = ctx->ba;
	BASolve* S;
	if ((rc = solve_state(ctx, &S))) return rc;
	const bool debug = HFinal_out || bFinal_out;
	if ((rc = enqueue_solve(ctx, S, iteration, lambda, solverMode, debug))) return rc;
	const int N = 8 * W->P.nF + CPARS;
	cudaStream_t s = ctx->ba_stream;
	CK(cudaMemcpyAsync(S->h_x, S->d_x, N * sizeof(double), cudaMemcpyDeviceToHost, s));
	if (pointStep_out) CK(cudaMemcpyAsync(pointStep_out, S->d_step, (size_t)W->P.nP * sizeof(float), cudaMemcpyDeviceToHost, s));
	if (HFinal_out) CK(cudaMemcpyAsync(HFinal_out, S->d_HF, (size_t)N * N * sizeof(double), cudaMemcpyDeviceToHost, s));
	if (bFinal_out) CK(cudaMemcpyAsync(bFinal_out, S->d_bF, N * sizeof(double), cudaMemcpyDeviceToHost, s));
	CK(cudaStreamSynchronize(s));
	if (x_out) memcpy(x_out, S->h_x, N * sizeof(double));
	if (S->d_dbg)
	{
		long long t[11];
		CK(cudaMemcpy(t, S->d_dbg, sizeof(t), cudaMemcpyDeviceToHost));
		fprintf(stderr, "ba_solve_kernel cycles: stageHM %lld rhs %lld assemble %lld scale %lld pivots %lld load %lld ldlt %lld subst %lld orth %lld xAd %lld | total %lld | %lld\n",
		        t[1] - t[0], 0LL, t[2] - t[1], t[3] - t[2], t[4] - t[3], t[5] - t[4], t[6] - t[5], t[7] - t[6], t[8] - t[7], t[9] - t[8], t[9] - t[0], t[10]);
	}
	return 0;
}

// One Gauss-Newton iteration of FullSystem::optimize as far as it lives on the device: linearizeAll(false) + applyRes
// (the caller decided to accept the previous step), solveSystemF, and the point half of doStepFromBackup.
extern "C" int dso_b200_ba_iteration(dso_b200_ctx* ctx, int iteration, double lambda, int solverMode, float stepfacD,
                                     double* x_out, double* energy_sum, float* step_sums2)
{
	int rc = ba_check_ready(ctx, __func__);
	if (rc) return rc;
	CK(cudaSetDevice(ctx->device));
	BAWorkspace* W = ctx->ba;
	BASolve* S;
	if ((rc = solve_state(ctx, &S))) return rc;
	if ((rc = ensure_point_buffers(ctx, S))) return rc;
	cudaStream_t s = ctx->ba_stream;
	const int N = 8 * W->P.nF + CPARS;
	// by-value kernel arguments of the captured sequence: orthogonalisation on/off, lambda, stepfacD, solver mode, L path
	const int orth = ((solverMode & SOLVER_ORTHOGONALIZE_X) || (iteration >= 2 && (solverMode & SOLVER_ORTHOGONALIZE_X_LATER))) ? 1 : 0;
	const BAGraphKey key = { lambda, stepfacD, solverMode | (S->haveFirstFrame << 20), W->nLin != 0 ? 1 : 0, S->haveP ? 1 : 0, S->d_dbg ? 1 : 0 };
	rc = ba_run_graphed(ctx, W->g_iter[orth], key, [&]() -> int {
		int r;
		if ((r = ba_enqueue_linearize(ctx, true, true))) return r;        // linearize + applyRes in one kernel
		// solve, then resubstitute + backup + point step + step sums in one launch
		if ((r = enqueue_solve(ctx, S, orth ? 2 : 0, lambda, solverMode, false, true, stepfacD))) return r;
		CK(cudaMemcpyAsync(S->h_x, S->d_x, N * sizeof(double), cudaMemcpyDeviceToHost, s));
		CK(cudaMemcpyAsync(S->h_x + NMAX, W->d_scalars, 6 * sizeof(double), cudaMemcpyDeviceToHost, s));
		return 0;
	});
	if (rc) return rc;
	W->top_done = true;
	CK(cudaStreamSynchronize(s));
	if (x_out) memcpy(x_out, S->h_x, N * sizeof(double));
	if (energy_sum) *energy_sum = S->h_x[NMAX];
	if (step_sums2) { step_sums2[0] = (float)S->h_x[NMAX + 4]; step_sums2[1] = (float)S->h_x[NMAX + 5]; }
	W->lin_done = false;
	return 0;
}

extern "C" int dso_b200_ba_points_backup(dso_b200_ctx* ctx)
{
	int rc = ba_check_ready(ctx, __func__);
	if (rc) return rc;
	CK(cudaSetDevice(ctx->device));
	BAWorkspace* W = ctx->ba;
	BASolve* S;
	if ((rc = solve_state(ctx, &S))) return rc;
	if ((rc = ensure_point_buffers(ctx, S))) return rc;
	ba_points_backup_kernel<<<(W->P.nP + 255) / 256, 256, 0, ctx->ba_stream>>>(W->P.nP, W->d_pt, S->d_idepth_backup, 0, nullptr, nullptr);
	LAUNCHED(ctx, 1);
	CK(cudaGetLastError());
	return 0;
}

extern "C" int dso_b200_ba_points_step(dso_b200_ctx* ctx, float stepfacD, float* sums2_out, float* idepth_out)
{
	int rc = ba_check_ready(ctx, __func__);
	if (rc) return rc;
	CK(cudaSetDevice(ctx->device));
	BAWorkspace* W = ctx->ba;
	BASolve* S;
	if ((rc = solve_state(ctx, &S))) return rc;
	BA_CHECK(S->capP >= W->P.nP, "call dso_b200_ba_points_backup and dso_b200_ba_solve first");
	cudaStream_t s = ctx->ba_stream;
	const int nP = W->P.nP, nb = (nP + 255) / 256;
	ba_points_step_kernel<<<nb, 256, 0, s>>>(nP, stepfacD, S->d_step, S->d_idepth_backup, W->d_pt, W->d_deltaF, S->d_stepBlocks);
	ba_points_step_final_kernel<<<1, 32, 0, s>>>(nb, S->d_stepBlocks, W->d_scalars + 4);
	LAUNCHED(ctx, 2);
	CK(cudaGetLastError());
	CK(cudaMemcpyAsync(S->h_x + NMAX, W->d_scalars + 4, 2 * sizeof(double), cudaMemcpyDeviceToHost, s));
	if (idepth_out)
	{
		// idepth is component z of the packed point record
		CK(cudaMemcpy2DAsync(idepth_out, sizeof(float), (const char*)W->d_pt + 8, sizeof(float4), sizeof(float), nP, cudaMemcpyDeviceToHost, s));
	}
	CK(cudaStreamSynchronize(s));
	if (sums2_out) { sums2_out[0] = (float)S->h_x[NMAX]; sums2_out[1] = (float)S->h_x[NMAX + 1]; }
	W->lin_done = false; W->top_done = false;
	return 0;
}

extern "C" int dso_b200_ba_points_restore(dso_b200_ctx* ctx)
{
	int rc = ba_check_ready(ctx, __func__);
	if (rc) return rc;
	CK(cudaSetDevice(ctx->device));
	BAWorkspace* W = ctx->ba;
	BASolve* S;
	if ((rc = solve_state(ctx, &S))) return rc;
	BA_CHECK(S->capP >= W->P.nP, "call dso_b200_ba_points_backup first");
	ba_points_backup_kernel<<<(W->P.nP + 255) / 256, 256, 0, ctx->ba_stream>>>(W->P.nP, W->d_pt, S->d_idepth_backup, 1, W->d_pt, W->d_deltaF);
	LAUNCHED(ctx, 1);
	CK(cudaGetLastError());
	W->lin_done = false; W->top_done = false;
	return 0;
}

// =========================================================================================================
// rest of N2: marginalizePointsF (EnergyFunctional.cpp:642-703), calcLEnergyF_MT (:346-428), calcMEnergyF (:334-343)
// =========================================================================================================
__global__ void ba_marg_prior_kernel(int nP, const unsigned char* __restrict__ pmask, float fac, float* priorF)
{
	const int p = blockIdx.x * blockDim.x + threadIdx.x;
	if (p < nP && pmask[p]) priorF[p] = fmul(priorF[p], fac);          // :656 p->priorF *= setting_idepthFixPriorMargFac
}

// H = M - Msc, b = Mb - Mbsc (:681-682); HM += w H, bM += w b (:694-695)
__global__ void ba_marg_combine_kernel(int N, double w, const double* __restrict__ M, const double* __restrict__ Mb,
                                       const double* __restrict__ Msc, const double* __restrict__ Mbsc,
                                       double* H, double* b, double* HM, double* bM)
{
	const int e = blockIdx.x * blockDim.x + threadIdx.x;
	if (e < N * N)
	{
		const double h = __dsub_rn(M[e], Msc[e]);
		H[e] = h;
		HM[e] = __dadd_rn(HM[e], __dmul_rn(w, h));
	}
	if (e < N)
	{
		const double v = __dsub_rn(Mb[e], Mbsc[e]);
		b[e] = v;
		bM[e] = __dadd_rn(bM[e], __dmul_rn(w, v));
	}
}

// calcLEnergyPt (:346-408) per point: the float terms of the reference (Jdelta * (2 res_toZero + Jdelta) per pattern pixel of
// every linearised active residual, deltaF^2 priorF), summed in FP64 by a fixed tree instead of the Accumulator11 float tiers
__global__ void __launch_bounds__(256) ba_lenergy_kernel(int nP, int nF, const int* __restrict__ res_begin, const int* __restrict__ rtarget,
                                                         const int* __restrict__ phost, const unsigned char* __restrict__ isAct,
                                                         const unsigned char* __restrict__ isLin, const unsigned char* __restrict__ efsel,
                                                         const float* __restrict__ J0, const float* __restrict__ J1,
                                                         const float* __restrict__ resToZero, const float* __restrict__ adHTdeltaF,
                                                         const float* __restrict__ cDeltaF, const float* __restrict__ deltaF,
                                                         const float* __restrict__ priorF, double* blocks)
{
	__shared__ double s[256];
	const int p = blockIdx.x * 256 + threadIdx.x;
	double e = 0.0;
	if (p < nP)
	{
		const float dd = deltaF[p];
		const int h = phost[p];
		for (int r = res_begin[p]; r < res_begin[p + 1]; r++)
		{
			if (!isLin[r] || !isAct[r]) continue;
			const float* J = (efsel[r] ? J1 : J0) + (size_t)r * JSZ;
			float d[4];
			jp_delta(J, adHTdeltaF + 8 * (h + nF * rtarget[r]), cDeltaF, dd, d);
#pragma unroll
			for (int i = 0; i < 8; i++)
			{
				float Jd = fmul(J[J_IDX0 + i], d[0]);
				Jd = fadd(Jd, fmul(J[J_IDX1 + i], d[1]));
				Jd = fadd(Jd, fmul(J[J_AB0 + i], d[2]));
				Jd = fadd(Jd, fmul(J[J_AB1 + i], d[3]));
				float r0 = resToZero[8 * (size_t)r + i];
				r0 = fadd(r0, r0);
				r0 = fadd(r0, Jd);
				e += (double)fmul(Jd, r0);
			}
		}
		e += (double)fmul(fmul(dd, dd), priorF[p]);
	}
	s[threadIdx.x] = e;
	__syncthreads();
	for (int o = 128; o > 0; o >>= 1)
	{
		if (threadIdx.x < o) s[threadIdx.x] += s[threadIdx.x + o];
		__syncthreads();
	}
	if (threadIdx.x == 0) blocks[blockIdx.x] = s[0];
}

// out[0] = calcLEnergyF_MT: frame and calibration prior terms (:416-420) + the point blocks; out[1] = calcMEnergyF
__global__ void __launch_bounds__(128) ba_energy_final_kernel2(int nF, int nb, const double* __restrict__ blocks, const double* __restrict__ priorH,
                                                               const double* __restrict__ priorD, const double* __restrict__ delta,
                                                               const float* __restrict__ cDeltaF, const double* __restrict__ HM,
                                                               const double* __restrict__ bM, double* out)
{
	__shared__ double s[128];
	const int tid = threadIdx.x, N = 8 * nF + CPARS;
	double m = 0.0;                                            // delta . (2 bM + HM delta)
	if (tid < N)
	{
		double acc = 0.0;
		for (int j = 0; j < N; j++) acc = __dadd_rn(acc, __dmul_rn(HM[(size_t)tid * N + j], j < CPARS ? (double)cDeltaF[j] : delta[j]));
		const double di = tid < CPARS ? (double)cDeltaF[tid] : delta[tid];
		m = di * (2.0 * bM[tid] + acc);
	}
	s[tid] = m;
	__syncthreads();
	if (tid == 0)
	{
		double t = 0;
		for (int i = 0; i < N; i++) t += s[i];
		out[1] = t;
		double E = 0.0;
		for (int h = 0; h < nF; h++)
		{
			double q[8];
			for (int i = 0; i < 8; i++) { const int k = CPARS + 8 * h + i; q[i] = (priorD[k] * priorH[k]) * priorD[k]; }
			E += ((q[0] + q[1]) + (q[2] + q[3])) + ((q[4] + q[5]) + (q[6] + q[7]));
		}
		float c[4];
		for (int i = 0; i < 4; i++) c[i] = fmul(fmul(cDeltaF[i], (float)priorH[i]), cDeltaF[i]);
		E += (double)fadd(fadd(c[0], c[1]), fadd(c[2], c[3]));
		double pts = 0.0;
		for (int i = 0; i < nb; i++) pts += blocks[i];
		out[0] = E + pts;
	}
}

static int marg_buffers(dso_b200_ctx* ctx, BASolve* S)
{
	BAWorkspace* W = ctx->ba;
	if (!S->d_margH) CK(cudaMalloc(&S->d_margH, ((size_t)NMAX * NMAX + NMAX) * sizeof(double)));
	if (W->P.nP > S->capPm)
	{
		if (S->capPm) { cudaFree(S->d_pmask); cudaFree(S->d_eblocks); }
		CK(cudaMalloc(&S->d_pmask, (size_t)W->P.nP));
		CK(cudaMalloc(&S->d_eblocks, ((size_t)W->P.nP / 256 + 1) * sizeof(double)));
		S->capPm = W->P.nP;
	}
	return 0;
}

extern "C" int dso_b200_ba_marginalize_points(dso_b200_ctx* ctx, const unsigned char* point_flags, float idepthFixPriorMargFac,
                                              double margWeightFac, double* H_out, double* b_out, int* resInM, float* idepth_hessian_out)
{
	int rc = ba_check_ready(ctx, __func__);
	if (rc) return rc;
	ARG_CHECK(point_flags, "null flags");
	CK(cudaSetDevice(ctx->device));
	BAWorkspace* W = ctx->ba;
	BASolve* S;
	if ((rc = solve_state(ctx, &S))) return rc;
	if ((rc = marg_buffers(ctx, S))) return rc;
	const int nP = W->P.nP, N = 8 * W->P.nF + CPARS;
	cudaStream_t s = ctx->ba_stream;
	CK(cudaMemcpyAsync(S->d_pmask, point_flags, (size_t)nP, cudaMemcpyHostToDevice, s));
	ba_marg_prior_kernel<<<(nP + 255) / 256, 256, 0, s>>>(nP, S->d_pmask, idepthFixPriorMargFac, W->d_priorF);
	LAUNCHED(ctx, 1);
	if ((rc = ba_enqueue_top(ctx, 2, S->d_pmask))) return rc;      // accSSE_top_A->addPoint<2> + stitchDouble -> d_HL, d_bL
	if ((rc = ba_enqueue_schur(ctx, S->d_pmask))) return rc;       // accSSE_bot->addPoint(p, false) + stitchDouble -> d_HS, d_bS
	double* dH = S->d_margH; double* db = S->d_margH + (size_t)NMAX * NMAX;
	const int orthPt = (S->margSolverMode & SOLVER_ORTHOGONALIZE_POINTMARG) && !S->haveFirstFrame, orthFull = S->margSolverMode & SOLVER_ORTHOGONALIZE_FULL;
	if (!orthPt && !orthFull)
	{
		ba_marg_combine_kernel<<<(N * N + 255) / 256, 256, 0, s>>>(N, margWeightFac, W->d_HL, W->d_bL, W->d_HS, W->d_bS, dH, db, S->d_HM, S->d_bM);
		LAUNCHED(ctx, 1);
	}
	else
	{
		// :681-699 with SOLVER_ORTHOGONALIZE_POINTMARG (H, b of the marginalised points projected, first frame gone) and / or
		// SOLVER_ORTHOGONALIZE_FULL (the accumulated prior projected)
		BA_CHECK(S->haveP, "the orthogonalising marginalisation modes need dso_b200_ba_set_nullspaces");
		if ((rc = nondefault_buffers(S))) return rc;
		double* T = S->d_work;
		ba_marg_diff_kernel<<<(N * N + 255) / 256, 256, 0, s>>>(N, W->d_HL, W->d_bL, W->d_HS, W->d_bS, dH, db);
		LAUNCHED(ctx, 1);
		if (orthPt)
		{
			ba_orth_stage1_kernel<<<N + 1, 128, 0, s>>>(N, S->d_P, dH, db, T);
			ba_orth_stage2_kernel<<<N + 1, 128, 0, s>>>(N, S->d_P, dH, db, T);
			LAUNCHED(ctx, 2);
		}
		ba_marg_accumulate_kernel<<<(N * N + 255) / 256, 256, 0, s>>>(N, margWeightFac, dH, db, S->d_HM, S->d_bM);
		LAUNCHED(ctx, 1);
		if (orthFull)
		{
			ba_orth_stage1_kernel<<<N + 1, 128, 0, s>>>(N, S->d_P, S->d_HM, S->d_bM, T);
			ba_orth_stage2_kernel<<<N + 1, 128, 0, s>>>(N, S->d_P, S->d_HM, S->d_bM, T);
			LAUNCHED(ctx, 2);
		}
	}
	CK(cudaGetLastError());
	if (H_out) CK(cudaMemcpyAsync(H_out, dH, (size_t)N * N * sizeof(double), cudaMemcpyDeviceToHost, s));
	if (b_out) CK(cudaMemcpyAsync(b_out, db, N * sizeof(double), cudaMemcpyDeviceToHost, s));
	if (idepth_hessian_out) CK(cudaMemcpyAsync(idepth_hessian_out, W->d_idh, (size_t)nP * sizeof(float), cudaMemcpyDeviceToHost, s));
	CK(cudaMemcpyAsync(S->h_x + NMAX, W->d_scalars + 2, 2 * sizeof(double), cudaMemcpyDeviceToHost, s));
	CK(cudaStreamSynchronize(s));
	if (resInM) *resInM = (int)S->h_x[NMAX + 1];
	W->top_done = false; W->lin_done = false;
	ba_graphs_invalidate(W);
	return 0;
}

extern "C" int dso_b200_ba_get_marg_prior(dso_b200_ctx* ctx, double* HM_out, double* bM_out)
{
	ARG_CHECK(ctx && ctx->ba && HM_out && bM_out, "null argument");
	BAWorkspace* W = ctx->ba;
	BA_CHECK(W->window_set, "call dso_b200_ba_set_window first");
	CK(cudaSetDevice(ctx->device));
	BASolve* S;
	int rc = solve_state(ctx, &S);
	if (rc) return rc;
	const int N = 8 * W->P.nF + CPARS;
	CK(cudaMemcpyAsync(HM_out, S->d_HM, (size_t)N * N * sizeof(double), cudaMemcpyDeviceToHost, ctx->ba_stream));
	CK(cudaMemcpyAsync(bM_out, S->d_bM, N * sizeof(double), cudaMemcpyDeviceToHost, ctx->ba_stream));
	CK(cudaStreamSynchronize(ctx->ba_stream));
	return 0;
}

extern "C" int dso_b200_ba_calc_energies(dso_b200_ctx* ctx, double* L_energy, double* M_energy)
{
	int rc = ba_check_ready(ctx, __func__);
	if (rc) return rc;
	CK(cudaSetDevice(ctx->device));
	BAWorkspace* W = ctx->ba;
	BASolve* S;
	if ((rc = solve_state(ctx, &S))) return rc;
	if ((rc = marg_buffers(ctx, S))) return rc;
	const int nP = W->P.nP, nb = (nP + 255) / 256;
	cudaStream_t s = ctx->ba_stream;
	ba_lenergy_kernel<<<nb, 256, 0, s>>>(nP, W->P.nF, W->d_res_begin, W->d_rtarget, W->d_phost, W->d_isAct, W->d_isLin, W->d_efsel, W->d_J[0], W->d_J[1],
	                                     W->d_resToZero, W->d_adHTdeltaF, W->d_cDeltaF, W->d_deltaF, W->d_priorF, S->d_eblocks);
	ba_energy_final_kernel2<<<1, 128, 0, s>>>(W->P.nF, nb, S->d_eblocks, W->d_priorH, W->d_priorD, W->d_delta, W->d_cDeltaF, S->d_HM, S->d_bM, W->d_scalars + 6);
	LAUNCHED(ctx, 2);
	CK(cudaGetLastError());
	CK(cudaMemcpyAsync(S->h_x + NMAX, W->d_scalars + 6, 2 * sizeof(double), cudaMemcpyDeviceToHost, s));
	CK(cudaStreamSynchronize(s));
	if (L_energy) *L_energy = S->h_x[NMAX];
	if (M_energy) *M_energy = S->h_x[NMAX + 1];
	return 0;
}
