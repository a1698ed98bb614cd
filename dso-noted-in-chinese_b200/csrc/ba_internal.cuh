// Shared definitions of the windowed-BA kernels (ba.cu: linearize / accumulate; ba_solve.cu: solve / resubstitute).
#pragma once
#include "common.cuh"
#include <vector>

#define CPARS 4
#define BA_OUT_DOUBLES (2 * (size_t)(8 * MAXF + CPARS) * (8 * MAXF + CPARS) + 2 * (8 * MAXF + CPARS) + 8)
#define CHUNK_R 64
#define CHUNK_P 16
#define NTOP 91           // 55 + 30 + 6 unique entries of the 13x13 block
#define NTOP_PAD 96
#define JSZ 74
#define ST_IN 0
#define ST_OOB 1
#define ST_OUTLIER 2
#define SCALE_IDEPTH 1.0f
#define SCALE_F 50.0f
#define SCALE_C 50.0f

// offsets into the flattened RawResidualJacobian (RawResidualJacobian.h:40-79)
#define J_RESF 0
#define J_PDXI0 8
#define J_PDXI1 14
#define J_PDC0 20
#define J_PDC1 24
#define J_PDD 28
#define J_IDX0 30
#define J_IDX1 38
#define J_AB0 46
#define J_AB1 54
#define J_IDX2 62
#define J_ABIDX 66
#define J_AB2 70


struct BAParams
{
	int nF, w, h, nR, nP;
	float fx, fy, cx, cy, fxi, fyi, wM3G, hM3G;
	float frameEnergyTH[MAXF];
	const float4* img[MAXF];
	int fixA, fixB;            // setting_affineOptModeA/B < 0 (Residuals.cpp:257-258)
	float huberTH, outlierTHSum;   // setting_huberTH, setting_outlierTHSumComponent (dso_b200_set_settings)
};

// inputs of the linearised-residual mode (addPoint<1>, fixLinearizationF)
struct BALinArgs
{
	const int* rpoint; const int* rtarget; const int* phost; const float* deltaF;
	const float* adHTdeltaF; const float* cDeltaF; const float* resToZero;
	const unsigned char* pmask;        // mode 2: points flagged PS_MARGINALIZE
	int nF;
};

struct BASolve;    // ba_solve.cu

// A fixed launch sequence captured once as a CUDA graph and replayed (dso_b200_ba_step, dso_b200_ba_iteration): the
// kernels are 3-15 us each, so launch gaps would otherwise be ~40 % of the step.  Invalidated by every call that
// changes a kernel argument (window, points, residuals, affine modes); `key` carries the by-value scalars.
struct BAGraphKey       // every by-value argument baked into the captured launches, compared field by field
{
	double lambda; float stepfacD; int solverMode, nLin, haveP, dbg;
	bool operator==(const BAGraphKey& o) const
	{
		return memcmp(&lambda, &o.lambda, sizeof(double)) == 0 && memcmp(&stepfacD, &o.stepfacD, sizeof(float)) == 0 &&
		       solverMode == o.solverMode && nLin == o.nLin && haveP == o.haveP && dbg == o.dbg;
	}
};
struct BAGraph
{
	cudaGraphExec_t exec; bool valid; BAGraphKey key; long long nlaunch;
};

struct BAWorkspace
{
	BAParams P;
	bool window_set, glue_set, points_set, res_set, lin_done, top_done;
	float* d_precalc; double* d_adHost; double* d_adTarget;
	// points
	int capP;
	float4* d_pt; float* d_color; float* d_weights; int* d_phost; float* d_priorF; float* d_deltaF; int* d_res_begin;
	float* d_Hdd; float* d_bd; float* d_Hcd; float* d_HdiF; float* d_bdSumF; float* d_idh;
	// residuals
	int capR;
	int* d_rhost; int* d_rpoint; int* d_rtarget; int* d_state; int* d_newState; double* d_energy; float* d_newEnergy; float* d_newEnergyWO;
	unsigned char* d_isLin; unsigned char* d_isAct; unsigned char* d_efsel;
	float* d_cpt; float* d_proj; float* d_J[2]; float* d_JpJdF; float* d_contrib; double* d_resEnergy;
	int* d_perm; int* d_subset;      // d_subset: index list of dso_b200_ba_linearize_residuals
	// chunks (top): residuals sorted by pair; (schur): points by host
	int nChunks, capChunks; int* d_chunk_start; int* d_chunk_cnt; int* d_pair_chunk_begin;   // [nF*nF+1]
	float* d_topPartials; int* d_chunk_nres;
	int nPC, capPC; int* d_pc_start; int* d_pc_cnt; int* d_pc_host; int* d_host_pc_begin;     // [nF+1]
	float* d_scPartials; int scN;
	double* d_scContrib;       // stage-1 contributions
	double* d_scRed; double* d_topRed; double* d_topContrib;
	// outputs
	double* d_outblk;          // one allocation holding the five buffers below
	double* d_HA; double* d_bA; double* d_HS; double* d_bS; double* d_energy_blocks; double* d_scalars;   // [0] energy sum, [1] resInA
	double* h_out;             // pinned: HA, bA, HS, bS, scalars
	std::vector<int> h_rpoint, h_res_begin, h_phost;
	// linearised residuals (accumulateLF_MT) and priors
	float* d_adHTdeltaF; float* d_cDeltaF; float h_cDeltaF[4];                   // EnergyFunctional::setDeltaF
	float* d_resToZero;                                                          // EFResidual::res_toZeroF [nR*8]
	float* d_Hdd_L; float* d_bd_L; float* d_Hcd_L; float* d_HcdSum;
	float* d_HdiF_m; float* d_bdSumF_m;                                          // scratch of the marginalisation pass (per point)              // per point: *_accLF, Hcd_accAF + Hcd_accLF
	double* d_HL; double* d_bL;                                                  // d_scalars[3] = resInL
	double* d_priorH; double* d_priorD; double* d_delta;                         // [8nF+4]: prior, delta_prior, delta (entries 0..3 of the deltas: cDeltaF)
	int nLin;                                                                    // linearised residuals (-1: unknown)
	BASolve* S;
	BAGraph g_step, g_iter[2];
	bool sums_in_schur;                // the next Schur accumulation computes the per-point sums + preparation itself
	bool contrib_from_lin;             // d_contrib holds addPoint<0>'s per-point terms already (written by the fused linearisation)
	cudaStream_t side_stream; cudaEvent_t ev_fork, ev_join; bool side_busy;     // parallel branches of the step (ba_fork / ba_join_side)
	bool use_graphs;                                                             // DSO_B200_NO_GRAPH=1 turns them off
};

void ba_graphs_invalidate(BAWorkspace* W);

// Runs `enqueue` (kernel launches and async copies on ctx->ba_stream only) through graph slot g.
template <class F>
static int ba_run_graphed(dso_b200_ctx* ctx, BAGraph& g, const BAGraphKey& key, F enqueue)
{
	BAWorkspace* W = ctx->ba;
	if (!W->use_graphs) return enqueue();
	if (!g.valid || !(g.key == key))
	{
		if (g.exec) { cudaGraphExecDestroy(g.exec); g.exec = nullptr; }
		g.valid = false;
		// launches issued while capturing are counted per thread (another host thread may be launching tracker work)
		t_capture_launches = 0; t_capturing = true;
		cudaError_t eb = cudaStreamBeginCapture(ctx->ba_stream, cudaStreamCaptureModeThreadLocal);
		if (eb != cudaSuccess) { t_capturing = false; CK(eb); }
		const int rc = enqueue();
		cudaGraph_t graph = nullptr;
		const cudaError_t e = cudaStreamEndCapture(ctx->ba_stream, &graph);
		t_capturing = false;
		g.nlaunch = t_capture_launches;
		if (rc) { if (graph) cudaGraphDestroy(graph); return rc; }
		CK(e);
		CK(cudaGraphInstantiate(&g.exec, graph, 0));
		CK(cudaGraphDestroy(graph));
		g.valid = true; g.key = key;
	}
	CK(cudaGraphLaunch(g.exec, ctx->ba_stream));
	LAUNCHED(ctx, g.nlaunch);
	return 0;
}

int ba_check_ready(dso_b200_ctx* ctx, const char* fn);
int ba_enqueue_linearize(dso_b200_ctx* ctx, bool fork = false, bool apply = false);
int ba_join_side(dso_b200_ctx* ctx);
int ba_enqueue_apply(dso_b200_ctx* ctx);
int ba_enqueue_top(dso_b200_ctx* ctx, int mode, const unsigned char* d_pmask = nullptr, bool fork = false, bool prep = false);
int ba_enqueue_schur(dso_b200_ctx* ctx, const unsigned char* d_pmask = nullptr, bool prep_done = false);
void ba_solve_destroy(BAWorkspace* W);

#define BA_CHECK(cond, msg) do { if (!(cond)) { dso_set_error("%s: %s", __func__, msg); return DSO_B200_E_STATE; } } while (0)
