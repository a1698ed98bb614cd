/* TEST INFRASTRUCTURE ONLY.
 * C interface of the CPU oracle: a restatement of the reference's (akawincent/DSO-noted-in-chinese)
 * CPU/SSE arithmetic for the photometric Gauss-Newton hot path. Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may load this library.
 *
 * PARITY UNPINNED: the reference ships no tests, golden vectors or fixtures for this path, and its
 * Eigen/Sophus/Boost dependencies are absent from this image, so the reference itself cannot be run
 * here ("Oracle"). Each function cites the reference lines it follows.
 *
 * All matrices row-major. Poses are 12 doubles: R (3x3 row-major) then t.
 */
#ifndef DSO_ORACLE_H
#define DSO_ORACLE_H
#ifdef __cplusplus
extern "C" {
#endif

#define DSO_ORACLE_MAX_LEVELS 6

/* ---- a1: FrameHessian::makeImages, HessianBlocks.cpp:137-211 -------------------------------- */
/* dIp[l]: 3*w_l*h_l floats (AoS I,dx,dy), absgrad[l]: w_l*h_l floats. Only the elements the
 * reference writes are written (rows 0 and h-1 of dx,dy,abs are left untouched). B may be NULL. */
void dso_oracle_make_images(const float* color, int w, int h, int levels, const float* B,
                            float* const* dIp, float* const* absgrad);

/* setting_affineOptModeA/B (util/settings.cpp:119-120): run-time globals of the reference; < 0 fixes the parameter
 * (CoarseTracker.cpp:621-643, :743-744; Residuals.cpp:257-258).  Process-wide, like the reference's. */
void dso_oracle_set_affine_modes(float modeA, float modeB);

/* ---- tracker (a2..a6) ------------------------------------------------------------------------ */
typedef struct dso_oracle_tracker dso_oracle_tracker;
/* CoarseTracker ctor + makeK, CoarseTracker.cpp:59-131 */
dso_oracle_tracker* dso_oracle_tracker_create(int w, int h, int levels, float fx, float fy, float cx, float cy);
void dso_oracle_tracker_destroy(dso_oracle_tracker*);
void dso_oracle_tracker_get_K(const dso_oracle_tracker*, int lvl, float* fx, float* fy, float* cx, float* cy, float* Ki9);
/* setCoarseTrackingRef + makeCoarseDepthL0, CoarseTracker.cpp:135-292, 523-537.
 * One entry per point with an IN residual towards the reference frame: centerProjectedTo (u,v,idepth)
 * and the point's HdiF. ref_dIp = the reference frame's pyramid (AoS). */
void dso_oracle_tracker_set_ref(dso_oracle_tracker*, int npts, const float* cpt_u, const float* cpt_v,
                                const float* cpt_idepth, const float* HdiF, const float* const* ref_dIp,
                                double ref_aff_a, double ref_aff_b, float ref_ab_exposure);
/* direct injection of the per-level clouds (bypasses makeCoarseDepthL0) */
void dso_oracle_tracker_set_cloud(dso_oracle_tracker*, int lvl, int n, const float* u, const float* v,
                                  const float* idepth, const float* color);
void dso_oracle_tracker_set_ref_aff(dso_oracle_tracker*, double ref_aff_a, double ref_aff_b, float ref_ab_exposure);
int dso_oracle_tracker_get_cloud(const dso_oracle_tracker*, int lvl, float* u, float* v, float* idepth, float* color);
void dso_oracle_tracker_set_new_frame(dso_oracle_tracker*, const float* const* dIp, float ab_exposure);
/* calcRes, CoarseTracker.cpp:359-520 */
void dso_oracle_calc_res(dso_oracle_tracker*, int lvl, const double* refToNew12, double aff_a, double aff_b,
                         float cutoffTH, double* out6);
/* calcGSSSE, CoarseTracker.cpp:296-354 (+ Accumulator9, MatrixAccumulators.h:958-1322) */
void dso_oracle_calc_gs(dso_oracle_tracker*, int lvl, double aff_a, double aff_b, double* H64, double* b8);
/* warped buffers of the last calcRes: idepth,u,v,dx,dy,residual,weight,refColor; returns buf_warped_n */
int dso_oracle_get_warped(const dso_oracle_tracker*, float* const* bufs8);
/* per-point status of the last calcRes: 0 = skipped (OOB / non-finite), 1 = saturated, 2 = in warped buffer */
void dso_oracle_get_point_status(const dso_oracle_tracker*, unsigned char* status);

typedef struct
{
	int ok;                   /* return value of trackNewestCoarse */
	int aborted_level;        /* level at which :713 returned false, else -1 */
	double pose[12];          /* lastToNew_out (unchanged from input when aborted) */
	double aff_a, aff_b;
	double lastResiduals[5];  /* NaN where not reached */
	double lastFlowIndicators[3];
	int n_passes;             /* level passes executed (<= levels+1) */
	int pass_level[8];
	double pass_rmse[8];
	int n_iterations;         /* total LM iterations */
	int n_evals;              /* total calcRes calls */
	int trace_len;            /* accept(1)/reject(0) decisions, at most 256 kept */
	unsigned char trace[256];
	double point_evals;       /* sum over calcRes calls of pc_n[lvl] (work measure) */
	/* margin of the first 64 accept tests (:669), E'/n' < E/n cross-multiplied: trace_dE = E'n - En', trace_E = En';
	 * |trace_dE / trace_E| says how close a decision was (a flip is only legitimate when it is ~ float rounding) */
	float trace_dE[64], trace_E[64];
} dso_oracle_track_result;
/* trackNewestCoarse, CoarseTracker.cpp:551-747. settings: huber 9, cutoff 20, affine modes 1e12/1e8 (settings.cpp) */
void dso_oracle_track_newest_coarse(dso_oracle_tracker*, const double* pose12_in, double aff_a_in, double aff_b_in,
                                    int coarsestLvl, const double* minResForAbort5, dso_oracle_track_result* out);

/* ---- windowed BA (a7..a15) -------------------------------------------------------------------- */
typedef struct dso_oracle_ba dso_oracle_ba;
dso_oracle_ba* dso_oracle_ba_create(int w, int h, int nF, float fx, float fy, float cx, float cy);
void dso_oracle_ba_destroy(dso_oracle_ba*);
/* per frame: evalPT pose (12), state[10], state_zero[10] (unscaled, HessianBlocks.h:150-170), exposure,
 * frameEnergyTH, level-0 dI (AoS 3*w*h, kept by pointer). value_minus_value_zero -> cDeltaF. */
void dso_oracle_ba_set_frame(dso_oracle_ba*, int idx, const double* evalPT12, const double* state10,
                             const double* state_zero10, float ab_exposure, float frameEnergyTH, const float* dI);
void dso_oracle_ba_set_calib_delta(dso_oracle_ba*, const double* value_minus_value_zero4);
/* FrameFramePrecalc::set for all pairs (HessianBlocks.cpp:213-250), setAdjointsF + setDeltaF
 * (EnergyFunctional.cpp:47-114, 178-205) */
void dso_oracle_ba_prepare(dso_oracle_ba*);
/* flat copies of the prepared glue, in the layout the CUDA library's C ABI takes:
 * precalc: nF*nF records of 28 floats [RTll_0(9) tTll_0(3) KRKiTll(9) KtTll(3) aff_mode(2) b0(1) pad(1)], index h*nF+t
 * adHost/adTarget: nF*nF*64 doubles, index h+t*nF (reference indexing); adHTdeltaF: nF*nF*8 floats, same indexing */
void dso_oracle_ba_get_glue(const dso_oracle_ba*, float* precalc, double* adHost, double* adTarget,
                            float* adHTdeltaF, float* cDeltaF4);
/* points: u,v,idepth_scaled,idepth_zero_scaled,color[8],weights[8],host,priorF,deltaF */
void dso_oracle_ba_set_points(dso_oracle_ba*, int nP, const float* u, const float* v, const float* idepth_scaled,
                              const float* idepth_zero_scaled, const float* color8, const float* weights8,
                              const int* host, const float* priorF, const float* deltaF);
/* residuals, in per-point residualsAll order: point index, target frame, state_state (0 IN,1 OOB,2 OUTLIER),
 * state_energy, isLinearized, isActive (EFResidual flags before this step) */
void dso_oracle_ba_set_residuals(dso_oracle_ba*, int nR, const int* point, const int* target, const int* state_state,
                                 const double* state_energy, const unsigned char* isLinearized,
                                 const unsigned char* isActive);
/* PointFrameResidual::linearize over all residuals (Residuals.cpp:78-291); returns the energy sum
 * (linearizeAll_Reductor stats[0], FullSystemOptimize.cpp:52-60). Outputs may be NULL.
 * J: nR*74 floats [resF8 Jpdxi0(6) Jpdxi1(6) Jpdc0(4) Jpdc1(4) Jpdd(2) JIdx0(8) JIdx1(8) JabF0(8) JabF1(8) JIdx2(4) JabJIdx(4) Jab2(4)] */
double dso_oracle_ba_linearize(dso_oracle_ba*, int* newState, float* newEnergy, float* newEnergyWithOutlier,
                               float* centerProjectedTo3, float* projectedTo16, float* J74, double* energy_ret);
/* applyRes(true) for every residual (Residuals.cpp:323-347) + EFResidual::takeDataF
 * (EnergyFunctionalStructs.cpp:39-49). Outputs: isActive flags, JpJdF (nR*8). */
void dso_oracle_ba_apply_res(dso_oracle_ba*, unsigned char* isActive_out, float* JpJdF_out);
/* accumulateAF_MT (EnergyFunctional.cpp:209-232; MT selects the thread-pool branch with nthreads workers,
 * chunks of 50 points) = AccumulatedTopHessianSSE::addPoint<0> + stitchDoubleMT.
 * H: (8nF+4)^2 doubles, b: 8nF+4. per-point: Hdd_accAF, bd_accAF, Hcd_accAF[4]. */
void dso_oracle_ba_accumulate_A(dso_oracle_ba*, int MT, int nthreads, double* H, double* b, int* resInA,
                                float* Hdd, float* bd, float* Hcd4);
/* accumulateSCF_MT (EnergyFunctional.cpp:257-274) = AccumulatedSCHessianSSE::addPoint + stitchDoubleMT */
void dso_oracle_ba_accumulate_SC(dso_oracle_ba*, int MT, int nthreads, double* H, double* b,
                                 float* HdiF, float* bdSumF, float* idepth_hessian);
/* same, multi-threaded linearizeAll (static split ceil(n/nthreads), FullSystemOptimize.cpp:164) for timing */
double dso_oracle_ba_linearize_mt(dso_oracle_ba*, int nthreads);

/* ---- N1/N2 rows: solveSystemF tail, resubstituteF_MT, linearised residuals --------------------- */
/* accumulateLF_MT (EnergyFunctional.cpp:233-254): addPoint<1> over linearised active residuals, priors added in the
 * stitch (AccumulatedTopHessian.cpp:314-324). per-point outputs are Hdd_accLF, bd_accLF, Hcd_accLF. */
void dso_oracle_ba_accumulate_L(dso_oracle_ba*, int MT, int nthreads, double* H, double* b, int* resInL,
                                float* Hdd, float* bd, float* Hcd4);
/* EnergyFunctional::cPrior (4), frames[i]->prior / delta_prior / delta (8 per frame, EnergyFunctionalStructs.cpp:55-63),
 * HM (N*N row-major), bM (N), N = 8nF+4.  NULL = zeros. */
void dso_oracle_ba_set_priors(dso_oracle_ba*, const double* cPrior4, const double* framePrior, const double* frameDeltaPrior,
                              const double* frameDelta, const double* HM, const double* bM);
/* lastNullspaces_pose followed by lastNullspaces_scale: n vectors of N doubles (EnergyFunctional.cpp:745-747) */
void dso_oracle_ba_set_nullspaces(dso_oracle_ba*, const double* ns, int n);
/* EFResidual::fixLinearizationF (EnergyFunctionalStructs.cpp:89-115) on the active residuals flagged in `which`
 * (NULL = all active); res_toZeroF_out: nR*8 (may be NULL) */
void dso_oracle_ba_fix_linearization(dso_oracle_ba*, const unsigned char* which, float* res_toZeroF_out);
/* EnergyFunctional::solveSystemF (EnergyFunctional.cpp:802-924) including its three accumulate calls and
 * resubstituteF_MT (:276-330).  solverMode: setting_solverMode bits (settings.h:35-46; default 128|2048); the SVD and
 * ORTHOGONALIZE_SYSTEM branches are not restated (returns -1).  x_out: N (= lastX), pointStep_out: nP (PointHessian::step),
 * HFinal_out/bFinal_out: the damped Schur system that is solved (debug; may be NULL). */
/* the solve step alone (EnergyFunctional.cpp:905-908) on a given N x N system: diagonal scaling + Eigen-style LDL^T */
void dso_oracle_scaled_ldlt_solve(const double* HFinal, const double* bFinal, int N, double* x);
int dso_oracle_ba_solve(dso_oracle_ba*, int iteration, double lambda, int solverMode, double solverModeDelta, int MT, int nthreads,
                        double* x_out, float* pointStep_out, double* HFinal_out, double* bFinal_out);

/* EnergyFunctional::calcLEnergyF_MT (EnergyFunctional.cpp:411-428 + calcLEnergyPt :346-408, Accumulator11 with its 1k / 1M
 * tiers) and calcMEnergyF (:334-343).  nthreads <= 1: one accumulator over all points; otherwise one per chunk of 50. */
double dso_oracle_ba_calc_L_energy(dso_oracle_ba*, int nthreads);
double dso_oracle_ba_calc_M_energy(dso_oracle_ba*);
/* EnergyFunctional::marginalizePointsF (EnergyFunctional.cpp:642-703) for the flagged points, without the structural part
 * (removePoint / makeIDX): priorF *= priorMargFac, addPoint<2> + AccumulatedSCHessianSSE::addPoint(p, false), stitch,
 * HM += margWeightFac * (M - Msc), bM likewise.  H_out / b_out: M - Msc, Mb - Mbsc.  The residuals of flagged points must
 * have been linearised (dso_oracle_ba_fix_linearization).  The reference-build variant is destructive (call it last). */
void dso_oracle_ba_marginalize_points(dso_oracle_ba*, const unsigned char* flags, float priorMargFac, double margWeightFac,
                                      double* H_out, double* b_out, int* resInM, float* idepth_hessian_out);
void dso_oracle_ba_get_marg_prior(dso_oracle_ba*, double* HM, double* bM);

/* ---- N3: immature points ------------------------------------------------------------------------ */
/* ImmaturePoint::ImmaturePoint (FullSystem/ImmaturePoint.cpp:25-57): pattern colours, gradient weights, gradH, energyTH of
 * n points at integer pixels (u, v) of the host frame's level-0 AoS image. */
void dso_oracle_immature_init(int w, int h, const float* host_dI, int n, const int* u, const int* v,
                              float* color8, float* weights8, float* gradH4, float* energyTH);
/* ImmaturePoint::traceOn (FullSystem/ImmaturePoint.cpp:77-438) for n points of one host frame on one new frame (level-0
 * AoS image): KRKi, Kt, affine = hostToFrame_* of FullSystem::traceNewCoarse (FullSystem.cpp:497-508).
 * In/out: idepth_min, idepth_max, quality, status (ImmaturePointStatus), lastTraceUV (2n), lastTracePixelInterval. */
void dso_oracle_trace_on(int w, int h, const float* frame_dI, int n, const float* u, const float* v,
                         const float* color8, const float* weights8, const float* gradH4, const float* energyTH,
                         float* idepth_min, float* idepth_max, float* quality, int* status,
                         const float* KRKi9, const float* Kt3, const float* aff2,
                         float* lastTraceUV2, float* lastTracePixelInterval);

/* FullSystem::optimizeImmaturePoint (FullSystemOptPoint.cpp:51-209) for n immature points of one host frame against nT
 * target frames (level-0 AoS images; per target PRE_RTll (9), PRE_tTll (3), PRE_aff_mode (2) of the host's FrameFramePrecalc).
 * result: 0 = null, -1 = the "delete me" marker, 1 = activated with idepth_out; res_state_out[k*nT+t]: final ResState of the
 * temporary residual to target t (meaningful when result != 0). */
void dso_oracle_optimize_immature(int w, int h, float fx, float fy, float cx, float cy, int nT, const float* const* target_dI,
                                  const float* RTll9, const float* tTll3, const float* aff2,
                                  int n, const float* u, const float* v, const float* color8, const float* weights8,
                                  const float* energyTH, const float* idepth_min, const float* idepth_max, int minObs,
                                  int* result, float* idepth_out, int* res_state_out);

#ifdef __cplusplus
}
#endif
#endif
