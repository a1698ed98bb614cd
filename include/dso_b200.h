/* dso_b200 -- C ABI of the B200 (sm_100a) implementation of DSO's photometric Gauss-Newton hot path.
 *
 * The reference (akawincent/DSO-noted-in-chinese) has no FFI layer: the drop-in boundary is the C++
 * member-function surface listed below.  Every entry point here replaces one of those members and
 * cites it (paths relative to the reference tree).  The C++ classes in
 * dso-noted-in-chinese_b200/host/ keep the reference signatures and forward to these calls.
 *
 * Conventions
 *   - every function returns 0 on success, <0 on error (DSO_B200_E_*); dso_b200_last_error() gives the
 *     message for the calling thread.  There is no CPU fallback: without a CUDA device every compute
 *     call fails with DSO_B200_E_CUDA.
 *   - matrices are row-major; poses are 12 doubles: R (3x3 row-major) followed by t.
 *   - "host" pointers are ordinary (ideally pinned, see dso_b200_host_alloc) CPU memory; "dev"
 *     pointers are CUDA device memory owned by the caller.
 *   - streams and threads.  The reference runs tracking (FullSystem::addActiveFrame, under trackMutex, FullSystem.cpp:838)
 *     and mapping (FullSystem::makeKeyFrame, under mapMutex, :1098) on two threads at once and swaps its two CoarseTracker
 *     objects between them (:898-902, :1180-1182).  A context therefore has one stream per tracker slot, one for the
 *     mapping side (windowed BA, immature points) and one for makeImages; work of different groups never queues behind
 *     each other, and device-side events order a frame slot's build before its readers.  One host thread per group may call
 *     into the same context concurrently (tracker slot 0 / tracker slot 1 / mapping); calls within a group are not
 *     re-entrant.  Calls that return data to host memory synchronise their own stream, *_async/_dev variants do not.
 */
#ifndef DSO_B200_H
#define DSO_B200_H
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif

#define DSO_B200_MAX_LEVELS 6      /* PYR_LEVELS, util/settings.h:50 */
#define DSO_B200_MAX_FRAMES 8      /* window size at optimise time, FullSystem.cpp:1102-1152 */
#define DSO_B200_MAX_HYP 32        /* FullSystem.cpp:311-354 builds 31 pose priors */
#define DSO_B200_MAX_PASSES 8      /* level passes of one trackNewestCoarse (levels + one repeat) */
#define DSO_B200_TRACE_MARGINS 64  /* accept tests whose margin is kept in dso_b200_track_result */

#define DSO_B200_E_ARG (-1)
#define DSO_B200_E_CUDA (-2)
#define DSO_B200_E_STATE (-3)
#define DSO_B200_E_LIMIT (-4)

typedef struct dso_b200_ctx dso_b200_ctx;

/* ---- context ----------------------------------------------------------------------------------- */
/* One context per FullSystem (FullSystem/FullSystem.h:141): owns the stream, n_frame_slots device
 * pyramids of at most max_w x max_h, two tracker slots (coarseTracker / coarseTracker_forNewKF,
 * FullSystem.h:286-287) and one BA workspace. */
int dso_b200_create(int device, int max_w, int max_h, int n_frame_slots, dso_b200_ctx** out);
void dso_b200_destroy(dso_b200_ctx* ctx);
const char* dso_b200_last_error(void);
/* cudaStream_t of the context's main stream / of one of its groups (so callers can order their own work or record
 * events on it): MAIN makeImages and its _dev variant; TRACKER0/1 everything taking that `trk` index (tracker_set_ref*,
 * calc_res, track_*); MAPPING every dso_b200_ba_*, dso_b200_trace_* and dso_b200_optimize_immature call; UPLOAD
 * dso_b200_make_images_async. */
#define DSO_B200_STREAM_MAIN 0
#define DSO_B200_STREAM_TRACKER0 1
#define DSO_B200_STREAM_TRACKER1 2
#define DSO_B200_STREAM_MAPPING 3
#define DSO_B200_STREAM_UPLOAD 4
void* dso_b200_stream(dso_b200_ctx* ctx);
void* dso_b200_stream_for(dso_b200_ctx* ctx, int which);
int dso_b200_sync(dso_b200_ctx* ctx);
/* pinned host memory for image / result buffers (cudaHostAlloc / cudaFreeHost) */
void* dso_b200_host_alloc(size_t bytes);
void dso_b200_host_free(void* p);
/* number of kernels this library has launched on this context since creation */
long long dso_b200_launch_count(dso_b200_ctx* ctx);

/* The reference's global settings that the kernels of this library read (util/settings.cpp; the line of each definition is
 * given).  The reference reads its globals at call time, so a caller that changes them calls dso_b200_set_settings again;
 * a new context starts with the defaults below.  The tracker's own settings travel in dso_b200_track_opts. */
typedef struct
{
	float huberTH;                      /* setting_huberTH = 9 (:127): calcRes, PointFrameResidual::linearize, traceOn, linearizeResidual */
	float outlierTHSumComponent;        /* setting_outlierTHSumComponent = 50*50 (:91): PointFrameResidual::linearize */
	float maxPixSearch;                 /* setting_maxPixSearch = 0.027 (:156): traceOn */
	int minTraceTestRadius;             /* setting_minTraceTestRadius = 2 (:158) */
	int GNItsOnPointActivation;         /* setting_GNItsOnPointActivation = 3 (:159): optimizeImmaturePoint */
	float trace_stepsize;               /* setting_trace_stepsize = 1.0 (:160) */
	int trace_GNIterations;             /* setting_trace_GNIterations = 3 (:161) */
	float trace_GNThreshold;            /* setting_trace_GNThreshold = 0.1 (:162) */
	float trace_extraSlackOnTH;         /* setting_trace_extraSlackOnTH = 1.2 (:163) */
	float trace_slackInterval;          /* setting_trace_slackInterval = 1.5 (:164) */
	float trace_minImprovementFactor;   /* setting_trace_minImprovementFactor = 2 (:165) */
	float minIdepthH_act;               /* setting_minIdepthH_act = 100 (:67): optimizeImmaturePoint */
} dso_b200_settings;
void dso_b200_settings_default(dso_b200_settings* s);
int dso_b200_set_settings(dso_b200_ctx* ctx, const dso_b200_settings* s);

/* ---- K1: FrameHessian::makeImages (FullSystem/HessianBlocks.cpp:137-211, decl HessianBlocks.h:282) */
/* color: w*h irradiance floats (ImageAndExposure::image).  B_lut: 256 floats (CalibHessian::B,
 * HessianBlocks.h:399) or NULL (= the "HCalib==0 / setting_gammaWeightsPixelSelect!=1" branch).
 * dIp_out[l]: 3*w_l*h_l floats in the reference's Eigen::Vector3f AoS layout, absgrad_out[l]: w_l*h_l
 * floats; either array (or single entries) may be NULL to skip the device->host copy.  The device
 * pyramid stays resident in frame_slot for the tracker and BA.  Rows 0 and h_l-1 of dx, dy and
 * absSquaredGrad, which the reference leaves uninitialised, are written as 0. */
int dso_b200_make_images(dso_b200_ctx* ctx, int frame_slot, const float* color_host, int w, int h, int levels,
                         const float* B_lut, float* const* dIp_out, float* const* absgrad_out);
/* Asynchronous variant for pipelined front ends (the image of frame k+1 is usually in host memory while frame k is still
 * being tracked): enqueues the H2D copy and the pyramid kernels of FrameHessian::makeImages (HessianBlocks.cpp:137-211,
 * device pyramid only, no gamma LUT) on the context's second stream and returns.  The next tracker / BA call that reads
 * `frame_slot` waits for it on the device.  `color_host` must stay valid until then and should be pinned
 * (dso_b200_host_alloc); the slot must not be one a running tracking call still reads.  No counterpart in the reference,
 * whose makeImages is synchronous. */
int dso_b200_make_images_async(dso_b200_ctx* ctx, int frame_slot, const float* color_host, int w, int h, int levels);
/* same, input already in device memory, nothing copied back, no synchronisation */
int dso_b200_make_images_dev(dso_b200_ctx* ctx, int frame_slot, const float* color_dev, int w, int h, int levels,
                             const float* B_lut_host);

/* ---- K2: CoarseTracker (FullSystem/CoarseTracker.{h,cpp}) ---------------------------------------- */
/* CoarseTracker::makeK (CoarseTracker.cpp:102-131) + the result of setCoarseTrackingRef /
 * makeCoarseDepthL0 (:135-292, :523-537): per-level clouds pc_u/pc_v/pc_idepth/pc_color with pc_n,
 * level-0 intrinsics, and the reference frame's affine parameters and exposure. trk is 0 or 1. */
int dso_b200_tracker_set_ref(dso_b200_ctx* ctx, int trk, int w, int h, int levels,
                             float fx, float fy, float cx, float cy,
                             const int* pc_n, const float* const* pc_u, const float* const* pc_v,
                             const float* const* pc_idepth, const float* const* pc_color,
                             double ref_aff_a, double ref_aff_b, float ref_ab_exposure);
/* The same with CoarseTracker::makeCoarseDepthL0 (CoarseTracker.cpp:135-292) done on the device (next row N3): the
 * caller passes, for every active point with lastResiduals[0] IN (:147), the projection into the newest key frame
 * r->centerProjectedTo[0..2] (cpt_u, cpt_v, cpt_idepth) and ph->efPoint->HdiF; the key frame's pyramid must already be in
 * `ref_frame_slot` (its level-l intensities are the cloud's colours, :270).  Splat, 2x2 sum pyramid, dilation,
 * normalisation and raster-order compaction run on the context stream; pc_n_out (may be NULL) receives pc_n[lvl].
 * Points projecting outside the image are skipped (the reference does not check).  Results are bit-identical to the
 * reference's sequential loops, also where several points fall on one pixel. */
int dso_b200_tracker_set_ref_points(dso_b200_ctx* ctx, int trk, int ref_frame_slot, int levels,
                                    float fx, float fy, float cx, float cy, int n_points,
                                    const float* cpt_u, const float* cpt_v, const float* cpt_idepth, const float* HdiF,
                                    double ref_aff_a, double ref_aff_b, float ref_ab_exposure, int* pc_n_out);
/* Reads back a tracker slot's cloud of one level as (u, v, idepth, color) records (tests, debugPlot). */
int dso_b200_tracker_get_cloud(dso_b200_ctx* ctx, int trk, int lvl, int max_points, float* out4, int* n_out);


/* CoarseTracker::calcRes (CoarseTracker.cpp:359-520, decl CoarseTracker.h:100), fused with the
 * accumulation of calcGSSSE so that no warped buffers are materialised.
 * out6 = [E, numTermsInE, flowT, 0, flowRT, saturatedRatio] as the reference's Vec6. */
int dso_b200_calc_res(dso_b200_ctx* ctx, int trk, int frame_slot, int lvl, const double* refToNew12,
                      double aff_a, double aff_b, float new_ab_exposure, float cutoffTH, double* out6);
/* CoarseTracker::calcGSSSE (CoarseTracker.cpp:296-354, decl CoarseTracker.h:101): H (8x8) and b of
 * the points kept by the last dso_b200_calc_res on this tracker slot, with the 1/n and SCALE_*
 * factors applied exactly as :338-353 (n = padded buf_warped_n). */
int dso_b200_calc_gs(dso_b200_ctx* ctx, int trk, double* H64, double* b8);
/* calcRes + calcGSSSE for callers that cannot allocate GPU memory.  (No counterpart in the reference: it
 * exists because of how this library is embedded, not because of anything DSO does.)
 *
 * Who needs it: DSO's tracking thread has a per-frame deadline and shares the GPU with the mapping thread's
 * bundle-adjustment work.  cudaMalloc / cudaFree stall the whole device while they run and can fail when
 * another component (the mapper, a host application's memory pool) already owns the memory, so a tracker
 * embedded in such a process is often forbidden from allocating after start-up -- and creating a
 * dso_b200_ctx (which reserves pyramids, clouds and work buffers) counts as allocating.  This entry point
 * needs no context: the caller hands in its own device buffers, and the only work memory the kernel needs
 * (per-block partial sums and a completion counter) is a fixed array compiled into the library.  It calls
 * no cudaMalloc, cudaFree or other allocating routine.
 *
 * Inputs (device memory of the current CUDA device, 16-byte aligned):
 *   cloud_dev  n_points x 4 floats (u, v, idepth, color), the reference cloud of one pyramid level
 *   frame_dev  w*h x 4 floats (I, dx, dy, unused), that level of the new frame
 *   fx,fy,cx,cy intrinsics of THAT level; is_level0 != 0 enables the flow indicators (CoarseTracker.cpp:416)
 *   pose / affine / exposure / cut-off as for dso_b200_calc_res; stream is a cudaStream_t (NULL = default)
 * Output: sums52_dev, 52 floats in device memory, written asynchronously on `stream`:
 *   [0] E  [1] sumSquaredShiftT  [2] sumSquaredShiftRT  [3] sumSquaredShiftNum  [4] numTermsInE  [5] numSaturated
 *   [6] numTermsInWarped (unpadded)  [8..43] upper triangle (row-major, r<=c) of sum w*J*J^T  [44..51] sum w*J*r
 *   i.e. the reference's Vec6 is [E, [4], [1]/([3]+0.1), 0, [2]/([3]+0.1), [5]/[4]] and H, b follow from
 *   [8..51] times 1/n (n = [6] rounded up to a multiple of 4) and the SCALE_* factors of CoarseTracker.cpp:338-353.
 *
 * Limits and rules:
 *   - n_points <= DSO_B200_NOALLOC_MAX_POINTS; larger clouds are rejected with DSO_B200_E_LIMIT.
 *   - the fixed array (about 53 KB) is held by every GPU that loads the library until the process exits.
 *   - all calls share that array and it keeps the previous call's contents (this function clears the part
 *     that must start at zero); two calls that overlap on one GPU -- different streams or threads -- race.
 *   - the first call on each device looks up the array's device address; make that first call before
 *     starting any stream capture. */
#define DSO_B200_NOALLOC_MAX_POINTS 65536
int dso_b200_calc_res_noalloc(const float* cloud_dev, int n_points, const float* frame_dev, int w, int h,
                              int is_level0, float fx, float fy, float cx, float cy,
                              const double* refToNew12, double ref_aff_a, double ref_aff_b, float ref_ab_exposure,
                              double aff_a, double aff_b, float new_ab_exposure, float cutoffTH,
                              void* stream, float* sums52_dev);
/* Debug / parity: per reference point of level lvl of the last calc_res, in cloud order:
 * status (0 skipped, 1 saturated, 2 kept) and the 8 values the reference appends to buf_warped_*
 * (idepth,u,v,dx,dy,residual,weight,refColor; CoarseTracker.cpp:476-483).  Re-runs the evaluation. */
int dso_b200_calc_res_debug(dso_b200_ctx* ctx, int trk, int frame_slot, int lvl, const double* refToNew12,
                            double aff_a, double aff_b, float new_ab_exposure, float cutoffTH,
                            unsigned char* status_out, float* warped8_out);

typedef struct
{
	float huberTH;            /* setting_huberTH = 9 (util/settings.cpp:127) */
	float coarseCutoffTH;     /* setting_coarseCutoffTH = 20 (:138) */
	float affineOptModeA;     /* setting_affineOptModeA = 1e12 (:119) */
	float affineOptModeB;     /* setting_affineOptModeB = 1e8 (:120) */
	int maxIterations[5];     /* {10,20,50,50,50}, CoarseTracker.cpp:569 */
} dso_b200_track_opts;
void dso_b200_track_opts_default(dso_b200_track_opts* o);

typedef struct
{
	int ok;                          /* what trackNewestCoarse returns when minResForAbort is all-NaN */
	int reserved;
	double pose[12];                 /* lastToNew_out */
	double aff_a, aff_b;             /* aff_g2l_out */
	double lastResiduals[5];         /* CoarseTracker::lastResiduals (NaN where a level was not run) */
	double lastFlowIndicators[3];    /* CoarseTracker::lastFlowIndicators */
	int n_passes;                    /* level passes executed; the RMSE after each pass lets the caller */
	int pass_level[DSO_B200_MAX_PASSES];   /* replay the abort test of CoarseTracker.cpp:713 for any */
	double pass_rmse[DSO_B200_MAX_PASSES]; /* minResForAbort (FullSystem.cpp:380-442) */
	int n_iterations;
	int n_evals;
	int trace_len;
	unsigned char trace[256];        /* accept(1) / reject(0) per LM iteration */
	double point_evals;
	/* how close the first DSO_B200_TRACE_MARGINS accept tests (CoarseTracker.cpp:669, E'/n' < E/n, cross-multiplied) were:
	 * trace_dE[i] = E'n - En', trace_E[i] = En'.  The float energy sums are taken in a different order than the reference's
	 * sequential loop, so a decision with |trace_dE / trace_E| of the order of float rounding may legitimately differ. */
	float trace_dE[DSO_B200_TRACE_MARGINS], trace_E[DSO_B200_TRACE_MARGINS];
} dso_b200_track_result;

/* CoarseTracker::trackNewestCoarse (CoarseTracker.cpp:551-747, decl CoarseTracker.h:51-55) for nHyp
 * initial poses at once (the priors of FullSystem::trackNewCoarse, FullSystem.cpp:311-354): the whole
 * coarse-to-fine LM loop runs inside one persistent kernel, one thread-block cluster per hypothesis.
 * Every hypothesis is run to completion with no abort; use dso_b200_track_replay_abort for the
 * reference's sequential early-out semantics. */
int dso_b200_track_batched(dso_b200_ctx* ctx, int trk, int frame_slot, int nHyp, const double* pose12_init,
                           double aff_a_init, double aff_b_init, float new_ab_exposure, int coarsestLvl,
                           const dso_b200_track_opts* opts, dso_b200_track_result* out);
/* Asynchronous form of dso_b200_track_batched: enqueues the work on the tracker slot's stream and returns.
 * dso_b200_track_fetch hands out results in call order: each fetch waits for and returns the oldest call on trk whose
 * results have not been fetched yet.  At most 4 such calls may be in flight; a fifth is refused with DSO_B200_E_STATE.
 * With pose12_init == NULL only the kernel is enqueued: it re-uses the initial poses of the previous call and leaves its
 * results on the device (used to time the kernel alone); a fetch with no un-fetched call pending waits for the slot's
 * stream and returns those device-side results of the last launch. */
int dso_b200_track_batched_async(dso_b200_ctx* ctx, int trk, int frame_slot, int nHyp, const double* pose12_init,
                                 double aff_a_init, double aff_b_init, float new_ab_exposure, int coarsestLvl,
                                 const dso_b200_track_opts* opts);
int dso_b200_track_fetch(dso_b200_ctx* ctx, int trk, int nHyp, dso_b200_track_result* out);
/* The pipelined per-frame call of a live front end (FullSystem::addActiveFrame, FullSystem.cpp:860 + :905, with the next
 * image's makeImages taken off the critical path): dso_b200_track_batched_async on frame_slot (whose pyramid a previous
 * call built), dso_b200_make_images_async of the NEXT image into next_slot (skipped when next_slot < 0 or the pointer is
 * NULL; next_color_host must stay valid until that slot is consumed), then dso_b200_track_fetch.  Same results and errors
 * as the three calls; if the prefetch is refused, its error is returned and the tracking result stays fetchable. */
int dso_b200_track_and_prefetch(dso_b200_ctx* ctx, int trk, int frame_slot, int nHyp, const double* pose12_init,
                                double aff_a_init, double aff_b_init, float new_ab_exposure, int coarsestLvl,
                                const dso_b200_track_opts* opts, int next_slot, const float* next_color_host,
                                int w, int h, int levels, dso_b200_track_result* out);
/* One frame end to end -- what FullSystem::addActiveFrame does per frame (FullSystem.cpp:860 makeImages,
 * :905 trackNewCoarse -> :387 trackNewestCoarse): upload of the image from host memory, pyramid into
 * frame_slot, nHyp hypotheses, download of the results; one synchronisation at the end. */
int dso_b200_track_frame(dso_b200_ctx* ctx, int trk, int frame_slot, const float* color_host, int w, int h,
                         int levels, int nHyp, const double* pose12_init, double aff_a_init, double aff_b_init,
                         float new_ab_exposure, int coarsestLvl, const dso_b200_track_opts* opts,
                         dso_b200_track_result* out);
/* Profiling aid (no counterpart in the reference): with enable != 0 the persistent kernel records, for
 * thread 0 of CTA 0 of the last launch, the SM cycles spent in [point evaluation, CTA reduction, cluster
 * exchange + barrier, LM step], the number of evaluations, and the whole kernel in SM cycles and in ns;
 * cycles7 (120 values, may be NULL; [24+4e..27+4e] the four phases of evaluation e < 24; [7] prologue, [14] cloud staging, [15] epilogue cycles, [16..20] their split) receives them; [8..11] split the LM step into sums+control, solve,
 * exponential+pose update, parameter build. */
int dso_b200_track_phase_cycles(dso_b200_ctx* ctx, int enable, long long* cycles7);
/* Launch shape of the persistent tracker (no counterpart in the reference; tests and tuning): threads per CTA (128, 256),
 * CTAs per cluster (power of two <= 16) and, for single-hypothesis calls, the number of clusters the cloud is
 * spread over (1 = never split; by default clouds of >= 32768 points per level are split).  0 = automatic.  Results
 * change only through the order of the float sums. */
int dso_b200_track_set_tuning(dso_b200_ctx* ctx, int threads, int cluster_size, int n_clusters);
/* shape3 = { threads, cluster size, clusters per hypothesis } of the last persistent-tracker launch on this context.  With
 * more than one cluster per hypothesis the launch is cooperative (all clusters co-resident, guaranteed by the driver); if the
 * device cannot host them at once -- e.g. another context occupies SMs -- the call falls back to one cluster. */
int dso_b200_track_last_shape(dso_b200_ctx* ctx, int* shape3);
/* Profiling aid (no counterpart in the reference): runs the tracker's LM step (decision, damped 8x8 solve, SE3 update,
 * next EvalParams) `iters` times in one warp on a synthetic accepted evaluation; cycles7 = { total SM cycles, then the split
 * level, solve, exp+pose, params, sums->H, control }. */
int dso_b200_debug_lm_bench(dso_b200_ctx* ctx, int iters, long long* cycles7);
/* Applies CoarseTracker.cpp:713 to a finished result: returns 1 if trackNewestCoarse would have
 * returned true for this minResForAbort, 0 if it would have aborted (lastResiduals5_out then holds
 * what CoarseTracker::lastResiduals would contain at the abort). */
int dso_b200_track_replay_abort(const dso_b200_track_result* r, const double* minResForAbort5,
                                double* lastResiduals5_out);
/* Same LM loop driven from the host with one dso_b200_calc_res/calc_gs pair per evaluation
 * (validation path for the private-method drop-ins; one hypothesis, honours minResForAbort). */
int dso_b200_track_hostloop(dso_b200_ctx* ctx, int trk, int frame_slot, const double* pose12_init,
                            double aff_a_init, double aff_b_init, float new_ab_exposure, int coarsestLvl,
                            const double* minResForAbort5, const dso_b200_track_opts* opts,
                            dso_b200_track_result* out);

/* ---- K3-K5: windowed BA --------------------------------------------------------------------------- */
/* FrameFramePrecalc (HessianBlocks.h:73-103) fields read by PointFrameResidual::linearize. */
typedef struct
{
	float PRE_RTll_0[9];
	float PRE_tTll_0[3];
	float PRE_KRKiTll[9];
	float PRE_KtTll[3];
	float PRE_aff_mode[2];
	float PRE_b0_mode;
	float pad;
} dso_b200_precalc;

/* Window description: nF frames whose level-0 images live in frame_slots[i] (target->dI), calibration
 * CalibHessian::fxl..cyl (HessianBlocks.h:361-368), per-frame frameEnergyTH (HessianBlocks.h:136). */
int dso_b200_ba_set_window(dso_b200_ctx* ctx, int nF, const int* frame_slots, int w, int h,
                           float fx, float fy, float cx, float cy, const float* frameEnergyTH);
/* setting_affineOptModeA/B (util/settings.cpp:119-120, read at Residuals.cpp:257-258): a negative mode fixes that
 * affine parameter, i.e. the stored JabF row is zero.  Defaults (1e12, 1e8) optimise both.  The tracker takes the
 * same two settings through dso_b200_track_opts. */
int dso_b200_ba_set_affine_modes(dso_b200_ctx* ctx, float modeA, float modeB);
/* Per-iteration glue computed on the host exactly as the reference does (FrameFramePrecalc::set,
 * HessianBlocks.cpp:213-250; EnergyFunctional::setAdjointsF / setDeltaF, EnergyFunctional.cpp:47-114,
 * 178-205).  precalc[h*nF+t]; adHost/adTarget[(h+t*nF)*64] doubles; adHTdeltaF[(h+t*nF)*8]; cDeltaF[4]. */
int dso_b200_ba_set_glue(dso_b200_ctx* ctx, const dso_b200_precalc* precalc, const double* adHost,
                         const double* adTarget, const float* adHTdeltaF, const float* cDeltaF4);
/* The same glue computed inside the library from the frames' states, then uploaded: FrameHessian::setState
 * (HessianBlocks.h:183-199: state -> state_scaled, PRE_worldToCam = exp(w2c_leftEps) * worldToCam_evalPT),
 * FrameFramePrecalc::set for every host/target pair (HessianBlocks.cpp:213-250) and
 * EnergyFunctional::setAdjointsF (EnergyFunctional.cpp:47-114).  Per frame: worldToCam_evalPT (12 doubles),
 * state and state_zero (10 doubles each, unscaled), ab_exposure.  The *_out arrays (may be NULL) return the
 * computed glue in the layout of dso_b200_ba_set_glue. */
int dso_b200_ba_set_frames(dso_b200_ctx* ctx, const double* evalPT12, const double* state10, const double* state_zero10,
                           const float* ab_exposure, dso_b200_precalc* precalc_out, double* adHost_out, double* adTarget_out);
/* PointHessian fields (HessianBlocks.h:431-444) + EFPoint::priorF/deltaF (EnergyFunctionalStructs.h:111-112);
 * points must be grouped by host frame in EnergyFunctional::allPoints order. */
int dso_b200_ba_set_points(dso_b200_ctx* ctx, int nP, const float* u, const float* v, const float* idepth_scaled,
                           const float* idepth_zero_scaled, const float* color8, const float* weights8,
                           const int* host, const float* priorF, const float* deltaF);
/* PointFrameResidual graph (Residuals.h:56-112) in EFPoint::residualsAll order: point, target,
 * state_state (0 IN, 1 OOB, 2 OUTLIER), state_energy, EFResidual::isLinearized / isActive. */
int dso_b200_ba_set_residuals(dso_b200_ctx* ctx, int nR, const int* point, const int* target, const int* state_state,
                              const double* state_energy, const unsigned char* isLinearized,
                              const unsigned char* isActive);
/* PointFrameResidual::linearize (FullSystem/Residuals.cpp:78-291) over all residuals =
 * FullSystem::linearizeAll_Reductor (FullSystemOptimize.cpp:52-60).  Outputs (each may be NULL):
 * state_NewState, state_NewEnergy, state_NewEnergyWithOutlier, centerProjectedTo (nR*3),
 * projectedTo (nR*16), RawResidualJacobian (nR*74 floats: resF8 Jpdxi0(6) Jpdxi1(6) Jpdc0(4) Jpdc1(4)
 * Jpdd(2) JIdx0(8) JIdx1(8) JabF0(8) JabF1(8) JIdx2(4) JabJIdx(4) Jab2(4)), and the energy sum. */
int dso_b200_ba_linearize(dso_b200_ctx* ctx, int* newState, float* newEnergy, float* newEnergyWithOutlier,
                          float* centerProjectedTo3, float* projectedTo16, float* J74, double* energy_sum);
/* The per-object form of the two calls around this comment: PointFrameResidual::linearize (decl FullSystem/Residuals.h:95)
 * and, with apply != 0, PointFrameResidual::applyRes(true) (Residuals.h:110) for the k residuals listed in residual_idx
 * (indices into the array given to dso_b200_ba_set_residuals); with reset_oob != 0 PointFrameResidual::resetOOB
 * (Residuals.h:101-108) comes first.  This is the serial loop of FullSystem::flagPointsForRemoval (FullSystem.cpp:776-787)
 * as one batch; the other residuals are not touched.  Outputs (k entries each, may be NULL): state_NewState,
 * state_NewEnergy, EFResidual::isActive() after applyRes. */
int dso_b200_ba_linearize_residuals(dso_b200_ctx* ctx, int k, const int* residual_idx, int reset_oob, int apply,
                                    int* newState_out, float* newEnergy_out, unsigned char* isActive_out);
/* PointFrameResidual::applyRes(true) for all residuals (Residuals.cpp:323-347) including
 * EFResidual::takeDataF (EnergyFunctionalStructs.cpp:39-49).  Outputs may be NULL. */
int dso_b200_ba_apply_res(dso_b200_ctx* ctx, unsigned char* isActive_out, float* JpJdF8_out);
/* EnergyFunctional::accumulateAF_MT (OptimizationBackend/EnergyFunctional.cpp:209-232):
 * AccumulatedTopHessianSSE::addPoint<0> (AccumulatedTopHessian.cpp:39-177) + stitchDoubleMT
 * (AccumulatedTopHessian.h:91-139).  H: (8nF+4)^2 doubles, b: 8nF+4.  Per-point outputs may be NULL. */
int dso_b200_ba_accumulate_A(dso_b200_ctx* ctx, double* H, double* b, int* resInA,
                             float* Hdd_accAF, float* bd_accAF, float* Hcd_accAF4);
/* EnergyFunctional::accumulateSCF_MT (EnergyFunctional.cpp:257-274): AccumulatedSCHessianSSE::addPoint
 * (AccumulatedSCHessian.cpp:35-85) + stitchDoubleMT (AccumulatedSCHessian.h:93-135). */
int dso_b200_ba_accumulate_SC(dso_b200_ctx* ctx, double* H_sc, double* b_sc,
                              float* HdiF, float* bdSumF, float* idepth_hessian);
/* The per-point results of the last accumulation as they stand on the device (after dso_b200_ba_step, _iteration, _solve or
 * the piecewise calls above): EFPoint::Hdd_accAF, bd_accAF, Hcd_accAF (AccumulatedTopHessian.cpp:162-166), HdiF, bdSumF and
 * PointHessian::idepth_hessian (AccumulatedSCHessian.cpp:50-60), nP (4 nP for Hcd) floats each; any pointer may be NULL. */
int dso_b200_ba_get_point_terms(dso_b200_ctx* ctx, float* Hdd_accAF, float* bd_accAF, float* Hcd_accAF4,
                                float* HdiF, float* bdSumF, float* idepth_hessian);
/* linearize -> applyRes(true) -> accumulateAF -> accumulateSCF back to back with a single
 * synchronisation; H_A/b_A/H_sc/b_sc as above (one GN iteration's worth of FullSystem::optimize,
 * FullSystemOptimize.cpp:455-486 + EnergyFunctional.cpp:816-820). */
int dso_b200_ba_step(dso_b200_ctx* ctx, double* H_A, double* b_A, double* H_sc, double* b_sc,
                     double* energy_sum, int* resInA);

/* ---- next rows: solveSystemF tail, resubstituteF_MT, linearised residuals ---------------
 * The state these calls add lives next to the window's: priors, the marginalisation prior HM/bM, the null-space
 * projector, res_toZeroF, the per-point *_accLF sums, the solution x and the per-point steps. */

/* CalibHessian::value_minus_value_zero -> EnergyFunctional::cDeltaF (EnergyFunctional.cpp:193). Default zeros. */
int dso_b200_ba_set_calib_delta(dso_b200_ctx* ctx, const double* value_minus_value_zero4);
/* EnergyFunctional::cPrior (4), EFFrame::prior / delta_prior / delta (8 per frame; EnergyFunctionalStructs.cpp:55-63),
 * HM (N*N row-major) and bM (N), N = 8nF+4.  NULL: zeros for cPrior/framePrior/HM/bM; for the two delta arrays NULL keeps
 * what dso_b200_ba_set_frames derived from the frame states (delta = state - state_zero, delta_prior = state). */
int dso_b200_ba_set_priors(dso_b200_ctx* ctx, const double* cPrior4, const double* framePrior, const double* frameDeltaPrior,
                           const double* frameDelta, const double* HM, const double* bM);
/* lastNullspaces_pose followed by lastNullspaces_scale (n vectors of N doubles).  Builds the projector of
 * EnergyFunctional::orthogonalize (EnergyFunctional.cpp:741-785; solverModeDelta = setting_solverModeDelta) once, on the
 * host, and keeps it on the device.  n = 0 removes it. */
int dso_b200_ba_set_nullspaces(dso_b200_ctx* ctx, const double* nullspaces, int n, double solverModeDelta);
/* The global state the non-default solver branches read besides the solverMode argument: haveFirstFrame = "a frame with
 * frameID 0 is in the window" (EnergyFunctional.cpp:686-688, :833-835; decides whether SOLVER_ORTHOGONALIZE_SYSTEM /
 * _POINTMARG project at all), margSolverMode = setting_solverMode as dso_b200_ba_marginalize_points sees it, and
 * setting_solverModeDelta (util/settings.cpp:60-61).  Defaults: 1, 128|2048, 1e-5. */
int dso_b200_ba_set_solver_state(dso_b200_ctx* ctx, int haveFirstFrame, int margSolverMode, double solverModeDelta);
/* EFResidual::fixLinearizationF (EnergyFunctionalStructs.cpp:89-115) on the active residuals flagged in `which`
 * (nR bytes, NULL = every active residual): res_toZeroF = resF - J*delta, isLinearized = true.
 * res_toZeroF_out: nR*8 floats or NULL. */
int dso_b200_ba_fix_linearization(dso_b200_ctx* ctx, const unsigned char* which, float* res_toZeroF_out);
/* EnergyFunctional::accumulateLF_MT (EnergyFunctional.cpp:233-254): AccumulatedTopHessianSSE::addPoint<1>
 * (AccumulatedTopHessian.cpp:39-177) over the linearised active residuals + stitchDoubleMT with usePrior = true
 * (AccumulatedTopHessian.cpp:314-324). */
int dso_b200_ba_accumulate_L(dso_b200_ctx* ctx, double* H_L, double* b_L, int* resInL,
                             float* Hdd_accLF, float* bd_accLF, float* Hcd_accLF4);
/* EnergyFunctional::solveSystemF (EnergyFunctional.cpp:802-924) after linearize + applyRes: the three accumulate calls,
 * HFinal_top / bFinal_top, damping, the diagonally scaled LDL^T solve, orthogonalize(&x) when solverMode asks for it
 * (and null spaces were set), and resubstituteF_MT (EnergyFunctional.cpp:276-330).  solverMode = setting_solverMode
 * (util/settings.h:35-46; the reference default is 128|2048).  The non-default branches are there too:
 * SOLVER_ORTHOGONALIZE_SYSTEM (:832-851: HT_act / bT_act projected off the null spaces unless the first frame is in the
 * window, see dso_b200_ba_set_solver_state) and SOLVER_SVD / SOLVER_SVD_CUT7 (:870-900: one-sided Jacobi SVD of the scaled
 * system on the device; agrees with Eigen's JacobiSVD up to rounding of the singular triplets).  x_out: N doubles (= lastX; frame steps are -x segments, HCalib->step = -x.head<4>()),
 * pointStep_out: nP floats (PointHessian::step), both may be NULL (results stay on the device);
 * HFinal_out/bFinal_out: the system that was solved, for inspection (may be NULL). */
int dso_b200_ba_solve(dso_b200_ctx* ctx, int iteration, double lambda, int solverMode, double* x_out, float* pointStep_out,
                      double* HFinal_out, double* bFinal_out);
/* FullSystem::backupState / doStepFromBackup / loadSateBackup, point half (FullSystemOptimize.cpp:225-345):
 * backup keeps idepth; step sets idepth = idepth_zero = idepth_backup + stepfacD*step for every point and returns
 * sums2_out = { sum step^2, sum |idepth_backup| } (FP64 tree sums; the reference adds floats in point order) and,
 * if idepth_out != NULL, the nP new inverse depths; restore puts the backup back. */
int dso_b200_ba_points_backup(dso_b200_ctx* ctx);
int dso_b200_ba_points_step(dso_b200_ctx* ctx, float stepfacD, float* sums2_out, float* idepth_out);
int dso_b200_ba_points_restore(dso_b200_ctx* ctx);
/* linearize -> applyRes(true) -> solveSystemF -> backup + step of the points, one synchronisation: the device side of one
 * iteration of FullSystem::optimize (FullSystemOptimize.cpp:455-520).  The caller applies -x to the frame states and
 * calls dso_b200_ba_set_frames before the next one. */
int dso_b200_ba_iteration(dso_b200_ctx* ctx, int iteration, double lambda, int solverMode, float stepfacD,
                          double* x_out, double* energy_sum, float* step_sums2);
/* EnergyFunctional::marginalizePointsF (EnergyFunctional.cpp:642-703) for the points flagged PS_MARGINALIZE (point_flags, nP
 * bytes), whose active residuals have been linearised (dso_b200_ba_fix_linearization, as FullSystem::flagPointsForRemoval
 * does): priorF *= idepthFixPriorMargFac, AccumulatedTopHessianSSE::addPoint<2> + AccumulatedSCHessianSSE::addPoint(p, false),
 * both stitched, H = M - Msc, b = Mb - Mbsc, and on the device HM += margWeightFac * H, bM += margWeightFac * b.  H_out, b_out
 * (may be NULL) return H and b, resInM the residual count, idepth_hessian_out (nP, may be NULL) PointHessian::idepth_hessian.
 * The structural part (removePoint, makeIDX, connectivityMap) stays with the caller, who re-uploads the window afterwards.
 * SOLVER_ORTHOGONALIZE_POINTMARG (H, b projected off the null spaces when the first frame is gone, :684-692) and
 * SOLVER_ORTHOGONALIZE_FULL (the accumulated HM, bM projected, :698-699) follow the mode given to dso_b200_ba_set_solver_state. */
int dso_b200_ba_marginalize_points(dso_b200_ctx* ctx, const unsigned char* point_flags, float idepthFixPriorMargFac,
                                   double margWeightFac, double* H_out, double* b_out, int* resInM, float* idepth_hessian_out);
/* The marginalisation prior as it stands on the device (after dso_b200_ba_set_priors and any marginalisation). */
int dso_b200_ba_get_marg_prior(dso_b200_ctx* ctx, double* HM_out, double* bM_out);
/* EnergyFunctional::calcLEnergyF_MT (EnergyFunctional.cpp:411-428, calcLEnergyPt :346-408) and calcMEnergyF (:334-343), the
 * two terms FullSystem::optimize adds to the linearisation energy in its accept test.  The point terms are the reference's
 * float values summed in FP64 by a fixed tree (the reference uses float tiers and a dynamic thread split). */
int dso_b200_ba_calc_energies(dso_b200_ctx* ctx, double* L_energy, double* M_energy);

/* ---- next row N3, second half: immature points (FullSystem/ImmaturePoint.{h,cpp}) ----------------------------------
 * The immature points of the window live on the device between frames.  dso_b200_trace_set_points uploads them when the
 * set changes (points created by makeNewTraces, activated or dropped): per point u, v, color[8], weights[8], gradH (2x2,
 * row-major), energyTH (ImmaturePoint.h:60-75, filled by the constructor, ImmaturePoint.cpp:25-57), the index of its host
 * frame among the n_hosts key frames, and the trace state idepth_min, idepth_max, quality, lastTraceStatus
 * (ImmaturePointStatus as int), lastTraceUV, lastTracePixelInterval (the last two may be NULL). */
int dso_b200_trace_set_points(dso_b200_ctx* ctx, int n, const float* u, const float* v, const float* color8, const float* weights8,
                              const float* gradH4, const float* energyTH, const int* host_idx,
                              const float* idepth_min, const float* idepth_max, const float* quality, const int* status,
                              const float* lastTraceUV2, const float* lastTracePixelInterval);
/* FullSystem::traceNewCoarse (FullSystem.cpp:486-530): ImmaturePoint::traceOn (ImmaturePoint.cpp:77-438) for every uploaded
 * point on the new frame in `frame_slot` (device pyramid from dso_b200_make_images*).  Per host frame h < n_hosts:
 * hostToFrame_KRKi (9, row-major), hostToFrame_Kt (3), hostToFrame_affine (2) as computed at FullSystem.cpp:497-508.
 * The trace state is updated in place on the device; any non-NULL output array receives a copy (n entries, UV: 2n). */
int dso_b200_trace_new_coarse(dso_b200_ctx* ctx, int frame_slot, int n_hosts, const float* KRKi9, const float* Kt3, const float* aff2,
                              float* idepth_min_out, float* idepth_max_out, float* quality_out, int* status_out,
                              float* lastTraceUV2_out, float* lastTracePixelInterval_out);

/* FullSystem::optimizeImmaturePoint (FullSystemOptPoint.cpp:51-209, with ImmaturePoint::linearizeResidual,
 * ImmaturePoint.cpp:499-570) for n immature points of ONE host key frame: the other key frames are the targets
 * (target_slots: their device pyramids; per target PRE_RTll (9, row-major), PRE_tTll (3), PRE_aff_mode (2) of
 * host->targetPrecalc[target], HessianBlocks.h:73-99), fx..cy = HCalib->fxl() ... cyl().
 * result[k]: 0 = the reference returns null, -1 = its "(PointHessian*)-1" marker, 1 = the point is activated with
 * idepth_out[k] (NaN otherwise); res_state_out[k*n_targets + t] (may be NULL): final ResState of the temporary residual to
 * target t, i.e. which PointFrameResiduals to create (:185-205).  Called from FullSystem::activatePointsMT_Reductor. */
int dso_b200_optimize_immature(dso_b200_ctx* ctx, int n_targets, const int* target_slots, float fx, float fy, float cx, float cy,
                               const float* RTll9, const float* tTll3, const float* aff2,
                               int n, const float* u, const float* v, const float* color8, const float* weights8, const float* energyTH,
                               const float* idepth_min, const float* idepth_max, int minObs,
                               int* result, float* idepth_out, int* res_state_out);

/* ---- next row N4: the two-frame initialiser (FullSystem/CoarseInitializer.{h,cpp}) -------------------------------------
 * CoarseInitializer::calcResAndGS (CoarseInitializer.cpp:370-732, decl CoarseInitializer.h:166-171) on pyramid level lvl:
 * first_slot / new_slot hold the device pyramids of firstFrame / newFrame (dso_b200_make_images*), fx..cy are
 * HCalib->fxl() ... cyl() (the per-level intrinsics follow CoarseInitializer::makeK, :1126-1155, in double), the Pnt fields
 * (CoarseInitializer.h:45-85) come as arrays of npts entries: u, v, idepth_new, isGood, energy (2 per point), outlierTH,
 * iR; alphaW, alphaK, couplingWeight are the members trackFrame sets (:101-105); setting_huberTH comes from
 * dso_b200_set_settings.  Outputs: H_out, b_out, H_out_sc, b_out_sc (row-major 8x8 / 8 floats), res3 = the returned Vec3f
 * (E.A, alphaEnergy, E.num), and per point (each may be NULL) isGood_new, energy_new (2), maxstep, lastHessian_new and
 * JbBuffer_new (10).  The per-point outputs are bit-identical to the reference for the points it keeps (isGood_new); for the
 * others JbBuffer_new / lastHessian_new, which the reference leaves stale, are written as 0.  Points must keep the border
 * CoarseInitializer::setFirst gives them (pattern + 1 pixel); closer ones are dropped instead of read out of bounds. */
int dso_b200_init_calc_res_and_gs(dso_b200_ctx* ctx, int first_slot, int new_slot, int lvl,
                                  float fx, float fy, float cx, float cy,
                                  const double* refToNew12, double aff_a, double aff_b,
                                  int npts, const float* u, const float* v, const float* idepth_new, const unsigned char* isGood,
                                  const float* energy2, const float* outlierTH, const float* iR,
                                  float alphaW, float alphaK, float couplingWeight,
                                  float* H64, float* b8, float* Hsc64, float* bsc8, float* res3,
                                  unsigned char* isGood_new, float* energy_new2, float* maxstep, float* lastHessian_new,
                                  float* JbBuffer_new10);

#ifdef __cplusplus
}
#endif
#endif
