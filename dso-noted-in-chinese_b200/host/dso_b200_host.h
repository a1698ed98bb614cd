// Host-side mirror of the reference's C++ class surface for the hot path (namespace dso), forwarding to the
// C ABI of include/dso_b200.h.  Same member names, argument meaning and failure behaviour as the reference:
// there are no status codes on this surface -- a non-zero C-ABI return aborts (there is no CPU fallback).
//
// The reference builds on Eigen/Sophus types that are not available in this image, so the value types
// below (SE3, AffLight, Vec5, Mat88 ...) are minimal PODs with the reference's names; in the real tree the
// maintainer keeps Eigen/Sophus and uses the two-line converters shown in INTEGRATION.md instead.
#pragma once
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <vector>
#include "../../include/dso_b200.h"

namespace dso
{
#define DSO_B200_CHECK(call)                                                                      \
	do { int rc_ = (call); if (rc_ != 0) { fprintf(stderr, "dso_b200: %s failed (%d): %s\n", #call, rc_, dso_b200_last_error()); abort(); } } while (0)

struct SE3 { double R[9]; double t[3]; SE3() { for (int i = 0; i < 9; i++) R[i] = (i % 4 == 0); t[0] = t[1] = t[2] = 0; } };
struct AffLight { double a, b; AffLight(double a_ = 0, double b_ = 0) : a(a_), b(b_) {} };   // util/NumType.h:166-197
struct Vec5 { double v[5]; double& operator[](int i) { return v[i]; } const double& operator[](int i) const { return v[i]; } };
struct Vec3 { double v[3]; double& operator[](int i) { return v[i]; } };
struct Vec6 { double v[6]; double& operator[](int i) { return v[i]; } };
struct Vec8 { double v[8]; double& operator[](int i) { return v[i]; } };
struct Mat88 { double m[64]; double& operator()(int r, int c) { return m[r * 8 + c]; } };

static inline void pose12(const SE3& T, double* p) { memcpy(p, T.R, 72); memcpy(p + 9, T.t, 24); }
static inline void from12(const double* p, SE3& T) { memcpy(T.R, p, 72); memcpy(T.t, p + 9, 24); }

// CalibHessian (FullSystem/HessianBlocks.h:322-410): only what the hot path reads
struct CalibHessian
{
	float value_scaledf[4];
	float B[256];
	float& fxl() { return value_scaledf[0]; }
	float& fyl() { return value_scaledf[1]; }
	float& cxl() { return value_scaledf[2]; }
	float& cyl() { return value_scaledf[3]; }
};

// One device context per FullSystem (FullSystem.h:141); shared by the classes below.
struct DeviceContext
{
	dso_b200_ctx* ctx;
	int w, h, levels;
	DeviceContext(int device, int w_, int h_, int levels_, int n_frame_slots) : ctx(0), w(w_), h(h_), levels(levels_)
	{ DSO_B200_CHECK(dso_b200_create(device, w_, h_, n_frame_slots, &ctx)); }
	~DeviceContext() { dso_b200_destroy(ctx); }
};

// FrameHessian (FullSystem/HessianBlocks.h:105-320): pyramid ownership as in the reference (new[] / delete[])
struct FrameHessian
{
	DeviceContext* dev;
	int frame_slot;                 // device pyramid of this frame
	float* dIp[DSO_B200_MAX_LEVELS];           // Eigen::Vector3f* in the reference: 3 floats per pixel
	float* absSquaredGrad[DSO_B200_MAX_LEVELS];
	float* dI;
	float ab_exposure;
	float frameEnergyTH;
	FrameHessian(DeviceContext* d, int slot) : dev(d), frame_slot(slot), dI(0), ab_exposure(1), frameEnergyTH(8 * 8 * 8)
	{ for (int i = 0; i < DSO_B200_MAX_LEVELS; i++) { dIp[i] = 0; absSquaredGrad[i] = 0; } }
	~FrameHessian() { for (int i = 0; i < DSO_B200_MAX_LEVELS; i++) { delete[] dIp[i]; delete[] absSquaredGrad[i]; } }

	// void FrameHessian::makeImages(float* color, CalibHessian* HCalib)   HessianBlocks.cpp:137
	void makeImages(float* color, CalibHessian* HCalib)
	{
		for (int i = 0; i < dev->levels; i++)
		{
			size_t n = (size_t)(dev->w >> i) * (dev->h >> i);
			dIp[i] = new float[3 * n];
			absSquaredGrad[i] = new float[n];
		}
		dI = dIp[0];
		DSO_B200_CHECK(dso_b200_make_images(dev->ctx, frame_slot, color, dev->w, dev->h, dev->levels,
		                                    HCalib ? HCalib->B : 0, dIp, absSquaredGrad));
	}
};

// CoarseTracker (FullSystem/CoarseTracker.h:44-127)
class CoarseTracker
{
public:
	CoarseTracker(DeviceContext* d, int slot) : lastRef(0), newFrame(0), lastRef_aff_g2l(0, 0), firstCoarseRMSE(-1), dev(d), trk(slot)
	{ for (int i = 0; i < 5; i++) lastResiduals[i] = NAN; for (int i = 0; i < 3; i++) lastFlowIndicators[i] = 1000; fx0 = fy0 = cx0 = cy0 = 0; }

	// void makeK(CalibHessian* HCalib)   CoarseTracker.cpp:102 (the per-level K pyramid is derived inside the library)
	void makeK(CalibHessian* HCalib) { fx0 = HCalib->fxl(); fy0 = HCalib->fyl(); cx0 = HCalib->cxl(); cy0 = HCalib->cyl(); }

	// void setCoarseTrackingRef(std::vector<FrameHessian*>)   CoarseTracker.cpp:523: makeCoarseDepthL0 (:135-292) stays on the
	// host; its per-level outputs pc_u/pc_v/pc_idepth/pc_color/pc_n are uploaded here.
	void setCoarseTrackingRef(FrameHessian* lastRef_, AffLight ref_aff, const int* pc_n, const float* const* pc_u,
	                          const float* const* pc_v, const float* const* pc_idepth, const float* const* pc_color)
	{
		lastRef = lastRef_; lastRef_aff_g2l = ref_aff; firstCoarseRMSE = -1;
		DSO_B200_CHECK(dso_b200_tracker_set_ref(dev->ctx, trk, dev->w, dev->h, dev->levels, fx0, fy0, cx0, cy0, pc_n, pc_u, pc_v,
		                                        pc_idepth, pc_color, ref_aff.a, ref_aff.b, lastRef_->ab_exposure));
	}

	// bool trackNewestCoarse(FrameHessian*, SE3&, AffLight&, int coarsestLvl, Vec5 minResForAbort, Output3DWrapper* = 0)
	// CoarseTracker.cpp:551.  The whole LM loop runs in one persistent kernel; the abort test (:713) is replayed.
	bool trackNewestCoarse(FrameHessian* newFrameHessian, SE3& lastToNew_out, AffLight& aff_g2l_out, int coarsestLvl,
	                       Vec5 minResForAbort, void* wrap = 0)
	{
		(void)wrap;
		newFrame = newFrameHessian;
		double p[12];
		pose12(lastToNew_out, p);
		dso_b200_track_result r;
		DSO_B200_CHECK(dso_b200_track_batched(dev->ctx, trk, newFrame->frame_slot, 1, p, aff_g2l_out.a, aff_g2l_out.b,
		                                      newFrame->ab_exposure, coarsestLvl, 0, &r));
		double lr[5];
		int ok = dso_b200_track_replay_abort(&r, minResForAbort.v, lr);
		for (int i = 0; i < 5; i++) lastResiduals[i] = lr[i];
		for (int i = 0; i < 3; i++) lastFlowIndicators[i] = r.lastFlowIndicators[i];
		bool aborted = false;
		for (int q = 0; q < r.n_passes; q++) if (r.pass_rmse[q] > 1.5 * minResForAbort[r.pass_level[q]]) aborted = true;
		if (!aborted) { from12(r.pose, lastToNew_out); aff_g2l_out.a = r.aff_a; aff_g2l_out.b = r.aff_b; }   // :727-728
		return ok != 0;
	}

	// Vec6 calcRes(int lvl, const SE3& refToNew, AffLight aff_g2l, float cutoffTH)   CoarseTracker.cpp:359
	Vec6 calcRes(int lvl, const SE3& refToNew, AffLight aff_g2l, float cutoffTH)
	{
		double p[12];
		pose12(refToNew, p);
		Vec6 rs;
		DSO_B200_CHECK(dso_b200_calc_res(dev->ctx, trk, newFrame->frame_slot, lvl, p, aff_g2l.a, aff_g2l.b, newFrame->ab_exposure, cutoffTH, rs.v));
		return rs;
	}
	// void calcGSSSE(int lvl, Mat88& H_out, Vec8& b_out, const SE3& refToNew, AffLight aff_g2l)   CoarseTracker.cpp:296
	void calcGSSSE(int lvl, Mat88& H_out, Vec8& b_out, const SE3& refToNew, AffLight aff_g2l)
	{
		(void)lvl; (void)refToNew; (void)aff_g2l;   // always follows calcRes on the same state (:584-595, :667-687)
		DSO_B200_CHECK(dso_b200_calc_gs(dev->ctx, trk, H_out.m, b_out.v));
	}

	FrameHessian* lastRef;
	FrameHessian* newFrame;
	AffLight lastRef_aff_g2l;
	double lastResiduals[5];       // Vec5
	double lastFlowIndicators[3];  // Vec3
	double firstCoarseRMSE;
private:
	DeviceContext* dev;
	int trk;
	float fx0, fy0, cx0, cy0;
};

// ---- windowed BA: EnergyFunctional's accumulate / solve members (OptimizationBackend/EnergyFunctional.h:60-160) ----
struct VecX { std::vector<double> v; int size() const { return (int)v.size(); } double& operator[](int i) { return v[i]; } const double& operator[](int i) const { return v[i]; } };
struct MatXX { int n; std::vector<double> m; MatXX() : n(0) {} double& operator()(int r, int c) { return m[(size_t)r * n + c]; } int rows() const { return n; } int cols() const { return n; } };

// Flat description of the optimisation window, gathered by the caller from frameHessians / pointHessians / residuals
// in EnergyFunctional::makeIDX order (EnergyFunctional.cpp:928-948): frames by idx, points by host frame, residuals by point.
struct WindowArrays
{
	int nF, nP, nR;
	const int* frame_slots; const float* frameEnergyTH;                           // per frame
	const double* evalPT12; const double* state10; const double* state_zero10; const float* ab_exposure;
	const float *u, *v, *idepth_scaled, *idepth_zero_scaled, *color8, *weights8, *priorF, *deltaF; const int* host;   // per point
	const int *res_point, *res_target, *res_state; const double* res_energy; const unsigned char *res_isLinearized, *res_isActive;   // per residual
};

class EnergyFunctional;

// PointFrameResidual (FullSystem/Residuals.h:56-112): the per-object methods remain.  FullSystem::linearizeAll works on the
// whole window (EnergyFunctional::linearizeAll below); the serial callers of the per-object form -- flagPointsForRemoval,
// FullSystem.cpp:776-787 -- go through a ResidualBatch that forwards a batch of 1..k residuals in one call.
enum ResState { IN = 0, OOB, OUTLIER };             // Residuals.h:47
struct PointFrameResidual
{
	int idx;                        // position in EFPoint::residualsAll order, as uploaded by EnergyFunctional::setWindow
	ResState state_state, state_NewState;
	double state_energy, state_NewEnergy;
	bool isActive;                  // efResidual->isActive()
	explicit PointFrameResidual(int i) : idx(i), state_state(IN), state_NewState(OUTLIER), state_energy(0), state_NewEnergy(0), isActive(false) {}
};
class ResidualBatch
{
public:
	explicit ResidualBatch(DeviceContext* d) : dev(d) {}
	void add(PointFrameResidual* r) { rs.push_back(r); }
	// for every added residual: r->resetOOB(); r->linearize(HCalib); r->applyRes(true)   (Residuals.h:101-108, :95, :110)
	void resetOOB_linearize_applyRes()
	{
		const int k = (int)rs.size();
		if (!k) return;
		std::vector<int> idx(k), ns(k); std::vector<float> ne(k); std::vector<unsigned char> ia(k);
		for (int q = 0; q < k; q++) idx[q] = rs[q]->idx;
		DSO_B200_CHECK(dso_b200_ba_linearize_residuals(dev->ctx, k, idx.data(), 1, 1, ns.data(), ne.data(), ia.data()));
		for (int q = 0; q < k; q++)
		{
			PointFrameResidual* r = rs[q];
			r->state_NewState = (ResState)ns[q]; r->state_NewEnergy = ne[q];
			r->state_state = r->state_NewState; r->state_energy = r->state_NewEnergy;     // applyRes, Residuals.cpp:344-345
			r->isActive = ia[q] != 0;
		}
		rs.clear();
	}
private:
	DeviceContext* dev;
	std::vector<PointFrameResidual*> rs;
};

class EnergyFunctional
{
public:
	explicit EnergyFunctional(DeviceContext* d) : nFrames(0), nPoints(0), nResiduals(0), resInA(0), resInL(0), currentLambda(0), dev(d) {}

	// insertFrame / insertPoint / insertResidual / makeIDX (EnergyFunctional.cpp:432-520, 928-948) collapse into one upload
	void setWindow(const WindowArrays& W, CalibHessian* HCalib)
	{
		nFrames = W.nF; nPoints = W.nP; nResiduals = W.nR;
		DSO_B200_CHECK(dso_b200_ba_set_window(dev->ctx, W.nF, W.frame_slots, dev->w, dev->h, HCalib->fxl(), HCalib->fyl(), HCalib->cxl(), HCalib->cyl(), W.frameEnergyTH));
		setPrecalcValues(W);
		DSO_B200_CHECK(dso_b200_ba_set_points(dev->ctx, W.nP, W.u, W.v, W.idepth_scaled, W.idepth_zero_scaled, W.color8, W.weights8, W.host, W.priorF, W.deltaF));
		DSO_B200_CHECK(dso_b200_ba_set_residuals(dev->ctx, W.nR, W.res_point, W.res_target, W.res_state, W.res_energy, W.res_isLinearized, W.res_isActive));
	}
	// FullSystem::setPrecalcValues (FullSystemOptimize.cpp:413-422): FrameFramePrecalc::set for all pairs + ef->setDeltaF
	void setPrecalcValues(const WindowArrays& W)
	{
		DSO_B200_CHECK(dso_b200_ba_set_frames(dev->ctx, W.evalPT12, W.state10, W.state_zero10, W.ab_exposure, 0, 0, 0));
	}
	// FullSystem::linearizeAll(fixLinearization) (FullSystemOptimize.cpp:146-220): energy of the new linearisation;
	// fixLinearization = true also applies it (applyRes_Reductor(true, ...))
	double linearizeAll(bool fixLinearization)
	{
		double e = 0;
		DSO_B200_CHECK(dso_b200_ba_linearize(dev->ctx, 0, 0, 0, 0, 0, 0, &e));
		if (fixLinearization) applyRes();
		return e;
	}
	void applyRes() { DSO_B200_CHECK(dso_b200_ba_apply_res(dev->ctx, 0, 0)); }   // FullSystem::applyRes_Reductor(true, ...)

	// void accumulateAF_MT(MatXX &H, VecX &b, bool MT)   EnergyFunctional.cpp:209 (MT has no meaning on the device)
	void accumulateAF_MT(MatXX& H, VecX& b, bool MT) { (void)MT; size(H, b); DSO_B200_CHECK(dso_b200_ba_accumulate_A(dev->ctx, H.m.data(), b.v.data(), &resInA, 0, 0, 0)); }
	// void accumulateLF_MT(MatXX &H, VecX &b, bool MT)   EnergyFunctional.cpp:233
	void accumulateLF_MT(MatXX& H, VecX& b, bool MT) { (void)MT; size(H, b); DSO_B200_CHECK(dso_b200_ba_accumulate_L(dev->ctx, H.m.data(), b.v.data(), &resInL, 0, 0, 0)); }
	// void accumulateSCF_MT(MatXX &H, VecX &b, bool MT)  EnergyFunctional.cpp:257
	void accumulateSCF_MT(MatXX& H, VecX& b, bool MT) { (void)MT; size(H, b); DSO_B200_CHECK(dso_b200_ba_accumulate_SC(dev->ctx, H.m.data(), b.v.data(), 0, 0, 0)); }

	// priors / marginalisation prior / null spaces, set when they change (insertFrame, marginalizeFrame, FullSystem::optimize)
	void setPriors(const double* cPrior4, const double* framePrior8, const MatXX& HM, const VecX& bM)
	{ DSO_B200_CHECK(dso_b200_ba_set_priors(dev->ctx, cPrior4, framePrior8, 0, 0, HM.m.data(), bM.v.data())); }
	void setNullspaces(const double* ns, int n, double setting_solverModeDelta) { DSO_B200_CHECK(dso_b200_ba_set_nullspaces(dev->ctx, ns, n, setting_solverModeDelta)); }
	// the globals the non-default solver branches read: "for(EFFrame* f : frames) if(f->frameID==0) haveFirstFrame=true"
	// (EnergyFunctional.cpp:686-688, :833-835), setting_solverMode as marginalizePointsF sees it, setting_solverModeDelta
	void setSolverState(bool haveFirstFrame, int setting_solverMode, double setting_solverModeDelta)
	{ DSO_B200_CHECK(dso_b200_ba_set_solver_state(dev->ctx, haveFirstFrame ? 1 : 0, setting_solverMode, setting_solverModeDelta)); }

	// void solveSystemF(int iteration, double lambda, CalibHessian* HCalib)   EnergyFunctional.cpp:802; lastX as in :921,
	// frame / calibration steps are -x segments (:281-288); point steps stay on the device (doStepFromBackupPoints)
	void solveSystemF(int iteration, double lambda, int setting_solverMode, const double* value_minus_value_zero4)
	{
		DSO_B200_CHECK(dso_b200_ba_set_calib_delta(dev->ctx, value_minus_value_zero4));
		lastX.v.assign(8 * nFrames + 4, 0.0);
		currentLambda = lambda;
		DSO_B200_CHECK(dso_b200_ba_solve(dev->ctx, iteration, lambda, setting_solverMode, lastX.v.data(), 0, 0, 0));
		currentLambda = 0;
	}
	// void marginalizePointsF()   EnergyFunctional.cpp:642 (accumulation + HM/bM update; removePoint/makeIDX stay with the caller)
	void marginalizePointsF(const unsigned char* isMarginalizeFlag, float setting_idepthFixPriorMargFac, double setting_margWeightFac)
	{ DSO_B200_CHECK(dso_b200_ba_marginalize_points(dev->ctx, isMarginalizeFlag, setting_idepthFixPriorMargFac, setting_margWeightFac, 0, 0, &resInM, 0)); }
	// the per-point results of the last accumulation (EFPoint::Hdd_accAF / bd_accAF / Hcd_accAF / HdiF / bdSumF and
	// PointHessian::idepth_hessian, which FullSystem::flagPointsForRemoval and the point marginalisation read); nPoints floats
	// each (4 nPoints for Hcd_accAF), any pointer may be null
	void getPointTerms(float* Hdd_accAF, float* bd_accAF, float* Hcd_accAF4, float* HdiF, float* bdSumF, float* idepth_hessian)
	{ DSO_B200_CHECK(dso_b200_ba_get_point_terms(dev->ctx, Hdd_accAF, bd_accAF, Hcd_accAF4, HdiF, bdSumF, idepth_hessian)); }
	// double calcLEnergyF_MT() / double calcMEnergyF()   EnergyFunctional.cpp:411, :334
	double calcLEnergyF_MT() { double l = 0, m = 0; DSO_B200_CHECK(dso_b200_ba_calc_energies(dev->ctx, &l, &m)); return l; }
	double calcMEnergyF() { double l = 0, m = 0; DSO_B200_CHECK(dso_b200_ba_calc_energies(dev->ctx, &l, &m)); return m; }
	// point loops of FullSystem::backupState / doStepFromBackup / loadSateBackup (FullSystemOptimize.cpp:225-345)
	void backupStatePoints() { DSO_B200_CHECK(dso_b200_ba_points_backup(dev->ctx)); }
	void doStepFromBackupPoints(float stepfacD, float& sumID, float& sumNID) { float s[2]; DSO_B200_CHECK(dso_b200_ba_points_step(dev->ctx, stepfacD, s, 0)); sumID = s[0]; sumNID = s[1]; }
	void loadStateBackupPoints() { DSO_B200_CHECK(dso_b200_ba_points_restore(dev->ctx)); }

	int nFrames, nPoints, nResiduals;
	int resInA, resInL, resInM = 0;
	VecX lastX;
	double currentLambda;
private:
	void size(MatXX& H, VecX& b) const { const int N = 8 * nFrames + 4; H.n = N; H.m.assign((size_t)N * N, 0.0); b.v.assign(N, 0.0); }
	DeviceContext* dev;
};

// ---- CoarseInitializer (FullSystem/CoarseInitializer.h:45-223): the linearisation of the two-frame bootstrap ----
struct Vec3f { float v[3]; float& operator[](int i) { return v[i]; } };
struct Vec8f { float v[8]; float& operator[](int i) { return v[i]; } };
struct Mat88f { float m[64]; float& operator()(int r, int c) { return m[r * 8 + c]; } };
struct Pnt                      // the fields calcResAndGS reads and writes (CoarseInitializer.h:45-85)
{
	float u, v, idepth; bool isGood; float energy[2];
	float idepth_new; bool isGood_new; float energy_new[2];
	float iR, lastHessian, lastHessian_new, maxstep, outlierTH;
};
class CoarseInitializer
{
public:
	CoarseInitializer(DeviceContext* d, CalibHessian* HCalib) : firstFrame(0), newFrame(0), alphaK(2.5f * 2.5f), alphaW(150.f * 150.f), couplingWeight(1), dev(d), calib(HCalib)
	{ for (int i = 0; i < DSO_B200_MAX_LEVELS; i++) { points[i] = 0; numPoints[i] = 0; } }
	Pnt* points[DSO_B200_MAX_LEVELS];
	int numPoints[DSO_B200_MAX_LEVELS];
	FrameHessian* firstFrame;
	FrameHessian* newFrame;
	std::vector<float> JbBuffer_new;          // Vec10f per point of the level last linearised
	float alphaK, alphaW, couplingWeight;     // set by trackFrame (CoarseInitializer.cpp:101-105)

	// Vec3f calcResAndGS(int lvl, Mat88f&, Vec8f&, Mat88f&, Vec8f&, const SE3& refToNew, AffLight refToNew_aff, bool plot)
	// CoarseInitializer.cpp:370: one kernel over the points of the level; the Pnt fields are gathered / scattered here
	Vec3f calcResAndGS(int lvl, Mat88f& H_out, Vec8f& b_out, Mat88f& H_out_sc, Vec8f& b_out_sc, const SE3& refToNew, AffLight refToNew_aff, bool plot)
	{
		(void)plot;
		const int n = numPoints[lvl];
		Pnt* p = points[lvl];
		std::vector<float> u(n), v(n), id(n), e(2 * n), oth(n), iR(n), en(2 * n), ms(n), lh(n);
		std::vector<unsigned char> g(n), gn(n);
		for (int i = 0; i < n; i++)
		{
			u[i] = p[i].u; v[i] = p[i].v; id[i] = p[i].idepth_new; g[i] = p[i].isGood; e[2 * i] = p[i].energy[0]; e[2 * i + 1] = p[i].energy[1];
			oth[i] = p[i].outlierTH; iR[i] = p[i].iR;
		}
		JbBuffer_new.assign((size_t)n * 10, 0.f);
		double pose[12];
		pose12(refToNew, pose);
		Vec3f res;
		DSO_B200_CHECK(dso_b200_init_calc_res_and_gs(dev->ctx, firstFrame->frame_slot, newFrame->frame_slot, lvl, calib->fxl(), calib->fyl(),
		                                             calib->cxl(), calib->cyl(), pose, refToNew_aff.a, refToNew_aff.b, n, u.data(), v.data(), id.data(),
		                                             g.data(), e.data(), oth.data(), iR.data(), alphaW, alphaK, couplingWeight, H_out.m, b_out.v,
		                                             H_out_sc.m, b_out_sc.v, res.v, gn.data(), en.data(), ms.data(), lh.data(), JbBuffer_new.data()));
		for (int i = 0; i < n; i++)
		{
			p[i].isGood_new = gn[i] != 0; p[i].energy_new[0] = en[2 * i]; p[i].energy_new[1] = en[2 * i + 1]; p[i].maxstep = ms[i];
			if (p[i].isGood_new) p[i].lastHessian_new = lh[i];
		}
		return res;
	}
private:
	DeviceContext* dev;
	CalibHessian* calib;
};

}   // namespace dso
