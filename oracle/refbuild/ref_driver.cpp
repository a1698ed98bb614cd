// TEST INFRASTRUCTURE ONLY -- C driver around the UNMODIFIED reference translation units (compiled from
// /root/reference against the stand-in Eigen/Sophus/Boost headers in compat/, see README.md).  It builds the
// reference's own object graph (FrameHessian / PointHessian / PointFrameResidual / EnergyFunctional /
// CoarseTracker) from the same flat arrays the oracle takes and calls the reference's own member functions,
// so that oracle/oracle_*.cpp can be checked against the real code.  Same C signatures as dso_oracle.h with
// the prefix dso_ref_.
#define private public
#define protected public
#include "util/NumType.h"
#include "util/settings.h"
#include "util/globalCalib.h"
#include "util/FrameShell.h"
#include "FullSystem/FullSystem.h"
#include "FullSystem/HessianBlocks.h"
#include "FullSystem/Residuals.h"
#include "FullSystem/ImmaturePoint.h"
#include "FullSystem/CoarseTracker.h"
#include "FullSystem/CoarseInitializer.h"
#include "FullSystem/PixelSelector2.h"
#include "OptimizationBackend/EnergyFunctional.h"
#include "OptimizationBackend/EnergyFunctionalStructs.h"
#undef private
#undef protected
#include "../dso_oracle.h"
#include <cstring>
#include <vector>

namespace dso
{
// statics that live in translation units we do not compile (FullSystem.cpp:61-63)
int FrameHessian::instanceCounter = 0;
int PointHessian::instanceCounter = 0;
int CalibHessian::instanceCounter = 0;
}
using namespace dso;

// FullSystem/PixelSelector2.cpp does not compile in this annotated fork (comments swallowed code at :137 and :274), so the
// three members CoarseInitializer::setFirst links against are stubs: setFirst (point selection, kd-tree) is outside the
// pinned path, the driver fills the Pnt arrays itself.
namespace dso
{
PixelSelector::PixelSelector(int, int) { abort(); }
PixelSelector::~PixelSelector() {}
int PixelSelector::makeMaps(const FrameHessian* const, float*, float, int, bool, float) { abort(); return 0; }
}

static void set_calib(int w, int h, int levels, float fx, float fy, float cx, float cy)
{
	static int lw = -1, lh = -1, ll = -1; static float lfx = 0, lfy = 0, lcx = 0, lcy = 0;
	if (w == lw && h == lh && levels == ll && fx == lfx && fy == lfy && cx == lcx && cy == lcy) return;   // globals already describe this camera
	lw = w; lh = h; ll = levels; lfx = fx; lfy = fy; lcx = cx; lcy = cy;
	Eigen::Matrix3f K;
	K << fx, 0.0, cx, 0.0, fy, cy, 0.0, 0.0, 1.0;
	setGlobalCalib(w, h, K);          // util/globalCalib.cpp:45 (prints the level count)
	pyrLevelsUsed = levels;           // plain global, util/settings.cpp:32
	setting_debugout_runquiet = true;
	setting_render_displayCoarseTrackingFull = false;
	multiThreading = true;
}

static SE3 to_se3(const double* p)
{
	Eigen::Matrix<double, 3, 3> R;
	for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) R(r, c) = p[r * 3 + c];
	return SE3(R, Vec3(p[9], p[10], p[11]));
}
static void from_se3(const SE3& T, double* p)
{
	for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) p[r * 3 + c] = T.rotationMatrix()(r, c);
	for (int r = 0; r < 3; r++) p[9 + r] = T.translation()[r];
}
static ImmaturePoint* raw_immature(FrameHessian* host, float u, float v, float idepth, const float* color, const float* weights)
{
	ImmaturePoint* ip = (ImmaturePoint*)calloc(1, sizeof(ImmaturePoint));   // ctor lives in an uncompiled TU; PointHessian's ctor only reads fields
	ip->host = host; ip->u = u; ip->v = v; ip->idepth_min = ip->idepth_max = idepth; ip->my_type = 1; ip->energyTH = 0;
	for (int i = 0; i < patternNum; i++) { ip->color[i] = color ? color[i] : 0; ip->weights[i] = weights ? weights[i] : 1; }
	return ip;
}

// ---- makeImages -----------------------------------------------------------------------------------------
extern "C" void dso_ref_make_images(const float* color, int w, int h, int levels, const float* B, float* const* dIp, float* const* absgrad)
{
	set_calib(w, h, levels, 500, 500, w * 0.5f, h * 0.5f);
	CalibHessian HCalib;
	if (B) for (int i = 0; i < 256; i++) HCalib.B[i] = B[i];
	FrameHessian* fh = new FrameHessian();
	fh->makeImages(const_cast<float*>(color), B ? &HCalib : 0);
	for (int l = 0; l < levels; l++)
	{
		const int wl = w >> l, hl = h >> l;
		// copy exactly what the reference writes: channel 0 everywhere, gradients / abs for idx in [wl, wl*(hl-1))
		for (int i = 0; i < wl * hl; i++) dIp[l][3 * i] = fh->dIp[l][i][0];
		for (int i = wl; i < wl * (hl - 1); i++) { dIp[l][3 * i + 1] = fh->dIp[l][i][1]; dIp[l][3 * i + 2] = fh->dIp[l][i][2]; absgrad[l][i] = fh->absSquaredGrad[l][i]; }
	}
	delete fh;
}

// ---- tracker --------------------------------------------------------------------------------------------
struct dso_oracle_tracker
{
	int w, h, levels; float fx, fy, cx, cy;
	CalibHessian* HCalib;
	CoarseTracker* ct;
	FrameHessian* ref; FrameHessian* hostDummy; FrameHessian* newF;
	FrameShell refShell;
	std::vector<PointHessian*> pts; std::vector<PointFrameResidual*> res; std::vector<EFPoint*> efp; std::vector<EFResidual*> efr; std::vector<ImmaturePoint*> ips;
	void calib() { set_calib(w, h, levels, fx, fy, cx, cy); }
};

extern "C" void dso_ref_set_affine_modes(float modeA, float modeB) { setting_affineOptModeA = modeA; setting_affineOptModeB = modeB; }
// the reference's own globals (util/settings.cpp), same block as dso_oracle_set_settings; NULL restores the defaults
struct ref_settings { float huberTH, outlierTHSumComponent, maxPixSearch; int minTraceTestRadius, GNItsOnPointActivation; float trace_stepsize;
                      int trace_GNIterations; float trace_GNThreshold, trace_extraSlackOnTH, trace_slackInterval, trace_minImprovementFactor, minIdepthH_act; };
extern "C" void dso_ref_set_settings(const ref_settings* p)
{
	const ref_settings d = { 9, 50 * 50, 0.027f, 2, 3, 1.0f, 3, 0.1f, 1.2f, 1.5f, 2, 100 };
	const ref_settings s = p ? *p : d;
	setting_huberTH = s.huberTH; setting_outlierTHSumComponent = s.outlierTHSumComponent; setting_maxPixSearch = s.maxPixSearch;
	setting_minTraceTestRadius = s.minTraceTestRadius; setting_GNItsOnPointActivation = s.GNItsOnPointActivation;
	setting_trace_stepsize = s.trace_stepsize; setting_trace_GNIterations = s.trace_GNIterations; setting_trace_GNThreshold = s.trace_GNThreshold;
	setting_trace_extraSlackOnTH = s.trace_extraSlackOnTH; setting_trace_slackInterval = s.trace_slackInterval;
	setting_trace_minImprovementFactor = s.trace_minImprovementFactor; setting_minIdepthH_act = s.minIdepthH_act;
}

extern "C" dso_oracle_tracker* dso_ref_tracker_create(int w, int h, int levels, float fx, float fy, float cx, float cy)
{
	dso_oracle_tracker* T = new dso_oracle_tracker();
	T->w = w; T->h = h; T->levels = levels; T->fx = fx; T->fy = fy; T->cx = cx; T->cy = cy;
	T->calib();
	T->HCalib = new CalibHessian();
	T->ct = new CoarseTracker(w, h);
	T->ct->makeK(T->HCalib);
	T->ref = 0; T->hostDummy = 0; T->newF = 0;
	return T;
}
extern "C" void dso_ref_tracker_destroy(dso_oracle_tracker* T) { delete T->ct; delete T->HCalib; delete T; }   // frames/points leak on purpose (test process)
extern "C" void dso_ref_tracker_get_K(const dso_oracle_tracker* T, int lvl, float* fx, float* fy, float* cx, float* cy, float* Ki9)
{
	*fx = T->ct->fx[lvl]; *fy = T->ct->fy[lvl]; *cx = T->ct->cx[lvl]; *cy = T->ct->cy[lvl];
	if (Ki9) for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) Ki9[r * 3 + c] = T->ct->Ki[lvl](r, c);
}
static FrameHessian* frame_from_pyr(dso_oracle_tracker* T, const float* const* pyr)
{
	FrameHessian* fh = new FrameHessian();
	for (int l = 0; l < T->levels; l++)
	{
		const int n = (T->w >> l) * (T->h >> l);
		fh->dIp[l] = new Eigen::Vector3f[n];
		fh->absSquaredGrad[l] = new float[n];
		memcpy((void*)fh->dIp[l], pyr[l], sizeof(float) * 3 * n);
	}
	fh->dI = fh->dIp[0];
	return fh;
}
extern "C" void dso_ref_tracker_set_ref(dso_oracle_tracker* T, int npts, const float* cu, const float* cv, const float* cid, const float* HdiF,
                                        const float* const* ref_dIp, double ref_a, double ref_b, float ref_exposure)
{
	T->calib();
	T->ref = frame_from_pyr(T, ref_dIp);
	T->ref->ab_exposure = ref_exposure;
	T->ref->shell = &T->refShell; T->refShell.id = 1;
	Vec10 st = Vec10::Zero(); st[6] = ref_a / SCALE_A; st[7] = ref_b / SCALE_B;
	T->ref->worldToCam_evalPT = SE3();
	T->ref->setState(st);             // aff_g2l() reads state_scaled[6,7]
	T->hostDummy = new FrameHessian();
	for (int k = 0; k < npts; k++)
	{
		ImmaturePoint* ip = raw_immature(T->hostDummy, 10, 10, 1, 0, 0);
		PointHessian* ph = new PointHessian(ip, T->HCalib);
		PointFrameResidual* r = new PointFrameResidual(ph, T->hostDummy, T->ref);
		r->centerProjectedTo = Vec3f(cu[k], cv[k], cid[k]);
		EFPoint* ep = (EFPoint*)calloc(1, sizeof(EFPoint)); ep->HdiF = HdiF[k];
		ph->efPoint = ep;
		ph->lastResiduals[0].first = r; ph->lastResiduals[0].second = ResState::IN;
		T->hostDummy->pointHessians.push_back(ph);
		T->pts.push_back(ph); T->res.push_back(r); T->ips.push_back(ip);
	}
	std::vector<FrameHessian*> fhs; fhs.push_back(T->hostDummy); fhs.push_back(T->ref);
	T->ct->setCoarseTrackingRef(fhs);     // CoarseTracker.cpp:523 -> makeCoarseDepthL0 :135
}
extern "C" void dso_ref_tracker_set_ref_aff(dso_oracle_tracker* T, double a, double b, float e) { T->ct->lastRef_aff_g2l = AffLight(a, b); if (T->ref) T->ref->ab_exposure = e; }
extern "C" int dso_ref_tracker_get_cloud(const dso_oracle_tracker* T, int lvl, float* u, float* v, float* idepth, float* color)
{
	const int n = T->ct->pc_n[lvl];
	if (u) memcpy(u, T->ct->pc_u[lvl], n * 4);
	if (v) memcpy(v, T->ct->pc_v[lvl], n * 4);
	if (idepth) memcpy(idepth, T->ct->pc_idepth[lvl], n * 4);
	if (color) memcpy(color, T->ct->pc_color[lvl], n * 4);
	return n;
}
extern "C" void dso_ref_tracker_set_new_frame(dso_oracle_tracker* T, const float* const* dIp, float ab_exposure)
{
	T->newF = frame_from_pyr(T, dIp);    // leaks the previous one (test process)
	T->newF->ab_exposure = ab_exposure;
	T->ct->newFrame = T->newF;
}
extern "C" void dso_ref_calc_res(dso_oracle_tracker* T, int lvl, const double* pose12, double a, double b, float cutoffTH, double* out6)
{
	T->calib();
	T->ct->debugPlot = false;
	Vec6 r = T->ct->calcRes(lvl, to_se3(pose12), AffLight(a, b), cutoffTH);     // CoarseTracker.cpp:359
	for (int i = 0; i < 6; i++) out6[i] = r[i];
}
extern "C" void dso_ref_calc_gs(dso_oracle_tracker* T, int lvl, double a, double b, double* H64, double* b8)
{
	Mat88 H; Vec8 bb;
	T->ct->calcGSSSE(lvl, H, bb, SE3(), AffLight(a, b));                       // CoarseTracker.cpp:296
	for (int r = 0; r < 8; r++) { for (int c = 0; c < 8; c++) H64[r * 8 + c] = H(r, c); b8[r] = bb[r]; }
}
extern "C" int dso_ref_get_warped(const dso_oracle_tracker* T, float* const* bufs8)
{
	const int n = T->ct->buf_warped_n;
	float* src[8] = { T->ct->buf_warped_idepth, T->ct->buf_warped_u, T->ct->buf_warped_v, T->ct->buf_warped_dx, T->ct->buf_warped_dy,
	                  T->ct->buf_warped_residual, T->ct->buf_warped_weight, T->ct->buf_warped_refColor };
	for (int k = 0; k < 8; k++) if (bufs8[k]) memcpy(bufs8[k], src[k], sizeof(float) * n);
	return n;
}
extern "C" void dso_ref_track_newest_coarse(dso_oracle_tracker* T, const double* pose12_in, double a_in, double b_in, int coarsestLvl,
                                            const double* minRes5, dso_oracle_track_result* out)
{
	T->calib();
	memset(out, 0, sizeof(*out));
	SE3 pose = to_se3(pose12_in);
	AffLight aff(a_in, b_in);
	Vec5 mr; for (int i = 0; i < 5; i++) mr[i] = minRes5[i];
	bool ok = T->ct->trackNewestCoarse(T->newF, pose, aff, coarsestLvl, mr, 0);   // CoarseTracker.cpp:551
	out->ok = ok ? 1 : 0; out->aborted_level = -1;
	from_se3(pose, out->pose); out->aff_a = aff.a; out->aff_b = aff.b;
	for (int i = 0; i < 5; i++) out->lastResiduals[i] = T->ct->lastResiduals[i];
	for (int i = 0; i < 3; i++) out->lastFlowIndicators[i] = T->ct->lastFlowIndicators[i];
}

// ---- windowed BA ------------------------------------------------------------------------------------------
struct dso_oracle_ba
{
	int w, h, nF; float fx, fy, cx, cy;
	CalibHessian* HCalib;
	EnergyFunctional* ef;
	std::vector<FrameHessian*> frames;
	std::vector<FrameShell*> shells;
	std::vector<PointHessian*> pts;
	std::vector<PointFrameResidual*> res;
	IndexThreadReduce<Vec10>* red;
	bool frames_inserted;
	int margSolverMode; double margSolverModeDelta;
	void calib() { set_calib(w, h, 1, fx, fy, cx, cy); }
};

extern "C" dso_oracle_ba* dso_ref_ba_create(int w, int h, int nF, float fx, float fy, float cx, float cy)
{
	dso_oracle_ba* B = new dso_oracle_ba();
	B->w = w; B->h = h; B->nF = nF; B->fx = fx; B->fy = fy; B->cx = cx; B->cy = cy;
	B->calib();
	B->HCalib = new CalibHessian();
	B->ef = new EnergyFunctional();
	B->red = new IndexThreadReduce<Vec10>();     // NUM_THREADS = 6 persistent workers (util/IndexThreadReduce.h)
	B->ef->red = B->red;
	B->frames.assign(nF, (FrameHessian*)0);
	B->shells.assign(nF, (FrameShell*)0);
	B->frames_inserted = false;
	B->margSolverMode = setting_solverMode; B->margSolverModeDelta = setting_solverModeDelta;
	return B;
}
extern "C" void dso_ref_ba_destroy(dso_oracle_ba* B) { delete B->red; delete B; }   // the object graph leaks on purpose (test process)

extern "C" void dso_ref_ba_set_frame(dso_oracle_ba* B, int idx, const double* evalPT12, const double* state10, const double* state_zero10,
                                     float ab_exposure, float frameEnergyTH, const float* dI)
{
	B->calib();
	FrameHessian* fh = new FrameHessian();
	fh->shell = new FrameShell(); fh->shell->id = idx;
	fh->frameID = idx; fh->idx = idx;
	fh->ab_exposure = ab_exposure; fh->frameEnergyTH = frameEnergyTH;
	const int n = B->w * B->h;
	fh->dIp[0] = new Eigen::Vector3f[n];
	fh->absSquaredGrad[0] = new float[n];
	memcpy((void*)fh->dIp[0], dI, sizeof(float) * 3 * n);
	fh->dI = fh->dIp[0];
	Vec10 st, sz;
	for (int i = 0; i < 10; i++) { st[i] = state10[i]; sz[i] = state_zero10[i]; }
	fh->worldToCam_evalPT = to_se3(evalPT12);
	fh->setState(st);                 // HessianBlocks.h:183: state_scaled, PRE_worldToCam, PRE_camToWorld
	fh->state_zero = sz;              // setStateZero also builds null spaces (not on this path)
	B->frames[idx] = fh;
}
extern "C" void dso_ref_ba_set_calib_delta(dso_oracle_ba* B, const double* d)
{
	for (int i = 0; i < 4; i++) B->HCalib->value_minus_value_zero[i] = d[i];
}
extern "C" void dso_ref_ba_prepare(dso_oracle_ba* B)
{
	B->calib();
	if (!B->frames_inserted)
	{
		for (FrameHessian* fh : B->frames) B->ef->insertFrame(fh, B->HCalib);      // EnergyFunctional.cpp:447 (-> setAdjointsF)
		B->frames_inserted = true;
	}
	// FullSystem::setPrecalcValues (FullSystemOptimize.cpp): FrameFramePrecalc::set for all pairs, then setDeltaF
	for (FrameHessian* fh : B->frames)
	{
		fh->targetPrecalc.resize(B->frames.size());
		for (unsigned int i = 0; i < B->frames.size(); i++) fh->targetPrecalc[i].set(fh, B->frames[i], B->HCalib);
	}
	B->ef->setAdjointsF(B->HCalib);
	B->ef->setDeltaF(B->HCalib);
}
extern "C" void dso_ref_ba_get_glue(const dso_oracle_ba* B, float* precalc, double* adHost, double* adTarget, float* adHTdeltaF, float* cDeltaF4)
{
	const int nF = B->nF;
	for (int h = 0; h < nF; h++)
		for (int t = 0; t < nF; t++)
		{
			const FrameFramePrecalc& P = B->frames[h]->targetPrecalc[t];
			if (precalc)
			{
				float* o = precalc + 28 * (h * nF + t);
				for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) { o[r * 3 + c] = P.PRE_RTll_0(r, c); o[12 + r * 3 + c] = P.PRE_KRKiTll(r, c); }
				for (int r = 0; r < 3; r++) { o[9 + r] = P.PRE_tTll_0[r]; o[21 + r] = P.PRE_KtTll[r]; }
				o[24] = P.PRE_aff_mode[0]; o[25] = P.PRE_aff_mode[1]; o[26] = P.PRE_b0_mode; o[27] = 0;
			}
			const int idx = h + t * nF;
			for (int r = 0; r < 8; r++)
				for (int c = 0; c < 8; c++)
				{
					if (adHost) adHost[idx * 64 + r * 8 + c] = B->ef->adHost[idx](r, c);
					if (adTarget) adTarget[idx * 64 + r * 8 + c] = B->ef->adTarget[idx](r, c);
				}
			if (adHTdeltaF) for (int c = 0; c < 8; c++) adHTdeltaF[idx * 8 + c] = B->ef->adHTdeltaF[idx](0, c);
		}
	if (cDeltaF4) for (int i = 0; i < 4; i++) cDeltaF4[i] = B->ef->cDeltaF[i];
}
extern "C" void dso_ref_ba_set_points(dso_oracle_ba* B, int nP, const float* u, const float* v, const float* idepth_scaled,
                                      const float* idepth_zero_scaled, const float* color8, const float* weights8,
                                      const int* host, const float* priorF, const float* deltaF)
{
	B->calib();
	for (int i = 0; i < nP; i++)
	{
		FrameHessian* hf = B->frames[host[i]];
		ImmaturePoint* ip = raw_immature(hf, u[i], v[i], idepth_scaled[i], color8 + 8 * i, weights8 + 8 * i);
		PointHessian* ph = new PointHessian(ip, B->HCalib);
		ph->setIdepthScaled(idepth_scaled[i]);
		ph->setIdepthZero(idepth_zero_scaled[i] * SCALE_IDEPTH_INVERSE);
		ph->hasDepthPrior = priorF[i] != 0;                                   // EFPoint::takeData: priorF = hasDepthPrior ? setting_idepthFixPrior*SCALE^2 : 0
		ph->setPointStatus(PointHessian::ACTIVE);
		hf->pointHessians.push_back(ph);
		B->ef->insertPoint(ph);                                                // EnergyFunctional.cpp:486
		ph->efPoint->priorF = priorF[i];
		ph->efPoint->deltaF = deltaF[i];
		B->pts.push_back(ph);
	}
}
extern "C" void dso_ref_ba_set_residuals(dso_oracle_ba* B, int nR, const int* point, const int* target, const int* state_state,
                                         const double* state_energy, const unsigned char* isLinearized, const unsigned char* isActive)
{
	B->calib();
	for (int i = 0; i < nR; i++)
	{
		PointHessian* ph = B->pts[point[i]];
		PointFrameResidual* r = new PointFrameResidual(ph, ph->host, B->frames[target[i]]);   // resetOOB in the ctor
		r->setState((ResState)state_state[i]);
		r->state_energy = state_energy ? state_energy[i] : 0;
		r->state_NewEnergy = r->state_energy;
		ph->residuals.push_back(r);
		B->ef->insertResidual(r);                                                              // EnergyFunctional.cpp:432
		r->efResidual->isLinearized = isLinearized ? isLinearized[i] != 0 : false;
		r->efResidual->isActiveAndIsGoodNEW = isActive ? isActive[i] != 0 : false;
		B->res.push_back(r);
	}
	B->ef->makeIDX();
}
static void export_lin(dso_oracle_ba* B, int* newState, float* newEnergy, float* newEnergyWO, float* cpt, float* proj, float* J74)
{
	for (size_t i = 0; i < B->res.size(); i++)
	{
		PointFrameResidual* r = B->res[i];
		if (newState) newState[i] = (int)r->state_NewState;
		if (newEnergy) newEnergy[i] = (float)r->state_NewEnergy;
		if (newEnergyWO) newEnergyWO[i] = (float)r->state_NewEnergyWithOutlier;
		if (cpt) for (int k = 0; k < 3; k++) cpt[3 * i + k] = r->centerProjectedTo[k];
		if (proj) for (int k = 0; k < 8; k++) { proj[16 * i + 2 * k] = r->projectedTo[k][0]; proj[16 * i + 2 * k + 1] = r->projectedTo[k][1]; }
		if (J74)
		{
			const RawResidualJacobian* J = r->J;
			float* o = J74 + 74 * i;
			for (int k = 0; k < 8; k++) { o[k] = J->resF[k]; o[30 + k] = J->JIdx[0][k]; o[38 + k] = J->JIdx[1][k]; o[46 + k] = J->JabF[0][k]; o[54 + k] = J->JabF[1][k]; }
			for (int k = 0; k < 6; k++) { o[8 + k] = J->Jpdxi[0][k]; o[14 + k] = J->Jpdxi[1][k]; }
			for (int k = 0; k < 4; k++) { o[20 + k] = J->Jpdc[0][k]; o[24 + k] = J->Jpdc[1][k]; }
			o[28] = J->Jpdd[0]; o[29] = J->Jpdd[1];
			for (int a = 0; a < 2; a++) for (int c = 0; c < 2; c++) { o[62 + a * 2 + c] = J->JIdx2(a, c); o[66 + a * 2 + c] = J->JabJIdx(a, c); o[70 + a * 2 + c] = J->Jab2(a, c); }
		}
	}
}
extern "C" double dso_ref_ba_linearize(dso_oracle_ba* B, int* newState, float* newEnergy, float* newEnergyWO, float* cpt, float* proj, float* J74, double* energy_ret)
{
	B->calib();
	double stats0 = 0;
	for (size_t k = 0; k < B->res.size(); k++)
	{
		double e = B->res[k]->linearize(B->HCalib);           // Residuals.cpp:78
		if (energy_ret) energy_ret[k] = e;
		stats0 += e;
	}
	export_lin(B, newState, newEnergy, newEnergyWO, cpt, proj, J74);
	return stats0;
}
static void lin_range(dso_oracle_ba* B, int min, int max, Vec10* stats, int tid)
{
	(void)tid;
	for (int k = min; k < max; k++) (*stats)[0] += B->res[k]->linearize(B->HCalib);   // FullSystem::linearizeAll_Reductor, FullSystemOptimize.cpp:52-60
}
extern "C" double dso_ref_ba_linearize_mt(dso_oracle_ba* B, int nthreads)
{
	(void)nthreads;   // the reference's pool has NUM_THREADS = 6 workers, compile-time
	B->calib();
	B->red->reduce(boost::bind(&lin_range, B, _1, _2, _3, _4), 0, (int)B->res.size(), 0);   // FullSystemOptimize.cpp:164
	return B->red->stats[0];
}
extern "C" void dso_ref_ba_apply_res(dso_oracle_ba* B, unsigned char* isActive_out, float* JpJdF_out)
{
	for (size_t i = 0; i < B->res.size(); i++)
	{
		B->res[i]->applyRes(true);                             // Residuals.cpp:323
		if (isActive_out) isActive_out[i] = B->res[i]->efResidual->isActive();
		if (JpJdF_out) for (int k = 0; k < 8; k++) JpJdF_out[8 * i + k] = B->res[i]->efResidual->JpJdF[k];
	}
}
extern "C" void dso_ref_ba_linearize_residuals(dso_oracle_ba* B, int k, const int* idx, int reset_oob, int apply,
                                               int* newState, float* newEnergy, unsigned char* isActive)
{
	B->calib();
	for (int q = 0; q < k; q++)
	{
		PointFrameResidual* r = B->res[idx[q]];
		if (reset_oob) r->resetOOB();                          // Residuals.h:101
		r->linearize(B->HCalib);                               // Residuals.cpp:78
		if (apply) r->applyRes(true);                          // Residuals.cpp:323
		if (newState) newState[q] = (int)r->state_NewState;
		if (newEnergy) newEnergy[q] = (float)r->state_NewEnergy;
		if (isActive) isActive[q] = r->efResidual->isActive();
	}
}
static void export_sys(const MatXX& H, const VecX& b, double* Ho, double* bo)
{
	const int N = H.rows();
	for (int r = 0; r < N; r++) { for (int c = 0; c < N; c++) Ho[r * N + c] = H(r, c); bo[r] = b[r]; }
}
extern "C" void dso_ref_ba_accumulate_A(dso_oracle_ba* B, int MT, int nthreads, double* H, double* b, int* resInA, float* Hdd, float* bd, float* Hcd4)
{
	(void)nthreads;
	B->calib();
	MatXX HA; VecX bA;
	B->ef->accumulateAF_MT(HA, bA, MT != 0);                   // EnergyFunctional.cpp:209
	export_sys(HA, bA, H, b);
	if (resInA) *resInA = B->ef->resInA;
	for (size_t i = 0; i < B->pts.size(); i++)
	{
		EFPoint* p = B->pts[i]->efPoint;
		if (Hdd) Hdd[i] = p->Hdd_accAF;
		if (bd) bd[i] = p->bd_accAF;
		if (Hcd4) for (int k = 0; k < 4; k++) Hcd4[4 * i + k] = p->Hcd_accAF[k];
	}
}
extern "C" void dso_ref_ba_accumulate_SC(dso_oracle_ba* B, int MT, int nthreads, double* H, double* b, float* HdiF, float* bdSumF, float* idepth_hessian)
{
	(void)nthreads;
	B->calib();
	// accumulateSCF reads Hdd_accLF etc., which accumulateLF_MT initialises (EnergyFunctional.cpp:817)
	MatXX HL; VecX bL;
	B->ef->accumulateLF_MT(HL, bL, MT != 0);
	MatXX HS; VecX bS;
	B->ef->accumulateSCF_MT(HS, bS, MT != 0);                  // EnergyFunctional.cpp:257
	export_sys(HS, bS, H, b);
	for (size_t i = 0; i < B->pts.size(); i++)
	{
		EFPoint* p = B->pts[i]->efPoint;
		if (HdiF) HdiF[i] = p->HdiF;
		if (bdSumF) bdSumF[i] = p->bdSumF;
		if (idepth_hessian) idepth_hessian[i] = B->pts[i]->idepth_hessian;
	}
}

// ---- N1/N2: accumulateLF_MT, fixLinearizationF, solveSystemF + resubstituteF_MT (the reference's own code) ----
extern "C" void dso_ref_ba_accumulate_L(dso_oracle_ba* B, int MT, int nthreads, double* H, double* b, int* resInL, float* Hdd, float* bd, float* Hcd4)
{
	(void)nthreads;
	B->calib();
	MatXX HL; VecX bL;
	B->ef->accumulateLF_MT(HL, bL, MT != 0);                   // EnergyFunctional.cpp:233
	export_sys(HL, bL, H, b);
	if (resInL) *resInL = B->ef->resInL;
	for (size_t i = 0; i < B->pts.size(); i++)
	{
		EFPoint* p = B->pts[i]->efPoint;
		if (Hdd) Hdd[i] = p->Hdd_accLF;
		if (bd) bd[i] = p->bd_accLF;
		if (Hcd4) for (int k = 0; k < 4; k++) Hcd4[4 * i + k] = p->Hcd_accLF[k];
	}
}
extern "C" void dso_ref_ba_set_priors(dso_oracle_ba* B, const double* cPrior4, const double* framePrior, const double* frameDeltaPrior,
                                      const double* frameDelta, const double* HM, const double* bM)
{
	const int nF = B->nF, N = 8 * nF + CPARS;
	EnergyFunctional* ef = B->ef;
	for (int i = 0; i < 4; i++) ef->cPrior[i] = cPrior4 ? cPrior4[i] : 0.0;
	ef->cPriorF = ef->cPrior.cast<float>();
	for (int h = 0; h < nF; h++)
		for (int i = 0; i < 8; i++)
		{
			ef->frames[h]->prior[i] = framePrior ? framePrior[8 * h + i] : 0.0;
			ef->frames[h]->delta_prior[i] = frameDeltaPrior ? frameDeltaPrior[8 * h + i] : 0.0;
			ef->frames[h]->delta[i] = frameDelta ? frameDelta[8 * h + i] : 0.0;
		}
	ef->HM = MatXX::Zero(N, N); ef->bM = VecX::Zero(N);
	if (HM) for (int r = 0; r < N; r++) for (int c = 0; c < N; c++) ef->HM(r, c) = HM[r * N + c];
	if (bM) for (int r = 0; r < N; r++) ef->bM[r] = bM[r];
}
extern "C" void dso_ref_ba_set_nullspaces(dso_oracle_ba* B, const double* ns, int n)
{
	const int N = 8 * B->nF + CPARS;
	B->ef->lastNullspaces_pose.clear(); B->ef->lastNullspaces_scale.clear();
	for (int k = 0; k < n; k++)
	{
		VecX v(N);
		for (int i = 0; i < N; i++) v[i] = ns[(size_t)k * N + i];
		B->ef->lastNullspaces_pose.push_back(v);               // orthogonalize concatenates pose and scale vectors (:745-747)
	}
}
extern "C" void dso_ref_ba_fix_linearization(dso_oracle_ba* B, const unsigned char* which, float* res_toZeroF_out)
{
	for (size_t i = 0; i < B->res.size(); i++)
	{
		EFResidual* r = B->res[i]->efResidual;
		if (r->isActive() && (!which || which[i])) r->fixLinearizationF(B->ef);    // EnergyFunctionalStructs.cpp:89
		if (res_toZeroF_out) for (int k = 0; k < 8; k++) res_toZeroF_out[8 * i + k] = r->res_toZeroF[k];
	}
}
extern "C" int dso_ref_ba_solve(dso_oracle_ba* B, int iteration, double lambda, int solverMode, double solverModeDelta, int MT, int nthreads,
                                double* x_out, float* pointStep_out, double* HFinal_out, double* bFinal_out)
{
	(void)nthreads; (void)HFinal_out; (void)bFinal_out;
	B->calib();
	const int N = 8 * B->nF + CPARS;
	if ((int)B->ef->bM.size() != N) dso_ref_ba_set_priors(B, 0, 0, 0, 0, 0, 0);
	const bool mtOld = multiThreading; const int smOld = setting_solverMode; const double sdOld = setting_solverModeDelta;
	multiThreading = MT != 0; setting_solverMode = solverMode; setting_solverModeDelta = solverModeDelta;
	B->ef->solveSystemF(iteration, lambda, B->HCalib);         // EnergyFunctional.cpp:802
	multiThreading = mtOld; setting_solverMode = smOld; setting_solverModeDelta = sdOld;
	if (x_out) for (int i = 0; i < N; i++) x_out[i] = B->ef->lastX[i];
	if (pointStep_out) for (size_t i = 0; i < B->pts.size(); i++) pointStep_out[i] = B->pts[i]->step;
	return 0;
}

// the global state the non-default solver branches look at: frameID of the first frame of the window (0 = the gauge anchor is
// still there, EnergyFunctional.cpp:688, :835) and setting_solverMode / setting_solverModeDelta for marginalizePointsF
extern "C" void dso_ref_ba_set_solver_state(dso_oracle_ba* B, int haveFirstFrame, int margSolverMode, double margSolverModeDelta)
{
	B->calib();
	B->ef->frames[0]->frameID = haveFirstFrame ? 0 : 1000;
	B->margSolverMode = margSolverMode; B->margSolverModeDelta = margSolverModeDelta;
}

// ---- rest of N2: marginalizePointsF, calcLEnergyF_MT, calcMEnergyF (the reference's own code) ----
extern "C" double dso_ref_ba_calc_L_energy(dso_oracle_ba* B, int nthreads)
{
	(void)nthreads;      // the reference always goes through its 6-worker pool (chunks of 50 points)
	B->calib();
	return B->ef->calcLEnergyF_MT();                           // EnergyFunctional.cpp:411
}
extern "C" double dso_ref_ba_calc_M_energy(dso_oracle_ba* B)
{
	const int N = 8 * B->nF + CPARS;
	if ((int)B->ef->bM.size() != N) return 0.0;
	return B->ef->calcMEnergyF();                              // EnergyFunctional.cpp:334
}
extern "C" void dso_ref_ba_get_marg_prior(dso_oracle_ba* B, double* HM, double* bM)
{
	const int N = 8 * B->nF + CPARS;
	if ((int)B->ef->bM.size() != N) dso_ref_ba_set_priors(B, 0, 0, 0, 0, 0, 0);
	for (int r = 0; r < N; r++) { for (int c = 0; c < N; c++) HM[r * N + c] = B->ef->HM(r, c); bM[r] = B->ef->bM[r]; }
}
// destructive: marginalizePointsF removes the flagged points from the EnergyFunctional (call it last)
extern "C" void dso_ref_ba_marginalize_points(dso_oracle_ba* B, const unsigned char* flags, float priorMargFac, double margWeightFac,
                                              double* H_out, double* b_out, int* resInM, float* idepth_hessian_out)
{
	B->calib();
	const int N = 8 * B->nF + CPARS;
	if ((int)B->ef->bM.size() != N) dso_ref_ba_set_priors(B, 0, 0, 0, 0, 0, 0);
	MatXX HM0 = B->ef->HM; VecX bM0 = B->ef->bM;
	const float f0 = setting_idepthFixPriorMargFac, w0 = setting_margWeightFac;
	setting_idepthFixPriorMargFac = priorMargFac; setting_margWeightFac = (float)margWeightFac;
	for (size_t i = 0; i < B->pts.size(); i++)
		if (flags[i]) B->pts[i]->efPoint->stateFlag = EFPointStatus::PS_MARGINALIZE;
	const int res0 = B->ef->resInM;
	const int sm0 = setting_solverMode; const double sd0 = setting_solverModeDelta;
	setting_solverMode = B->margSolverMode; setting_solverModeDelta = B->margSolverModeDelta;
	B->ef->marginalizePointsF();                               // EnergyFunctional.cpp:642
	setting_solverMode = sm0; setting_solverModeDelta = sd0;
	setting_idepthFixPriorMargFac = f0; setting_margWeightFac = w0;
	if (resInM) *resInM = B->ef->resInM - res0;
	for (int r = 0; r < N; r++)
	{
		for (int c = 0; c < N; c++) if (H_out) H_out[r * N + c] = (B->ef->HM(r, c) - HM0(r, c)) / margWeightFac;
		if (b_out) b_out[r] = (B->ef->bM[r] - bM0[r]) / margWeightFac;
	}
	if (idepth_hessian_out) for (size_t i = 0; i < B->pts.size(); i++) idepth_hessian_out[i] = B->pts[i]->idepth_hessian;
}

// ---- N3: ImmaturePoint constructor and traceOn (FullSystem/ImmaturePoint.cpp, compiled unmodified) ----
static FrameHessian* frame_with_dI(int w, int h, const float* dI)
{
	FrameHessian* fh = new FrameHessian();
	fh->shell = new FrameShell(); fh->shell->id = 0;
	fh->dIp[0] = (Eigen::Vector3f*)const_cast<float*>(dI);     // borrowed for the duration of the call
	fh->dI = fh->dIp[0];
	(void)w; (void)h;
	return fh;
}
static void release_frame(FrameHessian* fh) { fh->dIp[0] = 0; fh->dI = 0; }   // the shell object leaks on purpose (test process; ~FrameHessian frees image arrays it never allocated here)

extern "C" void dso_ref_immature_init(int w, int h, const float* host_dI, int n, const int* u, const int* v,
                                      float* color8, float* weights8, float* gradH4, float* energyTH)
{
	set_calib(w, h, 1, 500, 500, w * 0.5f, h * 0.5f);
	CalibHessian HCalib;
	FrameHessian* host = frame_with_dI(w, h, host_dI);
	for (int k = 0; k < n; k++)
	{
		ImmaturePoint ip(u[k], v[k], host, 1, &HCalib);           // ImmaturePoint.cpp:25
		for (int i = 0; i < 8; i++) { color8[8 * k + i] = ip.color[i]; weights8[8 * k + i] = ip.weights[i]; }
		gradH4[4 * k] = ip.gradH(0, 0); gradH4[4 * k + 1] = ip.gradH(0, 1); gradH4[4 * k + 2] = ip.gradH(1, 0); gradH4[4 * k + 3] = ip.gradH(1, 1);
		energyTH[k] = ip.energyTH;
	}
	release_frame(host);
}
extern "C" void dso_ref_trace_on(int w, int h, const float* frame_dI, int n, const float* u, const float* v,
                                 const float* color8, const float* weights8, const float* gradH4, const float* energyTH,
                                 float* idepth_min, float* idepth_max, float* quality, int* status,
                                 const float* KRKi9, const float* Kt3, const float* aff2,
                                 float* lastTraceUV2, float* lastTracePixelInterval)
{
	set_calib(w, h, 1, 500, 500, w * 0.5f, h * 0.5f);
	CalibHessian HCalib;
	FrameHessian* frame = frame_with_dI(w, h, frame_dI);
	FrameHessian* host = frame_with_dI(w, h, frame_dI);
	host->shell->id = 1;
	Mat33f KRKi; Vec3f Kt; Vec2f aff;
	for (int r = 0; r < 3; r++) { for (int c = 0; c < 3; c++) KRKi(r, c) = KRKi9[r * 3 + c]; Kt[r] = Kt3[r]; }
	aff[0] = aff2[0]; aff[1] = aff2[1];
	for (int k = 0; k < n; k++)
	{
		ImmaturePoint* ip = raw_immature(host, u[k], v[k], 1, color8 + 8 * k, weights8 + 8 * k);
		ip->gradH(0, 0) = gradH4[4 * k]; ip->gradH(0, 1) = gradH4[4 * k + 1]; ip->gradH(1, 0) = gradH4[4 * k + 2]; ip->gradH(1, 1) = gradH4[4 * k + 3];
		ip->energyTH = energyTH[k]; ip->idepth_min = idepth_min[k]; ip->idepth_max = idepth_max[k]; ip->quality = quality[k];
		ip->lastTraceStatus = (ImmaturePointStatus)status[k];
		ip->lastTraceUV = Vec2f(lastTraceUV2[2 * k], lastTraceUV2[2 * k + 1]); ip->lastTracePixelInterval = lastTracePixelInterval[k];
		ip->traceOn(frame, KRKi, Kt, aff, &HCalib, false);        // ImmaturePoint.cpp:77
		idepth_min[k] = ip->idepth_min; idepth_max[k] = ip->idepth_max; quality[k] = ip->quality; status[k] = (int)ip->lastTraceStatus;
		lastTraceUV2[2 * k] = ip->lastTraceUV[0]; lastTraceUV2[2 * k + 1] = ip->lastTraceUV[1];
		lastTracePixelInterval[k] = ip->lastTracePixelInterval;
		free(ip);
	}
	release_frame(frame); release_frame(host);
}

// FullSystem::optimizeImmaturePoint (FullSystemOptPoint.cpp:51, compiled unmodified).  The FullSystem object is raw zeroed
// memory with only the members that function touches brought to life (its constructor starts threads and opens log files).
extern "C" void dso_ref_optimize_immature(int w, int h, float fx, float fy, float cx, float cy, int nT, const float* const* target_dI,
                                          const float* RTll9, const float* tTll3, const float* aff2,
                                          int n, const float* u, const float* v, const float* color8, const float* weights8,
                                          const float* energyTH, const float* idepth_min, const float* idepth_max, int minObs,
                                          int* result, float* idepth_out, int* res_state_out)
{
	set_calib(w, h, 1, fx, fy, cx, cy);
	FullSystem* fs = (FullSystem*)calloc(1, sizeof(FullSystem));
	new (&fs->Hcalib) CalibHessian();
	new (&fs->frameHessians) std::vector<FrameHessian*>();
	FrameHessian* host = frame_with_dI(w, h, target_dI[0]);
	host->idx = 0; host->frameID = 0;
	fs->frameHessians.push_back(host);
	host->targetPrecalc.resize(nT + 1);
	for (int t = 0; t < nT; t++)
	{
		FrameHessian* fh = frame_with_dI(w, h, target_dI[t]);
		fh->idx = t + 1; fh->frameID = t + 1; fh->shell->id = t + 1;
		fs->frameHessians.push_back(fh);
		FrameFramePrecalc& P = host->targetPrecalc[t + 1];
		for (int r = 0; r < 3; r++) { for (int c = 0; c < 3; c++) P.PRE_RTll(r, c) = RTll9[9 * t + r * 3 + c]; P.PRE_tTll[r] = tTll3[3 * t + r]; }
		P.PRE_aff_mode[0] = aff2[2 * t]; P.PRE_aff_mode[1] = aff2[2 * t + 1];
	}
	std::vector<ImmaturePointTemporaryResidual> tmp(nT + 1);
	for (int k = 0; k < n; k++)
	{
		ImmaturePoint* ip = raw_immature(host, u[k], v[k], 1, color8 + 8 * k, weights8 + 8 * k);
		ip->energyTH = energyTH[k]; ip->idepth_min = idepth_min[k]; ip->idepth_max = idepth_max[k];
		PointHessian* ph = fs->optimizeImmaturePoint(ip, minObs, tmp.data());
		result[k] = ph == 0 ? 0 : (ph == (PointHessian*)((long)(-1)) ? -1 : 1);
		idepth_out[k] = result[k] == 1 ? ph->idepth : NAN;
		for (int t = 0; t < nT; t++) res_state_out[k * nT + t] = (int)tmp[t].state_state;
	}
	for (FrameHessian* fh : fs->frameHessians) release_frame(fh);
}


// ---- N4: CoarseInitializer::calcResAndGS (FullSystem/CoarseInitializer.cpp:370-732, compiled unmodified) ----
extern "C" void dso_ref_init_make_K(float fx0, float fy0, float cx0, float cy0, int lvl, double* fxfycxcy4, double* Ki9)
{
	// makeK reads wG/hG/pyrLevelsUsed and the CalibHessian; any image size will do for the intrinsics
	set_calib(640, 480, lvl + 1, fx0, fy0, cx0, cy0);
	CalibHessian HCalib;
	CoarseInitializer ci(640, 480);
	ci.makeK(&HCalib);                                           // :1126
	fxfycxcy4[0] = ci.fx[lvl]; fxfycxcy4[1] = ci.fy[lvl]; fxfycxcy4[2] = ci.cx[lvl]; fxfycxcy4[3] = ci.cy[lvl];
	for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) Ki9[r * 3 + c] = ci.Ki[lvl](r, c);
}
extern "C" void dso_ref_init_calc_res_and_gs(int w0, int h0, int lvl, float fx0, float fy0, float cx0, float cy0,
                                             const float* colorRef, const float* colorNew,
                                             const double* refToNew12, double aff_a, double aff_b,
                                             int npts, const float* pu, const float* pv, const float* idepth_new, const unsigned char* isGood,
                                             const float* energy2, const float* outlierTH, const float* iR,
                                             float alphaW, float alphaK, float couplingWeight,
                                             float* H64, float* b8, float* Hsc64, float* bsc8, float* res3,
                                             unsigned char* isGood_new, float* energy_new2, float* maxstep_out, float* lastHessian_new,
                                             float* Jb10)
{
	set_calib(w0, h0, lvl + 1, fx0, fy0, cx0, cy0);
	CalibHessian HCalib;
	CoarseInitializer* ci = new CoarseInitializer(w0, h0);
	ci->makeK(&HCalib);
	FrameHessian* first = new FrameHessian(); FrameHessian* nw = new FrameHessian();
	first->dIp[lvl] = (Eigen::Vector3f*)colorRef; nw->dIp[lvl] = (Eigen::Vector3f*)colorNew;
	ci->firstFrame = first; ci->newFrame = nw;
	ci->alphaW = alphaW; ci->alphaK = alphaK; ci->couplingWeight = couplingWeight; ci->regWeight = 0.8f;
	Pnt* pts = new Pnt[npts];
	for (int i = 0; i < npts; i++)
	{
		Pnt& p = pts[i];
		memset((void*)&p, 0, sizeof(Pnt));
		p.u = pu[i]; p.v = pv[i]; p.idepth = p.idepth_new = idepth_new[i]; p.isGood = isGood[i] != 0;
		p.energy[0] = energy2[2 * i]; p.energy[1] = energy2[2 * i + 1];
		p.energy_new[0] = p.energy_new[1] = 0;
		p.iR = iR[i]; p.outlierTH = outlierTH[i]; p.my_type = 1;
	}
	ci->points[lvl] = pts; ci->numPoints[lvl] = npts;
	Mat88f H, Hsc; Vec8f b, bsc;
	Vec3f res = ci->calcResAndGS(lvl, H, b, Hsc, bsc, to_se3(refToNew12), AffLight(aff_a, aff_b), false);   // :370
	for (int r = 0; r < 8; r++)
	{
		for (int c = 0; c < 8; c++) { H64[r * 8 + c] = H(r, c); Hsc64[r * 8 + c] = Hsc(r, c); }
		b8[r] = b[r]; bsc8[r] = bsc[r];
	}
	for (int k = 0; k < 3; k++) res3[k] = res[k];
	for (int i = 0; i < npts; i++)
	{
		isGood_new[i] = pts[i].isGood_new; energy_new2[2 * i] = pts[i].energy_new[0]; energy_new2[2 * i + 1] = pts[i].energy_new[1];
		maxstep_out[i] = pts[i].maxstep; lastHessian_new[i] = pts[i].lastHessian_new;
		if (Jb10) for (int k = 0; k < 10; k++) Jb10[10 * i + k] = ci->JbBuffer_new[i][k];
	}
	ci->points[lvl] = 0; ci->numPoints[lvl] = 0;
	delete[] pts;
	first->dIp[lvl] = 0; nw->dIp[lvl] = 0;       // borrowed
	delete ci;                                    // frames leak on purpose (test process; ~FrameHessian needs more state)
}
