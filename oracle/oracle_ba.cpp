// TEST INFRASTRUCTURE ONLY -- CPU oracle (see dso_oracle.h). PARITY UNPINNED (no reference fixtures exist).
// Restates the windowed-BA part of the hot path: FrameFramePrecalc::set, setAdjointsF/setDeltaF,
// PointFrameResidual::linearize, applyRes/takeDataF, AccumulatedTopHessianSSE::addPoint<0> + stitch,
// AccumulatedSCHessianSSE::addPoint + stitch, with the reference's three-tier float accumulators.
// Compile with -ffp-contract=off.
#include "dso_oracle.h"
#include "oracle_math.h"
#include "oracle_settings.h"
#include <vector>
#include <algorithm>
#include <thread>
#include <mutex>
#include <condition_variable>
#include <atomic>
#include <functional>
#include <cstdio>

using namespace orc;

#define CPARS 4
#define patternNum 8
// settings.cpp:274 (pattern 8) / settings.h:215-218
static const int patternP[8][2] = { {0,-2}, {-1,-1}, {1,-1}, {-2,0}, {0,0}, {2,0}, {-1,1}, {0,2} };
// HessianBlocks.h:54-61
#define SCALE_IDEPTH 1.0f
#define SCALE_XI_ROT 1.0f
#define SCALE_XI_TRANS 0.5f
#define SCALE_F 50.0f
#define SCALE_C 50.0f
#define SCALE_A 10.0f
#define SCALE_B 1000.0f
// settings.cpp:103,110,124-125
// setting_outlierTHSumComponent, setting_huberTH: run-time values (oracle_settings.h)
extern float dso_oracle_setting_affineOptModeA, dso_oracle_setting_affineOptModeB;   // oracle_tracker.cpp (settings.cpp:119-120)
#define setting_affineOptModeA dso_oracle_setting_affineOptModeA
#define setting_affineOptModeB dso_oracle_setting_affineOptModeB

enum { ST_IN = 0, ST_OOB = 1, ST_OUTLIER = 2 };

// RawResidualJacobian, RawResidualJacobian.h:40-79, flattened to 74 floats
struct RawJ
{
	float resF[8];
	float Jpdxi[2][6];
	float Jpdc[2][4];
	float Jpdd[2];
	float JIdx[2][8];
	float JabF[2][8];
	float JIdx2[4];   // row-major 2x2
	float JabJIdx[4];
	float Jab2[4];
};
static_assert(sizeof(RawJ) == 74 * 4, "RawJ must be 74 packed floats");

// ---- accumulators ----------------------------------------------------------------------------
// AccumulatorApprox, MatrixAccumulators.h:579-956
struct AccApprox
{
	float Data[60], Data1k[60], Data1m[60];
	float TR[32], TR1k[32], TR1m[32];
	float BR[8], BR1k[8], BR1m[8];
	float numIn1, numIn1k, numIn1m;
	size_t num;
	float H[13][13];
	void initialize()
	{
		memset(Data, 0, sizeof(Data)); memset(Data1k, 0, sizeof(Data1k)); memset(Data1m, 0, sizeof(Data1m));
		memset(TR, 0, sizeof(TR)); memset(TR1k, 0, sizeof(TR1k)); memset(TR1m, 0, sizeof(TR1m));
		memset(BR, 0, sizeof(BR)); memset(BR1k, 0, sizeof(BR1k)); memset(BR1m, 0, sizeof(BR1m));
		num = 0; numIn1 = numIn1k = numIn1m = 0;
	}
	void shiftUp(bool force)
	{
		if (numIn1 > 1000 || force)
		{
			for (int i = 0; i < 60; i++) Data1k[i] = Data[i] + Data1k[i];
			for (int i = 0; i < 32; i++) TR1k[i] = TR[i] + TR1k[i];
			for (int i = 0; i < 8; i++) BR1k[i] = BR[i] + BR1k[i];
			numIn1k += numIn1; numIn1 = 0;
			memset(Data, 0, sizeof(Data)); memset(TR, 0, sizeof(TR)); memset(BR, 0, sizeof(BR));
		}
		if (numIn1k > 1000 || force)
		{
			for (int i = 0; i < 60; i++) Data1m[i] = Data1k[i] + Data1m[i];
			for (int i = 0; i < 32; i++) TR1m[i] = TR1k[i] + TR1m[i];
			for (int i = 0; i < 8; i++) BR1m[i] = BR1k[i] + BR1m[i];
			numIn1m += numIn1k; numIn1k = 0;
			memset(Data1k, 0, sizeof(Data1k)); memset(TR1k, 0, sizeof(TR1k)); memset(BR1k, 0, sizeof(BR1k));
		}
	}
	// update(x4,x6,y4,y6,a,b,c), :738-842
	void update(const float* x4, const float* x6, const float* y4, const float* y6, float a, float b, float c)
	{
		float x[10], y[10];
		for (int i = 0; i < 4; i++) { x[i] = x4[i]; y[i] = y4[i]; }
		for (int i = 0; i < 6; i++) { x[4 + i] = x6[i]; y[4 + i] = y6[i]; }
		int idx = 0;
		for (int r = 0; r < 10; r++)
			for (int cc = r; cc < 10; cc++)
			{
				Data[idx] += a * x[cc] * x[r] + c * y[cc] * y[r] + b * (x[cc] * y[r] + y[cc] * x[r]);
				idx++;
			}
		num++; numIn1++;
		shiftUp(false);
	}
	// updateTopRight, :845-897
	void updateTopRight(const float* x4, const float* x6, const float* y4, const float* y6,
	                    float TR00, float TR10, float TR01, float TR11, float TR02, float TR12)
	{
		for (int i = 0; i < 10; i++)
		{
			float xi = i < 4 ? x4[i] : x6[i - 4];
			float yi = i < 4 ? y4[i] : y6[i - 4];
			TR[3 * i + 0] += xi * TR00 + yi * TR10;
			TR[3 * i + 1] += xi * TR01 + yi * TR11;
			TR[3 * i + 2] += xi * TR02 + yi * TR12;
		}
	}
	// updateBotRight, :899-913
	void updateBotRight(float a00, float a01, float a02, float a11, float a12, float a22)
	{
		BR[0] += a00; BR[1] += a01; BR[2] += a02; BR[3] += a11; BR[4] += a12; BR[5] += a22;
	}
	// finish, :605-640
	void finish()
	{
		memset(H, 0, sizeof(H));
		shiftUp(true);
		int idx = 0;
		for (int r = 0; r < 10; r++) for (int c = r; c < 10; c++) { H[r][c] = H[c][r] = Data1m[idx]; idx++; }
		idx = 0;
		for (int r = 0; r < 10; r++) for (int c = 0; c < 3; c++) { H[r][c + 10] = H[c + 10][r] = TR1m[idx]; idx++; }
		H[10][10] = BR1m[0];
		H[10][11] = H[11][10] = BR1m[1];
		H[10][12] = H[12][10] = BR1m[2];
		H[11][11] = BR1m[3];
		H[11][12] = H[12][11] = BR1m[4];
		H[12][12] = BR1m[5];
		num = numIn1 + numIn1k + numIn1m;
	}
};

// AccumulatorXX<i,j> / AccumulatorX<i>, MatrixAccumulators.h:34-87, 172-232
template <int I, int Jn>
struct AccXX
{
	float A[I][Jn], A1k[I][Jn], A1m[I][Jn];
	float numIn1, numIn1k, numIn1m;
	size_t num;
	void initialize() { memset(A, 0, sizeof(A)); memset(A1k, 0, sizeof(A1k)); memset(A1m, 0, sizeof(A1m)); num = 0; numIn1 = numIn1k = numIn1m = 0; }
	void shiftUp(bool force)
	{
		if (numIn1 > 1000 || force)
		{
			for (int i = 0; i < I; i++) for (int j = 0; j < Jn; j++) { A1k[i][j] += A[i][j]; A[i][j] = 0; }
			numIn1k += numIn1; numIn1 = 0;
		}
		if (numIn1k > 1000 || force)
		{
			for (int i = 0; i < I; i++) for (int j = 0; j < Jn; j++) { A1m[i][j] += A1k[i][j]; A1k[i][j] = 0; }
			numIn1m += numIn1k; numIn1k = 0;
		}
	}
	void finish() { shiftUp(true); num = numIn1 + numIn1k + numIn1m; }
	// A += w*L*R^T : outer product of the evaluated (w*L) with R
	void update(const float* L, const float* R, float w)
	{
		float wL[I];
		for (int i = 0; i < I; i++) wL[i] = w * L[i];
		for (int j = 0; j < Jn; j++) for (int i = 0; i < I; i++) A[i][j] += R[j] * wL[i];
		numIn1++;
		shiftUp(false);
	}
};
template <int I>
struct AccX
{
	float A[I], A1k[I], A1m[I];
	float numIn1, numIn1k, numIn1m;
	size_t num;
	void initialize() { memset(A, 0, sizeof(A)); memset(A1k, 0, sizeof(A1k)); memset(A1m, 0, sizeof(A1m)); num = 0; numIn1 = numIn1k = numIn1m = 0; }
	void shiftUp(bool force)
	{
		if (numIn1 > 1000 || force) { for (int i = 0; i < I; i++) { A1k[i] += A[i]; A[i] = 0; } numIn1k += numIn1; numIn1 = 0; }
		if (numIn1k > 1000 || force) { for (int i = 0; i < I; i++) { A1m[i] += A1k[i]; A1k[i] = 0; } numIn1m += numIn1k; numIn1k = 0; }
	}
	void finish() { shiftUp(true); num = numIn1 + numIn1k + numIn1m; }
	void update(const float* L, float w) { for (int i = 0; i < I; i++) A[i] += w * L[i]; numIn1++; shiftUp(false); }
};

// ---- a small persistent worker pool standing in for IndexThreadReduce (util/IndexThreadReduce.h) ----
// Workers take chunks of `step` indices from [first,end) until the range is exhausted.
class Pool
{
public:
	explicit Pool(int n) : nthreads(n), stop(false), generation(0), pending(0)
	{
		for (int i = 0; i < n; i++) workers.emplace_back([this, i] { loop(i); });
	}
	~Pool()
	{
		{ std::unique_lock<std::mutex> lk(m); stop = true; generation++; }
		cv.notify_all();
		for (auto& t : workers) t.join();
	}
	void reduce(const std::function<void(int, int, int)>& fn, int first, int end, int step)
	{
		if (step == 0) step = (end - first + nthreads - 1) / nthreads;
		if (step <= 0) step = 1;
		{
			std::unique_lock<std::mutex> lk(m);
			job = fn; next = first; jend = end; jstep = step; pending = nthreads; generation++;
		}
		cv.notify_all();
		std::unique_lock<std::mutex> lk(m);
		done.wait(lk, [this] { return pending == 0; });
	}
	int nthreads;
private:
	void loop(int tid)
	{
		unsigned long seen = 0;
		while (true)
		{
			{
				std::unique_lock<std::mutex> lk(m);
				cv.wait(lk, [&] { return generation != seen; });
				seen = generation;
				if (stop) return;
			}
			while (true)
			{
				int lo;
				{
					std::unique_lock<std::mutex> lk(m);   // the reference also hands out chunks under a mutex
					lo = next; next += jstep;
				}
				if (lo >= jend) break;
				job(lo, std::min(lo + jstep, jend), tid);
			}
			{
				std::unique_lock<std::mutex> lk(m);
				if (--pending == 0) done.notify_all();
			}
		}
	}
	std::vector<std::thread> workers;
	std::mutex m;
	std::condition_variable cv, done;
	bool stop;
	unsigned long generation;
	int pending;
	std::function<void(int, int, int)> job;
	int next, jend, jstep;
};

// ---- state -------------------------------------------------------------------------------------
struct Frame
{
	SE3d evalPT, PRE_worldToCam, PRE_camToWorld;
	double state[10], state_zero[10], state_scaled[10];
	float ab_exposure, frameEnergyTH;
	const float* dI;
};
struct Precalc
{
	float RTll[9], tTll[3], RTll_0[9], tTll_0[3], KRKiTll[9], RKiTll[9], KtTll[3], aff_mode[2], b0_mode, distanceLL;
};
struct Pt
{
	float u, v, idepth_scaled, idepth_zero_scaled, color[8], weights[8];
	int host;
	float priorF, deltaF;
	std::vector<int> residualsAll;
	float Hdd_accAF, bd_accAF, Hcd_accAF[4];
	float Hdd_accLF, bd_accLF, Hcd_accLF[4];
	float HdiF, bdSumF, idepth_hessian;
};
struct Res
{
	int point, host, target;
	int state_state, state_NewState;
	double state_energy, state_NewEnergy, state_NewEnergyWithOutlier;
	RawJ J;      // PointFrameResidual::J
	RawJ efJ;    // EFResidual::J
	bool isLinearized, isActive;
	float JpJdF[8];
	float res_toZeroF[8];
	float centerProjectedTo[3];
	float projectedTo[8][2];
};

struct dso_oracle_ba
{
	int w, h, nF;
	float fxl, fyl, cxl, cyl, fxli, fyli, cxli, cyli;   // CalibHessian::value_scaledf / value_scaledi
	float wM3G, hM3G;
	double calibDelta[4];
	std::vector<Frame> frames;
	std::vector<Precalc> pre;                // [h*nF+t]
	std::vector<double> adHost, adTarget;    // [ (h+t*nF)*64 ]
	std::vector<float> adHostF, adTargetF, adHTdeltaF;
	float cDeltaF[4];
	std::vector<Pt> points;
	std::vector<Res> res;
	// priors and marginalisation state (EnergyFunctional members cPrior, frames[i]->prior/delta_prior/delta, HM, bM)
	double cPrior[4];
	std::vector<double> framePrior, frameDeltaPrior, frameDelta, HM, bM;
	std::vector<double> nullspaces; int nNullspaces;     // lastNullspaces_pose then _scale, each 8nF+4 long
	int haveFirstFrame;                                  // "for(EFFrame* f : frames) if(f->frameID==0)" (EnergyFunctional.cpp:688, :835)
	int margSolverMode; double margSolverModeDelta;      // setting_solverMode / setting_solverModeDelta seen by marginalizePointsF
	std::vector<float> pointStep;
	Pool* pool; int poolN;
	dso_oracle_ba() : nNullspaces(0), haveFirstFrame(1), margSolverMode(128 | 2048), margSolverModeDelta(1e-5), pool(0), poolN(0) { for (int i = 0; i < 4; i++) cPrior[i] = 0; }
	Pool* getPool(int n) { if (!pool || poolN != n) { delete pool; pool = new Pool(n); poolN = n; } return pool; }
};

extern "C" dso_oracle_ba* dso_oracle_ba_create(int w, int h, int nF, float fx, float fy, float cx, float cy)
{
	dso_oracle_ba* B = new dso_oracle_ba();
	B->w = w; B->h = h; B->nF = nF;
	// CalibHessian::setValueScaled, HessianBlocks.h:383-397
	B->fxl = fx; B->fyl = fy; B->cxl = cx; B->cyl = cy;
	B->fxli = 1.0f / fx; B->fyli = 1.0f / fy; B->cxli = -cx / fx; B->cyli = -cy / fy;
	B->wM3G = w - 3; B->hM3G = h - 3;   // globalCalib.cpp:70-71
	B->frames.resize(nF);
	for (int i = 0; i < 4; i++) B->calibDelta[i] = 0;
	return B;
}
extern "C" void dso_oracle_ba_destroy(dso_oracle_ba* B) { delete B->pool; delete B; }

extern "C" void dso_oracle_ba_set_frame(dso_oracle_ba* B, int idx, const double* evalPT12, const double* state10,
                                        const double* state_zero10, float ab_exposure, float frameEnergyTH, const float* dI)
{
	Frame& f = B->frames[idx];
	f.evalPT = SE3d::from12(evalPT12);
	memcpy(f.state, state10, 80); memcpy(f.state_zero, state_zero10, 80);
	// FrameHessian::setState, HessianBlocks.h:183-199
	for (int i = 0; i < 3; i++) f.state_scaled[i] = SCALE_XI_TRANS * f.state[i];
	for (int i = 3; i < 6; i++) f.state_scaled[i] = SCALE_XI_ROT * f.state[i];
	f.state_scaled[6] = SCALE_A * f.state[6]; f.state_scaled[7] = SCALE_B * f.state[7];
	f.state_scaled[8] = SCALE_A * f.state[8]; f.state_scaled[9] = SCALE_B * f.state[9];
	f.PRE_worldToCam = se3_exp(f.state_scaled) * f.evalPT;
	f.PRE_camToWorld = se3_inverse(f.PRE_worldToCam);
	f.ab_exposure = ab_exposure; f.frameEnergyTH = frameEnergyTH; f.dI = dI;
}
extern "C" void dso_oracle_ba_set_calib_delta(dso_oracle_ba* B, const double* d) { memcpy(B->calibDelta, d, 32); }

static void fromToVecExposure(float exposureF, float exposureT, double Fa, double Fb, double Ta, double Tb, double* out)
{
	if (exposureF == 0 || exposureT == 0) exposureT = exposureF = 1;
	double a = exp(Ta - Fa) * exposureT / exposureF;
	double b = Tb - a * Fb;
	out[0] = a; out[1] = b;
}

extern "C" void dso_oracle_ba_prepare(dso_oracle_ba* B)
{
	int nF = B->nF;
	B->pre.assign(nF * nF, Precalc());
	B->adHost.assign(nF * nF * 64, 0); B->adTarget.assign(nF * nF * 64, 0);
	B->adHostF.assign(nF * nF * 64, 0); B->adTargetF.assign(nF * nF * 64, 0);
	B->adHTdeltaF.assign(nF * nF * 8, 0);
	float K[9] = { B->fxl, 0, B->cxl, 0, B->fyl, B->cyl, 0, 0, 1 };
	float Kinv[9];
	mat33f_inverse(K, Kinv);
	for (int h = 0; h < nF; h++)
		for (int t = 0; t < nF; t++)
		{
			const Frame& host = B->frames[h]; const Frame& target = B->frames[t];
			// FrameFramePrecalc::set, HessianBlocks.cpp:213-250
			Precalc& P = B->pre[h * nF + t];
			SE3d leftToLeft_0 = target.evalPT * se3_inverse(host.evalPT);
			for (int i = 0; i < 9; i++) P.RTll_0[i] = (float)leftToLeft_0.R[i];
			for (int i = 0; i < 3; i++) P.tTll_0[i] = (float)leftToLeft_0.t[i];
			SE3d leftToLeft = target.PRE_worldToCam * host.PRE_camToWorld;
			for (int i = 0; i < 9; i++) P.RTll[i] = (float)leftToLeft.R[i];
			for (int i = 0; i < 3; i++) P.tTll[i] = (float)leftToLeft.t[i];
			P.distanceLL = (float)std::sqrt(leftToLeft.t[0] * leftToLeft.t[0] + leftToLeft.t[1] * leftToLeft.t[1] + leftToLeft.t[2] * leftToLeft.t[2]);
			float KR[9];
			mat33f_mul(K, P.RTll, KR);
			mat33f_mul(KR, Kinv, P.KRKiTll);
			mat33f_mul(P.RTll, Kinv, P.RKiTll);
			mat33f_vec(K, P.tTll, P.KtTll);
			double aff[2];
			fromToVecExposure(host.ab_exposure, target.ab_exposure, host.state_scaled[6], host.state_scaled[7],
			                  target.state_scaled[6], target.state_scaled[7], aff);
			P.aff_mode[0] = (float)aff[0]; P.aff_mode[1] = (float)aff[1];
			P.b0_mode = host.state_zero[7] * SCALE_B;   // aff_g2l_0().b

			// setAdjointsF, EnergyFunctional.cpp:57-94
			SE3d hostToTarget = leftToLeft_0;
			double Ad[36];
			se3_adj(hostToTarget, Ad);
			double* AH = &B->adHost[(h + t * nF) * 64];
			double* AT = &B->adTarget[(h + t * nF) * 64];
			for (int i = 0; i < 64; i++) { AH[i] = (i % 9 == 0); AT[i] = (i % 9 == 0); }
			for (int r = 0; r < 6; r++) for (int c = 0; c < 6; c++) AH[r * 8 + c] = -Ad[c * 6 + r];
			double a0[2];
			fromToVecExposure(host.ab_exposure, target.ab_exposure, host.state_zero[6] * SCALE_A, host.state_zero[7] * SCALE_B,
			                  target.state_zero[6] * SCALE_A, target.state_zero[7] * SCALE_B, a0);
			float affLL0 = (float)a0[0];
			AT[6 * 8 + 6] = -affLL0;
			AH[6 * 8 + 6] = affLL0;
			AT[7 * 8 + 7] = -1;
			AH[7 * 8 + 7] = affLL0;
			for (int c = 0; c < 8; c++)
			{
				for (int r = 0; r < 3; r++) { AH[r * 8 + c] *= SCALE_XI_TRANS; AT[r * 8 + c] *= SCALE_XI_TRANS; }
				for (int r = 3; r < 6; r++) { AH[r * 8 + c] *= SCALE_XI_ROT; AT[r * 8 + c] *= SCALE_XI_ROT; }
				AH[6 * 8 + c] *= SCALE_A; AT[6 * 8 + c] *= SCALE_A;
				AH[7 * 8 + c] *= SCALE_B; AT[7 * 8 + c] *= SCALE_B;
			}
			for (int i = 0; i < 64; i++) { B->adHostF[(h + t * nF) * 64 + i] = (float)AH[i]; B->adTargetF[(h + t * nF) * 64 + i] = (float)AT[i]; }
		}
	// setDeltaF, EnergyFunctional.cpp:178-193 (row vector times 8x8, float; each coefficient an unrolled 8-term redux)
	for (int h = 0; h < nF; h++)
		for (int t = 0; t < nF; t++)
		{
			int idx = h + t * nF;
			for (int c = 0; c < 8; c++)
			{
				float p0[8], p1[8];
				for (int r = 0; r < 8; r++)
				{
					p0[r] = (float)(B->frames[h].state[r] - B->frames[h].state_zero[r]) * B->adHostF[idx * 64 + r * 8 + c];
					p1[r] = (float)(B->frames[t].state[r] - B->frames[t].state_zero[r]) * B->adTargetF[idx * 64 + r * 8 + c];
				}
				const float s0 = ((p0[0] + p0[1]) + (p0[2] + p0[3])) + ((p0[4] + p0[5]) + (p0[6] + p0[7]));
				const float s1 = ((p1[0] + p1[1]) + (p1[2] + p1[3])) + ((p1[4] + p1[5]) + (p1[6] + p1[7]));
				B->adHTdeltaF[idx * 8 + c] = s0 + s1;
			}
		}
	for (int i = 0; i < 4; i++) B->cDeltaF[i] = (float)B->calibDelta[i];
}

extern "C" void dso_oracle_ba_get_glue(const dso_oracle_ba* B, float* precalc, double* adHost, double* adTarget,
                                       float* adHTdeltaF, float* cDeltaF4)
{
	int nF = B->nF;
	if (precalc)
		for (int i = 0; i < nF * nF; i++)
		{
			const Precalc& P = B->pre[i];
			float* o = precalc + 28 * i;
			memcpy(o, P.RTll_0, 36); memcpy(o + 9, P.tTll_0, 12); memcpy(o + 12, P.KRKiTll, 36); memcpy(o + 21, P.KtTll, 12);
			o[24] = P.aff_mode[0]; o[25] = P.aff_mode[1]; o[26] = P.b0_mode; o[27] = 0;
		}
	if (adHost) memcpy(adHost, B->adHost.data(), sizeof(double) * nF * nF * 64);
	if (adTarget) memcpy(adTarget, B->adTarget.data(), sizeof(double) * nF * nF * 64);
	if (adHTdeltaF) memcpy(adHTdeltaF, B->adHTdeltaF.data(), sizeof(float) * nF * nF * 8);
	if (cDeltaF4) memcpy(cDeltaF4, B->cDeltaF, 16);
}

extern "C" void dso_oracle_ba_set_points(dso_oracle_ba* B, int nP, const float* u, const float* v, const float* idepth_scaled,
                                         const float* idepth_zero_scaled, const float* color8, const float* weights8,
                                         const int* host, const float* priorF, const float* deltaF)
{
	B->points.assign(nP, Pt());
	for (int i = 0; i < nP; i++)
	{
		Pt& p = B->points[i];
		p.u = u[i]; p.v = v[i]; p.idepth_scaled = idepth_scaled[i]; p.idepth_zero_scaled = idepth_zero_scaled[i];
		memcpy(p.color, color8 + 8 * i, 32); memcpy(p.weights, weights8 + 8 * i, 32);
		p.host = host[i]; p.priorF = priorF[i]; p.deltaF = deltaF[i];
		p.Hdd_accAF = p.bd_accAF = 0; p.Hdd_accLF = p.bd_accLF = 0;
		for (int k = 0; k < 4; k++) p.Hcd_accAF[k] = p.Hcd_accLF[k] = 0;
		p.HdiF = p.bdSumF = p.idepth_hessian = 0;
	}
}

extern "C" void dso_oracle_ba_set_residuals(dso_oracle_ba* B, int nR, const int* point, const int* target, const int* state_state,
                                            const double* state_energy, const unsigned char* isLinearized, const unsigned char* isActive)
{
	B->res.assign(nR, Res());
	for (auto& p : B->points) p.residualsAll.clear();
	for (int i = 0; i < nR; i++)
	{
		Res& r = B->res[i];
		memset(&r, 0, sizeof(Res));
		r.point = point[i]; r.target = target[i]; r.host = B->points[point[i]].host;
		r.state_state = state_state[i];
		r.state_NewState = ST_OUTLIER;   // resetOOB, Residuals.h:97-105
		r.state_energy = state_energy ? state_energy[i] : 0;
		r.state_NewEnergy = r.state_energy;
		r.isLinearized = isLinearized ? isLinearized[i] != 0 : false;
		r.isActive = isActive ? isActive[i] != 0 : false;
		B->points[point[i]].residualsAll.push_back(i);
	}
}

// getInterpolatedElement33, globalFuncs.h:75-89
static inline void interp33(const float* mat, float x, float y, int width, float* out)
{
	int ix = (int)x;
	int iy = (int)y;
	float dx = x - ix;
	float dy = y - iy;
	float dxdy = dx * dy;
	const float* bp = mat + 3 * (ix + iy * width);
	float w11 = dxdy, w01 = dy - dxdy, w10 = dx - dxdy, w00 = 1 - dx - dy + dxdy;
	for (int c = 0; c < 3; c++)
		out[c] = w11 * bp[3 * (1 + width) + c] + w01 * bp[3 * width + c] + w10 * bp[3 + c] + w00 * bp[c];
}

// PointFrameResidual::linearize, Residuals.cpp:78-291
static double linearize(dso_oracle_ba* B, Res& R)
{
	R.state_NewEnergyWithOutlier = -1;
	if (R.state_state == ST_OOB) { R.state_NewState = ST_OOB; return R.state_energy; }
	const Pt& point = B->points[R.point];
	const Frame& host = B->frames[R.host];
	const Frame& target = B->frames[R.target];
	const Precalc* precalc = &B->pre[R.host * B->nF + R.target];
	float energyLeft = 0;
	const float* dIl = target.dI;
	const float* PRE_KRKiTll = precalc->KRKiTll;
	const float* PRE_KtTll = precalc->KtTll;
	const float* PRE_RTll_0 = precalc->RTll_0;
	const float* PRE_tTll_0 = precalc->tTll_0;
	const float* color = point.color;
	const float* weights = point.weights;
	float affLL[2] = { precalc->aff_mode[0], precalc->aff_mode[1] };
	float b0 = precalc->b0_mode;
	const float fxl = B->fxl, fyl = B->fyl, cxl = B->cxl, cyl = B->cyl, fxli = B->fxli, fyli = B->fyli;
#define R0(r, c) PRE_RTll_0[(r) * 3 + (c)]

	float d_xi_x[6], d_xi_y[6], d_C_x[4], d_C_y[4], d_d_x, d_d_y;
	{
		float drescale, u, v, new_idepth, Ku, Kv, KliP[3];
		// projectPoint (8-arg form), ResidualProjections.h:69-101, dx = dy = 0
		KliP[0] = (point.u + 0 - cxl) * fxli;
		KliP[1] = (point.v + 0 - cyl) * fyli;
		KliP[2] = 1;
		float ptp[3];
		for (int r = 0; r < 3; r++)
			ptp[r] = dot3_eigen(R0(r, 0), KliP[0], R0(r, 1), KliP[1], R0(r, 2), KliP[2]) + PRE_tTll_0[r] * point.idepth_zero_scaled;
		drescale = 1.0f / ptp[2];
		new_idepth = point.idepth_zero_scaled * drescale;
		bool ok = true;
		if (!(drescale > 0)) ok = false;
		if (ok)
		{
			u = ptp[0] * drescale;
			v = ptp[1] * drescale;
			Ku = u * fxl + cxl;
			Kv = v * fyl + cyl;
			ok = Ku > 1.1f && Kv > 1.1f && Ku < B->wM3G && Kv < B->hM3G;
		}
		if (!ok) { R.state_NewState = ST_OOB; return R.state_energy; }

		R.centerProjectedTo[0] = Ku; R.centerProjectedTo[1] = Kv; R.centerProjectedTo[2] = new_idepth;

		// :131-132
		d_d_x = drescale * (PRE_tTll_0[0] - PRE_tTll_0[2] * u) * SCALE_IDEPTH * fxl;
		d_d_y = drescale * (PRE_tTll_0[1] - PRE_tTll_0[2] * v) * SCALE_IDEPTH * fyl;
		// :138-156
		d_C_x[2] = drescale * (R0(2, 0) * u - R0(0, 0));
		d_C_x[3] = fxl * drescale * (R0(2, 1) * u - R0(0, 1)) * fyli;
		d_C_x[0] = KliP[0] * d_C_x[2];
		d_C_x[1] = KliP[1] * d_C_x[3];
		d_C_y[2] = fyl * drescale * (R0(2, 0) * v - R0(1, 0)) * fxli;
		d_C_y[3] = drescale * (R0(2, 1) * v - R0(1, 1));
		d_C_y[0] = KliP[0] * d_C_y[2];
		d_C_y[1] = KliP[1] * d_C_y[3];
		d_C_x[0] = (d_C_x[0] + u) * SCALE_F;
		d_C_x[1] *= SCALE_F;
		d_C_x[2] = (d_C_x[2] + 1) * SCALE_C;
		d_C_x[3] *= SCALE_C;
		d_C_y[0] *= SCALE_F;
		d_C_y[1] = (d_C_y[1] + v) * SCALE_F;
		d_C_y[2] *= SCALE_C;
		d_C_y[3] = (d_C_y[3] + 1) * SCALE_C;
		// :159-171
		d_xi_x[0] = new_idepth * fxl;
		d_xi_x[1] = 0;
		d_xi_x[2] = -new_idepth * u * fxl;
		d_xi_x[3] = -u * v * fxl;
		d_xi_x[4] = (1 + u * u) * fxl;
		d_xi_x[5] = -v * fxl;
		d_xi_y[0] = 0;
		d_xi_y[1] = new_idepth * fyl;
		d_xi_y[2] = -new_idepth * v * fyl;
		d_xi_y[3] = -(1 + v * v) * fyl;
		d_xi_y[4] = u * v * fyl;
		d_xi_y[5] = u * fyl;
	}
	RawJ* J = &R.J;
	for (int i = 0; i < 6; i++) { J->Jpdxi[0][i] = d_xi_x[i]; J->Jpdxi[1][i] = d_xi_y[i]; }
	for (int i = 0; i < 4; i++) { J->Jpdc[0][i] = d_C_x[i]; J->Jpdc[1][i] = d_C_y[i]; }
	J->Jpdd[0] = d_d_x; J->Jpdd[1] = d_d_y;

	float JIdxJIdx_00 = 0, JIdxJIdx_11 = 0, JIdxJIdx_10 = 0;
	float JabJIdx_00 = 0, JabJIdx_01 = 0, JabJIdx_10 = 0, JabJIdx_11 = 0;
	float JabJab_00 = 0, JabJab_01 = 0, JabJab_11 = 0;
	float wJI2_sum = 0;

	for (int idx = 0; idx < patternNum; idx++)
	{
		float Ku, Kv;
		// projectPoint (7-arg form), ResidualProjections.h:55-65
		{
			float up = point.u + patternP[idx][0], vp = point.v + patternP[idx][1];
			float ptp[3];
			for (int r = 0; r < 3; r++)
				ptp[r] = dot3_eigen(PRE_KRKiTll[r * 3], up, PRE_KRKiTll[r * 3 + 1], vp, PRE_KRKiTll[r * 3 + 2], 1.0f) + PRE_KtTll[r] * point.idepth_scaled;
			Ku = ptp[0] / ptp[2];
			Kv = ptp[1] / ptp[2];
			if (!(Ku > 1.1f && Kv > 1.1f && Ku < B->wM3G && Kv < B->hM3G))
			{ R.state_NewState = ST_OOB; return R.state_energy; }
		}
		R.projectedTo[idx][0] = Ku;
		R.projectedTo[idx][1] = Kv;

		float hitColor[3];
		interp33(dIl, Ku, Kv, B->w, hitColor);
		float residual = hitColor[0] - (float)(affLL[0] * color[idx] + affLL[1]);
		float drdA = (color[idx] - b0);
		if (!std::isfinite((float)hitColor[0]))
		{ R.state_NewState = ST_OOB; return R.state_energy; }

		float w = sqrtf(setting_outlierTHSumComponent / (setting_outlierTHSumComponent + (hitColor[1] * hitColor[1] + hitColor[2] * hitColor[2])));
		w = 0.5f * (w + weights[idx]);
		float hw = fabsf(residual) < setting_huberTH ? 1 : setting_huberTH / fabsf(residual);
		energyLeft += w * w * hw * residual * residual * (2 - hw);
		{
			if (hw < 1) hw = sqrtf(hw);
			hw = hw * w;
			hitColor[1] *= hw;
			hitColor[2] *= hw;
			J->resF[idx] = residual * hw;
			J->JIdx[0][idx] = hitColor[1];
			J->JIdx[1][idx] = hitColor[2];
			J->JabF[0][idx] = drdA * hw;
			J->JabF[1][idx] = hw;
			JIdxJIdx_00 += hitColor[1] * hitColor[1];
			JIdxJIdx_11 += hitColor[2] * hitColor[2];
			JIdxJIdx_10 += hitColor[1] * hitColor[2];
			JabJIdx_00 += drdA * hw * hitColor[1];
			JabJIdx_01 += drdA * hw * hitColor[2];
			JabJIdx_10 += hw * hitColor[1];
			JabJIdx_11 += hw * hitColor[2];
			JabJab_00 += drdA * drdA * hw * hw;
			JabJab_01 += drdA * hw * hw;
			JabJab_11 += hw * hw;
			wJI2_sum += hw * hw * (hitColor[1] * hitColor[1] + hitColor[2] * hitColor[2]);
			if (setting_affineOptModeA < 0) J->JabF[0][idx] = 0;
			if (setting_affineOptModeB < 0) J->JabF[1][idx] = 0;
		}
	}
	J->JIdx2[0] = JIdxJIdx_00; J->JIdx2[1] = JIdxJIdx_10; J->JIdx2[2] = JIdxJIdx_10; J->JIdx2[3] = JIdxJIdx_11;
	J->JabJIdx[0] = JabJIdx_00; J->JabJIdx[1] = JabJIdx_01; J->JabJIdx[2] = JabJIdx_10; J->JabJIdx[3] = JabJIdx_11;
	J->Jab2[0] = JabJab_00; J->Jab2[1] = JabJab_01; J->Jab2[2] = JabJab_01; J->Jab2[3] = JabJab_11;

	R.state_NewEnergyWithOutlier = energyLeft;
	float th = std::max<float>(host.frameEnergyTH, target.frameEnergyTH);
	if (energyLeft > th || wJI2_sum < 2)
	{
		energyLeft = th;
		R.state_NewState = ST_OUTLIER;
	}
	else
		R.state_NewState = ST_IN;
	R.state_NewEnergy = energyLeft;
	return energyLeft;
#undef R0
}

static void export_linearize(dso_oracle_ba* B, int* newState, float* newEnergy, float* newEnergyWithOutlier,
                             float* cpt, float* proj, float* J74)
{
	int nR = (int)B->res.size();
	for (int i = 0; i < nR; i++)
	{
		const Res& r = B->res[i];
		if (newState) newState[i] = r.state_NewState;
		if (newEnergy) newEnergy[i] = (float)r.state_NewEnergy;
		if (newEnergyWithOutlier) newEnergyWithOutlier[i] = (float)r.state_NewEnergyWithOutlier;
		if (cpt) memcpy(cpt + 3 * i, r.centerProjectedTo, 12);
		if (proj) memcpy(proj + 16 * i, r.projectedTo, 64);
		if (J74) memcpy(J74 + 74 * i, &r.J, sizeof(RawJ));
	}
}

extern "C" double dso_oracle_ba_linearize(dso_oracle_ba* B, int* newState, float* newEnergy, float* newEnergyWithOutlier,
                                          float* cpt, float* proj, float* J74, double* energy_ret)
{
	// linearizeAll_Reductor, FullSystemOptimize.cpp:52-60 (single-thread branch :171-176)
	double stats0 = 0;
	int nR = (int)B->res.size();
	for (int k = 0; k < nR; k++)
	{
		double e = linearize(B, B->res[k]);
		if (energy_ret) energy_ret[k] = e;
		stats0 += e;
	}
	export_linearize(B, newState, newEnergy, newEnergyWithOutlier, cpt, proj, J74);
	return stats0;
}

extern "C" double dso_oracle_ba_linearize_mt(dso_oracle_ba* B, int nthreads)
{
	// multi-thread branch, FullSystemOptimize.cpp:164: static split, step = ceil(n/NUM_THREADS)
	Pool* pool = B->getPool(nthreads);
	std::vector<double> stats(nthreads, 0.0);
	pool->reduce([&](int lo, int hi, int tid) { for (int k = lo; k < hi; k++) stats[tid] += linearize(B, B->res[k]); },
	             0, (int)B->res.size(), 0);
	double s = 0; for (double v : stats) s += v;
	return s;
}

// applyRes(true), Residuals.cpp:323-347 + EFResidual::takeDataF, EnergyFunctionalStructs.cpp:39-49
static void apply_one(Res& r);
extern "C" void dso_oracle_ba_apply_res(dso_oracle_ba* B, unsigned char* isActive_out, float* JpJdF_out)
{
	int nR = (int)B->res.size();
	for (int i = 0; i < nR; i++)
	{
		Res& r = B->res[i];
		apply_one(r);
		if (isActive_out) isActive_out[i] = r.isActive;
		if (JpJdF_out) memcpy(JpJdF_out + 8 * i, r.JpJdF, 32);
	}
}
// the per-object sequence of FullSystem::flagPointsForRemoval (FullSystem.cpp:776-787) for the listed residuals:
// resetOOB (Residuals.h:101-108), linearize (Residuals.cpp:78), applyRes(true) (Residuals.cpp:323)
extern "C" void dso_oracle_ba_linearize_residuals(dso_oracle_ba* B, int k, const int* idx, int reset_oob, int apply,
                                                  int* newState, float* newEnergy, unsigned char* isActive)
{
	for (int q = 0; q < k; q++)
	{
		Res& r = B->res[idx[q]];
		if (reset_oob) { r.state_NewEnergy = r.state_energy = 0; r.state_NewState = ST_OUTLIER; r.state_state = ST_IN; }
		linearize(B, r);
		if (apply) apply_one(r);
		if (newState) newState[q] = r.state_NewState;
		if (newEnergy) newEnergy[q] = (float)r.state_NewEnergy;
		if (isActive) isActive[q] = r.isActive;
	}
}
static void apply_one(Res& r)
{
	{
		{
		if (r.state_state != ST_OOB)
		{
			if (r.state_NewState == ST_IN)
			{
				r.isActive = true;
				std::swap(r.J, r.efJ);
				const RawJ* J = &r.efJ;
				float JI_JI_Jd[2];
				JI_JI_Jd[0] = J->JIdx2[0] * J->Jpdd[0] + J->JIdx2[1] * J->Jpdd[1];
				JI_JI_Jd[1] = J->JIdx2[2] * J->Jpdd[0] + J->JIdx2[3] * J->Jpdd[1];
				for (int k = 0; k < 6; k++)
					r.JpJdF[k] = J->Jpdxi[0][k] * JI_JI_Jd[0] + J->Jpdxi[1][k] * JI_JI_Jd[1];
				r.JpJdF[6] = J->JabJIdx[0] * J->Jpdd[0] + J->JabJIdx[1] * J->Jpdd[1];
				r.JpJdF[7] = J->JabJIdx[2] * J->Jpdd[0] + J->JabJIdx[3] * J->Jpdd[1];
			}
			else
				r.isActive = false;
			r.state_state = r.state_NewState;
			r.state_energy = r.state_NewEnergy;
		}
		}
	}
}

// Eigen's unrolled fixed-size redux (halves, then halves again): 6 terms (p0+(p1+p2))+(p3+(p4+p5)), 4 terms (p0+p1)+(p2+p3)
static inline float dot6_tree(const float* a, const float* b)
{
	return (a[0] * b[0] + (a[1] * b[1] + a[2] * b[2])) + (a[3] * b[3] + (a[4] * b[4] + a[5] * b[5]));
}
static inline float sum8_tree(const float* p) { return ((p[0] + p[1]) + (p[2] + p[3])) + ((p[4] + p[5]) + (p[6] + p[7])); }
static inline float dot4_tree(const float* a, const float* b) { return (a[0] * b[0] + a[1] * b[1]) + (a[2] * b[2] + a[3] * b[3]); }

// Jp*delta of fixLinearizationF / addPoint<1> (EnergyFunctionalStructs.cpp:92-100, AccumulatedTopHessian.cpp:82-87)
static inline void jp_delta(const dso_oracle_ba* B, const Res& r, const RawJ* J, float dd, float* out4)
{
	const float* dp = &B->adHTdeltaF[(r.host + B->nF * r.target) * 8];
	out4[0] = dot6_tree(J->Jpdxi[0], dp) + dot4_tree(J->Jpdc[0], B->cDeltaF) + J->Jpdd[0] * dd;
	out4[1] = dot6_tree(J->Jpdxi[1], dp) + dot4_tree(J->Jpdc[1], B->cDeltaF) + J->Jpdd[1] * dd;
	out4[2] = dp[6]; out4[3] = dp[7];
}

// AccumulatedTopHessianSSE::addPoint<mode>, AccumulatedTopHessian.cpp:39-177 (0 = active, 1 = linearised, 2 = marginalise)
static void addPointA(dso_oracle_ba* B, Pt& p, AccApprox* acc, int* nres, int mode = 0)
{
	int nF = B->nF;
	float bd_acc = 0, Hdd_acc = 0, Hcd_acc[4] = { 0, 0, 0, 0 };
	for (int ri : p.residualsAll)
	{
		Res& r = B->res[ri];
		if (mode == 0 && (r.isLinearized || !r.isActive)) continue;
		if (mode == 1 && (!r.isLinearized || !r.isActive)) continue;
		if (mode == 2 && !r.isActive) continue;
		const RawJ* rJ = &r.efJ;
		int htIDX = r.host + r.target * nF;
		float resLin[8];
		const float* resApprox = rJ->resF;
		if (mode == 2) resApprox = r.res_toZeroF;
		if (mode == 1)
		{
			float d[4];
			jp_delta(B, r, rJ, p.deltaF, d);
			for (int i = 0; i < patternNum; i++)
			{
				float rtz = r.res_toZeroF[i];              // :93-99, one mul + one add per term
				rtz = rtz + rJ->JIdx[0][i] * d[0];
				rtz = rtz + rJ->JIdx[1][i] * d[1];
				rtz = rtz + rJ->JabF[0][i] * d[2];
				rtz = rtz + rJ->JabF[1][i] * d[3];
				resLin[i] = rtz;
			}
			resApprox = resLin;
		}
		float JI_r[2] = { 0, 0 }, Jab_r[2] = { 0, 0 }, rr = 0;
		for (int i = 0; i < patternNum; i++)
		{
			JI_r[0] += resApprox[i] * rJ->JIdx[0][i];
			JI_r[1] += resApprox[i] * rJ->JIdx[1][i];
			Jab_r[0] += resApprox[i] * rJ->JabF[0][i];
			Jab_r[1] += resApprox[i] * rJ->JabF[1][i];
			rr += resApprox[i] * resApprox[i];
		}
		acc[htIDX].update(rJ->Jpdc[0], rJ->Jpdxi[0], rJ->Jpdc[1], rJ->Jpdxi[1], rJ->JIdx2[0], rJ->JIdx2[1], rJ->JIdx2[3]);
		acc[htIDX].updateBotRight(rJ->Jab2[0], rJ->Jab2[1], Jab_r[0], rJ->Jab2[3], Jab_r[1], rr);
		acc[htIDX].updateTopRight(rJ->Jpdc[0], rJ->Jpdxi[0], rJ->Jpdc[1], rJ->Jpdxi[1],
		                          rJ->JabJIdx[0], rJ->JabJIdx[1], rJ->JabJIdx[2], rJ->JabJIdx[3], JI_r[0], JI_r[1]);
		float Ji2_Jpdd[2];
		Ji2_Jpdd[0] = rJ->JIdx2[0] * rJ->Jpdd[0] + rJ->JIdx2[1] * rJ->Jpdd[1];
		Ji2_Jpdd[1] = rJ->JIdx2[2] * rJ->Jpdd[0] + rJ->JIdx2[3] * rJ->Jpdd[1];
		bd_acc += JI_r[0] * rJ->Jpdd[0] + JI_r[1] * rJ->Jpdd[1];
		Hdd_acc += Ji2_Jpdd[0] * rJ->Jpdd[0] + Ji2_Jpdd[1] * rJ->Jpdd[1];
		for (int k = 0; k < 4; k++) Hcd_acc[k] += rJ->Jpdc[0][k] * Ji2_Jpdd[0] + rJ->Jpdc[1][k] * Ji2_Jpdd[1];
		(*nres)++;
	}
	if (mode == 0)
	{
		p.Hdd_accAF = Hdd_acc; p.bd_accAF = bd_acc;
		for (int k = 0; k < 4; k++) p.Hcd_accAF[k] = Hcd_acc[k];
	}
	else
	{
		p.Hdd_accLF = Hdd_acc; p.bd_accLF = bd_acc;
		for (int k = 0; k < 4; k++) p.Hcd_accLF[k] = Hcd_acc[k];
	}
	if (mode == 2)
	{
		p.Hdd_accAF = 0; p.bd_accAF = 0;
		for (int k = 0; k < 4; k++) p.Hcd_accAF[k] = 0;
	}
}

static inline void mm8(const double* A, const double* Bm, double* C, bool transB)
{
	for (int i = 0; i < 8; i++)
		for (int j = 0; j < 8; j++)
		{
			double s = 0;
			for (int k = 0; k < 8; k++) s += A[i * 8 + k] * (transB ? Bm[j * 8 + k] : Bm[k * 8 + j]);
			C[i * 8 + j] = s;
		}
}

// stitchDoubleInternal, AccumulatedTopHessian.cpp:259-325 (usePrior=false for accSSE_top_A)
static void stitchA(dso_oracle_ba* B, std::vector<std::vector<AccApprox>>& accs, int toAggregate, double* H, double* b, int kmin, int kmax)
{
	int nF = B->nF, N = 8 * nF + CPARS;
	for (int k = kmin; k < kmax; k++)
	{
		int h = k % nF, t = k / nF;
		int hIdx = CPARS + h * 8, tIdx = CPARS + t * 8, aidx = h + nF * t;
		double accH[13][13];
		memset(accH, 0, sizeof(accH));
		for (int tid2 = 0; tid2 < toAggregate; tid2++)
		{
			accs[tid2][aidx].finish();
			if (accs[tid2][aidx].num == 0) continue;
			for (int i = 0; i < 13; i++) for (int j = 0; j < 13; j++) accH[i][j] += (double)accs[tid2][aidx].H[i][j];
		}
		const double* AH = &B->adHost[aidx * 64];
		const double* AT = &B->adTarget[aidx * 64];
		double A88[64], T1[64], T2[64];
		for (int i = 0; i < 8; i++) for (int j = 0; j < 8; j++) A88[i * 8 + j] = accH[CPARS + i][CPARS + j];
		mm8(AH, A88, T1, false); mm8(T1, AH, T2, true);
		for (int i = 0; i < 8; i++) for (int j = 0; j < 8; j++) H[(hIdx + i) * N + hIdx + j] += T2[i * 8 + j];
		mm8(AT, A88, T1, false); mm8(T1, AT, T2, true);
		for (int i = 0; i < 8; i++) for (int j = 0; j < 8; j++) H[(tIdx + i) * N + tIdx + j] += T2[i * 8 + j];
		mm8(AH, A88, T1, false); mm8(T1, AT, T2, true);
		for (int i = 0; i < 8; i++) for (int j = 0; j < 8; j++) H[(hIdx + i) * N + tIdx + j] += T2[i * 8 + j];
		for (int i = 0; i < 8; i++)
			for (int j = 0; j < CPARS; j++)
			{
				double sh = 0, st = 0;
				for (int q = 0; q < 8; q++) { sh += AH[i * 8 + q] * accH[CPARS + q][j]; st += AT[i * 8 + q] * accH[CPARS + q][j]; }
				H[(hIdx + i) * N + j] += sh;
				H[(tIdx + i) * N + j] += st;
			}
		for (int i = 0; i < CPARS; i++) for (int j = 0; j < CPARS; j++) H[i * N + j] += accH[i][j];
		for (int i = 0; i < 8; i++)
		{
			double sh = 0, st = 0;
			for (int q = 0; q < 8; q++) { sh += AH[i * 8 + q] * accH[CPARS + q][CPARS + 8]; st += AT[i * 8 + q] * accH[CPARS + q][CPARS + 8]; }
			b[hIdx + i] += sh;
			b[tIdx + i] += st;
		}
		for (int i = 0; i < CPARS; i++) b[i] += accH[i][CPARS + 8];
	}
}

// accumulateAF_MT, EnergyFunctional.cpp:209-232 + stitchDoubleMT, AccumulatedTopHessian.h:91-139
static void accumulate_top(dso_oracle_ba* B, int mode, bool usePrior, int MT, int nthreads, double* H, double* b, int* resInA,
                           float* Hdd, float* bd, float* Hcd4, const unsigned char* pointMask = nullptr)
{
	int nF = B->nF, N = 8 * nF + CPARS, nP = (int)B->points.size();
	int T = MT ? nthreads : 1;
	std::vector<std::vector<AccApprox>> accs(T, std::vector<AccApprox>(nF * nF));
	std::vector<int> nres(T, 0);
	for (auto& a : accs) for (auto& x : a) x.initialize();
	memset(H, 0, sizeof(double) * N * N); memset(b, 0, sizeof(double) * N);
	if (MT)
	{
		Pool* pool = B->getPool(nthreads);
		pool->reduce([&](int lo, int hi, int tid) { for (int i = lo; i < hi; i++) addPointA(B, B->points[i], accs[tid].data(), &nres[tid], mode); }, 0, nP, 50);
		std::vector<std::vector<double>> Hs(T, std::vector<double>(N * N, 0.0)), bs(T, std::vector<double>(N, 0.0));
		pool->reduce([&](int lo, int hi, int tid) { stitchA(B, accs, T, Hs[tid].data(), bs[tid].data(), lo, hi); }, 0, nF * nF, 0);
		for (int i = 0; i < T; i++)
		{
			for (int k = 0; k < N * N; k++) H[k] += Hs[i][k];
			for (int k = 0; k < N; k++) b[k] += bs[i][k];
			if (i > 0) nres[0] += nres[i];
		}
	}
	else
	{
		for (int i = 0; i < nP; i++) if (!pointMask || pointMask[i]) addPointA(B, B->points[i], accs[0].data(), &nres[0], mode);
		stitchA(B, accs, 1, H, b, 0, nF * nF);
	}
	if (usePrior)
	{
		// AccumulatedTopHessian.cpp:314-324 ("only do this on one thread")
		for (int i = 0; i < CPARS; i++) { H[i * N + i] += B->cPrior[i]; b[i] += B->cPrior[i] * (double)B->cDeltaF[i]; }
		if ((int)B->framePrior.size() == 8 * nF)
			for (int h = 0; h < nF; h++)
				for (int i = 0; i < 8; i++)
				{
					const int k = CPARS + h * 8 + i;
					H[k * N + k] += B->framePrior[h * 8 + i];
					b[k] += B->framePrior[h * 8 + i] * B->frameDeltaPrior[h * 8 + i];
				}
	}
	// "make diagonal by copying over parts", AccumulatedTopHessian.h:126-138
	for (int h = 0; h < nF; h++)
	{
		int hIdx = CPARS + h * 8;
		for (int i = 0; i < CPARS; i++) for (int j = 0; j < 8; j++) H[i * N + hIdx + j] = H[(hIdx + j) * N + i];
		for (int t = h + 1; t < nF; t++)
		{
			int tIdx = CPARS + t * 8;
			for (int i = 0; i < 8; i++) for (int j = 0; j < 8; j++) H[(hIdx + i) * N + tIdx + j] += H[(tIdx + j) * N + hIdx + i];
			for (int i = 0; i < 8; i++) for (int j = 0; j < 8; j++) H[(tIdx + i) * N + hIdx + j] = H[(hIdx + j) * N + tIdx + i];
		}
	}
	if (resInA) *resInA = nres[0];
	for (int i = 0; i < nP; i++)
	{
		const Pt& p = B->points[i];
		if (Hdd) Hdd[i] = mode == 0 ? p.Hdd_accAF : p.Hdd_accLF;
		if (bd) bd[i] = mode == 0 ? p.bd_accAF : p.bd_accLF;
		if (Hcd4) memcpy(Hcd4 + 4 * i, mode == 0 ? p.Hcd_accAF : p.Hcd_accLF, 16);
	}
}

extern "C" void dso_oracle_ba_accumulate_A(dso_oracle_ba* B, int MT, int nthreads, double* H, double* b, int* resInA,
                                           float* Hdd, float* bd, float* Hcd4)
{
	accumulate_top(B, 0, false, MT, nthreads, H, b, resInA, Hdd, bd, Hcd4);
}
// accumulateLF_MT, EnergyFunctional.cpp:233-254: addPoint<1> over the linearised residuals, stitch with usePrior = true
extern "C" void dso_oracle_ba_accumulate_L(dso_oracle_ba* B, int MT, int nthreads, double* H, double* b, int* resInL,
                                           float* Hdd, float* bd, float* Hcd4)
{
	accumulate_top(B, 1, true, MT, nthreads, H, b, resInL, Hdd, bd, Hcd4);
}

struct SCAcc
{
	std::vector<AccXX<8, CPARS>> accE;
	std::vector<AccX<8>> accEB;
	std::vector<AccXX<8, 8>> accD;
	AccXX<CPARS, CPARS> accHcc;
	AccX<CPARS> accbc;
	void setZero(int n)
	{
		accE.resize(n * n); accEB.resize(n * n); accD.resize(n * n * n);
		accHcc.initialize(); accbc.initialize();
		for (auto& a : accE) a.initialize();
		for (auto& a : accEB) a.initialize();
		for (auto& a : accD) a.initialize();
	}
};

// AccumulatedSCHessianSSE::addPoint, AccumulatedSCHessian.cpp:35-85
static void addPointSC(dso_oracle_ba* B, Pt& p, bool shiftPriorToZero, SCAcc& S)
{
	int nF = B->nF;
	int ngoodres = 0;
	for (int ri : p.residualsAll) if (B->res[ri].isActive) ngoodres++;
	if (ngoodres == 0)
	{
		p.HdiF = 0; p.bdSumF = 0; p.idepth_hessian = 0;
		return;
	}
	float H = p.Hdd_accAF + p.Hdd_accLF + p.priorF;
	if (H < 1e-10) H = 1e-10;
	p.idepth_hessian = H;
	p.HdiF = 1.0 / H;
	p.bdSumF = p.bd_accAF + p.bd_accLF;
	if (shiftPriorToZero) p.bdSumF += p.priorF * p.deltaF;
	float Hcd[4];
	for (int k = 0; k < 4; k++) Hcd[k] = p.Hcd_accAF[k] + p.Hcd_accLF[k];
	S.accHcc.update(Hcd, Hcd, p.HdiF);
	S.accbc.update(Hcd, p.bdSumF * p.HdiF);
	int nFrames2 = nF * nF;
	for (int r1i : p.residualsAll)
	{
		Res& r1 = B->res[r1i];
		if (!r1.isActive) continue;
		int r1ht = r1.host + r1.target * nF;
		for (int r2i : p.residualsAll)
		{
			Res& r2 = B->res[r2i];
			if (!r2.isActive) continue;
			S.accD[r1ht + r2.target * nFrames2].update(r1.JpJdF, r2.JpJdF, p.HdiF);
		}
		S.accE[r1ht].update(r1.JpJdF, Hcd, p.HdiF);
		S.accEB[r1ht].update(r1.JpJdF, p.HdiF * p.bdSumF);
	}
}

// stitchDoubleInternal, AccumulatedSCHessian.cpp:86-165
static void stitchSC(dso_oracle_ba* B, std::vector<SCAcc>& S, int toAggregate, double* H, double* b, int kmin, int kmax)
{
	int nf = B->nF, N = 8 * nf + CPARS, nframes2 = nf * nf;
	for (int k = kmin; k < kmax; k++)
	{
		int i = k % nf, j = k / nf;
		int iIdx = CPARS + i * 8, jIdx = CPARS + j * 8, ijIdx = i + nf * j;
		double Hpc[8][CPARS], bp[8];
		memset(Hpc, 0, sizeof(Hpc)); memset(bp, 0, sizeof(bp));
		for (int tid2 = 0; tid2 < toAggregate; tid2++)
		{
			S[tid2].accE[ijIdx].finish();
			S[tid2].accEB[ijIdx].finish();
			for (int r = 0; r < 8; r++) { for (int c = 0; c < CPARS; c++) Hpc[r][c] += (double)S[tid2].accE[ijIdx].A1m[r][c]; bp[r] += (double)S[tid2].accEB[ijIdx].A1m[r]; }
		}
		const double* AHij = &B->adHost[ijIdx * 64];
		const double* ATij = &B->adTarget[ijIdx * 64];
		for (int r = 0; r < 8; r++)
		{
			for (int c = 0; c < CPARS; c++)
			{
				double sh = 0, st = 0;
				for (int q = 0; q < 8; q++) { sh += AHij[r * 8 + q] * Hpc[q][c]; st += ATij[r * 8 + q] * Hpc[q][c]; }
				H[(iIdx + r) * N + c] += sh;
				H[(jIdx + r) * N + c] += st;
			}
			double sh = 0, st = 0;
			for (int q = 0; q < 8; q++) { sh += AHij[r * 8 + q] * bp[q]; st += ATij[r * 8 + q] * bp[q]; }
			b[iIdx + r] += sh;
			b[jIdx + r] += st;
		}
		for (int k2 = 0; k2 < nf; k2++)
		{
			int kIdx = CPARS + k2 * 8;
			int ijkIdx = ijIdx + k2 * nframes2;
			int ikIdx = i + nf * k2;
			double accDM[64];
			memset(accDM, 0, sizeof(accDM));
			for (int tid2 = 0; tid2 < toAggregate; tid2++)
			{
				S[tid2].accD[ijkIdx].finish();
				if (S[tid2].accD[ijkIdx].num == 0) continue;
				for (int r = 0; r < 8; r++) for (int c = 0; c < 8; c++) accDM[r * 8 + c] += (double)S[tid2].accD[ijkIdx].A1m[r][c];
			}
			const double* AHik = &B->adHost[ikIdx * 64];
			const double* ATik = &B->adTarget[ikIdx * 64];
			double T1[64], T2[64];
			mm8(AHij, accDM, T1, false); mm8(T1, AHik, T2, true);
			for (int r = 0; r < 8; r++) for (int c = 0; c < 8; c++) H[(iIdx + r) * N + iIdx + c] += T2[r * 8 + c];
			mm8(ATij, accDM, T1, false); mm8(T1, ATik, T2, true);
			for (int r = 0; r < 8; r++) for (int c = 0; c < 8; c++) H[(jIdx + r) * N + kIdx + c] += T2[r * 8 + c];
			mm8(ATij, accDM, T1, false); mm8(T1, AHik, T2, true);
			for (int r = 0; r < 8; r++) for (int c = 0; c < 8; c++) H[(jIdx + r) * N + iIdx + c] += T2[r * 8 + c];
			mm8(AHij, accDM, T1, false); mm8(T1, ATik, T2, true);
			for (int r = 0; r < 8; r++) for (int c = 0; c < 8; c++) H[(iIdx + r) * N + kIdx + c] += T2[r * 8 + c];
		}
	}
	if (kmin == 0)
	{
		for (int tid2 = 0; tid2 < toAggregate; tid2++)
		{
			S[tid2].accHcc.finish();
			S[tid2].accbc.finish();
			for (int r = 0; r < CPARS; r++) { for (int c = 0; c < CPARS; c++) H[r * N + c] += (double)S[tid2].accHcc.A1m[r][c]; b[r] += (double)S[tid2].accbc.A1m[r]; }
		}
	}
}

// accumulateSCF_MT, EnergyFunctional.cpp:257-274 + stitchDoubleMT, AccumulatedSCHessian.h:93-135
extern "C" void dso_oracle_ba_accumulate_SC(dso_oracle_ba* B, int MT, int nthreads, double* H, double* b,
                                            float* HdiF, float* bdSumF, float* idepth_hessian)
{
	int nF = B->nF, N = 8 * nF + CPARS, nP = (int)B->points.size();
	int T = MT ? nthreads : 1;
	std::vector<SCAcc> S(T);
	for (auto& s : S) s.setZero(nF);
	memset(H, 0, sizeof(double) * N * N); memset(b, 0, sizeof(double) * N);
	if (MT)
	{
		Pool* pool = B->getPool(nthreads);
		pool->reduce([&](int lo, int hi, int tid) { for (int i = lo; i < hi; i++) addPointSC(B, B->points[i], true, S[tid]); }, 0, nP, 50);
		std::vector<std::vector<double>> Hs(T, std::vector<double>(N * N, 0.0)), bs(T, std::vector<double>(N, 0.0));
		pool->reduce([&](int lo, int hi, int tid) { stitchSC(B, S, T, Hs[tid].data(), bs[tid].data(), lo, hi); }, 0, nF * nF, 0);
		for (int i = 0; i < T; i++)
		{
			for (int k = 0; k < N * N; k++) H[k] += Hs[i][k];
			for (int k = 0; k < N; k++) b[k] += bs[i][k];
		}
	}
	else
	{
		// EnergyFunctional.cpp:268-272: ST branch also passes shiftPriorToZero = true
		for (int i = 0; i < nP; i++) addPointSC(B, B->points[i], true, S[0]);
		stitchSC(B, S, 1, H, b, 0, nF * nF);
	}
	for (int h = 0; h < nF; h++)
	{
		int hIdx = CPARS + h * 8;
		for (int i = 0; i < CPARS; i++) for (int j = 0; j < 8; j++) H[i * N + hIdx + j] = H[(hIdx + j) * N + i];
	}
	for (int i = 0; i < nP; i++)
	{
		if (HdiF) HdiF[i] = B->points[i].HdiF;
		if (bdSumF) bdSumF[i] = B->points[i].bdSumF;
		if (idepth_hessian) idepth_hessian[i] = B->points[i].idepth_hessian;
	}
}

// =================================================================================================
// N1: solveSystemF tail + resubstituteF_MT (EnergyFunctional.cpp:802-924, 276-330) and the pieces of
// N2 it reads (fixLinearizationF, EnergyFunctionalStructs.cpp:89-115)
// =================================================================================================
extern "C" void dso_oracle_ba_set_priors(dso_oracle_ba* B, const double* cPrior4, const double* framePrior, const double* frameDeltaPrior,
                                         const double* frameDelta, const double* HM, const double* bM)
{
	const int nF = B->nF, N = 8 * nF + CPARS;
	for (int i = 0; i < 4; i++) B->cPrior[i] = cPrior4 ? cPrior4[i] : 0.0;
	B->framePrior.assign(8 * nF, 0.0); B->frameDeltaPrior.assign(8 * nF, 0.0); B->frameDelta.assign(8 * nF, 0.0);
	if (framePrior) B->framePrior.assign(framePrior, framePrior + 8 * nF);
	if (frameDeltaPrior) B->frameDeltaPrior.assign(frameDeltaPrior, frameDeltaPrior + 8 * nF);
	if (frameDelta) B->frameDelta.assign(frameDelta, frameDelta + 8 * nF);
	B->HM.assign((size_t)N * N, 0.0); B->bM.assign(N, 0.0);
	if (HM) B->HM.assign(HM, HM + (size_t)N * N);
	if (bM) B->bM.assign(bM, bM + N);
}
extern "C" void dso_oracle_ba_set_nullspaces(dso_oracle_ba* B, const double* ns, int n)
{
	const int N = 8 * B->nF + CPARS;
	B->nNullspaces = n;
	B->nullspaces.assign(ns, ns + (size_t)n * N);
}

// EFResidual::fixLinearizationF for the flagged residuals (which == NULL: every active one)
extern "C" void dso_oracle_ba_fix_linearization(dso_oracle_ba* B, const unsigned char* which, float* res_toZeroF_out)
{
	for (size_t i = 0; i < B->res.size(); i++)
	{
		Res& r = B->res[i];
		if (r.isActive && (!which || which[i]))
		{
			const RawJ* J = &r.efJ;
			float d[4];
			jp_delta(B, r, J, B->points[r.point].deltaF, d);
			for (int k = 0; k < patternNum; k++)
			{
				float rtz = J->resF[k];
				rtz = rtz - J->JIdx[0][k] * d[0];
				rtz = rtz - J->JIdx[1][k] * d[1];
				rtz = rtz - J->JabF[0][k] * d[2];
				rtz = rtz - J->JabF[1][k] * d[3];
				r.res_toZeroF[k] = rtz;
			}
			r.isLinearized = true;
		}
		if (res_toZeroF_out) memcpy(res_toZeroF_out + 8 * i, r.res_toZeroF, 32);
	}
}

// EnergyFunctional::orthogonalize(VecX* b, MatXX* H), :741-792: NNpiTS = 0.5 (N Npi^T + Npi N^T) from the normalised
// null-space vectors; b -= NNpiTS b; H -= NNpiTS H NNpiTS
static bool nullspace_projector(const dso_oracle_ba* B, double solverModeDelta, int N, std::vector<double>& PS)
{
	const int k = B->nNullspaces;
	if (k == 0) return false;
	std::vector<double> Nn((size_t)N * k), U((size_t)N * k), S(k), V((size_t)k * k);
	for (int j = 0; j < k; j++)
	{
		double n2 = 0;
		for (int i = 0; i < N; i++) n2 += B->nullspaces[(size_t)j * N + i] * B->nullspaces[(size_t)j * N + i];
		const double nrm = std::sqrt(n2);
		for (int i = 0; i < N; i++) Nn[(size_t)i * k + j] = B->nullspaces[(size_t)j * N + i] / nrm;
	}
	orc::jacobi_svd_thin(Nn.data(), N, k, U.data(), S.data(), V.data());
	double maxSv = 0;
	for (int i = 0; i < k; i++) if (S[i] > maxSv) maxSv = S[i];
	for (int i = 0; i < k; i++) S[i] = (S[i] > solverModeDelta * maxSv) ? 1.0 / S[i] : 0.0;
	// Npi = U diag(S) V^T  (N x k);  NNpiT = Nn Npi^T (N x N);  NNpiTS = 0.5 (NNpiT + NNpiT^T)
	std::vector<double> Npi((size_t)N * k, 0.0), P((size_t)N * N, 0.0);
	for (int i = 0; i < N; i++)
		for (int j = 0; j < k; j++)
		{
			double s = 0;
			for (int q = 0; q < k; q++) s += U[(size_t)i * k + q] * S[q] * V[(size_t)j * k + q];
			Npi[(size_t)i * k + j] = s;
		}
	for (int i = 0; i < N; i++)
		for (int j = 0; j < N; j++)
		{
			double s = 0;
			for (int q = 0; q < k; q++) s += Nn[(size_t)i * k + q] * Npi[(size_t)j * k + q];
			P[(size_t)i * N + j] = s;
		}
	PS.assign((size_t)N * N, 0.0);
	for (int i = 0; i < N; i++)
		for (int j = 0; j < N; j++) PS[(size_t)i * N + j] = 0.5 * (P[(size_t)i * N + j] + P[(size_t)j * N + i]);
	return true;
}
static void orthogonalize_sys(const dso_oracle_ba* B, double solverModeDelta, int N, double* b, double* H)
{
	std::vector<double> PS;
	if (!nullspace_projector(B, solverModeDelta, N, PS)) return;
	if (b)
	{
		std::vector<double> y(N);
		for (int i = 0; i < N; i++)
		{
			double s = 0;
			for (int j = 0; j < N; j++) s += PS[(size_t)i * N + j] * b[j];
			y[i] = s;
		}
		for (int i = 0; i < N; i++) b[i] -= y[i];
	}
	if (H)
	{
		std::vector<double> T((size_t)N * N), T2((size_t)N * N);
		for (int i = 0; i < N; i++)
			for (int j = 0; j < N; j++)
			{
				double s = 0;
				for (int q = 0; q < N; q++) s += PS[(size_t)i * N + q] * H[(size_t)q * N + j];
				T[(size_t)i * N + j] = s;
			}
		for (int i = 0; i < N; i++)
			for (int j = 0; j < N; j++)
			{
				double s = 0;
				for (int q = 0; q < N; q++) s += T[(size_t)i * N + q] * PS[(size_t)q * N + j];
				T2[(size_t)i * N + j] = s;
			}
		for (size_t i = 0; i < (size_t)N * N; i++) H[i] -= T2[i];
	}
}
static void orthogonalize_vec(const dso_oracle_ba* B, double solverModeDelta, std::vector<double>& x)
{
	orthogonalize_sys(B, solverModeDelta, (int)x.size(), x.data(), 0);
}

#define SOLVER_SVD 1
#define SOLVER_ORTHOGONALIZE_SYSTEM 2
#define SOLVER_USE_GN 64
#define SOLVER_FIX_LAMBDA 128
#define SOLVER_ORTHOGONALIZE_X 256
#define SOLVER_ORTHOGONALIZE_X_LATER 2048

// EnergyFunctional::resubstituteF_MT + resubstituteFPt, EnergyFunctional.cpp:276-330
static void resubstitute(dso_oracle_ba* B, const std::vector<double>& x)
{
	const int nF = B->nF, N = 8 * nF + CPARS, nP = (int)B->points.size();
	std::vector<float> xF(N), xAd((size_t)nF * nF * 8);
	for (int i = 0; i < N; i++) xF[i] = (float)x[i];
	for (int h = 0; h < nF; h++)
		for (int t = 0; t < nF; t++)
		{
			const float* AH = &B->adHostF[(h + nF * t) * 64];
			const float* AT = &B->adTargetF[(h + nF * t) * 64];
			for (int j = 0; j < 8; j++)
			{
				// (1x8)*(8x8) coefficient products, each an unrolled 8-term redux, then the two rows added (:291-292)
				float ph[8], pt[8];
				for (int i = 0; i < 8; i++) { ph[i] = xF[CPARS + 8 * h + i] * AH[i * 8 + j]; pt[i] = xF[CPARS + 8 * t + i] * AT[i * 8 + j]; }
				xAd[(size_t)(nF * h + t) * 8 + j] = sum8_tree(ph) + sum8_tree(pt);
			}
		}
	B->pointStep.assign(nP, 0.f);
	for (int k = 0; k < nP; k++)
	{
		Pt& p = B->points[k];
		int ngoodres = 0;
		for (int ri : p.residualsAll) if (B->res[ri].isActive) ngoodres++;
		if (ngoodres == 0) { B->pointStep[k] = 0; continue; }
		float b = p.bdSumF;
		float hc[4];
		for (int i = 0; i < 4; i++) hc[i] = p.Hcd_accAF[i] + p.Hcd_accLF[i];
		b -= dot4_tree(&xF[0], hc);
		for (int ri : p.residualsAll)
		{
			const Res& r = B->res[ri];
			if (!r.isActive) continue;
			const float* xa = &xAd[(size_t)(r.host * nF + r.target) * 8];
			float pr[8];
			for (int i = 0; i < 8; i++) pr[i] = xa[i] * r.JpJdF[i];
			b -= sum8_tree(pr);
		}
		B->pointStep[k] = -b * p.HdiF;
	}
}

// x = SVecI * (SVecI H SVecI).ldlt().solve(SVecI b), SVecI = 1/sqrt(diag(H) + 10), EnergyFunctional.cpp:905-908
extern "C" void dso_oracle_scaled_ldlt_solve(const double* HF, const double* bF, int N, double* x)
{
	std::vector<double> sv(N), Hs((size_t)N * N), bs(N), y(N);
	for (int i = 0; i < N; i++) sv[i] = 1.0 / std::sqrt(HF[(size_t)i * N + i] + 10);
	for (int i = 0; i < N; i++) for (int j = 0; j < N; j++) Hs[(size_t)i * N + j] = (sv[i] * HF[(size_t)i * N + j]) * sv[j];
	for (int i = 0; i < N; i++) bs[i] = sv[i] * bF[i];
	orc::ldlt_solve_n(Hs.data(), bs.data(), y.data(), N);
	for (int i = 0; i < N; i++) x[i] = sv[i] * y[i];
}

// EnergyFunctional::solveSystemF, :802-924.  Restated: the default solver (LDLT on the diagonally scaled system) with
// SOLVER_USE_GN / SOLVER_FIX_LAMBDA / SOLVER_ORTHOGONALIZE_X(_LATER), and the non-default SVD (+CUT7) and ORTHOGONALIZE_SYSTEM branches.
// x_out: 8nF+4 (= lastX; frame steps are -x segments, the calibration step -x.head<4>), pointStep_out: nP (PointHessian::step).
extern "C" int dso_oracle_ba_solve(dso_oracle_ba* B, int iteration, double lambda, int solverMode, double solverModeDelta, int MT, int nthreads,
                                   double* x_out, float* pointStep_out, double* HFinal_out, double* bFinal_out)
{
	if (solverMode & SOLVER_USE_GN) lambda = 0;
	if (solverMode & SOLVER_FIX_LAMBDA) lambda = 1e-5;
	const int nF = B->nF, N = 8 * nF + CPARS;
	if ((int)B->bM.size() != N) dso_oracle_ba_set_priors(B, B->cPrior, 0, 0, 0, 0, 0);
	std::vector<double> HA((size_t)N * N), bA(N), HL((size_t)N * N), bL(N), HS((size_t)N * N), bS(N);
	accumulate_top(B, 0, false, MT, nthreads, HA.data(), bA.data(), 0, 0, 0, 0);     // :816
	accumulate_top(B, 1, true, MT, nthreads, HL.data(), bL.data(), 0, 0, 0, 0);      // :817
	dso_oracle_ba_accumulate_SC(B, MT, nthreads, HS.data(), bS.data(), 0, 0, 0);      // :820
	std::vector<double> d(N), bMtop(N), HF((size_t)N * N), bF(N);
	for (int i = 0; i < CPARS; i++) d[i] = (double)B->cDeltaF[i];
	for (int i = 0; i < 8 * nF; i++) d[CPARS + i] = B->frameDelta[i];
	for (int i = 0; i < N; i++)
	{
		double s = 0;
		for (int j = 0; j < N; j++) s += B->HM[(size_t)i * N + j] * d[j];
		bMtop[i] = B->bM[i] + s;                                                       // :824
	}
	if (solverMode & SOLVER_ORTHOGONALIZE_SYSTEM)
	{
		// :832-851: HT_act = HL + HA - H_sc, bT_act = bL + bA - b_sc, projected off the null spaces unless the first frame
		// (the gauge anchor) is still in the window; then the marginalisation prior, then the damping
		for (size_t i = 0; i < (size_t)N * N; i++) HF[i] = (HL[i] + HA[i]) - HS[i];
		for (int i = 0; i < N; i++) bF[i] = (bL[i] + bA[i]) - bS[i];
		if (!B->haveFirstFrame) orthogonalize_sys(B, solverModeDelta, N, bF.data(), HF.data());
		for (size_t i = 0; i < (size_t)N * N; i++) HF[i] = HF[i] + B->HM[i];
		for (int i = 0; i < N; i++) bF[i] = bF[i] + bMtop[i];
		for (int i = 0; i < N; i++) HF[(size_t)i * N + i] *= (1 + lambda);
	}
	else
	{
	for (size_t i = 0; i < (size_t)N * N; i++) HF[i] = (HL[i] + B->HM[i]) + HA[i];     // :857
	for (int i = 0; i < N; i++) bF[i] = ((bL[i] + bMtop[i]) + bA[i]) - bS[i];          // :859
	for (int i = 0; i < N; i++) HF[(size_t)i * N + i] *= (1 + lambda);                  // :865
	const double isc = 1.0f / (1 + lambda);                                             // :866
	for (size_t i = 0; i < (size_t)N * N; i++) HF[i] -= HS[i] * isc;
	}
	if (HFinal_out) memcpy(HFinal_out, HF.data(), sizeof(double) * N * N);
	if (bFinal_out) memcpy(bFinal_out, bF.data(), sizeof(double) * N);
	std::vector<double> x(N);
	if (solverMode & SOLVER_SVD)
	{
		// :870-900: SVecI = 1/sqrt(diag) (no +10 here), thin SVD of the scaled matrix, singular values below
		// solverModeDelta * max zeroed (SVD_CUT7: also the last seven), x = SVecI V (U^T b ./ S)
		std::vector<double> sv(N), Hs((size_t)N * N), bs(N), U((size_t)N * N), S(N), V((size_t)N * N);
		for (int i = 0; i < N; i++) sv[i] = 1.0 / std::sqrt(HF[(size_t)i * N + i]);
		for (int i = 0; i < N; i++)
			for (int j = 0; j < N; j++) Hs[(size_t)i * N + j] = (sv[i] * HF[(size_t)i * N + j]) * sv[j];
		for (int i = 0; i < N; i++) bs[i] = sv[i] * bF[i];
		orc::jacobi_svd_thin(Hs.data(), N, N, U.data(), S.data(), V.data());
		double maxSv = 0;
		for (int i = 0; i < N; i++) if (S[i] > maxSv) maxSv = S[i];
		// Eigen::JacobiSVD returns the singular values in decreasing order (SVD_CUT7 drops "the last seven")
		std::vector<int> ord(N);
		for (int i = 0; i < N; i++) ord[i] = i;
		std::stable_sort(ord.begin(), ord.end(), [&](int a, int b) { return S[a] > S[b]; });
		std::vector<double> Ub(N);
		for (int i = 0; i < N; i++)
		{
			double t = 0;
			for (int j = 0; j < N; j++) t += U[(size_t)j * N + ord[i]] * bs[j];
			Ub[i] = t;
		}
		for (int i = 0; i < N; i++)
		{
			const double Si = S[ord[i]];
			if (Si < solverModeDelta * maxSv) Ub[i] = 0;
			if ((solverMode & 16) && (i >= N - 7)) Ub[i] = 0;                              // SOLVER_SVD_CUT7
			else Ub[i] /= Si;
		}
		for (int i = 0; i < N; i++)
		{
			double t = 0;
			for (int j = 0; j < N; j++) t += V[(size_t)i * N + ord[j]] * Ub[j];
			x[i] = sv[i] * t;
		}
	}
	else
		dso_oracle_scaled_ldlt_solve(HF.data(), bF.data(), N, x.data());                // :905-908
	if ((solverMode & SOLVER_ORTHOGONALIZE_X) || (iteration >= 2 && (solverMode & SOLVER_ORTHOGONALIZE_X_LATER)))
		orthogonalize_vec(B, solverModeDelta, x);                                       // :913-917
	resubstitute(B, x);                                                                 // :925
	if (x_out) memcpy(x_out, x.data(), sizeof(double) * N);
	if (pointStep_out) memcpy(pointStep_out, B->pointStep.data(), sizeof(float) * B->pointStep.size());
	return 0;
}

// =================================================================================================
// rest of N2: marginalizePointsF (EnergyFunctional.cpp:642-703), calcLEnergyF_MT / calcLEnergyPt (:346-428),
// calcMEnergyF (:334-343)
// =================================================================================================
// Accumulator11, MatrixAccumulators.h:89-170
struct Acc11
{
	float A; float d[4], d1k[4], d1m[4]; float n1, n1k, n1m;
	void initialize() { A = 0; for (int i = 0; i < 4; i++) d[i] = d1k[i] = d1m[i] = 0; n1 = n1k = n1m = 0; }
	void shiftUp(bool force)
	{
		if (n1 > 1000 || force) { for (int i = 0; i < 4; i++) { d1k[i] = d[i] + d1k[i]; d[i] = 0; } n1k += n1; n1 = 0; }
		if (n1k > 1000 || force) { for (int i = 0; i < 4; i++) { d1m[i] = d1k[i] + d1m[i]; d1k[i] = 0; } n1m += n1k; n1k = 0; }
	}
	void updateSingle(float v) { d[0] += v; n1++; shiftUp(false); }
	void updateSSENoShift(const float* v4) { for (int i = 0; i < 4; i++) d[i] = d[i] + v4[i]; n1++; }
	void finish() { shiftUp(true); A = d1m[0] + d1m[1] + d1m[2] + d1m[3]; }
};

static float lenergy_range(dso_oracle_ba* B, int lo, int hi)
{
	Acc11 E; E.initialize();
	for (int i = lo; i < hi; i++)
	{
		Pt& p = B->points[i];
		for (int ri : p.residualsAll)
		{
			Res& r = B->res[ri];
			if (!r.isLinearized || !r.isActive) continue;
			const RawJ* rJ = &r.efJ;
			float d[4];
			jp_delta(B, r, rJ, p.deltaF, d);
			for (int k = 0; k + 3 < patternNum; k += 4)
			{
				float v[4];
				for (int q = 0; q < 4; q++)
				{
					float Jdelta = rJ->JIdx[0][k + q] * d[0];
					Jdelta = Jdelta + rJ->JIdx[1][k + q] * d[1];
					Jdelta = Jdelta + rJ->JabF[0][k + q] * d[2];
					Jdelta = Jdelta + rJ->JabF[1][k + q] * d[3];
					float r0 = r.res_toZeroF[k + q];
					r0 = r0 + r0;
					r0 = r0 + Jdelta;
					v[q] = Jdelta * r0;
				}
				E.updateSSENoShift(v);
			}
		}
		E.updateSingle(p.deltaF * p.deltaF * p.priorF);
	}
	E.finish();
	return E.A;
}

// calcLEnergyF_MT: nthreads <= 1 sums all points with one accumulator (what the reference's reduce does for one worker);
// otherwise chunks of 50 points per accumulator, chunk results added in double in chunk order
extern "C" double dso_oracle_ba_calc_L_energy(dso_oracle_ba* B, int nthreads)
{
	const int nF = B->nF, nP = (int)B->points.size();
	double E = 0;
	if ((int)B->framePrior.size() == 8 * nF)
		for (int h = 0; h < nF; h++)
		{
			double e = 0;            // f->delta_prior.cwiseProduct(f->prior).dot(f->delta_prior), dynamic-free Vec8: unrolled redux
			double t[8];
			for (int i = 0; i < 8; i++) t[i] = (B->frameDeltaPrior[8 * h + i] * B->framePrior[8 * h + i]) * B->frameDeltaPrior[8 * h + i];
			e = ((t[0] + t[1]) + (t[2] + t[3])) + ((t[4] + t[5]) + (t[6] + t[7]));
			E += e;
		}
	{
		float t[4];
		for (int i = 0; i < 4; i++) t[i] = (B->cDeltaF[i] * (float)B->cPrior[i]) * B->cDeltaF[i];     // cDeltaF.cwiseProduct(cPriorF).dot(cDeltaF)
		E += (t[0] + t[1]) + (t[2] + t[3]);
	}
	double pts = 0;
	if (nthreads <= 1) pts = lenergy_range(B, 0, nP);
	else for (int lo = 0; lo < nP; lo += 50) pts += lenergy_range(B, lo, std::min(nP, lo + 50));
	return E + pts;
}

// calcMEnergyF: delta.dot(2*bM + HM*delta)
extern "C" double dso_oracle_ba_calc_M_energy(dso_oracle_ba* B)
{
	const int nF = B->nF, N = 8 * nF + CPARS;
	if ((int)B->bM.size() != N) return 0.0;
	std::vector<double> d(N);
	for (int i = 0; i < CPARS; i++) d[i] = (double)B->cDeltaF[i];
	for (int i = 0; i < 8 * nF; i++) d[CPARS + i] = B->frameDelta[i];
	double e = 0;
	for (int i = 0; i < N; i++)
	{
		double s = 0;
		for (int j = 0; j < N; j++) s += B->HM[(size_t)i * N + j] * d[j];
		e += d[i] * (2 * B->bM[i] + s);
	}
	return e;
}

// marginalizePointsF for the points flagged PS_MARGINALIZE: priorF *= setting_idepthFixPriorMargFac, addPoint<2> +
// AccumulatedSCHessianSSE::addPoint(p, false), H = M - Msc, HM += setting_margWeightFac * H (likewise b).  The structural
// part (removePoint, makeIDX) is the caller's.  H_out / b_out: the contribution M - Msc, Mb - Mbsc.
extern "C" void dso_oracle_ba_marginalize_points(dso_oracle_ba* B, const unsigned char* flags, float priorMargFac, double margWeightFac,
                                                 double* H_out, double* b_out, int* resInM, float* idepth_hessian_out)
{
	const int nF = B->nF, N = 8 * nF + CPARS, nP = (int)B->points.size();
	if ((int)B->bM.size() != N) dso_oracle_ba_set_priors(B, B->cPrior, 0, 0, 0, 0, 0);
	for (int i = 0; i < nP; i++) if (flags[i]) B->points[i].priorF *= priorMargFac;
	std::vector<double> M((size_t)N * N), Mb(N), Msc((size_t)N * N, 0.0), Mbsc(N, 0.0);
	accumulate_top(B, 2, false, 0, 1, M.data(), Mb.data(), resInM, 0, 0, 0, flags);
	std::vector<SCAcc> S(1);
	S[0].setZero(nF);
	for (int i = 0; i < nP; i++) if (flags[i]) addPointSC(B, B->points[i], false, S[0]);
	stitchSC(B, S, 1, Msc.data(), Mbsc.data(), 0, nF * nF);
	for (int h = 0; h < nF; h++)
	{
		int hIdx = CPARS + h * 8;
		for (int i = 0; i < CPARS; i++) for (int j = 0; j < 8; j++) Msc[i * N + hIdx + j] = Msc[(hIdx + j) * N + i];
	}
	std::vector<double> Hm((size_t)N * N), bm(N);
	for (size_t i = 0; i < (size_t)N * N; i++) Hm[i] = M[i] - Msc[i];
	for (int i = 0; i < N; i++) bm[i] = Mb[i] - Mbsc[i];
	if ((B->margSolverMode & 4) && !B->haveFirstFrame)                                    // SOLVER_ORTHOGONALIZE_POINTMARG, :684-692
		orthogonalize_sys(B, B->margSolverModeDelta, N, bm.data(), Hm.data());
	for (size_t i = 0; i < (size_t)N * N; i++)
	{
		if (H_out) H_out[i] = Hm[i];
		B->HM[i] += margWeightFac * Hm[i];
	}
	for (int i = 0; i < N; i++)
	{
		if (b_out) b_out[i] = bm[i];
		B->bM[i] += margWeightFac * bm[i];
	}
	if (B->margSolverMode & 8)                                                            // SOLVER_ORTHOGONALIZE_FULL, :698-699
		orthogonalize_sys(B, B->margSolverModeDelta, N, B->bM.data(), B->HM.data());
	if (idepth_hessian_out) for (int i = 0; i < nP; i++) idepth_hessian_out[i] = B->points[i].idepth_hessian;
}
// the pieces of global state the non-default solver branches look at: whether the frame with frameID 0 is still in the
// window (:688, :835) and setting_solverMode / setting_solverModeDelta as seen by marginalizePointsF
extern "C" void dso_oracle_ba_set_solver_state(dso_oracle_ba* B, int haveFirstFrame, int margSolverMode, double margSolverModeDelta)
{
	B->haveFirstFrame = haveFirstFrame; B->margSolverMode = margSolverMode; B->margSolverModeDelta = margSolverModeDelta;
}

extern "C" void dso_oracle_ba_get_marg_prior(dso_oracle_ba* B, double* HM, double* bM)
{
	const int N = 8 * B->nF + CPARS;
	if ((int)B->bM.size() != N) dso_oracle_ba_set_priors(B, B->cPrior, 0, 0, 0, 0, 0);
	memcpy(HM, B->HM.data(), sizeof(double) * N * N); memcpy(bM, B->bM.data(), sizeof(double) * N);
}
