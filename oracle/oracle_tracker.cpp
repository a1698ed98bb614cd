// TEST INFRASTRUCTURE ONLY -- CPU oracle (see dso_oracle.h). PARITY UNPINNED (no reference fixtures exist).
// Restates FrameHessian::makeImages and the CoarseTracker hot path of the reference in plain C++.
// Compile with -ffp-contract=off: every float operation below is a separate IEEE operation,
// written in the evaluation order of the reference expression it cites.
#include "dso_oracle.h"
#include "oracle_math.h"
#include <vector>
#include <cstdlib>
#include <cstdio>

#include "oracle_settings.h"
using namespace orc;

// settings.cpp defaults; run-time globals in the reference, same here (oracle_settings.h)
dso_oracle_settings dso_oracle_set = { 9, 50 * 50, 0.027f, 2, 3, 1.0f, 3, 0.1f, 1.2f, 1.5f, 2, 100 };
extern "C" void dso_oracle_set_settings(const dso_oracle_settings* s)
{
	const dso_oracle_settings d = { 9, 50 * 50, 0.027f, 2, 3, 1.0f, 3, 0.1f, 1.2f, 1.5f, 2, 100 };
	dso_oracle_set = s ? *s : d;
}
static const float setting_coarseCutoffTH = 20;
// settings.cpp:119-120 are run-time globals in the reference (-1: fixed, >= 0: optimised); same here
float dso_oracle_setting_affineOptModeA = 1e12f;
float dso_oracle_setting_affineOptModeB = 1e8f;
#define setting_affineOptModeA dso_oracle_setting_affineOptModeA
#define setting_affineOptModeB dso_oracle_setting_affineOptModeB
extern "C" void dso_oracle_set_affine_modes(float modeA, float modeB)
{
	dso_oracle_setting_affineOptModeA = modeA; dso_oracle_setting_affineOptModeB = modeB;
}
// HessianBlocks.h:54-61
#define SCALE_XI_ROT 1.0f
#define SCALE_XI_TRANS 0.5f
#define SCALE_A 10.0f
#define SCALE_B 1000.0f

// ---------------------------------------------------------------------------------------------
// HessianBlocks.cpp:137-211
extern "C" void dso_oracle_make_images(const float* color, int w, int h, int levels, const float* B,
                                       float* const* dIp, float* const* absgrad)
{
	// :155-156
	for (int i = 0; i < w * h; i++) dIp[0][3 * i] = color[i];
	for (int lvl = 0; lvl < levels; lvl++)
	{
		int wl = w >> lvl, hl = h >> lvl;
		float* dI_l = dIp[lvl];
		float* dabs_l = absgrad[lvl];
		if (lvl > 0)
		{
			int wlm1 = w >> (lvl - 1);
			const float* dI_lm = dIp[lvl - 1];
			// :175-183
			for (int y = 0; y < hl; y++)
				for (int x = 0; x < wl; x++)
				{
					dI_l[3 * (x + y * wl)] = 0.25f * (dI_lm[3 * (2 * x + 2 * y * wlm1)] +
					                                  dI_lm[3 * (2 * x + 1 + 2 * y * wlm1)] +
					                                  dI_lm[3 * (2 * x + 2 * y * wlm1 + wlm1)] +
					                                  dI_lm[3 * (2 * x + 1 + 2 * y * wlm1 + wlm1)]);
				}
		}
		// :188-209
		for (int idx = wl; idx < wl * (hl - 1); idx++)
		{
			float dx = 0.5f * (dI_l[3 * (idx + 1)] - dI_l[3 * (idx - 1)]);
			float dy = 0.5f * (dI_l[3 * (idx + wl)] - dI_l[3 * (idx - wl)]);
			if (!std::isfinite(dx)) dx = 0;
			if (!std::isfinite(dy)) dy = 0;
			dI_l[3 * idx + 1] = dx;
			dI_l[3 * idx + 2] = dy;
			dabs_l[idx] = dx * dx + dy * dy;
			if (B != 0)
			{
				// CalibHessian::getBGradOnly, HessianBlocks.h:402-410
				int c = dI_l[3 * idx] + 0.5f;
				if (c < 5) c = 5;
				if (c > 250) c = 250;
				float gw = B[c + 1] - B[c];
				dabs_l[idx] *= gw * gw;
			}
		}
	}
}

// ---------------------------------------------------------------------------------------------
// Accumulator9, MatrixAccumulators.h:958-1322 (4 SSE lanes, three tiers shifted every >1000 updates)
struct Acc9
{
	float d[45][4], d1k[45][4], d1m[45][4];
	float numIn1, numIn1k, numIn1m;
	float H[9][9];
	void initialize() { memset(d, 0, sizeof(d)); memset(d1k, 0, sizeof(d1k)); memset(d1m, 0, sizeof(d1m)); numIn1 = numIn1k = numIn1m = 0; }
	void shiftUp(bool force)
	{
		if (numIn1 > 1000 || force)
		{
			for (int i = 0; i < 45; i++) for (int l = 0; l < 4; l++) d1k[i][l] = d[i][l] + d1k[i][l];
			numIn1k += numIn1; numIn1 = 0; memset(d, 0, sizeof(d));
		}
		if (numIn1k > 1000 || force)
		{
			for (int i = 0; i < 45; i++) for (int l = 0; l < 4; l++) d1m[i][l] = d1k[i][l] + d1m[i][l];
			numIn1m += numIn1k; numIn1k = 0; memset(d1k, 0, sizeof(d1k));
		}
	}
	// updateSSE_eighted, :1067-1142: J[k][lane]
	void update_weighted(const float J[9][4], const float w[4])
	{
		int idx = 0;
		for (int r = 0; r < 9; r++)
		{
			float Jw[4];
			for (int l = 0; l < 4; l++) Jw[l] = J[r][l] * w[l];
			for (int c = r; c < 9; c++)
			{
				for (int l = 0; l < 4; l++) d[idx][l] = d[idx][l] + Jw[l] * J[c][l];
				idx++;
			}
		}
		numIn1++;
		shiftUp(false);
	}
	void finish()
	{
		shiftUp(true);
		int idx = 0;
		for (int r = 0; r < 9; r++)
			for (int c = r; c < 9; c++)
			{
				float v = d1m[idx][0] + d1m[idx][1] + d1m[idx][2] + d1m[idx][3];
				H[r][c] = H[c][r] = v;
				idx++;
			}
	}
};

struct dso_oracle_tracker
{
	int levels;
	int w[DSO_ORACLE_MAX_LEVELS], h[DSO_ORACLE_MAX_LEVELS];
	float fx[DSO_ORACLE_MAX_LEVELS], fy[DSO_ORACLE_MAX_LEVELS], cx[DSO_ORACLE_MAX_LEVELS], cy[DSO_ORACLE_MAX_LEVELS];
	float K[DSO_ORACLE_MAX_LEVELS][9], Ki[DSO_ORACLE_MAX_LEVELS][9];
	std::vector<float> idepth[DSO_ORACLE_MAX_LEVELS], weightSums[DSO_ORACLE_MAX_LEVELS], weightSums_bak[DSO_ORACLE_MAX_LEVELS];
	std::vector<float> pc_u[DSO_ORACLE_MAX_LEVELS], pc_v[DSO_ORACLE_MAX_LEVELS], pc_idepth[DSO_ORACLE_MAX_LEVELS], pc_color[DSO_ORACLE_MAX_LEVELS];
	int pc_n[DSO_ORACLE_MAX_LEVELS];
	std::vector<float> buf[8]; // idepth,u,v,dx,dy,residual,weight,refColor
	int buf_warped_n;
	std::vector<unsigned char> status;
	double ref_a, ref_b;     // lastRef_aff_g2l
	float ref_exposure;      // lastRef->ab_exposure
	const float* newI[DSO_ORACLE_MAX_LEVELS];
	float new_exposure;
	Acc9 acc;
	// stats
	long n_evals; double point_evals;
};

// AffLight::fromToVecExposure, NumType.h:176-189
static void fromToVecExposure(float exposureF, float exposureT, double Fa, double Fb, double Ta, double Tb, double* out)
{
	if (exposureF == 0 || exposureT == 0) exposureT = exposureF = 1;
	double a = exp(Ta - Fa) * exposureT / exposureF;
	double b = Tb - a * Fb;
	out[0] = a; out[1] = b;
}

// CoarseTracker ctor + makeK, CoarseTracker.cpp:59-131
extern "C" dso_oracle_tracker* dso_oracle_tracker_create(int ww, int hh, int levels, float fx0, float fy0, float cx0, float cy0)
{
	dso_oracle_tracker* t = new dso_oracle_tracker();
	t->levels = levels;
	t->w[0] = ww; t->h[0] = hh;
	t->fx[0] = fx0; t->fy[0] = fy0; t->cx[0] = cx0; t->cy[0] = cy0;
	for (int level = 1; level < levels; ++level)
	{
		t->w[level] = t->w[0] >> level;
		t->h[level] = t->h[0] >> level;
		t->fx[level] = t->fx[level - 1] * 0.5;   // float*double -> float, exact
		t->fy[level] = t->fy[level - 1] * 0.5;
		t->cx[level] = (t->cx[0] + 0.5) / ((int)1 << level) - 0.5;   // double arithmetic, stored to float (:118)
		t->cy[level] = (t->cy[0] + 0.5) / ((int)1 << level) - 0.5;
	}
	for (int level = 0; level < levels; ++level)
	{
		float* K = t->K[level];
		K[0] = t->fx[level]; K[1] = 0; K[2] = t->cx[level];
		K[3] = 0; K[4] = t->fy[level]; K[5] = t->cy[level];
		K[6] = 0; K[7] = 0; K[8] = 1;
		mat33f_inverse(K, t->Ki[level]);
		int n = t->w[level] * t->h[level];
		t->idepth[level].assign(n, 0); t->weightSums[level].assign(n, 0); t->weightSums_bak[level].assign(n, 0);
		t->pc_u[level].assign(n, 0); t->pc_v[level].assign(n, 0); t->pc_idepth[level].assign(n, 0); t->pc_color[level].assign(n, 0);
		t->pc_n[level] = 0;
		t->newI[level] = 0;
	}
	for (int i = 0; i < 8; i++) t->buf[i].assign(ww * hh + 8, 0);
	t->status.assign(ww * hh, 0);
	t->buf_warped_n = 0;
	t->ref_a = t->ref_b = 0; t->ref_exposure = 1; t->new_exposure = 1;
	t->n_evals = 0; t->point_evals = 0;
	return t;
}
extern "C" void dso_oracle_tracker_destroy(dso_oracle_tracker* t) { delete t; }
extern "C" void dso_oracle_tracker_get_K(const dso_oracle_tracker* t, int lvl, float* fx, float* fy, float* cx, float* cy, float* Ki9)
{
	*fx = t->fx[lvl]; *fy = t->fy[lvl]; *cx = t->cx[lvl]; *cy = t->cy[lvl];
	if (Ki9) memcpy(Ki9, t->Ki[lvl], 36);
}

// makeCoarseDepthL0, CoarseTracker.cpp:135-292
extern "C" void dso_oracle_tracker_set_ref(dso_oracle_tracker* t, int npts, const float* cpt_u, const float* cpt_v,
                                           const float* cpt_idepth, const float* HdiF, const float* const* ref_dIp,
                                           double ref_aff_a, double ref_aff_b, float ref_ab_exposure)
{
	const int* w = t->w; const int* h = t->h;
	memset(t->idepth[0].data(), 0, sizeof(float) * w[0] * h[0]);
	memset(t->weightSums[0].data(), 0, sizeof(float) * w[0] * h[0]);
	// :141-158
	for (int k = 0; k < npts; k++)
	{
		int u = cpt_u[k] + 0.5f;
		int v = cpt_v[k] + 0.5f;
		float new_idepth = cpt_idepth[k];
		float weight = sqrtf(1e-3 / (HdiF[k] + 1e-12));
		t->idepth[0][u + w[0] * v] += new_idepth * weight;
		t->weightSums[0][u + w[0] * v] += weight;
	}
	// :161-186
	for (int lvl = 1; lvl < t->levels; lvl++)
	{
		int lvlm1 = lvl - 1;
		int wl = w[lvl], hl = h[lvl], wlm1 = w[lvlm1];
		float* idepth_l = t->idepth[lvl].data();
		float* weightSums_l = t->weightSums[lvl].data();
		float* idepth_lm = t->idepth[lvlm1].data();
		float* weightSums_lm = t->weightSums[lvlm1].data();
		for (int y = 0; y < hl; y++)
			for (int x = 0; x < wl; x++)
			{
				int bidx = 2 * x + 2 * y * wlm1;
				idepth_l[x + y * wl] = idepth_lm[bidx] + idepth_lm[bidx + 1] + idepth_lm[bidx + wlm1] + idepth_lm[bidx + wlm1 + 1];
				weightSums_l[x + y * wl] = weightSums_lm[bidx] + weightSums_lm[bidx + 1] + weightSums_lm[bidx + wlm1] + weightSums_lm[bidx + wlm1 + 1];
			}
	}
	// :190-217 dilate (diagonal neighbours) on levels 0,1
	for (int lvl = 0; lvl < 2 && lvl < t->levels; lvl++)
	{
		int wh = w[lvl] * h[lvl] - w[lvl];
		int wl = w[lvl];
		float* weightSumsl = t->weightSums[lvl].data();
		float* weightSumsl_bak = t->weightSums_bak[lvl].data();
		memcpy(weightSumsl_bak, weightSumsl, w[lvl] * h[lvl] * sizeof(float));
		float* idepthl = t->idepth[lvl].data();
		for (int i = w[lvl]; i < wh; i++)
		{
			if (weightSumsl_bak[i] <= 0)
			{
				float sum = 0, num = 0, numn = 0;
				if (weightSumsl_bak[i + 1 + wl] > 0) { sum += idepthl[i + 1 + wl]; num += weightSumsl_bak[i + 1 + wl]; numn++; }
				if (weightSumsl_bak[i - 1 - wl] > 0) { sum += idepthl[i - 1 - wl]; num += weightSumsl_bak[i - 1 - wl]; numn++; }
				if (weightSumsl_bak[i + wl - 1] > 0) { sum += idepthl[i + wl - 1]; num += weightSumsl_bak[i + wl - 1]; numn++; }
				if (weightSumsl_bak[i - wl + 1] > 0) { sum += idepthl[i - wl + 1]; num += weightSumsl_bak[i - wl + 1]; numn++; }
				if (numn > 0) { idepthl[i] = sum / numn; weightSumsl[i] = num / numn; }
			}
		}
	}
	// :221-242 dilate (4-neighbours) on levels >= 2
	for (int lvl = 2; lvl < t->levels; lvl++)
	{
		int wh = w[lvl] * h[lvl] - w[lvl];
		int wl = w[lvl];
		float* weightSumsl = t->weightSums[lvl].data();
		float* weightSumsl_bak = t->weightSums_bak[lvl].data();
		memcpy(weightSumsl_bak, weightSumsl, w[lvl] * h[lvl] * sizeof(float));
		float* idepthl = t->idepth[lvl].data();
		for (int i = w[lvl]; i < wh; i++)
		{
			if (weightSumsl_bak[i] <= 0)
			{
				float sum = 0, num = 0, numn = 0;
				if (weightSumsl_bak[i + 1] > 0) { sum += idepthl[i + 1]; num += weightSumsl_bak[i + 1]; numn++; }
				if (weightSumsl_bak[i - 1] > 0) { sum += idepthl[i - 1]; num += weightSumsl_bak[i - 1]; numn++; }
				if (weightSumsl_bak[i + wl] > 0) { sum += idepthl[i + wl]; num += weightSumsl_bak[i + wl]; numn++; }
				if (weightSumsl_bak[i - wl] > 0) { sum += idepthl[i - wl]; num += weightSumsl_bak[i - wl]; numn++; }
				if (numn > 0) { idepthl[i] = sum / numn; weightSumsl[i] = num / numn; }
			}
		}
	}
	// :246-290 normalise + compact
	for (int lvl = 0; lvl < t->levels; lvl++)
	{
		float* weightSumsl = t->weightSums[lvl].data();
		float* idepthl = t->idepth[lvl].data();
		const float* dIRefl = ref_dIp[lvl];
		int wl = w[lvl], hl = h[lvl];
		int lpc_n = 0;
		float* lpc_u = t->pc_u[lvl].data();
		float* lpc_v = t->pc_v[lvl].data();
		float* lpc_idepth = t->pc_idepth[lvl].data();
		float* lpc_color = t->pc_color[lvl].data();
		for (int y = 2; y < hl - 2; y++)
			for (int x = 2; x < wl - 2; x++)
			{
				int i = x + y * wl;
				if (weightSumsl[i] > 0)
				{
					idepthl[i] /= weightSumsl[i];
					lpc_u[lpc_n] = x;
					lpc_v[lpc_n] = y;
					lpc_idepth[lpc_n] = idepthl[i];
					lpc_color[lpc_n] = dIRefl[3 * i];
					if (!std::isfinite(lpc_color[lpc_n]) || !(idepthl[i] > 0))
					{
						idepthl[i] = -1;
						continue;
					}
					lpc_n++;
				}
				else
					idepthl[i] = -1;
				weightSumsl[i] = 1;
			}
		t->pc_n[lvl] = lpc_n;
	}
	// setCoarseTrackingRef :532-533
	t->ref_a = ref_aff_a; t->ref_b = ref_aff_b; t->ref_exposure = ref_ab_exposure;
}

extern "C" void dso_oracle_tracker_set_cloud(dso_oracle_tracker* t, int lvl, int n, const float* u, const float* v,
                                             const float* idepth, const float* color)
{
	memcpy(t->pc_u[lvl].data(), u, n * 4); memcpy(t->pc_v[lvl].data(), v, n * 4);
	memcpy(t->pc_idepth[lvl].data(), idepth, n * 4); memcpy(t->pc_color[lvl].data(), color, n * 4);
	t->pc_n[lvl] = n;
}
extern "C" void dso_oracle_tracker_set_ref_aff(dso_oracle_tracker* t, double a, double b, float e) { t->ref_a = a; t->ref_b = b; t->ref_exposure = e; }
extern "C" int dso_oracle_tracker_get_cloud(const dso_oracle_tracker* t, int lvl, float* u, float* v, float* idepth, float* color)
{
	int n = t->pc_n[lvl];
	if (u) memcpy(u, t->pc_u[lvl].data(), n * 4);
	if (v) memcpy(v, t->pc_v[lvl].data(), n * 4);
	if (idepth) memcpy(idepth, t->pc_idepth[lvl].data(), n * 4);
	if (color) memcpy(color, t->pc_color[lvl].data(), n * 4);
	return n;
}
extern "C" void dso_oracle_tracker_set_new_frame(dso_oracle_tracker* t, const float* const* dIp, float ab_exposure)
{
	for (int l = 0; l < t->levels; l++) t->newI[l] = dIp[l];
	t->new_exposure = ab_exposure;
}

// getInterpolatedElement33, globalFuncs.h:75-89 (AoS Vector3f image)
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

// RKi*[x,y,1] + t*id in Eigen's evaluation order (product temporary, then sum)
static inline void proj3(const float* M, float x, float y, const float* t, float id, float* pt)
{
	for (int r = 0; r < 3; r++)
	{
		float p = dot3_eigen(M[r * 3], x, M[r * 3 + 1], y, M[r * 3 + 2], 1.0f);
		pt[r] = p + t[r] * id;
	}
}
static inline void proj3m(const float* M, float x, float y, const float* t, float id, float* pt)
{
	for (int r = 0; r < 3; r++)
	{
		float p = dot3_eigen(M[r * 3], x, M[r * 3 + 1], y, M[r * 3 + 2], 1.0f);
		pt[r] = p - t[r] * id;
	}
}

// calcRes, CoarseTracker.cpp:359-520
static void calcRes(dso_oracle_tracker* T, int lvl, const SE3d& refToNew, double aff_a, double aff_b, float cutoffTH, double* rs)
{
	float E = 0;
	int numTermsInE = 0;
	int numTermsInWarped = 0;
	int numSaturated = 0;
	int wl = T->w[lvl];
	int hl = T->h[lvl];
	const float* dINewl = T->newI[lvl];
	float fxl = T->fx[lvl], fyl = T->fy[lvl], cxl = T->cx[lvl], cyl = T->cy[lvl];

	// :375-378
	float Rf[9], RKi[9], t[3];
	for (int i = 0; i < 9; i++) Rf[i] = (float)refToNew.R[i];
	mat33f_mul(Rf, T->Ki[lvl], RKi);
	for (int i = 0; i < 3; i++) t[i] = (float)refToNew.t[i];
	double aff[2];
	fromToVecExposure(T->ref_exposure, T->new_exposure, T->ref_a, T->ref_b, aff_a, aff_b, aff);
	float affLL[2] = { (float)aff[0], (float)aff[1] };

	float sumSquaredShiftT = 0, sumSquaredShiftRT = 0, sumSquaredShiftNum = 0;
	float maxEnergy = 2 * setting_huberTH * cutoffTH - setting_huberTH * setting_huberTH;   // :385

	int nl = T->pc_n[lvl];
	const float* lpc_u = T->pc_u[lvl].data();
	const float* lpc_v = T->pc_v[lvl].data();
	const float* lpc_idepth = T->pc_idepth[lvl].data();
	const float* lpc_color = T->pc_color[lvl].data();
	float** B = 0; (void)B;
	float* b_id = T->buf[0].data(); float* b_u = T->buf[1].data(); float* b_v = T->buf[2].data();
	float* b_dx = T->buf[3].data(); float* b_dy = T->buf[4].data(); float* b_r = T->buf[5].data();
	float* b_w = T->buf[6].data(); float* b_c = T->buf[7].data();
	const float* Ki = T->Ki[lvl];

	for (int i = 0; i < nl; i++)
	{
		T->status[i] = 0;
		float id = lpc_idepth[i];
		float x = lpc_u[i];
		float y = lpc_v[i];
		float pt[3];
		proj3(RKi, x, y, t, id, pt);       // :408
		float u = pt[0] / pt[2];
		float v = pt[1] / pt[2];
		float Ku = fxl * u + cxl;
		float Kv = fyl * v + cyl;
		float new_idepth = id / pt[2];

		if (lvl == 0 && i % 32 == 0)   // :416-447
		{
			float ptT[3]; proj3(Ki, x, y, t, id, ptT);
			float uT = ptT[0] / ptT[2], vT = ptT[1] / ptT[2];
			float KuT = fxl * uT + cxl, KvT = fyl * vT + cyl;
			float ptT2[3]; proj3m(Ki, x, y, t, id, ptT2);
			float uT2 = ptT2[0] / ptT2[2], vT2 = ptT2[1] / ptT2[2];
			float KuT2 = fxl * uT2 + cxl, KvT2 = fyl * vT2 + cyl;
			float pt3[3]; proj3m(RKi, x, y, t, id, pt3);
			float u3 = pt3[0] / pt3[2], v3 = pt3[1] / pt3[2];
			float Ku3 = fxl * u3 + cxl, Kv3 = fyl * v3 + cyl;
			sumSquaredShiftT += (KuT - x) * (KuT - x) + (KvT - y) * (KvT - y);
			sumSquaredShiftT += (KuT2 - x) * (KuT2 - x) + (KvT2 - y) * (KvT2 - y);
			sumSquaredShiftRT += (Ku - x) * (Ku - x) + (Kv - y) * (Kv - y);
			sumSquaredShiftRT += (Ku3 - x) * (Ku3 - x) + (Kv3 - y) * (Kv3 - y);
			sumSquaredShiftNum += 2;
		}

		if (!(Ku > 2 && Kv > 2 && Ku < wl - 3 && Kv < hl - 3 && new_idepth > 0)) continue;   // :449

		float refColor = lpc_color[i];
		float hitColor[3];
		interp33(dINewl, Ku, Kv, wl, hitColor);
		if (!std::isfinite((float)hitColor[0])) continue;
		float residual = hitColor[0] - (float)(affLL[0] * refColor + affLL[1]);
		float hw = fabsf(residual) < setting_huberTH ? 1 : setting_huberTH / fabsf(residual);

		if (fabsf(residual) > cutoffTH)
		{
			E += maxEnergy;
			numTermsInE++;
			numSaturated++;
			T->status[i] = 1;
		}
		else
		{
			E += hw * residual * residual * (2 - hw);
			numTermsInE++;
			b_id[numTermsInWarped] = new_idepth;
			b_u[numTermsInWarped] = u;
			b_v[numTermsInWarped] = v;
			b_dx[numTermsInWarped] = hitColor[1];
			b_dy[numTermsInWarped] = hitColor[2];
			b_r[numTermsInWarped] = residual;
			b_w[numTermsInWarped] = hw;
			b_c[numTermsInWarped] = lpc_color[i];
			numTermsInWarped++;
			T->status[i] = 2;
		}
	}
	while (numTermsInWarped % 4 != 0)   // :488-499
	{
		for (int k = 0; k < 8; k++) T->buf[k][numTermsInWarped] = 0;
		numTermsInWarped++;
	}
	T->buf_warped_n = numTermsInWarped;

	rs[0] = E;
	rs[1] = numTermsInE;
	rs[2] = sumSquaredShiftT / (sumSquaredShiftNum + 0.1);
	rs[3] = 0;
	rs[4] = sumSquaredShiftRT / (sumSquaredShiftNum + 0.1);
	rs[5] = numSaturated / (float)numTermsInE;
	T->n_evals++; T->point_evals += nl;
}

// calcGSSSE, CoarseTracker.cpp:296-354
static void calcGSSSE(dso_oracle_tracker* T, int lvl, double* H_out, double* b_out, double aff_a, double aff_b)
{
	Acc9& acc = T->acc;
	acc.initialize();
	float fxl = T->fx[lvl], fyl = T->fy[lvl];
	float b0 = (float)T->ref_b;
	double aff[2];
	fromToVecExposure(T->ref_exposure, T->new_exposure, T->ref_a, T->ref_b, aff_a, aff_b, aff);
	float a = (float)aff[0];
	const float* b_id = T->buf[0].data(); const float* b_u = T->buf[1].data(); const float* b_v = T->buf[2].data();
	const float* b_dx = T->buf[3].data(); const float* b_dy = T->buf[4].data(); const float* b_r = T->buf[5].data();
	const float* b_w = T->buf[6].data(); const float* b_c = T->buf[7].data();
	int n = T->buf_warped_n;
	for (int i = 0; i < n; i += 4)
	{
		float J[9][4], w[4];
		for (int l = 0; l < 4; l++)
		{
			float dx = b_dx[i + l] * fxl;
			float dy = b_dy[i + l] * fyl;
			float u = b_u[i + l], v = b_v[i + l], id = b_id[i + l];
			J[0][l] = id * dx;
			J[1][l] = id * dy;
			J[2][l] = 0 - id * (u * dx + v * dy);
			J[3][l] = 0 - ((u * v) * dx + dy * (1 + v * v));
			J[4][l] = (u * v) * dy + dx * (1 + u * u);
			J[5][l] = u * dy - v * dx;
			J[6][l] = a * (b0 - b_c[i + l]);
			J[7][l] = -1;
			J[8][l] = b_r[i + l];
			w[l] = b_w[i + l];
		}
		acc.update_weighted(J, w);
	}
	acc.finish();
	// :338-339
	float invn = 1.0f / n;
	for (int r = 0; r < 8; r++)
	{
		for (int c = 0; c < 8; c++) H_out[r * 8 + c] = (double)acc.H[r][c] * invn;
		b_out[r] = (double)acc.H[r][8] * invn;
	}
	// :342-353
	const float sc[8] = { SCALE_XI_ROT, SCALE_XI_ROT, SCALE_XI_ROT, SCALE_XI_TRANS, SCALE_XI_TRANS, SCALE_XI_TRANS, SCALE_A, SCALE_B };
	for (int r = 0; r < 8; r++) for (int c = 0; c < 8; c++) H_out[r * 8 + c] *= sc[c];
	for (int r = 0; r < 8; r++) for (int c = 0; c < 8; c++) H_out[r * 8 + c] *= sc[r];
	for (int r = 0; r < 8; r++) b_out[r] *= sc[r];
}

extern "C" void dso_oracle_calc_res(dso_oracle_tracker* t, int lvl, const double* T12, double a, double b, float cutoffTH, double* out6)
{
	calcRes(t, lvl, SE3d::from12(T12), a, b, cutoffTH, out6);
}
extern "C" void dso_oracle_calc_gs(dso_oracle_tracker* t, int lvl, double a, double b, double* H64, double* b8)
{
	calcGSSSE(t, lvl, H64, b8, a, b);
}
extern "C" int dso_oracle_get_warped(const dso_oracle_tracker* t, float* const* bufs8)
{
	for (int k = 0; k < 8; k++) if (bufs8[k]) memcpy(bufs8[k], t->buf[k].data(), sizeof(float) * t->buf_warped_n);
	return t->buf_warped_n;
}
extern "C" void dso_oracle_get_point_status(const dso_oracle_tracker* t, unsigned char* status)
{
	memcpy(status, t->status.data(), t->w[0] * t->h[0]);
}

// trackNewestCoarse, CoarseTracker.cpp:551-747
extern "C" void dso_oracle_track_newest_coarse(dso_oracle_tracker* T, const double* pose12_in, double aff_a_in, double aff_b_in,
                                               int coarsestLvl, const double* minResForAbort, dso_oracle_track_result* out)
{
	memset(out, 0, sizeof(*out));
	out->aborted_level = -1;
	for (int i = 0; i < 5; i++) out->lastResiduals[i] = NAN;     // :563
	for (int i = 0; i < 3; i++) out->lastFlowIndicators[i] = 1000; // :564
	memcpy(out->pose, pose12_in, 96); out->aff_a = aff_a_in; out->aff_b = aff_b_in;
	long evals0 = T->n_evals; double pe0 = T->point_evals;

	int maxIterations[] = { 10, 20, 50, 50, 50 };
	float lambdaExtrapolationLimit = 0.001;
	SE3d refToNew_current = SE3d::from12(pose12_in);
	double a_cur = aff_a_in, b_cur = aff_b_in;
	bool haveRepeated = false;

	for (int lvl = coarsestLvl; lvl >= 0; lvl--)
	{
		double H[64], b[8];
		float levelCutoffRepeat = 1;
		double resOld[6];
		calcRes(T, lvl, refToNew_current, a_cur, b_cur, setting_coarseCutoffTH * levelCutoffRepeat, resOld);
		while (resOld[5] > 0.6 && levelCutoffRepeat < 50)   // :586-593
		{
			levelCutoffRepeat *= 2;
			calcRes(T, lvl, refToNew_current, a_cur, b_cur, setting_coarseCutoffTH * levelCutoffRepeat, resOld);
		}
		calcGSSSE(T, lvl, H, b, a_cur, b_cur);
		float lambda = 0.01;

		for (int iteration = 0; iteration < maxIterations[lvl]; iteration++)
		{
			double Hl[64];
			memcpy(Hl, H, sizeof(Hl));
			for (int i = 0; i < 8; i++) Hl[i * 8 + i] *= (1 + lambda);
			double nb[8], inc[8];
			for (int i = 0; i < 8; i++) nb[i] = -b[i];
			ldlt_solve(Hl, nb, inc, 8);   // :618
			// :621-643 fixed-affine variants (not reached with the default modes 1e12 / 1e8)
			if (setting_affineOptModeA < 0 && setting_affineOptModeB < 0)
			{ ldlt_solve(Hl, nb, inc, 6); inc[6] = inc[7] = 0; }
			if (!(setting_affineOptModeA < 0) && setting_affineOptModeB < 0)
			{ ldlt_solve(Hl, nb, inc, 7); inc[7] = 0; }
			if (setting_affineOptModeA < 0 && !(setting_affineOptModeB < 0))
			{
				double Hs[64], bs[8], is[8];
				memcpy(Hs, Hl, sizeof(Hs)); memcpy(bs, nb, sizeof(bs));
				for (int i = 0; i < 8; i++) Hs[i * 8 + 6] = Hs[i * 8 + 7];
				for (int i = 0; i < 8; i++) Hs[6 * 8 + i] = Hs[7 * 8 + i];
				bs[6] = bs[7];
				ldlt_solve(Hs, bs, is, 7);
				for (int i = 0; i < 6; i++) inc[i] = is[i];
				inc[6] = 0; inc[7] = is[6];
			}

			float extrapFac = 1;
			if (lambda < lambdaExtrapolationLimit) extrapFac = sqrtf(sqrtf(lambdaExtrapolationLimit / lambda));   // :650 (CoarseTracker.h:30 pulls in <math.h> => float overloads)
			for (int i = 0; i < 8; i++) inc[i] *= extrapFac;

			double incScaled[8];
			for (int i = 0; i < 8; i++) incScaled[i] = inc[i];
			for (int i = 0; i < 3; i++) incScaled[i] *= SCALE_XI_ROT;
			for (int i = 3; i < 6; i++) incScaled[i] *= SCALE_XI_TRANS;
			incScaled[6] *= SCALE_A;
			incScaled[7] *= SCALE_B;
			double s = 0; for (int i = 0; i < 8; i++) s += incScaled[i];
			if (!std::isfinite(s)) for (int i = 0; i < 8; i++) incScaled[i] = 0;

			SE3d refToNew_new = se3_exp(incScaled) * refToNew_current;   // :661
			double a_new = a_cur + incScaled[6];
			double b_new = b_cur + incScaled[7];

			double resNew[6];
			calcRes(T, lvl, refToNew_new, a_new, b_new, setting_coarseCutoffTH * levelCutoffRepeat, resNew);
			bool accept = (resNew[0] / resNew[1]) < (resOld[0] / resOld[1]);   // :669
			if (out->trace_len < 64)
			{
				out->trace_dE[out->trace_len] = (float)(resNew[0] * resOld[1] - resOld[0] * resNew[1]);
				out->trace_E[out->trace_len] = (float)(resOld[0] * resNew[1]);
			}
			if (out->trace_len < 256) out->trace[out->trace_len++] = accept ? 1 : 0;
			out->n_iterations++;

			if (accept)
			{
				calcGSSSE(T, lvl, H, b, a_new, b_new);
				memcpy(resOld, resNew, sizeof(resOld));
				a_cur = a_new; b_cur = b_new;
				refToNew_current = refToNew_new;
				lambda *= 0.5;
			}
			else
			{
				lambda *= 4;
				if (lambda < lambdaExtrapolationLimit) lambda = lambdaExtrapolationLimit;
			}
			double nrm = 0; for (int i = 0; i < 8; i++) nrm += inc[i] * inc[i];
			nrm = std::sqrt(nrm);
			if (!(nrm > 1e-3)) break;   // :700
		}

		out->lastResiduals[lvl] = sqrtf((float)(resOld[0] / resOld[1]));   // :710
		out->lastFlowIndicators[0] = resOld[2]; out->lastFlowIndicators[1] = resOld[3]; out->lastFlowIndicators[2] = resOld[4];
		if (out->n_passes < 8) { out->pass_level[out->n_passes] = lvl; out->pass_rmse[out->n_passes] = out->lastResiduals[lvl]; out->n_passes++; }
		if (out->lastResiduals[lvl] > 1.5 * minResForAbort[lvl])   // :713
		{
			out->ok = 0; out->aborted_level = lvl;
			out->n_evals = (int)(T->n_evals - evals0); out->point_evals = T->point_evals - pe0;
			return;
		}
		if (levelCutoffRepeat > 1 && !haveRepeated)   // :717-722
		{
			lvl++;
			haveRepeated = true;
		}
	}

	refToNew_current.to12(out->pose);   // :727-728
	out->aff_a = a_cur; out->aff_b = b_cur;
	out->n_evals = (int)(T->n_evals - evals0); out->point_evals = T->point_evals - pe0;
	out->ok = 1;
	if ((setting_affineOptModeA != 0 && (fabsf(out->aff_a) > 1.2)) || (setting_affineOptModeB != 0 && (fabsf(out->aff_b) > 200)))
	{ out->ok = 0; return; }
	double rel[2];
	fromToVecExposure(T->ref_exposure, T->new_exposure, T->ref_a, T->ref_b, out->aff_a, out->aff_b, rel);
	float relAff[2] = { (float)rel[0], (float)rel[1] };
	if ((setting_affineOptModeA == 0 && (fabsf(logf((float)relAff[0])) > 1.5)) || (setting_affineOptModeB == 0 && (fabsf((float)relAff[1]) > 200)))
	{ out->ok = 0; return; }
	if (setting_affineOptModeA < 0) out->aff_a = 0;
	if (setting_affineOptModeB < 0) out->aff_b = 0;
}
