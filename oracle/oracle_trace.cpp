// TEST INFRASTRUCTURE ONLY -- CPU oracle for the DSO photometric Gauss-Newton hot path (next row N3, second half).
// Nothing in the product (dso-noted-in-chinese_b200/, include/) may include or link this file.
//
// Restatement of ImmaturePoint's constructor (FullSystem/ImmaturePoint.cpp:25-57) and ImmaturePoint::traceOn
// (FullSystem/ImmaturePoint.cpp:77-438): the epipolar-line search that FullSystem::traceNewCoarse (FullSystem.cpp:486-530)
// runs for every immature point of every key frame on each new frame.  Plain float arithmetic, one rounding per operation
// (-ffp-contract=off), in the reference's order; pinned bit for bit against the reference build in
// tests/test_oracle_vs_reference.py.
#include "dso_oracle.h"
#include "oracle_math.h"
#include "oracle_settings.h"
#include <cmath>
#include <cstring>
#include <algorithm>

namespace {

const int patternNum = 8;
const int patternP[8][2] = { {0,-2}, {-1,-1}, {1,-1}, {-2,0}, {0,0}, {2,0}, {-1,1}, {0,2} };   // settings.cpp:274 (pattern 8)
// util/settings.cpp:90-91, 137, 156-165, 113
const float setting_outlierTH = 12 * 12;
const float setting_overallEnergyTHWeight = 1;
// the remaining settings of :91, :127, :156-165 are run-time values (oracle_settings.h)

enum { IPS_GOOD = 0, IPS_OOB, IPS_OUTLIER, IPS_SKIPPED, IPS_BADCONDITION, IPS_UNINITIALIZED };

// util/globalFuncs.h:126-140 (channel 0 only) and :75-89 (all three channels) on the AoS Vector3f image
inline float interp31(const float* mat, float x, float y, int width)
{
	int ix = (int)x, iy = (int)y;
	float dx = x - ix, dy = y - iy;
	float dxdy = dx * dy;
	const float* bp = mat + 3 * (ix + iy * width);
	return dxdy * bp[3 * (1 + width)] + (dy - dxdy) * bp[3 * width] + (dx - dxdy) * bp[3] + (1 - dx - dy + dxdy) * bp[0];
}
inline void interp33(const float* mat, float x, float y, int width, float* out)
{
	int ix = (int)x, iy = (int)y;
	float dx = x - ix, dy = y - iy;
	float dxdy = dx * dy;
	const float* bp = mat + 3 * (ix + iy * width);
	for (int c = 0; c < 3; c++)
		out[c] = dxdy * bp[3 * (1 + width) + c] + (dy - dxdy) * bp[3 * width + c] + (dx - dxdy) * bp[3 + c] + (1 - dx - dy + dxdy) * bp[c];
}
// util/globalFuncs.h:166-188
inline void interp33BiLin(const float* mat, float x, float y, int width, float* out)
{
	int ix = (int)x, iy = (int)y;
	const float* bp = mat + 3 * (ix + iy * width);
	float tl = bp[0], tr = bp[3], bl = bp[3 * width], br = bp[3 * (width + 1)];
	float dx = x - ix, dy = y - iy;
	float topInt = dx * tr + (1 - dx) * tl;
	float botInt = dx * br + (1 - dx) * bl;
	float leftInt = dy * bl + (1 - dy) * tl;
	float rightInt = dy * br + (1 - dy) * tr;
	out[0] = dx * rightInt + (1 - dx) * leftInt;
	out[1] = rightInt - leftInt;
	out[2] = botInt - topInt;
}

}   // namespace

// ImmaturePoint::ImmaturePoint(int u_, int v_, FrameHessian* host_, float type, CalibHessian*), ImmaturePoint.cpp:25-57
extern "C" void dso_oracle_immature_init(int w, int h, const float* host_dI, int n, const int* u, const int* v,
                                         float* color8, float* weights8, float* gradH4, float* energyTH)
{
	(void)h;
	for (int k = 0; k < n; k++)
	{
		float* color = color8 + 8 * k; float* weights = weights8 + 8 * k; float* G = gradH4 + 4 * k;
		G[0] = G[1] = G[2] = G[3] = 0;
		bool bad = false;
		for (int idx = 0; idx < patternNum && !bad; idx++)
		{
			int dx = patternP[idx][0], dy = patternP[idx][1];
			float ptc[3];
			interp33BiLin(host_dI, (float)(u[k] + dx), (float)(v[k] + dy), w, ptc);
			color[idx] = ptc[0];
			if (!std::isfinite(color[idx])) { energyTH[k] = NAN; bad = true; break; }
			G[0] += ptc[1] * ptc[1]; G[1] += ptc[1] * ptc[2]; G[2] += ptc[2] * ptc[1]; G[3] += ptc[2] * ptc[2];   // gradH += g g^T
			weights[idx] = sqrtf(setting_outlierTHSumComponent / (setting_outlierTHSumComponent + (ptc[1] * ptc[1] + ptc[2] * ptc[2])));
		}
		if (bad) continue;
		float e = patternNum * setting_outlierTH;
		e *= setting_overallEnergyTHWeight * setting_overallEnergyTHWeight;
		energyTH[k] = e;
	}
}

namespace {
struct IP
{
	float u, v, idepth_min, idepth_max, quality, energyTH;
	const float* gradH; const float* color; const float* weights;
	int lastTraceStatus; float lastTraceUV[2]; float lastTracePixelInterval;
};

inline int fail(IP& p, int status) { p.lastTraceUV[0] = -1; p.lastTraceUV[1] = -1; p.lastTracePixelInterval = 0; return p.lastTraceStatus = status; }

// ImmaturePoint::traceOn, ImmaturePoint.cpp:77-438
int trace_on(IP& p, const float* dI, int wG0, int hG0, const float* KRKi, const float* Kt, const float* aff)
{
	if (p.lastTraceStatus == IPS_OOB) return p.lastTraceStatus;                       // :79
	float maxPixSearch = (wG0 + hG0) * setting_maxPixSearch;                           // :85
	float uv1[3] = { p.u, p.v, 1 };
	float pr[3];
	orc::mat33f_vec(KRKi, uv1, pr);                                                    // :96
	float ptpMin[3];
	for (int i = 0; i < 3; i++) ptpMin[i] = pr[i] + Kt[i] * p.idepth_min;              // :98
	float uMin = ptpMin[0] / ptpMin[2];
	float vMin = ptpMin[1] / ptpMin[2];
	if (!(uMin > 4 && vMin > 4 && uMin < wG0 - 5 && vMin < hG0 - 5)) return fail(p, IPS_OOB);   // :103-110
	float dist, uMax, vMax, ptpMax[3];
	if (std::isfinite(p.idepth_max))
	{
		for (int i = 0; i < 3; i++) ptpMax[i] = pr[i] + Kt[i] * p.idepth_max;          // :119
		uMax = ptpMax[0] / ptpMax[2];
		vMax = ptpMax[1] / ptpMax[2];
		if (!(uMax > 4 && vMax > 4 && uMax < wG0 - 5 && vMax < hG0 - 5)) return fail(p, IPS_OOB);   // :124-130
		dist = (uMin - uMax) * (uMin - uMax) + (vMin - vMax) * (vMin - vMax);          // :135
		dist = sqrtf(dist);
		if (dist < setting_trace_slackInterval)                                        // :138-145
		{
			p.lastTraceUV[0] = (uMax + uMin) * 0.5f; p.lastTraceUV[1] = (vMax + vMin) * 0.5f;
			p.lastTracePixelInterval = dist;
			return p.lastTraceStatus = IPS_SKIPPED;
		}
	}
	else
	{
		dist = maxPixSearch;                                                           // :150
		for (int i = 0; i < 3; i++) ptpMax[i] = pr[i] + Kt[i] * 0.01f;                 // :153 (Vec3f * 0.01 -> float scalar)
		uMax = ptpMax[0] / ptpMax[2];
		vMax = ptpMax[1] / ptpMax[2];
		float dx = uMax - uMin, dy = vMax - vMin;
		float d = 1.0f / sqrtf(dx * dx + dy * dy);
		uMax = uMin + dist * dx * d;                                                   // :162
		vMax = vMin + dist * dy * d;
		if (!(uMax > 4 && vMax > 4 && uMax < wG0 - 5 && vMax < hG0 - 5)) return fail(p, IPS_OOB);   // :166-172
	}
	if (!(p.idepth_min < 0 || (ptpMin[2] > 0.75 && ptpMin[2] < 1.5))) return fail(p, IPS_OOB);       // :179-185
	float dx = setting_trace_stepsize * (uMax - uMin);                                 // :196
	float dy = setting_trace_stepsize * (vMax - vMin);
	const float* G = p.gradH;
	float a, b;
	{
		float t0 = dx * G[0] + dy * G[2], t1 = dx * G[1] + dy * G[3];                  // (dx,dy) gradH (dx,dy)^T  :199
		a = t0 * dx + t1 * dy;
		float s0 = dy * G[0] + (-dx) * G[2], s1 = dy * G[1] + (-dx) * G[3];            // (dy,-dx) gradH (dy,-dx)^T :200
		b = s0 * dy + s1 * (-dx);
	}
	float errorInPixel = 0.2f + 0.2f * (a + b) / a;                                    // :201
	if (errorInPixel * setting_trace_minImprovementFactor > dist && std::isfinite(p.idepth_max))   // :203-209
	{
		p.lastTraceUV[0] = (uMax + uMin) * 0.5f; p.lastTraceUV[1] = (vMax + vMin) * 0.5f;
		p.lastTracePixelInterval = dist;
		return p.lastTraceStatus = IPS_BADCONDITION;
	}
	if (errorInPixel > 10) errorInPixel = 10;
	dx /= dist;                                                                        // :215
	dy /= dist;
	if (dist > maxPixSearch)                                                           // :227-232
	{
		uMax = uMin + maxPixSearch * dx;
		vMax = vMin + maxPixSearch * dy;
		dist = maxPixSearch;
	}
	int numSteps = 1.9999f + dist / setting_trace_stepsize;                            // :234
	float R00 = KRKi[0], R01 = KRKi[1], R10 = KRKi[3], R11 = KRKi[4];                  // Rplane :236
	float randShift = uMin * 1000 - floorf(uMin * 1000);                               // :238
	float ptx = uMin - randShift * dx;
	float pty = vMin - randShift * dy;
	float rot[8][2];
	for (int idx = 0; idx < patternNum; idx++)
	{
		float px = (float)patternP[idx][0], py = (float)patternP[idx][1];
		rot[idx][0] = R00 * px + R01 * py;                                             // :245
		rot[idx][1] = R10 * px + R11 * py;
	}
	if (!std::isfinite(dx) || !std::isfinite(dy)) return fail(p, IPS_OOB);             // :250-256
	float errors[100];
	float bestU = 0, bestV = 0, bestEnergy = 1e10;
	int bestIdx = -1;
	if (numSteps >= 100) numSteps = 99;                                                // :264
	for (int i = 0; i < numSteps; i++)                                                 // :267-300 discrete search along the line
	{
		float energy = 0;
		for (int idx = 0; idx < patternNum; idx++)
		{
			float hitColor = interp31(dI, (float)(ptx + rot[idx][0]), (float)(pty + rot[idx][1]), wG0);
			if (!std::isfinite(hitColor)) { energy += 1e5; continue; }
			float residual = hitColor - (float)(aff[0] * p.color[idx] + aff[1]);
			float hw = std::fabs(residual) < setting_huberTH ? 1 : setting_huberTH / std::fabs(residual);
			energy += hw * residual * residual * (2 - hw);
		}
		errors[i] = energy;
		if (energy < bestEnergy) { bestU = ptx; bestV = pty; bestEnergy = energy; bestIdx = i; }
		ptx += dx;
		pty += dy;
	}
	float secondBest = 1e10;                                                           // :303-309
	for (int i = 0; i < numSteps; i++)
		if ((i < bestIdx - setting_minTraceTestRadius || i > bestIdx + setting_minTraceTestRadius) && errors[i] < secondBest)
			secondBest = errors[i];
	float newQuality = secondBest / bestEnergy;
	if (newQuality < p.quality || numSteps > 10) p.quality = newQuality;               // :311
	float uBak = bestU, vBak = bestV, gnstepsize = 1, stepBack = 0;                    // :315
	if (setting_trace_GNIterations > 0) bestEnergy = 1e5;
	for (int it = 0; it < setting_trace_GNIterations; it++)                            // :318-375 Gauss-Newton along the line
	{
		float H = 1, bb = 0, energy = 0;
		for (int idx = 0; idx < patternNum; idx++)
		{
			float hit[3];
			interp33(dI, (float)(bestU + rot[idx][0]), (float)(bestV + rot[idx][1]), wG0, hit);
			if (!std::isfinite((float)hit[0])) { energy += 1e5; continue; }
			float residual = hit[0] - (aff[0] * p.color[idx] + aff[1]);
			float dResdDist = dx * hit[1] + dy * hit[2];
			float hw = std::fabs(residual) < setting_huberTH ? 1 : setting_huberTH / std::fabs(residual);
			H += hw * dResdDist * dResdDist;
			bb += hw * residual * dResdDist;
			energy += p.weights[idx] * p.weights[idx] * hw * residual * residual * (2 - hw);
		}
		if (energy > bestEnergy)                                                       // :341-351
		{
			stepBack *= 0.5f;
			bestU = uBak + stepBack * dx;
			bestV = vBak + stepBack * dy;
		}
		else                                                                           // :352-372
		{
			float step = -gnstepsize * bb / H;
			if (step < -0.5) step = -0.5;
			else if (step > 0.5) step = 0.5;
			if (!std::isfinite(step)) step = 0;
			uBak = bestU;
			vBak = bestV;
			stepBack = step;
			bestU += step * dx;
			bestV += step * dy;
			bestEnergy = energy;
		}
		if (fabsf(stepBack) < setting_trace_GNThreshold) break;                        // :374
	}
	if (!(bestEnergy < p.energyTH * setting_trace_extraSlackOnTH))                     // :380-391
	{
		p.lastTracePixelInterval = 0;
		p.lastTraceUV[0] = -1; p.lastTraceUV[1] = -1;
		if (p.lastTraceStatus == IPS_OUTLIER) return p.lastTraceStatus = IPS_OOB;
		return p.lastTraceStatus = IPS_OUTLIER;
	}
	if (dx * dx > dy * dy)                                                             // :395-404 interval in inverse depth
	{
		p.idepth_min = (pr[2] * (bestU - errorInPixel * dx) - pr[0]) / (Kt[0] - Kt[2] * (bestU - errorInPixel * dx));
		p.idepth_max = (pr[2] * (bestU + errorInPixel * dx) - pr[0]) / (Kt[0] - Kt[2] * (bestU + errorInPixel * dx));
	}
	else
	{
		p.idepth_min = (pr[2] * (bestV - errorInPixel * dy) - pr[1]) / (Kt[1] - Kt[2] * (bestV - errorInPixel * dy));
		p.idepth_max = (pr[2] * (bestV + errorInPixel * dy) - pr[1]) / (Kt[1] - Kt[2] * (bestV + errorInPixel * dy));
	}
	if (p.idepth_min > p.idepth_max) std::swap(p.idepth_min, p.idepth_max);            // :405
	if (!std::isfinite(p.idepth_min) || !std::isfinite(p.idepth_max) || (p.idepth_max < 0))   // :408-415
		return fail(p, IPS_OUTLIER);
	p.lastTracePixelInterval = 2 * errorInPixel;                                       // :417
	p.lastTraceUV[0] = bestU; p.lastTraceUV[1] = bestV;
	return p.lastTraceStatus = IPS_GOOD;
}
}   // namespace

// traceOn for n points of one host frame against one new frame (the inner loop of FullSystem::traceNewCoarse,
// FullSystem.cpp:497-520).  In/out: idepth_min, idepth_max, quality, status; out: lastTraceUV (2n), lastTracePixelInterval.
extern "C" void dso_oracle_trace_on(int w, int h, const float* frame_dI, int n, const float* u, const float* v,
                                    const float* color8, const float* weights8, const float* gradH4, const float* energyTH,
                                    float* idepth_min, float* idepth_max, float* quality, int* status,
                                    const float* KRKi9, const float* Kt3, const float* aff2,
                                    float* lastTraceUV2, float* lastTracePixelInterval)
{
	for (int k = 0; k < n; k++)
	{
		IP p;
		p.u = u[k]; p.v = v[k]; p.idepth_min = idepth_min[k]; p.idepth_max = idepth_max[k]; p.quality = quality[k]; p.energyTH = energyTH[k];
		p.gradH = gradH4 + 4 * k; p.color = color8 + 8 * k; p.weights = weights8 + 8 * k;
		p.lastTraceStatus = status[k]; p.lastTraceUV[0] = lastTraceUV2[2 * k]; p.lastTraceUV[1] = lastTraceUV2[2 * k + 1];
		p.lastTracePixelInterval = lastTracePixelInterval[k];
		trace_on(p, frame_dI, w, h, KRKi9, Kt3, aff2);
		idepth_min[k] = p.idepth_min; idepth_max[k] = p.idepth_max; quality[k] = p.quality; status[k] = p.lastTraceStatus;
		lastTraceUV2[2 * k] = p.lastTraceUV[0]; lastTraceUV2[2 * k + 1] = p.lastTraceUV[1];
		lastTracePixelInterval[k] = p.lastTracePixelInterval;
	}
}

// ---------------------------------------------------------------------------------------------------------
// FullSystem::optimizeImmaturePoint (FullSystemOptPoint.cpp:51-209) with ImmaturePoint::linearizeResidual
// (ImmaturePoint.cpp:499-570), projectPoint (ResidualProjections.h:69-101) and derive_idepth (:39-47)
// ---------------------------------------------------------------------------------------------------------
namespace {
// setting_minIdepthH_act (settings.cpp:67) and setting_GNItsOnPointActivation (:159): oracle_settings.h
enum { RS_IN = 0, RS_OOB = 1, RS_OUTLIER = 2 };

struct Calib { float fxl, fyl, cxl, cyl, fxli, fyli, wM3G, hM3G; int w; };
struct Target { const float* dI; const float* RTll; const float* tTll; const float* aff; };
struct TmpRes { int state_state, state_NewState; float state_energy, state_NewEnergy; };   // the energies are doubles holding floats

// ImmaturePoint::linearizeResidual: adds to Hdd / bd as it goes (also when it leaves early), returns the energy
double linearize_residual(const Calib& C, const Target& T, float pu, float pv, const float* color, const float* weights, float energyTH,
                          float outlierTHSlack, TmpRes& tr, float& Hdd, float& bd, float idepth)
{
	if (tr.state_state == RS_OOB) { tr.state_NewState = RS_OOB; return tr.state_energy; }       // :505-506
	float energyLeft = 0;
	for (int idx = 0; idx < patternNum; idx++)
	{
		int dx = patternP[idx][0], dy = patternP[idx][1];
		// projectPoint, ResidualProjections.h:69-101
		float KliP[3] = { (pu + dx - C.cxl) * C.fxli, (pv + dy - C.cyl) * C.fyli, 1 };
		float Rk[3];
		orc::mat33f_vec(T.RTll, KliP, Rk);
		float ptp[3];
		for (int i = 0; i < 3; i++) ptp[i] = Rk[i] + T.tTll[i] * idepth;
		float drescale = 1.0f / ptp[2];
		bool ok = drescale > 0;
		float u = 0, v = 0, Ku = 0, Kv = 0;
		if (ok)
		{
			u = ptp[0] * drescale;
			v = ptp[1] * drescale;
			Ku = u * C.fxl + C.cxl;
			Kv = v * C.fyl + C.cyl;
			ok = Ku > 1.1f && Kv > 1.1f && Ku < C.wM3G && Kv < C.hM3G;
		}
		if (!ok) { tr.state_NewState = RS_OOB; return tr.state_energy; }                        // :524-526
		float hit[3];
		interp33(T.dI, Ku, Kv, C.w, hit);
		if (!std::isfinite((float)hit[0])) { tr.state_NewState = RS_OOB; return tr.state_energy; }   // :531
		float residual = hit[0] - (T.aff[0] * color[idx] + T.aff[1]);
		float hw = fabsf(residual) < setting_huberTH ? 1 : setting_huberTH / fabsf(residual);
		energyLeft += weights[idx] * weights[idx] * hw * residual * residual * (2 - hw);
		float dxInterp = hit[1] * C.fxl;
		float dyInterp = hit[2] * C.fyl;
		float d_idepth = (dxInterp * drescale * (T.tTll[0] - T.tTll[2] * u) + dyInterp * drescale * (T.tTll[1] - T.tTll[2] * v)) * 1.0f;   // SCALE_IDEPTH
		hw *= weights[idx] * weights[idx];
		Hdd += (hw * d_idepth) * d_idepth;
		bd += (hw * residual) * d_idepth;
	}
	if (energyLeft > energyTH * outlierTHSlack) { energyLeft = energyTH * outlierTHSlack; tr.state_NewState = RS_OUTLIER; }   // :555-563
	else tr.state_NewState = RS_IN;
	tr.state_NewEnergy = energyLeft;
	return energyLeft;
}
}   // namespace

// result per point: 0 = returned null (:90-97 / :117-124), -1 = "(PointHessian*)-1" (:160-176), 1 = a PointHessian would be
// created with idepth_out; res_state_out[k*nT + t] = final state_state of the residual to target t (ResState)
extern "C" void dso_oracle_optimize_immature(int w, int h, float fx, float fy, float cx, float cy, int nT, const float* const* target_dI,
                                             const float* RTll9, const float* tTll3, const float* aff2,
                                             int n, const float* u, const float* v, const float* color8, const float* weights8,
                                             const float* energyTH, const float* idepth_min, const float* idepth_max, int minObs,
                                             int* result, float* idepth_out, int* res_state_out)
{
	Calib C;
	C.fxl = fx; C.fyl = fy; C.cxl = cx; C.cyl = cy; C.fxli = 1.0f / fx; C.fyli = 1.0f / fy;     // CalibHessian::setValueScaled
	C.wM3G = (float)(w - 3); C.hM3G = (float)(h - 3); C.w = w;                                   // globalCalib.cpp:70-71
	for (int k = 0; k < n; k++)
	{
		const float* color = color8 + 8 * k; const float* weights = weights8 + 8 * k;
		TmpRes res[16];
		Target T[16];
		const int nres = nT;
		for (int t = 0; t < nres; t++)
		{
			res[t].state_NewEnergy = res[t].state_energy = 0; res[t].state_NewState = RS_OUTLIER; res[t].state_state = RS_IN;   // :57-66
			T[t].dI = target_dI[t]; T[t].RTll = RTll9 + 9 * t; T[t].tTll = tTll3 + 3 * t; T[t].aff = aff2 + 2 * t;
		}
		float lastEnergy = 0, lastHdd = 0, lastbd = 0;
		float currentIdepth = (idepth_max[k] + idepth_min[k]) * 0.5f;                            // :76
		for (int i = 0; i < nres; i++)
		{
			lastEnergy += linearize_residual(C, T[i], u[k], v[k], color, weights, energyTH[k], 1000, res[i], lastHdd, lastbd, currentIdepth);
			res[i].state_state = res[i].state_NewState;
			res[i].state_energy = res[i].state_NewEnergy;
		}
		result[k] = 1; idepth_out[k] = NAN;
		for (int t = 0; t < nres; t++) res_state_out[k * nT + t] = res[t].state_state;
		if (!std::isfinite(lastEnergy) || lastHdd < setting_minIdepthH_act) { result[k] = 0; continue; }    // :90-97
		float lambda = 0.1;
		bool null_ret = false;
		for (int iteration = 0; iteration < setting_GNItsOnPointActivation; iteration++)
		{
			float H = lastHdd;
			H *= 1 + lambda;
			float step = (1.0 / H) * lastbd;                                                     // :106 (double reciprocal, double product)
			float newIdepth = currentIdepth - step;
			float newHdd = 0, newbd = 0, newEnergy = 0;
			for (int i = 0; i < nres; i++)
				newEnergy += linearize_residual(C, T[i], u[k], v[k], color, weights, energyTH[k], 1, res[i], newHdd, newbd, newIdepth);
			if (!std::isfinite(lastEnergy) || newHdd < setting_minIdepthH_act) { null_ret = true; break; }   // :117-124
			if (newEnergy < lastEnergy)
			{
				currentIdepth = newIdepth; lastHdd = newHdd; lastbd = newbd; lastEnergy = newEnergy;
				for (int i = 0; i < nres; i++) { res[i].state_state = res[i].state_NewState; res[i].state_energy = res[i].state_NewEnergy; }
				lambda *= 0.5;
			}
			else lambda *= 5;
			if (fabsf(step) < 0.0001 * currentIdepth) break;                                     // :151
		}
		for (int t = 0; t < nres; t++) res_state_out[k * nT + t] = res[t].state_state;
		if (null_ret) { result[k] = 0; continue; }
		if (!std::isfinite(currentIdepth)) { result[k] = -1; continue; }                        // :155-159
		int numGoodRes = 0;
		for (int i = 0; i < nres; i++) if (res[i].state_state == RS_IN) numGoodRes++;
		if (numGoodRes < minObs) { result[k] = -1; continue; }                                   // :162-170
		if (!std::isfinite(energyTH[k])) { result[k] = -1; continue; }                           // :174 (PointHessian copies energyTH)
		idepth_out[k] = currentIdepth;                                                           // :181-182
	}
}
