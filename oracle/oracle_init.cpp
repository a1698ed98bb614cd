// TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's two-frame initialiser linearisation (next row N4):
//   CoarseInitializer::calcResAndGS   FullSystem/CoarseInitializer.cpp:370-732
//   CoarseInitializer::makeK          :1126-1155 (double-precision intrinsics pyramid)
//   Accumulator9::updateSSE / updateSingleWeighted / finish and Accumulator11 (OptimizationBackend/MatrixAccumulators.h)
// Used by tests/ and bench.py's CPU legs only; pinned bit for bit against the reference's own translation unit
// (oracle/refbuild, tests/test_oracle_vs_reference.py).  Compiled with -ffp-contract=off: every float operation is a
// separate IEEE operation in the reference's order.
#include "dso_oracle.h"
#include "oracle_math.h"
#include "oracle_settings.h"
#include <cmath>
#include <cstring>
#include <vector>

using namespace orc;

namespace {

const int patternP[8][2] = { {0,-2}, {-1,-1}, {1,-1}, {-2,0}, {0,0}, {2,0}, {-1,1}, {0,2} };   // settings.cpp:274 (pattern 8)

// util/globalFuncs.h:75-89 / :126-140 on the AoS Vector3f image
inline void interp33(const float* mat, float x, float y, int width, float* out)
{
	int ix = (int)x, iy = (int)y;
	float dx = x - ix, dy = y - iy, dxdy = dx * dy;
	const float* bp = mat + 3 * (ix + iy * width);
	for (int c = 0; c < 3; c++)
		out[c] = dxdy * bp[3 * (1 + width) + c] + (dy - dxdy) * bp[3 * width + c] + (dx - dxdy) * bp[3 + c] + (1 - dx - dy + dxdy) * bp[c];
}
inline float interp31(const float* mat, float x, float y, int width)
{
	int ix = (int)x, iy = (int)y;
	float dx = x - ix, dy = y - iy, dxdy = dx * dy;
	const float* bp = mat + 3 * (ix + iy * width);
	return dxdy * bp[3 * (1 + width)] + (dy - dxdy) * bp[3 * width] + (dx - dxdy) * bp[3] + (1 - dx - dy + dxdy) * bp[0];
}

// Accumulator9 (MatrixAccumulators.h): 45 x 4 float lanes in three tiers
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
	// updateSSE: J[k][lane], all four lanes
	void updateSSE(const float J[9][4])
	{
		int idx = 0;
		for (int r = 0; r < 9; r++)
			for (int c = r; c < 9; c++)
			{
				for (int l = 0; l < 4; l++) d[idx][l] = d[idx][l] + J[r][l] * J[c][l];
				idx++;
			}
		numIn1++;
		shiftUp(false);
	}
	// updateSingleWeighted: lane 0 only; row r: first J_r*J_r*w, then J_r *= w, then J_c*J_r
	void updateSingleWeighted(const float Jin[9], float w)
	{
		float J[9];
		for (int k = 0; k < 9; k++) J[k] = Jin[k];
		int idx = 0;
		for (int r = 0; r < 9; r++)
		{
			d[idx][0] += J[r] * J[r] * w; idx++;
			J[r] *= w;
			for (int c = r + 1; c < 9; c++) { d[idx][0] += J[c] * J[r]; idx++; }
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

// Accumulator11 (MatrixAccumulators.h:89-175)
struct Acc11
{
	float A; size_t num;
	float d[4], d1k[4], d1m[4];
	float numIn1, numIn1k, numIn1m;
	void initialize() { A = 0; memset(d, 0, 16); memset(d1k, 0, 16); memset(d1m, 0, 16); num = 0; numIn1 = numIn1k = numIn1m = 0; }
	void shiftUp(bool force)
	{
		if (numIn1 > 1000 || force)
		{
			for (int l = 0; l < 4; l++) d1k[l] = d[l] + d1k[l];
			numIn1k += numIn1; numIn1 = 0; memset(d, 0, 16);
		}
		if (numIn1k > 1000 || force)
		{
			for (int l = 0; l < 4; l++) d1m[l] = d1k[l] + d1m[l];
			numIn1m += numIn1k; numIn1k = 0; memset(d1k, 0, 16);
		}
	}
	void updateSingle(float val) { d[0] += val; num++; numIn1++; shiftUp(false); }
	void finish() { shiftUp(true); A = d1m[0] + d1m[1] + d1m[2] + d1m[3]; }
};

// Eigen's compute_inverse<Matrix3d> (cofactors of column 0, determinant by expansion along it), as mat33f_inverse in float
inline void mat33d_inverse(const double* m, double* out)
{
	auto cof = [&](int i, int j) { int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3; return m[i1 * 3 + j1] * m[i2 * 3 + j2] - m[i1 * 3 + j2] * m[i2 * 3 + j1]; };
	const double c0 = cof(0, 0), c1 = cof(1, 0), c2 = cof(2, 0);
	const double det = c0 * m[0] + (c1 * m[3] + c2 * m[6]);
	const double invdet = 1.0 / det;
	out[0] = c0 * invdet; out[1] = c1 * invdet; out[2] = c2 * invdet;
	out[3] = cof(0, 1) * invdet; out[4] = cof(1, 1) * invdet; out[5] = cof(2, 1) * invdet;
	out[6] = cof(0, 2) * invdet; out[7] = cof(1, 2) * invdet; out[8] = cof(2, 2) * invdet;
}

}   // namespace

// CoarseInitializer::makeK, CoarseInitializer.cpp:1126-1155, for one level: fx, fy, cx, cy and K^-1 in double
extern "C" void dso_oracle_init_make_K(float fx0, float fy0, float cx0, float cy0, int lvl, double* fxfycxcy4, double* Ki9)
{
	double fx = fx0, fy = fy0, cx = cx0, cy = cy0;
	const double cx00 = cx, cy00 = cy;
	for (int level = 1; level <= lvl; level++)
	{
		fx = fx * 0.5; fy = fy * 0.5;
		cx = (cx00 + 0.5) / ((int)1 << level) - 0.5;
		cy = (cy00 + 0.5) / ((int)1 << level) - 0.5;
	}
	fxfycxcy4[0] = fx; fxfycxcy4[1] = fy; fxfycxcy4[2] = cx; fxfycxcy4[3] = cy;
	const double K[9] = { fx, 0.0, cx, 0.0, fy, cy, 0.0, 0.0, 1.0 };
	mat33d_inverse(K, Ki9);
}

// CoarseInitializer::calcResAndGS, CoarseInitializer.cpp:370-732.  colorRef / colorNew: dIp[lvl] of the first and the new
// frame (AoS, 3 floats per pixel); the Pnt fields as arrays.  Outputs: H_out, b_out, H_out_sc, b_out_sc (row-major 8x8 / 8),
// the returned Vec3f, and per point isGood_new, energy_new, maxstep, lastHessian_new, JbBuffer_new.
extern "C" void dso_oracle_init_calc_res_and_gs(int w0, int h0, int lvl, float fx0, float fy0, float cx0, float cy0,
                                                const float* colorRef, const float* colorNew,
                                                const double* refToNew12, double aff_a, double aff_b,
                                                int npts, const float* pu, const float* pv, const float* idepth_new, const unsigned char* isGood,
                                                const float* energy2, const float* outlierTH, const float* iR,
                                                float alphaW, float alphaK, float couplingWeight,
                                                float* H64, float* b8, float* Hsc64, float* bsc8, float* res3,
                                                unsigned char* isGood_new, float* energy_new2, float* maxstep_out, float* lastHessian_new,
                                                float* Jb10)
{
	const int wl = w0 >> lvl, hl = h0 >> lvl;
	double kk[4], Kid[9];
	dso_oracle_init_make_K(fx0, fy0, cx0, cy0, lvl, kk, Kid);
	SE3d refToNew = SE3d::from12(refToNew12);
	double RKd[9];
	mat33d_mul(refToNew.R, Kid, RKd);
	float RKi[9], t[3];
	for (int i = 0; i < 9; i++) RKi[i] = (float)RKd[i];                        // :385
	for (int i = 0; i < 3; i++) t[i] = (float)refToNew.t[i];                    // :387
	const float r2new_aff[2] = { (float)std::exp(aff_a), (float)aff_b };        // :390
	const float fxl = (float)kk[0], fyl = (float)kk[1], cxl = (float)kk[2], cyl = (float)kk[3];   // :394-397

	Acc11 E; Acc9 acc9;
	acc9.initialize(); E.initialize();
	std::vector<float> Jb((size_t)npts * 10, 0.f);
	for (int i = 0; i < npts; i++)
	{
		float mstep = 1e10f;                                                     // :414
		if (!isGood[i])                                                          // :416-425
		{
			E.updateSingle(energy2[2 * i]);
			energy_new2[2 * i] = energy2[2 * i]; energy_new2[2 * i + 1] = energy2[2 * i + 1];
			isGood_new[i] = 0;
			maxstep_out[i] = mstep;
			continue;
		}
		float dp[8][8], dd[8], r[8];         // dp[k][idx]
		float* JbN = &Jb[(size_t)i * 10];
		for (int k = 0; k < 10; k++) JbN[k] = 0;                                 // :445
		bool good = true;
		float energy = 0;
		for (int idx = 0; idx < 8; idx++)
		{
			int dx = patternP[idx][0], dy = patternP[idx][1];
			const float x = pu[i] + dx, y = pv[i] + dy;
			float pt[3];
			for (int q = 0; q < 3; q++)                                          // :463 (row dot in Eigen's order c0 + (c1 + c2))
				pt[q] = dot3_eigen(RKi[q * 3], x, RKi[q * 3 + 1], y, RKi[q * 3 + 2], 1.f) + t[q] * idepth_new[i];
			float u = pt[0] / pt[2];
			float v = pt[1] / pt[2];
			float Ku = fxl * u + cxl;
			float Kv = fyl * v + cyl;
			float new_idepth = idepth_new[i] / pt[2];
			if (!(Ku > 1 && Kv > 1 && Ku < wl - 2 && Kv < hl - 2 && new_idepth > 0)) { good = false; break; }   // :473-477
			float hit[3];
			interp33(colorNew, Ku, Kv, wl, hit);                                 // :483
			float rlR = interp31(colorRef, x, y, wl);                            // :488
			if (!std::isfinite(rlR) || !std::isfinite(hit[0])) { good = false; break; }   // :491-495
			float residual = hit[0] - r2new_aff[0] * rlR - r2new_aff[1];        // :499
			float hw = fabs(residual) < setting_huberTH ? 1 : setting_huberTH / fabs(residual);   // :502
			energy += hw * residual * residual * (2 - hw);                       // :504
			float dxdd = (t[0] - t[2] * u) / pt[2];                              // :511-512
			float dydd = (t[1] - t[2] * v) / pt[2];
			if (hw < 1) hw = sqrtf(hw);                                          // :516
			float dxInterp = hw * hit[1] * fxl;
			float dyInterp = hw * hit[2] * fyl;
			dp[0][idx] = new_idepth * dxInterp;                                  // :529-534
			dp[1][idx] = new_idepth * dyInterp;
			dp[2][idx] = -new_idepth * (u * dxInterp + v * dyInterp);
			dp[3][idx] = -u * v * dxInterp - (1 + v * v) * dyInterp;
			dp[4][idx] = (1 + u * u) * dxInterp + u * v * dyInterp;
			dp[5][idx] = -v * dxInterp + u * dyInterp;
			dp[6][idx] = -hw * r2new_aff[0] * rlR;                               // :542-543
			dp[7][idx] = -hw * 1;
			dd[idx] = dxInterp * dxdd + dyInterp * dydd;                         // :546
			r[idx] = hw * residual;
			const float a = dxdd * fxl, b = dydd * fyl;
			float ms = 1.0f / std::sqrt(a * a + b * b);                          // :551
			if (ms < mstep) mstep = ms;
			for (int k = 0; k < 8; k++) JbN[k] += dp[k][idx] * dd[idx];          // :562-571
			JbN[8] += r[idx] * dd[idx];
			JbN[9] += dd[idx] * dd[idx];
		}
		maxstep_out[i] = mstep;
		if (!good || energy > outlierTH[i] * 20)                                 // :574-580
		{
			E.updateSingle(energy2[2 * i]);
			isGood_new[i] = 0;
			energy_new2[2 * i] = energy2[2 * i]; energy_new2[2 * i + 1] = energy2[2 * i + 1];
			continue;
		}
		E.updateSingle(energy);                                                  // :582-584
		isGood_new[i] = 1;
		energy_new2[2 * i] = energy;
		energy_new2[2 * i + 1] = energy2[2 * i + 1];                             // overwritten below (:627)
		for (int q = 0; q + 3 < 8; q += 4)                                       // :587-597 (patternNum = 8: two SSE updates, no tail)
		{
			float J[9][4];
			for (int k = 0; k < 8; k++) for (int l = 0; l < 4; l++) J[k][l] = dp[k][q + l];
			for (int l = 0; l < 4; l++) J[8][l] = r[q + l];
			acc9.updateSSE(J);
		}
	}
	E.finish();
	acc9.finish();

	// :617-633: the regulariser energies go into E (after its finish), not into EAlpha -- EAlpha.A stays 0 and E.num doubles
	Acc11 EAlpha;
	EAlpha.initialize();
	for (int i = 0; i < npts; i++)
	{
		if (!isGood_new[i]) E.updateSingle(energy2[2 * i + 1]);
		else
		{
			energy_new2[2 * i + 1] = (idepth_new[i] - 1) * (idepth_new[i] - 1);
			E.updateSingle(energy_new2[2 * i + 1]);
		}
	}
	EAlpha.finish();
	const double tsq = refToNew.t[0] * refToNew.t[0] + (refToNew.t[1] * refToNew.t[1] + refToNew.t[2] * refToNew.t[2]);
	float alphaEnergy = alphaW * (EAlpha.A + tsq * npts);                        // :635
	float alphaOpt;
	if (alphaEnergy > alphaK * npts) { alphaOpt = 0; alphaEnergy = alphaK * npts; }   // :640-649
	else alphaOpt = alphaW;

	Acc9 acc9SC;
	acc9SC.initialize();
	for (int i = 0; i < npts; i++)                                               // :652-675
	{
		lastHessian_new[i] = 0;
		if (!isGood_new[i]) continue;
		float* JbN = &Jb[(size_t)i * 10];
		lastHessian_new[i] = JbN[9];
		JbN[8] += alphaOpt * (idepth_new[i] - 1);
		JbN[9] += alphaOpt;
		if (alphaOpt == 0)
		{
			JbN[8] += couplingWeight * (idepth_new[i] - iR[i]);
			JbN[9] += couplingWeight;
		}
		JbN[9] = 1 / (1 + JbN[9]);
		acc9SC.updateSingleWeighted(JbN, JbN[9]);
	}
	acc9SC.finish();

	for (int r = 0; r < 8; r++)                                                  // :680-683
	{
		for (int c = 0; c < 8; c++) { H64[r * 8 + c] = acc9.H[r][c]; Hsc64[r * 8 + c] = acc9SC.H[r][c]; }
		b8[r] = acc9.H[r][8]; bsc8[r] = acc9SC.H[r][8];
	}
	H64[0] += alphaOpt * npts; H64[9] += alphaOpt * npts; H64[18] += alphaOpt * npts;   // :685-687
	double lg[6];
	se3_log(refToNew, lg);
	for (int k = 0; k < 3; k++) b8[k] += (float)lg[k] * alphaOpt * npts;         // :689-692
	res3[0] = E.A; res3[1] = alphaEnergy; res3[2] = (float)E.num;               // :694
	if (Jb10) memcpy(Jb10, Jb.data(), Jb.size() * sizeof(float));
}
