// Double-precision pieces of the coarse tracker's LM step, usable from host and device:
// 8x8 LDL^T solve (Eigen::LDLT semantics: diagonal pivoting, CoarseTracker.cpp:618), SE3 exponential
// and composition (Sophus SE3::exp / operator*, CoarseTracker.cpp:661), AffLight::fromToVecExposure
// (util/NumType.h:176-189) and the float conversions of CoarseTracker.cpp:375-378.
#pragma once
#include <cuda_runtime.h>
#include <math.h>

#define HD __host__ __device__ __forceinline__

struct Pose
{
	double R[9];
	double t[3];
};

HD void pose_identity(Pose& p)
{
	for (int i = 0; i < 9; i++) p.R[i] = (i % 4 == 0) ? 1.0 : 0.0;
	p.t[0] = p.t[1] = p.t[2] = 0.0;
}
HD void pose_from12(Pose& p, const double* v)
{
	for (int i = 0; i < 9; i++) p.R[i] = v[i];
	for (int i = 0; i < 3; i++) p.t[i] = v[9 + i];
}
HD void pose_to12(const Pose& p, double* v)
{
	for (int i = 0; i < 9; i++) v[i] = p.R[i];
	for (int i = 0; i < 3; i++) v[9 + i] = p.t[i];
}
// Three-term double products in the order Eigen's fixed-size redux takes them, c0 + (c1 + c2), every operation
// rounded separately
HD double dot3dd(double a0, double b0, double a1, double b1, double a2, double b2)
{
#ifdef __CUDA_ARCH__
	return __dadd_rn(__dmul_rn(a0, b0), __dadd_rn(__dmul_rn(a1, b1), __dmul_rn(a2, b2)));
#else
	volatile double p0 = a0 * b0, p1 = a1 * b1, p2 = a2 * b2;
	volatile double s12 = p1 + p2;
	return p0 + s12;
#endif
}
HD double dadd_r(double a, double b)
{
#ifdef __CUDA_ARCH__
	return __dadd_rn(a, b);
#else
	volatile double s = a + b;
	return s;
#endif
}
HD double dmul_r(double a, double b)
{
#ifdef __CUDA_ARCH__
	return __dmul_rn(a, b);
#else
	volatile double s = a * b;
	return s;
#endif
}
HD void mat33d_mul(const double* A, const double* B, double* C)
{
	double T[9];
	for (int r = 0; r < 3; r++)
		for (int c = 0; c < 3; c++)
			T[r * 3 + c] = dot3dd(A[r * 3 + 0], B[c], A[r * 3 + 1], B[3 + c], A[r * 3 + 2], B[6 + c]);
	for (int i = 0; i < 9; i++) C[i] = T[i];
}
// c = a * b  (Sophus SE3 operator*: R = Ra Rb, t = Ra tb + ta)
HD void pose_mul(const Pose& a, const Pose& b, Pose& c)
{
	Pose o;
	mat33d_mul(a.R, b.R, o.R);
	for (int r = 0; r < 3; r++)
		o.t[r] = dadd_r(dot3dd(a.R[r * 3 + 0], b.t[0], a.R[r * 3 + 1], b.t[1], a.R[r * 3 + 2], b.t[2]), a.t[r]);
	c = o;
}
// Sophus SE3::inverse: R^T, -R^T t
HD void pose_inverse(const Pose& a, Pose& o)
{
	Pose q;
	for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) q.R[r * 3 + c] = a.R[c * 3 + r];
	for (int r = 0; r < 3; r++)
		q.t[r] = -dot3dd(q.R[r * 3 + 0], a.t[0], q.R[r * 3 + 1], a.t[1], q.R[r * 3 + 2], a.t[2]);
	o = q;
}
// xi = (translation part u[3], rotation part w[3]);  R = I + A W + B W^2,  t = (I + B W + C W^2) u
HD void se3_exp(const double* xi, Pose& o)
{
	const double* u = xi; const double* w = xi + 3;
	const double th2 = dadd_r(dadd_r(dmul_r(w[0], w[0]), dmul_r(w[1], w[1])), dmul_r(w[2], w[2]));
	const double th = sqrt(th2);
	double A, B, C;
	if (th < 1e-5)
	{
		A = dadd_r(1.0 - th2 / 6.0, dmul_r(th2, th2) / 120.0);
		B = dadd_r(0.5 - th2 / 24.0, dmul_r(th2, th2) / 720.0);
		C = dadd_r(1.0 / 6.0 - th2 / 120.0, dmul_r(th2, th2) / 5040.0);
	}
	else
	{
		double s, c;
#ifdef __CUDA_ARCH__
		sincos(th, &s, &c);
#else
		s = sin(th); c = cos(th);
#endif
		A = s / th;
		B = (1.0 - c) / th2;
		C = (th - s) / dmul_r(th2, th);
	}
	const double W[9] = { 0, -w[2], w[1], w[2], 0, -w[0], -w[1], w[0], 0 };
	double W2[9];
	mat33d_mul(W, W, W2);
	for (int i = 0; i < 9; i++) o.R[i] = dadd_r(dadd_r((i % 4 == 0) ? 1.0 : 0.0, dmul_r(A, W[i])), dmul_r(B, W2[i]));
	for (int r = 0; r < 3; r++)
	{
		double V[3];
		for (int c = 0; c < 3; c++) V[c] = dadd_r(dadd_r((r == c) ? 1.0 : 0.0, dmul_r(B, W[r * 3 + c])), dmul_r(C, W2[r * 3 + c]));
		o.t[r] = dot3dd(V[0], u[0], V[1], u[1], V[2], u[2]);
	}
}
// Sophus SE3::Adj, 6x6 row-major: [[R, hat(t) R], [0, R]]
HD void pose_adjoint(const Pose& T, double* Ad)
{
	const double Th[9] = { 0, -T.t[2], T.t[1], T.t[2], 0, -T.t[0], -T.t[1], T.t[0], 0 };
	double ThR[9];
	mat33d_mul(Th, T.R, ThR);
	for (int i = 0; i < 36; i++) Ad[i] = 0;
	for (int r = 0; r < 3; r++)
		for (int c = 0; c < 3; c++)
		{
			Ad[r * 6 + c] = T.R[r * 3 + c];
			Ad[r * 6 + 3 + c] = ThR[r * 3 + c];
			Ad[(r + 3) * 6 + 3 + c] = T.R[r * 3 + c];
		}
}

// Sophus SE3::log (host only): xi = (V^-1 t, w) with the closed forms of SO3::log and V^-1 = I - W/2 + k W^2
static inline void pose_log(const Pose& T, double* xi)
{
	const double tr = T.R[0] + T.R[4] + T.R[8];
	double c = (tr - 1.0) * 0.5;
	c = c < -1.0 ? -1.0 : (c > 1.0 ? 1.0 : c);
	const double th = acos(c);
	double f;
	if (th < 1e-7) f = 0.5 + th * th / 12.0;
	else f = th / (2.0 * sin(th));
	const double w[3] = { f * (T.R[7] - T.R[5]), f * (T.R[2] - T.R[6]), f * (T.R[3] - T.R[1]) };
	const double th2 = th * th;
	const double W[9] = { 0, -w[2], w[1], w[2], 0, -w[0], -w[1], w[0], 0 };
	double W2[9];
	mat33d_mul(W, W, W2);
	double k;
	if (th < 1e-5) k = 1.0 / 12.0 + th2 / 720.0;
	else
	{
		const double A = sin(th) / th, B = (1.0 - cos(th)) / th2;
		k = (1.0 - A / (2.0 * B)) / th2;
	}
	for (int r = 0; r < 3; r++)
	{
		double V[3];
		for (int cc = 0; cc < 3; cc++) V[cc] = (r == cc ? 1.0 : 0.0) - 0.5 * W[r * 3 + cc] + k * W2[r * 3 + cc];
		xi[r] = dot3dd(V[0], T.t[0], V[1], T.t[1], V[2], T.t[2]);
	}
	xi[3] = w[0]; xi[4] = w[1]; xi[5] = w[2];
}

// AffLight::fromToVecExposure (NumType.h:176-189)
HD void aff_from_to(float exposureF, float exposureT, double Fa, double Fb, double Ta, double Tb, double& a, double& b)
{
	if (exposureF == 0 || exposureT == 0) exposureT = exposureF = 1;
	a = exp(Ta - Fa) * exposureT / exposureF;
	b = Tb - a * Fb;
}

// LDL^T with diagonal pivoting on the (not yet updated) diagonal, as Eigen's unblocked LDLT does;
// A row-major with row stride 8, n <= 8.  Solves A x = rhs.
HD void ldlt_solve(const double* Ain, const double* rhs, double* x, int n)
{
	double A[64];
	int tr[8];
	for (int i = 0; i < n; i++) for (int j = 0; j < n; j++) A[i * 8 + j] = Ain[i * 8 + j];
	for (int k = 0; k < n; k++)
	{
		int p = k;
		double best = fabs(A[k * 8 + k]);
		for (int i = k + 1; i < n; i++) { double v = fabs(A[i * 8 + i]); if (v > best) { best = v; p = i; } }
		tr[k] = p;
		if (p != k)
		{
			for (int j = 0; j < n; j++) { double t = A[k * 8 + j]; A[k * 8 + j] = A[p * 8 + j]; A[p * 8 + j] = t; }
			for (int i = 0; i < n; i++) { double t = A[i * 8 + k]; A[i * 8 + k] = A[i * 8 + p]; A[i * 8 + p] = t; }
		}
		for (int j = 0; j < k; j++) A[k * 8 + k] -= A[k * 8 + j] * A[k * 8 + j] * A[j * 8 + j];
		const double d = A[k * 8 + k];
		for (int i = k + 1; i < n; i++)
		{
			double s = A[i * 8 + k];
			for (int j = 0; j < k; j++) s -= A[i * 8 + j] * A[k * 8 + j] * A[j * 8 + j];
			s = (fabs(d) > 0.0) ? s / d : 0.0;
			A[i * 8 + k] = s;
			A[k * 8 + i] = s;
		}
	}
	double y[8];
	for (int i = 0; i < n; i++) y[i] = rhs[i];
	for (int k = 0; k < n; k++) { double t = y[k]; y[k] = y[tr[k]]; y[tr[k]] = t; }
	for (int i = 0; i < n; i++) for (int j = 0; j < i; j++) y[i] -= A[i * 8 + j] * y[j];
	for (int i = 0; i < n; i++)
	{
		const double d = A[i * 8 + i];
		y[i] = (fabs(d) > 2.2250738585072014e-308) ? y[i] / d : 0.0;
	}
	for (int i = n - 1; i >= 0; i--) for (int j = i + 1; j < n; j++) y[i] -= A[j * 8 + i] * y[j];
	for (int k = n - 1; k >= 0; k--) { double t = y[k]; y[k] = y[tr[k]]; y[tr[k]] = t; }
	for (int i = 0; i < n; i++) x[i] = y[i];
}

// Eigen's fixed-size float 3x3 product: each coefficient is c0 + (c1 + c2)
HD float dot3h(float a0, float b0, float a1, float b1, float a2, float b2)
{
#ifdef __CUDA_ARCH__
	return __fadd_rn(__fmul_rn(a0, b0), __fadd_rn(__fmul_rn(a1, b1), __fmul_rn(a2, b2)));
#else
	volatile float p0 = a0 * b0, p1 = a1 * b1, p2 = a2 * b2;   // volatile: keep the host compiler from contracting
	volatile float s12 = p1 + p2;
	return p0 + s12;
#endif
}
HD void mat33f_mul(const float* A, const float* B, float* C)
{
	float T[9];
	for (int r = 0; r < 3; r++)
		for (int c = 0; c < 3; c++)
			T[r * 3 + c] = dot3h(A[r * 3 + 0], B[c], A[r * 3 + 1], B[3 + c], A[r * 3 + 2], B[6 + c]);
	for (int i = 0; i < 9; i++) C[i] = T[i];
}
