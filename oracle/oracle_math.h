// TEST INFRASTRUCTURE ONLY -- CPU oracle for the DSO photometric Gauss-Newton hot path.
// Nothing in the product (dso-noted-in-chinese_b200/, include/) may include or link this file.
//
// Small linear-algebra helpers that stand in for the third-party arithmetic the
// reference reaches through Eigen and Sophus (neither is shipped under /root/reference,
// versions unpinned -> "parity unpinned" for these pieces):
//   * float 3x3 products / inverse in the evaluation order Eigen 3.3 uses for small
//     fixed-size types (coefficient-based lazy product, unrolled redux  c0 + (c1 + c2)),
//   * SE3 exp / log / inverse / compose / adjoint (Sophus' published formulas),
//   * LDL^T with diagonal pivoting for the 8x8 tracker system (Eigen::LDLT algorithm).
#pragma once
#include <cmath>
#include <cstring>
#include <algorithm>
#include <vector>

namespace orc {

// ---- float 3x3, row-major m[r*3+c] ------------------------------------------------
// Eigen fixed-size redux over 3 terms: func(c0, func(c1, c2)).
static inline float dot3_eigen(float a0, float b0, float a1, float b1, float a2, float b2)
{
	float p0 = a0 * b0, p1 = a1 * b1, p2 = a2 * b2;
	return p0 + (p1 + p2);
}
static inline void mat33f_mul(const float* A, const float* B, float* C)
{
	float T[9];
	for (int r = 0; r < 3; r++)
		for (int c = 0; c < 3; c++)
			T[r * 3 + c] = dot3_eigen(A[r * 3 + 0], B[0 * 3 + c], A[r * 3 + 1], B[1 * 3 + c], A[r * 3 + 2], B[2 * 3 + c]);
	memcpy(C, T, sizeof(T));
}
static inline void mat33f_vec(const float* A, const float* v, float* out)
{
	float T[3];
	for (int r = 0; r < 3; r++)
		T[r] = dot3_eigen(A[r * 3 + 0], v[0], A[r * 3 + 1], v[1], A[r * 3 + 2], v[2]);
	out[0] = T[0]; out[1] = T[1]; out[2] = T[2];
}
// Eigen's compute_inverse<Matrix3f>: cofactors of column 0, det by expansion along it.
static inline float cof33(const float* m, int i, int j)
{
	int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
	return m[i1 * 3 + j1] * m[i2 * 3 + j2] - m[i1 * 3 + j2] * m[i2 * 3 + j1];
}
static inline void mat33f_inverse(const float* m, float* out)
{
	float c0 = cof33(m, 0, 0), c1 = cof33(m, 1, 0), c2 = cof33(m, 2, 0);
	float det = dot3_eigen(c0, m[0], c1, m[3], c2, m[6]);
	float invdet = 1.0f / det;
	float R[9];
	R[0] = c0 * invdet; R[1] = c1 * invdet; R[2] = c2 * invdet;
	R[3] = cof33(m, 0, 1) * invdet; R[4] = cof33(m, 1, 1) * invdet; R[5] = cof33(m, 2, 1) * invdet;
	R[6] = cof33(m, 0, 2) * invdet; R[7] = cof33(m, 1, 2) * invdet; R[8] = cof33(m, 2, 2) * invdet;
	memcpy(out, R, sizeof(R));
}

// ---- SE3 in double: R row-major 3x3, t ---------------------------------------------
struct SE3d
{
	double R[9];
	double t[3];
	SE3d() { setIdentity(); }
	void setIdentity() { for (int i = 0; i < 9; i++) R[i] = (i % 4 == 0); t[0] = t[1] = t[2] = 0; }
	static SE3d from12(const double* p) { SE3d T; memcpy(T.R, p, 72); memcpy(T.t, p + 9, 24); return T; }
	void to12(double* p) const { memcpy(p, R, 72); memcpy(p + 9, t, 24); }
};
// 3-term products in the order Eigen's fixed-size redux uses (also in double): c0 + (c1 + c2)
static inline double dot3d(double a0, double b0, double a1, double b1, double a2, double b2)
{
	double p0 = a0 * b0, p1 = a1 * b1, p2 = a2 * b2;
	return p0 + (p1 + p2);
}
static inline void mat33d_mul(const double* A, const double* B, double* C)
{
	double T[9];
	for (int r = 0; r < 3; r++)
		for (int c = 0; c < 3; c++)
			T[r * 3 + c] = dot3d(A[r * 3 + 0], B[c], A[r * 3 + 1], B[3 + c], A[r * 3 + 2], B[6 + c]);
	memcpy(C, T, sizeof(T));
}
static inline SE3d operator*(const SE3d& a, const SE3d& b)
{
	SE3d o;
	mat33d_mul(a.R, b.R, o.R);
	for (int r = 0; r < 3; r++)
		o.t[r] = dot3d(a.R[r * 3 + 0], b.t[0], a.R[r * 3 + 1], b.t[1], a.R[r * 3 + 2], b.t[2]) + a.t[r];
	return o;
}
static inline SE3d se3_inverse(const SE3d& a)
{
	SE3d o;
	for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) o.R[r * 3 + c] = a.R[c * 3 + r];
	for (int r = 0; r < 3; r++)
		o.t[r] = -dot3d(o.R[r * 3 + 0], a.t[0], o.R[r * 3 + 1], a.t[1], o.R[r * 3 + 2], a.t[2]);
	return o;
}
static inline void hat3(const double* w, double* W)
{
	W[0] = 0; W[1] = -w[2]; W[2] = w[1];
	W[3] = w[2]; W[4] = 0; W[5] = -w[0];
	W[6] = -w[1]; W[7] = w[0]; W[8] = 0;
}
// Sophus convention: tangent = (upsilon[3] translation part, omega[3] rotation part).
// exp: R = I + A W + B W^2, t = V upsilon, V = I + B W + C W^2,
// A = sin(th)/th, B = (1-cos th)/th^2, C = (th - sin th)/th^3.
static inline SE3d se3_exp(const double* xi)
{
	const double* u = xi; const double* w = xi + 3;
	double th2 = w[0] * w[0] + w[1] * w[1] + w[2] * w[2];
	double th = std::sqrt(th2);
	double A, B, C;
	if (th < 1e-5)
	{
		A = 1.0 - th2 / 6.0 + th2 * th2 / 120.0;
		B = 0.5 - th2 / 24.0 + th2 * th2 / 720.0;
		C = 1.0 / 6.0 - th2 / 120.0 + th2 * th2 / 5040.0;
	}
	else
	{
		A = std::sin(th) / th;
		B = (1.0 - std::cos(th)) / th2;
		C = (th - std::sin(th)) / (th2 * th);
	}
	double W[9], W2[9];
	hat3(w, W);
	mat33d_mul(W, W, W2);
	SE3d o;
	for (int i = 0; i < 9; i++)
	{
		double I = (i % 4 == 0);
		o.R[i] = I + A * W[i] + B * W2[i];
	}
	for (int r = 0; r < 3; r++)
	{
		double V[3];
		for (int c = 0; c < 3; c++) V[c] = (r == c ? 1.0 : 0.0) + B * W[r * 3 + c] + C * W2[r * 3 + c];
		o.t[r] = dot3d(V[0], u[0], V[1], u[1], V[2], u[2]);
	}
	return o;
}
// log: inverse of the above.
static inline void se3_log(const SE3d& T, double* xi)
{
	double tr = T.R[0] + T.R[4] + T.R[8];
	double c = std::min(1.0, std::max(-1.0, (tr - 1.0) * 0.5));
	double th = std::acos(c);
	double w[3];
	double f;
	if (th < 1e-7) f = 0.5 + th * th / 12.0;
	else f = th / (2.0 * std::sin(th));
	w[0] = f * (T.R[7] - T.R[5]);
	w[1] = f * (T.R[2] - T.R[6]);
	w[2] = f * (T.R[3] - T.R[1]);
	double th2 = th * th;
	double W[9], W2[9];
	hat3(w, W);
	mat33d_mul(W, W, W2);
	// V^-1 = I - 0.5 W + k W^2, k = (1 - A/(2B))/th^2
	double k;
	if (th < 1e-5) k = 1.0 / 12.0 + th2 / 720.0;
	else
	{
		double A = std::sin(th) / th, B = (1.0 - std::cos(th)) / th2;
		k = (1.0 - A / (2.0 * B)) / th2;
	}
	for (int r = 0; r < 3; r++)
	{
		double V[3];
		for (int cc = 0; cc < 3; cc++) V[cc] = (r == cc ? 1.0 : 0.0) - 0.5 * W[r * 3 + cc] + k * W2[r * 3 + cc];
		xi[r] = dot3d(V[0], T.t[0], V[1], T.t[1], V[2], T.t[2]);
	}
	xi[3] = w[0]; xi[4] = w[1]; xi[5] = w[2];
}
// Adjoint, Sophus layout: [[R, hat(t) R], [0, R]] (6x6 row-major).
static inline void se3_adj(const SE3d& T, double* Ad)
{
	double Th[9], ThR[9];
	hat3(T.t, Th);
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

// ---- LDL^T with diagonal pivoting (Eigen::LDLT<MatrixXd>::compute + solve), n <= 8 ----
// A row-major symmetric n x n (lower triangle is read), solves A x = rhs.
static inline void ldlt_solve(const double* Ain, const double* rhs, double* x, int n)
{
	double A[64];
	int perm[8];
	for (int i = 0; i < n; i++) for (int j = 0; j < n; j++) A[i * n + j] = Ain[i * 8 + j];
	for (int i = 0; i < n; i++) perm[i] = i;
	int transp[8];
	for (int k = 0; k < n; k++)
	{
		// largest |diagonal| in the trailing part
		int p = k; double best = std::fabs(A[k * n + k]);
		for (int i = k + 1; i < n; i++) { double v = std::fabs(A[i * n + i]); if (v > best) { best = v; p = i; } }
		transp[k] = p;
		if (p != k)
		{
			// symmetric swap of rows/cols k and p on the lower triangle
			for (int j = 0; j < n; j++) std::swap(A[k * n + j], A[p * n + j]);
			for (int i = 0; i < n; i++) std::swap(A[i * n + k], A[i * n + p]);
		}
		// A(k,k) -= sum_j L(k,j)^2 D(j);  column update
		for (int j = 0; j < k; j++) A[k * n + k] -= A[k * n + j] * A[k * n + j] * A[j * n + j];
		double d = A[k * n + k];
		for (int i = k + 1; i < n; i++)
		{
			double s = A[i * n + k];
			for (int j = 0; j < k; j++) s -= A[i * n + j] * A[k * n + j] * A[j * n + j];
			A[i * n + k] = (std::fabs(d) > 0) ? s / d : 0.0;
			A[k * n + i] = A[i * n + k];
		}
	}
	double y[8];
	for (int i = 0; i < n; i++) y[i] = rhs[i];
	for (int k = 0; k < n; k++) std::swap(y[k], y[transp[k]]);
	// L y' = y
	for (int i = 0; i < n; i++) for (int j = 0; j < i; j++) y[i] -= A[i * n + j] * y[j];
	// D
	for (int i = 0; i < n; i++)
	{
		double d = A[i * n + i];
		y[i] = (std::fabs(d) > 2.2250738585072014e-308) ? y[i] / d : 0.0;
	}
	// L^T
	for (int i = n - 1; i >= 0; i--) for (int j = i + 1; j < n; j++) y[i] -= A[j * n + i] * y[j];
	for (int k = n - 1; k >= 0; k--) std::swap(y[k], y[transp[k]]);
	for (int i = 0; i < n; i++) x[i] = y[i];
	(void)perm;
}

// ---- dynamic-size pieces for the solveSystemF tail (EnergyFunctional.cpp:802-924) ----
// Eigen::LDLT<MatrixXd> (diagonal pivoting), same algorithm as ldlt_solve above for any n; A row-major n x n.
static inline void ldlt_solve_n(const double* Ain, const double* rhs, double* x, int n)
{
	std::vector<double> A(Ain, Ain + (size_t)n * n), y(rhs, rhs + n);
	std::vector<int> transp(n);
	for (int k = 0; k < n; k++)
	{
		int p = k; double best = std::fabs(A[k * n + k]);
		for (int i = k + 1; i < n; i++) { double v = std::fabs(A[i * n + i]); if (v > best) { best = v; p = i; } }
		transp[k] = p;
		if (p != k)
		{
			for (int j = 0; j < n; j++) std::swap(A[k * n + j], A[p * n + j]);
			for (int i = 0; i < n; i++) std::swap(A[i * n + k], A[i * n + p]);
		}
		for (int j = 0; j < k; j++) A[k * n + k] -= A[k * n + j] * A[k * n + j] * A[j * n + j];
		const double d = A[k * n + k];
		for (int i = k + 1; i < n; i++)
		{
			double s = A[i * n + k];
			for (int j = 0; j < k; j++) s -= A[i * n + j] * A[k * n + j] * A[j * n + j];
			s = (std::fabs(d) > 0) ? s / d : 0.0;
			A[i * n + k] = s; A[k * n + i] = s;
		}
	}
	for (int k = 0; k < n; k++) std::swap(y[k], y[transp[k]]);
	for (int i = 0; i < n; i++) for (int j = 0; j < i; j++) y[i] -= A[i * n + j] * y[j];
	for (int i = 0; i < n; i++)
	{
		const double d = A[i * n + i];
		y[i] = (std::fabs(d) > 2.2250738585072014e-308) ? y[i] / d : 0.0;
	}
	for (int i = n - 1; i >= 0; i--) for (int j = i + 1; j < n; j++) y[i] -= A[j * n + i] * y[j];
	for (int k = n - 1; k >= 0; k--) std::swap(y[k], y[transp[k]]);
	for (int i = 0; i < n; i++) x[i] = y[i];
}

// Thin SVD of an m x k matrix (k <= m) by one-sided Jacobi rotations (Hestenes): N = U diag(S) V^T.
// Stands in for Eigen::JacobiSVD in EnergyFunctional::orthogonalize (:767); the projector built from it does not
// depend on the SVD algorithm beyond rounding.  N, U: m x k row-major; V: k x k row-major.
static inline void jacobi_svd_thin(const double* N, int m, int k, double* U, double* S, double* V)
{
	std::vector<double> A(N, N + (size_t)m * k);
	for (int i = 0; i < k; i++) for (int j = 0; j < k; j++) V[i * k + j] = (i == j) ? 1.0 : 0.0;
	for (int sweep = 0; sweep < 60; sweep++)
	{
		double off = 0;
		for (int p = 0; p < k - 1; p++)
			for (int q = p + 1; q < k; q++)
			{
				double alpha = 0, beta = 0, gamma = 0;
				for (int i = 0; i < m; i++) { alpha += A[i * k + p] * A[i * k + p]; beta += A[i * k + q] * A[i * k + q]; gamma += A[i * k + p] * A[i * k + q]; }
				if (gamma == 0 || std::fabs(gamma) <= 1e-300) continue;
				off = std::max(off, std::fabs(gamma) / std::sqrt(alpha * beta));
				if (std::fabs(gamma) <= 1e-17 * std::sqrt(alpha * beta)) continue;
				const double zeta = (beta - alpha) / (2.0 * gamma);
				const double t = (zeta >= 0 ? 1.0 : -1.0) / (std::fabs(zeta) + std::sqrt(1.0 + zeta * zeta));
				const double c = 1.0 / std::sqrt(1.0 + t * t), s = c * t;
				for (int i = 0; i < m; i++)
				{
					const double ap = A[i * k + p], aq = A[i * k + q];
					A[i * k + p] = c * ap - s * aq; A[i * k + q] = s * ap + c * aq;
				}
				for (int i = 0; i < k; i++)
				{
					const double vp = V[i * k + p], vq = V[i * k + q];
					V[i * k + p] = c * vp - s * vq; V[i * k + q] = s * vp + c * vq;
				}
			}
		if (off < 1e-15) break;
	}
	for (int j = 0; j < k; j++)
	{
		double n2 = 0;
		for (int i = 0; i < m; i++) n2 += A[i * k + j] * A[i * k + j];
		S[j] = std::sqrt(n2);
		for (int i = 0; i < m; i++) U[i * k + j] = S[j] > 0 ? A[i * k + j] / S[j] : 0.0;
	}
}

} // namespace orc
