// TEST INFRASTRUCTURE ONLY -- stand-in for the part of Sophus (SE3d / SO3d / Sim3d) that the reference's hot-path
// files use.  Published formulas: exp/log via the Rodrigues coefficients, Adj = [[R, hat(t) R], [0, R]];
// tangent order (translation, rotation).  Rotation is kept as a matrix (Sophus keeps a quaternion).
#ifndef DSO_COMPAT_SOPHUS_SE3
#define DSO_COMPAT_SOPHUS_SE3
#include <Eigen/Core>
namespace Sophus
{
typedef Eigen::Matrix<double, 3, 3> Mat3;
typedef Eigen::Matrix<double, 3, 1> V3;
typedef Eigen::Matrix<double, 6, 1> V6;
typedef Eigen::Matrix<double, 6, 6> Mat6;

inline Mat3 hat(const V3& w) { Mat3 W; W << 0, -w[2], w[1], w[2], 0, -w[0], -w[1], w[0], 0; return W; }

struct Quaterniond
{
	double qw, qx, qy, qz;
	Quaterniond(double w_, double x_, double y_, double z_) : qw(w_), qx(x_), qy(y_), qz(z_) {}
	Mat3 toRotationMatrix() const
	{
		double n = std::sqrt(qw * qw + qx * qx + qy * qy + qz * qz), w = qw / n, x = qx / n, y = qy / n, z = qz / n;
		Mat3 R;
		R << 1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w), 2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w),
			2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y);
		return R;
	}
};

class SO3d
{
public:
	Mat3 R;
	SO3d() { R.setIdentity(); }
	explicit SO3d(const Mat3& R_) : R(R_) {}
	const Mat3& matrix() const { return R; }
	static SO3d exp(const V3& w)
	{
		const double th2 = w.squaredNorm(), th = std::sqrt(th2);
		double A, B;
		if (th < 1e-5) { A = 1.0 - th2 / 6.0 + th2 * th2 / 120.0; B = 0.5 - th2 / 24.0 + th2 * th2 / 720.0; }
		else { A = std::sin(th) / th; B = (1.0 - std::cos(th)) / th2; }
		Mat3 W = hat(w);
		return SO3d(Mat3(Mat3::Identity() + W * A + (W * W) * B));
	}
	V3 log() const
	{
		const double c = std::min(1.0, std::max(-1.0, (R.trace() - 1.0) * 0.5));
		const double th = std::acos(c);
		const double f = th < 1e-7 ? 0.5 + th * th / 12.0 : th / (2.0 * std::sin(th));
		return V3(f * (R(2, 1) - R(1, 2)), f * (R(0, 2) - R(2, 0)), f * (R(1, 0) - R(0, 1)));
	}
	SO3d inverse() const { return SO3d(Mat3(R.transpose())); }
	SO3d operator*(const SO3d& o) const { return SO3d(Mat3(R * o.R)); }
	V3 operator*(const V3& p) const { return R * p; }
};

class SE3d
{
public:
	Mat3 R; V3 t;
	SE3d() { R.setIdentity(); t.setZero(); }
	SE3d(const Mat3& R_, const V3& t_) : R(R_), t(t_) {}
	SE3d(const SO3d& so3, const V3& t_) : R(so3.R), t(t_) {}
	SE3d(const Quaterniond& q, const V3& t_) : R(q.toRotationMatrix()), t(t_) {}
	const Mat3& rotationMatrix() const { return R; }
	V3& translation() { return t; }
	const V3& translation() const { return t; }
	SO3d so3() const { return SO3d(R); }
	void setRotationMatrix(const Mat3& R_) { R = R_; }
	Eigen::Matrix<double, 3, 4> matrix3x4() const { Eigen::Matrix<double, 3, 4> m; m.block<3, 3>(0, 0) = R; m.block<3, 1>(0, 3) = t; return m; }
	Eigen::Matrix<double, 4, 4> matrix() const { Eigen::Matrix<double, 4, 4> m; m.setIdentity(); m.block<3, 3>(0, 0) = R; m.block<3, 1>(0, 3) = t; return m; }
	SE3d inverse() const { Mat3 Rt = R.transpose(); return SE3d(Rt, V3(-(Rt * t))); }
	SE3d operator*(const SE3d& o) const { return SE3d(Mat3(R * o.R), V3(R * o.t + t)); }
	SE3d& operator*=(const SE3d& o) { *this = *this * o; return *this; }
	V3 operator*(const V3& p) const { return R * p + t; }
	static SE3d exp(const V6& xi)
	{
		V3 u = xi.head<3>(), w = xi.tail<3>();
		const double th2 = w.squaredNorm(), th = std::sqrt(th2);
		double A, B, C;
		if (th < 1e-5) { A = 1.0 - th2 / 6.0 + th2 * th2 / 120.0; B = 0.5 - th2 / 24.0 + th2 * th2 / 720.0; C = 1.0 / 6.0 - th2 / 120.0 + th2 * th2 / 5040.0; }
		else { A = std::sin(th) / th; B = (1.0 - std::cos(th)) / th2; C = (th - std::sin(th)) / (th2 * th); }
		Mat3 W = hat(w), W2 = W * W;
		Mat3 Rr = Mat3::Identity() + W * A + W2 * B;
		Mat3 V = Mat3::Identity() + W * B + W2 * C;
		return SE3d(Rr, V3(V * u));
	}
	V6 log() const
	{
		V3 w = SO3d(R).log();
		const double th2 = w.squaredNorm(), th = std::sqrt(th2);
		double k;
		if (th < 1e-5) k = 1.0 / 12.0 + th2 / 720.0;
		else { const double A = std::sin(th) / th, B = (1.0 - std::cos(th)) / th2; k = (1.0 - A / (2.0 * B)) / th2; }
		Mat3 W = hat(w);
		Mat3 Vinv = Mat3::Identity() - W * 0.5 + (W * W) * k;
		V6 xi; xi.head<3>() = Vinv * t; xi.tail<3>() = w;
		return xi;
	}
	Mat6 Adj() const
	{
		Mat6 A; A.setZero();
		A.block<3, 3>(0, 0) = R; A.block<3, 3>(3, 3) = R; A.block<3, 3>(0, 3) = hat(t) * R;
		return A;
	}
};
typedef SE3d SE3;
typedef SO3d SO3;
}   // namespace Sophus
#endif
