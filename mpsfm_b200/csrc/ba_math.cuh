// ba_math.cuh -- per-observation arithmetic of the bundle-adjustment hot path.
//
// Everything here is __host__ __device__ so that tests can compile the very same arithmetic with
// g++ and compare it with the numpy oracle on the CPU box (tests/hostcheck); the product only ever
// runs it inside the sm_100a kernels of ba_kernels.cu.
//
// What is computed (reference call sites under /root/reference/mpsfm/sfm/mapper/bundle_adjustment.py):
//   :104      reprojection residual  r = ImgFromCam(params, q*X + t) - x_obs   (COLMAP cost functor)
//   :163-176  log-depth prior        r = log(exp(s) d + b) - log((q*X + t)_z)  (fork-only functor)
//   :98-102, :160-161  ScaledLoss(loss(a), w) on each residual block, Ceres corrector
//   :114-122  tangent of an image = (half-angle rotation vector, translation)
//
// With p_rot = R(q) X, p = p_rot + t and a_k = d r_k / d p (rows of the 3x3 matrix A), every
// Jacobian block of an observation factors through three small objects:
//     Jc = A' G,   G = [ -2 [p_rot]x | I3 ]     (image tangent, 3x6)
//     Jp = A' R                                  (point, 3x3)
// where A' = diag(sqrt(rho'_k)) A is the robustified A.  The kernels therefore only need
//     N = A'^T A'  (3x3 symmetric),   h = A'^T r'  (3),   p_rot,   R
// per observation: Hcc += G^T N G, gc += G^T h, Hpp += R^T N R, gp += R^T h, W = G^T N R.
#pragma once

#include <float.h>
#include <math.h>
#include <stddef.h>
#include <stdint.h>

#ifdef __CUDACC__
#define MPBA_HD __host__ __device__ __forceinline__
#else
#define MPBA_HD inline
#endif

namespace mpba {

// ------------------------------------------------------------------------------------------------
// forward-mode dual number over (u, v): the analytic derivative carrier for the camera models that
// have no short closed form (fisheye family, FOV, FULL_OPENCV, THIN_PRISM).
struct Dual2 {
  double a, du, dv;
};
MPBA_HD Dual2 mk(double a) { return Dual2{a, 0.0, 0.0}; }
MPBA_HD Dual2 operator+(Dual2 x, Dual2 y) { return Dual2{x.a + y.a, x.du + y.du, x.dv + y.dv}; }
MPBA_HD Dual2 operator-(Dual2 x, Dual2 y) { return Dual2{x.a - y.a, x.du - y.du, x.dv - y.dv}; }
MPBA_HD Dual2 operator*(Dual2 x, Dual2 y) {
  return Dual2{x.a * y.a, x.a * y.du + y.a * x.du, x.a * y.dv + y.a * x.dv};
}
MPBA_HD Dual2 operator/(Dual2 x, Dual2 y) {
  const double inv = 1.0 / y.a, q = x.a * inv;
  return Dual2{q, (x.du - q * y.du) * inv, (x.dv - q * y.dv) * inv};
}
MPBA_HD Dual2 operator+(Dual2 x, double y) { return Dual2{x.a + y, x.du, x.dv}; }
MPBA_HD Dual2 operator+(double y, Dual2 x) { return Dual2{x.a + y, x.du, x.dv}; }
MPBA_HD Dual2 operator-(Dual2 x, double y) { return Dual2{x.a - y, x.du, x.dv}; }
MPBA_HD Dual2 operator-(double y, Dual2 x) { return Dual2{y - x.a, -x.du, -x.dv}; }
MPBA_HD Dual2 operator*(Dual2 x, double y) { return Dual2{x.a * y, x.du * y, x.dv * y}; }
MPBA_HD Dual2 operator*(double y, Dual2 x) { return Dual2{x.a * y, x.du * y, x.dv * y}; }
MPBA_HD Dual2 operator/(Dual2 x, double y) {
  const double inv = 1.0 / y;
  return Dual2{x.a * inv, x.du * inv, x.dv * inv};
}
MPBA_HD Dual2 dsqrt(Dual2 x) {
  const double s = sqrt(x.a), k = 0.5 / s;
  return Dual2{s, k * x.du, k * x.dv};
}
MPBA_HD Dual2 datan(Dual2 x) {
  const double k = 1.0 / (1.0 + x.a * x.a);
  return Dual2{atan(x.a), k * x.du, k * x.dv};
}
MPBA_HD double dsqrt(double x) { return sqrt(x); }
MPBA_HD double datan(double x) { return atan(x); }
MPBA_HD double val(double x) { return x; }
MPBA_HD double val(Dual2 x) { return x.a; }
MPBA_HD double one_like(double) { return 1.0; }
MPBA_HD Dual2 one_like(Dual2) { return Dual2{1.0, 0.0, 0.0}; }

// ------------------------------------------------------------------------------------------------
// Ceres loss functions (upstream loss_function.cc): rho and rho' of `type` with scale a at squared
// norm s, multiplied by `weight` as ceres::ScaledLoss does.  All four have rho'' <= 0, so the
// corrector (upstream corrector.cc) always takes its "alpha = 0" branch: rows scale by sqrt(rho').
template <bool WANT_RHO>
MPBA_HD void loss_eval(int type, double a, double weight, double s, double& rho, double& rho1) {
  if (type == 0) {  // TRIVIAL
    rho = s;
    rho1 = 1.0;
  } else {
    const double b = a * a, c = 1.0 / b;
    if (type == 1) {  // SOFT_L1
      const double sum = 1.0 + s * c, tmp = sqrt(sum);
      rho = 2.0 * b * (tmp - 1.0);
      rho1 = fmax(DBL_MIN, 1.0 / tmp);
    } else if (type == 2) {  // CAUCHY
      const double sum = 1.0 + s * c, inv = 1.0 / sum;
      rho = WANT_RHO ? b * log(sum) : 0.0;
      rho1 = fmax(DBL_MIN, inv);
    } else {  // HUBER
      if (s > b) {
        const double r = sqrt(s);
        rho = 2.0 * a * r - b;
        rho1 = fmax(DBL_MIN, a / r);
      } else {
        rho = s;
        rho1 = 1.0;
      }
    }
  }
  rho *= weight;
  rho1 *= weight;
}

// ------------------------------------------------------------------------------------------------
// COLMAP ImgFromCam for normalised coordinates (u, v) = (X/Z, Y/Z); T = double or Dual2.
template <class T>
MPBA_HD T fisheye_scale(T u, T v, double k1, double k2, double k3, double k4, bool plain_theta) {
  // theta_d / r with theta = atan(r);  1 where r <= eps (COLMAP leaves the point undistorted)
  const double r2v = val(u) * val(u) + val(v) * val(v);
  if (sqrt(r2v) <= DBL_EPSILON) return one_like(u);
  const T r = dsqrt(u * u + v * v);
  const T theta = datan(r);
  if (plain_theta) return theta / r;
  const T t2 = theta * theta;
  const T poly = 1.0 + t2 * (k1 + t2 * (k2 + t2 * (k3 + t2 * k4)));
  return theta * poly / r;
}

template <class T>
MPBA_HD void img_from_cam_generic(int model, const double* p, T u, T v, T& x, T& y) {
  switch (model) {
    case 0:  // SIMPLE_PINHOLE f, cx, cy
      x = p[0] * u + p[1];
      y = p[0] * v + p[2];
      break;
    case 1:  // PINHOLE fx, fy, cx, cy
      x = p[0] * u + p[2];
      y = p[1] * v + p[3];
      break;
    case 2: {  // SIMPLE_RADIAL f, cx, cy, k
      const T r2 = u * u + v * v;
      const T radial = p[3] * r2;
      x = p[0] * (u + u * radial) + p[1];
      y = p[0] * (v + v * radial) + p[2];
    } break;
    case 3: {  // RADIAL f, cx, cy, k1, k2
      const T r2 = u * u + v * v;
      const T radial = p[3] * r2 + p[4] * r2 * r2;
      x = p[0] * (u + u * radial) + p[1];
      y = p[0] * (v + v * radial) + p[2];
    } break;
    case 4: {  // OPENCV fx, fy, cx, cy, k1, k2, p1, p2
      const T u2 = u * u, v2 = v * v, uv = u * v;
      const T r2 = u2 + v2;
      const T radial = p[4] * r2 + p[5] * r2 * r2;
      const T du = u * radial + 2.0 * p[6] * uv + p[7] * (r2 + 2.0 * u2);
      const T dv = v * radial + 2.0 * p[7] * uv + p[6] * (r2 + 2.0 * v2);
      x = p[0] * (u + du) + p[2];
      y = p[1] * (v + dv) + p[3];
    } break;
    case 5: {  // OPENCV_FISHEYE fx, fy, cx, cy, k1..k4
      const T s = fisheye_scale(u, v, p[4], p[5], p[6], p[7], false);
      x = p[0] * (u * s) + p[2];
      y = p[1] * (v * s) + p[3];
    } break;
    case 6: {  // FULL_OPENCV fx, fy, cx, cy, k1, k2, p1, p2, k3, k4, k5, k6
      const T u2 = u * u, v2 = v * v, uv = u * v;
      const T r2 = u2 + v2, r4 = r2 * r2, r6 = r4 * r2;
      const T radial = (1.0 + p[4] * r2 + p[5] * r4 + p[8] * r6) / (1.0 + p[9] * r2 + p[10] * r4 + p[11] * r6);
      const T du = u * radial + 2.0 * p[6] * uv + p[7] * (r2 + 2.0 * u2) - u;
      const T dv = v * radial + 2.0 * p[7] * uv + p[6] * (r2 + 2.0 * v2) - v;
      x = p[0] * (u + du) + p[2];
      y = p[1] * (v + dv) + p[3];
    } break;
    case 7: {  // FOV fx, fy, cx, cy, omega
      const double omega = p[4], omega2 = omega * omega, kEps = 1e-4;
      const T r2 = u * u + v * v;
      T factor;
      if (omega2 < kEps) {
        factor = (omega2 * r2) / 3.0 - omega2 / 12.0 + 1.0;
      } else {
        const double th = tan(omega / 2.0);
        if (val(r2) < kEps) {
          factor = (-2.0 * th * (4.0 * r2 * th * th - 3.0)) / (3.0 * omega);
        } else {
          const T radius = dsqrt(r2);
          factor = datan(radius * 2.0 * th) / (radius * omega);
        }
      }
      x = p[0] * (u * factor) + p[2];
      y = p[1] * (v * factor) + p[3];
    } break;
    case 8: {  // SIMPLE_RADIAL_FISHEYE f, cx, cy, k
      const T s = fisheye_scale(u, v, p[3], 0.0, 0.0, 0.0, false);
      x = p[0] * (u * s) + p[1];
      y = p[0] * (v * s) + p[2];
    } break;
    case 9: {  // RADIAL_FISHEYE f, cx, cy, k1, k2
      const T s = fisheye_scale(u, v, p[3], p[4], 0.0, 0.0, false);
      x = p[0] * (u * s) + p[1];
      y = p[0] * (v * s) + p[2];
    } break;
    default: {  // 10 THIN_PRISM_FISHEYE fx, fy, cx, cy, k1, k2, p1, p2, k3, k4, sx1, sy1
      const T s = fisheye_scale(u, v, 0.0, 0.0, 0.0, 0.0, true);
      const T uu = u * s, vv = v * s;
      const T u2 = uu * uu, v2 = vv * vv, uv = uu * vv;
      const T t2 = u2 + v2, t4 = t2 * t2;
      const T radial = p[4] * t2 + p[5] * t4 + p[8] * t4 * t2 + p[9] * t4 * t4;
      const T du = uu * radial + 2.0 * p[6] * uv + p[7] * (t2 + 2.0 * u2) + p[10] * t2;
      const T dv = vv * radial + 2.0 * p[7] * uv + p[6] * (t2 + 2.0 * v2) + p[11] * t2;
      x = p[0] * (uu + du) + p[2];
      y = p[1] * (vv + dv) + p[3];
    } break;
  }
}

// value + 2x2 Jacobian D = d(x,y)/d(u,v) (row-major).  Closed forms for the five models the
// reference pipeline can meet in practice, duals for the rest.
MPBA_HD void img_from_cam_jac(int model, const double* p, double u, double v, double& x, double& y, double D[4]) {
  switch (model) {
    case 0:
      x = p[0] * u + p[1];
      y = p[0] * v + p[2];
      D[0] = p[0]; D[1] = 0.0; D[2] = 0.0; D[3] = p[0];
      return;
    case 1:
      x = p[0] * u + p[2];
      y = p[1] * v + p[3];
      D[0] = p[0]; D[1] = 0.0; D[2] = 0.0; D[3] = p[1];
      return;
    case 2: {
      const double r2 = u * u + v * v, k = p[3], f = p[0];
      const double radial = k * r2;
      x = f * (u + u * radial) + p[1];
      y = f * (v + v * radial) + p[2];
      const double g = 1.0 + radial, kuv2 = 2.0 * k * u * v;
      D[0] = f * (g + 2.0 * k * u * u); D[1] = f * kuv2; D[2] = f * kuv2; D[3] = f * (g + 2.0 * k * v * v);
      return;
    }
    case 3: {
      const double r2 = u * u + v * v, f = p[0];
      const double radial = p[3] * r2 + p[4] * r2 * r2;
      x = f * (u + u * radial) + p[1];
      y = f * (v + v * radial) + p[2];
      const double g = 1.0 + radial, gp = p[3] + 2.0 * p[4] * r2, c = 2.0 * u * v * gp;
      D[0] = f * (g + 2.0 * u * u * gp); D[1] = f * c; D[2] = f * c; D[3] = f * (g + 2.0 * v * v * gp);
      return;
    }
    case 4: {
      const double k1 = p[4], k2 = p[5], p1 = p[6], p2 = p[7];
      const double u2 = u * u, v2 = v * v, uv = u * v, r2 = u2 + v2;
      const double radial = k1 * r2 + k2 * r2 * r2;
      const double du = u * radial + 2.0 * p1 * uv + p2 * (r2 + 2.0 * u2);
      const double dv = v * radial + 2.0 * p2 * uv + p1 * (r2 + 2.0 * v2);
      x = p[0] * (u + du) + p[2];
      y = p[1] * (v + dv) + p[3];
      const double gp = k1 + 2.0 * k2 * r2;  // d radial / d r2
      const double cross = 2.0 * (p1 * u + p2 * v + uv * gp);
      D[0] = p[0] * (1.0 + radial + 2.0 * u2 * gp + 2.0 * p1 * v + 6.0 * p2 * u);
      D[1] = p[0] * cross;
      D[2] = p[1] * cross;
      D[3] = p[1] * (1.0 + radial + 2.0 * v2 * gp + 6.0 * p1 * v + 2.0 * p2 * u);
      return;
    }
    default: {
      Dual2 xd, yd;
      img_from_cam_generic<Dual2>(model, p, Dual2{u, 1.0, 0.0}, Dual2{v, 0.0, 1.0}, xd, yd);
      x = xd.a;
      y = yd.a;
      D[0] = xd.du; D[1] = xd.dv; D[2] = yd.du; D[3] = yd.dv;
      return;
    }
  }
}

// ------------------------------------------------------------------------------------------------
// small dense helpers; symmetric 3x3 stored as (xx, xy, xz, yy, yz, zz)
MPBA_HD void quat_to_rotmat(const double q[4], double R[9]) {
  const double x = q[0], y = q[1], z = q[2], w = q[3];
  R[0] = 1.0 - 2.0 * (y * y + z * z); R[1] = 2.0 * (x * y - z * w);       R[2] = 2.0 * (x * z + y * w);
  R[3] = 2.0 * (x * y + z * w);       R[4] = 1.0 - 2.0 * (x * x + z * z); R[5] = 2.0 * (y * z - x * w);
  R[6] = 2.0 * (x * z - y * w);       R[7] = 2.0 * (y * z + x * w);       R[8] = 1.0 - 2.0 * (x * x + y * y);
}

// Eigen's Quaternion::_transformVector: v + 2w (u x v) + 2 u x (u x v)
MPBA_HD void quat_rotate(const double q[4], const double v[3], double out[3]) {
  double uvx = q[1] * v[2] - q[2] * v[1], uvy = q[2] * v[0] - q[0] * v[2], uvz = q[0] * v[1] - q[1] * v[0];
  uvx += uvx; uvy += uvy; uvz += uvz;
  out[0] = v[0] + q[3] * uvx + (q[1] * uvz - q[2] * uvy);
  out[1] = v[1] + q[3] * uvy + (q[2] * uvx - q[0] * uvz);
  out[2] = v[2] + q[3] * uvz + (q[0] * uvy - q[1] * uvx);
}

// ceres::EigenQuaternionManifold::Plus: q (+) d = [sin|d| d/|d|, cos|d|] (x) q
MPBA_HD void quat_plus(const double q[4], const double d[3], double out[4]) {
  const double n2 = d[0] * d[0] + d[1] * d[1] + d[2] * d[2];
  if (n2 == 0.0) {
    out[0] = q[0]; out[1] = q[1]; out[2] = q[2]; out[3] = q[3];
    return;
  }
  const double n = sqrt(n2), sbd = sin(n) / n;
  const double ax = d[0] * sbd, ay = d[1] * sbd, az = d[2] * sbd, aw = cos(n);
  const double bx = q[0], by = q[1], bz = q[2], bw = q[3];
  out[0] = aw * bx + ax * bw + ay * bz - az * by;
  out[1] = aw * by - ax * bz + ay * bw + az * bx;
  out[2] = aw * bz + ax * by - ay * bx + az * bw;
  out[3] = aw * bw - ax * bx - ay * by - az * bz;
}

MPBA_HD void cross3(const double a[3], const double b[3], double c[3]) {
  c[0] = a[1] * b[2] - a[2] * b[1];
  c[1] = a[2] * b[0] - a[0] * b[2];
  c[2] = a[0] * b[1] - a[1] * b[0];
}

MPBA_HD void sym3_mul_vec(const double S[6], const double v[3], double o[3]) {
  o[0] = S[0] * v[0] + S[1] * v[1] + S[2] * v[2];
  o[1] = S[1] * v[0] + S[3] * v[1] + S[4] * v[2];
  o[2] = S[2] * v[0] + S[4] * v[1] + S[5] * v[2];
}

// Q = N R  (N symmetric 3x3, R row-major 3x3)
MPBA_HD void sym3_mul_mat(const double N[6], const double R[9], double Q[9]) {
  for (int j = 0; j < 3; ++j) {
    const double r0 = R[j], r1 = R[3 + j], r2 = R[6 + j];
    Q[j] = N[0] * r0 + N[1] * r1 + N[2] * r2;
    Q[3 + j] = N[1] * r0 + N[3] * r1 + N[4] * r2;
    Q[6 + j] = N[2] * r0 + N[4] * r1 + N[5] * r2;
  }
}

// Cholesky of a symmetric positive definite 3x3 (what Ceres' Schur eliminator does with Eigen's LLT
// on the e-block): returns Li = L^-1 (lower: i00, i10, i11, i20, i21, i22) so that H^-1 = Li^T Li;
// false if H is not positive definite.
// The factor is formed through reciprocal square roots (three rsqrt instead of three sqrt + six divisions: the
// per-point algebra sits on the critical path of k_points).
MPBA_HD double mpba_rsqrt(double x) {
#ifdef __CUDA_ARCH__
  return rsqrt(x);
#else
  return 1.0 / sqrt(x);
#endif
}
MPBA_HD bool sym3_chol_inverse(const double H[6], double Li[6]) {
  const double a = H[0], b = H[1], c = H[2], d = H[3], e = H[4], f = H[5];
  if (!(a > 0.0)) return false;
  const double i00 = mpba_rsqrt(a), l10 = b * i00, l20 = c * i00;
  const double t11 = d - l10 * l10;
  if (!(t11 > 0.0)) return false;
  const double i11 = mpba_rsqrt(t11), l21 = (e - l20 * l10) * i11;
  const double t22 = f - l20 * l20 - l21 * l21;
  if (!(t22 > 0.0)) return false;
  const double i22 = mpba_rsqrt(t22);
  const double i10 = -l10 * i00 * i11;
  const double i21 = -l21 * i11 * i22;
  const double i20 = -(l20 * i00 + l21 * i10) * i22;
  Li[0] = i00; Li[1] = i10; Li[2] = i11; Li[3] = i20; Li[4] = i21; Li[5] = i22;
  return true;
}

// H^-1 = Li^T Li as (xx, xy, xz, yy, yz, zz)
MPBA_HD void sym3_from_chol_inverse(const double Li[6], double Hi[6]) {
  const double i00 = Li[0], i10 = Li[1], i11 = Li[2], i20 = Li[3], i21 = Li[4], i22 = Li[5];
  Hi[0] = i00 * i00 + i10 * i10 + i20 * i20;
  Hi[1] = i10 * i11 + i20 * i21;
  Hi[2] = i20 * i22;
  Hi[3] = i11 * i11 + i21 * i21;
  Hi[4] = i21 * i22;
  Hi[5] = i22 * i22;
}

MPBA_HD bool sym3_inverse(const double H[6], double Hi[6]) {
  double Li[6];
  if (!sym3_chol_inverse(H, Li)) return false;
  sym3_from_chol_inverse(Li, Hi);
  return true;
}

// Z = Q Li^T (3x3 row-major), so that Za Zb^T = Qa H^-1 Qb^T
MPBA_HD void mul_chol_inverse_t(const double Q[9], const double Li[6], double Z[9]) {
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    const double q0 = Q[3 * r], q1 = Q[3 * r + 1], q2 = Q[3 * r + 2];
    Z[3 * r] = q0 * Li[0];
    Z[3 * r + 1] = q0 * Li[1] + q1 * Li[2];
    Z[3 * r + 2] = q0 * Li[3] + q1 * Li[4] + q2 * Li[5];
  }
}

// ------------------------------------------------------------------------------------------------
struct LossSpec {
  int reproj_loss, depth_loss;
  double reproj_scale, reproj_weight;
  // 0 (default): COLMAP >= 3.11's ReprojErrorCostFunctor -- ImgFromCam fails for a point at z < epsilon in the camera
  // frame and the functor then returns a ZERO residual (and, through the jets, a zero Jacobian) instead of inf;
  // 1: strict -- the projection is evaluated as written (inf / nan residual, the solve reports it)
  int behind_strict;
  // loss scale of the point-to-plane normal-prior rows (EXTENSION, see linearize_obs); the loss type is depth_loss
  double plane_scale;
};

struct ObsLin {
  double N[6];     // A'^T A'
  double h[3];     // A'^T r'
  double prot[3];  // R(q) X
  double rho;      // rho_reproj + rho_depth of the two blocks (not halved)
  double r[3];     // raw residuals (x, y, log-depth); r[2] = 0 without a prior
  double rn;       // raw point-to-plane residual of the normal-prior row (extension); 0 without one
};

// One observation.  pose = {qx,qy,qz,qw,tx,ty,tz}, R = rotmat(q), X the point, (d, w, a) the depth
// prior (w == 0: none), (b, es) the image's shift and exp(log-scale) (the kernels keep exp(s) in slot 7 of the
// pose record: one exp per image and state instead of one per observation).
//
// plane != nullptr: EXTENSION without a counterpart in the reference (the reference has no Ceres
// residual for normals; they act through the depth-integration solver).  One more 1-D row per observation, the
// point-to-plane distance to the monocular surface patch at the keypoint,
//     r_n = m . p_cam - c,      m = sqrt(w_n) n,   c = sqrt(w_n) n . q,
// n the prior unit normal in the camera frame, q the prior surface point (prior depth times the keypoint's ray), both
// folded by the host into plane = {m_x, m_y, m_z, c} (m = 0: no row).  Loss: ScaledLoss(depth_loss(plane_scale), 1).
// Linear in p_cam: dr_n / dp_cam = m.  Defined in oracle/residuals.py; parity "n/a" (nothing to compare with upstream).
template <bool WANT_JAC, bool WANT_RHO>
MPBA_HD void linearize_obs(const double* pose, const double R[9], const double X[3], int model, const double* prm,
                           double ox, double oy, double d, double w, double a, double b, double es,
                           const LossSpec& ls, ObsLin& o, const double* plane = nullptr) {
  o.prot[0] = R[0] * X[0] + R[1] * X[1] + R[2] * X[2];
  o.prot[1] = R[3] * X[0] + R[4] * X[1] + R[5] * X[2];
  o.prot[2] = R[6] * X[0] + R[7] * X[1] + R[8] * X[2];
  const double px = o.prot[0] + pose[4], py = o.prot[1] + pose[5], pz = o.prot[2] + pose[6];
  const double iz = 1.0 / pz, u = px * iz, v = py * iz;
  const bool behind = !ls.behind_strict && pz < DBL_EPSILON;
  double r0 = 0.0, r1 = 0.0, rho = 0.0, rho1 = 0.0;
  double x, y, D[4] = {0.0, 0.0, 0.0, 0.0};
  if (!behind) {
    if (WANT_JAC) {
      img_from_cam_jac(model, prm, u, v, x, y, D);
    } else {
      img_from_cam_generic<double>(model, prm, u, v, x, y);
    }
    r0 = x - ox;
    r1 = y - oy;
    loss_eval<WANT_RHO>(ls.reproj_loss, ls.reproj_scale, ls.reproj_weight, r0 * r0 + r1 * r1, rho, rho1);
  }
  o.r[0] = r0;
  o.r[1] = r1;
  o.rho = rho;
  double a0[3], a1[3];
  if (WANT_JAC) {
    if (behind) {
#pragma unroll
      for (int j = 0; j < 6; ++j) o.N[j] = 0.0;
      o.h[0] = o.h[1] = o.h[2] = 0.0;
    } else {
      // rows of d(x,y)/dp = D * [[1/z, 0, -u/z], [0, 1/z, -v/z]]
      a0[0] = D[0] * iz; a0[1] = D[1] * iz; a0[2] = -(D[0] * u + D[1] * v) * iz;
      a1[0] = D[2] * iz; a1[1] = D[3] * iz; a1[2] = -(D[2] * u + D[3] * v) * iz;
      o.N[0] = rho1 * (a0[0] * a0[0] + a1[0] * a1[0]);
      o.N[1] = rho1 * (a0[0] * a0[1] + a1[0] * a1[1]);
      o.N[2] = rho1 * (a0[0] * a0[2] + a1[0] * a1[2]);
      o.N[3] = rho1 * (a0[1] * a0[1] + a1[1] * a1[1]);
      o.N[4] = rho1 * (a0[1] * a0[2] + a1[1] * a1[2]);
      o.N[5] = rho1 * (a0[2] * a0[2] + a1[2] * a1[2]);
      o.h[0] = rho1 * (a0[0] * r0 + a1[0] * r1);
      o.h[1] = rho1 * (a0[1] * r0 + a1[1] * r1);
      o.h[2] = rho1 * (a0[2] * r0 + a1[2] * r1);
    }
  }
  o.r[2] = 0.0;
  if (w > 0.0) {
    const double r2 = log((es * d + b) * iz);
    o.r[2] = r2;
    double rhod, rhod1;
    loss_eval<WANT_RHO>(ls.depth_loss, a, w, r2 * r2, rhod, rhod1);
    o.rho += rhod;
    if (WANT_JAC) {
      // a2 = (0, 0, -1/z)
      o.N[5] += rhod1 * iz * iz;
      o.h[2] -= rhod1 * iz * r2;
    }
  }
  o.rn = 0.0;
  if (plane != nullptr) {
    const double m0 = plane[0], m1 = plane[1], m2 = plane[2];
    if (m0 != 0.0 || m1 != 0.0 || m2 != 0.0) {
      const double rn = m0 * px + m1 * py + m2 * pz - plane[3];
      o.rn = rn;
      double rhon, rhon1;
      loss_eval<WANT_RHO>(ls.depth_loss, ls.plane_scale, 1.0, rn * rn, rhon, rhon1);
      o.rho += rhon;
      if (WANT_JAC) {
        o.N[0] += rhon1 * m0 * m0; o.N[1] += rhon1 * m0 * m1; o.N[2] += rhon1 * m0 * m2;
        o.N[3] += rhon1 * m1 * m1; o.N[4] += rhon1 * m1 * m2; o.N[5] += rhon1 * m2 * m2;
        o.h[0] += rhon1 * m0 * rn; o.h[1] += rhon1 * m1 * rn; o.h[2] += rhon1 * m2 * rn;
      }
    }
  }
}

// G^T h = [2 p x h ; h]
MPBA_HD void gt_vec(const double p[3], const double h[3], double o[6]) {
  double c[3];
  cross3(p, h, c);
  o[0] = 2.0 * c[0]; o[1] = 2.0 * c[1]; o[2] = 2.0 * c[2];
  o[3] = h[0]; o[4] = h[1]; o[5] = h[2];
}

// G y = y_t - 2 p x y_r
MPBA_HD void g_vec(const double p[3], const double y[6], double o[3]) {
  double c[3];
  cross3(p, y, c);
  o[0] = y[3] - 2.0 * c[0];
  o[1] = y[4] - 2.0 * c[1];
  o[2] = y[5] - 2.0 * c[2];
}

// explicit G (3x6 row-major) = [-2 [p]x | I]
MPBA_HD void g_matrix(const double p[3], double G[18]) {
  G[0] = 0.0;          G[1] = 2.0 * p[2];   G[2] = -2.0 * p[1];  G[3] = 1.0; G[4] = 0.0; G[5] = 0.0;
  G[6] = -2.0 * p[2];  G[7] = 0.0;          G[8] = 2.0 * p[0];   G[9] = 0.0; G[10] = 1.0; G[11] = 0.0;
  G[12] = 2.0 * p[1];  G[13] = -2.0 * p[0]; G[14] = 0.0;         G[15] = 0.0; G[16] = 0.0; G[17] = 1.0;
}

// upper triangle (21 values, row-major: (0,0..5),(1,1..5),...) of G^T N G with G = [P | I], P = -2 [p]x:
//   [ P^T N P   P^T N ]        T = N P (3x3),  P^T N = T^T,  P^T N P = P^T T
//   [   N P       N   ]
// written out so that the structural zeros / ones of G cost nothing (about 45 flops instead of 117).
MPBA_HD void gt_n_g_upper(const double p[3], const double N[6], double H[21]) {
  const double qx = 2.0 * p[0], qy = 2.0 * p[1], qz = 2.0 * p[2];
  // columns of P: (0, -qz, qy), (qz, 0, -qx), (-qy, qx, 0);  rows of the symmetric N: (N0 N1 N2), (N1 N3 N4), (N2 N4 N5)
  double T[9];  // T[3 r + c] = sum_k N[r][k] P[k][c]
  T[0] = qy * N[2] - qz * N[1]; T[1] = qz * N[0] - qx * N[2]; T[2] = qx * N[1] - qy * N[0];
  T[3] = qy * N[4] - qz * N[3]; T[4] = qz * N[1] - qx * N[4]; T[5] = qx * N[3] - qy * N[1];
  T[6] = qy * N[5] - qz * N[4]; T[7] = qz * N[2] - qx * N[5]; T[8] = qx * N[4] - qy * N[2];
  // (P^T T)[a][b] = sum_k P[k][a] T[k][b]
  H[0] = qy * T[6] - qz * T[3];   // (0,0)
  H[1] = qy * T[7] - qz * T[4];   // (0,1)
  H[2] = qy * T[8] - qz * T[5];   // (0,2)
  H[3] = T[0]; H[4] = T[3]; H[5] = T[6];  // (0,3..5) = T^T row 0
  H[6] = qz * T[1] - qx * T[7];   // (1,1)
  H[7] = qz * T[2] - qx * T[8];   // (1,2)
  H[8] = T[1]; H[9] = T[4]; H[10] = T[7];  // (1,3..5)
  H[11] = qx * T[5] - qy * T[2];  // (2,2)
  H[12] = T[2]; H[13] = T[5]; H[14] = T[8];  // (2,3..5)
  H[15] = N[0]; H[16] = N[1]; H[17] = N[2];  // (3,3..5)
  H[18] = N[3]; H[19] = N[4];                // (4,4..5)
  H[20] = N[5];                              // (5,5)
}

// M (6x6 row-major) = Ga^T K Gb for a 3x3 K (row-major)
MPBA_HD void gt_k_g(const double pa[3], const double K[9], const double pb[3], double M[36]) {
  double Ga[18], Gb[18], T[18];
  g_matrix(pa, Ga);
  g_matrix(pb, Gb);
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 6; ++j) T[6 * i + j] = K[3 * i] * Gb[j] + K[3 * i + 1] * Gb[6 + j] + K[3 * i + 2] * Gb[12 + j];
  for (int i = 0; i < 6; ++i)
    for (int j = 0; j < 6; ++j) M[6 * i + j] = Ga[i] * T[j] + Ga[6 + i] * T[6 + j] + Ga[12 + i] * T[12 + j];
}

// Ceres LM damping of one tangent coordinate in unscaled variables:
//   scaled system adds clamp(c^2 H_jj, lo, hi) / radius to the diagonal of diag(c) H diag(c),
//   i.e. the unscaled one adds clamp(c^2 H_jj) / (radius c^2).
// same value up to rounding without a division in the common (unclamped) case: clamp(c^2 h) / (radius c^2) = h / radius
MPBA_HD double lm_damping_fast(double hjj, double c, double radius, double inv_radius, double lo, double hi) {
  const double cc = c * c, t = cc * hjj;
  if (t < lo) return lo / (radius * cc);
  if (t > hi) return hi / (radius * cc);
  return hjj * inv_radius;
}

MPBA_HD double lm_damping(double hjj, double c, double radius, double lo, double hi) {
  const double cc = c * c;
  double dj = cc * hjj;
  dj = dj < lo ? lo : (dj > hi ? hi : dj);
  return dj / (radius * cc);
}

// ------------------------------------------------------------------------------------------------
// Post-BA track filter.  The reference calls pycolmap's
// ObservationManager.filter_points3D(max_reproj_error, min_tri_angle, point3D_ids) after every bundle
// (mpsfm/sfm/mapper/base.py:700-711, 772-797); this restates COLMAP's published two-phase rule per track:
//   1. an observation whose squared reprojection error exceeds max_err^2 is deleted (a point at
//      z < DBL_EPSILON in the camera frame has error DBL_MAX); a track left with fewer than two
//      observations deletes its point; a surviving point's error is the mean of the kept errors,
//   2. a surviving point whose kept observations have no pair of projection centres c = -R^T t that
//      subtends at least min_tri_angle at the point is deleted.
// num_filtered counts observations as COLMAP does (whole track when the point goes).

// squared reprojection error of X in the image with pose = {qx,qy,qz,qw,tx,ty,tz}
MPBA_HD double reproj_error2(const double* pose, int model, const double* prm, const double X[3], double ox,
                             double oy) {
  double R[9];
  quat_to_rotmat(pose, R);
  const double px = R[0] * X[0] + R[1] * X[1] + R[2] * X[2] + pose[4];
  const double py = R[3] * X[0] + R[4] * X[1] + R[5] * X[2] + pose[5];
  const double pz = R[6] * X[0] + R[7] * X[1] + R[8] * X[2] + pose[6];
  if (pz < DBL_EPSILON) return DBL_MAX;
  double x, y;
  img_from_cam_generic<double>(model, prm, px / pz, py / pz, x, y);
  const double r0 = x - ox, r1 = y - oy;
  return r0 * r0 + r1 * r1;
}

MPBA_HD void projection_center(const double* pose, double c[3]) {
  double R[9];
  quat_to_rotmat(pose, R);
  c[0] = -(R[0] * pose[4] + R[3] * pose[5] + R[6] * pose[6]);
  c[1] = -(R[1] * pose[4] + R[4] * pose[5] + R[7] * pose[6]);
  c[2] = -(R[2] * pose[4] + R[5] * pose[5] + R[8] * pose[6]);
}

// angle at X between the rays to c1 and c2, folded into [0, pi/2]; law of cosines on squared lengths, the ratio is
// NOT clamped (a ratio pushed past 1 by rounding gives nan, which fails every ">= min angle" test, as in COLMAP)
MPBA_HD double triangulation_angle(const double c1[3], const double c2[3], const double X[3]) {
  double b2 = 0.0, r1 = 0.0, r2 = 0.0;
  for (int j = 0; j < 3; ++j) {
    const double b = c1[j] - c2[j], u = X[j] - c1[j], v = X[j] - c2[j];
    b2 += b * b;
    r1 += u * u;
    r2 += v * v;
  }
  const double den = 2.0 * sqrt(r1 * r2);
  if (den == 0.0) return 0.0;
  const double ang = fabs(acos((r1 + r2 - b2) / den));
  const double oth = 3.14159265358979323846 - ang;
  return oth < ang ? oth : ang;
}

// One track: observations k0 .. k1 (track order) of point X.  pose8 [C][8] (unit quaternions), keep [N] gets 1 / 0
// per observation.  Returns true when the point survives; err = mean kept reprojection error (-1 when it does not),
// filtered = observations removed.
MPBA_HD bool filter_track(const double* pose8, const int32_t* img_camera, const int32_t* cam_model,
                          const double* cam_params, int max_prm, const int32_t* pm_img, const double* pm_xy, int k0,
                          int k1, const double X[3], double max_err2, double min_angle, uint8_t* keep, double& err,
                          int& filtered) {
  const int len = k1 - k0;
  int kept = 0;
  double sum = 0.0;
  for (int k = k0; k < k1; ++k) {
    const int img = pm_img[k], cam = img_camera[img];
    const double e2 = reproj_error2(pose8 + 8 * (size_t)img, cam_model[cam], cam_params + (size_t)max_prm * cam, X,
                                    pm_xy[2 * (size_t)k], pm_xy[2 * (size_t)k + 1]);
    const bool ok = !(e2 > max_err2);
    keep[k] = ok ? 1 : 0;
    if (ok) {
      sum += sqrt(e2);
      ++kept;
    }
  }
  bool alive = kept >= 2;  // COLMAP: deleted when (len - kept) >= len - 1, which covers len < 2
  if (alive) {
    filtered = len - kept;
    err = sum / kept;
    alive = false;
    for (int a = k0 + 1; a < k1 && !alive; ++a) {
      if (!keep[a]) continue;
      double ca[3];
      projection_center(pose8 + 8 * (size_t)pm_img[a], ca);
      for (int b = k0; b < a; ++b) {
        if (!keep[b]) continue;
        double cb[3];
        projection_center(pose8 + 8 * (size_t)pm_img[b], cb);
        if (triangulation_angle(ca, cb, X) >= min_angle) {
          alive = true;
          break;
        }
      }
    }
    if (!alive) filtered += kept;
  } else {
    filtered = len;
  }
  if (!alive) {
    for (int k = k0; k < k1; ++k) keep[k] = 0;
    err = -1.0;
  }
  return alive;
}

}  // namespace mpba
