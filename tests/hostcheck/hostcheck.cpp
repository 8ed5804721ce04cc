// hostcheck.cpp -- TEST-ONLY: compiles the product's per-observation arithmetic (mpsfm_b200/csrc/ba_math.cuh,
// the same source the sm_100a kernels inline) with g++ so that the CPU suite can compare it with the
// numpy oracle.  Nothing in mpsfm_b200/ loads this library.
#include "../../mpsfm_b200/csrc/ba_math.cuh"

using namespace mpba;

extern "C" {

// per observation: pose [n][8], X [n][3], model [n], prm [n][12], xy [n][2], dwa [n][3], bs [n][2]
// out [n][40]: N(6) h(3) prot(3) rho(1) r(3) Hcc_upper(21) gc(6)  = 43 -> use 48 stride
// plane [n][4] = {m, c} of the normal-prior extension or NULL; out[43] = its raw residual
void hc_obs_plane(int n, const double* pose, const double* X, const int* model, const double* prm, const double* xy,
                  const double* dwa, const double* bs, int reproj_loss, int depth_loss, double reproj_scale,
                  double reproj_weight, int want_jac, const double* plane, double plane_scale, double* out) {
  LossSpec ls{reproj_loss, depth_loss, reproj_scale, reproj_weight};
  ls.plane_scale = plane_scale;
  for (int i = 0; i < n; ++i) {
    double R[9];
    quat_to_rotmat(pose + 8 * i, R);
    ObsLin o;
    for (int k = 0; k < 6; ++k) o.N[k] = 0;
    for (int k = 0; k < 3; ++k) o.h[k] = 0;
    if (want_jac)
      linearize_obs<true, true>(pose + 8 * i, R, X + 3 * i, model[i], prm + 12 * i, xy[2 * i], xy[2 * i + 1], dwa[3 * i],
                                dwa[3 * i + 1], dwa[3 * i + 2], bs[2 * i], exp(bs[2 * i + 1]), ls, o, plane ? plane + 4 * i : nullptr);
    else
      linearize_obs<false, true>(pose + 8 * i, R, X + 3 * i, model[i], prm + 12 * i, xy[2 * i], xy[2 * i + 1], dwa[3 * i],
                                 dwa[3 * i + 1], dwa[3 * i + 2], bs[2 * i], exp(bs[2 * i + 1]), ls, o, plane ? plane + 4 * i : nullptr);
    double* d = out + 48 * i;
    for (int k = 0; k < 6; ++k) d[k] = o.N[k];
    for (int k = 0; k < 3; ++k) d[6 + k] = o.h[k];
    for (int k = 0; k < 3; ++k) d[9 + k] = o.prot[k];
    d[12] = o.rho;
    for (int k = 0; k < 3; ++k) d[13 + k] = o.r[k];
    if (want_jac) {
      gt_n_g_upper(o.prot, o.N, d + 16);
      gt_vec(o.prot, o.h, d + 37);
    }
    d[43] = o.rn;
  }
}

void hc_obs(int n, const double* pose, const double* X, const int* model, const double* prm, const double* xy,
            const double* dwa, const double* bs, int reproj_loss, int depth_loss, double reproj_scale,
            double reproj_weight, int want_jac, double* out) {
  hc_obs_plane(n, pose, X, model, prm, xy, dwa, bs, reproj_loss, depth_loss, reproj_scale, reproj_weight, want_jac,
               nullptr, 1.0, out);
}

void hc_gt_k_g(const double* pa, const double* K, const double* pb, double* M) { gt_k_g(pa, K, pb, M); }
int hc_sym3_inverse(const double* H, double* Hi) { return sym3_inverse(H, Hi) ? 1 : 0; }
void hc_quat_plus(const double* q, const double* d, double* out) { quat_plus(q, d, out); }
double hc_lm_damping(double hjj, double c, double radius, double lo, double hi) { return lm_damping(hjj, c, radius, lo, hi); }
void hc_g_vec(const double* p, const double* y, double* o) { g_vec(p, y, o); }
void hc_rotmat(const double* q, double* R) { quat_to_rotmat(q, R); }
void hc_project(int model, const double* prm, double u, double v, double* xy, double* D) {
  img_from_cam_jac(model, prm, u, v, xy[0], xy[1], D);
  double x2, y2;
  img_from_cam_generic<double>(model, prm, u, v, x2, y2);
  xy[2] = x2;
  xy[3] = y2;
}
// the post-BA track filter exactly as k_filter_tracks runs it (ba_math.cuh: filter_track), track-order arrays
long long hc_filter_tracks(int P, const int* pt_ptr, const double* pose8, const int* img_camera, const int* cam_model,
                           const double* cam_params, const int* pm_img, const double* pm_xy, const double* X,
                           double max_err, double min_deg, const unsigned char* select, unsigned char* obs_keep,
                           unsigned char* pt_keep, double* pt_error) {
  long long total = 0;
  for (int p = 0; p < P; ++p) {
    const int k0 = pt_ptr[p], k1 = pt_ptr[p + 1];
    bool alive = true;
    double err = -1.0;
    int filtered = 0;
    if (!select || select[p]) {
      alive = filter_track(pose8, img_camera, cam_model, cam_params, 12, pm_img, pm_xy, k0, k1, X + 3 * (size_t)p,
                           max_err * max_err, min_deg * (3.14159265358979323846 / 180.0), obs_keep, err, filtered);
    } else {
      for (int k = k0; k < k1; ++k) obs_keep[k] = 1;
    }
    pt_keep[p] = alive ? 1 : 0;
    pt_error[p] = err;
    total += filtered;
  }
  return total;
}
double hc_reproj_error2(const double* pose, int model, const double* prm, const double* X, double ox, double oy) {
  return reproj_error2(pose, model, prm, X, ox, oy);
}
double hc_triangulation_angle(const double* c1, const double* c2, const double* X) {
  return triangulation_angle(c1, c2, X);
}
}
