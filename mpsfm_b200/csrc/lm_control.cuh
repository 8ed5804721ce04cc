// lm_control.cuh -- one Levenberg-Marquardt decision on the device (see lm_loop.cu); shared by k_lm_control and the
// fused image tail kernel of ba_kernels.cu.
#pragma once

#include <cfloat>

#include "ba_kernels.cuh"

namespace mpba {

// One LM decision.  Mirrors solve_impl's loop body in ba_solver.cu line by line (the comments there carry the Ceres
// references); `msg` codes are decoded by lm_message().
// Called by ONE thread; returns true while the loop goes on.
__device__ __forceinline__ bool lm_control_step(double* ld, mpba_iteration* rows, double* scal, const LmControl& o) {
  const double* s = scal;
  int nrows = (int)ld[LD_NROWS];
  mpba_iteration last = rows[nrows - 1];
  int termination = -1, msg = 0, accept = 0;
  double radius = ld[LD_RADIUS], decrease = ld[LD_DECREASE], x_cost = ld[LD_X_COST];
  const double fixed_cost = ld[LD_FIXED_COST];
  int iteration = (int)ld[LD_ITERATION];
  bool done = false;

  if (ld[LD_NEED_GMAX] != 0.0) {
    const double g = fmax(ld[LD_GMAX_C], s[S_GMAX_P]);
    last.gradient_max_norm = g;
    rows[nrows - 1].gradient_max_norm = g;
    ld[LD_NEED_GMAX] = 0.0;
    if (g <= o.gradient_tolerance) {
      termination = MPBA_CONVERGENCE;
      msg = LM_MSG_GRADIENT;
      done = true;
    }
  }
  if (!done) {
    mpba_iteration it = {};
    it.cost = last.cost;
    it.gradient_max_norm = last.gradient_max_norm;
    it.pcg_iterations = (int32_t)s[S_PCG_ITERS];
    ld[LD_TOTAL_PCG] += (double)it.pcg_iterations;
    ++iteration;
    const double mcc = s[S_MCC] + s[S_MCC_C];
    const bool ok = s[S_PCG_FLAG] != 2.0 && s[S_BAD_POINTS] == 0.0 && isfinite(mcc) && isfinite(s[S_STEP2_P]) &&
                    isfinite(s[S_STEP2_C]) && mcc > 0.0;
    it.step_is_valid = ok ? 1 : 0;
    if (!ok) {
      const int invalid_run = (int)ld[LD_INVALID_RUN] + 1;
      ld[LD_INVALID_RUN] = invalid_run;
      radius /= decrease;
      decrease *= 2.0;
      it.trust_region_radius = radius;
      ld[LD_N_BAD] += 1.0;
      rows[nrows++] = it;
      if (invalid_run >= o.max_consecutive_invalid) {
        termination = MPBA_FAILURE;
        msg = LM_MSG_INVALID;
        done = true;
      }
    } else {
      ld[LD_INVALID_RUN] = 0.0;
      double cand_cost = 0.5 * s[S_COST_FREE];
      if (!isfinite(cand_cost)) cand_cost = DBL_MAX;
      it.step_norm = sqrt(s[S_STEP2_P] + s[S_STEP2_C]);
      const double x_norm = sqrt(s[S_XNORM2_P] + s[S_XNORM2_C]);
      if (it.step_norm <= o.parameter_tolerance * (x_norm + o.parameter_tolerance)) {
        termination = MPBA_CONVERGENCE;
        msg = LM_MSG_PARAMETER;
        rows[nrows++] = it;
        done = true;
      } else {
        it.cost_change = x_cost - cand_cost;
        if (fabs(it.cost_change) <= o.function_tolerance * x_cost) {
          termination = MPBA_CONVERGENCE;
          msg = LM_MSG_FUNCTION;
          rows[nrows++] = it;
          done = true;
        } else {
          it.relative_decrease = it.cost_change / mcc;
          if (it.relative_decrease > o.min_relative_decrease) {
            accept = 1;
            x_cost = cand_cost;
            ld[LD_GMAX_C] = s[S_GMAX_C];
            ld[LD_NEED_GMAX] = 1.0;
            it.cost = x_cost + fixed_cost;
            it.step_is_successful = 1;
            const double t = __dsub_rn(__dmul_rn(2.0, it.relative_decrease), 1.0);
            const double t3 = __dmul_rn(__dmul_rn(t, t), t);
            radius = fmin(o.max_trust_region_radius, radius / fmax(1.0 / 3.0, __dsub_rn(1.0, t3)));
            decrease = 2.0;
            ld[LD_N_OK] += 1.0;
            if (x_cost < ld[LD_MIN_COST]) ld[LD_MIN_COST] = x_cost;
          } else {
            radius /= decrease;
            decrease *= 2.0;
            ld[LD_N_BAD] += 1.0;
          }
          it.trust_region_radius = radius;
          rows[nrows++] = it;
        }
      }
    }
    if (!done) {
      if (iteration >= o.max_num_iterations) {
        termination = MPBA_NO_CONVERGENCE;
        msg = LM_MSG_MAX_ITERATIONS;
      } else if (radius <= o.min_trust_region_radius) {
        termination = MPBA_CONVERGENCE;
        msg = LM_MSG_MIN_RADIUS;
      }
    }
  }
  ld[LD_RADIUS] = radius;
  ld[LD_INV_RADIUS] = 1.0 / radius;
  ld[LD_FIRST] = 0.0;
  ld[LD_DECREASE] = decrease;
  ld[LD_X_COST] = x_cost;
  ld[LD_ITERATION] = iteration;
  ld[LD_TERMINATION] = termination;
  ld[LD_MSG] = msg;
  ld[LD_ACCEPT] = accept;
  ld[LD_NROWS] = nrows;
  scal[S_BAD_POINTS] = 0.0;  // the failure counter of the next k_points (the host loop's memset)
  return termination < 0;
}


}  // namespace mpba
