// C entry points over the reference's OWN mjpc sources (compiled in place from
// /root/reference/mujoco_mpc/mjpc/{norm.cc,spline/spline.cc}; see oracle/Makefile `ref`).
// Used to pin the oracle's restatement of spline + norm.  Test infrastructure only.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include <mujoco/mujoco.h>
#include "mjpc/norm.h"
#include "mjpc/spline/spline.h"

// the five MuJoCo utility symbols norm.cc needs (libmujoco is absent): plain definitions
extern "C" {
mjtNum mju_dot(const mjtNum* a, const mjtNum* b, int n) { mjtNum s = 0; for (int i = 0; i < n; i++) s += a[i] * b[i]; return s; }
void mju_error(const char* msg, ...) { std::fprintf(stderr, "mju_error: %s\n", msg); std::abort(); }
void mju_scl(mjtNum* r, const mjtNum* v, mjtNum s, int n) { for (int i = 0; i < n; i++) r[i] = v[i] * s; }
mjtNum mju_sign(mjtNum x) { return x < 0 ? -1 : (x > 0 ? 1 : 0); }
void mju_zero(mjtNum* r, int n) { for (int i = 0; i < n; i++) r[i] = 0; }

double ref_norm(const double* x, const double* params, int n, int type) {
  return mjpc::Norm(nullptr, nullptr, x, params, n, (mjpc::NormType)type);
}

// value + gradient (norm_test.cc checks gradients by finite differences)
double ref_norm_grad(const double* x, const double* params, int n, int type, double* g) {
  return mjpc::Norm(g, nullptr, x, params, n, (mjpc::NormType)type);
}

void ref_spline_sample(const double* times, const double* values, int nnode, int dim, int interpolation,
                       double time, double* out) {
  mjpc::spline::TimeSpline s(dim, (mjpc::spline::SplineInterpolation)interpolation);
  for (int i = 0; i < nnode; i++) s.AddNode(times[i], absl::MakeConstSpan(values + (size_t)i * dim, dim));
  s.Sample(time, absl::MakeSpan(out, dim));
}

// DiscardBefore semantics (spline_test.cc:161-231): returns remaining node count, writes remaining times
int ref_spline_discard_before(const double* times, const double* values, int nnode, int dim, int interpolation,
                              double time, double* out_times) {
  mjpc::spline::TimeSpline s(dim, (mjpc::spline::SplineInterpolation)interpolation);
  for (int i = 0; i < nnode; i++) s.AddNode(times[i], absl::MakeConstSpan(values + (size_t)i * dim, dim));
  s.DiscardBefore(time);
  int n = (int)s.Size();
  for (int i = 0; i < n; i++) out_times[i] = s.NodeAt(i).time();
  return n;
}
}
