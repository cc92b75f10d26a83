// rollouts_shim.cc — the C++ side of seam B2, compiled (g++, no CUDA headers) against include/hmj_b200.h and linked to
// libhmj_b200.so.  It is the code INTEGRATION.md §2 tells a maintainer to put into mjpc's SamplingPlanner
// (mjpc/planners/sampling/planner.cc:342-380 Rollouts, :190-208 OptimizePolicy, :525-534 CopyCandidateToPolicy),
// carried by a stand-in class with the members those bodies touch (mjpc itself needs MuJoCo + abseil, absent here).
// tests/test_gpu_cpp_shim.py drives it through the extern "C" functions at the bottom and checks that a C++ host gets
// the same bits as the Python host.
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "hmj_b200.h"

namespace shim {

constexpr int kNu = 12;
constexpr int kNState = 37;

struct Trajectory {          // the two fields Rollouts() writes (mjpc/trajectory.h:74-86)
  double total_return = 0.0;
  bool failure = false;
};

class SamplingPlanner {
 public:
  SamplingPlanner(const void* model_blob, int num_trajectory, int device, int num_spline_points, int interpolation)
      : num_trajectory_(num_trajectory), num_spline_points_(num_spline_points), interpolation_(interpolation) {
    if (hmj_batch_create(static_cast<const hmj_model_t*>(model_blob), num_trajectory, device, &batch_) != HMJ_OK)
      throw std::runtime_error(std::string("hmj_batch_create: ") + hmj_last_error());   // planner.cc:69-72 uses mju_error
    trajectory.resize(num_trajectory);
    values_.resize(static_cast<size_t>(num_trajectory) * num_spline_points * kNu);
    returns_.resize(num_trajectory);
    failure_.resize(num_trajectory);
    node_times_.resize(num_spline_points);
    state.assign(kNState, 0.0);
    winner_values_.resize(static_cast<size_t>(num_spline_points) * kNu);
  }
  ~SamplingPlanner() { hmj_batch_destroy(batch_); }

  // planner timestep (Agent::PlanIteration writes model->opt.timestep, mjpc/agent.cc:275)
  void SetTimestep(double dt) {
    std::vector<hmj_params_t> p(num_trajectory_);
    Check(hmj_batch_get_params(batch_, p.data(), 0, num_trajectory_), "get_params");
    for (auto& q : p) q.timestep = dt;
    Check(hmj_batch_set_params(batch_, p.data(), 0, num_trajectory_), "set_params");
  }

  // SamplingPlanner::Rollouts with the pool.Schedule fan-out (planner.cc:350-378) replaced by one C call.
  // candidate_values: [N][P][nu], nominal in sample 0, noise already added and clamped (AddNoiseToPolicy :313-339).
  void Rollouts(int num_trajectory, int horizon, const double* candidate_values, int record) {
    std::copy(candidate_values, candidate_values + values_.size(), values_.begin());
    if (hmj_plan_rollouts(batch_, state.data(), time, node_times_.data(), values_.data(), num_spline_points_, interpolation_,
                          horizon, record, returns_.data(), failure_.data()) != HMJ_OK)
      throw std::runtime_error(std::string("hmj_plan_rollouts: ") + hmj_last_error());
    for (int i = 0; i < num_trajectory; i++) {
      trajectory[i].total_return = returns_[i];      // 1e6 on failure (trajectory.cc:183-186)
      trajectory[i].failure = failure_[i] != 0;
    }
  }

  // OptimizePolicy's device half + the winner exchange: candidates built on the device from the nominal spline,
  // rollouts, arg-min, exchange over `comm` (NULL: one process), CopyCandidateToPolicy from the returned record.
  void OptimizePolicyCandidates(int horizon, const double* nominal_values, const double* noise, double noise_std,
                                unsigned long long seed, unsigned long long iteration, int first_global, hmj_comm_t* comm) {
    Check(hmj_plan_iterate(batch_, state.data(), time, node_times_.data(), nominal_values, num_spline_points_, interpolation_,
                           horizon, noise, noise_std, 0.0, seed, iteration, first_global, nullptr), "hmj_plan_iterate");
    std::vector<double> rec(hmj_plan_record_size(num_spline_points_));
    Check(hmj_plan_exchange(batch_, comm, num_spline_points_, rec.data()), "hmj_plan_exchange");
    winner_return = rec[0];
    winner = static_cast<int>(rec[1]);
    nominal_return = rec[2];
    std::copy(rec.begin() + 3, rec.end(), winner_values_.begin());   // CopyCandidateToPolicy (planner.cc:525-534)
  }

  std::vector<double> state;
  double time = 0.0;
  std::vector<Trajectory> trajectory;
  std::vector<double> node_times_, winner_values_;
  int winner = -1;
  double winner_return = 0.0, nominal_return = 0.0;
  hmj_batch_t* batch_ = nullptr;

 private:
  static void Check(int rc, const char* what) {
    if (rc != HMJ_OK) throw std::runtime_error(std::string(what) + ": " + hmj_last_error());
  }
  int num_trajectory_, num_spline_points_, interpolation_;
  std::vector<double> values_, returns_;
  std::vector<int32_t> failure_;
};

}  // namespace shim

// ---------------------------------------------------------------- flat driver for the pytest (ctypes)
static thread_local std::string g_shim_err;
#define SHIM_TRY(body) try { body; return 0; } catch (const std::exception& e) { g_shim_err = e.what(); return -1; }

extern "C" {
const char* shim_last_error() { return g_shim_err.c_str(); }

void* shim_create(const void* model_blob, int n, int device, int npoints, int interpolation, double timestep) {
  try {
    auto* p = new shim::SamplingPlanner(model_blob, n, device, npoints, interpolation);
    p->SetTimestep(timestep);
    return p;
  } catch (const std::exception& e) {
    g_shim_err = e.what();
    return nullptr;
  }
}
void shim_destroy(void* h) { delete static_cast<shim::SamplingPlanner*>(h); }

int shim_set_state(void* h, const double* state37, double time, const double* node_times, int npoints) {
  auto* p = static_cast<shim::SamplingPlanner*>(h);
  SHIM_TRY(p->state.assign(state37, state37 + shim::kNState); p->time = time; p->node_times_.assign(node_times, node_times + npoints));
}
int shim_rollouts(void* h, int n, int horizon, const double* values, int record, double* returns, int32_t* failure) {
  auto* p = static_cast<shim::SamplingPlanner*>(h);
  SHIM_TRY(p->Rollouts(n, horizon, values, record);
           for (int i = 0; i < n; i++) { returns[i] = p->trajectory[i].total_return; failure[i] = p->trajectory[i].failure; });
}
int shim_optimize(void* h, int horizon, const double* nominal, const double* noise, double noise_std, unsigned long long seed,
                  unsigned long long iteration, int first_global, void* comm, double* out /* winner, return, nominal, values */) {
  auto* p = static_cast<shim::SamplingPlanner*>(h);
  SHIM_TRY(p->OptimizePolicyCandidates(horizon, nominal, noise, noise_std, seed, iteration, first_global,
                                       static_cast<hmj_comm_t*>(comm));
           out[0] = p->winner; out[1] = p->winner_return; out[2] = p->nominal_return;
           std::copy(p->winner_values_.begin(), p->winner_values_.end(), out + 3));
}
}
