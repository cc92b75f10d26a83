/* hmj_b200.h — C ABI of the B200-native batched humanoid step / rollout engine.
 *
 * Drop-in boundary for the two seams of mcgill-robotics/Humanoid-MuJoCo:
 *
 *  B1  Python env:  `mujoco.mj_step(self.model, self.data)` at simulation/cpu_env.py:684 plus the
 *      reads around it (cpu_env.py:418-440,465-616) -> hmj_batch_* (N lock-step environments).
 *  B2  mjpc planner: `SamplingPlanner::Rollouts` (mjpc/planners/sampling/planner.cc:342-380) ->
 *      `Trajectory::Rollout` (mjpc/trajectory.cc:100-210) -> hmj_plan_* (N samples x H knots).
 *
 * Plain C: opaque handle, int status codes (0 = ok, <0 = error, see hmj_last_error), caller-owned
 * buffers, no exceptions and no torch / C++ types across the boundary.  One host thread per handle.
 * Host pointers unless the name says `_dev`.  All floating-point data is fp64 (mjtNum = double,
 * simulation/mujoco/include/mujoco/mjtnum.h:27); hfield samples are fp32 (mjmodel.h:908).
 */
#ifndef HMJ_B200_H_
#define HMJ_B200_H_

#include <stddef.h>
#include <stdint.h>

#include "hmj_model.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct hmj_batch_ hmj_batch_t;

#define HMJ_OK 0
#define HMJ_ERR_CUDA (-1)
#define HMJ_ERR_ARG (-2)
#define HMJ_ERR_MODEL (-3)

/* stages for hmj_batch_run_stage (one kernel each) */
#define HMJ_STAGE_SMOOTH 1   /* kinematics .. qfrc_smooth            (mujoco.h:310-349,253) */
#define HMJ_STAGE_COLLIDE 2  /* mj_collision narrowphase             (mujoco.h:355) */
#define HMJ_STAGE_SOLVE 3    /* constraints + Newton, no integration (mujoco.h:358,367,259) */
#define HMJ_STAGE_SOLVE_INTEGRATE 4 /* ... + mj_Euler                (mujoco.h:262) */

const char* hmj_last_error(void);
int hmj_device_count(void);
int hmj_version(void);

/* measured FP64 FMA peak of the device in TFLOP/s (roofline denominator; MEASURED_PEAKS.json has none) */
int hmj_fp64_peak(int device, double* tflops);

/* ---- lifecycle.  `model` is the blob written by tools/compile_model.py (replaces
 * MjModel.from_xml_path, cpu_env.py:86 / mj_loadXML in mjpc/agent.cc). */
int hmj_batch_create(const hmj_model_t* model, int n_envs, int device, hmj_batch_t** out);
void hmj_batch_destroy(hmj_batch_t* b);
int hmj_batch_size(const hmj_batch_t* b);
int hmj_batch_sync(hmj_batch_t* b);

/* ---- per-env model parameters (the fields CPUEnv.reset randomises, cpu_env.py:189-280) */
int hmj_default_params(const hmj_model_t* model, hmj_params_t* out);
int hmj_batch_set_params(hmj_batch_t* b, const hmj_params_t* params, int first, int count);
int hmj_batch_get_params(hmj_batch_t* b, hmj_params_t* params, int first, int count);

/* ---- state (MjData.qpos/qvel/time/qacc_warmstart/ctrl/xfrc_applied). NULL = leave unchanged. */
int hmj_batch_set_state(hmj_batch_t* b, const double* qpos, const double* qvel, const double* time,
                        const double* qacc_warmstart);
int hmj_batch_get_state(hmj_batch_t* b, double* qpos, double* qvel, double* time, double* qacc_warmstart);
int hmj_batch_set_ctrl(hmj_batch_t* b, const double* ctrl);
/* one pushed body per env (cpu_env.py:618-654): body id (0 = none) and a 6-D wrench [force, torque] */
int hmj_batch_set_xfrc(hmj_batch_t* b, const int32_t* body, const double* wrench);

/* ---- stepping: mj_step x nsteps.  `ctrl` (host, [N][12]) may be NULL to keep the current controls.
 * hmj_batch_step copies ctrl H2D, runs, and synchronises; hmj_batch_step_dev takes a DEVICE pointer
 * (or NULL) and only enqueues on the batch stream. */
int hmj_batch_step(hmj_batch_t* b, const double* ctrl, int nsteps);
int hmj_batch_step_dev(hmj_batch_t* b, const double* ctrl_dev, int nsteps);
int hmj_batch_forward(hmj_batch_t* b); /* mj_forward */
int hmj_batch_run_stage(hmj_batch_t* b, int stage);

/* ---- results */
int hmj_batch_get_contacts(hmj_batch_t* b, int32_t* ncon, hmj_contact_t* contacts /* [N][HMJ_MAXCON] */);
/* named array download / upload (row-major [N][...]); names: qpos qvel time warmstart ctrl xfrc
 * geom_xpos geom_xmat cdof com qM qfrc_smooth qacc_smooth qacc qfrc_constraint sens residual
 * cost_sum trace (double)  ncon nefc niter status xfrc_body (int32) */
int hmj_batch_get_array(hmj_batch_t* b, const char* name, void* out, size_t bytes);
int hmj_batch_set_array(hmj_batch_t* b, const char* name, const void* in, size_t bytes);
void* hmj_batch_device_ptr(hmj_batch_t* b, const char* name);
void* hmj_batch_stream(hmj_batch_t* b);
int hmj_batch_kernel_launches(const hmj_batch_t* b); /* kernels launched since create */

/* Support-function candidate tables of the model's convex hulls (the engine's counterpart of MuJoCo's mesh_graph
 * hill-climb, simulation/mujoco/include/mujoco/mjmodel.h:867,876): directions are binned into 6 x G x G cube-map cells and
 * every cell lists the hull vertices that can be the support point for a direction in it, so a support call scans a
 * handful of vertices instead of the hull and still returns the exhaustive scan's vertex.  Host-only (no GPU needed);
 * start is [7][6*G*G + 2] offsets into vert, capacities are checked. */
int hmj_support_cells(const hmj_model_t* model, uint16_t* start, int start_cap, uint16_t* vert, int vert_cap, int* grid,
                      int* entries);

/* ---- CPUEnv.step epilogue on device (cpu_env.py:465-616 + reward_functions.py:247-374):
 * obs [N][30], reward [N], terminated/truncated/self_collision [N] (int32).  Reads the one-step-stale
 * xquat/xpos/qfrc/contacts exactly like the reference.
 * prev_action/latest_action are [N][12] (already divided by nothing; the kernel applies /(pi/2)). */
int hmj_env_obs_reward(hmj_batch_t* b, const double* prev_action, const double* latest_action,
                       const double* noise /* [N][31]: q12 qd12 gyro3 quat4, NULL = 0 */,
                       double max_sim_time, double* obs, double* reward, int32_t* terminated,
                       int32_t* truncated, int32_t* self_collision);

/* hmj_env_obs_reward without any allocation (no counterpart in the reference: added
 * for callers that cannot call cudaMalloc — graph capture, serving loops, harnesses that forbid it).  Same
 * arguments and results; the device staging block is a static array of the library instead of an allocation, so the
 * batch may hold at most hmj_env_obs_reward_noalloc_max_envs() = 4096 envs (larger: HMJ_ERR_ARG, see hmj_last_error).
 * The array is shared by all batches of a device: calls that overlap on the GPU race; the call returns synchronised. */
int hmj_env_obs_reward_noalloc(hmj_batch_t* b, const double* prev_action, const double* latest_action,
                               const double* noise, double max_sim_time, double* obs, double* reward,
                               int32_t* terminated, int32_t* truncated, int32_t* self_collision);
int hmj_env_obs_reward_noalloc_max_envs(void);

/* same, all pointers DEVICE pointers, enqueue only on the batch stream */
int hmj_env_obs_reward_dev(hmj_batch_t* b, const double* prev_action_dev, const double* latest_action_dev,
                           const double* noise_dev, double max_sim_time, double* obs_dev, double* reward_dev,
                           int32_t* terminated_dev, int32_t* truncated_dev, int32_t* self_collision_dev);

/* CPUEnv.step (cpu_env.py:676-693) for all N envs in ONE call: ctrl / push wrench H2D (NULL = keep), nsteps x mj_step
 * (PHYSICS_STEPS_PER_CONTROL_STEP, simulation_parameters.py:82), observation + reward epilogue, results D2H, one sync.
 * time_out [N] = data.time, root_quat_out [N][4] = data.xquat[1] (one step stale); either may be NULL. */
int hmj_env_step(hmj_batch_t* b, const double* ctrl, const int32_t* xfrc_body, const double* xfrc,
                 const double* prev_action, const double* latest_action, const double* noise, double max_sim_time,
                 int nsteps, double* obs, double* reward, int32_t* terminated, int32_t* truncated,
                 int32_t* self_collision, double* time_out, double* root_quat_out);

/* ---- CPUEnv.step with the per-step bookkeeping ON THE DEVICE (cpu_env.py:656-674 action noise + delay FIFO,
 * :618-654 push scheduler, :528-547 observation delay FIFOs, :516 torso_quat).  The per-env bookkeeping arrays are
 * named device arrays ("env_act_noise" [N][12], "env_obs_noise" [N][31], "env_act_delay" [N] i32, "env_obs_delay"
 * [N][3] i32, "env_act_count" / "env_obs_count" [N] i32, "env_prev_action" / "env_latest_action" [N][12],
 * "env_force_sched" [N][5], "env_force_body" / "env_force_draws" [N] i32), written by the host at reset through
 * hmj_batch_set_array.  hmj_env_io returns views into ONE pinned host block: inputs action [N][12], mask [N] i32;
 * outputs obs [N][30], reward [N], time [N], torso_quat [N][4], terminated / truncated / self_collision [N] i32.
 * hmj_env_step_fused: one H2D copy, k_env_pre, nsteps x mj_step, obs/reward, delay FIFOs, one D2H copy, one sync.
 * has_action = 0 replays data.ctrl (CPUEnv.step(None)); has_mask != 0 limits the bookkeeping to mask[env] != 0.
 * Pushes draw from a counter-based stream splitmix64(seed, env, draw) instead of the reference's unseeded `random`. */
int hmj_env_io(hmj_batch_t* b, double** action, int32_t** mask, double** obs, double** reward, double** time,
               double** torso_quat, int32_t** terminated, int32_t** truncated, int32_t** self_collision);
int hmj_env_step_fused(hmj_batch_t* b, int has_action, int has_mask, double rf, double max_sim_time,
                       unsigned long long seed, int nsteps);
/* The same step with the controls read straight from the caller's host array (action_host [N][12], NULL = replay
 * data.ctrl) instead of the IO block: no staging copy on the host.  Pinned memory (hmj_pinned_alloc, or any
 * cudaHostRegister'ed / framework-pinned buffer) makes the H2D copy a plain DMA; pageable memory works and is staged
 * by the CUDA driver.  The array must stay valid until the call returns (it synchronises). */
int hmj_env_step_fused_from(hmj_batch_t* b, const double* action_host, int has_mask, double rf, double max_sim_time,
                            unsigned long long seed, int nsteps);
/* ---- CPUEnv.reset (cpu_env.py:374-416) and DummyVecEnv's auto-reset (rl/train.py:134-151) ON THE DEVICE.
 * A batch owns S spare rows behind its N envs (hmj_batch_spare_rows; N/8 clamped to [32, 1024]).  Fresh episodes —
 * domain randomisation of the model (:189-280), delays (:135-187), constant noise vectors, start pose (:282-328), the
 * probe step with action None and the reference's rejection rule (:411-414: first-step contact / terminated /
 * truncated -> draw again) — are generated into the spare rows in the background on a side stream, from the
 * counter-based stream (seed, row, generation).  hmj_env_pool_config(enable=1, ...) makes hmj_env_step_fused[_from]
 * replace every env whose episode ended (terminated | truncated) with a ready episode inside the same call: its obs /
 * time / torso_quat rows then hold the FIRST observation of the new episode (reward and flags stay the ended
 * episode's), and the terminal observations come back through hmj_env_reset_io.  No bystander env is stepped or
 * copied.  hmj_env_reset resets the listed envs now (ids == NULL: all) and blocks until every one is served. */
int hmj_env_pool_config(hmj_batch_t* b, int enable, double rf, double max_sim_time, unsigned long long seed, double timestep);
int hmj_env_reset(hmj_batch_t* b, const int32_t* ids, int count);
int hmj_env_reset_io(hmj_batch_t* b, int32_t** counts, int32_t** ended_ids, double** term_obs);
/* the envs that ended in the last pooled control step (env order) + their terminal observations [count][30]; pinned
 * host memory owned by the batch, valid until the next step */
int hmj_env_terminal(hmj_batch_t* b, int32_t** ids, double** obs, int32_t* count);
int hmj_batch_spare_rows(const hmj_batch_t* b);
/* host-side counters: generation rounds launched, blocking drains, drain iterations, pooled control steps */
int hmj_env_pool_stats(const hmj_batch_t* b, long long* out4);
/* Page-locked host memory for step inputs (cudaHostAlloc / cudaFreeHost). */
int hmj_pinned_alloc(size_t bytes, void** out);
int hmj_pinned_free(void* p);

/* ---- per-kernel device timing (CUDA events on the batch stream, one pair around every launch of a step;
 * MuJoCo's own equivalent is the mjTIMER_* profiler, mjdata.h:68-93).  enable resets the counters. */
int hmj_batch_profile(hmj_batch_t* b, int enable);
int hmj_batch_profile_read(hmj_batch_t* b, double* ms_smooth, double* ms_collide, double* ms_solve, int* nsteps);

/* ---- planner rollouts (SamplingPlanner::Rollouts): every env is one sample.
 * state [37] = qpos,qvel shared by all samples (State::CopyTo, mjpc/states/state.cc:128-135);
 * spline_times [P], spline_values [N][P][12] = candidate policies (noise already added and clamped
 * by the caller: AddNoiseToPolicy planner.cc:313-339 is entropy-seeded in the reference, so noise is
 * an explicit input); interpolation 0/1/2 = zero/linear/cubic (spline.h).  horizon = number of knots
 * (H-1 steps + 1 forward).  Outputs: total_return [N] (1e6 on failure), failure [N] (int32).
 * If record != 0 the full trajectories are kept on the device for hmj_plan_fetch. */
int hmj_plan_set_task(hmj_batch_t* b, int nterm, const int32_t* dim, const int32_t* norm, const double* weight,
                      const double* normprm /* [nterm][3] */, double risk, double height_goal);
int hmj_plan_rollouts(hmj_batch_t* b, const double* state, double time, const double* spline_times,
                      const double* spline_values, int nnode, int interpolation, int horizon, int record,
                      double* total_return, int32_t* failure);
/* Trajectory::NoisyRollout (mjpc/trajectory.cc:100-210 with xfrc_std > 0) for every sample: the same rollouts with an
 * Ornstein-Uhlenbeck wrench on EVERY body, refreshed before each step: x <- rate x + N(0, scale), rate = exp(-dt / xfrc_rate),
 * scale = xfrc_std sqrt(1 - rate^2) (:147-155).  This is the robustness pass of RobustPlanner::OptimizePolicy
 * (mjpc/planners/robust/robust_planner.cc:96-151).  xfrc_noise [N][H-1][90]: unit normals injected by the caller (NULL:
 * drawn on the device from the counter-based stream (seed, call counter, first_global + sample)). */
int hmj_plan_rollouts_noisy(hmj_batch_t* b, const double* state, double time, const double* spline_times,
                            const double* spline_values, int nnode, int interpolation, int horizon, double xfrc_std,
                            double xfrc_rate, const double* xfrc_noise, unsigned long long seed, int first_global,
                            double* total_return, int32_t* failure);
/* device-resident variant: spline_values_dev [N][P][12], returns_dev [N]; enqueue only */
int hmj_plan_rollouts_dev(hmj_batch_t* b, const double* state, double time, const double* spline_times,
                          const double* spline_values_dev, int nnode, int interpolation, int horizon,
                          double* returns_dev);
/* local argmin of the last rollouts (lowest index wins ties): device result {cost, index}; also copied to host */
int hmj_plan_argmin(hmj_batch_t* b, double* best_cost, int32_t* best_index);
/* recorded trajectory of one sample (Trajectory buffers, mjpc/trajectory.h:74-86) */
int hmj_plan_fetch(hmj_batch_t* b, int sample, int horizon, double* states /* [H][37] */, double* actions /* [H][12] */,
                   double* times /* [H] */, double* residual /* [H][28] */, double* costs /* [H] */, double* trace /* [H][3] */);

/* ---- one predictive-sampling iteration, candidates built on the device (SamplingPlanner::OptimizePolicy's device half:
 * AddNoiseToPolicy planner.cc:313-339, Rollouts :342-380, the arg-min of :151-187).  nominal_values [P][12] is the
 * resampled nominal spline (UpdateNominalPolicy stays on the host: P x 12 numbers).  noise [N][P][12] is added as given
 * (NULL: drawn on the device from the counter-based stream (seed, iteration, global sample); std = noise_std, or
 * noise_std2 with probability 0.2 when noise_std2 > 0, scaled by half the control range).  The sample with GLOBAL index
 * first_global + s == 0 keeps the nominal (planner.cc:361).  The whole batch of rollouts replays as ONE CUDA graph.
 * Enqueue only: the result is this rank's winner record {cost, global index, return of local sample 0, winner spline
 * values [P][12]} = hmj_plan_record_size(P) doubles at record_dev (NULL: hmj_plan_record_ptr(b)).
 * Multi-GPU: all-gather the records of all ranks ON hmj_batch_stream(b) (ncclAllGather, or the built-in
 * hmj_comm_allgather below), then hmj_plan_select picks the global winner on the device (lexicographic (cost, global
 * index); NaN never wins) and returns its record to the host — one synchronisation per iteration, no host round trip
 * between the collective and the selection.  gathered_dev == NULL / nranks <= 1: single-GPU, selects the local record. */
int hmj_plan_iterate(hmj_batch_t* b, const double* state, double time, const double* spline_times,
                     const double* nominal_values, int nnode, int interpolation, int horizon, const double* noise,
                     double noise_std, double noise_std2, unsigned long long seed, unsigned long long iteration,
                     int first_global, double* record_dev);
/* The same with a per-parameter ABSOLUTE noise standard deviation noise_std_vec [P][12] (NULL: hmj_plan_iterate): the
 * cross-entropy planner's sampling distribution, max(sqrt(variance), std_min) per spline parameter without control-range
 * scaling (mjpc/planners/cross_entropy/planner.cc:346-378).  Elite selection and the mean / variance refit (:213-283) are
 * done by the caller from hmj_plan_population. */
int hmj_plan_iterate_ex(hmj_batch_t* b, const double* state, double time, const double* spline_times,
                        const double* nominal_values, int nnode, int interpolation, int horizon, const double* noise,
                        double noise_std, double noise_std2, const double* noise_std_vec, unsigned long long seed,
                        unsigned long long iteration, int first_global, double* record_dev);
int hmj_plan_record_size(int nnode);
void* hmj_plan_record_ptr(hmj_batch_t* b);
int hmj_plan_select(hmj_batch_t* b, const double* gathered_dev, int nranks, int nnode, double* record_host);
/* ---- mj_inverse for every env: qfrc_inverse = M qacc + qfrc_bias - qfrc_passive - qfrc_constraint with the constraint
 * forces of the GIVEN acceleration (no solver).  Replaces the 200 mj_inverse calls of find_ideal_distance_to_ground
 * (controllers/mpc_utils.py:36-55) by one batched call.  qacc [N][18] (NULL: zeros), qfrc_inverse [N][18] out. */
int hmj_batch_inverse(hmj_batch_t* b, const double* qacc, double* qfrc_inverse);
/* ---- batched finite-difference transition matrices: mjd_transitionFD for T knots in ONE batched step, as mjpc's
 * ModelDerivatives::Compute needs them for iLQG / gradient planners (mjpc/planners/model_derivatives.cc:51-96).  The batch
 * must hold >= 97 T environments (1 + 2 x 48 nudged copies per knot); their state is overwritten, their parameters are
 * the caller's.  x [T][37], u [T][12], time [T], warmstart [T][18] or NULL.  Row-major outputs per knot (NULL = skip):
 * A [T][36][36] = dy/dx, B [T][36][12] = dy/du (y: next state in tangent coordinates), C [T][28][36], D [T][28][12]
 * (s: the 28 task residuals of that step's forward pass). */
int hmj_fd_transition(hmj_batch_t* b, int T, const double* x, const double* u, const double* time, const double* warmstart,
                      double eps, int centered, double* A, double* B, double* C, double* D);
/* ---- built-in winner exchange for hosts without a communicator of their own (a C++ SamplingPlanner): NCCL is
 * dlopen'ed at run time (libnccl.so.2; HMJ_NCCL_LIB overrides).  Rank 0 calls hmj_comm_unique_id (128 bytes) and hands
 * the id to the other processes by any means; every process calls hmj_comm_create with its rank and device.
 * hmj_plan_exchange = ncclAllGather of the winner records on hmj_batch_stream + hmj_plan_select: the planner's one
 * collective per iteration (the parallel form of planner.cc:151-187 + CopyCandidateToPolicy :525-534). */
typedef struct hmj_comm_ hmj_comm_t;
int hmj_comm_unique_id(void* id128);
int hmj_comm_create(const void* id128, int rank, int nranks, int device, hmj_comm_t** out);
void hmj_comm_destroy(hmj_comm_t* c);
int hmj_comm_rank(const hmj_comm_t* c);
int hmj_comm_size(const hmj_comm_t* c);
int hmj_plan_exchange(hmj_batch_t* b, hmj_comm_t* c, int nnode, double* record_host);
/* returns [N] / failure flags [N] / candidate values [N][P][12] of the last hmj_plan_iterate (any may be NULL) */
int hmj_plan_population(hmj_batch_t* b, int nnode, double* total_return, int32_t* failure, double* values);
/* API-level launches so far (a CUDA-graph replay counts once; hmj_batch_kernel_launches counts kernel nodes) */
int hmj_batch_host_launches(const hmj_batch_t* b);

#ifdef __cplusplus
}
#endif
#endif /* HMJ_B200_H_ */
