/* hmj_oracle_batch.c — multi-threaded drivers of the oracle, used ONLY to time the CPU baseline
 * (bench.py `cpu_baseline` / `--impl reference`) and to generate large batches of oracle outputs
 * for parity tests.  TEST INFRASTRUCTURE ONLY (see hmj_oracle.h).
 *
 * Methodology follows the vendored MuJoCo sample simulation/mujoco/sample/testspeed.cc:84-103,203-220:
 * one thread per host core, private per-thread data, disjoint slices of environments / samples,
 * wall clock over all threads.  The planner variant replaces mjpc's ThreadPool fan-out
 * (mjpc/planners/sampling/planner.cc:342-380, mjpc/threadpool.cc:51-85).
 */
#define _POSIX_C_SOURCE 200809L
#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "hmj_oracle_internal.h"

static double now_s(void) {
  struct timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

typedef struct {
  const hmj_model_t* m;
  const hmj_params_t* params;
  orc_data_t* data;
  const double* ctrl; /* [ctrl_steps][n_envs][12] or NULL */
  int n_envs, ctrl_steps, nsteps;
  int* next; /* shared work counter: environments are claimed one at a time (dynamic balance) */
} step_job_t;

static void* step_worker(void* arg) {
  step_job_t* j = (step_job_t*)arg;
  for (;;) {
    const int e = __atomic_fetch_add(j->next, 1, __ATOMIC_RELAXED);
    if (e >= j->n_envs) break;
    for (int s = 0; s < j->nsteps; s++) {
      if (j->ctrl) {
        const double* c = j->ctrl + ((size_t)(s % j->ctrl_steps) * j->n_envs + e) * HMJ_NU;
        memcpy(j->data[e].ctrl, c, HMJ_NU * sizeof(double));
      }
      orc_step(j->m, j->params + e, j->data + e);
    }
  }
  return NULL;
}

/* n_envs environments x nsteps mj_step on nthreads threads; returns wall seconds */
double orc_step_batch(const hmj_model_t* m, const hmj_params_t* params, orc_data_t* data, int n_envs,
                      const double* ctrl, int ctrl_steps, int nsteps, int nthreads) {
  if (nthreads < 1) nthreads = 1;
  if (nthreads > n_envs) nthreads = n_envs;
  pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * (size_t)nthreads);
  step_job_t* jobs = (step_job_t*)malloc(sizeof(step_job_t) * (size_t)nthreads);
  int next = 0;
  const double t0 = now_s();
  for (int t = 0; t < nthreads; t++) {
    step_job_t j = {m, params, data, ctrl, n_envs, ctrl_steps > 0 ? ctrl_steps : 1, nsteps, &next};
    jobs[t] = j;
    pthread_create(&th[t], NULL, step_worker, &jobs[t]);
  }
  for (int t = 0; t < nthreads; t++) pthread_join(th[t], NULL);
  const double t1 = now_s();
  free(th);
  free(jobs);
  return t1 - t0;
}

typedef struct {
  const hmj_model_t* m;
  const hmj_params_t* p;
  const double* state;
  double time;
  const double* spline_times;
  const double* spline_values; /* [n][nnode][12] */
  int nnode, interpolation, horizon, n;
  int* next;
  double* total_return;
  int32_t* failure;
} roll_job_t;

static void* roll_worker(void* arg) {
  roll_job_t* j = (roll_job_t*)arg;
  const int H = j->horizon;
  double* states = (double*)malloc(sizeof(double) * (size_t)H * (HMJ_NSTATE + HMJ_NU + 1 + HMJ_NRESIDUAL + 1 + 3));
  double* actions = states + (size_t)H * HMJ_NSTATE;
  double* times = actions + (size_t)H * HMJ_NU;
  double* residual = times + H;
  double* costs = residual + (size_t)H * HMJ_NRESIDUAL;
  double* trace = costs + H;
  for (;;) {
    const int i = __atomic_fetch_add(j->next, 1, __ATOMIC_RELAXED);
    if (i >= j->n) break;
    double tot = 0;
    int f = orc_rollout(j->m, j->p, j->state, j->time, j->spline_times,
                        j->spline_values + (size_t)i * j->nnode * HMJ_NU, j->nnode, j->interpolation, H, states,
                        actions, times, residual, costs, trace, &tot);
    j->total_return[i] = tot;
    if (j->failure) j->failure[i] = f;
  }
  free(states);
  return NULL;
}

/* SamplingPlanner::Rollouts over n candidate splines on nthreads threads; returns wall seconds */
double orc_rollout_batch(const hmj_model_t* m, const hmj_params_t* p, const double* state, double time,
                         const double* spline_times, const double* spline_values, int n, int nnode,
                         int interpolation, int horizon, double* total_return, int32_t* failure, int nthreads) {
  if (nthreads < 1) nthreads = 1;
  if (nthreads > n) nthreads = n;
  pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * (size_t)nthreads);
  roll_job_t* jobs = (roll_job_t*)malloc(sizeof(roll_job_t) * (size_t)nthreads);
  int next = 0;
  const double t0 = now_s();
  for (int t = 0; t < nthreads; t++) {
    roll_job_t j = {m, p, state, time, spline_times, spline_values, nnode, interpolation, horizon, n, &next, total_return, failure};
    jobs[t] = j;
    pthread_create(&th[t], NULL, roll_worker, &jobs[t]);
  }
  for (int t = 0; t < nthreads; t++) pthread_join(th[t], NULL);
  const double t1 = now_s();
  free(th);
  free(jobs);
  return t1 - t0;
}
