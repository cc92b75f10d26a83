// hmj_api.cu — host side of the C ABI declared in include/hmj_b200.h.
// Owns device buffers, derived lookup tables and one CUDA stream per batch; every entry point
// returns an int status and never throws.  The product path: there is no CPU fallback — if CUDA
// is unavailable every call fails with HMJ_ERR_CUDA.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <tuple>
#include <vector>

#include "../../include/hmj_b200.h"
#include "hmj_dev.cuh"

static thread_local char g_err[512] = "";

static int fail(int code, const char* fmt, const char* detail = "") {
  snprintf(g_err, sizeof(g_err), fmt, detail);
  return code;
}

#define CK(call)                                                                    \
  do {                                                                              \
    cudaError_t e_ = (call);                                                        \
    if (e_ != cudaSuccess) return fail(HMJ_ERR_CUDA, "CUDA error: %s", cudaGetErrorString(e_)); \
  } while (0)

struct ArrayInfo { void* ptr; size_t bytes; };

struct hmj_batch_ {
  int n = 0;
  int device = 0;
  cudaStream_t stream = nullptr;
  hmj_tables_t* tab_dev = nullptr;
  hmj_task_t* task_dev = nullptr;
  hmj_tables_t* tab_host = nullptr;
  hmj_buffers_t buf{};
  std::vector<void*> allocs;
  std::map<std::string, ArrayInfo> arrays;
  // planner scratch
  double* state_dev = nullptr;       // [37]
  double* spline_times_dev = nullptr;// [64]
  double* spline_values_dev = nullptr; size_t spline_values_cap = 0;
  double* returns_dev = nullptr;     // [N]
  int32_t* failure_dev = nullptr;    // [N]
  double* argmin_dev = nullptr;      // [2]
  double* nominal_dev = nullptr;     // [64][12] nominal spline values of hmj_plan_iterate
  double* stdvec_dev = nullptr;      // [64][12] per-parameter noise std (cross-entropy planner)
  unsigned long long* seed_iter_dev = nullptr;   // [2] seed, call counter of the OU force noise
  double* ou_noise_dev = nullptr; size_t ou_noise_cap = 0;   // injected unit normals [N][H-1][90]
  unsigned long long ou_calls = 0;
  double* noise_dev = nullptr;       // [N][nnode][12] caller-provided noise (same capacity as spline_values_dev)
  double* record_dev = nullptr;      // [3 + 64*12] this rank's winner record
  double* select_dev = nullptr;      // [3 + 64*12] global winner record
  double* plan_stage = nullptr;      // pinned [3072]: state|time @0, spline times @64, nominal @128, record out @1024, std vector @2048
  int plan_nnode = 0;
  int host_launches = 0;             // API-level launches (a graph replay counts once)
  struct RolloutKey {
    int nnode, interp, horizon, record;
    const void* values; void* returns; void* failure; void* rec;
    bool operator<(const RolloutKey& o) const {
      return std::tie(nnode, interp, horizon, record, values, returns, failure, rec) <
             std::tie(o.nnode, o.interp, o.horizon, o.record, o.values, o.returns, o.failure, o.rec);
    }
  };
  std::map<RolloutKey, cudaGraphExec_t> rollout_graphs;
  hmj_record_t rec{};
  int rec_cap = 0;
  // env scratch
  double* env_stage = nullptr; size_t env_stage_cap = 0;
  // device-resident CPUEnv bookkeeping + one pinned host / device IO block for the fused control step
  hmj_envstate_t env{};
  char* io_host = nullptr;   // pinned: [in: action N*12 f64 | mask N i32][out: obs N*30, reward N, time N, quat N*4 f64 | term, trunc, selfc N i32]
  char* io_dev = nullptr;
  size_t io_in_bytes = 0, io_out_bytes = 0;
  int launches = 0;
  // episode pool (hmj_reset.cu): S spare rows behind the N envs, generated on their own stream
  int n_spare = 0;
  hmj_buffers_t spare{};             // view of rows [N, N + S)
  hmj_envstate_t env_spare{};
  hmj_pool_t pool{};
  cudaStream_t spare_stream = nullptr;
  hmj_solve_ctx_t spare_ctx{};
  cudaEvent_t ev_main = nullptr, round_done = nullptr;
  cudaGraphExec_t round_graph = nullptr;
  bool pool_on = false, round_inflight = false, pool_need_round = true;
  long long step_idx = 0, round_start = 0;
  double pool_rf = 0, pool_max_sim_time = 0, pool_timestep = 0.002;
  unsigned long long pool_seed = 0;
  long long pool_rounds = 0, pool_drains = 0, pool_drain_iters = 0;
  double* fd_dev = nullptr; size_t fd_cap = 0;   // scratch of hmj_fd_transition
  char* term_host = nullptr;         // pinned: terminal observations [N][30] f64 | env ids [N] i32 of a step with more than 64 endings
  size_t io_ext_offset = 0;          // byte offset of the pool's result block inside the IO out block
  hmj_solve_ctx_t solve_ctx{};       // side streams + events that overlap the solver classes
  // Pipes: the physics chain of a step (smooth -> collide -> solve) is enqueued once per contiguous block of envs, each on
  // its own stream, forked from / joined to the main stream inside the call.  Every stage is a single wave whose length is
  // its slowest env, so stages of different pipes fill each other's tails (envs are independent: no data-path exchange).
  struct Pipe {
    int start = 0;                   // first env of the pipe
    hmj_buffers_t view{};
    cudaStream_t stream = nullptr;
    cudaEvent_t done = nullptr;
    hmj_solve_ctx_t ctx{};
  };
  std::vector<Pipe> pipes;
  cudaEvent_t pipe_fork = nullptr;
  // step graphs: [0] forward only, [1] with integration
  cudaGraphExec_t step_graph[2] = {nullptr, nullptr};
  // control-step graph of hmj_env_step_fused_from (prologue, physics, epilogue and result copies inside the pipes)
  struct CtrlKey {
    int has_action = -1, split = 0;
    double rf = 0, mst = 0;
    unsigned long long seed = 0;
    bool operator==(const CtrlKey& o) const {
      return has_action == o.has_action && split == o.split && rf == o.rf && mst == o.mst && seed == o.seed;
    }
  };
  cudaGraphExec_t ctrl_graph = nullptr;
  CtrlKey ctrl_key{};
  // the action upload of a control step runs on a stream of its own; the graph's prologue kernels wait for this event
  // (an event-wait node), everything that does not need the actions (k_kin, the collision kernels) starts at once
  cudaStream_t copy_stream = nullptr;
  cudaEvent_t h2d_event = nullptr;
  bool ctrl_graph_waits_h2d = false;
  int ctrl_mode = 2;                 // 2: control-step graph, action upload on the copy stream beside k_kin; 1: upload in front of the graph;
                                     // 0: prologue / epilogue as whole-batch kernels around the step graph (HMJ_E2E)
  int graph_mode = 1;
  bool warmed = false;
  // per-kernel CUDA-event profiling (hmj_batch_profile): events bracket every launch of a step
  bool profile = false;
  std::vector<cudaEvent_t> prof_events;  // 4 per step: before smooth, collide, solve, after solve
  size_t prof_used = 0;
};

static int pool_step(hmj_batch_t* b, const int32_t* term, const int32_t* trunc, bool drain);
static int pool_drain(hmj_batch_t* b);
static int32_t* io_counts_host(hmj_batch_t* b);

extern "C" const char* hmj_last_error(void) { return g_err; }
extern "C" int hmj_version(void) { return HMJ_VERSION; }
extern "C" int hmj_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
  return n;
}

extern "C" int hmj_default_params(const hmj_model_t* m, hmj_params_t* p) {
  if (!m || !p) return fail(HMJ_ERR_ARG, "null argument%s");
  memset(p, 0, sizeof(*p));
  p->timestep = m->timestep;
  memcpy(p->body_mass, m->body_mass, sizeof(p->body_mass));
  memcpy(p->dof_armature, m->dof_armature, sizeof(p->dof_armature));
  memcpy(p->jnt_range, m->jnt_range, sizeof(p->jnt_range));
  memcpy(p->jnt_margin, m->jnt_margin, sizeof(p->jnt_margin));
  memcpy(p->jnt_stiffness, m->jnt_stiffness, sizeof(p->jnt_stiffness));
  memcpy(p->act_gainprm0, m->act_gainprm0, sizeof(p->act_gainprm0));
  memcpy(p->act_forcerange, m->act_forcerange, sizeof(p->act_forcerange));
  memcpy(p->floor_friction, m->geom_friction[0], sizeof(p->floor_friction));
  memcpy(p->hfield_data, m->hfield_data, sizeof(p->hfield_data));
  return HMJ_OK;
}

// Support candidates per direction cell.  A direction d with major axis `ax` (largest |component|, lowest axis on ties) and
// sign sgn lies on cube face 2 * ax + (sgn < 0) at (u, v) = (d[a], d[b]) / |d[ax]|, a = (ax + 1) % 3, b = (ax + 2) % 3;
// cell (iu, iv) of the face covers u in [-1 + 2 iu / G, -1 + 2 (iu + 1) / G], likewise v, widened by HMJ_SUP_DELTA on every
// side (the kernels find the cell in fp32).  Vertex w is dropped from a cell when ONE other vertex u beats it by more than
// `margin` at the four corner directions of the widened cell: (p_u - p_w) . d is affine in (u, v), so u then beats w by more
// than margin * |d[ax]| for every direction of the cell — ten million times the rounding error of the kernels' fp64 dot
// products (coordinates < 1 m) — and w can be neither the exact nor the computed arg-max there.  Everything else stays.
static void build_support_cells(const hmj_model_t* m, hmj_tables_t* t) {
  const int G = HMJ_SUP_G;
  const double margin = 1e-9;
  t->sup_ok = 0;
  static std::mutex mu;
  static std::vector<unsigned char> cache_key;
  static std::vector<uint16_t> cache_start, cache_vert;
  std::lock_guard<std::mutex> lock(mu);
  std::vector<unsigned char> key(sizeof(m->mesh_vert) + sizeof(m->mesh_vertadr) + sizeof(m->mesh_vertnum));
  memcpy(key.data(), m->mesh_vert, sizeof(m->mesh_vert));
  memcpy(key.data() + sizeof(m->mesh_vert), m->mesh_vertadr, sizeof(m->mesh_vertadr));
  memcpy(key.data() + sizeof(m->mesh_vert) + sizeof(m->mesh_vertadr), m->mesh_vertnum, sizeof(m->mesh_vertnum));
  if (key != cache_key) {   // built once per process and model (a few ms)
    std::vector<uint16_t> start, vert;
    std::vector<double> D;
    std::vector<int> keep;
    for (int k = 0; k < HMJ_NMESH; k++) {
      const int nv = m->mesh_vertnum[k], va = m->mesh_vertadr[k];
      D.assign((size_t)nv * 4, 0.0);
      for (int face = 0; face < 6; face++) {
        const int ax = face >> 1, a = (ax + 1) % 3, b = (ax + 2) % 3;
        const double sgn = (face & 1) ? -1.0 : 1.0;
        for (int iu = 0; iu < G; iu++)
          for (int iv = 0; iv < G; iv++) {
            start.push_back((uint16_t)vert.size());
            const double ub[2] = {-1.0 + 2.0 * iu / G - HMJ_SUP_DELTA, -1.0 + 2.0 * (iu + 1) / G + HMJ_SUP_DELTA};
            const double vb[2] = {-1.0 + 2.0 * iv / G - HMJ_SUP_DELTA, -1.0 + 2.0 * (iv + 1) / G + HMJ_SUP_DELTA};
            int lead = 0;
            double lead_sum = -1e300;
            for (int i = 0; i < nv; i++) {
              const double p[3] = {(double)m->mesh_vert[va + i][0], (double)m->mesh_vert[va + i][1], (double)m->mesh_vert[va + i][2]};
              double sum = 0;
              for (int c = 0; c < 4; c++) {
                const double d = sgn * p[ax] + ub[c >> 1] * p[a] + vb[c & 1] * p[b];
                D[(size_t)i * 4 + c] = d;
                sum += d;
              }
              if (sum > lead_sum) { lead_sum = sum; lead = i; }
            }
            auto beats = [&](int u, int w) {
              for (int c = 0; c < 4; c++)
                if (!(D[(size_t)u * 4 + c] - D[(size_t)w * 4 + c] > margin)) return false;
              return true;
            };
            keep.clear();
            for (int w = 0; w < nv; w++) {
              bool out = w != lead && beats(lead, w);   // the vertex best at the cell's centre removes almost everyone
              for (int u = 0; u < nv && !out; u++)
                if (u != w && beats(u, w)) out = true;
              if (!out) keep.push_back(w);
            }
            for (int w : keep) vert.push_back((uint16_t)w);
          }
      }
      start.push_back((uint16_t)vert.size());   // entry HMJ_SUP_CELLS of this mesh
      start.push_back((uint16_t)vert.size());   // padding entry
      if (vert.size() > HMJ_SUP_CAP) return;    // does not fit: sup_ok stays 0, the kernels scan whole hulls
    }
    cache_key = key; cache_start = start; cache_vert = vert;
  }
  memcpy(t->sup_start, cache_start.data(), sizeof(t->sup_start));
  memset(t->sup_vert, 0, sizeof(t->sup_vert));
  memcpy(t->sup_vert, cache_vert.data(), cache_vert.size() * sizeof(uint16_t));
  t->sup_ok = getenv("HMJ_SUP_CELLS_OFF") ? 0 : 1;   // developer switch: whole-hull scans
}

// The candidate tables of a model, for inspection (tests/test_host_logic.py checks them against whole-hull arg-max on the
// CPU; no GPU needed).  start: [HMJ_NMESH][cells + 2] offsets into vert; returns the cells-per-face-edge count G in *grid.
extern "C" int hmj_support_cells(const hmj_model_t* model, uint16_t* start, int start_cap, uint16_t* vert, int vert_cap, int* grid,
                                 int* entries) {
  if (!model || !start || !vert || start_cap < HMJ_NMESH * (HMJ_SUP_CELLS + 2) || vert_cap < HMJ_SUP_CAP)
    return fail(HMJ_ERR_ARG, "bad argument to support_cells%s");
  hmj_tables_t* t = new hmj_tables_t;
  memset(t, 0, sizeof(*t));
  build_support_cells(model, t);
  const int ok = t->sup_ok || getenv("HMJ_SUP_CELLS_OFF");
  memcpy(start, t->sup_start, sizeof(t->sup_start));
  memcpy(vert, t->sup_vert, sizeof(t->sup_vert));
  if (grid) *grid = HMJ_SUP_G;
  if (entries) *entries = t->sup_start[HMJ_NMESH - 1][HMJ_SUP_CELLS];
  delete t;
  return ok ? HMJ_OK : fail(HMJ_ERR_ARG, "candidate tables exceed their capacity%s");
}

static void build_tables(const hmj_model_t* m, hmj_tables_t* t) {
  memset(t, 0, sizeof(*t));
  t->m = *m;
  // levels
  int maxd = 0;
  for (int b = 0; b < HMJ_NBODY; b++) maxd = m->body_depth[b] > maxd ? m->body_depth[b] : maxd;
  t->nlevel = maxd + 1;
  for (int b = 0; b < HMJ_NBODY; b++) {
    int L = m->body_depth[b];
    t->level_body[L][t->level_n[L]++] = b;
  }
  // subtree ranges (DFS numbering: descendants are contiguous)
  for (int b = 0; b < HMJ_NBODY; b++) {
    int e = b + 1;
    while (e < HMJ_NBODY) {
      int a = e;
      while (a > b) a = m->body_parentid[a];
      if (a != b) break;
      e++;
    }
    t->subtree_end[b] = e;
  }
  for (int i = 0; i < HMJ_NV; i++) {
    int adr = m->dof_Madr[i];
    for (int j = i; j >= 0; j = m->dof_parentid[j]) { t->M_row[adr] = i; t->M_col[adr] = j; adr++; }
    int jid = m->dof_jntid[i];
    t->dof_hinge_jnt[i] = (m->jnt_type[jid] == HMJ_JNT_HINGE) ? jid : -1;
  }
  for (int b = 1; b < HMJ_NBODY; b++) {
    int bb = b;
    while (bb && !m->body_dofnum[bb]) bb = m->body_parentid[bb];
    uint32_t mask = 0;
    if (bb) {
      int da = m->body_dofadr[bb] + m->body_dofnum[bb] - 1;
      while (da >= 0) { mask |= 1u << da; da = m->dof_parentid[da]; }
    }
    t->body_dofmask[b] = mask;
  }
  int slot = 0;
  for (int k = 0; k < m->npair; k++) {
    t->pair_slot[k] = slot;
    int n = (m->pair_geom1[k] == 0) ? HMJ_MAXCONPAIR : 1;
    for (int s = 0; s < n; s++) t->slot_pair[slot + s] = k;
    slot += n;
  }
  t->pair_slot[m->npair] = slot;
  {  // narrowphase task groups and their capacities (units of N tasks)
    int cap[64] = {0};
    int nmm = 0;
    for (int k = 0; k < m->npair; k++) {
      if (m->pair_geom1[k] == 0) {
        const int g = m->geom_meshid[m->pair_geom2[k]];
        t->pair_group[k] = g;
        cap[g] += HMJ_MAXCONPAIR;
      } else {
        const int g = HMJ_NMESH + nmm++;
        t->pair_group[k] = g;
        cap[g] = 1;
      }
    }
    int base = 0;
    for (int g = 0; g < HMJ_NGROUP; g++) { t->group_base[g] = base; base += cap[g]; }
    t->group_base[HMJ_NGROUP] = base;
    // cost of one MPR trip of a group ~ vertices scanned per support pair; order groups by it (descending) and cap the
    // item size of the expensive ones so that their items stay short (k_narrow's header explains the schedule)
    int verts[64] = {0};
    for (int k = 0; k < m->npair; k++) {
      const int g = t->pair_group[k];
      const int g1 = m->pair_geom1[k], g2 = m->pair_geom2[k];
      verts[g] = (g1 == 0 ? 0 : m->mesh_vertnum[m->geom_meshid[g1]]) + m->mesh_vertnum[m->geom_meshid[g2]];
    }
    for (int g = 0; g < 64; g++) { t->group_order[g] = g; t->group_shiftcap[g] = 5; }
    for (int a = 1; a < HMJ_NGROUP; a++) {   // insertion sort, stable
      const int g = t->group_order[a];
      int bpos = a;
      while (bpos > 0 && verts[t->group_order[bpos - 1]] < verts[g]) { t->group_order[bpos] = t->group_order[bpos - 1]; bpos--; }
      t->group_order[bpos] = g;
    }
    {  // k_broad's CTA order: the expensive (pair, 128-env) CTAs first, so that the round-robin CTA placement gives every
       // SM one of them before the cheap ones fill the remaining slots
      int cost[HMJ_NPAIR + 1];
      for (int k = 0; k < m->npair; k++)
        cost[k] = (m->pair_geom1[k] == 0 ? 100000 + m->mesh_vertnum[m->geom_meshid[m->pair_geom2[k]]] : 0);
      for (int k = 0; k <= HMJ_NPAIR; k++) t->broad_order[k] = k;
      for (int a = 1; a < m->npair; a++) {
        const int k = t->broad_order[a];
        int bpos = a;
        while (bpos > 0 && cost[t->broad_order[bpos - 1]] < cost[k]) { t->broad_order[bpos] = t->broad_order[bpos - 1]; bpos--; }
        t->broad_order[bpos] = k;
      }
    }
    // with the candidate-list support functions a trip no longer costs more on a big hull: no caps (measured: all 32-query
    // items, settled regime +14 %, C3 unchanged); the override stays for experiments
    int thr[3] = {250, 180, 50}, capv[3] = {5, 5, 5};
    if (const char* e = getenv("HMJ_NARROW_CAP"))   // developer override: "v0:s0,v1:s1,v2:s2" (vertex threshold : log2 item size)
      sscanf(e, "%d:%d,%d:%d,%d:%d", &thr[0], &capv[0], &thr[1], &capv[1], &thr[2], &capv[2]);
    for (int g = 0; g < HMJ_NGROUP; g++)
      t->group_shiftcap[g] = verts[g] >= thr[0] ? capv[0] : (verts[g] >= thr[1] ? capv[1] : (verts[g] >= thr[2] ? capv[2] : 5));
  }
  int adr = 0;
  for (int k = 0; k < HMJ_NMESH; k++) {
    int nv = m->mesh_vertnum[k], va = m->mesh_vertadr[k];
    int npad = (nv + HMJ_MESH_PAD - 1) / HMJ_MESH_PAD * HMJ_MESH_PAD;
    t->mesh_adr[k] = adr;
    t->mesh_npad[k] = npad;
    for (int q = 0; q < 3; q++) { t->mesh_lo[k][q] = 1e30; t->mesh_hi[k][q] = -1e30; }
    for (int i = 0; i < nv; i++)
      for (int q = 0; q < 3; q++) {
        const double c = m->mesh_vert[va + i][q];
        if (c < t->mesh_lo[k][q]) t->mesh_lo[k][q] = c;
        if (c > t->mesh_hi[k][q]) t->mesh_hi[k][q] = c;
      }
    for (int i = 0; i < npad; i++) {
      int src = i < nv ? i : 0;
      t->vx[adr + i] = m->mesh_vert[va + src][0];
      t->vy[adr + i] = m->mesh_vert[va + src][1];
      t->vz[adr + i] = m->mesh_vert[va + src][2];
      t->dvx[adr + i] = (double)t->vx[adr + i];
      t->dvy[adr + i] = (double)t->vy[adr + i];
      t->dvz[adr + i] = (double)t->vz[adr + i];
    }
    adr += npad;
  }
  t->mesh_adr[HMJ_NMESH] = adr;
  build_support_cells(m, t);
  {   // the structure the tree Cholesky is written for: root dofs 0-5 in a line, four serial 3-dof chains off dof 5
    static const int expect[HMJ_NV] = {-1, 0, 1, 2, 3, 4, 5, 6, 7, 5, 9, 10, 5, 12, 13, 5, 15, 16};
    bool ok = true;
    for (int d = 0; d < HMJ_NV; d++) ok = ok && m->dof_parentid[d] == expect[d];
    t->tree_chol = (ok && !getenv("HMJ_TREE_CHOL_OFF")) ? 1 : 0;   // developer switch: dense factorisations everywhere
  }
  int e = 0;
  for (int i = 0; i < HMJ_NV; i++)
    for (int j = 0; j <= i; j++) { t->tri_i[e] = i; t->tri_j[e] = j; e++; }
}

// `count` elements (per-env arrays pass rows x width with rows = N + S: the spare rows of the episode pool sit behind the
// N envs; the named-array accessors see the first N rows only).
template <typename Tp>
static int alloc_array(hmj_batch_* b, const char* name, Tp** out, size_t count, bool per_env = true) {
  void* p = nullptr;
  CK(cudaMalloc(&p, count * sizeof(Tp)));
  b->allocs.push_back(p);
  CK(cudaMemsetAsync(p, 0, count * sizeof(Tp), b->stream));
  const size_t rows = (size_t)b->n + b->n_spare;
  if (name) b->arrays[name] = ArrayInfo{p, per_env ? count / rows * (size_t)b->n * sizeof(Tp) : count * sizeof(Tp)};
  *out = (Tp*)p;
  return HMJ_OK;
}

// inside hmj_batch_create after `b` exists: a failure must release what was built so far
#define CKB(call)                                                                   \
  do {                                                                              \
    cudaError_t e_ = (call);                                                        \
    if (e_ != cudaSuccess) {                                                        \
      hmj_batch_destroy(b);                                                         \
      return fail(HMJ_ERR_CUDA, "CUDA error: %s", cudaGetErrorString(e_));          \
    }                                                                               \
  } while (0)

#define ALLOC(name, field, count)                                   \
  do {                                                              \
    int rc_ = alloc_array(b, name, &b->buf.field, (size_t)(count)); \
    if (rc_) { hmj_batch_destroy(b); return rc_; }                  \
  } while (0)
#define ALLOC_FLAT(name, field, count)                                     \
  do {                                                                     \
    int rc_ = alloc_array(b, name, &b->buf.field, (size_t)(count), false); \
    if (rc_) { hmj_batch_destroy(b); return rc_; }                         \
  } while (0)


// Buffers of envs [start, start + count) as a batch of their own (every per-env array is env-major).
static hmj_buffers_t pipe_view(const hmj_buffers_t& B, int start, int count, int pipe, int group_total) {
  hmj_buffers_t V = B;
  const size_t s = (size_t)start;
  V.n = count;
  V.qpos += s * HMJ_NQ; V.qvel += s * HMJ_NV; V.time += s; V.warmstart += s * HMJ_NV; V.ctrl += s * HMJ_NU;
  V.xfrc += s * 6; V.xfrc_body += s; V.xfrc_all += s * HMJ_NBODY * 6; V.params += s;
  V.geom_xpos += s * HMJ_NGEOM * 3; V.geom_xmat += s * HMJ_NGEOM * 9; V.cdof += s * HMJ_NV * 6; V.com += s * 3;
  V.xipos += s * HMJ_NBODY * 3; V.ximat += s * HMJ_NBODY * 9;
  V.qM += s * HMJ_NM; V.qfrc_smooth += s * HMJ_NV; V.euler_factor += s * HMJ_FAC_DOUBLES; V.sens += s * SENS_NUM; V.slots += s * HMJ_NSLOT * HMJ_SLOT_DOUBLES; V.slot_mask += s * 4;
  V.wq_count += 64 * pipe; V.wq_task += s * group_total;
  V.solve_count += 4 * pipe; V.solve_list += 4 * s;
  V.qacc += s * HMJ_NV; V.qacc_smooth += s * HMJ_NV; V.qfrc_constraint += s * HMJ_NV; V.qfrc_ext += s * HMJ_NV; V.qfrc_inverse += s * HMJ_NV; V.contact += s * HMJ_MAXCON;
  V.ncon += s; V.nefc += s; V.niter += s; V.status += s;
  V.residual += s * HMJ_NRESIDUAL; V.cost_sum += s; V.cost_last += s; V.trace += s * 3;
  return V;
}

#define HMJ_MAX_PIPES 8
static int pipe_count_for(int n) {
  const char* e = getenv("HMJ_PIPES");   // developer override
  int p = e ? atoi(e) : (n >= 2048 ? 2 : 1);   // measured at 4096 envs: 2 pipes +4 %, 4 and 8 none (profiles/README.md)
  if (p < 1) p = 1;
  if (p > HMJ_MAX_PIPES) p = HMJ_MAX_PIPES;
  while (p > 1 && n / p < 128) p--;
  return p;
}

static int make_pipes(hmj_batch_* b) {
  const int P = pipe_count_for(b->n);
  if (P <= 1) return HMJ_OK;
  const int group_total = b->tab_host->group_base[HMJ_NGROUP];
  CK(cudaEventCreateWithFlags(&b->pipe_fork, cudaEventDisableTiming));
  b->pipes.resize(P);
  const int base = b->n / P, rem = b->n % P;
  int start = 0;
  for (int p = 0; p < P; p++) {
    hmj_batch_::Pipe& pp = b->pipes[p];
    const int count = base + (p < rem ? 1 : 0);
    pp.view = pipe_view(b->buf, start, count, p, group_total);
    pp.start = start;
    start += count;
    CK(cudaStreamCreateWithFlags(&pp.stream, cudaStreamNonBlocking));
    CK(cudaEventCreateWithFlags(&pp.done, cudaEventDisableTiming));
    for (int k = 0; k < 2; k++) {
      CK(cudaStreamCreateWithFlags(&pp.ctx.side[k], cudaStreamNonBlocking));
      CK(cudaEventCreateWithFlags(&pp.ctx.join[k], cudaEventDisableTiming));
    }
    CK(cudaEventCreateWithFlags(&pp.ctx.fork, cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&pp.ctx.fac_fork, cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&pp.ctx.fac_done, cudaEventDisableTiming));
  }
  return HMJ_OK;
}

extern "C" void hmj_batch_destroy(hmj_batch_t* b) {
  if (!b) return;
  cudaSetDevice(b->device);
  if (b->stream) cudaStreamSynchronize(b->stream);
  for (void* p : b->allocs) cudaFree(p);
  for (cudaEvent_t e : b->prof_events) cudaEventDestroy(e);
  if (b->io_host) cudaFreeHost(b->io_host);
  if (b->plan_stage) cudaFreeHost(b->plan_stage);
  if (b->term_host) cudaFreeHost(b->term_host);
  for (auto& kv : b->rollout_graphs) cudaGraphExecDestroy(kv.second);
  if (b->tab_dev) cudaFree(b->tab_dev);
  if (b->task_dev) cudaFree(b->task_dev);
  for (int k = 0; k < 2; k++)
    if (b->step_graph[k]) cudaGraphExecDestroy(b->step_graph[k]);
  if (b->ctrl_graph) cudaGraphExecDestroy(b->ctrl_graph);
  if (b->copy_stream) { cudaStreamSynchronize(b->copy_stream); cudaStreamDestroy(b->copy_stream); }
  if (b->h2d_event) cudaEventDestroy(b->h2d_event);
  for (auto& pp : b->pipes) {
    for (int k = 0; k < 2; k++) {
      if (pp.ctx.side[k]) cudaStreamDestroy(pp.ctx.side[k]);
      if (pp.ctx.join[k]) cudaEventDestroy(pp.ctx.join[k]);
    }
    if (pp.ctx.fork) cudaEventDestroy(pp.ctx.fork);
    if (pp.ctx.fac_fork) cudaEventDestroy(pp.ctx.fac_fork);
    if (pp.ctx.fac_done) cudaEventDestroy(pp.ctx.fac_done);
    if (pp.done) cudaEventDestroy(pp.done);
    if (pp.stream) cudaStreamDestroy(pp.stream);
  }
  if (b->pipe_fork) cudaEventDestroy(b->pipe_fork);
  if (b->spare_stream) cudaStreamSynchronize(b->spare_stream);
  if (b->round_graph) cudaGraphExecDestroy(b->round_graph);
  for (int k = 0; k < 2; k++) {
    if (b->spare_ctx.side[k]) cudaStreamDestroy(b->spare_ctx.side[k]);
    if (b->spare_ctx.join[k]) cudaEventDestroy(b->spare_ctx.join[k]);
  }
  if (b->spare_ctx.fork) cudaEventDestroy(b->spare_ctx.fork);
  if (b->spare_ctx.fac_fork) cudaEventDestroy(b->spare_ctx.fac_fork);
  if (b->spare_ctx.fac_done) cudaEventDestroy(b->spare_ctx.fac_done);
  if (b->ev_main) cudaEventDestroy(b->ev_main);
  if (b->round_done) cudaEventDestroy(b->round_done);
  if (b->spare_stream) cudaStreamDestroy(b->spare_stream);
  for (int k = 0; k < 2; k++) {
    if (b->solve_ctx.side[k]) cudaStreamDestroy(b->solve_ctx.side[k]);
    if (b->solve_ctx.join[k]) cudaEventDestroy(b->solve_ctx.join[k]);
  }
  if (b->solve_ctx.fork) cudaEventDestroy(b->solve_ctx.fork);
  if (b->solve_ctx.fac_fork) cudaEventDestroy(b->solve_ctx.fac_fork);
  if (b->solve_ctx.fac_done) cudaEventDestroy(b->solve_ctx.fac_done);
  if (b->stream) cudaStreamDestroy(b->stream);
  delete b->tab_host;
  delete b;
}

extern "C" int hmj_batch_create(const hmj_model_t* model, int n_envs, int device, hmj_batch_t** out) {
  if (!model || !out || n_envs <= 0) return fail(HMJ_ERR_ARG, "bad argument to hmj_batch_create%s");
  if (model->magic != HMJ_MAGIC || model->version != HMJ_VERSION || model->sizeof_model != (int)sizeof(hmj_model_t))
    return fail(HMJ_ERR_MODEL, "model blob magic/version/size mismatch%s");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
    return fail(HMJ_ERR_CUDA, "no CUDA device: the hmj_b200 engine has no CPU fallback%s");
  if (device < 0 || device >= ndev) return fail(HMJ_ERR_ARG, "bad device index%s");
  CK(cudaSetDevice(device));
  hmj_batch_* b = new hmj_batch_();
  b->n = n_envs;
  b->device = device;
  b->buf.n = n_envs;
  CKB(cudaStreamCreateWithFlags(&b->stream, cudaStreamNonBlocking));
  CKB(cudaStreamCreateWithFlags(&b->copy_stream, cudaStreamNonBlocking));
  CKB(cudaEventCreateWithFlags(&b->h2d_event, cudaEventDisableTiming));
  for (int k = 0; k < 2; k++) {
    CKB(cudaStreamCreateWithFlags(&b->solve_ctx.side[k], cudaStreamNonBlocking));
    CKB(cudaEventCreateWithFlags(&b->solve_ctx.join[k], cudaEventDisableTiming));
  }
  CKB(cudaEventCreateWithFlags(&b->solve_ctx.fork, cudaEventDisableTiming));
  CKB(cudaEventCreateWithFlags(&b->solve_ctx.fac_fork, cudaEventDisableTiming));
  CKB(cudaEventCreateWithFlags(&b->solve_ctx.fac_done, cudaEventDisableTiming));
  b->tab_host = new hmj_tables_t;
  build_tables(model, b->tab_host);
  CKB(cudaMalloc((void**)&b->tab_dev, sizeof(hmj_tables_t)));
  CKB(cudaMemcpyAsync(b->tab_dev, b->tab_host, sizeof(hmj_tables_t), cudaMemcpyHostToDevice, b->stream));
  hmj_task_t task;
  memset(&task, 0, sizeof(task));
  task.nterm = model->task_nterm;
  for (int k = 0; k < model->task_nterm; k++) {
    task.dim[k] = model->task_dim[k]; task.norm[k] = model->task_norm[k]; task.weight[k] = model->task_weight[k];
    for (int q = 0; q < HMJ_MAXNORMPRM; q++) task.normprm[k][q] = model->task_normprm[k][q];
  }
  task.risk = model->task_risk;
  task.height_goal = model->task_height_goal;
  CKB(cudaMalloc((void**)&b->task_dev, sizeof(hmj_task_t)));
  CKB(cudaMemcpyAsync(b->task_dev, &task, sizeof(task), cudaMemcpyHostToDevice, b->stream));

  {   // spare rows of the episode pool: N / 8 clamped to [32, 1024] (HMJ_SPARE overrides; 0 = no pool)
    int S = n_envs / 8;
    S = S < 32 ? 32 : (S > 1024 ? 1024 : S);
    if (const char* e = getenv("HMJ_SPARE")) S = atoi(e);
    if (S < 0) S = 0;
    if (S > 1024) S = 1024;
    b->n_spare = S;
  }
  const size_t N = (size_t)n_envs + b->n_spare;   // rows of every per-env array
  ALLOC("qpos", qpos, N * HMJ_NQ);
  ALLOC("qvel", qvel, N * HMJ_NV);
  ALLOC("time", time, N);
  ALLOC("warmstart", warmstart, N * HMJ_NV);
  ALLOC("ctrl", ctrl, N * HMJ_NU);
  ALLOC("xfrc", xfrc, N * 6);
  ALLOC("xfrc_body", xfrc_body, N);
  ALLOC("xfrc_all", xfrc_all, N * HMJ_NBODY * 6);
  ALLOC("params", params, N);
  ALLOC("geom_xpos", geom_xpos, N * HMJ_NGEOM * 3);
  ALLOC("geom_xmat", geom_xmat, N * HMJ_NGEOM * 9);
  ALLOC("cdof", cdof, N * HMJ_NV * 6);
  ALLOC("com", com, N * 3);
  ALLOC("xipos", xipos, N * HMJ_NBODY * 3);
  ALLOC("ximat", ximat, N * HMJ_NBODY * 9);
  ALLOC("qM", qM, N * HMJ_NM);
  ALLOC("qfrc_smooth", qfrc_smooth, N * HMJ_NV);
  ALLOC("euler_factor", euler_factor, N * HMJ_FAC_DOUBLES);
  ALLOC("sens", sens, N * SENS_NUM);
  ALLOC("slots", slots, N * HMJ_NSLOT * HMJ_SLOT_DOUBLES);
  ALLOC("slot_mask", slot_mask, N * 4);
  ALLOC_FLAT("wq_count", wq_count, 64 * (HMJ_MAX_PIPES + 1));
  ALLOC("wq_task", wq_task, N * (size_t)b->tab_host->group_base[HMJ_NGROUP]);
  ALLOC_FLAT("solve_count", solve_count, 4 * (HMJ_MAX_PIPES + 1));
  ALLOC("solve_list", solve_list, 4 * N);
  ALLOC("qacc", qacc, N * HMJ_NV);
  ALLOC("qacc_smooth", qacc_smooth, N * HMJ_NV);
  ALLOC("qfrc_constraint", qfrc_constraint, N * HMJ_NV);
  ALLOC("qfrc_ext", qfrc_ext, N * HMJ_NV);
  ALLOC("qfrc_inverse", qfrc_inverse, N * HMJ_NV);
  ALLOC("contact", contact, N * HMJ_MAXCON);
  ALLOC("ncon", ncon, N);
  ALLOC("nefc", nefc, N);
  ALLOC("niter", niter, N);
  ALLOC("status", status, N);
  ALLOC("residual", residual, N * HMJ_NRESIDUAL);
  ALLOC("cost_sum", cost_sum, N);
  ALLOC("cost_last", cost_last, N);
  ALLOC("trace", trace, N * 3);
  {
    int rc;
    if ((rc = alloc_array(b, nullptr, &b->state_dev, HMJ_NSTATE + 1, false)) || (rc = alloc_array(b, nullptr, &b->nominal_dev, 64 * HMJ_NU, false)) || (rc = alloc_array(b, nullptr, &b->stdvec_dev, 64 * HMJ_NU, false)) || (rc = alloc_array(b, nullptr, &b->seed_iter_dev, 2, false)) ||
        (rc = alloc_array(b, nullptr, &b->record_dev, 3 + 64 * HMJ_NU, false)) || (rc = alloc_array(b, nullptr, &b->select_dev, 3 + 64 * HMJ_NU, false)) || (rc = alloc_array(b, nullptr, &b->spline_times_dev, 64, false)) ||
        (rc = alloc_array(b, "returns", &b->returns_dev, N)) || (rc = alloc_array(b, nullptr, &b->failure_dev, N)) ||
        (rc = alloc_array(b, "argmin", &b->argmin_dev, 2, false))) {
      hmj_batch_destroy(b);
      return rc;
    }
  }
  {  // CPUEnv bookkeeping arrays (named, so the host side sets them at reset with hmj_batch_set_array)
    int rc = 0;
#define EALLOC(name, field, count) if (!rc) rc = alloc_array(b, name, &b->env.field, (size_t)(count))
    EALLOC("env_act_noise", act_noise, N * HMJ_NU);
    EALLOC("env_obs_noise", obs_noise, N * 31);
    EALLOC("env_act_delay", act_delay, N);
    EALLOC("env_obs_delay", obs_delay, N * 3);
    EALLOC("env_act_count", act_count, N);
    EALLOC("env_obs_count", obs_count, N);
    EALLOC("env_act_hist", act_hist, N * HMJ_ENV_HIST * HMJ_NU);
    EALLOC("env_obs_hist", obs_hist, N * HMJ_ENV_HIST * 30);
    EALLOC("env_prev_action", prev_action, N * HMJ_NU);
    EALLOC("env_latest_action", latest_action, N * HMJ_NU);
    EALLOC("env_force_sched", force_sched, N * 5);
    EALLOC("env_force_body", force_body, N);
    EALLOC("env_force_draws", force_draws, N);
    EALLOC("env_torso_quat", torso_quat, N * 4);
#undef EALLOC
    const size_t NE = (size_t)n_envs;
    b->io_in_bytes = NE * HMJ_NU * 8 + NE * 4;
    b->io_in_bytes = (b->io_in_bytes + 63) / 64 * 64;
    b->io_out_bytes = NE * (30 + 1 + 1 + 4) * 8 + 3 * NE * 4;
    // the episode pool's per-step result block rides behind the flags: counts[8] i32 | ended_ids[64] i32 | term_obs[64][30] f64
    b->io_ext_offset = (b->io_out_bytes + 7) / 8 * 8;
    b->io_out_bytes = b->io_ext_offset + 8 * 4 + HMJ_POOL_MAXTERM * 4 + HMJ_POOL_MAXTERM * 30 * 8;
    if (!rc && cudaHostAlloc((void**)&b->io_host, b->io_in_bytes + b->io_out_bytes, cudaHostAllocDefault) != cudaSuccess)
      rc = fail(HMJ_ERR_CUDA, "cudaHostAlloc failed%s");
    if (!rc) { memset(b->io_host, 0, b->io_in_bytes + b->io_out_bytes); rc = alloc_array(b, nullptr, &b->io_dev, b->io_in_bytes + b->io_out_bytes); }
    if (!rc && cudaHostAlloc((void**)&b->plan_stage, 3072 * sizeof(double), cudaHostAllocDefault) != cudaSuccess)
      rc = fail(HMJ_ERR_CUDA, "cudaHostAlloc failed%s");
    if (!rc && b->n_spare > 0) {   // episode pool
      const size_t S = (size_t)b->n_spare;
      hmj_pool_t& P = b->pool;
      P.S = (int32_t)S;
#define PALLOC(field, count) if (!rc) rc = alloc_array(b, nullptr, &P.field, (size_t)(count), false)
      PALLOC(state, S); PALLOC(active, S); PALLOC(gen, S); PALLOC(obs, S * 30); PALLOC(reward, S);
      PALLOC(term, S); PALLOC(trunc, S); PALLOC(selfc, S);
      PALLOC(pending, NE); PALLOC(assign, NE); PALLOC(served_list, NE);
      PALLOC(ended_ids_full, NE); PALLOC(term_obs_full, NE * 30);
#undef PALLOC
      if (!rc) rc = alloc_array(b, "env_episode_src", &P.episode_src, N);
      if (!rc && cudaHostAlloc((void**)&b->term_host, NE * (30 * 8 + 4), cudaHostAllocDefault) != cudaSuccess)
        rc = fail(HMJ_ERR_CUDA, "cudaHostAlloc failed%s");
      if (!rc) {
        char* ext = b->io_dev + b->io_in_bytes + b->io_ext_offset;
        P.counts = (int32_t*)ext;
        P.ended_ids = P.counts + 8;
        P.term_obs = (double*)(ext + 8 * 4 + HMJ_POOL_MAXTERM * 4);
        b->arrays["env_terminal_obs_compact"] = ArrayInfo{P.term_obs_full, NE * 30 * 8};
        b->arrays["env_ended_ids_compact"] = ArrayInfo{P.ended_ids_full, NE * 4};
        b->arrays["pool_state"] = ArrayInfo{P.state, S * 4};
      }
      // highest stream priority: a round is a chain of ~15 small kernels that must get SM slots ahead of the main batch's
      // large grids, or its latency under load (not its work) would stall the step that joins it
      int prio_lo = 0, prio_hi = 0;
      cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi);
      if (!rc && (cudaStreamCreateWithPriority(&b->spare_stream, cudaStreamNonBlocking, prio_hi) != cudaSuccess ||
                  cudaEventCreateWithFlags(&b->ev_main, cudaEventDisableTiming) != cudaSuccess ||
                  cudaEventCreateWithFlags(&b->round_done, cudaEventDisableTiming) != cudaSuccess))
        rc = fail(HMJ_ERR_CUDA, "stream/event creation failed%s");
      for (int k = 0; k < 2 && !rc; k++)
        if (cudaStreamCreateWithPriority(&b->spare_ctx.side[k], cudaStreamNonBlocking, prio_hi) != cudaSuccess ||
            cudaEventCreateWithFlags(&b->spare_ctx.join[k], cudaEventDisableTiming) != cudaSuccess)
          rc = fail(HMJ_ERR_CUDA, "stream/event creation failed%s");
      if (!rc && (cudaEventCreateWithFlags(&b->spare_ctx.fork, cudaEventDisableTiming) != cudaSuccess ||
                  cudaEventCreateWithFlags(&b->spare_ctx.fac_fork, cudaEventDisableTiming) != cudaSuccess ||
                  cudaEventCreateWithFlags(&b->spare_ctx.fac_done, cudaEventDisableTiming) != cudaSuccess))
        rc = fail(HMJ_ERR_CUDA, "stream/event creation failed%s");
    }
    if (rc) { hmj_batch_destroy(b); return rc; }
  }
  // default params and state (qpos0) for every env
  {
    std::vector<hmj_params_t> hp(N);
    hmj_default_params(model, &hp[0]);
    for (size_t i = 1; i < N; i++) hp[i] = hp[0];
    CKB(cudaMemcpyAsync(b->buf.params, hp.data(), N * sizeof(hmj_params_t), cudaMemcpyHostToDevice, b->stream));
    std::vector<double> q(N * HMJ_NQ);
    for (size_t i = 0; i < N; i++) memcpy(&q[i * HMJ_NQ], model->qpos0, sizeof(model->qpos0));
    CKB(cudaMemcpyAsync(b->buf.qpos, q.data(), q.size() * sizeof(double), cudaMemcpyHostToDevice, b->stream));
    CKB(cudaStreamSynchronize(b->stream));
  }
  if (const char* e = getenv("HMJ_GRAPH")) b->graph_mode = atoi(e);   // developer override: 0 = direct launches
  if (const char* e = getenv("HMJ_E2E")) b->ctrl_mode = atoi(e);      // developer override: 0 = whole-batch prologue / epilogue
  {
    int rc = make_pipes(b);
    if (rc) { hmj_batch_destroy(b); return rc; }
  }
  if (b->n_spare > 0) {   // the spare rows as a batch of their own, masked to the rows a generation round works on
    b->spare = pipe_view(b->buf, b->n, b->n_spare, HMJ_MAX_PIPES, b->tab_host->group_base[HMJ_NGROUP]);
    b->spare.active = b->pool.active;
    const size_t s0 = (size_t)b->n;
    hmj_envstate_t& E = b->env_spare;
    E = b->env;
    E.act_noise += s0 * HMJ_NU; E.obs_noise += s0 * 31; E.act_delay += s0; E.obs_delay += s0 * 3; E.act_count += s0; E.obs_count += s0;
    E.act_hist += s0 * HMJ_ENV_HIST * HMJ_NU; E.obs_hist += s0 * HMJ_ENV_HIST * 30; E.prev_action += s0 * HMJ_NU;
    E.latest_action += s0 * HMJ_NU; E.force_sched += s0 * 5; E.force_body += s0; E.force_draws += s0; E.torso_quat += s0 * 4;
  }
  *out = b;
  return HMJ_OK;
}

extern "C" int hmj_batch_size(const hmj_batch_t* b) { return b ? b->n : 0; }
extern "C" int hmj_batch_kernel_launches(const hmj_batch_t* b) { return b ? b->launches : 0; }
extern "C" int hmj_batch_host_launches(const hmj_batch_t* b) { return b ? b->host_launches : 0; }
extern "C" void* hmj_batch_stream(hmj_batch_t* b) { return b ? (void*)b->stream : nullptr; }

extern "C" int hmj_batch_sync(hmj_batch_t* b) {
  if (!b) return fail(HMJ_ERR_ARG, "null batch%s");
  CK(cudaSetDevice(b->device));
  CK(cudaStreamSynchronize(b->stream));
  CK(cudaGetLastError());
  return HMJ_OK;
}

static int h2d(hmj_batch_t* b, void* dst, const void* src, size_t bytes) {
  if (!src) return HMJ_OK;
  CK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, b->stream));
  return HMJ_OK;
}
static int d2h(hmj_batch_t* b, void* dst, const void* src, size_t bytes) {
  if (!dst) return HMJ_OK;
  CK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, b->stream));
  return HMJ_OK;
}

extern "C" int hmj_batch_set_params(hmj_batch_t* b, const hmj_params_t* params, int first, int count) {
  if (!b || !params || first < 0 || count < 0 || first + count > b->n) return fail(HMJ_ERR_ARG, "bad argument to set_params%s");
  CK(cudaSetDevice(b->device));
  int rc = h2d(b, b->buf.params + first, params, (size_t)count * sizeof(hmj_params_t));
  if (rc) return rc;
  CK(cudaStreamSynchronize(b->stream));
  return HMJ_OK;
}

extern "C" int hmj_batch_get_params(hmj_batch_t* b, hmj_params_t* params, int first, int count) {
  if (!b || !params || first < 0 || count < 0 || first + count > b->n) return fail(HMJ_ERR_ARG, "bad argument to get_params%s");
  CK(cudaSetDevice(b->device));
  int rc = d2h(b, params, b->buf.params + first, (size_t)count * sizeof(hmj_params_t));
  if (rc) return rc;
  CK(cudaStreamSynchronize(b->stream));
  return HMJ_OK;
}

extern "C" int hmj_batch_set_state(hmj_batch_t* b, const double* qpos, const double* qvel, const double* time, const double* ws) {
  if (!b) return fail(HMJ_ERR_ARG, "null batch%s");
  CK(cudaSetDevice(b->device));
  const size_t N = b->n;
  int rc;
  if ((rc = h2d(b, b->buf.qpos, qpos, N * HMJ_NQ * 8)) || (rc = h2d(b, b->buf.qvel, qvel, N * HMJ_NV * 8)) ||
      (rc = h2d(b, b->buf.time, time, N * 8)) || (rc = h2d(b, b->buf.warmstart, ws, N * HMJ_NV * 8)))
    return rc;
  CK(cudaStreamSynchronize(b->stream));
  return HMJ_OK;
}

extern "C" int hmj_batch_get_state(hmj_batch_t* b, double* qpos, double* qvel, double* time, double* ws) {
  if (!b) return fail(HMJ_ERR_ARG, "null batch%s");
  CK(cudaSetDevice(b->device));
  const size_t N = b->n;
  int rc;
  if ((rc = d2h(b, qpos, b->buf.qpos, N * HMJ_NQ * 8)) || (rc = d2h(b, qvel, b->buf.qvel, N * HMJ_NV * 8)) ||
      (rc = d2h(b, time, b->buf.time, N * 8)) || (rc = d2h(b, ws, b->buf.warmstart, N * HMJ_NV * 8)))
    return rc;
  CK(cudaStreamSynchronize(b->stream));
  return HMJ_OK;
}

extern "C" int hmj_batch_set_ctrl(hmj_batch_t* b, const double* ctrl) {
  if (!b || !ctrl) return fail(HMJ_ERR_ARG, "bad argument to set_ctrl%s");
  CK(cudaSetDevice(b->device));
  int rc = h2d(b, b->buf.ctrl, ctrl, (size_t)b->n * HMJ_NU * 8);
  if (rc) return rc;
  CK(cudaStreamSynchronize(b->stream));
  return HMJ_OK;
}

extern "C" int hmj_batch_set_xfrc(hmj_batch_t* b, const int32_t* body, const double* wrench) {
  if (!b || !body || !wrench) return fail(HMJ_ERR_ARG, "bad argument to set_xfrc%s");
  CK(cudaSetDevice(b->device));
  int rc;
  if ((rc = h2d(b, b->buf.xfrc_body, body, (size_t)b->n * 4)) || (rc = h2d(b, b->buf.xfrc, wrench, (size_t)b->n * 6 * 8))) return rc;
  CK(cudaStreamSynchronize(b->stream));
  return HMJ_OK;
}

static cudaEvent_t prof_event(hmj_batch_t* b) {
  if (b->prof_used == b->prof_events.size()) {
    cudaEvent_t e;
    cudaEventCreate(&e);
    b->prof_events.push_back(e);
  }
  cudaEvent_t e = b->prof_events[b->prof_used++];
  cudaEventRecord(e, b->stream);
  return e;
}

static void enqueue_step(hmj_batch_t* b, int integrate) {
  const bool prof = b->profile;
  if (prof) prof_event(b);
  hmj_launch_smooth(b->tab_dev, b->buf, b->task_dev, 0, nullptr, nullptr, 0, 0, 0, b->stream, &b->solve_ctx);
  if (prof) prof_event(b);
  hmj_launch_factor(b->tab_dev, b->buf, b->stream, &b->solve_ctx);
  hmj_launch_collide(b->tab_dev, b->buf, b->stream);
  if (prof) prof_event(b);
  hmj_launch_solve(b->tab_dev, b->buf, integrate, b->stream, &b->solve_ctx);
  if (prof) prof_event(b);
  b->launches += 10; b->host_launches += 10;  // kin, dyn, factor, broad, narrow, classify, solve x 4 classes
}

// The launches of ONE physics step of the whole batch, on the main stream (forking pipes / side streams and joining
// them back before returning): this is what gets captured into the step graph.
static void enqueue_one_step(hmj_batch_t* b, int integrate) {
  if (b->pipes.empty()) {
    hmj_launch_smooth(b->tab_dev, b->buf, b->task_dev, 0, nullptr, nullptr, 0, 0, 0, b->stream, &b->solve_ctx);
    hmj_launch_factor(b->tab_dev, b->buf, b->stream, &b->solve_ctx);
    hmj_launch_collide(b->tab_dev, b->buf, b->stream, false);
    hmj_launch_solve(b->tab_dev, b->buf, integrate, b->stream, &b->solve_ctx, false);
    return;
  }
  cudaEventRecord(b->pipe_fork, b->stream);
  for (auto& pp : b->pipes) cudaStreamWaitEvent(pp.stream, b->pipe_fork, 0);
  for (auto& pp : b->pipes) hmj_launch_smooth(b->tab_dev, pp.view, b->task_dev, 0, nullptr, nullptr, 0, 0, 0, pp.stream, &pp.ctx);
  for (auto& pp : b->pipes) hmj_launch_factor(b->tab_dev, pp.view, pp.stream, &pp.ctx);
  for (auto& pp : b->pipes) hmj_launch_collide(b->tab_dev, pp.view, pp.stream, false);
  for (auto& pp : b->pipes) hmj_launch_solve(b->tab_dev, pp.view, integrate, pp.stream, &pp.ctx, false);
  for (auto& pp : b->pipes) {
    cudaEventRecord(pp.done, pp.stream);
    cudaStreamWaitEvent(b->stream, pp.done, 0);
  }
}

// ONE control step of the whole batch as a chain on the main stream (captured into b->ctrl_graph): every pipe runs its
// own prologue (k_env_pre on its rows), physics step and epilogue (observation, reward, flags, delay FIFOs in one kernel)
// and, with `to_host`, copies its observation rows to the pinned result block as soon as they exist — prologue, epilogue
// and copy of one pipe overlap the physics of the other; the small result arrays follow in one copy after the join.
// Without `to_host` the caller copies the block itself (the episode pool patches it first).
static void enqueue_control_chain(hmj_batch_t* b, int has_action, double rf, double mst, unsigned long long seed, bool to_host) {
  const size_t N = b->n;
  const double* d_action = has_action ? (const double*)b->io_dev : nullptr;
  double* d_out = (double*)(b->io_dev + b->io_in_bytes);
  char* h_out = b->io_host + b->io_in_bytes;
  hmj_results_t R{};
  R.obs = d_out; R.reward = d_out + N * 30; R.time = d_out + N * 31; R.quat = d_out + N * 32;
  R.terminated = (int32_t*)(d_out + N * 36); R.truncated = R.terminated + N; R.self_collision = R.terminated + 2 * N;
  struct Part { hmj_buffers_t view; int start; cudaStream_t stream; const hmj_solve_ctx_t* ctx; };
  std::vector<Part> parts;
  if (b->pipes.empty()) parts.push_back(Part{b->buf, 0, b->stream, &b->solve_ctx});
  else {
    cudaEventRecord(b->pipe_fork, b->stream);
    for (auto& pp : b->pipes) {
      cudaStreamWaitEvent(pp.stream, b->pipe_fork, 0);
      parts.push_back(Part{pp.view, pp.start, pp.stream, &pp.ctx});
    }
  }
  // the prologue (action noise + FIFO -> ctrl, push scheduler -> xfrc) feeds k_dyn, not k_kin: it goes to the part's side
  // stream, where k_dyn follows it, and runs beside k_kin (with the action upload in front of it: launch_control_graph)
  for (auto& q : parts) {
    cudaEventRecord(q.ctx->fork, q.stream);
    cudaStreamWaitEvent(q.ctx->side[0], q.ctx->fork, 0);
    hmj_launch_env_pre_rows(b->buf, b->env, d_action, rf, seed, q.start, q.view.n, q.ctx->side[0]);
  }
  for (auto& q : parts) hmj_launch_smooth(b->tab_dev, q.view, b->task_dev, 0, nullptr, nullptr, 0, 0, 0, q.stream, q.ctx);
  for (auto& q : parts) hmj_launch_factor(b->tab_dev, q.view, q.stream, q.ctx);
  for (auto& q : parts) hmj_launch_collide(b->tab_dev, q.view, q.stream, false);
  for (auto& q : parts) hmj_launch_solve(b->tab_dev, q.view, 1, q.stream, q.ctx, false);
  for (auto& q : parts) {
    hmj_launch_env_epilogue(b->tab_dev, b->buf, b->env, mst, R, q.start, q.view.n, q.stream);
    if (to_host)
      cudaMemcpyAsync(h_out + (size_t)q.start * 30 * 8, R.obs + (size_t)q.start * 30, (size_t)q.view.n * 30 * 8,
                      cudaMemcpyDeviceToHost, q.stream);
  }
  for (auto& pp : b->pipes) {
    cudaEventRecord(pp.done, pp.stream);
    cudaStreamWaitEvent(b->stream, pp.done, 0);
  }
  if (to_host) cudaMemcpyAsync(h_out + N * 30 * 8, R.reward, b->io_ext_offset - N * 30 * 8, cudaMemcpyDeviceToHost, b->stream);
}

// Replays (capturing it first when the scalars baked into it changed) the control-step graph.  false: not available.
static bool launch_control_graph(hmj_batch_t* b, int has_action, double rf, double mst, unsigned long long seed, bool split,
                                 bool side_copy) {
  hmj_batch_::CtrlKey key;
  key.has_action = has_action; key.split = (split ? 1 : 0) + (side_copy ? 2 : 0); key.rf = rf; key.mst = mst; key.seed = seed;
  if (!b->ctrl_graph || !(key == b->ctrl_key)) {
    if (b->ctrl_graph) { cudaGraphExecDestroy(b->ctrl_graph); b->ctrl_graph = nullptr; }
    cudaGraph_t g = nullptr;
    if (cudaStreamBeginCapture(b->stream, cudaStreamCaptureModeThreadLocal) != cudaSuccess) { cudaGetLastError(); return false; }
    enqueue_control_chain(b, has_action, rf, mst, seed, split);
    bool ok = cudaStreamEndCapture(b->stream, &g) == cudaSuccess && g;
    b->ctrl_graph_waits_h2d = false;
    if (ok && side_copy) {
      // an event-wait node in front of every prologue kernel: the action upload is recorded into h2d_event before each launch
      size_t nn = 0;
      std::vector<cudaGraphNode_t> nodes;
      if (cudaGraphGetNodes(g, nullptr, &nn) == cudaSuccess && nn) {
        nodes.resize(nn);
        cudaGraphGetNodes(g, nodes.data(), &nn);
      }
      cudaGraphNode_t wait_node = nullptr;
      int wired = 0;
      if (cudaGraphAddEventWaitNode(&wait_node, g, nullptr, 0, b->h2d_event) == cudaSuccess) {
        for (cudaGraphNode_t nd : nodes) {
          cudaGraphNodeType ty;
          cudaKernelNodeParams kp;
          if (cudaGraphNodeGetType(nd, &ty) != cudaSuccess || ty != cudaGraphNodeTypeKernel) continue;
          if (cudaGraphKernelNodeGetParams(nd, &kp) != cudaSuccess || kp.func != hmj_env_pre_kernel()) continue;
          if (cudaGraphAddDependencies(g, &wait_node, &nd, 1) == cudaSuccess) wired++;
        }
      }
      const int want = b->pipes.empty() ? 1 : (int)b->pipes.size();
      if (wired == want) b->ctrl_graph_waits_h2d = true;
      else { ok = false; cudaGetLastError(); }   // do not run a half-wired graph
    }
    if (!ok || cudaGraphInstantiate(&b->ctrl_graph, g, 0) != cudaSuccess) {
      if (g) cudaGraphDestroy(g);
      b->ctrl_graph = nullptr;
      cudaGetLastError();
      return false;
    }
    cudaGraphDestroy(g);
    b->ctrl_key = key;
  }
  if (cudaGraphLaunch(b->ctrl_graph, b->stream) != cudaSuccess) { cudaGetLastError(); return false; }
  const int P = b->pipes.empty() ? 1 : (int)b->pipes.size();
  b->launches += 12 * P;   // prologue + 10 step kernels + epilogue per pipe
  b->host_launches += 1;
  return true;
}

// nsteps physics steps of the whole batch.  The step chain (8 kernels per pipe, forks and joins; the per-step counters are zeroed inside k_smooth) is
// captured once per `integrate` flavour into a CUDA graph and replayed: one launch per step instead of ~15 API calls,
// and the inter-kernel gaps shrink to the graph's dependency latency.  The first step of a batch runs un-captured
// (it performs the launchers' one-time attribute / occupancy queries, which must not happen inside a capture).
static void enqueue_steps(hmj_batch_t* b, int integrate, int nsteps) {
  if (nsteps <= 0) return;
  if (b->profile) {
    for (int s = 0; s < nsteps; s++) enqueue_step(b, integrate);
    return;
  }
  const int per_step = 10 * (b->pipes.empty() ? 1 : (int)b->pipes.size());
  const int f = integrate ? 1 : 0;
  for (int s = 0; s < nsteps; s++) {
    b->launches += per_step;
    if (b->graph_mode == 0 || !b->warmed) {   // graphs disabled, or first step ever
      b->host_launches += per_step;
      enqueue_one_step(b, integrate);
      b->warmed = true;
      continue;
    }
    if (!b->step_graph[f]) {
      cudaGraph_t g = nullptr;
      if (cudaStreamBeginCapture(b->stream, cudaStreamCaptureModeThreadLocal) == cudaSuccess) {
        enqueue_one_step(b, integrate);
        if (cudaStreamEndCapture(b->stream, &g) == cudaSuccess && g &&
            cudaGraphInstantiate(&b->step_graph[f], g, 0) == cudaSuccess) {
          cudaGraphDestroy(g);
        } else {
          if (g) cudaGraphDestroy(g);
          b->step_graph[f] = nullptr;
          b->graph_mode = 0;   // capture unavailable: keep launching directly
          cudaGetLastError();
          enqueue_one_step(b, integrate);
          continue;
        }
      } else {
        b->graph_mode = 0;
        cudaGetLastError();
        enqueue_one_step(b, integrate);
        continue;
      }
    }
    cudaGraphLaunch(b->step_graph[f], b->stream);
    b->host_launches += 1;
  }
}

// Per-kernel device time of the steps enqueued while profiling was on (CUDA events on the batch stream).
extern "C" int hmj_batch_profile(hmj_batch_t* b, int enable) {
  if (!b) return fail(HMJ_ERR_ARG, "null batch%s");
  CK(cudaSetDevice(b->device));
  CK(cudaStreamSynchronize(b->stream));
  b->profile = enable != 0;
  b->prof_used = 0;
  return HMJ_OK;
}

extern "C" int hmj_batch_profile_read(hmj_batch_t* b, double* ms_smooth, double* ms_collide, double* ms_solve, int* nsteps) {
  if (!b) return fail(HMJ_ERR_ARG, "null batch%s");
  CK(cudaSetDevice(b->device));
  CK(cudaStreamSynchronize(b->stream));
  double acc[3] = {0, 0, 0};
  const size_t n = b->prof_used / 4;
  for (size_t i = 0; i < n; i++)
    for (int k = 0; k < 3; k++) {
      float ms = 0;
      CK(cudaEventElapsedTime(&ms, b->prof_events[4 * i + k], b->prof_events[4 * i + k + 1]));
      acc[k] += ms;
    }
  if (ms_smooth) *ms_smooth = acc[0];
  if (ms_collide) *ms_collide = acc[1];
  if (ms_solve) *ms_solve = acc[2];
  if (nsteps) *nsteps = (int)n;
  return HMJ_OK;
}

extern "C" int hmj_batch_step_dev(hmj_batch_t* b, const double* ctrl_dev, int nsteps) {
  if (!b || nsteps < 0) return fail(HMJ_ERR_ARG, "bad argument to step%s");
  CK(cudaSetDevice(b->device));
  if (ctrl_dev && ctrl_dev != b->buf.ctrl)
    CK(cudaMemcpyAsync(b->buf.ctrl, ctrl_dev, (size_t)b->n * HMJ_NU * 8, cudaMemcpyDeviceToDevice, b->stream));
  enqueue_steps(b, 1, nsteps);
  CK(cudaGetLastError());
  return HMJ_OK;
}

extern "C" int hmj_batch_step(hmj_batch_t* b, const double* ctrl, int nsteps) {
  if (!b || nsteps < 0) return fail(HMJ_ERR_ARG, "bad argument to step%s");
  CK(cudaSetDevice(b->device));
  int rc = h2d(b, b->buf.ctrl, ctrl, (size_t)b->n * HMJ_NU * 8);
  if (rc) return rc;
  enqueue_steps(b, 1, nsteps);
  CK(cudaStreamSynchronize(b->stream));
  CK(cudaGetLastError());
  return HMJ_OK;
}

extern "C" int hmj_batch_forward(hmj_batch_t* b) {
  if (!b) return fail(HMJ_ERR_ARG, "null batch%s");
  CK(cudaSetDevice(b->device));
  enqueue_steps(b, 0, 1);
  CK(cudaStreamSynchronize(b->stream));
  CK(cudaGetLastError());
  return HMJ_OK;
}

extern "C" int hmj_batch_run_stage(hmj_batch_t* b, int stage) {
  if (!b) return fail(HMJ_ERR_ARG, "null batch%s");
  CK(cudaSetDevice(b->device));
  switch (stage) {
    case HMJ_STAGE_SMOOTH: hmj_launch_smooth(b->tab_dev, b->buf, b->task_dev, 0, nullptr, nullptr, 0, 0, 0, b->stream); break;
    case HMJ_STAGE_COLLIDE: hmj_launch_collide(b->tab_dev, b->buf, b->stream); break;
    case HMJ_STAGE_SOLVE:
    case HMJ_STAGE_SOLVE_INTEGRATE:
      hmj_launch_factor(b->tab_dev, b->buf, b->stream, &b->solve_ctx);
      hmj_launch_solve(b->tab_dev, b->buf, stage == HMJ_STAGE_SOLVE_INTEGRATE ? 1 : 0, b->stream, &b->solve_ctx);
      break;
    default: return fail(HMJ_ERR_ARG, "unknown stage%s");
  }
  b->launches += 1; b->host_launches += 1;
  CK(cudaStreamSynchronize(b->stream));
  CK(cudaGetLastError());
  return HMJ_OK;
}

extern "C" int hmj_batch_get_contacts(hmj_batch_t* b, int32_t* ncon, hmj_contact_t* contacts) {
  if (!b) return fail(HMJ_ERR_ARG, "null batch%s");
  CK(cudaSetDevice(b->device));
  int rc;
  if ((rc = d2h(b, ncon, b->buf.ncon, (size_t)b->n * 4)) ||
      (rc = d2h(b, contacts, b->buf.contact, (size_t)b->n * HMJ_MAXCON * sizeof(hmj_contact_t))))
    return rc;
  CK(cudaStreamSynchronize(b->stream));
  return HMJ_OK;
}

extern "C" void* hmj_batch_device_ptr(hmj_batch_t* b, const char* name) {
  if (!b || !name) return nullptr;
  auto it = b->arrays.find(name);
  return it == b->arrays.end() ? nullptr : it->second.ptr;
}

extern "C" int hmj_batch_get_array(hmj_batch_t* b, const char* name, void* out, size_t bytes) {
  if (!b || !name || !out) return fail(HMJ_ERR_ARG, "bad argument to get_array%s");
  auto it = b->arrays.find(name);
  if (it == b->arrays.end()) return fail(HMJ_ERR_ARG, "unknown array '%s'", name);
  if (bytes != it->second.bytes) return fail(HMJ_ERR_ARG, "size mismatch for array '%s'", name);
  CK(cudaSetDevice(b->device));
  CK(cudaMemcpyAsync(out, it->second.ptr, bytes, cudaMemcpyDeviceToHost, b->stream));
  CK(cudaStreamSynchronize(b->stream));
  return HMJ_OK;
}

extern "C" int hmj_batch_set_array(hmj_batch_t* b, const char* name, const void* in, size_t bytes) {
  if (!b || !name || !in) return fail(HMJ_ERR_ARG, "bad argument to set_array%s");
  auto it = b->arrays.find(name);
  if (it == b->arrays.end()) return fail(HMJ_ERR_ARG, "unknown array '%s'", name);
  if (bytes != it->second.bytes) return fail(HMJ_ERR_ARG, "size mismatch for array '%s'", name);
  CK(cudaSetDevice(b->device));
  CK(cudaMemcpyAsync(it->second.ptr, in, bytes, cudaMemcpyHostToDevice, b->stream));
  CK(cudaStreamSynchronize(b->stream));
  return HMJ_OK;
}

// ------------------------------------------------------------------ CPUEnv epilogue
extern "C" int hmj_env_obs_reward_dev(hmj_batch_t* b, const double* prev_action_dev, const double* latest_action_dev,
                                      const double* noise_dev, double max_sim_time, double* obs_dev, double* reward_dev,
                                      int32_t* terminated_dev, int32_t* truncated_dev, int32_t* self_collision_dev) {
  if (!b || !prev_action_dev || !latest_action_dev || !obs_dev || !reward_dev || !terminated_dev || !truncated_dev)
    return fail(HMJ_ERR_ARG, "bad argument to env_obs_reward_dev%s");
  CK(cudaSetDevice(b->device));
  hmj_launch_env_obs_reward(b->tab_dev, b->buf, prev_action_dev, latest_action_dev, noise_dev, max_sim_time, obs_dev, reward_dev,
                            terminated_dev, truncated_dev, self_collision_dev, b->stream);
  b->launches += 1; b->host_launches += 1;
  CK(cudaGetLastError());
  return HMJ_OK;
}

extern "C" int hmj_env_obs_reward(hmj_batch_t* b, const double* prev_action, const double* latest_action, const double* noise,
                                  double max_sim_time, double* obs, double* reward, int32_t* terminated, int32_t* truncated,
                                  int32_t* self_collision) {
  if (!b || !prev_action || !latest_action || !obs || !reward || !terminated || !truncated)
    return fail(HMJ_ERR_ARG, "bad argument to env_obs_reward%s");
  CK(cudaSetDevice(b->device));
  const size_t N = b->n;
  // staging layout (doubles): prev 12N | latest 12N | noise 31N | obs 30N | reward N | then int32: term N, trunc N, selfc N
  const size_t nd = N * (12 + 12 + 31 + 30 + 1);
  const size_t bytes = nd * 8 + 3 * N * 4;
  if (b->env_stage_cap < bytes) {
    void* p = nullptr;
    CK(cudaMalloc(&p, bytes));
    b->allocs.push_back(p);
    b->env_stage = (double*)p;
    b->env_stage_cap = bytes;
  }
  double* d_prev = b->env_stage;
  double* d_lat = d_prev + 12 * N;
  double* d_noise = d_lat + 12 * N;
  double* d_obs = d_noise + 31 * N;
  double* d_rew = d_obs + 30 * N;
  int32_t* d_term = (int32_t*)(d_rew + N);
  int32_t* d_trunc = d_term + N;
  int32_t* d_self = d_trunc + N;
  int rc;
  if ((rc = h2d(b, d_prev, prev_action, 12 * N * 8)) || (rc = h2d(b, d_lat, latest_action, 12 * N * 8)) ||
      (rc = h2d(b, d_noise, noise, 31 * N * 8)))
    return rc;
  rc = hmj_env_obs_reward_dev(b, d_prev, d_lat, noise ? d_noise : nullptr, max_sim_time, d_obs, d_rew, d_term, d_trunc, d_self);
  if (rc) return rc;
  if ((rc = d2h(b, obs, d_obs, 30 * N * 8)) || (rc = d2h(b, reward, d_rew, N * 8)) || (rc = d2h(b, terminated, d_term, N * 4)) ||
      (rc = d2h(b, truncated, d_trunc, N * 4)) || (rc = d2h(b, self_collision, d_self, N * 4)))
    return rc;
  CK(cudaStreamSynchronize(b->stream));
  return HMJ_OK;
}

// CPUEnv.step as ONE call (cpu_env.py:676-693): host controls / push wrench in, nsteps x mj_step, observation and
// reward epilogue, host results out, one synchronisation.  The per-env staging lives in one device block.
extern "C" int hmj_env_step(hmj_batch_t* b, const double* ctrl, const int32_t* xfrc_body, const double* xfrc,
                            const double* prev_action, const double* latest_action, const double* noise, double max_sim_time,
                            int nsteps, double* obs, double* reward, int32_t* terminated, int32_t* truncated,
                            int32_t* self_collision, double* time_out, double* root_quat_out) {
  if (!b || !prev_action || !latest_action || !obs || !reward || !terminated || !truncated || nsteps < 0)
    return fail(HMJ_ERR_ARG, "bad argument to env_step%s");
  CK(cudaSetDevice(b->device));
  const size_t N = b->n;
  const size_t nd = N * (12 + 12 + 31 + 30 + 1);
  const size_t bytes = nd * 8 + 3 * N * 4;
  if (b->env_stage_cap < bytes) {
    void* p = nullptr;
    CK(cudaMalloc(&p, bytes));
    b->allocs.push_back(p);
    b->env_stage = (double*)p;
    b->env_stage_cap = bytes;
  }
  double* d_prev = b->env_stage;
  double* d_lat = d_prev + 12 * N;
  double* d_noise = d_lat + 12 * N;
  double* d_obs = d_noise + 31 * N;
  double* d_rew = d_obs + 30 * N;
  int32_t* d_term = (int32_t*)(d_rew + N);
  int32_t* d_trunc = d_term + N;
  int32_t* d_self = d_trunc + N;
  int rc;
  if ((rc = h2d(b, b->buf.ctrl, ctrl, N * HMJ_NU * 8)) || (rc = h2d(b, b->buf.xfrc_body, xfrc_body, N * 4)) ||
      (rc = h2d(b, b->buf.xfrc, xfrc, N * 6 * 8)) || (rc = h2d(b, d_prev, prev_action, 12 * N * 8)) ||
      (rc = h2d(b, d_lat, latest_action, 12 * N * 8)) || (rc = h2d(b, d_noise, noise, 31 * N * 8)))
    return rc;
  enqueue_steps(b, 1, nsteps);
  hmj_launch_env_obs_reward(b->tab_dev, b->buf, d_prev, d_lat, noise ? d_noise : nullptr, max_sim_time, d_obs, d_rew, d_term,
                            d_trunc, d_self, b->stream);
  b->launches += 1; b->host_launches += 1;
  if ((rc = d2h(b, obs, d_obs, 30 * N * 8)) || (rc = d2h(b, reward, d_rew, N * 8)) || (rc = d2h(b, terminated, d_term, N * 4)) ||
      (rc = d2h(b, truncated, d_trunc, N * 4)) || (rc = d2h(b, self_collision, d_self, N * 4)) ||
      (rc = d2h(b, time_out, b->buf.time, N * 8)))
    return rc;
  if (root_quat_out)
    CK(cudaMemcpy2DAsync(root_quat_out, 4 * 8, b->buf.sens + SENS_ROOT_XQUAT, SENS_NUM * 8, 4 * 8, N, cudaMemcpyDeviceToHost, b->stream));
  CK(cudaStreamSynchronize(b->stream));
  CK(cudaGetLastError());
  return HMJ_OK;
}

// Pinned host IO block of the fused control step (views for the host language; valid until hmj_batch_destroy).
extern "C" int hmj_env_io(hmj_batch_t* b, double** action, int32_t** mask, double** obs, double** reward, double** time,
                          double** torso_quat, int32_t** terminated, int32_t** truncated, int32_t** self_collision) {
  if (!b || !b->io_host) return fail(HMJ_ERR_ARG, "null batch%s");
  const size_t N = b->n;
  char* in = b->io_host;
  char* out = b->io_host + b->io_in_bytes;
  if (action) *action = (double*)in;
  if (mask) *mask = (int32_t*)(in + N * HMJ_NU * 8);
  double* o = (double*)out;
  if (obs) *obs = o;
  if (reward) *reward = o + N * 30;
  if (time) *time = o + N * 31;
  if (torso_quat) *torso_quat = o + N * 32;
  int32_t* f = (int32_t*)(o + N * 36);
  if (terminated) *terminated = f;
  if (truncated) *truncated = f + N;
  if (self_collision) *self_collision = f + 2 * N;
  return HMJ_OK;
}

// CPUEnv.step (cpu_env.py:676-693) with ALL per-step bookkeeping on the device: action noise + delay FIFO, push
// scheduler, nsteps x mj_step, observation / reward, observation delay FIFOs.  Inputs are read from / results
// written to the pinned IO block of hmj_env_io: one H2D copy, one D2H copy, one synchronisation per control step.
// has_action == 0 replays data.ctrl (CPUEnv.step(None)); has_mask != 0 restricts the bookkeeping to mask[env] != 0
// (physics is lock-step for everyone; CPUEnv.reset uses this for its first-step rejection loop).
extern "C" int hmj_env_step_fused_from(hmj_batch_t* b, const double* action_host, int has_mask, double rf, double max_sim_time,
                                       unsigned long long seed, int nsteps) {
  if (!b) return fail(HMJ_ERR_ARG, "bad argument to env_step_fused%s");  // nsteps < 0: observation only (no prologue, no physics)
  CK(cudaSetDevice(b->device));
  const size_t N = b->n;
  const int has_action = action_host != nullptr;
  double* d_action = (double*)b->io_dev;
  int32_t* d_mask = (int32_t*)(b->io_dev + N * HMJ_NU * 8);
  double* d_out = (double*)(b->io_dev + b->io_in_bytes);
  double* d_obs = d_out;
  double* d_rew = d_out + N * 30;
  double* d_time = d_out + N * 31;
  double* d_quat = d_out + N * 32;
  int32_t* d_flags = (int32_t*)(d_out + N * 36);
  // the common call (one physics step, every env) replays the control-step graph; there the action upload goes to the copy
  // stream and only the graph's prologue kernels wait for it (ctrl_mode 2)
  const bool graph_path = b->ctrl_mode && !has_mask && nsteps == 1 && !b->profile && b->warmed && b->graph_mode != 0;
  const bool side_copy = graph_path && has_action && b->ctrl_mode == 2;
  if (side_copy) {
    CK(cudaMemcpyAsync(d_action, action_host, N * HMJ_NU * 8, cudaMemcpyHostToDevice, b->copy_stream));
    CK(cudaEventRecord(b->h2d_event, b->copy_stream));
  } else if (has_action && has_mask && (const char*)action_host == b->io_host) {
    CK(cudaMemcpyAsync(b->io_dev, b->io_host, N * HMJ_NU * 8 + N * 4, cudaMemcpyHostToDevice, b->stream));  // action + mask in one copy
  } else {
    if (has_action) CK(cudaMemcpyAsync(d_action, action_host, N * HMJ_NU * 8, cudaMemcpyHostToDevice, b->stream));
    if (has_mask) CK(cudaMemcpyAsync(d_mask, b->io_host + N * HMJ_NU * 8, N * 4, cudaMemcpyHostToDevice, b->stream));
  }
  const int32_t* mask = has_mask ? d_mask : nullptr;
  const bool pool = b->pool_on && nsteps >= 0 && !has_mask;
  // the graph's result copies are inside unless the episode pool still has to look at (and patch) the results
  bool copied = false, done = false;
  if (graph_path) {
    done = launch_control_graph(b, has_action, rf, max_sim_time, seed, !pool, side_copy);
    if (done) copied = !pool;
    else {
      b->ctrl_mode = 0;   // capture unavailable: whole-batch kernels around the step graph from now on
      if (side_copy) CK(cudaStreamWaitEvent(b->stream, b->h2d_event, 0));   // the kernels below need the upload
    }
  }
  if (!done) {
    if (nsteps >= 0) hmj_launch_env_pre(b->buf, b->env, has_action ? d_action : nullptr, mask, rf, seed, b->stream);
    enqueue_steps(b, 1, nsteps);
    hmj_launch_env_obs_reward(b->tab_dev, b->buf, b->env.prev_action, b->env.latest_action, b->env.obs_noise, max_sim_time, d_obs,
                              d_rew, d_flags, d_flags + N, d_flags + 2 * N, b->stream);
    hmj_launch_env_post(b->buf, b->env, mask, d_obs, d_time, d_quat, b->stream);
    b->launches += 3; b->host_launches += 3;
  }
  if (pool) {
    int rc = pool_step(b, d_flags, d_flags + N, false);
    if (rc) return rc;
  }
  if (!copied)
    CK(cudaMemcpyAsync(b->io_host + b->io_in_bytes, d_out, pool ? b->io_out_bytes : b->io_ext_offset, cudaMemcpyDeviceToHost, b->stream));
  CK(cudaStreamSynchronize(b->stream));
  CK(cudaGetLastError());
  if (pool) {
    int32_t* counts = io_counts_host(b);
    b->pool_need_round = counts[6] > 0;
    if (counts[0] > HMJ_POOL_MAXTERM) {   // more endings than the result block carries: fetch exactly those rows
      const size_t cnt = (size_t)counts[0];
      CK(cudaMemcpyAsync(b->term_host, b->pool.term_obs_full, cnt * 30 * 8, cudaMemcpyDeviceToHost, b->stream));
      CK(cudaMemcpyAsync(b->term_host + N * 30 * 8, b->pool.ended_ids_full, cnt * 4, cudaMemcpyDeviceToHost, b->stream));
      CK(cudaStreamSynchronize(b->stream));
    }
    if (counts[2] > 0) {   // more episodes ended than the pool held: serve the rest now, then refresh their result rows
      const int32_t needed = counts[0], ended_now = counts[4];
      int rc = pool_drain(b);
      if (rc) return rc;
      CK(cudaMemcpyAsync(b->io_host + b->io_in_bytes, d_out, b->io_ext_offset, cudaMemcpyDeviceToHost, b->stream));
      CK(cudaStreamSynchronize(b->stream));
      counts[0] = needed; counts[4] = ended_now;
    }
  }
  return HMJ_OK;
}

extern "C" int hmj_env_step_fused(hmj_batch_t* b, int has_action, int has_mask, double rf, double max_sim_time,
                                  unsigned long long seed, int nsteps) {
  if (!b) return fail(HMJ_ERR_ARG, "bad argument to env_step_fused%s");
  return hmj_env_step_fused_from(b, has_action ? (const double*)b->io_host : nullptr, has_mask, rf, max_sim_time, seed, nsteps);
}

extern "C" int hmj_pinned_alloc(size_t bytes, void** out) {
  if (!out || bytes == 0) return fail(HMJ_ERR_ARG, "bad argument to pinned_alloc%s");
  CK(cudaHostAlloc(out, bytes, cudaHostAllocDefault));
  return HMJ_OK;
}

extern "C" int hmj_pinned_free(void* p) {
  if (p) CK(cudaFreeHost(p));
  return HMJ_OK;
}

// ------------------------------------------------------------------ episode pool: CPUEnv.reset / DummyVecEnv auto-reset on the device
// (kernels and the design: hmj_reset.cu)
#define HMJ_POOL_JOIN_LAG 2   /* a round launched behind step t is joined (and its rows promoted) at step t + 2 */

// One generation round on the spare stream: draw episodes into the EMPTY rows, CPUEnv.reset's probe step() with action
// None, the second _get_obs() that reset returns, the rejection rule.
static void enqueue_round_chain(hmj_batch_t* b) {
  cudaStream_t s = b->spare_stream;
  const hmj_buffers_t& V = b->spare;
  const hmj_envstate_t& E = b->env_spare;
  hmj_launch_env_reset(b->tab_dev, V, E, b->pool, b->pool_rf, b->pool_timestep, b->pool_seed, b->n, s);
  hmj_launch_env_pre(V, E, nullptr, b->pool.active, b->pool_rf, b->pool_seed, s, b->n);
  hmj_launch_smooth(b->tab_dev, V, b->task_dev, 0, nullptr, nullptr, 0, 0, 0, s, &b->spare_ctx);
  hmj_launch_factor(b->tab_dev, V, s, &b->spare_ctx);
  hmj_launch_collide(b->tab_dev, V, s, false);
  hmj_launch_solve(b->tab_dev, V, 1, s, &b->spare_ctx, false);
  for (int k = 0; k < 2; k++) {
    hmj_launch_env_obs_reward(b->tab_dev, V, E.prev_action, E.latest_action, E.obs_noise, b->pool_max_sim_time, b->pool.obs,
                              b->pool.reward, b->pool.term, b->pool.trunc, b->pool.selfc, s);
    hmj_launch_env_post(V, E, b->pool.active, b->pool.obs, nullptr, nullptr, s);
  }
  hmj_launch_pool_validate(V, b->pool, s);
}

// Launch a round behind everything the main stream has enqueued so far (the rows emptied by this step's copies are seen).
static int launch_round(hmj_batch_t* b) {
  CK(cudaEventRecord(b->ev_main, b->stream));
  CK(cudaStreamWaitEvent(b->spare_stream, b->ev_main, 0));
  b->launches += 17;
  if (b->graph_mode == 0 || !b->warmed) {
    enqueue_round_chain(b);
    b->warmed = true;
    b->host_launches += 17;
  } else {
    if (!b->round_graph) {
      cudaGraph_t g = nullptr;
      CK(cudaStreamBeginCapture(b->spare_stream, cudaStreamCaptureModeThreadLocal));
      enqueue_round_chain(b);
      cudaError_t e = cudaStreamEndCapture(b->spare_stream, &g);
      if (e == cudaSuccess && g) e = cudaGraphInstantiate(&b->round_graph, g, 0);
      if (g) cudaGraphDestroy(g);
      if (e != cudaSuccess || !b->round_graph) return fail(HMJ_ERR_CUDA, "pool round graph capture failed: %s", cudaGetErrorString(e));
    }
    CK(cudaGraphLaunch(b->round_graph, b->spare_stream));
    b->host_launches += 1;
  }
  CK(cudaEventRecord(b->round_done, b->spare_stream));
  b->pool_rounds++;
  b->round_inflight = true;
  b->round_start = b->step_idx;
  return HMJ_OK;
}

static double* io_out_dev(hmj_batch_t* b) { return (double*)(b->io_dev + b->io_in_bytes); }
static int32_t* io_counts_host(hmj_batch_t* b) { return (int32_t*)(b->io_host + b->io_in_bytes + b->io_ext_offset); }

// Inside a control step, after the epilogue: join a due round, hand ready episodes to the envs that ended, start the
// next round.  Every decision uses host-known numbers only (step index, the previous step's counts): deterministic.
static int pool_step(hmj_batch_t* b, const int32_t* term, const int32_t* trunc, bool drain) {
  const size_t N = b->n;
  double* d_out = io_out_dev(b);
  int promote = 0;
  if (b->round_inflight && (drain || b->step_idx - b->round_start >= HMJ_POOL_JOIN_LAG)) {
    CK(cudaStreamWaitEvent(b->stream, b->round_done, 0));
    b->round_inflight = false;
    promote = 1;
  }
  const bool launching = !b->round_inflight && (b->pool_need_round || drain);
  hmj_launch_autoreset(b->buf, b->env, b->spare, b->env_spare, b->pool, term, trunc, d_out, d_out + N * 31, d_out + N * 32,
                       promote | (drain ? 2 : 0) | (launching ? 4 : 0), b->stream);
  b->launches += 2; b->host_launches += 2;
  if (launching) {
    int rc = launch_round(b);
    if (rc) return rc;
  }
  if (!drain) b->step_idx++;
  return HMJ_OK;
}

// Serve every env that still waits for an episode (a burst larger than the pool: the first reset of a run, the
// curriculum's reset-everything, a mass time-out): generation rounds back to back, blocking.
static int pool_drain(hmj_batch_t* b) {
  int32_t* counts = io_counts_host(b);
  b->pool_drains++;
  for (int it = 0; it < 100000; it++) {
    b->pool_drain_iters++;
    int rc = pool_step(b, nullptr, nullptr, true);
    if (rc) return rc;
    CK(cudaMemcpyAsync(counts, b->pool.counts, 8 * sizeof(int32_t), cudaMemcpyDeviceToHost, b->stream));
    CK(cudaStreamSynchronize(b->stream));
    if (counts[2] == 0) {
      b->pool_need_round = counts[6] > 0;
      return HMJ_OK;
    }
  }
  return fail(HMJ_ERR_CUDA, "episode pool: environments still waiting after 100000 generation rounds%s");
}

/* CPUEnv.reset + DummyVecEnv auto-reset on the device.  enable != 0: hmj_env_step_fused[_from] (without mask) replaces
 * every env whose episode ended (terminated | truncated) with a fresh pre-validated episode inside the same call. */
extern "C" int hmj_env_pool_config(hmj_batch_t* b, int enable, double rf, double max_sim_time, unsigned long long seed,
                                   double timestep) {
  if (!b) return fail(HMJ_ERR_ARG, "null batch%s");
  if (b->n_spare <= 0) return fail(HMJ_ERR_ARG, "this batch was created without spare rows (HMJ_SPARE=0)%s");
  CK(cudaSetDevice(b->device));
  CK(cudaStreamSynchronize(b->stream));
  CK(cudaStreamSynchronize(b->spare_stream));
  const bool changed = rf != b->pool_rf || max_sim_time != b->pool_max_sim_time || seed != b->pool_seed || timestep != b->pool_timestep;
  if (changed) {
    if (b->round_graph) { cudaGraphExecDestroy(b->round_graph); b->round_graph = nullptr; }
    b->pool_rf = rf; b->pool_max_sim_time = max_sim_time; b->pool_seed = seed; b->pool_timestep = timestep;
    // episodes drawn under the old settings are discarded (the curriculum raises rf and resets everything)
    CK(cudaMemsetAsync(b->pool.state, 0, (size_t)b->n_spare * 4, b->stream));
    CK(cudaMemsetAsync(b->pool.active, 0, (size_t)b->n_spare * 4, b->stream));
    CK(cudaStreamSynchronize(b->stream));
    b->round_inflight = false;
    b->pool_need_round = true;
  }
  b->pool_on = enable != 0;
  return HMJ_OK;
}

/* CPUEnv.reset for the listed envs (ids == NULL: all): blocking.  The first observation of each new episode lands in
 * the IO block's obs / time / torso_quat rows of those envs (hmj_env_io); other rows are untouched. */
extern "C" int hmj_env_reset(hmj_batch_t* b, const int32_t* ids, int count) {
  if (!b || (ids && count < 0)) return fail(HMJ_ERR_ARG, "bad argument to env_reset%s");
  if (b->n_spare <= 0) return fail(HMJ_ERR_ARG, "this batch was created without spare rows (HMJ_SPARE=0)%s");
  CK(cudaSetDevice(b->device));
  for (int i = 0; ids && i < count; i++)
    if (ids[i] < 0 || ids[i] >= b->n) return fail(HMJ_ERR_ARG, "env id out of range%s");
  const int32_t* ids_dev = nullptr;
  if (ids) {
    if (count == 0) return HMJ_OK;
    CK(cudaMemcpyAsync(b->pool.served_list, ids, (size_t)count * 4, cudaMemcpyHostToDevice, b->stream));   // staging: rewritten by the next assign
    ids_dev = b->pool.served_list;
  }
  hmj_launch_pool_mark_pending(b->pool, ids_dev, count, b->n, b->stream);
  b->launches += 1; b->host_launches += 1;
  int rc = pool_drain(b);
  if (rc) return rc;
  CK(cudaMemcpyAsync(b->io_host + b->io_in_bytes, io_out_dev(b), b->io_out_bytes, cudaMemcpyDeviceToHost, b->stream));
  CK(cudaStreamSynchronize(b->stream));
  CK(cudaGetLastError());
  return HMJ_OK;
}

/* Views of the pool's per-step result block (pinned): counts[8] = {needed, served, still waiting (0 after the call),
 * ready rows left, ended this step, S, empty rows, 0}; ended_ids[64] / term_obs[64][30] = the first 64 envs that ended
 * this step in env order with their terminal observations (more: named arrays "env_ended_ids_compact" /
 * "env_terminal_obs_compact", N rows). */
extern "C" int hmj_env_reset_io(hmj_batch_t* b, int32_t** counts, int32_t** ended_ids, double** term_obs) {
  if (!b || !b->io_host) return fail(HMJ_ERR_ARG, "null batch%s");
  char* ext = b->io_host + b->io_in_bytes + b->io_ext_offset;
  if (counts) *counts = (int32_t*)ext;
  if (ended_ids) *ended_ids = (int32_t*)ext + 8;
  if (term_obs) *term_obs = (double*)(ext + 8 * 4 + HMJ_POOL_MAXTERM * 4);
  return HMJ_OK;
}
/* The envs that ended in the last pooled control step (env order) and their terminal observations: `count` entries
 * (pinned host memory, valid until the next step). */
extern "C" int hmj_env_terminal(hmj_batch_t* b, int32_t** ids, double** obs, int32_t* count) {
  if (!b || !b->io_host || !b->term_host) return fail(HMJ_ERR_ARG, "no episode pool%s");
  const int32_t* counts = io_counts_host(b);
  char* ext = b->io_host + b->io_in_bytes + b->io_ext_offset;
  const bool big = counts[0] > HMJ_POOL_MAXTERM;
  if (ids) *ids = big ? (int32_t*)(b->term_host + (size_t)b->n * 30 * 8) : (int32_t*)ext + 8;
  if (obs) *obs = big ? (double*)b->term_host : (double*)(ext + 8 * 4 + HMJ_POOL_MAXTERM * 4);
  if (count) *count = counts[0];
  return HMJ_OK;
}
extern "C" int hmj_batch_spare_rows(const hmj_batch_t* b) { return b ? b->n_spare : 0; }
/* host-side counters of the pool: generation rounds launched, blocking drains, drain iterations, control steps */
extern "C" int hmj_env_pool_stats(const hmj_batch_t* b, long long* out4) {
  if (!b || !out4) return fail(HMJ_ERR_ARG, "null argument%s");
  out4[0] = b->pool_rounds; out4[1] = b->pool_drains; out4[2] = b->pool_drain_iters; out4[3] = b->step_idx;
  return HMJ_OK;
}

// ------------------------------------------------------------------ planner
extern "C" int hmj_plan_set_task(hmj_batch_t* b, int nterm, const int32_t* dim, const int32_t* norm, const double* weight,
                                 const double* normprm, double risk, double height_goal) {
  if (!b || nterm < 0 || nterm > HMJ_NTERM || !dim || !norm || !weight || !normprm) return fail(HMJ_ERR_ARG, "bad argument to set_task%s");
  int tot = 0;
  for (int k = 0; k < nterm; k++) tot += dim[k];
  if (tot != HMJ_NRESIDUAL) return fail(HMJ_ERR_ARG, "cost terms must cover the 28 residuals%s");
  CK(cudaSetDevice(b->device));
  hmj_task_t task;
  memset(&task, 0, sizeof(task));
  task.nterm = nterm;
  for (int k = 0; k < nterm; k++) {
    task.dim[k] = dim[k]; task.norm[k] = norm[k]; task.weight[k] = weight[k];
    for (int q = 0; q < HMJ_MAXNORMPRM; q++) task.normprm[k][q] = normprm[k * HMJ_MAXNORMPRM + q];
  }
  task.risk = risk;
  task.height_goal = height_goal;
  CK(cudaMemcpyAsync(b->task_dev, &task, sizeof(task), cudaMemcpyHostToDevice, b->stream));
  CK(cudaStreamSynchronize(b->stream));
  return HMJ_OK;
}

static int ensure_record(hmj_batch_t* b, int horizon) {
  if (b->rec_cap >= horizon) { b->rec.horizon = horizon; return HMJ_OK; }
  if (b->rec_cap > 0) {   // the horizon grew: release the superseded buffers (no kernel uses them: the stream is drained)
    CK(cudaStreamSynchronize(b->stream));
    double* old[6] = {b->rec.states, b->rec.actions, b->rec.times, b->rec.residual, b->rec.costs, b->rec.trace};
    for (double* p : old) {
      for (size_t k = 0; k < b->allocs.size(); k++)
        if (b->allocs[k] == (void*)p) { b->allocs.erase(b->allocs.begin() + k); break; }
      cudaFree(p);
    }
    b->rec = hmj_record_t{};
    b->rec_cap = 0;
  }
  const size_t HN = (size_t)horizon * b->n;
  int rc;
  if ((rc = alloc_array(b, nullptr, &b->rec.states, HN * HMJ_NSTATE)) || (rc = alloc_array(b, nullptr, &b->rec.actions, HN * HMJ_NU)) ||
      (rc = alloc_array(b, nullptr, &b->rec.times, HN)) || (rc = alloc_array(b, nullptr, &b->rec.residual, HN * HMJ_NRESIDUAL)) ||
      (rc = alloc_array(b, nullptr, &b->rec.costs, HN)) || (rc = alloc_array(b, nullptr, &b->rec.trace, HN * 3)))
    return rc;
  b->rec_cap = horizon;
  b->rec.horizon = horizon;
  return HMJ_OK;
}

// The launches of one batch of rollouts (Trajectory::NoisyRollout for every sample, trajectory.cc:100-210), no host data:
// init from the state block, H-1 x [policy + smooth + residual/cost -> factor || collide -> solve + Euler], the final
// mj_forward (sensors/residual only; ctrl keeps the last action, trajectory.cc:190-206), returns.
struct RolloutNoise { double rate = 0, scale = 0; const double* noise_dev = nullptr; int first_global = 0; bool on = false; };

static void enqueue_rollout_chain(hmj_batch_t* b, const double* values_dev, int nnode, int interpolation, int horizon, int record,
                                  double* returns_dev, int32_t* failure_dev, const RolloutNoise& ou = RolloutNoise()) {
  hmj_launch_plan_init(b->buf, b->state_dev, b->stream);
  const int xf = ou.on ? HMJ_PLAN_XFRC : 0;
  for (int t = 0; t < horizon - 1; t++) {
    if (ou.on) hmj_launch_plan_ou(b->buf, t, horizon - 1, ou.rate, ou.scale, ou.noise_dev, b->seed_iter_dev, ou.first_global, b->stream);
    hmj_launch_smooth(b->tab_dev, b->buf, b->task_dev, HMJ_PLAN_COST | HMJ_PLAN_POLICY | xf, b->spline_times_dev, values_dev, nnode,
                      interpolation, 1, b->stream, &b->solve_ctx);
    // the knot is recorded behind k_dyn (residual, cost) on the side stream; k_solve joins that stream before it moves the state
    if (record) hmj_launch_plan_record(b->buf, b->rec, t, b->solve_ctx.side[0]);
    hmj_launch_factor(b->tab_dev, b->buf, b->stream, &b->solve_ctx);
    hmj_launch_collide(b->tab_dev, b->buf, b->stream, false);
    hmj_launch_solve(b->tab_dev, b->buf, 1, b->stream, &b->solve_ctx, false);
  }
  hmj_launch_smooth(b->tab_dev, b->buf, b->task_dev, HMJ_PLAN_COST | xf, nullptr, nullptr, 0, 0, 0, b->stream);
  if (record) hmj_launch_plan_record(b->buf, b->rec, horizon - 1, b->stream);
  hmj_launch_plan_finish(b->buf, horizon, returns_dev, failure_dev, b->stream);
}

// One batch of rollouts = ONE graph launch.  Everything that changes between planning iterations (state, time, spline
// nodes, candidate values) lives in device buffers at fixed addresses, so the chain is captured once per
// (nnode, interpolation, horizon, record, buffers) and replayed; a horizon of 100 is ~800 kernel nodes that would
// otherwise be ~1500 API calls per iteration.
static int launch_rollouts(hmj_batch_t* b, const double* values_dev, int nnode, int interpolation, int horizon, int record,
                           double* returns_dev, int32_t* failure_dev) {
  b->launches += 1 + (horizon - 1) * (10 + (record ? 1 : 0)) + 2 + (record ? 1 : 0) + 1;
  if (b->graph_mode == 0 || !b->warmed) {   // first use of the launchers performs their one-time attribute queries
    enqueue_rollout_chain(b, values_dev, nnode, interpolation, horizon, record, returns_dev, failure_dev);
    b->warmed = true;
    b->host_launches += 1 + (horizon - 1) * (10 + (record ? 1 : 0)) + 2 + (record ? 1 : 0) + 1;
    CK(cudaGetLastError());
    return HMJ_OK;
  }
  const hmj_batch_::RolloutKey key{nnode, interpolation, horizon, record, (const void*)values_dev, (void*)returns_dev,
                                   (void*)failure_dev, (void*)b->rec.states};
  auto it = b->rollout_graphs.find(key);
  if (it == b->rollout_graphs.end()) {
    if (b->rollout_graphs.size() >= 32) {   // callers that sweep horizons: drop the cache rather than grow without bound
      for (auto& kv : b->rollout_graphs) cudaGraphExecDestroy(kv.second);
      b->rollout_graphs.clear();
    }
    cudaGraph_t g = nullptr;
    cudaGraphExec_t ge = nullptr;
    CK(cudaStreamBeginCapture(b->stream, cudaStreamCaptureModeThreadLocal));
    enqueue_rollout_chain(b, values_dev, nnode, interpolation, horizon, record, returns_dev, failure_dev);
    cudaError_t e = cudaStreamEndCapture(b->stream, &g);
    if (e == cudaSuccess && g) e = cudaGraphInstantiate(&ge, g, 0);
    if (g) cudaGraphDestroy(g);
    if (e != cudaSuccess || !ge) return fail(HMJ_ERR_CUDA, "rollout graph capture failed: %s", cudaGetErrorString(e));
    it = b->rollout_graphs.emplace(key, ge).first;
  }
  CK(cudaGraphLaunch(it->second, b->stream));
  b->host_launches += 1;
  return HMJ_OK;
}

static int upload_plan_inputs(hmj_batch_t* b, const double* state, double time, const double* spline_times, int nnode, int horizon) {
  if (nnode < 0 || nnode > 64 || horizon < 1 || horizon > 512) return fail(HMJ_ERR_ARG, "bad nnode/horizon%s");  // kMaxTrajectoryHorizon
  double* st = b->plan_stage;   // pinned: state | time | spline times
  memcpy(st, state, HMJ_NSTATE * 8);
  st[HMJ_NSTATE] = time;
  CK(cudaMemcpyAsync(b->state_dev, st, (HMJ_NSTATE + 1) * 8, cudaMemcpyHostToDevice, b->stream));
  if (nnode) {
    memcpy(st + 64, spline_times, (size_t)nnode * 8);
    CK(cudaMemcpyAsync(b->spline_times_dev, st + 64, (size_t)nnode * 8, cudaMemcpyHostToDevice, b->stream));
  }
  return HMJ_OK;
}

static int enqueue_rollouts(hmj_batch_t* b, const double* state, double time, const double* spline_times,
                            const double* spline_values_dev, int nnode, int interpolation, int horizon, int record,
                            double* returns_dev, int32_t* failure_dev) {
  // the pinned staging block is reused by the next call: wait for the previous copies out of it
  CK(cudaStreamSynchronize(b->stream));
  int rc = upload_plan_inputs(b, state, time, spline_times, nnode, horizon);
  if (rc) return rc;
  if (record && (rc = ensure_record(b, horizon))) return rc;
  return launch_rollouts(b, spline_values_dev, nnode, interpolation, horizon, record, returns_dev, failure_dev);
}

static int ensure_values(hmj_batch_t* b, int nnode) {
  const size_t nvals = (size_t)b->n * nnode * HMJ_NU;
  if (b->spline_values_cap >= nvals) return HMJ_OK;
  CK(cudaStreamSynchronize(b->stream));
  for (void** pp : {(void**)&b->spline_values_dev, (void**)&b->noise_dev}) {
    if (*pp) {
      for (size_t k = 0; k < b->allocs.size(); k++)
        if (b->allocs[k] == *pp) { b->allocs.erase(b->allocs.begin() + k); break; }
      cudaFree(*pp);
      *pp = nullptr;
    }
    void* p = nullptr;
    CK(cudaMalloc(&p, nvals * 8));
    b->allocs.push_back(p);
    *pp = p;
  }
  b->spline_values_cap = nvals;
  return HMJ_OK;
}

extern "C" int hmj_plan_rollouts_dev(hmj_batch_t* b, const double* state, double time, const double* spline_times,
                                     const double* spline_values_dev, int nnode, int interpolation, int horizon, double* returns_dev) {
  if (!b || !state || !spline_times || !spline_values_dev) return fail(HMJ_ERR_ARG, "bad argument to plan_rollouts_dev%s");
  CK(cudaSetDevice(b->device));
  return enqueue_rollouts(b, state, time, spline_times, spline_values_dev, nnode, interpolation, horizon, 0,
                          returns_dev ? returns_dev : b->returns_dev, b->failure_dev);
}

extern "C" int hmj_plan_rollouts(hmj_batch_t* b, const double* state, double time, const double* spline_times,
                                 const double* spline_values, int nnode, int interpolation, int horizon, int record,
                                 double* total_return, int32_t* failure) {
  if (!b || !state || !spline_times || !spline_values || !total_return) return fail(HMJ_ERR_ARG, "bad argument to plan_rollouts%s");
  CK(cudaSetDevice(b->device));
  int rc = ensure_values(b, nnode);
  if (rc) return rc;
  const size_t nvals = (size_t)b->n * nnode * HMJ_NU;
  CK(cudaMemcpyAsync(b->spline_values_dev, spline_values, nvals * 8, cudaMemcpyHostToDevice, b->stream));
  rc = enqueue_rollouts(b, state, time, spline_times, b->spline_values_dev, nnode, interpolation, horizon, record,
                        b->returns_dev, b->failure_dev);
  if (rc) return rc;
  if ((rc = d2h(b, total_return, b->returns_dev, (size_t)b->n * 8)) || (rc = d2h(b, failure, b->failure_dev, (size_t)b->n * 4))) return rc;
  CK(cudaStreamSynchronize(b->stream));
  CK(cudaGetLastError());
  return HMJ_OK;
}

/* Trajectory::NoisyRollout for every sample (mjpc/trajectory.cc:100-210 with xfrc_std > 0; RobustPlanner::OptimizePolicy,
 * mjpc/planners/robust/robust_planner.cc:96-151): the rollouts of hmj_plan_rollouts with an Ornstein-Uhlenbeck wrench on
 * every body, x <- rate x + N(0, scale), rate = exp(-dt / xfrc_rate), scale = xfrc_std sqrt(1 - rate^2), refreshed before every
 * step.  xfrc_noise [N][H-1][90] = unit normals injected by the caller (NULL: drawn on the device from (seed, call counter,
 * first_global + sample)).  dt is the timestep of env 0's parameters. */
extern "C" int hmj_plan_rollouts_noisy(hmj_batch_t* b, const double* state, double time, const double* spline_times,
                                       const double* spline_values, int nnode, int interpolation, int horizon, double xfrc_std,
                                       double xfrc_rate, const double* xfrc_noise, unsigned long long seed, int first_global,
                                       double* total_return, int32_t* failure) {
  if (!b || !state || !spline_times || !spline_values || !total_return || !(xfrc_rate > 0) || xfrc_std < 0)
    return fail(HMJ_ERR_ARG, "bad argument to plan_rollouts_noisy%s");
  CK(cudaSetDevice(b->device));
  int rc = ensure_values(b, nnode);
  if (rc) return rc;
  const size_t nvals = (size_t)b->n * nnode * HMJ_NU;
  CK(cudaMemcpyAsync(b->spline_values_dev, spline_values, nvals * 8, cudaMemcpyHostToDevice, b->stream));
  CK(cudaStreamSynchronize(b->stream));
  if ((rc = upload_plan_inputs(b, state, time, spline_times, nnode, horizon))) return rc;
  hmj_params_t p0;
  CK(cudaMemcpyAsync(&p0, b->buf.params, sizeof(p0), cudaMemcpyDeviceToHost, b->stream));
  const unsigned long long si[2] = {seed, b->ou_calls++};
  CK(cudaMemcpyAsync(b->seed_iter_dev, si, sizeof(si), cudaMemcpyHostToDevice, b->stream));
  const size_t nn = (size_t)b->n * (horizon > 1 ? horizon - 1 : 0) * HMJ_NBODY * 6;
  if (xfrc_noise && nn) {
    if (b->ou_noise_cap < nn) {
      CK(cudaStreamSynchronize(b->stream));
      if (b->ou_noise_dev) {
        for (size_t k = 0; k < b->allocs.size(); k++)
          if (b->allocs[k] == (void*)b->ou_noise_dev) { b->allocs.erase(b->allocs.begin() + k); break; }
        cudaFree(b->ou_noise_dev);
        b->ou_noise_dev = nullptr;
      }
      void* p = nullptr;
      CK(cudaMalloc(&p, nn * 8));
      b->allocs.push_back(p);
      b->ou_noise_dev = (double*)p;
      b->ou_noise_cap = nn;
    }
    CK(cudaMemcpyAsync(b->ou_noise_dev, xfrc_noise, nn * 8, cudaMemcpyHostToDevice, b->stream));
  }
  CK(cudaStreamSynchronize(b->stream));
  RolloutNoise ou;
  ou.on = true;
  ou.rate = exp(-p0.timestep / xfrc_rate);
  ou.scale = xfrc_std * sqrt(1.0 - ou.rate * ou.rate);
  ou.noise_dev = xfrc_noise ? b->ou_noise_dev : nullptr;
  ou.first_global = first_global;
  enqueue_rollout_chain(b, b->spline_values_dev, nnode, interpolation, horizon, 0, b->returns_dev, b->failure_dev, ou);
  b->warmed = true;
  const int nl = 1 + (horizon - 1) * 11 + 2 + 1;
  b->launches += nl; b->host_launches += nl;
  if ((rc = d2h(b, total_return, b->returns_dev, (size_t)b->n * 8)) || (rc = d2h(b, failure, b->failure_dev, (size_t)b->n * 4))) return rc;
  CK(cudaStreamSynchronize(b->stream));
  CK(cudaGetLastError());
  return HMJ_OK;
}

// ---- one predictive-sampling iteration with the candidates built on the device (SamplingPlanner::OptimizePolicy's
// device half, planner.cc:190-208, 313-380): candidates = clamp(nominal + noise), N rollouts (one graph launch), local
// arg-min, winner record.  Nothing is read back: the caller all-gathers the record (hmj_plan_record_size doubles at
// `record_dev`, or the batch's own buffer when NULL) over its communicator ON hmj_batch_stream and finishes with
// hmj_plan_select.  noise == NULL draws the noise on the device (seed, iteration, global sample index).
extern "C" int hmj_plan_iterate_ex(hmj_batch_t* b, const double* state, double time, const double* spline_times,
                                   const double* nominal_values, int nnode, int interpolation, int horizon, const double* noise,
                                   double noise_std, double noise_std2, const double* noise_std_vec, unsigned long long seed,
                                   unsigned long long iteration, int first_global, double* record_dev) {
  if (!b || !state || !spline_times || !nominal_values || nnode < 1) return fail(HMJ_ERR_ARG, "bad argument to plan_iterate%s");
  CK(cudaSetDevice(b->device));
  int rc = ensure_values(b, nnode);
  if (rc) return rc;
  CK(cudaStreamSynchronize(b->stream));   // pinned staging reuse
  if ((rc = upload_plan_inputs(b, state, time, spline_times, nnode, horizon))) return rc;
  double* nom = b->plan_stage + 128;
  memcpy(nom, nominal_values, (size_t)nnode * HMJ_NU * 8);
  CK(cudaMemcpyAsync(b->nominal_dev, nom, (size_t)nnode * HMJ_NU * 8, cudaMemcpyHostToDevice, b->stream));
  if (noise_std_vec) {
    double* sv = b->plan_stage + 2048;
    memcpy(sv, noise_std_vec, (size_t)nnode * HMJ_NU * 8);
    CK(cudaMemcpyAsync(b->stdvec_dev, sv, (size_t)nnode * HMJ_NU * 8, cudaMemcpyHostToDevice, b->stream));
  }
  const size_t nvals = (size_t)b->n * nnode * HMJ_NU;
  if (noise) CK(cudaMemcpyAsync(b->noise_dev, noise, nvals * 8, cudaMemcpyHostToDevice, b->stream));
  hmj_launch_plan_candidates(b->tab_dev, b->n, nnode, b->nominal_dev, noise ? b->noise_dev : nullptr, noise_std, noise_std2, seed,
                             iteration, first_global, b->spline_values_dev, b->stream, noise_std_vec ? b->stdvec_dev : nullptr);
  if ((rc = launch_rollouts(b, b->spline_values_dev, nnode, interpolation, horizon, 0, b->returns_dev, b->failure_dev))) return rc;
  hmj_launch_plan_winner(b->returns_dev, b->n, b->spline_values_dev, nnode, first_global, b->argmin_dev,
                         record_dev ? record_dev : b->record_dev, b->stream);
  b->launches += 2;
  b->host_launches += 2;
  b->plan_nnode = nnode;
  CK(cudaGetLastError());
  return HMJ_OK;
}

extern "C" int hmj_plan_iterate(hmj_batch_t* b, const double* state, double time, const double* spline_times,
                                const double* nominal_values, int nnode, int interpolation, int horizon, const double* noise,
                                double noise_std, double noise_std2, unsigned long long seed, unsigned long long iteration,
                                int first_global, double* record_dev) {
  return hmj_plan_iterate_ex(b, state, time, spline_times, nominal_values, nnode, interpolation, horizon, noise, noise_std,
                             noise_std2, nullptr, seed, iteration, first_global, record_dev);
}

extern "C" int hmj_plan_record_size(int nnode) { return 3 + nnode * HMJ_NU; }
extern "C" void* hmj_plan_record_ptr(hmj_batch_t* b) { return b ? (void*)b->record_dev : nullptr; }

// Second half of the exchange (the parallel form of the partial_sort + CopyCandidateToPolicy pair, planner.cc:151-187,
// 525-534): `gathered_dev` = nranks winner records, rank-major (NULL / nranks <= 1: this batch's own record).  Picks the
// global winner on the batch stream and returns its record to the host: {cost, global index, nominal return, spline values}.
extern "C" int hmj_plan_select(hmj_batch_t* b, const double* gathered_dev, int nranks, int nnode, double* record_host) {
  if (!b || !record_host || nnode < 1 || nnode > 64) return fail(HMJ_ERR_ARG, "bad argument to plan_select%s");
  CK(cudaSetDevice(b->device));
  const int stride = hmj_plan_record_size(nnode);
  const double* src = b->record_dev;
  if (gathered_dev && nranks > 1) {
    hmj_launch_plan_select(gathered_dev, nranks, stride, b->select_dev, b->stream);
    b->launches += 1;
    b->host_launches += 1;
    src = b->select_dev;
  } else if (gathered_dev) {
    src = gathered_dev;
  }
  double* out = b->plan_stage + 1024;
  CK(cudaMemcpyAsync(out, src, (size_t)stride * 8, cudaMemcpyDeviceToHost, b->stream));
  CK(cudaStreamSynchronize(b->stream));
  CK(cudaGetLastError());
  memcpy(record_host, out, (size_t)stride * 8);
  return HMJ_OK;
}

// Returns / failure flags / candidate values of the last hmj_plan_iterate (any may be NULL), for callers that want the
// whole population (the reference keeps trajectory[i].total_return).
extern "C" int hmj_plan_population(hmj_batch_t* b, int nnode, double* total_return, int32_t* failure, double* values) {
  if (!b) return fail(HMJ_ERR_ARG, "null batch%s");
  CK(cudaSetDevice(b->device));
  int rc;
  if ((rc = d2h(b, total_return, b->returns_dev, (size_t)b->n * 8)) || (rc = d2h(b, failure, b->failure_dev, (size_t)b->n * 4))) return rc;
  if (values) {
    if ((size_t)b->n * nnode * HMJ_NU > b->spline_values_cap) return fail(HMJ_ERR_ARG, "no candidates of that size%s");
    if ((rc = d2h(b, values, b->spline_values_dev, (size_t)b->n * nnode * HMJ_NU * 8))) return rc;
  }
  CK(cudaStreamSynchronize(b->stream));
  return HMJ_OK;
}

extern "C" int hmj_plan_argmin(hmj_batch_t* b, double* best_cost, int32_t* best_index) {
  if (!b) return fail(HMJ_ERR_ARG, "null batch%s");
  CK(cudaSetDevice(b->device));
  hmj_launch_argmin(b->returns_dev, b->n, b->argmin_dev, b->stream);
  b->launches += 1; b->host_launches += 1;
  double out[2];
  CK(cudaMemcpyAsync(out, b->argmin_dev, sizeof(out), cudaMemcpyDeviceToHost, b->stream));
  CK(cudaStreamSynchronize(b->stream));
  if (best_cost) *best_cost = out[0];
  if (best_index) *best_index = (int32_t)out[1];
  return HMJ_OK;
}

extern "C" int hmj_plan_fetch(hmj_batch_t* b, int sample, int horizon, double* states, double* actions, double* times,
                              double* residual, double* costs, double* trace) {
  if (!b || sample < 0 || sample >= b->n || horizon < 1 || horizon > b->rec.horizon || !b->rec.states)
    return fail(HMJ_ERR_ARG, "no recorded trajectory for this sample/horizon%s");
  CK(cudaSetDevice(b->device));
  const size_t N = b->n;
  auto fetch = [&](double* dst, const double* src, int dim) -> int {
    if (!dst) return HMJ_OK;
    CK(cudaMemcpy2DAsync(dst, (size_t)dim * 8, src + (size_t)sample * dim, N * dim * 8, (size_t)dim * 8, horizon,
                         cudaMemcpyDeviceToHost, b->stream));
    return HMJ_OK;
  };
  int rc;
  if ((rc = fetch(states, b->rec.states, HMJ_NSTATE)) || (rc = fetch(actions, b->rec.actions, HMJ_NU)) ||
      (rc = fetch(times, b->rec.times, 1)) || (rc = fetch(residual, b->rec.residual, HMJ_NRESIDUAL)) ||
      (rc = fetch(costs, b->rec.costs, 1)) || (rc = fetch(trace, b->rec.trace, 3)))
    return rc;
  CK(cudaStreamSynchronize(b->stream));
  if (actions && horizon > 1)  // final action is a copy of the previous one (trajectory.cc:190-192)
    memcpy(actions + (size_t)(horizon - 1) * HMJ_NU, actions + (size_t)(horizon - 2) * HMJ_NU, HMJ_NU * 8);
  return HMJ_OK;
}

/* mj_inverse for every env (simulation/mujoco/include/mujoco/mujoco.h mj_inverse; the MPC workaround
 * find_ideal_distance_to_ground runs it 200 times per control step, controllers/mpc_utils.py:36-55): position and velocity
 * stages, collision, constraint rows, the constraint forces of the GIVEN acceleration (mj_invConstraint: no solver), and
 * qfrc_inverse = M qacc + qfrc_bias - qfrc_passive - qfrc_constraint.  qacc [N][18] (NULL: zeros); qfrc_inverse [N][18] out. */
extern "C" int hmj_batch_inverse(hmj_batch_t* b, const double* qacc, double* qfrc_inverse) {
  if (!b || !qfrc_inverse) return fail(HMJ_ERR_ARG, "bad argument to batch_inverse%s");
  CK(cudaSetDevice(b->device));
  const size_t N = b->n;
  if (qacc) CK(cudaMemcpyAsync(b->buf.qacc, qacc, N * HMJ_NV * 8, cudaMemcpyHostToDevice, b->stream));
  else CK(cudaMemsetAsync(b->buf.qacc, 0, N * HMJ_NV * 8, b->stream));
  hmj_launch_smooth(b->tab_dev, b->buf, b->task_dev, 0, nullptr, nullptr, 0, 0, 0, b->stream, &b->solve_ctx);
  hmj_launch_factor(b->tab_dev, b->buf, b->stream, &b->solve_ctx);
  hmj_launch_collide(b->tab_dev, b->buf, b->stream, false);
  hmj_launch_solve(b->tab_dev, b->buf, 2, b->stream, &b->solve_ctx, false);
  b->warmed = true;
  b->launches += 10; b->host_launches += 10;
  CK(cudaMemcpyAsync(qfrc_inverse, b->buf.qfrc_inverse, N * HMJ_NV * 8, cudaMemcpyDeviceToHost, b->stream));
  CK(cudaStreamSynchronize(b->stream));
  CK(cudaGetLastError());
  return HMJ_OK;
}

// ------------------------------------------------------------------ batched finite-difference derivatives (hmj_fd.cu)
/* mjd_transitionFD for T knots in one batched step (mjpc ModelDerivatives::Compute, model_derivatives.cc:51-96): the batch
 * must hold at least 97 T environments; their state is overwritten (parameters are the caller's: set the planner timestep
 * with hmj_batch_set_params).  x [T][37], u [T][12], time [T], warmstart [T][18] or NULL (zeros); eps as mjpc passes it
 * (1e-6 forward in iLQG); outputs row-major per knot: A [T][36][36], B [T][36][12], C [T][28][36], D [T][28][12] (any NULL). */
extern "C" int hmj_fd_transition(hmj_batch_t* b, int T, const double* x, const double* u, const double* time,
                                 const double* warmstart, double eps, int centered, double* A, double* Bm, double* C, double* D) {
  if (!b || T < 1 || !x || !u || !time || !(eps > 0)) return fail(HMJ_ERR_ARG, "bad argument to fd_transition%s");
  const int slots = 97, ndx = 2 * HMJ_NV;
  if ((long long)T * slots > b->n) return fail(HMJ_ERR_ARG, "fd_transition needs 97 environments per knot%s");
  CK(cudaSetDevice(b->device));
  const size_t nin = (size_t)T * (HMJ_NSTATE + HMJ_NU + 1 + HMJ_NV);
  const size_t nout = (size_t)T * (ndx * ndx + ndx * HMJ_NU + HMJ_NRESIDUAL * ndx + HMJ_NRESIDUAL * HMJ_NU);
  const size_t bytes = (nin + nout) * 8 + (size_t)T * slots * 4;
  if (b->fd_cap < bytes) {
    CK(cudaStreamSynchronize(b->stream));
    if (b->fd_dev) {
      for (size_t k = 0; k < b->allocs.size(); k++)
        if (b->allocs[k] == (void*)b->fd_dev) { b->allocs.erase(b->allocs.begin() + k); break; }
      cudaFree(b->fd_dev);
      b->fd_dev = nullptr;
    }
    void* p = nullptr;
    CK(cudaMalloc(&p, bytes));
    b->allocs.push_back(p);
    b->fd_dev = (double*)p;
    b->fd_cap = bytes;
  }
  double* d_x = b->fd_dev;
  double* d_u = d_x + (size_t)T * HMJ_NSTATE;
  double* d_t = d_u + (size_t)T * HMJ_NU;
  double* d_w = d_t + T;
  double* d_A = d_w + (size_t)T * HMJ_NV;
  double* d_B = d_A + (size_t)T * ndx * ndx;
  double* d_C = d_B + (size_t)T * ndx * HMJ_NU;
  double* d_D = d_C + (size_t)T * HMJ_NRESIDUAL * ndx;
  int32_t* d_valid = (int32_t*)(d_D + (size_t)T * HMJ_NRESIDUAL * HMJ_NU);
  int rc;
  if ((rc = h2d(b, d_x, x, (size_t)T * HMJ_NSTATE * 8)) || (rc = h2d(b, d_u, u, (size_t)T * HMJ_NU * 8)) ||
      (rc = h2d(b, d_t, time, (size_t)T * 8)) || (rc = h2d(b, d_w, warmstart, (size_t)T * HMJ_NV * 8)))
    return rc;
  hmj_buffers_t V = pipe_view(b->buf, 0, T * slots, 0, b->tab_host->group_base[HMJ_NGROUP]);
  hmj_launch_fd_setup(b->tab_dev, V, T, d_x, d_u, d_t, warmstart ? d_w : nullptr, eps, centered, d_valid, b->stream);
  hmj_launch_smooth(b->tab_dev, V, b->task_dev, HMJ_PLAN_COST, nullptr, nullptr, 0, 0, 0, b->stream, &b->solve_ctx);
  hmj_launch_factor(b->tab_dev, V, b->stream, &b->solve_ctx);
  hmj_launch_collide(b->tab_dev, V, b->stream, false);
  hmj_launch_solve(b->tab_dev, V, 1, b->stream, &b->solve_ctx, false);
  hmj_launch_fd_diff(b->tab_dev, V, T, d_valid, eps, d_A, d_B, d_C, d_D, b->stream);
  b->warmed = true;
  b->launches += 12; b->host_launches += 12;
  if ((rc = d2h(b, A, d_A, (size_t)T * ndx * ndx * 8)) || (rc = d2h(b, Bm, d_B, (size_t)T * ndx * HMJ_NU * 8)) ||
      (rc = d2h(b, C, d_C, (size_t)T * HMJ_NRESIDUAL * ndx * 8)) || (rc = d2h(b, D, d_D, (size_t)T * HMJ_NRESIDUAL * HMJ_NU * 8)))
    return rc;
  CK(cudaStreamSynchronize(b->stream));
  CK(cudaGetLastError());
  return HMJ_OK;
}

// ------------------------------------------------------------------ built-in winner exchange over NCCL
// A C++ SamplingPlanner shards its samples over GPUs without Python: one communicator per process (NCCL is loaded at
// run time with dlopen — the library itself links only cudart — so single-GPU users never need it), and the planner's
// ONE exchange per iteration is an all-gather of the winner records on the batch stream followed by the device selection.
#include <dlfcn.h>
struct hmj_nccl_id { char internal[128]; };
typedef int (*nccl_get_id_fn)(hmj_nccl_id*);
typedef int (*nccl_init_rank_fn)(void**, int, hmj_nccl_id, int);
typedef int (*nccl_destroy_fn)(void*);
typedef int (*nccl_allgather_fn)(const void*, void*, size_t, int, void*, cudaStream_t);
typedef const char* (*nccl_errstr_fn)(int);
static struct {
  void* handle;
  nccl_get_id_fn get_id;
  nccl_init_rank_fn init_rank;
  nccl_destroy_fn destroy;
  nccl_allgather_fn allgather;
  nccl_errstr_fn errstr;
} g_nccl;

static int load_nccl() {
  if (g_nccl.handle) return HMJ_OK;
  const char* names[] = {getenv("HMJ_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
  void* h = nullptr;
  for (const char* nm : names)
    if (nm && (h = dlopen(nm, RTLD_NOW | RTLD_GLOBAL))) break;
  if (!h) return fail(HMJ_ERR_CUDA, "NCCL not found (dlopen libnccl.so.2): %s", dlerror());
  g_nccl.get_id = (nccl_get_id_fn)dlsym(h, "ncclGetUniqueId");
  g_nccl.init_rank = (nccl_init_rank_fn)dlsym(h, "ncclCommInitRank");
  g_nccl.destroy = (nccl_destroy_fn)dlsym(h, "ncclCommDestroy");
  g_nccl.allgather = (nccl_allgather_fn)dlsym(h, "ncclAllGather");
  g_nccl.errstr = (nccl_errstr_fn)dlsym(h, "ncclGetErrorString");
  if (!g_nccl.get_id || !g_nccl.init_rank || !g_nccl.destroy || !g_nccl.allgather) {
    dlclose(h);
    return fail(HMJ_ERR_CUDA, "NCCL library lacks the expected symbols%s");
  }
  g_nccl.handle = h;
  return HMJ_OK;
}

struct hmj_comm_ {
  void* comm = nullptr;
  int rank = 0, nranks = 1, device = 0;
  double* gathered = nullptr;   // [nranks][3 + 64*12]
};

#define NCK(call)                                                                                         \
  do {                                                                                                    \
    int r_ = (call);                                                                                      \
    if (r_ != 0) return fail(HMJ_ERR_CUDA, "NCCL error: %s", g_nccl.errstr ? g_nccl.errstr(r_) : "?");   \
  } while (0)

extern "C" int hmj_comm_unique_id(void* id128) {
  if (!id128) return fail(HMJ_ERR_ARG, "null id%s");
  int rc = load_nccl();
  if (rc) return rc;
  NCK(g_nccl.get_id((hmj_nccl_id*)id128));
  return HMJ_OK;
}

extern "C" int hmj_comm_create(const void* id128, int rank, int nranks, int device, hmj_comm_t** out) {
  if (!id128 || !out || nranks < 1 || rank < 0 || rank >= nranks) return fail(HMJ_ERR_ARG, "bad argument to comm_create%s");
  int rc = load_nccl();
  if (rc) return rc;
  CK(cudaSetDevice(device));
  hmj_comm_* c = new hmj_comm_();
  c->rank = rank; c->nranks = nranks; c->device = device;
  hmj_nccl_id id;
  memcpy(&id, id128, sizeof(id));
  int r = g_nccl.init_rank(&c->comm, nranks, id, rank);
  if (r != 0) { delete c; return fail(HMJ_ERR_CUDA, "ncclCommInitRank: %s", g_nccl.errstr ? g_nccl.errstr(r) : "?"); }
  if (cudaMalloc((void**)&c->gathered, (size_t)nranks * (3 + 64 * HMJ_NU) * 8) != cudaSuccess) {
    g_nccl.destroy(c->comm);
    delete c;
    return fail(HMJ_ERR_CUDA, "cudaMalloc failed%s");
  }
  *out = c;
  return HMJ_OK;
}

extern "C" void hmj_comm_destroy(hmj_comm_t* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  if (c->gathered) cudaFree(c->gathered);
  if (c->comm && g_nccl.destroy) g_nccl.destroy(c->comm);
  delete c;
}

extern "C" int hmj_comm_rank(const hmj_comm_t* c) { return c ? c->rank : 0; }
extern "C" int hmj_comm_size(const hmj_comm_t* c) { return c ? c->nranks : 1; }

// all-gather of this rank's winner record (written by hmj_plan_iterate) on the batch stream + device selection:
// the whole exchange of one planning iteration.  comm == NULL: single process.
extern "C" int hmj_plan_exchange(hmj_batch_t* b, hmj_comm_t* c, int nnode, double* record_host) {
  if (!b || !record_host || nnode < 1 || nnode > 64) return fail(HMJ_ERR_ARG, "bad argument to plan_exchange%s");
  if (!c || c->nranks == 1) return hmj_plan_select(b, nullptr, 1, nnode, record_host);
  if (c->device != b->device) return fail(HMJ_ERR_ARG, "communicator and batch live on different devices%s");
  CK(cudaSetDevice(b->device));
  const int stride = hmj_plan_record_size(nnode);
  NCK(g_nccl.allgather(b->record_dev, c->gathered, (size_t)stride, /*ncclFloat64*/ 8, c->comm, b->stream));
  return hmj_plan_select(b, c->gathered, c->nranks, nnode, record_host);
}

// ------------------------------------------------------------------ allocation-free variant of hmj_env_obs_reward
// No counterpart in the reference; added for callers that must not reach
// cudaMalloc (code under graph capture, tight serving loops, harnesses that forbid it).  hmj_env_obs_reward stages its host
// inputs and results in a device block it allocates on first use; this variant keeps that block in a file-scope device
// array of fixed size, so the call allocates nothing on the host or the device: async copies, one memset, one kernel.
//   * limit: batches of at most HMJ_NOALLOC_MAX_ENVS envs (700 bytes of scratch per env, 2.8 MB in total; every GPU that
//     loads the library holds the array until the process exits);
//   * ONE array per device, shared by every batch and every call: calls that overlap on the GPU (two batches = two
//     streams, or two host threads) race on it; the call synchronises its stream before it returns, so calls made one
//     after the other from one thread are safe.  The array keeps the previous call's contents; the result part is
//     cleared at the start of every call (rows of envs the kernel skips read as zero, never as the last call's).
#define HMJ_NOALLOC_MAX_ENVS 4096
#define HMJ_NOALLOC_DOUBLES_PER_ENV (12 + 12 + 31 + 30 + 1)
#define HMJ_NOALLOC_BYTES ((size_t)HMJ_NOALLOC_MAX_ENVS * (HMJ_NOALLOC_DOUBLES_PER_ENV * 8 + 3 * 4))
static __device__ __align__(256) unsigned char g_noalloc_scratch[HMJ_NOALLOC_BYTES];
// In host code `g_noalloc_scratch` is only a handle for the runtime, not a device address: the address each device gave
// the array is looked up once per device (with that device current) and cached here.
static unsigned char* g_noalloc_addr[HMJ_MAX_DEVICES] = {nullptr};

static int noalloc_scratch(int device, unsigned char** out) {
  if (device < 0 || device >= HMJ_MAX_DEVICES) return fail(HMJ_ERR_ARG, "device index beyond the scratch table%s");
  if (!g_noalloc_addr[device]) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);   // the current device's copy of the array
    if (e != cudaSuccess || !p) {
      cudaGetLastError();
      return fail(HMJ_ERR_CUDA, "cudaGetSymbolAddress(static scratch) failed: %s", e != cudaSuccess ? cudaGetErrorString(e) : "null address");
    }
    g_noalloc_addr[device] = (unsigned char*)p;
  }
  *out = g_noalloc_addr[device];
  return HMJ_OK;
}

extern "C" int hmj_env_obs_reward_noalloc_max_envs(void) { return HMJ_NOALLOC_MAX_ENVS; }

extern "C" int hmj_env_obs_reward_noalloc(hmj_batch_t* b, const double* prev_action, const double* latest_action,
                                          const double* noise, double max_sim_time, double* obs, double* reward,
                                          int32_t* terminated, int32_t* truncated, int32_t* self_collision) {
  if (!b || !prev_action || !latest_action || !obs || !reward || !terminated || !truncated)
    return fail(HMJ_ERR_ARG, "bad argument to env_obs_reward_noalloc%s");
  if (b->n > HMJ_NOALLOC_MAX_ENVS) {
    snprintf(g_err, sizeof(g_err), "env_obs_reward_noalloc: batch of %d envs exceeds the static scratch (at most %d envs); use hmj_env_obs_reward",
             b->n, HMJ_NOALLOC_MAX_ENVS);
    return HMJ_ERR_ARG;
  }
  CK(cudaSetDevice(b->device));
  unsigned char* base = nullptr;
  int rc = noalloc_scratch(b->device, &base);
  if (rc) return rc;
  const size_t N = b->n;
  // layout as in hmj_env_obs_reward (doubles): prev 12N | latest 12N | noise 31N | obs 30N | reward N | int32: term N, trunc N, selfc N
  double* d_prev = (double*)base;
  double* d_lat = d_prev + 12 * N;
  double* d_noise = d_lat + 12 * N;
  double* d_obs = d_noise + 31 * N;
  double* d_rew = d_obs + 30 * N;
  int32_t* d_term = (int32_t*)(d_rew + N);
  int32_t* d_trunc = d_term + N;
  int32_t* d_self = d_trunc + N;
  CK(cudaMemsetAsync(d_obs, 0, 31 * N * 8 + 3 * N * 4, b->stream));   // results start at zero, whatever the last call left
  if ((rc = h2d(b, d_prev, prev_action, 12 * N * 8)) || (rc = h2d(b, d_lat, latest_action, 12 * N * 8)) ||
      (rc = h2d(b, d_noise, noise, 31 * N * 8)))
    return rc;
  hmj_launch_env_obs_reward(b->tab_dev, b->buf, d_prev, d_lat, noise ? d_noise : nullptr, max_sim_time, d_obs, d_rew, d_term,
                            d_trunc, d_self, b->stream);
  b->launches += 1; b->host_launches += 1;
  CK(cudaGetLastError());
  if ((rc = d2h(b, obs, d_obs, 30 * N * 8)) || (rc = d2h(b, reward, d_rew, N * 8)) || (rc = d2h(b, terminated, d_term, N * 4)) ||
      (rc = d2h(b, truncated, d_trunc, N * 4)) || (rc = d2h(b, self_collision, d_self, N * 4)))
    return rc;
  CK(cudaStreamSynchronize(b->stream));
  return HMJ_OK;
}
