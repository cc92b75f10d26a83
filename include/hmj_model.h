/* hmj_model.h — compiled-model blob for the Humanoid-MuJoCo humanoid (POD, C ABI).
 *
 * One flat, versioned struct produced offline by tools/compile_model.py from the
 * reference assets (simulation/assets/world.xml, humanoid.xml, humanoid_urdf STL files and
 * mujoco_mpc/mjpc/tasks/humanoid_cap/stand/task.xml) and consumed unchanged by the CPU
 * oracle (oracle/) and the CUDA library (humanoid-mujoco_b200/csrc/).  It plays the
 * role of MuJoCo's mjModel for this one robot; field names follow mjModel
 * (simulation/mujoco/include/mujoco/mjmodel.h:560-1040) so parity can be read off.
 *
 * Dimensions are compile-time constants: the engine is specialised to this kinematic
 * tree (15 bodies, free joint + 12 hinges, 12 motors, 9 convex meshes + 1 height field).
 */
#ifndef HMJ_MODEL_H_
#define HMJ_MODEL_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HMJ_MAGIC 0x484d4a31 /* "HMJ1" */
#define HMJ_VERSION 4

#define HMJ_NBODY 15
#define HMJ_NJNT 13
#define HMJ_NQ 19
#define HMJ_NV 18
#define HMJ_NU 12
#define HMJ_NM 117    /* sparse inertia entries (mjModel.nM) */
#define HMJ_NGEOM 10  /* colliding geoms: 0 = floor hfield, 1..9 = meshes */
#define HMJ_NMESH 7   /* distinct convex hulls referenced by the 9 mesh geoms */
#define HMJ_NSITE 2
#define HMJ_NPAIR 37  /* candidate geom pairs after exclude/parent filters */
#define HMJ_MAXVERT 1024
#define HMJ_HF_NROW 8
#define HMJ_HF_NCOL 8
#define HMJ_NHFDATA 64

/* planner task (humanoid_cap stand) */
#define HMJ_NRESIDUAL 28
#define HMJ_NTERM 5
#define HMJ_MAXNORMPRM 3
#define HMJ_NSTATE (HMJ_NQ + HMJ_NV) /* na = 0 */

/* capacities (per environment) */
#define HMJ_MAXCON 16               /* contacts kept per env; overflow -> HMJ_STATUS_CONTACTFULL */
#define HMJ_MAXCONPAIR 8            /* contacts kept per mesh-hfield pair (MuJoCo: mjMAXCONPAIR 50) */
#define HMJ_MAXEFC (12 * 2 + 10 * HMJ_MAXCON)

/* joint types (mjtJoint) */
#define HMJ_JNT_FREE 0
#define HMJ_JNT_HINGE 3

/* geom types (mjtGeom) */
#define HMJ_GEOM_HFIELD 1
#define HMJ_GEOM_MESH 7

typedef struct hmj_model_ {
  int32_t magic, version, sizeof_model, reserved0;

  /* ---- mjOption (defaults; neither XML has an <option>) ---- */
  double timestep;
  double gravity[3];
  double tolerance;      /* solver tolerance 1e-8 */
  double ls_tolerance;   /* 0.01 */
  double mpr_tolerance;  /* 1e-6 */
  double meaninertia;    /* mjModel.stat.meaninertia */
  int32_t iterations;    /* 100 */
  int32_t ls_iterations; /* 50 */
  int32_t mpr_iterations;/* 50 */
  int32_t reserved1;

  /* ---- bodies ---- */
  int32_t body_parentid[HMJ_NBODY];
  int32_t body_rootid[HMJ_NBODY];
  int32_t body_weldid[HMJ_NBODY];
  int32_t body_jntadr[HMJ_NBODY];
  int32_t body_jntnum[HMJ_NBODY];
  int32_t body_dofadr[HMJ_NBODY];
  int32_t body_dofnum[HMJ_NBODY];
  int32_t body_sameframe[HMJ_NBODY]; /* inertial frame == body frame */
  int32_t body_depth[HMJ_NBODY];
  int32_t reserved2;
  double body_pos[HMJ_NBODY][3];
  double body_quat[HMJ_NBODY][4];
  double body_ipos[HMJ_NBODY][3];
  double body_iquat[HMJ_NBODY][4];
  double body_mass[HMJ_NBODY];
  double body_subtreemass[HMJ_NBODY];
  double body_inertia[HMJ_NBODY][3];
  double body_invweight0[HMJ_NBODY][2];

  /* ---- joints ---- */
  int32_t jnt_type[HMJ_NJNT];
  int32_t jnt_qposadr[HMJ_NJNT];
  int32_t jnt_dofadr[HMJ_NJNT];
  int32_t jnt_bodyid[HMJ_NJNT];
  int32_t jnt_limited[HMJ_NJNT];
  int32_t reserved3;
  double jnt_pos[HMJ_NJNT][3];
  double jnt_axis[HMJ_NJNT][3];
  double jnt_stiffness[HMJ_NJNT];
  double jnt_range[HMJ_NJNT][2];
  double jnt_margin[HMJ_NJNT];
  double jnt_solref[HMJ_NJNT][2];
  double jnt_solimp[HMJ_NJNT][5];
  double qpos0[HMJ_NQ];
  double qpos_spring[HMJ_NQ];

  /* ---- dofs ---- */
  int32_t dof_bodyid[HMJ_NV];
  int32_t dof_jntid[HMJ_NV];
  int32_t dof_parentid[HMJ_NV];
  int32_t dof_Madr[HMJ_NV + 1];
  int32_t reserved4;
  double dof_armature[HMJ_NV];
  double dof_damping[HMJ_NV];
  double dof_invweight0[HMJ_NV];

  /* ---- actuators (motors on hinge joints) ---- */
  int32_t act_dofadr[HMJ_NU];
  int32_t act_qposadr[HMJ_NU];
  double act_gear[HMJ_NU];
  double act_gainprm0[HMJ_NU];
  double act_ctrlrange[HMJ_NU][2];
  double act_forcerange[HMJ_NU][2];

  /* ---- colliding geoms ---- */
  int32_t geom_type[HMJ_NGEOM];
  int32_t geom_bodyid[HMJ_NGEOM];
  int32_t geom_meshid[HMJ_NGEOM]; /* -1 for the hfield */
  int32_t geom_mjid[HMJ_NGEOM];   /* geom id in the MuJoCo model (for contact.geom1/2) */
  double geom_pos[HMJ_NGEOM][3];
  double geom_quat[HMJ_NGEOM][4];
  double geom_rbound[HMJ_NGEOM];
  double geom_friction[HMJ_NGEOM][3];
  double geom_solref[HMJ_NGEOM][2];
  double geom_solimp[HMJ_NGEOM][5];
  double geom_solmix[HMJ_NGEOM];

  /* ---- candidate collision pairs, in MuJoCo contact order ---- */
  int32_t npair;
  int32_t reserved5;
  int32_t pair_geom1[HMJ_NPAIR + 1];
  int32_t pair_geom2[HMJ_NPAIR + 1];

  /* ---- convex hull vertices (float32, mesh CoM/principal frame) ---- */
  int32_t mesh_vertadr[HMJ_NMESH + 1];
  int32_t mesh_vertnum[HMJ_NMESH + 1];
  float mesh_vert[HMJ_MAXVERT][3];

  /* ---- height field ---- */
  double hfield_size[4]; /* x half-extent, y half-extent, z top scale, z bottom */
  int32_t hfield_nrow, hfield_ncol;
  float hfield_data[HMJ_NHFDATA];

  /* ---- sites (feet) ---- */
  int32_t site_bodyid[HMJ_NSITE];
  double site_pos[HMJ_NSITE][3];

  /* ---- planner task: humanoid_cap stand (task.xml user sensors) ---- */
  int32_t task_nterm, task_nresidual;
  int32_t task_dim[HMJ_NTERM + 1];
  int32_t task_norm[HMJ_NTERM + 1];
  double task_weight[HMJ_NTERM + 1];
  double task_normprm[HMJ_NTERM + 1][HMJ_MAXNORMPRM];
  double task_risk;
  double task_height_goal;    /* residual_Height Goal */
  double agent_timestep;      /* 0.02 */
  double agent_horizon;       /* 0.35 */
  double sampling_exploration;/* 0.05 */
  int32_t sampling_spline_points; /* 3 */
  int32_t sampling_trajectories;  /* 10 (default) */
} hmj_model_t;


/* Per-environment parameters: the mjModel fields CPUEnv.reset domain-randomises
 * (simulation/cpu_env.py:189-280) plus the timestep the caller runs at
 * (0.002 env, cpu_env.py:87; 0.02 planner, mjpc/agent.cc:275). */
typedef struct hmj_params_ {
  double timestep;
  double body_mass[HMJ_NBODY];
  double dof_armature[HMJ_NV];
  double jnt_range[HMJ_NJNT][2];
  double jnt_margin[HMJ_NJNT];
  double jnt_stiffness[HMJ_NJNT];
  double act_gainprm0[HMJ_NU];
  double act_forcerange[HMJ_NU][2];
  double floor_friction[3];
  float hfield_data[HMJ_NHFDATA];
} hmj_params_t;

#define HMJ_NPARAM_DOUBLES (1 + HMJ_NBODY + HMJ_NV + 2 * HMJ_NJNT + HMJ_NJNT + HMJ_NJNT + HMJ_NU + 2 * HMJ_NU + 3)

/* One contact, MuJoCo mjContact subset (mjdata.h:99-130). geom ids are MuJoCo geom ids. */
typedef struct hmj_contact_ {
  double dist;
  double pos[3];
  double frame[9];
  double friction[5];
  double solref[2];
  double solimp[5];
  int32_t geom1, geom2; /* MuJoCo geom ids (contact.geom1/geom2) */
  int32_t g1, g2;       /* colliding-geom indices (0..HMJ_NGEOM-1) */
} hmj_contact_t;

/* status bits (per env / per sample) */
#define HMJ_STATUS_BADQPOS 1
#define HMJ_STATUS_BADQVEL 2
#define HMJ_STATUS_BADQACC 4
#define HMJ_STATUS_CONTACTFULL 8
#define HMJ_STATUS_BADCTRL 16
#define HMJ_STATUS_INERTIA 32

#ifdef __cplusplus
}
#endif
#endif /* HMJ_MODEL_H_ */
