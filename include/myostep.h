/* libmyostep.so -- C ABI of the B200-native batched musculoskeletal step.
 *
 * This is the drop-in boundary for the physics step the reference reaches through third-party code:
 *   mjx.put_model / mjx.put_data / mjx.step / mjx.get_data   (reference: myosuite/mjx/play.py:10-11,41,47,49)
 *   brax pipeline_init / pipeline_step                       (reference: myosuite/mjx/myodm_v0.py:159,287)
 *   dm_control Physics.step(n) -> n x mj_step, sim.forward() (reference: myosuite/physics/mj_sim_scene.py:55,
 *                                                             myosuite/robot/robot.py:880-882,976-979)
 * Plain pointers and sizes only; no torch / jax types.  All device work is enqueued on the caller's CUDA
 * stream (passed as void* = cudaStream_t); nothing here synchronises unless stated.
 *
 * Batched layout: struct-of-arrays with the WORLD AXIS FASTEST.  A field with k elements per world is one dense
 * buffer [k][nworld] (element i of world w at base[i * nworld + w]); the logical shape a caller sees is the vmapped
 * mjx.Data leaf [nworld, k...] as a strided view (torch: buffer.t(); XLA FFI: operand layout minor_to_major with
 * the world axis minor-most).  One GPU thread owns one world, so the 32 lanes of a warp read and write one
 * coalesced 128-byte line per element.
 */
#ifndef MYOSTEP_H_
#define MYOSTEP_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MYO_NPRM 10      /* mjNGAIN / mjNBIAS / mjNDYN */
#define MYO_NEQDATA 11   /* mjNEQDATA */
#define MYO_NIMP 5
#define MYO_NREF 2

/* mjModel-shaped host arrays (float64 / int32, row-major, field names and enum values of MuJoCo 3.2.8's
 * mjModel).  Filled by mjx.put_model from our MJCF compiler or from a real mujoco.MjModel. */
typedef struct myo_model_desc {
  /* sizes */
  int32_t nq, nv, nu, na, nbody, njnt, ngeom, nsite, ntendon, nwrap, neq, npair, nexclude, nM;
  /* mjOption */
  int32_t opt_iterations, opt_ls_iterations, opt_disableflags, opt_cone, opt_solver, opt_integrator;
  double opt_timestep, opt_tolerance, opt_ls_tolerance, opt_impratio;
  double opt_gravity[3];
  double stat_meaninertia;
  /* bodies */
  const int32_t *body_parentid, *body_rootid, *body_weldid, *body_jntnum, *body_jntadr, *body_dofnum, *body_dofadr,
      *body_geomnum, *body_geomadr;
  const double *body_pos, *body_quat, *body_ipos, *body_iquat, *body_mass, *body_subtreemass, *body_inertia,
      *body_invweight0;
  /* joints and dofs */
  const int32_t *jnt_type, *jnt_qposadr, *jnt_dofadr, *jnt_bodyid, *jnt_limited;
  const double *jnt_pos, *jnt_axis, *jnt_stiffness, *jnt_range, *jnt_margin, *jnt_solref, *jnt_solimp;
  const double *qpos0, *qpos_spring;
  const int32_t *dof_bodyid, *dof_jntid, *dof_parentid, *dof_Madr;
  const double *dof_armature, *dof_damping, *dof_invweight0;
  /* geoms and sites */
  const int32_t *geom_type, *geom_contype, *geom_conaffinity, *geom_condim, *geom_bodyid, *geom_priority;
  const double *geom_size, *geom_rbound, *geom_pos, *geom_quat, *geom_friction, *geom_solmix, *geom_solref,
      *geom_solimp, *geom_margin, *geom_gap;
  const int32_t *site_bodyid;
  const double *site_pos, *site_quat;
  /* tendons */
  const int32_t *tendon_adr, *tendon_num, *tendon_limited;
  const double *tendon_range, *tendon_margin, *tendon_stiffness, *tendon_damping, *tendon_lengthspring,
      *tendon_solref_lim, *tendon_solimp_lim, *tendon_invweight0;
  const int32_t *wrap_type, *wrap_objid;
  const double *wrap_prm;
  /* actuators */
  const int32_t *actuator_trntype, *actuator_trnid, *actuator_dyntype, *actuator_gaintype, *actuator_biastype,
      *actuator_ctrllimited, *actuator_forcelimited, *actuator_actlimited, *actuator_actadr;
  const double *actuator_dynprm, *actuator_gainprm, *actuator_biasprm, *actuator_ctrlrange, *actuator_forcerange,
      *actuator_actrange, *actuator_lengthrange, *actuator_gear, *actuator_acc0;
  /* equality constraints (joint type only) */
  const int32_t *eq_type, *eq_obj1id, *eq_obj2id, *eq_active0;
  const double *eq_data, *eq_solref, *eq_solimp;
  /* explicit contact pairs and excludes */
  const int32_t *pair_dim, *pair_geom1, *pair_geom2, *exclude_signature;
  const double *pair_solref, *pair_solimp, *pair_margin, *pair_gap, *pair_friction;
  /* touch sensors (mjSENS_TOUCH only; myolegs_assets.xml:106-110): the sensor's site, and every site's shape for the zone test */
  int32_t nsensor;
  const int32_t *sensor_objid, *site_type;
  const double *site_size;
  /* convex hulls of the mesh geoms (row n2: mesh-convex collision of the object_sim contact*.stl pieces, reference
   * myosuite/simhive/object_sim/common.xml:10-20): geom_dataid[g] = mesh of geom g or -1; mesh m's hull vertices are
   * mesh_vert[3 * mesh_vertadr[m] ...), mesh_vertnum[m] of them, in the geom's own frame (MuJoCo: re-centred on the centre of
   * mass / principal axes).  Collision uses the hull, as MuJoCo does. */
  int32_t nmesh, nmeshvert;
  const int32_t *mesh_vertadr, *mesh_vertnum, *geom_dataid;
  const double *mesh_vert;
} myo_model_desc;

/* status codes */
enum {
  MYO_OK = 0,
  MYO_ERR_ARG = -1,          /* bad argument / unknown field / size mismatch */
  MYO_ERR_UNSUPPORTED = -2,  /* model uses a feature outside the supported subset (raised at model create) */
  MYO_ERR_CUDA = -3,         /* CUDA runtime error; no CPU fallback exists */
  MYO_ERR_UNBOUND = -4       /* a required state field has no bound device buffer */
};

/* per-world status bits written by the step kernels into the `status` field.
 * DEVIATION from MuJoCo, on purpose: mj_checkPos / mj_checkVel / mj_checkAcc call mj_resetData (qpos0, zero qvel / act / ctrl,
 * time = 0, a warning counter) and carry on; here the bad world is reset to qpos0 with zero qvel / act / warm start, its `time`
 * and `ctrl` are KEPT (the batch shares one control loop), it is not integrated on the step that found the bad qacc, and the
 * event is recorded as a bit in `status` (sticky until the caller clears the field) instead of mjData.warning. */
enum {
  MYO_WORLD_BADQACC = 1,     /* NaN/Inf or huge qacc: world was reset to qpos0 (mj_checkAcc rule) */
  MYO_WORLD_NCON_OVERFLOW = 2,
  MYO_WORLD_NEFC_OVERFLOW = 4,
  MYO_WORLD_SOLVER_MAXITER = 8,
  MYO_WORLD_BADQPOS = 16,
  MYO_WORLD_BADQVEL = 32
};

/* Batched per-world fields.  State fields are read and written in place by myo_step; the others are
 * outputs (mjData names).  Shapes are per world ([k] below means a [k][nworld] buffer); dtype float32 unless noted. */
typedef enum myo_field {
  /* state */
  MYO_F_QPOS = 0,        /* [nq] */
  MYO_F_QVEL,            /* [nv] */
  MYO_F_ACT,             /* [na] */
  MYO_F_CTRL,            /* [nu]   input */
  MYO_F_TIME,            /* [1] */
  MYO_F_QACC_WARMSTART,  /* [nv] */
  /* outputs */
  MYO_F_QACC,            /* [nv] */
  MYO_F_QACC_SMOOTH,     /* [nv] */
  MYO_F_ACT_DOT,         /* [na] */
  MYO_F_XPOS,            /* [nbody,3] */
  MYO_F_XQUAT,           /* [nbody,4] */
  MYO_F_XMAT,            /* [nbody,9] */
  MYO_F_XIPOS,           /* [nbody,3] */
  MYO_F_XIMAT,           /* [nbody,9] */
  MYO_F_XANCHOR,         /* [njnt,3] */
  MYO_F_XAXIS,           /* [njnt,3] */
  MYO_F_GEOM_XPOS,       /* [ngeom,3] */
  MYO_F_GEOM_XMAT,       /* [ngeom,9] */
  MYO_F_SITE_XPOS,       /* [nsite,3] */
  MYO_F_SITE_XMAT,       /* [nsite,9] */
  MYO_F_SUBTREE_COM,     /* [nbody,3] */
  MYO_F_CINERT,          /* [nbody,10] */
  MYO_F_CRB,             /* [nbody,10] */
  MYO_F_CDOF,            /* [nv,6] */
  MYO_F_CDOF_DOT,        /* [nv,6] */
  MYO_F_CVEL,            /* [nbody,6] */
  MYO_F_TEN_LENGTH,      /* [ntendon] */
  MYO_F_TEN_VELOCITY,    /* [ntendon] */
  MYO_F_TEN_J,           /* [ntendon,nv] */
  MYO_F_TEN_WRAPADR,     /* [ntendon] int32 */
  MYO_F_TEN_WRAPNUM,     /* [ntendon] int32 */
  MYO_F_WRAP_OBJ,        /* [nwrap,2] int32 */
  MYO_F_WRAP_XPOS,       /* [nwrap,6] */
  MYO_F_ACTUATOR_LENGTH,   /* [nu] */
  MYO_F_ACTUATOR_VELOCITY, /* [nu] */
  MYO_F_ACTUATOR_MOMENT,   /* [nu,nv] */
  MYO_F_ACTUATOR_FORCE,    /* [nu] */
  MYO_F_QM,              /* [nv,nv] dense symmetric (MJX layout for nv<60) */
  MYO_F_QLD,             /* [nv,nv] dense: strict lower part = L (unit, L^T D L), diagonal = D */
  MYO_F_QLDIAGINV,       /* [nv] */
  MYO_F_QFRC_BIAS,       /* [nv] */
  MYO_F_QFRC_PASSIVE,    /* [nv] */
  MYO_F_QFRC_SPRING,     /* [nv] */
  MYO_F_QFRC_DAMPER,     /* [nv] */
  MYO_F_QFRC_ACTUATOR,   /* [nv] */
  MYO_F_QFRC_SMOOTH,     /* [nv] */
  MYO_F_QFRC_CONSTRAINT, /* [nv] */
  MYO_F_QFRC_INVERSE,    /* [nv] */
  /* constraints and contacts (padded to njmax / nconmax rows per world) */
  MYO_F_NCON,            /* [1] int32 */
  MYO_F_NEFC,            /* [1] int32 */
  MYO_F_CONTACT_GEOM,    /* [nconmax,2] int32   (geom1, geom2) */
  MYO_F_CONTACT_DIM,     /* [nconmax] int32 */
  MYO_F_CONTACT_EFC_ADDRESS, /* [nconmax] int32 */
  MYO_F_CONTACT_DIST,    /* [nconmax] */
  MYO_F_CONTACT_POS,     /* [nconmax,3] */
  MYO_F_CONTACT_FRAME,   /* [nconmax,9] */
  MYO_F_CONTACT_INCLUDEMARGIN, /* [nconmax] */
  MYO_F_CONTACT_FRICTION,/* [nconmax,5] */
  MYO_F_CONTACT_SOLREF,  /* [nconmax,2] */
  MYO_F_CONTACT_SOLIMP,  /* [nconmax,5] */
  MYO_F_EFC_J,           /* [njmax,nv] */
  MYO_F_EFC_POS,         /* [njmax] */
  MYO_F_EFC_MARGIN,      /* [njmax] */
  MYO_F_EFC_D,           /* [njmax] */
  MYO_F_EFC_AREF,        /* [njmax] */
  MYO_F_EFC_FORCE,       /* [njmax] */
  MYO_F_EFC_TYPE,        /* [njmax] int32 (mjtConstraint) */
  MYO_F_SOLVER_NITER,    /* [1] int32 */
  MYO_F_STATUS,          /* [1] int32, MYO_WORLD_* bits, sticky until myo_reset */
  MYO_F_QFRC_APPLIED,    /* [nv]   optional input (mjData.qfrc_applied): added to qfrc_smooth when bound */
  MYO_F_SENSORDATA,      /* [nsensor] touch sensors (mj_sensorAcc): summed normal force of the contacts of the site's body whose
                            normal ray meets the site's volume */
  /* Optional per-world MODEL overrides (inputs).  The compiled model is one immutable table set shared by every world; these
   * cover what the reference's envs rewrite in mjModel per episode: the manipulated object's shape
   * (myosuite/envs/myo/myobase/reorient_sar_v0.py:207-245: geom_type, geom_size, geom_pos) and muscle weakness
   * (myosuite/envs/myo/base_v0.py:64-68: actuator_gainprm[:, 2]).  An unbound field, or an entry holding its "keep" value,
   * means the compiled model's value. */
  MYO_F_GEOM_TYPE_OVR,   /* [ngeom] int32 mjtGeom of a collision geom (sphere 2, capsule 3, ellipsoid 4, cylinder 5, box 6); < 0 = keep */
  MYO_F_GEOM_SIZE_OVR,   /* [ngeom,3] geom_size; size[0] <= 0 = keep */
  MYO_F_GEOM_POS_OVR,    /* [ngeom,3] geom_pos in its body's frame; NaN in x = keep */
  MYO_F_ACT_GAINSCALE,   /* [nu] factor on the gain of each actuator (a muscle's peak force); bound buffers start at 1 */
  MYO_F_COUNT
} myo_field;

typedef struct myo_model myo_model;   /* device-resident compiled model (immutable) */
typedef struct myo_batch myo_batch;   /* nworld worlds of one model on one device */

/* mjx.put_model: copy + preprocess the model onto `device`.  Fails with MYO_ERR_UNSUPPORTED like MJX's
 * put_model raises NotImplementedError (reference: myosuite/mjx/play.py:10). */
int myo_model_create(const myo_model_desc *desc, int device, myo_model **out);
void myo_model_destroy(myo_model *m);

/* Batch of `nworld` worlds; nconmax / njmax are per-world capacities (0 = model-derived default). */
int myo_batch_create(const myo_model *m, int nworld, int nconmax, int njmax, myo_batch **out);
void myo_batch_destroy(myo_batch *b);
int myo_batch_nconmax(const myo_batch *b);
int myo_batch_njmax(const myo_batch *b);
int myo_batch_smem_bytes(const myo_batch *b);   /* bytes of per-world scratch in the library-owned global buffer */
int myo_batch_team(const myo_batch *b);         /* lanes per world: always 1 (lane = world; the warps of a tile CTA share the items) */
int myo_batch_ctas_per_sm(const myo_batch *b);  /* always 1 (kept for ABI stability with round 1) */
int myo_batch_engine(const myo_batch *b);       /* always 1: the phase pipeline of lane = world kernels (round 1's one-launch kernel, 0, is gone) */

/* Per-world element count / element size of a field for this batch (0 elements = field absent for the model). */
int64_t myo_field_count(const myo_batch *b, int field);
int myo_field_itemsize(int field);
int myo_field_is_int(int field);   /* 1: int32 elements, 0: float32 */
const char *myo_field_name(int field);

/* Bind a caller-owned device buffer (e.g. a DLPack-exported torch / jax array) of
 * myo_field_count(field) * nworld elements, laid out [count][nworld], to `field`; zero-copy, the library never frees it.
 * dev_ptr == NULL unbinds (an unbound output field is simply not materialised). */
int myo_batch_bind(myo_batch *b, int field, void *dev_ptr, int64_t nbytes);

/* mj_forward on every world (reference: sim.forward() at myosuite/robot/robot.py:979; brax pipeline_init at
 * myosuite/mjx/myodm_v0.py:159): all outputs from the bound state, state not advanced. */
int myo_forward(myo_batch *b, void *cuda_stream);

/* mj_forward on the worlds whose mask entry is non-zero only (NULL = all); the others are not touched. */
int myo_forward_masked(myo_batch *b, const int32_t *world_mask_dev, void *cuda_stream);

/* mj_step1-style refresh of the masked worlds (NULL = all): the position- and velocity-dependent fields (kinematics, com,
 * tendon lengths / moment arms / velocities, actuator_length / actuator_velocity, cvel, cdof_dot, qfrc_bias, qfrc_passive)
 * are recomputed from the current state; contacts, constraint rows, actuator_force, qacc and the warm start keep the values
 * of the last full pass.  This is what the classic path's observations see: dm_control's Physics.step finishes with
 * mj_step1 after the last mj_step (reference: self.sim.step(substeps), myosuite/physics/mj_sim_scene.py:55), so xpos,
 * xipos, cvel, actuator_length ... describe the NEW state while actuator_force is the last step's. */
int myo_forward_posvel(myo_batch *b, const int32_t *world_mask_dev, void *cuda_stream);

/* nsubsteps x mj_step on every world with ctrl held (reference: jit_step at myosuite/mjx/play.py:47;
 * sim.advance(substeps=n_frames) at myosuite/robot/robot.py:880-882; pipeline_step at myodm_v0.py:287).
 * Output fields describe the last substep's forward pass, as mjData does after mj_step. */
int myo_step(myo_batch *b, int nsubsteps, void *cuda_stream);

/* Count-aware outputs (opt-in, off by default).  MJX and this library pad contact_* and efc_* to nconmax / njmax and rewrite the
 * padding (zeros; efc_type / contact_geom / contact_efc_address: -1) on every step, which is 60 % of those fields' bytes in
 * the MyoHand rollout.  A caller that reads them only up to ncon / nefc -- as mjData consumers do -- can switch the padding
 * off: with on != 0 the records of the contacts / rows that exist are written and the rest of the buffers is left alone. */
int myo_batch_set_count_aware(myo_batch *b, int on);

/* Kept for ABI stability with round 1 (whose separate wide-output kernel could be pipelined under the next step): every
 * field is now written, as full lines, by the pipeline phase that produces it, and is complete in stream order when
 * myo_step returns (mjx.step semantics).  Both calls are accepted and do nothing. */
int myo_batch_set_pipelined(myo_batch *b, int on);
int myo_outputs_join(myo_batch *b, void *cuda_stream);

/* mj_resetData + mj_forward on worlds whose mask entry is non-zero (NULL = all): qpos=qpos0, qvel=act=0,
 * time=0, warmstart=0, status=0 (reference: myosuite/robot/robot.py:976-979). world_mask is int32[nworld] on device. */
int myo_reset(myo_batch *b, const int32_t *world_mask_dev, void *cuda_stream);

/* Host-buffer convenience used by the end-to-end path: H2D of qpos/qvel/act/ctrl (any may be NULL), step,
 * D2H of qpos/qvel/act, all in the [k][nworld] layout.  Pinned host memory makes the copies asynchronous;
 * when wide output fields are bound the D2H copies wait for the step kernel only and run on a library-owned side
 * stream under the wide-output kernel (the state is final by then); the caller's stream joins it and is
 * synchronised at the end. */
int myo_step_host(myo_batch *b, int nsubsteps, const float *qpos, const float *qvel, const float *act,
                  const float *ctrl, float *qpos_out, float *qvel_out, float *act_out, void *cuda_stream);

/* ------------------------------------------------------------------------------------------------------------------
 * Fused env layer (SURVEY 8f rows n1 + n3): the elementwise work either side of the physics step of the reference's
 * pose environments, as two small kernels around myo_step -- no host round trip per env step.
 *   pre  : action -> ctrl      BaseV0.step, myosuite/envs/myo/base_v0.py:81-110 (muscle squash 1/(1+exp(-5(a-.5))),
 *                               fatigue 3CC-r, reafferentation EIP->EPL) and Robot.process_actuator "default_robot"
 *                               normalisation, myosuite/robot/robot.py:771-783
 *          fatigue              CumulativeFatigue.compute_act, myosuite/envs/myo/fatigue.py:61-108
 *   step : frame_skip x mj_step with ctrl held   myosuite/robot/robot.py:880-882
 *   post : obs / reward terms / done             PoseEnvV0.get_obs_dict / get_reward_dict,
 *                                                myosuite/envs/myo/myobase/pose_v0.py:98-138
 *          masked auto-reset with target resampling   PoseEnvV0.reset / get_target_pose, pose_v0.py:141-168,190-251
 * Random draws are a counter-based hash of (seed, global world index, episode, stream, index), so a world's episode
 * sequence does not depend on the batch size or on the number of GPUs.  All env buffers are caller-owned device
 * memory in the [k][nworld] layout, bound with myo_env_bind. */
enum { MYO_COND_NONE = 0, MYO_COND_SARCOPENIA = 1, MYO_COND_FATIGUE = 2, MYO_COND_REAFFERENTATION = 3 };
enum { MYO_ENV_POSE = 0, MYO_ENV_WALK = 1, MYO_ENV_REORIENT = 2 };   /* PoseEnvV0 (pose_v0.py), WalkEnvV0 (walk_v0.py), ProprioceptiveEnvV0 (reorient_sar_v0.py) */
enum {   /* obs_dict keys: pose_v0.py:106-116, walk_v0.py:263-283 */
  MYO_OBS_TIME = 0, MYO_OBS_QPOS = 1, MYO_OBS_QVEL = 2, MYO_OBS_ACT = 3, MYO_OBS_POSE_ERR = 4,
  MYO_OBS_QPOS_WITHOUT_XY = 5, MYO_OBS_COM_VEL = 6, MYO_OBS_TORSO_ANGLE = 7, MYO_OBS_FEET_HEIGHTS = 8, MYO_OBS_HEIGHT = 9,
  MYO_OBS_FEET_REL_POSITIONS = 10, MYO_OBS_PHASE_VAR = 11, MYO_OBS_MUSCLE_LENGTH = 12, MYO_OBS_MUSCLE_VELOCITY = 13,
  MYO_OBS_MUSCLE_FORCE = 14,
  /* reorient_sar_v0.py:115-137 */
  MYO_OBS_HAND_JNT = 15, MYO_OBS_OBJ_POS = 16, MYO_OBS_OBJ_VEL = 17, MYO_OBS_OBJ_ROT = 18, MYO_OBS_OBJ_DES_ROT = 19,
  MYO_OBS_OBJ_ERR_POS = 20, MYO_OBS_OBJ_ERR_ROT = 21, MYO_OBS_MLEN = 22, MYO_OBS_MVEL = 23, MYO_OBS_MFORCE = 24
};
#define MYO_MAX_OBS_KEYS 16
enum { MYO_TARGET_FIXED = 0, MYO_TARGET_GENERATE = 1 };
enum { MYO_RESET_INIT = 0, MYO_RESET_RANDOM = 1 };
/* slots of MYO_E_RWD_TERMS.  pose: pose, bonus, penalty, act_reg, sparse, solved, done, -.
 * walk (walk_v0.py:297-310): vel_reward, cyclic_hip, ref_rot, joint_angle_rew, act_mag, sparse, solved, done. */
enum { MYO_RWD_POSE = 0, MYO_RWD_BONUS, MYO_RWD_PENALTY, MYO_RWD_ACT_REG, MYO_RWD_SPARSE, MYO_RWD_SOLVED, MYO_RWD_DONE, MYO_RWD_SPARE, MYO_RWD_COUNT };
/* reorient (reorient_sar_v0.py:139-168): pos_align, rot_align, act_reg, drop, bonus, sparse, solved, done */
enum { MYO_RWD_R_POS_ALIGN = 0, MYO_RWD_R_ROT_ALIGN, MYO_RWD_R_ACT_REG, MYO_RWD_R_DROP, MYO_RWD_R_BONUS, MYO_RWD_R_SPARSE, MYO_RWD_R_SOLVED, MYO_RWD_R_DONE };
enum { MYO_RWD_W_VEL = 0, MYO_RWD_W_CYCLIC_HIP, MYO_RWD_W_REF_ROT, MYO_RWD_W_JOINT_ANGLE, MYO_RWD_W_ACT_MAG, MYO_RWD_W_SPARSE, MYO_RWD_W_SOLVED, MYO_RWD_W_DONE };

typedef enum myo_env_field {
  MYO_E_ACTION = 0,   /* [nu]            input of myo_env_step */
  MYO_E_OBS,          /* [nobs]          concatenation of the obs_keys entries, in order */
  MYO_E_REWARD,       /* [1]             rwd_dict["dense"] */
  MYO_E_DONE,         /* [1]             1.0 where the episode ended on this step (done or horizon) */
  MYO_E_RWD_TERMS,    /* [MYO_RWD_COUNT] pose, bonus, penalty, act_reg, sparse, solved */
  MYO_E_TARGET,       /* [ntarget]       target_jnt_value per world */
  MYO_E_FATIGUE,      /* [3, nmuscle]    MA, MR, MF (only with MYO_COND_FATIGUE) */
  MYO_E_STEPS,        /* [1] int32       env steps since the world's last reset */
  MYO_E_EPISODE,      /* [1] int32       episodes started by the world (the RNG counter) */
  MYO_E_COUNT
} myo_env_field;

/* host arrays; copied at myo_env_create */
typedef struct myo_env_desc {
  int32_t frame_skip, normalize_act, muscle_condition, epl_id, eip_id;
  int32_t nobs_keys, obs_keys[MYO_MAX_OBS_KEYS];
  int32_t ntarget, target_type, reset_type, auto_reset, horizon, world_offset;
  uint64_t seed;
  float pose_thd, far_th, w_pose, w_bonus, w_act_reg, w_penalty;
  float fatigue_F, fatigue_R, fatigue_r;
  const int32_t *target_qposadr;   /* [ntarget] */
  const float *target_range;       /* [ntarget,2] (generate) */
  const float *target_value;       /* [ntarget]   (fixed) */
  const float *init_qpos, *init_qvel;        /* [nq], [nv] */
  const float *reset_lo, *reset_hi;          /* [nq] uniform reset range (jnt_range) */
  const int32_t *act_is_muscle;              /* [nu] actuator_dyntype == mjDYN_MUSCLE */
  const float *ctrl_lo, *ctrl_hi;            /* [nu] actuator_ctrlrange */
  const float *tau_act, *tau_deact;          /* [nu] actuator_dynprm[:,0], [:,1] */
  /* ---- kind == MYO_ENV_WALK (myosuite/envs/myo/myobase/walk_v0.py:229-257 setup, :263-310 obs / reward, :312-350 reset,
   * :352-478 helpers).  Needs the fields xpos, xquat, xipos, cvel, actuator_length / velocity / force bound on the batch. */
  int32_t kind;
  int32_t foot_l_bid, foot_r_bid, pelvis_bid, torso_bid;   /* talus_l, talus_r, pelvis, torso */
  int32_t hip_flex_adr[2];                                 /* qpos address of hip_flexion_l, hip_flexion_r */
  int32_t hip_angle_adr[4];                                /* hip_adduction_l / _r, hip_rotation_l / _r */
  float min_height, max_rot, hip_period, target_x_vel, target_y_vel, target_rot[4], reset_noise;
  float w_vel, w_done, w_cyclic, w_ref_rot, w_joint_angle;
  const float *body_mass;                                  /* [nbody] */
  const float *alt_qpos, *alt_qvel;                        /* [nq], [nv]: second start state of reset_type random (key 3) */
  /* ---- kind == MYO_ENV_REORIENT (myosuite/envs/myo/myobase/reorient_sar_v0.py:67-173).  Needs xpos, site_xpos, geom_xpos,
   * actuator_length / velocity / force bound on the batch.  MYO_E_TARGET holds obj_des_rot [3] per world: the XML's target
   * pose (target_value) or, with target_type generate, the orientation drawn at reset as Geometries8EnvV0.reset does
   * (euler (U(-1,1), U(-0.8,1.2), 0) -> quat, :251-257) applied to the target's top - bottom axis (target_axis). */
  int32_t obj_bid, eps_ball_sid, obj_top_gid, obj_bot_gid;
  float pen_length, tar_length, target_axis[3];
  float w_pos_align, w_rot_align, w_r_act_reg, w_drop, w_r_bonus;
  /* Geometries8EnvV0.reset (reorient_sar_v0.py:178-249): with randomize_shape every new episode draws the object's collision
   * geom (obj_gid) as a capsule, ellipsoid, cylinder or box in one of two sizes and moves its top / bottom marker geoms
   * accordingly, PER WORLD, through the batch's MYO_F_GEOM_TYPE_OVR / _SIZE_OVR / _POS_OVR buffers (which must be bound). */
  int32_t randomize_shape, obj_gid;
} myo_env_desc;

typedef struct myo_env myo_env;
int myo_env_create(myo_batch *b, const myo_env_desc *desc, myo_env **out);
void myo_env_destroy(myo_env *e);
int64_t myo_env_field_count(const myo_env *e, int field);   /* elements per world */
int myo_env_bind(myo_env *e, int field, void *dev_ptr, int64_t nbytes);
int myo_env_seed(myo_env *e, uint64_t seed);
/* reset the masked worlds (NULL = all): new target, new initial state, zero reward / done, fresh observation.
 * Derived mjData fields are refreshed by the next step (call myo_forward for them earlier). */
int myo_env_reset(myo_env *e, const int32_t *world_mask_dev, void *cuda_stream);
/* one env step of every world: pre kernel, frame_skip fused mj_steps, post kernel (3 or 4 launches, no sync).
 * The walk env adds the refresh of the position / velocity fields its observation reads (myo_forward_posvel) and a
 * masked mj_forward of the worlds it reset. */
int myo_env_step(myo_env *e, void *cuda_stream);
/* MujocoEnv.forward (obs + reward terms + done from the CURRENT state; no physics, no counters, no reset):
 * what the reference evaluates after reset and after every step (myosuite/envs/myo/base_v0.py:120). */
int myo_env_forward(myo_env *e, void *cuda_stream);

/* Known-answer hook for the device rotation helpers used by kinematics and free-joint integration (reference KATs:
 * myosuite/tests/mjx/test_quat_math.py against myosuite/utils/quat_math.py:15 mulQuat, :160 quat2mat, :209 rotVecQuat).
 * Device pointers: qa, qb [n,4] (w,x,y,z); prod [n,4] = qa*qb; mat [n,9] = quat2mat(normalise(qa));
 * rotated [n,3] = mat * qb[1:4]. */
int myo_kat_quat(int n, const float *qa, const float *qb, float *prod, float *mat, float *rotated, void *cuda_stream);

/* Allocation-free forward kinematics.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Runs mj_kinematics + mj_comPos of `nworld` worlds of model `m` from `qpos` and writes the requested position-stage
 * outputs -- the same arithmetic, in the same order, as myo_forward on a batch (the kin + posout kernels fused into one
 * launch), so the results are bit-identical to the xpos / xquat / xmat / site_xpos / geom_xpos fields of a batch.
 *   qpos      [nq][nworld]        device, float32, world axis fastest (the layout of every field above)
 *   xpos      [3 nbody][nworld]   device or NULL      xquat [4 nbody][nworld] or NULL      xmat [9 nbody][nworld] or NULL
 *   site_xpos [3 nsite][nworld]   device or NULL      geom_xpos [3 ngeom][nworld] or NULL  (at least one output)
 * The call itself allocates nothing (no cudaMalloc / cudaFree, no allocating library routine): the per-world scratch a
 * batch would own lives in one fixed 8 MiB device array inside the library, which every GPU that loads the library holds
 * until exit.  LIMIT: ceil(nworld / 32) * scratch floats per 32-world tile * 128 B <= 8 MiB (MyoHand: about 1300 worlds);
 * a larger nworld is rejected with MYO_ERR_ARG and a message that states the model's maximum.  The model itself was
 * allocated by myo_model_create, before.  Enqueued on `cuda_stream` (a memset of the scratch in use, the parameter upload,
 * one kernel); no synchronisation.  All calls share the array and one parameter slot: host-side calls are serialised by a
 * mutex, but calls that OVERLAP ON THE GPU (different streams, no ordering between them) race -- order them. */
int myo_kinematics_noalloc(const myo_model *m, int nworld, const float *qpos, float *xpos, float *xquat, float *xmat,
                           float *site_xpos, float *geom_xpos, void *cuda_stream);

/* number of kernels launched by this library since load (for bench.py's gpu_launches) */
int64_t myo_launch_count(void);

/* thread-local message for the last non-zero status */
const char *myo_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* MYOSTEP_H_ */
