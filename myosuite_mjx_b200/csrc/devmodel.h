// Device-side compiled model ("plan") of the batched musculoskeletal step.  Built once on the host from the
// mjModel-shaped myo_model_desc (plan.cu) and shared read-only by every world; all tables below are small
// (tens of KB) and stay L1/L2 resident.
//
// Design notes (what differs from the array-per-stage layout of MJX / MuJoCo; DESIGN.md has the full story):
//  * weld groups are folded at compile time: only bodies with joints ("frames", DFS order, frame 0 = world)
//    are moved by kinematics; every body / geom / site carries a constant offset to its frame;
//  * each frame carries the composite inertia of its weld group, so CRB / RNE run over frames;
//  * tendon paths are flattened into "units" (site-site, or site-geom-site), their straight segments into
//    slots (point, direction) and the moment arms into a static sparse pattern (nnz list) whose entries are
//    Pluecker-line . cdof products summed in a fixed order (deterministic, no atomics);
//  * the LDL^T factorisation is a precomputed list of multiply-subtract ops per column;
//  * collision candidates are a static, ordered pair list with pre-mixed contact parameters.
#ifndef MYO_DEVMODEL_H_
#define MYO_DEVMODEL_H_

#include <stdint.h>

#include "../../include/myostep.h"

#define MYO_MAXROOT 4

enum { UNIT_SITE = 0, UNIT_SPHERE = 4, UNIT_CYLINDER = 5 };          // = mjtWrap values for the geom kinds
enum { ENT_ALWAYS = 0, ENT_NOWRAP = 1, ENT_WRAP = 2 };

struct DevModel {
  // ---- sizes
  int nq, nv, nu, na, nbody, njnt, ngeom, nsite, nten, nwrap, nM;
  int nsensor;                                 // touch sensors
  const int *sens_site, *sens_body, *sens_type; // [nsensor]: site id, its body, site shape (mjtGeom)
  const float *sens_size;                      // [nsensor,3]
  int nframe, nlevel, nroot, ndlevel;
  int nunit, nslot, nentry, nnz;
  int ncg, npair, nmprpair, neq, nlimj, nlimt;   // nmprpair: candidate pairs whose types need the convex narrow phase
  int nldlop;
  int nconmax, njmax, maxbase;
  // ---- options
  float timestep, gravity[3], tolerance, ls_tolerance, meaninertia, impratio;
  int iterations, ls_iterations, disableflags;
  int any_damping;
  // ---- frames (moving bodies; DFS order)
  const int *fr_parent, *fr_body, *fr_jntadr, *fr_jntnum, *fr_dofadr, *fr_dofnum, *fr_subend, *fr_root;
  const float *fr_pre_pos, *fr_pre_quat;       // parent frame -> this body before its joints
  const float *fr_mass, *fr_com, *fr_inertia;  // composite of the weld group, frame-local (I: xx yy zz xy xz yz about com)
  const int *fr_mask;                          // [nframe][2] bit i set: dof i moves this frame
  const int *lvl_start, *lvl_frames;           // frames grouped by depth
  const float *root_const;                     // [nroot][4]: sum(m*xipos) of world-welded bodies of the tree, total mass
  const int *root_body;                        // [nroot] body id of the root body
  const int *root_fstart, *root_fend;          // frame range of each root tree
  // ---- joints / dofs
  const int *jnt_type, *jnt_qposadr, *jnt_dofadr, *jnt_frame;
  const float *jnt_pos, *jnt_axis, *qpos0;
  const int *dof_frame, *dof_jnt, *dof_Madr, *dof_parent, *dof_rowlen;
  const float *dof_armature, *dof_damping, *dof_invweight0;
  const int *M_row, *M_col;                    // [nM] sparse lower-triangular pattern (row dof, ancestor dof)
  const int *dof_prevcount;                    // [nv] number of leading chain entries (self first) to skip for cdof_dot (1, or up to 3 inside a free joint's rotation block)
  // LDL plan: columns processed nv-1..0; ops of column k are [ldl_start[k], ldl_start[k+1])
  const int *ldl_start, *ldl_dst, *ldl_a, *ldl_b;
  // solve plan: dofs grouped by depth; descendants list per dof
  const int *dlvl_start, *dlvl_dofs;           // [ndlevel+1], [nv]
  const int *desc_start, *desc_dof, *desc_idx; // per dof j: (descendant i, index of L[i,j] in qLD)
  // ---- bodies / geoms / sites (offsets relative to their frame)
  const int *body_frame, *body_root, *body_subend, *body_parent;
  const float *body_off_pos, *body_off_quat, *body_ipos, *body_iquat, *body_mass, *body_inertia, *body_subtreemass;
  const int *geom_frame, *site_frame, *geom_body;
  const float *geom_lpos, *geom_lquat, *site_lpos, *site_lquat;
  const float *body_off_mat, *body_imat, *geom_lmat, *site_lmat;   // 3x3 row-major rotations of the offsets above
  // ---- tendons
  const int *unit_tendon, *unit_type, *unit_s0, *unit_s1, *unit_geom, *unit_side, *unit_slot;
  const int *unit_woff, *unit_gord, *unit_end, *gunit, *pul_off, *pul_gord, *ten_woff, *ten_gord;   // path-point outputs (plan.cu)
  int ngunit, npulley;
  const int *unit_order;                       // [nunit] site-site units first, then the units that wrap a geom
  const float *unit_div, *unit_radius;
  const int *ten_ustart;                       // [nten+1]
  const int *ent_slot, *ent_dof, *ent_kind, *ent_unit;
  const float *ent_scale;                      // +-1/divisor
  const int *nz_estart, *nz_tendon, *nz_dof;   // [nnz+1], [nnz], [nnz]
  const int *ten_nzstart;                      // [nten+1]
  const int *ten_dense, *act_dense;            // [nten][nv], [nu][nv]: source of each dense ten_J / actuator_moment entry (sparse index, -1 zero, -2 gear)
  const int *col_start, *col_nz;               // [nv+1], [nnz]  column view of the pattern
  const float *ten_range, *ten_stiffness, *ten_damping, *ten_lengthspring;
  // ---- actuators
  const int *act_trntype, *act_trnid, *act_dyntype, *act_gaintype, *act_biastype, *act_flags, *act_actadr;
  const float *act_gear, *act_dynprm, *act_gainprm, *act_biasprm, *act_ctrlrange, *act_forcerange, *act_actrange,
      *act_lengthrange, *act_acc0;
  const float *jnt_stiffness, *qpos_spring;
  const int *act_same_prm;                     // gainprm == biasprm (one muscle curve evaluation gives gain and bias)
  const int *tact_start, *tact_list;           // [nten+1]: actuators driving each tendon
  const int *dact_start, *dact_list;           // [nv+1]: joint-transmission actuators on each dof
  int any_tendon_passive;
  // ---- collision
  const int *cg_geom, *cg_type, *cg_frame, *cg_body;
  const float *cg_lpos, *cg_lmat, *cg_size, *cg_rbound, *cg_invweight;   // invweight: [ncg][2]
  const int *cg_meshadr, *cg_meshnum;          // [ncg] hull vertices of a mesh collision geom in mesh_vert (0, 0 for primitives)
  const float *mesh_vert;                      // [nmeshvert][3]
  const int *pair_cg1, *pair_cg2, *pair_dim;
  const float *pair_prm;                       // [npair][16]: margin, gap, friction[5], solref[2], sanitised solimp[5], K, B
  // ---- constraints
  const int *eq_dof1, *eq_dof2, *eq_q1, *eq_q2;
  const float *eq_data, *eq_prm;               // eq_prm [neq][8]: K, B, solimp[5], invweight
  const int *limj_q, *limj_dof, *limj_jnt;
  const float *limj_prm;                       // [nlimj][12]: lo, hi, margin, K, B, solimp[5], invweight, pad
  const int *limt_ten;
  const int *dof_limidx, *ten_limidx;          // [nv], [nten]: index into the limited-joint / limited-tendon lists or -1
  const int *dofeq_start, *dofeq_list;         // [nv+1]: equality rows touching each dof, entry = 2*row + (0: dof1, 1: dof2)
  const float *limt_prm;                       // [nlimt][12]
};

// bit flags in act_flags
enum { ACT_CTRLLIMITED = 1, ACT_FORCELIMITED = 2, ACT_ACTLIMITED = 4 };

struct FieldPtrs {
  void *p[MYO_F_COUNT];
};

// per-world scratch layout (float offsets into the world's column of the [total][stride] scratch), computed on the host
struct SmemLayout {
  int total;      // floats per world
  // persistent region
  int qpos, qvel, act, ctrl, frames, xanchor, xaxis, rootcom, cinert, crb, cdof, cdof_dot, cvel, cacc, qM, qLD, qLDinv, qH,
      qHinv, tenlen, tenvel, tenJ, actlen, actvel, actforce, actdot, bias, passive, qfrc_act, smooth, qfrc_con, qacc_smooth,
      qacc, warm, tmpv, tforce;
  int counts;    // ncon, nefc, solver_niter of the last forward pass (int bits)
  int wstate;    // per-world registers that cross kernel boundaries in the phase pipeline: status (int bits), time, skip flag of mj_checkAcc (int bits)
  // overlay region A (tendon stage)
  int slots, ulen, uwrap, uwpt, gpre, upts, tmpbody;
  // overlay region B (collision / constraint / solver stage)
  int cgpose, cgshape, con, conJ, conU, conW, efc, H, Hsave, nvec, limrow, basev, cand, ctmp;
  int nv;        // copy of DevModel.nv (the launch sizes the shared-memory Hessian tiles from it)
  int hshared;   // set by the kernel variant at upload: Newton Hessian tiles live in shared memory
  int wantw;     // set at upload: a path-point field is bound, the tendon stage keeps tangent points and wrap counts
};

// batch-owned work buffers the kernels need besides the scratch
struct TileAux {
  int2* mpr_list;   // (world, survivor position) of the candidate pairs that need the convex narrow phase this pass
  int* mpr_count;   // entries appended so far (zeroed before every collision pass)
  int mpr_cap;
  long long* dbg;   // developer timing of the solver phases ([tiles][32] clock64 sums; MYO_NEWTON_TIMING), else null
  int tl_off;   // developer timeline (MYO_NEWTON_TIMING): dbg + tl_off = [kernel][start, end] in globaltimer ns
};

#endif  // MYO_DEVMODEL_H_
