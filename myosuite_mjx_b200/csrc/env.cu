// Fused env layer around the physics step (include/myostep.h "Fused env layer"; SURVEY 8f rows n1 + n3).
// Reference behaviour restated here (formulas, not code):
//   action -> ctrl            myosuite/envs/myo/base_v0.py:81-110, myosuite/robot/robot.py:771-783
//   fatigue (3CC-r)           myosuite/envs/myo/fatigue.py:61-108
//   obs / reward terms / done myosuite/envs/myo/myobase/pose_v0.py:98-138
//   reset, target sampling    myosuite/envs/myo/myobase/pose_v0.py:141-168,190-251
// Layout: every env buffer is [k][nworld] with the world axis fastest, like the mjData fields, so the lanes of a warp
// (32 consecutive worlds) read and write one full line per element.  All arithmetic is float32.
#include <cuda_runtime.h>

#include <cstring>
#include <string>
#include <vector>

#include "devmodel.h"
#include "internal.h"

struct EnvTables {   // device copies of the myo_env_desc arrays
  const int* target_qposadr; const float *target_range, *target_value, *init_qpos, *init_qvel, *reset_lo, *reset_hi;
  const int* act_is_muscle; const int* muscle_slot;   // muscle_slot[a]: index among the muscle actuators, -1 otherwise
  const float *ctrl_lo, *ctrl_hi, *tau_act, *tau_deact;
  const float *body_mass, *alt_qpos, *alt_qvel;   // walk
};
struct EnvParams {
  int nworld, nq, nv, na, nu, nmuscle;
  int frame_skip, normalize_act, cond, epl, eip;
  int nobs_keys, obs_keys[MYO_MAX_OBS_KEYS], nobs;
  int ntarget, target_type, reset_type, auto_reset, horizon, world_offset;
  unsigned long long seed;
  float dt, pose_thd, far_th, w_pose, w_bonus, w_act_reg, w_penalty, fat_F, fat_R, fat_r;
  EnvTables T;
  void* E[MYO_E_COUNT];   // env buffers
  float *qpos, *qvel, *act, *ctrl, *time, *warm; int* status;   // bound state fields of the batch
  // walk env
  int kind, nbody, foot_l, foot_r, pelvis, torso, hip_flex[2], hip_angle[4];
  float min_height, max_rot, hip_period, target_x_vel, target_y_vel, target_rot[4], reset_noise, w_vel, w_done, w_cyclic, w_ref_rot, w_joint_angle;
  const float *xpos, *xquat, *xipos, *cvel, *actlen, *actvel, *actfrc;   // bound derived fields the walk observation reads
  // reorient env
  int obj_bid, eps_sid, top_gid, bot_gid, shape8, obj_gid;
  int* gtype_ovr; float *gsize_ovr, *gpos_ovr;   // the batch's per-world geom override buffers (randomize_shape)
  float pen_length, tar_length, target_axis[3], w_pos_align, w_rot_align, w_r_act_reg, w_drop, w_r_bonus;
  const float *site_xpos, *geom_xpos;
  int* reset_mask;   // [nworld], library-owned: worlds reset by the last step (they need a full mj_forward)
};

struct myo_env {
  myo_batch* b;
  EnvParams P;
  std::vector<void*> owned;
};

namespace {
// counter-based uniform in [0, 1): splitmix64 finaliser of a key built from (seed, world, episode, stream, index)
__host__ __device__ inline float envUniform(unsigned long long seed, unsigned gworld, unsigned episode, unsigned stream, unsigned idx) {
  unsigned long long z = seed ^ ((unsigned long long)gworld * 0xD1B54A32D192ED03ull) ^ ((unsigned long long)episode * 0x8CB92BA72F3D8DD7ull) ^
                         ((((unsigned long long)stream << 32) | idx) * 0x9E3779B97F4A7C15ull);
  z += 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  z ^= z >> 31;
  return (float)(z >> 40) * (1.0f / 16777216.0f);
}

#define EF(field, k) ((float*)P.E[field])[(size_t)(k) * nw + w]
#define EI(field, k) ((int*)P.E[field])[(size_t)(k) * nw + w]
#define ST(ptr, k) (ptr)[(size_t)(k) * nw + w]

// ---- pre: action -> ctrl, one thread per (actuator, world)
__global__ void __launch_bounds__(256) envPreKernel(const EnvParams P) {
  const size_t nw = (size_t)P.nworld;
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nw * (size_t)P.nu) return;
  const int a = (int)(i / nw);
  const size_t w = i - (size_t)a * nw;
  const bool muscle = P.T.act_is_muscle[a] != 0;
  const bool squash = P.na > 0 && P.normalize_act;   // base_v0.py:85: only then are muscle entries squashed
  int src = a;
  bool zero = false;
  if (P.cond == MYO_COND_REAFFERENTATION) {          // base_v0.py:104-108: EPL takes EIP's command, EIP is silenced
    if (a == P.epl) src = P.eip;
    else if (a == P.eip) zero = true;
  }
  float x = EF(MYO_E_ACTION, src);
  if (squash) {
    if (P.T.act_is_muscle[src]) x = 1.0f / (1.0f + expf(-5.0f * (x - 0.5f)));
  } else if (P.normalize_act) {                       // robot.py:771-783 (default_robot, "pos"): mean + a * range / 2
    x = 0.5f * (P.T.ctrl_lo[src] + P.T.ctrl_hi[src]) + x * (0.5f * (P.T.ctrl_hi[src] - P.T.ctrl_lo[src]));
  }
  if (zero) x = 0.f;
  if (P.cond == MYO_COND_FATIGUE && muscle) {         // fatigue.py:61-108, target load TL = x
    const int m = P.T.muscle_slot[a], nm = P.nmuscle;
    float MA = EF(MYO_E_FATIGUE, m), MR = EF(MYO_E_FATIGUE, nm + m), MF = EF(MYO_E_FATIGUE, 2 * nm + m);
    const float TL = x, dt = P.dt, F = P.fat_F;
    const float LD = 1.0f / P.T.tau_act[a] * (0.5f + 1.5f * MA), LR = (0.5f + 1.5f * MA) / P.T.tau_deact[a];
    float C, rR;
    if (MA < TL) { C = MR > TL - MA ? LD * (TL - MA) : LD * MR; rR = P.fat_R; }
    else { C = LR * (TL - MA); rR = P.fat_r * P.fat_R; }
    const float lo = fmaxf(-MA / dt + F * MA, (MR - 1.0f) / dt + rR * MF);
    const float hi = fminf((1.0f - MA) / dt + F * MA, MR / dt + rR * MF);
    C = fminf(fmaxf(C, lo), hi);                      // np.clip(C, lo, hi)
    const float dMA = (C - F * MA) * dt, dMR = (-C + rR * MF) * dt, dMF = (F * MA - rR * MF) * dt;
    MA += dMA; MR += dMR; MF += dMF;
    EF(MYO_E_FATIGUE, m) = MA; EF(MYO_E_FATIGUE, nm + m) = MR; EF(MYO_E_FATIGUE, 2 * nm + m) = MF;
    x = MA;
  }
  ST(P.ctrl, a) = x;
}

// new episode for world w: target, initial state, counters (pose_v0.py:190-251; robot.py:976-979 for the state write)
__device__ void envResetWorld(const EnvParams& P, size_t w) {
  const size_t nw = (size_t)P.nworld;
  const unsigned gw = (unsigned)(P.world_offset + (int)w);
  const unsigned ep = (unsigned)EI(MYO_E_EPISODE, 0);
  if (P.kind == MYO_ENV_REORIENT) {
    // obj_des_rot: the XML's target axis, or a fresh desired orientation (reorient_sar_v0.py:251-257: euler (U(-1,1),
    // U(-0.8,1.2), 0) -> quat by myosuite/utils/quat_math.py:77-93, applied to the target's top - bottom axis)
    float d[3] = {P.T.target_value[0], P.T.target_value[1], P.T.target_value[2]};
    if (P.target_type == MYO_TARGET_GENERATE) {
      const float e0 = __fadd_rn(-1.0f, __fmul_rn(2.0f, envUniform(P.seed, gw, ep, 0u, 0u))), e1 = __fadd_rn(-0.8f, __fmul_rn(2.0f, envUniform(P.seed, gw, ep, 0u, 1u)));
      const float ai = 0.f, aj = -0.5f * e1, ak = 0.5f * e0;
      const float si = sinf(ai), sj = sinf(aj), sk = sinf(ak), ci = cosf(ai), cj = cosf(aj), ck = cosf(ak);
      const float cc = ci * ck, cs = ci * sk, sc = si * ck, ss = si * sk;
      const float qw = cj * cc + sj * ss, qz = cj * sc - sj * cs, qy = -(cj * ss + sj * cc), qx = cj * cs - sj * sc;
      const float ax = P.target_axis[0], ay = P.target_axis[1], az = P.target_axis[2];
      // rotate the axis by the unit quaternion: v + 2 q_v x (q_v x v + w v)
      const float tx = qy * az - qz * ay + qw * ax, ty = qz * ax - qx * az + qw * ay, tz = qx * ay - qy * ax + qw * az;
      d[0] = (ax + 2.f * (qy * tz - qz * ty)) / P.tar_length; d[1] = (ay + 2.f * (qz * tx - qx * tz)) / P.tar_length; d[2] = (az + 2.f * (qx * ty - qy * tx)) / P.tar_length;
    }
    for (int k = 0; k < 3; k++) EF(MYO_E_TARGET, k) = d[k];
    if (P.shape8) {
      // Geometries8EnvV0.reset (reorient_sar_v0.py:178-245): geom_type uniform in {3, 4, 5, 6}, one of two sizes, the top /
      // bottom markers at +-1.3 size[1] (capsule), +-size[1] (cylinder), +-size[2] (ellipsoid, box) on the object's axis
      const int type = 3 + min(3, (int)(4.f * envUniform(P.seed, gw, ep, 5u, 0u))), big = envUniform(P.seed, gw, ep, 5u, 1u) < 0.5f ? 0 : 1;
      float s0, s1, s2;
      if (type == 3) { s0 = big ? 0.019f : 0.013f; s1 = s2 = big ? 0.040f : 0.025f; }
      else if (type == 4) { s0 = big ? 0.019f : 0.011f; s1 = s2 = big ? 0.040f : 0.025f; }
      else if (type == 5) { s0 = big ? 0.019f : 0.013f; s1 = s2 = big ? 0.040f : 0.025f; }
      else { s0 = s1 = s2 = big ? 0.023f : 0.017f; }
      const float z = type == 3 ? 1.3f * s1 : (type == 5 ? s1 : s2);
      ST(P.gtype_ovr, P.obj_gid) = type;
      ST(P.gsize_ovr, 3 * P.obj_gid) = s0; ST(P.gsize_ovr, 3 * P.obj_gid + 1) = s1; ST(P.gsize_ovr, 3 * P.obj_gid + 2) = s2;
      ST(P.gpos_ovr, 3 * P.top_gid) = 0.f; ST(P.gpos_ovr, 3 * P.top_gid + 1) = 0.f; ST(P.gpos_ovr, 3 * P.top_gid + 2) = z;
      ST(P.gpos_ovr, 3 * P.bot_gid) = 0.f; ST(P.gpos_ovr, 3 * P.bot_gid + 1) = 0.f; ST(P.gpos_ovr, 3 * P.bot_gid + 2) = -z;
    }
  } else
  for (int k = 0; k < P.ntarget; k++) {
    float t;
    if (P.target_type == MYO_TARGET_GENERATE) {
      const float lo = P.T.target_range[2 * k], hi = P.T.target_range[2 * k + 1];
      t = __fadd_rn(lo, __fmul_rn(hi - lo, envUniform(P.seed, gw, ep, 0u, (unsigned)k)));   // no fma: bit-equal to numpy float32
    } else t = P.T.target_value[k];
    EF(MYO_E_TARGET, k) = t;
  }
  if (P.kind == MYO_ENV_WALK) {
    // walk_v0.py:312-350: "random" starts from one of two keyframes (left or right knee flexed) plus N(0, 0.02) noise on
    // every coordinate but the root height and orientation; otherwise the given start state
    const bool alt = P.reset_type == MYO_RESET_RANDOM && !(envUniform(P.seed, gw, ep, 2u, 0u) < 0.5f);
    const float* q0 = alt ? P.T.alt_qpos : P.T.init_qpos;
    const float* v0 = alt ? P.T.alt_qvel : P.T.init_qvel;
    for (int j = 0; j < P.nq; j++) {
      float q = q0[j];
      if (P.reset_type == MYO_RESET_RANDOM && !(j >= 2 && j < 7)) {
        const float u1 = 1.0f - envUniform(P.seed, gw, ep, 3u, (unsigned)j), u2 = envUniform(P.seed, gw, ep, 4u, (unsigned)j);
        q += P.reset_noise * sqrtf(-2.0f * logf(u1)) * cosf(6.2831853071795865f * u2);   // Box-Muller
      }
      ST(P.qpos, j) = q;
    }
    for (int j = 0; j < P.nv; j++) { ST(P.qvel, j) = v0[j]; ST(P.warm, j) = 0.f; }
  } else {
    for (int j = 0; j < P.nq; j++) {
      float q = P.T.init_qpos[j];
      if (P.reset_type == MYO_RESET_RANDOM) {
        const float lo = P.T.reset_lo[j], hi = P.T.reset_hi[j];
        q = __fadd_rn(lo, __fmul_rn(hi - lo, envUniform(P.seed, gw, ep, 1u, (unsigned)j)));
      }
      ST(P.qpos, j) = q;
    }
    for (int j = 0; j < P.nv; j++) { ST(P.qvel, j) = P.T.init_qvel[j]; ST(P.warm, j) = 0.f; }
  }
  for (int j = 0; j < P.na; j++) ST(P.act, j) = 0.f;
  for (int j = 0; j < P.nu; j++) ST(P.ctrl, j) = 0.f;
  ST(P.time, 0) = 0.f;
  if (P.status) ST(P.status, 0) = 0;
  if (P.cond == MYO_COND_FATIGUE)
    for (int m = 0; m < P.nmuscle; m++) { EF(MYO_E_FATIGUE, m) = 0.f; EF(MYO_E_FATIGUE, P.nmuscle + m) = 1.f; EF(MYO_E_FATIGUE, 2 * P.nmuscle + m) = 0.f; }
  EI(MYO_E_STEPS, 0) = 0;
  EI(MYO_E_EPISODE, 0) = (int)(ep + 1u);
}

// walk_v0.py:436-478: mass-weighted centre-of-mass velocity (x, y of -cvel's linear part) and height (z of xipos)
__device__ void walkCom(const EnvParams& P, size_t w, float* vx, float* vy, float* h) {
  const size_t nw = (size_t)P.nworld;
  float sx = 0.f, sy = 0.f, sz = 0.f, sm_ = 0.f;
  for (int b = 0; b < P.nbody; b++) {
    const float m = P.T.body_mass[b];
    sx = fmaf(m, -ST(P.cvel, 6 * b + 3), sx); sy = fmaf(m, -ST(P.cvel, 6 * b + 4), sy); sz = fmaf(m, ST(P.xipos, 3 * b + 2), sz);
    sm_ += m;
  }
  *vx = sx / sm_; *vy = sy / sm_; *h = sz / sm_;
}
__device__ __forceinline__ float walkPhase(const EnvParams& P, int steps) { const float x = (float)steps / P.hip_period; return x - floorf(x); }

// obs vector from the state (pose_v0.py:98-116): concatenation of the obs_keys entries
__device__ void envWriteObs(const EnvParams& P, size_t w) {
  const size_t nw = (size_t)P.nworld;
  int o = 0;
  for (int q = 0; q < P.nobs_keys; q++) {
    switch (P.obs_keys[q]) {
      case MYO_OBS_TIME: EF(MYO_E_OBS, o++) = ST(P.time, 0); break;
      case MYO_OBS_QPOS: for (int j = 0; j < P.nq; j++) EF(MYO_E_OBS, o++) = ST(P.qpos, j); break;
      case MYO_OBS_QVEL: for (int j = 0; j < P.nv; j++) EF(MYO_E_OBS, o++) = ST(P.qvel, j) * P.dt; break;
      case MYO_OBS_ACT:
        if (P.na > 0) { for (int j = 0; j < P.na; j++) EF(MYO_E_OBS, o++) = ST(P.act, j); }
        else { for (int j = 0; j < P.nq; j++) EF(MYO_E_OBS, o++) = 0.f; }   // pose_v0.py:108-112: zeros_like(qpos)
        break;
      case MYO_OBS_POSE_ERR: for (int k = 0; k < P.ntarget; k++) EF(MYO_E_OBS, o++) = EF(MYO_E_TARGET, k) - ST(P.qpos, P.T.target_qposadr[k]); break;
      // ---- walk_v0.py:263-283
      case MYO_OBS_QPOS_WITHOUT_XY: for (int j = 2; j < P.nq; j++) EF(MYO_E_OBS, o++) = ST(P.qpos, j); break;
      case MYO_OBS_COM_VEL: { float vx, vy, h; walkCom(P, w, &vx, &vy, &h); EF(MYO_E_OBS, o++) = vx; EF(MYO_E_OBS, o++) = vy; } break;
      case MYO_OBS_TORSO_ANGLE: for (int k = 0; k < 4; k++) EF(MYO_E_OBS, o++) = ST(P.xquat, 4 * P.torso + k); break;
      case MYO_OBS_FEET_HEIGHTS: EF(MYO_E_OBS, o++) = ST(P.xpos, 3 * P.foot_l + 2); EF(MYO_E_OBS, o++) = ST(P.xpos, 3 * P.foot_r + 2); break;
      case MYO_OBS_HEIGHT: { float vx, vy, h; walkCom(P, w, &vx, &vy, &h); EF(MYO_E_OBS, o++) = h; } break;
      case MYO_OBS_FEET_REL_POSITIONS:
        for (int k = 0; k < 3; k++) EF(MYO_E_OBS, o++) = ST(P.xpos, 3 * P.foot_l + k) - ST(P.xpos, 3 * P.pelvis + k);
        for (int k = 0; k < 3; k++) EF(MYO_E_OBS, o++) = ST(P.xpos, 3 * P.foot_r + k) - ST(P.xpos, 3 * P.pelvis + k);
        break;
      case MYO_OBS_PHASE_VAR: EF(MYO_E_OBS, o++) = walkPhase(P, max(EI(MYO_E_STEPS, 0) - 1, 0)); break;   // self.steps lags the step count by one (walk_v0.py:335-338)
      case MYO_OBS_MUSCLE_LENGTH: for (int j = 0; j < P.nu; j++) EF(MYO_E_OBS, o++) = ST(P.actlen, j); break;
      case MYO_OBS_MUSCLE_VELOCITY: for (int j = 0; j < P.nu; j++) EF(MYO_E_OBS, o++) = fminf(fmaxf(ST(P.actvel, j), -100.f), 100.f); break;
      case MYO_OBS_MUSCLE_FORCE: for (int j = 0; j < P.nu; j++) EF(MYO_E_OBS, o++) = fminf(fmaxf(ST(P.actfrc, j) / 1000.f, -100.f), 100.f); break;
      // ---- reorient_sar_v0.py:115-137
      case MYO_OBS_HAND_JNT: for (int j = 0; j < P.nq - 6; j++) EF(MYO_E_OBS, o++) = ST(P.qpos, j); break;
      case MYO_OBS_OBJ_POS: for (int k = 0; k < 3; k++) EF(MYO_E_OBS, o++) = ST(P.xpos, 3 * P.obj_bid + k); break;
      case MYO_OBS_OBJ_VEL: for (int j = P.nv - 6; j < P.nv; j++) EF(MYO_E_OBS, o++) = ST(P.qvel, j) * P.dt; break;
      case MYO_OBS_OBJ_ROT: for (int k = 0; k < 3; k++) EF(MYO_E_OBS, o++) = (ST(P.geom_xpos, 3 * P.top_gid + k) - ST(P.geom_xpos, 3 * P.bot_gid + k)) / P.pen_length; break;
      case MYO_OBS_OBJ_DES_ROT: for (int k = 0; k < 3; k++) EF(MYO_E_OBS, o++) = EF(MYO_E_TARGET, k); break;
      case MYO_OBS_OBJ_ERR_POS: for (int k = 0; k < 3; k++) EF(MYO_E_OBS, o++) = ST(P.xpos, 3 * P.obj_bid + k) - ST(P.site_xpos, 3 * P.eps_sid + k); break;
      case MYO_OBS_OBJ_ERR_ROT: for (int k = 0; k < 3; k++) EF(MYO_E_OBS, o++) = (ST(P.geom_xpos, 3 * P.top_gid + k) - ST(P.geom_xpos, 3 * P.bot_gid + k)) / P.pen_length - EF(MYO_E_TARGET, k); break;
      case MYO_OBS_MLEN: for (int j = 0; j < P.nu; j++) EF(MYO_E_OBS, o++) = ST(P.actlen, j); break;
      case MYO_OBS_MVEL: for (int j = 0; j < P.nu; j++) EF(MYO_E_OBS, o++) = ST(P.actvel, j); break;
      case MYO_OBS_MFORCE: for (int j = 0; j < P.nu; j++) EF(MYO_E_OBS, o++) = ST(P.actfrc, j); break;
    }
  }
}

// ---- post: reward terms, done, auto-reset, obs; one thread per world
__global__ void __launch_bounds__(128) envPostKernel(const EnvParams P, const int advance) {
  const size_t nw = (size_t)P.nworld;
  const size_t w = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (w >= nw) return;
  float d2 = 0.f;
  for (int k = 0; k < P.ntarget; k++) { const float e = EF(MYO_E_TARGET, k) - ST(P.qpos, P.T.target_qposadr[k]); d2 = fmaf(e, e, d2); }
  const float dist = sqrtf(d2);
  float a2 = 0.f;
  for (int j = 0; j < P.na; j++) { const float a = ST(P.act, j); a2 = fmaf(a, a, a2); }
  float act_mag = sqrtf(a2);
  if (P.na != 0) act_mag = act_mag / (float)P.na;
  const float pose = -dist, bonus = (dist < P.pose_thd ? 1.f : 0.f) + (dist < 1.5f * P.pose_thd ? 1.f : 0.f);
  const float penalty = dist > P.far_th ? -1.f : 0.f, act_reg = -act_mag;
  const bool done = dist > P.far_th;
  EF(MYO_E_RWD_TERMS, MYO_RWD_POSE) = pose; EF(MYO_E_RWD_TERMS, MYO_RWD_BONUS) = bonus; EF(MYO_E_RWD_TERMS, MYO_RWD_PENALTY) = penalty;
  EF(MYO_E_RWD_TERMS, MYO_RWD_ACT_REG) = act_reg; EF(MYO_E_RWD_TERMS, MYO_RWD_SPARSE) = -dist;
  EF(MYO_E_RWD_TERMS, MYO_RWD_SOLVED) = dist < P.pose_thd ? 1.f : 0.f;
  EF(MYO_E_RWD_TERMS, MYO_RWD_DONE) = done ? 1.f : 0.f; EF(MYO_E_RWD_TERMS, MYO_RWD_SPARE) = 0.f;
  // dense = sum over rwd_keys_wt in its dict order: pose, bonus, act_reg, penalty (pose_v0.py:16-21,134-136)
  EF(MYO_E_REWARD, 0) = ((P.w_pose * pose + P.w_bonus * bonus) + P.w_act_reg * act_reg) + P.w_penalty * penalty;
  const int steps = EI(MYO_E_STEPS, 0) + advance;
  if (advance) EI(MYO_E_STEPS, 0) = steps;
  const bool ended = done || (P.horizon > 0 && steps >= P.horizon);
  EF(MYO_E_DONE, 0) = ended ? 1.f : 0.f;
  if (ended && advance && P.auto_reset) envResetWorld(P, w);
  envWriteObs(P, w);
}

// ---- walk: reward terms, done, reset decision (walk_v0.py:285-310, 357-478); one thread per world.  Runs after the
// position / velocity refresh, so xipos, cvel and qpos describe the stepped state as they do on the classic path.
__global__ void __launch_bounds__(128) envWalkRewardKernel(const EnvParams P, const int advance) {
  const size_t nw = (size_t)P.nworld;
  const size_t w = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (w >= nw) return;
  const int steps0 = EI(MYO_E_STEPS, 0);               // the reference's self.steps while it evaluates this step's reward
  float vx, vy, height;
  walkCom(P, w, &vx, &vy, &height);
  const float ey = P.target_y_vel - vy, ex = P.target_x_vel - vx;
  const float vel_reward = expf(-(ey * ey)) + expf(-(ex * ex));
  const float ph = walkPhase(P, advance ? steps0 : max(steps0 - 1, 0));
  const float twopi = 6.2831853071795865f;
  const float d0 = 0.8f * cosf(ph * twopi + 3.14159265358979324f) - ST(P.qpos, P.hip_flex[0]), d1 = 0.8f * cosf(ph * twopi) - ST(P.qpos, P.hip_flex[1]);
  const float cyclic = sqrtf(d0 * d0 + d1 * d1);
  float r2 = 0.f;
  for (int k = 0; k < 4; k++) { const float e = 5.0f * (ST(P.qpos, 3 + k) - P.target_rot[k]); r2 = fmaf(e, e, r2); }
  const float ref_rot = expf(-sqrtf(r2));
  float mag = 0.f;
  for (int k = 0; k < 4; k++) mag += fabsf(ST(P.qpos, P.hip_angle[k]));
  const float joint_angle = expf(-5.0f * (mag * 0.25f));
  float a2 = 0.f;
  for (int j = 0; j < P.na; j++) { const float a = ST(P.act, j); a2 = fmaf(a, a, a2); }
  const float act_mag = P.na != 0 ? sqrtf(a2) / (float)P.na : 0.f;
  // done: centre of mass too low, or the root frame's x axis turned away (quat2mat(q)[0, 0], myosuite/utils/quat_math.py:160-183)
  const float qw = ST(P.qpos, 3), qx = ST(P.qpos, 4), qy = ST(P.qpos, 5), qz = ST(P.qpos, 6);
  const float Nq = qw * qw + qx * qx + qy * qy + qz * qz;
  const float r00 = Nq > 2.220446049250313e-16f ? 1.0f - (qy * (qy * (2.0f / Nq)) + qz * (qz * (2.0f / Nq))) : 1.0f;
  const bool done = height < P.min_height || fabsf(r00) > P.max_rot;
  EF(MYO_E_RWD_TERMS, MYO_RWD_W_VEL) = vel_reward; EF(MYO_E_RWD_TERMS, MYO_RWD_W_CYCLIC_HIP) = cyclic; EF(MYO_E_RWD_TERMS, MYO_RWD_W_REF_ROT) = ref_rot;
  EF(MYO_E_RWD_TERMS, MYO_RWD_W_JOINT_ANGLE) = joint_angle; EF(MYO_E_RWD_TERMS, MYO_RWD_W_ACT_MAG) = act_mag;
  EF(MYO_E_RWD_TERMS, MYO_RWD_W_SPARSE) = vel_reward; EF(MYO_E_RWD_TERMS, MYO_RWD_W_SOLVED) = vel_reward >= 1.0f ? 1.f : 0.f;
  EF(MYO_E_RWD_TERMS, MYO_RWD_W_DONE) = done ? 1.f : 0.f;
  // dense in the order of DEFAULT_RWD_KEYS_AND_WEIGHTS (walk_v0.py:200-206): vel_reward, done, cyclic_hip, ref_rot, joint_angle_rew
  EF(MYO_E_REWARD, 0) = (((P.w_vel * vel_reward + P.w_done * (done ? 1.f : 0.f)) + P.w_cyclic * cyclic) + P.w_ref_rot * ref_rot) + P.w_joint_angle * joint_angle;
  const int steps = steps0 + advance;
  if (advance) EI(MYO_E_STEPS, 0) = steps;
  const bool ended = done || (P.horizon > 0 && steps >= P.horizon);
  EF(MYO_E_DONE, 0) = ended ? 1.f : 0.f;
  const bool reset = ended && advance && P.auto_reset;
  if (reset) envResetWorld(P, w);
  if (P.reset_mask) P.reset_mask[w] = reset ? 1 : 0;
}
// ---- reorient: reward terms, done, reset decision (reorient_sar_v0.py:139-168); one thread per world, after the refresh
__global__ void __launch_bounds__(128) envReorientRewardKernel(const EnvParams P, const int advance) {
  const size_t nw = (size_t)P.nworld;
  const size_t w = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (w >= nw) return;
  float e2 = 0.f, dotv = 0.f, n1 = 0.f, n2 = 0.f;
  for (int k = 0; k < 3; k++) {
    const float e = ST(P.xpos, 3 * P.obj_bid + k) - ST(P.site_xpos, 3 * P.eps_sid + k); e2 = fmaf(e, e, e2);
    const float r = (ST(P.geom_xpos, 3 * P.top_gid + k) - ST(P.geom_xpos, 3 * P.bot_gid + k)) / P.pen_length, d = EF(MYO_E_TARGET, k);
    dotv = fmaf(r, d, dotv); n1 = fmaf(r, r, n1); n2 = fmaf(d, d, n2);
  }
  const float pos_align = sqrtf(e2);
  const float rot_align = dotv / (sqrtf(n1) * sqrtf(n2));   // calculate_cosine, myosuite/utils/vector_math.py:11-36
  const bool dropped = pos_align > 0.075f;
  float a2 = 0.f;
  for (int j = 0; j < P.na; j++) { const float a = ST(P.act, j); a2 = fmaf(a, a, a2); }
  const float act_mag = P.na != 0 ? sqrtf(a2) / (float)P.na : 0.f;
  const float close = pos_align < 0.075f ? 1.f : 0.f;
  const float bonus = 1.0f * (rot_align > 0.9f ? 1.f : 0.f) * close + 5.0f * (rot_align > 0.95f ? 1.f : 0.f) * close;
  EF(MYO_E_RWD_TERMS, MYO_RWD_R_POS_ALIGN) = -pos_align; EF(MYO_E_RWD_TERMS, MYO_RWD_R_ROT_ALIGN) = rot_align; EF(MYO_E_RWD_TERMS, MYO_RWD_R_ACT_REG) = -act_mag;
  EF(MYO_E_RWD_TERMS, MYO_RWD_R_DROP) = dropped ? -1.f : 0.f; EF(MYO_E_RWD_TERMS, MYO_RWD_R_BONUS) = bonus;
  EF(MYO_E_RWD_TERMS, MYO_RWD_R_SPARSE) = -pos_align + rot_align;
  EF(MYO_E_RWD_TERMS, MYO_RWD_R_SOLVED) = (rot_align > 0.95f && !dropped) ? 1.f : 0.f; EF(MYO_E_RWD_TERMS, MYO_RWD_R_DONE) = dropped ? 1.f : 0.f;
  // dense in the order of DEFAULT_RWD_KEYS_AND_WEIGHTS (reorient_sar_v0.py:37-43): pos_align, rot_align, act_reg, drop, bonus
  EF(MYO_E_REWARD, 0) = (((P.w_pos_align * -pos_align + P.w_rot_align * rot_align) + P.w_r_act_reg * -act_mag) + P.w_drop * (dropped ? -1.f : 0.f)) + P.w_r_bonus * bonus;
  const int steps = EI(MYO_E_STEPS, 0) + advance;
  if (advance) EI(MYO_E_STEPS, 0) = steps;
  const bool ended = dropped || (P.horizon > 0 && steps >= P.horizon);
  EF(MYO_E_DONE, 0) = ended ? 1.f : 0.f;
  const bool reset = ended && advance && P.auto_reset;
  if (reset) envResetWorld(P, w);
  if (P.reset_mask) P.reset_mask[w] = reset ? 1 : 0;
}
__global__ void __launch_bounds__(128) envObsKernel(const EnvParams P) {
  const size_t w = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (w >= (size_t)P.nworld) return;
  envWriteObs(P, w);
}

__global__ void __launch_bounds__(128) envResetKernel(const EnvParams P, const int* __restrict__ mask) {
  const size_t nw = (size_t)P.nworld;
  const size_t w = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (w >= nw) return;
  if (mask && !mask[w]) return;
  envResetWorld(P, w);
  EF(MYO_E_REWARD, 0) = 0.f; EF(MYO_E_DONE, 0) = 0.f;
  for (int k = 0; k < MYO_RWD_COUNT; k++) EF(MYO_E_RWD_TERMS, k) = 0.f;
  if (P.kind == MYO_ENV_POSE) envWriteObs(P, w);   // the walk / reorient observations need mj_forward of the new state first
}

template <class T> const T* toDevice(const T* host, size_t n, std::vector<void*>& owned) {
  if (!host || n == 0) return nullptr;
  void* p = nullptr;
  if (cudaMalloc(&p, n * sizeof(T)) != cudaSuccess) throw std::string("cudaMalloc failed");
  if (cudaMemcpy(p, host, n * sizeof(T), cudaMemcpyHostToDevice) != cudaSuccess) { cudaFree(p); throw std::string("cudaMemcpy failed"); }
  owned.push_back(p);
  return (const T*)p;
}
int64_t envFieldCount(const EnvParams& P, int field) {
  switch (field) {
    case MYO_E_ACTION: return P.nu;
    case MYO_E_OBS: return P.nobs;
    case MYO_E_REWARD: case MYO_E_DONE: case MYO_E_STEPS: case MYO_E_EPISODE: return 1;
    case MYO_E_RWD_TERMS: return MYO_RWD_COUNT;
    case MYO_E_TARGET: return P.ntarget;
    case MYO_E_FATIGUE: return P.cond == MYO_COND_FATIGUE ? 3 * P.nmuscle : 0;
  }
  return 0;
}
}  // namespace

extern "C" {

int myo_env_create(myo_batch* b, const myo_env_desc* d, myo_env** out) {
  if (!b || !d || !out) return myo_fail(MYO_ERR_ARG, "bad argument");
  const DevModel& D = b->D;
  if (d->frame_skip < 1) return myo_fail(MYO_ERR_ARG, "frame_skip must be >= 1");
  if (d->nobs_keys < 0 || d->nobs_keys > MYO_MAX_OBS_KEYS) return myo_fail(MYO_ERR_ARG, "too many obs keys");
  if (d->kind != MYO_ENV_POSE && d->kind != MYO_ENV_WALK && d->kind != MYO_ENV_REORIENT) return myo_fail(MYO_ERR_ARG, "unknown env kind");
  if (d->kind == MYO_ENV_REORIENT) {
    if (D.nq < 7 || D.nv < 6) return myo_fail(MYO_ERR_ARG, "the reorient env needs an object with six trailing coordinates");
    if (d->obj_bid < 0 || d->obj_bid >= D.nbody || d->eps_ball_sid < 0 || d->eps_ball_sid >= D.nsite || d->obj_top_gid < 0 || d->obj_top_gid >= D.ngeom ||
        d->obj_bot_gid < 0 || d->obj_bot_gid >= D.ngeom) return myo_fail(MYO_ERR_ARG, "reorient body / site / geom id out of range");
    if (!(d->pen_length > 0.f) || !(d->tar_length > 0.f) || d->ntarget != 3 || !d->target_value) return myo_fail(MYO_ERR_ARG, "reorient target description missing");
  }
  if (d->kind == MYO_ENV_WALK) {
    if (D.nq < 7 || !d->body_mass) return myo_fail(MYO_ERR_ARG, "the walk env needs a free root joint and body_mass");
    if (d->reset_type == MYO_RESET_RANDOM && (!d->alt_qpos || !d->alt_qvel)) return myo_fail(MYO_ERR_ARG, "alt_qpos / alt_qvel missing");
    const int bids[4] = {d->foot_l_bid, d->foot_r_bid, d->pelvis_bid, d->torso_bid};
    for (int k = 0; k < 4; k++) if (bids[k] < 0 || bids[k] >= D.nbody) return myo_fail(MYO_ERR_ARG, "walk body id out of range");
    for (int k = 0; k < 2; k++) if (d->hip_flex_adr[k] < 0 || d->hip_flex_adr[k] >= D.nq) return myo_fail(MYO_ERR_ARG, "hip_flex_adr out of range");
    for (int k = 0; k < 4; k++) if (d->hip_angle_adr[k] < 0 || d->hip_angle_adr[k] >= D.nq) return myo_fail(MYO_ERR_ARG, "hip_angle_adr out of range");
    if (!(d->hip_period > 0.f)) return myo_fail(MYO_ERR_ARG, "hip_period must be positive");
  }
  if (d->ntarget < 0 || (d->kind == MYO_ENV_POSE && d->ntarget > 0 && !d->target_qposadr)) return myo_fail(MYO_ERR_ARG, "target_qposadr missing");
  if (d->kind == MYO_ENV_POSE && d->target_type == MYO_TARGET_GENERATE && d->ntarget > 0 && !d->target_range) return myo_fail(MYO_ERR_ARG, "target_range missing");
  if (d->target_type == MYO_TARGET_FIXED && d->ntarget > 0 && !d->target_value) return myo_fail(MYO_ERR_ARG, "target_value missing");
  if (!d->init_qpos || !d->init_qvel) return myo_fail(MYO_ERR_ARG, "init_qpos / init_qvel missing");
  if (d->kind == MYO_ENV_POSE && d->reset_type == MYO_RESET_RANDOM && (!d->reset_lo || !d->reset_hi)) return myo_fail(MYO_ERR_ARG, "reset range missing");
  if (D.nu > 0 && (!d->act_is_muscle || !d->ctrl_lo || !d->ctrl_hi)) return myo_fail(MYO_ERR_ARG, "actuator tables missing");
  if (d->muscle_condition == MYO_COND_FATIGUE && (!d->tau_act || !d->tau_deact)) return myo_fail(MYO_ERR_ARG, "tau_act / tau_deact missing");
  if (d->muscle_condition == MYO_COND_REAFFERENTATION && (d->epl_id < 0 || d->epl_id >= D.nu || d->eip_id < 0 || d->eip_id >= D.nu))
    return myo_fail(MYO_ERR_ARG, "reafferentation needs the EPL and EIP actuator ids");
  if (d->kind == MYO_ENV_POSE)
    for (int k = 0; k < d->ntarget; k++) if (d->target_qposadr[k] < 0 || d->target_qposadr[k] >= D.nq) return myo_fail(MYO_ERR_ARG, "target_qposadr out of range");
  if (cudaSetDevice(b->device) != cudaSuccess) return myo_fail(MYO_ERR_CUDA, "cudaSetDevice failed: libmyostep has no CPU fallback");
  myo_env* e = new myo_env();
  e->b = b;
  EnvParams& P = e->P;
  memset(&P, 0, sizeof(P));
  P.nworld = b->nworld; P.nq = D.nq; P.nv = D.nv; P.na = D.na; P.nu = D.nu;
  P.frame_skip = d->frame_skip; P.normalize_act = d->normalize_act; P.cond = d->muscle_condition; P.epl = d->epl_id; P.eip = d->eip_id;
  P.nobs_keys = d->nobs_keys;
  P.ntarget = d->ntarget; P.target_type = d->target_type; P.reset_type = d->reset_type; P.auto_reset = d->auto_reset; P.horizon = d->horizon;
  P.world_offset = d->world_offset; P.seed = d->seed;
  P.dt = D.timestep * (float)d->frame_skip;
  P.pose_thd = d->pose_thd; P.far_th = d->far_th; P.w_pose = d->w_pose; P.w_bonus = d->w_bonus; P.w_act_reg = d->w_act_reg; P.w_penalty = d->w_penalty;
  P.fat_F = d->fatigue_F; P.fat_R = d->fatigue_R; P.fat_r = d->fatigue_r;
  P.kind = d->kind; P.nbody = D.nbody;
  P.shape8 = d->randomize_shape; P.obj_gid = d->obj_gid;
  if (P.shape8 && (d->kind != MYO_ENV_REORIENT || d->obj_gid < 0 || d->obj_gid >= D.ngeom)) return myo_fail(MYO_ERR_ARG, "randomize_shape needs the reorient env and a valid obj_gid");
  P.obj_bid = d->obj_bid; P.eps_sid = d->eps_ball_sid; P.top_gid = d->obj_top_gid; P.bot_gid = d->obj_bot_gid; P.pen_length = d->pen_length; P.tar_length = d->tar_length;
  for (int k = 0; k < 3; k++) P.target_axis[k] = d->target_axis[k];
  P.w_pos_align = d->w_pos_align; P.w_rot_align = d->w_rot_align; P.w_r_act_reg = d->w_r_act_reg; P.w_drop = d->w_drop; P.w_r_bonus = d->w_r_bonus;
  P.foot_l = d->foot_l_bid; P.foot_r = d->foot_r_bid; P.pelvis = d->pelvis_bid; P.torso = d->torso_bid;
  for (int k = 0; k < 2; k++) P.hip_flex[k] = d->hip_flex_adr[k];
  for (int k = 0; k < 4; k++) { P.hip_angle[k] = d->hip_angle_adr[k]; P.target_rot[k] = d->target_rot[k]; }
  P.min_height = d->min_height; P.max_rot = d->max_rot; P.hip_period = d->hip_period; P.target_x_vel = d->target_x_vel; P.target_y_vel = d->target_y_vel;
  P.reset_noise = d->reset_noise; P.w_vel = d->w_vel; P.w_done = d->w_done; P.w_cyclic = d->w_cyclic; P.w_ref_rot = d->w_ref_rot; P.w_joint_angle = d->w_joint_angle;
  for (int q = 0; q < d->nobs_keys; q++) {
    const int key = d->obs_keys[q];
    P.obs_keys[q] = key;
    if (key >= MYO_OBS_QPOS_WITHOUT_XY && key <= MYO_OBS_MUSCLE_FORCE && d->kind != MYO_ENV_WALK) { delete e; return myo_fail(MYO_ERR_ARG, "walk obs key on another env kind"); }
    if (key >= MYO_OBS_HAND_JNT && d->kind != MYO_ENV_REORIENT) { delete e; return myo_fail(MYO_ERR_ARG, "reorient obs key on another env kind"); }
    if (key == MYO_OBS_POSE_ERR && d->kind != MYO_ENV_POSE) { delete e; return myo_fail(MYO_ERR_ARG, "pose obs key on a walk env"); }
    switch (key) {
      case MYO_OBS_TIME: P.nobs += 1; break;
      case MYO_OBS_QPOS: P.nobs += D.nq; break;
      case MYO_OBS_QVEL: P.nobs += D.nv; break;
      case MYO_OBS_ACT: P.nobs += D.na > 0 ? D.na : D.nq; break;
      case MYO_OBS_POSE_ERR: P.nobs += d->ntarget; break;
      case MYO_OBS_QPOS_WITHOUT_XY: P.nobs += D.nq - 2; break;
      case MYO_OBS_COM_VEL: case MYO_OBS_FEET_HEIGHTS: P.nobs += 2; break;
      case MYO_OBS_TORSO_ANGLE: P.nobs += 4; break;
      case MYO_OBS_HEIGHT: case MYO_OBS_PHASE_VAR: P.nobs += 1; break;
      case MYO_OBS_FEET_REL_POSITIONS: P.nobs += 6; break;
      case MYO_OBS_MUSCLE_LENGTH: case MYO_OBS_MUSCLE_VELOCITY: case MYO_OBS_MUSCLE_FORCE: P.nobs += D.nu; break;
      case MYO_OBS_HAND_JNT: P.nobs += D.nq - 6; break;
      case MYO_OBS_OBJ_VEL: P.nobs += 6; break;
      case MYO_OBS_OBJ_POS: case MYO_OBS_OBJ_ROT: case MYO_OBS_OBJ_DES_ROT: case MYO_OBS_OBJ_ERR_POS: case MYO_OBS_OBJ_ERR_ROT: P.nobs += 3; break;
      case MYO_OBS_MLEN: case MYO_OBS_MVEL: case MYO_OBS_MFORCE: P.nobs += D.nu; break;
      default: delete e; return myo_fail(MYO_ERR_ARG, "unknown obs key");
    }
  }
  std::vector<int> slot(D.nu > 0 ? D.nu : 1, -1);
  for (int a = 0; a < D.nu; a++) if (d->act_is_muscle[a]) slot[a] = P.nmuscle++;
  try {
    P.T.target_qposadr = toDevice(d->target_qposadr, d->ntarget, e->owned);
    P.T.target_range = toDevice(d->target_range, 2 * (size_t)d->ntarget, e->owned);
    P.T.target_value = toDevice(d->target_value, d->ntarget, e->owned);
    P.T.init_qpos = toDevice(d->init_qpos, D.nq, e->owned);
    P.T.init_qvel = toDevice(d->init_qvel, D.nv, e->owned);
    P.T.reset_lo = toDevice(d->reset_lo, D.nq, e->owned);
    P.T.reset_hi = toDevice(d->reset_hi, D.nq, e->owned);
    P.T.act_is_muscle = toDevice(d->act_is_muscle, D.nu, e->owned);
    P.T.muscle_slot = toDevice(slot.data(), D.nu, e->owned);
    P.T.ctrl_lo = toDevice(d->ctrl_lo, D.nu, e->owned);
    P.T.ctrl_hi = toDevice(d->ctrl_hi, D.nu, e->owned);
    P.T.tau_act = toDevice(d->tau_act, D.nu, e->owned);
    P.T.tau_deact = toDevice(d->tau_deact, D.nu, e->owned);
    if (d->kind == MYO_ENV_WALK) {
      P.T.body_mass = toDevice(d->body_mass, D.nbody, e->owned);
      P.T.alt_qpos = toDevice(d->alt_qpos, D.nq, e->owned);
      P.T.alt_qvel = toDevice(d->alt_qvel, D.nv, e->owned);
    }
    if (d->kind != MYO_ENV_POSE) {
      void* m = nullptr;
      if (cudaMalloc(&m, (size_t)b->nworld * sizeof(int)) != cudaSuccess || cudaMemset(m, 0, (size_t)b->nworld * sizeof(int)) != cudaSuccess) throw std::string("cudaMalloc failed");
      e->owned.push_back(m); P.reset_mask = (int*)m;
    }
  } catch (const std::string& msg) { myo_env_destroy(e); return myo_fail(MYO_ERR_CUDA, msg); }
  *out = e;
  return MYO_OK;
}
void myo_env_destroy(myo_env* e) {
  if (!e) return;
  for (void* p : e->owned) cudaFree(p);
  delete e;
}
int64_t myo_env_field_count(const myo_env* e, int field) { return e ? envFieldCount(e->P, field) : 0; }
int myo_env_bind(myo_env* e, int field, void* dev_ptr, int64_t nbytes) {
  if (!e || field < 0 || field >= MYO_E_COUNT) return myo_fail(MYO_ERR_ARG, "bad env field id");
  if (dev_ptr) {
    const int64_t need = envFieldCount(e->P, field) * (int64_t)e->P.nworld * 4;
    if (need == 0) return myo_fail(MYO_ERR_ARG, "env field is empty for this configuration");
    if (nbytes != need) return myo_fail(MYO_ERR_ARG, "size mismatch binding an env field");
    if (((uintptr_t)dev_ptr & 3) != 0) return myo_fail(MYO_ERR_ARG, "buffer must be 4-byte aligned");
  }
  e->P.E[field] = dev_ptr;
  return MYO_OK;
}
int myo_env_seed(myo_env* e, uint64_t seed) {
  if (!e) return myo_fail(MYO_ERR_ARG, "bad argument");
  e->P.seed = seed;
  return MYO_OK;
}
static int envReady(myo_env* e) {
  if (!e) return myo_fail(MYO_ERR_ARG, "bad argument");
  EnvParams& P = e->P;
  for (int f = 0; f < MYO_E_COUNT; f++)
    if (envFieldCount(P, f) > 0 && !P.E[f]) return myo_fail(MYO_ERR_UNBOUND, "an env field has no bound device buffer");
  const FieldPtrs& F = e->b->F;   // the state buffers may have been re-bound since the last call
  P.qpos = (float*)F.p[MYO_F_QPOS]; P.qvel = (float*)F.p[MYO_F_QVEL]; P.act = (float*)F.p[MYO_F_ACT]; P.ctrl = (float*)F.p[MYO_F_CTRL];
  P.time = (float*)F.p[MYO_F_TIME]; P.warm = (float*)F.p[MYO_F_QACC_WARMSTART]; P.status = (int*)F.p[MYO_F_STATUS];
  if (!P.qpos || !P.qvel || !P.time || !P.warm || (P.na && !P.act) || (P.nu && !P.ctrl)) return myo_fail(MYO_ERR_UNBOUND, "a state field has no bound device buffer");
  if (P.kind == MYO_ENV_WALK) {
    P.xpos = (const float*)F.p[MYO_F_XPOS]; P.xquat = (const float*)F.p[MYO_F_XQUAT]; P.xipos = (const float*)F.p[MYO_F_XIPOS]; P.cvel = (const float*)F.p[MYO_F_CVEL];
    P.actlen = (const float*)F.p[MYO_F_ACTUATOR_LENGTH]; P.actvel = (const float*)F.p[MYO_F_ACTUATOR_VELOCITY]; P.actfrc = (const float*)F.p[MYO_F_ACTUATOR_FORCE];
    if (!P.xpos || !P.xquat || !P.xipos || !P.cvel || !P.actlen || !P.actvel || !P.actfrc)
      return myo_fail(MYO_ERR_UNBOUND, "the walk env reads xpos, xquat, xipos, cvel, actuator_length, actuator_velocity and actuator_force: bind them on the batch");
  }
  if (P.kind == MYO_ENV_REORIENT) {
    P.xpos = (const float*)F.p[MYO_F_XPOS]; P.site_xpos = (const float*)F.p[MYO_F_SITE_XPOS]; P.geom_xpos = (const float*)F.p[MYO_F_GEOM_XPOS];
    P.actlen = (const float*)F.p[MYO_F_ACTUATOR_LENGTH]; P.actvel = (const float*)F.p[MYO_F_ACTUATOR_VELOCITY]; P.actfrc = (const float*)F.p[MYO_F_ACTUATOR_FORCE];
    if (!P.xpos || !P.site_xpos || !P.geom_xpos || !P.actlen || !P.actvel || !P.actfrc)
      return myo_fail(MYO_ERR_UNBOUND, "the reorient env reads xpos, site_xpos, geom_xpos, actuator_length, actuator_velocity and actuator_force: bind them on the batch");
    if (P.shape8) {
      P.gtype_ovr = (int*)F.p[MYO_F_GEOM_TYPE_OVR]; P.gsize_ovr = (float*)F.p[MYO_F_GEOM_SIZE_OVR]; P.gpos_ovr = (float*)F.p[MYO_F_GEOM_POS_OVR];
      if (!P.gtype_ovr || !P.gsize_ovr || !P.gpos_ovr)
        return myo_fail(MYO_ERR_UNBOUND, "randomize_shape writes geom_type_ovr, geom_size_ovr and geom_pos_ovr: bind them on the batch");
    }
  }
  if (cudaSetDevice(e->b->device) != cudaSuccess) return myo_fail(MYO_ERR_CUDA, "cudaSetDevice failed: libmyostep has no CPU fallback");
  return MYO_OK;
}
static void envRewardLaunch(const EnvParams& P, int advance, void* stream) {
  if (P.kind == MYO_ENV_WALK) envWalkRewardKernel<<<(P.nworld + 127) / 128, 128, 0, (cudaStream_t)stream>>>(P, advance);
  else envReorientRewardKernel<<<(P.nworld + 127) / 128, 128, 0, (cudaStream_t)stream>>>(P, advance);
  myo_count_launch(1);
}
static int envLaunched() {
  return cudaPeekAtLastError() == cudaSuccess ? MYO_OK : myo_fail(MYO_ERR_CUDA, std::string("kernel launch failed: ") + cudaGetErrorString(cudaGetLastError()));
}
int myo_env_reset(myo_env* e, const int32_t* world_mask_dev, void* stream) {
  int rc = envReady(e);
  if (rc) return rc;
  const EnvParams& P = e->P;
  envResetKernel<<<(P.nworld + 127) / 128, 128, 0, (cudaStream_t)stream>>>(P, world_mask_dev);
  myo_count_launch(1);
  if (P.kind != MYO_ENV_POSE) {   // robot.reset ends with sim.forward() (robot.py:976-979); the observation reads its outputs
    rc = myo_forward_masked(e->b, world_mask_dev, stream);
    if (rc) return rc;
    envObsKernel<<<(P.nworld + 127) / 128, 128, 0, (cudaStream_t)stream>>>(P);
    myo_count_launch(1);
  }
  return envLaunched();
}
int myo_env_step(myo_env* e, void* stream) {
  int rc = envReady(e);
  if (rc) return rc;
  const EnvParams& P = e->P;
  if (P.nu > 0) {
    const size_t n = (size_t)P.nworld * P.nu;
    envPreKernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(P);
    myo_count_launch(1);
  }
  rc = myo_step(e->b, P.frame_skip, stream);
  if (rc) return rc;
  if (P.kind != MYO_ENV_POSE) {
    // classic-path observation semantics: position / velocity fields of the stepped state (dm_control Physics.step ends
    // with mj_step1), then reward / done / reset, a full mj_forward of the worlds that were reset, then the observation
    rc = myo_forward_posvel(e->b, nullptr, stream);
    if (rc) return rc;
    envRewardLaunch(P, 1, stream);
    if (P.auto_reset) {
      rc = myo_forward_masked(e->b, P.reset_mask, stream);
      if (rc) return rc;
    }
    envObsKernel<<<(P.nworld + 127) / 128, 128, 0, (cudaStream_t)stream>>>(P);
    myo_count_launch(1);
    return envLaunched();
  }
  envPostKernel<<<(P.nworld + 127) / 128, 128, 0, (cudaStream_t)stream>>>(P, 1);
  myo_count_launch(1);
  return envLaunched();
}
int myo_env_forward(myo_env* e, void* stream) {
  int rc = envReady(e);
  if (rc) return rc;
  if (e->P.kind != MYO_ENV_POSE) {   // reads the derived fields as they are: call myo_forward / myo_forward_posvel first if the state was edited
    envRewardLaunch(e->P, 0, stream);
    envObsKernel<<<(e->P.nworld + 127) / 128, 128, 0, (cudaStream_t)stream>>>(e->P);
    myo_count_launch(1);
    return envLaunched();
  }
  envPostKernel<<<(e->P.nworld + 127) / 128, 128, 0, (cudaStream_t)stream>>>(e->P, 0);
  myo_count_launch(1);
  return envLaunched();
}

}  // extern "C"
