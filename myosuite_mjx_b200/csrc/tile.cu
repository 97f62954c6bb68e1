// TILE-mode kernels of the batched mj_step for sm_100a: LANE = WORLD.
//
// A CTA owns a tile of 32 consecutive worlds (one per lane) and brings NW warps to it; the items of every model-driven
// loop (frames, sites, tendon units, moment-arm entries, dofs, candidate pairs ...) are striped over the warps.  So
//  * every model-table index is warp-uniform (one broadcast load, no divergence from the model's structure),
//  * every access to the per-world scratch and to a [k][nworld] mjData field is one full, coalesced 128-byte line,
//  * all 32 lanes of every instruction do useful work; lanes differ only on DATA (wrap / no wrap, contact or not),
//  * a CTA barrier (not a kernel boundary) orders dependent item sets inside a phase; a kernel boundary orders phases.
// The regular phases are the same source as the TEAM build (common.cuh, compiled here with TILE defined); the collision
// phase is written for this mode (per-lane candidate lists, ordered compaction by per-warp chunks).
//
// One mj_step = kin -> tendon -> inertia -> collide -> velact -> [team.cu solveKernel] -> finish   (api.cu).
#define TILE 1   // common.cuh: TEAM = t.nw, the warps per tile -- a launch parameter (blockDim.x / 32), not a build constant
#include "common.cuh"

namespace {

enum { KIN_LOAD = 1, KIN_RESET = 2, KIN_CHECK = 4 };
enum { FIN_EULER = 1, FIN_STATE = 2, FIN_WARM = 4, FIN_ACC = 8 };
enum { WS_STATUS = 0, WS_TIME = 1, WS_SKIP = 2 };

struct Ctx { SP sm; Tm t; World W; bool live; };
__device__ __forceinline__ Ctx tileCtx(int nworld, int stride, const int* mask, float* scratch) {
  Ctx c;
  const int lane = threadIdx.x & 31;
  c.t.tl = threadIdx.x >> 5; c.t.nw = blockDim.x >> 5;
  const size_t w = (size_t)blockIdx.x * 32 + lane;
  c.live = w < (size_t)nworld && !(mask && !mask[w]);
  c.W.w = w; c.W.nw = (size_t)nworld;
  c.sm.p = scratch + w; c.sm.s = stride;
  return c;
}

// ------------------------------------------------------------------------------------------ kin: state in, mj_kinematics, mj_comPos
__global__ void __launch_bounds__(512) kinKernel(int slot, int fslot, int nworld, int stride, const int* __restrict__ mask, int flags,
                                                 float* __restrict__ scratch) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L; const FieldPtrs& F = tP[slot].F;
  Ctx c = tileCtx(nworld, stride, mask, scratch);
  if (!c.live) return;   // exited lanes are not waited for by the CTA barriers below
  SP sm = c.sm; Tm t = c.t; World W = c.W;
  __shared__ int sBad[32];
  const int lane = threadIdx.x & 31;
  const int nq = D.nq, nv = D.nv, na = D.na, nu = D.nu;
  if (flags & KIN_RESET) {
    TFOR(i, nq) sm[L.qpos + i] = D.qpos0[i];
    TFOR(i, nv) { sm[L.qvel + i] = 0.f; sm[L.warm + i] = 0.f; }
    TFOR(i, na) sm[L.act + i] = 0.f;
    TFOR(i, nu) sm[L.ctrl + i] = 0.f;
    if (t.tl == 0) { sm[L.wstate + WS_STATUS] = __int_as_float(0); sm[L.wstate + WS_TIME] = 0.f; }
  } else if (flags & KIN_LOAD) {
    void* gq = F.p[MYO_F_QPOS]; void* gv = F.p[MYO_F_QVEL]; void* ga = F.p[MYO_F_ACT]; void* gc = F.p[MYO_F_CTRL];
    void* gw = F.p[MYO_F_QACC_WARMSTART]; void* gt = F.p[MYO_F_TIME]; void* gs = F.p[MYO_F_STATUS];
    TFOR(i, nq) sm[L.qpos + i] = OUTF(gq, i);
    TFOR(i, nv) { sm[L.qvel + i] = OUTF(gv, i); sm[L.warm + i] = OUTF(gw, i); }
    TFOR(i, na) sm[L.act + i] = OUTF(ga, i);
    TFOR(i, nu) sm[L.ctrl + i] = OUTF(gc, i);
    if (t.tl == 0) { sm[L.wstate + WS_STATUS] = __int_as_float(gs ? OUTI(gs, 0) : 0); sm[L.wstate + WS_TIME] = OUTF(gt, 0); }
  }
  if (t.tl == 0) { sm[L.wstate + WS_SKIP] = __int_as_float(0); sBad[lane] = 0; }
  TSYNC();
  if (flags & KIN_CHECK) {   // mj_checkPos / mj_checkVel: a world with a bad state is reset and flagged
    bool bq = false, bv = false;
    TFOR(i, nq) bq |= !(fabsf(sm[L.qpos + i]) < 1e10f);
    TFOR(i, nv) bv |= !(fabsf(sm[L.qvel + i]) < 1e10f);
    if (bq) atomicOr(&sBad[lane], MYO_WORLD_BADQPOS);
    if (bv) atomicOr(&sBad[lane], MYO_WORLD_BADQVEL);
    TSYNC();
    const int bad = sBad[lane];
    if (bad) {
      TFOR(i, nq) sm[L.qpos + i] = D.qpos0[i];
      TFOR(i, nv) { sm[L.qvel + i] = 0.f; sm[L.warm + i] = 0.f; }
      TFOR(i, na) sm[L.act + i] = 0.f;
      if (t.tl == 0) sm[L.wstate + WS_STATUS] = __int_as_float(__float_as_int(sm[L.wstate + WS_STATUS]) | bad);
    }
    TSYNC();
  }
  kinematics(slot, sm, t);
  comPos(slot, sm, t);
  if (fslot != MYO_NSLOT) writePositionOutputs(slot, fslot, sm, t, W);
}

// ------------------------------------------------------------------------------------------ tendon: mj_tendon (+ path points)
__global__ void __launch_bounds__(512) tendonKernel(int slot, int fslot, int nworld, int stride, const int* __restrict__ mask, float* __restrict__ scratch) {
  Ctx c = tileCtx(nworld, stride, mask, scratch);
  if (!c.live) return;
  tendons(slot, fslot, c.sm, c.t, c.W);
}

// ------------------------------------------------------------------------------------------ inertia: mj_crb, mj_factorM (+ Euler's M + h B factor)
__global__ void __launch_bounds__(512) inertiaKernel(int slot, int fslot, int nworld, int stride, const int* __restrict__ mask, int euler,
                                                     float* __restrict__ scratch) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L;
  Ctx c = tileCtx(nworld, stride, mask, scratch);
  if (!c.live) return;
  SP sm = c.sm; Tm t = c.t;
  inertia(slot, fslot, sm, t, c.W);
  if (euler && D.any_damping) {   // mj_Euler's implicit joint damping: factor of M + h diag(damping), used by finishKernel
    const float h = D.timestep;
    TSYNC();
    TFOR(e, D.nM) sm[L.qH + e] = sm[L.qM + e] + (D.M_row[e] == D.M_col[e] ? h * D.dof_damping[D.M_row[e]] : 0.f);
    TSYNC();
    factorLDL(slot, sm + L.qH, sm + L.qHinv, t);
  }
}

// ------------------------------------------------------------------------------------------ velact: mj_fwdVelocity, mj_fwdActuation, mj_fwdAcceleration
__global__ void __launch_bounds__(512) velactKernel(int slot, int fslot, int nworld, int stride, const int* __restrict__ mask, int posvel,
                                                    float* __restrict__ scratch) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L;
  Ctx c = tileCtx(nworld, stride, mask, scratch);
  if (!c.live) return;
  SP sm = c.sm; Tm t = c.t; World W = c.W;
  velocity(slot, fslot, sm, t, W);
  if (posvel) {   // mj_transmission outputs of the refreshed tendon stage (actuation() writes them on a full pass)
    const FieldPtrs& F = tP[fslot].F;
    void* olen = F.p[MYO_F_ACTUATOR_LENGTH]; void* ovel = F.p[MYO_F_ACTUATOR_VELOCITY];
    TFOR(i, D.nu) {
      const int id = D.act_trnid[i];
      const float gear = D.act_gear[i];
      float len, vel;
      if (D.act_trntype[i] == 3) { len = gear * sm[L.tenlen + id]; vel = gear * sm[L.tenvel + id]; }
      else { len = gear * sm[L.qpos + D.jnt_qposadr[id]]; vel = gear * sm[L.qvel + D.jnt_dofadr[id]]; }
      sm[L.actlen + i] = len; sm[L.actvel + i] = vel;
      if (olen) OUTF(olen, i) = len;
      if (ovel) OUTF(ovel, i) = vel;
    }
    return;
  }
  actuation(slot, fslot, sm, t, W);
}

// ------------------------------------------------------------------------------------------ collide: mj_collision
// Broad phase over the static, ordered candidate-pair list; narrow phase over each world's survivors; hits compacted in
// pair order, so contact order -- hence geom1 / geom2 / efc_address -- is MuJoCo's.  Lane = world: each lane keeps its own
// survivor list in its scratch column.  Ordered compaction without a serial walk: warp j owns a contiguous chunk of the
// item range, counts its survivors per lane, the per-warp counts meet in shared memory, and every warp then writes its
// survivors starting at its lane's exclusive prefix.  The convex narrow phase (MPR, float64, long) is listed per lane
// first and run afterwards, so the lanes of a warp that have an MPR pair run it at the same time.
#define MAXW 16   // warps per tile CTA (blockDim <= 512)
__device__ __forceinline__ bool broadPass(const DevModel& D, const SmemLayout& L, SP sm, int pr) {
  const int c1 = D.pair_cg1[pr], c2 = D.pair_cg2[pr];
  auto P1 = sm + L.cgpose + 12 * c1; auto P2 = sm + L.cgpose + 12 * c2;
  const f3 d = ld3(P2) - ld3(P1);
  const float margin = D.pair_prm[16 * pr];
  if (D.cg_type[c1] == G_PLANE) return !(d.x * P1[5] + d.y * P1[8] + d.z * P1[11] > D.cg_rbound[c2] + margin);
  const float r = D.cg_rbound[c1] + D.cg_rbound[c2] + margin;
  return !(dot(d, d) > r * r);
}
__device__ __forceinline__ int warpMax(int v) { return __reduce_max_sync(__activemask(), v); }

__global__ void __launch_bounds__(512) collideKernel(int slot, int nworld, int stride, const int* __restrict__ mask, float* __restrict__ scratch) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L;
  Ctx c = tileCtx(nworld, stride, mask, scratch);
  if (!c.live) return;
  SP sm = c.sm; Tm t = c.t;
  const int lane = threadIdx.x & 31, nw = t.nw;
  __shared__ int sCnt[MAXW][32];
  __shared__ int sNm[32];
  if (D.npair == 0 || (D.disableflags & (DSBL_CONTACT | DSBL_CONSTRAINT))) {
    if (t.tl == 0) sm[L.counts] = __int_as_float(0);
    return;
  }
  TFOR(g, D.ncg) {
    const int f = D.cg_frame[g];
    auto o = sm + L.cgpose + 12 * g;
    st3(o, ld3(FR_POS(sm, L, f)) + mulM(FR_MAT(sm, L, f), ld3(D.cg_lpos + 3 * g)));
    float M[9]; mulMM(M, FR_MAT(sm, L, f), D.cg_lmat + 9 * g);
#pragma unroll
    for (int k = 0; k < 9; k++) o[3 + k] = M[k];
  }
  TSYNC();
  // ---- broad phase: count, prefix over the warps, write (the test is cheap enough to run twice)
  SPI cand = asInt(sm + L.cand);
  int ncand;
  {
    const int chunk = (D.npair + nw - 1) / nw, p0 = t.tl * chunk, p1 = min(D.npair, p0 + chunk);
    int cnt = 0;
    for (int pr = p0; pr < p1; pr++) cnt += broadPass(D, L, sm, pr) ? 1 : 0;
    sCnt[t.tl][lane] = cnt;
    TSYNC();
    int off = 0; ncand = 0;
    for (int j = 0; j < nw; j++) { const int v = sCnt[j][lane]; if (j < t.tl) off += v; ncand += v; }
    for (int pr = p0; pr < p1; pr++) if (broadPass(D, L, sm, pr)) cand[off++] = pr;
    TSYNC();
  }
  // ---- narrow phase.  Per-survivor results live in the solver's scratch (unused until makeConstraint): 16 floats =
  // count, two hits, one MPR-list slot.
  auto tmp = sm + L.conJ;
  const int cap = max(1, (L.limrow - L.conJ) / 16);
  const int maxcand = warpMax(ncand);
  int ncon = 0;
  for (int c0 = 0; c0 < maxcand; c0 += cap) {   // one round unless a world has more survivors than the scratch holds
    const int cn = max(0, min(cap, ncand - c0));
    if (t.tl == 0) sNm[lane] = 0;
    TSYNC();
    for (int k = t.tl; k < cn; k += nw) {   // closed-form pairs; MPR pairs are only listed
      const int pr = cand[c0 + k];
      const int t1 = D.cg_type[D.pair_cg1[pr]], t2 = D.cg_type[D.pair_cg2[pr]];
      const bool mp = t1 != G_PLANE && !((t1 == G_SPHERE || t1 == G_CAPSULE) && (t2 == G_SPHERE || t2 == G_CAPSULE) && !(t1 == G_CAPSULE && t2 == G_SPHERE));
      if (!mp) {
        Hit h[2];
        const int n = collidePair(slot, sm, pr, h);
        auto o = tmp + 16 * k;
        o[0] = __int_as_float(n);
        for (int q = 0; q < n; q++) { o[1 + 7 * q] = h[q].dist; st3(o + 2 + 7 * q, h[q].pos); st3(o + 5 + 7 * q, h[q].nrm); }
      } else {
        const int i = atomicAdd(&sNm[lane], 1);   // list order is irrelevant: results go to the survivor's own record
        tmp[16 * i + 15] = __int_as_float(k);
      }
    }
    TSYNC();
    const int nm = sNm[lane];
    for (int i = t.tl; i < nm; i += nw) {
      const int k = __float_as_int(tmp[16 * i + 15]);
      Hit h[2];
      const int n = collidePair(slot, sm, cand[c0 + k], h);
      auto o = tmp + 16 * k;
      o[0] = __int_as_float(n);
      if (n) { o[1] = h[0].dist; st3(o + 2, h[0].pos); st3(o + 5, h[0].nrm); }
    }
    TSYNC();
    // in-order compaction of the hits
    const int kc = (warpMax(cn) + nw - 1) / nw, k0 = t.tl * kc, k1 = min(cn, k0 + kc);
    int cnt = 0;
    for (int k = k0; k < k1; k++) cnt += __float_as_int(tmp[16 * k]);
    sCnt[t.tl][lane] = cnt;
    TSYNC();
    int at = ncon, total = 0;
    for (int j = 0; j < nw; j++) { const int v = sCnt[j][lane]; if (j < t.tl) at += v; total += v; }
    for (int k = k0; k < k1; k++) {
      const int n = __float_as_int(tmp[16 * k]);
      for (int q = 0; q < n; q++, at++) {
        if (at >= D.nconmax) continue;
        const int pr = cand[c0 + k];
        auto cr = sm + L.con + 12 * at; auto o = tmp + 16 * k + 1 + 7 * q;
        for (int e = 0; e < 7; e++) cr[e] = o[e];
        cr[7] = __int_as_float(pr); cr[8] = __int_as_float(-1); cr[9] = __int_as_float(D.pair_dim[pr]);
        // dofs that move exactly one of the two bodies (static per pair): the only non-zero columns of the contact's rows
        const int f1 = D.cg_frame[D.pair_cg1[pr]], f2 = D.cg_frame[D.pair_cg2[pr]];
        cr[10] = __int_as_float(D.fr_mask[2 * f1] ^ D.fr_mask[2 * f2]); cr[11] = __int_as_float(D.fr_mask[2 * f1 + 1] ^ D.fr_mask[2 * f2 + 1]);
      }
    }
    ncon += total;
    TSYNC();
  }
  if (t.tl == 0) {
    if (ncon > D.nconmax) {
      sm[L.wstate + WS_STATUS] = __int_as_float(__float_as_int(sm[L.wstate + WS_STATUS]) | MYO_WORLD_NCON_OVERFLOW);
      ncon = D.nconmax;
    }
    sm[L.counts] = __int_as_float(ncon);
  }
}

// ------------------------------------------------------------------------------------------ finish: mj_Euler, state out
// mj_Euler with implicit joint damping (the factor of M + h B comes from inertiaKernel), quaternion integration on the
// manifold; then the state and the per-dof results of the solve go out as full lines.
__global__ void __launch_bounds__(512) finishKernel(int slot, int fslot, int nworld, int stride, const int* __restrict__ mask, int flags,
                                                    float* __restrict__ scratch) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L; const FieldPtrs& F = tP[slot].F;
  Ctx c = tileCtx(nworld, stride, mask, scratch);
  if (!c.live) return;
  SP sm = c.sm; Tm t = c.t; World W = c.W;
  const int nq = D.nq, nv = D.nv, na = D.na, nu = D.nu;
  if (flags & FIN_ACC) {   // results of the solve (read-only here; mj_Euler below does not change them)
    const FieldPtrs& Fo = tP[fslot].F;
    void* o;
    if ((o = Fo.p[MYO_F_QACC])) TFOR(i, nv) OUTF(o, i) = sm[L.qacc + i];
    if ((o = Fo.p[MYO_F_QFRC_CONSTRAINT])) TFOR(i, nv) OUTF(o, i) = sm[L.qfrc_con + i];
    if ((o = Fo.p[MYO_F_QFRC_INVERSE])) TFOR(i, nv) OUTF(o, i) = 0.f;
  }
  if (flags & FIN_EULER) {
    const bool skip = __float_as_int(sm[L.wstate + WS_SKIP]) != 0;   // mj_checkAcc reset this world: nothing to integrate
    const float h = D.timestep;
    auto qa = sm + L.tmpv;
    if (D.any_damping) {
      TFOR(i, nv) qa[i] = sm[L.smooth + i] + sm[L.qfrc_con + i];
      TSYNC();
      solveLDL(slot, sm + L.qH, sm + L.qHinv, qa, t);
    } else {
      TFOR(i, nv) qa[i] = sm[L.qacc + i];
      TSYNC();
    }
    if (!skip) {
      if (!(D.disableflags & DSBL_ACTUATION))
        TFOR(i, nu) {
          const int aa = D.act_actadr[i];
          if (aa < 0) continue;
          float a = fmaf(h, sm[L.actdot + aa], sm[L.act + aa]);
          if (D.act_flags[i] & ACT_ACTLIMITED) a = clipf(a, D.act_actrange[2 * i], D.act_actrange[2 * i + 1]);
          sm[L.act + aa] = a;
        }
      TFOR(i, nv) sm[L.qvel + i] = fmaf(h, qa[i], sm[L.qvel + i]);
    }
    TSYNC();
    if (!skip) {
      TFOR(j, D.njnt) {
        const int qa_ = D.jnt_qposadr[j], da = D.jnt_dofadr[j];
        if (D.jnt_type[j] == JNT_FREE) {
          for (int k = 0; k < 3; k++) sm[L.qpos + qa_ + k] = fmaf(h, sm[L.qvel + da + k], sm[L.qpos + qa_ + k]);
          const f3 wv = ld3(sm + L.qvel + da + 3);
          const float ang = norm3(wv) * h;
          if (ang > 0.f) {
            const f3 ax = unit3(wv);
            float sn, cs; sincos_(0.5f * ang, &sn, &cs);
            f4 dq; dq.w = cs; dq.x = ax.x * sn; dq.y = ax.y * sn; dq.z = ax.z * sn;
            st4(sm + L.qpos + qa_ + 3, qnormalize(qmul(ld4(sm + L.qpos + qa_ + 3), dq)));
          }
        } else sm[L.qpos + qa_] = fmaf(h, sm[L.qvel + da], sm[L.qpos + qa_]);
      }
      if (t.tl == 0) sm[L.wstate + WS_TIME] += h;
    }
    TSYNC();
  }
  if (flags & FIN_STATE) {
    void* gq = F.p[MYO_F_QPOS]; void* gv = F.p[MYO_F_QVEL]; void* ga = F.p[MYO_F_ACT]; void* gt = F.p[MYO_F_TIME];
    TFOR(i, nq) OUTF(gq, i) = sm[L.qpos + i];
    TFOR(i, nv) OUTF(gv, i) = sm[L.qvel + i];
    TFOR(i, na) OUTF(ga, i) = sm[L.act + i];
    if (t.tl == 0) OUTF(gt, 0) = sm[L.wstate + WS_TIME];
  }
  if (flags & FIN_WARM) { void* gw = F.p[MYO_F_QACC_WARMSTART]; TFOR(i, nv) OUTF(gw, i) = sm[L.warm + i]; }
  void* gs = F.p[MYO_F_STATUS];
  if (gs && t.tl == 0) OUTI(gs, 0) = __float_as_int(sm[L.wstate + WS_STATUS]);
}

// ------------------------------------------------------------------------------------------ conout: contact / constraint records
// contact_* and efc_* (padded to nconmax / njmax, efc_J expanded to dense rows): pure stores, no ordering between items, so
// the items are striped over the warps of gridDim.y CTAs per tile.
__global__ void __launch_bounds__(256) conoutKernel(int slot, int fslot, int nworld, int stride, const int* __restrict__ mask, float* __restrict__ scratch) {
  const SmemLayout& L = tP[slot].L;
  Ctx c = tileCtx(nworld, stride, mask, scratch);
  if (!c.live) return;
  Tm t; t.nw = c.t.nw * gridDim.y; t.tl = blockIdx.y * c.t.nw + c.t.tl;
  SP sm = c.sm;
  writeConstraintOutputs(slot, fslot, sm, t, c.W, __float_as_int(sm[L.counts]), __float_as_int(sm[L.counts + 1]), __float_as_int(sm[L.counts + 2]));
}

// known-answer hook for the rotation helpers above (tests/: reference KATs of myosuite/utils/quat_math.py)
__global__ void quatKatKernel(int n, const float* __restrict__ qa, const float* __restrict__ qb, float* __restrict__ prod,
                              float* __restrict__ mat, float* __restrict__ rotated) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const f4 a = ld4(qa + 4 * i), b = ld4(qb + 4 * i);
  st4(prod + 4 * i, qmul(a, b));
  float R[9]; q2m(R, qnormalize(a));
  for (int k = 0; k < 9; k++) mat[9 * i + k] = R[k];
  st3(rotated + 3 * i, mulM(R, mk3(b.x, b.y, b.z)));
}


}  // namespace

// ---------------------------------------------------------------------------------------------- host side
int myo_tile_upload(int slot, const DevModel& D, const SmemLayout& L, const FieldPtrs& F, void* stream) {
  static TParams staging[MYO_NSLOT];   // must outlive the async copy
  if (slot < 0 || slot >= MYO_NSLOT) return -1;
  TParams& P = staging[slot];
  P.D = D; P.L = L; P.F = F;
  return cudaMemcpyToSymbolAsync(tP, &P, sizeof(TParams), (size_t)slot * sizeof(TParams), cudaMemcpyHostToDevice, (cudaStream_t)stream) == cudaSuccess ? 0 : -1;
}
static int envInt(const char* name, int dflt) { const char* e = getenv(name); return e ? atoi(e) : dflt; }
// Warps per tile of each kernel (developer knobs MYO_NW_*): enough warps per SM to hide the scratch latency, few enough
// that the item loops still have work for every warp.
int myo_tile_launch(int kernel, int slot, int fslot, int nworld, int stride, const int* mask, int flags, float* scratch, void* stream) {
  static const int nwKin = envInt("MYO_NW_KIN", 16), nwTen = envInt("MYO_NW_TEN", 16), nwIne = envInt("MYO_NW_INE", 8),
                   nwCol = envInt("MYO_NW_COL", 16), nwVel = envInt("MYO_NW_VEL", 8), nwFin = envInt("MYO_NW_FIN", 8),
                   nwOut = envInt("MYO_NW_OUT", 8), yOut = envInt("MYO_Y_OUT", 4);
  const int tiles = (nworld + 31) / 32;
  cudaStream_t st = (cudaStream_t)stream;
  switch (kernel) {
    case MYO_K_KIN: kinKernel<<<tiles, 32 * nwKin, 0, st>>>(slot, fslot, nworld, stride, mask, flags, scratch); break;
    case MYO_K_TENDON: tendonKernel<<<tiles, 32 * nwTen, 0, st>>>(slot, fslot, nworld, stride, mask, scratch); break;
    case MYO_K_INERTIA: inertiaKernel<<<tiles, 32 * nwIne, 0, st>>>(slot, fslot, nworld, stride, mask, flags, scratch); break;
    case MYO_K_COLLIDE: collideKernel<<<tiles, 32 * nwCol, 0, st>>>(slot, nworld, stride, mask, scratch); break;
    case MYO_K_VELACT: velactKernel<<<tiles, 32 * nwVel, 0, st>>>(slot, fslot, nworld, stride, mask, flags, scratch); break;
    case MYO_K_FINISH: finishKernel<<<tiles, 32 * nwFin, 0, st>>>(slot, fslot, nworld, stride, mask, flags, scratch); break;
    case MYO_K_CONOUT: conoutKernel<<<dim3(tiles, yOut), 32 * nwOut, 0, st>>>(slot, fslot, nworld, stride, mask, scratch); break;
    default: return -1;
  }
  return cudaPeekAtLastError() == cudaSuccess ? 0 : -1;
}
int myo_launch_quat_kat(int n, const float* qa, const float* qb, float* prod, float* mat, float* rotated, void* stream) {
  quatKatKernel<<<(n + 127) / 128, 128, 0, (cudaStream_t)stream>>>(n, qa, qb, prod, mat, rotated);
  return cudaPeekAtLastError() == cudaSuccess ? 0 : -1;
}
