// TILE-mode kernels of the batched mj_step for sm_100a: LANE = WORLD.
//
// A CTA owns a tile of 32 consecutive worlds (one per lane) and brings NW warps to it; the items of every model-driven
// loop (frames, sites, tendon units, moment-arm entries, dofs, candidate pairs ...) are striped over the warps.  So
//  * every model-table index is warp-uniform (one broadcast load, no divergence from the model's structure),
//  * every access to the per-world scratch and to a [k][nworld] mjData field is one full, coalesced 128-byte line,
//  * all 32 lanes of every instruction do useful work; lanes differ only on DATA (wrap / no wrap, contact or not),
//  * a CTA barrier (not a kernel boundary) orders dependent item sets inside a phase; a kernel boundary orders phases.
// The collision phase keeps per-lane candidate lists and compacts in order by per-warp chunks; the constraint solve runs the
// tile's 32 worlds in lock step (per-lane activity masks).
//
// One mj_step (api.cu launchPhases): kin, then three branches on their own streams --
//   collide -> collideMpr -> collideCompact -> makecon   |   tendon -> velact   |   inertia (+ body velocities, bias force)
// -- plus the output stream (posout after kin, conout after makecon), then newton, whose tail is mj_Euler + the state outputs.
// Measured (profiles/r2f_*): every kernel is bound by the latency of its dependent instructions, and the register file holds
// at most two or three kernels' tiles at once, so the phase-2 kernels are capped at 56 registers (three CTAs per SM).
#include "common.cuh"
#include <cstring>
#include <mutex>
#include <string>

#ifndef MINB_KIN
#define MINB_KIN 2
#endif
#ifndef MINB_TENDON
#define MINB_TENDON 3
#endif
#ifndef MINB_INERTIA
#define MINB_INERTIA 3
#endif
#ifndef MINB_VELACT
#define MINB_VELACT 3
#endif
#ifndef MINB_COLLIDE
#define MINB_COLLIDE 3
#endif
#ifndef MINB_NEWTON
#define MINB_NEWTON 2
#endif
#ifndef MINB_FINISH
#define MINB_FINISH 2
#endif
// (syrk block size: 6 x 6 = ten blocks, one round of 12 warps, was +2 % step time (registers); 8 x 8 +3.5 %)
#ifndef SYRK_UNROLL
#define SYRK_UNROLL 1   // (contacts per iteration of the syrk: 24 J loads each; 2: +0.5 % step time)
#endif
namespace {

enum { KIN_LOAD = 1, KIN_RESET = 2, KIN_CHECK = 4 };
enum { FIN_EULER = 1, FIN_STATE = 2, FIN_WARM = 4, FIN_ACC = 8, FIN_LIVE = 16 };   // FIN_LIVE: efc_force rows up to nefc only (count-aware batch)
enum { WS_STATUS = 0, WS_TIME = 1, WS_SKIP = 2 };

struct Ctx { SP sm; Tm t; World W; bool live; };
// developer timeline (MYO_NEWTON_TIMING): first start and last end of each kernel's CTAs, globaltimer ns
struct Timeline {
  unsigned long long* p;
  __device__ __forceinline__ static unsigned long long now() { unsigned long long v; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(v)); return v; }
  __device__ __forceinline__ Timeline(int slot, int k) {
    long long* d = tP[slot].X.dbg;
    p = (d && threadIdx.x == 0) ? (unsigned long long*)d + tP[slot].X.tl_off + 2 * k : nullptr;
    if (p) atomicMin(p, now());
  }
  __device__ __forceinline__ ~Timeline() { if (p) atomicMax(p + 1, now()); }
};
__device__ __forceinline__ Ctx tileCtxAt(int slot, int nworld, const int* mask, float* scratch, int tile) {
  Ctx c;
  const int lane = threadIdx.x & 31;
  c.t.tl = threadIdx.x >> 5; c.t.nw = blockDim.x >> 5;
  const size_t w = (size_t)tile * 32 + lane;
  c.live = w < (size_t)nworld && !(mask && !mask[w]);
  c.W.w = w; c.W.nw = (size_t)nworld; c.W.w32 = (unsigned)w; c.W.nw32 = (unsigned)nworld;
  c.sm.p = scratch + (size_t)tile * tP[slot].L.total * SCR + lane;   // this tile's block, this lane's column
  return c;
}
__device__ __forceinline__ Ctx tileCtx(int slot, int nworld, const int* mask, float* scratch) { return tileCtxAt(slot, nworld, mask, scratch, blockIdx.x); }
// Output kernels off the critical path (posout, tenout) run on a LIMITED grid, each CTA walking tiles blockIdx.x, + gridDim.x,
// ...: a kernel of one CTA per tile holds its SM slots for its whole duration (all 256 tiles are resident at once), and stream
// priorities only decide between CTAs that are still waiting -- so a full-grid output kernel made the solver's feeders
// (collideCompact, makecon) wait tens of microseconds for slots.  The lanes of a partial last tile exit there: it is the last
// tile of the CTA that gets it.  With a world mask any tile can have dead lanes: the host then launches one CTA per tile.
#define TILE_LOOP(tile) for (int tile = blockIdx.x, ntiles_ = (nworld + 31) >> 5; tile < ntiles_; tile += gridDim.x)

// ------------------------------------------------------------------------------------------ kin: state in, mj_kinematics, mj_comPos
__global__ void __launch_bounds__(32 * MAXW, MINB_KIN) kinKernel(int slot, int fslot, int nworld, const int* __restrict__ mask, int flags,
                                                 float* __restrict__ scratch) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L; const FieldPtrs& F = tP[slot].F;
  Timeline tline(slot, MYO_K_KIN);
  if (blockIdx.x == 0 && threadIdx.x == 0 && tP[slot].X.mpr_count) *tP[slot].X.mpr_count = 0;   // the convex narrow phase's work list (collideKernel appends)
  Ctx c = tileCtx(slot, nworld, mask, scratch);
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
}
// The position stage's output fields (xpos ... site_xmat, cinert, subtree_com): everything after kin waits for kin, so the
// copies (4/5 of the old kin kernel's time) are their own launch on the low-priority output stream, beside the next phases.
#ifndef MINB_POSOUT
#define MINB_POSOUT 3
#endif
__global__ void __launch_bounds__(32 * MAXW, MINB_POSOUT) posoutKernel(int slot, int fslot, int nworld, const int* __restrict__ mask, float* __restrict__ scratch) {
  Timeline tline(slot, MYO_K_POSOUT);
  TILE_LOOP(tile) {
    Ctx c = tileCtxAt(slot, nworld, mask, scratch, tile);
    if (!c.live) return;
    writePositionOutputs(slot, fslot, c.sm, c.t, c.W);
    TSYNC();
  }
}

// ------------------------------------------------------------------------------------------ tendon: mj_tendon (+ path points)
__global__ void __launch_bounds__(32 * MAXW, MINB_TENDON) tendonKernel(int slot, int fslot, int nworld, const int* __restrict__ mask, float* __restrict__ scratch) {
  Timeline tline(slot, MYO_K_TENDON);
  Ctx c = tileCtx(slot, nworld, mask, scratch);
  if (!c.live) return;
  tendons(slot, fslot, c.sm, c.t, c.W);
}

// ------------------------------------------------------------------------------------------ inertia: mj_crb, mj_factorM (+ Euler's M + h B factor)
__global__ void __launch_bounds__(32 * MAXW, MINB_INERTIA) inertiaKernel(int slot, int fslot, int nworld, const int* __restrict__ mask, int euler,
                                                     int flags2, float* __restrict__ scratch) {
  Timeline tline(slot, MYO_K_INERTIA);
  Ctx c = tileCtx(slot, nworld, mask, scratch);
  if (!c.live) return;
  SP sm = c.sm; Tm t = c.t;
  extern __shared__ float tileShared[];
  SP stage; stage.p = (flags2 & 1) ? tileShared + (threadIdx.x & 31) : nullptr;
  inertia(slot, fslot, sm, t, c.W, stage, euler);
  TSYNC();
  velocityBody(slot, fslot, sm, t, c.W);   // (cdof_dot was inertia()'s scratch up to here)
}

// ------------------------------------------------------------------------------------------ velact: mj_fwdVelocity, mj_fwdActuation, mj_fwdAcceleration
__global__ void __launch_bounds__(32 * MAXW, MINB_VELACT) velactKernel(int slot, int fslot, int nworld, const int* __restrict__ mask, int posvel,
                                                    int flags2, float* __restrict__ scratch) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L;
  Timeline tline(slot, MYO_K_VELACT);
  Ctx c = tileCtx(slot, nworld, mask, scratch);
  if (!c.live) return;
  SP sm = c.sm; Tm t = c.t; World W = c.W;
  if (posvel) velocityBody(slot, fslot, sm, t, W);   // (a full pass has it from inertiaKernel)
  velocityTendon(slot, fslot, sm, t, W);
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
  extern __shared__ float tileShared[];
  SP stage; stage.p = (flags2 & 1) ? tileShared + (threadIdx.x & 31) : nullptr;
  actuation(slot, fslot, sm, t, W, stage);
}

// ------------------------------------------------------------------------------------------ collide: mj_collision
// Broad phase over the static, ordered candidate-pair list; narrow phase over each world's survivors; hits compacted in
// pair order, so contact order -- hence geom1 / geom2 / efc_address -- is MuJoCo's.  Lane = world: each lane keeps its own
// survivor list in its scratch column.  Ordered compaction without a serial walk: warp j owns a contiguous chunk of the
// item range, counts its survivors per lane, the per-warp counts meet in shared memory, and every warp then writes its
// survivors starting at its lane's exclusive prefix.  The convex narrow phase (MPR, float64, long) is listed per lane
// first and run afterwards, so the lanes of a warp that have an MPR pair run it at the same time.

__device__ __forceinline__ bool broadPass(const DevModel& D, const SmemLayout& L, SP sm, bool ovr, int pr) {
  const int c1 = D.pair_cg1[pr], c2 = D.pair_cg2[pr];
  auto P1 = sm + L.cgpose + 12 * c1; auto P2 = sm + L.cgpose + 12 * c2;
  const f3 d = ld3(P2) - ld3(P1);
  const float margin = D.pair_prm[16 * pr];
  const float r1 = ovr ? sm[L.cgshape + 5 * c1 + 4] : D.cg_rbound[c1], r2 = ovr ? sm[L.cgshape + 5 * c2 + 4] : D.cg_rbound[c2];
  if (D.cg_type[c1] == G_PLANE) return !(d.x * P1[5] + d.y * P1[8] + d.z * P1[11] > r2 + margin);   // (a plane stays a plane)
  const float r = r1 + r2 + margin;
  return !(dot(d, d) > r * r);
}

__device__ __forceinline__ bool isMprPair(const DevModel& D, const SmemLayout& L, SP sm, bool ovr, int pr) {
  const int c1 = D.pair_cg1[pr], c2 = D.pair_cg2[pr];
  const int t1 = ovr ? __float_as_int(sm[L.cgshape + 5 * c1]) : D.cg_type[c1], t2 = ovr ? __float_as_int(sm[L.cgshape + 5 * c2]) : D.cg_type[c2];
  return t1 != G_PLANE && !((t1 == G_SPHERE || t1 == G_CAPSULE) && (t2 == G_SPHERE || t2 == G_CAPSULE) && !(t1 == G_CAPSULE && t2 == G_SPHERE));
}
// Support value h(d) = max over the shape of (x - centre) . d, for a unit direction d in world coordinates (float32).
__device__ __forceinline__ float supportValue(int type, f3 size, SP R, f3 d) {
  const f3 l = mk3(R[0] * d.x + R[3] * d.y + R[6] * d.z, R[1] * d.x + R[4] * d.y + R[7] * d.z, R[2] * d.x + R[5] * d.y + R[8] * d.z);   // R^T d
  switch (type) {
    case G_SPHERE: return size.x;
    case G_CAPSULE: return size.x + size.y * fabsf(l.z);
    case G_ELLIPSOID: return sqrtf(size.x * size.x * l.x * l.x + size.y * size.y * l.y * l.y + size.z * size.z * l.z * l.z);
    case G_CYLINDER: return size.x * sqrtf(l.x * l.x + l.y * l.y) + size.y * fabsf(l.z);
    case G_BOX: return size.x * fabsf(l.x) + size.y * fabsf(l.y) + size.z * fabsf(l.z);
    default: return 3.4e38f;   // unknown to this test: never separated
  }
}
// The portal search's first exit, in float32 and ahead of the work list: along the line of centres u the margin-inflated shapes
// are apart when |c2 - c1| >= h1(u) + h2(-u) + margin (mprPenetration: dot(s1.v, dir) <= 0).  With a float32 safety band the
// test is conservative: a pair it removes is one the float64 search would drop at its first step; the rest go on the list.
__device__ __forceinline__ bool centreAxisApart(const DevModel& D, const SmemLayout& L, SP sm, bool ovr, int pr) {
  const int c1 = D.pair_cg1[pr], c2 = D.pair_cg2[pr];
  const CgShape S1 = cgShape(D, L, sm, ovr, c1), S2 = cgShape(D, L, sm, ovr, c2);
  auto P1 = sm + L.cgpose + 12 * c1; auto P2 = sm + L.cgpose + 12 * c2;
  const f3 v0 = ld3(P2) - ld3(P1);
  const float dist = norm3(v0);
  if (!(dist > 1e-6f)) return false;
  const f3 u = v0 * (1.f / dist);
  const float h1 = supportValue(S1.type, S1.size, P1 + 3, u), h2 = supportValue(S2.type, S2.size, P2 + 3, mk3(-u.x, -u.y, -u.z));
  const float reach = h1 + h2 + D.pair_prm[16 * pr];
  return dist > reach + 1e-5f * (dist + reach) + 1e-6f;
}
// collide A: geom poses, broad phase (ordered survivor list per lane), closed-form narrow phase.  Survivors that need the
// convex narrow phase (MPR: float64, long, data-dependent) are appended to a batch-wide work list instead: a tile has a
// handful of them per world at different survivor positions, so inside the tile they would run at a few active lanes
// while the other warps wait; collideMprKernel runs them one pair per LANE.  Per-survivor result record in L.ctmp:
// 32 floats = hit count, up to MAXHIT hits (dist, pos, normal).
__global__ void __launch_bounds__(32 * MAXW, MINB_COLLIDE) collideKernel(int slot, int nworld, const int* __restrict__ mask, float* __restrict__ scratch) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L; const TileAux& X = tP[slot].X;
  Timeline tline(slot, MYO_K_COLLIDE);
  Ctx c = tileCtx(slot, nworld, mask, scratch);
  if (!c.live) return;
  SP sm = c.sm; Tm t = c.t;
  const int lane = threadIdx.x & 31, nw = t.nw;
  __shared__ int sCnt[MAXW][32];
  if (D.npair == 0 || (D.disableflags & (DSBL_CONTACT | DSBL_CONSTRAINT))) {
    if (t.tl == 0) { sm[L.counts] = __int_as_float(0); sm[L.counts + 3] = __int_as_float(0); }
    return;
  }
  const FieldPtrs& Fin = tP[slot].F;
  const void* tovr = Fin.p[MYO_F_GEOM_TYPE_OVR]; const void* sovr = Fin.p[MYO_F_GEOM_SIZE_OVR]; const void* povr = Fin.p[MYO_F_GEOM_POS_OVR];
  const bool ovr = tovr != nullptr;   // shapes are per world: the narrow phase reads this world's cgshape records
  const World W = c.W;
  TFOR(g, D.ncg) {
    const int f = D.cg_frame[g], gid = D.cg_geom[g];
    auto o = sm + L.cgpose + 12 * g;
    st3(o, ld3(FR_POS(sm, L, f)) + mulM(FR_MAT(sm, L, f), povr ? geomLocalPos(D, povr, c.W, gid) : ld3(D.cg_lpos + 3 * g)));
    if (ovr) {
      int type = D.cg_type[g]; f3 size = ld3(D.cg_size + 3 * g); float rb = D.cg_rbound[g];
      const int to = OUTI(tovr, gid);
      if (to >= 0 && type != G_PLANE) type = to;
      if (sovr && OUTF(sovr, 3 * gid) > 0.f) size = mk3(OUTF(sovr, 3 * gid), OUTF(sovr, 3 * gid + 1), OUTF(sovr, 3 * gid + 2));
      if (to >= 0 || (sovr && OUTF(sovr, 3 * gid) > 0.f)) rb = geomRbound(type, size);
      auto sh = sm + L.cgshape + 5 * g;
      sh[0] = __int_as_float(type); st3(sh + 1, size); sh[4] = rb;
    }
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
    for (int pr = p0; pr < p1; pr++) cnt += broadPass(D, L, sm, ovr, pr) ? 1 : 0;
    sCnt[t.tl][lane] = cnt;
    TSYNC();
    int off = 0; ncand = 0;
    for (int j = 0; j < nw; j++) { const int v = sCnt[j][lane]; if (j < t.tl) off += v; ncand += v; }
    for (int pr = p0; pr < p1; pr++) if (broadPass(D, L, sm, ovr, pr)) cand[off++] = pr;
    TSYNC();
  }
  if (t.tl == 0) sm[L.counts + 3] = __int_as_float(ncand);
  // ---- narrow phase of the closed-form pairs; MPR pairs go to the work list (one atomic per warp and survivor position)
  auto tmp = sm + L.ctmp;
  const int w = (int)c.W.w;
  for (int k = t.tl; k < ncand; k += nw) {
    const int pr = cand[k];
    bool mp = isMprPair(D, L, sm, ovr, pr);
    if (mp && centreAxisApart(D, L, sm, ovr, pr)) { mp = false; tmp[32 * k] = __int_as_float(0); }   // no contact: not listed
    else if (!mp) {
      Hit h[MAXHIT];
      int sw;
      const int n = collidePair<false>(slot, sm, pr, h, &sw);
      auto o = tmp + 32 * k;
      o[0] = __int_as_float(n | (sw << 8));
      for (int q = 0; q < n; q++) { o[1 + 7 * q] = h[q].dist; st3(o + 2 + 7 * q, h[q].pos); st3(o + 5 + 7 * q, h[q].nrm); }
    }
    const unsigned act = __activemask();
    const unsigned mm = __ballot_sync(act, mp);
    if (mm) {
      const int leader = __ffs(mm) - 1;
      int base = 0;
      if (lane == leader) base = atomicAdd(X.mpr_count, __popc(mm));
      base = __shfl_sync(act, base, leader);
      if (mp) {
        const int at = base + __popc(mm & ((1u << lane) - 1u));
        if (at < X.mpr_cap) X.mpr_list[at] = make_int2(w, k);
      }
    }
  }
}
// collide MPR: one listed (world, survivor) pair per lane, all lanes busy; the result goes to the survivor's own record, so
// the (atomic) order of the list does not matter
#define MPR_BATCH 512
__global__ void __launch_bounds__(128) collideMprKernel(int slot, float* __restrict__ scratch) {
  const SmemLayout& L = tP[slot].L; const TileAux& X = tP[slot].X;
  Timeline tline(slot, MYO_K_COLLIDE_MPR);
  const int n = min(*X.mpr_count, X.mpr_cap);
  // Two passes per batch of MPR_BATCH pairs: most pairs of the list are separated and leave the portal search after one or
  // two support evaluations, a few run tens of refinement steps.  One pair per lane straight through leaves ~5 of 32 lanes
  // busy (ncu: 5.25 threads per instruction); so the batch is filtered first, and the pairs that have a portal are
  // compacted in shared memory and run in full by consecutive lanes.  (Each record has its own slot: order is free.)
#ifdef MPR_FILTER
  __shared__ int sList[MPR_BATCH];
  __shared__ int sCount;
#endif
  auto record = [&](int i, bool filter) {
    const int2 e = X.mpr_list[i];
    SP sm; sm.p = scratch + (size_t)(e.x >> 5) * L.total * SCR + (e.x & 31);
    const int pr = __float_as_int(sm[L.cand + e.y]);
    Hit h[MAXHIT];
    int sw;
    const int nh = collidePair<true>(slot, sm, pr, h, &sw, filter);
    if (nh < 0) return false;
    auto o = sm + L.ctmp + 32 * e.y;
    o[0] = __int_as_float(nh | (sw << 8));
    if (nh) { o[1] = h[0].dist; st3(o + 2, h[0].pos); st3(o + 5, h[0].nrm); }
    return true;
  };
  const int batch = min(MPR_BATCH, max((int)blockDim.x, (int)(((n + gridDim.x - 1) / gridDim.x + blockDim.x - 1) / blockDim.x * blockDim.x)));
  for (int base = blockIdx.x * batch; base < n; base += gridDim.x * batch) {
#ifdef MPR_FILTER
    if (threadIdx.x == 0) sCount = 0;
    __syncthreads();
#endif
#ifndef MPR_FILTER
    // With the float32 centre-axis test in collideKernel four of five listed pairs do have a portal: a filter pass would only
    // repeat their portal search (-0.9 % step time without it).  -DMPR_FILTER restores it.
    for (int k = threadIdx.x; k < batch && base + k < n; k += blockDim.x) record(base + k, false);
#else
    for (int k = threadIdx.x; k < batch && base + k < n; k += blockDim.x)
      if (!record(base + k, true)) sList[atomicAdd(&sCount, 1)] = base + k;
    __syncthreads();
    const int m = sCount;
    if (X.dbg && threadIdx.x == 0) {   // developer counters: pairs listed, pairs with a portal
      unsigned long long* q = (unsigned long long*)X.dbg + X.tl_off + 48;
      atomicAdd(q, (unsigned long long)min(batch, n - base)); atomicAdd(q + 1, (unsigned long long)m);
    }
    for (int k = threadIdx.x; k < m; k += blockDim.x) record(sList[k], false);
    __syncthreads();
#endif
  }
}
// collide B: in-order compaction of the hits into contact records -- contact order (hence geom1 / geom2 / efc_address) is
// MuJoCo's
__device__ __noinline__ void compactTile(int slot, SP sm, Tm t, unsigned lmask) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L;
  const int lane = threadIdx.x & 31, nw = t.nw;
  __shared__ int sCnt[MAXW][32];
  if (D.npair == 0 || (D.disableflags & (DSBL_CONTACT | DSBL_CONSTRAINT))) return;
  SPI cand = asInt(sm + L.cand);
  auto tmp = sm + L.ctmp;
  const int ncand = __float_as_int(sm[L.counts + 3]);
  const int kc = (__reduce_max_sync(lmask, ncand) + nw - 1) / nw, k0 = min(ncand, t.tl * kc), k1 = min(ncand, k0 + kc);
  int cnt = 0;
  for (int kb = k0; kb < k1; kb += 8) {   // (hit counts of eight survivors loaded together)
    int h[8];
#pragma unroll
    for (int u = 0; u < 8; u++) h[u] = kb + u < k1 ? __float_as_int(tmp[32 * (kb + u)]) : 0;
#pragma unroll
    for (int u = 0; u < 8; u++) cnt += h[u] & 0xff;
  }
  sCnt[t.tl][lane] = cnt;
  TSYNC();
  int at = 0, ncon = 0;
  for (int j = 0; j < nw; j++) { const int v = sCnt[j][lane]; if (j < t.tl) at += v; ncon += v; }
  for (int k = k0; k < k1; k++) {
    const int nsw = __float_as_int(tmp[32 * k]), n = nsw & 0xff, sw = nsw >> 8;
    for (int q = 0; q < n; q++, at++) {
      if (at >= D.nconmax) continue;
      const int pr = cand[k];
      auto cr = sm + L.con + 12 * at; auto o = tmp + 32 * k + 1 + 7 * q;
      float rec[7];
#pragma unroll
      for (int e = 0; e < 7; e++) rec[e] = o[e];
#pragma unroll
      for (int e = 0; e < 7; e++) cr[e] = rec[e];
      cr[7] = __int_as_float(pr | (sw << 30)); cr[8] = __int_as_float(-1); cr[9] = __int_as_float(D.pair_dim[pr]);
      // dofs that move exactly one of the two bodies (static per pair): the only non-zero columns of the contact's rows
      const int f1 = D.cg_frame[D.pair_cg1[pr]], f2 = D.cg_frame[D.pair_cg2[pr]];
      cr[10] = __int_as_float(D.fr_mask[2 * f1] ^ D.fr_mask[2 * f2]); cr[11] = __int_as_float(D.fr_mask[2 * f1 + 1] ^ D.fr_mask[2 * f2 + 1]);
    }
  }
  if (t.tl == 0) {
    if (ncon > D.nconmax) {
      sm[L.wstate + WS_STATUS] = __int_as_float(__float_as_int(sm[L.wstate + WS_STATUS]) | MYO_WORLD_NCON_OVERFLOW);
      ncon = D.nconmax;
    }
    sm[L.counts] = __int_as_float(ncon);
  }
}
__global__ void __launch_bounds__(32 * MAXW, 2) collideCompactKernel(int slot, int nworld, const int* __restrict__ mask, float* __restrict__ scratch) {
  Timeline tline(slot, MYO_K_COLLIDE_COMPACT);
  Ctx c = tileCtx(slot, nworld, mask, scratch);
  const unsigned lmask = __ballot_sync(0xffffffffu, c.live);
  if (!c.live) return;
  compactTile(slot, c.sm, c.t, lmask);
}

// ------------------------------------------------------------------------------------------ newton: mj_makeConstraint, mj_fwdConstraint
// The data-dependent middle of the step with LANE = WORLD as well.  Loops over a world's rows / contacts take the lane's
// own count as bound (lanes with fewer rows idle, no barrier inside such a loop); every barrier and every shared-memory
// sum sits in CTA-uniform control flow, whose bounds are the tile's maxima (warp reductions of per-lane values that are
// identical in every warp of the tile).  A world that has converged is masked out and the tile iterates until its slowest
// world is done.  Per-lane scalars (cost, step size, iteration count, flags) are replicated in every warp: they come from
// sums that are bit-identical in all warps, so all warps take the same branches.
// One shared buffer for every tile-wide scan / minimum / sum (each use is bracketed by its own pair of barriers).  Small on
// purpose: with 4 sums per call the solver kernel's two CTAs fit the 132 KB shared-memory configuration and L1 keeps 124 KB
// instead of 92 (12 sums per call: 18 KB more per CTA).
#ifndef MAXSUM
#define MAXSUM 4
#endif
__device__ __forceinline__ float* tileRedBuf() { __shared__ float red[MAXSUM * MAXW * 32]; return red; }
struct TileScan { int off, total; };
// exclusive prefix of a per-(warp, lane) count over the warps of the tile, in warp order: warp j owns a contiguous chunk
// of an item range, so per-lane ordered compaction needs no serial walk
__device__ __noinline__ TileScan tileScan(Tm t, int cnt) {
  int (*sc)[32] = (int (*)[32])tileRedBuf();
  const int lane = threadIdx.x & 31;
  sc[t.tl][lane] = cnt;
  __syncthreads();
  TileScan r; r.off = 0; r.total = 0;
  for (int j = 0; j < t.nw; j++) { const int v = sc[j][lane]; if (j < t.tl) r.off += v; r.total += v; }
  __syncthreads();
  return r;
}
// The tendon stage's output fields (ten_length, dense ten_J, the path points), from what tendons() left in the scratch: their
// own launch on the output stream (tenoutKernel) -- a quarter of the tendon kernel's time was these stores, and velact and the
// solver wait for that kernel.
__device__ __noinline__ void tendonOutputs(int slot, int fslot, SP sm, Tm t, World W) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L; const FieldPtrs& F = tP[fslot].F;
  void* wx = F.p[MYO_F_WRAP_XPOS]; void* wo = F.p[MYO_F_WRAP_OBJ]; void* wa = F.p[MYO_F_TEN_WRAPADR]; void* wn = F.p[MYO_F_TEN_WRAPNUM];
  const bool wrout = wx || wo || wa || wn;
  if (wrout) {
    // Path points in mj_tendon's order: per unit its first site, the two tangent points if it wraps (wrap_obj = geom id),
    // the last site at a tendon / pulley-branch end (wrap_obj -1), a marker per pulley (wrap_obj -2, position unset = 0).
    // Offsets = static offsets + 2 x (wrapping units before the item); entries past the last point are zero.
    // gpre[k] = wrapping units among the first k geom units: counted by per-warp chunks, prefix over the warps
    const int ngu = D.ngunit;
    const int chunk = (ngu + TEAM - 1) / TEAM, k0 = min(ngu, t.tl * chunk), k1 = min(ngu, k0 + chunk);
    int cnt = 0;
    for (int k = k0; k < k1; k++) cnt += sm[L.uwrap + D.gunit[k]] != 0.f ? 1 : 0;
    const TileScan sc = tileScan(t, cnt);
    int acc = sc.off;
    for (int k = k0; k < k1; k++) { sm[L.gpre + k] = __int_as_float(acc); acc += sm[L.uwrap + D.gunit[k]] != 0.f ? 1 : 0; }
    if (t.tl == 0) sm[L.gpre + ngu] = __int_as_float(sc.total);
    TSYNC();
    auto put = [&](int i, f3 p, int obj) {
      if (wx) { OUTF(wx, 3 * i) = p.x; OUTF(wx, 3 * i + 1) = p.y; OUTF(wx, 3 * i + 2) = p.z; }
      if (wo) OUTI(wo, i) = obj;
    };
    TFOR(u, D.nunit) {
      const int go = D.unit_gord[u], g = D.unit_geom[u];
      const bool wrapped = sm[L.uwrap + u] != 0.f;
      const int off = D.unit_woff[u] + 2 * __float_as_int(sm[L.gpre + go]);
      put(off, ld3(sm + L.upts + 6 * u), -1);
      if (wrapped) { auto wp = sm + L.uwpt + 6 * go; put(off + 1, ld3(wp), g); put(off + 2, ld3(wp + 3), g); }
      if (D.unit_end[u]) put(off + (wrapped ? 3 : 1), ld3(sm + L.upts + 6 * u + 3), -1);
    }
    TFOR(i, D.npulley) put(D.pul_off[i] + 2 * __float_as_int(sm[L.gpre + D.pul_gord[i]]), mk3(0.f, 0.f, 0.f), -2);
    TFOR(k, D.nten) {
      const int g0 = __float_as_int(sm[L.gpre + D.ten_gord[k]]), g1 = __float_as_int(sm[L.gpre + D.ten_gord[k + 1]]);
      if (wa) OUTI(wa, k) = D.ten_woff[k] + 2 * g0;
      if (wn) OUTI(wn, k) = D.ten_woff[k + 1] - D.ten_woff[k] + 2 * (g1 - g0);
    }
    // entries past the last point: only those a previous step can have written (all geom units wrapped); the rest of the
    // buffer keeps the zeros it was given when it was bound (api.cu)
    for (int i = D.ten_woff[D.nten] + 2 * __float_as_int(sm[L.gpre + ngu]) + t.tl; i < D.ten_woff[D.nten] + 2 * ngu; i += TEAM) put(i, mk3(0.f, 0.f, 0.f), 0);
  }
  void* tl = F.p[MYO_F_TEN_LENGTH]; void* tj = F.p[MYO_F_TEN_J];
  if (tl) TFOR(k, D.nten) OUTF(tl, k) = sm[L.tenlen + k];
  // dense ten_J and actuator_moment rows from the compile-time source map of every entry (plan.cu: sparse index, -1 zero,
  // -2 the gear): independent items, four per iteration -- the serial merge of a sparse row into a dense one put a table
  // lookup and a branch on every element's chain
  const int nv = D.nv;
  if (tj) {
    const int n = D.nten * nv;
    for (int i0 = t.tl; i0 < n; i0 += 4 * TEAM) {
      float v[4];
#pragma unroll
      for (int u = 0; u < 4; u++) { const int i = i0 + u * TEAM; const int src = i < n ? D.ten_dense[i] : -1; v[u] = src >= 0 ? sm[L.tenJ + src] : 0.f; }
#pragma unroll
      for (int u = 0; u < 4; u++) if (i0 + u * TEAM < n) OUTF(tj, i0 + u * TEAM) = v[u];
    }
  }
  void* am = F.p[MYO_F_ACTUATOR_MOMENT];
  if (am) {   // mj_transmission: needs the tendon rows and the model only
    const int n = D.nu * nv;
    for (int i0 = t.tl; i0 < n; i0 += 4 * TEAM) {
      float v[4];
#pragma unroll
      for (int u = 0; u < 4; u++) {
        const int i = i0 + u * TEAM; const int src = i < n ? D.act_dense[i] : -1;
        const float gear = src != -1 ? D.act_gear[i / nv] : 0.f;
        v[u] = src >= 0 ? gear * sm[L.tenJ + src] : gear;   // (-2: the gear itself; -1: gear = 0)
      }
#pragma unroll
      for (int u = 0; u < 4; u++) if (i0 + u * TEAM < n) OUTF(am, i0 + u * TEAM) = v[u];
    }
  }
}
#ifndef MINB_TENOUT
#define MINB_TENOUT 3
#endif
__global__ void __launch_bounds__(32 * MAXW, MINB_TENOUT) tenoutKernel(int slot, int fslot, int nworld, const int* __restrict__ mask, float* __restrict__ scratch) {
  Timeline tline(slot, MYO_K_TENOUT);
  TILE_LOOP(tile) {
    Ctx c = tileCtxAt(slot, nworld, mask, scratch, tile);
    if (!c.live) return;
    tendonOutputs(slot, fslot, c.sm, c.t, c.W);
    TSYNC();
  }
}
__device__ __noinline__ int tileMinI(Tm t, int v) {
  int (*sc)[32] = (int (*)[32])tileRedBuf();
  const int lane = threadIdx.x & 31;
  sc[t.tl][lane] = v;
  __syncthreads();
  int m = 0x7fffffff;
  for (int j = 0; j < t.nw; j++) m = min(m, sc[j][lane]);
  __syncthreads();
  return m;
}
// N <= MAXSUM sums at once (one pair of barriers)
template <int N> __device__ __forceinline__ void tileSumN(Tm t, float* v) {
  static_assert(N <= MAXSUM, "tileSumN: raise MAXSUM");
  float* red = tileRedBuf() + (threadIdx.x & 31);
#pragma unroll
  for (int k = 0; k < N; k++) red[(k * MAXW + t.tl) * 32] = v[k];
  __syncthreads();
#pragma unroll
  for (int k = 0; k < N; k++) { float s = 0.f; for (int j = 0; j < t.nw; j++) s += red[(k * MAXW + j) * 32]; v[k] = s; }
  __syncthreads();
}

__device__ __forceinline__ void limitState(const DevModel& D, const SmemLayout& L, SP sm, int li, float* dlo, float* dhi, float* margin, const float** prm) {
  const bool isj = li < D.nlimj;
  *prm = isj ? D.limj_prm + 12 * li : D.limt_prm + 12 * (li - D.nlimj);
  const float val = isj ? sm[L.qpos + D.limj_q[li]] : sm[L.tenlen + D.limt_ten[li - D.nlimj]];
  *margin = (*prm)[2]; *dlo = val - (*prm)[0]; *dhi = (*prm)[1] - val;
}
__device__ __forceinline__ int contactRows(const DevModel& D, const SmemLayout& L, SP sm, int c) {
  auto cr = sm + L.con + 12 * c;
  const int pr = CON_PAIR(__float_as_int(cr[7])), dim = __float_as_int(cr[9]);
  const float inc = D.pair_prm[16 * pr] - D.pair_prm[16 * pr + 1];
  return cr[0] < inc ? (dim == 1 ? 1 : 2 * (dim - 1)) : 0;
}
// mj_makeConstraint: rows in MuJoCo's order (equality, joint limits, tendon limits, contacts); returns the lane's nefc
__device__ __noinline__ int makeConstraintTile(int slot, SP sm, Tm t, int ncon, int nconT, int* status) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L;
  const int nv = D.nv, njmax = D.njmax, nw = t.nw;
  auto qpos = sm + L.qpos; auto qvel = sm + L.qvel;
  if (D.disableflags & DSBL_CONSTRAINT) return 0;
  TFOR(e, D.neq) {
    const int d1 = D.eq_dof1[e], d2 = D.eq_dof2[e], q1 = D.eq_q1[e], q2 = D.eq_q2[e];
    auto dt = D.eq_data + 5 * e;
    float pos, dpoly = 0.f, vel = qvel[d1];
    if (d2 >= 0) {
      const float x = qpos[q2] - D.qpos0[q2];
      pos = qpos[q1] - D.qpos0[q1] - (dt[0] + x * (dt[1] + x * (dt[2] + x * (dt[3] + x * dt[4]))));
      dpoly = dt[1] + x * (2.f * dt[2] + x * (3.f * dt[3] + x * 4.f * dt[4]));
      vel -= dpoly * qvel[d2];
    } else pos = qpos[q1] - D.qpos0[q1] - dt[0];
    setRow(slot, sm, e, ROW_EQ, e, -dpoly, 0, pos, 0.f, D.eq_prm[8 * e + 7], D.eq_prm + 8 * e, vel, 0.f);
  }
  int nefc = D.neq;
  // ---- joint limits, then tendon limits
  const int nlim = D.nlimj + D.nlimt;
  if (nlim > 0) {
    const int chunk = (nlim + nw - 1) / nw, l0 = min(nlim, t.tl * chunk), l1 = min(nlim, l0 + chunk);
    int cnt = 0;
    for (int li = l0; li < l1; li++) {
      float dlo, dhi, margin; const float* prm;
      limitState(D, L, sm, li, &dlo, &dhi, &margin, &prm);
      cnt += (dlo < margin) + (dhi < margin);
    }
    const TileScan sc = tileScan(t, cnt);
    int at = nefc + sc.off;
    for (int li = l0; li < l1; li++) {
      float dlo, dhi, margin; const float* prm;
      limitState(D, L, sm, li, &dlo, &dhi, &margin, &prm);
      int r0 = -1, r1 = -1;
      if ((dlo < margin) || (dhi < margin)) {
        const bool isj = li < D.nlimj;
        const float v = isj ? qvel[D.limj_dof[li]] : sm[L.tenvel + D.limt_ten[li - D.nlimj]];
        const int id = isj ? li : li - D.nlimj, type = isj ? ROW_LIMJ : ROW_LIMT;
        if (dlo < margin) { if (at < njmax) { r0 = at; setRow(slot, sm, at, type, id, 1.f, 0, dlo, margin, prm[10], prm + 3, v, 0.f); } at++; }
        if (dhi < margin) { if (at < njmax) { r1 = at; setRow(slot, sm, at, type, id, -1.f, 0, dhi, margin, prm[10], prm + 3, -v, 0.f); } at++; }
      }
      sm[L.limrow + 2 * li] = __int_as_float(r0); sm[L.limrow + 2 * li + 1] = __int_as_float(r1);
    }
    nefc += sc.total;
  }
  // ---- contacts: row addresses, base Jacobian rows, pyramid rows
  if (nconT > 0) {
    const int nb = D.maxbase;
    const int chunk = (nconT + nw - 1) / nw, c0 = min(nconT, t.tl * chunk), c1 = min(nconT, c0 + chunk);
    int cnt = 0;
    for (int c = c0; c < c1; c++) cnt += c < ncon ? contactRows(D, L, sm, c) : 0;
    const TileScan sc = tileScan(t, cnt);
    int at = nefc + sc.off, unfit = 0x7fffffff;
    for (int c = c0; c < c1; c++) {
      if (c >= ncon) break;
      const int rows = contactRows(D, L, sm, c);
      sm[L.con + 12 * c + 8] = __int_as_float((rows > 0 && at + rows <= njmax) ? at : -1);
      if (rows > 0 && at + rows > njmax) unfit = min(unfit, at);
      at += rows;
    }
    nefc += sc.total;
    // a contact that does not fit is dropped whole, and so is every later one (their addresses lie beyond njmax): the
    // rows that exist are [0, unfit), all written below
    unfit = tileMinI(t, unfit);
    if (unfit < nefc) { *status |= MYO_WORLD_NEFC_OVERFLOW; nefc = min(unfit, njmax); }
    // base Jacobian rows: one item = (contact, dof); contacts past the lane's own count up to the tile's get zero rows
    // (the Hessian product below runs over the tile's count)
    for (int c = t.tl; c < nconT; c += nw) {
      auto cr = sm + L.con + 12 * c;
      auto Jc = sm + L.conJ + (c * nb) * nv;
      if (c >= ncon) {
        for (int k = 0; k < nb * nv; k++) Jc[k] = 0.f;
        continue;
      }
      const unsigned mlo = (unsigned)__float_as_int(cr[10]), mhi = (unsigned)__float_as_int(cr[11]);
      const int prw = __float_as_int(cr[7]);
      const int f1 = D.cg_frame[CON_SWAP(prw) ? D.pair_cg2[CON_PAIR(prw)] : D.pair_cg1[CON_PAIR(prw)]];
      const unsigned f1lo = (unsigned)D.fr_mask[2 * f1], f1hi = (unsigned)D.fr_mask[2 * f1 + 1];
      const f3 pos = ld3(cr + 1), n = ld3(cr + 4);
      f3 y, z; makeFrame(n, &y, &z);
      // four dofs per iteration, their cdof / tree-centre loads issued together (one dof at a time was a chain of ~2 nv
      // scratch latencies per contact: the bulk of this kernel)
      for (int k0 = 0; k0 < nv; k0 += 4) {
        float cdv[4][6], rc[4][3]; bool on[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
          const int k = k0 + u;
          on[u] = k < nv && (((k < 32 ? mlo : mhi) >> (k & 31)) & 1u);   // off: dof k moves both bodies or neither
          if (on[u]) {
            auto cd = sm + L.cdof + 6 * k; auto rt = sm + L.rootcom + 3 * D.fr_root[D.dof_frame[k]];
#pragma unroll
            for (int e = 0; e < 6; e++) cdv[u][e] = cd[e];
#pragma unroll
            for (int e = 0; e < 3; e++) rc[u][e] = rt[e];
          }
        }
#pragma unroll
        for (int u = 0; u < 4; u++) {
          const int k = k0 + u;
          if (k >= nv) break;
          auto J = Jc + k;
          if (!on[u]) { J[0] = 0.f; if (nb > 1) { J[nv] = 0.f; J[2 * nv] = 0.f; } if (nb > 3) J[3 * nv] = 0.f; continue; }
          const float sgn = (((k < 32 ? f1lo : f1hi) >> (k & 31)) & 1u) ? -1.f : 1.f;   // it moves exactly one of them: the first (-) or the second (+)
          const f3 w = mk3(cdv[u][0], cdv[u][1], cdv[u][2]);
          const f3 jr = w * sgn;
          const f3 jp = (mk3(cdv[u][3], cdv[u][4], cdv[u][5]) + cross(w, pos - mk3(rc[u][0], rc[u][1], rc[u][2]))) * sgn;
          J[0] = dot(n, jp);
          if (nb > 1) { J[nv] = dot(y, jp); J[2 * nv] = dot(z, jp); }
          if (nb > 3) J[3 * nv] = dot(n, jr);
        }
      }
    }
    TSYNC();
    auto bx = sm + L.basev;
    baseMul(slot, sm, t, ncon, qvel, bx);
    TSYNC();
    for (int c = t.tl; c < ncon; c += nw) {
      auto cr = sm + L.con + 12 * c;
      const int adr = __float_as_int(cr[8]);
      if (adr < 0) continue;
      const int pr = CON_PAIR(__float_as_int(cr[7])), dim = __float_as_int(cr[9]);
      auto pp = D.pair_prm + 16 * pr;
      const float margin = pp[0] - pp[1], dist = cr[0];
      const int g1 = D.pair_cg1[pr], g2 = D.pair_cg2[pr];
      const float tran = D.cg_invweight[2 * g1] + D.cg_invweight[2 * g2];
      const float kb[7] = {pp[14], pp[15], pp[9], pp[10], pp[11], pp[12], pp[13]};
      const int bn = c * nb;
      if (dim == 1) setRow(slot, sm, adr, ROW_CON, bn, 0.f, -1, dist, margin, tran, kb, bx[bn], 0.f);
      else {
        const float mu0 = pp[2] * rsqrtf(D.impratio);
        const float dA0 = tran + pp[2] * pp[2] * tran;   // first pyramid row sets the shared regulariser
        for (int k = 1; k < dim; k++) {
          const float mu = pp[1 + k];
          const int bt = bn + k;
          setRow(slot, sm, adr + 2 * (k - 1), ROW_CON, bn, mu, bt, dist, margin, dA0, kb, bx[bn] + mu * bx[bt], 2.f * mu0 * mu0);
          setRow(slot, sm, adr + 2 * (k - 1) + 1, ROW_CON, bn, -mu, bt, dist, margin, dA0, kb, bx[bn] - mu * bx[bt], 2.f * mu0 * mu0);
        }
      }
    }
  }
  if (nefc > njmax) { *status |= MYO_WORLD_NEFC_OVERFLOW; nefc = njmax; }
  TSYNC();
  return nefc;
}

// H += Jb^T (W Jb) over the tile's contacts, in 4 x 4 register blocks of the lower triangle, one block per warp at a time.
// The weighted rows U = W Jb (W: the contact's symmetric weight matrix in base-row space, conW) are formed in registers from
// the NB base rows of each contact -- as many loads as reading a stored U, no U sweep, and half the Jacobian footprint in L2.
// The block's columns are immediate offsets from two row pointers (clamping them to nv - 1 cost 1.7 % of the step in index
// arithmetic); columns past nv read the next row / region of the scratch (finite or not, never stored).
template <int NB> __device__ __forceinline__ void syrkBlocks(int slot, SP sm, Tm t, SP H, int nconT) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L;
  const int nv = D.nv, nw = t.nw;
  const int nblk = (nv + 3) >> 2, nB = (nblk * (nblk + 1)) >> 1;
  for (int blk = t.tl; blk < nB; blk += nw) {
    int R = 0, rem = blk;
    while (rem > R) { rem -= R + 1; R++; }   // blk -> (R, C), C <= R
    const int C = rem;
    float acc[4][4];
#pragma unroll
    for (int a = 0; a < 4; a++)
#pragma unroll
      for (int b = 0; b < 4; b++) acc[a][b] = 0.f;
    auto Jr = sm + L.conJ + 4 * R; auto Jc = sm + L.conJ + 4 * C;
    TFOR_PRAGMA(SYRK_UNROLL)
    for (int c = 0; c < nconT; c++) {
      auto Wi = sm + L.conW + 8 * c;
      float jr[NB][4], jc[NB][4], Wn[NB], Wt[NB];
#pragma unroll
      for (int b = 0; b < NB; b++) {
        Wn[b] = Wi[b]; Wt[b] = b > 0 ? Wi[3 + b] : 0.f;   // Wn = (Wnn, Wnt[0..]), Wt = (-, Wtt[0..])
#pragma unroll
        for (int a = 0; a < 4; a++) {
#ifdef SYRK_CLAMP
          jr[b][a] = 4 * R + a < nv ? LDJ(Jr[(c * NB + b) * nv + a]) : 0.f; jc[b][a] = 4 * C + a < nv ? LDJ(Jc[(c * NB + b) * nv + a]) : 0.f;
#else
          jr[b][a] = LDJ(Jr[(c * NB + b) * nv + a]); jc[b][a] = LDJ(Jc[(c * NB + b) * nv + a]);
#endif
        }
      }
#pragma unroll
      for (int a = 0; a < 4; a++) {
        float u = Wn[0] * jr[0][a];   // U row of the normal: Wnn jn + sum_q Wnt[q] jq
#pragma unroll
        for (int b = 1; b < NB; b++) u = fmaf(Wn[b], jr[b][a], u);
#pragma unroll
        for (int q = 0; q < 4; q++) acc[a][q] = fmaf(u, jc[0][q], acc[a][q]);
#pragma unroll
        for (int b = 1; b < NB; b++) {
          const float ub = fmaf(Wn[b], jr[0][a], Wt[b] * jr[b][a]);   // U row of tangent b: Wnt jn + Wtt jb
#pragma unroll
          for (int q = 0; q < 4; q++) acc[a][q] = fmaf(ub, jc[b][q], acc[a][q]);
        }
      }
    }
#pragma unroll
    for (int a = 0; a < 4; a++)
#pragma unroll
      for (int b = 0; b < 4; b++) {
        const int r = 4 * R + a, c = 4 * C + b;
        if (r < nv && c <= r) H[TRI(r) + c] += acc[a][b];
      }
  }
}
// H = M + J^T D_active J (lower triangle, row-packed); the unfactored matrix is also saved for the later passes' updates.
// Contacts enter through their base rows (normal, tangents, torsion): with W the small symmetric weight matrix of a
// contact's active pyramid rows in base-row space, J^T D J = Jb^T (W Jb), a dense [nv x nbase] x [nbase x nv] product over
// the tile's nbase = maxbase * nconT base rows, done in 4 x 4 register blocks of H (syrkBlocks: one block per warp at a
// time, W Jb formed in registers, the warps stream the same rows of Jb through L1 together).
__device__ __noinline__ void hessianAssemble(int slot, SP sm, Tm t, SP H, int ncon, int nconT, int nefc) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L;
  const int nv = D.nv, nb = D.maxbase, nw = t.nw;
  TFOR(i, TRI(nv)) H[i] = 0.f;
  TSYNC();
  TFOR(e, D.nM) H[TRI(D.M_row[e]) + D.M_col[e]] = sm[L.qM + e];
  TSYNC();
  TFOR(k, nv) {
    const int li = D.dof_limidx[k];
    if (li < 0) continue;
    for (int q = 0; q < 2; q++) {
      const int r = __float_as_int(sm[L.limrow + 2 * li + q]);
      if (r >= 0 && EFC(E_JAR, r) < 0.f) H[TRI(k) + k] += EFC(E_D, r);
    }
  }
  TSYNC();
  if (t.tl == 0)
    for (int e = 0; e < D.neq; e++) {
      const int d1 = D.eq_dof1[e], d2 = D.eq_dof2[e];
      const float Dr = EFC(E_D, e), c = EFC(E_COEF, e);
      H[TRI(d1) + d1] += Dr;
      if (d2 >= 0) { H[TRI(max(d1, d2)) + min(d1, d2)] += Dr * c; H[TRI(d2) + d2] += Dr * c * c; }
    }
  for (int lt = 0; lt < D.nlimt; lt++)
    for (int q = 0; q < 2; q++) {
      TSYNC();
      const int r = __float_as_int(sm[L.limrow + 2 * (D.nlimj + lt) + q]);
      if (r < 0 || !(EFC(E_JAR, r) < 0.f)) continue;   // per-lane; no barrier below in this iteration
      const int tn = D.limt_ten[lt], n0 = D.ten_nzstart[tn], cnt = D.ten_nzstart[tn + 1] - n0;
      const float Dr = EFC(E_D, r);
      TFOR(i, cnt * cnt) {
        const int a = i / cnt, b = i - a * cnt;
        const int da = D.nz_dof[n0 + a], db = D.nz_dof[n0 + b];
        if (da >= db) H[TRI(da) + db] += Dr * sm[L.tenJ + n0 + a] * sm[L.tenJ + n0 + b];
      }
    }
  TSYNC();
  if (nconT > 0) {   // tile-uniform
    // weights in base-row space, once per contact: W = sum_p D_p w_p w_p^T, w_p = e_n +- mu e_t  ->  conW[8 c] = Wnn, Wnt[3], Wtt[3]
    // (the row records sit at per-lane addresses: these are the only uncoalesced loads of the assembly)
    for (int c = t.tl; c < nconT; c += nw) {
      auto cr = sm + L.con + 12 * c;
      float Wnn = 0.f, Wnt[3] = {0.f, 0.f, 0.f}, Wtt[3] = {0.f, 0.f, 0.f};
      if (c < ncon) {
        const int adr = __float_as_int(cr[8]), dim = __float_as_int(cr[9]);
        if (adr >= 0) {
          if (dim == 1) { if (EFC(E_JAR, adr) < 0.f) Wnn = EFC(E_D, adr); }
          else
            for (int q = 1; q < dim; q++) {
              const int r0 = adr + 2 * (q - 1);
              const float mu = EFC(E_COEF, r0);
              const float Dp = EFC(E_JAR, r0) < 0.f ? EFC(E_D, r0) : 0.f, Dm = EFC(E_JAR, r0 + 1) < 0.f ? EFC(E_D, r0 + 1) : 0.f;
              Wnn += Dp + Dm; Wnt[q - 1] = mu * (Dp - Dm); Wtt[q - 1] = mu * mu * (Dp + Dm);
            }
        }
      }
      auto Wo = sm + L.conW + 8 * c;
      Wo[0] = Wnn; Wo[1] = Wnt[0]; Wo[2] = Wnt[1]; Wo[3] = Wnt[2]; Wo[4] = Wtt[0]; Wo[5] = Wtt[1]; Wo[6] = Wtt[2];
    }
    TSYNC();
    if (nb == 3) syrkBlocks<3>(slot, sm, t, H, nconT);
    else if (nb == 4) syrkBlocks<4>(slot, sm, t, H, nconT);
    else syrkBlocks<1>(slot, sm, t, H, nconT);
    TSYNC();
  }
  TFOR(i, TRI(nv)) sm[L.Hsave + i] = H[i];   // keep the unfactored matrix: later passes of the tile update it (hessianUpdate)
}
// Later Newton passes of a tile: only the rows whose active state flipped in the last line search change H, each by a
// rank-1 term +-D_r J_r^T J_r (the flags were left in E_JV by the row update): per-lane flip lists, then the tile's warps
// apply them two at a time as dense vectors staged in shared memory (va, vb: two nv-vectors that are free here).
__device__ __noinline__ void hessianUpdate(int slot, SP sm, Tm t, SP H, int nefc, SP va, SP vb) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L;
  const int nv = D.nv, nw = t.nw;
  for (int i0 = t.tl; i0 < TRI(nv); i0 += 8 * nw) {   // (eight loads in flight: a plain striped loop pays one scratch latency per element)
    float h[8];
#pragma unroll
    for (int u = 0; u < 8; u++) h[u] = i0 + u * nw < TRI(nv) ? sm[L.Hsave + i0 + u * nw] : 0.f;
#pragma unroll
    for (int u = 0; u < 8; u++) if (i0 + u * nw < TRI(nv)) H[i0 + u * nw] = h[u];
  }
  // Each lane's flipped rows as a list (E_FORCE is free between the force evaluations): the lanes of a warp then take their
  // i-th flipped row together -- walking all rows instead would run the update body once per row that flipped in ANY of the
  // 32 worlds, with one or two lanes active each time.
  // (built in order by per-warp chunks of the row range: count, prefix over the warps, write)
  const int nrT = __reduce_max_sync(__activemask(), nefc) - D.neq, chunk = (max(nrT, 0) + nw - 1) / nw;
  const int r0 = min(nefc, D.neq + t.tl * chunk), r1 = min(nefc, r0 + chunk);
  int cnt = 0;
  for (int r = r0; r < r1; r++) cnt += EFC(E_JV, r) != 0.f ? 1 : 0;
  const TileScan sc = tileScan(t, cnt);
  int at = sc.off;
  for (int r = r0; r < r1; r++) if (EFC(E_JV, r) != 0.f) { EFC(E_FORCE, at) = __int_as_float(r); at++; }
  const int nflip = sc.total;
  TSYNC();
  // Two flipped rows per round: the tile's warps expand them into dense vectors v (nv entries per lane, shared memory; one
  // round of coalesced J loads), then add s v v^T to the rows of H they own (rows a and nv - 1 - a: equal work per warp).
  // A lane whose list is shorter rides along with s = 0.  (Walking the rows' dof bit sets per warp instead put ~10
  // dependent pairs of J loads on every warp's chain for each flip: 52 K cycles per update against 97 K for a full assembly.)
  const int nflipT = __reduce_max_sync(__activemask(), nflip);
  SP v[2] = {va, vb};
  for (int i0 = 0; i0 < nflipT; i0 += 2) {
    float sc2[2];
#pragma unroll
    for (int q = 0; q < 2; q++) {
      const bool on = i0 + q < nflip;
      const int r = on ? EFCI(E_FORCE, i0 + q) : 0;
      sc2[q] = on ? EFC(E_JV, r) * EFC(E_D, r) : 0.f;
      const int d0 = on ? EFCI(E_DESC0, r) : ROW_LIMJ, type = d0 & 3, id = d0 >> 2;
      if (type == ROW_CON) {
        const int bt = EFCI(E_DESC1, r);
        const float ct = bt >= 0 ? EFC(E_COEF, r) : 0.f;
        auto Jn = sm + L.conJ + id * nv; auto Jt = sm + L.conJ + (bt >= 0 ? bt : id) * nv;
        TFOR(k, nv) v[q][k] = fmaf(ct, Jt[k], Jn[k]);
      } else if (type == ROW_LIMJ) {
        const int kd = on ? D.limj_dof[id] : -1;
        TFOR(k, nv) v[q][k] = k == kd ? 1.f : 0.f;
      } else {   // tendon limit: the tendon's sparse row (coef^2 = 1)
        const int tn = D.limt_ten[id], n0 = D.ten_nzstart[tn], n1 = D.ten_nzstart[tn + 1];
        TFOR(k, nv) { float a = 0.f; for (int n = n0; n < n1; n++) if (D.nz_dof[n] == k) a += sm[L.tenJ + n]; v[q][k] = a; }
      }
    }
    TSYNC();
    for (int a0 = t.tl; a0 < (nv + 1) / 2; a0 += nw) {
#pragma unroll
      for (int h = 0; h < 2; h++) {
        const int a = h ? nv - 1 - a0 : a0;
        if (h && a == a0) break;
        const float wa = sc2[0] * v[0][a], wb = sc2[1] * v[1][a];
        auto Ha = H + TRI(a);
        HOT
        for (int b = 0; b <= a; b++) Ha[b] = fmaf(wb, v[1][b], fmaf(wa, v[0][b], Ha[b]));
      }
    }
    TSYNC();
  }
  TFOR(i, TRI(nv)) sm[L.Hsave + i] = H[i];
}
// Blocked left-looking Cholesky of H in place (L L^T = H), panels of CHB columns, two barriers per panel: (1) every row at
// or below the panel takes the finished columns out of its panel entries (rows over warps); (2) every warp factors the
// CHB x CHB diagonal block in registers -- redundantly, which saves the barrier a single owner would need -- and solves
// its rows below against it; the factored block itself is written back after the panel's closing barrier (see there).
// The factor's diagonal goes to Hd (the solve divides by it).  All branches are warp-uniform.
#define CHB 6
#define CHT(a, b) ((a) * ((a) + 1) / 2 + (b))
__device__ __noinline__ void choleskyTile(int nv, SP H, SP Hd, Tm t) {
  const int nw = t.nw;
  TSYNC();
  for (int j0 = 0; j0 < nv; j0 += CHB) {
    const int pb = min(CHB, nv - j0);
    if (j0 > 0) {
      for (int r = j0 + t.tl; r < nv; r += nw) {
        auto Hr = H + TRI(r);
        const int cn = min(r - j0 + 1, pb);
        float acc[CHB];
#pragma unroll
        for (int c = 0; c < CHB; c++) acc[c] = c < cn ? Hr[j0 + c] : 0.f;
        SP Hc[CHB];
#pragma unroll
        for (int c = 0; c < CHB; c++) Hc[c] = H + TRI(min(j0 + c, nv - 1));   // (past the panel: computed, not stored)
#pragma unroll 2
        for (int k = 0; k < j0; k++) {
          const float a = Hr[k];
#pragma unroll
          for (int c = 0; c < CHB; c++) acc[c] = fmaf(-a, Hc[c][k], acc[c]);
        }
#pragma unroll
        for (int c = 0; c < CHB; c++) if (c < cn) Hr[j0 + c] = acc[c];
      }
      TSYNC();
    }
    float l[CHB * (CHB + 1) / 2], inv[CHB];
#pragma unroll
    for (int a = 0; a < CHB; a++)
#pragma unroll
      for (int b = 0; b <= a; b++) l[CHT(a, b)] = a < pb ? H[TRI(j0 + a) + j0 + b] : (a == b ? 1.f : 0.f);
#pragma unroll
    for (int a = 0; a < CHB; a++)
#pragma unroll
      for (int b = 0; b <= a; b++) {
        float sacc = l[CHT(a, b)];
#pragma unroll
        for (int k = 0; k < b; k++) sacc = fmaf(-l[CHT(a, k)], l[CHT(b, k)], sacc);
        if (b < a) l[CHT(a, b)] = sacc * inv[b];
        else { const float d = sqrtf(fmaxf(sacc, MINVAL)); l[CHT(a, a)] = d; inv[a] = 1.f / d; }
      }
    for (int r = j0 + pb + t.tl; r < nv; r += nw) {
      auto Hr = H + TRI(r);
      float x[CHB];
#pragma unroll
      for (int b = 0; b < CHB; b++) {
        float sacc = b < pb ? Hr[j0 + b] : 0.f;
#pragma unroll
        for (int k = 0; k < b; k++) sacc = fmaf(-x[k], l[CHT(b, k)], sacc);
        x[b] = sacc * inv[b];
      }
#pragma unroll
      for (int b = 0; b < CHB; b++) if (b < pb) Hr[j0 + b] = x[b];
    }
    TSYNC();
    // The factored diagonal block goes back only now: every warp has read the unfactored block by the barrier above (writing
    // it before that barrier raced with the slower warps' loads -- a whole tile's solve then ran on a half-factored block),
    // and nothing reads these entries or Hd before cholSolveTile (the later panels' updates use the rows BELOW the blocks).
    if (t.tl == 0) {
#pragma unroll
      for (int a = 0; a < CHB; a++) if (a < pb) {
#pragma unroll
        for (int b = 0; b < a; b++) H[TRI(j0 + a) + j0 + b] = l[CHT(a, b)];
        Hd[j0 + a] = l[CHT(a, a)];
      }
    }
  }
  TSYNC();
}
// x <- (L L^T)^-1 x with the factor above; y is a scratch vector.  Same panels, one barrier per panel each way: every warp
// solves the panel's own unknowns redundantly, then takes them out of the rows it owns.
__device__ __noinline__ void cholSolveTile(int nv, SP H, SP Hd, SP x, SP y, Tm t) {
  const int nw = t.nw;
  for (int j0 = 0; j0 < nv; j0 += CHB) {
    const int pb = min(CHB, nv - j0);
    float yb[CHB];
#pragma unroll
    for (int b = 0; b < CHB; b++) {
      const int jb = min(j0 + b, nv - 1);
      float sacc = x[jb];
#pragma unroll
      for (int k = 0; k < b; k++) sacc = fmaf(-H[TRI(jb) + j0 + k], yb[k], sacc);
      yb[b] = b < pb ? sacc / Hd[jb] : 0.f;
    }
    if (t.tl == 0) {
#pragma unroll
      for (int b = 0; b < CHB; b++) if (b < pb) y[j0 + b] = yb[b];
    }
    for (int r = j0 + pb + t.tl; r < nv; r += nw) {
      auto Hr = H + TRI(r) + j0;
      float sacc = x[r];
#pragma unroll
      for (int b = 0; b < CHB; b++) if (b < pb) sacc = fmaf(-Hr[b], yb[b], sacc);
      x[r] = sacc;
    }
    TSYNC();
  }
  for (int j0 = ((nv - 1) / CHB) * CHB; j0 >= 0; j0 -= CHB) {
    const int pb = min(CHB, nv - j0);
    float zb[CHB];
#pragma unroll
    for (int bb = 0; bb < CHB; bb++) {
      const int b = CHB - 1 - bb;
      zb[b] = 0.f;
      if (b < pb) {
        float sacc = y[j0 + b];
#pragma unroll
        for (int k = b + 1; k < CHB; k++) if (k < pb) sacc = fmaf(-H[TRI(j0 + k) + j0 + b], zb[k], sacc);
        zb[b] = sacc / Hd[j0 + b];
      }
    }
    if (t.tl == 0) {
#pragma unroll
      for (int b = 0; b < CHB; b++) if (b < pb) x[j0 + b] = zb[b];
    }
    for (int r = t.tl; r < j0; r += nw) {
      float sacc = y[r];
#pragma unroll
      for (int b = 0; b < CHB; b++) if (b < pb) sacc = fmaf(-H[TRI(j0 + b) + r], zb[b], sacc);
      y[r] = sacc;
    }
    TSYNC();
  }
}
// K points of the 1-D cost along the search direction in one pass over the lane's rows (`tol`: the float32 noise floor of the derivative test, LS_NOISE)
#ifndef LS_NOISE
#define LS_NOISE 2e-6f
#endif
template <int K> __device__ __forceinline__ void lsEvalTile(int slot, SP sm, Tm t, int nefc, const float* a, const float* qg, float gtol, LsPt* out) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L;
  float q[4 * K];
#pragma unroll
  for (int i = 0; i < 4 * K; i++) q[i] = 0.f;
  const int neq = D.neq;
  for (int rb = t.tl; rb < nefc; rb += RB * t.nw) {
    float jarv[RB], jvv[RB], Drv[RB];
#pragma unroll
    for (int u = 0; u < RB; u++) {
      const int r = rb + u * t.nw; const bool ok = r < nefc;
      jarv[u] = ok ? EFC(E_JAR, r) : 0.f; jvv[u] = ok ? EFC(E_JV, r) : 0.f; Drv[u] = ok ? EFC(E_D, r) : 0.f;
    }
#pragma unroll
    for (int u = 0; u < RB; u++) {
      const int r = rb + u * t.nw;
      const float jar = jarv[u], jv = jvv[u], Dr = Drv[u];
      if (r < nefc) {
#pragma unroll
        for (int k = 0; k < K; k++) {
          const float x = fmaf(a[k], jv, jar);
          if (r < neq || x < 0.f) {   // equality rows come first and are always active
            const float dj = Dr * jar;
            q[4 * k] = fmaf(0.5f * dj, jar, q[4 * k]); q[4 * k + 1] = fmaf(dj, jv, q[4 * k + 1]);
            q[4 * k + 2] = fmaf(0.5f * Dr * jv, jv, q[4 * k + 2]); q[4 * k + 3] = fmaf(fabsf(dj), fabsf(jv), q[4 * k + 3]);
          }
        }
      }
    }
  }
#pragma unroll
  for (int k = 0; k < K; k++) tileSumN<4>(t, q + 4 * k);   // (K = 3 only in the bracket refinement, a few percent of the evaluations)
#pragma unroll
  for (int k = 0; k < K; k++) {
    const float q0 = q[4 * k] + qg[0], q1 = q[4 * k + 1] + qg[1], q2 = q[4 * k + 2] + qg[2], qa = q[4 * k + 3] + fabsf(qg[1]);
    LsPt p; p.alpha = a[k]; p.cost = a[k] * a[k] * q2 + a[k] * q1 + q0; p.d0 = 2.f * a[k] * q2 + q1; p.d1 = 2.f * q2;
    p.tol = fmaxf(gtol, LS_NOISE * (qa + 2.f * fabsf(a[k]) * q2));
    if (p.d1 <= 0.f) p.d1 = MINVAL;
    out[k] = p;
  }
}

struct NewtonMem { SP H, Hd, qacc, Ma, grad, search, Mv, tmp; };
// mj_solNewton for the 32 worlds of the tile in lock step; returns the lane's iteration count
// developer timing (MYO_NEWTON_TIMING): thread 0 of the tile adds the clocks since the last mark to slot k of its tile's row
#define TICK(k) do { if (dbg && threadIdx.x == 0) { const long long now_ = clock64(); dbg[blockIdx.x * 32 + (k)] += now_ - tlast; tlast = now_; } } while (0)
__device__ __noinline__ int newtonTile(int slot, SP sm, Tm t, const NewtonMem& M, unsigned lmask, int ncon, int nconT, int nefc, int* status) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L;
  long long* dbg = tP[slot].X.dbg; long long tlast = dbg ? clock64() : 0;
  const int nv = D.nv, nw = t.nw;
  SP qacc = M.qacc, Ma = M.Ma, grad = M.grad, search = M.search, Mv = M.Mv, tmp = M.tmp;
  auto smooth = sm + L.smooth; auto qas = sm + L.qacc_smooth; auto warm = sm + L.warm;
  const float scale = 1.f / (D.meaninertia * (float)max(1, nv));
  // warm start: cheaper of qacc_warmstart and qacc_smooth.  The warm start is evaluated last, so when every world of the
  // tile keeps it (the steady state of a rollout) its row residuals, M qacc and cost are already in place.
  float costS, costW = 3.4e38f, gaussW = 0.f;
  if (!(D.disableflags & DSBL_WARMSTART)) {
    mulMsparse(slot, sm + L.qM, warm, Ma, t);
    TSYNC();
    float g = 0.f;
    TFOR(k, nv) g = fmaf(0.5f * (Ma[k] - smooth[k]), warm[k] - qas[k], g);
    gaussW = TSUM(g);
    float c2[2];
    residualCost2(slot, sm, t, ncon, nefc, qas, warm, c2);   // both candidates in one sweep over J
    costS = c2[0]; costW = gaussW + c2[1];
  } else costS = residualCost(slot, sm, t, ncon, nefc, qas);
  const bool useWarm = !(costW > costS);
  TFOR(k, nv) qacc[k] = useWarm ? warm[k] : qas[k];
  TSYNC();
  float gauss = gaussW, cost = costW;
  if (__ballot_sync(lmask, !useWarm)) {   // tile-uniform: some world starts from qacc_smooth
    mulMsparse(slot, sm + L.qM, qacc, Ma, t);
    TSYNC();
    float g = 0.f;
    TFOR(k, nv) g = fmaf(0.5f * (Ma[k] - smooth[k]), qacc[k] - qas[k], g);
    gauss = TSUM(g);
    cost = gauss + residualCost(slot, sm, t, ncon, nefc, qacc);
  }
  TICK(1);
  int iter = 0, pass = 0;   // pass: Newton passes of the tile so far (tile-uniform)
  bool active = nefc > 0;
  // Rows and contacts of the worlds that are still iterating: a world that has converged drops out of every per-row and
  // per-contact loop of the later passes (their trip counts are per lane), so a pass that only serves the tile's stragglers
  // runs to THEIR row counts and loads only their lines.
  int nefcA = nefc, nconA = ncon;
  bool forceOut = nefc == 0;   // (a world without rows: the caller zeroes its qfrc_constraint)
  const int lsmax = D.ls_iterations;
  for (;;) {
    // gradient and termination tests
    constraintForce(slot, sm, t, nconA, nefcA, tmp);
    TICK(15);
    float gg[2] = {0.f, 0.f};
    TFOR(k, nv) {
      const float g = Ma[k] - smooth[k] - tmp[k]; grad[k] = g; search[k] = g; gg[0] = fmaf(g, g, gg[0]);
      const float a = fabsf(Ma[k]) + fabsf(smooth[k]) + fabsf(tmp[k]); gg[1] = fmaf(a, a, gg[1]);
    }
    tileSumN<2>(t, gg);
    const float gn = sqrtf(gg[0]), gr = sqrtf(gg[1]);
    // MuJoCo's test (scaled gradient < tolerance = 1e-8) cannot be met in float32: the gradient is a difference of
    // O(gr) terms and bottoms out near eps * gr, after which further iterations only chase rounding noise.
    if (active && iter > 0 && (scale * gn < D.tolerance || gn < GRAD_NOISE * gr)) active = false;
    if (active && iter >= D.iterations) { *status |= MYO_WORLD_SOLVER_MAXITER; active = false; }
    TICK(2);
    // A world that has just stopped iterating (here, or at the end of the last pass) has its final J^T f in tmp -- this pass's
    // force evaluation still covered it -- and its final efc_force in the rows: qfrc_constraint goes out now, and no force
    // evaluation is needed after the loop (the loop only exits here, right after one).
    if (!active && !forceOut) { TFOR(k, nv) sm[L.qfrc_con + k] = tmp[k]; forceOut = true; }
    if (!__ballot_sync(lmask, active)) break;   // the tile is done when its slowest world is
    if (!active) { nefcA = 0; nconA = 0; }
#ifdef NO_HUPDATE
    if (true) { hessianAssemble(slot, sm, t, M.H, ncon, nconT, nefc); TICK(3); } else {
#else
    if (pass == 0) { hessianAssemble(slot, sm, t, M.H, ncon, nconT, nefc); TICK(3); } else {
#endif
      hessianUpdate(slot, sm, t, M.H, nefcA, M.grad, M.Mv); TICK(4); }   // (grad, Mv: free between the gradient and the solve)
    choleskyTile(nv, M.H, M.Hd, t);
    TICK(5);
    pass++;
    if (dbg && threadIdx.x == 0) dbg[blockIdx.x * 32 + 11] += 1;
    cholSolveTile(nv, M.H, M.Hd, search, tmp, t);
    TICK(6);
    float sn = 0.f;
    TFOR(k, nv) { const float s = -search[k]; search[k] = s; sn = fmaf(s, s, sn); }
    sn = sqrtf(TSUM(sn));
    if (active && !(sn >= MINVAL)) active = false;
    // exact line search along `search`
    mulMsparse(slot, sm + L.qM, search, Mv, t);
    TICK(12);
    auto bx = sm + L.basev;
    baseMul(slot, sm, t, nconA, search, bx);
    TSYNC();
    TICK(13);
    for (int rb = t.tl; rb < nefcA; rb += RB * nw) {
      float jv[RB];
      rowMulBatch(slot, sm, rb, nw, nefcA, search, bx, jv);
#pragma unroll
      for (int u = 0; u < RB; u++) if (rb + u * nw < nefcA) EFC(E_JV, rb + u * nw) = jv[u];
    }
    float g12[2] = {0.f, 0.f};
    TFOR(k, nv) { g12[0] = fmaf(search[k], Ma[k] - smooth[k], g12[0]); g12[1] = fmaf(0.5f * search[k], Mv[k], g12[1]); }
    tileSumN<2>(t, g12);   // its barriers also publish E_JV
    TICK(7);
    const float qg[3] = {gauss, g12[0], g12[1]};
    const float gtol = D.tolerance * D.ls_tolerance * sn / scale;
    int lsiter = 0;
    LsPt p0, p1;
    { const float a0 = 0.f; lsEvalTile<1>(slot, sm, t, nefcA, &a0, qg, gtol, &p0); lsiter++; }
    { const float a1 = p0.alpha - p0.d0 / p0.d1; lsEvalTile<1>(slot, sm, t, nefcA, &a1, qg, gtol, &p1); lsiter++; }
    if (p0.cost < p1.cost) p1 = p0;
    bool done = !active || fabsf(p1.d0) < p1.tol;
    const float dir = p1.d0 < 0.f ? 1.f : -1.f;
    LsPt p2 = p1; bool have2 = false;
    for (;;) {   // Newton steps on the 1-D cost until the derivative changes sign
      const bool want = !done && p1.d0 * dir <= -p1.tol && lsiter < lsmax;
      if (!__ballot_sync(lmask, want)) break;
      const float a = want ? p1.alpha - p1.d0 / p1.d1 : 0.f;
      LsPt pn; lsEvalTile<1>(slot, sm, t, nefcA, &a, qg, gtol, &pn);
      if (want) { p2 = p1; have2 = true; p1 = pn; lsiter++; if (fabsf(p1.d0) < p1.tol) done = true; }
    }
    bool refine = !done && lsiter < lsmax && have2;
    LsPt lo = p2, hi = p1;
    for (;;) {   // bracket refinement: Newton steps from both ends and the midpoint, three points per pass
      const bool want = refine && lsiter < lsmax;
      if (!__ballot_sync(lmask, want)) break;
      const float ca[3] = {want ? lo.alpha - lo.d0 / lo.d1 : 0.f, want ? hi.alpha - hi.d0 / hi.d1 : 0.f, want ? 0.5f * (lo.alpha + hi.alpha) : 0.f};
      LsPt cand[3];
      lsEvalTile<3>(slot, sm, t, nefcA, ca, qg, gtol, cand);
      if (want) {
        lsiter += 3;
        bool moved = false, hit = false;
        LsPt best = lo.cost < hi.cost ? lo : hi;
#pragma unroll
        for (int q = 0; q < 3; q++) {
          if (hit) continue;
          const LsPt c = cand[q];
          if (fabsf(c.d0) < c.tol) { p1 = c; hit = true; continue; }
          if (!((c.alpha - lo.alpha) * dir > 0.f && (hi.alpha - c.alpha) * dir > 0.f)) continue;
          if (c.d0 * dir < 0.f) lo = c; else hi = c;
          moved = true;
          if (c.cost < best.cost) best = c;
        }
        if (hit) { done = true; refine = false; }
        else if (!moved) { p1 = best; refine = false; }
        else p1 = lo.cost < hi.cost ? lo : hi;
      }
    }
    TICK(8);
    const float alpha = p1.alpha;
    if (active && alpha == 0.f) active = false;
    if (active) TFOR(k, nv) { qacc[k] = fmaf(alpha, search[k], qacc[k]); Ma[k] = fmaf(alpha, Mv[k], Ma[k]); }
    float gc[2] = {0.f, 0.f};
    for (int rb = t.tl; rb < nefcA; rb += RB * nw) {
      float jarv[RB], jvv[RB], Drv[RB]; int d0v[RB];
#pragma unroll
      for (int u = 0; u < RB; u++) {
        const int r = rb + u * nw; const bool ok = r < nefcA;
        jarv[u] = ok ? EFC(E_JAR, r) : 0.f; jvv[u] = ok ? EFC(E_JV, r) : 0.f; Drv[u] = ok ? EFC(E_D, r) : 0.f; d0v[u] = ok ? EFCI(E_DESC0, r) : ROW_CON;
      }
#pragma unroll
      for (int u = 0; u < RB; u++) {
        const int r = rb + u * nw;
        if (r < nefcA) {
          float jar = jarv[u], flip = 0.f;
          if (active) {
            const float old = jar;
            jar = fmaf(alpha, jvv[u], jar); EFC(E_JAR, r) = jar;
            flip = (jar < 0.f ? 1.f : 0.f) - (old < 0.f ? 1.f : 0.f);   // the row joins (+1) or leaves (-1) the active set: hessianUpdate
          }
          EFC(E_JV, r) = flip;
          const bool on = rowOn(d0v[u], jar);
          EFC(E_FORCE, r) = on ? -Drv[u] * jar : 0.f;   // (the flip list of hessianUpdate borrowed these entries; the next force evaluation reads them)
          if (on) gc[1] = fmaf(0.5f * Drv[u] * jar, jar, gc[1]);
        }
      }
    }
    TSYNC();
    TFOR(k, nv) gc[0] = fmaf(0.5f * (Ma[k] - smooth[k]), qacc[k] - qas[k], gc[0]);
    tileSumN<2>(t, gc);
    if (active) {
      gauss = gc[0];
      const float newcost = gauss + gc[1];
      iter++;
      const float improvement = scale * (cost - newcost);
      cost = newcost;
      if (improvement < D.tolerance) active = false;
    }
    TICK(9);
  }
  TICK(10);
  return iter;
}

// Shared memory of a tile: Hessian (lower triangle), its factor's diagonal and six nv-vectors, [element][32 lanes].  When a
// model's block does not fit (nv beyond ~55), the same kernel runs with everything in the world's scratch column.
// ------------------------------------------------------------------------------------------ finish: mj_Euler, state out
// mj_Euler with implicit joint damping (the factor of M + h B comes from inertiaKernel), quaternion integration on the
// manifold; then the state and the per-dof results of the solve go out as full lines.
// (a device function: the solver kernel runs it as its own tail -- each tile goes on to integrate as soon as ITS solve is
// done, instead of the whole batch meeting at a kernel boundary first; finishKernel remains for the paths without a solve)
// `stage`: a shared-memory block of >= nM + 2 nv elements per lane (the solver kernel's, free after the solve), or null
__device__ __noinline__ void finishTile(int slot, int fslot, SP sm, Tm t, World W, int flags, SP stage) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L; const FieldPtrs& F = tP[slot].F;
  const int nq = D.nq, nv = D.nv, na = D.na, nu = D.nu;
  if (flags & FIN_ACC) {   // results of the solve (read-only here; mj_Euler below does not change them)
    const FieldPtrs& Fo = tP[fslot].F;
    void* o;
    if ((o = Fo.p[MYO_F_QACC])) TFOR(i, nv) OUTF(o, i) = sm[L.qacc + i];
    if ((o = Fo.p[MYO_F_QFRC_CONSTRAINT])) TFOR(i, nv) OUTF(o, i) = sm[L.qfrc_con + i];
    if ((o = Fo.p[MYO_F_QFRC_INVERSE])) TFOR(i, nv) OUTF(o, i) = 0.f;
    if ((o = Fo.p[MYO_F_EFC_FORCE])) {
      const int nefc = __float_as_int(sm[L.counts + 1]);
      const int rtop = (flags & FIN_LIVE) ? nefc : D.njmax;
      for (int r0 = t.tl; r0 < rtop; r0 += 8 * t.nw) {   // (eight loads in flight)
        float v[8];
#pragma unroll
        for (int u = 0; u < 8; u++) v[u] = r0 + u * t.nw < nefc ? EFC(E_FORCE, r0 + u * t.nw) : 0.f;
#pragma unroll
        for (int u = 0; u < 8; u++) if (r0 + u * t.nw < rtop) OUTF(o, r0 + u * t.nw) = v[u];
      }
    }
    if ((o = Fo.p[MYO_F_SOLVER_NITER]) && t.tl == 0) OUTI(o, 0) = __float_as_int(sm[L.counts + 2]);
  }
  if (flags & FIN_EULER) {
    const bool skip = __float_as_int(sm[L.wstate + WS_SKIP]) != 0;   // mj_checkAcc reset this world: nothing to integrate
    const float h = D.timestep;
    auto qa = sm + L.tmpv;
    if (D.any_damping && stage.p) {
      // the level-by-level solve is a chain of dependent loads with a barrier per level: run it on a shared-memory copy of
      // the factor (measured: 5 % of the solver kernel's time when it reads the factor from the scratch)
      SP LD = stage, dinv = stage + D.nM; qa = dinv + nv;
      const LdlTab T = ldlTabStage(D, (int*)(stage.p - (threadIdx.x & 31) + (D.nM + 2 * nv) * SCR), t);   // (the tables follow the lanes' block)
      for (int e0 = t.tl; e0 < D.nM; e0 += 8 * t.nw) {   // (eight loads in flight)
        float v[8];
#pragma unroll
        for (int u = 0; u < 8; u++) v[u] = e0 + u * t.nw < D.nM ? sm[L.qH + e0 + u * t.nw] : 0.f;
#pragma unroll
        for (int u = 0; u < 8; u++) if (e0 + u * t.nw < D.nM) LD[e0 + u * t.nw] = v[u];
      }
      TFOR(i, nv) { dinv[i] = sm[L.qHinv + i]; qa[i] = sm[L.smooth + i] + sm[L.qfrc_con + i]; }
      TSYNC();
      solveLDL(slot, LD, dinv, qa, t, T);
    } else if (D.any_damping) {
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
__global__ void __launch_bounds__(32 * MAXW, MINB_FINISH) finishKernel(int slot, int fslot, int nworld, const int* __restrict__ mask, int flags,
                                                    float* __restrict__ scratch) {
  Timeline tline(slot, MYO_K_FINISH);
  Ctx c = tileCtx(slot, nworld, mask, scratch);
  if (!c.live) return;
  SP none; none.p = nullptr;
  finishTile(slot, fslot, c.sm, c.t, c.W, flags, none);
}

// mj_makeConstraint as its own launch: it needs the contacts and the velocity, not the smooth forces, so it runs on the
// collision stream beside the tendon / actuation stages (models with tendon limits: after them, api.cu).
#ifndef MINB_MAKECON
#define MINB_MAKECON 2   // (80 registers: the batched Jacobian loads spill at 56; -0.5 % step time)
#endif
__global__ void __launch_bounds__(32 * MAXW, MINB_MAKECON) makeconKernel(int slot, int nworld, const int* __restrict__ mask, float* __restrict__ scratch) {
  const SmemLayout& L = tP[slot].L;
  Timeline tline(slot, MYO_K_MAKECON);
  Ctx c = tileCtx(slot, nworld, mask, scratch);
  const unsigned lmask = __ballot_sync(0xffffffffu, c.live);
  if (!c.live) return;
  SP sm = c.sm; Tm t = c.t;
  int status = __float_as_int(sm[L.wstate + WS_STATUS]);
  const int ncon = __float_as_int(sm[L.counts]);
  const int nconT = __reduce_max_sync(lmask, ncon);
  const int nefc = makeConstraintTile(slot, sm, t, ncon, nconT, &status);
  if (t.tl == 0) { sm[L.counts + 1] = __int_as_float(nefc); sm[L.wstate + WS_STATUS] = __int_as_float(status); }
}
// collideCompact + makecon in one launch (models without tendon limits): both work on the tile's own records, and the kernel
// boundary between them cost the solver's critical path ~40 us of waiting for SM slots.
__global__ void __launch_bounds__(32 * MAXW, MINB_MAKECON) compactMakeconKernel(int slot, int nworld, const int* __restrict__ mask, float* __restrict__ scratch) {
  const SmemLayout& L = tP[slot].L;
  Timeline tline(slot, MYO_K_MAKECON);
  Ctx c = tileCtx(slot, nworld, mask, scratch);
  const unsigned lmask = __ballot_sync(0xffffffffu, c.live);
  if (!c.live) return;
  SP sm = c.sm; Tm t = c.t;
  compactTile(slot, sm, t, lmask);
  TSYNC();
  int status = __float_as_int(sm[L.wstate + WS_STATUS]);
  const int ncon = __float_as_int(sm[L.counts]);
  const int nconT = __reduce_max_sync(lmask, ncon);
  const int nefc = makeConstraintTile(slot, sm, t, ncon, nconT, &status);
  if (t.tl == 0) { sm[L.counts + 1] = __int_as_float(nefc); sm[L.wstate + WS_STATUS] = __int_as_float(status); }
}
static inline size_t newtonSmemFloats(int nv) { return (size_t)(TRI(nv) + 7 * nv) * 32; }   // (+ the LDL solve's index tables, myo_tile_launch)
#define NEWTON_SMEM_MAX (190 * 1024)   // + ~9 KB of static reduction / scan buffers <= 227 KB per CTA
__global__ void __launch_bounds__(32 * MAXW, MINB_NEWTON) newtonKernel(int slot, int fslot, int nworld, const int* __restrict__ mask, int check_acc,
                                                    int use_smem, int fin, float* __restrict__ scratch) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L;
  extern __shared__ float tileShared[];
  __shared__ int sBad[32];
  const int lane = threadIdx.x & 31;
  Timeline tline(slot, MYO_K_NEWTON);
  Ctx c = tileCtx(slot, nworld, mask, scratch);
  const unsigned lmask = __ballot_sync(0xffffffffu, c.live);
  if (!c.live) return;
  SP sm = c.sm; Tm t = c.t;
  const int nv = D.nv;
  int status = __float_as_int(sm[L.wstate + WS_STATUS]);
  const int ncon = __float_as_int(sm[L.counts]);
  const int nconT = __reduce_max_sync(lmask, ncon);
  if (t.tl == 0) sBad[lane] = 0;
  long long* dbg = tP[slot].X.dbg; long long tlast = dbg ? clock64() : 0;
  const int nefc = __float_as_int(sm[L.counts + 1]);   // rows: makeconKernel
  const int nefcT = __reduce_max_sync(lmask, nefc);
  int niter = 0;
  if (nefcT == 0) {   // tile-uniform
    TFOR(i, nv) { const float a = sm[L.qacc_smooth + i]; sm[L.qacc + i] = a; sm[L.warm + i] = a; sm[L.qfrc_con + i] = 0.f; }
  } else {
    NewtonMem M;
    if (use_smem) {
      float* p = tileShared + lane;
      auto take = [&](int n) { SP r; r.p = p; p += n * SCR; return r; };
      M.H = take(TRI(nv)); M.Hd = take(nv); M.qacc = take(nv); M.Ma = take(nv); M.grad = take(nv); M.search = take(nv); M.Mv = take(nv); M.tmp = take(nv);
    } else {
      M.H = sm + L.H; M.Hd = sm + L.H + TRI(nv); M.qacc = sm + L.nvec; M.Ma = M.qacc + nv; M.grad = M.Ma + nv; M.search = M.grad + nv;
      M.Mv = M.search + nv; M.tmp = M.Mv + nv;
    }
    niter = newtonTile(slot, sm, t, M, lmask, ncon, nconT, nefc, &status);
    // a world without rows in a tile that has some: its warm start and its only candidate coincide, qacc = qacc_smooth
    TFOR(i, nv) { const float a = nefc > 0 ? M.qacc[i] : sm[L.qacc_smooth + i]; sm[L.qacc + i] = a; sm[L.warm + i] = a; if (nefc == 0) sm[L.qfrc_con + i] = 0.f; }
  }
  TSYNC();
  // mj_checkAcc: a world whose qacc went bad is reset (qpos0, zero velocity / activation) and flagged; finishKernel does
  // not integrate it, its next forward pass starts from the reset state.
  int skip = 0;
  if (check_acc) {
    bool bad = false;
    TFOR(i, nv) bad |= !(fabsf(sm[L.qacc + i]) < 1e10f);
    if (bad) sBad[lane] = 1;
    TSYNC();
    if (sBad[lane]) {
      TFOR(i, D.nq) sm[L.qpos + i] = D.qpos0[i];
      TFOR(i, nv) { sm[L.qvel + i] = 0.f; sm[L.warm + i] = 0.f; }
      TFOR(i, D.na) sm[L.act + i] = 0.f;
      status |= MYO_WORLD_BADQACC; skip = 1;
    }
  }
  if (fslot != MYO_NSLOT && D.nsensor > 0) touchSensors(slot, fslot, sm, t, c.W, nefc > 0 ? ncon : 0);
  if (t.tl == 0) {
    sm[L.counts + 2] = __int_as_float(niter);
    sm[L.wstate + WS_STATUS] = __int_as_float(status); sm[L.wstate + WS_SKIP] = __int_as_float(skip);
  }
  if (fin >= 0) {   // mj_Euler and the state / result outputs of this tile (finishTile)
    TSYNC();
    SP stage; stage.p = use_smem ? tileShared + lane : nullptr;   // (TRI(nv) + 7 nv elements per lane >= nM + 2 nv)
    finishTile(slot, fslot, sm, t, c.W, fin, stage);
  }
}

// ------------------------------------------------------------------------------------------ conout: contact / constraint records
// contact_* and efc_* (padded to nconmax / njmax, efc_J expanded to dense rows): pure stores, no ordering between items, so
// the items are striped over the warps of gridDim.y CTAs per tile.
__global__ void __launch_bounds__(256) conoutKernel(int slot, int fslot, int nworld, const int* __restrict__ mask, int live_only, float* __restrict__ scratch) {
  const SmemLayout& L = tP[slot].L;
  Timeline tline(slot, MYO_K_CONOUT);
  Ctx c = tileCtx(slot, nworld, mask, scratch);
  if (!c.live) return;
  Tm t; t.nw = c.t.nw * gridDim.y; t.tl = blockIdx.y * c.t.nw + c.t.tl;
  SP sm = c.sm;
  writeConstraintOutputs(slot, fslot, sm, t, c.W, __float_as_int(sm[L.counts]), __float_as_int(sm[L.counts + 1]), live_only != 0);
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
int myo_tile_upload(int slot, const DevModel& D, const SmemLayout& L, const FieldPtrs& F, const TileAux& X, void* stream) {
  static TParams staging[MYO_NSLOT];   // must outlive the async copy
  if (slot < 0 || slot >= MYO_NSLOT) return -1;
  TParams& P = staging[slot];
  P.D = D; P.L = L; P.F = F; P.X = X;
  return cudaMemcpyToSymbolAsync(tP, &P, sizeof(TParams), (size_t)slot * sizeof(TParams), cudaMemcpyHostToDevice, (cudaStream_t)stream) == cudaSuccess ? 0 : -1;
}
// An empty launch: api.cu puts one in front of the tendon and inertia kernels so that the collision kernel, which becomes
// ready at the same instant (end of kin) and heads the longest branch, gets its CTAs onto the SMs first.
__global__ void nopKernel() {}
int myo_nop_launch(void* stream) { nopKernel<<<1, 32, 0, (cudaStream_t)stream>>>(); return cudaGetLastError() == cudaSuccess ? 0 : -1; }
static int envInt(const char* name, int dflt) { const char* e = getenv(name); return e ? atoi(e) : dflt; }
// Warps per tile of each kernel (developer knobs MYO_NW_*): enough warps per SM to hide the scratch latency, few enough
// that the item loops still have work for every warp.
int myo_tile_launch(int kernel, int slot, int fslot, int nworld, const int* mask, int flags, float* scratch, void* stream) {
  // Measured on 8192 MyoHand worlds (profiles/r2_warps_per_tile.txt): every kernel is bound by the dependent-instruction
  // latency of its warps, and a tile's warps split its items, so more warps per tile shorten the chains -- until the
  // register file (two tiles per SM must stay resident: 256 tiles on 148 SMs) forces spills.  12 is the sweet spot for the
  // solver (8: +6.6 % step time, 10: +3.9 %, 16 at <= 64 registers: +1.8 %).  With the 56-register caps of the phase-2 kernels
  // (three CTAs per SM) 12 warps is also best for tendon, inertia and the output kernels (8: +2 % step time, +4 % in the
  // contact-poor first steps).
  static const int nwKin = envInt("MYO_NW_KIN", 12), nwTen = envInt("MYO_NW_TEN", 12), nwIne = envInt("MYO_NW_INE", 12),
                   nwCol = envInt("MYO_NW_COL", 12), nwVel = envInt("MYO_NW_VEL", 12), nwFin = envInt("MYO_NW_FIN", 8),
                   nwOut = std::min(8, envInt("MYO_NW_OUT", 8)), yOut = envInt("MYO_Y_OUT", 4), nwNew = envInt("MYO_NW_NEWTON", 12),
                   nwMak = envInt("MYO_NW_MAKECON", 12), nwPos = envInt("MYO_NW_POSOUT", 12), outGrid = std::max(1, envInt("MYO_OUT_GRID", 1 << 30));   // (limited grids measured: 148 CTAs +3 % step time, 74 +4 %, 48 +15 % -- the output kernels need the whole machine to finish inside the step)
  const int tiles = (nworld + 31) / 32;
  cudaStream_t st = (cudaStream_t)stream;
  switch (kernel) {
    case MYO_K_KIN: kinKernel<<<tiles, 32 * nwKin, 0, st>>>(slot, fslot, nworld, mask, flags, scratch); break;
    case MYO_K_TENDON: tendonKernel<<<tiles, 32 * nwTen, 0, st>>>(slot, fslot, nworld, mask, scratch); break;
    case MYO_K_INERTIA: {   // flags: euler | nM << 8 | ints of the factorisation's index tables << 20; shared memory: two copies of
      // the factor's pattern per lane + the tables (<= 48 KB, else the scratch path)
      const size_t sh = (size_t)2 * ((flags >> 8) & 0xfff) * 32 * sizeof(float) + (size_t)(flags >> 20) * sizeof(int);
      const int staged = sh > 0 && sh <= 48 * 1024;
      inertiaKernel<<<tiles, 32 * nwIne, staged ? sh : 0, st>>>(slot, fslot, nworld, mask, flags & 1, staged, scratch);
    } break;
    case MYO_K_COLLIDE: collideKernel<<<tiles, 32 * nwCol, 0, st>>>(slot, nworld, mask, scratch); break;
    case MYO_K_COLLIDE_MPR: collideMprKernel<<<148 * 4, 128, 0, st>>>(slot, scratch); break;
    case MYO_K_COLLIDE_COMPACT: collideCompactKernel<<<tiles, 32 * nwCol, 0, st>>>(slot, nworld, mask, scratch); break;
    case MYO_K_VELACT: {   // flags: posvel | (nM + 2 nv) << 8 | ints of the LDL index tables << 20
      const size_t sh = (size_t)((flags >> 8) & 0xfff) * 32 * sizeof(float) + (size_t)(flags >> 20) * sizeof(int);
      const int staged = sh > 0 && sh <= 48 * 1024;
      velactKernel<<<tiles, 32 * nwVel, staged ? sh : 0, st>>>(slot, fslot, nworld, mask, flags & 1, staged, scratch);
    } break;
    case MYO_K_FINISH: finishKernel<<<tiles, 32 * nwFin, 0, st>>>(slot, fslot, nworld, mask, flags, scratch); break;
    case MYO_K_NEWTON: {
      const int nv = (flags >> 8) & 0xff;   // the host passes nv and the size of the LDL index tables in the upper bits (the model lives in constant memory)
      size_t shmem = newtonSmemFloats(nv) * sizeof(float) + (size_t)(flags >> 16) * sizeof(int);
      const int use_smem = shmem <= NEWTON_SMEM_MAX ? 1 : 0;
      if (!use_smem) shmem = 0;
      static size_t configured[64] = {0};   // per device: largest dynamic shared-memory size opted in to so far
      int dev = 0;
      if (shmem > 48 * 1024) {
        if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return -1;
        if (configured[dev] < shmem) {
          if (cudaFuncSetAttribute(newtonKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)shmem) != cudaSuccess) return -1;
          configured[dev] = shmem;
        }
      }
      newtonKernel<<<tiles, 32 * nwNew, shmem, st>>>(slot, fslot, nworld, mask, flags & 1, use_smem, (flags & 32) ? (((flags >> 1) & 15) | ((flags & 64) ? 16 : 0)) : -1, scratch);
    } break;
    case MYO_K_TENOUT: tenoutKernel<<<mask ? tiles : std::min(tiles, outGrid), 32 * nwPos, 0, st>>>(slot, fslot, nworld, mask, scratch); break;
    case MYO_K_POSOUT: posoutKernel<<<mask ? tiles : std::min(tiles, outGrid), 32 * nwPos, 0, st>>>(slot, fslot, nworld, mask, scratch); break;
    case MYO_K_COMPACT_MAKECON: compactMakeconKernel<<<tiles, 32 * nwMak, 0, st>>>(slot, nworld, mask, scratch); break;
    case MYO_K_MAKECON: makeconKernel<<<tiles, 32 * nwMak, 0, st>>>(slot, nworld, mask, scratch); break;
    case MYO_K_CONOUT: conoutKernel<<<dim3(tiles, yOut), 32 * nwOut, 0, st>>>(slot, fslot, nworld, mask, flags & 1, scratch); break;
    default: return -1;
  }
  return cudaPeekAtLastError() == cudaSuccess ? 0 : -1;
}
int myo_launch_quat_kat(int n, const float* qa, const float* qb, float* prod, float* mat, float* rotated, void* stream) {
  quatKatKernel<<<(n + 127) / 128, 128, 0, (cudaStream_t)stream>>>(n, qa, qb, prod, mat, rotated);
  return cudaPeekAtLastError() == cudaSuccess ? 0 : -1;
}

// ---------------------------------------------------------------------------------------------- allocation-free entry point
// myo_kinematics_noalloc (include/myostep.h): mj_kinematics + mj_comPos and the position-stage outputs for callers that
// cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  No counterpart in the
// reference.  The kernel is the kin + posout pair of the pipeline fused into
// one launch; its per-world scratch (the front of the pipeline's layout: state, frames, anchors, axes, tree centres of mass,
// body inertias, cdof, + the body-level records of the output stage) lives in a file-scope device array of fixed size
// instead of a batch's cudaMalloc'ed buffer.  Every GPU that loads the library holds the array until exit; all calls share
// it, so calls that overlap on the GPU race (see the header).
#define MYO_NOALLOC_SCRATCH_BYTES (8u << 20)
namespace {
static __device__ __align__(256) float g_noallocScratch[MYO_NOALLOC_SCRATCH_BYTES / sizeof(float)];
__global__ void __launch_bounds__(32 * MAXW, 2) kinNoallocKernel(int slot, int nworld, const float* __restrict__ qpos, float* __restrict__ scratch) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L;
  Ctx c = tileCtx(slot, nworld, nullptr, scratch);
  if (!c.live) return;
  SP sm = c.sm; Tm t = c.t; World W = c.W;
  TFOR(i, D.nq) sm[L.qpos + i] = OUTF(qpos, i);
  TSYNC();
  kinematics(slot, sm, t);
  comPos(slot, sm, t);
  writePositionOutputs(slot, slot, sm, t, W);
}
std::mutex g_noallocMutex;          // one parameter slot and one scratch array: host-side calls are serialised
float* g_noallocPtr[64] = {nullptr};  // device address of g_noallocScratch, per device (the symbol's name is only a host-side stand-in)
TParams g_noallocParams;            // staging of the parameter upload (outlives the asynchronous copy)
}  // namespace

int myo_kinematics_noalloc(const myo_model* m, int nworld, const float* qpos, float* xpos, float* xquat, float* xmat,
                           float* site_xpos, float* geom_xpos, void* stream) {
  if (!m || !qpos || nworld <= 0) return myo_fail(MYO_ERR_ARG, "myo_kinematics_noalloc: null model / qpos or nworld <= 0");
  if (!xpos && !xquat && !xmat && !site_xpos && !geom_xpos) return myo_fail(MYO_ERR_ARG, "myo_kinematics_noalloc: no output buffer given");
  if (m->device < 0 || m->device >= 64) return myo_fail(MYO_ERR_ARG, "myo_kinematics_noalloc: device index out of range");
  if (cudaSetDevice(m->device) != cudaSuccess) return myo_fail(MYO_ERR_CUDA, "cudaSetDevice failed");
  std::lock_guard<std::mutex> lk(g_noallocMutex);
  // the pipeline's layout, cut after the regions this stage touches (everything before cdof_dot), + the output stage's records
  TParams& P = g_noallocParams;
  memset(&P, 0, sizeof(P));
  P.D = m->D;
  try { myo_make_layout(P.D, 0, 0, P.L); } catch (const std::string& e) { return myo_fail(MYO_ERR_ARG, e); }
  P.L.tmpbody = P.L.cdof_dot;
  P.L.total = P.L.tmpbody + ((13 * P.D.nbody + 3) & ~3);
  const size_t tiles = ((size_t)nworld + 31) / 32, tileBytes = (size_t)P.L.total * 32 * sizeof(float);
  if (tiles * tileBytes > MYO_NOALLOC_SCRATCH_BYTES)
    return myo_fail(MYO_ERR_ARG, "myo_kinematics_noalloc: nworld " + std::to_string(nworld) + " exceeds the fixed scratch (8 MiB = " +
                                     std::to_string(MYO_NOALLOC_SCRATCH_BYTES / tileBytes * 32) + " worlds of this model); use a batch (myo_batch_create + myo_forward)");
  float*& ptr = g_noallocPtr[m->device];
  if (!ptr) {   // once per device, with that device current
    void* p = nullptr;
    if (cudaGetSymbolAddress(&p, g_noallocScratch) != cudaSuccess || !p) return myo_fail(MYO_ERR_CUDA, std::string("cudaGetSymbolAddress failed: ") + cudaGetErrorString(cudaGetLastError()));
    ptr = (float*)p;
  }
  P.F.p[MYO_F_XPOS] = xpos; P.F.p[MYO_F_XQUAT] = xquat; P.F.p[MYO_F_XMAT] = xmat; P.F.p[MYO_F_SITE_XPOS] = site_xpos; P.F.p[MYO_F_GEOM_XPOS] = geom_xpos;
  cudaStream_t st = (cudaStream_t)stream;
  // the array keeps the last call's contents: the part this call uses starts from zeros, like a batch's scratch
  if (cudaMemsetAsync(ptr, 0, tiles * tileBytes, st) != cudaSuccess) return myo_fail(MYO_ERR_CUDA, "cudaMemsetAsync failed");
  if (cudaMemcpyToSymbolAsync(tP, &P, sizeof(TParams), (size_t)MYO_NOALLOC_SLOT * sizeof(TParams), cudaMemcpyHostToDevice, st) != cudaSuccess)
    return myo_fail(MYO_ERR_CUDA, "parameter upload failed");
  myo_count_launch(1);
  kinNoallocKernel<<<(unsigned)tiles, 32 * 8, 0, st>>>(MYO_NOALLOC_SLOT, nworld, qpos, ptr);
  if (cudaPeekAtLastError() != cudaSuccess) return myo_fail(MYO_ERR_CUDA, std::string("kernel launch failed: ") + cudaGetErrorString(cudaGetLastError()));
  return MYO_OK;
}
