// TEAM-mode kernels of the batched mj_step for sm_100a: a team of TEAM lanes owns one world (32 / TEAM worlds per warp).
//
//  * solveKernel: the DATA-DEPENDENT part of a step -- constraint rows (mj_makeConstraint), the Newton solve
//    (mj_fwdConstraint) and the contact / constraint outputs -- for the phase pipeline (tile.cu runs the regular phases
//    with lane = world; api.cu strings the kernels together).  Worlds of a warp diverge freely on contacts and solver
//    iterations; nothing here waits for another world.
//  * stepKernel: the whole step fused in one launch (round-1 design, kept as the reference build: MYO_ENGINE=mono), the
//    CTA walking the phases together.
// Both run on the world-fastest [k][NW] scratch shared with the TILE-mode kernels (common.cuh).
//
// Replaces (third-party code the reference calls; call sites in include/myostep.h): mj_collision, mj_makeConstraint,
// mj_fwdConstraint (Newton), mj_sensorAcc (touch), mj_checkAcc; with common.cuh the whole of mj_step.
#include "common.cuh"

namespace {
// Dense Newton Hessian (nv x nv) of this lane's world.  It is the most re-read block of the solver (assembly, Cholesky,
// two triangular solves per iteration), so it lives in shared memory -- interleaved over the HS worlds of a warp, one
// [TRI(nv)][HS] tile per warp -- whenever the CTA's tiles fit (hshared, set at upload), else in the scratch.
// Only the lower triangle is stored, row-packed: element (r, c <= r) at TRI(r) + c.
#define HS (32 / TEAM)
extern __shared__ float ctaShared[];
__device__ __forceinline__ SP hessian(SP sm, int offH, int nv, int hshared) {
  if (!hshared) return sm + offH;
  SP r; r.p = ctaShared + (threadIdx.x >> 5) * (TRI(nv) * HS) + (threadIdx.x & 31) / TEAM; r.s = HS;
  return r;
}
__device__ __forceinline__ int teamMinI(Tm t, int v) {
#pragma unroll
  for (int o = TEAM / 2; o > 0; o >>= 1) v = min(v, __shfl_xor_sync(t.mask, v, o));
  return v;
}
__device__ __forceinline__ bool teamAny(Tm t, bool p) {
#if TEAM == 1
  return p;
#else
  return (__ballot_sync(t.mask, p) & t.mask) != 0u;
#endif
}
// exclusive prefix sum over the team's lanes (lane order), *total = team sum
__device__ __forceinline__ int teamExclScan(Tm t, int v, int* total) {
  int incl = v;
#pragma unroll
  for (int o = 1; o < TEAM; o <<= 1) { const int u = __shfl_up_sync(t.mask, incl, o, TEAM); if (t.tl >= o) incl += u; }
  *total = __shfl_sync(t.mask, incl, TEAM - 1, TEAM);
  return incl - v;
}
__device__ __forceinline__ float teamBcast(Tm t, float v, int src) { return __shfl_sync(t.mask, v, src, TEAM); }

// returns the number of contacts; contact record: dist, pos[3], normal[3], pair, efc_address, dim, spare[2]
// Pass 1: bounding-sphere / plane-distance filter over the static candidate-pair list, TEAM pairs at a time, survivors
// compacted in pair order by a team prefix sum.  Pass 2: narrow phase over the survivors, TEAM at a time, hits
// compacted in the same order -- so contact order (hence geom1/geom2/efc_address) is MuJoCo's.
__device__ __noinline__ int collision(int slot, SP sm, Tm t, int* status) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L;
  if (D.npair == 0) return 0;
  TFOR(c, D.ncg) {
    const int f = D.cg_frame[c];
    auto o = sm + L.cgpose + 12 * c;
    st3(o, ld3(FR_POS(sm, L, f)) + mulM(FR_MAT(sm, L, f), ld3(D.cg_lpos + 3 * c)));
    float M[9]; mulMM(M, FR_MAT(sm, L, f), D.cg_lmat + 9 * c);
#pragma unroll
    for (int k = 0; k < 9; k++) o[3 + k] = M[k];
  }
  TSYNC();
  SPI cand = asInt(sm + L.cand);
  int ncand = 0;
  for (int base = 0; base < D.npair; base += TEAM) {
    const int pr = base + t.tl;
    bool pass = false;
    if (pr < D.npair) {
      const int c1 = D.pair_cg1[pr], c2 = D.pair_cg2[pr];
      auto P1 = sm + L.cgpose + 12 * c1; auto P2 = sm + L.cgpose + 12 * c2;
      const f3 d = ld3(P2) - ld3(P1);
      const float margin = D.pair_prm[16 * pr];
      if (D.cg_type[c1] == G_PLANE) pass = !(d.x * P1[5] + d.y * P1[8] + d.z * P1[11] > D.cg_rbound[c2] + margin);
      else { const float r = D.cg_rbound[c1] + D.cg_rbound[c2] + margin; pass = !(dot(d, d) > r * r); }
    }
    int total; const int at = ncand + teamExclScan(t, pass ? 1 : 0, &total);
    if (pass) cand[at] = pr;
    ncand += total;
  }
  TSYNC();
  // Narrow phase in three sweeps per chunk of survivors, so that the lanes running the (long, branchy, float64) MPR do it
  // at the same time instead of idling the lanes whose pair had a closed form: (a) closed-form pairs, MPR pairs only
  // listed; (b) the listed MPR pairs, TEAM at a time; (c) in-order compaction of all hits.  Per-survivor results live in
  // the solver's scratch (unused until makeConstraint): 16 floats = count, two hits, one list slot.
  auto tmp = sm + L.conJ;
  const int cap = max(TEAM, (L.limrow - L.conJ) / 16);
  int ncon = 0;
  for (int c0 = 0; c0 < ncand; c0 += cap) {
    const int cn = min(cap, ncand - c0);
    int nm = 0;
    for (int base = 0; base < cn; base += TEAM) {
      const int k = base + t.tl;
      bool mp = false;
      if (k < cn) {
        const int pr = cand[c0 + k];
        const int t1 = D.cg_type[D.pair_cg1[pr]], t2 = D.cg_type[D.pair_cg2[pr]];
        mp = t1 != G_PLANE && !((t1 == G_SPHERE || t1 == G_CAPSULE) && (t2 == G_SPHERE || t2 == G_CAPSULE) && !(t1 == G_CAPSULE && t2 == G_SPHERE));
        if (!mp) {
          Hit h[2];
          const int n = collidePair(slot, sm, pr, h);
          auto o = tmp + 16 * k;
          o[0] = __int_as_float(n);
          for (int q = 0; q < n; q++) { o[1 + 7 * q] = h[q].dist; st3(o + 2 + 7 * q, h[q].pos); st3(o + 5 + 7 * q, h[q].nrm); }
        }
      }
      int total; const int at = nm + teamExclScan(t, mp ? 1 : 0, &total);
      if (mp) tmp[16 * at + 15] = __int_as_float(k);
      nm += total;
    }
    TSYNC();
    for (int base = 0; base < nm; base += TEAM) {
      const int j = base + t.tl;
      if (j < nm) {
        const int k = __float_as_int(tmp[16 * j + 15]);
        Hit h[2];
        const int n = collidePair(slot, sm, cand[c0 + k], h);
        auto o = tmp + 16 * k;
        o[0] = __int_as_float(n);
        if (n) { o[1] = h[0].dist; st3(o + 2, h[0].pos); st3(o + 5, h[0].nrm); }
      }
    }
    TSYNC();
    for (int base = 0; base < cn; base += TEAM) {
      const int k = base + t.tl;
      const int n = k < cn ? __float_as_int(tmp[16 * k]) : 0;
      int total; int at = ncon + teamExclScan(t, n, &total);
      for (int q = 0; q < n; q++, at++) {
        if (at >= D.nconmax) continue;
        const int pr = cand[c0 + k];
        auto c = sm + L.con + 12 * at; auto o = tmp + 16 * k + 1 + 7 * q;
        for (int e = 0; e < 7; e++) c[e] = o[e];
        c[7] = __int_as_float(pr); c[8] = __int_as_float(-1); c[9] = __int_as_float(D.pair_dim[pr]);
        // dofs that move exactly one of the two bodies (static per pair): the only non-zero columns of the contact's rows
        const int f1 = D.cg_frame[D.pair_cg1[pr]], f2 = D.cg_frame[D.pair_cg2[pr]];
        c[10] = __int_as_float(D.fr_mask[2 * f1] ^ D.fr_mask[2 * f2]); c[11] = __int_as_float(D.fr_mask[2 * f1 + 1] ^ D.fr_mask[2 * f2 + 1]);
      }
      ncon += total;
    }
    TSYNC();
  }
  if (ncon > D.nconmax) { *status |= MYO_WORLD_NCON_OVERFLOW; ncon = D.nconmax; }
  TSYNC();
  return ncon;
}
__device__ __noinline__ int makeConstraint(int slot, SP sm, Tm t, int ncon, int* status) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L;
  const int nv = D.nv, njmax = D.njmax;
  auto qpos = sm + L.qpos; auto qvel = sm + L.qvel;
  if (D.disableflags & DSBL_CONSTRAINT) return 0;
  // ---- equality rows (always active)
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
  // ---- joint limits, then tendon limits (row order by a team prefix sum)
  const int nlim = D.nlimj + D.nlimt;
  for (int base = 0; base < nlim; base += TEAM) {
    const int li = base + t.tl;
    float dlo = 1.f, dhi = 1.f, margin = 0.f; int cnt = 0;
    const bool isj = li < D.nlimj;
    const float* prm = nullptr;
    if (li < nlim) {
      prm = isj ? D.limj_prm + 12 * li : D.limt_prm + 12 * (li - D.nlimj);
      const float val = isj ? qpos[D.limj_q[li]] : sm[L.tenlen + D.limt_ten[li - D.nlimj]];
      margin = prm[2]; dlo = val - prm[0]; dhi = prm[1] - val;
      cnt = (dlo < margin) + (dhi < margin);
    }
    int total; int at = nefc + teamExclScan(t, cnt, &total);
    if (li < nlim) {
      int r0 = -1, r1 = -1;
      if (cnt) {
        const float v = isj ? qvel[D.limj_dof[li]] : sm[L.tenvel + D.limt_ten[li - D.nlimj]];
        const int id = isj ? li : li - D.nlimj, type = isj ? ROW_LIMJ : ROW_LIMT;
        if (dlo < margin) { if (at < njmax) { r0 = at; setRow(slot, sm, at, type, id, 1.f, 0, dlo, margin, prm[10], prm + 3, v, 0.f); } at++; }
        if (dhi < margin) { if (at < njmax) { r1 = at; setRow(slot, sm, at, type, id, -1.f, 0, dhi, margin, prm[10], prm + 3, -v, 0.f); } at++; }
      }
      sm[L.limrow + 2 * li] = __int_as_float(r0); sm[L.limrow + 2 * li + 1] = __int_as_float(r1);
    }
    nefc += total;
  }
  // ---- contacts: row addresses, base Jacobian rows, pyramid rows
  if (ncon > 0) {
    const int nb = D.maxbase;
    int unfit = 0x7fffffff;   // address of the first contact whose rows do not fit: rows are written contiguously below it
    for (int base = 0; base < ncon; base += TEAM) {
      const int c = base + t.tl;
      int rows = 0;
      if (c < ncon) {
        auto cr = sm + L.con + 12 * c;
        const int pr = __float_as_int(cr[7]), dim = __float_as_int(cr[9]);
        const float inc = D.pair_prm[16 * pr] - D.pair_prm[16 * pr + 1];
        if (cr[0] < inc) rows = dim == 1 ? 1 : 2 * (dim - 1);
      }
      int total; const int at = nefc + teamExclScan(t, rows, &total);
      if (c < ncon) sm[L.con + 12 * c + 8] = __int_as_float((rows > 0 && at + rows <= njmax) ? at : -1);
      if (c < ncon && rows > 0 && at + rows > njmax) unfit = min(unfit, at);
      nefc += total;
    }
    unfit = teamMinI(t, unfit);
    // a contact that does not fit is dropped whole, and so is every later one (their addresses lie beyond njmax): the rows
    // that exist are [0, unfit), all written below
    if (unfit < nefc) { *status |= MYO_WORLD_NEFC_OVERFLOW; nefc = min(unfit, njmax); }
    // base Jacobian rows: one item = (contact, dof)
    TFOR(idx, ncon * nv) {
      const int c = idx / nv, k = idx - c * nv;
      auto cr = sm + L.con + 12 * c;
      auto J = sm + L.conJ + (c * nb) * nv + k;
      if (!(((unsigned)__float_as_int(cr[10 + (k >> 5)]) >> (k & 31)) & 1u)) {   // dof k moves both bodies or neither
        J[0] = 0.f; if (nb > 1) { J[nv] = 0.f; J[2 * nv] = 0.f; } if (nb > 3) J[3 * nv] = 0.f; continue;
      }
      const int f1 = D.cg_frame[D.pair_cg1[__float_as_int(cr[7])]];
      const int in1 = (D.fr_mask[2 * f1 + (k >> 5)] >> (k & 31)) & 1;   // it moves exactly one of them: the first (-) or the second (+)
      const f3 pos = ld3(cr + 1), n = ld3(cr + 4);
      f3 y, z; makeFrame(n, &y, &z);
      const float sgn = in1 ? -1.f : 1.f;
      auto cd = sm + L.cdof + 6 * k;
      const f3 jr = ld3(cd) * sgn;
      const f3 jp = (ld3(cd + 3) + cross(ld3(cd), pos - ld3(sm + L.rootcom + 3 * D.fr_root[D.dof_frame[k]]))) * sgn;
      J[0] = dot(n, jp);
      if (nb > 1) { J[nv] = dot(y, jp); J[2 * nv] = dot(z, jp); }
      if (nb > 3) J[3 * nv] = dot(n, jr);
    }
    TSYNC();
    auto bx = sm + L.basev;
    baseMul(slot, sm, t, ncon, qvel, bx);
    TSYNC();
    TFOR(c, ncon) {
      auto cr = sm + L.con + 12 * c;
      const int adr = __float_as_int(cr[8]);
      if (adr < 0) continue;
      const int pr = __float_as_int(cr[7]), dim = __float_as_int(cr[9]);
      auto pp = D.pair_prm + 16 * pr;
      const float margin = pp[0] - pp[1], dist = cr[0];
      const int c1 = D.pair_cg1[pr], c2 = D.pair_cg2[pr];
      const float tran = D.cg_invweight[2 * c1] + D.cg_invweight[2 * c2];
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

// H = M + J^T D_active J (lower triangle), Cholesky in place
__device__ __noinline__ void hessianFactor(int slot, SP sm, Tm t, int ncon, int nefc) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L;
  const int nv = D.nv, nb = D.maxbase;
  auto H = hessian(sm, L.H, nv, L.hshared);
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
      const int r = __float_as_int(sm[L.limrow + 2 * (D.nlimj + lt) + q]);
      if (r < 0 || !(EFC(E_JAR, r) < 0.f)) continue;
      TSYNC();
      const int tn = D.limt_ten[lt], n0 = D.ten_nzstart[tn], cnt = D.ten_nzstart[tn + 1] - n0;
      const float Dr = EFC(E_D, r);
      TFOR(i, cnt * cnt) {
        const int a = i / cnt, b = i - a * cnt;
        const int da = D.nz_dof[n0 + a], db = D.nz_dof[n0 + b];
        if (da >= db) H[TRI(da) + db] += Dr * sm[L.tenJ + n0 + a] * sm[L.tenJ + n0 + b];
      }
    }
  TSYNC();
  for (int c = 0; c < ncon; c++) {
    auto cr = sm + L.con + 12 * c;
    const int adr = __float_as_int(cr[8]), dim = __float_as_int(cr[9]);
    if (adr < 0) continue;
    // weights in base-row space: W = sum_p D_p w_p w_p^T, w_p = e_n +- mu e_t
    float Wnn = 0.f, Wnt[3] = {0.f, 0.f, 0.f}, Wtt[3] = {0.f, 0.f, 0.f};
    if (dim == 1) { if (EFC(E_JAR, adr) < 0.f) Wnn = EFC(E_D, adr); }
    else
      for (int k = 1; k < dim; k++) {
        const int r0 = adr + 2 * (k - 1);
        const float mu = EFC(E_COEF, r0);
        const float Dp = EFC(E_JAR, r0) < 0.f ? EFC(E_D, r0) : 0.f, Dm = EFC(E_JAR, r0 + 1) < 0.f ? EFC(E_D, r0 + 1) : 0.f;
        Wnn += Dp + Dm; Wnt[k - 1] = mu * (Dp - Dm); Wtt[k - 1] = mu * mu * (Dp + Dm);
      }
    if (Wnn == 0.f) continue;
    auto Jn = sm + L.conJ + (c * nb) * nv;
    // The contact Jacobian is non-zero only on dofs that move exactly one of the two bodies: that bit set is static per
    // pair.  Lane tl owns the tl-th, (tl + TEAM)-th ... set bit as a row of H and finds it directly (__fns), so the
    // lanes of a team work on their rows at the same time; columns are the set bits up to the row.
    const unsigned mlo = (unsigned)__float_as_int(cr[10]), mhi = (unsigned)__float_as_int(cr[11]);   // collision() stored the pair's dof set
    const int nlo = __popc(mlo), nrow = nlo + __popc(mhi);
    #pragma unroll 1
    for (int a = t.tl; a < nrow; a += TEAM) {
      const int r = a < nlo ? (int)__fns(mlo, 0, a + 1) : 32 + (int)__fns(mhi, 0, a + 1 - nlo);
      const float jn = Jn[r];
      float un = Wnn * jn, ut[3] = {0.f, 0.f, 0.f};
      #pragma unroll 1
      for (int k = 1; k < dim; k++) {
        const float jt = Jn[k * nv + r];
        un = fmaf(Wnt[k - 1], jt, un); ut[k - 1] = fmaf(Wnt[k - 1], jn, Wtt[k - 1] * jt);
      }
      auto Hr = H + TRI(r);
      #pragma unroll 1
      for (int half = 0; half < 2; half++) {
        unsigned w;
        if (half == 0) w = r >= 32 ? mlo : (mlo & ((2u << r) - 1u));
        else w = r >= 32 ? (mhi & ((2u << (r - 32)) - 1u)) : 0u;
        #pragma unroll 1
        while (w) {
          const int cc = 32 * half + __ffs((int)w) - 1;
          w &= w - 1u;
          auto Jc = Jn + cc;
          float acc = fmaf(un, Jc[0], Hr[cc]);
          if (dim > 1) { acc = fmaf(ut[0], Jc[nv], acc); acc = fmaf(ut[1], Jc[2 * nv], acc); }
          if (dim > 3) acc = fmaf(ut[2], Jc[3 * nv], acc);
          Hr[cc] = acc;
        }
      }
    }
    TSYNC();   // row ownership changes with the next contact's bit set
  }
  TSYNC();
  // left-looking Cholesky, lanes over rows
  for (int j = 0; j < nv; j++) {
    float sj = 0.f;
    for (int r = j + t.tl; r < nv; r += TEAM) {
      auto Hr = H + TRI(r); auto Hj = H + TRI(j);
      float s0 = Hr[j], s1 = 0.f;
      int k = 0;
      for (; k + 1 < j; k += 2) { s0 = fmaf(-Hr[k], Hj[k], s0); s1 = fmaf(-Hr[k + 1], Hj[k + 1], s1); }
      if (k < j) s0 = fmaf(-Hr[k], Hj[k], s0);
      s0 += s1;
      if (r == j) sj = s0; else Hr[j] = s0;
    }
    sj = teamBcast(t, sj, 0);
    const float d = sqrtf(fmaxf(sj, MINVAL)), inv = 1.f / d;
    for (int r = j + t.tl; r < nv; r += TEAM) { if (r == j) H[TRI(j) + j] = d; else H[TRI(r) + j] *= inv; }
    TSYNC();
  }
}
__device__ __noinline__ LsPt lsEval(int slot, SP sm, Tm t, int nefc, float a, const float* qg, float gtol) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L;
  float q0 = 0.f, q1 = 0.f, q2 = 0.f, qa = 0.f;
  const int neq = D.neq;
  TFOR(r, nefc) {
    const float jar = EFC(E_JAR, r), jv = EFC(E_JV, r), x = fmaf(a, jv, jar);
    if (r < neq || x < 0.f) {   // equality rows come first and are always active
      const float Dr = EFC(E_D, r), dj = Dr * jar;
      q0 = fmaf(0.5f * dj, jar, q0); q1 = fmaf(dj, jv, q1); q2 = fmaf(0.5f * Dr * jv, jv, q2); qa = fmaf(fabsf(dj), fabsf(jv), qa);
    }
  }
  q0 = teamSum(t, q0) + qg[0]; q1 = teamSum(t, q1) + qg[1]; q2 = teamSum(t, q2) + qg[2]; qa = teamSum(t, qa) + fabsf(qg[1]);
  LsPt p; p.alpha = a; p.cost = a * a * q2 + a * q1 + q0; p.d0 = 2.f * a * q2 + q1; p.d1 = 2.f * q2;
  p.tol = fmaxf(gtol, LS_NOISE * (qa + 2.f * fabsf(a) * q2));
  if (p.d1 <= 0.f) p.d1 = MINVAL;
  return p;
}
__device__ __noinline__ int solveNewton(int slot, SP sm, Tm t, int ncon, int nefc, int* status) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L;
  const int nv = D.nv;
  auto qacc = sm + L.qacc; auto Ma = sm + L.nvec; auto grad = Ma + nv; auto search = grad + nv; auto Mv = search + nv;
  auto tmp = Mv + nv;
  auto smooth = sm + L.smooth; auto qas = sm + L.qacc_smooth;
  const float scale = 1.f / (D.meaninertia * (float)max(1, nv));
  // warm start: cheaper of qacc_warmstart and qacc_smooth
  float costW = 3.4e38f;
  if (!(D.disableflags & DSBL_WARMSTART)) {
    mulMsparse(slot, sm + L.qM, sm + L.warm, Ma, t);
    TSYNC();
    float g = 0.f;
    TFOR(k, nv) g = fmaf(0.5f * (Ma[k] - smooth[k]), sm[L.warm + k] - qas[k], g);
    costW = teamSum(t, g) + residualCost(slot, sm, t, ncon, nefc, sm + L.warm);
  }
  const float costS = residualCost(slot, sm, t, ncon, nefc, qas);
  const bool useWarm = !(costW > costS);
  TFOR(k, nv) qacc[k] = useWarm ? sm[L.warm + k] : qas[k];
  TSYNC();
  mulMsparse(slot, sm + L.qM, qacc, Ma, t);
  TSYNC();
  float gauss = 0.f;
  TFOR(k, nv) gauss = fmaf(0.5f * (Ma[k] - smooth[k]), qacc[k] - qas[k], gauss);
  gauss = teamSum(t, gauss);
  float cost = gauss + residualCost(slot, sm, t, ncon, nefc, qacc);
  int iter = 0;
  // #pragma unroll 1
  for (;;) {
    // gradient and termination tests
    constraintForce(slot, sm, t, ncon, nefc, tmp);
    float gn = 0.f, gr = 0.f;
    TFOR(k, nv) {
      const float g = Ma[k] - smooth[k] - tmp[k]; grad[k] = g; search[k] = g; gn = fmaf(g, g, gn);
      const float a = fabsf(Ma[k]) + fabsf(smooth[k]) + fabsf(tmp[k]); gr = fmaf(a, a, gr);
    }
    gn = sqrtf(teamSum(t, gn)); gr = sqrtf(teamSum(t, gr));
    TSYNC();
    // MuJoCo's test (scaled gradient < tolerance = 1e-8) cannot be met in float32: the gradient is a difference of
    // O(gr) terms and bottoms out near eps * gr, after which further iterations only chase rounding noise.
    if (iter > 0 && (scale * gn < D.tolerance || gn < GRAD_NOISE * gr)) break;
    if (iter >= D.iterations) { *status |= MYO_WORLD_SOLVER_MAXITER; break; }
    hessianFactor(slot, sm, t, ncon, nefc);
    cholSolve(slot, hessian(sm, L.H, nv, L.hshared), search, t);
    float sn = 0.f;
    TFOR(k, nv) { const float s = -search[k]; search[k] = s; sn = fmaf(s, s, sn); }
    sn = sqrtf(teamSum(t, sn));
    TSYNC();
    if (sn < MINVAL) break;
    // exact line search along `search`
    mulMsparse(slot, sm + L.qM, search, Mv, t);
    auto bx = sm + L.basev;
    baseMul(slot, sm, t, ncon, search, bx);
    TSYNC();
    TFOR(r, nefc) EFC(E_JV, r) = rowMul(slot, sm, r, search, bx);
    float g1 = 0.f, g2 = 0.f;
    TFOR(k, nv) { g1 = fmaf(search[k], Ma[k] - smooth[k], g1); g2 = fmaf(0.5f * search[k], Mv[k], g2); }
    TSYNC();
    const float qg[3] = {gauss, teamSum(t, g1), teamSum(t, g2)};
    const float gtol = D.tolerance * D.ls_tolerance * sn / scale;
    int lsiter = 0;
    LsPt p0 = lsEval(slot, sm, t, nefc, 0.f, qg, gtol); lsiter++;
    LsPt p1 = lsEval(slot, sm, t, nefc, p0.alpha - p0.d0 / p0.d1, qg, gtol); lsiter++;
    if (p0.cost < p1.cost) p1 = p0;
    bool done = fabsf(p1.d0) < p1.tol;
    if (!done) {
      const float dir = p1.d0 < 0.f ? 1.f : -1.f;
      LsPt p2 = p1; bool have2 = false;
      // #pragma unroll 1
      while (p1.d0 * dir <= -p1.tol && lsiter < D.ls_iterations) {
        p2 = p1; have2 = true;
        p1 = lsEval(slot, sm, t, nefc, p1.alpha - p1.d0 / p1.d1, qg, gtol); lsiter++;
        if (fabsf(p1.d0) < p1.tol) { done = true; break; }
      }
      if (!done && lsiter < D.ls_iterations && have2) {
        LsPt lo = p2, hi = p1;
        // #pragma unroll 1
        while (lsiter < D.ls_iterations) {
          bool moved = false;
          LsPt best = lo.cost < hi.cost ? lo : hi;
          const float ca[3] = {lo.alpha - lo.d0 / lo.d1, hi.alpha - hi.d0 / hi.d1, 0.5f * (lo.alpha + hi.alpha)};
          LsPt cand[3];
          // #pragma unroll 1
          for (int q = 0; q < 3; q++) { cand[q] = lsEval(slot, sm, t, nefc, ca[q], qg, gtol); lsiter++; }
          // #pragma unroll 1
          for (int q = 0; q < 3; q++) {
            if (done) break;
            const LsPt c = cand[q];
            if (fabsf(c.d0) < c.tol) { p1 = c; done = true; break; }
            if (!((c.alpha - lo.alpha) * dir > 0.f && (hi.alpha - c.alpha) * dir > 0.f)) continue;
            if (c.d0 * dir < 0.f) lo = c; else hi = c;
            moved = true;
            if (c.cost < best.cost) best = c;
          }
          if (done) break;
          if (!moved) { p1 = best; break; }
          p1 = lo.cost < hi.cost ? lo : hi;
        }
      }
    }
    const float alpha = p1.alpha;
    if (alpha == 0.f) break;
    TFOR(k, nv) { qacc[k] = fmaf(alpha, search[k], qacc[k]); Ma[k] = fmaf(alpha, Mv[k], Ma[k]); }
    float c2 = 0.f;
    TFOR(r, nefc) {
      const float jar = fmaf(alpha, EFC(E_JV, r), EFC(E_JAR, r));
      EFC(E_JAR, r) = jar;
      if (rowOn(EFCI(E_DESC0, r), jar)) c2 = fmaf(0.5f * EFC(E_D, r) * jar, jar, c2);
    }
    TSYNC();
    float g = 0.f;
    TFOR(k, nv) g = fmaf(0.5f * (Ma[k] - smooth[k]), qacc[k] - qas[k], g);
    gauss = teamSum(t, g);
    const float newcost = gauss + teamSum(t, c2);
    iter++;
    const float improvement = scale * (cost - newcost);
    cost = newcost;
    if (improvement < D.tolerance) break;
  }
  constraintForce(slot, sm, t, ncon, nefc, sm + L.qfrc_con);
  return iter;
}

__device__ __forceinline__ bool anyBad(SP v, int n, Tm t) {
  bool bad = false;
  TFOR(i, n) bad |= !(fabsf(v[i]) < 1e10f);
  return teamAny(t, bad);
}
// ------------------------------------------------------------------------------------------ solveKernel (phase pipeline)
// The data-dependent middle of a step: constraint rows, Newton solve, mj_checkAcc, touch sensors.  Reads what the TILE
// kernels left in the scratch (contacts and their count, smooth forces, qacc_smooth, the mass matrix and its factor) and
// leaves qacc, qfrc_constraint, the warm start, the row records and the counts there for finishKernel / conoutKernel.
// Small CTAs, no barriers: a world that needs three Newton iterations holds up nobody but the three other worlds of its warp.
enum { WS_STATUS = 0, WS_TIME = 1, WS_SKIP = 2 };
#ifndef SOLVE_WARPS
#define SOLVE_WARPS 4
#endif
static inline size_t solveTileBytes(int nv) { return (size_t)SOLVE_WARPS * HS * TRI(nv) * sizeof(float); }
__global__ void __launch_bounds__(32 * SOLVE_WARPS) solveKernel(int slot, int fslot, int nworld, int stride, const int* __restrict__ mask, int check_acc,
                                                                float* __restrict__ scratch) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L;
  const int lane = threadIdx.x & 31, g = lane / TEAM;
  const size_t warp = (size_t)blockIdx.x * SOLVE_WARPS + (threadIdx.x >> 5);
  Tm t; t.tl = lane % TEAM; t.mask = TEAM == 32 ? 0xffffffffu : (((1u << TEAM) - 1u) << (g * TEAM));
  World W; W.w = warp * HS + g; W.nw = (size_t)nworld;
  if (!(W.w < W.nw && !(mask && !mask[W.w]))) return;
  SP sm; sm.p = scratch + W.w; sm.s = stride;
  const int nv = D.nv;
  int status = __float_as_int(sm[L.wstate + WS_STATUS]);
  const int ncon = __float_as_int(sm[L.counts]);
  int nefc = 0, niter = 0;
  TSYNC();
  nefc = makeConstraint(slot, sm, t, ncon, &status);
  if (nefc == 0) {
    TFOR(i, nv) { const float a = sm[L.qacc_smooth + i]; sm[L.qacc + i] = a; sm[L.warm + i] = a; sm[L.qfrc_con + i] = 0.f; }
  } else {
    niter = solveNewton(slot, sm, t, ncon, nefc, &status);
    TFOR(i, nv) sm[L.warm + i] = sm[L.qacc + i];
  }
  TSYNC();
  // mj_checkAcc: a world whose qacc went bad is reset (qpos0, zero velocity / activation) and flagged; finishKernel does
  // not integrate it, its next forward pass starts from the reset state.
  int skip = 0;
  if (check_acc && anyBad(sm + L.qacc, nv, t)) {
    TSYNC();
    resetState(slot, sm, t);
    status |= MYO_WORLD_BADQACC;
    skip = 1;
  }
  if (fslot != MYO_NSLOT && D.nsensor > 0) touchSensors(slot, fslot, sm, t, W, nefc > 0 ? ncon : 0);
  if (t.tl == 0) {
    sm[L.counts + 1] = __int_as_float(nefc); sm[L.counts + 2] = __int_as_float(niter);
    sm[L.wstate + WS_STATUS] = __int_as_float(status); sm[L.wstate + WS_SKIP] = __int_as_float(skip);
  }
}

// ------------------------------------------------------------------------------------------ stepKernel (whole step, one launch)
// The warps of a CTA walk the phases of the step together (a CTA barrier between phases): the step is several
// hundred KB of SASS, far beyond the instruction caches, and keeping an SM's warps inside the same few KB of code
// removes most instruction-fetch stalls.  Teams still diverge freely inside a phase (contacts, solver iterations).
#ifdef NO_PHASE_SYNC
#define PHASE_SYNC() do { } while (0)
#else
#define PHASE_SYNC() __syncthreads()
#endif
// mj_forward for one world: everything from the state in the scratch
// posvel: mj_step1-style refresh -- only the position- and velocity-dependent stages run (kinematics, com, tendons,
// transmission lengths / velocities, body velocities, bias and passive forces); contacts, constraint rows, actuator
// forces, accelerations and the warm start keep the values of the last full pass.
__device__ __forceinline__ void forwardWorld(int slot, int fslot, SP sm, Tm t, World W, bool live, bool last, bool posvel, int* status, int* ncon_o, int* nefc_o, int* niter_o) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L;
  const int nv = D.nv;
  if (live) { kinematics(slot, sm, t); comPos(slot, sm, t); }
  PHASE_SYNC();
  if (live && last) writePositionOutputs(slot, fslot, sm, t, W);
  PHASE_SYNC();
  if (live) tendons(slot, fslot, sm, t, W);
  PHASE_SYNC();
  if (live && !posvel) inertia(slot, fslot, sm, t, W);
  PHASE_SYNC();
  int ncon = 0, nefc = 0, niter = 0;
  if (live && !posvel && !(D.disableflags & (DSBL_CONTACT | DSBL_CONSTRAINT))) ncon = collision(slot, sm, t, status);
  PHASE_SYNC();
  if (live) velocity(slot, fslot, sm, t, W);
  PHASE_SYNC();
  if (posvel) {   // CTA-uniform
    if (live) {   // mj_transmission outputs of the refreshed tendon stage (actuation() writes them on a full pass)
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
    }
    *ncon_o = 0; *nefc_o = 0; *niter_o = 0;
    return;
  }
  if (live) actuation(slot, fslot, sm, t, W);
  PHASE_SYNC();
  if (live) nefc = makeConstraint(slot, sm, t, ncon, status);
  PHASE_SYNC();
  if (live) {
    if (nefc == 0) {
      TFOR(i, nv) { const float a = sm[L.qacc_smooth + i]; sm[L.qacc + i] = a; sm[L.warm + i] = a; sm[L.qfrc_con + i] = 0.f; }
    } else {
      niter = solveNewton(slot, sm, t, ncon, nefc, status);
      TFOR(i, nv) sm[L.warm + i] = sm[L.qacc + i];
    }
    TSYNC();
  }
  if (live && t.tl == 0) { sm[L.counts] = __int_as_float(ncon); sm[L.counts + 1] = __int_as_float(nefc); sm[L.counts + 2] = __int_as_float(niter); }
#ifndef NO_SYNC_AFTER_NEWTON
  PHASE_SYNC();
#endif
  *ncon_o = ncon; *nefc_o = nefc; *niter_o = niter;
}

// One warp = HS worlds, TEAM lanes each; CTA_WARPS warps per CTA.  `scratch`: [SmemLayout.total][stride], world fastest.
// 14 warps, one CTA per SM: 8192 worlds x 8 lanes = 2048 warps = 147 CTAs, one wave on the 148 SMs with every SM
// equally loaded (8-warp CTAs left 40 SMs with half the warps of the rest); <= 144 registers per thread.
#ifndef CTA_WARPS
#define CTA_WARPS 14
#endif
#ifndef CTA_MINBLOCKS
#define CTA_MINBLOCKS 1
#endif
#ifndef HSHARED_MAX_BYTES
#define HSHARED_MAX_BYTES (200 * 1024)
#endif
static inline size_t hessianTileBytes(int nv) { return (size_t)CTA_WARPS * HS * TRI(nv) * sizeof(float); }
__global__ void __launch_bounds__(32 * CTA_WARPS, CTA_MINBLOCKS) stepKernel(int slot, int nworld, int stride, int nsub, int forward_only, const int* __restrict__ mask,
                                                             int do_reset, float* __restrict__ scratch) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L; const FieldPtrs& F = tP[slot].F;
  const int lane = threadIdx.x & 31, g = lane / TEAM;
  const size_t warp = (size_t)blockIdx.x * CTA_WARPS + (threadIdx.x >> 5);
  Tm t; t.tl = lane % TEAM; t.mask = TEAM == 32 ? 0xffffffffu : (((1u << TEAM) - 1u) << (g * TEAM));
  World W; W.w = warp * HS + g; W.nw = (size_t)nworld;
  const bool live = W.w < W.nw && !(mask && !mask[W.w]);   // a mask restricts any launch to the flagged worlds
  SP sm; sm.p = scratch + (live ? W.w : 0); sm.s = stride;   // a dead team never touches its scratch
  if (!live) W.w = 0;
  const int nq = D.nq, nv = D.nv, na = D.na, nu = D.nu;
  void* gq = F.p[MYO_F_QPOS]; void* gv = F.p[MYO_F_QVEL]; void* ga = F.p[MYO_F_ACT]; void* gc = F.p[MYO_F_CTRL];
  void* gw = F.p[MYO_F_QACC_WARMSTART]; void* gt = F.p[MYO_F_TIME]; void* gs = F.p[MYO_F_STATUS];
  int status = 0;
  float time = 0.f;
  if (live) {
    status = gs ? OUTI(gs, 0) : 0;
    time = OUTF(gt, 0);
    if (do_reset) {
      resetState(slot, sm, t);
      TFOR(i, nu) sm[L.ctrl + i] = 0.f;
      time = 0.f; status = 0;
    } else {
      TFOR(i, nq) sm[L.qpos + i] = OUTF(gq, i);
      TFOR(i, nv) { sm[L.qvel + i] = OUTF(gv, i); sm[L.warm + i] = OUTF(gw, i); }
      TFOR(i, na) sm[L.act + i] = OUTF(ga, i);
      TFOR(i, nu) sm[L.ctrl + i] = OUTF(gc, i);
    }
    TSYNC();
  }
  const bool posvel = forward_only == 2;
  const int nloop = forward_only ? 1 : nsub;
  for (int sub = 0; sub < nloop; sub++) {
    const bool last = sub == nloop - 1;
    const int fslot = last ? slot : MYO_NSLOT;
    if (live && !forward_only) {   // mj_checkPos / mj_checkVel
      if (anyBad(sm + L.qpos, nq, t)) { TSYNC(); resetState(slot, sm, t); status |= MYO_WORLD_BADQPOS; }
      if (anyBad(sm + L.qvel, nv, t)) { TSYNC(); resetState(slot, sm, t); status |= MYO_WORLD_BADQVEL; }
    }
    int ncon = 0, nefc = 0, niter = 0;
    forwardWorld(slot, fslot, sm, t, W, live, last, posvel, &status, &ncon, &nefc, &niter);
    // mj_checkAcc: a world whose qacc went bad is reset (qpos0, zero velocity / activation) and flagged; this substep
    // does not integrate it, its next forward pass starts from the reset state.
    bool skip = false;
    if (live && !forward_only && anyBad(sm + L.qacc, nv, t)) {
      TSYNC();
      resetState(slot, sm, t);
      status |= MYO_WORLD_BADQACC;
      skip = true;
    }
    if (live && last && !posvel) {
      writeConstraintOutputs(slot, fslot, sm, t, W, ncon, nefc, niter);
      if (D.nsensor > 0) touchSensors(slot, fslot, sm, t, W, nefc > 0 ? ncon : 0);
      void* o;
      if ((o = F.p[MYO_F_QACC])) TFOR(i, nv) OUTF(o, i) = sm[L.qacc + i];
      if ((o = F.p[MYO_F_QFRC_CONSTRAINT])) TFOR(i, nv) OUTF(o, i) = sm[L.qfrc_con + i];
      if ((o = F.p[MYO_F_QFRC_INVERSE])) TFOR(i, nv) OUTF(o, i) = 0.f;
    }
    if (forward_only) break;
    if (live && !skip) {
    // ---- mj_Euler with implicit joint damping
    const float h = D.timestep;
    auto qa = sm + L.tmpv;
    if (D.any_damping) {
      TFOR(e, D.nM) sm[L.qH + e] = sm[L.qM + e] + (D.M_row[e] == D.M_col[e] ? h * D.dof_damping[D.M_row[e]] : 0.f);
      TFOR(i, nv) qa[i] = sm[L.smooth + i] + sm[L.qfrc_con + i];
      TSYNC();
      factorLDL(slot, sm + L.qH, sm + L.qHinv, t);
      solveLDL(slot, sm + L.qH, sm + L.qHinv, qa, t);
    } else {
      TFOR(i, nv) qa[i] = sm[L.qacc + i];
      TSYNC();
    }
    if (!(D.disableflags & DSBL_ACTUATION))
      TFOR(i, nu) {
        const int aa = D.act_actadr[i];
        if (aa < 0) continue;
        float a = fmaf(h, sm[L.actdot + aa], sm[L.act + aa]);
        if (D.act_flags[i] & ACT_ACTLIMITED) a = clipf(a, D.act_actrange[2 * i], D.act_actrange[2 * i + 1]);
        sm[L.act + aa] = a;
      }
    TFOR(i, nv) sm[L.qvel + i] = fmaf(h, qa[i], sm[L.qvel + i]);
    TSYNC();
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
    time += h;
    TSYNC();
    }
#ifndef NO_SYNC_AFTER_EULER
    PHASE_SYNC();
#endif
  }
  // ---- state out
  if (!live) return;
  if (!forward_only || do_reset) {
    TFOR(i, nq) OUTF(gq, i) = sm[L.qpos + i];
    TFOR(i, nv) OUTF(gv, i) = sm[L.qvel + i];
    TFOR(i, na) OUTF(ga, i) = sm[L.act + i];
    if (t.tl == 0) OUTF(gt, 0) = time;
  }
  if (!posvel) TFOR(i, nv) OUTF(gw, i) = sm[L.warm + i];
  if (gs && t.tl == 0) OUTI(gs, 0) = status;
}

}  // namespace

#define CAT2_(a, b) a##b
#define CAT2(a, b) CAT2_(a, b)
#ifndef VSUFFIX   // symbol suffix of this build of the kernel: the team width, or e.g. 8x2 for two co-resident CTAs per SM
#define VSUFFIX TEAM
#endif
#define VFN(name) CAT2(name, VSUFFIX)

int VFN(myo_team_upload_)(int slot, const DevModel& D, const SmemLayout& L, const FieldPtrs& F, void* stream) {
  static TParams staging[MYO_NSLOT];   // must outlive the async copy
  if (slot < 0 || slot >= MYO_NSLOT) return -1;
  TParams& P = staging[slot];
  P.D = D; P.L = L; P.F = F;
  P.L.hshared = hessianTileBytes(D.nv) <= HSHARED_MAX_BYTES ? 1 : 0;
  return cudaMemcpyToSymbolAsync(tP, &P, sizeof(TParams), (size_t)slot * sizeof(TParams), cudaMemcpyHostToDevice, (cudaStream_t)stream) == cudaSuccess ? 0 : -1;
}
static int optInShared(const void* fn, size_t bytes, bool* done) {
  int dev = 0;
  if (bytes <= 48 * 1024) return 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return -1;
  if (done[dev]) return 0;
  if (cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)HSHARED_MAX_BYTES) != cudaSuccess) return -1;
  done[dev] = true;
  return 0;
}
// whole step in one launch (MYO_ENGINE=mono)
int VFN(myo_team_launch_)(int slot, const SmemLayout& L, int nworld, int stride, int nsub, int forward_only, const int* reset_mask, int do_reset,
                          float* scratch, void* stream) {
  const int nwarp = (nworld + HS - 1) / HS;
  size_t shmem = hessianTileBytes(L.nv);
  if (shmem > HSHARED_MAX_BYTES) shmem = 0;
  static bool configured[64] = {false};   // per device: opt in to > 48 KB of dynamic shared memory once
  if (optInShared((const void*)stepKernel, shmem, configured) != 0) return -1;
  stepKernel<<<(nwarp + CTA_WARPS - 1) / CTA_WARPS, 32 * CTA_WARPS, shmem, (cudaStream_t)stream>>>(slot, nworld, stride, nsub, forward_only, reset_mask, do_reset, scratch);
  return cudaPeekAtLastError() == cudaSuccess ? 0 : -1;
}
// constraint rows + Newton of the phase pipeline
int VFN(myo_team_solve_)(int slot, int fslot, const SmemLayout& L, int nworld, int stride, const int* mask, int check_acc, float* scratch, void* stream) {
  const int nwarp = (nworld + HS - 1) / HS;
  // same rule as the upload (hshared): the Hessian tiles live in shared memory iff the one-launch kernel's tiles fit
  size_t shmem = hessianTileBytes(L.nv) <= HSHARED_MAX_BYTES ? solveTileBytes(L.nv) : 0;
  static bool configured[64] = {false};
  if (optInShared((const void*)solveKernel, shmem, configured) != 0) return -1;
  solveKernel<<<(nwarp + SOLVE_WARPS - 1) / SOLVE_WARPS, 32 * SOLVE_WARPS, shmem, (cudaStream_t)stream>>>(slot, fslot, nworld, stride, mask, check_acc, scratch);
  return cudaPeekAtLastError() == cudaSuccess ? 0 : -1;
}
