// Wide mjData outputs, written by a second kernel with one THREAD per (world, item) and the world axis across the
// lanes: every store of a warp is one full 128-byte line of the [k][nworld] field, and every model-table load is
// warp-uniform.  The step kernel (team.cu) keeps T lanes per world, so its own stores cover only 32/T consecutive
// worlds (16 bytes per element at T = 8); the fields below are 70 % of the bytes of a full mjx.Data and need nothing
// but the frames / sparse rows that the step kernel leaves in the per-world scratch:
//   site_xpos, site_xmat, geom_xpos, geom_xmat, xpos, xquat, xmat, xipos, ximat   (mj_kinematics outputs)
//   ten_J, actuator_moment                                                      (mj_tendon / mj_transmission, dense rows)
//   qM                                                                          (mj_crb, dense symmetric)
//   efc_J                                                                       (mj_makeConstraint, dense rows)
//   contact_* , efc_pos/margin/D/aref/force/type, xanchor, xaxis, cdof          (small per-row copies)
//   ten_wrapadr, ten_wrapnum, wrap_obj, wrap_xpos                               (mj_tendon's path points)
// Same arithmetic, in the same order, as the step kernel's own writers for these fields (which remain in team.cu and
// run when this kernel is switched off with MYO_WIDE_OUT=0).
#include <cuda_runtime.h>

#include <cstdlib>

#include "devmodel.h"
#include "internal.h"

struct OutParams { DevModel D; SmemLayout L; FieldPtrs F; };
__constant__ OutParams oP[MYO_PARAM_SLOTS];

namespace {
enum { ROW_EQ = 0, ROW_LIMJ = 1, ROW_LIMT = 2, ROW_CON = 3 };
enum { E_D = 0, E_AREF, E_JAR, E_JV, E_FORCE, E_DESC0, E_COEF, E_DESC1, E_POS, E_MARGIN };   // as in team.cu

// the scratch is [SmemLayout.total][stride] with the world axis fastest: lanes = consecutive worlds read full lines
struct CtxT { const float* sm; size_t w, nw; int stride; };
__device__ __forceinline__ float S(const CtxT& c, int k) { return c.sm[k * c.stride]; }
// `n` consecutive elements of a [k][nworld] field, first one at element index k0: pointer stepping, no multiplies
struct Row {
  float* p; size_t nw;
  __device__ __forceinline__ Row(void* base, size_t k0, size_t nw_, size_t w) : p((float*)base + k0 * nw_ + w), nw(nw_) {}
#ifndef OUT_NO_STCS   // streaming (evict-first) stores: the outputs are written once and not re-read by the library
  __device__ __forceinline__ void put(float v) { __stcs(p, v); p += nw; }
  __device__ __forceinline__ void puti(int v) { __stcs((int*)p, v); p += nw; }
#else
  __device__ __forceinline__ void put(float v) { *p = v; p += nw; }
  __device__ __forceinline__ void puti(int v) { *(int*)p = v; p += nw; }
#endif
};

// world pose of an object fixed to frame f: p = x_f + R_f lpos, M = R_f lmat
__device__ __forceinline__ void framePose(const CtxT& c, const SmemLayout& L, int f, const float* lpos, const float* lmat, float* p, float* M, float* R) {
  const int o = L.frames + 16 * f;
#pragma unroll
  for (int k = 0; k < 9; k++) R[k] = S(c, o + 7 + k);
#pragma unroll
  for (int i = 0; i < 3; i++) p[i] = S(c, o + i) + fmaf(R[3 * i], lpos[0], fmaf(R[3 * i + 1], lpos[1], R[3 * i + 2] * lpos[2]));
  if (M)
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
      for (int j = 0; j < 3; j++) M[3 * i + j] = fmaf(R[3 * i], lmat[j], fmaf(R[3 * i + 1], lmat[3 + j], R[3 * i + 2] * lmat[6 + j]));
}

#ifndef OUT_THREADS
#define OUT_THREADS 128
#endif
#ifndef OUT_MINBLOCKS
#define OUT_MINBLOCKS 12
#endif
__global__ void __launch_bounds__(OUT_THREADS, OUT_MINBLOCKS) outKernel(int slot, int nworld, int STRIDE, const int* __restrict__ mask, int posvel,
                                                 const float* __restrict__ scratch) {
  const DevModel& D = oP[slot].D; const SmemLayout& L = oP[slot].L; const FieldPtrs& F = oP[slot].F;
  CtxT c; c.w = (size_t)blockIdx.x * OUT_THREADS + threadIdx.x; c.nw = (size_t)nworld; c.stride = STRIDE;
  if (c.w >= c.nw) return;
  if (mask && !mask[c.w]) return;   // a mask restricts the launch to the flagged worlds
  c.sm = scratch + c.w;
  const int nv = D.nv;
  void* sp = F.p[MYO_F_SITE_XPOS]; void* smt = F.p[MYO_F_SITE_XMAT]; void* gp = F.p[MYO_F_GEOM_XPOS]; void* gm = F.p[MYO_F_GEOM_XMAT];
  void* xpos = F.p[MYO_F_XPOS]; void* xquat = F.p[MYO_F_XQUAT]; void* xmat = F.p[MYO_F_XMAT]; void* xipos = F.p[MYO_F_XIPOS];
  void* ximat = F.p[MYO_F_XIMAT]; void* tj = F.p[MYO_F_TEN_J]; void* am = F.p[MYO_F_ACTUATOR_MOMENT]; void* qMo = F.p[MYO_F_QM];
  void* eJ = F.p[MYO_F_EFC_J];
  void* cg = F.p[MYO_F_CONTACT_GEOM]; void* cdim = F.p[MYO_F_CONTACT_DIM]; void* cadr = F.p[MYO_F_CONTACT_EFC_ADDRESS];
  void* cdist = F.p[MYO_F_CONTACT_DIST]; void* cpos = F.p[MYO_F_CONTACT_POS]; void* cfr = F.p[MYO_F_CONTACT_FRAME];
  void* cinc = F.p[MYO_F_CONTACT_INCLUDEMARGIN]; void* cfri = F.p[MYO_F_CONTACT_FRICTION]; void* csr = F.p[MYO_F_CONTACT_SOLREF];
  void* csi = F.p[MYO_F_CONTACT_SOLIMP];
  void* epos = F.p[MYO_F_EFC_POS]; void* emar = F.p[MYO_F_EFC_MARGIN]; void* eD = F.p[MYO_F_EFC_D]; void* earef = F.p[MYO_F_EFC_AREF];
  void* efrc = F.p[MYO_F_EFC_FORCE]; void* etyp = F.p[MYO_F_EFC_TYPE];
  void* xa = F.p[MYO_F_XANCHOR]; void* xx = F.p[MYO_F_XAXIS]; void* cdo = F.p[MYO_F_CDOF];
  void* wxp = F.p[MYO_F_WRAP_XPOS]; void* wob = F.p[MYO_F_WRAP_OBJ]; void* wad = F.p[MYO_F_TEN_WRAPADR]; void* wnu = F.p[MYO_F_TEN_WRAPNUM];
  const bool wpts = wxp || wob;
  void* qLDo = F.p[MYO_F_QLD]; void* cvo = F.p[MYO_F_CVEL]; void* cdd = F.p[MYO_F_CDOF_DOT];
  const int n0 = (sp || smt) ? D.nsite : 0, n1 = n0 + ((gp || gm) ? D.ngeom : 0),
            n2 = n1 + ((xpos || xquat || xmat || xipos || ximat) ? D.nbody : 0), n3 = n2 + (tj ? D.nten : 0), n4 = n3 + (am ? D.nu : 0),
            n5 = n4 + (qMo ? nv : 0), n6 = n5 + (eJ ? D.njmax : 0),
            n7 = n6 + ((cg || cdim || cadr || cdist || cpos || cfr || cinc || cfri || csr || csi) ? D.nconmax : 0),
            n8 = n7 + ((epos || emar || eD || earef || efrc || etyp) ? D.njmax : 0), n9 = n8 + ((xa || xx) ? D.njnt : 0),
            n10 = n9 + (cdo ? nv : 0),
            // tendon path points: one item per path unit, pulley marker, tendon, and clearable entry past the last point
            n11 = n10 + (wpts ? D.nunit : 0), n12 = n11 + (wpts ? D.npulley : 0), n13 = n12 + ((wad || wnu) ? D.nten : 0),
            n14 = n13 + (wpts ? 2 * D.ngunit : 0),
            n15 = n14 + (cvo ? D.nbody : 0), n16 = n15 + (cdd ? nv : 0), n17 = n16 + (qLDo ? nv : 0);
  const int ncon = __float_as_int(S(c, L.counts));
  const int nefc = __float_as_int(S(c, L.counts + 1));   // left by the step kernel: ncon, nefc, niter
  const size_t nw = c.nw, w = c.w;
  for (int it = blockIdx.y; it < n17; it += gridDim.y) {
    if (posvel && ((it >= n4 && it < n8) || it >= n16)) continue;   // position / velocity refresh: inertia, contacts and constraint rows are not recomputed
    if (it < n1) {   // site or geom pose
      const bool site = it < n0;
      const int q = site ? it : it - n0;
      void* po = site ? sp : gp; void* mo = site ? smt : gm;
      float p[3], M[9], R[9];
      framePose(c, L, site ? D.site_frame[q] : D.geom_frame[q], (site ? D.site_lpos : D.geom_lpos) + 3 * q, (site ? D.site_lmat : D.geom_lmat) + 9 * q,
                p, mo ? M : nullptr, R);
      if (po) { Row o(po, 3 * (size_t)q, nw, w); o.put(p[0]); o.put(p[1]); o.put(p[2]); }
      if (mo) {
        Row o(mo, 9 * (size_t)q, nw, w);
#pragma unroll
        for (int k = 0; k < 9; k++) o.put(M[k]);
      }
    } else if (it < n2) {
      const int b = it - n1;
      float p[3], M[9], R[9];
      const int f = D.body_frame[b];
      framePose(c, L, f, D.body_off_pos + 3 * b, D.body_off_mat + 9 * b, p, M, R);
      if (xpos) { Row o(xpos, 3 * (size_t)b, nw, w); o.put(p[0]); o.put(p[1]); o.put(p[2]); }
      if (xquat) {
        const int o_ = L.frames + 16 * f + 3;
        const float aw = S(c, o_), ax = S(c, o_ + 1), ay = S(c, o_ + 2), az = S(c, o_ + 3);
        const float* bq = D.body_off_quat + 4 * b;
        float qw = aw * bq[0] - ax * bq[1] - ay * bq[2] - az * bq[3], qx = aw * bq[1] + ax * bq[0] + ay * bq[3] - az * bq[2],
              qy = aw * bq[2] - ax * bq[3] + ay * bq[0] + az * bq[1], qz = aw * bq[3] + ax * bq[2] - ay * bq[1] + az * bq[0];
        const float n = sqrtf(qw * qw + qx * qx + qy * qy + qz * qz);
        if (n < 1e-15f) { qw = 1.f; qx = qy = qz = 0.f; } else { const float s = 1.f / n; qw *= s; qx *= s; qy *= s; qz *= s; }
        Row o(xquat, 4 * (size_t)b, nw, w); o.put(qw); o.put(qx); o.put(qy); o.put(qz);
      }
      if (xmat) {
        Row o(xmat, 9 * (size_t)b, nw, w);
#pragma unroll
        for (int k = 0; k < 9; k++) o.put(M[k]);
      }
      if (xipos) {
        const float* ip = D.body_ipos + 3 * b;
        Row o(xipos, 3 * (size_t)b, nw, w);
#pragma unroll
        for (int i = 0; i < 3; i++) o.put(p[i] + fmaf(M[3 * i], ip[0], fmaf(M[3 * i + 1], ip[1], M[3 * i + 2] * ip[2])));
      }
      if (ximat) {
        const float* im = D.body_imat + 9 * b;
        Row o(ximat, 9 * (size_t)b, nw, w);
#pragma unroll
        for (int i = 0; i < 3; i++)
#pragma unroll
          for (int j = 0; j < 3; j++) o.put(fmaf(M[3 * i], im[j], fmaf(M[3 * i + 1], im[3 + j], M[3 * i + 2] * im[6 + j])));
      }
    } else if (it < n4) {   // dense ten_J / actuator_moment row = zeros with the sparse entries (sorted by dof) merged in
      const bool isten = it < n3;
      const int a = isten ? it - n2 : it - n3;
      int q = 0, qend = 0, jd = -1;
      float gear = 1.f;
      if (isten) { q = D.ten_nzstart[a]; qend = D.ten_nzstart[a + 1]; }
      else {           // the tendon's sparse row scaled by the gear, or the joint's dof
        const int id = D.act_trnid[a];
        gear = D.act_gear[a];
        if (D.act_trntype[a] == 3) { q = D.ten_nzstart[id]; qend = D.ten_nzstart[id + 1]; } else jd = D.jnt_dofadr[id];
      }
      Row o(isten ? tj : am, (size_t)a * nv, nw, w);
      int next = q < qend ? D.nz_dof[q] : -1;
      const float* src = c.sm + (L.tenJ + q) * STRIDE;
      for (int i = 0; i < nv; i++) {
        float v = 0.f;
        if (i == next) { v = gear * *src; src += STRIDE; q++; next = q < qend ? D.nz_dof[q] : -1; }
        if (i == jd) v = gear;
        o.put(v);
      }
    } else if (it < n5) {   // dense symmetric qM row: zeros, then the dof's ancestor chain and its descendants
      const int i = it - n4;
      { Row o(qMo, (size_t)i * nv, nw, w); for (int k = 0; k < nv; k++) o.put(0.f); }
      float* row = (float*)qMo + (size_t)i * nv * nw + w;
      const int a = D.dof_Madr[i], d = D.dof_rowlen[i];
      for (int q = 0; q < d; q++) row[(size_t)D.M_col[a + q] * nw] = S(c, L.qM + a + q);
      for (int e = D.desc_start[i]; e < D.desc_start[i + 1]; e++) row[(size_t)D.desc_dof[e] * nw] = S(c, L.qM + D.desc_idx[e]);
    } else if (it < n6) {   // dense efc_J row from the implied representation (see team.cu "constraints")
      const int r = it - n5;
      const bool on = r < nefc;
      const int njm = D.njmax;
      const int d0 = on ? __float_as_int(S(c, L.efc + E_DESC0 * njm + r)) : 0, type = d0 & 3, id = d0 >> 2;
      const float coef = on ? S(c, L.efc + E_COEF * njm + r) : 0.f;
      Row o(eJ, (size_t)r * nv, nw, w);
      if (!on) { for (int k = 0; k < nv; k++) o.put(0.f); }
      else if (type == ROW_CON) {
        const int t1 = __float_as_int(S(c, L.efc + E_DESC1 * njm + r));
        const float* p0 = c.sm + (L.conJ + id * nv) * STRIDE;
        if (t1 >= 0) {
          const float* p1 = c.sm + (L.conJ + t1 * nv) * STRIDE;
#pragma unroll 4
          for (int k = 0; k < nv; k++) o.put(p0[k * STRIDE] + coef * p1[k * STRIDE]);
        } else {
#pragma unroll 4
          for (int k = 0; k < nv; k++) o.put(p0[k * STRIDE]);
        }
      } else if (type == ROW_LIMJ) {
        const int dof = D.limj_dof[id];
        for (int k = 0; k < nv; k++) o.put(k == dof ? coef : 0.f);
      } else if (type == ROW_EQ) {
        const int d1 = D.eq_dof1[id], d2 = D.eq_dof2[id];
        for (int k = 0; k < nv; k++) o.put(k == d1 ? 1.f : (k == d2 ? coef : 0.f));
      } else {
        const int tn = D.limt_ten[id];
        int q = D.ten_nzstart[tn];
        const int qend = D.ten_nzstart[tn + 1];
        for (int k = 0; k < nv; k++) {
          float v = 0.f;
          if (q < qend && D.nz_dof[q] == k) { v = coef * S(c, L.tenJ + q); q++; }
          o.put(v);
        }
      }
    } else if (it < n7) {   // contact record (padded to nconmax)
      const int q = it - n6;
      const bool on = q < ncon;
      const int cr = L.con + 12 * q;
      const int pr = on ? __float_as_int(S(c, cr + 7)) : 0;
      const float* pp = D.pair_prm + 16 * pr;
      if (cg) { Row o(cg, 2 * (size_t)q, nw, w); o.puti(on ? D.cg_geom[D.pair_cg1[pr]] : -1); o.puti(on ? D.cg_geom[D.pair_cg2[pr]] : -1); }
      if (cdim) Row(cdim, q, nw, w).puti(on ? __float_as_int(S(c, cr + 9)) : 0);
      if (cadr) Row(cadr, q, nw, w).puti(on ? __float_as_int(S(c, cr + 8)) : -1);
      if (cdist) Row(cdist, q, nw, w).put(on ? S(c, cr) : 0.f);
      if (cpos) { Row o(cpos, 3 * (size_t)q, nw, w); for (int k = 0; k < 3; k++) o.put(on ? S(c, cr + 1 + k) : 0.f); }
      if (cfr) {
        float n[3] = {0.f, 0.f, 0.f}, y[3] = {0.f, 0.f, 0.f}, z[3] = {0.f, 0.f, 0.f};
        if (on) {   // contact frame: normal, then the tangent basis of team.cu makeFrame
          for (int k = 0; k < 3; k++) n[k] = S(c, cr + 4 + k);
          float t[3] = {0.f, 0.f, 0.f};
          if (n[1] < 0.5f && n[1] > -0.5f) t[1] = 1.f; else t[2] = 1.f;
          const float d = fmaf(n[0], t[0], fmaf(n[1], t[1], n[2] * t[2]));
          for (int k = 0; k < 3; k++) t[k] -= n[k] * d;
          const float nn = sqrtf(fmaf(t[0], t[0], fmaf(t[1], t[1], t[2] * t[2])));
          if (nn < 1e-15f) { t[0] = 1.f; t[1] = 0.f; t[2] = 0.f; } else { const float s = 1.f / nn; t[0] *= s; t[1] *= s; t[2] *= s; }
          y[0] = t[0]; y[1] = t[1]; y[2] = t[2];
          z[0] = n[1] * t[2] - n[2] * t[1]; z[1] = n[2] * t[0] - n[0] * t[2]; z[2] = n[0] * t[1] - n[1] * t[0];
        }
        Row o(cfr, 9 * (size_t)q, nw, w);
        for (int k = 0; k < 3; k++) o.put(n[k]);
        for (int k = 0; k < 3; k++) o.put(y[k]);
        for (int k = 0; k < 3; k++) o.put(z[k]);
      }
      if (cinc) Row(cinc, q, nw, w).put(on ? pp[0] - pp[1] : 0.f);
      if (cfri) { Row o(cfri, 5 * (size_t)q, nw, w); for (int k = 0; k < 5; k++) o.put(on ? pp[2 + k] : 0.f); }
      if (csr) { Row o(csr, 2 * (size_t)q, nw, w); for (int k = 0; k < 2; k++) o.put(on ? pp[7 + k] : 0.f); }
      if (csi) { Row o(csi, 5 * (size_t)q, nw, w); for (int k = 0; k < 5; k++) o.put(on ? pp[9 + k] : 0.f); }
    } else if (it < n8) {   // per-row constraint scalars (padded to njmax)
      const int r = it - n7, njm = D.njmax;
      const bool on = r < nefc;
      if (epos) Row(epos, r, nw, w).put(on ? S(c, L.efc + E_POS * njm + r) : 0.f);
      if (emar) Row(emar, r, nw, w).put(on ? S(c, L.efc + E_MARGIN * njm + r) : 0.f);
      if (eD) Row(eD, r, nw, w).put(on ? S(c, L.efc + E_D * njm + r) : 0.f);
      if (earef) Row(earef, r, nw, w).put(on ? S(c, L.efc + E_AREF * njm + r) : 0.f);
      if (efrc) Row(efrc, r, nw, w).put(on ? S(c, L.efc + E_FORCE * njm + r) : 0.f);
      if (etyp) {
        const int d0 = on ? __float_as_int(S(c, L.efc + E_DESC0 * njm + r)) : 0, type = d0 & 3;
        Row(etyp, r, nw, w).puti(!on ? -1 : (type == ROW_EQ ? 0 : (type == ROW_LIMJ ? 3 : (type == ROW_LIMT ? 4 : (__float_as_int(S(c, L.efc + E_DESC1 * njm + r)) < 0 ? 5 : 6)))));
      }
    } else if (it < n9) {
      const int j = it - n8;
      if (xa) { Row o(xa, 3 * (size_t)j, nw, w); for (int k = 0; k < 3; k++) o.put(S(c, L.xanchor + 3 * j + k)); }
      if (xx) { Row o(xx, 3 * (size_t)j, nw, w); for (int k = 0; k < 3; k++) o.put(S(c, L.xaxis + 3 * j + k)); }
    } else if (it < n10) {
      const int i = it - n9;
      Row o(cdo, 6 * (size_t)i, nw, w);
#pragma unroll
      for (int k = 0; k < 6; k++) o.put(S(c, L.cdof + 6 * i + k));
    } else if (it >= n14) {
      if (it < n15) {          // cvel of body b = that of its frame
        const int b = it - n14, o_ = L.cvel + 6 * D.body_frame[b];
        Row o(cvo, 6 * (size_t)b, nw, w);
#pragma unroll
        for (int k = 0; k < 6; k++) o.put(S(c, o_ + k));
      } else if (it < n16) {
        const int i = it - n15;
        Row o(cdd, 6 * (size_t)i, nw, w);
#pragma unroll
        for (int k = 0; k < 6; k++) o.put(S(c, L.cdof_dot + 6 * i + k));
      } else {                 // dense qLD row: zeros, then the dof's ancestor chain (lower triangle of the L^T D L factor)
        const int i = it - n16;
        { Row o(qLDo, (size_t)i * nv, nw, w); for (int k = 0; k < nv; k++) o.put(0.f); }
        float* row = (float*)qLDo + (size_t)i * nv * nw + w;
        const int a = D.dof_Madr[i], d = D.dof_rowlen[i];
        for (int q = 0; q < d; q++) row[(size_t)D.M_col[a + q] * nw] = S(c, L.qLD + a + q);
      }
    } else {
      // mj_tendon's path points in MuJoCo's compacted order (team.cu tendons() has the same writer for MYO_WIDE_OUT=0):
      // entry index = static offset (every unit unwrapped) + 2 x wrapped geom units before the item (scratch gpre)
      auto putpt = [&](int i, float x, float y, float z, int obj) {
        if (wxp) { Row o(wxp, 3 * (size_t)i, nw, w); o.put(x); o.put(y); o.put(z); }
        if (wob) Row(wob, i, nw, w).puti(obj);
      };
      if (it < n11) {
        const int u = it - n10, go = D.unit_gord[u], g = D.unit_geom[u];
        const int pre = __float_as_int(S(c, L.gpre + go));
        const bool wrapped = g >= 0 && __float_as_int(S(c, L.gpre + go + 1)) != pre;
        const int off = D.unit_woff[u] + 2 * pre;
        float p[3], R[9];
        const int s0 = D.unit_s0[u];
        framePose(c, L, D.site_frame[s0], D.site_lpos + 3 * s0, nullptr, p, nullptr, R);
        putpt(off, p[0], p[1], p[2], -1);
        if (wrapped) {
          const int wp = L.uwpt + 6 * go;
          putpt(off + 1, S(c, wp), S(c, wp + 1), S(c, wp + 2), g);
          putpt(off + 2, S(c, wp + 3), S(c, wp + 4), S(c, wp + 5), g);
        }
        if (D.unit_end[u]) {
          const int s1 = D.unit_s1[u];
          framePose(c, L, D.site_frame[s1], D.site_lpos + 3 * s1, nullptr, p, nullptr, R);
          putpt(off + (wrapped ? 3 : 1), p[0], p[1], p[2], -1);
        }
      } else if (it < n12) {
        const int i = it - n11;
        putpt(D.pul_off[i] + 2 * __float_as_int(S(c, L.gpre + D.pul_gord[i])), 0.f, 0.f, 0.f, -2);
      } else if (it < n13) {
        const int k = it - n12;
        const int g0 = __float_as_int(S(c, L.gpre + D.ten_gord[k])), g1 = __float_as_int(S(c, L.gpre + D.ten_gord[k + 1]));
        if (wad) Row(wad, k, nw, w).puti(D.ten_woff[k] + 2 * g0);
        if (wnu) Row(wnu, k, nw, w).puti(D.ten_woff[k + 1] - D.ten_woff[k] + 2 * (g1 - g0));
      } else {   // entries a fuller wrap state of an earlier step can have written
        const int i = it - n13;
        if (i >= 2 * __float_as_int(S(c, L.gpre + D.ngunit))) putpt(D.ten_woff[D.nten] + i, 0.f, 0.f, 0.f, 0);
      }
    }
  }
}
}  // namespace

// fields this kernel writes (the step kernel is given null pointers for them when the wide path is on)
const int kWideOutFields[] = {MYO_F_SITE_XPOS, MYO_F_SITE_XMAT, MYO_F_GEOM_XPOS, MYO_F_GEOM_XMAT, MYO_F_XPOS, MYO_F_XQUAT, MYO_F_XMAT,
                              MYO_F_XIPOS, MYO_F_XIMAT, MYO_F_TEN_J, MYO_F_ACTUATOR_MOMENT, MYO_F_QM, MYO_F_EFC_J,
                              MYO_F_CONTACT_GEOM, MYO_F_CONTACT_DIM, MYO_F_CONTACT_EFC_ADDRESS, MYO_F_CONTACT_DIST, MYO_F_CONTACT_POS,
                              MYO_F_CONTACT_FRAME, MYO_F_CONTACT_INCLUDEMARGIN, MYO_F_CONTACT_FRICTION, MYO_F_CONTACT_SOLREF,
                              MYO_F_CONTACT_SOLIMP, MYO_F_EFC_POS, MYO_F_EFC_MARGIN, MYO_F_EFC_D, MYO_F_EFC_AREF, MYO_F_EFC_FORCE,
                              MYO_F_EFC_TYPE, MYO_F_XANCHOR, MYO_F_XAXIS, MYO_F_CDOF, MYO_F_WRAP_XPOS, MYO_F_WRAP_OBJ, MYO_F_TEN_WRAPADR, MYO_F_TEN_WRAPNUM,
                              MYO_F_CVEL, MYO_F_CDOF_DOT, MYO_F_QLD};
const int kNumWideOutFields = sizeof(kWideOutFields) / sizeof(int);

int myo_out_upload(int slot, const DevModel& D, const SmemLayout& L, const FieldPtrs& F, void* stream) {
  static OutParams staging[MYO_PARAM_SLOTS];   // must outlive the async copy
  if (slot < 0 || slot >= MYO_PARAM_SLOTS) return -1;
  OutParams& P = staging[slot];
  P.D = D; P.L = L;
  for (int f = 0; f < MYO_F_COUNT; f++) P.F.p[f] = nullptr;
  for (int i = 0; i < kNumWideOutFields; i++) P.F.p[kWideOutFields[i]] = F.p[kWideOutFields[i]];
  return cudaMemcpyToSymbolAsync(oP, &P, sizeof(OutParams), (size_t)slot * sizeof(OutParams), cudaMemcpyHostToDevice, (cudaStream_t)stream) == cudaSuccess ? 0 : -1;
}
int myo_out_launch(int slot, int nworld, int stride, const int* mask, int posvel, const float* scratch, void* stream) {
  dim3 grid((nworld + OUT_THREADS - 1) / OUT_THREADS, 1);
  // OUT_MINBLOCKS CTAs are resident per SM; many short CTAs (8 waves) let the hardware scheduler even out the item costs
  static const int want_env = getenv("MYO_OUT_BLOCKS") ? atoi(getenv("MYO_OUT_BLOCKS")) : 0;   // developer knob
  const int want = want_env > 0 ? want_env : 148 * OUT_MINBLOCKS * 8;
  grid.y = (unsigned)((want + grid.x - 1) / grid.x);
  if (grid.y < 1) grid.y = 1;
  if (grid.y > 512) grid.y = 512;
  cudaStream_t st = (cudaStream_t)stream;
  outKernel<<<grid, OUT_THREADS, 0, st>>>(slot, nworld, stride, mask, posvel, scratch);
  return cudaPeekAtLastError() == cudaSuccess ? 0 : -1;
}
