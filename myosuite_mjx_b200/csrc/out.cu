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

struct Ctx { const float* sm; size_t stride, w, nw; };
__device__ __forceinline__ float S(const Ctx& c, int k) { return c.sm[(size_t)k * c.stride]; }
#define OF(ptr, k) ((float*)(ptr))[(size_t)(k) * c.nw + c.w]

// world pose of an object fixed to frame f: p = x_f + R_f lpos, M = R_f lmat
__device__ __forceinline__ void framePose(const Ctx& c, const SmemLayout& L, int f, const float* lpos, const float* lmat, float* p, float* M, float* R) {
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

__global__ void __launch_bounds__(128, 8) outKernel(int slot, int nworld, int stride, const int* __restrict__ mask, int do_reset,
                                                 const float* __restrict__ scratch) {
  const DevModel& D = oP[slot].D; const SmemLayout& L = oP[slot].L; const FieldPtrs& F = oP[slot].F;
  Ctx c; c.w = (size_t)blockIdx.x * 128 + threadIdx.x; c.nw = (size_t)nworld; c.stride = (size_t)stride;
  if (c.w >= c.nw) return;
  if (do_reset && mask && !mask[c.w]) return;
  c.sm = scratch + (c.w / c.stride) * (size_t)L.total * c.stride + (c.w % c.stride);
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
  const int n0 = (sp || smt) ? D.nsite : 0, n1 = n0 + ((gp || gm) ? D.ngeom : 0),
            n2 = n1 + ((xpos || xquat || xmat || xipos || ximat) ? D.nbody : 0), n3 = n2 + (tj ? D.nten : 0), n4 = n3 + (am ? D.nu : 0),
            n5 = n4 + (qMo ? nv : 0), n6 = n5 + (eJ ? D.njmax : 0),
            n7 = n6 + ((cg || cdim || cadr || cdist || cpos || cfr || cinc || cfri || csr || csi) ? D.nconmax : 0),
            n8 = n7 + ((epos || emar || eD || earef || efrc || etyp) ? D.njmax : 0), n9 = n8 + ((xa || xx) ? D.njnt : 0),
            n10 = n9 + (cdo ? nv : 0);
  const int ncon = __float_as_int(S(c, L.counts));
  const int nefc = __float_as_int(S(c, L.counts + 1));   // left by the step kernel: ncon, nefc, niter
  for (int it = blockIdx.y; it < n10; it += gridDim.y) {
    if (it < n0) {
      const int q = it;
      float p[3], M[9], R[9];
      framePose(c, L, D.site_frame[q], D.site_lpos + 3 * q, D.site_lmat + 9 * q, p, smt ? M : nullptr, R);
      if (sp) { OF(sp, 3 * q) = p[0]; OF(sp, 3 * q + 1) = p[1]; OF(sp, 3 * q + 2) = p[2]; }
      if (smt)
#pragma unroll
        for (int k = 0; k < 9; k++) OF(smt, 9 * q + k) = M[k];
    } else if (it < n1) {
      const int g = it - n0;
      float p[3], M[9], R[9];
      framePose(c, L, D.geom_frame[g], D.geom_lpos + 3 * g, D.geom_lmat + 9 * g, p, gm ? M : nullptr, R);
      if (gp) { OF(gp, 3 * g) = p[0]; OF(gp, 3 * g + 1) = p[1]; OF(gp, 3 * g + 2) = p[2]; }
      if (gm)
#pragma unroll
        for (int k = 0; k < 9; k++) OF(gm, 9 * g + k) = M[k];
    } else if (it < n2) {
      const int b = it - n1;
      float p[3], M[9], R[9];
      const int f = D.body_frame[b];
      framePose(c, L, f, D.body_off_pos + 3 * b, D.body_off_mat + 9 * b, p, M, R);
      if (xpos) { OF(xpos, 3 * b) = p[0]; OF(xpos, 3 * b + 1) = p[1]; OF(xpos, 3 * b + 2) = p[2]; }
      if (xquat) {
        const int o = L.frames + 16 * f + 3;
        const float aw = S(c, o), ax = S(c, o + 1), ay = S(c, o + 2), az = S(c, o + 3);
        const float* bq = D.body_off_quat + 4 * b;
        float qw = aw * bq[0] - ax * bq[1] - ay * bq[2] - az * bq[3], qx = aw * bq[1] + ax * bq[0] + ay * bq[3] - az * bq[2],
              qy = aw * bq[2] - ax * bq[3] + ay * bq[0] + az * bq[1], qz = aw * bq[3] + ax * bq[2] - ay * bq[1] + az * bq[0];
        const float n = sqrtf(qw * qw + qx * qx + qy * qy + qz * qz);
        if (n < 1e-15f) { qw = 1.f; qx = qy = qz = 0.f; } else { const float s = 1.f / n; qw *= s; qx *= s; qy *= s; qz *= s; }
        OF(xquat, 4 * b) = qw; OF(xquat, 4 * b + 1) = qx; OF(xquat, 4 * b + 2) = qy; OF(xquat, 4 * b + 3) = qz;
      }
      if (xmat)
#pragma unroll
        for (int k = 0; k < 9; k++) OF(xmat, 9 * b + k) = M[k];
      if (xipos) {
        const float* ip = D.body_ipos + 3 * b;
#pragma unroll
        for (int i = 0; i < 3; i++) OF(xipos, 3 * b + i) = p[i] + fmaf(M[3 * i], ip[0], fmaf(M[3 * i + 1], ip[1], M[3 * i + 2] * ip[2]));
      }
      if (ximat) {
        const float* im = D.body_imat + 9 * b;
#pragma unroll
        for (int i = 0; i < 3; i++)
#pragma unroll
          for (int j = 0; j < 3; j++) OF(ximat, 9 * b + 3 * i + j) = fmaf(M[3 * i], im[j], fmaf(M[3 * i + 1], im[3 + j], M[3 * i + 2] * im[6 + j]));
      }
    } else if (it < n3) {   // dense ten_J row = zeros with the sparse entries (sorted by dof) merged in
      const int k = it - n2;
      int q = D.ten_nzstart[k];
      const int qend = D.ten_nzstart[k + 1];
      for (int i = 0; i < nv; i++) {
        float v = 0.f;
        if (q < qend && D.nz_dof[q] == i) { v = S(c, L.tenJ + q); q++; }
        OF(tj, k * nv + i) = v;
      }
    } else if (it < n4) {   // dense actuator_moment row = the tendon's sparse row scaled by the gear, or the joint's dof
      const int a = it - n3;
      const int id = D.act_trnid[a];
      const bool ten = D.act_trntype[a] == 3;
      int k = ten ? D.ten_nzstart[id] : 0;
      const int kend = ten ? D.ten_nzstart[id + 1] : 0;
      const int jd = ten ? -1 : D.jnt_dofadr[id];
      const float gear = D.act_gear[a];
      for (int i = 0; i < nv; i++) {
        float v = 0.f;
        if (k < kend && D.nz_dof[k] == i) { v = gear * S(c, L.tenJ + k); k++; }
        if (i == jd) v = gear;
        OF(am, a * nv + i) = v;
      }
    } else if (it < n5) {   // dense symmetric qM row: zeros, then the dof's ancestor chain and its descendants
      const int i = it - n4;
      for (int k = 0; k < nv; k++) OF(qMo, i * nv + k) = 0.f;
      const int a = D.dof_Madr[i], d = D.dof_rowlen[i];
      for (int q = 0; q < d; q++) OF(qMo, i * nv + D.M_col[a + q]) = S(c, L.qM + a + q);
      for (int e = D.desc_start[i]; e < D.desc_start[i + 1]; e++) OF(qMo, i * nv + D.desc_dof[e]) = S(c, L.qM + D.desc_idx[e]);
    } else if (it < n6) {   // dense efc_J row from the implied representation (see team.cu "constraints")
      const int r = it - n5;
      const bool on = r < nefc;
      const int njm = D.njmax;
      const int d0 = on ? __float_as_int(S(c, L.efc + E_DESC0 * njm + r)) : 0, type = d0 & 3, id = d0 >> 2;
      const float coef = on ? S(c, L.efc + E_COEF * njm + r) : 0.f;
      const int t1 = on && type == ROW_CON ? __float_as_int(S(c, L.efc + E_DESC1 * njm + r)) : -1;
      for (int k = 0; k < nv; k++) {
        float v = 0.f;
        if (on) {
          if (type == ROW_LIMJ) v = D.limj_dof[id] == k ? coef : 0.f;
          else if (type == ROW_EQ) v = D.eq_dof1[id] == k ? 1.f : (D.eq_dof2[id] == k ? coef : 0.f);
          else if (type == ROW_CON) v = S(c, L.conJ + id * nv + k) + (t1 >= 0 ? coef * S(c, L.conJ + t1 * nv + k) : 0.f);
          else { const int tn = D.limt_ten[id]; for (int q = D.ten_nzstart[tn]; q < D.ten_nzstart[tn + 1]; q++) if (D.nz_dof[q] == k) v = coef * S(c, L.tenJ + q); }
        }
        OF(eJ, r * nv + k) = v;
      }
    } else if (it < n7) {   // contact record (padded to nconmax)
      const int q = it - n6;
      const bool on = q < ncon;
      const int cr = L.con + 12 * q;
      const int pr = on ? __float_as_int(S(c, cr + 7)) : 0;
      const float* pp = D.pair_prm + 16 * pr;
      if (cg) { ((int*)cg)[(size_t)(2 * q) * c.nw + c.w] = on ? D.cg_geom[D.pair_cg1[pr]] : -1; ((int*)cg)[(size_t)(2 * q + 1) * c.nw + c.w] = on ? D.cg_geom[D.pair_cg2[pr]] : -1; }
      if (cdim) ((int*)cdim)[(size_t)q * c.nw + c.w] = on ? __float_as_int(S(c, cr + 9)) : 0;
      if (cadr) ((int*)cadr)[(size_t)q * c.nw + c.w] = on ? __float_as_int(S(c, cr + 8)) : -1;
      if (cdist) OF(cdist, q) = on ? S(c, cr) : 0.f;
      if (cpos) for (int k = 0; k < 3; k++) OF(cpos, 3 * q + k) = on ? S(c, cr + 1 + k) : 0.f;
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
        for (int k = 0; k < 3; k++) { OF(cfr, 9 * q + k) = n[k]; OF(cfr, 9 * q + 3 + k) = y[k]; OF(cfr, 9 * q + 6 + k) = z[k]; }
      }
      if (cinc) OF(cinc, q) = on ? pp[0] - pp[1] : 0.f;
      if (cfri) for (int k = 0; k < 5; k++) OF(cfri, 5 * q + k) = on ? pp[2 + k] : 0.f;
      if (csr) for (int k = 0; k < 2; k++) OF(csr, 2 * q + k) = on ? pp[7 + k] : 0.f;
      if (csi) for (int k = 0; k < 5; k++) OF(csi, 5 * q + k) = on ? pp[9 + k] : 0.f;
    } else if (it < n8) {   // per-row constraint scalars (padded to njmax)
      const int r = it - n7, njm = D.njmax;
      const bool on = r < nefc;
      if (epos) OF(epos, r) = on ? S(c, L.efc + E_POS * njm + r) : 0.f;
      if (emar) OF(emar, r) = on ? S(c, L.efc + E_MARGIN * njm + r) : 0.f;
      if (eD) OF(eD, r) = on ? S(c, L.efc + E_D * njm + r) : 0.f;
      if (earef) OF(earef, r) = on ? S(c, L.efc + E_AREF * njm + r) : 0.f;
      if (efrc) OF(efrc, r) = on ? S(c, L.efc + E_FORCE * njm + r) : 0.f;
      if (etyp) {
        const int d0 = on ? __float_as_int(S(c, L.efc + E_DESC0 * njm + r)) : 0, type = d0 & 3;
        ((int*)etyp)[(size_t)r * c.nw + c.w] = !on ? -1 : (type == ROW_EQ ? 0 : (type == ROW_LIMJ ? 3 : (type == ROW_LIMT ? 4 : (__float_as_int(S(c, L.efc + E_DESC1 * njm + r)) < 0 ? 5 : 6))));
      }
    } else if (it < n9) {
      const int j = it - n8;
      if (xa) for (int k = 0; k < 3; k++) OF(xa, 3 * j + k) = S(c, L.xanchor + 3 * j + k);
      if (xx) for (int k = 0; k < 3; k++) OF(xx, 3 * j + k) = S(c, L.xaxis + 3 * j + k);
    } else {
      const int i = it - n9;
      for (int k = 0; k < 6; k++) OF(cdo, 6 * i + k) = S(c, L.cdof + 6 * i + k);
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
                              MYO_F_EFC_TYPE, MYO_F_XANCHOR, MYO_F_XAXIS, MYO_F_CDOF};
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
int myo_out_launch(int slot, int nworld, int stride, const int* reset_mask, int do_reset, const float* scratch, void* stream) {
  dim3 grid((nworld + 127) / 128, 1);
  // 8 CTAs are resident per SM (64 registers); many short CTAs (8 waves) let the hardware scheduler even out the item costs
  static const int want_env = getenv("MYO_OUT_BLOCKS") ? atoi(getenv("MYO_OUT_BLOCKS")) : 0;   // developer knob
  const int want = want_env > 0 ? want_env : 148 * 8 * 8;
  grid.y = (unsigned)((want + grid.x - 1) / grid.x);
  if (grid.y < 1) grid.y = 1;
  if (grid.y > 512) grid.y = 512;
  outKernel<<<grid, 128, 0, (cudaStream_t)stream>>>(slot, nworld, stride, reset_mask, do_reset, scratch);
  return cudaPeekAtLastError() == cudaSuccess ? 0 : -1;
}
