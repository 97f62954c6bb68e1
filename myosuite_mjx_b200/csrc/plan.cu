// Host-side model compilation: myo_model_desc (mjModel-shaped, float64) -> DevModel tables on the GPU.
// Runs once per model inside myo_model_create (the mjx.put_model replacement; reference call site
// myosuite/mjx/play.py:10).  Nothing here runs per step.
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <map>
#include <set>
#include <string>
#include <vector>

#include "devmodel.h"
#include "internal.h"

namespace {

const double kMinVal = 1e-15;
enum { JNT_FREE = 0, JNT_BALL = 1, JNT_SLIDE = 2, JNT_HINGE = 3 };
enum { G_PLANE = 0, G_HFIELD, G_SPHERE, G_CAPSULE, G_ELLIPSOID, G_CYLINDER, G_BOX, G_MESH };
enum { W_PULLEY = 2, W_SITE = 3, W_SPHERE = 4, W_CYLINDER = 5 };
enum { DSBL_FILTERPARENT = 1 << 9, DSBL_REFSAFE = 1 << 11, DSBL_EULERDAMP = 1 << 14 };

struct Q { double w, x, y, z; };
struct P3 { double x, y, z; };
Q qmul(Q a, Q b) {
  return {a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z, a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y,
          a.w * b.y - a.x * b.z + a.y * b.w + a.z * b.x, a.w * b.z + a.x * b.y - a.y * b.x + a.z * b.w};
}
void q2m(Q q, double* R) {
  double w = q.w, x = q.x, y = q.y, z = q.z;
  R[0] = w * w + x * x - y * y - z * z; R[1] = 2 * (x * y - w * z); R[2] = 2 * (x * z + w * y);
  R[3] = 2 * (x * y + w * z); R[4] = w * w - x * x + y * y - z * z; R[5] = 2 * (y * z - w * x);
  R[6] = 2 * (x * z - w * y); R[7] = 2 * (y * z + w * x); R[8] = w * w - x * x - y * y + z * z;
}
P3 rot(Q q, P3 v) {
  double R[9]; q2m(q, R);
  return {R[0] * v.x + R[1] * v.y + R[2] * v.z, R[3] * v.x + R[4] * v.y + R[5] * v.z, R[6] * v.x + R[7] * v.y + R[8] * v.z};
}
Q qnorm(Q q) { double n = std::sqrt(q.w * q.w + q.x * q.x + q.y * q.y + q.z * q.z); return {q.w / n, q.x / n, q.y / n, q.z / n}; }
Q ldq(const double* p) { return {p[0], p[1], p[2], p[3]}; }
P3 ldp(const double* p) { return {p[0], p[1], p[2]}; }
P3 add(P3 a, P3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }

template <typename T> T* upload(const std::vector<T>& v, std::vector<void*>& owned) {
  void* d = nullptr;
  size_t n = std::max<size_t>(v.size(), 1) * sizeof(T);
#ifdef MYO_PLAN_HOST_ONLY   // tools/plan_stats: table sizes without a device (developer tool, never part of the library)
  d = malloc(n); if (!v.empty()) memcpy(d, v.data(), v.size() * sizeof(T)); owned.push_back(d); return (T*)d;
#endif
  if (cudaMalloc(&d, n) != cudaSuccess) throw std::string("cudaMalloc failed");
  if (!v.empty() && cudaMemcpy(d, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice) != cudaSuccess)
    throw std::string("cudaMemcpy failed");
  owned.push_back(d);
  return (T*)d;
}
void pushp(std::vector<float>& v, P3 p) { v.push_back((float)p.x); v.push_back((float)p.y); v.push_back((float)p.z); }
void pushq(std::vector<float>& v, Q q) { v.push_back((float)q.w); v.push_back((float)q.x); v.push_back((float)q.y); v.push_back((float)q.z); }

// K, B of a constraint row from solref (with the refsafe clamp) and sanitised solimp -> prm[0..6] = K,B,solimp[5]
void rowParams(const double* solref_in, const double* solimp_in, double timestep, bool refsafe, float* out) {
  double solref[2] = {solref_in[0], solref_in[1]}, si[5];
  for (int k = 0; k < 5; k++) si[k] = solimp_in[k];
  si[0] = std::min(0.9999, std::max(0.0001, si[0])); si[1] = std::min(0.9999, std::max(0.0001, si[1]));
  si[2] = std::max(0.0, si[2]); si[3] = std::min(0.9999, std::max(0.0001, si[3])); si[4] = std::max(1.0, si[4]);
  if (solref[0] > 0 && refsafe) solref[0] = std::max(solref[0], 2 * timestep);
  double K, B, dmax = si[1];
  if (solref[0] > 0) {
    K = 1 / std::max(kMinVal, dmax * dmax * solref[0] * solref[0] * solref[1] * solref[1]);
    B = 2 / std::max(kMinVal, dmax * solref[0]);
  } else {
    K = -solref[0] / std::max(kMinVal, dmax * dmax);
    B = -solref[1] / std::max(kMinVal, dmax);
  }
  out[0] = (float)K; out[1] = (float)B;
  for (int k = 0; k < 5; k++) out[2 + k] = (float)si[k];
}

}  // namespace

// Builds every table; throws std::string on unsupported models.
void myo_build_plan(const myo_model_desc& m, int nconmax_req, int njmax_req, DevModel& D, SmemLayout& L,
                    std::vector<void*>& owned) {
  memset(&D, 0, sizeof(D));
  const int nbody = m.nbody, nv = m.nv, njnt = m.njnt;
  if (m.opt_cone != 0) throw std::string("elliptic friction cones are not supported");
  if (m.opt_integrator != 0) throw std::string("only the Euler integrator is supported");
  if (nv > 64) throw std::string("nv > 64 is not supported");
  D.nq = m.nq; D.nv = nv; D.nu = m.nu; D.na = m.na; D.nbody = nbody; D.njnt = njnt; D.ngeom = m.ngeom; D.nsite = m.nsite;
  D.nten = m.ntendon; D.nwrap = m.nwrap; D.nM = m.nM;
  D.timestep = (float)m.opt_timestep;
  for (int k = 0; k < 3; k++) D.gravity[k] = (float)m.opt_gravity[k];
  D.tolerance = (float)m.opt_tolerance; D.ls_tolerance = (float)m.opt_ls_tolerance; D.meaninertia = (float)m.stat_meaninertia;
  D.impratio = (float)m.opt_impratio; D.iterations = m.opt_iterations; D.ls_iterations = m.opt_ls_iterations;
  D.disableflags = m.opt_disableflags;

  // ------------------------------------------------------------------ frames and constant offsets
  std::vector<int> body_frame(nbody, 0), fr_parent(1, 0), fr_body(1, 0), fr_jntadr(1, 0), fr_jntnum(1, 0), fr_dofadr(1, 0),
      fr_dofnum(1, 0), fr_level(1, 0), fr_root(1, 0);
  std::vector<P3> off_pos(nbody, P3{0, 0, 0}), pre_pos(1, P3{0, 0, 0});
  std::vector<Q> off_quat(nbody, Q{1, 0, 0, 0}), pre_quat(1, Q{1, 0, 0, 0});
  for (int b = 1; b < nbody; b++) {
    int p = m.body_parentid[b];
    P3 pos = add(off_pos[p], rot(off_quat[p], ldp(&m.body_pos[3 * b])));
    Q quat = qnorm(qmul(off_quat[p], ldq(&m.body_quat[4 * b])));
    if (m.body_jntnum[b] > 0) {
      int f = (int)fr_parent.size();
      for (int j = m.body_jntadr[b]; j < m.body_jntadr[b] + m.body_jntnum[b]; j++) {
        if (m.jnt_type[j] == JNT_BALL) throw std::string("ball joints are not supported");
        if (m.jnt_type[j] == JNT_FREE && m.body_jntnum[b] != 1) throw std::string("free joint must be alone in its body");
      }
      fr_parent.push_back(body_frame[p]); fr_body.push_back(b); fr_jntadr.push_back(m.body_jntadr[b]);
      fr_jntnum.push_back(m.body_jntnum[b]); fr_dofadr.push_back(m.body_dofadr[b]); fr_dofnum.push_back(m.body_dofnum[b]);
      fr_level.push_back(fr_level[body_frame[p]] + 1);
      pre_pos.push_back(pos); pre_quat.push_back(quat);
      body_frame[b] = f; off_pos[b] = {0, 0, 0}; off_quat[b] = {1, 0, 0, 0};
    } else {
      body_frame[b] = body_frame[p]; off_pos[b] = pos; off_quat[b] = quat;
    }
  }
  const int nframe = (int)fr_parent.size();
  D.nframe = nframe;
  std::vector<int> fr_subend(nframe), body_subend(nbody);
  for (int f = 0; f < nframe; f++) fr_subend[f] = f + 1;
  for (int f = nframe - 1; f >= 1; f--) if (fr_parent[f] > 0) fr_subend[fr_parent[f]] = std::max(fr_subend[fr_parent[f]], fr_subend[f]);
  fr_subend[0] = nframe;
  for (int b = 0; b < nbody; b++) body_subend[b] = b + 1;
  for (int b = nbody - 1; b >= 1; b--) { int p = m.body_parentid[b]; body_subend[p] = std::max(body_subend[p], body_subend[b]); }
  // levels
  int nlevel = 0;
  for (int f = 1; f < nframe; f++) nlevel = std::max(nlevel, fr_level[f]);
  D.nlevel = nlevel;
  std::vector<int> lvl_start(nlevel + 1, 0), lvl_frames;
  for (int l = 1; l <= nlevel; l++) {
    lvl_start[l - 1] = (int)lvl_frames.size();
    for (int f = 1; f < nframe; f++) if (fr_level[f] == l) lvl_frames.push_back(f);
  }
  lvl_start[nlevel] = (int)lvl_frames.size();
  // roots (children of the world), constant part of each tree's centre of mass
  std::vector<int> root_body, body_root(nbody, 0), root_fstart, root_fend;
  std::map<int, int> rootidx;
  for (int b = 1; b < nbody; b++) {
    int r = m.body_rootid[b];
    if (!rootidx.count(r)) { rootidx[r] = (int)root_body.size(); root_body.push_back(r); }
    body_root[b] = rootidx[r];
  }
  const int nroot = std::max<int>(1, (int)root_body.size());
  if (root_body.empty()) root_body.push_back(0);
  D.nroot = nroot;
  std::vector<float> root_const(4 * nroot, 0.f);
  std::vector<double> rc(4 * nroot, 0.0);
  for (int b = 1; b < nbody; b++) {
    int r = body_root[b];
    if (body_frame[b] == 0) {
      P3 c = add(off_pos[b], rot(off_quat[b], ldp(&m.body_ipos[3 * b])));
      rc[4 * r] += m.body_mass[b] * c.x; rc[4 * r + 1] += m.body_mass[b] * c.y; rc[4 * r + 2] += m.body_mass[b] * c.z;
    }
    rc[4 * r + 3] += m.body_mass[b];
  }
  for (int i = 0; i < 4 * nroot; i++) root_const[i] = (float)rc[i];
  root_fstart.assign(nroot, 0); root_fend.assign(nroot, 0);
  for (int f = 1; f < nframe; f++) {
    int r = body_root[fr_body[f]];
    fr_root.push_back(r);
    if (root_fend[r] == 0) root_fstart[r] = f;
    root_fend[r] = f + 1;
  }
  // composite inertia of each weld group in frame-local coordinates
  std::vector<float> fr_mass(nframe, 0.f), fr_com(3 * nframe, 0.f), fr_inertia(6 * nframe, 0.f);
  for (int f = 1; f < nframe; f++) {
    double M = 0, c[3] = {0, 0, 0};
    for (int b = 1; b < nbody; b++) if (body_frame[b] == f) {
      P3 cb = add(off_pos[b], rot(off_quat[b], ldp(&m.body_ipos[3 * b])));
      M += m.body_mass[b]; c[0] += m.body_mass[b] * cb.x; c[1] += m.body_mass[b] * cb.y; c[2] += m.body_mass[b] * cb.z;
    }
    if (M > kMinVal) for (int k = 0; k < 3; k++) c[k] /= M;
    double I[9] = {0};
    for (int b = 1; b < nbody; b++) if (body_frame[b] == f) {
      P3 cb = add(off_pos[b], rot(off_quat[b], ldp(&m.body_ipos[3 * b])));
      double R[9]; q2m(qnorm(qmul(off_quat[b], ldq(&m.body_iquat[4 * b]))), R);
      double d[3] = {cb.x - c[0], cb.y - c[1], cb.z - c[2]}, mb = m.body_mass[b];
      double dd = d[0] * d[0] + d[1] * d[1] + d[2] * d[2];
      for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) {
        double s = 0;
        for (int k = 0; k < 3; k++) s += R[3 * i + k] * m.body_inertia[3 * b + k] * R[3 * j + k];
        I[3 * i + j] += s + mb * ((i == j ? dd : 0.0) - d[i] * d[j]);
      }
    }
    fr_mass[f] = (float)M;
    for (int k = 0; k < 3; k++) fr_com[3 * f + k] = (float)c[k];
    fr_inertia[6 * f] = (float)I[0]; fr_inertia[6 * f + 1] = (float)I[4]; fr_inertia[6 * f + 2] = (float)I[8];
    fr_inertia[6 * f + 3] = (float)I[1]; fr_inertia[6 * f + 4] = (float)I[2]; fr_inertia[6 * f + 5] = (float)I[5];
  }
  std::vector<int> fr_mask(2 * nframe, 0);  // ancestor-dof bit masks (filled once the dof chains are known)
  std::vector<float> fpp, fpq;
  for (int f = 0; f < nframe; f++) { pushp(fpp, pre_pos[f]); pushq(fpq, pre_quat[f]); }
  D.fr_parent = upload(fr_parent, owned); D.fr_body = upload(fr_body, owned); D.fr_jntadr = upload(fr_jntadr, owned);
  D.fr_jntnum = upload(fr_jntnum, owned); D.fr_dofadr = upload(fr_dofadr, owned); D.fr_dofnum = upload(fr_dofnum, owned);
  D.fr_subend = upload(fr_subend, owned); D.fr_root = upload(fr_root, owned); D.fr_pre_pos = upload(fpp, owned);
  D.fr_pre_quat = upload(fpq, owned); D.fr_mass = upload(fr_mass, owned); D.fr_com = upload(fr_com, owned);
  D.fr_inertia = upload(fr_inertia, owned); D.lvl_start = upload(lvl_start, owned); D.lvl_frames = upload(lvl_frames, owned);
  D.root_const = upload(root_const, owned); D.root_body = upload(root_body, owned);
  D.root_fstart = upload(root_fstart, owned); D.root_fend = upload(root_fend, owned);

  // ------------------------------------------------------------------ joints, dofs, sparse inertia pattern
  {
    std::vector<int> jt(m.jnt_type, m.jnt_type + njnt), jq(m.jnt_qposadr, m.jnt_qposadr + njnt), jd(m.jnt_dofadr, m.jnt_dofadr + njnt), jf(njnt);
    std::vector<float> jp(3 * njnt), ja(3 * njnt), q0(m.nq), jstiff(njnt), qspring(m.nq);
    for (int j = 0; j < njnt; j++) jf[j] = body_frame[m.jnt_bodyid[j]];
    for (int i = 0; i < 3 * njnt; i++) { jp[i] = (float)m.jnt_pos[i]; ja[i] = (float)m.jnt_axis[i]; }
    for (int i = 0; i < m.nq; i++) { q0[i] = (float)m.qpos0[i]; qspring[i] = (float)m.qpos_spring[i]; }
    for (int j = 0; j < njnt; j++) jstiff[j] = (float)m.jnt_stiffness[j];
    D.jnt_type = upload(jt, owned); D.jnt_qposadr = upload(jq, owned); D.jnt_dofadr = upload(jd, owned); D.jnt_frame = upload(jf, owned);
    D.jnt_pos = upload(jp, owned); D.jnt_axis = upload(ja, owned); D.qpos0 = upload(q0, owned);
    D.jnt_stiffness = upload(jstiff, owned); D.qpos_spring = upload(qspring, owned);
  }
  std::vector<int> dof_rowlen(nv), dof_depth(nv);
  std::vector<std::vector<int>> chain(nv);  // self first, then ancestors
  {
    std::vector<int> dframe(nv), djnt(m.dof_jntid, m.dof_jntid + nv), dadr(m.dof_Madr, m.dof_Madr + nv), dpar(m.dof_parentid, m.dof_parentid + nv),
        Mrow, Mcol, prevc(nv);
    std::vector<float> arm(nv), damp(nv), invw(nv);
    int any = 0;
    for (int i = 0; i < nv; i++) {
      dframe[i] = body_frame[m.dof_bodyid[i]];
      for (int j = i; j >= 0; j = m.dof_parentid[j]) { chain[i].push_back(j); Mrow.push_back(i); Mcol.push_back(j); }
      dof_rowlen[i] = (int)chain[i].size(); dof_depth[i] = dof_rowlen[i] - 1;
      arm[i] = (float)m.dof_armature[i]; damp[i] = (float)m.dof_damping[i]; invw[i] = (float)m.dof_invweight0[i];
      if (m.dof_damping[i] > 0) any = 1;
      int j = m.dof_jntid[i], k = i - m.jnt_dofadr[j];
      prevc[i] = (m.jnt_type[j] == JNT_FREE) ? (k < 3 ? -1 : k - 3 + 1) : 1;
    }
    if ((int)Mrow.size() != m.nM) throw std::string("nM mismatch");
    D.any_damping = any && !(m.opt_disableflags & DSBL_EULERDAMP);
    D.dof_frame = upload(dframe, owned); D.dof_jnt = upload(djnt, owned); D.dof_Madr = upload(dadr, owned); D.dof_parent = upload(dpar, owned);
    D.dof_rowlen = upload(dof_rowlen, owned); D.dof_armature = upload(arm, owned); D.dof_damping = upload(damp, owned);
    D.dof_invweight0 = upload(invw, owned); D.M_row = upload(Mrow, owned); D.M_col = upload(Mcol, owned); D.dof_prevcount = upload(prevc, owned);
    // LDL ops, columns nv-1 .. 0
    std::vector<int> ls(nv + 1, 0), ld, la, lb;
    for (int pos = 0; pos < nv; pos++) {
      int k = nv - 1 - pos, d = dof_rowlen[k], akk = m.dof_Madr[k];
      ls[pos] = (int)ld.size();
      for (int s = 1; s < d; s++)
        for (int p = 0; p + s < d; p++) { ld.push_back(m.dof_Madr[chain[k][s]] + p); la.push_back(akk + s); lb.push_back(akk + s + p); }
    }
    ls[nv] = (int)ld.size();
    for (int f = 1; f < nframe; f++)
      for (int i : chain[fr_dofadr[f] + fr_dofnum[f] - 1]) fr_mask[2 * f + (i >> 5)] |= (int)(1u << (i & 31));
    D.fr_mask = upload(fr_mask, owned);
    D.nldlop = (int)ld.size();
    D.ldl_start = upload(ls, owned); D.ldl_dst = upload(ld, owned); D.ldl_a = upload(la, owned); D.ldl_b = upload(lb, owned);
    // solve plan
    int ndl = 0;
    for (int i = 0; i < nv; i++) ndl = std::max(ndl, dof_depth[i] + 1);
    D.ndlevel = ndl;
    std::vector<int> dls(ndl + 1, 0), dld, ds(nv + 1, 0), dd, di;
    for (int l = 0; l < ndl; l++) { dls[l] = (int)dld.size(); for (int i = 0; i < nv; i++) if (dof_depth[i] == l) dld.push_back(i); }
    dls[ndl] = (int)dld.size();
    for (int j = 0; j < nv; j++) {
      ds[j] = (int)dd.size();
      for (int i = j + 1; i < nv; i++)
        for (int s = 1; s < dof_rowlen[i]; s++) if (chain[i][s] == j) { dd.push_back(i); di.push_back(m.dof_Madr[i] + s); }
    }
    ds[nv] = (int)dd.size();
    D.dlvl_start = upload(dls, owned); D.dlvl_dofs = upload(dld, owned); D.desc_start = upload(ds, owned);
    D.desc_dof = upload(dd, owned); D.desc_idx = upload(di, owned);
  }

  // ------------------------------------------------------------------ bodies, geoms, sites
  {
    std::vector<float> bop, boq, bip(3 * nbody), biq(4 * nbody), bm(nbody), bin(3 * nbody), bsm(nbody);
    std::vector<int> bpar(m.body_parentid, m.body_parentid + nbody);
    for (int b = 0; b < nbody; b++) {
      pushp(bop, off_pos[b]); pushq(boq, off_quat[b]);
      for (int k = 0; k < 3; k++) { bip[3 * b + k] = (float)m.body_ipos[3 * b + k]; bin[3 * b + k] = (float)m.body_inertia[3 * b + k]; }
      for (int k = 0; k < 4; k++) biq[4 * b + k] = (float)m.body_iquat[4 * b + k];
      bm[b] = (float)m.body_mass[b]; bsm[b] = (float)m.body_subtreemass[b];
    }
    D.body_frame = upload(body_frame, owned); D.body_root = upload(body_root, owned); D.body_subend = upload(body_subend, owned);
    D.body_parent = upload(bpar, owned); D.body_off_pos = upload(bop, owned); D.body_off_quat = upload(boq, owned);
    D.body_ipos = upload(bip, owned); D.body_iquat = upload(biq, owned); D.body_mass = upload(bm, owned);
    D.body_inertia = upload(bin, owned); D.body_subtreemass = upload(bsm, owned);
    std::vector<float> bom, bim, glm, slm;
    auto pushm = [](std::vector<float>& v, Q q) { double R[9]; q2m(q, R); for (int k = 0; k < 9; k++) v.push_back((float)R[k]); };
    for (int b = 0; b < nbody; b++) { pushm(bom, off_quat[b]); pushm(bim, qnorm(ldq(&m.body_iquat[4 * b]))); }
    for (int g = 0; g < m.ngeom; g++) pushm(glm, qnorm(qmul(off_quat[m.geom_bodyid[g]], ldq(&m.geom_quat[4 * g]))));
    for (int s = 0; s < m.nsite; s++) pushm(slm, qnorm(qmul(off_quat[m.site_bodyid[s]], ldq(&m.site_quat[4 * s]))));
    D.body_off_mat = upload(bom, owned); D.body_imat = upload(bim, owned); D.geom_lmat = upload(glm, owned); D.site_lmat = upload(slm, owned);
    std::vector<int> gf(m.ngeom), sf(m.nsite);
    std::vector<float> gp, gq, sp, sq;
    for (int g = 0; g < m.ngeom; g++) {
      int b = m.geom_bodyid[g]; gf[g] = body_frame[b];
      pushp(gp, add(off_pos[b], rot(off_quat[b], ldp(&m.geom_pos[3 * g])))); pushq(gq, qnorm(qmul(off_quat[b], ldq(&m.geom_quat[4 * g]))));
    }
    for (int s = 0; s < m.nsite; s++) {
      int b = m.site_bodyid[s]; sf[s] = body_frame[b];
      pushp(sp, add(off_pos[b], rot(off_quat[b], ldp(&m.site_pos[3 * s])))); pushq(sq, qnorm(qmul(off_quat[b], ldq(&m.site_quat[4 * s]))));
    }
    { std::vector<int> gb(m.geom_bodyid, m.geom_bodyid + m.ngeom); D.geom_body = upload(gb, owned); }
    D.geom_frame = upload(gf, owned); D.site_frame = upload(sf, owned); D.geom_lpos = upload(gp, owned); D.geom_lquat = upload(gq, owned);
    D.site_lpos = upload(sp, owned); D.site_lquat = upload(sq, owned);
  }

  // ------------------------------------------------------------------ tendons: units, slots, moment-arm pattern
  auto ancset = [&](int b) {
    std::set<int> s;
    int f = body_frame[b];
    if (f == 0) return s;
    for (int i : chain[fr_dofadr[f] + fr_dofnum[f] - 1]) s.insert(i);
    return s;
  };
  struct Ent { int tendon, dof, unit, slot, kind; float scale; };
  std::vector<Ent> ents;
  std::vector<int> u_t, u_type, u_s0, u_s1, u_g, u_side, u_slot, ten_ustart(m.ntendon + 1, 0);
  std::vector<float> u_div, u_rad;
  // mjData path points (ten_wrapadr / ten_wrapnum / wrap_obj / wrap_xpos): a unit's first point sits at its STATIC offset
  // (every unit unwrapped: one point per unit, one per branch end, one per pulley marker) plus two for every unit before
  // it that currently wraps its geom.  *_gord = number of geom units before the item.
  std::vector<int> u_woff, u_gord, u_end, gunit, pul_off, pul_gord, ten_woff(m.ntendon + 1, 0), ten_gord(m.ntendon + 1, 0);
  int wstat = 0;
  int nslot = 0;
  for (int t = 0; t < m.ntendon; t++) {
    ten_ustart[t] = (int)u_t.size();
    ten_woff[t] = wstat; ten_gord[t] = (int)gunit.size();
    int adr = m.tendon_adr[t], num = m.tendon_num[t], k = 0;
    double div = 1;
    while (k < num - 1) {
      int t0 = m.wrap_type[adr + k], t1 = m.wrap_type[adr + k + 1];
      if (t0 == W_PULLEY) { div = m.wrap_prm[adr + k]; pul_off.push_back(wstat++); pul_gord.push_back((int)gunit.size()); k++; continue; }
      if (t1 == W_PULLEY) { k++; continue; }
      if (t0 != W_SITE) throw std::string("tendon path must alternate sites and wrap geoms");
      int u = (int)u_t.size(), s0 = m.wrap_objid[adr + k], s1, g = -1, side = -1, type = UNIT_SITE;
      if (t1 == W_SITE) { s1 = m.wrap_objid[adr + k + 1]; k += 1; }
      else if (t1 == W_SPHERE || t1 == W_CYLINDER) {
        if (k + 2 >= num || m.wrap_type[adr + k + 2] != W_SITE) throw std::string("wrap geom must be followed by a site");
        g = m.wrap_objid[adr + k + 1]; s1 = m.wrap_objid[adr + k + 2]; side = (int)std::lround(m.wrap_prm[adr + k + 1]); type = t1; k += 2;
      } else throw std::string("unsupported tendon wrap type (only spatial tendons)");
      int b0 = m.site_bodyid[s0], b1 = m.site_bodyid[s1], bg = g >= 0 ? m.geom_bodyid[g] : -1;
      int slot0 = -1, slot1 = -1;
      auto addseg = [&](int ba, int bb, int kind, int which) {
        std::set<int> A = ancset(ba), B = ancset(bb);
        std::vector<std::pair<int, float>> dl;
        for (int i : B) if (!A.count(i)) dl.push_back({i, (float)(1.0 / div)});
        for (int i : A) if (!B.count(i)) dl.push_back({i, (float)(-1.0 / div)});
        if (dl.empty()) return;
        int& slot = which == 0 ? slot0 : slot1;
        if (slot < 0) slot = nslot++;
        for (auto& e : dl) ents.push_back({t, e.first, u, slot, kind, e.second});
      };
      if (g < 0) addseg(b0, b1, ENT_ALWAYS, 0);
      else { addseg(b0, b1, ENT_NOWRAP, 0); addseg(b0, bg, ENT_WRAP, 0); addseg(bg, b1, ENT_WRAP, 1); }
      u_t.push_back(t); u_type.push_back(type); u_s0.push_back(s0); u_s1.push_back(s1); u_g.push_back(g); u_side.push_back(side);
      u_slot.push_back(slot0); u_slot.push_back(slot1); u_div.push_back((float)div); u_rad.push_back(g >= 0 ? (float)m.geom_size[3 * g] : 0.f);
      u_woff.push_back(wstat++); u_gord.push_back((int)gunit.size());
      if (g >= 0) gunit.push_back(u);
      const bool end = k == num - 1 || m.wrap_type[adr + k + 1] == W_PULLEY;   // last site of the tendon or of a pulley branch
      u_end.push_back(end ? 1 : 0);
      if (end) wstat++;
    }
  }
  ten_ustart[m.ntendon] = (int)u_t.size();
  ten_woff[m.ntendon] = wstat; ten_gord[m.ntendon] = (int)gunit.size();
  D.ngunit = (int)gunit.size(); D.npulley = (int)pul_off.size();
  if (wstat + 2 * D.ngunit > 2 * m.nwrap) throw std::string("tendon path points exceed 2 * nwrap");
  D.unit_woff = upload(u_woff, owned); D.unit_gord = upload(u_gord, owned); D.unit_end = upload(u_end, owned); D.gunit = upload(gunit, owned);
  D.pul_off = upload(pul_off, owned); D.pul_gord = upload(pul_gord, owned); D.ten_woff = upload(ten_woff, owned); D.ten_gord = upload(ten_gord, owned);
  D.nunit = (int)u_t.size(); D.nslot = nslot; D.nentry = (int)ents.size();
  std::stable_sort(ents.begin(), ents.end(), [](const Ent& a, const Ent& b) { return a.tendon != b.tendon ? a.tendon < b.tendon : a.dof < b.dof; });
  std::vector<int> e_slot, e_dof, e_kind, e_unit, nz_estart, nz_t, nz_d, ten_nzstart(m.ntendon + 1, 0);
  std::vector<float> e_scale;
  for (size_t e = 0; e < ents.size(); e++) {
    if (e == 0 || ents[e].tendon != ents[e - 1].tendon || ents[e].dof != ents[e - 1].dof) { nz_estart.push_back((int)e); nz_t.push_back(ents[e].tendon); nz_d.push_back(ents[e].dof); }
    e_slot.push_back(ents[e].slot); e_dof.push_back(ents[e].dof); e_kind.push_back(ents[e].kind); e_unit.push_back(ents[e].unit); e_scale.push_back(ents[e].scale);
  }
  const int nnz = (int)nz_t.size();
  nz_estart.push_back((int)ents.size());
  D.nnz = nnz;
  { int n = 0; for (int t = 0; t <= m.ntendon; t++) { while (n < nnz && nz_t[n] < t) n++; ten_nzstart[t] = n; } }
  std::vector<int> col_start(nv + 1, 0), col_nz;
  for (int i = 0; i < nv; i++) { col_start[i] = (int)col_nz.size(); for (int n = 0; n < nnz; n++) if (nz_d[n] == i) col_nz.push_back(n); }
  col_start[nv] = (int)col_nz.size();
  D.unit_tendon = upload(u_t, owned); D.unit_type = upload(u_type, owned); D.unit_s0 = upload(u_s0, owned); D.unit_s1 = upload(u_s1, owned);
  {  // processing order: site-site units first, wrapping units last, so the lanes of a team run the wrap code together
    std::vector<int> order;
    for (size_t u = 0; u < u_g.size(); u++) if (u_g[u] < 0) order.push_back((int)u);
    for (int ty : {(int)UNIT_SPHERE, (int)UNIT_CYLINDER})
      for (size_t u = 0; u < u_g.size(); u++) if (u_g[u] >= 0 && u_type[u] == ty) order.push_back((int)u);
    for (size_t u = 0; u < u_g.size(); u++) if (u_g[u] >= 0 && u_type[u] != UNIT_SPHERE && u_type[u] != UNIT_CYLINDER) order.push_back((int)u);
    D.unit_order = upload(order, owned);
  }
  D.unit_geom = upload(u_g, owned); D.unit_side = upload(u_side, owned); D.unit_slot = upload(u_slot, owned); D.unit_div = upload(u_div, owned);
  D.unit_radius = upload(u_rad, owned); D.ten_ustart = upload(ten_ustart, owned); D.ent_slot = upload(e_slot, owned); D.ent_dof = upload(e_dof, owned);
  D.ent_kind = upload(e_kind, owned); D.ent_unit = upload(e_unit, owned); D.ent_scale = upload(e_scale, owned); D.nz_estart = upload(nz_estart, owned);
  D.nz_tendon = upload(nz_t, owned); D.nz_dof = upload(nz_d, owned); D.ten_nzstart = upload(ten_nzstart, owned);
  D.col_start = upload(col_start, owned); D.col_nz = upload(col_nz, owned);
  {
    std::vector<float> tr(2 * m.ntendon), ts(m.ntendon), td(m.ntendon), tl(2 * m.ntendon);
    for (int t = 0; t < m.ntendon; t++) {
      tr[2 * t] = (float)m.tendon_range[2 * t]; tr[2 * t + 1] = (float)m.tendon_range[2 * t + 1]; ts[t] = (float)m.tendon_stiffness[t];
      td[t] = (float)m.tendon_damping[t]; tl[2 * t] = (float)m.tendon_lengthspring[2 * t]; tl[2 * t + 1] = (float)m.tendon_lengthspring[2 * t + 1];
    }
    D.ten_range = upload(tr, owned); D.ten_stiffness = upload(ts, owned); D.ten_damping = upload(td, owned); D.ten_lengthspring = upload(tl, owned);
  }

  // ------------------------------------------------------------------ actuators
  {
    const int nu = m.nu;
    std::vector<int> tt(nu), ti(nu), dt(nu), gt(nu), bt(nu), fl(nu), aa(nu);
    std::vector<float> gear(nu), dp(3 * nu), gp(9 * nu), bp(9 * nu), cr(2 * nu), fr(2 * nu), ar(2 * nu), lr(2 * nu), acc0(nu);
    for (int i = 0; i < nu; i++) {
      tt[i] = m.actuator_trntype[i]; ti[i] = m.actuator_trnid[2 * i];
      if (tt[i] == 3) { /* tendon */ } else if (tt[i] == 0) {
        int j = ti[i];
        if (m.jnt_type[j] != JNT_HINGE && m.jnt_type[j] != JNT_SLIDE) throw std::string("joint transmission needs a hinge or slide joint");
      } else throw std::string("unsupported actuator transmission");
      dt[i] = m.actuator_dyntype[i]; gt[i] = m.actuator_gaintype[i]; bt[i] = m.actuator_biastype[i]; aa[i] = m.actuator_actadr[i];
      fl[i] = (m.actuator_ctrllimited[i] ? ACT_CTRLLIMITED : 0) | (m.actuator_forcelimited[i] ? ACT_FORCELIMITED : 0) | (m.actuator_actlimited[i] ? ACT_ACTLIMITED : 0);
      gear[i] = (float)m.actuator_gear[6 * i]; acc0[i] = (float)m.actuator_acc0[i];
      for (int k = 0; k < 3; k++) dp[3 * i + k] = (float)m.actuator_dynprm[MYO_NPRM * i + k];
      for (int k = 0; k < 9; k++) { gp[9 * i + k] = (float)m.actuator_gainprm[MYO_NPRM * i + k]; bp[9 * i + k] = (float)m.actuator_biasprm[MYO_NPRM * i + k]; }
      for (int k = 0; k < 2; k++) {
        cr[2 * i + k] = (float)m.actuator_ctrlrange[2 * i + k]; fr[2 * i + k] = (float)m.actuator_forcerange[2 * i + k];
        ar[2 * i + k] = (float)m.actuator_actrange[2 * i + k]; lr[2 * i + k] = (float)m.actuator_lengthrange[2 * i + k];
      }
    }
    std::vector<int> same(nu, 0);
    for (int i = 0; i < nu; i++) { same[i] = 1; for (int k = 0; k < 9; k++) if (gp[9 * i + k] != bp[9 * i + k]) same[i] = 0; }
    D.act_same_prm = upload(same, owned);
    D.act_trntype = upload(tt, owned); D.act_trnid = upload(ti, owned); D.act_dyntype = upload(dt, owned); D.act_gaintype = upload(gt, owned);
    {   // dense-row source maps of the tendon output kernel: entry (row, dof) -> index of the sparse moment-arm entry, -1 = zero,
        // -2 = the gear itself (joint transmission): ten_J [nten][nv], actuator_moment [nu][nv]
      std::vector<int> td((size_t)m.ntendon * nv, -1), ad((size_t)nu * nv, -1);
      for (int t = 0; t < m.ntendon; t++) for (int n = ten_nzstart[t]; n < ten_nzstart[t + 1]; n++) td[(size_t)t * nv + nz_d[n]] = n;
      for (int i = 0; i < nu; i++) {
        if (tt[i] == 3) { const int t = ti[i]; for (int n = ten_nzstart[t]; n < ten_nzstart[t + 1]; n++) ad[(size_t)i * nv + nz_d[n]] = n; }
        else ad[(size_t)i * nv + m.jnt_dofadr[ti[i]]] = -2;
      }
      D.ten_dense = upload(td, owned); D.act_dense = upload(ad, owned);
    }
    D.act_biastype = upload(bt, owned); D.act_flags = upload(fl, owned); D.act_actadr = upload(aa, owned); D.act_gear = upload(gear, owned);
    D.act_dynprm = upload(dp, owned); D.act_gainprm = upload(gp, owned); D.act_biasprm = upload(bp, owned); D.act_ctrlrange = upload(cr, owned);
    std::vector<int> tas(m.ntendon + 1, 0), tal, das(nv + 1, 0), dal;
    for (int t = 0; t < m.ntendon; t++) { tas[t] = (int)tal.size(); for (int i = 0; i < nu; i++) if (tt[i] == 3 && ti[i] == t) tal.push_back(i); }
    tas[m.ntendon] = (int)tal.size();
    for (int d = 0; d < nv; d++) { das[d] = (int)dal.size(); for (int i = 0; i < nu; i++) if (tt[i] == 0 && m.jnt_dofadr[ti[i]] == d) dal.push_back(i); }
    das[nv] = (int)dal.size();
    D.tact_start = upload(tas, owned); D.tact_list = upload(tal, owned); D.dact_start = upload(das, owned); D.dact_list = upload(dal, owned);
    D.any_tendon_passive = 0;
    for (int t = 0; t < m.ntendon; t++) if (m.tendon_stiffness[t] > 0 || m.tendon_damping[t] > 0) D.any_tendon_passive = 1;
    D.act_forcerange = upload(fr, owned); D.act_actrange = upload(ar, owned); D.act_lengthrange = upload(lr, owned); D.act_acc0 = upload(acc0, owned);
  }

  // ------------------------------------------------------------------ collision candidates (static filters, MuJoCo order)
  int maxbase = 1;
  {
    std::map<int, int> cgidx;
    std::vector<int> cg_geom, cg_type, cg_frame, cg_body, cg_meshadr, cg_meshnum, p1, p2, pdim;
    std::vector<float> cg_lpos, cg_lmat, cg_size, cg_rbound, cg_invw, pprm;
    auto cg = [&](int g) {
      auto it = cgidx.find(g);
      if (it != cgidx.end()) return it->second;
      int id = (int)cg_geom.size(), b = m.geom_bodyid[g];
      cgidx[g] = id; cg_geom.push_back(g); cg_type.push_back(m.geom_type[g]); cg_frame.push_back(body_frame[b]); cg_body.push_back(b);
      pushp(cg_lpos, add(off_pos[b], rot(off_quat[b], ldp(&m.geom_pos[3 * g]))));
      double R[9]; q2m(qnorm(qmul(off_quat[b], ldq(&m.geom_quat[4 * g]))), R);
      for (int k = 0; k < 9; k++) cg_lmat.push_back((float)R[k]);
      for (int k = 0; k < 3; k++) cg_size.push_back((float)m.geom_size[3 * g + k]);
      cg_rbound.push_back((float)m.geom_rbound[g]);
      const int mesh = m.geom_type[g] == G_MESH ? m.geom_dataid[g] : -1;
      if (m.geom_type[g] == G_MESH && (mesh < 0 || mesh >= m.nmesh || m.mesh_vertnum[mesh] < 4)) throw std::string("mesh collision geom without a convex hull (geom_dataid / mesh_vert)");
      cg_meshadr.push_back(mesh >= 0 ? m.mesh_vertadr[mesh] : 0); cg_meshnum.push_back(mesh >= 0 ? m.mesh_vertnum[mesh] : 0);
      cg_invw.push_back((float)m.body_invweight0[2 * b]); cg_invw.push_back((float)m.body_invweight0[2 * b + 1]);
      return id;
    };
    auto addpair = [&](int g1, int g2, int dim, const double* fr5, const double* solref, const double* solimp, double margin, double gap) {
      int t1 = m.geom_type[g1], t2 = m.geom_type[g2];
      if (t1 > t2) { std::swap(g1, g2); std::swap(t1, t2); }
      if (t1 == G_HFIELD || t2 == G_HFIELD) return;  // height fields: not supported, pair dropped
      if (t1 == G_PLANE && t2 == G_PLANE) return;
      if (dim != 1 && dim != 3 && dim != 4) throw std::string("contact condim must be 1, 3 or 4");
      maxbase = std::max(maxbase, dim);
      p1.push_back(cg(g1)); p2.push_back(cg(g2)); pdim.push_back(dim);
      pprm.push_back((float)margin); pprm.push_back((float)gap);
      for (int k = 0; k < 5; k++) pprm.push_back((float)fr5[k]);
      pprm.push_back((float)solref[0]); pprm.push_back((float)solref[1]);
      float kb[8]; rowParams(solref, solimp, m.opt_timestep, !(m.opt_disableflags & DSBL_REFSAFE), kb);
      for (int k = 0; k < 5; k++) pprm.push_back(kb[2 + k]);   // sanitised solimp
      pprm.push_back(kb[0]); pprm.push_back(kb[1]);            // K, B
    };
    auto mixpair = [&](int g1, int g2) {
      double margin = std::max(m.geom_margin[g1], m.geom_margin[g2]), gap = std::max(m.geom_gap[g1], m.geom_gap[g2]);
      int pr1 = m.geom_priority[g1], pr2 = m.geom_priority[g2], dim;
      double f[3], solref[2], solimp[5];
      if (pr1 != pr2) {
        int g = pr1 > pr2 ? g1 : g2;
        dim = m.geom_condim[g];
        for (int k = 0; k < 3; k++) f[k] = m.geom_friction[3 * g + k];
        for (int k = 0; k < 2; k++) solref[k] = m.geom_solref[2 * g + k];
        for (int k = 0; k < 5; k++) solimp[k] = m.geom_solimp[5 * g + k];
      } else {
        dim = std::max(m.geom_condim[g1], m.geom_condim[g2]);
        for (int k = 0; k < 3; k++) f[k] = std::max(m.geom_friction[3 * g1 + k], m.geom_friction[3 * g2 + k]);
        double s1 = m.geom_solmix[g1], s2 = m.geom_solmix[g2], mix;
        if (s1 >= kMinVal && s2 >= kMinVal) mix = s1 / (s1 + s2);
        else if (s1 < kMinVal && s2 < kMinVal) mix = 0.5;
        else if (s1 < kMinVal) mix = 0.0;
        else mix = 1.0;
        if (m.geom_solref[2 * g1] > 0 && m.geom_solref[2 * g2] > 0)
          for (int k = 0; k < 2; k++) solref[k] = mix * m.geom_solref[2 * g1 + k] + (1 - mix) * m.geom_solref[2 * g2 + k];
        else
          for (int k = 0; k < 2; k++) solref[k] = std::min(m.geom_solref[2 * g1 + k], m.geom_solref[2 * g2 + k]);
        for (int k = 0; k < 5; k++) solimp[k] = mix * m.geom_solimp[5 * g1 + k] + (1 - mix) * m.geom_solimp[5 * g2 + k];
      }
      double fr5[5] = {f[0], f[0], f[1], f[2], f[2]};
      addpair(g1, g2, dim, fr5, solref, solimp, margin, gap);
    };
    const bool contacts_on = !(m.opt_disableflags & ((1 << 4) | 1));
    if (contacts_on) {
      std::vector<std::pair<int, int>> expl;
      for (int p = 0; p < m.npair; p++) {
        int b1 = m.geom_bodyid[m.pair_geom1[p]], b2 = m.geom_bodyid[m.pair_geom2[p]];
        expl.push_back({(std::min(b1, b2) << 16) + std::max(b1, b2), p});
      }
      std::stable_sort(expl.begin(), expl.end());
      size_t ei = 0;
      auto doExplicit = [&](int upto) {
        for (; ei < expl.size() && expl[ei].first <= upto; ei++) {
          int p = expl[ei].second;
          addpair(m.pair_geom1[p], m.pair_geom2[p], m.pair_dim[p], &m.pair_friction[5 * p], &m.pair_solref[2 * p], &m.pair_solimp[5 * p],
                  m.pair_margin[p], m.pair_gap[p]);
        }
      };
      for (int b1 = 0; b1 < nbody; b1++) {
        if (m.body_geomnum[b1] == 0) continue;
        for (int b2 = b1 + 1; b2 < nbody; b2++) {
          if (m.body_geomnum[b2] == 0) continue;
          int sig = (b1 << 16) + b2;
          doExplicit(sig - 1);
          int w1 = m.body_weldid[b1], w2 = m.body_weldid[b2];
          bool skip = (w1 == w2);
          int wp1 = m.body_weldid[m.body_parentid[w1]], wp2 = m.body_weldid[m.body_parentid[w2]];
          if (!(m.opt_disableflags & DSBL_FILTERPARENT) && w1 != 0 && w2 != 0 && (w1 == wp2 || w2 == wp1)) skip = true;
          for (int e = 0; e < m.nexclude; e++) if (m.exclude_signature[e] == sig) skip = true;
          if (!skip)
            for (int g1 = m.body_geomadr[b1]; g1 < m.body_geomadr[b1] + m.body_geomnum[b1]; g1++)
              for (int g2 = m.body_geomadr[b2]; g2 < m.body_geomadr[b2] + m.body_geomnum[b2]; g2++) {
                if (!((m.geom_contype[g1] & m.geom_conaffinity[g2]) || (m.geom_contype[g2] & m.geom_conaffinity[g1]))) continue;
                bool isexpl = false;
                for (int p = 0; p < m.npair; p++)
                  if ((m.pair_geom1[p] == g1 && m.pair_geom2[p] == g2) || (m.pair_geom1[p] == g2 && m.pair_geom2[p] == g1)) isexpl = true;
                if (!isexpl) mixpair(g1, g2);
              }
          doExplicit(sig);
        }
      }
      doExplicit(1 << 30);
    }
    {   // touch sensors: zone = the site's shape (box or sphere; the leg model's zones are boxes)
      std::vector<int> ss, sb, st; std::vector<float> sz;
      for (int i = 0; i < m.nsensor; i++) {
        const int sid = m.sensor_objid[i];
        if (sid < 0 || sid >= m.nsite) throw std::string("touch sensor refers to an unknown site");
        const int ty = m.site_type[sid];
        if (ty != 6 && ty != 2) throw std::string("unsupported touch-sensor zone (site type must be box or sphere)");
        ss.push_back(sid); sb.push_back(m.site_bodyid[sid]); st.push_back(ty);
        for (int k = 0; k < 3; k++) sz.push_back((float)m.site_size[3 * sid + k]);
      }
      D.nsensor = m.nsensor;
      D.sens_site = upload(ss, owned); D.sens_body = upload(sb, owned); D.sens_type = upload(st, owned); D.sens_size = upload(sz, owned);
    }
    D.ncg = (int)cg_geom.size(); D.npair = (int)p1.size();
    D.nmprpair = 0;
    for (size_t i = 0; i < p1.size(); i++) {   // same rule as tile.cu isMprPair
      const int t1 = cg_type[p1[i]], t2 = cg_type[p2[i]];
      const bool simple1 = t1 == G_SPHERE || t1 == G_CAPSULE, simple2 = t2 == G_SPHERE || t2 == G_CAPSULE;
      if (t1 != G_PLANE && !(simple1 && simple2 && !(t1 == G_CAPSULE && t2 == G_SPHERE))) D.nmprpair++;
    }
    D.cg_geom = upload(cg_geom, owned); D.cg_type = upload(cg_type, owned); D.cg_frame = upload(cg_frame, owned); D.cg_body = upload(cg_body, owned);
    D.cg_lpos = upload(cg_lpos, owned); D.cg_lmat = upload(cg_lmat, owned); D.cg_size = upload(cg_size, owned); D.cg_rbound = upload(cg_rbound, owned);
    D.cg_meshadr = upload(cg_meshadr, owned); D.cg_meshnum = upload(cg_meshnum, owned);
    { std::vector<float> mv((size_t)3 * std::max(0, m.nmeshvert)); for (size_t i = 0; i < mv.size(); i++) mv[i] = (float)m.mesh_vert[i]; D.mesh_vert = upload(mv, owned); }
    D.cg_invweight = upload(cg_invw, owned); D.pair_cg1 = upload(p1, owned); D.pair_cg2 = upload(p2, owned); D.pair_dim = upload(pdim, owned);
    D.pair_prm = upload(pprm, owned);
  }

  // ------------------------------------------------------------------ equality / limit rows
  {
    const bool refsafe = !(m.opt_disableflags & DSBL_REFSAFE);
    const bool con_on = !(m.opt_disableflags & 1), eq_on = con_on && !(m.opt_disableflags & 2), lim_on = con_on && !(m.opt_disableflags & 8);
    std::vector<int> ed1, ed2, eq1, eq2, lq, ld, lj, lt;
    std::vector<float> edata, eprm, lprm, tprm;
    for (int e = 0; e < m.neq && eq_on; e++) {
      if (!m.eq_active0[e]) continue;
      if (m.eq_type[e] != 2) throw std::string("only joint equality constraints are supported");
      int j1 = m.eq_obj1id[e], j2 = m.eq_obj2id[e];
      ed1.push_back(m.jnt_dofadr[j1]); eq1.push_back(m.jnt_qposadr[j1]);
      ed2.push_back(j2 >= 0 ? m.jnt_dofadr[j2] : -1); eq2.push_back(j2 >= 0 ? m.jnt_qposadr[j2] : -1);
      for (int k = 0; k < 5; k++) edata.push_back((float)m.eq_data[MYO_NEQDATA * e + k]);
      float prm[8]; rowParams(&m.eq_solref[2 * e], &m.eq_solimp[5 * e], m.opt_timestep, refsafe, prm);
      prm[7] = (float)(m.dof_invweight0[m.jnt_dofadr[j1]] + (j2 >= 0 ? m.dof_invweight0[m.jnt_dofadr[j2]] : 0.0));
      eprm.insert(eprm.end(), prm, prm + 8);
    }
    for (int j = 0; j < njnt && lim_on; j++) {
      if (!m.jnt_limited[j] || (m.jnt_type[j] != JNT_HINGE && m.jnt_type[j] != JNT_SLIDE)) continue;
      lq.push_back(m.jnt_qposadr[j]); ld.push_back(m.jnt_dofadr[j]); lj.push_back(j);
      float prm[12] = {(float)m.jnt_range[2 * j], (float)m.jnt_range[2 * j + 1], (float)m.jnt_margin[j]};
      rowParams(&m.jnt_solref[2 * j], &m.jnt_solimp[5 * j], m.opt_timestep, refsafe, prm + 3);
      prm[10] = (float)m.dof_invweight0[m.jnt_dofadr[j]]; prm[11] = 0;
      lprm.insert(lprm.end(), prm, prm + 12);
    }
    for (int t = 0; t < m.ntendon && lim_on; t++) {
      if (!m.tendon_limited[t]) continue;
      lt.push_back(t);
      float prm[12] = {(float)m.tendon_range[2 * t], (float)m.tendon_range[2 * t + 1], (float)m.tendon_margin[t]};
      rowParams(&m.tendon_solref_lim[2 * t], &m.tendon_solimp_lim[5 * t], m.opt_timestep, refsafe, prm + 3);
      prm[10] = (float)m.tendon_invweight0[t]; prm[11] = 0;
      tprm.insert(tprm.end(), prm, prm + 12);
    }
    std::vector<int> dli(nv, -1), tli(m.ntendon, -1), des(nv + 1, 0), del;
    for (size_t i = 0; i < ld.size(); i++) dli[ld[i]] = (int)i;
    for (size_t i = 0; i < lt.size(); i++) tli[lt[i]] = (int)i;
    for (int d = 0; d < nv; d++) {
      des[d] = (int)del.size();
      for (size_t e = 0; e < ed1.size(); e++) { if (ed1[e] == d) del.push_back(2 * (int)e); if (ed2[e] == d) del.push_back(2 * (int)e + 1); }
    }
    des[nv] = (int)del.size();
    D.dof_limidx = upload(dli, owned); D.ten_limidx = upload(tli, owned); D.dofeq_start = upload(des, owned); D.dofeq_list = upload(del, owned);
    D.neq = (int)ed1.size(); D.nlimj = (int)lq.size(); D.nlimt = (int)lt.size();
    D.eq_dof1 = upload(ed1, owned); D.eq_dof2 = upload(ed2, owned); D.eq_q1 = upload(eq1, owned); D.eq_q2 = upload(eq2, owned);
    D.eq_data = upload(edata, owned); D.eq_prm = upload(eprm, owned); D.limj_q = upload(lq, owned); D.limj_dof = upload(ld, owned);
    D.limj_jnt = upload(lj, owned); D.limj_prm = upload(lprm, owned); D.limt_ten = upload(lt, owned); D.limt_prm = upload(tprm, owned);
  }

  D.maxbase = std::min(maxbase, 4);
  D.nslot = nslot;
  myo_make_layout(D, nconmax_req, njmax_req, L);
}

// Per-batch capacities (nconmax, njmax) and the per-world shared-memory layout that follows from them.
void myo_make_layout(DevModel& D, int nconmax_req, int njmax_req, SmemLayout& L) {
  const int nv = D.nv;
  int maxcon = 4 * D.npair;  // up to 4 contacts per plane-box pair, 2 per plane-capsule pair, 1 otherwise
  // defaults sized from the MyoHand random-ctrl rollout (fist clench): max 27 contacts / 123 rows over 8192 worlds x 1000 steps
  D.nconmax = nconmax_req > 0 ? nconmax_req : std::min(maxcon, 32);
  int maxrows = D.neq + 2 * D.nlimj + 2 * D.nlimt + D.nconmax * (D.maxbase == 1 ? 1 : 2 * (D.maxbase - 1));
  D.njmax = njmax_req > 0 ? njmax_req : std::min(maxrows, 160);
  if (D.njmax < D.neq) throw std::string("njmax smaller than the number of equality rows");
  const int nb = D.maxbase;  // base Jacobian rows per contact (normal, 2 tangents, torsion)
  int o = 0;
  auto take = [&](int n) { int r = o; o += (n + 3) & ~3; return r; };
  L.qpos = take(D.nq); L.qvel = take(nv); L.act = take(D.na); L.ctrl = take(D.nu);
  L.frames = take(16 * D.nframe); L.xanchor = take(3 * D.njnt); L.xaxis = take(3 * D.njnt); L.rootcom = take(3 * D.nroot);
  L.cinert = take(10 * D.nframe); L.crb = take(10 * D.nframe); L.cdof = take(6 * nv); L.cdof_dot = take(6 * nv);
  L.cvel = take(6 * D.nframe); L.cacc = take(6 * D.nframe);
  L.qM = take(D.nM); L.qLD = take(D.nM); L.qLDinv = take(nv); L.qH = take(D.nM); L.qHinv = take(nv);   // qH: Euler's M + h B factor; qLD stays for the wide-output kernel
  L.tenlen = take(D.nten); L.tenvel = take(D.nten); L.tenJ = take(D.nnz);
  L.actlen = take(D.nu); L.actvel = take(D.nu); L.actforce = take(D.nu); L.actdot = take(D.na);
  L.bias = take(nv); L.passive = take(nv); L.qfrc_act = take(nv); L.smooth = take(nv); L.qfrc_con = take(nv);
  L.qacc_smooth = take(nv); L.qacc = take(nv); L.warm = take(nv); L.tmpv = take(nv); L.tforce = take(D.nten);
  L.counts = take(4); L.wstate = take(4);
  // path points (persistent: the wide-output kernel reads them after the step): tangent points of the wrapped units and the
  // running count of wrapped geom units
  L.uwpt = take(6 * D.ngunit); L.gpre = take(D.ngunit + 1);
  L.wantw = 0;
  // overlay A: tendon stage
  L.slots = take(6 * D.nslot); L.ulen = take(D.nunit); L.uwrap = take(D.nunit);
  L.upts = take(6 * D.nunit);   // the units' site positions, for the step kernel's own path-point writer   // tangent points / wrapped-unit prefix counts for the path-point outputs
  L.tmpbody = take(13 * D.nbody);   // posoutKernel's body-level records (it runs beside the tendon stage: no overlay)
  int endA = o;
  o = endA;   // no overlay: the tendon, inertia and collision phases of a step run concurrently on separate streams (api.cu)
  // region B: collision, constraints, solver
  L.cgpose = take(12 * D.ncg); L.cgshape = take(5 * D.ncg);   // cgshape: effective type, size[3], rbound (per-world overrides)
  L.con = take(12 * D.nconmax); L.conJ = take(nb * D.nconmax * nv); L.efc = take(10 * D.njmax);
  // H: lower triangle + the factor's diagonal (tile.cu newtonKernel when its block does not fit shared memory)
  L.H = take(nv * nv + nv);
  L.Hsave = take(nv * (nv + 1) / 2);   // the unfactored Hessian, updated by the later Newton passes of a tile
  L.nvec = take(6 * nv);
  L.limrow = take(2 * (D.nlimj + D.nlimt));
  L.basev = take(2 * nb * D.nconmax);
  L.cand = take(D.npair); L.ctmp = take(32 * D.npair);   // survivor list and per-survivor narrow-phase records
#ifdef KEEP_CONU
  L.conU = take(nb * D.nconmax * nv);   // (debug: the old layout, region unused)
#else
  L.conU = 0;   // (round-2 first builds kept U = W Jb here; syrkBlocks forms it in registers)
#endif
  L.conW = take(8 * D.nconmax);         // per-contact weights of the active pyramid rows in base-row space
  int endB = o;
  L.total = std::max(endA, endB);
  L.nv = nv; L.hshared = 0;
}
