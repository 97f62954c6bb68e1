// TEST INFRASTRUCTURE -- NOT PRODUCT CODE.  PARITY UNPINNED (see below).
//
// CPU restatement (single world, array-of-structs, readable, templated on the scalar type) of the MuJoCo 3.2.8
// `mj_step` / `mj_forward` semantics that the reference reaches through third-party packages:
//   mjx.step            <- myosuite/mjx/play.py:41,47          (jit_step = jax.jit(mjx.step))
//   pipeline_step/init  <- myosuite/mjx/myodm_v0.py:159,287    (brax -> mjx.step / mjx.forward)
//   Physics.step(n)     <- myosuite/physics/mj_sim_scene.py:55 (dm_control -> MuJoCo C mj_step)
//   sim.forward()       <- myosuite/robot/robot.py:979
// Neither MuJoCo, MJX, brax nor dm_control is vendored in /root/reference or installed in this image
// (mujoco is pinned only in prose: 3.2.8, README.md:40), and the reference's own tests hold no golden vector
// for a physics step (myosuite/tests/test_envs.py:40-125 is a self-consistency check).  This file therefore
// restates the *published algorithm* (MuJoCo documentation, "Computation" chapter) for the
// feature subset of the five MyoSuite configs, and is anchored by analytic known-answer tests in tests/.
// PARITY UNPINNED: no MuJoCo output is available to pin it against.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this
// library.  The product (myosuite_mjx_b200/csrc) never links or calls it.
//
// Pipeline order (one mj_step, Euler integrator, Newton solver, pyramidal cones):
//   kinematics -> comPos -> tendon -> crb -> factorM -> collision -> makeConstraint -> transmission
//   -> fwdVelocity (ten/actuator velocity, comVel, passive, rne) -> fwdActuation -> fwdAcceleration
//   -> fwdConstraint (warmstart choice, Newton) -> Euler (implicit joint damping) .
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <atomic>
#include <thread>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <map>
#include <string>
#include <limits>
#include <memory>
#include <vector>

#include "../include/myostep.h"

namespace {

constexpr double kMinVal = 1e-15;  // mjMINVAL
constexpr double kMinImp = 0.0001, kMaxImp = 0.9999;
constexpr double kMaxVal = 1e10;   // mjMAXVAL (bad-state threshold)
enum { JNT_FREE = 0, JNT_BALL = 1, JNT_SLIDE = 2, JNT_HINGE = 3 };
enum { G_PLANE = 0, G_HFIELD, G_SPHERE, G_CAPSULE, G_ELLIPSOID, G_CYLINDER, G_BOX, G_MESH };
enum { W_NONE = 0, W_JOINT, W_PULLEY, W_SITE, W_SPHERE, W_CYLINDER };
enum { DYN_NONE = 0, DYN_INTEGRATOR, DYN_FILTER, DYN_FILTEREXACT, DYN_MUSCLE };
enum { GAIN_FIXED = 0, GAIN_AFFINE, GAIN_MUSCLE };
enum { BIAS_NONE = 0, BIAS_AFFINE, BIAS_MUSCLE };
enum { TRN_JOINT = 0, TRN_TENDON = 3 };
enum { CNSTR_EQUALITY = 0, CNSTR_FRICTION_DOF, CNSTR_FRICTION_TENDON, CNSTR_LIMIT_JOINT, CNSTR_LIMIT_TENDON,
       CNSTR_CONTACT_FRICTIONLESS, CNSTR_CONTACT_PYRAMIDAL, CNSTR_CONTACT_ELLIPTIC };
enum { DSBL_CONSTRAINT = 1 << 0, DSBL_EQUALITY = 1 << 1, DSBL_FRICTIONLOSS = 1 << 2, DSBL_LIMIT = 1 << 3,
       DSBL_CONTACT = 1 << 4, DSBL_PASSIVE = 1 << 5, DSBL_GRAVITY = 1 << 6, DSBL_CLAMPCTRL = 1 << 7,
       DSBL_WARMSTART = 1 << 8, DSBL_FILTERPARENT = 1 << 9, DSBL_ACTUATION = 1 << 10, DSBL_REFSAFE = 1 << 11,
       DSBL_EULERDAMP = 1 << 14 };

// ------------------------------------------------------------------------------------------ small math
template <typename T> struct V3 { T x, y, z; };
template <typename T> inline V3<T> operator+(V3<T> a, V3<T> b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
template <typename T> inline V3<T> operator-(V3<T> a, V3<T> b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
template <typename T> inline V3<T> operator*(V3<T> a, T s) { return {a.x * s, a.y * s, a.z * s}; }
template <typename T> inline T dot(V3<T> a, V3<T> b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
template <typename T> inline V3<T> cross(V3<T> a, V3<T> b) {
  return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
template <typename T> inline T norm(V3<T> a) { return std::sqrt(dot(a, a)); }
template <typename T> inline V3<T> ld3(const T* p) { return {p[0], p[1], p[2]}; }
template <typename T> inline V3<T> ld3d(const double* p) { return {T(p[0]), T(p[1]), T(p[2])}; }
template <typename T> inline void st3(T* p, V3<T> v) { p[0] = v.x; p[1] = v.y; p[2] = v.z; }
template <typename T> inline T normalize(V3<T>& a) {
  T n = norm(a);
  if (n < T(kMinVal)) { a = {T(1), T(0), T(0)}; } else { a = a * (T(1) / n); }
  return n;
}
// rotation matrices are row-major 3x3; matvec and transposed matvec
template <typename T> inline V3<T> mulM(const T* R, V3<T> v) {
  return {R[0] * v.x + R[1] * v.y + R[2] * v.z, R[3] * v.x + R[4] * v.y + R[5] * v.z, R[6] * v.x + R[7] * v.y + R[8] * v.z};
}
template <typename T> inline V3<T> mulMT(const T* R, V3<T> v) {
  return {R[0] * v.x + R[3] * v.y + R[6] * v.z, R[1] * v.x + R[4] * v.y + R[7] * v.z, R[2] * v.x + R[5] * v.y + R[8] * v.z};
}
template <typename T> inline void quatMul(T* r, const T* a, const T* b) {
  T w = a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3];
  T x = a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2];
  T y = a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1];
  T z = a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0];
  r[0] = w; r[1] = x; r[2] = y; r[3] = z;
}
template <typename T> inline void quatNormalize(T* q) {
  T n = std::sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  if (n < T(kMinVal)) { q[0] = 1; q[1] = q[2] = q[3] = 0; } else { for (int i = 0; i < 4; i++) q[i] /= n; }
}
template <typename T> inline void quat2mat(T* R, const T* q) {
  T w = q[0], x = q[1], y = q[2], z = q[3];
  R[0] = w * w + x * x - y * y - z * z; R[1] = 2 * (x * y - w * z); R[2] = 2 * (x * z + w * y);
  R[3] = 2 * (x * y + w * z); R[4] = w * w - x * x + y * y - z * z; R[5] = 2 * (y * z - w * x);
  R[6] = 2 * (x * z - w * y); R[7] = 2 * (y * z + w * x); R[8] = w * w - x * x - y * y + z * z;
}
template <typename T> inline void mulMM(T* C, const T* A, const T* B) {
  for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) C[3 * i + j] = A[3 * i] * B[j] + A[3 * i + 1] * B[3 + j] + A[3 * i + 2] * B[6 + j];
}
template <typename T> inline void axisAngle2Quat(T* q, V3<T> axis, T ang) {
  T s = std::sin(ang / 2);
  q[0] = std::cos(ang / 2); q[1] = axis.x * s; q[2] = axis.y * s; q[3] = axis.z * s;
}
template <typename T> inline T clip(T x, T lo, T hi) { return x < lo ? lo : (x > hi ? hi : x); }

// spatial vectors are [angular(3); linear(3)]; inertia "cinert" is (Ixx Iyy Izz Ixy Ixz Iyz, m*cx m*cy m*cz, m)
template <typename T> inline void inertMulVec(T* r, const T* I, const T* v) {
  V3<T> w{v[0], v[1], v[2]}, l{v[3], v[4], v[5]}, mc{I[6], I[7], I[8]};
  V3<T> Iw{I[0] * w.x + I[3] * w.y + I[4] * w.z, I[3] * w.x + I[1] * w.y + I[5] * w.z, I[4] * w.x + I[5] * w.y + I[2] * w.z};
  V3<T> a = Iw + cross(mc, l);          // angular momentum about the reference point
  V3<T> p = l * I[9] + cross(w, mc);    // linear momentum  m (v + w x c)
  r[0] = a.x; r[1] = a.y; r[2] = a.z; r[3] = p.x; r[4] = p.y; r[5] = p.z;
}
template <typename T> inline void crossMotion(T* r, const T* vel, const T* v) {
  V3<T> w{vel[0], vel[1], vel[2]}, l{vel[3], vel[4], vel[5]}, a{v[0], v[1], v[2]}, b{v[3], v[4], v[5]};
  V3<T> ra = cross(w, a), rl = cross(w, b) + cross(l, a);
  r[0] = ra.x; r[1] = ra.y; r[2] = ra.z; r[3] = rl.x; r[4] = rl.y; r[5] = rl.z;
}
template <typename T> inline void crossForce(T* r, const T* vel, const T* f) {
  V3<T> w{vel[0], vel[1], vel[2]}, l{vel[3], vel[4], vel[5]}, a{f[0], f[1], f[2]}, b{f[3], f[4], f[5]};
  V3<T> ra = cross(w, a) + cross(l, b), rl = cross(w, b);
  r[0] = ra.x; r[1] = ra.y; r[2] = ra.z; r[3] = rl.x; r[4] = rl.y; r[5] = rl.z;
}

// ------------------------------------------------------------------------------------------ tendon wrapping
// 2-D helpers of mju_wrap (MuJoCo docs, "Tendon wrapping")
// Do the straight parts site0 -> tangent0 and site1 -> tangent1 cross?  MuJoCo tests the closed parameter square [0,1]^2.
// When the straight path grazes the circle the two tangent points coincide and the parts meet exactly at their END points
// (a = b = 1): whether that counts as a crossing is then rounding noise, and it flips the wrap between a zero arc and the
// long way round (2 cm of tendon length on the finger at its joint limits, differently in float32 and float64).  A meeting
// within 1e-4 of both end points is therefore not a crossing -- the same branch in every precision.
template <typename T> bool segIntersect(const T* p1, const T* p2, const T* p3, const T* p4) {
  T det = (p4[1] - p3[1]) * (p2[0] - p1[0]) - (p4[0] - p3[0]) * (p2[1] - p1[1]);
  if (std::fabs(det) < T(kMinVal)) return false;
  T a = ((p4[0] - p3[0]) * (p1[1] - p3[1]) - (p4[1] - p3[1]) * (p1[0] - p3[0])) / det;
  T b = ((p2[0] - p1[0]) * (p1[1] - p3[1]) - (p2[1] - p1[1]) * (p1[0] - p3[0])) / det;
  if (a > T(1) - T(1e-4) && b > T(1) - T(1e-4)) return false;
  return a >= 0 && a <= 1 && b >= 0 && b <= 1;
}
template <typename T> T arcAngle(const T* p0, const T* p1, int ind) {
  T n0 = std::max(T(kMinVal), std::sqrt(p0[0] * p0[0] + p0[1] * p0[1]));
  T n1 = std::max(T(kMinVal), std::sqrt(p1[0] * p1[0] + p1[1] * p1[1]));
  T c = clip((p0[0] * p1[0] + p0[1] * p1[1]) / (n0 * n1), T(-1), T(1));
  T ang = std::acos(c);
  T cr = p0[1] * p1[0] - p0[0] * p1[1];
  // Grazing path (coincident tangent points): cr is pure rounding noise there and a sign test on it would turn a ~0 arc
  // into a ~2 pi one -- a measure-zero discontinuity of the published recipe (hit e.g. by a joint sitting exactly on
  // its range limit).  Both this restatement and the CUDA kernel ignore cr below 1e-5 * |p0| |p1|.
  if (std::fabs(cr) > T(1e-5) * n0 * n1 && ((cr > 0 && ind) || (cr < 0 && !ind))) ang = T(2 * M_PI) - ang;
  return ang;
}
// returns arc length or -1; pnt = two tangent points (2-D)
template <typename T> T wrapCircle(T* pnt, const T* d, const T* sd, T rad) {
  T sq0 = d[0] * d[0] + d[1] * d[1], sq1 = d[2] * d[2] + d[3] * d[3], sqr = rad * rad;
  T dif[2] = {d[2] - d[0], d[3] - d[1]};
  T dd = dif[0] * dif[0] + dif[1] * dif[1];
  T a = -(dif[0] * d[0] + dif[1] * d[1]) / std::max(T(kMinVal), dd);
  a = clip(a, T(0), T(1));
  T near[2] = {d[0] + a * dif[0], d[1] + a * dif[1]};
  // Side-site rule: a straight segment that clears the circle still wraps when it passes on the far side of the side site.
  // At a pose where the side site lies exactly on the segment's normal (the finger at its joint limits: qpos = 0) the sign
  // of this dot product is rounding noise and flips a wrap of ~2 cm on or off, differently in float32 and float64.  Ties
  // within 1e-5 of the product's scale (|sd| = rad) count as "same side": the same branch in every precision.
  const T nn = near[0] * near[0] + near[1] * near[1];
  if (nn > sqr && (!sd || sd[0] * near[0] + sd[1] * near[1] >= T(-1e-5) * rad * std::sqrt(nn))) return -1;
  if (sq0 < sqr || sq1 < sqr) return -1;
  T s0 = std::sqrt(sq0 - sqr), s1 = std::sqrt(sq1 - sqr);
  T sol[2][4], good[2];
  for (int i = 0; i < 2; i++) {
    T sg = i == 0 ? T(1) : T(-1);
    sol[i][0] = (d[0] * sqr + sg * rad * d[1] * s0) / sq0;
    sol[i][1] = (d[1] * sqr - sg * rad * d[0] * s0) / sq0;
    sol[i][2] = (d[2] * sqr - sg * rad * d[3] * s1) / sq1;
    sol[i][3] = (d[3] * sqr + sg * rad * d[2] * s1) / sq1;
    if (sd) {
      T t[2] = {sol[i][0] + sol[i][2], sol[i][1] + sol[i][3]};
      T n = std::max(T(kMinVal), std::sqrt(t[0] * t[0] + t[1] * t[1]));
      good[i] = (t[0] * sd[0] + t[1] * sd[1]) / n;
    } else {
      T t[2] = {sol[i][0] - sol[i][2], sol[i][1] - sol[i][3]};
      good[i] = -(t[0] * t[0] + t[1] * t[1]);
    }
    if (segIntersect(d, sol[i], d + 2, sol[i] + 2)) good[i] = -10000;
  }
  int i = good[0] > good[1] ? 0 : 1;
  for (int k = 0; k < 4; k++) pnt[k] = sol[i][k];
  if (segIntersect(d, pnt, d + 2, pnt + 2)) return -1;
  return rad * arcAngle(pnt, pnt + 2, i);
}
// side site inside the wrap circle: single tangent point found by Newton iteration; returns 0 or -1
template <typename T> T wrapInside(T* pnt, const T* d, T rad) {
  T l0 = std::sqrt(d[0] * d[0] + d[1] * d[1]), l1 = std::sqrt(d[2] * d[2] + d[3] * d[3]);
  if (l0 <= rad || l1 <= rad) return -1;
  T dif[2] = {d[2] - d[0], d[3] - d[1]};
  T dd = dif[0] * dif[0] + dif[1] * dif[1];
  if (dd > T(kMinVal)) {
    T a = -(dif[0] * d[0] + dif[1] * d[1]) / dd;
    if (a > 0 && a < 1) {
      T t[2] = {d[0] + a * dif[0], d[1] + a * dif[1]};
      if (std::sqrt(t[0] * t[0] + t[1] * t[1]) <= rad) return -1;
    }
  }
  // default: point on the circle closest to the midpoint direction
  T mid[2] = {T(0.5) * (d[0] + d[2]), T(0.5) * (d[1] + d[3])};
  T mn = std::max(T(kMinVal), std::sqrt(mid[0] * mid[0] + mid[1] * mid[1]));
  pnt[0] = pnt[2] = rad * mid[0] / mn; pnt[1] = pnt[3] = rad * mid[1] / mn;
  T A = rad / l0, B = rad / l1;
  T cosG = (l0 * l0 + l1 * l1 - dd) / (2 * l0 * l1);
  if (cosG < -1 + T(kMinVal)) return -1;
  if (cosG > 1 - T(kMinVal)) return 0;
  T G = std::acos(cosG);
  T z = T(1 - 1e-7);
  auto fn = [&](T zz) { return std::asin(A * zz) + std::asin(B * zz) - 2 * std::asin(zz) + G; };
  T f = fn(z);
  if (f > 0) return 0;
  int it = 0;
  for (; it < 20 && std::fabs(f) > T(1e-6); it++) {
    T df = A / std::max(T(kMinVal), std::sqrt(1 - z * z * A * A)) + B / std::max(T(kMinVal), std::sqrt(1 - z * z * B * B)) -
           2 / std::max(T(kMinVal), std::sqrt(1 - z * z));
    if (df > -T(kMinVal)) return 0;
    T z1 = z - f / df;
    if (z1 > z) return 0;
    z = z1;
    f = fn(z);
    if (f > T(1e-6)) return 0;
  }
  if (it >= 20) return 0;
  T vec[2], ang;
  if (d[0] * d[3] - d[1] * d[2] > 0) { vec[0] = d[0] / l0; vec[1] = d[1] / l0; ang = std::asin(z) - std::asin(A * z); }
  else { vec[0] = d[2] / l1; vec[1] = d[3] / l1; ang = std::asin(z) - std::asin(B * z); }
  T c = std::cos(ang), s = std::sin(ang);
  pnt[0] = pnt[2] = rad * (c * vec[0] - s * vec[1]);
  pnt[1] = pnt[3] = rad * (s * vec[0] + c * vec[1]);
  return 0;
}
// 3-D: wrap x0->x1 around a sphere / cylinder (geom frame gpos,gmat; radius); wpnt = two world tangent points
template <typename T> T wrapGeom(T* wpnt, V3<T> x0, V3<T> x1, V3<T> gpos, const T* gmat, T radius, int type, const V3<T>* side) {
  if (radius < T(kMinVal)) return -1;
  V3<T> p0 = mulMT(gmat, x0 - gpos), p1 = mulMT(gmat, x1 - gpos);
  if (norm(p0) < T(kMinVal) || norm(p1) < T(kMinVal)) return -1;
  V3<T> ax0, ax1;
  if (type == W_SPHERE) {
    ax0 = p0; normalize(ax0);
    V3<T> nrm = cross(p0, p1);
    T n = norm(nrm);
    if (n < T(kMinVal)) {
      T a[3] = {std::fabs(ax0.x), std::fabs(ax0.y), std::fabs(ax0.z)};
      int k = (a[0] >= a[1] && a[0] >= a[2]) ? 0 : (a[1] >= a[2] ? 1 : 2);
      V3<T> o{T(k != 0), T(k != 1), T(k != 2)};
      nrm = cross(ax0, o);
      normalize(nrm);
    } else {
      nrm = nrm * (T(1) / n);
    }
    ax1 = cross(nrm, ax0); normalize(ax1);
  } else {
    ax0 = {T(1), T(0), T(0)}; ax1 = {T(0), T(1), T(0)};
  }
  T d[4] = {dot(p0, ax0), dot(p0, ax1), dot(p1, ax0), dot(p1, ax1)};
  T sd[2];
  bool inside = false;
  if (side) {
    V3<T> s = mulMT(gmat, *side - gpos);
    sd[0] = dot(s, ax0); sd[1] = dot(s, ax1);
    T n = std::max(T(kMinVal), std::sqrt(sd[0] * sd[0] + sd[1] * sd[1]));
    sd[0] *= radius / n; sd[1] *= radius / n;
    inside = norm(s) < radius;
  }
  T pnt[4];
  T wlen = inside ? wrapInside(pnt, d, radius) : wrapCircle(pnt, d, side ? sd : (const T*)nullptr, radius);
  if (wlen < 0) return -1;
  V3<T> r0 = ax0 * pnt[0] + ax1 * pnt[1], r1 = ax0 * pnt[2] + ax1 * pnt[3];
  if (type == W_CYLINDER) {
    T L0 = std::sqrt((p0.x - r0.x) * (p0.x - r0.x) + (p0.y - r0.y) * (p0.y - r0.y));
    T L1 = std::sqrt((p1.x - r1.x) * (p1.x - r1.x) + (p1.y - r1.y) * (p1.y - r1.y));
    T tot = std::max(T(kMinVal), L0 + wlen + L1);
    r0.z = p0.z + (p1.z - p0.z) * L0 / tot;
    r1.z = p0.z + (p1.z - p0.z) * (L0 + wlen) / tot;
    T h = r1.z - r0.z;
    wlen = std::sqrt(wlen * wlen + h * h);
  }
  st3(wpnt, mulM(gmat, r0) + gpos);
  st3(wpnt + 3, mulM(gmat, r1) + gpos);
  return wlen;
}

// ------------------------------------------------------------------------------------------ muscle model
template <typename T> T muscleDynamics(T ctrl, T act, const double* prm) {
  T c = clip(ctrl, T(0), T(1)), a = clip(act, T(0), T(1));
  T ta = T(prm[0]) * (T(0.5) + T(1.5) * a), td = T(prm[1]) / (T(0.5) + T(1.5) * a);
  T sw = T(prm[2]), dctrl = c - act, tau;
  if (sw < T(kMinVal)) tau = dctrl > 0 ? ta : td;
  else {
    T x = clip(dctrl / sw + T(0.5), T(0), T(1));
    T s = x * x * x * (x * (6 * x - 15) + 10);
    tau = td + (ta - td) * s;
  }
  return dctrl / std::max(T(kMinVal), tau);
}
template <typename T> T muscleFL(T L, T lmin, T lmax) {
  if (L < lmin || L > lmax) return 0;
  T a = T(0.5) * (lmin + 1), b = T(0.5) * (1 + lmax), x;
  if (L <= a) { x = (L - lmin) / std::max(T(kMinVal), a - lmin); return T(0.5) * x * x; }
  if (L <= 1) { x = (1 - L) / std::max(T(kMinVal), 1 - a); return 1 - T(0.5) * x * x; }
  if (L <= b) { x = (L - 1) / std::max(T(kMinVal), b - 1); return 1 - T(0.5) * x * x; }
  x = (lmax - L) / std::max(T(kMinVal), lmax - b);
  return T(0.5) * x * x;
}
template <typename T> void muscleGainBias(T len, T vel, const double* lr, T acc0, const double* prm, T* gain, T* bias) {
  T r0 = T(prm[0]), r1 = T(prm[1]), force = T(prm[2]), scale = T(prm[3]), lmin = T(prm[4]), lmax = T(prm[5]),
    vmax = T(prm[6]), fpmax = T(prm[7]), fvmax = T(prm[8]);
  if (force < 0) force = scale / std::max(T(kMinVal), acc0);
  T L0 = (T(lr[1]) - T(lr[0])) / std::max(T(kMinVal), r1 - r0);
  T L = r0 + (len - T(lr[0])) / std::max(T(kMinVal), L0);
  T V = vel / std::max(T(kMinVal), L0 * vmax);
  T FL = muscleFL(L, lmin, lmax);
  T y = fvmax - 1, FV;
  if (V <= -1) FV = 0;
  else if (V <= 0) FV = (V + 1) * (V + 1);
  else if (V <= y) FV = fvmax - (y - V) * (y - V) / std::max(T(kMinVal), y);
  else FV = fvmax;
  *gain = -force * FL * FV;
  T b = T(0.5) * (1 + lmax);
  if (L <= 1) *bias = 0;
  else if (L <= b) { T x = (L - 1) / std::max(T(kMinVal), b - 1); *bias = -force * fpmax * T(0.5) * x * x; }
  else { T x = (L - b) / std::max(T(kMinVal), b - 1); *bias = -force * fpmax * (T(0.5) + x); }
}

// ------------------------------------------------------------------------------------------ convex support + MPR
template <typename T> struct Shape {
  int type; V3<T> pos; const T* mat; V3<T> size; T margin;  // margin: inflation added along the query direction
  const double* verts = nullptr; int nvert = 0;              // convex hull of a mesh geom (geom frame), mjModel.mesh_vert
};
template <typename T> V3<T> support(const Shape<T>& s, V3<T> dir) {  // dir: unit, world frame
  V3<T> d = mulMT(s.mat, dir), r{0, 0, 0};
  switch (s.type) {
    case G_SPHERE: r = d * s.size.x; break;
    case G_CAPSULE: r = d * s.size.x; r.z += (d.z >= 0 ? s.size.y : -s.size.y); break;
    case G_ELLIPSOID: {
      V3<T> t{s.size.x * s.size.x * d.x, s.size.y * s.size.y * d.y, s.size.z * s.size.z * d.z};
      T n = std::sqrt(std::max(T(kMinVal), t.x * d.x + t.y * d.y + t.z * d.z));
      r = t * (T(1) / n);
    } break;
    case G_CYLINDER: {
      T n = std::sqrt(d.x * d.x + d.y * d.y);
      if (n > T(kMinVal)) { r.x = s.size.x * d.x / n; r.y = s.size.x * d.y / n; }
      r.z = d.z >= 0 ? s.size.y : -s.size.y;
    } break;
    case G_BOX: r = {d.x >= 0 ? s.size.x : -s.size.x, d.y >= 0 ? s.size.y : -s.size.y, d.z >= 0 ? s.size.z : -s.size.z}; break;
    case G_MESH: {  // support vertex of the convex hull (MuJoCo: hill climbing on the hull graph, mjc_meshSupport; same vertex)
      T best = T(-1e300);
      for (int i = 0; i < s.nvert; i++) {
        V3<T> v{T(s.verts[3 * i]), T(s.verts[3 * i + 1]), T(s.verts[3 * i + 2])};
        T val = dot(v, d);
        if (val > best) { best = val; r = v; }
      }
    } break;
    default: break;
  }
  return mulM(s.mat, r) + s.pos + dir * s.margin;
}
// Minkowski Portal Refinement on B - A.  Returns true on overlap; then depth>=0, nrm (A->B, unit), pos (midpoint of
// the witness points).  Our own statement of the XenoCollide algorithm that libccd (used by mjc_Convex) implements;
// depth/normal are those of the final portal plane.
template <typename T> bool mprPenetration(const Shape<T>& A, const Shape<T>& B, T tol, int maxit, T* depth, V3<T>* nrm, V3<T>* pos) {
  struct Sup { V3<T> v, a, b; };
  auto sup = [&](V3<T> d) { Sup s; s.a = support(A, d * T(-1)); s.b = support(B, d); s.v = s.b - s.a; return s; };
  V3<T> v0 = B.pos - A.pos;
  if (norm(v0) < T(1e-12)) v0 = {T(1e-5), 0, 0};
  V3<T> dir = v0 * T(-1); normalize(dir);
  Sup s1 = sup(dir);
  if (dot(s1.v, dir) <= 0) return false;
  dir = cross(s1.v, v0);
  if (norm(dir) < T(1e-12) * std::max(T(1), norm(s1.v) * norm(v0))) {
    // origin on the ray v0 -> v1: penetration along that ray
    V3<T> n = s1.v - v0; normalize(n);
    *depth = dot(s1.v, n); *nrm = n * T(-1); *pos = (s1.a + s1.b) * T(0.5);
    return true;
  }
  normalize(dir);
  Sup s2 = sup(dir);
  if (dot(s2.v, dir) <= 0) return false;
  dir = cross(s1.v - v0, s2.v - v0); normalize(dir);
  if (dot(dir, v0) > 0) { std::swap(s1, s2); dir = dir * T(-1); }
  Sup s3;
  for (int it = 0;; it++) {
    if (it > maxit) return false;
    s3 = sup(dir);
    if (dot(s3.v, dir) <= 0) return false;
    if (dot(cross(s1.v, s3.v), v0) < 0) { s2 = s3; dir = cross(s1.v - v0, s3.v - v0); normalize(dir); continue; }
    if (dot(cross(s3.v, s2.v), v0) < 0) { s1 = s3; dir = cross(s3.v - v0, s2.v - v0); normalize(dir); continue; }
    break;
  }
  bool hit = false;
  for (int it = 0;; it++) {
    dir = cross(s2.v - s1.v, s3.v - s1.v); normalize(dir);
    if (dot(dir, s1.v) >= 0) hit = true;
    Sup s4 = sup(dir);
    T reach = dot(s4.v, dir);
    if (reach < 0 && !hit) return false;
    T adv = std::min(std::min(dot(s4.v - s1.v, dir), dot(s4.v - s2.v, dir)), dot(s4.v - s3.v, dir));
    if (adv <= tol || it >= maxit) {
      if (!hit) return false;
      // barycentric coordinates of the origin's projection onto the portal
      T d = dot(dir, s1.v);
      V3<T> p = dir * d;
      V3<T> e0 = s2.v - s1.v, e1 = s3.v - s1.v, e2 = p - s1.v;
      T d00 = dot(e0, e0), d01 = dot(e0, e1), d11 = dot(e1, e1), d20 = dot(e2, e0), d21 = dot(e2, e1);
      T den = d00 * d11 - d01 * d01;
      T b1 = T(1), b2 = 0, b3 = 0;
      if (std::fabs(den) > T(kMinVal)) { b2 = (d11 * d20 - d01 * d21) / den; b3 = (d00 * d21 - d01 * d20) / den; b1 = 1 - b2 - b3; }
      b1 = clip(b1, T(0), T(1)); b2 = clip(b2, T(0), T(1)); b3 = clip(b3, T(0), T(1));
      T bs = std::max(T(kMinVal), b1 + b2 + b3);
      V3<T> pa = (s1.a * b1 + s2.a * b2 + s3.a * b3) * (T(1) / bs), pb = (s1.b * b1 + s2.b * b2 + s3.b * b3) * (T(1) / bs);
      *depth = d; *nrm = dir * T(-1); *pos = (pa + pb) * T(0.5);
      return true;
    }
    V3<T> c = cross(s4.v, v0);
    if (dot(s1.v, c) > 0) { if (dot(s2.v, c) > 0) s1 = s4; else s3 = s4; }
    else { if (dot(s3.v, c) > 0) s2 = s4; else s1 = s4; }
  }
}

// ------------------------------------------------------------------------------------------ the oracle
template <typename T> struct Contact {
  T dist; T pos[3]; T frame[9]; T includemargin; T friction[5]; T solref[2]; T solimp[5];
  int dim, geom1, geom2, efc_address;
};

template <typename T> struct Oracle {
  const myo_model_desc& m;
  int nq, nv, nu, na, nbody, njnt, ngeom, nsite, nten, nwrap, nM;
  std::vector<T> qpos, qvel, act, ctrl, qacc_warmstart, time;
  std::vector<T> xpos, xquat, xmat, xipos, ximat, xanchor, xaxis, geom_xpos, geom_xmat, site_xpos, site_xmat, subtree_com,
      cinert, crb, cdof, cdof_dot, cvel, ten_length, ten_velocity, ten_J, wrap_xpos, actuator_length, actuator_velocity,
      actuator_moment, actuator_force, act_dot, qM, qLD, qLDiagInv, qMdense, qLDdense, qfrc_bias, qfrc_passive, qfrc_spring,
      qfrc_damper, qfrc_actuator, qfrc_smooth, qfrc_constraint, qfrc_inverse, qacc, qacc_smooth, qfrc_applied, sensordata;
  std::vector<int32_t> ten_wrapadr, ten_wrapnum, wrap_obj;
  std::vector<Contact<T>> contacts;
  // exported flat contact / efc arrays
  std::vector<int32_t> ncon_v, nefc_v, contact_geom, contact_dim, contact_efc_address, efc_type, solver_niter, status;
  std::vector<T> contact_dist, contact_pos, contact_frame, contact_includemargin, contact_friction, contact_solref,
      contact_solimp, efc_J, efc_pos, efc_margin, efc_D, efc_R, efc_aref, efc_force, efc_diagApprox, efc_vel, solver_cost;
  int nefc = 0;
  std::vector<int> efc_id;  // object id per row (joint / tendon / eq / contact)
  std::map<std::string, std::pair<void*, std::pair<int64_t, int>>> fields;  // name -> (ptr, (count, is_int))

  explicit Oracle(const myo_model_desc& md) : m(md) {
    nq = m.nq; nv = m.nv; nu = m.nu; na = m.na; nbody = m.nbody; njnt = m.njnt; ngeom = m.ngeom; nsite = m.nsite;
    nten = m.ntendon; nwrap = m.nwrap; nM = m.nM;
    qpos.assign(nq, 0); qvel.assign(nv, 0); act.assign(na, 0); ctrl.assign(nu, 0); qacc_warmstart.assign(nv, 0); time.assign(1, 0);
    xpos.assign(3 * nbody, 0); xquat.assign(4 * nbody, 0); xmat.assign(9 * nbody, 0); xipos.assign(3 * nbody, 0);
    ximat.assign(9 * nbody, 0); xanchor.assign(3 * njnt, 0); xaxis.assign(3 * njnt, 0); geom_xpos.assign(3 * ngeom, 0);
    geom_xmat.assign(9 * ngeom, 0); site_xpos.assign(3 * nsite, 0); site_xmat.assign(9 * nsite, 0);
    subtree_com.assign(3 * nbody, 0); cinert.assign(10 * nbody, 0); crb.assign(10 * nbody, 0); cdof.assign(6 * nv, 0);
    cdof_dot.assign(6 * nv, 0); cvel.assign(6 * nbody, 0); ten_length.assign(nten, 0); ten_velocity.assign(nten, 0);
    ten_J.assign((size_t)nten * nv, 0); wrap_xpos.assign(6 * nwrap, 0); wrap_obj.assign(2 * nwrap, 0);
    ten_wrapadr.assign(nten, 0); ten_wrapnum.assign(nten, 0); actuator_length.assign(nu, 0); actuator_velocity.assign(nu, 0);
    actuator_moment.assign((size_t)nu * nv, 0); actuator_force.assign(nu, 0); act_dot.assign(na, 0); qM.assign(nM, 0);
    qLD.assign(nM, 0); qLDiagInv.assign(nv, 0); qMdense.assign((size_t)nv * nv, 0); qLDdense.assign((size_t)nv * nv, 0);
    for (auto* v : {&qfrc_bias, &qfrc_passive, &qfrc_spring, &qfrc_damper, &qfrc_actuator, &qfrc_smooth, &qfrc_constraint,
                    &qfrc_inverse, &qacc, &qacc_smooth, &qfrc_applied}) v->assign(nv, 0);
    sensordata.assign(m.nsensor, 0);
    ncon_v.assign(1, 0); nefc_v.assign(1, 0); solver_niter.assign(1, 0); status.assign(1, 0); solver_cost.assign(1, 0);
    reset();
#define REG(f) fields[#f] = {(void*)f.data(), {(int64_t)f.size(), 0}}
#define REGI(f) fields[#f] = {(void*)f.data(), {(int64_t)f.size(), 1}}
    REG(qpos); REG(qvel); REG(act); REG(ctrl); REG(qacc_warmstart); REG(time); REG(xpos); REG(xquat); REG(xmat); REG(xipos);
    REG(ximat); REG(xanchor); REG(xaxis); REG(geom_xpos); REG(geom_xmat); REG(site_xpos); REG(site_xmat); REG(subtree_com);
    REG(cinert); REG(crb); REG(cdof); REG(cdof_dot); REG(cvel); REG(ten_length); REG(ten_velocity); REG(ten_J); REG(wrap_xpos);
    REG(actuator_length); REG(actuator_velocity); REG(actuator_moment); REG(actuator_force); REG(act_dot); REG(qM); REG(qLD);
    REG(qLDiagInv); REG(qMdense); REG(qLDdense); REG(qfrc_bias); REG(qfrc_passive); REG(qfrc_spring); REG(qfrc_damper);
    REG(qfrc_actuator); REG(qfrc_applied); REG(qfrc_smooth); REG(qfrc_constraint); REG(qfrc_inverse); REG(qacc); REG(qacc_smooth);
    REGI(ten_wrapadr); REGI(ten_wrapnum); REGI(wrap_obj); REGI(ncon_v); REGI(nefc_v); REGI(solver_niter); REGI(status);
    REG(solver_cost); REG(sensordata);
  }
  void regDyn() {  // vectors whose size changes per step
    REGI(contact_geom); REGI(contact_dim); REGI(contact_efc_address); REGI(efc_type);
    REG(contact_dist); REG(contact_pos); REG(contact_frame); REG(contact_includemargin); REG(contact_friction);
    REG(contact_solref); REG(contact_solimp); REG(efc_J); REG(efc_pos); REG(efc_margin); REG(efc_D); REG(efc_R); REG(efc_aref);
    REG(efc_force); REG(efc_diagApprox); REG(efc_vel);
  }
  bool disabled(int flag) const { return (m.opt_disableflags & flag) != 0; }

  void reset() {
    for (int i = 0; i < nq; i++) qpos[i] = T(m.qpos0[i]);
    std::fill(qvel.begin(), qvel.end(), T(0)); std::fill(act.begin(), act.end(), T(0));
    std::fill(ctrl.begin(), ctrl.end(), T(0)); std::fill(qacc_warmstart.begin(), qacc_warmstart.end(), T(0));
    std::fill(qfrc_applied.begin(), qfrc_applied.end(), T(0));
    time[0] = 0; status[0] = 0;
  }

  // ---------------------------------------------------------------- mj_kinematics
  void kinematics() {
    for (int k = 0; k < 3; k++) xpos[k] = xipos[k] = 0;
    xquat[0] = 1; xquat[1] = xquat[2] = xquat[3] = 0;
    quat2mat(&xmat[0], &xquat[0]); quat2mat(&ximat[0], &xquat[0]);
    for (int b = 1; b < nbody; b++) {
      int p = m.body_parentid[b];
      int ja = m.body_jntadr[b], jn = m.body_jntnum[b];
      V3<T> pos; T quat[4];
      if (jn == 1 && m.jnt_type[ja] == JNT_FREE) {
        int qa = m.jnt_qposadr[ja];
        pos = ld3(&qpos[qa]);
        for (int k = 0; k < 4; k++) quat[k] = qpos[qa + 3 + k];
        quatNormalize(quat);
        for (int k = 0; k < 4; k++) qpos[qa + 3 + k] = quat[k];
        st3(&xanchor[3 * ja], pos); st3(&xaxis[3 * ja], ld3d<T>(&m.jnt_axis[3 * ja]));
      } else {
        pos = ld3(&xpos[3 * p]) + mulM(&xmat[9 * p], ld3d<T>(&m.body_pos[3 * b]));
        T bq[4] = {T(m.body_quat[4 * b]), T(m.body_quat[4 * b + 1]), T(m.body_quat[4 * b + 2]), T(m.body_quat[4 * b + 3])};
        quatMul(quat, &xquat[4 * p], bq);
        for (int j = ja; j < ja + jn; j++) {
          T R[9]; quat2mat(R, quat);
          V3<T> anchor = mulM(R, ld3d<T>(&m.jnt_pos[3 * j])) + pos;
          V3<T> axis = mulM(R, ld3d<T>(&m.jnt_axis[3 * j]));
          st3(&xanchor[3 * j], anchor); st3(&xaxis[3 * j], axis);
          int qa = m.jnt_qposadr[j];
          T d = qpos[qa] - T(m.qpos0[qa]);
          if (m.jnt_type[j] == JNT_SLIDE) pos = pos + axis * d;
          else if (m.jnt_type[j] == JNT_HINGE) {
            T ql[4], qn[4]; axisAngle2Quat(ql, ld3d<T>(&m.jnt_axis[3 * j]), d);
            quatMul(qn, quat, ql);
            for (int k = 0; k < 4; k++) quat[k] = qn[k];
            quat2mat(R, quat);
            pos = anchor - mulM(R, ld3d<T>(&m.jnt_pos[3 * j]));
          }
        }
      }
      quatNormalize(quat);
      st3(&xpos[3 * b], pos);
      for (int k = 0; k < 4; k++) xquat[4 * b + k] = quat[k];
      quat2mat(&xmat[9 * b], quat);
    }
    for (int b = 1; b < nbody; b++) {
      st3(&xipos[3 * b], ld3(&xpos[3 * b]) + mulM(&xmat[9 * b], ld3d<T>(&m.body_ipos[3 * b])));
      T iq[4] = {T(m.body_iquat[4 * b]), T(m.body_iquat[4 * b + 1]), T(m.body_iquat[4 * b + 2]), T(m.body_iquat[4 * b + 3])}, q[4];
      quatMul(q, &xquat[4 * b], iq); quat2mat(&ximat[9 * b], q);
    }
    for (int g = 0; g < ngeom; g++) {
      int b = m.geom_bodyid[g];
      st3(&geom_xpos[3 * g], ld3(&xpos[3 * b]) + mulM(&xmat[9 * b], ld3d<T>(&m.geom_pos[3 * g])));
      T gq[4] = {T(m.geom_quat[4 * g]), T(m.geom_quat[4 * g + 1]), T(m.geom_quat[4 * g + 2]), T(m.geom_quat[4 * g + 3])}, q[4];
      quatMul(q, &xquat[4 * b], gq); quat2mat(&geom_xmat[9 * g], q);
    }
    for (int s = 0; s < nsite; s++) {
      int b = m.site_bodyid[s];
      st3(&site_xpos[3 * s], ld3(&xpos[3 * b]) + mulM(&xmat[9 * b], ld3d<T>(&m.site_pos[3 * s])));
      T sq[4] = {T(m.site_quat[4 * s]), T(m.site_quat[4 * s + 1]), T(m.site_quat[4 * s + 2]), T(m.site_quat[4 * s + 3])}, q[4];
      quatMul(q, &xquat[4 * b], sq); quat2mat(&site_xmat[9 * s], q);
    }
  }

  // ---------------------------------------------------------------- mj_comPos
  void comPos() {
    for (int b = 0; b < nbody; b++) st3(&subtree_com[3 * b], ld3(&xipos[3 * b]) * T(m.body_mass[b]));
    for (int b = nbody - 1; b > 0; b--) {
      int p = m.body_parentid[b];
      for (int k = 0; k < 3; k++) subtree_com[3 * p + k] += subtree_com[3 * b + k];
    }
    for (int b = 0; b < nbody; b++) {
      T sm = T(m.body_subtreemass[b]);
      if (sm < T(kMinVal)) st3(&subtree_com[3 * b], ld3(&xipos[3 * b]));
      else for (int k = 0; k < 3; k++) subtree_com[3 * b + k] /= sm;
    }
    for (int b = 0; b < 10; b++) cinert[b] = 0;
    for (int b = 1; b < nbody; b++) {
      V3<T> off = ld3(&xipos[3 * b]) - ld3(&subtree_com[3 * m.body_rootid[b]]);
      const T* R = &ximat[9 * b];
      T mass = T(m.body_mass[b]);
      T I[9];
      for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) {
        T s = 0;
        for (int k = 0; k < 3; k++) s += R[3 * i + k] * T(m.body_inertia[3 * b + k]) * R[3 * j + k];
        I[3 * i + j] = s;
      }
      T* c = &cinert[10 * b];
      c[0] = I[0] + mass * (off.y * off.y + off.z * off.z);
      c[1] = I[4] + mass * (off.x * off.x + off.z * off.z);
      c[2] = I[8] + mass * (off.x * off.x + off.y * off.y);
      c[3] = I[1] - mass * off.x * off.y;
      c[4] = I[2] - mass * off.x * off.z;
      c[5] = I[5] - mass * off.y * off.z;
      c[6] = mass * off.x; c[7] = mass * off.y; c[8] = mass * off.z; c[9] = mass;
    }
    for (int j = 0; j < njnt; j++) {
      int b = m.jnt_bodyid[j], da = m.jnt_dofadr[j];
      V3<T> off = ld3(&subtree_com[3 * m.body_rootid[b]]) - ld3(&xanchor[3 * j]);
      V3<T> axis = ld3(&xaxis[3 * j]);
      auto setdof = [&](int d, V3<T> ang, V3<T> lin) { st3(&cdof[6 * d], ang); st3(&cdof[6 * d + 3], lin); };
      switch (m.jnt_type[j]) {
        case JNT_FREE:
          for (int k = 0; k < 3; k++) { V3<T> e{T(k == 0), T(k == 1), T(k == 2)}; setdof(da + k, {0, 0, 0}, e); }
          for (int k = 0; k < 3; k++) {
            V3<T> ax{xmat[9 * b + k], xmat[9 * b + 3 + k], xmat[9 * b + 6 + k]};
            setdof(da + 3 + k, ax, cross(ax, off));
          }
          break;
        case JNT_SLIDE: setdof(da, {0, 0, 0}, axis); break;
        case JNT_HINGE: setdof(da, axis, cross(axis, off)); break;
        default: break;
      }
    }
  }

  // point Jacobian (translational jp[3*nv], rotational jr[3*nv]) of a world point attached to `body`
  void jac(T* jp, T* jr, V3<T> point, int body) const {
    if (jp) std::fill(jp, jp + 3 * nv, T(0));
    if (jr) std::fill(jr, jr + 3 * nv, T(0));
    V3<T> off = point - ld3(&subtree_com[3 * m.body_rootid[body]]);
    while (body > 0 && m.body_dofnum[body] == 0) body = m.body_parentid[body];
    if (body == 0) return;
    for (int i = m.body_dofadr[body] + m.body_dofnum[body] - 1; i >= 0; i = m.dof_parentid[i]) {
      V3<T> ang = ld3(&cdof[6 * i]), lin = ld3(&cdof[6 * i + 3]);
      V3<T> v = lin + cross(ang, off);
      if (jp) { jp[i] = v.x; jp[nv + i] = v.y; jp[2 * nv + i] = v.z; }
      if (jr) { jr[i] = ang.x; jr[nv + i] = ang.y; jr[2 * nv + i] = ang.z; }
    }
  }

  // ---------------------------------------------------------------- mj_tendon (spatial only)
  void tendon() {
    std::fill(ten_J.begin(), ten_J.end(), T(0));
    std::vector<T> j0(3 * nv), j1(3 * nv);
    int wcount = 0;
    std::fill(wrap_obj.begin(), wrap_obj.end(), 0); std::fill(wrap_xpos.begin(), wrap_xpos.end(), T(0));
    for (int t = 0; t < nten; t++) {
      int adr = m.tendon_adr[t], num = m.tendon_num[t];
      T div = 1, len = 0;
      ten_wrapadr[t] = wcount; ten_wrapnum[t] = 0;
      auto addpt = [&](V3<T> p, int obj) {  // path points are recorded for rendering parity only
        if (wcount < nwrap * 2) { st3(&wrap_xpos[3 * wcount], p); wrap_obj[wcount] = obj; wcount++; ten_wrapnum[t]++; }
      };
      int k = 0;
      while (k < num - 1) {
        int ty0 = m.wrap_type[adr + k], ty1 = m.wrap_type[adr + k + 1];
        if (ty0 == W_PULLEY) { div = T(m.wrap_prm[adr + k]); addpt({0, 0, 0}, -2); k++; continue; }
        if (ty1 == W_PULLEY) { addpt(ld3(&site_xpos[3 * m.wrap_objid[adr + k]]), -1); k++; continue; }  // branch end
        int s0 = m.wrap_objid[adr + k];
        V3<T> pts[4]; int bod[4]; T wlen = -1; int npt = 2;
        pts[0] = ld3(&site_xpos[3 * s0]); bod[0] = m.site_bodyid[s0];
        int advance;
        if (ty1 == W_SITE) {
          int s1 = m.wrap_objid[adr + k + 1];
          pts[1] = ld3(&site_xpos[3 * s1]); bod[1] = m.site_bodyid[s1];
          advance = 1;
        } else {
          int g = m.wrap_objid[adr + k + 1], s1 = m.wrap_objid[adr + k + 2];
          int sideid = (int)std::lround(m.wrap_prm[adr + k + 1]);
          V3<T> side{0, 0, 0};
          if (sideid >= 0) side = ld3(&site_xpos[3 * sideid]);
          T wp[6];
          wlen = wrapGeom(wp, pts[0], ld3(&site_xpos[3 * s1]), ld3(&geom_xpos[3 * g]), &geom_xmat[9 * g], T(m.geom_size[3 * g]),
                          ty1, sideid >= 0 ? &side : nullptr);
          if (wlen < 0) { pts[1] = ld3(&site_xpos[3 * s1]); bod[1] = m.site_bodyid[s1]; }
          else {
            pts[1] = ld3(wp); pts[2] = ld3(wp + 3); bod[1] = bod[2] = m.geom_bodyid[g];
            pts[3] = ld3(&site_xpos[3 * s1]); bod[3] = m.site_bodyid[s1]; npt = 4;
          }
          advance = 2;
        }
        addpt(pts[0], -1);
        if (npt == 4) { addpt(pts[1], m.wrap_objid[adr + k + 1]); addpt(pts[2], m.wrap_objid[adr + k + 1]); }
        for (int a = 0; a < npt - 1; a += 2) {
          V3<T> dif = pts[a + 1] - pts[a];
          T L = norm(dif);
          len += L / div;
          if (bod[a] != bod[a + 1] && L > T(kMinVal)) {
            dif = dif * (T(1) / L);
            jac(j0.data(), nullptr, pts[a], bod[a]); jac(j1.data(), nullptr, pts[a + 1], bod[a + 1]);
            for (int i = 0; i < nv; i++)
              ten_J[(size_t)t * nv + i] += ((j1[i] - j0[i]) * dif.x + (j1[nv + i] - j0[nv + i]) * dif.y + (j1[2 * nv + i] - j0[2 * nv + i]) * dif.z) / div;
          }
        }
        if (npt == 4) len += wlen / div;
        k += advance;
        if (k == num - 1) addpt(ld3(&site_xpos[3 * m.wrap_objid[adr + k]]), -1);
      }
      ten_length[t] = len;
    }
  }

  // ---------------------------------------------------------------- mj_crb, mj_factorM, mj_solveM (App. B.5)
  int rowlen(int i) const { return (i + 1 < nv ? m.dof_Madr[i + 1] : nM) - m.dof_Madr[i]; }
  void crbM() {
    crb = cinert;
    for (int b = nbody - 1; b > 0; b--) {
      int p = m.body_parentid[b];
      if (p > 0) for (int k = 0; k < 10; k++) crb[10 * p + k] += crb[10 * b + k];
    }
    std::fill(qMdense.begin(), qMdense.end(), T(0));
    for (int i = 0; i < nv; i++) {
      T f[6]; inertMulVec(f, &crb[10 * m.dof_bodyid[i]], &cdof[6 * i]);
      int a = m.dof_Madr[i];
      for (int j = i; j >= 0; j = m.dof_parentid[j]) {
        T s = 0;
        for (int k = 0; k < 6; k++) s += cdof[6 * j + k] * f[k];
        if (j == i) s += T(m.dof_armature[i]);
        qM[a++] = s;
        qMdense[(size_t)i * nv + j] = qMdense[(size_t)j * nv + i] = s;
      }
    }
  }
  void factorSparse(std::vector<T>& LD, std::vector<T>& dinv) const {
    for (int k = nv - 1; k >= 0; k--) {
      int akk = m.dof_Madr[k], n = rowlen(k);
      T dkk = LD[akk];
      int a = akk + 1;
      for (int i = m.dof_parentid[k]; i >= 0; i = m.dof_parentid[i], a++) {
        T tmp = LD[a] / dkk;
        int ai = m.dof_Madr[i];
        for (int p = 0; a + p < akk + n; p++) LD[ai + p] -= LD[a + p] * tmp;
        LD[a] = tmp;
      }
      dinv[k] = T(1) / dkk;
    }
  }
  void solveSparse(const std::vector<T>& LD, const std::vector<T>& dinv, T* x) const {
    for (int i = nv - 1; i >= 0; i--) {
      int a = m.dof_Madr[i] + 1;
      for (int j = m.dof_parentid[i]; j >= 0; j = m.dof_parentid[j]) x[j] -= LD[a++] * x[i];
    }
    for (int i = 0; i < nv; i++) x[i] *= dinv[i];
    for (int i = 0; i < nv; i++) {
      int a = m.dof_Madr[i] + 1;
      for (int j = m.dof_parentid[i]; j >= 0; j = m.dof_parentid[j]) x[i] -= LD[a++] * x[j];
    }
  }
  void factorM() {
    qLD = qM;
    factorSparse(qLD, qLDiagInv);
    std::fill(qLDdense.begin(), qLDdense.end(), T(0));
    for (int i = 0; i < nv; i++) {
      int a = m.dof_Madr[i];
      for (int j = i; j >= 0; j = m.dof_parentid[j]) qLDdense[(size_t)i * nv + j] = qLD[a++];
    }
  }

  // ---------------------------------------------------------------- mj_transmission
  void transmission() {
    std::fill(actuator_moment.begin(), actuator_moment.end(), T(0));
    for (int i = 0; i < nu; i++) {
      T gear = T(m.actuator_gear[6 * i]);
      int id = m.actuator_trnid[2 * i];
      if (m.actuator_trntype[i] == TRN_TENDON) {
        actuator_length[i] = gear * ten_length[id];
        for (int k = 0; k < nv; k++) actuator_moment[(size_t)i * nv + k] = gear * ten_J[(size_t)id * nv + k];
      } else {  // hinge / slide joint
        actuator_length[i] = gear * qpos[m.jnt_qposadr[id]];
        actuator_moment[(size_t)i * nv + m.jnt_dofadr[id]] = gear;
      }
    }
  }

  // ---------------------------------------------------------------- collision (App. B.7)
  void makeFrame(T* frame, V3<T> n) const {
    normalize(n);
    V3<T> y{0, 0, 0};
    if (n.y < T(0.5) && n.y > T(-0.5)) y.y = 1; else y.z = 1;
    y = y - n * dot(n, y); normalize(y);
    V3<T> z = cross(n, y);
    st3(frame, n); st3(frame + 3, y); st3(frame + 6, z);
  }
  struct PairParam { int dim; T friction[5], solref[2], solimp[5], margin, gap; };
  void mixParams(int g1, int g2, PairParam& pp) const {
    pp.margin = T(std::max(m.geom_margin[g1], m.geom_margin[g2]));
    pp.gap = T(std::max(m.geom_gap[g1], m.geom_gap[g2]));
    int p1 = m.geom_priority[g1], p2 = m.geom_priority[g2];
    T f[3];
    if (p1 != p2) {
      int g = p1 > p2 ? g1 : g2;
      pp.dim = m.geom_condim[g];
      for (int k = 0; k < 3; k++) f[k] = T(m.geom_friction[3 * g + k]);
      for (int k = 0; k < 2; k++) pp.solref[k] = T(m.geom_solref[2 * g + k]);
      for (int k = 0; k < 5; k++) pp.solimp[k] = T(m.geom_solimp[5 * g + k]);
    } else {
      pp.dim = std::max(m.geom_condim[g1], m.geom_condim[g2]);
      for (int k = 0; k < 3; k++) f[k] = T(std::max(m.geom_friction[3 * g1 + k], m.geom_friction[3 * g2 + k]));
      double s1 = m.geom_solmix[g1], s2 = m.geom_solmix[g2], mix;
      if (s1 >= kMinVal && s2 >= kMinVal) mix = s1 / (s1 + s2);
      else if (s1 < kMinVal && s2 < kMinVal) mix = 0.5;
      else if (s1 < kMinVal) mix = 0.0;
      else mix = 1.0;
      if (m.geom_solref[2 * g1] > 0 && m.geom_solref[2 * g2] > 0)
        for (int k = 0; k < 2; k++) pp.solref[k] = T(mix * m.geom_solref[2 * g1 + k] + (1 - mix) * m.geom_solref[2 * g2 + k]);
      else
        for (int k = 0; k < 2; k++) pp.solref[k] = T(std::min(m.geom_solref[2 * g1 + k], m.geom_solref[2 * g2 + k]));
      for (int k = 0; k < 5; k++) pp.solimp[k] = T(mix * m.geom_solimp[5 * g1 + k] + (1 - mix) * m.geom_solimp[5 * g2 + k]);
    }
    pp.friction[0] = pp.friction[1] = f[0]; pp.friction[2] = f[1]; pp.friction[3] = pp.friction[4] = f[2];
  }
  void addContact(int g1, int g2, const PairParam& pp, T dist, V3<T> pos, V3<T> nrm) {
    if (!(dist < pp.margin)) return;
    Contact<T> c;
    c.dist = dist; st3(c.pos, pos); makeFrame(c.frame, nrm);
    c.includemargin = pp.margin - pp.gap;
    for (int k = 0; k < 5; k++) { c.friction[k] = pp.friction[k]; c.solimp[k] = pp.solimp[k]; }
    c.solref[0] = pp.solref[0]; c.solref[1] = pp.solref[1];
    c.dim = pp.dim; c.geom1 = g1; c.geom2 = g2; c.efc_address = -1;
    contacts.push_back(c);
  }
  void sphereSphere(int g1, int g2, const PairParam& pp, V3<T> c1, T r1, V3<T> c2, T r2) {
    V3<T> d = c2 - c1;
    T n = norm(d);
    V3<T> nrm = n < T(kMinVal) ? V3<T>{1, 0, 0} : d * (T(1) / n);
    T dist = n - r1 - r2;
    addContact(g1, g2, pp, dist, c1 + nrm * (r1 + dist / 2), nrm);
  }
  void meshOf(int g, Shape<T>& s) const {
    if (m.geom_type[g] != G_MESH) return;
    const int id = m.geom_dataid[g];
    s.verts = m.mesh_vert + 3 * m.mesh_vertadr[id]; s.nvert = m.mesh_vertnum[id];
  }
  void collidePair(int g1, int g2, const PairParam& pp) {
    int t1 = m.geom_type[g1], t2 = m.geom_type[g2];
    if (t1 > t2) { std::swap(g1, g2); std::swap(t1, t2); }
    V3<T> p1 = ld3(&geom_xpos[3 * g1]), p2 = ld3(&geom_xpos[3 * g2]);
    const T *R1 = &geom_xmat[9 * g1], *R2 = &geom_xmat[9 * g2];
    V3<T> sz1 = ld3d<T>(&m.geom_size[3 * g1]), sz2 = ld3d<T>(&m.geom_size[3 * g2]);
    T rb1 = T(m.geom_rbound[g1]), rb2 = T(m.geom_rbound[g2]);
    V3<T> ax1{R1[2], R1[5], R1[8]}, ax2{R2[2], R2[5], R2[8]};
    if (t1 == G_PLANE) {
      if (t2 == G_PLANE || t2 == G_HFIELD) return;
      T cd = dot(p2 - p1, ax1);
      if (cd > rb2 + pp.margin) return;  // bounding-sphere cull
      auto planeSphere = [&](V3<T> c, T r) {
        T dist = dot(c - p1, ax1) - r;
        addContact(g1, g2, pp, dist, c - ax1 * (r + dist / 2), ax1);
      };
      if (t2 == G_SPHERE) planeSphere(p2, sz2.x);
      else if (t2 == G_CAPSULE) { planeSphere(p2 + ax2 * sz2.y, sz2.x); planeSphere(p2 - ax2 * sz2.y, sz2.x); }
      else if (t2 == G_BOX) {
        // mjc_PlaneBox (MuJoCo engine_collision_primitive.c, restated): the corners below the box centre that lie within the
        // margin become contacts, in corner order (bit 0 of the index = x), at most four
        T dist = dot(p2 - p1, ax1);
        int cnt = 0;
        for (int i = 0; i < 8 && cnt < 4; i++) {
          V3<T> v{(i & 1) ? sz2.x : -sz2.x, (i & 2) ? sz2.y : -sz2.y, (i & 4) ? sz2.z : -sz2.z};
          V3<T> corner = mulM(R2, v);
          T ld = dot(ax1, corner);
          if (dist + ld > pp.margin || ld > 0) continue;
          T cd = dist + ld;
          addContact(g1, g2, pp, cd, p2 + corner - ax1 * (cd / 2), ax1);
          cnt++;
        }
      }
      else {  // ellipsoid, cylinder, mesh hull: deepest support point along -normal (single contact; MuJoCo's mjc_PlaneConvex adds
              // up to three more hull vertices near the plane for a mesh -- a documented difference)
        Shape<T> s{t2, p2, R2, sz2, 0};
        meshOf(g2, s);
        V3<T> sp = support(s, ax1 * T(-1));
        T dist = dot(sp - p1, ax1);
        addContact(g1, g2, pp, dist, sp - ax1 * (dist / 2), ax1);
      }
      return;
    }
    if (t1 == G_HFIELD) return;
    if (norm(p2 - p1) > rb1 + rb2 + pp.margin) return;  // bounding-sphere cull
    if (t1 == G_SPHERE && t2 == G_SPHERE) { sphereSphere(g1, g2, pp, p1, sz1.x, p2, sz2.x); return; }
    if (t1 == G_SPHERE && t2 == G_CAPSULE) {
      T x = clip(dot(p1 - p2, ax2), -sz2.y, sz2.y);
      sphereSphere(g1, g2, pp, p1, sz1.x, p2 + ax2 * x, sz2.x);
      return;
    }
    if (t1 == G_CAPSULE && t2 == G_CAPSULE) {
      // closest points of the two axis segments, parameters x1,x2 in [-1,1] of the half-length vectors
      V3<T> a1 = ax1 * sz1.y, a2 = ax2 * sz2.y, dif = p1 - p2;
      T ma = dot(a1, a1), mb = -dot(a1, a2), mc = dot(a2, a2), u = -dot(a1, dif), v = dot(a2, dif);
      T det = ma * mc - mb * mb, x1, x2;
      if (std::fabs(det) >= T(kMinVal)) {
        x1 = (mc * u - mb * v) / det; x2 = (ma * v - mb * u) / det;
        if (x1 > 1) { x1 = 1; x2 = (v - mb) / mc; } else if (x1 < -1) { x1 = -1; x2 = (v + mb) / mc; }
        if (x2 > 1) { x2 = 1; x1 = clip((u - mb) / ma, T(-1), T(1)); } else if (x2 < -1) { x2 = -1; x1 = clip((u + mb) / ma, T(-1), T(1)); }
      } else {  // parallel axes: midpoint of the overlap
        x1 = clip(u / std::max(T(kMinVal), ma), T(-1), T(1));
        x2 = clip((v - mb * x1) / std::max(T(kMinVal), mc), T(-1), T(1));
      }
      sphereSphere(g1, g2, pp, p1 + a1 * x1, sz1.x, p2 + a2 * x2, sz2.x);
      return;
    }
    // general convex pair: MPR on shapes inflated by margin/2 each
    Shape<T> A{t1, p1, R1, sz1, pp.margin / 2}, B{t2, p2, R2, sz2, pp.margin / 2};
    meshOf(g1, A); meshOf(g2, B);
    T depth; V3<T> nrm, pos;
    if (mprPenetration(A, B, T(1e-6), 50, &depth, &nrm, &pos)) addContact(g1, g2, pp, pp.margin - depth, pos, nrm);
  }
  void collision() {
    contacts.clear();
    if (disabled(DSBL_CONTACT) || disabled(DSBL_CONSTRAINT)) return;
    // explicit pairs keyed by body-pair signature, merged into the body-pair sweep
    std::vector<std::pair<int, int>> expl;  // (signature, pair index)
    for (int p = 0; p < m.npair; p++) {
      int b1 = m.geom_bodyid[m.pair_geom1[p]], b2 = m.geom_bodyid[m.pair_geom2[p]];
      expl.push_back({(std::min(b1, b2) << 16) + std::max(b1, b2), p});
    }
    std::stable_sort(expl.begin(), expl.end());
    size_t ei = 0;
    auto doExplicit = [&](int upto) {
      for (; ei < expl.size() && expl[ei].first <= upto; ei++) {
        int p = expl[ei].second;
        PairParam pp; pp.dim = m.pair_dim[p]; pp.margin = T(m.pair_margin[p]); pp.gap = T(m.pair_gap[p]);
        for (int k = 0; k < 5; k++) { pp.friction[k] = T(m.pair_friction[5 * p + k]); pp.solimp[k] = T(m.pair_solimp[5 * p + k]); }
        pp.solref[0] = T(m.pair_solref[2 * p]); pp.solref[1] = T(m.pair_solref[2 * p + 1]);
        collidePair(m.pair_geom1[p], m.pair_geom2[p], pp);
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
        if (!disabled(DSBL_FILTERPARENT) && w1 != 0 && w2 != 0 && (w1 == wp2 || w2 == wp1)) skip = true;
        for (int e = 0; e < m.nexclude; e++) if (m.exclude_signature[e] == sig) skip = true;
        if (!skip)
          for (int g1 = m.body_geomadr[b1]; g1 < m.body_geomadr[b1] + m.body_geomnum[b1]; g1++)
            for (int g2 = m.body_geomadr[b2]; g2 < m.body_geomadr[b2] + m.body_geomnum[b2]; g2++) {
              if (!((m.geom_contype[g1] & m.geom_conaffinity[g2]) || (m.geom_contype[g2] & m.geom_conaffinity[g1]))) continue;
              bool isexpl = false;
              for (int p = 0; p < m.npair; p++)
                if ((m.pair_geom1[p] == g1 && m.pair_geom2[p] == g2) || (m.pair_geom1[p] == g2 && m.pair_geom2[p] == g1)) isexpl = true;
              if (isexpl) continue;
              PairParam pp; mixParams(g1, g2, pp);
              collidePair(g1, g2, pp);
            }
        doExplicit(sig);
      }
    }
    doExplicit(1 << 30);
  }

  // ---------------------------------------------------------------- mj_makeConstraint (App. B.6)
  void addRow(int type, int id, const T* J, T pos, T margin, T diagApprox) {
    efc_type.push_back(type); efc_id.push_back(id);
    efc_J.insert(efc_J.end(), J, J + nv);
    efc_pos.push_back(pos); efc_margin.push_back(margin); efc_diagApprox.push_back(diagApprox);
    nefc++;
  }
  void makeConstraint() {
    nefc = 0;
    efc_type.clear(); efc_id.clear(); efc_J.clear(); efc_pos.clear(); efc_margin.clear(); efc_diagApprox.clear();
    for (auto& c : contacts) c.efc_address = -1;
    if (!disabled(DSBL_CONSTRAINT)) {
      std::vector<T> J(nv);
      // equality: joint coupling  q1 - q1_0 = poly(q2 - q2_0)
      if (!disabled(DSBL_EQUALITY))
        for (int e = 0; e < m.neq; e++) {
          if (!m.eq_active0[e] || m.eq_type[e] != 2) continue;
          int j1 = m.eq_obj1id[e], j2 = m.eq_obj2id[e];
          const double* d = &m.eq_data[MYO_NEQDATA * e];
          std::fill(J.begin(), J.end(), T(0));
          int q1 = m.jnt_qposadr[j1], d1 = m.jnt_dofadr[j1];
          T pos, dA = T(m.dof_invweight0[d1]);
          if (j2 >= 0) {
            int q2 = m.jnt_qposadr[j2], d2 = m.jnt_dofadr[j2];
            T x = qpos[q2] - T(m.qpos0[q2]);
            T poly = T(d[0]) + x * (T(d[1]) + x * (T(d[2]) + x * (T(d[3]) + x * T(d[4]))));
            T dpoly = T(d[1]) + x * (2 * T(d[2]) + x * (3 * T(d[3]) + x * 4 * T(d[4])));
            pos = qpos[q1] - T(m.qpos0[q1]) - poly;
            J[d1] = 1; J[d2] = -dpoly;
            dA += T(m.dof_invweight0[d2]);
          } else {
            pos = qpos[q1] - T(m.qpos0[q1]) - T(d[0]);
            J[d1] = 1;
          }
          addRow(CNSTR_EQUALITY, e, J.data(), pos, 0, dA);
        }
      // joint limits
      if (!disabled(DSBL_LIMIT)) {
        for (int j = 0; j < njnt; j++) {
          if (!m.jnt_limited[j]) continue;
          int ty = m.jnt_type[j];
          if (ty != JNT_HINGE && ty != JNT_SLIDE) continue;
          T q = qpos[m.jnt_qposadr[j]], margin = T(m.jnt_margin[j]);
          for (int side = -1; side <= 1; side += 2) {
            T dist = side * (T(m.jnt_range[2 * j + (side + 1) / 2]) - q);
            if (dist < margin) {
              std::fill(J.begin(), J.end(), T(0));
              J[m.jnt_dofadr[j]] = T(-side);
              addRow(CNSTR_LIMIT_JOINT, j, J.data(), dist, margin, T(m.dof_invweight0[m.jnt_dofadr[j]]));
            }
          }
        }
        for (int t = 0; t < nten; t++) {
          if (!m.tendon_limited[t]) continue;
          T margin = T(m.tendon_margin[t]);
          for (int side = -1; side <= 1; side += 2) {
            T dist = side * (T(m.tendon_range[2 * t + (side + 1) / 2]) - ten_length[t]);
            if (dist < margin) {
              for (int k = 0; k < nv; k++) J[k] = -side * ten_J[(size_t)t * nv + k];
              addRow(CNSTR_LIMIT_TENDON, t, J.data(), dist, margin, T(m.tendon_invweight0[t]));
            }
          }
        }
      }
      // contacts (pyramidal cones)
      if (!disabled(DSBL_CONTACT)) {
        std::vector<T> jp1(3 * nv), jr1(3 * nv), jp2(3 * nv), jr2(3 * nv), Jc(6 * nv), row(nv);
        for (size_t ci = 0; ci < contacts.size(); ci++) {
          Contact<T>& c = contacts[ci];
          if (!(c.dist < c.includemargin)) continue;
          int b1 = m.geom_bodyid[c.geom1], b2 = m.geom_bodyid[c.geom2];
          jac(jp1.data(), jr1.data(), ld3(c.pos), b1); jac(jp2.data(), jr2.data(), ld3(c.pos), b2);
          for (int r = 0; r < 3; r++)
            for (int k = 0; k < nv; k++) {
              T sp = 0, sr = 0;
              for (int a = 0; a < 3; a++) {
                sp += c.frame[3 * r + a] * (jp2[a * nv + k] - jp1[a * nv + k]);
                sr += c.frame[3 * r + a] * (jr2[a * nv + k] - jr1[a * nv + k]);
              }
              Jc[r * nv + k] = sp; Jc[(3 + r) * nv + k] = sr;
            }
          T tran = T(m.body_invweight0[2 * b1] + m.body_invweight0[2 * b2]);
          T rot = T(m.body_invweight0[2 * b1 + 1] + m.body_invweight0[2 * b2 + 1]);
          c.efc_address = nefc;
          if (c.dim == 1) {
            addRow(CNSTR_CONTACT_FRICTIONLESS, (int)ci, Jc.data(), c.dist, c.includemargin, tran);
          } else {
            for (int k = 1; k < c.dim; k++) {
              T mu = c.friction[k - 1];
              T dA = tran + mu * mu * (k < 3 ? tran : rot);
              for (int sgn = 1; sgn >= -1; sgn -= 2) {
                for (int i = 0; i < nv; i++) row[i] = Jc[i] + sgn * mu * Jc[k * nv + i];
                addRow(CNSTR_CONTACT_PYRAMIDAL, (int)ci, row.data(), c.dist, c.includemargin, dA);
              }
            }
          }
        }
      }
    }
    // impedance, regularisation, reference acceleration
    efc_D.assign(nefc, 0); efc_R.assign(nefc, 0); efc_aref.assign(nefc, 0); efc_vel.assign(nefc, 0); efc_force.assign(nefc, 0);
    std::vector<T> imp(nefc), KB(2 * nefc);
    for (int i = 0; i < nefc; i++) {
      T solref[2], solimp[5];
      int id = efc_id[i];
      switch (efc_type[i]) {
        case CNSTR_EQUALITY: for (int k = 0; k < 2; k++) solref[k] = T(m.eq_solref[2 * id + k]); for (int k = 0; k < 5; k++) solimp[k] = T(m.eq_solimp[5 * id + k]); break;
        case CNSTR_LIMIT_JOINT: for (int k = 0; k < 2; k++) solref[k] = T(m.jnt_solref[2 * id + k]); for (int k = 0; k < 5; k++) solimp[k] = T(m.jnt_solimp[5 * id + k]); break;
        case CNSTR_LIMIT_TENDON: for (int k = 0; k < 2; k++) solref[k] = T(m.tendon_solref_lim[2 * id + k]); for (int k = 0; k < 5; k++) solimp[k] = T(m.tendon_solimp_lim[5 * id + k]); break;
        default: for (int k = 0; k < 2; k++) solref[k] = contacts[id].solref[k]; for (int k = 0; k < 5; k++) solimp[k] = contacts[id].solimp[k]; break;
      }
      // sanitise
      solimp[0] = clip(solimp[0], T(kMinImp), T(kMaxImp)); solimp[1] = clip(solimp[1], T(kMinImp), T(kMaxImp));
      solimp[2] = std::max(T(0), solimp[2]); solimp[3] = clip(solimp[3], T(kMinImp), T(kMaxImp)); solimp[4] = std::max(T(1), solimp[4]);
      if (solref[0] > 0 && !disabled(DSBL_REFSAFE)) solref[0] = std::max(solref[0], T(2 * m.opt_timestep));
      // impedance d(x)
      T dmin = solimp[0], dmax = solimp[1], width = solimp[2], mid = solimp[3], power = solimp[4], d;
      if (dmin == dmax || width <= T(kMinVal)) d = T(0.5) * (dmin + dmax);
      else {
        T x = std::fabs(efc_pos[i] - efc_margin[i]) / width;
        if (x >= 1) d = dmax;
        else if (x <= 0) d = dmin;
        else {
          T y;
          if (power == 1) y = x;
          else if (x <= mid) y = std::pow(x, power) / std::pow(mid, power - 1);
          else y = 1 - std::pow(1 - x, power) / std::pow(1 - mid, power - 1);
          d = dmin + y * (dmax - dmin);
        }
      }
      imp[i] = d;
      efc_R[i] = std::max(T(kMinVal), (1 - d) * efc_diagApprox[i] / d);
      if (solref[0] > 0) {
        KB[2 * i] = 1 / std::max(T(kMinVal), dmax * dmax * solref[0] * solref[0] * solref[1] * solref[1]);
        KB[2 * i + 1] = 2 / std::max(T(kMinVal), dmax * solref[0]);
      } else {
        KB[2 * i] = -solref[0] / std::max(T(kMinVal), dmax * dmax);
        KB[2 * i + 1] = -solref[1] / std::max(T(kMinVal), dmax);
      }
    }
    // pyramidal contacts share one regulariser  R = 2 mu^2 R_first
    for (auto& c : contacts) {
      if (c.efc_address < 0 || c.dim == 1) continue;
      T mu = c.friction[0] * std::sqrt(T(1) / T(m.opt_impratio));
      T Rpy = 2 * mu * mu * efc_R[c.efc_address];
      for (int k = 0; k < 2 * (c.dim - 1); k++) efc_R[c.efc_address + k] = std::max(T(kMinVal), Rpy);
    }
    for (int i = 0; i < nefc; i++) {
      efc_D[i] = 1 / efc_R[i];
      T v = 0;
      for (int k = 0; k < nv; k++) v += efc_J[(size_t)i * nv + k] * qvel[k];
      efc_vel[i] = v;
      efc_aref[i] = -KB[2 * i + 1] * v - KB[2 * i] * imp[i] * (efc_pos[i] - efc_margin[i]);
    }
  }

  // ---------------------------------------------------------------- mj_fwdVelocity
  void fwdVelocity() {
    for (int t = 0; t < nten; t++) { T s = 0; for (int k = 0; k < nv; k++) s += ten_J[(size_t)t * nv + k] * qvel[k]; ten_velocity[t] = s; }
    for (int i = 0; i < nu; i++) { T s = 0; for (int k = 0; k < nv; k++) s += actuator_moment[(size_t)i * nv + k] * qvel[k]; actuator_velocity[i] = s; }
    // comVel
    for (int k = 0; k < 6; k++) cvel[k] = 0;
    for (int b = 1; b < nbody; b++) {
      T v[6];
      for (int k = 0; k < 6; k++) v[k] = cvel[6 * m.body_parentid[b] + k];
      int da = m.body_dofadr[b];
      for (int j = m.body_jntadr[b]; j < m.body_jntadr[b] + m.body_jntnum[b]; j++) {
        int d = m.jnt_dofadr[j];
        if (m.jnt_type[j] == JNT_FREE) {
          for (int k = 0; k < 18; k++) cdof_dot[6 * d + k] = 0;
          for (int a = 0; a < 3; a++) for (int k = 0; k < 6; k++) v[k] += cdof[6 * (d + a) + k] * qvel[d + a];
          for (int a = 3; a < 6; a++) crossMotion(&cdof_dot[6 * (d + a)], v, &cdof[6 * (d + a)]);
          for (int a = 3; a < 6; a++) for (int k = 0; k < 6; k++) v[k] += cdof[6 * (d + a) + k] * qvel[d + a];
        } else {
          crossMotion(&cdof_dot[6 * d], v, &cdof[6 * d]);
          for (int k = 0; k < 6; k++) v[k] += cdof[6 * d + k] * qvel[d];
        }
      }
      (void)da;
      for (int k = 0; k < 6; k++) cvel[6 * b + k] = v[k];
    }
    // passive
    for (int i = 0; i < nv; i++) qfrc_spring[i] = qfrc_damper[i] = 0;
    if (!disabled(DSBL_PASSIVE)) {
      for (int j = 0; j < njnt; j++) {
        T k = T(m.jnt_stiffness[j]);
        if (k == 0) continue;
        if (m.jnt_type[j] == JNT_HINGE || m.jnt_type[j] == JNT_SLIDE)
          qfrc_spring[m.jnt_dofadr[j]] = -k * (qpos[m.jnt_qposadr[j]] - T(m.qpos_spring[m.jnt_qposadr[j]]));
      }
      for (int i = 0; i < nv; i++) qfrc_damper[i] = -T(m.dof_damping[i]) * qvel[i];
      for (int t = 0; t < nten; t++) {
        T k = T(m.tendon_stiffness[t]), b = T(m.tendon_damping[t]);
        if (k > 0) {
          T lo = T(m.tendon_lengthspring[2 * t]), hi = T(m.tendon_lengthspring[2 * t + 1]), L = ten_length[t], f = 0;
          if (L > hi) f = k * (hi - L); else if (L < lo) f = k * (lo - L);
          for (int i = 0; i < nv; i++) qfrc_spring[i] += ten_J[(size_t)t * nv + i] * f;
        }
        if (b > 0) for (int i = 0; i < nv; i++) qfrc_damper[i] += -b * ten_velocity[t] * ten_J[(size_t)t * nv + i];
      }
    }
    for (int i = 0; i < nv; i++) qfrc_passive[i] = qfrc_spring[i] + qfrc_damper[i];
    // rne without acceleration
    std::vector<T> cacc(6 * nbody, 0), cfrc(6 * nbody, 0);
    if (!disabled(DSBL_GRAVITY)) for (int k = 0; k < 3; k++) cacc[3 + k] = -T(m.opt_gravity[k]);
    for (int b = 1; b < nbody; b++) {
      for (int k = 0; k < 6; k++) cacc[6 * b + k] = cacc[6 * m.body_parentid[b] + k];
      for (int d = m.body_dofadr[b]; d >= 0 && d < m.body_dofadr[b] + m.body_dofnum[b]; d++)
        for (int k = 0; k < 6; k++) cacc[6 * b + k] += cdof_dot[6 * d + k] * qvel[d];
      T Ia[6], Iv[6], cf[6];
      inertMulVec(Ia, &cinert[10 * b], &cacc[6 * b]);
      inertMulVec(Iv, &cinert[10 * b], &cvel[6 * b]);
      crossForce(cf, &cvel[6 * b], Iv);
      for (int k = 0; k < 6; k++) cfrc[6 * b + k] = Ia[k] + cf[k];
    }
    for (int b = nbody - 1; b > 0; b--) {
      int p = m.body_parentid[b];
      if (p > 0) for (int k = 0; k < 6; k++) cfrc[6 * p + k] += cfrc[6 * b + k];
    }
    for (int i = 0; i < nv; i++) {
      T s = 0;
      for (int k = 0; k < 6; k++) s += cdof[6 * i + k] * cfrc[6 * m.dof_bodyid[i] + k];
      qfrc_bias[i] = s;
    }
  }

  // ---------------------------------------------------------------- mj_fwdActuation (App. B.3)
  void fwdActuation() {
    std::fill(qfrc_actuator.begin(), qfrc_actuator.end(), T(0));
    std::fill(actuator_force.begin(), actuator_force.end(), T(0));
    std::fill(act_dot.begin(), act_dot.end(), T(0));
    if (disabled(DSBL_ACTUATION)) return;
    for (int i = 0; i < nu; i++) {
      T c = ctrl[i];
      if (m.actuator_ctrllimited[i] && !disabled(DSBL_CLAMPCTRL)) c = clip(c, T(m.actuator_ctrlrange[2 * i]), T(m.actuator_ctrlrange[2 * i + 1]));
      int aa = m.actuator_actadr[i];
      const double* dp = &m.actuator_dynprm[MYO_NPRM * i];
      T a = c;
      if (aa >= 0) {
        a = act[aa];
        switch (m.actuator_dyntype[i]) {
          case DYN_INTEGRATOR: act_dot[aa] = c; break;
          case DYN_FILTER: case DYN_FILTEREXACT: act_dot[aa] = (c - a) / std::max(T(kMinVal), T(dp[0])); break;
          case DYN_MUSCLE: act_dot[aa] = muscleDynamics(c, a, dp); break;
          default: break;
        }
      }
      const double *gp = &m.actuator_gainprm[MYO_NPRM * i], *bp = &m.actuator_biasprm[MYO_NPRM * i];
      T len = actuator_length[i], vel = actuator_velocity[i], gain = 0, bias = 0, mg = 0, mb = 0;
      bool needm = m.actuator_gaintype[i] == GAIN_MUSCLE || m.actuator_biastype[i] == BIAS_MUSCLE;
      if (needm) muscleGainBias(len, vel, &m.actuator_lengthrange[2 * i], T(m.actuator_acc0[i]), gp, &mg, &mb);
      switch (m.actuator_gaintype[i]) {
        case GAIN_FIXED: gain = T(gp[0]); break;
        case GAIN_AFFINE: gain = T(gp[0]) + T(gp[1]) * len + T(gp[2]) * vel; break;
        case GAIN_MUSCLE: gain = mg; break;
      }
      switch (m.actuator_biastype[i]) {
        case BIAS_AFFINE: bias = T(bp[0]) + T(bp[1]) * len + T(bp[2]) * vel; break;
        case BIAS_MUSCLE: {  // bias uses its own parameter block
          T g2, b2; muscleGainBias(len, vel, &m.actuator_lengthrange[2 * i], T(m.actuator_acc0[i]), bp, &g2, &b2); bias = b2;
        } break;
        default: break;
      }
      (void)mb;
      T f = gain * a + bias;
      if (m.actuator_forcelimited[i]) f = clip(f, T(m.actuator_forcerange[2 * i]), T(m.actuator_forcerange[2 * i + 1]));
      actuator_force[i] = f;
      for (int k = 0; k < nv; k++) qfrc_actuator[k] += actuator_moment[(size_t)i * nv + k] * f;
    }
  }

  void fwdAcceleration() {
    for (int i = 0; i < nv; i++) { qfrc_smooth[i] = qfrc_passive[i] - qfrc_bias[i] + qfrc_actuator[i] + qfrc_applied[i]; qacc_smooth[i] = qfrc_smooth[i]; }
    solveSparse(qLD, qLDiagInv, qacc_smooth.data());
  }

  // ---------------------------------------------------------------- mj_fwdConstraint: Newton solver on the primal problem
  bool rowIsEquality(int i) const { return efc_type[i] == CNSTR_EQUALITY; }
  // cost of the constraint part and forces at residual jar = J a - aref
  T constraintCost(const std::vector<T>& jar, std::vector<T>* force, std::vector<char>* active) const {
    T cost = 0;
    for (int i = 0; i < nefc; i++) {
      bool on = rowIsEquality(i) || jar[i] < 0;
      if (force) (*force)[i] = on ? -efc_D[i] * jar[i] : T(0);
      if (active) (*active)[i] = on;
      if (on) cost += T(0.5) * efc_D[i] * jar[i] * jar[i];
    }
    return cost;
  }
  void mulM_dense(T* res, const T* v) const {
    for (int i = 0; i < nv; i++) { T s = 0; for (int k = 0; k < nv; k++) s += qMdense[(size_t)i * nv + k] * v[k]; res[i] = s; }
  }
  void fwdConstraint() {
    solver_niter[0] = 0;
    if (nefc == 0) {
      qacc = qacc_smooth; qacc_warmstart = qacc_smooth;
      std::fill(qfrc_constraint.begin(), qfrc_constraint.end(), T(0));
      solver_cost[0] = 0;
      return;
    }
    const int n = nv;
    std::vector<T> Ma(n), jar(nefc), grad(n), Mgrad(n), search(n), Mv(n), Jv(nefc), H((size_t)n * n);
    std::vector<char> active(nefc);
    auto evalAt = [&](const std::vector<T>& a, T* gauss) {
      mulM_dense(Ma.data(), a.data());
      for (int i = 0; i < nefc; i++) { T s = 0; for (int k = 0; k < n; k++) s += efc_J[(size_t)i * n + k] * a[k]; jar[i] = s - efc_aref[i]; }
      T g = 0;
      for (int k = 0; k < n; k++) g += T(0.5) * (Ma[k] - qfrc_smooth[k]) * (a[k] - qacc_smooth[k]);
      *gauss = g;
      return constraintCost(jar, &efc_force, &active) + g;
    };
    // warm start: the cheaper of qacc_warmstart and qacc_smooth
    T gauss, cost;
    if (!disabled(DSBL_WARMSTART)) {
      T cw = evalAt(qacc_warmstart, &gauss);
      T cs = evalAt(qacc_smooth, &gauss);
      qacc = cw > cs ? qacc_smooth : qacc_warmstart;
    } else qacc = qacc_smooth;
    cost = evalAt(qacc, &gauss);
    const T scale = T(1) / (T(m.stat_meaninertia) * std::max(1, n));
    const T tolerance = T(m.opt_tolerance), lstol = T(m.opt_ls_tolerance);
    auto gradientAndDirection = [&]() {
      for (int k = 0; k < n; k++) {
        T s = Ma[k] - qfrc_smooth[k];
        for (int i = 0; i < nefc; i++) s -= efc_J[(size_t)i * n + k] * efc_force[i];
        grad[k] = s;
      }
      // H = M + J^T D_active J, Cholesky, Mgrad = H^-1 grad
      H = qMdense;
      for (int i = 0; i < nefc; i++) if (active[i]) {
        const T* Ji = &efc_J[(size_t)i * n];
        for (int r = 0; r < n; r++) if (Ji[r] != 0) for (int c = 0; c <= r; c++) H[(size_t)r * n + c] += efc_D[i] * Ji[r] * Ji[c];
      }
      for (int r = 0; r < n; r++) {   // in-place lower Cholesky
        for (int c = 0; c <= r; c++) {
          T s = H[(size_t)r * n + c];
          for (int k = 0; k < c; k++) s -= H[(size_t)r * n + k] * H[(size_t)c * n + k];
          H[(size_t)r * n + c] = (r == c) ? std::sqrt(std::max(T(kMinVal), s)) : s / H[(size_t)c * n + c];
        }
      }
      for (int r = 0; r < n; r++) { T s = grad[r]; for (int k = 0; k < r; k++) s -= H[(size_t)r * n + k] * Mgrad[k]; Mgrad[r] = s / H[(size_t)r * n + r]; }
      for (int r = n - 1; r >= 0; r--) { T s = Mgrad[r]; for (int k = r + 1; k < n; k++) s -= H[(size_t)k * n + r] * Mgrad[k]; Mgrad[r] = s / H[(size_t)r * n + r]; }
      for (int k = 0; k < n; k++) search[k] = -Mgrad[k];
    };
    gradientAndDirection();
    int iter = 0;
    while (iter < m.opt_iterations) {
      // ---- exact line search on the piecewise-quadratic 1-D cost
      T snorm = 0; for (int k = 0; k < n; k++) snorm += search[k] * search[k]; snorm = std::sqrt(snorm);
      T alpha = 0;
      if (snorm >= T(kMinVal)) {
        mulM_dense(Mv.data(), search.data());
        for (int i = 0; i < nefc; i++) { T s = 0; for (int k = 0; k < n; k++) s += efc_J[(size_t)i * n + k] * search[k]; Jv[i] = s; }
        T qg[3] = {gauss, 0, 0};
        for (int k = 0; k < n; k++) { qg[1] += search[k] * (Ma[k] - qfrc_smooth[k]); qg[2] += T(0.5) * search[k] * Mv[k]; }
        const T gtol = tolerance * lstol * snorm / scale;
        // tol: MuJoCo's gtol, or the rounding-noise floor of the derivative sum in the scalar type when that is larger
        // (16 eps * magnitude of the cancelling terms: 2e-6 in float32 -- the CUDA kernel's LS_NOISE -- and 3.6e-15 in
        // double, where it never binds).  Without it the float32 search runs to ls_iterations on rounding noise.
        struct Pt { T alpha, cost, d0, d1, tol; };
        const T lsnoise = T(16) * std::numeric_limits<T>::epsilon();
        int lsiter = 0;
        auto eval = [&](T a) {
          T q0 = qg[0], q1 = qg[1], q2 = qg[2], qa = std::fabs(qg[1]);
          for (int i = 0; i < nefc; i++) {
            T x = jar[i] + a * Jv[i];
            if (rowIsEquality(i) || x < 0) {
              q0 += T(0.5) * efc_D[i] * jar[i] * jar[i]; q1 += efc_D[i] * jar[i] * Jv[i]; q2 += T(0.5) * efc_D[i] * Jv[i] * Jv[i];
              qa += std::fabs(efc_D[i] * jar[i] * Jv[i]);
            }
          }
          lsiter++;
          Pt p{a, a * a * q2 + a * q1 + q0, 2 * a * q2 + q1, 2 * q2, std::max(gtol, lsnoise * (qa + 2 * std::fabs(a) * q2))};
          if (p.d1 <= 0) p.d1 = T(kMinVal);
          return p;
        };
        Pt p0 = eval(0);
        Pt p1 = eval(p0.alpha - p0.d0 / p0.d1);
        if (p0.cost < p1.cost) p1 = p0;
        bool done = std::fabs(p1.d0) < p1.tol;
        if (!done) {
          const T dir = p1.d0 < 0 ? T(1) : T(-1);
          Pt p2 = p1; bool have2 = false;
          while (p1.d0 * dir <= -p1.tol && lsiter < m.opt_ls_iterations) {
            p2 = p1; have2 = true;
            p1 = eval(p1.alpha - p1.d0 / p1.d1);
            if (std::fabs(p1.d0) < p1.tol) { done = true; break; }
          }
          if (!done && lsiter < m.opt_ls_iterations && have2) {
            // bracket: p2 still descends along dir, p1 has overshot
            Pt lo = p2, hi = p1;
            while (lsiter < m.opt_ls_iterations) {
              Pt cand[3] = {eval(lo.alpha - lo.d0 / lo.d1), eval(hi.alpha - hi.d0 / hi.d1), eval(T(0.5) * (lo.alpha + hi.alpha))};
              bool moved = false;
              Pt best = lo.cost < hi.cost ? lo : hi;
              for (auto& c : cand) {
                if (std::fabs(c.d0) < c.tol) { p1 = c; done = true; break; }
                bool insideBr = (c.alpha - lo.alpha) * dir > 0 && (hi.alpha - c.alpha) * dir > 0;
                if (!insideBr) continue;
                if (c.d0 * dir < 0) { lo = c; moved = true; } else { hi = c; moved = true; }
                if (c.cost < best.cost) best = c;
              }
              if (done) break;
              if (!moved) { p1 = best; break; }
              p1 = lo.cost < hi.cost ? lo : hi;
            }
          }
        }
        alpha = p1.alpha;
        if (getenv("MYO_ORC_TRACE")) fprintf(stderr, "  linesearch evals %d alpha %.9g d0 %.3g gtol %.3g\n", lsiter, (double)alpha, (double)p1.d0, (double)gtol);
      }
      if (alpha == 0) break;
      for (int k = 0; k < n; k++) { qacc[k] += alpha * search[k]; Ma[k] += alpha * Mv[k]; }
      for (int i = 0; i < nefc; i++) jar[i] += alpha * Jv[i];
      T oldcost = cost;
      gauss = 0;
      for (int k = 0; k < n; k++) gauss += T(0.5) * (Ma[k] - qfrc_smooth[k]) * (qacc[k] - qacc_smooth[k]);
      cost = constraintCost(jar, &efc_force, &active) + gauss;
      gradientAndDirection();
      iter++;
      T gn = 0; for (int k = 0; k < n; k++) gn += grad[k] * grad[k];
      T improvement = scale * (oldcost - cost), gradient = scale * std::sqrt(gn);
      if (getenv("MYO_ORC_TRACE")) fprintf(stderr, "  it %d cost %.9g impr/tol %.3g grad/tol %.3g relimpr %.3g\n", iter, (double)cost, (double)(improvement / tolerance), (double)(gradient / tolerance), (double)((oldcost - cost) / (std::fabs(cost) + 1e-30)));
      if (improvement < tolerance || gradient < tolerance) break;
    }
    solver_niter[0] = iter;
    solver_cost[0] = cost;
    for (int k = 0; k < n; k++) { T s = 0; for (int i = 0; i < nefc; i++) s += efc_J[(size_t)i * n + k] * efc_force[i]; qfrc_constraint[k] = s; }
    qacc_warmstart = qacc;
  }

  void exportContacts() {
    size_t nc = contacts.size();
    ncon_v[0] = (int)nc; nefc_v[0] = nefc;
    contact_geom.resize(2 * nc); contact_dim.resize(nc); contact_efc_address.resize(nc); contact_dist.resize(nc);
    contact_pos.resize(3 * nc); contact_frame.resize(9 * nc); contact_includemargin.resize(nc); contact_friction.resize(5 * nc);
    contact_solref.resize(2 * nc); contact_solimp.resize(5 * nc);
    for (size_t i = 0; i < nc; i++) {
      const Contact<T>& c = contacts[i];
      contact_geom[2 * i] = c.geom1; contact_geom[2 * i + 1] = c.geom2; contact_dim[i] = c.dim; contact_efc_address[i] = c.efc_address;
      contact_dist[i] = c.dist; contact_includemargin[i] = c.includemargin;
      for (int k = 0; k < 3; k++) contact_pos[3 * i + k] = c.pos[k];
      for (int k = 0; k < 9; k++) contact_frame[9 * i + k] = c.frame[k];
      for (int k = 0; k < 5; k++) { contact_friction[5 * i + k] = c.friction[k]; contact_solimp[5 * i + k] = c.solimp[k]; }
      contact_solref[2 * i] = c.solref[0]; contact_solref[2 * i + 1] = c.solref[1];
    }
    regDyn();
  }

  void forward() {
    kinematics(); comPos(); tendon(); crbM(); factorM(); collision(); makeConstraint(); transmission();
    fwdVelocity(); fwdActuation(); fwdAcceleration(); fwdConstraint();
    for (int i = 0; i < nv; i++) qfrc_inverse[i] = 0;
    sensorTouch();
    exportContacts();
  }

  // mj_sensorAcc, mjSENS_TOUCH (MuJoCo 3.2.8 engine_sensor.c semantics): for every active contact that involves the
  // sensor site's body, normal force = first component of mj_contactForce (pyramidal: the sum of the contact's row
  // forces, mju_decodePyramid); it is added when positive and when the ray from the contact point along the normal
  // (flipped if the site's body is the contact's second body) meets the site's zone (mju_rayGeom >= 0).
  void sensorTouch() {
    for (int s = 0; s < m.nsensor; s++) {
      const int sid = m.sensor_objid[s], body = m.site_bodyid[sid], type = m.site_type[sid];
      const T* sp = &site_xpos[3 * sid]; const T* R = &site_xmat[9 * sid];
      const double* size = m.site_size + 3 * sid;
      T total = 0;
      for (const Contact<T>& c : contacts) {
        if (c.efc_address < 0) continue;
        const int b1 = m.geom_bodyid[c.geom1], b2 = m.geom_bodyid[c.geom2];
        if (body != b1 && body != b2) continue;
        const int rows = c.dim == 1 ? 1 : 2 * (c.dim - 1);
        T fn = 0;
        for (int q = 0; q < rows; q++) fn += efc_force[c.efc_address + q];
        if (!(fn > 0)) continue;
        T dir[3] = {c.frame[0] * fn, c.frame[1] * fn, c.frame[2] * fn};
        const T nn = std::sqrt(dir[0] * dir[0] + dir[1] * dir[1] + dir[2] * dir[2]);
        for (int k = 0; k < 3; k++) dir[k] = nn < T(1e-15) ? T(k == 0) : dir[k] / nn;
        if (body == b2) for (int k = 0; k < 3; k++) dir[k] = -dir[k];
        T p[3], v[3];
        for (int k = 0; k < 3; k++) {   // ray in the site frame: R^T (x - site_xpos)
          p[k] = R[k] * (c.pos[0] - sp[0]) + R[3 + k] * (c.pos[1] - sp[1]) + R[6 + k] * (c.pos[2] - sp[2]);
          v[k] = R[k] * dir[0] + R[3 + k] * dir[1] + R[6 + k] * dir[2];
        }
        bool hit = true;
        if (type == 2) {          // sphere
          const T b = p[0] * v[0] + p[1] * v[1] + p[2] * v[2], cc = p[0] * p[0] + p[1] * p[1] + p[2] * p[2] - T(size[0] * size[0]), disc = b * b - cc;
          hit = disc >= 0 && (-b + std::sqrt(disc)) >= 0;
        } else {                  // box (validated at plan time by the library; the oracle supports the same two shapes)
          T t0 = 0, t1 = std::numeric_limits<T>::max();
          for (int k = 0; k < 3; k++) {
            if (std::fabs(v[k]) < T(1e-15)) { if (std::fabs(p[k]) > T(size[k])) hit = false; }
            else {
              T a = (T(-size[k]) - p[k]) / v[k], b = (T(size[k]) - p[k]) / v[k];
              if (a > b) std::swap(a, b);
              t0 = std::max(t0, a); t1 = std::min(t1, b);
            }
          }
          hit = hit && t0 <= t1;
        }
        if (hit) total += fn;
      }
      sensordata[s] = total;
    }
  }

  bool bad(const std::vector<T>& v) const {
    for (T x : v) if (!(std::fabs(x) < T(kMaxVal))) return true;
    return false;
  }
  void step() {
    if (bad(qpos)) { reset(); status[0] |= MYO_WORLD_BADQPOS; }
    if (bad(qvel)) { reset(); status[0] |= MYO_WORLD_BADQVEL; }
    forward();
    if (bad(qacc)) { T t = time[0]; int st = status[0]; reset(); forward(); time[0] = t; status[0] = st | MYO_WORLD_BADQACC; }
    // Euler with implicit joint damping
    const T h = T(m.opt_timestep);
    std::vector<T> qa(qacc);
    bool damp = false;
    for (int i = 0; i < nv; i++) if (m.dof_damping[i] > 0) damp = true;
    if (damp && !disabled(DSBL_EULERDAMP)) {
      std::vector<T> H(qM), Hd(nv);
      for (int i = 0; i < nv; i++) H[m.dof_Madr[i]] += h * T(m.dof_damping[i]);
      factorSparse(H, Hd);
      for (int i = 0; i < nv; i++) qa[i] = qfrc_smooth[i] + qfrc_constraint[i];
      solveSparse(H, Hd, qa.data());
    }
    if (!disabled(DSBL_ACTUATION))
      for (int i = 0; i < nu; i++) {
        int aa = m.actuator_actadr[i];
        if (aa < 0) continue;
        T a = act[aa] + h * act_dot[aa];
        if (m.actuator_actlimited[i]) a = clip(a, T(m.actuator_actrange[2 * i]), T(m.actuator_actrange[2 * i + 1]));
        act[aa] = a;
      }
    for (int i = 0; i < nv; i++) qvel[i] += h * qa[i];
    for (int j = 0; j < njnt; j++) {
      int qa_ = m.jnt_qposadr[j], da = m.jnt_dofadr[j];
      if (m.jnt_type[j] == JNT_FREE) {
        for (int k = 0; k < 3; k++) qpos[qa_ + k] += h * qvel[da + k];
        V3<T> w{qvel[da + 3], qvel[da + 4], qvel[da + 5]};
        T ang = norm(w) * h;
        if (ang > 0) {
          V3<T> ax = w; normalize(ax);
          T dq[4], qn[4]; axisAngle2Quat(dq, ax, ang);
          quatMul(qn, &qpos[qa_ + 3], dq); quatNormalize(qn);
          for (int k = 0; k < 4; k++) qpos[qa_ + 3 + k] = qn[k];
        }
      } else {
        qpos[qa_] += h * qvel[da];
      }
    }
    time[0] += h;
  }
};

typedef Oracle<double> OrcD;
typedef Oracle<float> OrcF;
struct Handle { int is_float; OrcD* d; OrcF* f; };

// xorshift64* stream used by the CPU-baseline rollout only
inline double urand(uint64_t& s) { s ^= s >> 12; s ^= s << 25; s ^= s >> 27; return double((s * 2685821657736338717ull) >> 11) / 9007199254740992.0; }

}  // namespace

extern "C" {

void* orc_create(const myo_model_desc* desc, int use_float) {
  Handle* h = new Handle{use_float, nullptr, nullptr};
  if (use_float) h->f = new OrcF(*desc); else h->d = new OrcD(*desc);
  return h;
}
void orc_destroy(void* hv) { Handle* h = (Handle*)hv; delete h->d; delete h->f; delete h; }
void orc_reset(void* hv) { Handle* h = (Handle*)hv; if (h->is_float) h->f->reset(); else h->d->reset(); }
void orc_forward(void* hv) { Handle* h = (Handle*)hv; if (h->is_float) h->f->forward(); else h->d->forward(); }
void orc_step(void* hv, int n) { Handle* h = (Handle*)hv; for (int i = 0; i < n; i++) { if (h->is_float) h->f->step(); else h->d->step(); } }
// returns pointer to the field's storage; *count = elements, *is_int = 1 for int32.  NULL if unknown.
void* orc_field(void* hv, const char* name, int64_t* count, int* is_int) {
  Handle* h = (Handle*)hv;
  auto& f = h->is_float ? h->f->fields : h->d->fields;
  auto it = f.find(name);
  if (it == f.end()) return nullptr;
  *count = it->second.second.first; *is_int = it->second.second.second;
  static double dummy[2];
  return it->second.first ? it->second.first : (void*)dummy;
}
int orc_scalar_size(void* hv) { return ((Handle*)hv)->is_float ? 4 : 8; }

// wrap-geometry known-answer entry: returns wlen (or -1) and the two tangent points
double orc_wrap(const double* x0, const double* x1, const double* gpos, const double* gmat, double radius, int type,
                const double* side, double* wpnt) {
  V3<double> s{0, 0, 0};
  if (side) s = ld3(side);
  return wrapGeom<double>(wpnt, ld3(x0), ld3(x1), ld3(gpos), gmat, radius, type, side ? &s : nullptr);
}
void orc_muscle(double len, double vel, const double* lengthrange, double acc0, const double* prm, double* gain, double* bias) {
  muscleGainBias<double>(len, vel, lengthrange, acc0, prm, gain, bias);
}
double orc_muscle_dynamics(double ctrl, double act, const double* prm) { return muscleDynamics<double>(ctrl, act, prm); }

// CPU baseline: `nworld` independent worlds, `nsteps` mj_steps each, ctrl ~ U[0,1] redrawn every `hold` steps,
// std::thread pool over worlds.  qpos_init: [nworld][nq] initial positions.  Returns final qpos into qpos_out (may be NULL).
int orc_rollout(const myo_model_desc* desc, int nworld, int nsteps, int hold, const double* qpos_init, uint64_t seed,
                int nthreads, double* qpos_out, double* qvel_out) {
  if (nthreads < 1) nthreads = 1;
  std::atomic<int> next(0);
  auto worker = [&]() {
    for (;;) {
      int w = next.fetch_add(1);
      if (w >= nworld) return;
      OrcD o(*desc);
      if (qpos_init) for (int i = 0; i < o.nq; i++) o.qpos[i] = qpos_init[(size_t)w * o.nq + i];
      uint64_t s = seed * 0x9E3779B97F4A7C15ull + (uint64_t)w * 0xBF58476D1CE4E5B9ull + 1;
      for (int t = 0; t < nsteps; t++) {
        if (t % hold == 0) for (int i = 0; i < o.nu; i++) o.ctrl[i] = urand(s);
        o.step();
      }
      if (qpos_out) for (int i = 0; i < o.nq; i++) qpos_out[(size_t)w * o.nq + i] = o.qpos[i];
      if (qvel_out) for (int i = 0; i < o.nv; i++) qvel_out[(size_t)w * o.nv + i] = o.qvel[i];
    }
  };
  std::vector<std::thread> pool;
  for (int t = 1; t < nthreads; t++) pool.emplace_back(worker);
  worker();
  for (auto& t : pool) t.join();
  return 0;
}

// CPU baseline with state carried across calls (one continuing rollout): a pool of worlds, advanced `nsteps` mj_steps per
// call on `nthreads` threads, ctrl ~ U[0,1] redrawn every `hold` steps from a per-world stream.
struct OrcPool { std::vector<std::unique_ptr<OrcD>> w; std::vector<uint64_t> rng; std::vector<int> t; };
void* orc_pool_create(const myo_model_desc* desc, int nworld, const double* qpos_init, uint64_t seed) {
  OrcPool* p = new OrcPool();
  for (int i = 0; i < nworld; i++) {
    p->w.emplace_back(new OrcD(*desc));
    OrcD& o = *p->w.back();
    if (qpos_init) for (int k = 0; k < o.nq; k++) o.qpos[k] = qpos_init[(size_t)i * o.nq + k];
    p->rng.push_back(seed * 0x9E3779B97F4A7C15ull + (uint64_t)i * 0xBF58476D1CE4E5B9ull + 1);
    p->t.push_back(0);
  }
  return p;
}
void orc_pool_destroy(void* pv) { delete (OrcPool*)pv; }
int orc_pool_advance(void* pv, int nsteps, int hold, int nthreads) {
  OrcPool* p = (OrcPool*)pv;
  if (nthreads < 1) nthreads = 1;
  std::atomic<int> next(0);
  const int nworld = (int)p->w.size();
  auto worker = [&]() {
    for (;;) {
      int i = next.fetch_add(1);
      if (i >= nworld) return;
      OrcD& o = *p->w[i];
      for (int k = 0; k < nsteps; k++, p->t[i]++) {
        if (p->t[i] % hold == 0) for (int a = 0; a < o.nu; a++) o.ctrl[a] = urand(p->rng[i]);
        o.step();
      }
    }
  };
  std::vector<std::thread> pool;
  for (int t = 1; t < nthreads; t++) pool.emplace_back(worker);
  worker();
  for (auto& t : pool) t.join();
  return 0;
}
void orc_pool_state(void* pv, double* qpos_out, double* qvel_out) {
  OrcPool* p = (OrcPool*)pv;
  for (size_t i = 0; i < p->w.size(); i++) {
    OrcD& o = *p->w[i];
    for (int k = 0; k < o.nq; k++) qpos_out[i * o.nq + k] = o.qpos[k];
    for (int k = 0; k < o.nv; k++) qvel_out[i * o.nv + k] = o.qvel[k];
  }
}

}  // extern "C"
