// Shared device code of the batched mj_step for sm_100a: small math, the per-world scratch accessor, and the REGULAR phases
// of the step (model-driven loops with no data-dependent trip counts): mj_kinematics, mj_comPos, the position outputs,
// mj_tendon (+mju_wrap), mj_crb / mj_factorM / mj_solveM, mj_fwdVelocity, mj_fwdActuation; plus the collision primitives
// (closed-form pairs, MPR) and the muscle / impedance curves.
//
// LANE = WORLD: a CTA of NW warps owns a tile of 32 consecutive worlds; the items of every loop
// (frames, sites, tendon units, dofs, rows ...) are striped over the WARPS (t.tl = warp index, TFOR), TSYNC is a CTA
// barrier.  Every model-table load is warp-uniform and every access to the scratch or to a [k][nworld] field is one
// coalesced 128-byte line with all 32 lanes useful.
// Per-world scratch, tile-blocked: scratch[tile][k][32] -- element k of the tile's lane l at tile_base + 32 k + l, the
// float offsets k (SmemLayout) computed by the host plan.  The element stride is the compile-time constant 32, so the
// neighbours of an element are immediate offsets of one address register.
//
// Replaces (third-party code the reference calls; call sites in include/myostep.h): mj_kinematics, mj_comPos, mj_tendon
// (+mju_wrap), mj_crb, mj_factorM, mj_transmission, mj_fwdVelocity, mj_fwdActuation, mj_fwdAcceleration.
#ifndef MYO_COMMON_CUH_
#define MYO_COMMON_CUH_
#include <cuda_runtime.h>
#include <math_constants.h>

#include <cstdlib>

#include "devmodel.h"
#include "internal.h"
#define MINVAL 1e-15f
#ifndef GRAD_NOISE
#define GRAD_NOISE 2e-6f   // float32 noise floor of the Newton gradient relative to the magnitude of its terms
#endif
#define TEAM (t.nw)   // item stride of a loop = the NW warps of the tile's CTA (NW = blockDim.x / 32, a launch parameter)
struct TParams { DevModel D; SmemLayout L; FieldPtrs F; TileAux X; };
#define MYO_NSLOT MYO_PARAM_SLOTS
#define MYO_NOALLOC_SLOT (MYO_NSLOT + 1)   // parameters of myo_kinematics_noalloc (tile.cu): no batch, so no batch slot
static_assert(sizeof(TParams) * (MYO_NSLOT + 2) <= 60 * 1024, "parameter slots exceed constant memory");
__constant__ TParams tP[MYO_NSLOT + 2];   // slot MYO_NSLOT stays zero: "materialise nothing" (inner substeps)
namespace {

enum { JNT_FREE = 0, JNT_BALL = 1, JNT_SLIDE = 2, JNT_HINGE = 3 };
enum { G_PLANE = 0, G_HFIELD, G_SPHERE, G_CAPSULE, G_ELLIPSOID, G_CYLINDER, G_BOX, G_MESH };
enum { DYN_NONE = 0, DYN_INTEGRATOR, DYN_FILTER, DYN_FILTEREXACT, DYN_MUSCLE };
enum { GAIN_FIXED = 0, GAIN_AFFINE, GAIN_MUSCLE };
enum { BIAS_NONE = 0, BIAS_AFFINE, BIAS_MUSCLE };
enum { DSBL_CONSTRAINT = 1 << 0, DSBL_EQUALITY = 1 << 1, DSBL_LIMIT = 1 << 3, DSBL_CONTACT = 1 << 4, DSBL_PASSIVE = 1 << 5,
       DSBL_GRAVITY = 1 << 6, DSBL_CLAMPCTRL = 1 << 7, DSBL_WARMSTART = 1 << 8, DSBL_ACTUATION = 1 << 10 };
enum { ROW_EQ = 0, ROW_LIMJ = 1, ROW_LIMT = 2, ROW_CON = 3 };

// ------------------------------------------------------------------------------------------ small math
struct f3 { float x, y, z; };
__device__ __forceinline__ f3 mk3(float x, float y, float z) { f3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ f3 operator+(f3 a, f3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ f3 operator-(f3 a, f3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ f3 operator*(f3 a, float s) { return mk3(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ float dot(f3 a, f3 b) { return fmaf(a.x, b.x, fmaf(a.y, b.y, a.z * b.z)); }
__device__ __forceinline__ f3 cross(f3 a, f3 b) { return mk3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
__device__ __forceinline__ float norm3(f3 a) { return sqrtf(dot(a, a)); }
template <class P> __device__ __forceinline__ f3 ld3(P p) { return mk3(p[0], p[1], p[2]); }
template <class P> __device__ __forceinline__ void st3(P p, f3 v) { p[0] = v.x; p[1] = v.y; p[2] = v.z; }
__device__ __forceinline__ f3 unit3(f3 a) {
  float n = norm3(a);
  return n < MINVAL ? mk3(1.f, 0.f, 0.f) : a * (1.f / n);
}
template <class P> __device__ __forceinline__ f3 mulM(P R, f3 v) {
  return mk3(fmaf(R[0], v.x, fmaf(R[1], v.y, R[2] * v.z)), fmaf(R[3], v.x, fmaf(R[4], v.y, R[5] * v.z)),
             fmaf(R[6], v.x, fmaf(R[7], v.y, R[8] * v.z)));
}
template <class P> __device__ __forceinline__ f3 mulMT(P R, f3 v) {
  return mk3(fmaf(R[0], v.x, fmaf(R[3], v.y, R[6] * v.z)), fmaf(R[1], v.x, fmaf(R[4], v.y, R[7] * v.z)),
             fmaf(R[2], v.x, fmaf(R[5], v.y, R[8] * v.z)));
}
struct f4 { float w, x, y, z; };
template <class P> __device__ __forceinline__ f4 ld4(P p) { f4 q; q.w = p[0]; q.x = p[1]; q.y = p[2]; q.z = p[3]; return q; }
template <class P> __device__ __forceinline__ void st4(P p, f4 q) { p[0] = q.w; p[1] = q.x; p[2] = q.y; p[3] = q.z; }
__device__ __forceinline__ f4 qmul(f4 a, f4 b) {
  f4 r;
  r.w = a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z;
  r.x = a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y;
  r.y = a.w * b.y - a.x * b.z + a.y * b.w + a.z * b.x;
  r.z = a.w * b.z + a.x * b.y - a.y * b.x + a.z * b.w;
  return r;
}
__device__ __forceinline__ f4 qnormalize(f4 q) {
  float n = sqrtf(q.w * q.w + q.x * q.x + q.y * q.y + q.z * q.z);
  if (n < MINVAL) { q.w = 1.f; q.x = q.y = q.z = 0.f; return q; }
  float s = 1.f / n;
  q.w *= s; q.x *= s; q.y *= s; q.z *= s;
  return q;
}
__device__ __forceinline__ void q2m(float* R, f4 q) {
  float w = q.w, x = q.x, y = q.y, z = q.z;
  R[0] = w * w + x * x - y * y - z * z; R[1] = 2.f * (x * y - w * z); R[2] = 2.f * (x * z + w * y);
  R[3] = 2.f * (x * y + w * z); R[4] = w * w - x * x + y * y - z * z; R[5] = 2.f * (y * z - w * x);
  R[6] = 2.f * (x * z - w * y); R[7] = 2.f * (y * z + w * x); R[8] = w * w - x * x - y * y + z * z;
}
// C = A * B (row-major 3x3)
template <class PC, class PA, class PB> __device__ __forceinline__ void mulMM(PC C, PA A, PB B) {
#pragma unroll
  for (int i = 0; i < 3; i++)
#pragma unroll
    for (int j = 0; j < 3; j++) C[3 * i + j] = fmaf(A[3 * i], B[j], fmaf(A[3 * i + 1], B[3 + j], A[3 * i + 2] * B[6 + j]));
}
__device__ __forceinline__ float clipf(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }
// spatial algebra on [angular; linear] 6-vectors; inertia = (Ixx Iyy Izz Ixy Ixz Iyz, m*cx m*cy m*cz, m)
template <class PR, class PI, class PV> __device__ __forceinline__ void inertMulVec(PR r, PI I, PV v) {
  f3 w = ld3(v), l = ld3(v + 3), mc = ld3(I + 6);
  f3 Iw = mk3(I[0] * w.x + I[3] * w.y + I[4] * w.z, I[3] * w.x + I[1] * w.y + I[5] * w.z, I[4] * w.x + I[5] * w.y + I[2] * w.z);
  st3(r, Iw + cross(mc, l));
  st3(r + 3, l * I[9] + cross(w, mc));
}
template <class PR, class PA, class PB> __device__ __forceinline__ void crossMotion(PR r, PA vel, PB v) {
  f3 w = ld3(vel), l = ld3(vel + 3), a = ld3(v), b = ld3(v + 3);
  st3(r, cross(w, a)); st3(r + 3, cross(w, b) + cross(l, a));
}
template <class PR, class PA, class PB> __device__ __forceinline__ void crossForce(PR r, PA vel, PB f) {
  f3 w = ld3(vel), l = ld3(vel + 3), a = ld3(f), b = ld3(f + 3);
  st3(r, cross(w, a) + cross(l, b)); st3(r + 3, cross(w, b));
}

// one out-of-line copy of each heavy libm routine
__device__ __noinline__ void sincos_(float x, float* s, float* c) { sincosf(x, s, c); }
__device__ __noinline__ float asin_(float x) { return asinf(x); }
__device__ __noinline__ float acos_(float x) { return acosf(x); }
__device__ __noinline__ float pow_(float x, float y) { return powf(x, y); }
__device__ __noinline__ float cos_(float x) { return cosf(x); }
__device__ __noinline__ float sin_(float x) { return sinf(x); }

// One world's column of the tile-blocked scratch (or of a [n][32] shared-memory block): element i at p[32 i].
#define SCR 32
struct SP {
  float* p;
  __device__ __forceinline__ float& operator[](int i) const { return p[i * SCR]; }
  __device__ __forceinline__ SP operator+(int k) const { SP r; r.p = p + k * SCR; return r; }
};
// Loads of the contact Jacobian blocks (conJ, conU: ~200 KB per tile and pass).  Measured (8192 MyoHand worlds): loading them
// around L1 (ld.global.cg, -DSTREAM_J) is 1 % SLOWER than plain loads -- the blocks are re-read soon enough for L1 to help.
#ifdef STREAM_J
#define LDJ(ref) __ldcg(&(ref))
#else
#define LDJ(ref) (ref)
#endif
// the same choice for the sweeps that read each element once (J x, J^T f, U = W Jb) and gain nothing from L1
#ifndef NO_STREAM_SWEEPS
#define LDJS(ref) __ldcg(&(ref))   // -0.5 % step time (two A/B runs)
#else
#define LDJS(ref) LDJ(ref)
#endif
struct SPI {
  int* p;
  __device__ __forceinline__ int& operator[](int i) const { return p[i * SCR]; }
};
__device__ __forceinline__ SPI asInt(SP s) { SPI r; r.p = (int*)s.p; return r; }

// t.tl = warp index inside the CTA, t.nw = warps of the CTA; a CTA barrier orders the items of one phase step against the next
struct Tm { int tl, nw; };
#define TSYNC() __syncthreads()
#ifndef TFOR_UNROLL
#define TFOR_UNROLL 1
#endif
#define TFOR_PRAGMA_(x) _Pragma(#x)
#define TFOR_PRAGMA(n) TFOR_PRAGMA_(unroll n)
#define TFOR(i, n) TFOR_PRAGMA(TFOR_UNROLL) for (int i = t.tl; i < (n); i += TEAM)
// the warps are bound by the latency of dependent loads: the hot inner loops are unrolled so that several loads are in flight
#ifndef HOT_UNROLL
#define HOT_UNROLL 2   // 1: +4 % step time, 4: same, 8: +1.3 % (profiles/r2_warps_per_tile.txt)
#endif
#define HOT TFOR_PRAGMA(HOT_UNROLL)


// frame record in the per-world scratch: pos[3], quat[4], mat[9]
#define FR_POS(sm, L, f) ((sm) + (L).frames + 16 * (f))
#define FR_QUAT(sm, L, f) ((sm) + (L).frames + 16 * (f) + 3)
#define FR_MAT(sm, L, f) ((sm) + (L).frames + 16 * (f) + 7)

// batched field access, struct-of-arrays, world fastest: element k of world w lives at base[k * nworld + w]
struct World { size_t w, nw; unsigned w32, nw32; };   // (32-bit copies: field indices k * nworld + w fit 32 bits, myo_batch_create checks)
// Element k of a [k][nworld] field for this lane's world.  Stores are streaming (st.global.cs, evict-first): the 640 MB of
// output fields a step writes are never read back by it, and at plain priority they push the solver's working set (the
// tiles' Jacobian blocks, ~100 MB at 8192 worlds, about the size of the L2) out to DRAM.  Reads (input fields) are plain.
#ifndef NO_STREAM_OUT
struct FRef { float* p; __device__ __forceinline__ void operator=(float v) const { __stcs(p, v); } __device__ __forceinline__ operator float() const { return *p; } };
struct IRef { int* p; __device__ __forceinline__ void operator=(int v) const { __stcs(p, v); } __device__ __forceinline__ operator int() const { return *p; } };
// (the element index in unsigned 32-bit arithmetic: one IMAD instead of the 64-bit multiply's four or five instructions per
// access -- the output kernels are issue-bound on exactly this address arithmetic)
#define OUTF(ptr, k) (FRef{(float*)(ptr) + ((unsigned)(k) * W.nw32 + W.w32)})
#define OUTI(ptr, k) (IRef{(int*)(ptr) + ((unsigned)(k) * W.nw32 + W.w32)})
#else
#define OUTF(ptr, k) ((float*)(ptr))[(size_t)(k) * W.nw + W.w]
#define OUTI(ptr, k) ((int*)(ptr))[(size_t)(k) * W.nw + W.w]
#endif
// ------------------------------------------------------------------------------------------ kinematics
__device__ __noinline__ void kinematics(int slot, SP sm, Tm t) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L;
  if (t.tl == 0) {
    auto p = FR_POS(sm, L, 0);
    p[0] = p[1] = p[2] = 0.f; p[3] = 1.f; p[4] = p[5] = p[6] = 0.f;
    p[7] = 1.f; p[8] = 0.f; p[9] = 0.f; p[10] = 0.f; p[11] = 1.f; p[12] = 0.f; p[13] = 0.f; p[14] = 0.f; p[15] = 1.f;
  }
  TSYNC();
  auto qpos = sm + L.qpos;
  for (int l = 0; l < D.nlevel; l++) {   // frames of one tree depth are independent
    for (int idx = D.lvl_start[l] + t.tl; idx < D.lvl_start[l + 1]; idx += TEAM) {
      const int f = D.lvl_frames[idx], p = D.fr_parent[f], ja = D.fr_jntadr[f], jn = D.fr_jntnum[f];
      f3 pos; f4 quat; float R[9];
      if (jn == 1 && D.jnt_type[ja] == JNT_FREE) {
        const int qa = D.jnt_qposadr[ja];
        pos = ld3(qpos + qa);
        quat = qnormalize(ld4(qpos + qa + 3));
        st4(qpos + qa + 3, quat);
        st3(sm + L.xanchor + 3 * ja, pos); st3(sm + L.xaxis + 3 * ja, ld3(D.jnt_axis + 3 * ja));
      } else {
        pos = ld3(FR_POS(sm, L, p)) + mulM(FR_MAT(sm, L, p), ld3(D.fr_pre_pos + 3 * f));
        quat = qmul(ld4(FR_QUAT(sm, L, p)), ld4(D.fr_pre_quat + 4 * f));
        q2m(R, quat);
        for (int j = ja; j < ja + jn; j++) {
          const f3 lpos = ld3(D.jnt_pos + 3 * j), laxis = ld3(D.jnt_axis + 3 * j);
          const f3 anchor = mulM(R, lpos) + pos, axis = mulM(R, laxis);
          st3(sm + L.xanchor + 3 * j, anchor); st3(sm + L.xaxis + 3 * j, axis);
          const int qa = D.jnt_qposadr[j];
          const float d = qpos[qa] - D.qpos0[qa];
          if (D.jnt_type[j] == JNT_SLIDE) pos = pos + axis * d;
          else {
            float sn, cs; sincos_(0.5f * d, &sn, &cs);
            f4 ql; ql.w = cs; ql.x = laxis.x * sn; ql.y = laxis.y * sn; ql.z = laxis.z * sn;
            quat = qmul(quat, ql);
            q2m(R, quat);
            pos = anchor - mulM(R, lpos);
          }
        }
      }
      quat = qnormalize(quat);
      q2m(R, quat);
      auto o = FR_POS(sm, L, f);
      st3(o, pos); st4(o + 3, quat);
#pragma unroll
      for (int k = 0; k < 9; k++) o[7 + k] = R[k];
    }
    TSYNC();
  }
}

// composite inertia of every frame about its tree's centre of mass, motion axes (mj_comPos over weld groups)
__device__ __noinline__ void comPos(int slot, SP sm, Tm t) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L;
  auto cin = sm + L.cinert;
  TFOR(f, D.nframe) {
    f3 cw = ld3(FR_POS(sm, L, f)) + mulM(FR_MAT(sm, L, f), ld3(D.fr_com + 3 * f));
    st3(cin + 10 * f + 6, cw); cin[10 * f + 9] = D.fr_mass[f];
  }
  TSYNC();
  TFOR(r, D.nroot) {
    f3 acc = ld3(D.root_const + 4 * r);
    const float M = D.root_const[4 * r + 3];
    // #pragma unroll 1
    for (int f = D.root_fstart[r]; f < D.root_fend[r]; f++) acc = acc + ld3(cin + 10 * f + 6) * cin[10 * f + 9];
    if (M < MINVAL) acc = mk3(0.f, 0.f, 0.f); else acc = acc * (1.f / M);
    st3(sm + L.rootcom + 3 * r, acc);
  }
  TSYNC();
  TFOR(f, D.nframe) {
    auto c = cin + 10 * f;
    if (f == 0) { for (int k = 0; k < 10; k++) c[k] = 0.f; continue; }
    auto R = FR_MAT(sm, L, f);
    auto I = D.fr_inertia + 6 * f;
    const float m = c[9];
    const f3 off = ld3(c + 6) - ld3(sm + L.rootcom + 3 * D.fr_root[f]);
    float T[9];
#pragma unroll
    for (int i = 0; i < 3; i++) {
      T[3 * i] = R[3 * i] * I[0] + R[3 * i + 1] * I[3] + R[3 * i + 2] * I[4];
      T[3 * i + 1] = R[3 * i] * I[3] + R[3 * i + 1] * I[1] + R[3 * i + 2] * I[5];
      T[3 * i + 2] = R[3 * i] * I[4] + R[3 * i + 1] * I[5] + R[3 * i + 2] * I[2];
    }
    const float xx = T[0] * R[0] + T[1] * R[1] + T[2] * R[2], yy = T[3] * R[3] + T[4] * R[4] + T[5] * R[5],
                zz = T[6] * R[6] + T[7] * R[7] + T[8] * R[8], xy = T[0] * R[3] + T[1] * R[4] + T[2] * R[5],
                xz = T[0] * R[6] + T[1] * R[7] + T[2] * R[8], yz = T[3] * R[6] + T[4] * R[7] + T[5] * R[8];
    c[0] = xx + m * (off.y * off.y + off.z * off.z); c[1] = yy + m * (off.x * off.x + off.z * off.z);
    c[2] = zz + m * (off.x * off.x + off.y * off.y); c[3] = xy - m * off.x * off.y; c[4] = xz - m * off.x * off.z;
    c[5] = yz - m * off.y * off.z; c[6] = m * off.x; c[7] = m * off.y; c[8] = m * off.z;
  }
  TFOR(j, D.njnt) {
    const int f = D.jnt_frame[j], da = D.jnt_dofadr[j];
    const f3 off = ld3(sm + L.rootcom + 3 * D.fr_root[f]) - ld3(sm + L.xanchor + 3 * j);
    const f3 axis = ld3(sm + L.xaxis + 3 * j);
    auto cd = sm + L.cdof + 6 * da;
    const int ty = D.jnt_type[j];
    if (ty == JNT_FREE) {
      auto R = FR_MAT(sm, L, f);
#pragma unroll
      for (int k = 0; k < 3; k++) {
        st3(cd + 6 * k, mk3(0.f, 0.f, 0.f)); st3(cd + 6 * k + 3, mk3(k == 0, k == 1, k == 2));
        const f3 ax = mk3(R[k], R[3 + k], R[6 + k]);
        st3(cd + 6 * (3 + k), ax); st3(cd + 6 * (3 + k) + 3, cross(ax, off));
      }
    } else if (ty == JNT_SLIDE) {
      st3(cd, mk3(0.f, 0.f, 0.f)); st3(cd + 3, axis);
    } else {
      st3(cd, axis); st3(cd + 3, cross(axis, off));
    }
  }
  TSYNC();
}

// ------------------------------------------------------------------------------------------ per-world model overrides
// geom_pos of geom g relative to its FRAME (the moving body its weld group hangs on): the compiled table, or -- when the
// caller bound MYO_F_GEOM_POS_OVR and this world's entry is not NaN -- the world's own body-local position carried through
// the body's constant offset to its frame (reference: reorient_sar_v0.py:241-245 rewrites geom_pos per episode).
__device__ __forceinline__ f3 geomLocalPos(const DevModel& D, const void* posovr, World W, int g) {
  if (posovr) {
    const float x = OUTF(posovr, 3 * g);
    if (x == x) {
      const int b = D.geom_body[g];
      return ld3(D.body_off_pos + 3 * b) + mulM(D.body_off_mat + 9 * b, mk3(x, OUTF(posovr, 3 * g + 1), OUTF(posovr, 3 * g + 2)));
    }
  }
  return ld3(D.geom_lpos + 3 * g);
}
// bounding-sphere radius of a primitive (mj_setConst rule for geom_rbound)
__device__ __forceinline__ float geomRbound(int type, f3 s) {
  switch (type) {
    case G_SPHERE: return s.x;
    case G_CAPSULE: return s.x + s.y;
    case G_ELLIPSOID: return fmaxf(s.x, fmaxf(s.y, s.z));
    case G_CYLINDER: return sqrtf(s.x * s.x + s.y * s.y);
    case G_BOX: return norm3(s);
    default: return 0.f;
  }
}

// ------------------------------------------------------------------------------------------ mjData position outputs
// Written straight to global memory (struct-of-arrays, coalesced across the warp's worlds); nothing here feeds the dynamics.
__device__ __noinline__ void writePositionOutputs(int slot, int fslot, SP sm, Tm t, World W) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L; const FieldPtrs& F = tP[fslot].F;
  void* xpos = F.p[MYO_F_XPOS]; void* xquat = F.p[MYO_F_XQUAT]; void* xmat = F.p[MYO_F_XMAT];
  void* xipos = F.p[MYO_F_XIPOS]; void* ximat = F.p[MYO_F_XIMAT]; void* cinert = F.p[MYO_F_CINERT];
  void* scom = F.p[MYO_F_SUBTREE_COM]; void* crb = F.p[MYO_F_CRB];
  const int nb = D.nbody;
  auto tb = sm + L.tmpbody;   // [nbody][13]: m * xipos (3) + body-level cinert (10)
  if (xpos || xquat || xmat || xipos || ximat || cinert || scom || crb) {
    TFOR(b, nb) {
      const int f = D.body_frame[b];
      auto R = FR_MAT(sm, L, f);
      const f3 pos = ld3(FR_POS(sm, L, f)) + mulM(R, ld3(D.body_off_pos + 3 * b));
      float M[9]; mulMM(M, R, D.body_off_mat + 9 * b);
      if (xpos) { OUTF(xpos, 3 * b) = pos.x; OUTF(xpos, 3 * b + 1) = pos.y; OUTF(xpos, 3 * b + 2) = pos.z; }
      if (xquat) {
        const f4 q = qnormalize(qmul(ld4(FR_QUAT(sm, L, f)), ld4(D.body_off_quat + 4 * b)));
        OUTF(xquat, 4 * b) = q.w; OUTF(xquat, 4 * b + 1) = q.x; OUTF(xquat, 4 * b + 2) = q.y; OUTF(xquat, 4 * b + 3) = q.z;
      }
      if (xmat)
#pragma unroll
        for (int k = 0; k < 9; k++) OUTF(xmat, 9 * b + k) = M[k];
      const f3 ip = pos + mulM(M, ld3(D.body_ipos + 3 * b));
      if (xipos) { OUTF(xipos, 3 * b) = ip.x; OUTF(xipos, 3 * b + 1) = ip.y; OUTF(xipos, 3 * b + 2) = ip.z; }
      const float m = D.body_mass[b];
      auto tt = tb + 13 * b;
      tt[0] = m * ip.x; tt[1] = m * ip.y; tt[2] = m * ip.z;
      if (ximat || cinert || crb) {
        float IM[9]; mulMM(IM, M, D.body_imat + 9 * b);
        if (ximat)
#pragma unroll
          for (int k = 0; k < 9; k++) OUTF(ximat, 9 * b + k) = IM[k];
        float c[10];
        if (b == 0) { for (int k = 0; k < 10; k++) c[k] = 0.f; }
        else {
          const float i0 = D.body_inertia[3 * b], i1 = D.body_inertia[3 * b + 1], i2 = D.body_inertia[3 * b + 2];
          const f3 off = ip - ld3(sm + L.rootcom + 3 * D.body_root[b]);
          c[0] = IM[0] * IM[0] * i0 + IM[1] * IM[1] * i1 + IM[2] * IM[2] * i2 + m * (off.y * off.y + off.z * off.z);
          c[1] = IM[3] * IM[3] * i0 + IM[4] * IM[4] * i1 + IM[5] * IM[5] * i2 + m * (off.x * off.x + off.z * off.z);
          c[2] = IM[6] * IM[6] * i0 + IM[7] * IM[7] * i1 + IM[8] * IM[8] * i2 + m * (off.x * off.x + off.y * off.y);
          c[3] = IM[0] * IM[3] * i0 + IM[1] * IM[4] * i1 + IM[2] * IM[5] * i2 - m * off.x * off.y;
          c[4] = IM[0] * IM[6] * i0 + IM[1] * IM[7] * i1 + IM[2] * IM[8] * i2 - m * off.x * off.z;
          c[5] = IM[3] * IM[6] * i0 + IM[4] * IM[7] * i1 + IM[5] * IM[8] * i2 - m * off.y * off.z;
          c[6] = m * off.x; c[7] = m * off.y; c[8] = m * off.z; c[9] = m;
        }
#pragma unroll
        for (int k = 0; k < 10; k++) { tt[3 + k] = c[k]; if (cinert) OUTF(cinert, 10 * b + k) = c[k]; }
      }
    }
    if (scom || crb) {
      TSYNC();
      // subtree sums: bodies are depth-first, a subtree is the contiguous range [b, body_subend[b])
      TFOR(b, nb) {
        const int e = D.body_subend[b];
        if (scom) {
          const float Ms = D.body_subtreemass[b];
          f3 c;
          if (Ms < MINVAL) {   // MuJoCo: subtree_com of a massless subtree = xipos
            const int f = D.body_frame[b];
            auto R = FR_MAT(sm, L, f);
            float M[9]; mulMM(M, R, D.body_off_mat + 9 * b);
            c = ld3(FR_POS(sm, L, f)) + mulM(R, ld3(D.body_off_pos + 3 * b)) + mulM(M, ld3(D.body_ipos + 3 * b));
          } else {
            f3 acc = mk3(0.f, 0.f, 0.f);
            for (int q = b; q < e; q++) acc = acc + ld3(tb + 13 * q);
            c = acc * (1.f / Ms);
          }
          OUTF(scom, 3 * b) = c.x; OUTF(scom, 3 * b + 1) = c.y; OUTF(scom, 3 * b + 2) = c.z;
        }
        if (crb) {
          float acc[10];
#pragma unroll
          for (int k = 0; k < 10; k++) acc[k] = 0.f;
          if (b > 0)
            for (int q = b; q < e; q++)
#pragma unroll
              for (int k = 0; k < 10; k++) acc[k] += tb[13 * q + 3 + k];
#pragma unroll
          for (int k = 0; k < 10; k++) OUTF(crb, 10 * b + k) = acc[k];
        }
      }
    }
  }
  void* xa = F.p[MYO_F_XANCHOR]; void* xx = F.p[MYO_F_XAXIS];
  if (xa) TFOR(i, 3 * D.njnt) OUTF(xa, i) = sm[L.xanchor + i];
  if (xx) TFOR(i, 3 * D.njnt) OUTF(xx, i) = sm[L.xaxis + i];
  void* cd = F.p[MYO_F_CDOF];
  if (cd) TFOR(i, 6 * D.nv) OUTF(cd, i) = sm[L.cdof + i];
  void* gp = F.p[MYO_F_GEOM_XPOS]; void* gm = F.p[MYO_F_GEOM_XMAT];
  if (gp || gm) {
    TFOR(g, D.ngeom) {
      const int f = D.geom_frame[g];
      auto R = FR_MAT(sm, L, f);
      if (gp) {
        const f3 p = ld3(FR_POS(sm, L, f)) + mulM(R, geomLocalPos(D, tP[slot].F.p[MYO_F_GEOM_POS_OVR], W, g));
        OUTF(gp, 3 * g) = p.x; OUTF(gp, 3 * g + 1) = p.y; OUTF(gp, 3 * g + 2) = p.z;
      }
      if (gm) {
        float M[9]; mulMM(M, R, D.geom_lmat + 9 * g);
#pragma unroll
        for (int k = 0; k < 9; k++) OUTF(gm, 9 * g + k) = M[k];
      }
    }
  }
  void* sp = F.p[MYO_F_SITE_XPOS]; void* smt = F.p[MYO_F_SITE_XMAT];
  if (sp || smt) {
    TFOR(q, D.nsite) {
      const int f = D.site_frame[q];
      auto R = FR_MAT(sm, L, f);
      if (sp) {
        const f3 p = ld3(FR_POS(sm, L, f)) + mulM(R, ld3(D.site_lpos + 3 * q));
        OUTF(sp, 3 * q) = p.x; OUTF(sp, 3 * q + 1) = p.y; OUTF(sp, 3 * q + 2) = p.z;
      }
      if (smt) {
        float M[9]; mulMM(M, R, D.site_lmat + 9 * q);
#pragma unroll
        for (int k = 0; k < 9; k++) OUTF(smt, 9 * q + k) = M[k];
      }
    }
  }
  TSYNC();   // tmpbody overlays the tendon stage's scratch
}

// ------------------------------------------------------------------------------------------ tendon wrapping (mju_wrap)
__device__ __forceinline__ bool segIntersect(const float* p1, const float* p2, const float* p3, const float* p4) {
  const float det = (p4[1] - p3[1]) * (p2[0] - p1[0]) - (p4[0] - p3[0]) * (p2[1] - p1[1]);
  if (fabsf(det) < MINVAL) return false;
  const float a = ((p4[0] - p3[0]) * (p1[1] - p3[1]) - (p4[1] - p3[1]) * (p1[0] - p3[0])) / det;
  const float b = ((p2[0] - p1[0]) * (p1[1] - p3[1]) - (p2[1] - p1[1]) * (p1[0] - p3[0])) / det;
  if (a > 1.f - 1e-4f && b > 1.f - 1e-4f) return false;   // the parts only meet at their end points (grazing path): not a crossing, see oracle/mjstep_oracle.cpp
  return a >= 0.f && a <= 1.f && b >= 0.f && b <= 1.f;
}
__device__ __noinline__ float wrapCircle(float* pnt, const float* d, const float* sd, float rad) {
  const float sq0 = d[0] * d[0] + d[1] * d[1], sq1 = d[2] * d[2] + d[3] * d[3], sqr = rad * rad;
  const float dx = d[2] - d[0], dy = d[3] - d[1];
  const float dd = dx * dx + dy * dy;
  float a = -(dx * d[0] + dy * d[1]) / fmaxf(MINVAL, dd);
  a = clipf(a, 0.f, 1.f);
  const float nx = d[0] + a * dx, ny = d[1] + a * dy;
  // side-site rule with a tie band (oracle/mjstep_oracle.cpp wrapCircle has the reasoning): within 1e-5 of the product's scale
  // the side site counts as "same side", so float32 and float64 take the same branch at joint-limit poses
  const float nn = nx * nx + ny * ny;
  if (nn > sqr && (!sd || sd[0] * nx + sd[1] * ny >= -1e-5f * rad * sqrtf(nn))) return -1.f;
  if (sq0 < sqr || sq1 < sqr) return -1.f;
  const float s0 = sqrtf(sq0 - sqr), s1 = sqrtf(sq1 - sqr);
  float sol[2][4], good[2];
#pragma unroll
  for (int i = 0; i < 2; i++) {
    const float sg = i == 0 ? 1.f : -1.f;
    sol[i][0] = (d[0] * sqr + sg * rad * d[1] * s0) / sq0;
    sol[i][1] = (d[1] * sqr - sg * rad * d[0] * s0) / sq0;
    sol[i][2] = (d[2] * sqr - sg * rad * d[3] * s1) / sq1;
    sol[i][3] = (d[3] * sqr + sg * rad * d[2] * s1) / sq1;
    if (sd) {
      const float tx = sol[i][0] + sol[i][2], ty = sol[i][1] + sol[i][3];
      good[i] = (tx * sd[0] + ty * sd[1]) / fmaxf(MINVAL, sqrtf(tx * tx + ty * ty));
    } else {
      const float tx = sol[i][0] - sol[i][2], ty = sol[i][1] - sol[i][3];
      good[i] = -(tx * tx + ty * ty);
    }
    if (segIntersect(d, sol[i], d + 2, sol[i] + 2)) good[i] = -10000.f;
  }
  const int i = good[0] > good[1] ? 0 : 1;
#pragma unroll
  for (int k = 0; k < 4; k++) pnt[k] = i == 0 ? sol[0][k] : sol[1][k];
  if (segIntersect(d, pnt, d + 2, pnt + 2)) return -1.f;
  const float n0 = fmaxf(MINVAL, sqrtf(pnt[0] * pnt[0] + pnt[1] * pnt[1])), n1 = fmaxf(MINVAL, sqrtf(pnt[2] * pnt[2] + pnt[3] * pnt[3]));
  float ang = acos_(clipf((pnt[0] * pnt[2] + pnt[1] * pnt[3]) / (n0 * n1), -1.f, 1.f));
  const float cr = pnt[1] * pnt[2] - pnt[0] * pnt[3];
  // coincident tangent points (grazing path): cr is rounding noise there and must not flip a ~0 arc into a ~2 pi one
  if (fabsf(cr) > 1e-5f * sqr && ((cr > 0.f && i) || (cr < 0.f && !i))) ang = 2.f * CUDART_PI_F - ang;
  return rad * ang;
}
__device__ __noinline__ float wrapInside(float* pnt, const float* d, float rad) {
  const float l0 = sqrtf(d[0] * d[0] + d[1] * d[1]), l1 = sqrtf(d[2] * d[2] + d[3] * d[3]);
  if (l0 <= rad || l1 <= rad) return -1.f;
  const float dx = d[2] - d[0], dy = d[3] - d[1];
  const float dd = dx * dx + dy * dy;
  if (dd > MINVAL) {
    const float a = -(dx * d[0] + dy * d[1]) / dd;
    if (a > 0.f && a < 1.f) {
      const float tx = d[0] + a * dx, ty = d[1] + a * dy;
      if (sqrtf(tx * tx + ty * ty) <= rad) return -1.f;
    }
  }
  const float mx = 0.5f * (d[0] + d[2]), my = 0.5f * (d[1] + d[3]);
  const float mn = fmaxf(MINVAL, sqrtf(mx * mx + my * my));
  pnt[0] = pnt[2] = rad * mx / mn; pnt[1] = pnt[3] = rad * my / mn;
  const float A = rad / l0, B = rad / l1;
  const float cosG = (l0 * l0 + l1 * l1 - dd) / (2.f * l0 * l1);
  if (cosG < -1.f + 1e-7f) return -1.f;
  if (cosG > 1.f - 1e-7f) return 0.f;
  const float G = acos_(cosG);
  // Root of asin(A z) + asin(B z) - 2 asin(z) + G on z in (0,1).  MuJoCo iterates on z from z = 1 - 1e-7, where
  // d/dz blows up and float32 cannot resolve the root; we iterate on t = acos(z) instead (same function, same
  // root, same bail-out tests translated: df/dz > 0 <=> dg/dt < 0, z increasing <=> t decreasing).
  float t = 4.4721e-4f;  // acos(1 - 1e-7)
  float ct = cos_(t);
  float f = asin_(A * ct) + asin_(B * ct) + 2.f * t - CUDART_PI_F + G;
  if (f > 0.f) return 0.f;
  int it = 0;
  // #pragma unroll 1
  for (; it < 20 && fabsf(f) > 1e-6f; it++) {
    const float st = sin_(t);
    const float dg = 2.f - st * (A / fmaxf(1e-12f, sqrtf(1.f - ct * ct * A * A)) + B / fmaxf(1e-12f, sqrtf(1.f - ct * ct * B * B)));
    if (dg < 1e-12f) return 0.f;
    const float t1 = t - f / dg;
    if (t1 < t) return 0.f;
    t = t1; ct = cos_(t);
    f = asin_(A * ct) + asin_(B * ct) + 2.f * t - CUDART_PI_F + G;
    if (f > 1e-5f) return 0.f;
  }
  if (it >= 20) return 0.f;
  float vx, vy, ang;
  if (d[0] * d[3] - d[1] * d[2] > 0.f) { vx = d[0] / l0; vy = d[1] / l0; ang = 0.5f * CUDART_PI_F - t - asin_(A * ct); }
  else { vx = d[2] / l1; vy = d[3] / l1; ang = 0.5f * CUDART_PI_F - t - asin_(B * ct); }
  float sn, cs; sincos_(ang, &sn, &cs);
  pnt[0] = pnt[2] = rad * (cs * vx - sn * vy);
  pnt[1] = pnt[3] = rad * (sn * vx + cs * vy);
  return 0.f;
}
// returns wrap length or -1; w0/w1 = world tangent points
__device__ __noinline__ float wrapGeom(f3 x0, f3 x1, f3 gpos, const float* gmat, float radius, int type, bool hasSide, f3 side, f3* w0, f3* w1) {
  if (radius < MINVAL) return -1.f;
  const f3 p0 = mulMT(gmat, x0 - gpos), p1 = mulMT(gmat, x1 - gpos);
  if (norm3(p0) < MINVAL || norm3(p1) < MINVAL) return -1.f;
  f3 ax0, ax1;
  if (type == UNIT_SPHERE) {
    ax0 = unit3(p0);
    f3 nrm = cross(p0, p1);
    const float n = norm3(nrm);
    if (n < MINVAL) {
      const float a0 = fabsf(ax0.x), a1 = fabsf(ax0.y), a2 = fabsf(ax0.z);
      const int k = (a0 >= a1 && a0 >= a2) ? 0 : (a1 >= a2 ? 1 : 2);
      nrm = unit3(cross(ax0, mk3(k != 0, k != 1, k != 2)));
    } else nrm = nrm * (1.f / n);
    ax1 = unit3(cross(nrm, ax0));
  } else { ax0 = mk3(1.f, 0.f, 0.f); ax1 = mk3(0.f, 1.f, 0.f); }
  float d[4] = {dot(p0, ax0), dot(p0, ax1), dot(p1, ax0), dot(p1, ax1)};
  float sd[2] = {0.f, 0.f};
  bool inside = false;
  if (hasSide) {
    const f3 s = mulMT(gmat, side - gpos);
    sd[0] = dot(s, ax0); sd[1] = dot(s, ax1);
    const float n = radius / fmaxf(MINVAL, sqrtf(sd[0] * sd[0] + sd[1] * sd[1]));
    sd[0] *= n; sd[1] *= n;
    inside = norm3(s) < radius;
  }
  float pnt[4];
  float wlen = inside ? wrapInside(pnt, d, radius) : wrapCircle(pnt, d, hasSide ? sd : nullptr, radius);
  if (wlen < 0.f) return -1.f;
  f3 r0 = ax0 * pnt[0] + ax1 * pnt[1], r1 = ax0 * pnt[2] + ax1 * pnt[3];
  if (type == UNIT_CYLINDER) {
    const float L0 = sqrtf((p0.x - r0.x) * (p0.x - r0.x) + (p0.y - r0.y) * (p0.y - r0.y));
    const float L1 = sqrtf((p1.x - r1.x) * (p1.x - r1.x) + (p1.y - r1.y) * (p1.y - r1.y));
    const float tot = fmaxf(MINVAL, L0 + wlen + L1);
    r0.z = p0.z + (p1.z - p0.z) * L0 / tot;
    r1.z = p0.z + (p1.z - p0.z) * (L0 + wlen) / tot;
    const float h = r1.z - r0.z;
    wlen = sqrtf(wlen * wlen + h * h);
  }
  *w0 = mulM(gmat, r0) + gpos; *w1 = mulM(gmat, r1) + gpos;
  return wlen;
}

// mj_tendon: path lengths + sparse moment arms
__device__ __noinline__ void tendons(int slot, int fslot, SP sm, Tm t, World W) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L; const FieldPtrs& F = tP[fslot].F;
  void* wx = F.p[MYO_F_WRAP_XPOS]; void* wo = F.p[MYO_F_WRAP_OBJ]; void* wa = F.p[MYO_F_TEN_WRAPADR]; void* wn = F.p[MYO_F_TEN_WRAPNUM];
  const bool wrout = wx || wo || wa || wn;                 // path points requested: tendonOutputs (tenoutKernel) writes them
  const bool wantw = wrout;                                 // ... from the tangent points and unit sites kept here
  TFOR(uo, D.nunit) {
    const int u = D.unit_order[uo];
    const int s0 = D.unit_s0[u], s1 = D.unit_s1[u], g = D.unit_geom[u];
    const int f0 = D.site_frame[s0], f1 = D.site_frame[s1];
    const f3 p0 = ld3(FR_POS(sm, L, f0)) + mulM(FR_MAT(sm, L, f0), ld3(D.site_lpos + 3 * s0));
    const f3 p1 = ld3(FR_POS(sm, L, f1)) + mulM(FR_MAT(sm, L, f1), ld3(D.site_lpos + 3 * s1));
    float wlen = -1.f; f3 w0 = p0, w1 = p1;
    if (g >= 0) {
      const int fg = D.geom_frame[g], sid = D.unit_side[u];
      const f3 gpos = ld3(FR_POS(sm, L, fg)) + mulM(FR_MAT(sm, L, fg), ld3(D.geom_lpos + 3 * g));
      float gm[9]; mulMM(gm, FR_MAT(sm, L, fg), D.geom_lmat + 9 * g);
      f3 side = mk3(0.f, 0.f, 0.f);
      if (sid >= 0) { const int fs = D.site_frame[sid]; side = ld3(FR_POS(sm, L, fs)) + mulM(FR_MAT(sm, L, fs), ld3(D.site_lpos + 3 * sid)); }
      wlen = wrapGeom(p0, p1, gpos, gm, D.unit_radius[u], D.unit_type[u], sid >= 0, side, &w0, &w1);
      if (wantw && wlen >= 0.f) { auto wp = sm + L.uwpt + 6 * D.unit_gord[u]; st3(wp, w0); st3(wp + 3, w1); }
    }
    const int sl0 = D.unit_slot[2 * u], sl1 = D.unit_slot[2 * u + 1];
    float len;
    if (wlen < 0.f) {
      const f3 dif = p1 - p0; const float n = norm3(dif);
      len = n;
      if (sl0 >= 0) { st3(sm + L.slots + 6 * sl0, p0); st3(sm + L.slots + 6 * sl0 + 3, n > MINVAL ? dif * (1.f / n) : mk3(0.f, 0.f, 0.f)); }
    } else {
      const f3 d0 = w0 - p0, d1 = p1 - w1; const float n0 = norm3(d0), n1 = norm3(d1);
      len = n0 + wlen + n1;
      if (sl0 >= 0) { st3(sm + L.slots + 6 * sl0, p0); st3(sm + L.slots + 6 * sl0 + 3, n0 > MINVAL ? d0 * (1.f / n0) : mk3(0.f, 0.f, 0.f)); }
      if (sl1 >= 0) { st3(sm + L.slots + 6 * sl1, w1); st3(sm + L.slots + 6 * sl1 + 3, n1 > MINVAL ? d1 * (1.f / n1) : mk3(0.f, 0.f, 0.f)); }
    }
    sm[L.ulen + u] = len / D.unit_div[u];
    sm[L.uwrap + u] = wlen < 0.f ? 0.f : 1.f;
    if (wrout) { st3(sm + L.upts + 6 * u, p0); if (D.unit_end[u]) st3(sm + L.upts + 6 * u + 3, p1); }   // the unit's sites, for the path points
  }
  TSYNC();
  TFOR(n, D.nnz) {
    float acc = 0.f;
    for (int e = D.nz_estart[n]; e < D.nz_estart[n + 1]; e++) {
      const int kind = D.ent_kind[e];
      if (kind != ENT_ALWAYS) {
        const bool wrapped = sm[L.uwrap + D.ent_unit[e]] != 0.f;
        if ((kind == ENT_WRAP) != wrapped) continue;
      }
      const int i = D.ent_dof[e], sl = D.ent_slot[e];
      const f3 p = ld3(sm + L.slots + 6 * sl), dir = ld3(sm + L.slots + 6 * sl + 3);
      const f3 com = ld3(sm + L.rootcom + 3 * D.fr_root[D.dof_frame[i]]);
      auto cd = sm + L.cdof + 6 * i;
      acc += D.ent_scale[e] * (dot(ld3(cd), cross(p - com, dir)) + dot(ld3(cd + 3), dir));
    }
    sm[L.tenJ + n] = acc;
  }
  TFOR(k, D.nten) {
    float acc = 0.f;
    for (int u = D.ten_ustart[k]; u < D.ten_ustart[k + 1]; u++) acc += sm[L.ulen + u];
    sm[L.tenlen + k] = acc;
  }
  TSYNC();
}
// ------------------------------------------------------------------------------------------ inertia: mj_crb, mj_factorM, mj_solveM
// sparse L^T D L in place over the dof-tree pattern; one precomputed multiply-subtract list per column
// The factorisation's index tables (see LdlTab below: same reason, same staging).
struct LdlFacTab { const int *ldl_start, *ldl_dst, *ldl_a, *ldl_b, *dof_Madr, *dof_rowlen; };
__device__ __forceinline__ LdlFacTab ldlFacTabGlobal(const DevModel& D) {
  LdlFacTab T; T.ldl_start = D.ldl_start; T.ldl_dst = D.ldl_dst; T.ldl_a = D.ldl_a; T.ldl_b = D.ldl_b; T.dof_Madr = D.dof_Madr; T.dof_rowlen = D.dof_rowlen; return T;
}
__device__ __noinline__ void factorLDL(int slot, SP LD, SP dinv, Tm t, const LdlFacTab& T) {
  const DevModel& D = tP[slot].D;
  const int nv = D.nv;
  for (int pos = 0; pos < nv; pos++) {
    const int k = nv - 1 - pos, akk = T.dof_Madr[k];
    const float inv = 1.f / LD[akk];
    for (int o = T.ldl_start[pos] + t.tl; o < T.ldl_start[pos + 1]; o += TEAM) LD[T.ldl_dst[o]] -= LD[T.ldl_a[o]] * LD[T.ldl_b[o]] * inv;
    TSYNC();
    const int d = T.dof_rowlen[k];
    for (int q = 1 + t.tl; q < d; q += TEAM) LD[akk + q] *= inv;
    if (t.tl == 0) dinv[k] = inv;
    TSYNC();
  }
}
__device__ __forceinline__ void factorLDL(int slot, SP LD, SP dinv, Tm t) { factorLDL(slot, LD, dinv, t, ldlFacTabGlobal(tP[slot].D)); }
// two matrices of the same pattern (M and Euler's M + h B) in one pass over the columns: half the barriers
__device__ __noinline__ void factorLDL2(int slot, SP LDa, SP dinva, SP LDb, SP dinvb, Tm t, const LdlFacTab& T) {
  const DevModel& D = tP[slot].D;
  const int nv = D.nv;
  for (int pos = 0; pos < nv; pos++) {
    const int k = nv - 1 - pos, akk = T.dof_Madr[k];
    const float inva = 1.f / LDa[akk], invb = 1.f / LDb[akk];
    for (int o = T.ldl_start[pos] + t.tl; o < T.ldl_start[pos + 1]; o += TEAM) {
      const int dst = T.ldl_dst[o], ia = T.ldl_a[o], ib = T.ldl_b[o];
      LDa[dst] -= LDa[ia] * LDa[ib] * inva; LDb[dst] -= LDb[ia] * LDb[ib] * invb;
    }
    TSYNC();
    const int d = T.dof_rowlen[k];
    for (int q = 1 + t.tl; q < d; q += TEAM) { LDa[akk + q] *= inva; LDb[akk + q] *= invb; }
    if (t.tl == 0) { dinva[k] = inva; dinvb[k] = invb; }
    TSYNC();
  }
}
// The solve's index tables.  Each level step is a chain of dependent table lookups (level -> dof -> entry list -> entry) in
// front of very little arithmetic; from global memory, with L1 swept by the Jacobian streams, that chain is what the other
// warps wait for at every level's barrier (66 % barrier stalls).  A kernel that has shared memory to spare copies the tables
// there once (ldlTabStage, ldlTabInts ints) and the solve reads them at shared-memory latency.
struct LdlTab { const int *dlvl_start, *dlvl_dofs, *desc_start, *desc_dof, *desc_idx, *dof_Madr, *dof_rowlen, *M_col; };
__device__ __forceinline__ LdlTab ldlTabGlobal(const DevModel& D) {
  LdlTab T; T.dlvl_start = D.dlvl_start; T.dlvl_dofs = D.dlvl_dofs; T.desc_start = D.desc_start; T.desc_dof = D.desc_dof; T.desc_idx = D.desc_idx;
  T.dof_Madr = D.dof_Madr; T.dof_rowlen = D.dof_rowlen; T.M_col = D.M_col; return T;
}
__host__ __device__ __forceinline__ int ldlTabInts(int nv, int nM, int ndlevel) { return (ndlevel + 1) + nv + (nv + 1) + 2 * (nM - nv) + 2 * nv + nM; }
// cooperative copy by the whole CTA; the caller synchronises before the first solve
// (by the LIVE lanes of the tile's warps: the threads of a partial or masked tile's other lanes have exited)
__device__ __forceinline__ LdlTab ldlTabStage(const DevModel& D, int* buf, Tm t) {
  const int nv = D.nv, nd = D.nM - nv;
  int* p = buf;
  const unsigned live = __activemask();
  const int nlive = __popc(live), rank = __popc(live & ((1u << (threadIdx.x & 31)) - 1u));
  auto put = [&](const int* src, int n) { for (int i = t.tl * nlive + rank; i < n; i += t.nw * nlive) p[i] = src[i]; const int* r = p; p += n; return r; };
  LdlTab T;
  T.dlvl_start = put(D.dlvl_start, D.ndlevel + 1); T.dlvl_dofs = put(D.dlvl_dofs, nv); T.desc_start = put(D.desc_start, nv + 1);
  T.desc_dof = put(D.desc_dof, nd); T.desc_idx = put(D.desc_idx, nd); T.dof_Madr = put(D.dof_Madr, nv); T.dof_rowlen = put(D.dof_rowlen, nv);
  T.M_col = put(D.M_col, D.nM);
  return T;
}
__host__ __device__ __forceinline__ int ldlFacTabInts(int nv, int nldlop) { return (nv + 1) + 3 * nldlop + 2 * nv; }
__device__ __forceinline__ LdlFacTab ldlFacTabStage(const DevModel& D, int* buf, Tm t) {
  const int nv = D.nv;
  int* p = buf;
  const unsigned live = __activemask();
  const int nlive = __popc(live), rank = __popc(live & ((1u << (threadIdx.x & 31)) - 1u));
  auto put = [&](const int* src, int n) { for (int i = t.tl * nlive + rank; i < n; i += t.nw * nlive) p[i] = src[i]; const int* r = p; p += n; return r; };
  LdlFacTab T;
  T.ldl_start = put(D.ldl_start, nv + 1); T.ldl_dst = put(D.ldl_dst, D.nldlop); T.ldl_a = put(D.ldl_a, D.nldlop); T.ldl_b = put(D.ldl_b, D.nldlop);
  T.dof_Madr = put(D.dof_Madr, nv); T.dof_rowlen = put(D.dof_rowlen, nv);
  return T;
}
// x <- (L^T D L)^-1 x, by dof-tree levels
__device__ __noinline__ void solveLDL(int slot, SP LD, SP dinv, SP x, Tm t, const LdlTab& T) {
  const DevModel& D = tP[slot].D;
  for (int l = D.ndlevel - 1; l >= 0; l--) {
    for (int idx = T.dlvl_start[l] + t.tl; idx < T.dlvl_start[l + 1]; idx += TEAM) {
      const int j = T.dlvl_dofs[idx];
      float acc = x[j];
      for (int e = T.desc_start[j]; e < T.desc_start[j + 1]; e++) acc = fmaf(-LD[T.desc_idx[e]], x[T.desc_dof[e]], acc);
      x[j] = acc;
    }
    TSYNC();
  }
  TFOR(i, D.nv) x[i] *= dinv[i];
  TSYNC();
  for (int l = 1; l < D.ndlevel; l++) {
    for (int idx = T.dlvl_start[l] + t.tl; idx < T.dlvl_start[l + 1]; idx += TEAM) {
      const int i = T.dlvl_dofs[idx], a = T.dof_Madr[i], d = T.dof_rowlen[i];
      float acc = x[i];
      for (int q = 1; q < d; q++) acc = fmaf(-LD[a + q], x[T.M_col[a + q]], acc);
      x[i] = acc;
    }
    TSYNC();
  }
}
__device__ __forceinline__ void solveLDL(int slot, SP LD, SP dinv, SP x, Tm t) { solveLDL(slot, LD, dinv, x, t, ldlTabGlobal(tP[slot].D)); }
// `stage`: shared memory for 2 nM elements per lane, or null.  The column-by-column factorisations are chains of dependent
// loads with two barriers per column: on a shared-memory copy they cost a fraction of what they do on the scratch (30 % of
// this kernel's samples before), and M and Euler's M + h B (euler != 0) share one pass.
__device__ __noinline__ void inertia(int slot, int fslot, SP sm, Tm t, World W, SP stage, int euler) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L; const FieldPtrs& F = tP[fslot].F;
  const int nv = D.nv;
  // composite inertia of each frame's subtree (frames are depth-first: subtree = contiguous range)
  TFOR(i, 10 * D.nframe) {
    const int f = i / 10, k = i - 10 * f;
    float acc = 0.f;
    if (f > 0) for (int g = f; g < D.fr_subend[f]; g++) acc += sm[L.cinert + 10 * g + k];
    sm[L.crb + i] = acc;
  }
  TSYNC();
  auto f6 = sm + L.cdof_dot;  // scratch until the velocity stage
  TFOR(i, nv) {
    float r[6];
    inertMulVec(r, sm + L.crb + 10 * D.dof_frame[i], sm + L.cdof + 6 * i);
#pragma unroll
    for (int k = 0; k < 6; k++) f6[6 * i + k] = r[k];
  }
  TSYNC();
  TFOR(e, D.nM) {
    const int r = D.M_row[e], c = D.M_col[e];
    auto a = sm + L.cdof + 6 * c; auto b = f6 + 6 * r;
    float v = a[0] * b[0] + a[1] * b[1] + a[2] * b[2] + a[3] * b[3] + a[4] * b[4] + a[5] * b[5];
    if (r == c) v += D.dof_armature[r];
    sm[L.qM + e] = v; sm[L.qLD + e] = v;
  }
  TSYNC();
  const bool damp = euler && D.any_damping;   // mj_Euler's implicit joint damping: factor of M + h diag(damping), used by finishTile
  const float h = D.timestep;
  if (stage.p) {
    SP A = stage, B = stage + D.nM;
    const LdlFacTab T = ldlFacTabStage(D, (int*)(stage.p - (threadIdx.x & 31) + 2 * D.nM * SCR), t);   // (the tables follow the lanes' block)
    TFOR(e, D.nM) { const float v = sm[L.qM + e]; A[e] = v; B[e] = v + (D.M_row[e] == D.M_col[e] ? h * D.dof_damping[D.M_row[e]] : 0.f); }
    TSYNC();
    if (damp) factorLDL2(slot, A, sm + L.qLDinv, B, sm + L.qHinv, t, T); else factorLDL(slot, A, sm + L.qLDinv, t, T);
    TFOR(e, D.nM) { sm[L.qLD + e] = A[e]; if (damp) sm[L.qH + e] = B[e]; }
    TSYNC();
  } else {
    factorLDL(slot, sm + L.qLD, sm + L.qLDinv, t);
    if (damp) {
      TFOR(e, D.nM) sm[L.qH + e] = sm[L.qM + e] + (D.M_row[e] == D.M_col[e] ? h * D.dof_damping[D.M_row[e]] : 0.f);
      TSYNC();
      factorLDL(slot, sm + L.qH, sm + L.qHinv, t);
    }
  }
  void* qMo = F.p[MYO_F_QM]; void* qLDo = F.p[MYO_F_QLD]; void* qdi = F.p[MYO_F_QLDIAGINV];
  if (qMo || qLDo) {
    TFOR(i, nv * nv) { if (qMo) OUTF(qMo, i) = 0.f; if (qLDo) OUTF(qLDo, i) = 0.f; }
    TSYNC();
    TFOR(e, D.nM) {
      const int r = D.M_row[e], c = D.M_col[e];
      if (qMo) { OUTF(qMo, r * nv + c) = sm[L.qM + e]; OUTF(qMo, c * nv + r) = sm[L.qM + e]; }
      if (qLDo) OUTF(qLDo, r * nv + c) = sm[L.qLD + e];
    }
  }
  if (qdi) TFOR(i, nv) OUTF(qdi, i) = sm[L.qLDinv + i];
}

// ------------------------------------------------------------------------------------------ mj_fwdVelocity
// The part that needs only the kinematics and qvel (body velocities, cdof_dot, mj_rne's bias force): it runs in
// inertiaKernel, beside the tendon stage; the tendon / passive part below follows the tendon stage in velactKernel.
__device__ __noinline__ void velocityBody(int slot, int fslot, SP sm, Tm t, World W) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L; const FieldPtrs& F = tP[fslot].F;
  const int nv = D.nv;
  auto qvel = sm + L.qvel;
  // body velocities: sum of the ancestor dof motions (all expressed about the tree's centre of mass)
  TFOR(f, D.nframe) {
    float v[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (f > 0) {
      const int last = D.fr_dofadr[f] + D.fr_dofnum[f] - 1, a = D.dof_Madr[last], d = D.dof_rowlen[last];
      for (int q = d - 1; q >= 0; q--) {
        const int j = D.M_col[a + q]; const float qv = qvel[j]; auto cd = sm + L.cdof + 6 * j;
#pragma unroll
        for (int k = 0; k < 6; k++) v[k] = fmaf(cd[k], qv, v[k]);
      }
    }
#pragma unroll
    for (int k = 0; k < 6; k++) sm[L.cvel + 6 * f + k] = v[k];
  }
  TFOR(i, nv) {
    const int pc = D.dof_prevcount[i], a = D.dof_Madr[i], d = D.dof_rowlen[i];
    auto out = sm + L.cdof_dot + 6 * i;
    if (pc < 0) { for (int k = 0; k < 6; k++) out[k] = 0.f; continue; }
    float v[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int q = d - 1; q >= pc; q--) {
      const int j = D.M_col[a + q]; const float qv = qvel[j]; auto cd = sm + L.cdof + 6 * j;
#pragma unroll
      for (int k = 0; k < 6; k++) v[k] = fmaf(cd[k], qv, v[k]);
    }
    float r[6];
    crossMotion(r, v, sm + L.cdof + 6 * i);
#pragma unroll
    for (int k = 0; k < 6; k++) out[k] = r[k];
  }
  TSYNC();
  const bool grav = !(D.disableflags & DSBL_GRAVITY);
  TFOR(f, D.nframe) {
    if (f == 0) { for (int k = 0; k < 6; k++) sm[L.cacc + k] = 0.f; continue; }
    float a6[6] = {0.f, 0.f, 0.f, grav ? -D.gravity[0] : 0.f, grav ? -D.gravity[1] : 0.f, grav ? -D.gravity[2] : 0.f};
    const int last = D.fr_dofadr[f] + D.fr_dofnum[f] - 1, a = D.dof_Madr[last], d = D.dof_rowlen[last];
    for (int q = d - 1; q >= 0; q--) {
      const int j = D.M_col[a + q]; const float qv = qvel[j]; auto cd = sm + L.cdof_dot + 6 * j;
#pragma unroll
      for (int k = 0; k < 6; k++) a6[k] = fmaf(cd[k], qv, a6[k]);
    }
    float Ia[6], Iv[6], cf[6];
    inertMulVec(Ia, sm + L.cinert + 10 * f, a6);
    inertMulVec(Iv, sm + L.cinert + 10 * f, sm + L.cvel + 6 * f);
    crossForce(cf, sm + L.cvel + 6 * f, Iv);
#pragma unroll
    for (int k = 0; k < 6; k++) sm[L.cacc + 6 * f + k] = Ia[k] + cf[k];   // cacc now holds the body force cfrc
  }
  TSYNC();
  void* obi = F.p[MYO_F_QFRC_BIAS];
  TFOR(i, nv) {
    const int f = D.dof_frame[i];
    float acc[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int g = f; g < D.fr_subend[f]; g++)
#pragma unroll
      for (int k = 0; k < 6; k++) acc[k] += sm[L.cacc + 6 * g + k];
    auto cd = sm + L.cdof + 6 * i;
    const float bias = cd[0] * acc[0] + cd[1] * acc[1] + cd[2] * acc[2] + cd[3] * acc[3] + cd[4] * acc[4] + cd[5] * acc[5];
    sm[L.bias + i] = bias;
    if (obi) OUTF(obi, i) = bias;
  }
  void* o;
  if ((o = F.p[MYO_F_CDOF_DOT])) TFOR(i, 6 * nv) OUTF(o, i) = sm[L.cdof_dot + i];
  if ((o = F.p[MYO_F_CVEL])) TFOR(i, 6 * D.nbody) OUTF(o, i) = sm[L.cvel + 6 * D.body_frame[i / 6] + i % 6];
  TSYNC();
}
// tendon velocities and mj_passive
__device__ __noinline__ void velocityTendon(int slot, int fslot, SP sm, Tm t, World W) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L; const FieldPtrs& F = tP[fslot].F;
  const int nv = D.nv;
  auto qvel = sm + L.qvel;
  TFOR(k, D.nten) {
    float acc = 0.f;
    for (int n = D.ten_nzstart[k]; n < D.ten_nzstart[k + 1]; n++) acc = fmaf(sm[L.tenJ + n], qvel[D.nz_dof[n]], acc);
    sm[L.tenvel + k] = acc;
  }
  if (D.any_tendon_passive) TSYNC();
  const bool passive_on = !(D.disableflags & DSBL_PASSIVE);
  void* osp = F.p[MYO_F_QFRC_SPRING]; void* odm = F.p[MYO_F_QFRC_DAMPER]; void* ops = F.p[MYO_F_QFRC_PASSIVE];
  TFOR(i, nv) {
    float spring = 0.f, damper = 0.f;
    if (passive_on) {
      const int j = D.dof_jnt[i];
      const float k = D.jnt_stiffness[j];
      if (k != 0.f && D.jnt_type[j] >= JNT_SLIDE) spring = -k * (sm[L.qpos + D.jnt_qposadr[j]] - D.qpos_spring[D.jnt_qposadr[j]]);
      damper = -D.dof_damping[i] * qvel[i];
      if (D.any_tendon_passive)
        for (int c = D.col_start[i]; c < D.col_start[i + 1]; c++) {
          const int n = D.col_nz[c], tn = D.nz_tendon[n];
          const float ks = D.ten_stiffness[tn], kd = D.ten_damping[tn], len = sm[L.tenlen + tn];
          if (ks > 0.f) {
            const float lo = D.ten_lengthspring[2 * tn], hi = D.ten_lengthspring[2 * tn + 1];
            const float fs = len > hi ? ks * (hi - len) : (len < lo ? ks * (lo - len) : 0.f);
            spring += sm[L.tenJ + n] * fs;
          }
          if (kd > 0.f) damper -= kd * sm[L.tenvel + tn] * sm[L.tenJ + n];
        }
    }
    sm[L.passive + i] = spring + damper;
    if (osp) OUTF(osp, i) = spring;
    if (odm) OUTF(odm, i) = damper;
    if (ops) OUTF(ops, i) = spring + damper;
  }
  void* o;
  if ((o = F.p[MYO_F_TEN_VELOCITY])) TFOR(k, D.nten) OUTF(o, k) = sm[L.tenvel + k];
  TSYNC();
}

// ------------------------------------------------------------------------------------------ mj_fwdActuation
__device__ __forceinline__ float muscleFL(float Ln, float lmin, float lmax) {
  if (Ln < lmin || Ln > lmax) return 0.f;
  const float a = 0.5f * (lmin + 1.f), b = 0.5f * (1.f + lmax);
  float x;
  if (Ln <= a) { x = (Ln - lmin) / fmaxf(MINVAL, a - lmin); return 0.5f * x * x; }
  if (Ln <= 1.f) { x = (1.f - Ln) / fmaxf(MINVAL, 1.f - a); return 1.f - 0.5f * x * x; }
  if (Ln <= b) { x = (Ln - 1.f) / fmaxf(MINVAL, b - 1.f); return 1.f - 0.5f * x * x; }
  x = (lmax - Ln) / fmaxf(MINVAL, lmax - b);
  return 0.5f * x * x;
}
__device__ __noinline__ void muscleGainBias(float len, float vel, const float* lr, float acc0, const float* prm, float* gain, float* bias) {
  const float r0 = prm[0], r1 = prm[1], scale = prm[3], lmin = prm[4], lmax = prm[5], vmax = prm[6], fpmax = prm[7], fvmax = prm[8];
  float force = prm[2];
  if (force < 0.f) force = scale / fmaxf(MINVAL, acc0);
  const float L0 = (lr[1] - lr[0]) / fmaxf(MINVAL, r1 - r0);
  const float Ln = r0 + (len - lr[0]) / fmaxf(MINVAL, L0);
  const float V = vel / fmaxf(MINVAL, L0 * vmax);
  const float FL = muscleFL(Ln, lmin, lmax);
  const float y = fvmax - 1.f;
  float FV;
  if (V <= -1.f) FV = 0.f;
  else if (V <= 0.f) FV = (V + 1.f) * (V + 1.f);
  else if (V <= y) FV = fvmax - (y - V) * (y - V) / fmaxf(MINVAL, y);
  else FV = fvmax;
  *gain = -force * FL * FV;
  const float b = 0.5f * (1.f + lmax);
  if (Ln <= 1.f) *bias = 0.f;
  else if (Ln <= b) { const float x = (Ln - 1.f) / fmaxf(MINVAL, b - 1.f); *bias = -force * fpmax * 0.5f * x * x; }
  else { const float x = (Ln - b) / fmaxf(MINVAL, b - 1.f); *bias = -force * fpmax * (0.5f + x); }
}
__device__ __noinline__ void actuation(int slot, int fslot, SP sm, Tm t, World W, SP stage) {   // stage: shared memory for nM + 2 nv elements per lane, or null
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L; const FieldPtrs& F = tP[fslot].F;
  const int nu = D.nu, nv = D.nv;
  const bool on = !(D.disableflags & DSBL_ACTUATION);
  void* olen = F.p[MYO_F_ACTUATOR_LENGTH]; void* ovel = F.p[MYO_F_ACTUATOR_VELOCITY]; void* ofrc = F.p[MYO_F_ACTUATOR_FORCE];
  void* odot = F.p[MYO_F_ACT_DOT];
  const void* gsc = tP[slot].F.p[MYO_F_ACT_GAINSCALE];   // input: always the batch's own slot
  TFOR(i, nu) {
    const int id = D.act_trnid[i];
    const float gear = D.act_gear[i];
    float len, vel;
    if (D.act_trntype[i] == 3) { len = gear * sm[L.tenlen + id]; vel = gear * sm[L.tenvel + id]; }
    else { len = gear * sm[L.qpos + D.jnt_qposadr[id]]; vel = gear * sm[L.qvel + D.jnt_dofadr[id]]; }
    float force = 0.f, adot = 0.f;
    const int aa = D.act_actadr[i];
    if (on) {
      float c = sm[L.ctrl + i];
      const int fl = D.act_flags[i];
      if ((fl & ACT_CTRLLIMITED) && !(D.disableflags & DSBL_CLAMPCTRL)) c = clipf(c, D.act_ctrlrange[2 * i], D.act_ctrlrange[2 * i + 1]);
      float a = c;
      if (aa >= 0) {
        a = sm[L.act + aa];
        auto dp = D.act_dynprm + 3 * i;
        const int dt = D.act_dyntype[i];
        if (dt == DYN_INTEGRATOR) adot = c;
        else if (dt == DYN_FILTER || dt == DYN_FILTEREXACT) adot = (c - a) / fmaxf(MINVAL, dp[0]);
        else if (dt == DYN_MUSCLE) {
          const float cc = clipf(c, 0.f, 1.f), ac = clipf(a, 0.f, 1.f);
          const float ta = dp[0] * (0.5f + 1.5f * ac), td = dp[1] / (0.5f + 1.5f * ac), dctrl = cc - a;
          float tau;
          if (dp[2] < MINVAL) tau = dctrl > 0.f ? ta : td;
          else { const float x = clipf(dctrl / dp[2] + 0.5f, 0.f, 1.f); tau = td + (ta - td) * (x * x * x * (x * (6.f * x - 15.f) + 10.f)); }
          adot = dctrl / fmaxf(MINVAL, tau);
        }
      }
      auto gp = D.act_gainprm + 9 * i; auto bp = D.act_biasprm + 9 * i;
      float gain = 0.f, bias = 0.f, tmp;
      const int gt = D.act_gaintype[i], bt = D.act_biastype[i];
      if (gt == GAIN_MUSCLE && bt == BIAS_MUSCLE && D.act_same_prm[i]) muscleGainBias(len, vel, D.act_lengthrange + 2 * i, D.act_acc0[i], gp, &gain, &bias);
      else {
        if (gt == GAIN_FIXED) gain = gp[0];
        else if (gt == GAIN_AFFINE) gain = gp[0] + gp[1] * len + gp[2] * vel;
        else muscleGainBias(len, vel, D.act_lengthrange + 2 * i, D.act_acc0[i], gp, &gain, &tmp);
        if (bt == BIAS_AFFINE) bias = bp[0] + bp[1] * len + bp[2] * vel;
        else if (bt == BIAS_MUSCLE) muscleGainBias(len, vel, D.act_lengthrange + 2 * i, D.act_acc0[i], bp, &tmp, &bias);
      }
      if (gsc) gain *= OUTF(gsc, i);   // per-world muscle weakness (reference: base_v0.py:64-68 scales gainprm[:, 2])
      force = gain * a + bias;
      if (fl & ACT_FORCELIMITED) force = clipf(force, D.act_forcerange[2 * i], D.act_forcerange[2 * i + 1]);
    }
    sm[L.actforce + i] = force;
    if (aa >= 0) { sm[L.actdot + aa] = adot; if (odot) OUTF(odot, aa) = adot; }
    if (olen) OUTF(olen, i) = len;
    if (ovel) OUTF(ovel, i) = vel;
    if (ofrc) OUTF(ofrc, i) = force;
  }
  TSYNC();
  TFOR(k, D.nten) {
    float acc = 0.f;
    for (int q = D.tact_start[k]; q < D.tact_start[k + 1]; q++) { const int a = D.tact_list[q]; acc += D.act_gear[a] * sm[L.actforce + a]; }
    sm[L.tforce + k] = acc;
  }
  TSYNC();
  // qfrc_actuator = moment^T force through the sparse tendon rows (column view), deterministic gathers
  const void* applied = tP[slot].F.p[MYO_F_QFRC_APPLIED];   // input: always the batch's own slot
  TFOR(i, nv) {
    float acc = 0.f;
    for (int c = D.col_start[i]; c < D.col_start[i + 1]; c++) { const int n = D.col_nz[c]; acc = fmaf(sm[L.tenJ + n], sm[L.tforce + D.nz_tendon[n]], acc); }
    for (int q = D.dact_start[i]; q < D.dact_start[i + 1]; q++) { const int a = D.dact_list[q]; acc += D.act_gear[a] * sm[L.actforce + a]; }
    sm[L.qfrc_act + i] = acc;
    const float sf = sm[L.passive + i] - sm[L.bias + i] + acc + (applied ? OUTF(applied, i) : 0.f);
    sm[L.smooth + i] = sf; sm[L.qacc_smooth + i] = sf;
  }
  TSYNC();
  if (stage.p) {   // the level-by-level solve on a shared-memory copy of the factor (27 % of this kernel's samples on the scratch)
    SP LD = stage, dinv = stage + D.nM, x = dinv + nv;
    const LdlTab T = ldlTabStage(D, (int*)(stage.p - (threadIdx.x & 31) + (D.nM + 2 * nv) * SCR), t);   // (the tables follow the lanes' block)
    TFOR(e, D.nM) LD[e] = sm[L.qLD + e];
    TFOR(i, nv) { dinv[i] = sm[L.qLDinv + i]; x[i] = sm[L.qacc_smooth + i]; }
    TSYNC();
    solveLDL(slot, LD, dinv, x, t, T);
    TFOR(i, nv) sm[L.qacc_smooth + i] = x[i];
    TSYNC();
  } else solveLDL(slot, sm + L.qLD, sm + L.qLDinv, sm + L.qacc_smooth, t);
  void* o;
  if ((o = F.p[MYO_F_QFRC_ACTUATOR])) TFOR(i, nv) OUTF(o, i) = sm[L.qfrc_act + i];
  if ((o = F.p[MYO_F_QFRC_SMOOTH])) TFOR(i, nv) OUTF(o, i) = sm[L.smooth + i];
  if ((o = F.p[MYO_F_QACC_SMOOTH])) TFOR(i, nv) OUTF(o, i) = sm[L.qacc_smooth + i];
}
// ------------------------------------------------------------------------------------------ collision
// The convex narrow phase (Minkowski portal refinement on B - A) runs in float64: its result is defined by an
// iteration with a 1e-6 stopping rule, and in float32 the portal degenerates near convergence (contact points of
// ellipsoids moved by millimetres).  Few pairs reach it per world and B200 runs FP64 at half the FP32 rate.
struct d3 { double x, y, z; };
__device__ __forceinline__ d3 mkd(double x, double y, double z) { d3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ d3 operator+(d3 a, d3 b) { return mkd(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ d3 operator-(d3 a, d3 b) { return mkd(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ d3 operator*(d3 a, double s) { return mkd(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ double dot(d3 a, d3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ d3 cross(d3 a, d3 b) { return mkd(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
__device__ __forceinline__ double norm3(d3 a) { return sqrt(dot(a, a)); }
__device__ __forceinline__ d3 unit3(d3 a) { const double q = dot(a, a); return q < 1e-30 ? mkd(1.0, 0.0, 0.0) : a * rsqrt(q); }   // (one rsqrt instead of sqrt + divide: these sit on the portal iteration's dependent chain)
// (the rotation is copied into the record once: the support function runs tens of times per pair; it and mprSup are
// inlined so that the portal's support points stay in registers instead of going through the stack)
struct Shape { int type; d3 pos; float mat[9]; d3 size; double margin; const float* verts; int nvert; };   // verts: hull of a mesh geom
__device__ __forceinline__ void shapeMat(Shape& s, SP R) {
#pragma unroll
  for (int i = 0; i < 9; i++) s.mat[i] = R[i];
}
__device__ __forceinline__ d3 supportPt(const Shape& s, d3 dir) {
  const float* R = s.mat;
  const d3 d = mkd(R[0] * dir.x + R[3] * dir.y + R[6] * dir.z, R[1] * dir.x + R[4] * dir.y + R[7] * dir.z, R[2] * dir.x + R[5] * dir.y + R[8] * dir.z);
  d3 r = mkd(0.0, 0.0, 0.0);
  switch (s.type) {
    case G_SPHERE: r = d * s.size.x; break;
    case G_CAPSULE: r = d * s.size.x; r.z += d.z >= 0.0 ? s.size.y : -s.size.y; break;
    case G_ELLIPSOID: {
      const d3 t = mkd(s.size.x * s.size.x * d.x, s.size.y * s.size.y * d.y, s.size.z * s.size.z * d.z);
      r = t * rsqrt(fmax(1e-30, t.x * d.x + t.y * d.y + t.z * d.z));
    } break;
    case G_CYLINDER: {
      const double q = d.x * d.x + d.y * d.y;
      if (q > 1e-24) { const double in = s.size.x * rsqrt(q); r.x = d.x * in; r.y = d.y * in; }
      r.z = d.z >= 0.0 ? s.size.y : -s.size.y;
    } break;
    case G_BOX: r = mkd(d.x >= 0.0 ? s.size.x : -s.size.x, d.y >= 0.0 ? s.size.y : -s.size.y, d.z >= 0.0 ? s.size.z : -s.size.z); break;
    case G_MESH: {   // the hull vertex farthest along d (MuJoCo climbs the hull's graph to the same vertex; first one wins a tie)
      double best = -1e300;
      for (int i = 0; i < s.nvert; i++) {
        const double vx = s.verts[3 * i], vy = s.verts[3 * i + 1], vz = s.verts[3 * i + 2];
        const double v = vx * d.x + vy * d.y + vz * d.z;
        if (v > best) { best = v; r = mkd(vx, vy, vz); }
      }
    } break;
    default: break;
  }
  return mkd(R[0] * r.x + R[1] * r.y + R[2] * r.z, R[3] * r.x + R[4] * r.y + R[5] * r.z, R[6] * r.x + R[7] * r.y + R[8] * r.z) + s.pos + dir * s.margin;
}
struct Sup { d3 v, a, b; };
__device__ __forceinline__ Sup mprSup(const Shape& A, const Shape& B, d3 d) {
  Sup s; s.a = supportPt(A, d * -1.0); s.b = supportPt(B, d); s.v = s.b - s.a; return s;
}
// portal_only: stop once a portal exists (true) -- every early "no intersection" exit lies before that point or in the first
// refinement steps, so this is the cheap filter collideMprKernel compacts its work list with
__device__ __forceinline__ bool mprPenetration(const Shape& A, const Shape& B, double tol, int maxit, double* depth, d3* nrm, d3* pos, bool portal_only) {
  d3 v0 = B.pos - A.pos;
  if (norm3(v0) < 1e-12) v0 = mkd(1e-5, 0.0, 0.0);
  d3 dir = unit3(v0 * -1.0);
  Sup s1 = mprSup(A, B, dir);
  if (dot(s1.v, dir) <= 0.0) return false;
  dir = cross(s1.v, v0);
  if (norm3(dir) < 1e-12 * fmax(1.0, norm3(s1.v) * norm3(v0))) {
    const d3 n = unit3(s1.v - v0);
    if (portal_only) return true;
    *depth = dot(s1.v, n); *nrm = n * -1.0; *pos = (s1.a + s1.b) * 0.5;
    return true;
  }
  dir = unit3(dir);
  Sup s2 = mprSup(A, B, dir);
  if (dot(s2.v, dir) <= 0.0) return false;
  dir = unit3(cross(s1.v - v0, s2.v - v0));
  if (dot(dir, v0) > 0.0) { Sup t = s1; s1 = s2; s2 = t; dir = dir * -1.0; }
  Sup s3;
  // #pragma unroll 1
  for (int it = 0;; it++) {
    if (it > maxit) return false;
    s3 = mprSup(A, B, dir);
    if (dot(s3.v, dir) <= 0.0) return false;
    if (dot(cross(s1.v, s3.v), v0) < 0.0) { s2 = s3; dir = unit3(cross(s1.v - v0, s3.v - v0)); continue; }
    if (dot(cross(s3.v, s2.v), v0) < 0.0) { s1 = s3; dir = unit3(cross(s3.v - v0, s2.v - v0)); continue; }
    break;
  }
  if (portal_only) return true;
  bool hit = false;
  // #pragma unroll 1
  for (int it = 0;; it++) {
    dir = unit3(cross(s2.v - s1.v, s3.v - s1.v));
    if (dot(dir, s1.v) >= 0.0) hit = true;
    const Sup s4 = mprSup(A, B, dir);
    if (dot(s4.v, dir) < 0.0 && !hit) return false;
    const double adv = fmin(fmin(dot(s4.v - s1.v, dir), dot(s4.v - s2.v, dir)), dot(s4.v - s3.v, dir));
    if (adv <= tol || it >= maxit) {
      if (!hit) return false;
      const double d = dot(dir, s1.v);
      const d3 p = dir * d, e0 = s2.v - s1.v, e1 = s3.v - s1.v, e2 = p - s1.v;
      const double d00 = dot(e0, e0), d01 = dot(e0, e1), d11 = dot(e1, e1), d20 = dot(e2, e0), d21 = dot(e2, e1);
      const double den = d00 * d11 - d01 * d01;
      double b1 = 1.0, b2 = 0.0, b3 = 0.0;
      if (fabs(den) > 1e-30) { b2 = (d11 * d20 - d01 * d21) / den; b3 = (d00 * d21 - d01 * d20) / den; b1 = 1.0 - b2 - b3; }
      b1 = fmin(fmax(b1, 0.0), 1.0); b2 = fmin(fmax(b2, 0.0), 1.0); b3 = fmin(fmax(b3, 0.0), 1.0);
      const double bs = 1.0 / fmax(1e-30, b1 + b2 + b3);
      const d3 pa = (s1.a * b1 + s2.a * b2 + s3.a * b3) * bs, pb = (s1.b * b1 + s2.b * b2 + s3.b * b3) * bs;
      *depth = d; *nrm = dir * -1.0; *pos = (pa + pb) * 0.5;
      return true;
    }
    const d3 c = cross(s4.v, v0);
    if (dot(s1.v, c) > 0.0) { if (dot(s2.v, c) > 0.0) s1 = s4; else s3 = s4; }
    else { if (dot(s3.v, c) > 0.0) s2 = s4; else s1 = s4; }
  }
}
struct Hit { float dist; f3 pos, nrm; };
__device__ __forceinline__ Hit sphereSphere(f3 c1, float r1, f3 c2, float r2) {
  const f3 d = c2 - c1; const float n = norm3(d);
  Hit h; h.nrm = n < MINVAL ? mk3(1.f, 0.f, 0.f) : d * (1.f / n);
  h.dist = n - r1 - r2; h.pos = c1 + h.nrm * (r1 + 0.5f * h.dist);
  return h;
}
// effective shape of a collision geom: the compiled tables, or this world's cgshape record when shape overrides are bound
struct CgShape { int type; f3 size; float rbound; };
__device__ __forceinline__ CgShape cgShape(const DevModel& D, const SmemLayout& L, SP sm, bool ovr, int c) {
  CgShape r;
  if (ovr) { auto o = sm + L.cgshape + 5 * c; r.type = __float_as_int(o[0]); r.size = ld3(o + 1); r.rbound = o[4]; }
  else { r.type = D.cg_type[c]; r.size = ld3(D.cg_size + 3 * c); r.rbound = D.cg_rbound[c]; }
  return r;
}
// contact record word 7: the candidate pair, with bit 30 set when this world's shape overrides put the pair's geoms in the
// other order (MuJoCo calls the collision function with the lower geom type first: geom1 / geom2 and the normal follow)
#define CON_PAIR(x) ((x) & 0x3fffffff)
#define CON_SWAP(x) (((x) >> 30) & 1)
#define MAXHIT 4   // contacts of one candidate pair: plane - box gives up to four, plane - capsule two, everything else one
// narrow phase for one candidate pair; returns the number of contacts (0..MAXHIT) with dist < margin
// (filter: convex pairs only run the portal search; -1 = a portal exists, run the pair in full)
// (CONVEX = false: the tile kernel's instance, which never sees a convex pair -- isMprPair sends those to the work list --
// and should not carry the MPR code's registers)
template <bool CONVEX> __device__ __noinline__ int collidePair(int slot, SP sm, int pr, Hit* h, int* swapped, bool filter = false) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L;
  int c1 = D.pair_cg1[pr], c2 = D.pair_cg2[pr];
  const bool ovr = tP[slot].F.p[MYO_F_GEOM_TYPE_OVR] != nullptr;
  CgShape S1 = cgShape(D, L, sm, ovr, c1), S2 = cgShape(D, L, sm, ovr, c2);
  *swapped = 0;
  if (S1.type > S2.type) { const int c = c1; c1 = c2; c2 = c; const CgShape S = S1; S1 = S2; S2 = S; *swapped = 1; }   // only through an override: the compiled list is ordered
  const int t1 = S1.type, t2 = S2.type;
  const float margin = D.pair_prm[16 * pr];
  auto P1 = sm + L.cgpose + 12 * c1; auto P2 = sm + L.cgpose + 12 * c2;
  const f3 p1 = ld3(P1), p2 = ld3(P2);
  const SP R1 = P1 + 3, R2 = P2 + 3;
  const f3 sz1 = S1.size, sz2 = S2.size;
  const f3 ax1 = mk3(R1[2], R1[5], R1[8]), ax2 = mk3(R2[2], R2[5], R2[8]);
  int n = 0;
  if (t1 == G_PLANE) {
    if (dot(p2 - p1, ax1) > S2.rbound + margin) return 0;
    if (t2 == G_SPHERE || t2 == G_CAPSULE) {
      const int cnt = t2 == G_SPHERE ? 1 : 2;
      // #pragma unroll 1
      for (int k = 0; k < cnt; k++) {
        const f3 c = t2 == G_SPHERE ? p2 : (k == 0 ? p2 + ax2 * sz2.y : p2 - ax2 * sz2.y);
        const float dist = dot(c - p1, ax1) - sz2.x;
        if (dist < margin) { h[n].dist = dist; h[n].pos = c - ax1 * (sz2.x + 0.5f * dist); h[n].nrm = ax1; n++; }
      }
    } else if (t2 == G_BOX) {
      // mjc_PlaneBox: the corners below the box centre that are within the margin, in corner order, at most four
      const float dist = dot(p2 - p1, ax1);
      for (int i = 0; i < 8 && n < MAXHIT; i++) {
        const f3 v = mk3((i & 1) ? sz2.x : -sz2.x, (i & 2) ? sz2.y : -sz2.y, (i & 4) ? sz2.z : -sz2.z);
        const f3 corner = mulM(R2, v);
        const float ld = dot(ax1, corner);
        if (dist + ld > margin || ld > 0.f) continue;
        const float cd = dist + ld;
        h[n].dist = cd; h[n].pos = p2 + corner - ax1 * (0.5f * cd); h[n].nrm = ax1; n++;
      }
    } else {
      Shape s; s.type = t2; s.pos = mkd(p2.x, p2.y, p2.z); shapeMat(s, R2); s.size = mkd(sz2.x, sz2.y, sz2.z); s.margin = 0.0;
      s.verts = D.mesh_vert + 3 * D.cg_meshadr[c2]; s.nvert = D.cg_meshnum[c2];
      const d3 spd = supportPt(s, mkd(-ax1.x, -ax1.y, -ax1.z));
      const f3 sp = mk3((float)spd.x, (float)spd.y, (float)spd.z);
      const float dist = (float)((spd.x - p1.x) * ax1.x + (spd.y - p1.y) * ax1.y + (spd.z - p1.z) * ax1.z);
      if (dist < margin) { h[0].dist = dist; h[0].pos = sp - ax1 * (0.5f * dist); h[0].nrm = ax1; n = 1; }
    }
    return n;
  }
  if (norm3(p2 - p1) > S1.rbound + S2.rbound + margin) return 0;
  if (t1 == G_SPHERE && t2 == G_SPHERE) h[0] = sphereSphere(p1, sz1.x, p2, sz2.x);
  else if (t1 == G_SPHERE && t2 == G_CAPSULE) h[0] = sphereSphere(p1, sz1.x, p2 + ax2 * clipf(dot(p1 - p2, ax2), -sz2.y, sz2.y), sz2.x);
  else if (t1 == G_CAPSULE && t2 == G_CAPSULE) {
    const f3 a1 = ax1 * sz1.y, a2 = ax2 * sz2.y, dif = p1 - p2;
    const float ma = dot(a1, a1), mb = -dot(a1, a2), mc = dot(a2, a2), u = -dot(a1, dif), v = dot(a2, dif);
    const float det = ma * mc - mb * mb;
    float x1, x2;
    if (fabsf(det) >= 1e-30f) {
      x1 = (mc * u - mb * v) / det; x2 = (ma * v - mb * u) / det;
      if (x1 > 1.f) { x1 = 1.f; x2 = (v - mb) / mc; } else if (x1 < -1.f) { x1 = -1.f; x2 = (v + mb) / mc; }
      if (x2 > 1.f) { x2 = 1.f; x1 = clipf((u - mb) / ma, -1.f, 1.f); } else if (x2 < -1.f) { x2 = -1.f; x1 = clipf((u + mb) / ma, -1.f, 1.f); }
    } else {
      x1 = clipf(u / fmaxf(1e-30f, ma), -1.f, 1.f);
      x2 = clipf((v - mb * x1) / fmaxf(1e-30f, mc), -1.f, 1.f);
    }
    h[0] = sphereSphere(p1 + a1 * x1, sz1.x, p2 + a2 * x2, sz2.x);
  } else if (!CONVEX) return 0;
  else {
    Shape A, B;
    A.type = t1; A.pos = mkd(p1.x, p1.y, p1.z); shapeMat(A, R1); A.size = mkd(sz1.x, sz1.y, sz1.z); A.margin = 0.5 * (double)margin;
    B.type = t2; B.pos = mkd(p2.x, p2.y, p2.z); shapeMat(B, R2); B.size = mkd(sz2.x, sz2.y, sz2.z); B.margin = 0.5 * (double)margin;
    A.verts = D.mesh_vert + 3 * D.cg_meshadr[c1]; A.nvert = D.cg_meshnum[c1]; B.verts = D.mesh_vert + 3 * D.cg_meshadr[c2]; B.nvert = D.cg_meshnum[c2];
    double depth; d3 nrm, pos;
    if (!mprPenetration(A, B, 1e-6, 50, &depth, &nrm, &pos, filter)) return 0;
    if (filter) return -1;
    h[0].dist = (float)((double)margin - depth); h[0].pos = mk3((float)pos.x, (float)pos.y, (float)pos.z);
    h[0].nrm = mk3((float)nrm.x, (float)nrm.y, (float)nrm.z);
  }
  return h[0].dist < margin ? 1 : 0;
}
__device__ __forceinline__ void makeFrame(f3 n, f3* y, f3* z) {
  f3 t = mk3(0.f, 0.f, 0.f);
  if (n.y < 0.5f && n.y > -0.5f) t.y = 1.f; else t.z = 1.f;
  t = unit3(t - n * dot(n, t));
  *y = t; *z = cross(n, t);
}
__device__ __noinline__ void resetState(int slot, SP sm, Tm t) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L;
  TFOR(i, D.nq) sm[L.qpos + i] = D.qpos0[i];
  TFOR(i, D.nv) { sm[L.qvel + i] = 0.f; sm[L.warm + i] = 0.f; }
  TFOR(i, D.na) sm[L.act + i] = 0.f;
  TSYNC();
}

// ------------------------------------------------------------------------------------------ constraint rows (shared record layout)
// Rows are never stored densely: a row is (type, id, coef) and its Jacobian is implied (limit rows 1-sparse, equality rows 2-sparse, tendon-limit rows the tendon's sparse row, contact rows +- base rows).
enum { E_D = 0, E_AREF, E_JAR, E_JV, E_FORCE, E_DESC0, E_COEF, E_DESC1, E_POS, E_MARGIN };
#define EFC(k, r) sm[L.efc + (k) * D.njmax + (r)]
#define EFCI(k, r) __float_as_int(sm[L.efc + (k) * D.njmax + (r)])

// ------------------------------------------------------------------------------------------ contact / efc outputs
// (efc_force and solver_niter, the results of the solve, are written by finishKernel: these records are complete once the
// rows exist, and go out while the solver runs)
// live_only (myo_batch_set_count_aware): only the records of the contacts / rows that exist are written, the padding up to
// nconmax / njmax keeps whatever the buffers held
__device__ __noinline__ void writeConstraintOutputs(int slot, int fslot, SP sm, Tm t, World W, int ncon, int nefc, bool live_only) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L; const FieldPtrs& F = tP[fslot].F;
  const int nv = D.nv, ncm = D.nconmax, njm = D.njmax;
  void* p;
  if (t.tl == 0) {
    if ((p = F.p[MYO_F_NCON])) OUTI(p, 0) = ncon;
    if ((p = F.p[MYO_F_NEFC])) OUTI(p, 0) = nefc;
  }
  void* cg = F.p[MYO_F_CONTACT_GEOM]; void* cdim = F.p[MYO_F_CONTACT_DIM]; void* cadr = F.p[MYO_F_CONTACT_EFC_ADDRESS];
  void* cdist = F.p[MYO_F_CONTACT_DIST]; void* cpos = F.p[MYO_F_CONTACT_POS]; void* cfr = F.p[MYO_F_CONTACT_FRAME];
  void* cinc = F.p[MYO_F_CONTACT_INCLUDEMARGIN]; void* cfri = F.p[MYO_F_CONTACT_FRICTION]; void* csr = F.p[MYO_F_CONTACT_SOLREF];
  void* csi = F.p[MYO_F_CONTACT_SOLIMP];
  if (cg || cdim || cadr || cdist || cpos || cfr || cinc || cfri || csr || csi) {
    TFOR(c, live_only ? ncon : ncm) {
      const bool on = c < ncon;
      auto cr = sm + L.con + 12 * c;
      const int prw = on ? __float_as_int(cr[7]) : 0, pr = CON_PAIR(prw), sw = CON_SWAP(prw);
      auto pp = D.pair_prm + 16 * pr;
      if (cg) {
        const int ga = on ? D.cg_geom[D.pair_cg1[pr]] : -1, gb = on ? D.cg_geom[D.pair_cg2[pr]] : -1;
        OUTI(cg, 2 * c) = sw ? gb : ga; OUTI(cg, 2 * c + 1) = sw ? ga : gb;
      }
      if (cdim) OUTI(cdim, c) = on ? __float_as_int(cr[9]) : 0;
      if (cadr) OUTI(cadr, c) = on ? __float_as_int(cr[8]) : -1;
      if (cdist) OUTF(cdist, c) = on ? cr[0] : 0.f;
      if (cpos) for (int k = 0; k < 3; k++) OUTF(cpos, 3 * c + k) = on ? cr[1 + k] : 0.f;
      if (cfr) {
        f3 n = mk3(0.f, 0.f, 0.f), y = n, z = n;
        if (on) { n = ld3(cr + 4); makeFrame(n, &y, &z); }
        OUTF(cfr, 9 * c) = n.x; OUTF(cfr, 9 * c + 1) = n.y; OUTF(cfr, 9 * c + 2) = n.z; OUTF(cfr, 9 * c + 3) = y.x; OUTF(cfr, 9 * c + 4) = y.y;
        OUTF(cfr, 9 * c + 5) = y.z; OUTF(cfr, 9 * c + 6) = z.x; OUTF(cfr, 9 * c + 7) = z.y; OUTF(cfr, 9 * c + 8) = z.z;
      }
      if (cinc) OUTF(cinc, c) = on ? pp[0] - pp[1] : 0.f;
      if (cfri) for (int k = 0; k < 5; k++) OUTF(cfri, 5 * c + k) = on ? pp[2 + k] : 0.f;
      if (csr) for (int k = 0; k < 2; k++) OUTF(csr, 2 * c + k) = on ? pp[7 + k] : 0.f;
      if (csi) for (int k = 0; k < 5; k++) OUTF(csi, 5 * c + k) = on ? pp[9 + k] : 0.f;
    }
  }
  void* epos = F.p[MYO_F_EFC_POS]; void* emar = F.p[MYO_F_EFC_MARGIN]; void* eD = F.p[MYO_F_EFC_D]; void* earef = F.p[MYO_F_EFC_AREF];
  void* etyp = F.p[MYO_F_EFC_TYPE]; void* eJ = F.p[MYO_F_EFC_J];
  if (epos || emar || eD || earef || etyp || eJ) {
    TFOR(r, live_only ? nefc : njm) {
      const bool on = r < nefc;
      if (epos) OUTF(epos, r) = on ? EFC(E_POS, r) : 0.f;
      if (emar) OUTF(emar, r) = on ? EFC(E_MARGIN, r) : 0.f;
      if (eD) OUTF(eD, r) = on ? EFC(E_D, r) : 0.f;
      if (earef) OUTF(earef, r) = on ? EFC(E_AREF, r) : 0.f;
      const int d0 = on ? EFCI(E_DESC0, r) : 0, type = d0 & 3, id = d0 >> 2;
      if (etyp) OUTI(etyp, r) = !on ? -1 : (type == ROW_EQ ? 0 : (type == ROW_LIMJ ? 3 : (type == ROW_LIMT ? 4 : (EFCI(E_DESC1, r) < 0 ? 5 : 6))));
      if (eJ) {
        const float coef = on ? EFC(E_COEF, r) : 0.f;
        const int t1 = on && type == ROW_CON ? EFCI(E_DESC1, r) : -1;
        if (on && type == ROW_CON) {
          // contact rows (most live rows): eight columns of the two base rows loaded together -- one dependent pair of scratch
          // loads per element was 54 % of this kernel's samples
          auto Jn = sm + L.conJ + id * nv; auto Jt = sm + L.conJ + (t1 >= 0 ? t1 : id) * nv;
          for (int k0 = 0; k0 < nv; k0 += 8) {
            float a[8], bb[8];
#pragma unroll
            for (int u = 0; u < 8; u++) { a[u] = k0 + u < nv ? Jn[k0 + u] : 0.f; bb[u] = (t1 >= 0 && k0 + u < nv) ? Jt[k0 + u] : 0.f; }
#pragma unroll
            for (int u = 0; u < 8; u++) if (k0 + u < nv) OUTF(eJ, r * nv + k0 + u) = a[u] + (t1 >= 0 ? coef * bb[u] : 0.f);
          }
        } else {
          for (int k = 0; k < nv; k++) {
            float v = 0.f;
            if (on) {
              if (type == ROW_LIMJ) v = D.limj_dof[id] == k ? coef : 0.f;
              else if (type == ROW_EQ) v = D.eq_dof1[id] == k ? 1.f : (D.eq_dof2[id] == k ? coef : 0.f);
              else { const int tn = D.limt_ten[id]; for (int q = D.ten_nzstart[tn]; q < D.ten_nzstart[tn + 1]; q++) if (D.nz_dof[q] == k) v = coef * sm[L.tenJ + q]; }
            }
            OUTF(eJ, r * nv + k) = v;
          }
        }
      }
    }
  }
}

#define TRI(r) (((r) * ((r) + 1)) >> 1)
// Sum of a per-item partial over the items of a world: the partials of the NW warps meet in shared memory and every warp
// adds them in the same (warp) order, so all warps of a tile hold bit-identical totals and take the same branches; must
// be called in CTA-uniform control flow.
#ifndef MAXW
#define MAXW 12   // warps per tile CTA (blockDim <= 32 MAXW); with two CTAs per SM the kernels get <= 85 registers
#endif
__device__ __noinline__ float tileSum(Tm t, float v) {
  __shared__ float red[MAXW][32];
  const int lane = threadIdx.x & 31;
  red[t.tl][lane] = v;
  __syncthreads();
  float s = 0.f;
  for (int j = 0; j < t.nw; j++) s += red[j][lane];
  __syncthreads();
  return s;
}
#define TSUM(v) tileSum(t, v)
// ------------------------------------------------------------------------------------------ constraints
// Rows are never stored densely: a row is (type, id, coef) and its Jacobian is implied -- limit rows are 1-sparse,
// equality rows 2-sparse, tendon-limit rows reuse the tendon's sparse moment-arm row, contact rows are +- combinations
// of the contact's base rows (normal, 2 tangents, torsion) held in conJ.

__device__ __noinline__ float impedance(const float* si, float x) {  // si = sanitised solimp, x = |pos - margin|
  const float dmin = si[0], dmax = si[1], width = si[2], mid = si[3], power = si[4];
  if (dmin == dmax || width <= MINVAL) return 0.5f * (dmin + dmax);
  x = x / width;
  if (x >= 1.f) return dmax;
  if (x <= 0.f) return dmin;
  float y;
  if (power == 1.f) y = x;
  else if (x <= mid) y = pow_(x, power) / pow_(mid, power - 1.f);
  else y = 1.f - pow_(1.f - x, power) / pow_(1.f - mid, power - 1.f);
  return dmin + y * (dmax - dmin);
}
__device__ __noinline__ void setRow(int slot, SP sm, int r, int type, int id, float coef, int desc1,
                                    float pos, float margin, float dA, const float* kb, float vel, float Rscale) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L;
  const float imp = impedance(kb + 2, fabsf(pos - margin));   // kb: K, B, solimp[5]
  float R = fmaxf(MINVAL, (1.f - imp) * dA / imp);
  if (Rscale > 0.f) R = fmaxf(MINVAL, Rscale * R);
  EFC(E_D, r) = 1.f / R;
  EFC(E_AREF, r) = -kb[1] * vel - kb[0] * imp * (pos - margin);
  EFC(E_DESC0, r) = __int_as_float(type | (id << 2)); EFC(E_COEF, r) = coef; EFC(E_DESC1, r) = __int_as_float(desc1);
  EFC(E_POS, r) = pos; EFC(E_MARGIN, r) = margin; EFC(E_FORCE, r) = 0.f;
}
// base-row products  out[b] = conJ[b] . x   for the nb*ncon contact base rows
// out[n] = Jb x[n] for NX right-hand sides in one sweep over the base rows.  Two rows per pass; the loads of a chunk of JKB
// columns of both rows are issued before the first use (the dot products are bound by the latency of the J loads, which
// miss L1: ~16 of them in flight per warp instead of 4).  Even / odd columns keep separate accumulators (summation order
// as before the chunking).
#ifndef JKB
#define JKB 8
#endif
template <int NX> __device__ __forceinline__ void baseMulT(int slot, SP sm, Tm t, int ncon, const SP* x, const SP* out) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L;
  const int nbr = D.maxbase * ncon, nv = D.nv;
  for (int b = t.tl; b < nbr; b += 2 * TEAM) {
    const int b2 = b + TEAM;
    auto J0 = sm + L.conJ + b * nv; auto J1 = sm + L.conJ + min(b2, nbr - 1) * nv;
    float a[NX][2], c[NX][2];
#pragma unroll
    for (int n = 0; n < NX; n++) { a[n][0] = a[n][1] = c[n][0] = c[n][1] = 0.f; }
    for (int k0 = 0; k0 < nv; k0 += JKB) {
      float j0[JKB], j1[JKB];
#pragma unroll
      for (int i = 0; i < JKB; i++) { j0[i] = LDJS(J0[k0 + i]); j1[i] = LDJS(J1[k0 + i]); }   // (past the row end: the next row or region of the scratch, masked below)
#pragma unroll
      for (int i = 0; i < JKB; i++)
        if (k0 + i < nv) {
#pragma unroll
          for (int n = 0; n < NX; n++) { const float xv = x[n][k0 + i]; a[n][i & 1] = fmaf(j0[i], xv, a[n][i & 1]); c[n][i & 1] = fmaf(j1[i], xv, c[n][i & 1]); }
        }
    }
#pragma unroll
    for (int n = 0; n < NX; n++) { out[n][b] = a[n][0] + a[n][1]; if (b2 < nbr) out[n][b2] = c[n][0] + c[n][1]; }
  }
}
__device__ __noinline__ void baseMul(int slot, SP sm, Tm t, int ncon, SP x, SP out) { baseMulT<1>(slot, sm, t, ncon, &x, &out); }
__device__ __noinline__ void baseMul2(int slot, SP sm, Tm t, int ncon, SP x0, SP x1, SP out0, SP out1) {
  const SP x[2] = {x0, x1}; const SP out[2] = {out0, out1};
  baseMulT<2>(slot, sm, t, ncon, x, out);
}
// J_r . x for one efc row (bx = base products of x)
__device__ __forceinline__ float rowMul(int slot, SP sm, int r, SP x, SP bx) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L;
  const int d0 = EFCI(E_DESC0, r), type = d0 & 3, id = d0 >> 2;
  const float coef = EFC(E_COEF, r);
  if (type == ROW_LIMJ) return coef * x[D.limj_dof[id]];
  if (type == ROW_CON) { const int q = EFCI(E_DESC1, r); return bx[id] + (q >= 0 ? coef * bx[q] : 0.f); }
  if (type == ROW_EQ) { const int d2 = D.eq_dof2[id]; return x[D.eq_dof1[id]] + (d2 >= 0 ? coef * x[d2] : 0.f); }
  const int tn = D.limt_ten[id];
  float acc = 0.f;
  for (int n = D.ten_nzstart[tn]; n < D.ten_nzstart[tn + 1]; n++) acc += sm[L.tenJ + n] * x[D.nz_dof[n]];
  return coef * acc;
}

// Row loops of the solver: a warp owns rows r = tl, tl + nw, ... and is bound by the latency of their loads (scratch, mostly
// L1 misses), so the loops take RB rows per iteration and issue all their loads before the first use.  (The compiler cannot
// do it: it cannot prove that the store to one row's record does not alias the next row's loads.)  Sums keep the row order.
#ifndef RB
#define RB 4
#endif
// J_r . x for RB rows (r0, r0 + nw, ...); contact rows -- the common case -- with their base products loaded together
__device__ __forceinline__ void rowMulBatch(int slot, SP sm, int r0, int nw, int nefc, SP x, SP bx, float* out) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L;
  int d0[RB], q1[RB]; float cf[RB];
#pragma unroll
  for (int u = 0; u < RB; u++) {
    const int r = r0 + u * nw; const bool ok = r < nefc;
    d0[u] = ok ? EFCI(E_DESC0, r) : ROW_CON; cf[u] = ok ? EFC(E_COEF, r) : 0.f; q1[u] = ok ? EFCI(E_DESC1, r) : -1;
  }
  float ba[RB], bb[RB];
#pragma unroll
  for (int u = 0; u < RB; u++) {
    const bool con = (d0[u] & 3) == ROW_CON;
    ba[u] = bx[con ? d0[u] >> 2 : 0]; bb[u] = bx[con && q1[u] >= 0 ? q1[u] : 0];
  }
#pragma unroll
  for (int u = 0; u < RB; u++) {
    const int r = r0 + u * nw;
    if ((d0[u] & 3) == ROW_CON) out[u] = ba[u] + (q1[u] >= 0 ? cf[u] * bb[u] : 0.f);
    else out[u] = r < nefc ? rowMul(slot, sm, r, x, bx) : 0.f;
  }
}

// ------------------------------------------------------------------------------------------ Newton solver (mj_fwdConstraint)
__device__ __noinline__ void mulMsparse(int slot, SP qM, SP x, SP out, Tm t) {
  const DevModel& D = tP[slot].D;
  TFOR(i, D.nv) {
    const int a = D.dof_Madr[i], d = D.dof_rowlen[i];
    float acc = 0.f;
    // (eight entries of M loaded together: one scratch latency per row part instead of one per entry; same summation order)
    for (int q0 = 0; q0 < d; q0 += 8) {
      float m[8];
#pragma unroll
      for (int u = 0; u < 8; u++) m[u] = q0 + u < d ? qM[a + q0 + u] : 0.f;
#pragma unroll
      for (int u = 0; u < 8; u++) if (q0 + u < d) acc = fmaf(m[u], x[D.M_col[a + q0 + u]], acc);
    }
    const int e1 = D.desc_start[i + 1];
    for (int e0 = D.desc_start[i]; e0 < e1; e0 += 8) {
      float m[8];
#pragma unroll
      for (int u = 0; u < 8; u++) m[u] = e0 + u < e1 ? qM[D.desc_idx[e0 + u]] : 0.f;
#pragma unroll
      for (int u = 0; u < 8; u++) if (e0 + u < e1) acc = fmaf(m[u], x[D.desc_dof[e0 + u]], acc);
    }
    out[i] = acc;
  }
}
__device__ __forceinline__ bool rowOn(int d0, float jar) { return ((d0 & 3) == ROW_EQ) || jar < 0.f; }
// jar = J x - aref for all rows, returns the (team-uniform) constraint cost; force/active are implied by jar
__device__ __noinline__ float residualCost(int slot, SP sm, Tm t, int ncon, int nefc, SP x) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L;
  auto bx = sm + L.basev;
  baseMul(slot, sm, t, ncon, x, bx);
  TSYNC();
  float cost = 0.f;
  TFOR(r, nefc) {
    const float jar = rowMul(slot, sm, r, x, bx) - EFC(E_AREF, r);
    EFC(E_JAR, r) = jar;
    const bool on = rowOn(EFCI(E_DESC0, r), jar);
    const float Dr = EFC(E_D, r);
    EFC(E_FORCE, r) = on ? -Dr * jar : 0.f;   // (every writer of E_JAR leaves the row's force beside it: constraintForce reads it)
    if (on) cost = fmaf(0.5f * Dr * jar, jar, cost);
  }
  TSYNC();
  return TSUM(cost);
}
// The two warm-start candidates in one sweep over J and the rows: cost[0] of x0 (rows not kept), cost[1] of x1, whose row
// residuals stay in E_JAR.
__device__ __noinline__ void residualCost2(int slot, SP sm, Tm t, int ncon, int nefc, SP x0, SP x1, float* cost) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L;
  auto bx0 = sm + L.basev + D.maxbase * D.nconmax; auto bx1 = sm + L.basev;   // (the base-force half of basev is free before the first force evaluation)
  baseMul2(slot, sm, t, ncon, x0, x1, bx0, bx1);
  TSYNC();
  float c[2] = {0.f, 0.f};
  for (int rb = t.tl; rb < nefc; rb += RB * TEAM) {
    float aref[RB], Dr[RB], j0[RB], j1[RB]; int d0[RB];
#pragma unroll
    for (int u = 0; u < RB; u++) {
      const int r = rb + u * TEAM; const bool ok = r < nefc;
      aref[u] = ok ? EFC(E_AREF, r) : 0.f; Dr[u] = ok ? EFC(E_D, r) : 0.f; d0[u] = ok ? EFCI(E_DESC0, r) : ROW_CON;
    }
    rowMulBatch(slot, sm, rb, TEAM, nefc, x0, bx0, j0);
    rowMulBatch(slot, sm, rb, TEAM, nefc, x1, bx1, j1);
#pragma unroll
    for (int u = 0; u < RB; u++) {
      const int r = rb + u * TEAM;
      if (r < nefc) {
        const float jar0 = j0[u] - aref[u], jar1 = j1[u] - aref[u];
        EFC(E_JAR, r) = jar1;
        const bool on1 = rowOn(d0[u], jar1);
        EFC(E_FORCE, r) = on1 ? -Dr[u] * jar1 : 0.f;
        if (rowOn(d0[u], jar0)) c[0] = fmaf(0.5f * Dr[u] * jar0, jar0, c[0]);
        if (on1) c[1] = fmaf(0.5f * Dr[u] * jar1, jar1, c[1]);
      }
    }
  }
  TSYNC();
  cost[0] = TSUM(c[0]); cost[1] = TSUM(c[1]);
}
// the non-contact part of (J^T f)[k]: equality, joint-limit and tendon-limit rows that touch dof k
__device__ __forceinline__ float sparseRowForce(int slot, SP sm, int k) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L;
  float acc = 0.f;
  for (int e = D.dofeq_start[k]; e < D.dofeq_start[k + 1]; e++) {
    const int v = D.dofeq_list[e], r = v >> 1;
    acc += (v & 1) ? EFC(E_COEF, r) * EFC(E_FORCE, r) : EFC(E_FORCE, r);
  }
  const int li = D.dof_limidx[k];
  if (li >= 0)
    for (int q = 0; q < 2; q++) { const int r = __float_as_int(sm[L.limrow + 2 * li + q]); if (r >= 0) acc += EFC(E_COEF, r) * EFC(E_FORCE, r); }
  if (D.nlimt > 0)
    for (int c = D.col_start[k]; c < D.col_start[k + 1]; c++) {
      const int n = D.col_nz[c], lt = D.ten_limidx[D.nz_tendon[n]];
      if (lt < 0) continue;
      for (int q = 0; q < 2; q++) {
        const int r = __float_as_int(sm[L.limrow + 2 * (D.nlimj + lt) + q]);
        if (r >= 0) acc += EFC(E_COEF, r) * EFC(E_FORCE, r) * sm[L.tenJ + n];
      }
    }
  return acc;
}
// out = sparse rows + Jb^T bf: G dofs per warp and pass (they share the loads of bf), the J loads of four base rows issued
// before their first use; even / odd base rows keep separate accumulators.  G = 2 when two passes' worth of warps cover nv
// (a third dof would be a clamped duplicate: a third of the loads wasted at nv = 23, 12 warps).
#ifndef JTB
#define JTB 4
#endif
template <int G> __device__ __forceinline__ void jtForce(int slot, SP sm, Tm t, int nbr, SP bf, SP out) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L;
  const int nv = D.nv;
  for (int kb = t.tl; kb < nv; kb += G * TEAM) {
    float res[G]; int kk[G];
#pragma unroll
    for (int g = 0; g < G; g++) { const int k = kb + g * TEAM; kk[g] = min(k, nv - 1); res[g] = k < nv ? sparseRowForce(slot, sm, k) : 0.f; }
    auto Jk = sm + L.conJ;
    float a[G][2];
#pragma unroll
    for (int g = 0; g < G; g++) a[g][0] = a[g][1] = 0.f;
    for (int b0 = 0; b0 < nbr; b0 += JTB) {
      float j[JTB][G], f[JTB];
#pragma unroll
      for (int i = 0; i < JTB; i++) {   // (rows past nbr: other rows / the next region of the scratch, masked below)
        f[i] = bf[b0 + i];
#pragma unroll
        for (int g = 0; g < G; g++) j[i][g] = LDJS(Jk[(b0 + i) * nv + kk[g]]);
      }
#pragma unroll
      for (int i = 0; i < JTB; i++)
        if (b0 + i < nbr) {
#pragma unroll
          for (int g = 0; g < G; g++) a[g][i & 1] = fmaf(j[i][g], f[i], a[g][i & 1]);
        }
    }
#pragma unroll
    for (int g = 0; g < G; g++) if (kb + g * TEAM < nv) out[kb + g * TEAM] = res[g] + a[g][0] + a[g][1];
  }
}
// out = J^T f from the rows' forces  (deterministic gathers, no atomics)
__device__ __noinline__ void constraintForce(int slot, SP sm, Tm t, int ncon, int nefc, SP out) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L;
  const int nv = D.nv, nb = D.maxbase;
#ifdef CF_TICKS
  long long* dbg = tP[slot].X.dbg; long long tlast = dbg ? clock64() : 0;
#define CFT(k) do { if (dbg && threadIdx.x == 0) { const long long now_ = clock64(); dbg[blockIdx.x * 32 + (k)] += now_ - tlast; tlast = now_; } } while (0)
#else
#define CFT(k)
#endif
  // (the rows' forces are already in E_FORCE: residualCost / residualCost2 and the solver's row update write them with E_JAR,
  // behind barriers of their own)
  CFT(16);
  auto bf = sm + L.basev + nb * D.nconmax;
  // base-row forces of a contact from its pyramid rows (mju_decodePyramid, transposed); two contacts per iteration, their
  // record loads and then their row loads issued together
  for (int cb = t.tl; cb < ncon; cb += 2 * TEAM) {
    int adr[2], dim[2];
#pragma unroll
    for (int u = 0; u < 2; u++) {
      const int c = cb + u * TEAM; const bool ok = c < ncon;
      auto cr = sm + L.con + 12 * (ok ? c : cb);
      adr[u] = ok ? __float_as_int(cr[8]) : -1; dim[u] = __float_as_int(cr[9]);
    }
    float fr[2][6], mu[2][3];   // rows of a contact: 2 (dim - 1) <= 6, friction coefficients of its tangent pairs
#pragma unroll
    for (int u = 0; u < 2; u++) {
      const int n = adr[u] < 0 ? 0 : (dim[u] == 1 ? 1 : 2 * (dim[u] - 1));
#pragma unroll
      for (int k = 0; k < 6; k++) fr[u][k] = k < n ? EFC(E_FORCE, adr[u] + k) : 0.f;
#pragma unroll
      for (int k = 0; k < 3; k++) mu[u][k] = (dim[u] > 1 && 2 * k < n) ? EFC(E_COEF, adr[u] + 2 * k) : 0.f;
    }
#pragma unroll
    for (int u = 0; u < 2; u++) {
      const int c = cb + u * TEAM;
      if (c >= ncon) continue;
      float fb[4] = {0.f, 0.f, 0.f, 0.f};
      if (adr[u] >= 0) {
        if (dim[u] == 1) fb[0] = fr[u][0];
        else
#pragma unroll
          for (int k = 1; k < 4; k++)
            if (k < dim[u]) { const float fp = fr[u][2 * (k - 1)], fm = fr[u][2 * (k - 1) + 1]; fb[0] += fp + fm; fb[k] = mu[u][k - 1] * (fp - fm); }
      }
      for (int k = 0; k < nb; k++) bf[c * nb + k] = fb[k];
    }
  }
  TSYNC();
  CFT(17);
  const int nbr = nb * ncon;
  if (nv <= 2 * TEAM) jtForce<2>(slot, sm, t, nbr, bf, out); else jtForce<3>(slot, sm, t, nbr, bf, out);
  TSYNC();
  CFT(18);
}
// One point of the 1-D cost along the search direction.  `tol` is the convergence threshold for |d0| at this point:
// MuJoCo's gtol, or the float32 noise floor of the derivative sum when that is larger (the sum cancels to ~eps * mag
// at the minimiser; without the floor the search burns all ls_iterations chasing rounding noise -- measured on the
// float build of the oracle: 34 evaluations per search instead of the double build's 2.8).
#ifndef LS_NOISE
#define LS_NOISE 2e-6f
#endif
struct LsPt { float alpha, cost, d0, d1, tol; };
// ------------------------------------------------------------------------------------------ touch sensors (mj_sensorAcc)
// sensordata[s] = sum of the normal forces of the active contacts that involve the sensor site's body and whose
// normal ray (from the contact point, pointing into that body's side) meets the site's zone.  The pyramidal normal
// force is the sum of the contact's row forces (mju_decodePyramid).
__device__ __noinline__ void touchSensors(int slot, int fslot, SP sm, Tm t, World W, int ncon) {
  const DevModel& D = tP[slot].D; const SmemLayout& L = tP[slot].L; const FieldPtrs& F = tP[fslot].F;
  void* out = F.p[MYO_F_SENSORDATA];
  if (!out) return;
  TFOR(s, D.nsensor) {
    const int sid = D.sens_site[s], body = D.sens_body[s], f = D.site_frame[sid];
    const f3 spos = ld3(FR_POS(sm, L, f)) + mulM(FR_MAT(sm, L, f), ld3(D.site_lpos + 3 * sid));
    float R[9]; mulMM(R, FR_MAT(sm, L, f), D.site_lmat + 9 * sid);
    const f3 size = ld3(D.sens_size + 3 * s);
    float total = 0.f;
    for (int c = 0; c < ncon; c++) {
      auto cr = sm + L.con + 12 * c;
      const int adr = __float_as_int(cr[8]), dim = __float_as_int(cr[9]), prw = __float_as_int(cr[7]), pr = CON_PAIR(prw);
      if (adr < 0) continue;
      const int b1 = D.cg_body[CON_SWAP(prw) ? D.pair_cg2[pr] : D.pair_cg1[pr]], b2 = D.cg_body[CON_SWAP(prw) ? D.pair_cg1[pr] : D.pair_cg2[pr]];
      if (body != b1 && body != b2) continue;
      float fn = 0.f;
      const int rows = dim == 1 ? 1 : 2 * (dim - 1);
      for (int q = 0; q < rows; q++) fn += EFC(E_FORCE, adr + q);
      if (!(fn > 0.f)) continue;
      f3 dir = unit3(ld3(cr + 4) * fn);
      if (body == b2) dir = dir * -1.f;
      const f3 p = mulMT(R, ld3(cr + 1) - spos), v = mulMT(R, dir);   // ray in the site's frame
      bool hit;
      if (D.sens_type[s] == G_SPHERE) {
        const float b = dot(p, v), cc = dot(p, p) - size.x * size.x, disc = b * b - cc;
        hit = disc >= 0.f && (-b + sqrtf(disc)) >= 0.f;
      } else {   // box: slab test, the ray t >= 0 meets the solid box
        float t0 = 0.f, t1 = 3.4e38f;
        const float pp[3] = {p.x, p.y, p.z}, vv[3] = {v.x, v.y, v.z}, hh[3] = {size.x, size.y, size.z};
        hit = true;
        for (int k = 0; k < 3; k++) {
          if (fabsf(vv[k]) < MINVAL) { if (fabsf(pp[k]) > hh[k]) hit = false; }
          else {
            const float inv = 1.f / vv[k];
            float a = (-hh[k] - pp[k]) * inv, b = (hh[k] - pp[k]) * inv;
            if (a > b) { const float tmp = a; a = b; b = tmp; }
            t0 = fmaxf(t0, a); t1 = fminf(t1, b);
          }
        }
        hit = hit && t0 <= t1;
      }
      if (hit) total += fn;
    }
    OUTF(out, s) = total;
  }
}

}  // namespace
#endif  // MYO_COMMON_CUH_
