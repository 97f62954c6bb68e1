// C ABI of libmyostep.so (see include/myostep.h).  Thin: argument checks, buffer binding, kernel launch.
// There is NO CPU fallback: every entry that computes needs a CUDA device and fails with MYO_ERR_CUDA without one.
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <atomic>
#include <functional>
#include <mutex>
#include <string>
#include <vector>

#include "devmodel.h"
#include "internal.h"


namespace {
std::atomic<int64_t> g_launches{0};
thread_local std::string g_err;
bool g_slot_used[64] = {false};   // guarded by g_slot_mutex: batches may be created and destroyed from several host threads
std::mutex g_slot_mutex;
void releaseSlot(int slot) { if (slot >= 0) { std::lock_guard<std::mutex> lk(g_slot_mutex); g_slot_used[slot] = false; } }
int fail(int code, const std::string& msg) { g_err = msg; return code; }
}  // namespace
int myo_fail(int code, const std::string& msg) { return fail(code, msg); }
void myo_count_launch(int n) { g_launches += n; }
namespace {

const char* kFieldNames[MYO_F_COUNT] = {
    "qpos", "qvel", "act", "ctrl", "time", "qacc_warmstart", "qacc", "qacc_smooth", "act_dot", "xpos", "xquat", "xmat", "xipos",
    "ximat", "xanchor", "xaxis", "geom_xpos", "geom_xmat", "site_xpos", "site_xmat", "subtree_com", "cinert", "crb", "cdof",
    "cdof_dot", "cvel", "ten_length", "ten_velocity", "ten_J", "ten_wrapadr", "ten_wrapnum", "wrap_obj", "wrap_xpos",
    "actuator_length", "actuator_velocity", "actuator_moment", "actuator_force", "qM", "qLD", "qLDiagInv", "qfrc_bias",
    "qfrc_passive", "qfrc_spring", "qfrc_damper", "qfrc_actuator", "qfrc_smooth", "qfrc_constraint", "qfrc_inverse", "ncon",
    "nefc", "contact_geom", "contact_dim", "contact_efc_address", "contact_dist", "contact_pos", "contact_frame",
    "contact_includemargin", "contact_friction", "contact_solref", "contact_solimp", "efc_J", "efc_pos", "efc_margin", "efc_D",
    "efc_aref", "efc_force", "efc_type", "solver_niter", "status", "qfrc_applied", "sensordata", "geom_type_ovr", "geom_size_ovr",
    "geom_pos_ovr", "act_gainscale"};

bool isIntField(int f) {
  switch (f) {
    case MYO_F_TEN_WRAPADR: case MYO_F_TEN_WRAPNUM: case MYO_F_WRAP_OBJ: case MYO_F_NCON: case MYO_F_NEFC: case MYO_F_CONTACT_GEOM:
    case MYO_F_CONTACT_DIM: case MYO_F_CONTACT_EFC_ADDRESS: case MYO_F_EFC_TYPE: case MYO_F_SOLVER_NITER: case MYO_F_STATUS:
    case MYO_F_GEOM_TYPE_OVR: return true;
    default: return false;
  }
}
int64_t fieldCount(const DevModel& D, int f) {
  const int64_t nv = D.nv, nb = D.nbody, nc = D.nconmax, nj = D.njmax;
  switch (f) {
    case MYO_F_QPOS: return D.nq;
    case MYO_F_QVEL: case MYO_F_QACC_WARMSTART: case MYO_F_QACC: case MYO_F_QACC_SMOOTH: case MYO_F_QLDIAGINV: case MYO_F_QFRC_BIAS:
    case MYO_F_QFRC_PASSIVE: case MYO_F_QFRC_SPRING: case MYO_F_QFRC_DAMPER: case MYO_F_QFRC_ACTUATOR: case MYO_F_QFRC_SMOOTH:
    case MYO_F_QFRC_CONSTRAINT: case MYO_F_QFRC_INVERSE: case MYO_F_QFRC_APPLIED: return nv;
    case MYO_F_ACT: case MYO_F_ACT_DOT: return D.na;
    case MYO_F_CTRL: case MYO_F_ACTUATOR_LENGTH: case MYO_F_ACTUATOR_VELOCITY: case MYO_F_ACTUATOR_FORCE: case MYO_F_ACT_GAINSCALE: return D.nu;
    case MYO_F_GEOM_TYPE_OVR: return D.ngeom;
    case MYO_F_GEOM_SIZE_OVR: case MYO_F_GEOM_POS_OVR: return 3 * (int64_t)D.ngeom;
    case MYO_F_SENSORDATA: return D.nsensor;
    case MYO_F_TIME: case MYO_F_NCON: case MYO_F_NEFC: case MYO_F_SOLVER_NITER: case MYO_F_STATUS: return 1;
    case MYO_F_XPOS: case MYO_F_XIPOS: case MYO_F_SUBTREE_COM: return 3 * nb;
    case MYO_F_XQUAT: return 4 * nb;
    case MYO_F_XMAT: case MYO_F_XIMAT: return 9 * nb;
    case MYO_F_XANCHOR: case MYO_F_XAXIS: return 3 * (int64_t)D.njnt;
    case MYO_F_GEOM_XPOS: return 3 * (int64_t)D.ngeom;
    case MYO_F_GEOM_XMAT: return 9 * (int64_t)D.ngeom;
    case MYO_F_SITE_XPOS: return 3 * (int64_t)D.nsite;
    case MYO_F_SITE_XMAT: return 9 * (int64_t)D.nsite;
    case MYO_F_CINERT: case MYO_F_CRB: return 10 * nb;
    case MYO_F_CDOF: case MYO_F_CDOF_DOT: return 6 * nv;
    case MYO_F_CVEL: return 6 * nb;
    case MYO_F_TEN_LENGTH: case MYO_F_TEN_VELOCITY: return D.nten;
    case MYO_F_TEN_J: return (int64_t)D.nten * nv;
    case MYO_F_TEN_WRAPADR: case MYO_F_TEN_WRAPNUM: return D.nten;
    case MYO_F_WRAP_OBJ: return 2 * (int64_t)D.nwrap;     // one entry per path point (mjData: [nwrap, 2])
    case MYO_F_WRAP_XPOS: return 6 * (int64_t)D.nwrap;    // three per path point (mjData: [nwrap, 6])
    case MYO_F_ACTUATOR_MOMENT: return (int64_t)D.nu * nv;
    case MYO_F_QM: case MYO_F_QLD: return nv * nv;
    case MYO_F_CONTACT_GEOM: case MYO_F_CONTACT_SOLREF: return 2 * nc;
    case MYO_F_CONTACT_DIM: case MYO_F_CONTACT_EFC_ADDRESS: case MYO_F_CONTACT_DIST: case MYO_F_CONTACT_INCLUDEMARGIN: return nc;
    case MYO_F_CONTACT_POS: return 3 * nc;
    case MYO_F_CONTACT_FRAME: return 9 * nc;
    case MYO_F_CONTACT_FRICTION: case MYO_F_CONTACT_SOLIMP: return 5 * nc;
    case MYO_F_EFC_J: return nj * nv;
    case MYO_F_EFC_POS: case MYO_F_EFC_MARGIN: case MYO_F_EFC_D: case MYO_F_EFC_AREF: case MYO_F_EFC_FORCE: case MYO_F_EFC_TYPE: return nj;
    default: return 0;
  }
}
const int kStateFields[] = {MYO_F_QPOS, MYO_F_QVEL, MYO_F_ACT, MYO_F_CTRL, MYO_F_TIME, MYO_F_QACC_WARMSTART, MYO_F_STATUS};

int checkReady(myo_batch* b) {
  if (!b) return fail(MYO_ERR_ARG, "null batch");
  for (int f : kStateFields)
    if (!b->F.p[f]) return fail(MYO_ERR_UNBOUND, std::string("state field not bound: ") + kFieldNames[f]);
  if (cudaSetDevice(b->device) != cudaSuccess) return fail(MYO_ERR_CUDA, "cudaSetDevice failed");
  return MYO_OK;
}
}  // namespace

extern "C" {

const char* myo_last_error(void) { return g_err.c_str(); }
int64_t myo_launch_count(void) { return g_launches; }
const char* myo_field_name(int field) { return (field >= 0 && field < MYO_F_COUNT) ? kFieldNames[field] : ""; }
int myo_field_itemsize(int field) { (void)field; return 4; }
int myo_field_is_int(int field) { return isIntField(field) ? 1 : 0; }

int myo_model_create(const myo_model_desc* desc, int device, myo_model** out) {
  if (!desc || !out) return fail(MYO_ERR_ARG, "null argument");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
    return fail(MYO_ERR_CUDA, "no CUDA device: libmyostep has no CPU fallback");
  if (device < 0 || device >= ndev || cudaSetDevice(device) != cudaSuccess) return fail(MYO_ERR_CUDA, "bad CUDA device index");
  myo_model* m = new myo_model();
  m->device = device;
  try {
    myo_build_plan(*desc, 0, 0, m->D, m->L, m->owned);
  } catch (const std::string& e) {
    for (void* p : m->owned) cudaFree(p);
    delete m;
    const bool cuda = e.find("cuda") != std::string::npos;
    return fail(cuda ? MYO_ERR_CUDA : MYO_ERR_UNSUPPORTED, e);
  }
  *out = m;
  return MYO_OK;
}
void myo_model_destroy(myo_model* m) {
  if (!m) return;
  cudaSetDevice(m->device);
  for (void* p : m->owned) cudaFree(p);
  delete m;
}

int myo_batch_create(const myo_model* m, int nworld, int nconmax, int njmax, myo_batch** out) {
  if (!m || !out || nworld <= 0) return fail(MYO_ERR_ARG, "bad argument");
  if (cudaSetDevice(m->device) != cudaSuccess) return fail(MYO_ERR_CUDA, "cudaSetDevice failed");
  myo_batch* b = new myo_batch();
  b->model = m; b->D = m->D; b->nworld = nworld; b->slot = -1; b->dirty = true; b->device = m->device;
  {
    std::lock_guard<std::mutex> lk(g_slot_mutex);
    for (int i = 0; i < MYO_PARAM_SLOTS; i++) if (!g_slot_used[i]) { b->slot = i; g_slot_used[i] = true; break; }
  }
  if (b->slot < 0) { delete b; return fail(MYO_ERR_ARG, "too many live batches (16 constant-memory parameter slots)"); }
  memset(&b->F, 0, sizeof(b->F));
  try {
    myo_make_layout(b->D, nconmax, njmax, b->L);
  } catch (const std::string& e) { releaseSlot(b->slot); delete b; return fail(MYO_ERR_ARG, e); }
  for (int f = 0; f < MYO_F_COUNT; f++)   // the kernels index fields with 32-bit element offsets k * nworld + w
    if (fieldCount(b->D, f) * (int64_t)nworld >= ((int64_t)1 << 32)) {
      releaseSlot(b->slot); delete b;
      return fail(MYO_ERR_ARG, std::string("batch too large: field ") + kFieldNames[f] + " would exceed 2^32 elements (split the worlds over more batches)");
    }
  {
    void* p = nullptr;
    // tile-blocked scratch [tiles][L.total][32]; zeroed once: a record the kernels have not written yet (rows past an
    // overflow, lanes of a partial tile) reads as zeros
    const size_t bytes = (size_t)((nworld + 31) / 32) * b->L.total * 32 * sizeof(float);
    if (cudaMalloc(&p, bytes) != cudaSuccess || cudaMemset(p, 0, bytes) != cudaSuccess) { releaseSlot(b->slot); delete b; return fail(MYO_ERR_CUDA, "cudaMalloc of the per-world scratch failed"); }
    b->owned.push_back(p); b->scratch = (float*)p;
    // work list of the convex narrow phase: at most every MPR-type candidate pair of every world
    b->X.mpr_cap = (int)std::min<int64_t>((int64_t)nworld * std::max(1, b->D.npair), (int64_t)1 << 30);   // (a type override can turn any pair convex)
    void* q = nullptr;
    if (cudaMalloc(&q, (size_t)b->X.mpr_cap * sizeof(int2) + 256) != cudaSuccess || cudaMemset(q, 0, 256) != cudaSuccess) { myo_batch_destroy(b); return fail(MYO_ERR_CUDA, "cudaMalloc of the narrow-phase work list failed"); }
    b->owned.push_back(q); b->X.mpr_count = (int*)q; b->X.mpr_list = (int2*)((char*)q + 256);
    b->X.dbg = nullptr;
    if (getenv("MYO_NEWTON_TIMING")) {   // developer knob: per-tile clock sums of the solver's phases (myo_debug_timing)
      void* t = nullptr;
      const size_t tb = (size_t)((nworld + 31) / 32) * 32 * sizeof(long long);
      if (cudaMalloc(&t, tb + 64 * sizeof(long long)) == cudaSuccess && cudaMemset(t, 0, tb + 64 * sizeof(long long)) == cudaSuccess) {
        b->owned.push_back(t); b->X.dbg = (long long*)t; b->X.tl_off = (int)(tb / sizeof(long long));
      }
    }
  }
  // library-owned default state buffers (bind overrides them)
  for (int f : kStateFields) {
    void* p = nullptr;
    size_t n = (size_t)nworld * std::max<int64_t>(1, fieldCount(b->D, f)) * 4;
    if (cudaMalloc(&p, n) != cudaSuccess || cudaMemset(p, 0, n) != cudaSuccess) { myo_batch_destroy(b); return fail(MYO_ERR_CUDA, "cudaMalloc failed"); }
    b->owned.push_back(p); b->F.p[f] = p;
  }
  *out = b;
  return MYO_OK;
}
void myo_batch_destroy(myo_batch* b) {
  if (!b) return;
  if (cudaSetDevice(b->device) != cudaSuccess) cudaGetLastError();   // (b->model may be gone already; never leave an error pending for the next launch check)
  for (void* p : b->owned) cudaFree(p);
  for (int i = 0; i < 5; i++) if (b->side[i]) { cudaStreamSynchronize((cudaStream_t)b->side[i]); cudaStreamDestroy((cudaStream_t)b->side[i]); }
  for (int i = 0; i < 10; i++) if (b->ev[i]) cudaEventDestroy((cudaEvent_t)b->ev[i]);
  releaseSlot(b->slot);
  delete b;
}
int myo_batch_nconmax(const myo_batch* b) { return b ? b->D.nconmax : 0; }
int myo_batch_njmax(const myo_batch* b) { return b ? b->D.njmax : 0; }
int myo_batch_smem_bytes(const myo_batch* b) { return b ? b->L.total * 4 : 0; }
int myo_batch_team(const myo_batch* b) { return b ? 1 : 0; }
int myo_batch_ctas_per_sm(const myo_batch* b) { return b ? 1 : 0; }
int myo_batch_engine(const myo_batch* b) { return b ? 1 : -1; }
int64_t myo_field_count(const myo_batch* b, int field) { return (b && field >= 0 && field < MYO_F_COUNT) ? fieldCount(b->D, field) : 0; }

int myo_batch_bind(myo_batch* b, int field, void* dev_ptr, int64_t nbytes) {
  if (!b || field < 0 || field >= MYO_F_COUNT) return fail(MYO_ERR_ARG, "bad field id");
  if (dev_ptr) {
    const int64_t need = (int64_t)b->nworld * fieldCount(b->D, field) * 4;
    if (need == 0) return fail(MYO_ERR_ARG, std::string("field is empty for this model: ") + kFieldNames[field]);
    if (nbytes != need) return fail(MYO_ERR_ARG, std::string("size mismatch binding ") + kFieldNames[field]);
    if (((uintptr_t)dev_ptr & 3) != 0) return fail(MYO_ERR_ARG, "buffer must be 4-byte aligned");
  } else {
    for (int f : kStateFields) if (f == field) return fail(MYO_ERR_ARG, "state fields cannot be unbound");
  }
  if (b->F.p[field] == dev_ptr) return MYO_OK;   // re-binding the same buffer (a caller that binds on every call) changes nothing
  b->F.p[field] = dev_ptr;
  b->dirty = true;
  return MYO_OK;
}

// Path-point buffers: the kernels clear only the entries a wrap state can reach; the rest of a newly bound buffer is
// zeroed here once (mj_makeData's zeros), on the caller's stream.
static int clearPathPoints(myo_batch* b, void* stream) {
  const int wf[2] = {MYO_F_WRAP_OBJ, MYO_F_WRAP_XPOS};
  for (int i = 0; i < 2; i++) {
    void* p = b->F.p[wf[i]];
    if (!p || p == b->wrap_cleared[i]) continue;
    if (cudaMemsetAsync(p, 0, (size_t)b->nworld * fieldCount(b->D, wf[i]) * 4, (cudaStream_t)stream) != cudaSuccess)
      return fail(MYO_ERR_CUDA, "clearing the path-point buffers failed");
    b->wrap_cleared[i] = p;
  }
  return MYO_OK;
}

// The phase pipeline (tile.cu, lane = world): one mj_step = kin, then  collide > collideMpr > collideCompact > makecon  |
// tendon > velact  |  inertia  on three library streams, the output fields of the position stage (posout) and the contact /
// constraint records (conout) on a fourth, low-priority one; all join before newton on the caller's stream, whose tail is
// mj_Euler and the state outputs.  Everything is ordered behind the caller's stream on entry and joined into it on exit, so
// the call is stream-ordered like a single kernel.  MYO_CONCURRENT=0 (developer knob) or a partial pass (posvel) runs
// the same kernels in order on the caller's stream.  Inner substeps of a fused launch materialise nothing.
static inline int ldlTabIntsHost(const DevModel& D) { return (D.ndlevel + 1) + D.nv + (D.nv + 1) + 2 * (D.nM - D.nv) + 2 * D.nv + D.nM; }   // = ldlTabInts (common.cuh)
// velactKernel's shared-memory request: (nM + 2 nv) floats per lane << 8 | ints of the LDL index tables << 20; 0 = not staged
static inline int velactStage(const DevModel& D) { const int f = D.nM + 2 * D.nv, ti = ldlTabIntsHost(D); return (f > 0xfff || ti > 0xfff) ? 0 : (f << 8) | (ti << 20); }
// inertiaKernel's shared-memory request: nM << 8 | ints of the LDL factorisation tables << 20; 0 = not staged
static inline int inertiaStage(const DevModel& D) { const int ti = (D.nv + 1) + 3 * D.nldlop + 2 * D.nv; return (D.nM > 0xfff || ti > 0x7ff) ? 0 : (D.nM << 8) | (ti << 20); }
static int launchPhases(myo_batch* b, int nsub, int forward_only, const int32_t* mask, int do_reset, void* stream) {
  int rc;
  if (b->dirty) {
    if (myo_tile_upload(b->slot, b->D, b->L, b->F, b->X, stream) != 0) return fail(MYO_ERR_CUDA, "parameter upload failed");
    static const int conFields[] = {MYO_F_NCON, MYO_F_NEFC, MYO_F_SOLVER_NITER, MYO_F_CONTACT_GEOM, MYO_F_CONTACT_DIM, MYO_F_CONTACT_EFC_ADDRESS,
                                    MYO_F_CONTACT_DIST, MYO_F_CONTACT_POS, MYO_F_CONTACT_FRAME, MYO_F_CONTACT_INCLUDEMARGIN, MYO_F_CONTACT_FRICTION,
                                    MYO_F_CONTACT_SOLREF, MYO_F_CONTACT_SOLIMP, MYO_F_EFC_J, MYO_F_EFC_POS, MYO_F_EFC_MARGIN, MYO_F_EFC_D,
                                    MYO_F_EFC_AREF, MYO_F_EFC_FORCE, MYO_F_EFC_TYPE};
    b->conout = false;
    for (int f : conFields) b->conout |= b->F.p[f] != nullptr;
    if ((rc = clearPathPoints(b, stream)) != MYO_OK) return rc;
    b->dirty = false;
  }
  static const bool concurrent = !(getenv("MYO_CONCURRENT") && atoi(getenv("MYO_CONCURRENT")) == 0);   // developer knob
  if (concurrent && !b->side[0]) {
    // side streams: 0 = collision chain + constraint rows (the longest branch: highest priority, so its CTAs get the first
    // free slots), 1 = tendon -> velact, 2 = inertia, 3 = contact / constraint records (lowest priority: they only have to
    // be out by the end of the step and must not take slots from the solver)
    int least = 0, greatest = 0;
    cudaDeviceGetStreamPriorityRange(&least, &greatest);
    static const bool prio = !(getenv("MYO_PRIORITY") && atoi(getenv("MYO_PRIORITY")) == 0);   // developer knob
    const int pr[5] = {prio ? greatest : least, prio ? std::min(least, greatest + 1) : least, prio ? std::min(least, greatest + 2) : least, least, least};
    for (int i = 0; i < 5; i++) { cudaStream_t cs; if (cudaStreamCreateWithPriority(&cs, cudaStreamNonBlocking, pr[i]) != cudaSuccess) return fail(MYO_ERR_CUDA, "side stream creation failed"); b->side[i] = cs; }
    for (int i = 0; i < 10; i++) { cudaEvent_t e; if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) return fail(MYO_ERR_CUDA, "event creation failed"); b->ev[i] = e; }
  }
  const bool posvel = forward_only == 2;
  const int nloop = forward_only ? 1 : nsub;
  const int slot = b->slot, nw = b->nworld;
  float* scr = b->scratch;
  static const bool debug_sync = getenv("MYO_DEBUG_SYNC") != nullptr;   // developer knob: synchronise after every kernel and name the one that failed
  void* cur = stream;   // the stream the next kernel goes to
  auto tile = [&](int k, int fslot, int flags) {
    g_launches++;
    int r = myo_tile_launch(k, slot, fslot, nw, mask, flags, scr, cur);
    if (r != 0 && debug_sync) fprintf(stderr, "[myostep] launch of tile kernel %d failed (flags %d): %s\n", k, flags, cudaGetErrorString(cudaPeekAtLastError()));
    if (debug_sync && r == 0) {
      const cudaError_t e = cudaStreamSynchronize((cudaStream_t)cur);
      if (e != cudaSuccess) { fprintf(stderr, "[myostep] tile kernel %d failed: %s\n", k, cudaGetErrorString(e)); r = -1; }
    }
    return r;
  };
  bool ok = true; bool reported_ = false;
#define DBGL() do { if (!ok && debug_sync && !reported_) { fprintf(stderr, "[myostep] launchPhases: first failure before api.cu line %d: %s\n", __LINE__, cudaGetErrorString(cudaPeekAtLastError())); reported_ = true; } } while (0)
  // mj_collision = broad phase + closed-form pairs, the listed convex pairs one per lane, ordered compaction
  // at_mark(): after kernel `mark` of the chain (1 = broad phase, 2 = convex narrow phase); fuse_rows: the compaction runs as
  // the head of the constraint-row kernel (MYO_K_COMPACT_MAKECON), launched by the caller
  auto collide = [&](int fslot, int mark = 0, const std::function<void()>& at_mark = nullptr, bool fuse_rows = false) {
    bool r = true;
    r &= tile(MYO_K_COLLIDE, fslot, 0) == 0;
    if (mark == 1) at_mark();
    if (b->D.npair > 0) {
      if (b->D.nmprpair > 0 || b->F.p[MYO_F_GEOM_TYPE_OVR]) r &= tile(MYO_K_COLLIDE_MPR, fslot, 0) == 0;
      if (mark == 2) at_mark();
      if (!fuse_rows) r &= tile(MYO_K_COLLIDE_COMPACT, fslot, 0) == 0;
    } else if (mark == 2) at_mark();
    return r;
  };
  static const int debug_stop = getenv("MYO_DEBUG_STOP") ? atoi(getenv("MYO_DEBUG_STOP")) : -1;   // developer knob: end the pipeline after this many kernels
  int nk = 0;
#define STOP_HERE() if (debug_stop >= 0 && ++nk >= debug_stop) break
  for (int sub = 0; sub < nloop && ok; sub++) {
    const bool first = sub == 0, last = sub == nloop - 1;
    const int fslot = last ? slot : MYO_PARAM_SLOTS;   // slot MYO_PARAM_SLOTS: no field bound, "materialise nothing"
    ok &= tile(MYO_K_KIN, fslot, (first ? (do_reset ? MYO_KIN_RESET : MYO_KIN_LOAD) : 0) | (forward_only ? 0 : MYO_KIN_CHECK)) == 0; STOP_HERE();
    // mj_makeConstraint needs the contacts and qvel only -- unless the model has tendon limits (tendon length / velocity)
    const bool makecon_early = b->D.nlimt == 0;
    const bool posout = fslot != MYO_PARAM_SLOTS;   // (field tests are per pointer inside the kernel)
    const bool conout = last && b->conout;
    if (concurrent && !posvel && debug_stop < 0) {
      // fork after kin: collision chain (+ constraint rows, + records out) | tendon -> velact | inertia; join before the solver
      auto S = [&](int i) { return (cudaStream_t)b->side[i]; };
      auto E = [&](int i) { return (cudaEvent_t)b->ev[i]; };
      const cudaStream_t main = (cudaStream_t)stream;
      ok &= cudaEventRecord(E(0), main) == cudaSuccess;
      DBGL();
      for (int i = 0; i < 4; i++) ok &= cudaStreamWaitEvent(S(i), E(0), 0) == cudaSuccess;
      // output stream (lowest priority): the position stage's fields now, the contact / constraint records once the rows exist
      static const int ndelay = getenv("MYO_DELAY") ? atoi(getenv("MYO_DELAY")) : 1;   // developer knob
      static const int posout_after = getenv("MYO_POSOUT_AFTER") ? atoi(getenv("MYO_POSOUT_AFTER")) : 0;   // developer knob: 0 = after kin (best), 1 = after collide (+0.4 %), 2 = after collideMpr (+1.4 %)
      // developer knob MYO_OUT_LATE: bit 0 = the tendon outputs, bit 1 = the position outputs start with the solver (stream 4)
      static const int out_late = getenv("MYO_OUT_LATE") ? atoi(getenv("MYO_OUT_LATE")) : 2;   // 0: +3 % step time, 1: +1 %, 3: +2.7 %
      if (posout && posout_after == 0 && !(out_late & 2)) {
        for (int i = 0; i < 2 * ndelay; i++) { g_launches++; ok &= myo_nop_launch(b->side[3]) == 0; }   // (behind tendon / inertia, see below)
        cur = b->side[3]; ok &= tile(MYO_K_POSOUT, fslot, 0) == 0;
      }
      static const bool conout_late = getenv("MYO_CONOUT_LATE") && atoi(getenv("MYO_CONOUT_LATE"));   // developer knob
      static const bool fuse_compact = getenv("MYO_FUSE_COMPACT") && atoi(getenv("MYO_FUSE_COMPACT"));   // developer knob (measured: the fused launch is 0.4 % slower at steady state, 3 % in the first steps)
      const bool fuse_rows = fuse_compact && makecon_early && b->D.npair > 0;
      auto rows = [&](cudaStream_t on) {   // on the stream `cur` = `on`
        ok &= tile(fuse_rows ? MYO_K_COMPACT_MAKECON : MYO_K_MAKECON, fslot, 0) == 0;
      DBGL();
        if (conout && !conout_late) {
          ok &= cudaEventRecord(E(3), on) == cudaSuccess && cudaStreamWaitEvent(S(3), E(3), 0) == cudaSuccess;
      DBGL();
          void* keep = cur; cur = b->side[3]; ok &= tile(MYO_K_CONOUT, fslot, b->count_aware ? 1 : 0) == 0; cur = keep;
        }
      };
      // the position outputs start behind the broad phase: the convex narrow phase (201 registers per thread) then takes the
      // SM slots the first kernels free before the output kernel does (stream priorities decide between waiting CTAs only)
      int collide_mark = posout && !(out_late & 2) ? posout_after : 0;
      cur = b->side[0]; ok &= collide(fslot, collide_mark, [&]() {
        ok &= cudaEventRecord(E(7), S(0)) == cudaSuccess && cudaStreamWaitEvent(S(3), E(7), 0) == cudaSuccess;
      DBGL();
        void* keep = cur; cur = b->side[3]; ok &= tile(MYO_K_POSOUT, fslot, 0) == 0; cur = keep;
      }, fuse_rows);
      for (int i = 0; i < ndelay; i++) { g_launches += 2; ok &= myo_nop_launch(b->side[1]) == 0 && myo_nop_launch(b->side[2]) == 0; }
      cur = b->side[1]; ok &= tile(MYO_K_TENDON, fslot, 0) == 0;
      if (posout && !(out_late & 1)) {   // ten_length, ten_J, path points: on the output stream, behind the tendon stage
        ok &= cudaEventRecord(E(6), S(1)) == cudaSuccess && cudaStreamWaitEvent(S(3), E(6), 0) == cudaSuccess;
      DBGL();
        cur = b->side[3]; ok &= tile(MYO_K_TENOUT, fslot, 0) == 0; cur = b->side[1];
      }
      cur = b->side[2]; ok &= tile(MYO_K_INERTIA, fslot, (forward_only ? 0 : 1) | inertiaStage(b->D)) == 0;
      ok &= cudaEventRecord(E(2), S(2)) == cudaSuccess && cudaStreamWaitEvent(S(1), E(2), 0) == cudaSuccess;
      DBGL();
      cur = b->side[1]; ok &= tile(MYO_K_VELACT, fslot, velactStage(b->D)) == 0;
      ok &= cudaEventRecord(E(5), S(1)) == cudaSuccess;
      DBGL();
      cur = b->side[0];   // (after the tendon branch in host order: the output stream then holds posout, tenout, conout in that order)
      if (makecon_early) rows(S(0));
      ok &= cudaEventRecord(E(1), S(0)) == cudaSuccess;
      DBGL();
      cur = stream;
      ok &= cudaStreamWaitEvent(main, E(1), 0) == cudaSuccess && cudaStreamWaitEvent(main, E(5), 0) == cudaSuccess;
      DBGL();
      if (!makecon_early) rows(main);
      if (posout && out_late) {
        ok &= cudaEventRecord(E(8), main) == cudaSuccess && cudaStreamWaitEvent((cudaStream_t)b->side[4], E(8), 0) == cudaSuccess;
      DBGL();
      }
      const int fin = (forward_only ? 0 : MYO_FIN_EULER) | (last && (!forward_only || do_reset) ? MYO_FIN_STATE : 0) | (last ? MYO_FIN_WARM | MYO_FIN_ACC | (b->count_aware ? MYO_FIN_LIVE : 0) : 0);
      static const bool fuse_finish = !(getenv("MYO_FUSE_FINISH") && atoi(getenv("MYO_FUSE_FINISH")) == 0);   // developer knob
      ok &= tile(MYO_K_NEWTON, fslot, (forward_only ? 0 : 1) | (fuse_finish ? (fin << 1) | 32 : 0) | (b->count_aware ? 64 : 0) | (b->D.nv << 8)) == 0;
      DBGL();
      if (conout && conout_late) {
        ok &= cudaEventRecord(E(3), main) == cudaSuccess && cudaStreamWaitEvent(S(3), E(3), 0) == cudaSuccess;
      DBGL();
        cur = b->side[3]; ok &= tile(MYO_K_CONOUT, fslot, b->count_aware ? 1 : 0) == 0; cur = stream;
      }
      if (!fuse_finish) ok &= tile(MYO_K_FINISH, fslot, fin) == 0;
      if (posout && out_late) {
        cur = b->side[4];
        if (out_late & 1) ok &= tile(MYO_K_TENOUT, fslot, 0) == 0;
        if (out_late & 2) ok &= tile(MYO_K_POSOUT, fslot, 0) == 0;
        cur = stream;
        ok &= cudaEventRecord(E(9), (cudaStream_t)b->side[4]) == cudaSuccess && cudaStreamWaitEvent(main, E(9), 0) == cudaSuccess;
      DBGL();
      }
      if (conout || posout) ok &= cudaEventRecord(E(4), S(3)) == cudaSuccess && cudaStreamWaitEvent(main, E(4), 0) == cudaSuccess;
      continue;
    }
    if (posout) { ok &= tile(MYO_K_POSOUT, fslot, 0) == 0; STOP_HERE(); }
    ok &= tile(MYO_K_TENDON, fslot, 0) == 0; STOP_HERE();
    if (posout) { ok &= tile(MYO_K_TENOUT, fslot, 0) == 0; STOP_HERE(); }
    if (!posvel) {
      ok &= tile(MYO_K_INERTIA, fslot, (forward_only ? 0 : 1) | inertiaStage(b->D)) == 0; STOP_HERE();
      ok &= collide(fslot); STOP_HERE();
    }
    ok &= tile(MYO_K_VELACT, fslot, (posvel ? 1 : 0) | (velactStage(b->D))) == 0; STOP_HERE();
    if (posvel) break;
    ok &= tile(MYO_K_MAKECON, fslot, 0) == 0; STOP_HERE();
    if (conout) { ok &= tile(MYO_K_CONOUT, fslot, b->count_aware ? 1 : 0) == 0; STOP_HERE(); }
    ok &= tile(MYO_K_NEWTON, fslot, (forward_only ? 0 : 1) | (b->D.nv << 8)) == 0; STOP_HERE();
    ok &= tile(MYO_K_FINISH, fslot, (forward_only ? 0 : MYO_FIN_EULER) | (last && (!forward_only || do_reset) ? MYO_FIN_STATE : 0) |
                                        (last ? MYO_FIN_WARM | MYO_FIN_ACC | (b->count_aware ? MYO_FIN_LIVE : 0) : 0)) == 0;
  }
  if (!ok) return fail(MYO_ERR_CUDA, std::string("kernel launch failed: ") + cudaGetErrorString(cudaGetLastError()));
  return MYO_OK;
}

static int launch(myo_batch* b, int nsub, int forward_only, const int32_t* mask, int do_reset, void* stream) {
  int rc = checkReady(b);
  if (rc) return rc;
  return launchPhases(b, nsub, forward_only, mask, do_reset, stream);
}
// Round 1 had a separate wide-output kernel that could be pipelined under the next step.  Every field is now written by
// the phase that produces it, so there is nothing to pipeline or join: both calls are accepted and do nothing.
int myo_batch_set_count_aware(myo_batch* b, int on) {
  if (!b) return fail(MYO_ERR_ARG, "null batch");
  b->count_aware = on != 0;
  return MYO_OK;
}
int myo_batch_set_pipelined(myo_batch* b, int on) { (void)on; return b ? MYO_OK : fail(MYO_ERR_ARG, "null batch"); }
int myo_outputs_join(myo_batch* b, void* stream) { (void)stream; return b ? MYO_OK : fail(MYO_ERR_ARG, "null batch"); }
int myo_forward(myo_batch* b, void* stream) { return launch(b, 0, 1, nullptr, 0, stream); }
int myo_step(myo_batch* b, int nsubsteps, void* stream) {
  if (nsubsteps < 1) return fail(MYO_ERR_ARG, "nsubsteps must be >= 1");
  return launch(b, nsubsteps, 0, nullptr, 0, stream);
}
int myo_reset(myo_batch* b, const int32_t* world_mask_dev, void* stream) { return launch(b, 0, 1, world_mask_dev, 1, stream); }
int myo_forward_masked(myo_batch* b, const int32_t* world_mask_dev, void* stream) { return launch(b, 0, 1, world_mask_dev, 0, stream); }
int myo_forward_posvel(myo_batch* b, const int32_t* world_mask_dev, void* stream) { return launch(b, 0, 2, world_mask_dev, 0, stream); }

int myo_step_host(myo_batch* b, int nsubsteps, const float* qpos, const float* qvel, const float* act, const float* ctrl,
                  float* qpos_out, float* qvel_out, float* act_out, void* stream) {
  int rc = checkReady(b);
  if (rc) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  const size_t n = b->nworld;
  const DevModel& D = b->D;
  bool ok = true;
  if (qpos) ok &= cudaMemcpyAsync(b->F.p[MYO_F_QPOS], qpos, n * D.nq * 4, cudaMemcpyHostToDevice, s) == cudaSuccess;
  if (qvel) ok &= cudaMemcpyAsync(b->F.p[MYO_F_QVEL], qvel, n * D.nv * 4, cudaMemcpyHostToDevice, s) == cudaSuccess;
  if (act && D.na) ok &= cudaMemcpyAsync(b->F.p[MYO_F_ACT], act, n * D.na * 4, cudaMemcpyHostToDevice, s) == cudaSuccess;
  if (ctrl && D.nu) ok &= cudaMemcpyAsync(b->F.p[MYO_F_CTRL], ctrl, n * D.nu * 4, cudaMemcpyHostToDevice, s) == cudaSuccess;
  if (!ok) return fail(MYO_ERR_CUDA, "host-to-device copy failed");
  if (nsubsteps < 1) return fail(MYO_ERR_ARG, "nsubsteps must be >= 1");
  rc = launch(b, nsubsteps, 0, nullptr, 0, stream);
  if (rc) return rc;
  if (qpos_out) ok &= cudaMemcpyAsync(qpos_out, b->F.p[MYO_F_QPOS], n * D.nq * 4, cudaMemcpyDeviceToHost, s) == cudaSuccess;
  if (qvel_out) ok &= cudaMemcpyAsync(qvel_out, b->F.p[MYO_F_QVEL], n * D.nv * 4, cudaMemcpyDeviceToHost, s) == cudaSuccess;
  if (act_out && D.na) ok &= cudaMemcpyAsync(act_out, b->F.p[MYO_F_ACT], n * D.na * 4, cudaMemcpyDeviceToHost, s) == cudaSuccess;
  if (!ok || cudaStreamSynchronize(s) != cudaSuccess) return fail(MYO_ERR_CUDA, std::string("device-to-host copy failed: ") + cudaGetErrorString(cudaGetLastError()));
  return MYO_OK;
}

// developer hooks (not in include/myostep.h): the scratch layout (SmemLayout as ints) and a copy of the scratch to the host
int myo_debug_layout(const myo_batch* b, int* out, int n) {
  const int m = (int)(sizeof(SmemLayout) / sizeof(int));
  for (int i = 0; i < n && i < m; i++) out[i] = ((const int*)&b->L)[i];
  return m;
}
int64_t myo_debug_scratch(myo_batch* b, float* host_out) {
  const int64_t n = (int64_t)((b->nworld + 31) / 32) * b->L.total * 32;
  if (host_out && (cudaDeviceSynchronize() != cudaSuccess || cudaMemcpy(host_out, b->scratch, n * sizeof(float), cudaMemcpyDeviceToHost) != cudaSuccess)) return -1;
  return n;
}
int64_t myo_debug_timing(myo_batch* b, long long* host_out) {
  const int64_t n = (int64_t)((b->nworld + 31) / 32) * 32;
  if (!b->X.dbg) return 0;
  if (host_out && (cudaDeviceSynchronize() != cudaSuccess || cudaMemcpy(host_out, b->X.dbg, n * sizeof(long long), cudaMemcpyDeviceToHost) != cudaSuccess)) return -1;
  return n;
}
// developer: per-kernel [start, end] globaltimer marks since the last call (32 kernels x 2); the call re-arms them
int myo_debug_timeline(myo_batch* b, unsigned long long* host_out) {
  if (!b->X.dbg) return 0;
  unsigned long long* p = (unsigned long long*)b->X.dbg + b->X.tl_off;
  if (cudaDeviceSynchronize() != cudaSuccess) return -1;
  if (host_out && cudaMemcpy(host_out, p, 64 * sizeof(long long), cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
  unsigned long long arm[64];
  for (int i = 0; i < 64; i++) arm[i] = ((i & 1) || i >= 48) ? 0ull : ~0ull;
  return cudaMemcpy(p, arm, sizeof(arm), cudaMemcpyHostToDevice) == cudaSuccess ? 64 : -1;
}
int myo_kat_quat(int n, const float* qa, const float* qb, float* prod, float* mat, float* rotated, void* stream) {
  if (n <= 0 || !qa || !qb || !prod || !mat || !rotated) return fail(MYO_ERR_ARG, "bad argument");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return fail(MYO_ERR_CUDA, "no CUDA device: libmyostep has no CPU fallback");
  g_launches++;
  return myo_launch_quat_kat(n, qa, qb, prod, mat, rotated, stream) == 0 ? MYO_OK : fail(MYO_ERR_CUDA, "kernel launch failed");
}

}  // extern "C"
