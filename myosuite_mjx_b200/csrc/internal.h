// Internal declarations shared by plan.cu / tile.cu / api.cu / env.cu (not part of the public C ABI).
#ifndef MYO_INTERNAL_H_
#define MYO_INTERNAL_H_
#include <string>
#include <vector>

#include "devmodel.h"

void myo_build_plan(const myo_model_desc& m, int nconmax_req, int njmax_req, DevModel& D, SmemLayout& L,
                    std::vector<void*>& owned);

void myo_make_layout(DevModel& D, int nconmax_req, int njmax_req, SmemLayout& L);

#define MYO_PARAM_SLOTS 16   // live batches per process: one constant-memory parameter slot each (17 x ~2.3 KB of the 64 KB)
// The kernels of the pipeline (tile.cu, lane = world): one entry per phase
enum { MYO_K_KIN = 0, MYO_K_TENDON, MYO_K_INERTIA, MYO_K_COLLIDE, MYO_K_VELACT, MYO_K_FINISH, MYO_K_CONOUT, MYO_K_NEWTON, MYO_K_COLLIDE_MPR, MYO_K_COLLIDE_COMPACT, MYO_K_MAKECON, MYO_K_POSOUT, MYO_K_TENOUT, MYO_K_COMPACT_MAKECON };
enum { MYO_KIN_LOAD = 1, MYO_KIN_RESET = 2, MYO_KIN_CHECK = 4 };                 // flags of MYO_K_KIN
enum { MYO_FIN_EULER = 1, MYO_FIN_STATE = 2, MYO_FIN_WARM = 4, MYO_FIN_ACC = 8, MYO_FIN_LIVE = 16 };  // flags of MYO_K_FINISH
int myo_tile_upload(int slot, const DevModel& D, const SmemLayout& L, const FieldPtrs& F, const TileAux& X, void* stream);
int myo_nop_launch(void* stream);
int myo_tile_launch(int kernel, int slot, int fslot, int nworld, const int* mask, int flags, float* scratch, void* stream);
struct myo_model {
  DevModel D;
  SmemLayout L;
  std::vector<void*> owned;
  int device;
};
struct myo_batch {
  const myo_model* model;
  DevModel D;   // copy with this batch's nconmax / njmax
  SmemLayout L;
  FieldPtrs F;
  int nworld;
  int slot;      // constant-memory parameter slot
  int device = 0;   // the model's device, kept here: a batch may be destroyed after its model (garbage collectors give no order)
  bool dirty;    // parameters changed since the last upload
  bool conout;   // some contact / constraint record field is bound: conoutKernel runs
  bool count_aware = false;   // myo_batch_set_count_aware: the padding of contact_* / efc_* past ncon / nefc is not written
  void* wrap_cleared[2] = {nullptr, nullptr};   // wrap_obj / wrap_xpos buffers already zeroed once (api.cu)
  float* scratch; // per-world intermediates, tile-blocked: [tiles][L.total][32]
  TileAux X;      // work list of the convex narrow phase
  // the phases that depend only on the kinematics (tendon, inertia, collide) and the two tails of a step (conout, finish) run
  // concurrently: two library-owned side streams fork from and join the caller's stream with events (graph-capturable)
  void* side[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};   // collision chain (high priority), tendon -> velact, inertia, records out (low priority)
  void* ev[10] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  std::vector<void*> owned;
};

// helpers shared by api.cu and env.cu
int myo_fail(int code, const std::string& msg);
void myo_count_launch(int n);
int myo_launch_quat_kat(int n, const float* qa, const float* qb, float* prod, float* mat, float* rotated, void* stream);

#endif
