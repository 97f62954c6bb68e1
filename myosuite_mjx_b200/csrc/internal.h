// Internal declarations shared by plan.cu / team.cu / tile.cu / api.cu (not part of the public C ABI).
#ifndef MYO_INTERNAL_H_
#define MYO_INTERNAL_H_
#include <string>
#include <vector>

#include "devmodel.h"

void myo_build_plan(const myo_model_desc& m, int nconmax_req, int njmax_req, DevModel& D, SmemLayout& L,
                    std::vector<void*>& owned);

void myo_make_layout(DevModel& D, int nconmax_req, int njmax_req, SmemLayout& L);

#define MYO_PARAM_SLOTS 8
// The TEAM-mode kernels (team.cu) are compiled once per team width (lanes per world); every variant has its own
// constant-memory parameter slots.
struct TeamVariant {
  int team;
  int (*upload)(int slot, const DevModel& D, const SmemLayout& L, const FieldPtrs& F, void* stream);
  // whole step in one launch: nsub substeps (forward_only: mj_forward; do_reset: mj_resetData + mj_forward on the masked worlds)
  int (*launch)(int slot, const SmemLayout& L, int nworld, int stride, int nsub, int forward_only, const int* reset_mask, int do_reset,
                float* scratch, void* stream);
  // phase pipeline: constraint rows + Newton solve (+ mj_checkAcc when stepping)
  int (*solve)(int slot, int fslot, const SmemLayout& L, int nworld, int stride, const int* mask, int check_acc, float* scratch, void* stream);
};
#define MYO_DECLARE_TEAM(T)                                                                                                    \
  int myo_team_upload_##T(int slot, const DevModel& D, const SmemLayout& L, const FieldPtrs& F, void* stream);                 \
  int myo_team_launch_##T(int slot, const SmemLayout& L, int nworld, int stride, int nsub, int forward_only, const int* reset_mask, \
                          int do_reset, float* scratch, void* stream);                                                         \
  int myo_team_solve_##T(int slot, int fslot, const SmemLayout& L, int nworld, int stride, const int* mask, int check_acc, float* scratch, void* stream);
MYO_DECLARE_TEAM(4)
MYO_DECLARE_TEAM(8)
MYO_DECLARE_TEAM(16)
// TILE-mode kernels (tile.cu, lane = world): one entry per phase of the pipeline
enum { MYO_K_KIN = 0, MYO_K_TENDON, MYO_K_INERTIA, MYO_K_COLLIDE, MYO_K_VELACT, MYO_K_FINISH, MYO_K_CONOUT, MYO_K_NEWTON };
enum { MYO_KIN_LOAD = 1, MYO_KIN_RESET = 2, MYO_KIN_CHECK = 4 };                 // flags of MYO_K_KIN
enum { MYO_FIN_EULER = 1, MYO_FIN_STATE = 2, MYO_FIN_WARM = 4, MYO_FIN_ACC = 8 };  // flags of MYO_K_FINISH
int myo_tile_upload(int slot, const DevModel& D, const SmemLayout& L, const FieldPtrs& F, void* stream);
int myo_tile_launch(int kernel, int slot, int fslot, int nworld, int stride, const int* mask, int flags, float* scratch, void* stream);
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
  int stride;    // worlds rounded up to 32: the world axis of the scratch
  int slot;      // constant-memory parameter slot
  const TeamVariant* kern;   // TEAM-mode variant (lanes per world) chosen for this batch size
  bool mono;     // whole step in one launch (round-1 kernel, MYO_ENGINE=mono) instead of the phase pipeline
  bool dirty;    // parameters changed since the last upload
  bool wide;     // mono engine: some wide output field is bound, the output kernel runs after the step kernel
  // myo_step_host: the state is final when the step ends, so its device-to-host copies run on a side stream
  void* copy_stream = nullptr; void* ev_step = nullptr; void* ev_copy = nullptr;
  void* wrap_cleared[2] = {nullptr, nullptr};   // wrap_obj / wrap_xpos buffers already zeroed once (api.cu launch)
  float* scratch; // per-world intermediates, [L.total][stride]
  // Pipelined outputs (mono engine, myo_batch_set_pipelined): the wide-output kernel of step k runs on the side stream under
  // the step kernel of step k+1.  Two scratch buffers alternate; `cur` holds the last pass; ev_out[i] marks the end of the
  // last output kernel that read buffer i; `out_pending`: an output kernel may still be running.
  float* scratch_buf[2] = {nullptr, nullptr};
  void* ev_out[2] = {nullptr, nullptr};
  bool ev_out_set[2] = {false, false};
  int cur = 0;
  bool pipelined = false, out_pending = false;
  std::vector<void*> owned;
};

// helpers shared by api.cu and env.cu
int myo_fail(int code, const std::string& msg);
void myo_count_launch(int n);
// out.cu (mono engine): wide mjData outputs, one thread per (world, item), written after the step kernel from its scratch
extern const int kWideOutFields[];
extern const int kNumWideOutFields;
int myo_out_upload(int slot, const DevModel& D, const SmemLayout& L, const FieldPtrs& F, void* stream);
int myo_out_launch(int slot, int nworld, int stride, const int* mask, int posvel, const float* scratch, void* stream);
int myo_launch_quat_kat(int n, const float* qa, const float* qb, float* prod, float* mat, float* rotated, void* stream);

#endif
