// Internal declarations shared by plan.cu / team.cu / api.cu (not part of the public C ABI).
#ifndef MYO_INTERNAL_H_
#define MYO_INTERNAL_H_
#include <string>
#include <vector>

#include "devmodel.h"

void myo_build_plan(const myo_model_desc& m, int nconmax_req, int njmax_req, DevModel& D, SmemLayout& L,
                    std::vector<void*>& owned);

void myo_make_layout(DevModel& D, int nconmax_req, int njmax_req, SmemLayout& L);

#define MYO_PARAM_SLOTS 8
// The step kernel (team.cu) is compiled once per team width (lanes per world); every variant has its own
// constant-memory parameter slots.
struct TeamVariant {
  int team;
  int ctas_per_sm;   // co-resident 14-warp CTAs per SM this build allows (1: <= 128 registers per thread, 2: <= 72)
  int (*upload)(int slot, const DevModel& D, const SmemLayout& L, const FieldPtrs& F, void* stream);
  // nsub substeps (forward_only: mj_forward; do_reset: mj_resetData + mj_forward on the masked worlds)
  int (*launch)(int slot, const SmemLayout& L, int nworld, int nsub, int forward_only, const int* reset_mask, int do_reset,
                float* scratch, void* stream);
  size_t (*scratch_floats)(const SmemLayout& L, int nworld);
};
#define MYO_DECLARE_TEAM(T)                                                                                              \
  int myo_team_upload_##T(int slot, const DevModel& D, const SmemLayout& L, const FieldPtrs& F, void* stream);           \
  int myo_team_launch_##T(int slot, const SmemLayout& L, int nworld, int nsub, int forward_only, const int* reset_mask, \
                          int do_reset, float* scratch, void* stream);                                                   \
  size_t myo_team_scratch_##T(const SmemLayout& L, int nworld);
MYO_DECLARE_TEAM(1)
MYO_DECLARE_TEAM(2)
MYO_DECLARE_TEAM(4)
MYO_DECLARE_TEAM(8)
MYO_DECLARE_TEAM(16)
MYO_DECLARE_TEAM(8x2)
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
  const TeamVariant* kern;   // step kernel variant (lanes per world) chosen for this batch size
  // the build that runs launches of three or more fused substeps: two co-resident CTAs drift into different phases over
  // the substeps and lose the shared instruction stream (16384 worlds x 10 substeps: 27.1 ms for one CTA per SM in two
  // waves, 31.4 ms for two co-resident CTAs; one substep: 2.06 against 1.82 ms, profiles/r1zz_fused_variants.txt)
  const TeamVariant* kern_fused = nullptr;
  bool dirty;    // parameters changed since the last upload
  bool wide;     // some wide output field is bound: the output kernel runs after the step kernel
  // myo_step_host: the state is final when the step kernel ends, so its device-to-host copies run on a side stream
  // under the wide-output kernel (created on first use)
  void* copy_stream = nullptr; void* ev_step = nullptr; void* ev_copy = nullptr;
  void* wrap_cleared[2] = {nullptr, nullptr};   // wrap_obj / wrap_xpos buffers already zeroed once (api.cu launch)
  float* scratch; // per-world intermediates, [nwarps][L.total][worlds per warp]
  // Pipelined outputs (myo_batch_set_pipelined): the wide-output kernel of step k runs on the side stream under the step
  // kernel of step k+1.  Two scratch buffers alternate; `cur` holds the last pass; ev_out[i] marks the end of the last
  // output kernel that read buffer i; `out_pending`: an output kernel may still be running (joined by the next call that
  // is not itself a pipelined step, or by myo_outputs_join).
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
// out.cu: wide mjData outputs, one thread per (world, item), written after the step kernel from its scratch
extern const int kWideOutFields[];
extern const int kNumWideOutFields;
int myo_out_upload(int slot, const DevModel& D, const SmemLayout& L, const FieldPtrs& F, void* stream);
int myo_out_launch(int slot, int nworld, int stride, const int* mask, int posvel, const float* scratch, void* stream);
int myo_launch_quat_kat(int n, const float* qa, const float* qb, float* prod, float* mat, float* rotated, void* stream);

#endif
