// Internal declarations shared by plan.cu / kernels.cu / api.cu (not part of the public C ABI).
#ifndef MYO_INTERNAL_H_
#define MYO_INTERNAL_H_
#include <string>
#include <vector>

#include "devmodel.h"

void myo_build_plan(const myo_model_desc& m, int nconmax_req, int njmax_req, DevModel& D, SmemLayout& L,
                    std::vector<void*>& owned);

void myo_make_layout(DevModel& D, int nconmax_req, int njmax_req, SmemLayout& L);

// lane = world kernel (tpw.cu): parameters live in constant-memory slots, one per live batch
int myo_tpw_param_slots();
size_t myo_tpw_scratch_floats(const SmemLayout& L, int nworld);   // floats of global scratch a batch needs
int myo_tpw_upload_params(int slot, const DevModel& D, const SmemLayout& L, const FieldPtrs& F, void* stream);
// nsub substeps (forward_only: mj_forward; do_reset: mj_resetData + mj_forward on the masked worlds) for nworld worlds
int myo_tpw_launch_step(int slot, const SmemLayout& L, int nworld, int nsub, int forward_only, const int* reset_mask, int do_reset,
                        float* scratch, void* stream);
int myo_tpw_launch_quat_kat(int n, const float* qa, const float* qb, float* prod, float* mat, float* rotated, void* stream);
int64_t myo_tpw_launch_counter();

#endif
