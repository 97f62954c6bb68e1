// Internal declarations shared by plan.cu / kernels.cu / api.cu (not part of the public C ABI).
#ifndef MYO_INTERNAL_H_
#define MYO_INTERNAL_H_
#include <string>
#include <vector>

#include "devmodel.h"

void myo_build_plan(const myo_model_desc& m, int nconmax_req, int njmax_req, DevModel& D, SmemLayout& L,
                    std::vector<void*>& owned);

void myo_make_layout(DevModel& D, int nconmax_req, int njmax_req, SmemLayout& L);

// kernel parameters live in constant-memory slots, one per live batch
int myo_param_slots();
int myo_upload_params(int slot, const DevModel& D, const SmemLayout& L, const FieldPtrs& F, void* stream);
int myo_pick_warps(const SmemLayout& L);
// launches the fused step kernel: nsub substeps (forward_only: mj_forward) for nworld worlds, `warps` worlds per CTA
int myo_launch_step(int slot, const SmemLayout& L, int nworld, int nsub, int forward_only, const int* reset_mask, int do_reset,
                    int warps, void* stream);
int myo_kernel_smem_bytes(const SmemLayout& L);
int myo_launch_quat_kat(int n, const float* qa, const float* qb, float* prod, float* mat, float* rotated, void* stream);

#endif
