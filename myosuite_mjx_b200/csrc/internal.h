// Internal declarations shared by plan.cu / kernels.cu / api.cu (not part of the public C ABI).
#ifndef MYO_INTERNAL_H_
#define MYO_INTERNAL_H_
#include <string>
#include <vector>

#include "devmodel.h"

void myo_build_plan(const myo_model_desc& m, int nconmax_req, int njmax_req, DevModel& D, SmemLayout& L,
                    std::vector<void*>& owned);

void myo_make_layout(DevModel& D, int nconmax_req, int njmax_req, SmemLayout& L);

// launches the fused step kernel: nsub substeps (0 = forward only) for nworld worlds
int myo_launch_step(const DevModel& D, const SmemLayout& L, const FieldPtrs& F, int nworld, int nsub, int forward_only,
                    const int* reset_mask, int do_reset, void* stream);
int myo_kernel_smem_bytes(const SmemLayout& L);

#endif
