// XLA FFI custom-call targets for jax.ffi ("compiled only where the header exists").
//
// Replaces the call `mjx.step(model, data)` inside a jitted program (reference: myosuite/mjx/play.py:47, brax
// pipeline_step behind myosuite/mjx/myodm_v0.py:287): the operands are the state leaves of the vmapped mjx.Data with
// the world axis minor-most (layout [k][nworld], see include/myostep.h), the results alias them
// (input_output_aliases), so XLA's buffers are bound zero-copy and stepped in place on XLA's stream.
//
// This image has no jaxlib / XLA headers (`xla/ffi/api/ffi.h` ships inside the jaxlib wheel: jax.ffi.include_dir()),
// so this translation unit is not part of the default build; `make xla_ffi.o XLA_INC=$(python -c
// 'import jax.ffi; print(jax.ffi.include_dir())')` compiles it where jax exists and the Makefile then links it
// into libmyostep.so.  Nothing else in the library depends on it.
#if defined(__has_include)
#if __has_include("xla/ffi/api/ffi.h")
#define MYO_HAVE_XLA_FFI 1
#endif
#endif

#ifdef MYO_HAVE_XLA_FFI
#include <cuda_runtime_api.h>

#include <cstdint>

#include "../../include/myostep.h"
#include "xla/ffi/api/ffi.h"

namespace ffi = xla::ffi;

namespace {
template <class B> int bindBuf(myo_batch* b, int field, B& buf) {
  return myo_batch_bind(b, field, buf.untyped_data(), (int64_t)buf.size_bytes());
}

// operands: qpos qvel act ctrl time qacc_warmstart; results (aliased): qpos qvel act time qacc_warmstart
ffi::Error StepImpl(cudaStream_t stream, int64_t batch, int32_t nsub, ffi::Buffer<ffi::F32> qpos, ffi::Buffer<ffi::F32> qvel,
                    ffi::Buffer<ffi::F32> act, ffi::Buffer<ffi::F32> ctrl, ffi::Buffer<ffi::F32> time, ffi::Buffer<ffi::F32> warm,
                    ffi::ResultBuffer<ffi::F32> qpos_o, ffi::ResultBuffer<ffi::F32> qvel_o, ffi::ResultBuffer<ffi::F32> act_o,
                    ffi::ResultBuffer<ffi::F32> time_o, ffi::ResultBuffer<ffi::F32> warm_o) {
  myo_batch* b = reinterpret_cast<myo_batch*>(batch);
  // Without aliasing XLA hands out distinct result buffers: copy the state across first, then step the results in place.
  struct { ffi::Buffer<ffi::F32>* in; ffi::ResultBuffer<ffi::F32>* out; int field; } io[] = {
      {&qpos, &qpos_o, MYO_F_QPOS}, {&qvel, &qvel_o, MYO_F_QVEL}, {&act, &act_o, MYO_F_ACT}, {&time, &time_o, MYO_F_TIME},
      {&warm, &warm_o, MYO_F_QACC_WARMSTART}};
  for (auto& x : io) {
    void* dst = (*x.out)->untyped_data();
    if (dst != x.in->untyped_data() && x.in->size_bytes() &&
        cudaMemcpyAsync(dst, x.in->untyped_data(), x.in->size_bytes(), cudaMemcpyDeviceToDevice, stream) != cudaSuccess)
      return ffi::Error::Internal("myo_step: state copy failed");
    if (x.in->size_bytes() && bindBuf(b, x.field, **x.out) != MYO_OK) return ffi::Error::Internal(myo_last_error());
  }
  if (ctrl.size_bytes() && bindBuf(b, MYO_F_CTRL, ctrl) != MYO_OK) return ffi::Error::Internal(myo_last_error());
  if (myo_step(b, nsub, stream) != MYO_OK) return ffi::Error::Internal(myo_last_error());
  return ffi::Error::Success();
}

// mj_forward on the bound state (reference: mjx.forward after a reset, myosuite/mjx/myodm_v0.py:159 pipeline_init)
ffi::Error ForwardImpl(cudaStream_t stream, int64_t batch, ffi::Buffer<ffi::F32> qpos, ffi::Buffer<ffi::F32> qvel,
                       ffi::Buffer<ffi::F32> act, ffi::Buffer<ffi::F32> ctrl, ffi::Buffer<ffi::F32> time, ffi::Buffer<ffi::F32> warm,
                       ffi::ResultBuffer<ffi::F32> warm_o) {
  myo_batch* b = reinterpret_cast<myo_batch*>(batch);
  if (warm_o->untyped_data() != warm.untyped_data() && warm.size_bytes() &&
      cudaMemcpyAsync(warm_o->untyped_data(), warm.untyped_data(), warm.size_bytes(), cudaMemcpyDeviceToDevice, stream) != cudaSuccess)
    return ffi::Error::Internal("myo_forward: state copy failed");
  if (bindBuf(b, MYO_F_QPOS, qpos) || bindBuf(b, MYO_F_QVEL, qvel) || (act.size_bytes() && bindBuf(b, MYO_F_ACT, act)) ||
      (ctrl.size_bytes() && bindBuf(b, MYO_F_CTRL, ctrl)) || bindBuf(b, MYO_F_TIME, time) ||
      bindBuf(b, MYO_F_QACC_WARMSTART, *warm_o))
    return ffi::Error::Internal(myo_last_error());
  if (myo_forward(b, stream) != MYO_OK) return ffi::Error::Internal(myo_last_error());
  return ffi::Error::Success();
}
}  // namespace

XLA_FFI_DEFINE_HANDLER_SYMBOL(MyoStep, StepImpl,
                              ffi::Ffi::Bind()
                                  .Ctx<ffi::PlatformStream<cudaStream_t>>()
                                  .Attr<int64_t>("batch")
                                  .Attr<int32_t>("nsub")
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>());

XLA_FFI_DEFINE_HANDLER_SYMBOL(MyoForward, ForwardImpl,
                              ffi::Ffi::Bind()
                                  .Ctx<ffi::PlatformStream<cudaStream_t>>()
                                  .Attr<int64_t>("batch")
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>());
#else
// no XLA headers on this machine: the translation unit is empty and libmyostep.so exports no MyoStep / MyoForward
// (myosuite_mjx_b200/lib.py reports that through has_xla_ffi()).
#endif
