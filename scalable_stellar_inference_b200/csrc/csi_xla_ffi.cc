// csi_xla_ffi.cc -- XLA typed-FFI handlers over the C ABI (include/csi.h): the custom calls a JAX host registers so
// that numpyro's NUTS drives the CUDA path unchanged (BASELINE north_star; plug points of SURVEY 8b).
//
//   csi_potential    : z f32[L]                 -> out6 f64[6], grad f32[L]   (value AND gradient: jax.custom_vjp.fwd
//                                                                              keeps grad as the residual, bwd scales it)
//   csi_net_forward  : x f32[n, n_in]           -> y f32[n, n_out]            (Emulator / BC emulator, attr which)
//   csi_net_vjp      : x f32[n, n_in], gy       -> gx f32[n, n_in]
//   csi_grid_forward : x f32[n, 4]              -> y f32[n, bands]            (BC table interpolation)
//   csi_grid_vjp     : x f32[n, 4], gy          -> gx f32[n, 4]
//
// The handle travels as an int64 attribute (the integer value of the csi_handle* returned by csi_create in the same
// process); the stream is XLA's.  Every call is stream-ordered, allocation-free and never synchronises the host, which
// is what an XLA custom call must be.  Built only where jaxlib's headers exist:
//
//     make -C scalable_stellar_inference_b200/csrc xla_ffi      # needs `python -c "import jax; jax.ffi.include_dir()"`
//
// Python side: jax.ffi.register_ffi_target("csi_potential",
//     jax.ffi.pycapsule(ctypes.CDLL("libcsi_xla_ffi.so").CsiPotential), platform="CUDA").
#include <cuda_runtime.h>

#include <cstdint>

#include "../../include/csi.h"
#include "xla/ffi/api/ffi.h"

namespace ffi = xla::ffi;

static ffi::Error status(int rc) {
  return rc == CSI_OK ? ffi::Error::Success() : ffi::Error(ffi::ErrorCode::kInternal, csi_last_error());
}
static csi_handle *handle_of(int64_t h) { return reinterpret_cast<csi_handle *>(static_cast<intptr_t>(h)); }

static ffi::Error PotentialImpl(cudaStream_t stream, int64_t handle, ffi::Buffer<ffi::F32> z,
                                ffi::ResultBuffer<ffi::F64> out6, ffi::ResultBuffer<ffi::F32> grad) {
  if (out6->element_count() < 6 || grad->element_count() != z.element_count())
    return ffi::Error(ffi::ErrorCode::kInvalidArgument, "csi_potential: out6 must hold 6 doubles, grad must match z");
  return status(csi_potential(handle_of(handle), z.typed_data(), out6->typed_data(), grad->typed_data(), stream));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(CsiPotential, PotentialImpl,
                              ffi::Ffi::Bind()
                                  .Ctx<ffi::PlatformStream<cudaStream_t>>()
                                  .Attr<int64_t>("handle")
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F64>>()
                                  .Ret<ffi::Buffer<ffi::F32>>());

static ffi::Error NetForwardImpl(cudaStream_t stream, int64_t handle, int64_t which, ffi::Buffer<ffi::F32> x,
                                 ffi::ResultBuffer<ffi::F32> y) {
  const int64_t n = x.dimensions().size() ? x.dimensions()[0] : 0;
  return status(csi_net_forward(handle_of(handle), (int32_t)which, x.typed_data(), y->typed_data(), n, stream));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(CsiNetForward, NetForwardImpl,
                              ffi::Ffi::Bind()
                                  .Ctx<ffi::PlatformStream<cudaStream_t>>()
                                  .Attr<int64_t>("handle")
                                  .Attr<int64_t>("which")
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>());

static ffi::Error NetVjpImpl(cudaStream_t stream, int64_t handle, int64_t which, ffi::Buffer<ffi::F32> x,
                             ffi::Buffer<ffi::F32> gy, ffi::ResultBuffer<ffi::F32> gx) {
  const int64_t n = x.dimensions().size() ? x.dimensions()[0] : 0;
  return status(csi_net_vjp(handle_of(handle), (int32_t)which, x.typed_data(), gy.typed_data(), gx->typed_data(), n, stream));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(CsiNetVjp, NetVjpImpl,
                              ffi::Ffi::Bind()
                                  .Ctx<ffi::PlatformStream<cudaStream_t>>()
                                  .Attr<int64_t>("handle")
                                  .Attr<int64_t>("which")
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>());

static ffi::Error GridForwardImpl(cudaStream_t stream, int64_t handle, ffi::Buffer<ffi::F32> x,
                                  ffi::ResultBuffer<ffi::F32> y) {
  const int64_t n = x.dimensions().size() ? x.dimensions()[0] : 0;
  return status(csi_grid_forward(handle_of(handle), x.typed_data(), y->typed_data(), n, stream));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(CsiGridForward, GridForwardImpl,
                              ffi::Ffi::Bind()
                                  .Ctx<ffi::PlatformStream<cudaStream_t>>()
                                  .Attr<int64_t>("handle")
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>());

static ffi::Error GridVjpImpl(cudaStream_t stream, int64_t handle, ffi::Buffer<ffi::F32> x, ffi::Buffer<ffi::F32> gy,
                              ffi::ResultBuffer<ffi::F32> gx) {
  const int64_t n = x.dimensions().size() ? x.dimensions()[0] : 0;
  return status(csi_grid_vjp(handle_of(handle), x.typed_data(), gy.typed_data(), gx->typed_data(), n, stream));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(CsiGridVjp, GridVjpImpl,
                              ffi::Ffi::Bind()
                                  .Ctx<ffi::PlatformStream<cudaStream_t>>()
                                  .Attr<int64_t>("handle")
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>());
