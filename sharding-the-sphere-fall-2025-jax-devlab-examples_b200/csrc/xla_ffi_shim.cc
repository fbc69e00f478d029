// XLA FFI adapter: exposes the C ABI of libcubesphere_b200.so as jax.ffi custom-call targets.
//
// NOT built by default and NOT exercised by the test-suite: neither JAX nor the XLA FFI headers exist in
// this image (or on the GPU box).  Build where JAX is installed:
//     make -C csrc shim XLA_FFI_INCLUDE=$(python -c "import jax.ffi; print(jax.ffi.include_dir())")
// It replaces the traced halo closure of the reference (JAX-DevLab-Examples.py:238-246) and the step that
// the reference does not ship by one custom call each; state buffers are aliased input -> output, so the
// step runs in place on XLA's buffers.  The context handle (cs_ctx*, created once per device through
// ctypes) travels as an int64 attribute.
#include <cstdint>

#include "cubesphere.h"
#include "xla/ffi/api/ffi.h"

namespace ffi = xla::ffi;

static ffi::Error to_error(int rc, const cs_ctx* ctx) {
    if (rc == CS_OK) return ffi::Error::Success();
    const char* msg = cs_last_error(ctx);
    if (rc == CS_E_INVALID) return ffi::Error::InvalidArgument(msg ? msg : "invalid argument");
    if (rc == CS_E_UNSUPPORTED) return ffi::Error(ffi::ErrorCode::kUnimplemented, msg ? msg : "unsupported");
    return ffi::Error::Internal(msg ? msg : "CUDA failure");
}

static ffi::Error StepImpl(cudaStream_t stream, int64_t ctx_handle, double dt, ffi::AnyBuffer h, ffi::AnyBuffer u,
                           ffi::AnyBuffer v, ffi::Result<ffi::AnyBuffer> ho, ffi::Result<ffi::AnyBuffer> uo,
                           ffi::Result<ffi::AnyBuffer> vo) {
    auto* ctx = reinterpret_cast<cs_ctx*>(ctx_handle);
    // input_output_aliases={0:0, 1:1, 2:2}: the results ARE the operands
    if (ho->untyped_data() != h.untyped_data() || uo->untyped_data() != u.untyped_data() || vo->untyped_data() != v.untyped_data())
        return ffi::Error::InvalidArgument("cs_step needs input_output_aliases={0: 0, 1: 1, 2: 2}");
    return to_error(cs_step(ctx, ho->untyped_data(), uo->untyped_data(), vo->untyped_data(), dt, nullptr, stream), ctx);
}

static ffi::Error HaloFillStateImpl(cudaStream_t stream, int64_t ctx_handle, ffi::AnyBuffer h, ffi::AnyBuffer u,
                                    ffi::AnyBuffer v, ffi::Result<ffi::AnyBuffer> ho, ffi::Result<ffi::AnyBuffer> uo,
                                    ffi::Result<ffi::AnyBuffer> vo) {
    auto* ctx = reinterpret_cast<cs_ctx*>(ctx_handle);
    if (ho->untyped_data() != h.untyped_data() || uo->untyped_data() != u.untyped_data() || vo->untyped_data() != v.untyped_data())
        return ffi::Error::InvalidArgument("cs_halo_fill_state needs input_output_aliases={0: 0, 1: 1, 2: 2}");
    return to_error(cs_halo_fill_state(ctx, ho->untyped_data(), uo->untyped_data(), vo->untyped_data(), stream), ctx);
}

static ffi::Error HaloFillScalarImpl(cudaStream_t stream, int64_t ctx_handle, ffi::AnyBuffer f, ffi::Result<ffi::AnyBuffer> fo) {
    auto* ctx = reinterpret_cast<cs_ctx*>(ctx_handle);
    if (fo->untyped_data() != f.untyped_data()) return ffi::Error::InvalidArgument("cs_halo_fill_scalar needs input_output_aliases={0: 0}");
    return to_error(cs_halo_fill_scalar(ctx, fo->untyped_data(), stream), ctx);
}

static ffi::Error DiagnosticsImpl(cudaStream_t stream, int64_t ctx_handle, ffi::AnyBuffer h, ffi::AnyBuffer u, ffi::AnyBuffer v,
                                  ffi::Result<ffi::Buffer<ffi::F64>> out) {
    auto* ctx = reinterpret_cast<cs_ctx*>(ctx_handle);
    return to_error(cs_diagnostics(ctx, h.untyped_data(), u.untyped_data(), v.untyped_data(), out->typed_data(), stream), ctx);
}

#define CS_BIND3() \
    ffi::Ffi::Bind().Ctx<ffi::PlatformStream<cudaStream_t>>().Attr<int64_t>("ctx")

XLA_FFI_DEFINE_HANDLER_SYMBOL(CsStep, StepImpl,
                              CS_BIND3().Attr<double>("dt").Arg<ffi::AnyBuffer>().Arg<ffi::AnyBuffer>().Arg<ffi::AnyBuffer>()
                                  .Ret<ffi::AnyBuffer>().Ret<ffi::AnyBuffer>().Ret<ffi::AnyBuffer>());
XLA_FFI_DEFINE_HANDLER_SYMBOL(CsHaloFillState, HaloFillStateImpl,
                              CS_BIND3().Arg<ffi::AnyBuffer>().Arg<ffi::AnyBuffer>().Arg<ffi::AnyBuffer>()
                                  .Ret<ffi::AnyBuffer>().Ret<ffi::AnyBuffer>().Ret<ffi::AnyBuffer>());
XLA_FFI_DEFINE_HANDLER_SYMBOL(CsHaloFillScalar, HaloFillScalarImpl, CS_BIND3().Arg<ffi::AnyBuffer>().Ret<ffi::AnyBuffer>());
XLA_FFI_DEFINE_HANDLER_SYMBOL(CsDiagnostics, DiagnosticsImpl,
                              CS_BIND3().Arg<ffi::AnyBuffer>().Arg<ffi::AnyBuffer>().Arg<ffi::AnyBuffer>().Ret<ffi::Buffer<ffi::F64>>());
