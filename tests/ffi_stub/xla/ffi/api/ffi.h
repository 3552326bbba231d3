// Minimal stand-in for jaxlib's xla/ffi/api/ffi.h -- ONLY so that csrc/csi_xla_ffi.cc can be syntax- and type-checked
// on machines without jaxlib (tests/test_abi.py).  It models the names the shim uses (Buffer / ResultBuffer with
// typed_data, element_count, dimensions; Error; PlatformStream; the Bind() chain; the handler macro) and checks the
// handler's argument list against the binding; it is not an implementation of the XLA FFI.
#pragma once
#include <cstddef>
#include <cstdint>
#include <string>
#include <tuple>
#include <type_traits>
#include <vector>

namespace xla::ffi {
enum DataType { F32, F64 };
template <DataType> struct NativeType;
template <> struct NativeType<F32> { using type = float; };
template <> struct NativeType<F64> { using type = double; };
template <DataType D> struct Buffer {
  using T = typename NativeType<D>::type;
  T *data = nullptr;
  std::vector<int64_t> dims;
  T *typed_data() const { return data; }
  size_t element_count() const { size_t n = 1; for (auto d : dims) n *= (size_t)d; return n; }
  const std::vector<int64_t> &dimensions() const { return dims; }
};
template <DataType D> struct ResultBuffer {
  Buffer<D> *p = nullptr;
  Buffer<D> *operator->() const { return p; }
};
enum class ErrorCode { kInternal, kInvalidArgument };
struct Error {
  bool ok = true; std::string msg;
  Error() = default;
  Error(ErrorCode, const std::string &m) : ok(false), msg(m) {}
  static Error Success() { return Error(); }
};
template <class T> struct PlatformStream { using type = T; };
template <class... Ts> struct Binding {
  template <class C> Binding<Ts..., typename C::type> Ctx() const { return {}; }
  template <class T> Binding<Ts..., T> Attr(const char *) const { return {}; }
  template <class T> Binding<Ts..., T> Arg() const { return {}; }
  template <class B> auto Ret() const { return RetImpl(static_cast<B *>(nullptr)); }
  template <DataType D> Binding<Ts..., ResultBuffer<D>> RetImpl(Buffer<D> *) const { return {}; }
  template <class F> static constexpr bool Matches() { return std::is_invocable_r_v<Error, F, Ts...>; }
};
struct Ffi { static Binding<> Bind() { return {}; } };
}  // namespace xla::ffi

#define XLA_FFI_DEFINE_HANDLER_SYMBOL(name, impl, binding)                                                \
  static_assert(decltype(binding)::template Matches<decltype(&impl)>(), #impl " does not match its binding"); \
  extern "C" void *name(void *frame) { (void)frame; return nullptr; }
