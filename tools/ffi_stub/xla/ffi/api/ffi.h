// COMPILE-CHECK STAND-IN for jaxlib's xla/ffi/api/ffi.h -- NOT the real header, never shipped, never linked.
//
// jaxlib (and with it the XLA FFI headers) cannot be installed in this image, so csrc/xla_ffi.cu would otherwise never
// see a compiler.  This file declares just the part of the public typed-FFI surface those handlers use, with the same
// names and shapes as the published API (xla::ffi::Ffi::Bind().Ctx<>().Arg<>().Ret<>().Attr<>(), Buffer<T>, AnyBuffer,
// ResultBuffer<T>, Error, ErrorCode, PlatformStream, XLA_FFI_DEFINE_HANDLER_SYMBOL), and makes the DEFINE macro
// static_assert that the implementation function is callable with the decoded argument list the binding describes.
// tools/check_ffi_compiles.sh compiles xla_ffi.cu against it (object file discarded).  It proves the handlers are
// well-formed C++ against this rendition of the API; it does not prove they register under a real jaxlib.
#pragma once
#include <cstddef>
#include <cstdint>
#include <string>
#include <type_traits>
#include <utility>
#include <vector>

namespace xla::ffi {

enum class ErrorCode { kOk, kInvalidArgument, kInternal };
enum DataType { F32 = 11, BF16 = 16 };

class Error {
 public:
  Error() = default;
  Error(ErrorCode code, std::string message) : code_(code), message_(std::move(message)) {}
  static Error Success() { return Error(); }

 private:
  ErrorCode code_ = ErrorCode::kOk;
  std::string message_;
};

struct Dims {
  std::vector<int64_t> v;
  int64_t operator[](size_t i) const { return v[i]; }
  int64_t back() const { return v.back(); }
  size_t size() const { return v.size(); }
};

class AnyBuffer {
 public:
  void* untyped_data() const { return nullptr; }
  size_t element_count() const { return 0; }
  Dims dimensions() const { return {}; }
  DataType element_type() const { return F32; }
};

template <DataType dtype>
class Buffer : public AnyBuffer {};

template <typename T>
class Result {
 public:
  T* operator->() { return &value_; }

 private:
  T value_;
};
template <DataType dtype>
using ResultBuffer = Result<Buffer<dtype>>;
using AnyResultBuffer = Result<AnyBuffer>;

template <typename T>
struct PlatformStream {};

namespace detail {
template <typename T> struct CtxDecoded;
template <typename T> struct CtxDecoded<PlatformStream<T>> { using type = T; };
template <typename T> struct RetDecoded { using type = Result<T>; };
}  // namespace detail

template <typename... Ts>
struct Binding {
  template <typename T> Binding<Ts..., typename detail::CtxDecoded<T>::type> Ctx() const { return {}; }
  template <typename T> Binding<Ts..., T> Arg() const { return {}; }
  template <typename T> Binding<Ts..., typename detail::RetDecoded<T>::type> Ret() const { return {}; }
  template <typename T> Binding<Ts..., T> Attr(const char*) const { return {}; }
  template <typename Fn>
  static constexpr bool Accepts() { return std::is_invocable_r_v<Error, Fn, Ts...>; }
};

struct Ffi {
  static Binding<> Bind() { return {}; }
};

}  // namespace xla::ffi

#define XLA_FFI_DEFINE_HANDLER_SYMBOL(symbol, impl, binding)                                                         \
  static_assert(decltype(binding)::Accepts<decltype(&impl)>(), #impl " does not take the arguments of its binding"); \
  extern "C" void* symbol(void* call_frame) { return call_frame ? nullptr : nullptr; }
