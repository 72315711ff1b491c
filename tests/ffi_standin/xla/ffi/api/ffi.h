// Compile-only stand-in for the PUBLIC C++ API surface of xla/ffi/api/ffi.h that ffi/spb_jax_ffi.cc uses, so the shim can
// be syntax- and type-checked on a box without JAX (tests/test_ffi_source.py: g++ -fsyntax-only).  Declarations only: no
// XLA struct layout, no call-frame decoding - nothing here can run, and a real build uses jax.ffi.include_dir() instead.
#pragma once
#include <cstddef>
#include <cstdint>
#include <string>
#include <utility>

namespace xla::ffi {

enum DataType { F32, BF16, U8, S32, S64 };
enum class ErrorCode { kInvalidArgument, kInternal };

template <typename T>
struct Span {
  const T* ptr;
  size_t n;
  size_t size() const { return n; }
  const T& operator[](size_t i) const { return ptr[i]; }
};

class Error {
 public:
  Error() = default;
  Error(ErrorCode, std::string) {}
  static Error Success() { return Error(); }
};

template <typename E>
struct Unexpected {
  explicit Unexpected(E e) : e(std::move(e)) {}
  E e;
};

template <typename T>
class ErrorOr {
 public:
  ErrorOr(T v) : v_(std::move(v)), ok_(true) {}
  ErrorOr(Unexpected<Error> u) : e_(std::move(u.e)), ok_(false) {}
  bool has_value() const { return ok_; }
  bool has_error() const { return !ok_; }
  const T& value() const { return v_; }
  const Error& error() const { return e_; }

 private:
  T v_{};
  Error e_;
  bool ok_;
};

class AnyBuffer {
 public:
  void* untyped_data() const;
  Span<int64_t> dimensions() const;
  DataType element_type() const;
  size_t size_bytes() const;
};

template <DataType dt> struct NativeOf { using type = void; };
template <> struct NativeOf<F32> { using type = float; };
template <> struct NativeOf<U8> { using type = uint8_t; };
template <> struct NativeOf<S32> { using type = int32_t; };

template <DataType dt>
class Buffer {
 public:
  typename NativeOf<dt>::type* typed_data() const;
  void* untyped_data() const;
  Span<int64_t> dimensions() const;
  size_t size_bytes() const;
};

template <typename T>
class Result {
 public:
  T* operator->() const;
  T& operator*() const;
};

template <typename T> struct PlatformStream {};

struct Binding {
  template <typename T> Binding& Ctx() { return *this; }
  template <typename T> Binding& Arg() { return *this; }
  template <typename T> Binding& Ret() { return *this; }
  template <typename T> Binding& Attr(const char*) { return *this; }
};
struct Ffi {
  static Binding Bind() { return Binding(); }
};

}  // namespace xla::ffi

#define XLA_FFI_DEFINE_HANDLER_SYMBOL(sym, impl, binding) \
  extern "C" void* sym() {                                \
    (void)(binding);                                      \
    return (void*)&impl;                                  \
  }
