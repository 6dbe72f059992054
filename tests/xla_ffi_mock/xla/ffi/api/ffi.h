// TEST INFRASTRUCTURE — a minimal stand-in for jaxlib's xla/ffi/api/ffi.h.
//
// The image has neither jax nor jaxlib (and no network), so the real XLA FFI headers are absent.
// This mock declares just the part of that API which parallax_b200/csrc/cotix_xla_ffi.cc uses —
// Buffer / ResultBuffer / Span / Error / Ffi::Bind() / XLA_FFI_DEFINE_HANDLER_SYMBOL — with the same
// names and member functions, so that the shim is COMPILED by the test-suite and its handler bodies
// are EXECUTED on the GPU (tests/test_xla_ffi_shim.py drives them through tests/xla_ffi_mock/driver.cc).
// What stays unverified here is only the real binding layer (argument decoding by XLA).
#pragma once
#include <cstddef>
#include <cstdint>
#include <string>
#include <utility>
#include <vector>

namespace xla {
namespace ffi {

enum class DataType : uint8_t { U8, U32, F32 };
inline constexpr DataType U8 = DataType::U8;
inline constexpr DataType U32 = DataType::U32;
inline constexpr DataType F32 = DataType::F32;

template <DataType>
struct NativeOf;
template <>
struct NativeOf<DataType::U8> { using type = uint8_t; };
template <>
struct NativeOf<DataType::U32> { using type = uint32_t; };
template <>
struct NativeOf<DataType::F32> { using type = float; };

template <typename T>
class Span {
 public:
  Span() : p_(nullptr), n_(0) {}
  Span(T* p, size_t n) : p_(p), n_(n) {}
  T* begin() const { return p_; }
  T* end() const { return p_ + n_; }
  size_t size() const { return n_; }
  T& operator[](size_t i) const { return p_[i]; }

 private:
  T* p_;
  size_t n_;
};

template <DataType dtype>
class Buffer {
 public:
  using T = typename NativeOf<dtype>::type;
  Buffer() : data_(nullptr) {}
  Buffer(void* data, std::vector<int64_t> dims) : data_(data), dims_(std::move(dims)) {}
  T* typed_data() const { return static_cast<T*>(data_); }
  void* untyped_data() const { return data_; }
  Span<const int64_t> dimensions() const { return Span<const int64_t>(dims_.data(), dims_.size()); }
  size_t element_count() const {
    size_t n = 1;
    for (int64_t d : dims_) n *= static_cast<size_t>(d);
    return n;
  }
  size_t size_bytes() const { return element_count() * sizeof(T); }

 private:
  void* data_;
  std::vector<int64_t> dims_;
};

template <typename T>
class Result {
 public:
  Result() = default;
  explicit Result(T v) : v_(std::move(v)) {}
  T* operator->() { return &v_; }
  T& operator*() { return v_; }

 private:
  T v_;
};
template <DataType dtype>
using ResultBuffer = Result<Buffer<dtype>>;

class Error {
 public:
  static Error Success() { return Error(true, ""); }
  static Error Internal(std::string m) { return Error(false, std::move(m)); }
  static Error InvalidArgument(std::string m) { return Error(false, std::move(m)); }
  bool success() const { return ok_; }
  const std::string& message() const { return msg_; }

 private:
  Error(bool ok, std::string m) : ok_(ok), msg_(std::move(m)) {}
  bool ok_;
  std::string msg_;
};

template <typename T>
struct PlatformStream {};
enum class Traits { kCmdBufferCompatible };

struct Binding {
  template <typename T>
  Binding Ctx() { return *this; }
  template <typename T>
  Binding Arg() { return *this; }
  template <typename T>
  Binding Ret() { return *this; }
  template <typename T>
  Binding Attr(const char*) { return *this; }
};
struct Ffi {
  static Binding Bind() { return Binding(); }
};

}  // namespace ffi
}  // namespace xla

// the real macro defines `extern "C" XLA_FFI_Error* name(XLA_FFI_CallFrame*)`; the mock exports a data
// symbol of the same name holding the address of the implementation, so that `nm` sees the handler
#define XLA_FFI_DEFINE_HANDLER_SYMBOL(name, impl, binding, ...) \
  extern "C" const void* const name = reinterpret_cast<const void*>(&impl)
