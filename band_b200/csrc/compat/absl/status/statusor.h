// Minimal absl::StatusOr stand-in (see status.h in this directory).
#ifndef BAND_COMPAT_ABSL_STATUS_STATUSOR_H_
#define BAND_COMPAT_ABSL_STATUS_STATUSOR_H_

#include <cstdlib>
#include <optional>
#include <utility>

#include "absl/status/status.h"

namespace absl {

template <typename T>
class StatusOr {
 public:
  StatusOr() : status_(StatusCode::kUnknown, "") {}
  StatusOr(const Status& s) : status_(s) {
    if (status_.ok()) status_ = InternalError("StatusOr constructed from OK status without a value");
  }
  StatusOr(Status&& s) : status_(std::move(s)) {
    if (status_.ok()) status_ = InternalError("StatusOr constructed from OK status without a value");
  }
  StatusOr(const T& v) : value_(v) {}
  StatusOr(T&& v) : value_(std::move(v)) {}
  StatusOr(const StatusOr&) = default;
  StatusOr(StatusOr&&) = default;
  StatusOr& operator=(const StatusOr&) = default;
  StatusOr& operator=(StatusOr&&) = default;

  bool ok() const { return status_.ok(); }
  const Status& status() const { return status_; }
  T& value() & { Check(); return *value_; }
  const T& value() const& { Check(); return *value_; }
  T&& value() && { Check(); return std::move(*value_); }
  T& operator*() & { return *value_; }
  const T& operator*() const& { return *value_; }
  T* operator->() { return &*value_; }
  const T* operator->() const { return &*value_; }

 private:
  void Check() const {
    if (!value_.has_value()) std::abort();  // -fno-exceptions build, like the reference (band/band.bzl:26-28)
  }
  Status status_;
  std::optional<T> value_;
};

}  // namespace absl
#endif
