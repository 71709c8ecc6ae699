// Minimal absl::Status stand-in so the reference's band/buffer signatures
// (absl::Status / absl::StatusOr returns, band/buffer/operator.h:31-78) compile
// unchanged. abseil itself is an un-vendored Bazel dependency of the reference
// (band/workspace_repo.bzl:24-58) and is absent from this image. Only the
// subset the path uses is provided; codes keep abseil's numeric values.
#ifndef BAND_COMPAT_ABSL_STATUS_STATUS_H_
#define BAND_COMPAT_ABSL_STATUS_STATUS_H_

#include <ostream>
#include <string>
#include <utility>

namespace absl {

enum class StatusCode : int {
  kOk = 0,
  kCancelled = 1,
  kUnknown = 2,
  kInvalidArgument = 3,
  kDeadlineExceeded = 4,
  kNotFound = 5,
  kAlreadyExists = 6,
  kPermissionDenied = 7,
  kResourceExhausted = 8,
  kFailedPrecondition = 9,
  kAborted = 10,
  kOutOfRange = 11,
  kUnimplemented = 12,
  kInternal = 13,
  kUnavailable = 14,
  kDataLoss = 15,
  kUnauthenticated = 16,
};

inline const char* StatusCodeToString(StatusCode c) {
  switch (c) {
    case StatusCode::kOk: return "OK";
    case StatusCode::kCancelled: return "CANCELLED";
    case StatusCode::kUnknown: return "UNKNOWN";
    case StatusCode::kInvalidArgument: return "INVALID_ARGUMENT";
    case StatusCode::kNotFound: return "NOT_FOUND";
    case StatusCode::kUnimplemented: return "UNIMPLEMENTED";
    case StatusCode::kInternal: return "INTERNAL";
    case StatusCode::kUnavailable: return "UNAVAILABLE";
    case StatusCode::kFailedPrecondition: return "FAILED_PRECONDITION";
    default: return "ERROR";
  }
}

class Status {
 public:
  Status() : code_(StatusCode::kOk) {}
  Status(StatusCode code, std::string msg) : code_(code), msg_(code == StatusCode::kOk ? "" : std::move(msg)) {}
  bool ok() const { return code_ == StatusCode::kOk; }
  StatusCode code() const { return code_; }
  const std::string& message() const { return msg_; }
  std::string ToString() const {
    return ok() ? "OK" : std::string(StatusCodeToString(code_)) + ": " + msg_;
  }
  friend bool operator==(const Status& a, const Status& b) { return a.code_ == b.code_ && a.msg_ == b.msg_; }
  friend bool operator!=(const Status& a, const Status& b) { return !(a == b); }

 private:
  StatusCode code_;
  std::string msg_;
};

inline std::ostream& operator<<(std::ostream& os, const Status& s) { return os << s.ToString(); }

inline Status OkStatus() { return Status(); }
inline Status InvalidArgumentError(std::string m) { return Status(StatusCode::kInvalidArgument, std::move(m)); }
inline Status InternalError(std::string m) { return Status(StatusCode::kInternal, std::move(m)); }
inline Status UnknownError(std::string m) { return Status(StatusCode::kUnknown, std::move(m)); }
inline Status UnimplementedError(std::string m) { return Status(StatusCode::kUnimplemented, std::move(m)); }
inline Status NotFoundError(std::string m) { return Status(StatusCode::kNotFound, std::move(m)); }
inline Status UnavailableError(std::string m) { return Status(StatusCode::kUnavailable, std::move(m)); }
inline Status FailedPreconditionError(std::string m) { return Status(StatusCode::kFailedPrecondition, std::move(m)); }

}  // namespace absl
#endif
