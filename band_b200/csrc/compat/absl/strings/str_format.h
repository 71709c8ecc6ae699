// Minimal absl::StrFormat stand-in: printf-style formatting into std::string.
// std::string arguments are accepted (passed as c_str()) like abseil's %s.
#ifndef BAND_COMPAT_ABSL_STRINGS_STR_FORMAT_H_
#define BAND_COMPAT_ABSL_STRINGS_STR_FORMAT_H_

#include <cstdio>
#include <string>
#include <type_traits>

namespace absl {
namespace str_format_internal {
template <typename T>
inline auto Arg(const T& v) {
  if constexpr (std::is_same_v<std::decay_t<T>, std::string>) return v.c_str();
  else if constexpr (std::is_enum_v<std::decay_t<T>>) return static_cast<long long>(v);
  else return v;
}
}  // namespace str_format_internal

template <typename... Args>
std::string StrFormat(const char* fmt, const Args&... args) {
  if constexpr (sizeof...(Args) == 0) {
    return std::string(fmt);
  } else {
    int n = std::snprintf(nullptr, 0, fmt, str_format_internal::Arg(args)...);
    if (n <= 0) return std::string();
    std::string out(static_cast<size_t>(n), '\0');
    std::snprintf(out.data(), static_cast<size_t>(n) + 1, fmt, str_format_internal::Arg(args)...);
    return out;
  }
}
}  // namespace absl
#endif
