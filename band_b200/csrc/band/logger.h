// band/logger.h -- same macro surface as the reference's logger
// (/root/reference/band/logger.h:43-98): singleton, severity filter, reporter
// callbacks, last message; BAND_LOG(severity, fmt, ...).
#ifndef BAND_LOGGER_H_
#define BAND_LOGGER_H_

#include <cstdarg>
#include <functional>
#include <map>
#include <mutex>
#include <string>

#include "absl/status/status.h"
#include "band/common.h"

namespace band {
class Logger {
 public:
  static Logger& Get();
  void SetVerbosity(LogSeverity severity);
  CallbackId SetReporter(std::function<void(LogSeverity, const char*)> reporter);
  absl::Status RemoveReporter(CallbackId callback_id);
  std::pair<LogSeverity, std::string> GetLastLog() const;
  void Log(LogSeverity severity, const char* format, ...);

 private:
  Logger() = default;
  Logger(const Logger&) = delete;
  Logger& operator=(const Logger&) = delete;
  mutable std::mutex mu_;
  CallbackId next_callback_id_ = 0;
  std::map<CallbackId, std::function<void(LogSeverity, const char*)>> reporters_;
  LogSeverity verbosity_ = LogSeverity::kInfo;
  std::pair<LogSeverity, std::string> last_message_;
};
}  // namespace band

#define BAND_LOG(severity, format, ...) band::Logger::Get().Log(severity, format, ##__VA_ARGS__);

#endif  // BAND_LOGGER_H_
