// glog/logging.h - the handful of glog symbols the reference benchmark uses
// (allreduce_over_mpi/benchmark.cpp:15,58-64,72-114,142-215), as a header-only stand-in: the image has
// no glog and no network. Messages go to stderr; FATAL aborts.
#ifndef FT_GLOG_SHIM_H_
#define FT_GLOG_SHIM_H_
#include <cstdlib>
#include <iostream>
#include <sstream>

static bool FLAGS_colorlogtostderr = false;
static bool FLAGS_logtostderr = true;

namespace google {
inline void InitGoogleLogging(const char*) { (void)FLAGS_colorlogtostderr; (void)FLAGS_logtostderr; }
inline void InstallFailureSignalHandler() {}
inline void ShutdownGoogleLogging() {}
enum { GLOG_INFO = 0, GLOG_WARNING = 1, GLOG_ERROR = 2, GLOG_FATAL = 3 };
class LogMessage {
 public:
  LogMessage(const char* file, int line, int sev) : sev_(sev) {
    static const char tag[] = "IWEF";
    ss_ << tag[sev] << ' ' << file << ':' << line << "] ";
  }
  ~LogMessage() {
    ss_ << '\n';
    std::cerr << ss_.str();
    if (sev_ == GLOG_FATAL) std::abort();
  }
  std::ostream& stream() { return ss_; }
 private:
  int sev_;
  std::ostringstream ss_;
};
struct Voidify { void operator&(std::ostream&) {} };
}  // namespace google

#define FT_GLOG_SEV_INFO 0
#define FT_GLOG_SEV_WARNING 1
#define FT_GLOG_SEV_ERROR 2
#define FT_GLOG_SEV_FATAL 3
#define LOG(sev) google::LogMessage(__FILE__, __LINE__, FT_GLOG_SEV_##sev).stream()
#define LOG_IF(sev, cond) !(cond) ? (void)0 : google::Voidify() & LOG(sev)
#define CHECK_GE(a, b) LOG_IF(FATAL, !((a) >= (b))) << "Check failed: " #a " >= " #b " "
#define CHECK(c) LOG_IF(FATAL, !(c)) << "Check failed: " #c " "
#endif
