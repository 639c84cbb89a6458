// oracle/shim/glog/logging.h -- TEST INFRASTRUCTURE ONLY.
// Minimal stand-in for <glog/logging.h> so that a handful of reference source
// files (which reach glog only through core/pbrt.h:63) can be compiled where
// they lie under /root/reference by oracle/Makefile, without running the
// reference's cmake build (glog's real header is cmake-generated).
// Provides only the macros the reference uses: CHECK*, DCHECK*, LOG(sev), VLOG(n).
#ifndef BRE_ORACLE_SHIM_GLOG_LOGGING_H
#define BRE_ORACLE_SHIM_GLOG_LOGGING_H
#include <cstdlib>
#include <iostream>
#include <sstream>
#include <string>

namespace bre_shim {
struct NullStream {
    template <typename T> NullStream &operator<<(const T &) { return *this; }
    NullStream &operator<<(std::ostream &(*)(std::ostream &)) { return *this; }
};
struct FatalStream {
    std::ostringstream ss;
    FatalStream(const char *file, int line, const char *what) {
        ss << file << ":" << line << ": " << what << " ";
    }
    template <typename T> FatalStream &operator<<(const T &v) { ss << v; return *this; }
    FatalStream &operator<<(std::ostream &(*f)(std::ostream &)) { ss << f; return *this; }
    [[noreturn]] ~FatalStream() { std::cerr << ss.str() << std::endl; std::abort(); }
};
struct ErrStream {
    std::ostringstream ss;
    template <typename T> ErrStream &operator<<(const T &v) { ss << v; return *this; }
    ErrStream &operator<<(std::ostream &(*f)(std::ostream &)) { ss << f; return *this; }
    ~ErrStream() { std::cerr << ss.str() << std::endl; }
};
// Lets "cond ? (void)0 : Voidify() & stream" type-check.
struct Voidify { template <typename T> void operator&(const T &) {} };
}  // namespace bre_shim

#define BRE_SHIM_CHECK(cond, text) \
    (cond) ? (void)0 : ::bre_shim::Voidify() & ::bre_shim::FatalStream(__FILE__, __LINE__, "Check failed: " text)

#define CHECK(c) BRE_SHIM_CHECK((c), #c)
#define CHECK_EQ(a, b) BRE_SHIM_CHECK((a) == (b), #a " == " #b)
#define CHECK_NE(a, b) BRE_SHIM_CHECK((a) != (b), #a " != " #b)
#define CHECK_LT(a, b) BRE_SHIM_CHECK((a) < (b), #a " < " #b)
#define CHECK_LE(a, b) BRE_SHIM_CHECK((a) <= (b), #a " <= " #b)
#define CHECK_GT(a, b) BRE_SHIM_CHECK((a) > (b), #a " > " #b)
#define CHECK_GE(a, b) BRE_SHIM_CHECK((a) >= (b), #a " >= " #b)
namespace bre_shim {
template <typename T> T CheckNotNull(const char *file, int line, const char *what, T &&t) {
    if (t == nullptr) { FatalStream(file, line, what); }
    return std::forward<T>(t);
}
}
#define CHECK_NOTNULL(p) ::bre_shim::CheckNotNull(__FILE__, __LINE__, "'" #p "' must be non NULL", (p))
#define DCHECK(c) CHECK(c)
#define DCHECK_EQ(a, b) CHECK_EQ(a, b)
#define DCHECK_NE(a, b) CHECK_NE(a, b)
#define DCHECK_LT(a, b) CHECK_LT(a, b)
#define DCHECK_LE(a, b) CHECK_LE(a, b)
#define DCHECK_GT(a, b) CHECK_GT(a, b)
#define DCHECK_GE(a, b) CHECK_GE(a, b)

#define BRE_SHIM_LOG_INFO ::bre_shim::NullStream()
#define BRE_SHIM_LOG_WARNING ::bre_shim::NullStream()
#define BRE_SHIM_LOG_ERROR ::bre_shim::ErrStream()
#define BRE_SHIM_LOG_FATAL ::bre_shim::FatalStream(__FILE__, __LINE__, "FATAL")
#define LOG(sev) BRE_SHIM_LOG_##sev
#define VLOG(n) ::bre_shim::NullStream()

// main/pbrt.cpp (the reference's command line, built unmodified into oracle/_ref/pbrt) sets glog's flags
namespace google { inline void InitGoogleLogging(const char *) {} }
static int FLAGS_stderrthreshold = 1, FLAGS_minloglevel = 0, FLAGS_v = 0;
static bool FLAGS_logtostderr = false;
static std::string FLAGS_log_dir;

#endif
