// Minimal stand-in for <glog/logging.h>, written for this repo (NOT reference code).
// TEST INFRASTRUCTURE ONLY: lets the reference's hot-path sources under
// /root/reference/src compile with plain g++ (no cmake, no vendored glog build).
// Semantics kept: CHECK* abort with a message, LOG(FATAL) aborts, other LOG/VLOG are no-ops.
#ifndef B200RT_ORACLE_GLOG_STUB_H
#define B200RT_ORACLE_GLOG_STUB_H
#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <sstream>
#include <array>
#include <cstdint>
#include <limits>

namespace glogstub {
struct NullStream {
    template <typename T> NullStream &operator<<(const T &) { return *this; }
    NullStream &operator<<(std::ostream &(*)(std::ostream &)) { return *this; }
};
struct FatalStream {
    std::ostringstream ss;
    const char *file; int line;
    FatalStream(const char *f, int l) : file(f), line(l) {}
    template <typename T> FatalStream &operator<<(const T &v) { ss << v; return *this; }
    FatalStream &operator<<(std::ostream &(*f)(std::ostream &)) { ss << f; return *this; }
    [[noreturn]] ~FatalStream() {
        std::fprintf(stderr, "[ref FATAL] %s:%d %s\n", file, line, ss.str().c_str());
        std::abort();
    }
};
struct Voidify { void operator&(const NullStream &) {} void operator&(const FatalStream &) {} };
}  // namespace glogstub

#define GLOGSTUB_INFO    ::glogstub::NullStream()
#define GLOGSTUB_WARNING ::glogstub::NullStream()
#define GLOGSTUB_ERROR   ::glogstub::NullStream()
#define GLOGSTUB_FATAL   ::glogstub::FatalStream(__FILE__, __LINE__)
#define LOG(sev) GLOGSTUB_##sev
#define VLOG(n) ::glogstub::NullStream()
#define CHECK(c) (c) ? (void)0 : ::glogstub::Voidify() & ::glogstub::FatalStream(__FILE__, __LINE__) << "Check failed: " #c " "
#define GLOGSTUB_OP(a, b, op) CHECK((a) op (b))
#define CHECK_EQ(a, b) GLOGSTUB_OP(a, b, ==)
#define CHECK_NE(a, b) GLOGSTUB_OP(a, b, !=)
#define CHECK_LT(a, b) GLOGSTUB_OP(a, b, <)
#define CHECK_LE(a, b) GLOGSTUB_OP(a, b, <=)
#define CHECK_GT(a, b) GLOGSTUB_OP(a, b, >)
#define CHECK_GE(a, b) GLOGSTUB_OP(a, b, >=)
#define CHECK_NOTNULL(p) (p)
// The reference is built Release (-DNDEBUG): DCHECKs compile to nothing.
#define DCHECK(c) while (false) CHECK(c)
#define DCHECK_EQ(a, b) while (false) CHECK_EQ(a, b)
#define DCHECK_NE(a, b) while (false) CHECK_NE(a, b)
#define DCHECK_LT(a, b) while (false) CHECK_LT(a, b)
#define DCHECK_LE(a, b) while (false) CHECK_LE(a, b)
#define DCHECK_GT(a, b) while (false) CHECK_GT(a, b)
#define DCHECK_GE(a, b) while (false) CHECK_GE(a, b)
namespace google { inline void InitGoogleLogging(const char *) {} }
// glog's command-line flags that src/main/pbrt.cpp sets (no-ops here)
#include <string>
static int FLAGS_stderrthreshold = 0, FLAGS_minloglevel = 0, FLAGS_v = 0;
static bool FLAGS_logtostderr = false;
static std::string FLAGS_log_dir;
#endif
