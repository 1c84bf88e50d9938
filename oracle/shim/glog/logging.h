// TEST INFRASTRUCTURE ONLY -- stand-in for glog (absent from this image) so the
// reference sources compile unchanged.  CHECK* aborts like glog's would;
// streamed messages are swallowed.
#ifndef NAUTILUS_ORACLE_SHIM_GLOG_
#define NAUTILUS_ORACLE_SHIM_GLOG_
#include <cstdio>
#include <cstdlib>
#include <iostream>
namespace nautilus_oracle_shim {
struct NullStream {
  template <typename T>
  NullStream& operator<<(const T&) { return *this; }
  NullStream& operator<<(std::ostream& (*)(std::ostream&)) { return *this; }
};
// Streams are swallowed; a failed check aborts when the temporary dies.
struct CheckSink {
  const char* expr;
  const char* file;
  int line;
  bool failed;
  CheckSink(bool ok, const char* e, const char* f, int l)
      : expr(e), file(f), line(l), failed(!ok) {}
  template <typename T>
  CheckSink& operator<<(const T&) { return *this; }
  CheckSink& operator<<(std::ostream& (*)(std::ostream&)) { return *this; }
  ~CheckSink() {
    if (failed) {
      std::fprintf(stderr, "CHECK failed: %s at %s:%d\n", expr, file, line);
      std::abort();
    }
  }
};
}  // namespace nautilus_oracle_shim
#define NAUTILUS_SHIM_CHECK(cond, text) \
  ::nautilus_oracle_shim::CheckSink((cond), text, __FILE__, __LINE__)
// The explicit casts keep -Wall quiet on unsigned >= 0 style comparisons that
// glog itself compiles without complaint.
#define CHECK(c) NAUTILUS_SHIM_CHECK((c), #c)
#define CHECK_LE(a, b) NAUTILUS_SHIM_CHECK(((a) <= (b)), #a " <= " #b)
#define CHECK_LT(a, b) NAUTILUS_SHIM_CHECK(((a) < (b)), #a " < " #b)
#define CHECK_GE(a, b) NAUTILUS_SHIM_CHECK(((a) >= (b)), #a " >= " #b)
#define CHECK_GT(a, b) NAUTILUS_SHIM_CHECK(((a) > (b)), #a " > " #b)
#define CHECK_EQ(a, b) NAUTILUS_SHIM_CHECK(((a) == (b)), #a " == " #b)
#define CHECK_NE(a, b) NAUTILUS_SHIM_CHECK(((a) != (b)), #a " != " #b)
#define CHECK_NOTNULL(p) (p)
#define LOG(sev) ::nautilus_oracle_shim::NullStream()
#define VLOG(n) ::nautilus_oracle_shim::NullStream()
#endif
