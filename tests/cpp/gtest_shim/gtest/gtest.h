// Minimal stand-in for googletest (absent from this image): just enough to compile
// and run the reference's third_party/csm/test/correlative_scan_matcher_tests.cc
// UNCHANGED against the drop-in header.  TEST registers itself; main() runs all.
#ifndef NAUTILUS_GTEST_SHIM_H_
#define NAUTILUS_GTEST_SHIM_H_
#include <cmath>
#include <cstdio>
#include <cstring>
#include <vector>

namespace testing {
struct TestInfo { const char* suite; const char* name; void (*fn)(bool*); };
inline std::vector<TestInfo>& Registry() { static std::vector<TestInfo> r; return r; }
struct Registrar { Registrar(const char* s, const char* n, void (*f)(bool*)) { Registry().push_back({s, n, f}); } };
// gtest's ASSERT_DOUBLE_EQ: equal within 4 ULPs
inline bool AlmostEqual(double a, double b) {
  if (std::isnan(a) || std::isnan(b)) return false;
  long long ia, ib;
  std::memcpy(&ia, &a, 8); std::memcpy(&ib, &b, 8);
  if (ia < 0) ia = (long long)0x8000000000000000ULL - ia;
  if (ib < 0) ib = (long long)0x8000000000000000ULL - ib;
  const long long d = ia > ib ? ia - ib : ib - ia;
  return d <= 4;
}
inline int RunAll() {
  int failed = 0;
  for (const TestInfo& t : Registry()) {
    bool ok = true;
    t.fn(&ok);
    std::printf("[ %s ] %s.%s\n", ok ? "    OK" : "FAILED", t.suite, t.name);
    failed += !ok;
  }
  std::printf("[==========] %zu tests ran, %d failed\n", Registry().size(), failed);
  return failed ? 1 : 0;
}
inline void InitGoogleTest(int*, char**) {}
}  // namespace testing

#define TEST(suite, name)                                                              \
  static void suite##_##name##_body(bool* gtest_ok_);                                   \
  static ::testing::Registrar suite##_##name##_reg(#suite, #name, suite##_##name##_body); \
  static void suite##_##name##_body(bool* gtest_ok_)
#define GTEST_FAIL_(what)                                                     \
  do {                                                                        \
    std::printf("%s:%d: Failure: %s\n", __FILE__, __LINE__, what);            \
    *gtest_ok_ = false;                                                       \
    return;                                                                   \
  } while (0)
#define ASSERT_DOUBLE_EQ(a, b) do { if (!::testing::AlmostEqual((double)(a), (double)(b))) GTEST_FAIL_("ASSERT_DOUBLE_EQ(" #a ", " #b ")"); } while (0)
#define ASSERT_EQ(a, b) do { if (!((a) == (b))) GTEST_FAIL_("ASSERT_EQ(" #a ", " #b ")"); } while (0)
#define ASSERT_LT(a, b) do { if (!((a) < (b))) GTEST_FAIL_("ASSERT_LT(" #a ", " #b ")"); } while (0)
#define ASSERT_TRUE(a) do { if (!(a)) GTEST_FAIL_("ASSERT_TRUE(" #a ")"); } while (0)
#define RUN_ALL_TESTS() ::testing::RunAll()
#endif  // NAUTILUS_GTEST_SHIM_H_
