// The reference's own unit tests (third_party/csm/test/
// correlative_scan_matcher_tests.cc:12-106) and one match, written against the
// drop-in header nautilus_b200/cpp/CorrelativeScanMatcher.h exactly as a
// reference caller would write them.  Built and run by tests/test_gpu_shim.py.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "CorrelativeScanMatcher.h"

using Eigen::Vector2f;
using std::vector;

static int failures = 0;
#define EXPECT(cond)                                                   \
  do {                                                                 \
    if (!(cond)) { std::printf("FAIL %s:%d %s\n", __FILE__, __LINE__, #cond); ++failures; } \
  } while (0)

int main(int argc, char** argv) {
  {  // BasicLookupTest, :12-20
    LookupTable table(4, 1);
    table.SetPointValue(Vector2f(0, 0), 0.5);
    EXPECT(table.GetPointValue(Vector2f(0, 0)) == 0.5);
    EXPECT(table.GetPointValue(Vector2f(.5, .5)) == 0.5);
    EXPECT(table.GetPointValue(Vector2f(-.5, -.5)) < 0.5);
  }
  {  // EdgeOfBoundsTest, :35-43
    LookupTable table(4, 1);
    table.SetPointValue(Vector2f(-4, -4), 0.23);
    EXPECT(table.GetPointValue(Vector2f(-4, -3.99)) == 0.23);
    table.SetPointValue(Vector2f(4, -4), 0.24);
    EXPECT(table.GetPointValue(Vector2f(3.99, -4)) < 0.24);
  }
  {  // GetLowResTableTest, :56-65
    LookupTable table(4, 0.5);
    table.SetPointValue(Vector2f(0, 0.1), 0.5);
    table.SetPointValue(Vector2f(0, 0.7), 0.56);
    CorrelativeScanMatcher matcher(4, 4, 1, 0.5);
    LookupTable low = matcher.GetLookupTableLowRes(table);
    EXPECT(low.GetPointValue(Vector2f(0, 0)) == 0.56);
  }
  {  // LowResTableInexpressibleTest, :81-106 (second half)
    LookupTable table(0.3, 0.03);
    table.SetPointValue(Vector2f(0, 0), 0.5);
    table.SetPointValue(Vector2f(0, -0.031), 0.56);
    table.SetPointValue(Vector2f(-0.061, -0.031), 0.56);
    table.SetPointValue(Vector2f(-0.031, -0.031), 0.75);
    CorrelativeScanMatcher matcher(0.3, 0.3, 0.3, 0.03);
    LookupTable low = matcher.GetLookupTableLowRes(table);
    EXPECT(low.width == 2 && low.height == 2);
    EXPECT(low.GetPointValue(Vector2f(0, 0)) == 0.5);
    EXPECT(low.GetPointValue(Vector2f(-0.03, -0.03)) == 0.75);
    EXPECT(low.GetPointValue(Vector2f(-0.03, 0.03)) == 1e-10);
    EXPECT(low.GetPointValue(Vector2f(0.0, -0.03)) == 0.56);
  }
  {  // MaxAreaTest / OutOfBoundsTest, :48-62
    LookupTable table(4, 1);
    table.SetPointValue(Vector2f(-3.5, -3), 0.2);
    table.SetPointValue(Vector2f(-3, 3), 0.1);
    table.SetPointValue(Vector2f(2, -4), 0.4);
    table.SetPointValue(Vector2f(2, 2), 0.5);
    table.SetPointValue(Vector2f(4, 0), 0.6);
    EXPECT(table.MaxArea(-4, -3, 3, 1) == 0.2);
    LookupTable t2(4, 1);
    t2.SetPointValue(Vector2f(-4, -5), 0.5);
    EXPECT(t2.MaxArea(-4, -4, 4, 4) < 0.5);
  }
  // one match on clouds read from argv[1] (binary: int32 na, int32 nb, floats)
  if (argc > 1) {
    FILE* f = std::fopen(argv[1], "rb");
    int n[2];
    if (!f || std::fread(n, sizeof(int), 2, f) != 2) return 2;
    vector<Vector2f> a(n[0]), b(n[1]);
    if (std::fread(a.data(), sizeof(Vector2f), a.size(), f) != a.size()) return 2;
    if (std::fread(b.data(), sizeof(Vector2f), b.size(), f) != b.size()) return 2;
    std::fclose(f);
    CorrelativeScanMatcher matcher(30, 2, 0.3, 0.03);
    auto r = matcher.GetTransformation(a, b, -0.5, 0.5);
    std::printf("MATCH %.17g %.9g %.9g %.9g\n", r.first, r.second.first.x(), r.second.first.y(),
                r.second.second);
    auto r5 = matcher.GetTransformation(a, b, 0.1, 0.3, M_PI / 2);
    std::printf("MATCH5 %.17g %.9g %.9g %.9g\n", r5.first, r5.second.first.x(),
                r5.second.first.y(), r5.second.second);
    LookupTable hi = matcher.GetLookupTableHighRes(b);
    double mx = 0;
    for (int i = 0; i < (int)(hi.width * hi.height); ++i) { const double v = hi.values.data()[i]; mx = v > mx ? v : mx; }
    std::printf("TABLEMAX %.17g %llu\n", mx, (unsigned long long)hi.width);
    Eigen::Matrix2f u2 = matcher.GetUncertaintyMatrix(a, b, 0.125);
    std::printf("UNC2 %.9g %.9g %.9g %.9g\n", u2(0, 0), u2(0, 1), u2(1, 0), u2(1, 1));
    std::vector<vector<Vector2f>> clouds(1, b);
    std::pair<double, double> st = matcher.GetLocalUncertaintyStats(clouds, a);
    std::printf("STATS %.9g %.9g\n", st.first, st.second);
    // GetLookupTable by hand, the way cc:21-30 spells it: SetPointValue, GaussianBlur(),
    // normalize() on a LookupTable -- must equal GetLookupTableHighRes of the same cloud
    LookupTable byhand(30, 0.03);
    for (const Vector2f& p : b) byhand.SetPointValue(p, 1);
    byhand.GaussianBlur();
    byhand.normalize();
    size_t diff = 0;
    for (size_t i = 0; i < (size_t)(hi.width * hi.height); ++i)
      diff += byhand.values.data()[i] != hi.values.data()[i];
    // hi comes from the cropped batch kernels (<= 1e-5 relative), byhand is the dense recursion
    double worst = 0;
    for (size_t i = 0; i < (size_t)(hi.width * hi.height); ++i) {
      const double r = byhand.values.data()[i], g = hi.values.data()[i];
      const double e = std::fabs(r - g) / (std::fabs(r) > 1e-10 ? std::fabs(r) : 1e-10);
      worst = e > worst ? e : worst;
    }
    std::printf("BYHAND %zu %.3g %.17g\n", diff, worst, byhand.values.max());
    // first use of one matcher from several threads at once (bag_avg_matcher.cc:104-111)
    CorrelativeScanMatcher shared(30, 2, 0.3, 0.03);
    double scores[8];
#pragma omp parallel for
    for (int t = 0; t < 8; ++t) scores[t] = shared.GetTransformation(a, b, -0.5, 0.5).first;
    int same = 1;
    for (int t = 0; t < 8; ++t) same &= scores[t] == r.first;
    std::printf("OMPFIRST %d\n", same);
  }
  std::printf(failures ? "SHIM FAILED %d\n" : "SHIM OK %d\n", failures);
  return failures ? 1 : 0;
}
