// Host build of butterfly_b200/csrc/bf_hankel.inl for tests/test_hankel_host.py (CPU check of the device math).
#define _GNU_SOURCE 1
#include <cmath>
struct double2 { double x, y; };
static inline double2 make_double2(double x, double y) { double2 r; r.x = x; r.y = y; return r; }
using std::fma; using std::log; using std::sqrt;
#define BF_HANKEL_FN
#define BF_HANKEL_TABLE static const
#include "bf_hankel.inl"
extern "C" void bfHankelHost(const double *x, long n, double *h0, double *h1) {
  for (long i = 0; i < n; ++i) {
    const double2 a = hankel0(x[i]), b = hankel1(x[i]);
    h0[2 * i] = a.x; h0[2 * i + 1] = a.y; h1[2 * i] = b.x; h1[2 * i + 1] = b.y;
  }
}
