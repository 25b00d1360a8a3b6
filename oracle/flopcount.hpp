// flopcount.hpp -- TEST INFRASTRUCTURE ONLY.  A float wrapper that counts every
// arithmetic operation; brax_step_oracle.c compiled as C++ with -DBXO_COUNT
// uses it as its scalar type `R`, so the float32 oracle itself reports the
// exact number of operations one System.step performs (SURVEY.md 8(d):
// "replace these estimates with an exact count").  Counters are per process;
// run single-threaded.
#ifndef BXO_FLOPCOUNT_HPP_
#define BXO_FLOPCOUNT_HPP_
#include <math.h>
enum { CNT_ADD, CNT_MUL, CNT_DIV, CNT_SQRT, CNT_TRIG, CNT_CMP, CNT_MINMAX, CNT_N };
extern "C" double bxo_counters[CNT_N];
struct CR {
  float v;
  CR() = default;
  CR(float x) : v(x) {}
  CR(double x) : v((float)x) {}
  CR(int x) : v((float)x) {}
  explicit operator float() const { return v; }
  explicit operator double() const { return v; }
};
#define BXO_BIN(op, cnt) \
  static inline CR operator op(CR a, CR b) { bxo_counters[cnt] += 1; return CR(a.v op b.v); }
BXO_BIN(+, CNT_ADD) BXO_BIN(-, CNT_ADD) BXO_BIN(*, CNT_MUL) BXO_BIN(/, CNT_DIV)
static inline CR operator-(CR a) { return CR(-a.v); }
static inline CR& operator+=(CR& a, CR b) { a = a + b; return a; }
static inline CR& operator-=(CR& a, CR b) { a = a - b; return a; }
static inline CR& operator*=(CR& a, CR b) { a = a * b; return a; }
static inline CR& operator/=(CR& a, CR b) { a = a / b; return a; }
#define BXO_CMP(op) \
  static inline bool operator op(CR a, CR b) { bxo_counters[CNT_CMP] += 1; return a.v op b.v; }
BXO_CMP(<) BXO_CMP(>) BXO_CMP(<=) BXO_CMP(>=) BXO_CMP(==) BXO_CMP(!=)
static inline CR cr_sqrt(CR a) { bxo_counters[CNT_SQRT] += 1; return CR(sqrtf(a.v)); }
static inline CR cr_sin(CR a) { bxo_counters[CNT_TRIG] += 1; return CR(sinf(a.v)); }
static inline CR cr_cos(CR a) { bxo_counters[CNT_TRIG] += 1; return CR(cosf(a.v)); }
static inline CR cr_acos(CR a) { bxo_counters[CNT_TRIG] += 1; return CR(acosf(a.v)); }
static inline CR cr_atan2(CR a, CR b) { bxo_counters[CNT_TRIG] += 1; return CR(atan2f(a.v, b.v)); }
static inline CR cr_fabs(CR a) { return CR(fabsf(a.v)); }
#endif
