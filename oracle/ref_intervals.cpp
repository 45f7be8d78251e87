// ORACLE — TEST INFRASTRUCTURE ONLY.
// Thin extern "C" wrapper around the REFERENCE's own Intervals class, compiled from the
// sources where they lie under /root/reference (never copied into this repo) into
// oracle/_ref/libref_intervals.so.  Used to pin oracle::Intervals.
#include <cstdlib>
#include "src/continuous-validation/intervals.hh"
using hpp::core::continuousValidation::Intervals;
using hpp::core::interval_t;
extern "C" int ref_intervals_union(const double* lo, const double* hi, size_t n, double* out_lo, double* out_hi, int cap) {
  Intervals I;
  for (size_t i = 0; i < n; ++i) I.unionInterval(interval_t(lo[i], hi[i]));
  int k = 0;
  for (auto& s : I.list()) { if (k < cap) { out_lo[k] = s.first; out_hi[k] = s.second; } ++k; }
  return k;
}
extern "C" int ref_intervals_contains(const double* lo, const double* hi, size_t n, double a, double b) {
  Intervals I;
  for (size_t i = 0; i < n; ++i) I.unionInterval(interval_t(lo[i], hi[i]));
  return I.contains(interval_t(a, b)) ? 1 : 0;
}
