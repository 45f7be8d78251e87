// Interval lists of the continuous validations (device side).
// Replaces the reference's Intervals (src/continuous-validation/intervals.hh:48-183) and the
// boost::icl::interval_set<closed> of BodyPairCollision (src/continuous-validation/body-pair-collision.cc:56-96).
#pragma once
#include "dev_model.cuh"

namespace hppb200 {

// Sorted disjoint closed-interval list, element j at v[j * stride]: same observable behaviour as the reference
// Intervals::unionInterval and as interval_set<closed>::insert (touching intervals merge).  False on overflow.
__device__ __forceinline__ bool wdUnion(double2* v, size_t stride, int& cnt, int cap, double a, double b) {
  int i = 0;
  while (i < cnt && v[i * stride].y < a) ++i;
  if (i == cnt) {
    if (cnt == cap) return false;
    v[i * stride] = make_double2(a, b);
    ++cnt;
    return true;
  }
  if (b < v[i * stride].x) {
    if (cnt == cap) return false;
    for (int k = cnt; k > i; --k) v[k * stride] = v[(k - 1) * stride];
    v[i * stride] = make_double2(a, b);
    ++cnt;
    return true;
  }
  double nlo = v[i * stride].x;
  if (a < nlo) nlo = a;
  double nhi = b;
  int j = i;
  while (j < cnt && v[j * stride].x <= b) {
    const double h = v[j * stride].y;
    if (h > nhi) nhi = h;
    ++j;
  }
  v[i * stride] = make_double2(nlo, nhi);
  const int removed = j - i - 1;
  if (removed > 0) {
    for (int k = i + 1; k + removed < cnt; ++k) v[k * stride] = v[(k + removed) * stride];
    cnt -= removed;
  }
  return true;
}
__device__ __forceinline__ bool wdContains(const double2* v, size_t stride, int cnt, double a, double b) {
  for (int i = 0; i < cnt; ++i) {
    const double2 x = v[i * stride];
    if (x.x <= a && b <= x.y) return true;
  }
  return false;
}

}  // namespace hppb200
