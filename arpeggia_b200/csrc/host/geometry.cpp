// Plane fit and plane angles on the host (feeds the ring / side-chain plane SoA and the sc_* columns),
// plus the cosine-domain boundaries that let the device reproduce acos-based comparisons exactly.
#include <cmath>
#include <cstring>
#include <limits>
#include <mutex>
#include <vector>

#include "common.hpp"

namespace arp {

namespace {
const double kPi = 3.14159265358979323846;
const double kHalfPi = 1.57079632679489661923;
inline double deg(double rad) { return rad * (180.0 / kPi); }  // f64::to_degrees
inline double dot(const double* a, const double* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
inline double len(const double* a) { return std::sqrt(dot(a, a)); }
inline double fold_deg(double c) {  // residues.rs:47-53 / 66-74
  double rad = std::acos(c);
  if (rad > kHalfPi) rad = kPi - rad;
  return deg(rad);
}
inline double plain_deg(double c) { return deg(std::acos(c)); }
}  // namespace

// residues.rs:270-298.  Centroid: column mean accumulated as sum(p * 1/n).  Normal: left singular vector of
// the centred 3 x n matrix with the smallest singular value, by one-sided Jacobi rotations of the three rows
// in the fixed cyclic order (x,y) (x,z) (y,z) until no pair needs rotating.
bool fit_plane(const double* pts, int64_t n, Plane& out) {
  if (n < 3) return false;
  const double w = 1.0 / (double)n;
  double c[3] = {0.0, 0.0, 0.0};
  for (int64_t i = 0; i < n; ++i)
    for (int k = 0; k < 3; ++k) c[k] = c[k] + w * pts[3 * i + k];
  // m[k*n + i]; residues have a dozen atoms at most: no heap traffic for them (this runs once or twice per residue)
  double m_small[3 * 32];
  std::vector<double> m_big;
  double* m = m_small;
  if (n > 32) {
    m_big.resize((size_t)(3 * n));
    m = m_big.data();
  }
  for (int k = 0; k < 3; ++k)
    for (int64_t i = 0; i < n; ++i) m[(size_t)(k * n + i)] = pts[3 * i + k] - c[k];
  double u[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
  const int pp[3] = {0, 0, 1}, qq[3] = {1, 2, 2};
  for (int sweep = 0; sweep < 60; ++sweep) {
    bool any = false;
    for (int s = 0; s < 3; ++s) {
      double* rp = &m[(size_t)(pp[s] * n)];
      double* rq = &m[(size_t)(qq[s] * n)];
      double a = 0.0, b = 0.0, g = 0.0;
      for (int64_t i = 0; i < n; ++i) {
        a = a + rp[i] * rp[i];
        b = b + rq[i] * rq[i];
        g = g + rp[i] * rq[i];
      }
      if (g == 0.0 || std::fabs(g) <= 1e-300 + 2.220446049250313e-16 * std::sqrt(a * b)) continue;
      any = true;
      const double zeta = (b - a) / (2.0 * g);
      const double t = (zeta >= 0.0 ? 1.0 : -1.0) / (std::fabs(zeta) + std::sqrt(1.0 + zeta * zeta));
      const double cs = 1.0 / std::sqrt(1.0 + t * t);
      const double sn = cs * t;
      for (int64_t i = 0; i < n; ++i) {
        const double vp = rp[i], vq = rq[i];
        rp[i] = cs * vp - sn * vq;
        rq[i] = sn * vp + cs * vq;
      }
      for (int k = 0; k < 3; ++k) {
        const double vp = u[k][pp[s]], vq = u[k][qq[s]];
        u[k][pp[s]] = cs * vp - sn * vq;
        u[k][qq[s]] = sn * vp + cs * vq;
      }
    }
    if (!any) break;
  }
  double e[3];
  for (int k = 0; k < 3; ++k) {
    e[k] = 0.0;
    for (int64_t i = 0; i < n; ++i) e[k] = e[k] + m[(size_t)(k * n + i)] * m[(size_t)(k * n + i)];
  }
  int lo = 0;
  if (e[1] < e[lo]) lo = 1;
  if (e[2] < e[lo]) lo = 2;
  for (int k = 0; k < 3; ++k) {
    out.c[k] = c[k];
    out.n[k] = u[k][lo];
  }
  return true;
}

double plane_point_dist(const Plane& p, const double pt[3]) {
  const double v[3] = {pt[0] - p.c[0], pt[1] - p.c[1], pt[2] - p.c[2]};
  return len(v);
}
double plane_point_angle(const Plane& p, const double pt[3]) {
  const double v[3] = {pt[0] - p.c[0], pt[1] - p.c[1], pt[2] - p.c[2]};
  return fold_deg(dot(p.n, v) / (len(p.n) * len(v)));
}
double plane_dihedral(const double n1[3], const double n2[3]) { return fold_deg(dot(n1, n2) / (len(n1) * len(n2))); }

// ---- boundary search ------------------------------------------------------------------------------
namespace {
// doubles <-> integers that order the same way
inline int64_t ord(double d) {
  int64_t i;
  std::memcpy(&i, &d, 8);
  return i < 0 ? (int64_t)0x8000000000000000LL - i : i;
}
inline double unord(int64_t i) {
  if (i < 0) i = (int64_t)0x8000000000000000LL - i;
  double d;
  std::memcpy(&d, &i, 8);
  return d;
}
// pred is true on a prefix [lo .. b] of the doubles in [lo, hi] and false after it: returns b
// (lo - 1ulp, i.e. "never", is returned as -inf-like sentinel when pred(lo) is false).
template <class P>
double last_true(P pred, double lo, double hi, double none) {
  int64_t a = ord(lo), b = ord(hi);
  if (!pred(lo)) return none;
  if (pred(hi)) return hi;
  while (b - a > 1) {  // invariant: pred(a) true, pred(b) false
    const int64_t mid = a + (b - a) / 2;
    if (pred(unord(mid)))
      a = mid;
    else
      b = mid;
  }
  return unord(a);
}
// pred is false on a prefix and true on the suffix [b .. hi]: returns b
template <class P>
double first_true(P pred, double lo, double hi, double none) {
  int64_t a = ord(lo), b = ord(hi);
  if (!pred(hi)) return none;
  if (pred(lo)) return lo;
  while (b - a > 1) {  // invariant: pred(a) false, pred(b) true
    const int64_t mid = a + (b - a) / 2;
    if (pred(unord(mid)))
      b = mid;
    else
      a = mid;
  }
  return unord(b);
}
}  // namespace

const AngleBounds& angle_bounds() {
  static AngleBounds B;
  static std::once_flag once;
  std::call_once(once, [] {
    const double never_le = -2.0;  // "c <= never_le" is false for every cosine
    const double never_ge = 2.0;   // "c >= never_ge" is false for every cosine
    const double m0 = -std::numeric_limits<double>::denorm_min();  // largest cosine below zero
    B.a_ge_90 = last_true([](double c) { return plain_deg(c) >= 90.0; }, -1.0, 1.0, never_le);
    B.a_ge_130 = last_true([](double c) { return plain_deg(c) >= 130.0; }, -1.0, 1.0, never_le);
    B.a_ge_60 = last_true([](double c) { return plain_deg(c) >= 60.0; }, -1.0, 1.0, never_le);
    B.a_le_120 = first_true([](double c) { return plain_deg(c) <= 120.0; }, -1.0, 1.0, never_ge);
    // folded angle: on [0,1] it decreases with c, on [-1,0) it increases with c
    B.f_le_30_pos = first_true([](double c) { return fold_deg(c) <= 30.0; }, 0.0, 1.0, never_ge);
    B.f_le_30_neg = last_true([](double c) { return fold_deg(c) <= 30.0; }, -1.0, m0, never_le);
    B.f_le_60_pos = first_true([](double c) { return fold_deg(c) <= 60.0; }, 0.0, 1.0, never_ge);
    B.f_le_60_neg = last_true([](double c) { return fold_deg(c) <= 60.0; }, -1.0, m0, never_le);
    B.f_lt_30_pos = first_true([](double c) { return fold_deg(c) < 30.0; }, 0.0, 1.0, never_ge);
    B.f_lt_30_neg = last_true([](double c) { return fold_deg(c) < 30.0; }, -1.0, m0, never_le);
    B.f_lt_60_pos = first_true([](double c) { return fold_deg(c) < 60.0; }, 0.0, 1.0, never_ge);
    B.f_lt_60_neg = last_true([](double c) { return fold_deg(c) < 60.0; }, -1.0, m0, never_le);
  });
  return B;
}

}  // namespace arp

extern "C" {
int arph_plane_fit(const double* xyz, int64_t n_points, double center[3], double normal[3]) {
  arp::Plane p;
  if (!arp::fit_plane(xyz, n_points, p)) {
    arp::last_error() = "plane fit needs at least 3 points";
    return ARP_ERR_ARG;
  }
  for (int k = 0; k < 3; ++k) {
    center[k] = p.c[k];
    normal[k] = p.n[k];
  }
  return ARP_OK;
}
double arph_plane_point_angle(const double center[3], const double normal[3], const double point[3]) {
  arp::Plane p;
  for (int k = 0; k < 3; ++k) {
    p.c[k] = center[k];
    p.n[k] = normal[k];
  }
  return arp::plane_point_angle(p, point);
}
double arph_plane_dihedral(const double n1[3], const double n2[3]) { return arp::plane_dihedral(n1, n2); }
}
