// pyi_axis_host.hpp -- host-side axis objects of the C-ABI: construction
// (the classification regular / irregular / periodic of the reference's
// math::Axis<T> constructors, cxx/include/pyinterp/math/axis.hpp:582-716) and
// the flat descriptor the kernels consume.  In drop-in mode the descriptor is
// read from a live reference Axis (pyi_axis_create_from_desc) and nothing here
// is re-derived.
#pragma once

#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/pyinterp_b200.h"
#include "pyi_device.cuh"

struct pyi_axis {
  pyi_axis_desc d{};
  std::vector<double> pf;   // irregular f64 points
  std::vector<int64_t> pi;  // irregular i64 points
  // uniform-bin lookup table of an irregular f64 axis (see irregular_find_f64)
  std::vector<int32_t> lut;
  double lut_min{0}, lut_inv{0};
  int lut_bins{0};
  // device mirrors, one per CUDA device
  struct Dev {
    void* points{nullptr};
    int32_t* lut{nullptr};
  };
  std::mutex mu;
  std::map<int, Dev> dev;

  pyi_axis() = default;
  pyi_axis(const pyi_axis& o)
      : d(o.d), pf(o.pf), pi(o.pi), lut(o.lut), lut_min(o.lut_min), lut_inv(o.lut_inv),
        lut_bins(o.lut_bins) {
    fix_points();
  }
  void fix_points() {
    d.points = d.kind == PYI_AXIS_IRREGULAR
                   ? (d.coord_type == PYI_COORD_F64 ? static_cast<const void*>(pf.data())
                                                    : static_cast<const void*>(pi.data()))
                   : nullptr;
  }
};

namespace pyi_host {

template <typename T>
struct Traits;
template <>
struct Traits<double> {
  static double& sc(pyi_scalar& s) { return s.f; }
  static double abs_(double v) { return std::fabs(v); }
};
template <>
struct Traits<int64_t> {
  static int64_t& sc(pyi_scalar& s) { return s.i; }
  static int64_t abs_(int64_t v) { return v < 0 ? -v : v; }
};

// Eigen's LinSpaced(n, low, high)[i] (the reference's is_evenly_spaced test
// compares against it, axis.hpp:582-597): floating version = low + i*step, or
// high - (n-1-i)*step when |high| < |low|, end points exact; integer version
// multiplies (or divides) by an integer step.
inline double linspaced(int64_t n, double low, double high, int64_t i) {
  const double step = (high - low) / static_cast<double>(n - 1);
  if (std::fabs(high) < std::fabs(low)) {
    return i == 0 ? low : high - static_cast<double>(n - 1 - i) * step;
  }
  return i == n - 1 ? high : low + static_cast<double>(i) * step;
}
inline int64_t linspaced(int64_t n, int64_t low, int64_t high, int64_t i) {
  const int64_t span = high - low;
  const int64_t aspan1 = (span < 0 ? -span : span) + 1;
  const int64_t mult = span / (n - 1);
  const int64_t div = ((high >= low ? n : -n) + span) / (aspan1 == 0 ? 1 : aspan1);
  return (aspan1 < n) ? low + i / div : low + i * mult;
}

inline bool is_same(double a, double b, double eps) {  // math.hpp:160-166
  const double diff = std::fabs(a - b);
  return diff <= eps || diff < std::fmax(std::fabs(a), std::fabs(b)) * eps;
}
inline bool is_same(int64_t a, int64_t b, int64_t eps) {  // math.hpp:154-158
  return (a - b < 0 ? b - a : a - b) <= eps;
}

template <typename T>
T coord(const pyi_axis& a, const std::vector<T>& pts, int64_t i) {
  pyi_axis_desc d = a.d;
  if (d.kind == PYI_AXIS_REGULAR) {
    if constexpr (std::is_same_v<T, double>) {
      return d.start.f + static_cast<double>(i) * d.step.f;
    } else {
      return d.start.i + i * d.step.i;
    }
  }
  return pts[static_cast<size_t>(i)];
}

// Uniform-bin table over an irregular f64 axis; must evaluate bin() exactly
// like the device does ((c - lut_min) * lut_inv with two roundings).
inline void build_lut(pyi_axis& a) {
  a.lut.clear();
  a.lut_bins = 0;
  const int64_t n = a.d.size;
  if (a.d.kind != PYI_AXIS_IRREGULAR || a.d.coord_type != PYI_COORD_F64 || n < 8) return;
  const double first = a.pf.front(), last = a.pf.back();
  const double span = last - first;
  if (!(std::fabs(span) > 0.0) || !std::isfinite(span)) return;
  const int bins = static_cast<int>(std::min<int64_t>(4 * n, 1 << 16));
  a.lut_min = first;
  a.lut_inv = static_cast<double>(bins) / span;
  if (!std::isfinite(a.lut_inv)) return;
  // bin -1 (f < 0) and bin `bins` (f >= bins) are the two open ends
  std::vector<int32_t> count(static_cast<size_t>(bins), 0);
  int32_t below0 = 0;
  for (int64_t k = 0; k < n; ++k) {
    volatile double diff = a.pf[static_cast<size_t>(k)] - a.lut_min;
    volatile double f = diff * a.lut_inv;
    if (f < 0.0) {
      below0 += 1;
    } else if (f < static_cast<double>(bins)) {
      count[static_cast<size_t>(static_cast<int>(f))] += 1;
    }
  }
  // lut[b] = number of points whose bin is < b
  a.lut.resize(static_cast<size_t>(bins) + 1);
  int32_t acc = below0;
  for (int b = 0; b <= bins; ++b) {
    a.lut[static_cast<size_t>(b)] = acc;
    if (b < bins) acc += count[static_cast<size_t>(b)];
  }
  a.lut_bins = bins;
}

// math::Axis<T>(values, epsilon, period) axis.hpp:696-716.  Returns nullptr
// and the reference's exception text in err.
template <typename T>
pyi_axis* make_axis(const T* values, int64_t n, T epsilon, bool has_period, T period,
                    std::string& err) {
  using Tr = Traits<T>;
  auto a = new pyi_axis();
  pyi_axis_desc& d = a->d;
  d.coord_type = std::is_same_v<T, double> ? PYI_COORD_F64 : PYI_COORD_I64;
  d.is_periodic = has_period ? 1 : 0;
  Tr::sc(d.period) = has_period ? period : T(0);
  if (n <= 0) {
    err = "Cannot create an Irregular axis with no points.";  // container.hpp:163-167
    delete a;
    return nullptr;
  }
  std::vector<T> v(values, values + n);
  const T per = Tr::sc(d.period);
  if (per != T(0) && n >= 2) {  // adjust_period axis.hpp:629-663
    const bool asc = v[0] < v[1];
    bool needs = false;
    for (int64_t k = 1; k < n && !needs; ++k) needs = asc ? v[k - 1] > v[k] : v[k - 1] < v[k];
    if (needs) {
      bool cross = false;
      for (int64_t k = 1; k < n; ++k) {
        if (!cross) cross = asc ? v[k - 1] > v[k] : v[k - 1] < v[k];
        if (cross) v[k] += asc ? per : -per;
      }
    }
  }
  bool regular = false;  // is_evenly_spaced axis.hpp:582-597
  if (n >= 2) {
    T maxdiff = T(0);
    for (int64_t k = 0; k < n; ++k) {
      const T df = Tr::abs_(v[k] - linspaced(n, v[0], v[n - 1], k));
      if (k == 0 || df > maxdiff) maxdiff = df;
    }
    regular = maxdiff <= epsilon;
  }
  d.size = n;
  if (regular) {  // Regular(values[0], values[n-1], n) container.hpp:415-436
    if (v[0] == v[n - 1]) {
      err = "An axis with a single value requires distinct start and stop values.";
      delete a;
      return nullptr;
    }
    d.kind = PYI_AXIS_REGULAR;
    Tr::sc(d.start) = v[0];
    if constexpr (std::is_same_v<T, double>) {
      d.step.f = (v[n - 1] - v[0]) / static_cast<double>(static_cast<int64_t>(n - 1));
      d.inv_step = 1.0 / d.step.f;  // container.hpp:547
    } else {
      d.step.i = (v[n - 1] - v[0]) / static_cast<int64_t>(n - 1);
      d.step_2 = d.step.i >> 1;  // container.hpp:585
    }
    d.is_ascending = coord<T>(*a, v, 0) < coord<T>(*a, v, 1);
  } else {
    d.kind = PYI_AXIS_IRREGULAR;
    Tr::sc(d.start) = v[0];
    d.is_ascending = n < 2 ? 1 : (v[0] < v[1]);
    bool mono = true;  // Irregular::is_monotonic container.hpp:189-201
    for (int64_t k = 1; k < n && mono; ++k) {
      if (v[k - 1] == v[k]) mono = false;
      if (d.is_ascending ? v[k] < v[k - 1] : v[k] > v[k - 1]) mono = false;
    }
    if (!mono) {
      err = "Axis values must be monotonic.";  // axis.hpp:606-608
      delete a;
      return nullptr;
    }
    if constexpr (std::is_same_v<T, double>) {
      a->pf = v;
    } else {
      a->pi = v;
    }
  }
  const T vmin = coord<T>(*a, v, d.is_ascending ? 0 : n - 1);
  const T vmax = coord<T>(*a, v, d.is_ascending ? n - 1 : 0);
  Tr::sc(d.min_value) = vmin;
  Tr::sc(d.max_value) = vmax;
  if (d.is_periodic) {  // compute_properties axis.hpp:609-623
    if (d.kind == PYI_AXIS_REGULAR) {
      if constexpr (std::is_same_v<T, double>) {
        d.is_periodic = is_same(std::fabs(d.step.f * static_cast<double>(n)), per, epsilon);
      } else {
        const int64_t span = static_cast<int64_t>(std::fabs(static_cast<double>(d.step.i * n)));
        d.is_periodic = is_same(span, per, epsilon);
      }
    } else if (n >= 2) {
      const T step = (v[n - 1] - v[0]) / static_cast<T>(n - 1);
      d.is_periodic = Tr::abs_((vmax - vmin) - per) <= step + epsilon;
    } else {
      d.is_periodic = 0;
    }
  }
  a->fix_points();
  build_lut(*a);
  return a;
}

// Axis::flip: Regular container.hpp:473-477,567-570 (step_2_ is left
// untouched), Irregular :184-187.
inline void flip_axis(pyi_axis& a) {
  pyi_axis_desc& d = a.d;
  if (d.kind == PYI_AXIS_REGULAR) {
    if (d.coord_type == PYI_COORD_F64) {
      d.start.f = d.start.f + static_cast<double>(d.size - 1) * d.step.f;
      d.step.f = -d.step.f;
      d.inv_step = -d.inv_step;
    } else {
      d.start.i = d.start.i + (d.size - 1) * d.step.i;
      d.step.i = -d.step.i;
    }
  } else if (d.coord_type == PYI_COORD_F64) {
    for (size_t i = 0, j = a.pf.size() - 1; i < j; ++i, --j) std::swap(a.pf[i], a.pf[j]);
    d.start.f = a.pf.front();
  } else {
    for (size_t i = 0, j = a.pi.size() - 1; i < j; ++i, --j) std::swap(a.pi[i], a.pi[j]);
    d.start.i = a.pi.front();
  }
  d.is_ascending = !d.is_ascending;
  // min / max are recomputed by the containers from the new start
  if (d.kind == PYI_AXIS_REGULAR) {
    if (d.coord_type == PYI_COORD_F64) {
      const double c0 = d.start.f;
      const double c1 = d.start.f + static_cast<double>(d.size - 1) * d.step.f;
      d.min_value.f = d.is_ascending ? c0 : c1;
      d.max_value.f = d.is_ascending ? c1 : c0;
    } else {
      const int64_t c0 = d.start.i;
      const int64_t c1 = d.start.i + (d.size - 1) * d.step.i;
      d.min_value.i = d.is_ascending ? c0 : c1;
      d.max_value.i = d.is_ascending ? c1 : c0;
    }
  }
  a.fix_points();
  build_lut(a);
}

}  // namespace pyi_host
