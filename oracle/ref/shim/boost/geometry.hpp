// TEST INFRASTRUCTURE ONLY -- not part of the product.
//
// Minimal stand-in for <boost/geometry.hpp> so that the reference headers
//   cxx/include/pyinterp/math/interpolate/geometric/bivariate.hpp
//   cxx/include/pyinterp/math/interpolate/geometric/multivariate.hpp
// compile UNMODIFIED in this container (Boost >= 1.90 is a system dependency
// of the reference, CMakeLists.txt:569, and is not installed here).
//
// Only what those two headers touch is provided:
//   model::point<T, N, CS>, cs::cartesian, cs::spherical_equatorial<degree>,
//   get<I>(p), comparable_distance(p, q).
//
// comparable_distance restates Boost.Geometry's default point-point strategies
// (published algorithm, boost/geometry/strategies/{cartesian/distance_pythagoras,
// spherical/distance_haversine}.hpp):
//   cartesian            : sum_i (p_i - q_i)^2
//   spherical_equatorial : comparable haversine on the unit sphere,
//        hav(lat2-lat1) + cos(lat1) * cos(lat2) * hav(lon2-lon1),
//        hav(t) = sin(t/2)^2, angles converted with d2r = pi/180 evaluated in
//        the coordinate type (calculation type = promote_floating_point of the
//        coordinate type: float stays float, double stays double).
// PARITY UNPINNED for the spherical branch: no reference test pins a value
// through it; tie/ordering behaviour is pinned by
// cxx/tests/math/interpolate/geometric/biavriate.cpp with cartesian points.
#pragma once

#include <cmath>
#include <cstddef>
#include <numbers>
#include <type_traits>

namespace boost::geometry {

struct degree {};
struct radian {};

namespace cs {
struct cartesian {};
template <typename Unit>
struct spherical_equatorial {};
template <typename Unit>
struct geographic {};
}  // namespace cs

namespace model {
template <typename T, std::size_t N, typename CS>
class point {
 public:
  using value_type = T;
  using coordinate_system = CS;
  static constexpr std::size_t dimension = N;

  constexpr point() = default;
  constexpr point(T v0, T v1) : v_{v0, v1} { static_assert(N >= 2); }

  template <std::size_t I>
  [[nodiscard]] constexpr auto get() const -> T {
    static_assert(I < N);
    return v_[I];
  }
  template <std::size_t I>
  constexpr void set(T v) {
    v_[I] = v;
  }

 private:
  T v_[N]{};
};
}  // namespace model

template <std::size_t I, typename T, std::size_t N, typename CS>
[[nodiscard]] constexpr auto get(const model::point<T, N, CS>& p) -> T {
  return p.template get<I>();
}

template <typename T, std::size_t N>
[[nodiscard]] constexpr auto comparable_distance(
    const model::point<T, N, cs::cartesian>& a,
    const model::point<T, N, cs::cartesian>& b) -> T {
  const T dx = get<0>(a) - get<0>(b);
  const T dy = get<1>(a) - get<1>(b);
  return dx * dx + dy * dy;
}

namespace detail {
template <typename T>
[[nodiscard]] inline auto hav(const T theta) -> T {
  const T s = std::sin(theta / T(2));
  return s * s;
}
}  // namespace detail

template <typename T>
[[nodiscard]] inline auto comparable_distance(
    const model::point<T, 2, cs::spherical_equatorial<degree>>& a,
    const model::point<T, 2, cs::spherical_equatorial<degree>>& b) -> T {
  const T d2r = std::numbers::pi_v<T> / T(180);
  const T lon1 = get<0>(a) * d2r;
  const T lat1 = get<1>(a) * d2r;
  const T lon2 = get<0>(b) * d2r;
  const T lat2 = get<1>(b) * d2r;
  return detail::hav(lat2 - lat1) +
         std::cos(lat1) * std::cos(lat2) * detail::hav(lon2 - lon1);
}

}  // namespace boost::geometry
