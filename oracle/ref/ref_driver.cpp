// TEST INFRASTRUCTURE ONLY -- not part of the product.
//
// liboracle_ref.so: the reference's own per-point math, compiled IN PLACE from
// the headers under /root/reference/cxx/include (nothing is copied into this
// repository), behind the C-ABI of ../orc_abi.h.
//
// Used directly from the reference tree:
//   math::Axis<double>, math::Axis<int64_t>          math/axis.hpp, math/axis/container.hpp
//   geometric::Bilinear / InverseDistanceWeighting / Nearest, trivariate(),
//   quadrivariate(), linear(), nearest()              math/interpolate/geometric/*.hpp,
//                                                     math/interpolate/{linear,nearest}.hpp
//   InterpolationCache                                math/interpolate/cache.hpp
//   bivariate::Bicubic / Bilinear, univariate::CSpline  math/interpolate/bivariate/*.hpp
//   parallel_for                                      parallel_for.hpp
// <boost/geometry.hpp> is satisfied by oracle/ref/shim (see its header).
//
// What cannot be compiled here (GCC 13 has no "deducing this"; nanobind grid
// container) is driven through a duck-typed grid, the same device the
// reference's own unit test uses (cxx/tests/math/interpolate/cache_loader.cpp:15-63):
//   - the per-point glue of pybind/geometric/{bi,tri,quadri}variate.hpp
//   - the window loader of math/interpolate/cache_loader.hpp:46-212
//   - the glue of pybind/windowed/bivariate.hpp:46-115
#include <algorithm>
#include <array>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <format>
#include <limits>
#include <memory>
#include <optional>
#include <stdexcept>
#include <string>
#include <tuple>
#include <utility>
#include <vector>

#include "pyinterp/math/axis.hpp"
#include "pyinterp/math/interpolate/bivariate/bicubic.hpp"
#include "pyinterp/math/interpolate/bivariate/bilinear.hpp"
#include "pyinterp/math/interpolate/cache.hpp"
#include "pyinterp/math/interpolate/geometric/bivariate.hpp"
#include "pyinterp/math/interpolate/geometric/multivariate.hpp"
#include "pyinterp/math/interpolate/univariate/cspline.hpp"
#include "pyinterp/parallel_for.hpp"

#include "../orc_abi.h"

namespace ref = pyinterp;
namespace geo = pyinterp::math::interpolate::geometric;
namespace itp = pyinterp::math::interpolate;

template <typename T>
using SphericalPoint = boost::geometry::model::point<
    T, 2, boost::geometry::cs::spherical_equatorial<boost::geometry::degree>>;
template <typename T>
using CartesianPoint =
    boost::geometry::model::point<T, 2, boost::geometry::cs::cartesian>;

// ---------------------------------------------------------------------------
// axis handle
// ---------------------------------------------------------------------------
struct orc_axis {
  int coord_type{PYI_COORD_F64};
  std::unique_ptr<ref::math::Axis<double>> f;
  std::unique_ptr<ref::math::Axis<int64_t>> i;
  double period_f{0};
  int64_t period_i{0};
  int64_t initial_step_i{0};  // Regular<int64>: step_ at construction (step_2_ = step_ >> 1)
  // storage behind pyi_axis_desc::points handed out by orc_axis_describe
  mutable std::vector<double> pts_f;
  mutable std::vector<int64_t> pts_i;
};

namespace {

void set_err(char* err, size_t len, const char* what) {
  if (err != nullptr && len > 0) {
    std::snprintf(err, len, "%s", what);
  }
}

// Raised through parallel_for exactly like the reference's bounds error; it
// additionally carries the point index and the axis.
struct OobError : std::invalid_argument {
  OobError(int axis_, const std::string& msg)
      : std::invalid_argument(msg), axis(axis_) {}
  int64_t index{-1};
  int axis;
};

constexpr std::array<std::string_view, 4> kLabels{"x", "y", "z", "u"};

// Duck-typed stand-in for pybind::Grid<DataType, Axes...> (pybind/grid.hpp).
template <typename DataType, typename ZT, size_t NDim>
struct DuckGrid {
  static constexpr size_t kNDim = NDim;
  const ref::math::Axis<double>* ax{};
  const ref::math::Axis<double>* ay{};
  const ref::math::Axis<ZT>* az{};
  const ref::math::Axis<double>* au{};
  const DataType* data{};
  int64_t n1{1}, n2{1}, n3{1};  // sizes of axes 1..3 (C order)

  template <size_t I>
  [[nodiscard]] auto axis() const -> const auto& {
    if constexpr (I == 0) {
      return *ax;
    } else if constexpr (I == 1) {
      return *ay;
    } else if constexpr (I == 2) {
      return *az;
    } else {
      return *au;
    }
  }
  // pybind/grid.hpp:340-345 (C-contiguous view)
  [[nodiscard]] auto value(int64_t i, int64_t j) const -> const DataType& {
    return data[i * n1 + j];
  }
  [[nodiscard]] auto value(int64_t i, int64_t j, int64_t k) const
      -> const DataType& {
    return data[(i * n1 + j) * n2 + k];
  }
  [[nodiscard]] auto value(int64_t i, int64_t j, int64_t k, int64_t l) const
      -> const DataType& {
    return data[((i * n1 + j) * n2 + k) * n3 + l];
  }
  // pybind/grid.hpp:362-371
  template <size_t I, typename C>
  [[nodiscard]] auto construct_bounds_error_description(const C& c) const
      -> std::string {
    const auto& a = axis<I>();
    return std::format("{} is out of bounds for axis {} [{}, ..., {}]",
                       a.coordinate_repr(c), kLabels[I],
                       a.coordinate_repr(a.min_value()),
                       a.coordinate_repr(a.max_value()));
  }
  // pybind/grid.hpp:388-398
  template <size_t I, typename C>
  [[nodiscard]] auto find_indexes(const C& c, bool bounds_error) const
      -> std::optional<std::pair<size_t, size_t>> {
    auto r = axis<I>().find_indexes(c);
    if (!r.has_value() && bounds_error) {
      throw OobError(static_cast<int>(I),
                     construct_bounds_error_description<I>(c));
    }
    if (!r) {
      return std::nullopt;
    }
    return std::make_pair(static_cast<size_t>(r->first),
                          static_cast<size_t>(r->second));
  }
};

// pybind/grid_dispatch.hpp:25-30
template <typename D>
using ResultOf = std::conditional_t<std::is_same_v<D, double> ||
                                        std::is_same_v<D, int64_t> ||
                                        std::is_same_v<D, uint64_t>,
                                    double, float>;

}  // namespace

// ---------------------------------------------------------------------------
// axes
// ---------------------------------------------------------------------------
extern "C" const char* orc_kind(void) { return "reference"; }

extern "C" orc_axis* orc_axis_new_f64(const double* values, int64_t n,
                                      double epsilon, int has_period,
                                      double period, char* err, size_t errlen) {
  try {
    auto h = std::make_unique<orc_axis>();
    h->coord_type = PYI_COORD_F64;
    Eigen::Map<const ref::Vector<double>> v(values, n);
    h->f = std::make_unique<ref::math::Axis<double>>(
        v, epsilon, has_period ? std::optional<double>(period) : std::nullopt);
    h->period_f = has_period ? period : 0.0;
    return h.release();
  } catch (const std::exception& e) {
    set_err(err, errlen, e.what());
    return nullptr;
  }
}

extern "C" orc_axis* orc_axis_new_i64(const int64_t* values, int64_t n,
                                      int64_t epsilon, int has_period,
                                      int64_t period, char* err,
                                      size_t errlen) {
  try {
    auto h = std::make_unique<orc_axis>();
    h->coord_type = PYI_COORD_I64;
    Eigen::Map<const ref::Vector<int64_t>> v(values, n);
    h->i = std::make_unique<ref::math::Axis<int64_t>>(
        v, epsilon,
        has_period ? std::optional<int64_t>(period) : std::nullopt);
    h->period_i = has_period ? period : 0;
    if (h->i->is_regular()) {
      h->initial_step_i = h->i->increment();
    }
    return h.release();
  } catch (const std::exception& e) {
    set_err(err, errlen, e.what());
    return nullptr;
  }
}

extern "C" void orc_axis_free(orc_axis* axis) { delete axis; }

extern "C" void orc_axis_flip(orc_axis* axis) {
  if (axis->f) axis->f->flip();
  if (axis->i) axis->i->flip();
}

namespace {
template <typename T>
const T* irregular_points(const ref::math::Axis<T>& a, std::vector<T>& store) {
  store.resize(static_cast<size_t>(a.size()));
  for (int64_t k = 0; k < a.size(); ++k) store[static_cast<size_t>(k)] = a(k);
  return store.data();
}
}  // namespace

extern "C" int orc_axis_describe(const orc_axis* axis, pyi_axis_desc* d) {
  std::memset(d, 0, sizeof(*d));
  d->coord_type = axis->coord_type;
  if (axis->coord_type == PYI_COORD_F64) {
    const auto& a = *axis->f;
    d->kind = a.is_regular() ? PYI_AXIS_REGULAR : PYI_AXIS_IRREGULAR;
    d->size = a.size();
    d->is_ascending = a.is_ascending();
    d->is_periodic = a.is_periodic();
    d->period.f = axis->period_f;
    d->start.f = a.front();
    d->min_value.f = a.min_value();
    d->max_value.f = a.max_value();
    if (a.is_regular()) {
      d->step.f = a.increment();
      d->inv_step = 1.0 / a.increment();
    } else {
      d->points = irregular_points(a, axis->pts_f);
    }
  } else {
    const auto& a = *axis->i;
    d->kind = a.is_regular() ? PYI_AXIS_REGULAR : PYI_AXIS_IRREGULAR;
    d->size = a.size();
    d->is_ascending = a.is_ascending();
    d->is_periodic = a.is_periodic();
    d->period.i = axis->period_i;
    d->start.i = a.front();
    d->min_value.i = a.min_value();
    d->max_value.i = a.max_value();
    if (a.is_regular()) {
      d->step.i = a.increment();
      d->step_2 = axis->initial_step_i >> 1;
    } else {
      d->points = irregular_points(a, axis->pts_i);
    }
  }
  return PYI_OK;
}

extern "C" void orc_axis_find_indexes(const orc_axis* axis, int64_t n,
                                      const void* coords, int64_t* out) {
  auto run = [&](const auto& a, const auto* c) {
    for (int64_t k = 0; k < n; ++k) {
      auto r = a.find_indexes(c[k]);
      out[2 * k] = r ? r->first : -1;
      out[2 * k + 1] = r ? r->second : -1;
    }
  };
  if (axis->coord_type == PYI_COORD_F64) {
    run(*axis->f, static_cast<const double*>(coords));
  } else {
    run(*axis->i, static_cast<const int64_t*>(coords));
  }
}

extern "C" void orc_axis_find_index(const orc_axis* axis, int64_t n,
                                    const void* coords, int bounded,
                                    int64_t* out) {
  auto run = [&](const auto& a, const auto* c) {
    for (int64_t k = 0; k < n; ++k) out[k] = a.find_index(c[k], bounded != 0);
  };
  if (axis->coord_type == PYI_COORD_F64) {
    run(*axis->f, static_cast<const double*>(coords));
  } else {
    run(*axis->i, static_cast<const int64_t*>(coords));
  }
}

namespace {
auto to_ref_boundary(int b) -> ref::math::axis::Boundary {
  switch (b) {
    case PYI_BOUNDARY_SHRINK:
      return ref::math::axis::kShrink;
    case ORC_BOUNDARY_EXPAND:
      return ref::math::axis::kExpand;
    case ORC_BOUNDARY_WRAP:
      return ref::math::axis::kWrap;
    case ORC_BOUNDARY_SYM:
      return ref::math::axis::kSym;
    default:
      return ref::math::axis::kUndef;
  }
}
}  // namespace

extern "C" int orc_axis_find_window(const orc_axis* axis, const void* coord,
                                    int half, int boundary, int64_t* idx,
                                    int* count, int* c0, int* c1) {
  auto run = [&](const auto& a, auto c) -> int {
    auto r = a.find_indexes(c, static_cast<size_t>(half),
                            to_ref_boundary(boundary));
    if (!r) return 0;
    *count = static_cast<int>(r->first.size());
    for (size_t k = 0; k < r->first.size(); ++k) idx[k] = r->first[k];
    *c0 = static_cast<int>(r->second.first);
    *c1 = static_cast<int>(r->second.second);
    return 1;
  };
  if (axis->coord_type == PYI_COORD_F64) {
    return run(*axis->f, *static_cast<const double*>(coord));
  }
  return run(*axis->i, *static_cast<const int64_t*>(coord));
}

// ---------------------------------------------------------------------------
// geometric drivers: glue of pybind/geometric/{bi,tri,quadri}variate.hpp over
// the duck-typed grid, evaluators straight from the reference headers.
// ---------------------------------------------------------------------------
namespace {

template <typename R>
using SpatialEval = geo::Bivariate<SphericalPoint, R>;

// Brackets + spatial corner coordinates shared by the three drivers
// (pybind/geometric/bivariate.hpp:52-89).
template <typename R>
struct Cell {
  size_t ix[2], iy[2];
  R x0, x1, y0, y1, px, py;
};

template <typename G, typename R>
auto locate_xy(const G& g, double x, double y, bool be, Cell<R>& c) -> bool {
  auto xi = g.template find_indexes<0>(x, be);
  if (!xi) return false;
  auto yi = g.template find_indexes<1>(y, be);
  if (!yi) return false;
  c.ix[0] = xi->first, c.ix[1] = xi->second;
  c.iy[0] = yi->first, c.iy[1] = yi->second;
  return true;
}

template <typename G, typename R>
void corner_coords(const G& g, double x, double y, Cell<R>& c) {
  const auto& xa = g.template axis<0>();
  const auto& ya = g.template axis<1>();
  c.x0 = static_cast<R>(xa.coordinate_value(c.ix[0]));
  c.x1 = static_cast<R>(
      xa.normalize_coordinate(xa.coordinate_value(c.ix[1]), c.x0));
  c.y0 = static_cast<R>(ya.coordinate_value(c.iy[0]));
  c.y1 = static_cast<R>(ya.coordinate_value(c.iy[1]));
  c.px = static_cast<R>(xa.normalize_coordinate(x, c.x0));
  c.py = static_cast<R>(y);
}

template <typename G, typename R>
auto point2d(const G& g, double x, double y, const SpatialEval<R>& sp, bool be)
    -> std::optional<R> {
  Cell<R> c;
  if (!locate_xy(g, x, y, be, c)) return std::nullopt;
  const auto v00 = static_cast<R>(g.value(c.ix[0], c.iy[0]));
  const auto v01 = static_cast<R>(g.value(c.ix[0], c.iy[1]));
  const auto v10 = static_cast<R>(g.value(c.ix[1], c.iy[0]));
  const auto v11 = static_cast<R>(g.value(c.ix[1], c.iy[1]));
  corner_coords(g, x, y, c);
  return static_cast<R>(sp.evaluate(SphericalPoint<R>(c.px, c.py),
                                    SphericalPoint<R>(c.x0, c.y0),
                                    SphericalPoint<R>(c.x1, c.y1), v00, v01,
                                    v10, v11));
}

template <typename G, typename R, typename ZT>
auto point3d(const G& g, double x, double y, ZT z, const SpatialEval<R>& sp,
             const geo::AxisInterpolator<ZT, R>& zi, bool be)
    -> std::optional<R> {
  Cell<R> c;
  if (!locate_xy(g, x, y, be, c)) return std::nullopt;
  auto zr = g.template find_indexes<2>(z, be);
  if (!zr) return std::nullopt;
  const size_t iz[2] = {zr->first, zr->second};
  R v[2][2][2];  // [x][y][z]
  for (int a = 0; a < 2; ++a)
    for (int b = 0; b < 2; ++b)
      for (int k = 0; k < 2; ++k)
        v[a][b][k] = static_cast<R>(g.value(c.ix[a], c.iy[b], iz[k]));
  corner_coords(g, x, y, c);
  const auto& za = g.template axis<2>();
  const auto z0 = static_cast<ZT>(za.coordinate_value(iz[0]));
  const auto z1 = static_cast<ZT>(za.coordinate_value(iz[1]));
  using P3 = geo::SpatialPoint3D<SphericalPoint, R, ZT>;
  const auto cube =
      geo::DataCube<R>(v[0][0][0], v[0][1][0], v[1][0][0], v[1][1][0],
                       v[0][0][1], v[0][1][1], v[1][0][1], v[1][1][1]);
  return static_cast<R>(geo::trivariate(
      P3(SphericalPoint<R>(c.px, c.py), z), P3(SphericalPoint<R>(c.x0, c.y0), z0),
      P3(SphericalPoint<R>(c.x1, c.y1), z1), cube, sp, zi));
}

template <typename G, typename R, typename ZT>
auto point4d(const G& g, double x, double y, ZT z, double u,
             const SpatialEval<R>& sp, const geo::AxisInterpolator<ZT, R>& zi,
             const geo::AxisInterpolator<double, R>& ui, bool be)
    -> std::optional<R> {
  Cell<R> c;
  if (!locate_xy(g, x, y, be, c)) return std::nullopt;
  auto zr = g.template find_indexes<2>(z, be);
  if (!zr) return std::nullopt;
  auto ur = g.template find_indexes<3>(u, be);
  if (!ur) return std::nullopt;
  const size_t iz[2] = {zr->first, zr->second};
  const size_t iu[2] = {ur->first, ur->second};
  R v[2][2][2][2];  // [x][y][z][u]
  for (int a = 0; a < 2; ++a)
    for (int b = 0; b < 2; ++b)
      for (int k = 0; k < 2; ++k)
        for (int l = 0; l < 2; ++l)
          v[a][b][k][l] =
              static_cast<R>(g.value(c.ix[a], c.iy[b], iz[k], iu[l]));
  corner_coords(g, x, y, c);
  const auto& za = g.template axis<2>();
  const auto& ua = g.template axis<3>();
  const auto z0 = static_cast<ZT>(za.coordinate_value(iz[0]));
  const auto z1 = static_cast<ZT>(za.coordinate_value(iz[1]));
  const auto u0 = static_cast<double>(ua.coordinate_value(iu[0]));
  const auto u1 = static_cast<double>(ua.coordinate_value(iu[1]));
  using P4 = geo::SpatialPoint4D<SphericalPoint, R, ZT, double>;
  // argument order of DataHypercube: multivariate.hpp:213-243
  const auto cube = geo::DataHypercube<R>(
      v[0][0][0][0], v[0][1][0][0], v[1][0][0][0], v[1][1][0][0],
      v[0][0][1][0], v[0][1][1][0], v[1][0][1][0], v[1][1][1][0],
      v[0][0][0][1], v[0][1][0][1], v[1][0][0][1], v[1][1][0][1],
      v[0][0][1][1], v[0][1][1][1], v[1][0][1][1], v[1][1][1][1]);
  return static_cast<R>(geo::quadrivariate(
      P4(SphericalPoint<R>(c.px, c.py), z, u),
      P4(SphericalPoint<R>(c.x0, c.y0), z0, u0),
      P4(SphericalPoint<R>(c.x1, c.y1), z1, u1), cube, sp, zi, ui));
}

auto to_axis_method(int m) -> geo::AxisMethod {
  return m == PYI_AXIS_NEAREST ? geo::AxisMethod::kNearest
                               : geo::AxisMethod::kLinear;
}

struct OobOut {
  int64_t* index;
  int* axis;
  char* msg;
  size_t msglen;
};

// Vectorised driver: NaN prefill + reference parallel_for
// (pybind/geometric/bivariate.hpp:141-158).
template <typename R, typename PointFn>
auto run_points(int64_t n, int num_threads, R* out, const OobOut& oob,
                PointFn&& fn) -> int {
  std::fill(out, out + n, std::numeric_limits<R>::quiet_NaN());
  try {
    ref::parallel_for(
        n,
        [&](const int64_t start, const int64_t end) {
          for (int64_t k = start; k < end; ++k) {
            try {
              auto r = fn(k);
              if (r.has_value()) out[k] = *r;
            } catch (OobError& e) {
              e.index = k;
              throw;
            }
          }
        },
        num_threads);
  } catch (const OobError& e) {
    if (oob.index) *oob.index = e.index;
    if (oob.axis) *oob.axis = e.axis;
    set_err(oob.msg, oob.msglen, e.what());
    return PYI_OOB;
  }
  return PYI_OK;
}

template <typename D, typename ZT>
auto geometric_typed(int ndim, const orc_axis* const* axes, const void* data,
                     const pyi_geometric_config* cfg, int64_t n,
                     const double* x, const double* y, const void* z,
                     const double* u, void* out, const OobOut& oob) -> int {
  using R = ResultOf<D>;
  auto sp = geo::make_interpolator<SphericalPoint, R>(
      static_cast<geo::InterpolationMethod>(cfg->method), cfg->idw_exp);
  const bool be = cfg->bounds_error != 0;
  auto* o = static_cast<R*>(out);
  const auto* d = static_cast<const D*>(data);
  auto zaxis = [&]() -> const ref::math::Axis<ZT>* {
    if (ndim < 3) return nullptr;
    if constexpr (std::is_same_v<ZT, double>) {
      return axes[2]->f.get();
    } else {
      return axes[2]->i.get();
    }
  };
  if (ndim == 2) {
    DuckGrid<D, ZT, 2> g{axes[0]->f.get(), axes[1]->f.get(), nullptr, nullptr,
                         d, axes[1]->f->size()};
    return run_points<R>(n, cfg->num_threads, o, oob, [&](int64_t k) {
      return point2d<decltype(g), R>(g, x[k], y[k], *sp, be);
    });
  }
  const auto zi = geo::get_axis_interpolator<ZT, R>(to_axis_method(cfg->z_method));
  const auto* zq = static_cast<const ZT*>(z);
  if (ndim == 3) {
    DuckGrid<D, ZT, 3> g{axes[0]->f.get(), axes[1]->f.get(), zaxis(), nullptr,
                         d, axes[1]->f->size(), zaxis()->size()};
    return run_points<R>(n, cfg->num_threads, o, oob, [&](int64_t k) {
      return point3d<decltype(g), R, ZT>(g, x[k], y[k], zq[k], *sp, zi, be);
    });
  }
  const auto ui =
      geo::get_axis_interpolator<double, R>(to_axis_method(cfg->u_method));
  DuckGrid<D, ZT, 4> g{axes[0]->f.get(), axes[1]->f.get(), zaxis(),
                       axes[3]->f.get(), d, axes[1]->f->size(),
                       zaxis()->size(), axes[3]->f->size()};
  return run_points<R>(n, cfg->num_threads, o, oob, [&](int64_t k) {
    return point4d<decltype(g), R, ZT>(g, x[k], y[k], zq[k], u[k], *sp, zi, ui,
                                       be);
  });
}

template <typename ZT, typename... A>
auto geometric_dtype(int dtype, A&&... a) -> int {
  switch (dtype) {
    case PYI_I8: return geometric_typed<int8_t, ZT>(a...);
    case PYI_U8: return geometric_typed<uint8_t, ZT>(a...);
    case PYI_I16: return geometric_typed<int16_t, ZT>(a...);
    case PYI_U16: return geometric_typed<uint16_t, ZT>(a...);
    case PYI_I32: return geometric_typed<int32_t, ZT>(a...);
    case PYI_U32: return geometric_typed<uint32_t, ZT>(a...);
    case PYI_I64: return geometric_typed<int64_t, ZT>(a...);
    case PYI_U64: return geometric_typed<uint64_t, ZT>(a...);
    case PYI_F32: return geometric_typed<float, ZT>(a...);
    case PYI_F64: return geometric_typed<double, ZT>(a...);
    default: return PYI_BAD_ARG;
  }
}

}  // namespace

extern "C" int orc_geometric(int ndim, const orc_axis* const* axes, int dtype,
                             const void* data, const pyi_geometric_config* cfg,
                             int64_t n, const double* x, const double* y,
                             const void* z, const double* u, void* out,
                             int64_t* oob_index, int* oob_axis, char* msg,
                             size_t msglen) {
  if (ndim < 2 || ndim > 4) return PYI_BAD_ARG;
  const OobOut oob{oob_index, oob_axis, msg, msglen};
  const bool zi64 = ndim >= 3 && axes[2]->coord_type == PYI_COORD_I64;
  if (zi64) {
    return geometric_dtype<int64_t>(dtype, ndim, axes, data, cfg, n, x, y, z, u,
                                    out, oob);
  }
  return geometric_dtype<double>(dtype, ndim, axes, data, cfg, n, x, y, z, u,
                                 out, oob);
}

// ---------------------------------------------------------------------------
// windowed 2-D driver: reference InterpolationCache + Bicubic/Bilinear; the
// loader (math/interpolate/cache_loader.hpp:46-212) and the per-point glue
// (pybind/windowed/bivariate.hpp:46-76) are driven over the duck grid.
// ---------------------------------------------------------------------------
namespace {

using Cache2D = itp::InterpolationCache<double, double, double>;

struct WinError : std::out_of_range {
  WinError(int axis_, const std::string& m) : std::out_of_range(m), axis(axis_) {}
  int64_t index{-1};
  int axis;
};

// cache_loader.hpp:46-172 for kNDim == 2. Returns false when the window
// cannot be built or holds a NaN; err/err_axis set only for a coordinate
// truly outside its axis with bounds_error.
template <typename G>
auto load_window(const G& g, double xq, double yq,
                 ref::math::axis::Boundary boundary, bool bounds_error,
                 Cache2D& cache, std::optional<std::string>& err,
                 int& err_axis) -> bool {
  const auto& ax = g.template axis<0>();
  const auto& ay = g.template axis<1>();
  auto wx = ax.find_indexes(xq, cache.template half_window<0>(), boundary);
  if (!wx) {
    if (bounds_error && !ax.contains(xq)) {
      err = g.template construct_bounds_error_description<0>(xq);
      err_axis = 0;
    }
    return false;
  }
  auto wy = ay.find_indexes(yq, cache.template half_window<1>(), boundary);
  if (!wy) {
    if (bounds_error && !ay.contains(yq)) {
      err = g.template construct_bounds_error_description<1>(yq);
      err_axis = 1;
    }
    return false;
  }
  std::array<size_t, 2> npts{wx->first.size(), wy->first.size()};
  if (npts[0] != cache.points_per_dim(0) || npts[1] != cache.points_per_dim(1)) {
    cache.resize(std::move(npts));
  }
  const double period = ax.is_periodic() ? ax.period().value() : 0.0;
  for (size_t k = 0; k < wx->first.size(); ++k) {
    auto raw = ax(wx->first[k]);
    if (period != 0.0) {
      raw = ref::math::normalize_period_half(raw, xq, period);
    }
    cache.template set_coord<0>(k, raw);
  }
  for (size_t k = 0; k < wy->first.size(); ++k) {
    cache.template set_coord<1>(k, ay(wy->first[k]));
  }
  auto& flat = cache.values_flat();
  size_t pos = 0;
  for (int64_t i : wx->first) {
    for (int64_t j : wy->first) {
      flat[pos++] = static_cast<double>(g.value(i, j));
    }
  }
  cache.finalize({wx->second, wy->second});
  return cache.is_valid();
}

template <typename G, typename R>
auto window_point(const G& g, double x, double y,
                  const pyi_windowed_config* cfg,
                  itp::BivariateBase<double>* interp, Cache2D& cache)
    -> std::optional<R> {
  // update_cache_if_needed, cache_loader.hpp:188-212
  bool success = true;
  bool was_updated = false;
  std::optional<std::string> err;
  int err_axis = -1;
  if (!(cache.has_domain() && cache.contains(x, y))) {
    was_updated = true;
    success = load_window(g, x, y, to_ref_boundary(cfg->boundary),
                          cfg->bounds_error != 0, cache, err, err_axis);
  }
  if (!success) {
    if (err.has_value()) throw WinError(err_axis, *err);
    return std::nullopt;
  }
  if (!cache.is_valid()) return std::nullopt;
  if (was_updated) {
    interp->prepare(cache.template coords_as_eigen<0>(),
                    cache.template coords_as_eigen<1>(), cache.matrix());
  }
  return static_cast<R>((*interp)(x, y));
}

auto make_window_interp(int method)
    -> std::unique_ptr<itp::BivariateBase<double>> {
  if (method == PYI_W_BILINEAR) {
    return std::make_unique<itp::bivariate::Bilinear<double>>();
  }
  return std::make_unique<itp::bivariate::Bicubic<double>>();
}

template <typename D>
auto windowed_typed(const orc_axis* const* axes, const void* data,
                    const pyi_windowed_config* cfg, int stateless, int64_t n,
                    const double* x, const double* y, void* out,
                    const OobOut& oob) -> int {
  using R = ResultOf<D>;
  auto* o = static_cast<R*>(out);
  DuckGrid<D, double, 2> g{axes[0]->f.get(), axes[1]->f.get(), nullptr, nullptr,
                           static_cast<const D*>(data), axes[1]->f->size()};
  std::fill(o, o + n, std::numeric_limits<R>::quiet_NaN());
  try {
    ref::parallel_for(
        n,
        [&](const int64_t start, const int64_t end) {
          auto cache = Cache2D(cfg->half_x, cfg->half_y);
          auto interp = make_window_interp(cfg->method);
          for (int64_t k = start; k < end; ++k) {
            if (stateless) {
              cache = Cache2D(cfg->half_x, cfg->half_y);
            }
            try {
              auto r = window_point<decltype(g), R>(g, x[k], y[k], cfg,
                                                    interp.get(), cache);
              if (r.has_value()) o[k] = *r;
            } catch (WinError& e) {
              e.index = k;
              throw;
            }
          }
        },
        cfg->num_threads);
  } catch (const WinError& e) {
    if (oob.index) *oob.index = e.index;
    if (oob.axis) *oob.axis = e.axis;
    set_err(oob.msg, oob.msglen, e.what());
    return PYI_OOB;
  }
  return PYI_OK;
}

}  // namespace

extern "C" int orc_windowed2d(const orc_axis* const* axes, int dtype,
                              const void* data, const pyi_windowed_config* cfg,
                              int stateless, int64_t n, const double* x,
                              const double* y, void* out, int64_t* oob_index,
                              int* oob_axis, char* msg, size_t msglen) {
  const OobOut oob{oob_index, oob_axis, msg, msglen};
  if (cfg->half_x < 1 || cfg->half_y < 1) return PYI_BAD_ARG;
#define ORC_W(T) windowed_typed<T>(axes, data, cfg, stateless, n, x, y, out, oob)
  switch (dtype) {
    case PYI_I8: return ORC_W(int8_t);
    case PYI_U8: return ORC_W(uint8_t);
    case PYI_I16: return ORC_W(int16_t);
    case PYI_U16: return ORC_W(uint16_t);
    case PYI_I32: return ORC_W(int32_t);
    case PYI_U32: return ORC_W(uint32_t);
    case PYI_I64: return ORC_W(int64_t);
    case PYI_U64: return ORC_W(uint64_t);
    case PYI_F32: return ORC_W(float);
    case PYI_F64: return ORC_W(double);
    default: return PYI_BAD_ARG;
  }
#undef ORC_W
}

// ---------------------------------------------------------------------------
// scalar probes (golden-vector tests)
// ---------------------------------------------------------------------------
namespace {
template <template <class> class P, typename T>
auto eval2d(int method, int idw_exp, T x, T y, T x0, T y0, T x1, T y1, T q00,
            T q01, T q10, T q11) -> T {
  auto e = geo::make_interpolator<P, T>(
      static_cast<geo::InterpolationMethod>(method), idw_exp);
  return e->evaluate(P<T>(x, y), P<T>(x0, y0), P<T>(x1, y1), q00, q01, q10, q11);
}
}  // namespace

extern "C" double orc_eval2d_f64(int method, int idw_exp, int cartesian,
                                 double x, double y, double x0, double y0,
                                 double x1, double y1, double q00, double q01,
                                 double q10, double q11) {
  return cartesian ? eval2d<CartesianPoint, double>(method, idw_exp, x, y, x0,
                                                    y0, x1, y1, q00, q01, q10,
                                                    q11)
                   : eval2d<SphericalPoint, double>(method, idw_exp, x, y, x0,
                                                    y0, x1, y1, q00, q01, q10,
                                                    q11);
}

extern "C" float orc_eval2d_f32(int method, int idw_exp, int cartesian, float x,
                                float y, float x0, float y0, float x1, float y1,
                                float q00, float q01, float q10, float q11) {
  return cartesian ? eval2d<CartesianPoint, float>(method, idw_exp, x, y, x0,
                                                   y0, x1, y1, q00, q01, q10,
                                                   q11)
                   : eval2d<SphericalPoint, float>(method, idw_exp, x, y, x0,
                                                   y0, x1, y1, q00, q01, q10,
                                                   q11);
}

extern "C" double orc_axis_interp_f64(int method, double x, double x0,
                                      double x1, double y0, double y1) {
  return method == PYI_AXIS_NEAREST
             ? itp::nearest<double, double>(x, x0, x1, y0, y1)
             : itp::linear<double, double>(x, x0, x1, y0, y1);
}

extern "C" double orc_axis_interp_i64(int method, int64_t x, int64_t x0,
                                      int64_t x1, double y0, double y1) {
  return method == PYI_AXIS_NEAREST
             ? itp::nearest<int64_t, double>(x, x0, x1, y0, y1)
             : itp::linear<int64_t, double>(x, x0, x1, y0, y1);
}

extern "C" float orc_axis_interp_i64_f32(int method, int64_t x, int64_t x0,
                                         int64_t x1, float y0, float y1) {
  return method == PYI_AXIS_NEAREST
             ? itp::nearest<int64_t, float>(x, x0, x1, y0, y1)
             : itp::linear<int64_t, float>(x, x0, x1, y0, y1);
}

extern "C" double orc_window_eval(int method, const double* xa, int nx,
                                  const double* ya, int ny, const double* za,
                                  double x, double y) {
  auto interp = make_window_interp(method);
  Eigen::Map<const ref::Vector<double>> vx(xa, nx);
  Eigen::Map<const ref::Vector<double>> vy(ya, ny);
  ref::Matrix<double> m(nx, ny);
  for (int i = 0; i < nx; ++i)
    for (int j = 0; j < ny; ++j) m(i, j) = za[i * ny + j];
  interp->prepare(vx, vy, m);
  return (*interp)(x, y);
}

extern "C" double orc_cspline(const double* xa, const double* ya, int n,
                              double x, int derivative) {
  itp::univariate::CSpline<double> s;
  ref::Vector<double> vx = Eigen::Map<const ref::Vector<double>>(xa, n);
  ref::Vector<double> vy = Eigen::Map<const ref::Vector<double>>(ya, n);
  return derivative ? s.derivative(vx, vy, x) : s(vx, vy, x);
}

// ---------------------------------------------------------------------------
// windowed N-D driver (1-D .. 4-D, every fitting model): the reference's
// InterpolationCache, factory(), Univariate / BivariateBase interpolators,
// linear() and nearest() used unmodified; restated over a duck grid (GCC 13):
//   - load_cache_generic                 math/interpolate/cache_loader.hpp:46-172
//   - univariate_single / _derivative    pybind/windowed/univariate.hpp:47-120
//   - bivariate_single                   pybind/windowed/bivariate.hpp:46-76
//   - trivariate_single                  pybind/windowed/trivariate.hpp:66-123
//   - quadrivariate_single               pybind/windowed/quadrivariate.hpp:50-148
// ---------------------------------------------------------------------------
#include "pyinterp/math/interpolate/factory.hpp"

namespace {

// Grid::value (pybind/grid.hpp:340-345) cast to the cache's double
// (cache_loader.hpp:150-153), any dtype.
auto fetch_f64(const void* data, int dtype, int64_t off) -> double {
  switch (dtype) {
    case PYI_I8: return static_cast<double>(static_cast<const int8_t*>(data)[off]);
    case PYI_U8: return static_cast<double>(static_cast<const uint8_t*>(data)[off]);
    case PYI_I16: return static_cast<double>(static_cast<const int16_t*>(data)[off]);
    case PYI_U16: return static_cast<double>(static_cast<const uint16_t*>(data)[off]);
    case PYI_I32: return static_cast<double>(static_cast<const int32_t*>(data)[off]);
    case PYI_U32: return static_cast<double>(static_cast<const uint32_t*>(data)[off]);
    case PYI_I64: return static_cast<double>(static_cast<const int64_t*>(data)[off]);
    case PYI_U64: return static_cast<double>(static_cast<const uint64_t*>(data)[off]);
    case PYI_F32: return static_cast<double>(static_cast<const float*>(data)[off]);
    default: return static_cast<const double*>(data)[off];
  }
}

template <typename ZT>
struct AnyGrid {
  const ref::math::Axis<double>* ax{};
  const ref::math::Axis<double>* ay{};
  const ref::math::Axis<ZT>* az{};
  const ref::math::Axis<double>* au{};
  const void* data{};
  int dtype{};
  int64_t n1{1}, n2{1}, n3{1};
};

template <typename A, typename C>
auto bounds_text(const A& a, int axis, const C& c) -> std::string {
  return std::format("{} is out of bounds for axis {} [{}, ..., {}]",
                     a.coordinate_repr(c), kLabels[axis],
                     a.coordinate_repr(a.min_value()),
                     a.coordinate_repr(a.max_value()));
}

template <typename ZT, size_t ND>
struct CacheOf;
template <typename ZT>
struct CacheOf<ZT, 1> {
  using type = itp::InterpolationCache<double, double>;
};
template <typename ZT>
struct CacheOf<ZT, 2> {
  using type = itp::InterpolationCache<double, double, double>;
};
template <typename ZT>
struct CacheOf<ZT, 3> {
  using type = itp::InterpolationCache<double, double, double, ZT>;
};
template <typename ZT>
struct CacheOf<ZT, 4> {
  using type = itp::InterpolationCache<double, double, double, ZT, double>;
};

using Window = std::pair<std::vector<int64_t>, std::pair<int64_t, int64_t>>;

// cache_loader.hpp:46-172
template <typename ZT, size_t ND>
auto load_window_nd(const AnyGrid<ZT>& g, double xq, double yq, ZT zq, double uq,
                    ref::math::axis::Boundary boundary, bool bounds_error,
                    typename CacheOf<ZT, ND>::type& cache,
                    std::optional<std::string>& err, int& err_axis) -> bool {
  std::array<Window, ND> w;
  auto find = [&](auto dim, const auto& axis, const auto& val) -> bool {
    constexpr size_t I = decltype(dim)::value;
    auto r = axis.find_indexes(val, cache.template half_window<I>(), boundary);
    if (!r) {
      if (bounds_error && !axis.contains(val)) {
        err = bounds_text(axis, static_cast<int>(I), val);
        err_axis = static_cast<int>(I);
      }
      return false;
    }
    w[I] = std::move(*r);
    return true;
  };
  if (!find(std::integral_constant<size_t, 0>{}, *g.ax, xq)) return false;
  if constexpr (ND >= 2) {
    if (!find(std::integral_constant<size_t, 1>{}, *g.ay, yq)) return false;
  }
  if constexpr (ND >= 3) {
    if (!find(std::integral_constant<size_t, 2>{}, *g.az, zq)) return false;
  }
  if constexpr (ND >= 4) {
    if (!find(std::integral_constant<size_t, 3>{}, *g.au, uq)) return false;
  }
  std::array<size_t, ND> npts;
  bool resized = false;
  for (size_t i = 0; i < ND; ++i) {
    npts[i] = w[i].first.size();
    resized = resized || npts[i] != cache.points_per_dim(i);
  }
  if (resized) cache.resize(std::move(npts));
  const double period = g.ax->is_periodic() ? g.ax->period().value() : 0.0;
  for (size_t k = 0; k < w[0].first.size(); ++k) {
    auto raw = (*g.ax)(w[0].first[k]);
    if (period != 0.0) raw = ref::math::normalize_period_half(raw, xq, period);
    cache.template set_coord<0>(k, raw);
  }
  if constexpr (ND >= 2) {
    for (size_t k = 0; k < w[1].first.size(); ++k)
      cache.template set_coord<1>(k, (*g.ay)(w[1].first[k]));
  }
  if constexpr (ND >= 3) {
    for (size_t k = 0; k < w[2].first.size(); ++k)
      cache.template set_coord<2>(k, (*g.az)(w[2].first[k]));
  }
  if constexpr (ND >= 4) {
    for (size_t k = 0; k < w[3].first.size(); ++k)
      cache.template set_coord<3>(k, (*g.au)(w[3].first[k]));
  }
  auto& flat = cache.values_flat();
  size_t pos = 0;
  const std::vector<int64_t> one{0};
  const auto& wy = ND >= 2 ? w[ND >= 2 ? 1 : 0].first : one;
  const auto& wz = ND >= 3 ? w[ND >= 3 ? 2 : 0].first : one;
  const auto& wu = ND >= 4 ? w[ND >= 4 ? 3 : 0].first : one;
  for (int64_t i : w[0].first)
    for (int64_t j : wy)
      for (int64_t k : wz)
        for (int64_t l : wu)
          flat[pos++] = fetch_f64(g.data, g.dtype, ((i * g.n1 + j) * g.n2 + k) * g.n3 + l);
  std::array<std::pair<int64_t, int64_t>, ND> brackets;
  for (size_t i = 0; i < ND; ++i) brackets[i] = w[i].second;
  cache.finalize(brackets);
  return cache.is_valid();
}

auto to_univariate_method(int m) -> itp::univariate::Method {
  using M = itp::univariate::Method;
  switch (m) {
    case PYI_S_AKIMA: return M::kAkima;
    case PYI_S_AKIMA_PERIODIC: return M::kAkimaPeriodic;
    case PYI_S_C_SPLINE: return M::kCSpline;
    case PYI_S_C_SPLINE_NOT_A_KNOT: return M::kCSplineNotAKnot;
    case PYI_S_C_SPLINE_PERIODIC: return M::kCSplinePeriodic;
    case PYI_S_LINEAR: return M::kLinear;
    case PYI_S_STEFFEN: return M::kSteffen;
    default: return M::kPolynomial;
  }
}

// config::windowed::Spatial::factory (config/windowed.hpp:118-123)
auto make_spatial(int method) -> std::unique_ptr<itp::BivariateBase<double>> {
  if (method == PYI_S_BICUBIC) return itp::factory<double>(itp::bivariate::Method::kBicubic);
  if (method == PYI_S_BILINEAR) return itp::factory<double>(itp::bivariate::Method::kBilinear);
  return itp::factory<double>(to_univariate_method(method));
}

template <typename ZT, size_t ND, typename R>
struct PointState {
  using Cache = typename CacheOf<ZT, ND>::type;
  Cache cache;
  std::unique_ptr<itp::Univariate<double>> uni;
  std::array<std::unique_ptr<itp::BivariateBase<double>>, 4> spatial;
  explicit PointState(const pyi_windowed_nd_config* cfg)
      : cache(static_cast<size_t>(cfg->half_x),
              static_cast<size_t>(ND >= 2 ? cfg->half_y : 1)) {
    if constexpr (ND == 1) {
      uni = itp::univariate::factory<double>(to_univariate_method(cfg->method));
    } else {
      for (size_t k = 0; k < (size_t{1} << (ND - 2)); ++k) spatial[k] = make_spatial(cfg->method);
    }
  }
};

template <typename ZT, size_t ND, typename R>
auto window_point_nd(const AnyGrid<ZT>& g, double x, double y, ZT z, double u,
                     const pyi_windowed_nd_config* cfg, PointState<ZT, ND, R>& st)
    -> std::optional<R> {
  auto& cache = st.cache;
  bool success = true;
  bool was_updated = false;
  std::optional<std::string> err;
  int err_axis = -1;
  bool inside = cache.has_domain();
  if (inside) {
    if constexpr (ND == 1) inside = cache.contains(x);
    if constexpr (ND == 2) inside = cache.contains(x, y);
    if constexpr (ND == 3) inside = cache.contains(x, y, z);
    if constexpr (ND == 4) inside = cache.contains(x, y, z, u);
  }
  if (!inside) {
    was_updated = true;
    success = load_window_nd<ZT, ND>(g, x, y, z, u, to_ref_boundary(cfg->boundary),
                                     cfg->bounds_error != 0, cache, err, err_axis);
  }
  if (!success) {
    if (err.has_value()) throw WinError(err_axis, *err);
    return std::nullopt;
  }
  if (!cache.is_valid()) return std::nullopt;
  if constexpr (ND == 1) {
    const auto xa = cache.template coords_as_eigen<0>();
    const double r = cfg->derivative ? st.uni->derivative(xa, cache.vector(), x)
                                     : (*st.uni)(xa, cache.vector(), x);
    return static_cast<R>(r);
  } else {
    const auto xa = cache.template coords_as_eigen<0>();
    const auto ya = cache.template coords_as_eigen<1>();
    const bool zlin = cfg->z_method != PYI_AXIS_NEAREST;
    const bool ulin = cfg->u_method != PYI_AXIS_NEAREST;
    if constexpr (ND == 2) {
      if (was_updated) st.spatial[0]->prepare(xa, ya, cache.matrix());
      return static_cast<R>((*st.spatial[0])(x, y));
    } else if constexpr (ND == 3) {
      if (was_updated) {
        st.spatial[0]->prepare(xa, ya, cache.matrix(0));
        st.spatial[1]->prepare(xa, ya, cache.matrix(1));
      }
      const auto z0 = cache.template coord<2>(0);
      const auto z1 = cache.template coord<2>(1);
      const auto f0 = (*st.spatial[0])(x, y);
      const auto f1 = (*st.spatial[1])(x, y);
      return static_cast<R>(zlin ? itp::linear(z, z0, z1, f0, f1)
                                 : itp::nearest(z, z0, z1, f0, f1));
    } else {
      if (was_updated) {
        st.spatial[0]->prepare(xa, ya, cache.matrix(0, 0));
        st.spatial[1]->prepare(xa, ya, cache.matrix(0, 1));
        st.spatial[2]->prepare(xa, ya, cache.matrix(1, 0));
        st.spatial[3]->prepare(xa, ya, cache.matrix(1, 1));
      }
      const auto z0 = cache.template coord<2>(0);
      const auto z1 = cache.template coord<2>(1);
      const auto u0 = cache.template coord<3>(0);
      const auto u1 = cache.template coord<3>(1);
      const auto f00 = (*st.spatial[0])(x, y);
      const auto f01 = (*st.spatial[1])(x, y);
      const auto f10 = (*st.spatial[2])(x, y);
      const auto f11 = (*st.spatial[3])(x, y);
      R f0;
      R f1;
      if (zlin) {
        f0 = static_cast<R>(itp::linear(z, z0, z1, f00, f10));
        f1 = static_cast<R>(itp::linear(z, z0, z1, f01, f11));
      } else {
        f0 = static_cast<R>(itp::nearest(z, z0, z1, f00, f10));
        f1 = static_cast<R>(itp::nearest(z, z0, z1, f01, f11));
      }
      return static_cast<R>(ulin ? itp::linear(u, u0, u1, f0, f1)
                                 : itp::nearest(u, u0, u1, f0, f1));
    }
  }
}

template <typename ZT, size_t ND, typename R>
auto windowed_nd_typed(const AnyGrid<ZT>& g, const pyi_windowed_nd_config* cfg,
                       int stateless, int64_t n, const double* x, const double* y,
                       const ZT* z, const double* u, R* o, const OobOut& oob) -> int {
  std::fill(o, o + n, std::numeric_limits<R>::quiet_NaN());
  try {
    ref::parallel_for(
        n,
        [&](const int64_t start, const int64_t end) {
          auto st = std::make_unique<PointState<ZT, ND, R>>(cfg);
          for (int64_t k = start; k < end; ++k) {
            if (stateless) st = std::make_unique<PointState<ZT, ND, R>>(cfg);
            try {
              auto r = window_point_nd<ZT, ND, R>(g, x[k], ND >= 2 ? y[k] : 0.0,
                                                  ND >= 3 ? z[k] : ZT{}, ND >= 4 ? u[k] : 0.0,
                                                  cfg, *st);
              if (r.has_value()) o[k] = *r;
            } catch (WinError& e) {
              e.index = k;
              throw;
            }
          }
        },
        cfg->num_threads);
  } catch (const WinError& e) {
    if (oob.index) *oob.index = e.index;
    if (oob.axis) *oob.axis = e.axis;
    set_err(oob.msg, oob.msglen, e.what());
    return PYI_OOB;
  }
  return PYI_OK;
}

template <typename ZT, size_t ND>
auto windowed_nd_z(const orc_axis* const* axes, int dtype, const void* data,
                   const pyi_windowed_nd_config* cfg, int stateless, int64_t n,
                   const double* x, const double* y, const void* z, const double* u,
                   void* out, const OobOut& oob) -> int {
  AnyGrid<ZT> g;
  g.ax = axes[0]->f.get();
  g.data = data;
  g.dtype = dtype;
  if constexpr (ND >= 2) {
    g.ay = axes[1]->f.get();
    g.n1 = g.ay->size();
  }
  if constexpr (ND >= 3) {
    if constexpr (std::is_same_v<ZT, int64_t>) {
      g.az = axes[2]->i.get();
    } else {
      g.az = axes[2]->f.get();
    }
    g.n2 = g.az->size();
  }
  if constexpr (ND >= 4) {
    g.au = axes[3]->f.get();
    g.n3 = g.au->size();
  }
  const bool r64 = dtype == PYI_F64 || dtype == PYI_I64 || dtype == PYI_U64;
  if (r64) {
    return windowed_nd_typed<ZT, ND, double>(g, cfg, stateless, n, x, y,
                                             static_cast<const ZT*>(z), u,
                                             static_cast<double*>(out), oob);
  }
  return windowed_nd_typed<ZT, ND, float>(g, cfg, stateless, n, x, y,
                                          static_cast<const ZT*>(z), u,
                                          static_cast<float*>(out), oob);
}

}  // namespace

extern "C" int orc_windowed_nd(int ndim, const orc_axis* const* axes, int dtype,
                               const void* data, const pyi_windowed_nd_config* cfg,
                               int stateless, int64_t n, const double* x,
                               const double* y, const void* z, const double* u,
                               void* out, int64_t* oob_index, int* oob_axis,
                               char* msg, size_t msglen) {
  const OobOut oob{oob_index, oob_axis, msg, msglen};
  if (cfg->half_x < 1 || (ndim >= 2 && cfg->half_y < 1)) return PYI_BAD_ARG;
  if (cfg->method < PYI_S_BICUBIC || cfg->method > PYI_S_POLYNOMIAL) return PYI_BAD_ARG;
  if (ndim == 1 && cfg->method < PYI_S_AKIMA) return PYI_BAD_ARG;
  if (dtype < PYI_I8 || dtype > PYI_F64) return PYI_BAD_ARG;
  const bool zi = ndim >= 3 && axes[2]->coord_type == PYI_COORD_I64;
#define ORC_ND(ZT, ND) \
  windowed_nd_z<ZT, ND>(axes, dtype, data, cfg, stateless, n, x, y, z, u, out, oob)
  switch (ndim) {
    case 1: return ORC_ND(double, 1);
    case 2: return ORC_ND(double, 2);
    case 3: return zi ? ORC_ND(int64_t, 3) : ORC_ND(double, 3);
    case 4: return zi ? ORC_ND(int64_t, 4) : ORC_ND(double, 4);
    default: return PYI_BAD_ARG;
  }
#undef ORC_ND
}

extern "C" double orc_univariate(int method, const double* xa, const double* ya,
                                 int n, double x, int derivative) {
  auto f = itp::univariate::factory<double>(to_univariate_method(method));
  ref::Vector<double> vx = Eigen::Map<const ref::Vector<double>>(xa, n);
  ref::Vector<double> vy = Eigen::Map<const ref::Vector<double>>(ya, n);
  return derivative ? f->derivative(vx, vy, x) : (*f)(vx, vy, x);
}

// ---------------------------------------------------------------------------
// dateutils::convert (dateutils.hpp:821-935), the reference's own function
// ---------------------------------------------------------------------------
#include "pyinterp/dateutils.hpp"

extern "C" int orc_datetime_convert(int src, int dst, int64_t n,
                                    const int64_t* in, int64_t* out, char* msg,
                                    size_t msglen) {
  static const char* units[] = {"Y",  "M",  "W",  "D",  "h",  "m", "s",
                                "ms", "us", "ns", "ps", "fs", "as"};
  if (src < 0 || src > 12 || dst < 0 || dst > 12) return PYI_BAD_ARG;
  try {
    const ref::dateutils::DType s(std::string("datetime64[") + units[src] + "]");
    const ref::dateutils::DType d(std::string("datetime64[") + units[dst] + "]");
    ref::Vector<int64_t> v = Eigen::Map<const ref::Vector<int64_t>>(in, n);
    ref::dateutils::convert(v, s, d);
    std::copy(v.data(), v.data() + n, out);
  } catch (const std::overflow_error& e) {
    set_err(msg, msglen, e.what());
    return PYI_OVERFLOW;
  } catch (const std::exception& e) {
    set_err(msg, msglen, e.what());
    return PYI_BAD_ARG;
  }
  return PYI_OK;
}

// TemporalAxis::coordinate_repr (math/temporal_axis.hpp:80-83,150-156): the
// reference's own dateutils::datetime64_to_string / timedelta64_to_string
// (dateutils.hpp:729-812), which is all format_temporal forwards to.
extern "C" int orc_temporal_repr(int is_timedelta, int unit, int64_t value, char* buf,
                                 size_t len) {
  static const char* units[] = {"Y",  "M",  "W",  "D",  "h",  "m", "s",
                                "ms", "us", "ns", "ps", "fs", "as"};
  if (unit < 0 || unit > 12 || !buf || len == 0) return -1;
  try {
    const ref::dateutils::DType d(std::string(is_timedelta ? "timedelta64[" : "datetime64[") +
                                  units[unit] + "]");
    const std::string s = is_timedelta ? ref::dateutils::timedelta64_to_string(value, d)
                                       : ref::dateutils::datetime64_to_string(value, d);
    set_err(buf, len, s.c_str());
    return static_cast<int>(s.size());
  } catch (const std::exception&) {
    return -1;
  }
}

// ---------------------------------------------------------------------------
// fill::gauss_seidel / fill::multigrid (fill/gauss_seidel.hpp:166-211,
// fill/multi_grid.hpp:276-334), the reference's own functions, in place on a
// row-major (nx, ny) grid.  config/fill.hpp and Boost.Accumulators come from
// ref/shim (see the headers there for why).
// ---------------------------------------------------------------------------
#include "pyinterp/fill/gauss_seidel.hpp"
#include "pyinterp/fill/multi_grid.hpp"

namespace {

template <typename T, typename Config, typename Fn>
int fill_run(void* grid, int64_t nx, int64_t ny, const Config& cfg, Fn&& fn,
             uint64_t* iterations, double* max_residual) {
  Eigen::Map<ref::RowMajorMatrix<T>> m(static_cast<T*>(grid), nx, ny);
  auto [it, res] = fn(ref::EigenDRef<ref::RowMajorMatrix<T>>(m), cfg);
  *iterations = it;
  *max_residual = static_cast<double>(res);
  return PYI_OK;
}

template <typename Config>
void fill_base(Config& c, int first_guess, int is_periodic,
               uint64_t max_iterations, double epsilon, int num_threads) {
  c.first_guess_ = static_cast<ref::config::fill::FirstGuess>(first_guess);
  c.is_periodic_ = is_periodic != 0;
  c.max_iterations_ = static_cast<uint32_t>(max_iterations);
  c.epsilon_ = epsilon;
  c.num_threads_ = num_threads <= 0 ? std::thread::hardware_concurrency()
                                    : static_cast<size_t>(num_threads);
}

}  // namespace

extern "C" int orc_fill_gauss_seidel(void* grid, int dtype, int64_t nx,
                                     int64_t ny, int first_guess,
                                     int is_periodic, uint64_t max_iterations,
                                     double epsilon, double relaxation,
                                     int num_threads, uint64_t* iterations,
                                     double* max_residual, char* msg,
                                     size_t msglen) {
  try {
    ref::config::fill::GaussSeidel cfg;
    fill_base(cfg, first_guess, is_periodic, max_iterations, epsilon,
              num_threads);
    cfg.relaxation_ = relaxation;
    if (dtype == PYI_F32) {
      return fill_run<float>(
          grid, nx, ny, cfg,
          [](auto g, const auto& c) { return ref::fill::gauss_seidel<float>(g, c); },
          iterations, max_residual);
    }
    if (dtype == PYI_F64) {
      return fill_run<double>(
          grid, nx, ny, cfg,
          [](auto g, const auto& c) { return ref::fill::gauss_seidel<double>(g, c); },
          iterations, max_residual);
    }
    set_err(msg, msglen, "grid must be float32 or float64");
    return PYI_BAD_ARG;
  } catch (const std::exception& e) {
    set_err(msg, msglen, e.what());
    return PYI_BAD_ARG;
  }
}

extern "C" int orc_fill_multigrid(void* grid, int dtype, int64_t nx, int64_t ny,
                                  int first_guess, int is_periodic,
                                  uint64_t max_iterations, double epsilon,
                                  uint32_t pre_smooth, uint32_t post_smooth,
                                  int num_threads, uint64_t* iterations,
                                  double* max_residual, char* msg,
                                  size_t msglen) {
  try {
    ref::config::fill::Multigrid cfg;
    fill_base(cfg, first_guess, is_periodic, max_iterations, epsilon,
              num_threads);
    cfg.pre_smooth_ = pre_smooth;
    cfg.post_smooth_ = post_smooth;
    if (dtype == PYI_F32) {
      return fill_run<float>(
          grid, nx, ny, cfg,
          [](auto g, const auto& c) { return ref::fill::multigrid<float>(g, c); },
          iterations, max_residual);
    }
    if (dtype == PYI_F64) {
      return fill_run<double>(
          grid, nx, ny, cfg,
          [](auto g, const auto& c) { return ref::fill::multigrid<double>(g, c); },
          iterations, max_residual);
    }
    set_err(msg, msglen, "grid must be float32 or float64");
    return PYI_BAD_ARG;
  } catch (const std::exception& e) {
    set_err(msg, msglen, e.what());
    return PYI_BAD_ARG;
  }
}

// ---------------------------------------------------------------------------
// fill::loess (fill/loess.hpp:207-253), the reference's own function: `in` is
// a row-major (nx, ny) grid, `out` receives the processed copy.  half_x /
// half_y are the half-window sizes (config nx / ny), value_type is
// LoessValueType (0 undefined, 1 defined, 2 all).
// ---------------------------------------------------------------------------
#include "pyinterp/fill/loess.hpp"

extern "C" int orc_fill_loess(const void* in, void* out, int dtype, int64_t nx, int64_t ny,
                              uint32_t half_x, uint32_t half_y, int value_type, int first_guess,
                              int is_periodic, uint64_t max_iterations, double epsilon,
                              int num_threads, char* msg, size_t msglen) {
  try {
    ref::config::fill::Loess cfg;
    fill_base(cfg, first_guess, is_periodic, max_iterations, epsilon, num_threads);
    cfg.nx_ = half_x;
    cfg.ny_ = half_y;
    cfg.value_type_ = static_cast<ref::config::fill::LoessValueType>(value_type);
    auto run = [&](auto tag) {
      using T = decltype(tag);
      Eigen::Map<const ref::RowMajorMatrix<T>> m(static_cast<const T*>(in), nx, ny);
      const ref::RowMajorMatrix<T> res =
          ref::fill::loess<T>(ref::EigenDRef<const ref::RowMajorMatrix<T>>(m), cfg);
      std::copy(res.data(), res.data() + nx * ny, static_cast<T*>(out));
    };
    if (dtype == PYI_F32) {
      run(float{});
    } else if (dtype == PYI_F64) {
      run(double{});
    } else {
      set_err(msg, msglen, "grid must be float32 or float64");
      return PYI_BAD_ARG;
    }
    return PYI_OK;
  } catch (const std::exception& e) {
    set_err(msg, msglen, e.what());
    return PYI_BAD_ARG;
  }
}
