/*
 * orc_abi.h -- C-ABI shared by the two CPU checkers of this repo.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may load these libraries, and only as the checker / the
 * timed CPU baseline.  The product (pangeo-pyinterp_b200/) never links,
 * imports or falls back to anything here.
 *
 * Two libraries export exactly this interface:
 *   oracle/liboracle.so            plain-C restatement of the reference algorithm
 *                                  (oracle/pyi_oracle.c), travels with the repo;
 *   oracle/_ref/liboracle_ref.so   the reference's OWN headers compiled in place
 *                                  from /root/reference (oracle/ref/ref_driver.cpp),
 *                                  used to pin the restatement and as the
 *                                  "reference" CPU baseline.
 *
 * The PODs (axis descriptor, configs, enums) are those of the product header so
 * that one parametrised test drives the oracle, the reference build and the
 * CUDA library with identical arguments.
 */
#ifndef ORC_ABI_H_
#define ORC_ABI_H_

#include <stddef.h>
#include <stdint.h>

#include "../include/pyinterp_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Boundary modes of math/axis.hpp:32-38 beyond the two the windowed config
 * exposes (PYI_BOUNDARY_UNDEF = 0, PYI_BOUNDARY_SHRINK = 1). */
#define ORC_BOUNDARY_EXPAND 2
#define ORC_BOUNDARY_WRAP 3
#define ORC_BOUNDARY_SYM 4

typedef struct orc_axis orc_axis;

/* "reference" for _ref/liboracle_ref.so, "port" for liboracle.so */
const char *orc_kind(void);

/* math::Axis<T>(values, epsilon, period)  math/axis.hpp:683-716.  NULL on
 * failure with the exception text in err. */
orc_axis *orc_axis_new_f64(const double *values, int64_t n, double epsilon,
                           int has_period, double period, char *err,
                           size_t errlen);
orc_axis *orc_axis_new_i64(const int64_t *values, int64_t n, int64_t epsilon,
                           int has_period, int64_t period, char *err,
                           size_t errlen);
void orc_axis_free(orc_axis *axis);
void orc_axis_flip(orc_axis *axis);
/* desc->points aliases memory owned by the axis (irregular only). */
int orc_axis_describe(const orc_axis *axis, pyi_axis_desc *desc);

/* Axis::find_indexes(T)  math/axis.hpp:720-778; (-1,-1) when nullopt. */
void orc_axis_find_indexes(const orc_axis *axis, int64_t n, const void *coords,
                           int64_t *out_pairs);
/* Axis::find_index(T, bounded)  math/axis.hpp:319-333. */
void orc_axis_find_index(const orc_axis *axis, int64_t n, const void *coords,
                         int bounded, int64_t *out);
/* Axis::find_indexes(T, half, boundary)  math/axis.hpp:814-881.
 * Returns 1 and fills idx[0..*count) and the centre pair, 0 for nullopt. */
int orc_axis_find_window(const orc_axis *axis, const void *coord, int half,
                         int boundary, int64_t *idx, int *count, int *c0,
                         int *c1);

/* geometric::pybind::{bi,tri,quadri}variate
 * (pybind/geometric/bivariate.hpp:127-161, trivariate.hpp:136-181,
 * quadrivariate.hpp:162-218) with Point = SphericalPoint and the reference's
 * parallel_for (cfg->num_threads, 0 = hardware_concurrency).
 * Returns PYI_OK, or PYI_OOB with the lowest offending point, its first
 * offending axis and the reference message text. */
int orc_geometric(int ndim, const orc_axis *const *axes, int dtype,
                  const void *data, const pyi_geometric_config *cfg, int64_t n,
                  const double *x, const double *y, const void *z,
                  const double *u, void *out, int64_t *oob_index,
                  int *oob_axis, char *msg, size_t msglen);

/* windowed::pybind::bivariate (pybind/windowed/bivariate.hpp:86-115).
 * stateless != 0: a fresh cache per point (what the GPU computes);
 * stateless == 0: the reference's per-thread cache with its "still inside the
 * central cell" fast path (history dependent). */
int orc_windowed2d(const orc_axis *const *axes, int dtype, const void *data,
                   const pyi_windowed_config *cfg, int stateless, int64_t n,
                   const double *x, const double *y, void *out,
                   int64_t *oob_index, int *oob_axis, char *msg,
                   size_t msglen);

/* windowed::pybind::univariate / univariate_derivative / bivariate /
 * trivariate / quadrivariate with every fitting model
 * (pybind/windowed/{uni,bi,tri,quadri}variate.hpp, math/interpolate/factory.hpp).
 * stateless == 0 (the reference's per-thread cache) is served by the reference
 * build only; the plain-C library returns PYI_BAD_ARG for it. */
int orc_windowed_nd(int ndim, const orc_axis *const *axes, int dtype,
                    const void *data, const pyi_windowed_nd_config *cfg,
                    int stateless, int64_t n, const double *x, const double *y,
                    const void *z, const double *u, void *out,
                    int64_t *oob_index, int *oob_axis, char *msg,
                    size_t msglen);
/* One univariate model on explicit nodes: value or derivative at x
 * (math/interpolate/univariate/ headers). method: pyi_spatial_method >= 2. */
double orc_univariate(int method, const double *xa, const double *ya, int n,
                      double x, int derivative);

/* dateutils::convert (dateutils.hpp:821-935) between pyi_resolution units.
 * PYI_OVERFLOW with the reference's message when its checked arithmetic
 * throws. */
int orc_datetime_convert(int src, int dst, int64_t n, const int64_t *in,
                         int64_t *out, char *msg, size_t msglen);

/* fill::loess (fill/loess.hpp:207-253): `out` receives the processed copy of the
 * row-major (nx, ny) grid `in`.  half_x / half_y: config nx / ny; value_type:
 * LoessValueType (0 undefined, 1 defined, 2 all). */
int orc_fill_loess(const void *in, void *out, int dtype, int64_t nx, int64_t ny,
                   uint32_t half_x, uint32_t half_y, int value_type,
                   int first_guess, int is_periodic, uint64_t max_iterations,
                   double epsilon, int num_threads, char *msg, size_t msglen);

/* TemporalAxis::coordinate_repr (math/temporal_axis.hpp:80-83,150-156;
 * dateutils::datetime64_to_string / timedelta64_to_string,
 * dateutils.hpp:729-812).  Returns the length written, -1 on a bad argument. */
int orc_temporal_repr(int is_timedelta, int unit, int64_t value, char *buf,
                      size_t len);

/* fill::gauss_seidel (fill/gauss_seidel.hpp:166-211) and fill::multigrid
 * (fill/multi_grid.hpp:276-334) in place on a row-major (nx, ny) grid of dtype
 * PYI_F32 / PYI_F64.  first_guess: pyi_first_guess.  num_threads <= 0: all
 * cores (the result does not depend on it). */
int orc_fill_gauss_seidel(void *grid, int dtype, int64_t nx, int64_t ny,
                          int first_guess, int is_periodic,
                          uint64_t max_iterations, double epsilon,
                          double relaxation, int num_threads,
                          uint64_t *iterations, double *max_residual, char *msg,
                          size_t msglen);
int orc_fill_multigrid(void *grid, int dtype, int64_t nx, int64_t ny,
                       int first_guess, int is_periodic,
                       uint64_t max_iterations, double epsilon,
                       uint32_t pre_smooth, uint32_t post_smooth,
                       int num_threads, uint64_t *iterations,
                       double *max_residual, char *msg, size_t msglen);

/* ---- scalar probes used by the golden-vector tests ------------------------ */
/* geometric evaluators (math/interpolate/geometric/bivariate.hpp).
 * cartesian != 0: the Cartesian point type the reference unit tests use;
 * cartesian == 0: SphericalPoint (the production path). */
double orc_eval2d_f64(int method, int idw_exp, int cartesian, double x, double y,
                      double x0, double y0, double x1, double y1, double q00,
                      double q01, double q10, double q11);
float orc_eval2d_f32(int method, int idw_exp, int cartesian, float x, float y,
                     float x0, float y0, float x1, float y1, float q00,
                     float q01, float q10, float q11);
/* interpolate/linear.hpp:22-33, nearest.hpp:21-32 (Axis = double / int64) */
double orc_axis_interp_f64(int method, double x, double x0, double x1, double y0,
                           double y1);
double orc_axis_interp_i64(int method, int64_t x, int64_t x0, int64_t x1,
                           double y0, double y1);
float orc_axis_interp_i64_f32(int method, int64_t x, int64_t x0, int64_t x1,
                              float y0, float y1);
/* windowed evaluators on an explicit window: za is [nx][ny] row-major.
 * method: pyi_windowed_method. (bivariate/bicubic.hpp, bivariate/bilinear.hpp) */
double orc_window_eval(int method, const double *xa, int nx, const double *ya,
                       int ny, const double *za, double x, double y);
/* natural cubic spline value (derivative == 0) or first derivative at x
 * (univariate/cspline.hpp, cspline_base.hpp). */
double orc_cspline(const double *xa, const double *ya, int n, double x,
                   int derivative);

#ifdef __cplusplus
}
#endif
#endif /* ORC_ABI_H_ */
