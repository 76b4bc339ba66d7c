/*
 * pyinterp_b200.h -- C-ABI of the B200-native regular-grid interpolation path.
 *
 * This is the drop-in boundary: the entry points below are what the reference's
 * C++ host layer (the templated "vectorised drivers" under
 * cxx/include/pyinterp/pybind/{geometric,windowed}/) binds instead of running
 * its own per-point loops.  Nothing here uses C++ or torch types: plain
 * pointers, sizes and POD structs only.
 *
 * Reference interfaces replaced (paths relative to the reference root):
 *   pyi_geometric          <- geometric::pybind::bivariate      cxx/include/pyinterp/pybind/geometric/bivariate.hpp:127-161
 *                             geometric::pybind::trivariate     cxx/include/pyinterp/pybind/geometric/trivariate.hpp:136-181
 *                             geometric::pybind::quadrivariate  cxx/include/pyinterp/pybind/geometric/quadrivariate.hpp:162-218
 *   pyi_windowed2d         <- windowed::pybind::bivariate       cxx/include/pyinterp/pybind/windowed/bivariate.hpp:86-115
 *   pyi_axis_find_indexes  <- pybind Axis.find_indexes          cxx/include/pyinterp/pybind/axis.hpp:84-101
 *                             (math::Axis<T>::find_indexes      cxx/include/pyinterp/math/axis.hpp:720-778)
 *   pyi_axis_find_index    <- math::Axis<T>::find_index         cxx/include/pyinterp/math/axis.hpp:319-333
 *   pyi_axis_create_*      <- math::Axis<T> constructors        cxx/include/pyinterp/math/axis.hpp:683-716
 *   pyi_axis_create_from_desc: descriptor read from a live reference Axis
 *                             (public accessors + pack(), axis.hpp:943-972)
 *   pyi_grid_create        <- pybind::Grid<DataType, Axes...>   cxx/include/pyinterp/pybind/grid.hpp:214-226,443-481
 *   pyi_format_bounds_error<- Grid::construct_bounds_error_description  pybind/grid.hpp:362-371
 *
 * All compute entry points run hand-written sm_100a CUDA kernels.  There is no
 * CPU fallback: without a usable CUDA device they return PYI_NO_DEVICE.
 */
#ifndef PYINTERP_B200_H_
#define PYINTERP_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PYI_ABI_VERSION 2

/* ---- status codes -------------------------------------------------------- */
typedef enum pyi_status {
  PYI_OK = 0,
  PYI_OOB = 1,        /* bounds_error set and a point is out of bounds: index/axis filled */
  PYI_BAD_SHAPE = 2,  /* inconsistent sizes (reference: std::invalid_argument) */
  PYI_BAD_ARG = 3,    /* invalid argument / unsupported combination */
  PYI_NO_DEVICE = 4,  /* no CUDA device: the product path refuses to run */
  PYI_CUDA_ERROR = 5, /* a CUDA call failed, see pyi_last_error() */
  PYI_NOMEM = 6,
  PYI_OVERFLOW = 7    /* checked int64 arithmetic overflowed (reference: std::overflow_error) */
} pyi_status;

/* ---- enumerations ---------------------------------------------------------- */
typedef enum pyi_axis_kind { PYI_AXIS_REGULAR = 0, PYI_AXIS_IRREGULAR = 1 } pyi_axis_kind;
typedef enum pyi_coord_type { PYI_COORD_F64 = 0, PYI_COORD_I64 = 1 } pyi_coord_type;

/* Grid element types: the ten dtypes of pybind/grid.hpp:604-606,819-879. */
typedef enum pyi_dtype {
  PYI_I8 = 0, PYI_U8 = 1, PYI_I16 = 2, PYI_U16 = 3, PYI_I32 = 4,
  PYI_U32 = 5, PYI_I64 = 6, PYI_U64 = 7, PYI_F32 = 8, PYI_F64 = 9
} pyi_dtype;

/* math/interpolate/geometric/bivariate.hpp:305-309 */
typedef enum pyi_geometric_method { PYI_BILINEAR = 0, PYI_IDW = 1, PYI_NEAREST = 2 } pyi_geometric_method;
/* math/interpolate/geometric/multivariate.hpp:24-27 */
typedef enum pyi_axis_method { PYI_AXIS_LINEAR = 0, PYI_AXIS_NEAREST = 1 } pyi_axis_method;
/* windowed 2-D evaluators on the path: bivariate/bicubic.hpp, bivariate/bilinear.hpp */
typedef enum pyi_windowed_method { PYI_W_BICUBIC = 0, PYI_W_BILINEAR = 1 } pyi_windowed_method;
/* every fitting model of the windowed path (config/windowed.hpp:21-56,
 * math/interpolate/factory.hpp:27-36,77-80): the two bivariate evaluators keep
 * the values of pyi_windowed_method, the others are the univariate models the
 * reference applies along y then along x (bivariate/spline.hpp:62-84). */
typedef enum pyi_spatial_method {
  PYI_S_BICUBIC = 0,
  PYI_S_BILINEAR = 1,
  PYI_S_AKIMA = 2,
  PYI_S_AKIMA_PERIODIC = 3,
  PYI_S_C_SPLINE = 4,
  PYI_S_C_SPLINE_NOT_A_KNOT = 5,
  PYI_S_C_SPLINE_PERIODIC = 6,
  PYI_S_LINEAR = 7,
  PYI_S_STEFFEN = 8,
  PYI_S_POLYNOMIAL = 9
} pyi_spatial_method;
/* the two boundary modes the reference exposes, config/windowed.hpp:88-98 */
typedef enum pyi_boundary { PYI_BOUNDARY_UNDEF = 0, PYI_BOUNDARY_SHRINK = 1 } pyi_boundary;
/* numpy datetime64 / timedelta64 units, values of dateutils::DType::Resolution
 * (dateutils.hpp:186-200) */
typedef enum pyi_resolution {
  PYI_RES_YEAR = 0,
  PYI_RES_MONTH = 1,
  PYI_RES_WEEK = 2,
  PYI_RES_DAY = 3,
  PYI_RES_HOUR = 4,
  PYI_RES_MINUTE = 5,
  PYI_RES_SECOND = 6,
  PYI_RES_MILLISECOND = 7,
  PYI_RES_MICROSECOND = 8,
  PYI_RES_NANOSECOND = 9,
  PYI_RES_PICOSECOND = 10,
  PYI_RES_FEMTOSECOND = 11,
  PYI_RES_ATTOSECOND = 12
} pyi_resolution;
/* config::fill::FirstGuess (config/fill.hpp:18-23) */
typedef enum pyi_first_guess {
  PYI_FIRST_GUESS_ZONAL_AVERAGE = 0,
  PYI_FIRST_GUESS_ZERO = 1
} pyi_first_guess;
/* How interpolated VALUES are computed (index selection is bit-exact with the
 * reference in every mode).  PYI_VALUES_EXACT: every rounding of the
 * reference's evaluator is reproduced, values are bit-identical.
 * PYI_VALUES_TOLERANCE: reciprocal-multiply and fused multiply-adds in the
 * weights (bilinear), precomputed spline-derivative weights (bicubic 6 x 6):
 * values within 1e-12 (float64) / 1e-5 (float32) relative of the reference,
 * NaN positions identical.  PYI_VALUES_DEFAULT: tolerance where a tolerance
 * kernel exists, the axes are well conditioned for it and the launch is large
 * enough to pay for its extra launches (bilinear: 2^22 points), else exact; the
 * environment variables PYINTERP_B200_BILINEAR / PYINTERP_B200_BICUBIC
 * (= "exact" | "tolerance") change what DEFAULT means for the process. */
typedef enum pyi_value_mode {
  PYI_VALUES_DEFAULT = 0,
  PYI_VALUES_EXACT = 1,
  PYI_VALUES_TOLERANCE = 2
} pyi_value_mode;
/* where caller buffers live */
typedef enum pyi_mem { PYI_MEM_HOST = 0, PYI_MEM_DEVICE = 1 } pyi_mem;

/* pyi_grid_create flags */
#define PYI_GRID_DEFAULT 0u
#define PYI_GRID_PACK_CELLS 1u /* also build the cell-packed copy (2^d corners contiguous) */
#define PYI_GRID_BORROW_DEVICE 2u /* device data: keep the caller's pointer, do not copy */

/* ---- PODs ---------------------------------------------------------------- */
typedef union pyi_scalar {
  double f;
  int64_t i;
} pyi_scalar;

/*
 * Flat mirror of a reference math::Axis<T> (T = double or int64_t).
 * Every field is either readable through the reference's public accessors
 * (size/front/increment/is_regular/is_ascending/is_periodic/min_value/
 * max_value) or, for `period` when is_periodic == 0, from pack()
 * (axis.hpp:959-960,967-968 write both unconditionally).
 * `inv_step` and `step_2` are the container's cached values
 * (container.hpp:547 `1.0/step`, :585 `step >> 1`; note flip() negates
 * inv_step but leaves step_2 untouched, :567-570 / :473-477).
 */
typedef struct pyi_axis_desc {
  int32_t kind;         /* pyi_axis_kind */
  int32_t coord_type;   /* pyi_coord_type */
  int64_t size;
  int32_t is_ascending;
  int32_t is_periodic;  /* is_periodic_ after compute_properties() */
  pyi_scalar period;    /* period_ (may be non-zero with is_periodic == 0) */
  pyi_scalar start;     /* regular: start_ ; irregular: points[0] */
  pyi_scalar step;      /* regular only */
  pyi_scalar min_value;
  pyi_scalar max_value;
  double inv_step;      /* regular f64 only */
  int64_t step_2;       /* regular i64 only */
  const void *points;   /* irregular only: `size` host values of coord_type */
} pyi_axis_desc;

/* config/geometric.hpp: spatial method, idw exponent, third/fourth axis
 * methods; config/common.hpp: bounds_error, num_threads. */
typedef struct pyi_geometric_config {
  int32_t method;       /* pyi_geometric_method */
  int32_t idw_exp;      /* default 2 */
  int32_t z_method;     /* pyi_axis_method (3-D / 4-D) */
  int32_t u_method;     /* pyi_axis_method (4-D) */
  int32_t bounds_error; /* 0: NaN, 1: report first out-of-bounds point */
  int32_t num_threads;  /* accepted for API parity, ignored on the GPU */
  int32_t value_mode;   /* pyi_value_mode */
} pyi_geometric_config;

/* config/windowed.hpp: method, half window sizes (default 3/3, :244-247),
 * boundary mode. */
typedef struct pyi_windowed_config {
  int32_t method;       /* pyi_windowed_method */
  int32_t half_x;
  int32_t half_y;
  int32_t boundary;     /* pyi_boundary */
  int32_t bounds_error;
  int32_t num_threads;
  int32_t value_mode;   /* pyi_value_mode */
} pyi_windowed_config;

/* config/windowed.hpp Univariate / Bivariate / Trivariate / Quadrivariate in
 * one POD: spatial model and window, third / fourth axis methods
 * (config/windowed.hpp:560-640), and the value-or-derivative switch of the 1-D
 * drivers (pybind/windowed/univariate.hpp:47-120). */
typedef struct pyi_windowed_nd_config {
  int32_t method;       /* pyi_spatial_method; 1-D grids: univariate models only */
  int32_t half_x;       /* half_window_size (1-D) / half_window_size_x */
  int32_t half_y;
  int32_t boundary;     /* pyi_boundary */
  int32_t bounds_error;
  int32_t num_threads;
  int32_t z_method;     /* pyi_axis_method (3-D / 4-D) */
  int32_t u_method;     /* pyi_axis_method (4-D) */
  int32_t derivative;   /* 1-D only: 0 value, 1 first derivative */
  int32_t value_mode;   /* pyi_value_mode */
} pyi_windowed_nd_config;

typedef struct pyi_axis pyi_axis; /* opaque */
typedef struct pyi_grid pyi_grid; /* opaque */

/* ---- library ------------------------------------------------------------- */
int pyi_abi_version(void);
/* Number of usable CUDA devices (0 when there is none). */
int pyi_device_count(void);
/* Bind the calling thread to `device`. PYI_NO_DEVICE when there is none. */
int pyi_set_device(int device);
/* Single-call multi-GPU sharding.  The reference's caller is ONE process whose
 * parallel_for (cxx/include/pyinterp/parallel_for.hpp:30-78) splits the points
 * of a call into contiguous chunks, one per worker.  After pyi_dist_init(n)
 * every HOST-memory call of pyi_geometric / pyi_windowed2d / pyi_windowed
 * splits its points the same way (n / G each, the first n % G shards one more)
 * over the first G = min(n, device count) CUDA devices (n <= 0: all of them):
 * the grid handle is replicated on each device on first use (peer copies over
 * NVLink, refreshed after pyi_grid_update), every shard runs its own copy /
 * kernel pipeline from its own host thread, results land in the caller's
 * output array, and the bounds-error report is the minimum over the shards,
 * i.e. the same point as on one device.  Small calls (fewer than 2^21 points
 * per device) use fewer devices.  Returns the device count in effect (0
 * without a CUDA device).  pyi_dist_size() reads it back. */
int pyi_dist_init(int n_devices);
int pyi_dist_size(void);
/* Thread-local text of the last failure. */
const char *pyi_last_error(void);

/* ---- axes (host side, no GPU needed) --------------------------------------- */
pyi_axis *pyi_axis_create_f64(const double *values, int64_t n, double epsilon,
                              int has_period, double period, int *status);
pyi_axis *pyi_axis_create_i64(const int64_t *values, int64_t n, int64_t epsilon,
                              int has_period, int64_t period, int *status);
/* Build from a descriptor read off a live reference Axis (drop-in mode). */
pyi_axis *pyi_axis_create_from_desc(const pyi_axis_desc *desc, int *status);
/* Reverse the axis in place (Axis::flip, axis.hpp:250). */
int pyi_axis_flip(pyi_axis *axis);
/* Fill `out`; out->points aliases storage owned by `axis`. */
int pyi_axis_describe(const pyi_axis *axis, pyi_axis_desc *out);
void pyi_axis_destroy(pyi_axis *axis);

/* ---- axis lookups (GPU) ---------------------------------------------------- */
/* coords: n values of the axis' coord type. out_pairs: n x 2 int64, (-1,-1)
 * when the coordinate cannot be bracketed. */
int pyi_axis_find_indexes(const pyi_axis *axis, int64_t n, const void *coords,
                          int64_t *out_pairs, int mem, void *stream);
/* out: n int64, -1 when outside and !bounded. */
int pyi_axis_find_index(const pyi_axis *axis, int64_t n, const void *coords,
                        int bounded, int64_t *out, int mem, void *stream);

/* ---- grids ----------------------------------------------------------------- */
/* data: C-contiguous array of `dtype`, shape[i] == axes[i]->size.
 * Periodic axes are only legal on axis 0; an int64 axis only on axis 2
 * (pybind/grid.hpp:450-464, src/pybind/grid.cpp:68-92). The grid copies the
 * axes; the caller may destroy them afterwards. */
pyi_grid *pyi_grid_create(int ndim, const pyi_axis *const *axes, int dtype,
                          const void *data, const int64_t *shape, int mem,
                          unsigned flags, int *status);
/* Re-upload the array (the reference borrows the caller's numpy buffer, which
 * may be mutated between calls). Rebuilds the packed copy if there is one. */
int pyi_grid_update(pyi_grid *grid, const void *data, int mem);
/* Build (or rebuild) the cell-packed copy (2^ndim corners per cell, result
 * type). Skipped, with PYI_OK, when it would take more than half of the free
 * device memory: the kernels then gather from the plain C-order array. */
int pyi_grid_pack(pyi_grid *grid);
/* The reference's Grid borrows the caller's numpy buffer (pybind/grid.hpp:214-226), which may be
 * written between calls (pyinterp/tests/core/windowed/test_bivariate.py:234-250).  Instead of
 * pyi_grid_update before every call, a binding can hand the buffer over once: `host_data` (the
 * grid's bytes, C order, alive and at the same address until pyi_grid_destroy) is page-locked in
 * place.  pyi_grid_mark_stale then says "the buffer may have changed": the next compute call either
 * reads it in place over PCIe (few points against a large array: no upload at all) or refreshes the
 * device copies first, at link speed.  Failure to page-lock only costs speed. */
int pyi_grid_track_host(pyi_grid *grid, const void *host_data);
int pyi_grid_mark_stale(pyi_grid *grid);
void pyi_grid_destroy(pyi_grid *grid);

/* ---- the hot path ---------------------------------------------------------- */
/*
 * x, y, u: n float64. z: n float64, or n int64 when axis 2 is an int64
 * (temporal) axis. out: n float64 for f64/i64/u64 grids, n float32 otherwise
 * (pybind/grid_dispatch.hpp:25-30). All n outputs are always written (NaN
 * where a point cannot be bracketed). With cfg->bounds_error the call returns
 * PYI_OOB and fills oob_index (lowest offending point) and oob_axis (first
 * offending axis of that point in x,y,z,u order).
 * mem == PYI_MEM_HOST: buffers are host memory, copies are staged and
 * pipelined internally.  mem == PYI_MEM_DEVICE: buffers are device memory and
 * the call only enqueues work on `stream` (a cudaStream_t, may be NULL); with
 * bounds_error it synchronises the stream to read the flag.
 */
int pyi_geometric(const pyi_grid *grid, const pyi_geometric_config *cfg,
                  int64_t n, const double *x, const double *y, const void *z,
                  const double *u, void *out, int mem, int64_t *oob_index,
                  int *oob_axis, void *stream);

/* Windowed 2-D interpolation (bicubic / windowed bilinear). With
 * bounds_error, PYI_OOB is reported only for points truly outside an axis
 * (cache_loader.hpp:69-80), mapped by the host to std::out_of_range. */
int pyi_windowed2d(const pyi_grid *grid, const pyi_windowed_config *cfg,
                   int64_t n, const double *x, const double *y, void *out,
                   int mem, int64_t *oob_index, int *oob_axis, void *stream);

/* Windowed interpolation on 1-D .. 4-D grids with every fitting model:
 * windowed::pybind::univariate / univariate_derivative / bivariate /
 * trivariate / quadrivariate (pybind/windowed/univariate.hpp:122-190,
 * bivariate.hpp:86-115, trivariate.hpp:136-165, quadrivariate.hpp:162-197).
 * y is unused for 1-D grids, z for < 3-D, u for < 4-D.  z is int64 when axis 2
 * is temporal.  Errors and memory spaces as pyi_windowed2d. */
int pyi_windowed(const pyi_grid *grid, const pyi_windowed_nd_config *cfg,
                 int64_t n, const double *x, const double *y, const void *z,
                 const double *u, void *out, int mem, int64_t *oob_index,
                 int *oob_axis, void *stream);

/* dateutils::convert (dateutils.hpp:821-935): n int64 datetime64 / timedelta64
 * values from resolution `src` to resolution `dst` (pyi_resolution), the
 * pre-step of TemporalAxis::cast_to_int64 (src/pybind/temporal_axis.cpp:78-98,
 * 216-220).  in == out is allowed.  Returns PYI_OVERFLOW when the reference's
 * checked arithmetic would throw: overflow_index = lowest offending value,
 * overflow_op = 1 multiplication / 2 addition (the reference's two messages). */
int pyi_datetime_convert(int src, int dst, int64_t n, const int64_t *in,
                         int64_t *out, int mem, int64_t *overflow_index,
                         int *overflow_op, void *stream);

/* config::fill::GaussSeidel / Multigrid (config/fill.hpp:67-160,255-324),
 * without num_threads (the device result does not depend on it, as the
 * reference's) */
typedef struct pyi_fill_config {
  int32_t first_guess;      /* pyi_first_guess */
  int32_t is_periodic;      /* x axis periodic (longitude) */
  uint64_t max_iterations;
  double epsilon;
  double relaxation;        /* Gauss-Seidel only */
  uint32_t pre_smooth;      /* multigrid only */
  uint32_t post_smooth;     /* multigrid only */
} pyi_fill_config;

/* fill::gauss_seidel (fill/gauss_seidel.hpp:166-211): the NaN cells of the
 * row-major (nx, ny) grid (dtype PYI_F32 or PYI_F64) are replaced IN PLACE by
 * the red-black relaxed solution of the Laplace equation.  iterations /
 * max_residual are the reference's returned tuple (max_residual is exact for
 * both dtypes).  Results are bit-identical to the reference whatever its
 * thread count. */
int pyi_fill_gauss_seidel(void *grid, int dtype, int64_t nx, int64_t ny,
                          const pyi_fill_config *cfg, int mem,
                          uint64_t *iterations, double *max_residual,
                          void *stream);

/* fill::multigrid (fill/multi_grid.hpp:276-334): V-cycles (red-black
 * Gauss-Seidel smoothing, 2 x 2 restriction, bilinear prolongation, 50
 * iterations on the coarsest grid) until max |residual| < epsilon.  Same
 * conventions and the same bit-identical guarantee as pyi_fill_gauss_seidel.
 * Grids hold at most 2^31 - 1 cells. */
int pyi_fill_multigrid(void *grid, int dtype, int64_t nx, int64_t ny,
                       const pyi_fill_config *cfg, int mem,
                       uint64_t *iterations, double *max_residual,
                       void *stream);

/* config::fill::Loess (config/fill.hpp:164-219) and LoessValueType (:26-30) */
typedef enum pyi_loess_value_type {
  PYI_LOESS_UNDEFINED = 0, /* fill only the undefined (NaN) cells */
  PYI_LOESS_DEFINED = 1,   /* smooth only the defined cells */
  PYI_LOESS_ALL = 2        /* smooth and fill every cell */
} pyi_loess_value_type;

typedef struct pyi_loess_config {
  int32_t first_guess;     /* pyi_first_guess (iterative mode) */
  int32_t is_periodic;     /* x axis periodic (longitude) */
  uint32_t max_iterations; /* 1: single pass; 0: only the first guess */
  uint32_t half_x;         /* config nx: half window size along x (rows), >= 1 */
  uint32_t half_y;         /* config ny: half window size along y (columns), >= 1 */
  int32_t value_type;      /* pyi_loess_value_type */
  double epsilon;          /* iteration stops when max |change| < epsilon */
} pyi_loess_config;

/* fill::loess (fill/loess.hpp:207-253): `result` receives the row-major (nx, ny)
 * grid `grid` (dtype PYI_F32 or PYI_F64, left untouched) with every processed
 * cell replaced by the tri-cube weighted mean of the defined values in its
 * (2 half_x + 1) x (2 half_y + 1) window; single pass, or iterated from the
 * first guess until the largest change is below epsilon.  Bit-identical to the
 * reference whatever its thread count. */
int pyi_fill_loess(const void *grid, void *result, int dtype, int64_t nx,
                   int64_t ny, const pyi_loess_config *cfg, int mem,
                   void *stream);

/* pyi_fill_loess for callers that may not allocate device memory (code under
 * CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
 * No counterpart in the reference.
 * Same arguments, kernels and results (bit for bit) as pyi_fill_loess with
 * PYI_MEM_DEVICE -- `grid` and `result` are device pointers on the current
 * device -- but the scratch of the iterative filling (the previous iterate and
 * two scalars) lives in a static device array of PYI_LOESS_STATIC_MAX_BYTES
 * (+ 256) bytes that every GPU loading the library holds until exit, and the
 * call makes no cudaMalloc / cudaFree / cudaMallocAsync nor any allocating
 * library call.  Grids of more than PYI_LOESS_STATIC_MAX_BYTES bytes
 * (1 048 576 float64 or 2 097 152 float32 cells) are rejected with
 * PYI_BAD_SHAPE.  All calls share the array: calls that overlap on one GPU
 * (two streams, two threads) race and must be serialised by the caller. */
#define PYI_LOESS_STATIC_MAX_BYTES (8u << 20)
int pyi_fill_loess_static(const void *grid, void *result, int dtype,
                          int64_t nx, int64_t ny, const pyi_loess_config *cfg,
                          void *stream);

/* Reference text "{coord} is out of bounds for axis {x|y|z|u} [{min}, ...,
 * {max}]" (pybind/grid.hpp:367-370). `coord` points to a double or an int64
 * according to the axis. Returns the length written (excluding NUL). */
int pyi_format_bounds_error(const pyi_grid *grid, int axis, const void *coord,
                            char *buf, size_t len);

/* TemporalAxis::coordinate_repr (math/temporal_axis.hpp:80-83,150-156;
 * dateutils::datetime64_to_string / timedelta64_to_string,
 * dateutils.hpp:729-812): the text the reference prints for an int64
 * coordinate of a temporal axis -- ISO-8601 in the axis resolution
 * (pyi_resolution) for datetime64, "<n> <unit>s" for timedelta64.  Used to
 * build the bounds-error message of a temporal z axis.  Returns the length
 * written (excluding NUL), -1 on a bad argument. */
int pyi_format_temporal(int is_timedelta, int resolution, int64_t value,
                        char *buf, size_t len);

/* The window weights the 6 x 6 tolerance kernels use for a fitting model that is linear in the node
 * values (pyi_spatial_method: bilinear, linear, the three c_spline models, polynomial), as polynomials
 * in the cell coordinate t of the centre cell of six uniformly spaced nodes `step` apart:
 * out[6 * j + d] = coefficient of t^d in the weight of node j, so that the reference's univariate
 * operator()(xa, ya, xa[2] + t step) (univariate/{linear,cspline*,polynomial}.hpp) equals
 * sum_j (sum_d out[6 j + d] t^d) ya[j] to rounding.  Host arithmetic only (no device needed): it
 * exists so that the tables can be checked against the reference without a GPU.  PYI_BAD_ARG for
 * the other models. */
int pyi_model_stencil(int method, double step, double *out36);

/* Number of kernel launches issued by this library in this process. */
int64_t pyi_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* PYINTERP_B200_H_ */
