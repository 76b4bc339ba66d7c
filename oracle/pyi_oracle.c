/*
 * pyi_oracle.c -- plain-C restatement of the reference algorithm for the
 * n-variate regular-grid interpolation path (CNES/pangeo-pyinterp, release
 * line 2026.7.0).  Exports the C-ABI of orc_abi.h as liboracle.so.
 *
 * TEST INFRASTRUCTURE ONLY.  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load this library, and
 * only as the checker or the timed CPU baseline.  The product never links,
 * imports or falls back to it.
 *
 * PARITY PINNING.  Every entry point is pinned against
 *   (1) the golden vectors of the reference's own unit tests (tests/golden/,
 *       tests/test_oracle_golden.py), and
 *   (2) oracle/_ref/liboracle_ref.so = the reference's own headers compiled in
 *       place (tests/test_oracle_vs_ref.py), bit for bit on randomized and
 *       adversarial inputs.
 * One arithmetic piece is a restatement of a dependency that is absent from
 * /root/reference: Boost.Geometry (>= 1.90, CMakeLists.txt:569)
 * comparable_distance on spherical_equatorial<degree> points, used by the idw
 * and nearest evaluators.  It follows the published comparable-haversine
 * strategy; no reference test pins a value through it => for idw / nearest on
 * the spherical point: PARITY UNPINNED (bilinear does not touch it).
 *
 * Each function cites the reference file:line it follows (paths relative to
 * cxx/include/pyinterp/ in the reference tree).
 *
 * Floating-point contract: compiled with -ffp-contract=off (the reference's
 * x86-64 wheels have no implicit FMA: no -march, CMakeLists.txt:139-141);
 * explicit fma() calls below are the reference's explicit std::fma calls.
 */
#define _GNU_SOURCE
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

#include "orc_abi.h"

struct orc_axis {
  int coord_type; /* pyi_coord_type */
  int kind;       /* pyi_axis_kind */
  int64_t size;
  int asc;
  int periodic; /* is_periodic_ */
  /* f64 axis */
  double period_f, start_f, step_f, inv_step;
  double *pts_f;
  /* i64 axis */
  int64_t period_i, start_i, step_i, step_2;
  int64_t *pts_i;
};

const char *orc_kind(void) { return "port"; }

static void set_err(char *err, size_t len, const char *what) {
  if (err != NULL && len > 0) {
    snprintf(err, len, "%s", what);
  }
}

/* ------------------------------------------------------------------------
 * math.hpp
 * ---------------------------------------------------------------------- */
/* math.hpp:93-97 (integral remainder, result takes the sign of y) */
static int64_t rem_i64(int64_t x, int64_t y) {
  int64_t r = x % y;
  return (r != 0 && (r ^ y) < 0) ? r + y : r;
}
/* math.hpp:104-111 */
static double rem_f64(double x, double y) {
  double r = remainder(x, y);
  if (r != 0.0 && signbit(r) != signbit(y)) {
    r += y;
  }
  return r;
}
/* math.hpp:120-123 */
static double normalize_period_f64(double x, double min, double period) {
  return rem_f64(x - min, period) + min;
}
static int64_t normalize_period_i64(int64_t x, int64_t min, int64_t period) {
  return rem_i64(x - min, period) + min;
}
/* math.hpp:131-152 (floating point branch) */
static double normalize_period_half_f64(double x, double min, double period) {
  const double half = period * 0.5;
  const double lo = min - half;
  const double hi = min + half;
  if (x >= lo && x < hi) {
    return x;
  }
  return normalize_period_f64(x, lo, period);
}
/* math.hpp:160-166 */
static int is_same_f64(double a, double b, double eps) {
  const double diff = fabs(a - b);
  return diff <= eps || diff < fmax(fabs(a), fabs(b)) * eps;
}
/* math.hpp:154-158 */
static int is_same_i64(int64_t a, int64_t b, int64_t eps) {
  return llabs(a - b) <= eps;
}

/* static_cast<int64_t>(double) as x86-64 executes it (cvttsd2si): NaN and
 * out-of-range values give INT64_MIN.  SURVEY.md 7.3-3. */
static int64_t cast_i64_x86(double v) {
  if (!(v >= -9223372036854775808.0 && v < 9223372036854775808.0)) {
    return INT64_MIN;
  }
  return (int64_t)v;
}

/* ------------------------------------------------------------------------
 * math/axis/container.hpp
 * ---------------------------------------------------------------------- */
/* AbstractRegular::coordinate_value container.hpp:480-483 / Irregular :206-209 */
static double coord_f64(const orc_axis *a, int64_t i) {
  if (a->kind == PYI_AXIS_REGULAR) {
    return a->start_f + (double)i * a->step_f;
  }
  return a->pts_f[i];
}
static int64_t coord_i64(const orc_axis *a, int64_t i) {
  if (a->kind == PYI_AXIS_REGULAR) {
    return a->start_i + i * a->step_i;
  }
  return a->pts_i[i];
}
/* min_value / max_value container.hpp:491-497 (regular), :221-227 (irregular) */
static double min_f64(const orc_axis *a) {
  return coord_f64(a, a->asc ? 0 : a->size - 1);
}
static double max_f64(const orc_axis *a) {
  return coord_f64(a, a->asc ? a->size - 1 : 0);
}
static int64_t min_i64(const orc_axis *a) {
  return coord_i64(a, a->asc ? 0 : a->size - 1);
}
static int64_t max_i64(const orc_axis *a) {
  return coord_i64(a, a->asc ? a->size - 1 : 0);
}

/* Irregular::find_index_impl container.hpp:382-404.  cmp(a,b) is a<b on an
 * ascending axis and a>b on a descending one. */
#define ORC_CMP(asc, a, b) ((asc) ? ((a) < (b)) : ((a) > (b)))
static int64_t irregular_find_f64(const orc_axis *a, double c, int bounded) {
  const double *p = a->pts_f;
  const int64_t n = a->size;
  int64_t lo = 0, len = n; /* std::lower_bound */
  while (len > 0) {
    const int64_t half = len >> 1;
    if (ORC_CMP(a->asc, p[lo + half], c)) {
      lo += half + 1;
      len -= half + 1;
    } else {
      len = half;
    }
  }
  if (lo == 0) {
    return ORC_CMP(a->asc, c, p[0]) ? (bounded ? 0 : -1) : 0;
  }
  if (lo == n) {
    return ORC_CMP(a->asc, p[n - 1], c) ? (bounded ? n - 1 : -1) : n - 1;
  }
  if (ORC_CMP(a->asc, c - p[lo - 1], p[lo] - c)) {
    return lo - 1;
  }
  return lo;
}
static int64_t irregular_find_i64(const orc_axis *a, int64_t c, int bounded) {
  const int64_t *p = a->pts_i;
  const int64_t n = a->size;
  int64_t lo = 0, len = n;
  while (len > 0) {
    const int64_t half = len >> 1;
    if (ORC_CMP(a->asc, p[lo + half], c)) {
      lo += half + 1;
      len -= half + 1;
    } else {
      len = half;
    }
  }
  if (lo == 0) {
    return ORC_CMP(a->asc, c, p[0]) ? (bounded ? 0 : -1) : 0;
  }
  if (lo == n) {
    return ORC_CMP(a->asc, p[n - 1], c) ? (bounded ? n - 1 : -1) : n - 1;
  }
  if (ORC_CMP(a->asc, c - p[lo - 1], p[lo] - c)) {
    return lo - 1;
  }
  return lo;
}

/* Regular<floating>::find_index container.hpp:550-564 */
static int64_t regular_find_f64(const orc_axis *a, double c, int bounded) {
  const int64_t idx = cast_i64_x86(round((c - a->start_f) * a->inv_step));
  if (idx < 0) {
    return bounded ? 0 : -1;
  }
  if (idx >= a->size) {
    return bounded ? a->size - 1 : -1;
  }
  return idx;
}
/* Regular<integral>::find_index + round container.hpp:588-615 */
static int64_t regular_find_i64(const orc_axis *a, int64_t c, int bounded) {
  const int64_t num = (int64_t)((uint64_t)c - (uint64_t)a->start_i);
  const int64_t idx =
      num > 0 ? (num + a->step_2) / a->step_i : (num - a->step_2) / a->step_i;
  if (idx < 0) {
    return bounded ? 0 : -1;
  }
  if (idx >= a->size) {
    return bounded ? a->size - 1 : -1;
  }
  return idx;
}
static int64_t container_find_f64(const orc_axis *a, double c, int bounded) {
  return a->kind == PYI_AXIS_REGULAR ? regular_find_f64(a, c, bounded)
                                     : irregular_find_f64(a, c, bounded);
}
static int64_t container_find_i64(const orc_axis *a, int64_t c, int bounded) {
  return a->kind == PYI_AXIS_REGULAR ? regular_find_i64(a, c, bounded)
                                     : irregular_find_i64(a, c, bounded);
}

/* ------------------------------------------------------------------------
 * math/axis.hpp : lookups
 * ---------------------------------------------------------------------- */
/* Axis::normalize_coordinate axis.hpp:294-310 (tests period_, not
 * is_periodic_) */
static double axis_normalize_f64(const orc_axis *a, double c, double min) {
  if (a->period_f != 0.0 && (c >= min + a->period_f || c < min)) {
    return normalize_period_f64(c, min, a->period_f);
  }
  return c;
}
static int64_t axis_normalize_i64(const orc_axis *a, int64_t c, int64_t min) {
  if (a->period_i != 0 && (c >= min + a->period_i || c < min)) {
    return normalize_period_i64(c, min, a->period_i);
  }
  return c;
}

/* Axis::find_index axis.hpp:319-333 */
static int64_t axis_find_index_f64(const orc_axis *a, double c, int bounded) {
  if (isnan(c)) { /* fill.hpp:36-40 */
    return -1;
  }
  c = axis_normalize_f64(a, c, min_f64(a));
  const int64_t r = container_find_f64(a, c, bounded);
  if (r == -1 && a->periodic) {
    return (c - max_f64(a)) < (min_f64(a) + a->period_f - c) ? a->size - 1 : 0;
  }
  return r;
}
static int64_t axis_find_index_i64(const orc_axis *a, int64_t c, int bounded) {
  if (c == INT64_MIN) { /* fill.hpp:43-47 */
    return -1;
  }
  c = axis_normalize_i64(a, c, min_i64(a));
  const int64_t r = container_find_i64(a, c, bounded);
  if (r == -1 && a->periodic) {
    return (c - max_i64(a)) < (min_i64(a) + a->period_i - c) ? a->size - 1 : 0;
  }
  return r;
}

/* Axis::find_indexes(T) axis.hpp:720-778.  Returns 1 and the pair, or 0. */
#define ORC_FIND_INDEXES_BODY(COORD, FIND)                                     \
  const int64_t n = a->size;                                                   \
  int64_t i0 = FIND(a, c, 0);                                                  \
  if (i0 == -1) {                                                              \
    if (a->periodic) {                                                         \
      *o0 = n - 1;                                                             \
      *o1 = 0;                                                                 \
      return 1;                                                                \
    }                                                                          \
    return 0;                                                                  \
  }                                                                            \
  const int below = c < COORD(a, i0);                                          \
  const int exact = c == COORD(a, i0);                                         \
  if (exact) {                                                                 \
    if (n == 1) {                                                              \
      return 0;                                                                \
    }                                                                          \
    const int64_t j = (i0 == n - 1) ? i0 - 1 : i0 + 1;                         \
    *o0 = i0 < j ? i0 : j;                                                     \
    *o1 = i0 < j ? j : i0;                                                     \
    return 1;                                                                  \
  }                                                                            \
  int64_t i1 = i0;                                                             \
  if (below) {                                                                 \
    i0 = a->asc ? i0 - 1 : i0 + 1;                                             \
  } else {                                                                     \
    i1 = a->asc ? i0 + 1 : i0 - 1;                                             \
  }                                                                            \
  if (a->periodic) {                                                           \
    *o0 = rem_i64(i0, n);                                                      \
    *o1 = rem_i64(i1, n);                                                      \
    return 1;                                                                  \
  }                                                                            \
  if (i0 >= 0 && i1 >= 0 && i0 < n && i1 < n) {                                \
    *o0 = i0;                                                                  \
    *o1 = i1;                                                                  \
    return 1;                                                                  \
  }                                                                            \
  return 0;

static int axis_find_indexes_f64(const orc_axis *a, double c, int64_t *o0,
                                 int64_t *o1) {
  if (isnan(c)) {
    return 0;
  }
  c = axis_normalize_f64(a, c, min_f64(a));
  /* delta = c - coord(i0); delta == 0 <=> c == coord, delta < 0 <=> c < coord
   * for finite IEEE doubles (and for +-inf neither holds, as in the
   * reference where delta would be nan/inf of matching sign). */
  ORC_FIND_INDEXES_BODY(coord_f64, container_find_f64)
}
static int axis_find_indexes_i64(const orc_axis *a, int64_t c, int64_t *o0,
                                 int64_t *o1) {
  if (c == INT64_MIN) {
    return 0;
  }
  c = axis_normalize_i64(a, c, min_i64(a));
  ORC_FIND_INDEXES_BODY(coord_i64, container_find_i64)
}

/* make_boundary_handler axis.hpp:784-810 */
static int64_t boundary_index(int boundary, int periodic, int64_t idx,
                              int64_t size) {
  if (periodic) {
    return rem_i64(idx, size);
  }
  switch (boundary) {
    case ORC_BOUNDARY_EXPAND:
      return idx < 0 ? 0 : (idx > size - 1 ? size - 1 : idx);
    case ORC_BOUNDARY_WRAP:
      return rem_i64(idx, size);
    case ORC_BOUNDARY_SYM:
      if (idx < 0) {
        return rem_i64(-idx, size);
      }
      return size - 2 - rem_i64(idx - size, size);
    default: /* undef, shrink */
      return -1;
  }
}

/* Axis::find_indexes(T, half, boundary) axis.hpp:814-881.  idx must hold
 * 2*half entries. */
static int axis_window_from_pair(const orc_axis *a, int found, int64_t p0,
                                 int64_t p1, int half, int boundary,
                                 int64_t *idx, int *count, int *c0, int *c1) {
  if (half == 0) {
    return 0;
  }
  if (!found) {
    if (a->size != 1) {
      return 0;
    }
    for (int k = 0; k < 2 * half; ++k) {
      idx[k] = 0;
    }
    *count = 2 * half;
    *c0 = 0;
    *c1 = 0;
    return 1;
  }
  const int64_t n = a->size;
  /* deque emulation: fill a 2*half buffer from the middle outwards */
  int64_t *buf = idx;
  int64_t tmp[2 * 64];
  if (half > 64) {
    return 0;
  }
  int front = half - 1, back = half; /* occupied range is [front, back] */
  tmp[front] = p0;
  tmp[back] = p1;
  int ci0 = 0, ci1 = 1;
  for (int shift = 1; shift < half; ++shift) {
    int64_t before = p0 - shift;
    if (before < 0) {
      before = boundary_index(boundary, a->periodic, before, n);
    }
    if (before >= 0) {
      tmp[--front] = before;
      ++ci0;
      ++ci1;
    } else if (boundary != PYI_BOUNDARY_SHRINK) {
      return 0;
    }
    int64_t after = p1 + shift;
    if (after >= n) {
      after = boundary_index(boundary, a->periodic, after, n);
    }
    if (after >= 0) {
      tmp[++back] = after;
    } else if (boundary != PYI_BOUNDARY_SHRINK) {
      return 0;
    }
  }
  *count = back - front + 1;
  for (int k = 0; k < *count; ++k) {
    buf[k] = tmp[front + k];
  }
  *c0 = ci0;
  *c1 = ci1;
  return 1;
}
static int axis_window_f64(const orc_axis *a, double c, int half, int boundary,
                           int64_t *idx, int *count, int *c0, int *c1) {
  int64_t p0 = 0, p1 = 0;
  const int found = axis_find_indexes_f64(a, c, &p0, &p1);
  return axis_window_from_pair(a, found, p0, p1, half, boundary, idx, count, c0,
                               c1);
}
static int axis_window_i64(const orc_axis *a, int64_t c, int half, int boundary,
                           int64_t *idx, int *count, int *c0, int *c1) {
  int64_t p0 = 0, p1 = 0;
  const int found = axis_find_indexes_i64(a, c, &p0, &p1);
  return axis_window_from_pair(a, found, p0, p1, half, boundary, idx, count, c0,
                               c1);
}

/* Axis::contains axis.hpp:261-272 */
static int axis_contains_f64(const orc_axis *a, double c) {
  if (isnan(c)) {
    return 0;
  }
  if (a->periodic) {
    return 1;
  }
  return c >= min_f64(a) && c <= max_f64(a);
}

/* ------------------------------------------------------------------------
 * math/axis.hpp : construction
 * ---------------------------------------------------------------------- */
/* Eigen LinSpaced(n, low, high)[i], floating point
 * (third_party/eigen/Eigen/src/Core/functors/NullaryFunctors.h:67-110; the
 * SSE2 packet path evaluates the same products and sums, no FMA). n >= 2. */
static double linspaced_f64(int64_t n, double low, double high, int64_t i) {
  const int64_t size1 = n - 1;
  const double step = (high - low) / (double)(n - 1);
  if (fabs(high) < fabs(low)) {
    return i == 0 ? low : high - (double)(size1 - i) * step;
  }
  return i == size1 ? high : low + (double)i * step;
}
/* integer LinSpaced, NullaryFunctors.h:113-133. n >= 2. */
static int64_t linspaced_i64(int64_t n, int64_t low, int64_t high, int64_t i) {
  const int64_t span = high - low;
  const int64_t aspan1 = llabs(span) + 1;
  const int64_t multiplier = span / (n - 1);
  const int64_t divisor = ((high >= low ? n : -n) + span) / (aspan1 == 0 ? 1 : aspan1);
  const int use_divisor = n > 1 && aspan1 < n;
  return use_divisor ? low + i / divisor : low + i * multiplier;
}

/* Axis::adjust_period axis.hpp:629-663.  Returns 1 when v was rewritten. */
#define ORC_ADJUST_PERIOD(v, n, period)                                        \
  do {                                                                         \
    if ((n) >= 2) {                                                            \
      const int ascending = (v)[0] < (v)[1];                                   \
      int needs = 0;                                                           \
      for (int64_t k = 1; k < (n) && !needs; ++k) {                            \
        needs = ascending ? (v)[k - 1] > (v)[k] : (v)[k - 1] < (v)[k];         \
      }                                                                        \
      if (needs) {                                                             \
        int cross = 0;                                                         \
        for (int64_t k = 1; k < (n); ++k) {                                    \
          if (!cross) {                                                        \
            cross = ascending ? (v)[k - 1] > (v)[k] : (v)[k - 1] < (v)[k];     \
          }                                                                    \
          if (cross) {                                                         \
            (v)[k] += ascending ? (period) : -(period);                        \
          }                                                                    \
        }                                                                      \
      }                                                                        \
    }                                                                          \
  } while (0)

/* Irregular::is_monotonic container.hpp:189-201 */
#define ORC_IS_MONOTONIC(p, n, asc, result)                                    \
  do {                                                                         \
    (result) = 1;                                                              \
    for (int64_t k = 1; k < (n) && (result); ++k) {                            \
      if ((p)[k - 1] == (p)[k]) (result) = 0;                                  \
    }                                                                          \
    for (int64_t k = 1; k < (n) && (result); ++k) {                            \
      if ((asc) ? ((p)[k] < (p)[k - 1]) : ((p)[k] > (p)[k - 1])) (result) = 0; \
    }                                                                          \
  } while (0)

orc_axis *orc_axis_new_f64(const double *values, int64_t n, double epsilon,
                           int has_period, double period, char *err,
                           size_t errlen) {
  orc_axis *a = (orc_axis *)calloc(1, sizeof(orc_axis));
  a->coord_type = PYI_COORD_F64;
  a->periodic = has_period != 0;            /* axis.hpp:699 */
  a->period_f = has_period ? period : 0.0;
  if (n <= 0) {
    /* Irregular ctor container.hpp:163-167 */
    set_err(err, errlen, "Cannot create an Irregular axis with no points.");
    free(a);
    return NULL;
  }
  double *v = (double *)malloc(sizeof(double) * (size_t)n);
  memcpy(v, values, sizeof(double) * (size_t)n);
  if (a->period_f != 0.0) { /* axis.hpp:706-712 */
    ORC_ADJUST_PERIOD(v, n, a->period_f);
  }
  /* is_evenly_spaced axis.hpp:582-597 + create_container_impl :669-677 */
  int regular = 0;
  if (n >= 2) {
    double maxdiff = 0.0; /* Eigen maxCoeff over |diff| */
    for (int64_t k = 0; k < n; ++k) {
      const double d = fabs(v[k] - linspaced_f64(n, v[0], v[n - 1], k));
      if (k == 0 || d > maxdiff) {
        maxdiff = d;
      }
    }
    regular = maxdiff <= epsilon;
  }
  a->size = n;
  if (regular) {
    /* AbstractRegular ctor container.hpp:415-436, Regular<float> :545-547 */
    if (v[0] == v[n - 1]) {
      set_err(err, errlen,
              "An axis with a single value requires distinct start and stop "
              "values.");
      free(v);
      free(a);
      return NULL;
    }
    a->kind = PYI_AXIS_REGULAR;
    a->start_f = v[0];
    a->step_f = (v[n - 1] - v[0]) / (double)(int64_t)(n - 1);
    a->inv_step = 1.0 / a->step_f;
    a->asc = coord_f64(a, 0) < coord_f64(a, 1); /* container.hpp:96-98 */
    free(v);
  } else {
    a->kind = PYI_AXIS_IRREGULAR;
    a->pts_f = v;
    a->start_f = v[0];
    a->asc = n < 2 ? 1 : v[0] < v[1];
    int mono;
    ORC_IS_MONOTONIC(v, n, a->asc, mono);
    if (!mono) { /* compute_properties axis.hpp:606-608 */
      set_err(err, errlen, "Axis values must be monotonic.");
      orc_axis_free(a);
      return NULL;
    }
  }
  /* compute_properties axis.hpp:609-623 */
  if (a->periodic) {
    if (a->kind == PYI_AXIS_REGULAR) {
      a->periodic = is_same_f64(fabs(a->step_f * (double)a->size), a->period_f,
                                epsilon);
    } else {
      const double step = (a->pts_f[n - 1] - a->pts_f[0]) / (double)(n - 1);
      a->periodic =
          fabs((max_f64(a) - min_f64(a)) - a->period_f) <= step + epsilon;
    }
  }
  return a;
}

orc_axis *orc_axis_new_i64(const int64_t *values, int64_t n, int64_t epsilon,
                           int has_period, int64_t period, char *err,
                           size_t errlen) {
  orc_axis *a = (orc_axis *)calloc(1, sizeof(orc_axis));
  a->coord_type = PYI_COORD_I64;
  a->periodic = has_period != 0;
  a->period_i = has_period ? period : 0;
  if (n <= 0) {
    set_err(err, errlen, "Cannot create an Irregular axis with no points.");
    free(a);
    return NULL;
  }
  int64_t *v = (int64_t *)malloc(sizeof(int64_t) * (size_t)n);
  memcpy(v, values, sizeof(int64_t) * (size_t)n);
  if (a->period_i != 0) {
    ORC_ADJUST_PERIOD(v, n, a->period_i);
  }
  int regular = 0;
  if (n >= 2) {
    int64_t maxdiff = 0;
    for (int64_t k = 0; k < n; ++k) {
      const int64_t d = llabs(v[k] - linspaced_i64(n, v[0], v[n - 1], k));
      if (k == 0 || d > maxdiff) {
        maxdiff = d;
      }
    }
    regular = maxdiff <= epsilon;
  }
  a->size = n;
  if (regular) {
    if (v[0] == v[n - 1]) {
      set_err(err, errlen,
              "An axis with a single value requires distinct start and stop "
              "values.");
      free(v);
      free(a);
      return NULL;
    }
    a->kind = PYI_AXIS_REGULAR;
    a->start_i = v[0];
    a->step_i = (v[n - 1] - v[0]) / (int64_t)(n - 1);
    a->step_2 = a->step_i >> 1; /* container.hpp:585 */
    a->asc = coord_i64(a, 0) < coord_i64(a, 1);
    free(v);
  } else {
    a->kind = PYI_AXIS_IRREGULAR;
    a->pts_i = v;
    a->start_i = v[0];
    a->asc = n < 2 ? 1 : v[0] < v[1];
    int mono;
    ORC_IS_MONOTONIC(v, n, a->asc, mono);
    if (!mono) {
      set_err(err, errlen, "Axis values must be monotonic.");
      orc_axis_free(a);
      return NULL;
    }
  }
  if (a->periodic) {
    if (a->kind == PYI_AXIS_REGULAR) {
      /* static_cast<T>(std::fabs(step * size())) goes through double */
      const int64_t span = (int64_t)fabs((double)(a->step_i * a->size));
      a->periodic = is_same_i64(span, a->period_i, epsilon);
    } else {
      const int64_t step = (a->pts_i[n - 1] - a->pts_i[0]) / (int64_t)(n - 1);
      a->periodic =
          llabs((max_i64(a) - min_i64(a)) - a->period_i) <= step + epsilon;
    }
  }
  return a;
}

void orc_axis_free(orc_axis *a) {
  if (a != NULL) {
    free(a->pts_f);
    free(a->pts_i);
    free(a);
  }
}

/* Axis::flip: Regular container.hpp:473-477,567-570 (step_2_ is NOT updated),
 * Irregular :184-187 */
void orc_axis_flip(orc_axis *a) {
  if (a->kind == PYI_AXIS_REGULAR) {
    if (a->coord_type == PYI_COORD_F64) {
      a->start_f = coord_f64(a, a->size - 1);
      a->step_f = -a->step_f;
      a->inv_step = -a->inv_step;
    } else {
      a->start_i = coord_i64(a, a->size - 1);
      a->step_i = -a->step_i;
    }
  } else {
    for (int64_t i = 0, j = a->size - 1; i < j; ++i, --j) {
      if (a->coord_type == PYI_COORD_F64) {
        const double t = a->pts_f[i];
        a->pts_f[i] = a->pts_f[j];
        a->pts_f[j] = t;
      } else {
        const int64_t t = a->pts_i[i];
        a->pts_i[i] = a->pts_i[j];
        a->pts_i[j] = t;
      }
    }
    if (a->coord_type == PYI_COORD_F64) {
      a->start_f = a->pts_f[0];
    } else {
      a->start_i = a->pts_i[0];
    }
  }
  a->asc = !a->asc;
}

int orc_axis_describe(const orc_axis *a, pyi_axis_desc *d) {
  memset(d, 0, sizeof(*d));
  d->kind = a->kind;
  d->coord_type = a->coord_type;
  d->size = a->size;
  d->is_ascending = a->asc;
  d->is_periodic = a->periodic;
  if (a->coord_type == PYI_COORD_F64) {
    d->period.f = a->period_f;
    d->start.f = a->start_f;
    d->step.f = a->step_f;
    d->min_value.f = min_f64(a);
    d->max_value.f = max_f64(a);
    d->inv_step = a->inv_step;
    d->points = a->pts_f;
  } else {
    d->period.i = a->period_i;
    d->start.i = a->start_i;
    d->step.i = a->step_i;
    d->min_value.i = min_i64(a);
    d->max_value.i = max_i64(a);
    d->step_2 = a->step_2;
    d->points = a->pts_i;
  }
  return PYI_OK;
}

void orc_axis_find_indexes(const orc_axis *a, int64_t n, const void *coords,
                           int64_t *out) {
  for (int64_t k = 0; k < n; ++k) {
    int64_t i0 = -1, i1 = -1;
    const int ok = a->coord_type == PYI_COORD_F64
                       ? axis_find_indexes_f64(a, ((const double *)coords)[k], &i0, &i1)
                       : axis_find_indexes_i64(a, ((const int64_t *)coords)[k], &i0, &i1);
    out[2 * k] = ok ? i0 : -1;
    out[2 * k + 1] = ok ? i1 : -1;
  }
}

void orc_axis_find_index(const orc_axis *a, int64_t n, const void *coords,
                         int bounded, int64_t *out) {
  for (int64_t k = 0; k < n; ++k) {
    out[k] = a->coord_type == PYI_COORD_F64
                 ? axis_find_index_f64(a, ((const double *)coords)[k], bounded)
                 : axis_find_index_i64(a, ((const int64_t *)coords)[k], bounded);
  }
}

int orc_axis_find_window(const orc_axis *a, const void *coord, int half,
                         int boundary, int64_t *idx, int *count, int *c0,
                         int *c1) {
  if (a->coord_type == PYI_COORD_F64) {
    return axis_window_f64(a, *(const double *)coord, half, boundary, idx,
                           count, c0, c1);
  }
  return axis_window_i64(a, *(const int64_t *)coord, half, boundary, idx, count,
                         c0, c1);
}

/* ------------------------------------------------------------------------
 * grid access + threading
 * ---------------------------------------------------------------------- */
/* Grid::value pybind/grid.hpp:340-345 (C-contiguous).  Every dtype converts
 * exactly to double except i64/u64, whose result type is double anyway. */
static double grid_value(int dtype, const void *data, int64_t off) {
  switch (dtype) {
    case PYI_I8: return (double)((const int8_t *)data)[off];
    case PYI_U8: return (double)((const uint8_t *)data)[off];
    case PYI_I16: return (double)((const int16_t *)data)[off];
    case PYI_U16: return (double)((const uint16_t *)data)[off];
    case PYI_I32: return (double)((const int32_t *)data)[off];
    case PYI_U32: return (double)((const uint32_t *)data)[off];
    case PYI_I64: return (double)((const int64_t *)data)[off];
    case PYI_U64: return (double)((const uint64_t *)data)[off];
    case PYI_F32: return (double)((const float *)data)[off];
    default: return ((const double *)data)[off];
  }
}
/* pybind/grid_dispatch.hpp:25-30 */
static int result_is_f64(int dtype) {
  return dtype == PYI_F64 || dtype == PYI_I64 || dtype == PYI_U64;
}

typedef struct chunk_status {
  int64_t oob_index; /* -1: none */
  int oob_axis;
} chunk_status;

typedef struct geo_job {
  int ndim;
  const orc_axis *const *axes;
  int64_t shape[4];
  int dtype;
  const void *data;
  const pyi_geometric_config *cfg;
  const double *x, *y, *u;
  const void *z;
  void *out;
} geo_job;

typedef void (*range_fn)(const void *job, int64_t start, int64_t end,
                         chunk_status *st);
typedef struct chunk_arg {
  range_fn fn;
  const void *job;
  int64_t start, end;
  chunk_status st;
} chunk_arg;
static void *chunk_main(void *p) {
  chunk_arg *c = (chunk_arg *)p;
  c->fn(c->job, c->start, c->end, &c->st);
  return NULL;
}
/* parallel_for parallel_for.hpp:30-78: static contiguous chunks; the error of
 * the lowest chunk wins. */
static chunk_status parallel_ranges(int64_t size, range_fn fn, const void *job,
                                    int64_t num_threads) {
  chunk_status none = {-1, -1};
  if (num_threads == 0) {
    num_threads = sysconf(_SC_NPROCESSORS_ONLN);
  }
  if (num_threads > size) {
    num_threads = size;
  }
  if (num_threads <= 1) {
    chunk_status st = none;
    fn(job, 0, size, &st);
    return st;
  }
  chunk_arg *args = (chunk_arg *)calloc((size_t)num_threads, sizeof(chunk_arg));
  pthread_t *th = (pthread_t *)calloc((size_t)num_threads, sizeof(pthread_t));
  const int64_t chunk = size / num_threads;
  const int64_t rem = size % num_threads;
  int64_t start = 0;
  for (int64_t t = 0; t < num_threads; ++t) {
    const int64_t end = start + chunk + (t < rem ? 1 : 0);
    args[t].fn = fn;
    args[t].job = job;
    args[t].start = start;
    args[t].end = end;
    args[t].st = none;
    pthread_create(&th[t], NULL, chunk_main, &args[t]);
    start = end;
  }
  chunk_status first = none;
  for (int64_t t = 0; t < num_threads; ++t) {
    pthread_join(th[t], NULL);
  }
  for (int64_t t = 0; t < num_threads; ++t) {
    if (args[t].st.oob_index >= 0) {
      first = args[t].st;
      break;
    }
  }
  free(args);
  free(th);
  return first;
}

/* "{coord} is out of bounds for axis {label} [{min}, ..., {max}]"
 * pybind/grid.hpp:362-371; doubles use the shortest round-trip form that
 * std::format("{}") produces, int64 (temporal) values are printed as
 * integers here (the reference prints ISO dates for TemporalAxis). */
static void fmt_f64(double v, char *buf, size_t len) {
  if (isnan(v)) {
    snprintf(buf, len, signbit(v) ? "-nan" : "nan");
    return;
  }
  if (isinf(v)) {
    snprintf(buf, len, v < 0 ? "-inf" : "inf");
    return;
  }
  /* shortest digits that round-trip */
  char e[40];
  int prec;
  for (prec = 0; prec < 17; ++prec) {
    snprintf(e, sizeof e, "%.*e", prec, v);
    if (strtod(e, NULL) == v) break;
  }
  /* std::to_chars picks fixed or scientific, whichever is shorter (fixed on
   * a tie) */
  char f[400];
  int fp;
  for (fp = 0; fp < 340; ++fp) {
    snprintf(f, sizeof f, "%.*f", fp, v);
    if (strtod(f, NULL) == v) break;
  }
  /* strip exponent leading zeros to the two digits to_chars writes */
  if (strlen(f) <= strlen(e)) {
    snprintf(buf, len, "%s", f);
  } else {
    snprintf(buf, len, "%s", e);
  }
}
static void bounds_message(const geo_job *J, int axis, int64_t k, char *msg,
                           size_t len) {
  static const char *labels[4] = {"x", "y", "z", "u"};
  const orc_axis *a = J->axes[axis];
  if (msg == NULL || len == 0) return;
  if (a->coord_type == PYI_COORD_I64) {
    snprintf(msg, len, "%lld is out of bounds for axis %s [%lld, ..., %lld]",
             (long long)((const int64_t *)J->z)[k], labels[axis],
             (long long)min_i64(a), (long long)max_i64(a));
    return;
  }
  const double c = axis == 0 ? J->x[k]
                   : axis == 1 ? J->y[k]
                   : axis == 2 ? ((const double *)J->z)[k]
                               : J->u[k];
  char b0[400], b1[400], b2[400];
  fmt_f64(c, b0, sizeof b0);
  fmt_f64(min_f64(a), b1, sizeof b1);
  fmt_f64(max_f64(a), b2, sizeof b2);
  snprintf(msg, len, "%s is out of bounds for axis %s [%s, ..., %s]", b0,
           labels[axis], b1, b2);
}

/* ------------------------------------------------------------------------
 * evaluators and per-point glue, once per result type
 * ---------------------------------------------------------------------- */
#define R double
#define SUF(n) n##_d
#define R_SIN sin
#define R_COS cos
#define R_EPS 2.220446049250313080847263336181640625e-16
#define R_PI 3.141592653589793238462643383279502884
#define R_NAN ((double)NAN)
#include "pyi_oracle_eval.inc"
#undef R
#undef SUF
#undef R_SIN
#undef R_COS
#undef R_EPS
#undef R_PI
#undef R_NAN

#define R float
#define SUF(n) n##_s
#define R_SIN sinf
#define R_COS cosf
#define R_EPS 1.1920928955078125e-07f
#define R_PI 3.141592653589793238462643383279502884f
#define R_NAN ((float)NAN)
#include "pyi_oracle_eval.inc"
#undef R
#undef SUF
#undef R_SIN
#undef R_COS
#undef R_EPS
#undef R_PI
#undef R_NAN

static void geo_range_d(const void *job, int64_t s, int64_t e, chunk_status *st) {
  geometric_range_d((const geo_job *)job, s, e, st);
}
static void geo_range_s(const void *job, int64_t s, int64_t e, chunk_status *st) {
  geometric_range_s((const geo_job *)job, s, e, st);
}

int orc_geometric(int ndim, const orc_axis *const *axes, int dtype,
                  const void *data, const pyi_geometric_config *cfg, int64_t n,
                  const double *x, const double *y, const void *z,
                  const double *u, void *out, int64_t *oob_index, int *oob_axis,
                  char *msg, size_t msglen) {
  if (ndim < 2 || ndim > 4 || dtype < PYI_I8 || dtype > PYI_F64) {
    return PYI_BAD_ARG;
  }
  geo_job J;
  memset(&J, 0, sizeof J);
  J.ndim = ndim;
  J.axes = axes;
  for (int k = 0; k < ndim; ++k) J.shape[k] = axes[k]->size;
  J.dtype = dtype;
  J.data = data;
  J.cfg = cfg;
  J.x = x;
  J.y = y;
  J.z = z;
  J.u = u;
  J.out = out;
  /* result.setConstant(NaN) bivariate.hpp:141-142 */
  if (result_is_f64(dtype)) {
    for (int64_t k = 0; k < n; ++k) ((double *)out)[k] = NAN;
  } else {
    for (int64_t k = 0; k < n; ++k) ((float *)out)[k] = NAN;
  }
  const chunk_status st = parallel_ranges(
      n, result_is_f64(dtype) ? geo_range_d : geo_range_s, &J, cfg->num_threads);
  if (st.oob_index >= 0) {
    if (oob_index) *oob_index = st.oob_index;
    if (oob_axis) *oob_axis = st.oob_axis;
    bounds_message(&J, st.oob_axis, st.oob_index, msg, msglen);
    return PYI_OOB;
  }
  return PYI_OK;
}

double orc_eval2d_f64(int method, int idw_exp, int cartesian, double x, double y,
                      double x0, double y0, double x1, double y1, double q00,
                      double q01, double q10, double q11) {
  return eval2d_d(method, idw_exp, cartesian, x, y, x0, y0, x1, y1, q00, q01,
                  q10, q11);
}
float orc_eval2d_f32(int method, int idw_exp, int cartesian, float x, float y,
                     float x0, float y0, float x1, float y1, float q00,
                     float q01, float q10, float q11) {
  return eval2d_s(method, idw_exp, cartesian, x, y, x0, y0, x1, y1, q00, q01,
                  q10, q11);
}
double orc_axis_interp_f64(int method, double x, double x0, double x1, double y0,
                           double y1) {
  return axis_interp_f64_d(method, x, x0, x1, y0, y1);
}
double orc_axis_interp_i64(int method, int64_t x, int64_t x0, int64_t x1,
                           double y0, double y1) {
  return axis_interp_i64_d(method, x, x0, x1, y0, y1);
}
float orc_axis_interp_i64_f32(int method, int64_t x, int64_t x0, int64_t x1,
                              float y0, float y1) {
  return axis_interp_i64_s(method, x, x0, x1, y0, y1);
}

/* ------------------------------------------------------------------------
 * windowed 2-D path (all arithmetic in double,
 * pybind/windowed/bivariate.hpp:32-33)
 * ---------------------------------------------------------------------- */
#define ORC_MAXW 128 /* 2 * max half window */

/* BracketFinder::search bracket_finder.hpp:40-78 */
static int bracket_search(const double *xa, int n, double x, int *o0, int *o1) {
  if (n == 0) {
    return 0;
  }
  const int ascending = n == 1 || xa[0] <= xa[n - 1];
  int lo = 0, len = n; /* std::lower_bound with < or > */
  while (len > 0) {
    const int half = len >> 1;
    if (ascending ? (xa[lo + half] < x) : (xa[lo + half] > x)) {
      lo += half + 1;
      len -= half + 1;
    } else {
      len = half;
    }
  }
  if (lo == n) {
    return 0;
  }
  if (lo == 0) {
    if (xa[0] == x && n > 1) {
      *o0 = 0;
      *o1 = 1;
      return 1;
    }
    return 0;
  }
  const int i0 = lo - 1, i1 = lo;
  const int ok = ascending ? (xa[i0] <= x && x <= xa[i1])
                           : (xa[i0] >= x && x >= xa[i1]);
  if (ok) {
    *o0 = i0;
    *o1 = i1;
    return 1;
  }
  return 0;
}

/* CSpline::compute_coefficients + solve_symmetric_tridiagonal
 * univariate/cspline.hpp:83-140,53-81.  c2[0..n) receives x_ (c2[0] =
 * c2[n-1] = 0, natural boundary).  Returns 0 when n < 4 (min_size). */
static int cspline_coefficients(const double *xa, const double *ya, int n,
                                double *c2) {
  if (n < 4) {
    return 0;
  }
  const int m = n - 2;
  double diag[ORC_MAXW], lower[ORC_MAXW], upper[ORC_MAXW], b[ORC_MAXW];
  double c[ORC_MAXW], d[ORC_MAXW];
  for (int i = 0; i < m; ++i) {
    const double x1 = xa[i + 1];
    const double y1 = ya[i + 1];
    const double h_i0 = x1 - xa[i];
    const double h_i1 = xa[i + 2] - x1;
    const double y_i0 = y1 - ya[i];
    const double y_i1 = ya[i + 2] - y1;
    const double g_i0 = (h_i0 != 0.0) ? 1.0 / h_i0 : 0.0;
    const double g_i1 = (h_i1 != 0.0) ? 1.0 / h_i1 : 0.0;
    lower[i] = h_i0;                 /* A(i, i-1), used for i > 0 */
    diag[i] = 2.0 * (h_i0 + h_i1);   /* A(i, i) */
    upper[i] = h_i1;                 /* A(i, i+1), used for i < m-1 */
    b[i] = 3.0 * fma(y_i1, g_i1, -y_i0 * g_i0);
  }
  double *x = c2 + 1;
  c2[0] = 0.0;
  c2[n - 1] = 0.0;
  const int m1 = m - 1, m2 = m - 2;
  double item = 1.0 / diag[0];
  c[0] = upper[0] * item;
  d[0] = b[0] * item;
  for (int i = 1; i < m1; ++i) {
    item = lower[i];
    const double inv_m = 1.0 / (diag[i] - item * c[i - 1]);
    c[i] = upper[i] * inv_m;
    d[i] = fma(-item, d[i - 1], b[i]) * inv_m;
  }
  item = lower[m1];
  d[m1] = (b[m1] - item * d[m2]) / (diag[m1] - item * c[m2]);
  x[m1] = d[m1];
  for (int i = m2; i >= 0; --i) {
    x[i] = fma(-c[i], x[i + 1], d[i]);
  }
  return 1;
}

/* CSplineCoefficients cspline_base.hpp:32-41 */
static void cspline_bcd(double c0, double c1, double dx, double dy, double *b,
                        double *c, double *d) {
  const double third = 1.0 / 3.0;
  *b = (dy / dx) - dx * (c1 + 2.0 * c0) * third;
  *c = c0;
  *d = (c1 - c0) / (3.0 * dx);
}
/* CSplineBase::interpolate_ / derivative_ cspline_base.hpp:87-142 */
static double cspline_eval(const double *xa, const double *ya, int n,
                           const double *c2, double x, int derivative) {
  int i0, i1;
  if (!bracket_search(xa, n, x, &i0, &i1)) {
    return NAN;
  }
  const double dx = xa[i1] - xa[i0];
  const double dy = ya[i1] - ya[i0];
  const double h = x - xa[i0];
  double b, c, d;
  cspline_bcd(c2[i0], c2[i1], dx, dy, &b, &c, &d);
  if (derivative) {
    return b + h * fma(3.0 * h, d, 2.0 * c);
  }
  return ya[i0] + h * fma(h, fma(h, d, c), b);
}
/* Univariate::derivative(xa, ya, Vector x) univariate.hpp:56-70 evaluated at
 * the nodes themselves, as Bicubic::compute_coefficients does. */
static void cspline_node_derivatives(const double *xa, const double *ya, int n,
                                     double *out) {
  double c2[ORC_MAXW];
  if (!cspline_coefficients(xa, ya, n, c2)) {
    for (int k = 0; k < n; ++k) out[k] = NAN;
    return;
  }
  for (int k = 0; k < n; ++k) {
    out[k] = cspline_eval(xa, ya, n, c2, xa[k], 1);
  }
}

double orc_cspline(const double *xa, const double *ya, int n, double x,
                   int derivative) {
  double c2[ORC_MAXW];
  if (n > ORC_MAXW || !cspline_coefficients(xa, ya, n, c2)) {
    return NAN;
  }
  return cspline_eval(xa, ya, n, c2, x, derivative);
}

/* A prepared window: BivariateBase::prepare + Bicubic::compute_coefficients
 * interpolate/bivariate.hpp:47-72, bivariate/bicubic.hpp:57-87. za, zx, zy,
 * zxy are [nx][ny] row-major. */
typedef struct window {
  int nx, ny;
  int valid;
  double xa[ORC_MAXW], ya[ORC_MAXW];
  double *za, *zx, *zy, *zxy; /* nx*ny each */
} window;

static void window_prepare(window *w, int method) {
  const int nx = w->nx, ny = w->ny;
  if (nx < 2 || ny < 2) { /* min_size() == 2 for both evaluators */
    w->valid = 0;
    return;
  }
  w->valid = 1;
  if (method != PYI_W_BICUBIC) {
    return;
  }
  double col[ORC_MAXW], der[ORC_MAXW];
  for (int j = 0; j < ny; ++j) { /* zx(:, j) */
    for (int i = 0; i < nx; ++i) col[i] = w->za[i * ny + j];
    cspline_node_derivatives(w->xa, col, nx, der);
    for (int i = 0; i < nx; ++i) w->zx[i * ny + j] = der[i];
  }
  for (int i = 0; i < nx; ++i) { /* zy(i, :) */
    cspline_node_derivatives(w->ya, &w->za[i * ny], ny, &w->zy[i * ny]);
  }
  for (int j = 0; j < ny; ++j) { /* zxy(:, j) */
    for (int i = 0; i < nx; ++i) col[i] = w->zy[i * ny + j];
    cspline_node_derivatives(w->xa, col, nx, der);
    for (int i = 0; i < nx; ++i) w->zxy[i * ny + j] = der[i];
  }
}

/* Bicubic::interpolate_ bivariate/bicubic.hpp:89-186 */
static double bicubic_eval(const window *w, double x, double y) {
  int i0, i1, j0, j1;
  if (!bracket_search(w->xa, w->nx, x, &i0, &i1) ||
      !bracket_search(w->ya, w->ny, y, &j0, &j1)) {
    return NAN;
  }
  const int ny = w->ny;
#define AT(m, i, j) ((m)[(i) * ny + (j)])
  const double x0 = w->xa[i0], x1 = w->xa[i1];
  const double y0 = w->ya[j0], y1 = w->ya[j1];
  const double dx = x1 - x0;
  const double dy = y1 - y0;
  const double dxdy = dx * dy;
  const double t = (x - x0) / dx;
  const double u = (y - y0) / dy;
  const double z00 = AT(w->za, i0, j0), z01 = AT(w->za, i0, j1);
  const double z10 = AT(w->za, i1, j0), z11 = AT(w->za, i1, j1);
  const double zx00 = AT(w->zx, i0, j0) * dx, zx01 = AT(w->zx, i0, j1) * dx;
  const double zx10 = AT(w->zx, i1, j0) * dx, zx11 = AT(w->zx, i1, j1) * dx;
  const double zy00 = AT(w->zy, i0, j0) * dy, zy01 = AT(w->zy, i0, j1) * dy;
  const double zy10 = AT(w->zy, i1, j0) * dy, zy11 = AT(w->zy, i1, j1) * dy;
  const double zxy00 = AT(w->zxy, i0, j0) * dxdy, zxy01 = AT(w->zxy, i0, j1) * dxdy;
  const double zxy10 = AT(w->zxy, i1, j0) * dxdy, zxy11 = AT(w->zxy, i1, j1) * dxdy;
#undef AT
  const double t2 = t * t, t3 = t2 * t;
  const double u2 = u * u, u3 = u2 * u;
  const double p0 = z00 + u * zy00 + u2 * fma(3.0, z01 - z00, -2.0 * zy00 - zy01) +
                    u3 * (2.0 * (z00 - z01) + zy00 + zy01);
  const double p1 = zx00 + u * zxy00 +
                    u2 * fma(3.0, zx01 - zx00, -2.0 * zxy00 - zxy01) +
                    u3 * (2.0 * (zx00 - zx01) + zxy00 + zxy01);
  const double a20 = fma(3.0, z10 - z00, -2.0 * zx00 - zx10);
  const double a21 = fma(3.0, zy10 - zy00, -2.0 * zxy00 - zxy10);
  const double a22 = 9.0 * (z00 - z01 - z10 + z11) +
                     6.0 * (zx00 - zx01 + zy00 - zy10) +
                     3.0 * (zx10 - zx11 + zy01 - zy11) + 4.0 * zxy00 +
                     2.0 * (zxy01 + zxy10) + zxy11;
  const double a23 = 6.0 * (z01 - z00 + z10 - z11) + 4.0 * (zx01 - zx00) +
                     3.0 * (zy10 - zy00 - zy01 + zy11) +
                     2.0 * (zx11 - zx10 - zxy00 - zxy01) - zxy10 - zxy11;
  const double p2 = a20 + u * a21 + u2 * a22 + u3 * a23;
  const double a30 = 2.0 * (z00 - z10) + zx00 + zx10;
  const double a31 = zxy00 + zxy10 + 2.0 * (zy00 - zy10);
  const double a32 = 6.0 * (z01 - z00 + z10 - z11) + 4.0 * (-zy00 + zy10) +
                     3.0 * (zx01 - zx00 - zx10 + zx11) +
                     2.0 * (zy11 - zy01 - zxy00 - zxy10) - zxy01 - zxy11;
  const double a33 = 4.0 * (z00 - z01 - z10 + z11) +
                     2.0 * (zx00 - zx01 + zx10 - zx11 + zy00 + zy01 - zy10 - zy11) +
                     zxy00 + zxy01 + zxy10 + zxy11;
  const double p3 = a30 + u * a31 + u2 * a32 + u3 * a33;
  return p0 + t * p1 + t2 * p2 + t3 * p3;
}

/* windowed Bilinear::interpolate_ bivariate/bilinear.hpp:34-83 */
static double wbilinear_eval(const window *w, double x, double y) {
  int i0, i1, j0, j1;
  if (!bracket_search(w->xa, w->nx, x, &i0, &i1) ||
      !bracket_search(w->ya, w->ny, y, &j0, &j1)) {
    return NAN;
  }
  const int ny = w->ny;
  const double x0 = w->xa[i0], x1 = w->xa[i1];
  const double y0 = w->ya[j0], y1 = w->ya[j1];
  const double t = (x - x0) / (x1 - x0);
  const double u = (y - y0) / (y1 - y0);
  const double tc = 1.0 - t, uc = 1.0 - u;
  const double z00 = w->za[i0 * ny + j0], z10 = w->za[i1 * ny + j0];
  const double z01 = w->za[i0 * ny + j1], z11 = w->za[i1 * ny + j1];
  const double bottom = fma(t, z10, tc * z00);
  const double top = fma(t, z11, tc * z01);
  return fma(u, top, uc * bottom);
}

static double window_eval(const window *w, int method, double x, double y) {
  if (!w->valid) { /* Bivariate::operator() interpolate/bivariate.hpp:133-138 */
    return NAN;
  }
  return method == PYI_W_BICUBIC ? bicubic_eval(w, x, y) : wbilinear_eval(w, x, y);
}

double orc_window_eval(int method, const double *xa, int nx, const double *ya,
                       int ny, const double *za, double x, double y) {
  if (nx > ORC_MAXW || ny > ORC_MAXW) return NAN;
  window w;
  w.nx = nx;
  w.ny = ny;
  memcpy(w.xa, xa, sizeof(double) * (size_t)nx);
  memcpy(w.ya, ya, sizeof(double) * (size_t)ny);
  const size_t cells = (size_t)nx * (size_t)ny;
  double *buf = (double *)malloc(sizeof(double) * 4 * cells);
  w.za = buf;
  w.zx = buf + cells;
  w.zy = buf + 2 * cells;
  w.zxy = buf + 3 * cells;
  memcpy(w.za, za, sizeof(double) * cells);
  window_prepare(&w, method);
  const double r = window_eval(&w, method, x, y);
  free(buf);
  return r;
}

typedef struct win_job {
  const orc_axis *const *axes;
  int dtype;
  const void *data;
  const pyi_windowed_config *cfg;
  int stateless;
  const double *x, *y;
  void *out;
} win_job;

/* Per-thread state of the reference: InterpolationCache + interpolator
 * (pybind/windowed/bivariate.hpp:99-102). */
typedef struct win_state {
  window w;
  int has_domain;              /* cache.has_domain() */
  double dmin[2], dmax[2];     /* DomainBounds cache.hpp:23-33 */
  int cache_valid;             /* cache.is_valid(): no NaN in the window */
  int npx, npy;                /* cache.points_per_dim */
} win_state;

/* load_cache_generic cache_loader.hpp:46-172 (2-D).  Returns success;
 * *err_axis >= 0 only when bounds_error and the coordinate is truly outside
 * that axis (:69-80). */
static int window_load(const win_job *J, win_state *S, double xq, double yq,
                       int *err_axis) {
  const orc_axis *ax = J->axes[0];
  const orc_axis *ay = J->axes[1];
  int64_t wx[ORC_MAXW], wy[ORC_MAXW];
  int cx = 0, cy = 0, x0, x1, y0, y1;
  if (!axis_window_f64(ax, xq, J->cfg->half_x, J->cfg->boundary, wx, &cx, &x0, &x1)) {
    if (J->cfg->bounds_error && !axis_contains_f64(ax, xq)) *err_axis = 0;
    return 0;
  }
  if (!axis_window_f64(ay, yq, J->cfg->half_y, J->cfg->boundary, wy, &cy, &y0, &y1)) {
    if (J->cfg->bounds_error && !axis_contains_f64(ay, yq)) *err_axis = 1;
    return 0;
  }
  window *w = &S->w;
  w->nx = cx;
  w->ny = cy;
  /* coordinates; x re-centred on the query when the axis is periodic
   * (:110-133) */
  const double period = ax->periodic ? ax->period_f : 0.0;
  for (int k = 0; k < cx; ++k) {
    double raw = coord_f64(ax, wx[k]);
    if (period != 0.0) {
      raw = normalize_period_half_f64(raw, xq, period);
    }
    w->xa[k] = raw;
  }
  for (int k = 0; k < cy; ++k) {
    w->ya[k] = coord_f64(ay, wy[k]);
  }
  const int64_t n1 = ay->size;
  int any_nan = 0;
  for (int i = 0; i < cx; ++i) {
    for (int j = 0; j < cy; ++j) {
      const double v = grid_value(J->dtype, J->data, wx[i] * n1 + wy[j]);
      w->za[i * cy + j] = v;
      any_nan |= isnan(v);
    }
  }
  /* finalize -> update_domain_for_axis cache.hpp:366-385 */
  const int vx = x0 >= 0 && x1 >= 0 && x0 < cx && x1 < cx;
  const int vy = y0 >= 0 && y1 >= 0 && y0 < cy && y1 < cy;
  S->has_domain = vx && vy;
  if (vx) {
    S->dmin[0] = w->xa[x0];
    S->dmax[0] = w->xa[x1];
  }
  if (vy) {
    S->dmin[1] = w->ya[y0];
    S->dmax[1] = w->ya[y1];
  }
  S->cache_valid = !any_nan;
  return S->cache_valid;
}

/* windowed::pybind::bivariate_single pybind/windowed/bivariate.hpp:46-76 +
 * update_cache_if_needed cache_loader.hpp:188-212.
 * Returns 1 with *res, 0 for "leave NaN", -1 for a bounds error. */
static int window_point(const win_job *J, win_state *S, double x, double y,
                        double *res, int *err_axis) {
  int was_updated = 0;
  int success = 1;
  const int in_domain = S->has_domain && x >= S->dmin[0] && x <= S->dmax[0] &&
                        y >= S->dmin[1] && y <= S->dmax[1];
  if (!in_domain) {
    was_updated = 1;
    *err_axis = -1;
    success = window_load(J, S, x, y, err_axis);
  }
  if (!success) {
    return *err_axis >= 0 ? -1 : 0;
  }
  if (!S->cache_valid) {
    return 0;
  }
  if (was_updated) {
    window_prepare(&S->w, J->cfg->method);
  }
  *res = window_eval(&S->w, J->cfg->method, x, y);
  return 1;
}

static void win_range(const void *job, int64_t start, int64_t end,
                      chunk_status *st) {
  const win_job *J = (const win_job *)job;
  const size_t cells = (size_t)(2 * J->cfg->half_x) * (size_t)(2 * J->cfg->half_y);
  double *buf = (double *)malloc(sizeof(double) * 4 * cells);
  win_state S;
  memset(&S, 0, sizeof S);
  S.w.za = buf;
  S.w.zx = buf + cells;
  S.w.zy = buf + 2 * cells;
  S.w.zxy = buf + 3 * cells;
  S.cache_valid = 0;
  const int f64 = result_is_f64(J->dtype);
  for (int64_t k = start; k < end; ++k) {
    if (J->stateless) {
      S.has_domain = 0;
    }
    double v = NAN;
    int err_axis = -1;
    const int rc = window_point(J, &S, J->x[k], J->y[k], &v, &err_axis);
    if (rc < 0) {
      st->oob_index = k;
      st->oob_axis = err_axis;
      break;
    }
    if (rc > 0) {
      if (f64) {
        ((double *)J->out)[k] = v;
      } else {
        ((float *)J->out)[k] = (float)v;
      }
    }
  }
  free(buf);
}

int orc_windowed2d(const orc_axis *const *axes, int dtype, const void *data,
                   const pyi_windowed_config *cfg, int stateless, int64_t n,
                   const double *x, const double *y, void *out,
                   int64_t *oob_index, int *oob_axis, char *msg,
                   size_t msglen) {
  if (cfg->half_x < 1 || cfg->half_y < 1 || 2 * cfg->half_x > ORC_MAXW ||
      2 * cfg->half_y > ORC_MAXW || dtype < PYI_I8 || dtype > PYI_F64) {
    return PYI_BAD_ARG;
  }
  win_job J;
  J.axes = axes;
  J.dtype = dtype;
  J.data = data;
  J.cfg = cfg;
  J.stateless = stateless;
  J.x = x;
  J.y = y;
  J.out = out;
  if (result_is_f64(dtype)) {
    for (int64_t k = 0; k < n; ++k) ((double *)out)[k] = NAN;
  } else {
    for (int64_t k = 0; k < n; ++k) ((float *)out)[k] = NAN;
  }
  const chunk_status st = parallel_ranges(n, win_range, &J, cfg->num_threads);
  if (st.oob_index >= 0) {
    if (oob_index) *oob_index = st.oob_index;
    if (oob_axis) *oob_axis = st.oob_axis;
    geo_job G;
    memset(&G, 0, sizeof G);
    G.axes = axes;
    G.x = x;
    G.y = y;
    bounds_message(&G, st.oob_axis, st.oob_index, msg, msglen);
    return PYI_OOB;
  }
  return PYI_OK;
}


/* the rest of the windowed path and dateutils::convert */
#include "pyi_oracle_wnd.inc"

/* grid filling */
#include "pyi_oracle_fill.inc"
