// pyi_datetime.cu -- K7: numpy datetime64 / timedelta64 resolution conversion
// on the device.
//
// Replaces dateutils::convert (cxx/include/pyinterp/dateutils.hpp:821-935), the
// serial pre-step the reference runs with the GIL held before every temporal
// trivariate / quadrivariate call (src/pybind/temporal_axis.cpp:78-98,216-220,
// pybind/grid_dispatch.hpp:190-195,244-249): every value is normalised to
// (days, seconds of day, fraction) and rebuilt in the target resolution with
// checked multiply / add (dateutils.hpp:50-65,94-118).  One thread per value;
// exact int64 arithmetic in the reference's own C types (the civil-calendar
// helpers narrow to int / unsigned, dateutils.hpp:540-607).  The first value
// (lowest index) whose checked arithmetic overflows is reported like the
// reference's std::overflow_error.
#include "pyi_common.cuh"

namespace pyi {

namespace {

constexpr long long kEpoch = 1970;
constexpr long long kDaysInWeek = 7;
constexpr long long kHoursInDay = 24;
constexpr long long kMinutesInDay = 1440;
constexpr long long kMonthsInYear = 12;
constexpr long long kSecondsInMinute = 60;
constexpr long long kSecondsInHour = 3600;
constexpr long long kSecondsInDay = 86400;

struct Date {
  int year;
  unsigned month;
  unsigned day;
};

// dateutils.hpp:560-582
__device__ __forceinline__ Date date_from_days(long long days) {
  days += 719468LL;
  const int era = static_cast<int>((days >= 0 ? days : days - 146096LL) / 146097LL);
  const unsigned doe = static_cast<unsigned>(days - era * 146097LL);
  const unsigned yoe = (doe - doe / 1460 + doe / 36524 - doe / 146096) / 365;
  const unsigned doy = doe - (365 * yoe + yoe / 4 - yoe / 100);
  const unsigned mon = (5 * doy + 2) / 153;
  const unsigned day = doy - (153 * mon + 2) / 5 + 1;
  const unsigned month = mon < 10 ? mon + 3 : mon - 9;
  Date d;
  d.year = (static_cast<int>(yoe) + era * 400) + static_cast<int>(month <= 2);
  d.month = month;
  d.day = day;
  return d;
}

// dateutils.hpp:588-607
__device__ __forceinline__ long long days_from_date(Date date) {
  int y = date.year;
  unsigned m = date.month;
  const unsigned d = date.day;
  if (m <= 2) {
    y -= 1;
    m += 12;
  }
  const int era = (y >= 0 ? y : y - 399) / 400;
  const unsigned yoe = static_cast<unsigned>(y - era * 400);
  const unsigned doy = (153 * (m - 3) + 2) / 5 + d - 1;
  const unsigned doe = yoe * 365 + yoe / 4 - yoe / 100 + doy;
  return era * 146097 + static_cast<long long>(doe) - 719468;
}

// dateutils.hpp:540-557
__device__ __forceinline__ Date date_from_years(long long years) {
  Date d;
  d.year = static_cast<int>(kEpoch + years);
  d.month = 1;
  d.day = 1;
  return d;
}
__device__ __forceinline__ Date date_from_months(long long months) {
  long long year = months / kMonthsInYear;
  long long month = months % kMonthsInYear;
  if (month != 0 && months < 0) {
    --year;
    month += kMonthsInYear;
  }
  Date d;
  d.year = static_cast<int>(kEpoch + year);
  d.month = static_cast<unsigned>(month + 1);
  d.day = 1;
  return d;
}

// detail::ckd_mul / ckd_add (dateutils.hpp:50-65,94-118); `bad` records the
// first failing operation: 1 multiplication, 2 addition
__device__ __forceinline__ long long ckd_mul(long long a, long long b, int& bad) {
  const long long lo = a * b;
  const long long hi = __mul64hi(a, b);
  if (hi != (lo >> 63) && bad == 0) bad = 1;
  return lo;
}
__device__ __forceinline__ long long ckd_add(long long a, long long b, int& bad) {
  const long long r = static_cast<long long>(static_cast<unsigned long long>(a) +
                                             static_cast<unsigned long long>(b));
  if ((((a ^ r) & (b ^ r)) < 0) && bad == 0) bad = 2;
  return r;
}

__device__ __forceinline__ long long order_of_magnitude(int resolution) {
  switch (resolution) {  // dateutils.hpp:320-339
    case PYI_RES_SECOND: return 1LL;
    case PYI_RES_MILLISECOND: return 1000LL;
    case PYI_RES_MICROSECOND: return 1000000LL;
    case PYI_RES_NANOSECOND: return 1000000000LL;
    case PYI_RES_PICOSECOND: return 1000000000000LL;
    case PYI_RES_FEMTOSECOND: return 1000000000000000LL;
    default: return 1000000000000000000LL;
  }
}

// a / 1000^k truncating, k = 0 .. 6 (warp-uniform switch): a division by a
// run-time 64-bit divisor is a ~100-instruction subroutine, by a constant a
// multiply-high -- the kernel went from 17-24 % to the HBM roofline's range
__device__ __forceinline__ long long div_pow1000(long long a, int k) {
  switch (k) {
    case 0: return a;
    case 1: return a / 1000LL;
    case 2: return a / 1000000LL;
    case 3: return a / 1000000000LL;
    case 4: return a / 1000000000000LL;
    case 5: return a / 1000000000000000LL;
    default: return a / 1000000000000000000LL;
  }
}

// one value of dateutils::convert (dateutils.hpp:845-933)
__device__ __forceinline__ long long convert_one(long long value, int src, int dst,
                                                 long long src_mag, long long dst_mag, int& bad) {
  long long days = 0, seconds = 0, fractional = 0;
  switch (src) {
    case PYI_RES_YEAR:
      days = days_from_date(date_from_years(value));
      break;
    case PYI_RES_MONTH:
      days = days_from_date(date_from_months(value));
      break;
    case PYI_RES_WEEK:
      days = value * kDaysInWeek;
      break;
    case PYI_RES_DAY:
      days = value;
      break;
    case PYI_RES_HOUR:
      days = value / kHoursInDay;
      seconds = (value % kHoursInDay) * kSecondsInHour;
      if (seconds < 0) {
        days--;
        seconds += kSecondsInDay;
      }
      break;
    case PYI_RES_MINUTE:
      days = value / kMinutesInDay;
      seconds = (value % kMinutesInDay) * kSecondsInMinute;
      if (seconds < 0) {
        days--;
        seconds += kSecondsInDay;
      }
      break;
    default: {
      // FractionalSeconds::epoch + days_since_epoch (dateutils.hpp:402-412,473-482)
      long long sec = div_pow1000(value, src - PYI_RES_SECOND);  // value / src_mag
      long long frac = value - sec * src_mag;                     // value % src_mag
      if (frac < 0) {
        frac += src_mag;
        sec -= 1;
      }
      days = sec / kSecondsInDay;
      seconds = sec % kSecondsInDay;
      if (seconds < 0) {
        --days;
        seconds += kSecondsInDay;
      }
      fractional = frac;
      break;
    }
  }
  switch (dst) {
    case PYI_RES_YEAR:
      return date_from_days(days).year - kEpoch;
    case PYI_RES_MONTH: {
      const Date date = date_from_days(days);
      return (date.year - kEpoch) * kMonthsInYear + (date.month - 1);
    }
    case PYI_RES_WEEK:
      return days / kDaysInWeek;
    case PYI_RES_DAY:
      return days;
    case PYI_RES_HOUR:
      return ckd_add(ckd_mul(days, kHoursInDay, bad), seconds / kSecondsInHour, bad);
    case PYI_RES_MINUTE:
      return ckd_add(ckd_mul(days, kMinutesInDay, bad), seconds / kSecondsInMinute, bad);
    default: {
      long long v = ckd_add(ckd_mul(ckd_mul(days, kSecondsInDay, bad), dst_mag, bad),
                            ckd_mul(seconds, dst_mag, bad), bad);
      if (src >= PYI_RES_SECOND) {
        // FractionalSeconds::cast (dateutils.hpp:422-427)
        // dst_mag / src_mag and src_mag / dst_mag are powers of 1000
        const long long c = src_mag <= dst_mag ? ckd_mul(order_of_magnitude(PYI_RES_SECOND + dst - src), fractional, bad)
                                               : div_pow1000(fractional, src - dst);
        v = ckd_add(v, c, bad);
      }
      return v;
    }
  }
}

// Closed form of convert_one between two FIXED-LENGTH units (day .. picosecond):
// the normalisation to (days, seconds, fraction) and back is value * F when the
// source unit is the longer one (F = ratio of the unit lengths) and
// floor(value / G) when it is the shorter one -- as long as none of the
// reference's checked operations can overflow, which `safe_abs` guarantees
// (every intermediate is bounded by |value| * max(F, 1) + 3 * 86400 * dst_mag).
// Values beyond it, and every other unit pair (calendar units, weeks, femto- /
// attoseconds whose day products overflow for any negative value), take the
// checked sequence.  KIND 0: none, 1: multiply, 2: floor-divide.
struct FastPlan {
  int kind;
  long long factor;
  unsigned long long safe_abs;
  double inv_factor;
};

__device__ __forceinline__ long long floor_div(long long a, long long g, double inv_g) {
  using U = unsigned long long;
  long long q = static_cast<long long>(floor(static_cast<double>(a) * inv_g));
  // |r| <= 2 g + 2^12: the quotient of the doubles is off by a few units at most; q g itself may
  // pass the int64 range by that much next to its ends, hence the wrapping (unsigned) products
  long long r = static_cast<long long>(static_cast<U>(a) - static_cast<U>(q) * static_cast<U>(g));
  const long long q2 = static_cast<long long>(floor(static_cast<double>(r) * inv_g));
  q += q2;
  r = static_cast<long long>(static_cast<U>(r) - static_cast<U>(q2) * static_cast<U>(g));
  if (r < 0) {
    --q;
  } else if (r >= g) {
    ++q;
  }
  return q;
}

template <int KIND>
__global__ void __launch_bounds__(kThreads)
datetime_convert_kernel(const long long* __restrict__ in, long long* __restrict__ out, long long n,
                        long long index_base, int src, int dst, const FastPlan plan,
                        unsigned long long* bad_key) {
  const long long src_mag = src >= PYI_RES_SECOND ? order_of_magnitude(src) : 1;
  const long long dst_mag = dst >= PYI_RES_SECOND ? order_of_magnitude(dst) : 0;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  unsigned long long key = kNoOob;
  for (long long k = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; k < n;
       k += stride) {
    int bad = 0;
    const long long value = __ldcs(in + k);
    long long v;
    const unsigned long long mag = value < 0 ? 0ULL - static_cast<unsigned long long>(value)
                                             : static_cast<unsigned long long>(value);
    if (KIND != 0 && mag <= plan.safe_abs) {
      v = KIND == 1 ? value * plan.factor : floor_div(value, plan.factor, plan.inv_factor);
    } else {
      v = convert_one(value, src, dst, src_mag, dst_mag, bad);
    }
    __stcs(out + k, v);
    if (bad != 0) {
      const unsigned long long kk =
          (static_cast<unsigned long long>(index_base + k) << 2) | static_cast<unsigned>(bad);
      key = kk < key ? kk : key;
    }
  }
  if (key != kNoOob) atomicMin(bad_key, key);
}

}  // namespace

cudaError_t launch_datetime_convert(const long long* in, long long* out, long long n,
                                    long long index_base, int src, int dst,
                                    unsigned long long* bad_key, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  const int blocks = grid_for(n, kThreads, 8);
  // unit lengths in picoseconds, day .. picosecond
  static const long long kLength[13] = {0, 0, 0, 86400000000000000LL, 3600000000000000LL, 60000000000000LL,
                                        1000000000000LL, 1000000000LL, 1000000LL, 1000LL, 1LL, 0, 0};
  FastPlan plan{0, 1, 0ULL, 1.0};
  if (src >= PYI_RES_DAY && src <= PYI_RES_PICOSECOND && dst >= PYI_RES_DAY && dst <= PYI_RES_PICOSECOND &&
      src != dst) {
    const long long ls = kLength[src], ld = kLength[dst];
    const long long dst_mag = dst >= PYI_RES_SECOND ? 1000000000000LL / ld : 1;
    const long long slack = 3 * 86400LL * dst_mag;  // <= 2.6e17
    const long long top = 9223372036854775807LL - slack;
    if (ls > ld) {
      plan.kind = 1;
      plan.factor = ls / ld;
      plan.safe_abs = static_cast<unsigned long long>(top / plan.factor);
    } else {
      plan.kind = 2;
      plan.factor = ld / ls;
      plan.safe_abs = static_cast<unsigned long long>(top);
    }
    plan.inv_factor = 1.0 / static_cast<double>(plan.factor);
  }
  if (plan.kind == 1) {
    datetime_convert_kernel<1><<<blocks, kThreads, 0, stream>>>(in, out, n, index_base, src, dst, plan, bad_key);
  } else if (plan.kind == 2) {
    datetime_convert_kernel<2><<<blocks, kThreads, 0, stream>>>(in, out, n, index_base, src, dst, plan, bad_key);
  } else {
    datetime_convert_kernel<0><<<blocks, kThreads, 0, stream>>>(in, out, n, index_base, src, dst, plan, bad_key);
  }
  g_launches.fetch_add(1);
  return cudaGetLastError();
}

}  // namespace pyi
