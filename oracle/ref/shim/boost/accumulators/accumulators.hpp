// TEST INFRASTRUCTURE ONLY -- not part of the product.
//
// Minimal stand-in for Boost.Accumulators so that the reference header
//   cxx/include/pyinterp/fill/helpers.hpp (set_zonal_average, lines 27-75)
// compiles UNMODIFIED in this container (Boost >= 1.90 is a system dependency
// of the reference, CMakeLists.txt:569, and is not installed here).
//
// Only accumulator_set<T, stats<tag::count, tag::mean>> is provided.  It
// restates the published algorithm of boost/accumulators/statistics/
// {count,sum,mean}.hpp: tag::mean is the LAZY mean (mean_impl), i.e.
//   sum   : starts at T(), `sum += sample` for every sample, in call order;
//   count : std::size_t incremented per sample;
//   mean  : numeric::fdiv(sum, count) = sum / count with count converted to
//           the sample type by the usual arithmetic conversions
//           (numeric/functional.hpp, fdiv_base<Left, Right>).
// (tag::immediate_mean, the running-update variant, is NOT what helpers.hpp
// asks for.)
#pragma once

#include <cstddef>

namespace boost::accumulators {

namespace tag {
struct count {};
struct mean {};
}  // namespace tag

template <typename... Stats>
struct stats {};

template <typename Sample, typename Features>
class accumulator_set {
 public:
  void operator()(const Sample& sample) {
    sum_ += sample;
    ++count_;
  }
  [[nodiscard]] auto count_value() const -> std::size_t { return count_; }
  [[nodiscard]] auto mean_value() const -> Sample { return sum_ / count_; }

 private:
  Sample sum_{};
  std::size_t count_{0};
};

template <typename Sample, typename Features>
auto count(const accumulator_set<Sample, Features>& acc) -> std::size_t {
  return acc.count_value();
}
template <typename Sample, typename Features>
auto mean(const accumulator_set<Sample, Features>& acc) -> Sample {
  return acc.mean_value();
}

}  // namespace boost::accumulators
