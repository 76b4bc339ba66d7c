// TEST INFRASTRUCTURE ONLY -- not part of the product.
//
// Stand-in for cxx/include/pyinterp/config/fill.hpp, found first on the include
// path of the oracle build.  The reference's header (and config/common.hpp
// below it) declares its builder methods with C++23 explicit object parameters
// ("this Derived self", config/fill.hpp:104-150, config/common.hpp:53-60),
// which g++ 13 -- the only host compiler of this image -- cannot parse.  The
// configuration objects hold no algorithm: this file gives the fill headers
//   cxx/include/pyinterp/fill/{gauss_seidel,multi_grid,loess}.hpp
// the accessors they read (same names, same default values, same enumerator
// order) so that THEY compile unmodified.
#pragma once

#include <cstddef>
#include <cstdint>

namespace pyinterp::config::fill {

// config/fill.hpp:18-23
enum class FirstGuess : uint8_t {
  kZonalAverage,
  kZero,
};

// config/fill.hpp:26-30
enum class LoessValueType : std::uint8_t {
  kUndefined,
  kDefined,
  kAll
};

// config/fill.hpp:67-160 (FillBase) flattened
struct FillBase {
  FirstGuess first_guess_{FirstGuess::kZonalAverage};
  bool is_periodic_{true};
  uint32_t max_iterations_{500};
  double epsilon_{1e-4};
  size_t num_threads_{1};
  [[nodiscard]] auto first_guess() const -> FirstGuess { return first_guess_; }
  [[nodiscard]] auto is_periodic() const -> bool { return is_periodic_; }
  [[nodiscard]] auto max_iterations() const -> uint32_t { return max_iterations_; }
  [[nodiscard]] auto epsilon() const -> double { return epsilon_; }
  [[nodiscard]] auto num_threads() const -> size_t { return num_threads_; }
};

// config/fill.hpp:255-279
struct GaussSeidel : FillBase {
  GaussSeidel() { max_iterations_ = 2000; }
  double relaxation_{1.0};
  [[nodiscard]] auto relaxation() const -> double { return relaxation_; }
};

// config/fill.hpp:282-324
struct Multigrid : FillBase {
  Multigrid() { max_iterations_ = 100; }
  uint32_t pre_smooth_{2};
  uint32_t post_smooth_{2};
  [[nodiscard]] auto pre_smooth() const -> uint32_t { return pre_smooth_; }
  [[nodiscard]] auto post_smooth() const -> uint32_t { return post_smooth_; }
};

// config/fill.hpp:164-219
struct Loess : FillBase {
  LoessValueType value_type_{LoessValueType::kUndefined};
  uint32_t nx_{3};
  uint32_t ny_{3};
  [[nodiscard]] auto value_type() const -> LoessValueType { return value_type_; }
  [[nodiscard]] auto nx() const -> uint32_t { return nx_; }
  [[nodiscard]] auto ny() const -> uint32_t { return ny_; }
};

}  // namespace pyinterp::config::fill
