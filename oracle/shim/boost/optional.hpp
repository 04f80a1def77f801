// Build shim (oracle/ test infrastructure): boost::optional as std::optional.
// Used by include/amber/prelude/algebra.h:32-37 and consumed at
// src/amber/scene/primitive_sphere.cc:82-88.  No arithmetic.
#pragma once
#include <optional>

namespace boost {
template <class T> using optional = std::optional<T>;
inline constexpr std::nullopt_t none = std::nullopt;
}  // namespace boost
