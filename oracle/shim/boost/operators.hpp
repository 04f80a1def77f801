// Build shim for compiling the UNMODIFIED reference without Boost installed.
// Test infrastructure only (oracle/): it contains no arithmetic of its own --
// every operator below is the textbook "binary op from compound assignment"
// that boost/operators.hpp provides.  Only the templates the reference names are
// present (include/amber/prelude/vector3.h:38, aabb.h:37-38, image.h:35-36,
// pixel.h:49, accumulator.h:35, include/amber/rendering/object.h:31).
#pragma once

namespace boost {

#define AMBER_SHIM_BINARY1(NAME, OP)                                  \
  template <class T> struct NAME {                                    \
    friend T operator OP(T lhs, const T& rhs) { lhs OP## = rhs; return lhs; } \
  };
AMBER_SHIM_BINARY1(addable, +)
AMBER_SHIM_BINARY1(subtractable, -)
AMBER_SHIM_BINARY1(multipliable, *)
AMBER_SHIM_BINARY1(dividable, /)
#undef AMBER_SHIM_BINARY1

template <class T>
struct field_operators : addable<T>, subtractable<T>, multipliable<T>, dividable<T> {};

// two-type forms; + and * are commutative so U op T is provided as well
template <class T, class U> struct addable2 {
  friend T operator+(T lhs, const U& rhs) { lhs += rhs; return lhs; }
  friend T operator+(const U& lhs, T rhs) { rhs += lhs; return rhs; }
};
template <class T, class U> struct subtractable2 {
  friend T operator-(T lhs, const U& rhs) { lhs -= rhs; return lhs; }
};
template <class T, class U> struct multipliable2 {
  friend T operator*(T lhs, const U& rhs) { lhs *= rhs; return lhs; }
  friend T operator*(const U& lhs, T rhs) { rhs *= lhs; return rhs; }
};
template <class T, class U> struct dividable2 {
  friend T operator/(T lhs, const U& rhs) { lhs /= rhs; return lhs; }
};
template <class T, class U>
struct multiplicative2 : multipliable2<T, U>, dividable2<T, U> {};
template <class T, class U>
struct field_operators2 : addable2<T, U>, subtractable2<T, U>, multipliable2<T, U>, dividable2<T, U> {};

template <class T> struct equality_comparable {
  friend bool operator!=(const T& lhs, const T& rhs) { return !(lhs == rhs); }
};

}  // namespace boost
