#pragma once
#include <cstdint>
#define CEREAL_NVP(x) x
#define CEREAL_CLASS_VERSION(T, v)
namespace cereal {
template <class T> T &&make_nvp(const char *, T &&t) { return static_cast<T &&>(t); }
} // namespace cereal
