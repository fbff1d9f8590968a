#pragma once
#include <string>
namespace fmt {
template <typename T, typename Char = char, typename Enable = void> struct formatter;
template <typename... Args> std::string format(const char *, Args &&...) { return {}; }
template <typename R> std::string join(R &&, const char *) { return {}; }
} // namespace fmt
