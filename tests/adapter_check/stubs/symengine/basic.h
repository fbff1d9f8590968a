// stand-in declarations of the SymEngine types sme/symbolic.hpp holds as members
#pragma once
#include <memory>
#include <vector>
namespace SymEngine {
class Basic {};
class Symbol : public Basic {};
template <class T> class RCP {
public:
  RCP() = default;
  T *operator->() const;
};
using vec_basic = std::vector<RCP<const Basic>>;
class LLVMDoubleVisitor {};
class SbmlParser {};
} // namespace SymEngine
