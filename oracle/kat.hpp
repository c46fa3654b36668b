// ORACLE — TEST INFRASTRUCTURE ONLY.
// Tiny known-answer-test harness: the reference's own #[test] functions (for the hot path),
// restated against the oracle. Each test cites the reference test it ports. `./kat` prints one
// line per test ("ok <name>" / "FAIL <name>: <why>") and exits non-zero on any failure;
// tests/test_oracle_kat.py runs it under pytest.
#pragma once
#include <cstdio>
#include <functional>
#include <sstream>
#include <string>
#include <vector>

#include "obj.hpp"
#include "world.hpp"

namespace kat {
struct Test {
  std::string name;
  std::function<void()> fn;
};
inline std::vector<Test>& registry() {
  static std::vector<Test> r;
  return r;
}
struct Reg {
  Reg(const char* n, std::function<void()> f) { registry().push_back({n, std::move(f)}); }
};
struct Failure {
  std::string msg;
};
inline void check(bool ok, const char* expr, const char* file, int line) {
  if (!ok) {
    std::ostringstream s;
    s << file << ":" << line << ": " << expr;
    throw Failure{s.str()};
  }
}
// copy-modify-store of a shape's material (the reference mutates get_base_mut().material in place)
template <class F>
inline void edit_material(orc::Shape& s, F f) {
  orc::Material m = s.material();
  f(m);
  s.set_material(m);
}
}  // namespace kat

#define KAT(name)                                   \
  static void kat_##name();                         \
  static kat::Reg kat_reg_##name(#name, kat_##name); \
  static void kat_##name()
#define CHECK(e) kat::check((e), #e, __FILE__, __LINE__)
#define CHECK_EQ(a, b) kat::check((a) == (b), #a " == " #b, __FILE__, __LINE__)
#define CHECK_NEAR(a, b) kat::check(orc::equal((a), (b)), #a " ~= " #b, __FILE__, __LINE__)
#define CHECK_THROWS(stmt)                                                      \
  do {                                                                          \
    bool threw_ = false;                                                        \
    try { stmt; } catch (const kat::Failure&) { throw; } catch (...) { threw_ = true; } \
    kat::check(threw_, "throws: " #stmt, __FILE__, __LINE__);                   \
  } while (0)
