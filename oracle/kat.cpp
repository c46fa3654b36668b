// ORACLE — TEST INFRASTRUCTURE ONLY. Runner for the known-answer tests (see kat.hpp).
#include "kat.hpp"

int main(int argc, char** argv) {
  std::string filter = argc > 1 ? argv[1] : "";
  int failed = 0, ran = 0;
  for (auto& t : kat::registry()) {
    if (!filter.empty() && t.name.find(filter) == std::string::npos) continue;
    ran++;
    try {
      t.fn();
      std::printf("ok %s\n", t.name.c_str());
    } catch (const kat::Failure& f) {
      failed++;
      std::printf("FAIL %s: %s\n", t.name.c_str(), f.msg.c_str());
    } catch (const std::exception& e) {
      failed++;
      std::printf("FAIL %s: exception: %s\n", t.name.c_str(), e.what());
    }
  }
  std::printf("# %d tests, %d failed\n", ran, failed);
  return failed ? 1 : 0;
}
