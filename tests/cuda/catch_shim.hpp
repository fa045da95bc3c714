// catch_shim.hpp -- the few pieces of the Catch2 v3 surface tests/cuda/test_flash_attention.cu uses
// (TEST_CASE with tags, SECTION, REQUIRE / CHECK, a "[tag]" or name-substring filter on the
// command line, --list-tests), so that the binary builds where Catch2 is not installed.  The
// reference vendors Catch2 v3.4.0 as a git submodule (Makefile:25-33 builds
// third_party/catch2/extras/catch_amalgamated.cpp); with a checkout at hand build against the real
// thing instead:  make CATCH2_DIR=/path/to/catch2 build/tests   (defines FA_USE_CATCH2).
// Not a copy of Catch2: no matchers, generators, reporters or exception translation.
#pragma once
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

namespace catch_shim {
struct Case {
  const char *name, *tags;
  void (*fn)();
};
inline std::vector<Case> &registry() {
  static std::vector<Case> r;
  return r;
}
struct Registrar {
  Registrar(const char *name, const char *tags, void (*fn)()) { registry().push_back({name, tags, fn}); }
};
struct State {
  int assertions = 0, failed = 0;
  bool abort_case = false;
};
inline State &state() {
  static State s;
  return s;
}
struct AbortCase {};
inline void check(bool ok, const char *expr, const char *file, int line, bool require) {
  ++state().assertions;
  if (ok) return;
  ++state().failed;
  std::printf("%s:%d: FAILED: %s( %s )\n", file, line, require ? "REQUIRE" : "CHECK", expr);
  if (require) throw AbortCase{};
}
}  // namespace catch_shim

#define FA_CAT2(a, b) a##b
#define FA_CAT(a, b) FA_CAT2(a, b)
#define FA_TEST_CASE_IMPL(fn, ...)                                   \
  static void fn();                                                  \
  static catch_shim::Registrar FA_CAT(fn, _reg)(__VA_ARGS__, &fn);   \
  static void fn()
#define TEST_CASE(...) FA_TEST_CASE_IMPL(FA_CAT(fa_test_case_, __LINE__), __VA_ARGS__)
#define SECTION(name) if (std::printf("  section: %s\n", name), true)
#define REQUIRE(expr) catch_shim::check(static_cast<bool>(expr), #expr, __FILE__, __LINE__, true)
#define CHECK(expr) catch_shim::check(static_cast<bool>(expr), #expr, __FILE__, __LINE__, false)

#ifdef FA_CATCH_SHIM_MAIN
int main(int argc, char **argv) {
  std::string filter = argc > 1 ? argv[1] : "";
  if (filter == "--list-tests") {
    for (auto &c : catch_shim::registry()) std::printf("%s\n    %s\n", c.name, c.tags);
    return 0;
  }
  int run = 0, failed_cases = 0;
  for (auto &c : catch_shim::registry()) {
    const bool by_tag = !filter.empty() && filter[0] == '[';
    // like Catch2: cases tagged [.] or [!benchmark]-style hidden tags run only when asked for by tag
    const bool hidden = std::strstr(c.tags, "[.") != nullptr;
    if (filter.empty() ? hidden : (by_tag ? std::strstr(c.tags, filter.c_str()) == nullptr
                                          : std::strstr(c.name, filter.c_str()) == nullptr))
      continue;
    ++run;
    std::printf("test case: %s %s\n", c.name, c.tags);
    const int before = catch_shim::state().failed;
    try {
      c.fn();
    } catch (const catch_shim::AbortCase &) {
    }
    if (catch_shim::state().failed != before) ++failed_cases;
  }
  std::printf("%d test cases, %d assertions; %d cases run, %d failed\n", run, catch_shim::state().assertions, run,
              failed_cases);
  return failed_cases;
}
#endif
