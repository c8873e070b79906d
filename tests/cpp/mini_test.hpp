// mini_test.hpp -- 40-line stand-in for GoogleTest (not on this machine): TEST_F-style fixtures are
// plain structs, EXPECT_* count failures, main() reports and returns non-zero on any failure.
#pragma once
#include <cstdio>
#include <functional>
#include <string>
#include <vector>

namespace mini {
struct registry {
    std::vector<std::pair<std::string, std::function<void()>>> tests;
    int failures = 0;
    static registry &get() { static registry r; return r; }
};
struct adder { adder(const char *n, std::function<void()> f) { registry::get().tests.emplace_back(n, std::move(f)); } };
inline int run_all() {
    auto &r = registry::get();
    for (auto &t : r.tests) {
        int before = r.failures;
        try { t.second(); } catch (const std::exception &e) { std::printf("  exception: %s\n", e.what()); ++r.failures; }
        std::printf("[%s] %s\n", r.failures == before ? " OK " : "FAIL", t.first.c_str());
    }
    std::printf("%zu tests, %d failures\n", r.tests.size(), r.failures);
    return r.failures ? 1 : 0;
}
} // namespace mini
#define MINI_TEST(name) static void name(); static mini::adder name##_adder(#name, name); static void name()
#define EXPECT_TRUE(x) do { if (!(x)) { std::printf("  %s:%d EXPECT_TRUE(%s)\n", __FILE__, __LINE__, #x); ++mini::registry::get().failures; } } while (0)
#define EXPECT_EQ(a, b) do { if (!((a) == (b))) { std::printf("  %s:%d EXPECT_EQ(%s, %s)\n", __FILE__, __LINE__, #a, #b); ++mini::registry::get().failures; } } while (0)
