// Minimal googletest-compatible header: just the surface the reference's suites use
// (TEST, ASSERT_EQ with `<<` messages, ASSERT_TRUE, a main like gtest_main).  googletest is
// fetched from the network by the reference's CMake (CMakeLists.txt.in:6-8); there is no
// network and no gtest on the build or GPU box, so the unmodified test sources are compiled
// against this instead.  Output mimics gtest's so logs read the same.
#ifndef HGB_MINI_GTEST_H
#define HGB_MINI_GTEST_H
#include <cstdio>
#include <iostream>
#include <sstream>
#include <string>
#include <vector>

namespace testing {
struct TestInfo { const char *suite; const char *name; void (*fn)(bool *); };
inline std::vector<TestInfo> &registry() { static std::vector<TestInfo> r; return r; }
struct Registrar {
    Registrar(const char *s, const char *n, void (*f)(bool *)) { registry().push_back({s, n, f}); }
};
// collects the streamed message; reports when the assertion statement ends
struct Failure {
    std::ostringstream msg;
    std::string head;
    bool *flag;
    Failure(bool *f, const std::string &h) : head(h), flag(f) {}
    Failure(const Failure &o) : head(o.head), flag(o.flag) { msg << o.msg.str(); }
    ~Failure() {
        *flag = true;
        std::cerr << head << msg.str() << std::endl;
    }
    template <class T> Failure &operator<<(const T &v) { msg << v; return *this; }
};
struct Voidify { void operator=(const Failure &) {} };
template <class A, class B>
std::string describe(const char *ea, const char *eb, const A &a, const B &b, const char *file, int line) {
    std::ostringstream o;
    o << file << ":" << line << ": Failure\nExpected equality of these values:\n  " << ea << "\n    Which is: "
      << a << "\n  " << eb << "\n    Which is: " << b << "\n";
    return o.str();
}
inline int run_all() {
    int failed = 0;
    std::printf("[==========] Running %zu tests.\n", registry().size());
    for (const TestInfo &t : registry()) {
        std::printf("[ RUN      ] %s.%s\n", t.suite, t.name);
        bool bad = false;
        t.fn(&bad);
        std::printf(bad ? "[  FAILED  ] %s.%s\n" : "[       OK ] %s.%s\n", t.suite, t.name);
        failed += bad;
    }
    std::printf("[==========] %zu tests ran.\n[  PASSED  ] %zu tests.\n", registry().size(),
                registry().size() - failed);
    if (failed) std::printf("[  FAILED  ] %d tests.\n", failed);
    return failed ? 1 : 0;
}
}  // namespace testing

#define TEST(suite, name)                                                                      \
    static void suite##_##name##_body(bool *gtest_failed_);                                    \
    static ::testing::Registrar suite##_##name##_reg(#suite, #name, suite##_##name##_body);    \
    static void suite##_##name##_body(bool *gtest_failed_)

// ASSERT_*: on failure report, mark the test failed and return from the test body
#define HGB_ASSERT_(cond, text)                                                                \
    if (cond) {                                                                                \
    } else                                                                                     \
        return ::testing::Voidify() = ::testing::Failure(gtest_failed_, text)

#define ASSERT_EQ(a, b) \
    HGB_ASSERT_(((a) == (b)), ::testing::describe(#a, #b, (a), (b), __FILE__, __LINE__))
#define ASSERT_TRUE(c) \
    HGB_ASSERT_((c), std::string(__FILE__) + ":" + std::to_string(__LINE__) + ": Failure\nValue of: " #c "\n  Actual: false\nExpected: true\n")

#ifndef HGB_GTEST_NO_MAIN
int main() { return ::testing::run_all(); }
#endif
#endif  // HGB_MINI_GTEST_H
