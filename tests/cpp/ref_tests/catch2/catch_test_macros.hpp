// Minimal stand-in for Catch2 v3 (the reference fetches it as a git submodule, .gitmodules:1-3; absent offline): TEST_CASE,
// REQUIRE and, in catch2/benchmark/catch_benchmark.hpp, BENCHMARK -- the three macros the reference's test sources use.
// TEST INFRASTRUCTURE ONLY.  The runner takes Catch2's "[tag]" filter as its first argument (AND of the tags given, a
// leading ~ excludes) and -d / yes are accepted and ignored; exit code = number of failed REQUIREs (capped at 255).
#pragma once
#include <chrono>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

namespace catch_stub {
struct Case {
	const char *name, *tags;
	void (*fn)();
};
inline std::vector<Case>& cases() {
	static std::vector<Case> c;
	return c;
}
inline int& failures() {
	static int f = 0;
	return f;
}
inline int& checks() {
	static int c = 0;
	return c;
}
struct Registrar {
	Registrar(const char* n, const char* t, void (*f)()) { cases().push_back({n, t, f}); }
};
inline void require(bool ok, const char* expr, const char* file, int line) {
	checks()++;
	if (!ok) {
		failures()++;
		std::printf("FAILED: REQUIRE( %s ) at %s:%d\n", expr, file, line);
	}
}
// Catch2 tag expression subset: "[a][b]" = both tags, "~[c]" excludes
inline bool selected(const char* tags, const char* filter) {
	if (!filter || !*filter) return true;
	std::string f(filter);
	size_t pos = 0;
	while ((pos = f.find('[', pos)) != std::string::npos) {
		const size_t end = f.find(']', pos);
		if (end == std::string::npos) break;
		const bool neg = pos > 0 && f[pos - 1] == '~';
		const bool has = std::strstr(tags, f.substr(pos, end - pos + 1).c_str()) != nullptr;
		if (has == neg) return false;
		pos = end + 1;
	}
	return true;
}
inline int run(int argc, char** argv) {
	const char* filter = nullptr;
	for (int i = 1; i < argc; i++)
		if (argv[i][0] == '[' || argv[i][0] == '~') filter = argv[i];
	int ran = 0;
	for (const Case& c : cases()) {
		if (!selected(c.tags, filter)) continue;
		const auto t0 = std::chrono::steady_clock::now();
		const int before = failures();
		c.fn();
		const double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
		std::printf("%.3f s: %s %s%s\n", s, c.name, c.tags, failures() == before ? "" : "  <-- FAILED");
		ran++;
	}
	std::printf("test cases: %d | assertions: %d | failed: %d\n", ran, checks(), failures());
	return failures() > 255 ? 255 : failures();
}
}  // namespace catch_stub

#define CATCH_STUB_CAT2(a, b) a##b
#define CATCH_STUB_CAT(a, b) CATCH_STUB_CAT2(a, b)
#define CATCH_STUB_TEST(fn, ...)                                                    \
	static void fn();                                                               \
	static catch_stub::Registrar CATCH_STUB_CAT(fn, _reg)(CATCH_STUB_FIRST(__VA_ARGS__, ""), CATCH_STUB_SECOND(__VA_ARGS__, "", ""), fn); \
	static void fn()
#define CATCH_STUB_FIRST(a, ...) a
#define CATCH_STUB_SECOND(a, b, ...) b
#define TEST_CASE(...) CATCH_STUB_TEST(CATCH_STUB_CAT(catch_stub_case_, __LINE__), __VA_ARGS__)
#define REQUIRE(...) catch_stub::require(static_cast<bool>(__VA_ARGS__), #__VA_ARGS__, __FILE__, __LINE__)

#if defined(CATCH_STUB_MAIN)
int main(int argc, char** argv) { return catch_stub::run(argc, argv); }
#endif
