// Stand-in for Catch2's BENCHMARK("name") { body };  -- runs the body a few times and prints the mean (no statistics).
#pragma once
#include <chrono>
#include <cstdio>
#include <type_traits>

namespace catch_stub {
struct Bench {
	const char* name;
	template <typename F>
	void operator=(F&& body) const {
		const int reps = 5;
		const auto t0 = std::chrono::steady_clock::now();
		for (int i = 0; i < reps; i++) {
			if constexpr (std::is_void_v<decltype(body())>) {
				body();
			} else {
				volatile auto keep = body();
				(void)keep;
			}
		}
		const double us = std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t0).count() / reps;
		std::printf("benchmark %-50s mean %.1f us\n", name, us);
	}
};
}  // namespace catch_stub
#define BENCHMARK(name) catch_stub::Bench{name} = [&]()
