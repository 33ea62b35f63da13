// Minimal stand-in for the slice of M.E. O'Neill's randutils that GHOST's public signatures name
// (reference thirdparty/randutils.hpp:605-807: mt19937_rng with uniform / pick / shuffle / variate).
// The GPU path draws from per-walker Philox streams; this class only keeps user code that mentions
// randutils::mt19937_rng (e.g. an overridden Objective::expert_heuristic_value) compiling.
#pragma once
#include <algorithm>
#include <iterator>
#include <random>

namespace randutils {
class mt19937_rng {
	std::mt19937 engine_;
public:
	mt19937_rng() : engine_(std::random_device{}()) {}
	template <typename Seed> explicit mt19937_rng(Seed s) : engine_(static_cast<std::mt19937::result_type>(s)) {}
	template <typename Seed> void seed(Seed s) { engine_.seed(static_cast<std::mt19937::result_type>(s)); }
	std::mt19937& engine() { return engine_; }
	template <typename N> N uniform(N lo, N hi) { return std::uniform_int_distribution<N>(lo, hi)(engine_); }
	template <typename R> auto pick(R&& range) -> decltype(*std::begin(range))
	{
		auto first = std::begin(range);
		auto n = std::distance(first, std::end(range));
		if (n > 1) std::advance(first, uniform<decltype(n)>(0, n - 1));
		return *first;
	}
	template <typename R> void shuffle(R&& range) { std::shuffle(std::begin(range), std::end(range), engine_); }
};
}
