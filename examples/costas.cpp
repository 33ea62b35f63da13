// Costas array of order n as a difference triangle: s is a permutation of 1..n and, for every lag d = 1..n-2, the
// differences s[i+d] - s[i] are pairwise different.  One AllDifferent plus n-2 constraints of the extension kind
// AllDifferentLagDiff on n variables (stock GHOST needs n(n-1)/2 auxiliary variables and as many linking equations:
// ghost_b200/problems.py costas(); that form is not solved even for n = 12).  BASELINE.json config 5 is n = 20.
// Optional third argument: the heuristic set (ghost::Options::strategy, 0 = the reference's default Adaptive Search,
// 1 = Antidote, 2 = Culprit projection, 3 = hill climbing, 4 = random walk).
//   g++ -std=c++17 -Iinclude examples/costas.cpp -Lghost_b200 -lghost_b200 -Wl,-rpath,$PWD/ghost_b200
#include <cstdlib>
#include <iostream>
#include <set>
#include <vector>

#include <ghost/solver.hpp>
#include <ghost/global_constraints/all_different.hpp>
#include <ghost/global_constraints/all_different_lag_diff.hpp>

class CostasBuilder : public ghost::ModelBuilder
{
	int n;
public:
	CostasBuilder(int n) : ghost::ModelBuilder(true), n(n) {}   // permutation problem: moves are swaps

	void declare_variables() override
	{
		for (int i = 0; i < n; ++i) variables.emplace_back(1, static_cast<std::size_t>(n), i);   // values 1..n, start value i + 1
	}

	void declare_constraints() override
	{
		using namespace ghost::global_constraints;
		std::vector<int> s;
		for (int i = 0; i < n; ++i) s.push_back(i);
		constraints.emplace_back(std::make_shared<AllDifferent>(s));
		for (int d = 1; d <= n - 2; ++d) constraints.emplace_back(std::make_shared<AllDifferentLagDiff>(s, d));
	}
};

int main(int argc, char** argv)
{
	const int n = argc > 1 ? std::atoi(argv[1]) : 14;
	const double budget_us = argc > 2 ? std::atof(argv[2]) * 1e6 : 10e6;
	CostasBuilder builder(n);
	ghost::Solver<CostasBuilder> solver(builder);
	ghost::Options options;
	if (argc > 3) options.strategy = std::atoi(argv[3]);
	double cost;
	std::vector<int> solution;
	bool found = solver.fast_search(cost, solution, budget_us, options);
	// an independent check of the Costas property
	bool ok = found;
	for (int d = 1; ok && d < n; ++d) {
		std::set<int> seen;
		for (int i = 0; ok && i + d < n; ++i) ok = seen.insert(solution[i + d] - solution[i]).second;
	}
	std::cout << (found ? "Solution found" : "Solution not found") << ", cost " << cost << (ok ? ", verified" : "") << "\n";
	if (found) { for (int v : solution) std::cout << v << " "; std::cout << "\n"; }
	return ok ? 0 : 1;
}
