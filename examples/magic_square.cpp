// Magic square n x n with stock global constraints only — BASELINE.json configs[1] at n = 30.
// This source compiles unchanged against the reference's headers (CPU solver) and against this
// repository's include/ghost (B200 solver):
//   g++ -std=c++17 -Iinclude examples/magic_square.cpp -Lghost_b200 -lghost_b200 -Wl,-rpath,$PWD/ghost_b200
#include <cstdlib>
#include <iostream>
#include <vector>

#include <ghost/solver.hpp>
#include <ghost/global_constraints/all_different.hpp>
#include <ghost/global_constraints/linear_equation_eq.hpp>

class MagicSquareBuilder : public ghost::ModelBuilder
{
	int n;
public:
	MagicSquareBuilder(int n) : ghost::ModelBuilder(true), n(n) {}   // permutation problem

	void declare_variables() override
	{
		for (int k = 0; k < n * n; ++k) variables.emplace_back(1, static_cast<std::size_t>(n * n), k); // x_k = k+1
	}

	void declare_constraints() override
	{
		using namespace ghost::global_constraints;
		const double magic = n * (n * n + 1) / 2;
		std::vector<int> all;
		for (int k = 0; k < n * n; ++k) all.push_back(k);
		constraints.emplace_back(std::make_shared<AllDifferent>(all));
		for (int r = 0; r < n; ++r) {
			std::vector<int> row;
			for (int c = 0; c < n; ++c) row.push_back(r * n + c);
			constraints.emplace_back(std::make_shared<LinearEquationEq>(row, magic));
		}
		for (int c = 0; c < n; ++c) {
			std::vector<int> col;
			for (int r = 0; r < n; ++r) col.push_back(r * n + c);
			constraints.emplace_back(std::make_shared<LinearEquationEq>(col, magic));
		}
		std::vector<int> d1, d2;
		for (int i = 0; i < n; ++i) { d1.push_back(i * n + i); d2.push_back(i * n + n - 1 - i); }
		constraints.emplace_back(std::make_shared<LinearEquationEq>(d1, magic));
		constraints.emplace_back(std::make_shared<LinearEquationEq>(d2, magic));
	}
};

int main(int argc, char** argv)
{
	const int n = argc > 1 ? std::atoi(argv[1]) : 6;
	const double budget_us = argc > 2 ? std::atof(argv[2]) * 1e6 : 10e6;
	MagicSquareBuilder builder(n);
	ghost::Solver<MagicSquareBuilder> solver(builder);
	double cost;
	std::vector<int> solution;
	bool found = solver.fast_search(cost, solution, budget_us);
	std::cout << (found ? "Solution found" : "Solution not found") << ", cost " << cost << "\n";
	if (found && n <= 12)
		for (int r = 0; r < n; ++r) {
			for (int c = 0; c < n; ++c) std::cout << solution[r * n + c] << "\t";
			std::cout << "\n";
		}
	return found ? 0 : 1;
}
