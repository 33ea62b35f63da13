// Every magic square of order n (n = 3: 8 squares, n = 4: 7040) with Solver::complete_search.
// The same source compiles against the reference's headers, where the enumeration is the AC-3 + depth-first
// search of include/solver.hpp:932-987 on the CPU (minutes for n = 3), and against this repository's
// include/ghost, where it runs on the GPU through ghost_b200_complete_search:
//   g++ -std=c++17 -Iinclude examples/all_magic_squares.cpp -Lghost_b200 -lghost_b200 -Wl,-rpath,$PWD/ghost_b200
#include <cstdlib>
#include <iostream>
#include <vector>

#include <ghost/solver.hpp>
#include <ghost/global_constraints/all_different.hpp>
#include <ghost/global_constraints/linear_equation_eq.hpp>

class SquareBuilder : public ghost::ModelBuilder
{
	int n;
public:
	SquareBuilder(int n) : ghost::ModelBuilder(), n(n) {}

	void declare_variables() override
	{
		for (int k = 0; k < n * n; ++k) variables.emplace_back(1, static_cast<std::size_t>(n * n), 0);
	}

	void declare_constraints() override
	{
		using namespace ghost::global_constraints;
		const double magic = n * (n * n + 1) / 2;
		std::vector<int> all, d1, d2;
		for (int k = 0; k < n * n; ++k) all.push_back(k);
		constraints.emplace_back(std::make_shared<AllDifferent>(all));
		for (int r = 0; r < n; ++r) {
			std::vector<int> row, col;
			for (int c = 0; c < n; ++c) { row.push_back(r * n + c); col.push_back(c * n + r); }
			constraints.emplace_back(std::make_shared<LinearEquationEq>(row, magic));
			constraints.emplace_back(std::make_shared<LinearEquationEq>(col, magic));
			d1.push_back(r * n + r);
			d2.push_back(r * n + n - 1 - r);
		}
		constraints.emplace_back(std::make_shared<LinearEquationEq>(d1, magic));
		constraints.emplace_back(std::make_shared<LinearEquationEq>(d2, magic));
	}
};

int main(int argc, char** argv)
{
	const int n = argc > 1 ? std::atoi(argv[1]) : 3;
	SquareBuilder builder(n);
	ghost::Solver<SquareBuilder> solver(builder);
	std::vector<double> costs;
	std::vector<std::vector<int>> solutions;
	const bool any = solver.complete_search(costs, solutions);
	std::cout << (any ? "Solutions: " : "No solution: ") << solutions.size() << "\n";
	for (std::size_t s = 0; s < solutions.size() && s < 2; ++s) {
		for (int r = 0; r < n; ++r) {
			for (int c = 0; c < n; ++c) std::cout << solutions[s][static_cast<std::size_t>(r * n + c)] << "\t";
			std::cout << "\n";
		}
		std::cout << "\n";
	}
	return any ? 0 : 1;
}
