// N-Queens as BASELINE.json words config 0, "3 x AllDifferent via ModelBuilder": q_i is the column of the queen in
// row i; AllDifferent(q), AllDifferent(q_i + i), AllDifferent(q_i - i).  The two diagonal constraints use the
// extension kind AllDifferentOffset (stock GHOST's AllDifferent has no offsets, so a stock model needs 2n auxiliary
// variables and 2n linking equations: ghost_b200/problems.py n_queens()).
//   g++ -std=c++17 -Iinclude examples/n_queens.cpp -Lghost_b200 -lghost_b200 -Wl,-rpath,$PWD/ghost_b200
#include <cstdlib>
#include <iostream>
#include <vector>

#include <ghost/solver.hpp>
#include <ghost/global_constraints/all_different.hpp>
#include <ghost/global_constraints/all_different_offset.hpp>

class QueensBuilder : public ghost::ModelBuilder
{
	int n;
public:
	QueensBuilder(int n) : ghost::ModelBuilder(), n(n) {}

	void declare_variables() override
	{
		for (int i = 0; i < n; ++i) variables.emplace_back(0, static_cast<std::size_t>(n), i);
	}

	void declare_constraints() override
	{
		using namespace ghost::global_constraints;
		std::vector<int> q, up, down;
		for (int i = 0; i < n; ++i) { q.push_back(i); up.push_back(i); down.push_back(-i); }
		constraints.emplace_back(std::make_shared<AllDifferent>(q));
		constraints.emplace_back(std::make_shared<AllDifferentOffset>(q, up));
		constraints.emplace_back(std::make_shared<AllDifferentOffset>(q, down));
	}
};

int main(int argc, char** argv)
{
	const int n = argc > 1 ? std::atoi(argv[1]) : 100;
	const double budget_us = argc > 2 ? std::atof(argv[2]) * 1e6 : 10e6;
	QueensBuilder builder(n);
	ghost::Solver<QueensBuilder> solver(builder);
	double cost;
	std::vector<int> solution;
	bool found = solver.fast_search(cost, solution, budget_us);
	// an independent check: no two queens share a column or a diagonal
	bool ok = found;
	for (int i = 0; ok && i < n; ++i)
		for (int j = i + 1; ok && j < n; ++j)
			if (solution[i] == solution[j] || solution[i] + i == solution[j] + j || solution[i] - i == solution[j] - j) ok = false;
	std::cout << (found ? "Solution found" : "Solution not found") << ", cost " << cost << (ok ? ", verified" : "") << "\n";
	return ok ? 0 : 1;
}
