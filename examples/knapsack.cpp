// Multi-dimensional knapsack as a COP — BASELINE.json configs[3]: binary variables,
// LinearEquationLeq capacity rows, LinearObjective (this build's addition) to maximise.
#include <cstdlib>
#include <iostream>
#include <random>
#include <vector>

#include <ghost/solver.hpp>

class KnapsackBuilder : public ghost::ModelBuilder
{
	int n, m;
	std::vector<std::vector<double>> w;
	std::vector<double> p, cap;
public:
	KnapsackBuilder(int n, int m, unsigned seed) : n(n), m(m), w(m, std::vector<double>(n)), p(n), cap(m)
	{
		std::mt19937 rng(seed);
		for (int j = 0; j < m; ++j) {
			double total = 0;
			for (int i = 0; i < n; ++i) { w[j][i] = 1 + rng() % 1000; total += w[j][i]; }
			cap[j] = std::floor(total / 2);
		}
		for (int i = 0; i < n; ++i) {
			double s = 0;
			for (int j = 0; j < m; ++j) s += w[j][i];
			p[i] = std::floor(s / m + rng() % 500);
		}
	}
	void declare_variables() override { create_n_variables(n, 0, 2); }
	void declare_constraints() override
	{
		for (int j = 0; j < m; ++j)
			constraints.emplace_back(std::make_shared<ghost::global_constraints::LinearEquationLeq>(variables, cap[j], w[j]));
	}
	void declare_objective() override
	{
		objective = std::make_shared<ghost::global_objectives::LinearObjective>(variables, p, true);
	}
};

int main(int argc, char** argv)
{
	const int n = argc > 1 ? std::atoi(argv[1]) : 100, m = argc > 2 ? std::atoi(argv[2]) : 5;
	KnapsackBuilder builder(n, m, 12345);
	ghost::Solver<KnapsackBuilder> solver(builder);
	double cost;
	std::vector<int> solution;
	bool found = solver.fast_search(cost, solution, argc > 3 ? std::atof(argv[3]) * 1e6 : 2e6);
	std::cout << (found ? "Feasible" : "Infeasible") << ", value " << cost << ", walkers " << solver.get_statistics().walkers << "\n";
	return found ? 0 : 1;
}
