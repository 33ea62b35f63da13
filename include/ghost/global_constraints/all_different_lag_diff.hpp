// AllDifferentLagDiff — extension kind (no counterpart in stock GHOST): the signed differences x[i+lag] - x[i] over
// the scope, in scope order, must be pairwise different — row `lag` of the difference triangle of a Costas array.
// Error = AllDifferent's measure, sum over difference values of C(count, 2)
// (reference src/global_constraints/all_different.cpp:59-75).  A Costas array of order n is a permutation problem
// with AllDifferent(s) and AllDifferentLagDiff(s, d) for d = 1..n-2: n variables and n - 1 constraints, where stock
// constraints need n(n-1)/2 auxiliary variables tied by as many LinearEquationEq (ghost_b200/problems.py).
#pragma once
#include <map>
#include <vector>

#include "../constraint.hpp"

namespace ghost
{
	namespace global_constraints
	{
		class AllDifferentLagDiff : public Constraint
		{
			int _lag;
			double required_error(const std::vector<Variable*>& variables) const override
			{
				std::map<int, long long> count;
				for (std::size_t i = 0; i + _lag < variables.size(); ++i) ++count[variables[i + _lag]->get_value() - variables[i]->get_value()];
				double pairs = 0;
				for (const auto& kv : count) pairs += static_cast<double>(kv.second * (kv.second - 1) / 2);
				return pairs;
			}
		public:
			AllDifferentLagDiff(const std::vector<int>& variables_index, int lag) : Constraint(variables_index), _lag(lag) {}
			AllDifferentLagDiff(const std::vector<Variable>& variables, int lag) : Constraint(variables), _lag(lag) {}
			int b200_kind() const override { return GHOST_KIND_ALL_DIFFERENT_LAG_DIFF; }
			void b200_describe(std::vector<double>& coefficients, double& parameter) const override { coefficients.clear(); parameter = _lag; }
		};
	}
}
