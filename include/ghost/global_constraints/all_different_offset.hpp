// AllDifferentOffset — extension kind (no counterpart in stock GHOST, whose AllDifferent has no offsets:
// reference include/global_constraints/all_different.hpp:46-74): the values x_i + o_i of the scope must be pairwise
// different.  Error = AllDifferent's own measure, sum over shifted values of C(count, 2)
// (reference src/global_constraints/all_different.cpp:59-75).  The diagonals of n-queens are
// AllDifferentOffset(q, {0,1,..,n-1}) and AllDifferentOffset(q, {0,-1,..,-(n-1)}): three constraints and n
// variables instead of 2n auxiliary variables tied by 2n LinearEquationEq.
#pragma once
#include <map>
#include <vector>

#include "../constraint.hpp"

namespace ghost
{
	namespace global_constraints
	{
		class AllDifferentOffset : public Constraint
		{
			std::vector<double> _offsets;
			double required_error(const std::vector<Variable*>& variables) const override
			{
				std::map<long long, long long> count;
				for (std::size_t i = 0; i < variables.size(); ++i) ++count[variables[i]->get_value() + static_cast<long long>(_offsets[i])];
				double pairs = 0;
				for (const auto& kv : count) pairs += static_cast<double>(kv.second * (kv.second - 1) / 2);
				return pairs;
			}
		public:
			AllDifferentOffset(const std::vector<int>& variables_index, const std::vector<int>& offsets)
				: Constraint(variables_index), _offsets(offsets.begin(), offsets.end()) {}
			AllDifferentOffset(const std::vector<Variable>& variables, const std::vector<int>& offsets)
				: Constraint(variables), _offsets(offsets.begin(), offsets.end()) {}
			int b200_kind() const override { return GHOST_KIND_ALL_DIFFERENT_OFFSET; }
			void b200_describe(std::vector<double>& coefficients, double& parameter) const override { coefficients = _offsets; parameter = 0.0; }
		};
	}
}
