// AllDifferentAbsDiff — extension kind (no counterpart in stock GHOST, which has no constraint with an
// absolute value): the |x[i+1] - x[i]| of the scope, in scope order, must be pairwise different;
// error = sum over difference values of C(count, 2).  Needed by the all-interval-series benchmark.
#pragma once
#include <cstdlib>
#include <map>
#include <vector>

#include "../constraint.hpp"

namespace ghost
{
	namespace global_constraints
	{
		class AllDifferentAbsDiff : public Constraint
		{
			double required_error(const std::vector<Variable*>& variables) const override
			{
				std::map<int, long long> count;
				for (std::size_t i = 0; i + 1 < variables.size(); ++i)
					++count[std::abs(variables[i + 1]->get_value() - variables[i]->get_value())];
				double pairs = 0;
				for (const auto& kv : count) pairs += static_cast<double>(kv.second * (kv.second - 1) / 2);
				return pairs;
			}
		public:
			AllDifferentAbsDiff(const std::vector<int>& variables_index) : Constraint(variables_index) {}
			AllDifferentAbsDiff(const std::vector<Variable>& variables) : Constraint(variables) {}
			int b200_kind() const override { return GHOST_KIND_ALL_DIFFERENT_ABS_DIFF; }
		};
	}
}
