// AllDifferent — SOFT_ALLDIFF error: sum over values of C(count, 2)
// (reference src/global_constraints/all_different.cpp:59-75).  Host side: checker + descriptor.
#pragma once
#include <map>
#include <vector>

#include "../constraint.hpp"

namespace ghost
{
	namespace global_constraints
	{
		class AllDifferent : public Constraint
		{
			double required_error(const std::vector<Variable*>& variables) const override
			{
				std::map<int, long long> count;
				for (const Variable* v : variables) ++count[v->get_value()];
				double pairs = 0;
				for (const auto& kv : count) pairs += static_cast<double>(kv.second * (kv.second - 1) / 2);
				return pairs;
			}
		public:
			AllDifferent(const std::vector<int>& variables_index) : Constraint(variables_index) {}
			AllDifferent(const std::vector<Variable>& variables) : Constraint(variables) {}
			int b200_kind() const override { return GHOST_KIND_ALL_DIFFERENT; }
		};
	}
}
