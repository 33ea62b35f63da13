// AllEqual — error: scope size minus the count of the most frequent value
// (reference src/global_constraints/all_equal.cpp:47-60).  Host side: checker + descriptor.
#pragma once
#include <algorithm>
#include <map>
#include <vector>

#include "../constraint.hpp"

namespace ghost
{
	namespace global_constraints
	{
		class AllEqual : public Constraint
		{
			double required_error(const std::vector<Variable*>& variables) const override
			{
				std::map<int, long long> count;
				long long lead = 0;
				for (const Variable* v : variables) lead = std::max(lead, ++count[v->get_value()]);
				return static_cast<double>(static_cast<long long>(variables.size()) - lead);
			}
		public:
			AllEqual(const std::vector<int>& variables_index) : Constraint(variables_index) {}
			AllEqual(const std::vector<Variable>& variables) : Constraint(variables) {}
			int b200_kind() const override { return GHOST_KIND_ALL_EQUAL; }
		};
	}
}
