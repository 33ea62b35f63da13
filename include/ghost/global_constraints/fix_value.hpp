// FixValue — every variable of the scope should equal `value`; error = sum |x - value|
// (reference src/global_constraints/fix_value.cpp:46-52).
#pragma once
#include <cstdlib>
#include <vector>

#include "../constraint.hpp"

namespace ghost
{
	namespace global_constraints
	{
		class FixValue : public Constraint
		{
			int _value;
			double required_error(const std::vector<Variable*>& variables) const override
			{
				double total = 0.;
				for (const Variable* v : variables) total += std::abs(v->get_value() - _value);
				return total;
			}
		public:
			FixValue(const std::vector<int>& variables_index, int value) : Constraint(variables_index), _value(value) {}
			FixValue(const std::vector<Variable>& variables, int value) : Constraint(variables), _value(value) {}
			int b200_kind() const override { return GHOST_KIND_FIX_VALUE; }
			void b200_describe(std::vector<double>& coefficients, double& parameter) const override { coefficients.clear(); parameter = _value; }
		};
	}
}
