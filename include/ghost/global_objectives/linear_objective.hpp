// LinearObjective — sum c_i x_i to minimise or maximise.  An ADDITION: GHOST ships no concrete
// objective (users write their own, cf. reference tutorial/wiki/src/objective_max_value.cpp:9-17), and a
// user-defined virtual cannot run inside the GPU walkers; this is the objective the B200 path can lower.
#pragma once
#include <vector>

#include "../objective.hpp"

namespace ghost
{
	namespace global_objectives
	{
		class LinearObjective : public Objective
		{
			std::vector<double> _coefficients;
			double required_cost(const std::vector<Variable*>& variables) const override
			{
				double total = 0.0;
				for (std::size_t i = 0; i < variables.size(); ++i) total += _coefficients[i] * variables[i]->get_value();
				return total;
			}
		public:
			LinearObjective(const std::vector<int>& variables_index, const std::vector<double>& coefficients, bool maximize,
			                const std::string& name = "LinearObjective")
				: Objective(variables_index, maximize, name), _coefficients(coefficients) {}
			LinearObjective(const std::vector<Variable>& variables, const std::vector<double>& coefficients, bool maximize,
			                const std::string& name = "LinearObjective")
				: Objective(variables, maximize, name), _coefficients(coefficients) {}
			int b200_kind() const override { return is_maximization() ? GHOST_OBJ_LINEAR_MAXIMIZE : GHOST_OBJ_LINEAR_MINIMIZE; }
			void b200_describe(std::vector<double>& coefficients) const override { coefficients = _coefficients; }
		};
	}
}
