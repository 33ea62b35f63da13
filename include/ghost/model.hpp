// ghost/model.hpp — what ModelBuilder::build_model returns (reference include/model.hpp:43-58).
#pragma once
#include <memory>
#include <vector>

#include "auxiliary_data.hpp"
#include "constraint.hpp"
#include "objective.hpp"
#include "variable.hpp"

namespace ghost
{
	struct Model final
	{
		std::vector<Variable> variables;
		std::vector<std::shared_ptr<Constraint>> constraints;
		std::shared_ptr<Objective> objective;
		std::shared_ptr<AuxiliaryData> auxiliary_data;
		bool permutation_problem = false;

		Model() = default;
		Model(std::vector<Variable>&& vars, const std::vector<std::shared_ptr<Constraint>>& ctrs,
		      const std::shared_ptr<Objective>& obj, const std::shared_ptr<AuxiliaryData>& aux, bool permutation)
			: variables(std::move(vars)), constraints(ctrs), objective(obj), auxiliary_data(aux), permutation_problem(permutation) {}
	};
}
