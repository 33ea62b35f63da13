// ghost/model_builder.hpp — users derive from ModelBuilder and declare variables / constraints /
// objective / auxiliary data (reference include/model_builder.hpp:62-179, src/model_builder.cpp:39-115).
#pragma once
#include <memory>
#include <string>
#include <vector>

#include "model.hpp"

namespace ghost
{
	class ModelBuilder
	{
		template <typename B> friend class Solver;

		//! Runs the declare_* hooks, numbers everything and wires the raw Variable pointers.
		Model build_model()
		{
			variables.clear();
			constraints.clear();
			declare_variables();
			for (std::size_t i = 0; i < variables.size(); ++i) {
				variables[i]._id = static_cast<int>(i);
				if (variables[i]._name.empty()) variables[i]._name = "v" + std::to_string(i);
			}
			declare_auxiliary_data();
			declare_constraints();
			declare_objective();
			if (constraints.empty()) constraints.emplace_back(std::make_shared<PureOptimization>(variables));

			Model model(std::move(variables), constraints, objective, auxiliary_data, permutation_problem);
			auto wire = [&model](std::vector<Variable*>& ptrs, const std::vector<int>& ids, std::map<int, int>& position) {
				ptrs.clear();
				position.clear();
				for (std::size_t k = 0; k < ids.size(); ++k) {
					ptrs.push_back(&model.variables.at(static_cast<std::size_t>(ids[k])));
					position[ids[k]] = static_cast<int>(k);
				}
			};
			for (std::size_t c = 0; c < model.constraints.size(); ++c) {
				model.constraints[c]->_id = static_cast<int>(c);
				wire(model.constraints[c]->_variables, model.constraints[c]->_variables_index, model.constraints[c]->_variables_position);
			}
			wire(model.auxiliary_data->_variables, model.auxiliary_data->_variables_index, model.auxiliary_data->_variables_position);
			wire(model.objective->_variables, model.objective->_variables_index, model.objective->_variables_position);
			variables.clear();
			return model;
		}

	protected:
		std::vector<Variable> variables;
		std::vector<std::shared_ptr<Constraint>> constraints;
		std::shared_ptr<Objective> objective;
		std::shared_ptr<AuxiliaryData> auxiliary_data;
		bool permutation_problem;

	public:
		ModelBuilder(bool permutation_problem = false) : permutation_problem(permutation_problem) {}
		virtual ~ModelBuilder() = default;

		virtual void declare_variables() = 0;
		virtual void declare_constraints() {}
		virtual void declare_objective() { objective = std::make_shared<NullObjective>(); }
		virtual void declare_auxiliary_data() { auxiliary_data = std::make_shared<NullAuxiliaryData>(); }

		void create_n_variables(int number, const std::vector<int>& domain, int index = 0)
		{
			for (int i = 0; i < number; ++i) variables.emplace_back(domain, index);
		}
		void create_n_variables(int number, int starting_value, std::size_t size, int index = 0)
		{
			for (int i = 0; i < number; ++i) variables.emplace_back(starting_value, size, index);
		}
		int get_number_variables() { return static_cast<int>(variables.size()); }
		int get_number_constraints() { return static_cast<int>(constraints.size()); }
	};
}
