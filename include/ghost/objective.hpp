// ghost/objective.hpp — objective functions, API of the reference's ghost::Objective / Minimize /
// Maximize / NullObjective (include/objective.hpp:61-457, src/objective.cpp:53-123).
// GHOST ships no concrete objective; the GPU path needs one it can describe, so this build adds
// ghost::global_objectives::LinearObjective (global_objectives/linear_objective.hpp).
#pragma once
#include <cmath>
#include <limits>
#include <map>
#include <ostream>
#include <stdexcept>
#include <string>
#include <vector>

#include "../ghost_b200.h"
#include "thirdparty/randutils.hpp"
#include "variable.hpp"

namespace ghost
{
	class Objective
	{
		friend class ModelBuilder;
		template <typename B> friend class Solver;
		friend class NullObjective;

		std::vector<Variable*> _variables;
		std::vector<int> _variables_index;
		std::map<int, int> _variables_position;
		bool _is_optimization = true;
		bool _is_maximization;
		std::string _name;

	protected:
		Objective(const std::vector<int>& variables_index, bool is_maximization, const std::string& name)
			: _variables_index(variables_index), _is_maximization(is_maximization), _name(name) {}
		Objective(const std::vector<Variable>& variables, bool is_maximization, const std::string& name)
			: _is_maximization(is_maximization), _name(name)
		{
			for (const auto& v : variables) _variables_index.push_back(v.get_id());
		}

		virtual double required_cost(const std::vector<Variable*>& variables) const = 0;
		virtual void conditional_update_data_structures(const std::vector<Variable*>&, int, int) {}

		//! Tie-break among values of equal constraint error: cheapest objective, random among equals.
		virtual int expert_heuristic_value(const std::vector<Variable*>& variables, int variable_index,
		                                   const std::vector<int>& possible_values, randutils::mt19937_rng& rng) const
		{
			Variable* var = variables[static_cast<std::size_t>(variable_index)];
			const int backup = var->get_value();
			double best = std::numeric_limits<double>::max();
			std::vector<int> keep;
			for (int v : possible_values) {
				var->set_value(v);
				double c = required_cost(variables);
				if (_is_maximization) c = -c;
				if (c < best) { best = c; keep.assign(1, v); }
				else if (c == best) keep.push_back(v);
			}
			var->set_value(backup);
			if (!keep.empty()) return rng.pick(keep);
			if (!possible_values.empty()) return rng.pick(possible_values);
			return backup;
		}
		virtual int expert_heuristic_value_permutation(const std::vector<Variable*>&, int, const std::vector<int>& bad_variables,
		                                               randutils::mt19937_rng& rng) const { return rng.pick(bad_variables); }
		virtual double expert_postprocess(const std::vector<Variable*>&, double best_cost) const { return best_cost; }

	public:
		struct nanException : std::runtime_error { using std::runtime_error::runtime_error; };

		Objective(const Objective&) = default;
		Objective(Objective&&) = default;
		Objective& operator=(const Objective&) = delete;
		Objective& operator=(Objective&&) = delete;
		virtual ~Objective() = default;

		//! Internal cost: negated for a maximisation (src/objective.cpp:53-64).
		double cost() const
		{
			double value = required_cost(_variables);
			if (std::isnan(value)) throw nanException("Objective required_cost returned a NaN value");
			return _is_maximization ? -value : value;
		}
		double postprocess(double best_cost) const { return expert_postprocess(_variables, best_cost); }

		std::string get_name() const { return _name; }
		bool is_optimization() const { return _is_optimization; }
		bool is_maximization() const { return _is_maximization; }
		std::vector<int> get_variable_ids() const { return _variables_index; }

		// ---- B200 lowering hooks (additions) ----
		virtual int b200_kind() const { return _is_optimization ? GHOST_OBJ_USER : GHOST_OBJ_NONE; }
		virtual void b200_describe(std::vector<double>& coefficients) const { coefficients.clear(); }

		friend std::ostream& operator<<(std::ostream& os, const Objective& o) { return os << "Objective name: " << o._name << "\n********"; }
	};

	class NullObjective : public Objective
	{
		double required_cost(const std::vector<Variable*>&) const override { return 0.0; }
	public:
		NullObjective() : Objective(std::vector<int>{0}, false, "nullObjective") { _is_optimization = false; }
	};

	class Minimize : public Objective
	{
	public:
		Minimize(const std::vector<int>& variables_index, const std::string& name = "Minimize") : Objective(variables_index, false, name) {}
		Minimize(const std::vector<Variable>& variables, const std::string& name = "Minimize") : Objective(variables, false, name) {}
		Minimize(const std::vector<int>& variables_index, const char* name) : Objective(variables_index, false, std::string(name)) {}
		Minimize(const std::vector<Variable>& variables, const char* name) : Objective(variables, false, std::string(name)) {}
	};

	class Maximize : public Objective
	{
	public:
		Maximize(const std::vector<int>& variables_index, const std::string& name = "Maximize") : Objective(variables_index, true, name) {}
		Maximize(const std::vector<Variable>& variables, const std::string& name = "Maximize") : Objective(variables, true, name) {}
		Maximize(const std::vector<int>& variables_index, const char* name) : Objective(variables_index, true, std::string(name)) {}
		Maximize(const std::vector<Variable>& variables, const char* name) : Objective(variables, true, std::string(name)) {}
	};
}
