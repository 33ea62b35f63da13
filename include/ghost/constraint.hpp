// ghost/constraint.hpp — abstract constraint, API of the reference's ghost::Constraint
// (include/constraint.hpp:62-312, src/constraint.cpp:57-125).
//
// On the B200 build the search itself runs on the GPU over a flat descriptor, so a constraint has two
// jobs on the host: describe itself to the lowering (b200_kind / b200_describe — the stock global
// constraints override them) and act as the checker that re-scores returned assignments (error()).
// A user-defined subclass keeps compiling, but Solver::fast_search refuses it: user-defined virtuals
// are outside the GPU path and there is no CPU fallback.
#pragma once
#include <cmath>
#include <map>
#include <memory>
#include <ostream>
#include <stdexcept>
#include <string>
#include <typeinfo>
#include <vector>

#include "../ghost_b200.h"
#include "variable.hpp"

namespace ghost
{
	class Constraint
	{
		friend class ModelBuilder;
		template <typename B> friend class Solver;

		std::vector<Variable*> _variables;
		std::vector<int> _variables_index;       // global ids of the scope, in scope order
		std::map<int, int> _variables_position;  // global id -> position in the scope
		int _id = 0;

	protected:
		//! Error of the current assignment of the scope; 0 iff satisfied, never negative or NaN.
		virtual double required_error(const std::vector<Variable*>& variables) const = 0;

		//! Error change if the scope positions `indexes` took `candidate_values` (optional override).
		virtual double optional_delta_error(const std::vector<Variable*>&, const std::vector<int>&, const std::vector<int>&) const
		{
			throw std::logic_error("Constraint::optional_delta_error() has not been user-defined.\n");
		}

		//! Hook called after a variable of the scope changed (optional override).
		virtual void conditional_update_data_structures(const std::vector<Variable*>&, int, int) {}

		int get_id() const { return _id; }

	public:
		struct nanException : std::runtime_error { using std::runtime_error::runtime_error; };
		struct negativeException : std::runtime_error { using std::runtime_error::runtime_error; };

		Constraint(const std::vector<int>& variables_index) : _variables_index(variables_index) {}
		Constraint(const std::vector<Variable>& variables)
		{
			for (const auto& v : variables) _variables_index.push_back(v.get_id());
		}
		Constraint(const Constraint&) = default;
		Constraint(Constraint&&) = default;
		Constraint& operator=(const Constraint&) = delete;
		Constraint& operator=(Constraint&&) = delete;
		virtual ~Constraint() = default;

		bool has_variable(int var_id) const { return _variables_position.count(var_id) > 0; }

		//! Checked error of the current assignment (src/constraint.cpp:57-65).
		double error() const
		{
			double value = required_error(_variables);
			if (std::isnan(value)) throw nanException("Constraint required_error returned a NaN value");
			if (value < 0) throw negativeException("Constraint required_error returned a negative value");
			return value;
		}

		std::vector<int> get_variable_ids() const { return _variables_index; }

		// ---- B200 lowering hooks (additions) ----
		//! GhostConstraintKind of a stock constraint; GHOST_KIND_USER for anything user-defined.
		virtual int b200_kind() const { return GHOST_KIND_USER; }
		//! Coefficients aligned with the scope (linear kinds) and the scalar parameter (rhs / fixed value).
		virtual void b200_describe(std::vector<double>& coefficients, double& parameter) const { coefficients.clear(); parameter = 0.0; }

		friend std::ostream& operator<<(std::ostream& os, const Constraint& c)
		{
			return os << "Constraint type: " << typeid(c).name() << "\nId: " << c._id << "\n########";
		}
	};

	//! Inserted by ModelBuilder when the user declares no constraint (pure optimisation runs).
	class PureOptimization : public Constraint
	{
		double required_error(const std::vector<Variable*>&) const override { return 0.; }
		double optional_delta_error(const std::vector<Variable*>&, const std::vector<int>&, const std::vector<int>&) const override { return 0.; }
	public:
		PureOptimization(const std::vector<Variable>& variables) : Constraint(variables) {}
		int b200_kind() const override { return GHOST_KIND_PURE_OPTIMIZATION; }
	};
}
