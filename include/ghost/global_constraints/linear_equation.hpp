// LinearEquation — common part of the six linear constraints: S = sum a_i x_i compared with rhs
// (reference src/global_constraints/linear_equation.cpp:50-57 and the compute_error of
// linear_equation_{eq,neq,leq,geq,l,g}.cpp:54-61).
#pragma once
#include <algorithm>
#include <cmath>
#include <vector>

#include "../constraint.hpp"

namespace ghost
{
	namespace global_constraints
	{
		class LinearEquation : public Constraint
		{
			std::vector<double> _coefficients;
		protected:
			double rhs;
			LinearEquation(const std::vector<int>& variables_index, double rhs, const std::vector<double>& coefficients)
				: Constraint(variables_index), _coefficients(coefficients), rhs(rhs) {}
			LinearEquation(const std::vector<Variable>& variables, double rhs, const std::vector<double>& coefficients)
				: Constraint(variables), _coefficients(coefficients), rhs(rhs) {}
			virtual double compute_error(double sum) const = 0;
			double required_error(const std::vector<Variable*>& variables) const override
			{
				double sum = 0.0;
				for (std::size_t i = 0; i < variables.size(); ++i) sum += _coefficients[i] * variables[i]->get_value();
				return compute_error(sum);
			}
		public:
			void b200_describe(std::vector<double>& coefficients, double& parameter) const override { coefficients = _coefficients; parameter = rhs; }
		};

#define GHOST_B200_LINEAR_KIND(NAME, KIND, ERROR_EXPR)                                                                         \
		class NAME : public LinearEquation                                                                                     \
		{                                                                                                                      \
			double compute_error(double sum) const override { return ERROR_EXPR; }                                              \
		public:                                                                                                                \
			NAME(const std::vector<int>& variables_index, double rhs, const std::vector<double>& coefficients)                 \
				: LinearEquation(variables_index, rhs, coefficients) {}                                                        \
			NAME(const std::vector<int>& variables_index, double rhs)                                                          \
				: LinearEquation(variables_index, rhs, std::vector<double>(variables_index.size(), 1.0)) {}                    \
			NAME(const std::vector<Variable>& variables, double rhs, const std::vector<double>& coefficients)                  \
				: LinearEquation(variables, rhs, coefficients) {}                                                              \
			NAME(const std::vector<Variable>& variables, double rhs)                                                           \
				: LinearEquation(variables, rhs, std::vector<double>(variables.size(), 1.0)) {}                                \
			int b200_kind() const override { return KIND; }                                                                    \
		};
	}
}
