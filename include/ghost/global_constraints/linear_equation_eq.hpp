// LinearEquationEq — sum = rhs; error |sum - rhs|
// (reference src/global_constraints/linear_equation_eq.cpp:54-61).
#pragma once
#include "linear_equation.hpp"

namespace ghost
{
	namespace global_constraints
	{
		GHOST_B200_LINEAR_KIND(LinearEquationEq, GHOST_KIND_LINEAR_EQ, std::abs(sum - rhs))
	}
}
