// LinearEquationLeq — sum <= rhs; error max(0, sum - rhs)
// (reference src/global_constraints/linear_equation_leq.cpp:54-61).
#pragma once
#include "linear_equation.hpp"

namespace ghost
{
	namespace global_constraints
	{
		GHOST_B200_LINEAR_KIND(LinearEquationLeq, GHOST_KIND_LINEAR_LEQ, std::max(0.0, sum - rhs))
	}
}
