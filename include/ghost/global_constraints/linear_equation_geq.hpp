// LinearEquationGeq — sum >= rhs; error max(0, rhs - sum)
// (reference src/global_constraints/linear_equation_geq.cpp:54-61).
#pragma once
#include "linear_equation.hpp"

namespace ghost
{
	namespace global_constraints
	{
		GHOST_B200_LINEAR_KIND(LinearEquationGeq, GHOST_KIND_LINEAR_GEQ, std::max(0.0, rhs - sum))
	}
}
