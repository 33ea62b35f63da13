// LinearEquationL — sum < rhs; error max(0, sum - rhs) + [sum == rhs]
// (reference src/global_constraints/linear_equation_l.cpp:54-61).
#pragma once
#include "linear_equation.hpp"

namespace ghost
{
	namespace global_constraints
	{
		GHOST_B200_LINEAR_KIND(LinearEquationL, GHOST_KIND_LINEAR_L, std::max(0.0, sum - rhs) + (rhs == sum ? 1.0 : 0.0))
	}
}
