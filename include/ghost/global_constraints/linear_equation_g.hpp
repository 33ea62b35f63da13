// LinearEquationG — sum > rhs; error max(0, rhs - sum) + [sum == rhs]
// (reference src/global_constraints/linear_equation_g.cpp:54-61).
#pragma once
#include "linear_equation.hpp"

namespace ghost
{
	namespace global_constraints
	{
		GHOST_B200_LINEAR_KIND(LinearEquationG, GHOST_KIND_LINEAR_G, std::max(0.0, rhs - sum) + (rhs == sum ? 1.0 : 0.0))
	}
}
