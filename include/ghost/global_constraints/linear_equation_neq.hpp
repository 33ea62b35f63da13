// LinearEquationNeq — sum != rhs; error 1 when equal
// (reference src/global_constraints/linear_equation_neq.cpp:54-61).
#pragma once
#include "linear_equation.hpp"

namespace ghost
{
	namespace global_constraints
	{
		GHOST_B200_LINEAR_KIND(LinearEquationNeq, GHOST_KIND_LINEAR_NEQ, (rhs == sum ? 1.0 : 0.0))
	}
}
