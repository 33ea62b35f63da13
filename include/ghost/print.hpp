// ghost/print.hpp — candidate pretty-printer hook (reference include/print.hpp:45-57).
#pragma once
#include <sstream>
#include <vector>

#include "variable.hpp"

namespace ghost
{
	class Print
	{
	public:
		virtual ~Print() = default;
		virtual std::stringstream print_candidate(const std::vector<Variable>& variables) const
		{
			std::stringstream out;
			out << "Variables:\n";
			for (const auto& v : variables) out << v.get_name() << "=" << v.get_value() << "\n";
			return out;
		}
	};
}
