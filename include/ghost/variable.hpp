// ghost/variable.hpp — integer decision variable, API of the reference's ghost::Variable
// (include/variable.hpp:56-228, src/variable.cpp:37-136), written from scratch for the B200 build.
#pragma once
#include <algorithm>
#include <cstddef>
#include <numeric>
#include <ostream>
#include <stdexcept>
#include <string>
#include <vector>

#include "thirdparty/randutils.hpp"

namespace ghost
{
	class ModelBuilder;
	template <typename B> class Solver;

	class Variable final
	{
		friend class ModelBuilder;
		template <typename B> friend class Solver;

		std::vector<int> _domain;
		int _id = 0;
		std::string _name;
		int _current_value = 0;
		int _min_value = 0;
		int _max_value = 0;

		void finish(int index)
		{
			_current_value = _domain.at(static_cast<std::size_t>(index));
			auto mm = std::minmax_element(_domain.begin(), _domain.end());
			_min_value = *mm.first;
			_max_value = *mm.second;
		}

	public:
		struct valueException : std::runtime_error
		{
			valueException(int value, int lo, int hi)
				: std::runtime_error("Wrong value " + std::to_string(value) + " passed to Variable::set_value: not in the domain ["
				                     + std::to_string(lo) + ", " + std::to_string(hi) + "].\n") {}
		};

		Variable() = default;

		//! Explicit domain (may have holes); `index` is the position of the initial value.
		Variable(const std::vector<int>& domain, int index = 0, const std::string& name = std::string())
			: _domain(domain), _name(name) { finish(index); }

		//! Contiguous domain [starting_value, starting_value + size).
		Variable(int starting_value, std::size_t size, int index = 0, const std::string& name = std::string())
			: _domain(size), _name(name)
		{
			std::iota(_domain.begin(), _domain.end(), starting_value);
			finish(index);
		}

		Variable(const std::vector<int>& domain, const std::string& name) : Variable(domain, 0, name) {}
		Variable(int starting_value, std::size_t size, const std::string& name) : Variable(starting_value, size, 0, name) {}

		std::vector<int> get_full_domain() const { return _domain; }

		//! `range` values around the current one, circularly, in domain order.
		std::vector<int> get_partial_domain(int range) const
		{
			const int n = static_cast<int>(_domain.size());
			if (range >= n) return _domain;
			if (range <= 0) return {};
			const int at = static_cast<int>(std::find(_domain.begin(), _domain.end(), _current_value) - _domain.begin());
			std::vector<int> positions(static_cast<std::size_t>(range));
			for (int t = 0; t < range; ++t) positions[static_cast<std::size_t>(t)] = ((at - range / 2 + t) % n + n) % n;
			std::sort(positions.begin(), positions.end());
			std::vector<int> out;
			out.reserve(positions.size());
			for (int p : positions) out.push_back(_domain[static_cast<std::size_t>(p)]);
			return out;
		}

		int get_value() const { return _current_value; }

		void set_value(int value)
		{
			if (std::find(_domain.begin(), _domain.end(), value) == _domain.end())
				throw valueException(value, _min_value, _max_value);
			_current_value = value;
		}

		std::size_t get_domain_size() const { return _domain.size(); }
		int get_domain_min_value() const { return _min_value; }
		int get_domain_max_value() const { return _max_value; }
		std::string get_name() const { return _name; }
		int get_id() const { return _id; }

		friend std::ostream& operator<<(std::ostream& os, const Variable& v)
		{
			os << "Variable name: " << v._name << "\nId: " << v._id << "\nValue: " << v._current_value << "\nDomain: ";
			for (int d : v._domain) os << d << ", ";
			return os << "\n--------";
		}
	};
}
