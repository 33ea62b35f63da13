// ghost/auxiliary_data.hpp — user hook called after value changes (reference include/auxiliary_data.hpp:
// 55-126).  Kept for API compatibility; the GPU path requires the null one (a user-defined update
// cannot run inside the walkers).
#pragma once
#include <map>
#include <vector>

#include "variable.hpp"

namespace ghost
{
	class AuxiliaryData
	{
		friend class ModelBuilder;
		template <typename B> friend class Solver;
		std::vector<Variable*> _variables;
		std::vector<int> _variables_index;
		std::map<int, int> _variables_position;

	protected:
		virtual void required_update(const std::vector<Variable*>& variables, int index, int new_value) = 0;

	public:
		AuxiliaryData() = default;
		AuxiliaryData(const std::vector<int>& variables_index) : _variables_index(variables_index) {}
		AuxiliaryData(const std::vector<Variable>& variables)
		{
			for (const auto& v : variables) _variables_index.push_back(v.get_id());
		}
		AuxiliaryData(const AuxiliaryData&) = default;
		AuxiliaryData(AuxiliaryData&&) = default;
		AuxiliaryData& operator=(const AuxiliaryData&) = delete;
		AuxiliaryData& operator=(AuxiliaryData&&) = delete;
		virtual ~AuxiliaryData() = default;
		//! true only for NullAuxiliaryData: the GPU path cannot call back into user code
		virtual bool b200_is_null() const { return false; }
	};

	class NullAuxiliaryData : public AuxiliaryData
	{
	public:
		NullAuxiliaryData() = default;
		void required_update(const std::vector<Variable*>&, int, int) override {}
		bool b200_is_null() const override { return true; }
	};
}
