// ghost/solver.hpp — ghost::Solver with the reference's public signatures (include/solver.hpp:118,
// 431-434, 851-902, 932-934, 1001, 1010), whose fast_search runs on the GPU.
//
// Reference: fast_search resolves option defaults (451-473), builds one Model + SearchUnit per thread
// and runs SearchUnit::local_search in each (492-587), then keeps the best (594-759).
// Here: build the Model once, lower it to the flat GhostModelDesc, call ghost_b200_search() — thousands
// of walkers instead of number_threads SearchUnits — and read the winner back.  A returned solution is
// re-scored with the host-side Constraint::error (the same checker role the reference's classes have);
// post-processing and the Maximize sign flip follow 734-754.  Models with user-defined constraints,
// objectives or auxiliary data are refused with the library's message: there is no CPU fallback.
#pragma once
#include <algorithm>
#include <chrono>
#include <cmath>
#include <functional>
#include <limits>
#include <random>
#include <stdexcept>
#include <string>
#include <vector>

#include "../ghost_b200.h"
#include "model_builder.hpp"
#include "options.hpp"
#include "global_constraints/all_different.hpp"
#include "global_constraints/all_different_abs_diff.hpp"
#include "global_constraints/all_different_offset.hpp"
#include "global_constraints/all_different_lag_diff.hpp"
#include "global_constraints/fix_value.hpp"
#include "global_constraints/linear_equation_eq.hpp"
#include "global_constraints/linear_equation_neq.hpp"
#include "global_constraints/linear_equation_leq.hpp"
#include "global_constraints/linear_equation_geq.hpp"
#include "global_constraints/linear_equation_l.hpp"
#include "global_constraints/linear_equation_g.hpp"
#include "global_objectives/linear_objective.hpp"

namespace ghost
{
	//! Counters of the last fast_search (the reference prints them under GHOST_BENCH, solver.hpp:764-833).
	struct SearchStatistics
	{
		GhostCounters winner{}, total{};
		double search_time_us = 0, kernel_time_us = 0;
		int walkers = 0, threads_per_walker = 0;
		long long kernel_launches = 0;
		double satisfaction_error = std::numeric_limits<double>::max();
	};

	template <typename ModelBuilderType> class Solver final
	{
		Model _model;
		ModelBuilderType _model_builder;
		SearchStatistics _stats;

		// Flat arrays of one lowered Model (they back the pointers of the GhostModelDesc).
		struct Lowered
		{
			std::vector<int32_t> dom_offset, dom_values, start_value, ctr_kind, ctr_offset, ctr_vars, obj_vars;
			std::vector<double> ctr_coefs, ctr_param, obj_coefs;
			GhostModelDesc desc{};
		};

		static void lower(const Model& model, Lowered& L)
		{
			if (!model.auxiliary_data->b200_is_null())
				throw std::runtime_error("ghost_b200: user-defined AuxiliaryData cannot run on the GPU path (no CPU fallback)");
			L.dom_offset.assign(1, 0);
			for (const auto& v : model.variables) {
				for (int d : v.get_full_domain()) L.dom_values.push_back(d);
				L.dom_offset.push_back(static_cast<int32_t>(L.dom_values.size()));
				L.start_value.push_back(v.get_value());
			}
			L.ctr_offset.assign(1, 0);
			for (const auto& c : model.constraints) {
				std::vector<double> coefs;
				double param = 0.0;
				c->b200_describe(coefs, param);
				const auto ids = c->get_variable_ids();
				L.ctr_kind.push_back(c->b200_kind());
				for (std::size_t i = 0; i < ids.size(); ++i) {
					L.ctr_vars.push_back(ids[i]);
					L.ctr_coefs.push_back(i < coefs.size() ? coefs[i] : 1.0);
				}
				L.ctr_offset.push_back(static_cast<int32_t>(L.ctr_vars.size()));
				L.ctr_param.push_back(param);
			}
			GhostModelDesc& d = L.desc;
			d.num_variables = static_cast<int32_t>(model.variables.size());
			d.num_constraints = static_cast<int32_t>(model.constraints.size());
			d.permutation_problem = model.permutation_problem ? 1 : 0;
			d.objective_kind = model.objective->b200_kind();
			if (d.objective_kind == GHOST_OBJ_LINEAR_MINIMIZE || d.objective_kind == GHOST_OBJ_LINEAR_MAXIMIZE) {
				model.objective->b200_describe(L.obj_coefs);
				for (int id : model.objective->get_variable_ids()) L.obj_vars.push_back(id);
				L.obj_coefs.resize(L.obj_vars.size(), 0.0);
			}
			d.num_objective_terms = static_cast<int32_t>(L.obj_vars.size());
			d.dom_offset = L.dom_offset.data(); d.dom_values = L.dom_values.data(); d.start_value = L.start_value.data();
			d.ctr_kind = L.ctr_kind.data(); d.ctr_offset = L.ctr_offset.data(); d.ctr_vars = L.ctr_vars.data();
			d.ctr_coefs = L.ctr_coefs.data(); d.ctr_param = L.ctr_param.data();
			d.obj_vars = L.obj_vars.data(); d.obj_coefs = L.obj_coefs.data();
		}

		double total_error() const
		{
			double e = 0.0;
			for (const auto& c : _model.constraints) e += c->error();
			return e;
		}

	public:
		Solver(const ModelBuilderType& model_builder) : _model_builder(model_builder) {}

		/*! Search for one solution (CSP) or a good one (COP) within `timeout` microseconds.
		 * Same contract as the reference (include/solver.hpp:431-434): returns true iff an assignment
		 * with satisfaction error 0 was found; final_cost is then 0 (CSP) or the objective value (COP,
		 * sign as the user sees it); otherwise final_cost is the best satisfaction error reached. */
		bool fast_search(double& final_cost, std::vector<int>& final_solution, double timeout, Options& options)
		{
			_model = _model_builder.build_model();
			Lowered L;
			lower(_model, L);
			GhostModel* gm = nullptr;
			if (ghost_b200_model_create(&L.desc, &gm) != GHOST_B200_OK) throw std::runtime_error(ghost_b200_last_error());

			GhostSearchOpts o;
			ghost_b200_default_opts(&o);
			o.custom_starting_point = options.custom_starting_point;
			o.resume_search = options.resume_search;
			o.enable_optimization_guidance = options.enable_optimization_guidance;
			o.tabu_time_local_min = options.tabu_time_local_min;
			o.tabu_time_selected = options.tabu_time_selected;
			o.percent_chance_force_trying_on_plateau = options.percent_chance_force_trying_on_plateau;
			o.reset_threshold = options.reset_threshold;
			o.restart_threshold = options.restart_threshold;
			o.number_variables_to_reset = options.number_variables_to_reset;
			o.number_start_samplings = options.number_start_samplings;
			o.number_walkers = options.number_walkers;
			o.threads_per_walker = options.threads_per_walker;
			o.epoch_iterations = options.epoch_iterations;
			o.strategy = options.strategy;
			o.elite_restart_percent = options.elite_restart_percent;
			o.seed = options.seed ? options.seed : (static_cast<uint64_t>(std::random_device{}()) << 32) ^ std::random_device{}();
			o.timeout_us = timeout;
			o.stop_on_first_solution = 1;

			final_solution.assign(_model.variables.size(), 0);
			GhostResult r{};
			r.solution = final_solution.data();
			// one GPU: the model handle created above; several: walkers sharded over the devices, one stop flag
			const int rc = options.number_gpus == 1 ? ghost_b200_search(gm, &o, &r)
			                                        : ghost_b200_search_devices(&L.desc, &o, options.number_gpus, &r);
			ghost_b200_model_destroy(gm);
			if (rc != GHOST_B200_OK) throw std::runtime_error(ghost_b200_last_error());

			_stats.winner = r.winner; _stats.total = r.total;
			_stats.search_time_us = r.search_time_us; _stats.kernel_time_us = r.kernel_time_us;
			_stats.walkers = r.walkers; _stats.threads_per_walker = r.threads_per_walker;
			_stats.kernel_launches = r.kernel_launches;
			_stats.satisfaction_error = r.best_sat_error;

			// the winning assignment becomes the model's state, and is re-scored by the host-side constraints
			for (std::size_t i = 0; i < final_solution.size(); ++i) _model.variables[i].set_value(final_solution[i]);
			const bool found = r.found != 0;
			if (found && total_error() != 0.0)
				throw std::logic_error("ghost_b200: the GPU reported a solution that the host-side constraints reject");

			const bool optimization = _model.objective->is_optimization();
			if (optimization) {
				double internal = r.best_opt_cost;                              // sign-flipped for Maximize
				if (found) internal = _model.objective->postprocess(internal);  // solver.hpp:734-741
				final_cost = _model.objective->is_maximization() ? -internal : internal; // 743-751
			} else
				final_cost = r.best_sat_error;
			return found;
		}

		bool fast_search(double& final_cost, std::vector<int>& final_solution, double timeout)
		{
			Options options;
			return fast_search(final_cost, final_solution, timeout, options);
		}
		bool fast_search(double& final_cost, std::vector<int>& final_solution, std::chrono::microseconds timeout, Options& options)
		{
			return fast_search(final_cost, final_solution, static_cast<double>(timeout.count()), options);
		}
		bool fast_search(double& final_cost, std::vector<int>& final_solution, std::chrono::microseconds timeout)
		{
			Options options;
			return fast_search(final_cost, final_solution, timeout, options);
		}

		/*! Enumerate every solution (same contract as the reference, include/solver.hpp:932-987): appends to
		 * final_solutions all assignments with error 0 in the reference's depth-first order (variable 0
		 * first, values in domain order) and to final_costs the objective value of each (user's sign; 0
		 * for a satisfaction problem); returns true iff there is at least one.  The enumeration runs on
		 * the GPU through ghost_b200_complete_search; every returned row is re-checked with the host-side
		 * constraints.  Models outside the stock constraint set are refused (no CPU fallback). */
		bool complete_search(std::vector<double>& final_costs, std::vector<std::vector<int>>& final_solutions, Options&)
		{
			_model = _model_builder.build_model();
			Lowered L;
			lower(_model, L);
			GhostModel* gm = nullptr;
			if (ghost_b200_model_create(&L.desc, &gm) != GHOST_B200_OK) throw std::runtime_error(ghost_b200_last_error());
			GhostSolutionSet* set = nullptr;
			const int rc = ghost_b200_complete_search(gm, 0, &set);
			ghost_b200_model_destroy(gm);
			if (rc != GHOST_B200_OK) throw std::runtime_error(ghost_b200_last_error());
			const int64_t n = ghost_b200_solutions_count(set);
			const std::size_t V = _model.variables.size();
			std::vector<int32_t> flat(static_cast<std::size_t>(n) * V);
			std::vector<double> costs(static_cast<std::size_t>(n));
			const int rr = n > 0 ? ghost_b200_solutions_read(set, 0, n, flat.data(), costs.data()) : GHOST_B200_OK;
			double kernel_us = 0; int64_t launches = 0;
			ghost_b200_solutions_stats(set, &kernel_us, &launches, nullptr, nullptr);
			ghost_b200_solutions_destroy(set);
			if (rr != GHOST_B200_OK) throw std::runtime_error(ghost_b200_last_error());
			_stats = SearchStatistics();
			_stats.kernel_time_us = kernel_us; _stats.kernel_launches = launches;
			for (int64_t i = 0; i < n; ++i) {
				std::vector<int> solution(flat.begin() + static_cast<std::ptrdiff_t>(i * V), flat.begin() + static_cast<std::ptrdiff_t>((i + 1) * V));
				for (std::size_t v = 0; v < V; ++v) _model.variables[v].set_value(solution[v]);
				if (total_error() != 0.0)
					throw std::logic_error("ghost_b200: the GPU enumerated an assignment that the host-side constraints reject");
				final_costs.push_back(costs[static_cast<std::size_t>(i)]);
				final_solutions.emplace_back(std::move(solution));
			}
			return n > 0;
		}
		bool complete_search(std::vector<double>& final_costs, std::vector<std::vector<int>>& final_solutions)
		{
			Options options;
			return complete_search(final_costs, final_solutions, options);
		}

		std::vector<Variable> get_variables() { return _model.variables; }
		//! Counters of the last fast_search (addition).
		const SearchStatistics& get_statistics() const { return _stats; }
	};

	template <typename B> Solver(const B&) -> Solver<B>;
}
