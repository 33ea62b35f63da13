// ghost/options.hpp — solver tunables: the fields of the reference's ghost::Options
// (include/options.hpp:44-58, defaults src/options.cpp:36-50; -1 = "resolved by the solver",
// include/solver.hpp:451-473) plus what a GPU population needs.
#pragma once
#include <algorithm>
#include <cstdint>
#include <memory>
#include <thread>

#include "print.hpp"

namespace ghost
{
	struct Options
	{
		bool custom_starting_point = false;
		bool resume_search = false;
		bool parallel_runs = false;             //!< kept for source compatibility: the GPU always runs parallel restarts
		bool enable_optimization_guidance = true;
		int number_threads = std::max(2, static_cast<int>(std::thread::hardware_concurrency()) / 2); //!< unused on the GPU
		std::shared_ptr<Print> print = std::make_shared<Print>();
		int tabu_time_local_min = -1;
		int tabu_time_selected = -1;
		int percent_chance_force_trying_on_plateau = -1;
		int reset_threshold = -1;
		int restart_threshold = -1;
		int number_variables_to_reset = -1;
		int number_start_samplings = -1;

		// ---- additions for the B200 path ----
		int number_walkers = 0;                 //!< parallel restarts on the GPU; 0 = fill it
		int threads_per_walker = 0;             //!< 0 = chosen from the model (32, 64, 128 or 256)
		int epoch_iterations = 0;               //!< iterations per kernel launch; 0 = adaptive
		std::uint64_t seed = 0;                 //!< Philox key; 0 = drawn from std::random_device
		int number_gpus = 1;                    //!< GPUs of this process to search on; 0 = every visible one
		int strategy = 0;                       //!< GhostStrategy: 0 = Adaptive Search heuristics (what the reference's Solver builds), 1 = Antidote heuristics, 2 = Culprit error projection, 3 = hill climbing, 4 = random walk (the reference's GHOST_HILL_CLIMBING / GHOST_RANDOM_WALK builds)
		int elite_restart_percent = 0;          //!< cooperative walkers: chance (%) that a restart begins from the search's best assignment
	};
}
