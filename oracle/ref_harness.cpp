/*
 * ref_harness.cpp — TEST INFRASTRUCTURE ONLY.  Glue (our own code) that drives the UNMODIFIED
 * reference, compiled from the sources where they lie under /root/reference by oracle/Makefile
 * into oracle/_ref/ (git-ignored).  No reference source is copied into this repository.
 *
 * It gives the tests and bench.py three things:
 *   1. ref_eval_moves(): constraint errors, projected errors and per-candidate cumulated deltas
 *      computed by the reference's own Constraint::error / simulate_delta / update and its own
 *      AdaptiveSearchErrorProjection, for a model described by the same GhostModelDesc the GPU
 *      path takes.  The private members are reached through an explicit specialisation of
 *      ghost::Solver<>, which Constraint / ModelBuilder / Objective befriend
 *      (include/constraint.hpp:65, include/model_builder.hpp:64, include/objective.hpp:63).
 *   2. ref_fast_search(): the reference's Solver::fast_search (include/solver.hpp:431) on that
 *      model — the timed CPU baseline.
 *   3. with -DREF_MAIN: an executable that runs fast_search on a descriptor file; compiled with
 *      -DGHOST_BENCH it prints the reference's own counters (include/solver.hpp:764-833), with
 *      -DGHOST_TRACE the reference's own per-iteration trace (include/search_unit.hpp:105-144,
 *      741-746, 808-819, 893-1003) used as golden trajectory.
 *
 * Stock constraints come from the reference.  Two things GHOST does not ship are defined here
 * against the reference's public base classes, exactly as a GHOST user would: a linear objective
 * (cf. tutorial/wiki/src/objective_max_value.cpp:9-17) and the AllDifferentAbsDiff extension kind.
 */
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <map>
#include <memory>
#include <numeric>
#include <string>
#include <vector>

#include "solver.hpp"
#include "global_constraints/all_different.hpp"
#include "global_constraints/fix_value.hpp"
#include "global_constraints/all_equal.hpp"
#include "global_constraints/linear_equation_eq.hpp"
#include "global_constraints/linear_equation_neq.hpp"
#include "global_constraints/linear_equation_leq.hpp"
#include "global_constraints/linear_equation_geq.hpp"
#include "global_constraints/linear_equation_l.hpp"
#include "global_constraints/linear_equation_g.hpp"
#include "algorithms/adaptive_search_error_projection_algorithm.hpp"
#include "algorithms/antidote_search_variable_heuristic.hpp"
#include "algorithms/culprit_search_error_projection_algorithm.hpp"
#include "algorithms/uniform_variable_heuristic.hpp"
#include "algorithms/adaptive_search_variable_candidates_heuristic.hpp"
#include "algorithms/adaptive_search_value_heuristic.hpp"
#include "algorithms/antidote_search_variable_candidates_heuristic.hpp"
#include "algorithms/all_free_variable_candidates_heuristic.hpp"
#include "algorithms/null_error_projection_algorithm.hpp"
#include "algorithms/random_walk_value_heuristic.hpp"
#include "algorithms/antidote_search_value_heuristic.hpp"

#include "../include/ghost_b200.h"

using namespace ghost;
using namespace ghost::global_constraints;

/* ---- user-side additions, written against the reference's base classes ------------------- */

class HarnessLinearMin : public Minimize
{
	std::vector<double> _coef;
	double required_cost( const std::vector<Variable*>& variables ) const override
	{
		double total = 0.0;
		for( size_t i = 0 ; i < variables.size() ; ++i )
			total += variables[i]->get_value() * _coef[i];
		return total;
	}
public:
	HarnessLinearMin( const std::vector<int>& ids, const std::vector<double>& coef )
		: Minimize( ids, "LinearObjective" ), _coef( coef ) { }
};

class HarnessLinearMax : public Maximize
{
	std::vector<double> _coef;
	double required_cost( const std::vector<Variable*>& variables ) const override
	{
		double total = 0.0;
		for( size_t i = 0 ; i < variables.size() ; ++i )
			total += variables[i]->get_value() * _coef[i];
		return total;
	}
public:
	HarnessLinearMax( const std::vector<int>& ids, const std::vector<double>& coef )
		: Maximize( ids, "LinearObjective" ), _coef( coef ) { }
};

/* AllDifferentAbsDiff: the |x[i+1]-x[i]| of the scope must be pairwise different; error is the
 * SOFT_ALLDIFF measure on those differences.  No cache: the delta recounts from the scope. */
class HarnessAllDifferentAbsDiff : public Constraint
{
	static double pairs( const std::vector<int>& values )
	{
		std::map<int,int> count;
		for( size_t i = 0 ; i + 1 < values.size() ; ++i )
			++count[ std::abs( values[i+1] - values[i] ) ];
		double e = 0;
		for( auto& kv : count )
			if( kv.second > 1 )
				e += static_cast<double>( kv.second ) * ( kv.second - 1 ) / 2;
		return e;
	}
	double required_error( const std::vector<Variable*>& variables ) const override
	{
		std::vector<int> values;
		for( auto v : variables ) values.push_back( v->get_value() );
		return pairs( values );
	}
	double optional_delta_error( const std::vector<Variable*>& variables,
	                             const std::vector<int>& indexes,
	                             const std::vector<int>& candidate_values ) const override
	{
		std::vector<int> values;
		for( auto v : variables ) values.push_back( v->get_value() );
		double before = pairs( values );
		for( size_t i = 0 ; i < indexes.size() ; ++i ) values[ indexes[i] ] = candidate_values[i];
		return pairs( values ) - before;
	}
public:
	HarnessAllDifferentAbsDiff( const std::vector<int>& ids ) : Constraint( ids ) { }
};

/* AllDifferentLagDiff: the x[i+lag] - x[i] of the scope must be pairwise different (one row of the difference triangle
 * of a Costas array); SOFT_ALLDIFF measure on those differences.  No cache: the delta recounts from the scope. */
class HarnessAllDifferentLagDiff : public Constraint
{
	int _lag;
	double pairs( const std::vector<int>& values ) const
	{
		std::map<int,int> count;
		for( size_t i = 0 ; i + _lag < values.size() ; ++i )
			++count[ values[i+_lag] - values[i] ];
		double e = 0;
		for( auto& kv : count )
			if( kv.second > 1 )
				e += static_cast<double>( kv.second ) * ( kv.second - 1 ) / 2;
		return e;
	}
	double required_error( const std::vector<Variable*>& variables ) const override
	{
		std::vector<int> values;
		for( auto v : variables ) values.push_back( v->get_value() );
		return pairs( values );
	}
	double optional_delta_error( const std::vector<Variable*>& variables,
	                             const std::vector<int>& indexes,
	                             const std::vector<int>& candidate_values ) const override
	{
		std::vector<int> values;
		for( auto v : variables ) values.push_back( v->get_value() );
		double before = pairs( values );
		for( size_t i = 0 ; i < indexes.size() ; ++i ) values[ indexes[i] ] = candidate_values[i];
		return pairs( values ) - before;
	}
public:
	HarnessAllDifferentLagDiff( const std::vector<int>& ids, int lag ) : Constraint( ids ), _lag( lag ) { }
};

/* AllDifferentOffset: the x_i + o_i of the scope must be pairwise different (n-queens diagonals); the error is
 * AllDifferent's own measure (src/global_constraints/all_different.cpp:59-75) on the shifted values.  No cache:
 * the delta recounts from the scope, as a user who does not override conditional_update_data_structures gets. */
class HarnessAllDifferentOffset : public Constraint
{
	std::vector<int> _offset;
	double pairs( const std::vector<int>& values ) const
	{
		std::map<int,int> count;
		for( size_t i = 0 ; i < values.size() ; ++i )
			++count[ values[i] + _offset[i] ];
		double e = 0;
		for( auto& kv : count )
			if( kv.second > 1 )
				e += static_cast<double>( kv.second ) * ( kv.second - 1 ) / 2;
		return e;
	}
	double required_error( const std::vector<Variable*>& variables ) const override
	{
		std::vector<int> values;
		for( auto v : variables ) values.push_back( v->get_value() );
		return pairs( values );
	}
	double optional_delta_error( const std::vector<Variable*>& variables,
	                             const std::vector<int>& indexes,
	                             const std::vector<int>& candidate_values ) const override
	{
		std::vector<int> values;
		for( auto v : variables ) values.push_back( v->get_value() );
		double before = pairs( values );
		for( size_t i = 0 ; i < indexes.size() ; ++i ) values[ indexes[i] ] = candidate_values[i];
		return pairs( values ) - before;
	}
public:
	HarnessAllDifferentOffset( const std::vector<int>& ids, const std::vector<double>& offsets )
		: Constraint( ids ), _offset( offsets.begin(), offsets.end() ) { }
};

/* ---- a ModelBuilder fed by the flat descriptor --------------------------------------------- */

struct DescCopy
{
	int V = 0, C = 0, permutation = 0, objective_kind = 0;
	std::vector<int> dom_offset, dom_values, start_value, ctr_kind, ctr_offset, ctr_vars, obj_vars;
	std::vector<double> ctr_coefs, ctr_param, obj_coefs;

	DescCopy() = default;
	explicit DescCopy( const GhostModelDesc* d )
	{
		V = d->num_variables; C = d->num_constraints; permutation = d->permutation_problem;
		objective_kind = d->objective_kind;
		dom_offset.assign( d->dom_offset, d->dom_offset + V + 1 );
		dom_values.assign( d->dom_values, d->dom_values + dom_offset[V] );
		start_value.assign( d->start_value, d->start_value + V );
		ctr_kind.assign( d->ctr_kind, d->ctr_kind + C );
		ctr_offset.assign( d->ctr_offset, d->ctr_offset + C + 1 );
		ctr_vars.assign( d->ctr_vars, d->ctr_vars + ctr_offset[C] );
		if( d->ctr_coefs ) ctr_coefs.assign( d->ctr_coefs, d->ctr_coefs + ctr_offset[C] );
		else ctr_coefs.assign( ctr_offset[C], 1.0 );
		if( d->ctr_param ) ctr_param.assign( d->ctr_param, d->ctr_param + C );
		else ctr_param.assign( C, 0.0 );
		if( d->num_objective_terms > 0 )
		{
			obj_vars.assign( d->obj_vars, d->obj_vars + d->num_objective_terms );
			obj_coefs.assign( d->obj_coefs, d->obj_coefs + d->num_objective_terms );
		}
	}

	/* whitespace-separated text form, written by ghost_b200/modeldesc.py (ModelDesc.dump) */
	bool load( const char* path )
	{
		std::ifstream in( path );
		std::string magic;
		if( !( in >> magic ) || magic != "GHOSTDESC1" ) return false;
		int ndom, nnz, nobj;
		in >> V >> C >> permutation >> objective_kind >> ndom >> nnz >> nobj;
		auto read_i = [&]( std::vector<int>& v, int n ){ v.resize( n ); for( auto& e : v ) in >> e; };
		auto read_d = [&]( std::vector<double>& v, int n ){ v.resize( n ); for( auto& e : v ) in >> e; };
		read_i( dom_offset, V + 1 ); read_i( dom_values, ndom ); read_i( start_value, V );
		read_i( ctr_kind, C ); read_i( ctr_offset, C + 1 ); read_i( ctr_vars, nnz );
		read_d( ctr_coefs, nnz ); read_d( ctr_param, C ); read_i( obj_vars, nobj ); read_d( obj_coefs, nobj );
		return static_cast<bool>( in );
	}
};

class DescBuilder : public ModelBuilder
{
public:
	std::shared_ptr<const DescCopy> d;

	explicit DescBuilder( std::shared_ptr<const DescCopy> desc )
		: ModelBuilder( desc->permutation != 0 ), d( std::move( desc ) ) { }

	void declare_variables() override
	{
		for( int v = 0 ; v < d->V ; ++v )
		{
			std::vector<int> domain( d->dom_values.begin() + d->dom_offset[v], d->dom_values.begin() + d->dom_offset[v+1] );
			auto it = std::find( domain.begin(), domain.end(), d->start_value[v] );
			int index = it == domain.end() ? 0 : static_cast<int>( it - domain.begin() );
			variables.emplace_back( domain, index );
		}
	}

	void declare_constraints() override
	{
		for( int c = 0 ; c < d->C ; ++c )
		{
			std::vector<int> ids( d->ctr_vars.begin() + d->ctr_offset[c], d->ctr_vars.begin() + d->ctr_offset[c+1] );
			std::vector<double> coef( d->ctr_coefs.begin() + d->ctr_offset[c], d->ctr_coefs.begin() + d->ctr_offset[c+1] );
			double p = d->ctr_param[c];
			switch( d->ctr_kind[c] )
			{
			case GHOST_KIND_ALL_DIFFERENT: constraints.emplace_back( std::make_shared<AllDifferent>( ids ) ); break;
			case GHOST_KIND_LINEAR_EQ:  constraints.emplace_back( std::make_shared<LinearEquationEq>( ids, p, coef ) ); break;
			case GHOST_KIND_LINEAR_NEQ: constraints.emplace_back( std::make_shared<LinearEquationNeq>( ids, p, coef ) ); break;
			case GHOST_KIND_LINEAR_LEQ: constraints.emplace_back( std::make_shared<LinearEquationLeq>( ids, p, coef ) ); break;
			case GHOST_KIND_LINEAR_GEQ: constraints.emplace_back( std::make_shared<LinearEquationGeq>( ids, p, coef ) ); break;
			case GHOST_KIND_LINEAR_L:   constraints.emplace_back( std::make_shared<LinearEquationL>( ids, p, coef ) ); break;
			case GHOST_KIND_LINEAR_G:   constraints.emplace_back( std::make_shared<LinearEquationG>( ids, p, coef ) ); break;
			case GHOST_KIND_FIX_VALUE:  constraints.emplace_back( std::make_shared<FixValue>( ids, static_cast<int>( p ) ) ); break;
			case GHOST_KIND_ALL_DIFFERENT_ABS_DIFF: constraints.emplace_back( std::make_shared<HarnessAllDifferentAbsDiff>( ids ) ); break;
			case GHOST_KIND_ALL_EQUAL: constraints.emplace_back( std::make_shared<AllEqual>( ids ) ); break;
			case GHOST_KIND_ALL_DIFFERENT_LAG_DIFF: constraints.emplace_back( std::make_shared<HarnessAllDifferentLagDiff>( ids, static_cast<int>( p ) ) ); break;
			case GHOST_KIND_ALL_DIFFERENT_OFFSET: constraints.emplace_back( std::make_shared<HarnessAllDifferentOffset>( ids, coef ) ); break;
			case GHOST_KIND_PURE_OPTIMIZATION: break; /* build_model() inserts PureOptimization when there is no constraint */
			default: break;
			}
		}
	}

	void declare_objective() override
	{
		if( d->objective_kind == GHOST_OBJ_LINEAR_MINIMIZE )
			objective = std::make_shared<HarnessLinearMin>( d->obj_vars, d->obj_coefs );
		else if( d->objective_kind == GHOST_OBJ_LINEAR_MAXIMIZE )
			objective = std::make_shared<HarnessLinearMax>( d->obj_vars, d->obj_coefs );
		else
			ModelBuilder::declare_objective();
	}
};

/* ---- friend access ------------------------------------------------------------------------- */

struct OracleTag { };

namespace ghost
{
	template<> class Solver<OracleTag>
	{
	public:
		static Model build( ModelBuilder& b ) { return b.build_model(); }
		static double error( Constraint& c ) { return c.error(); }
		static void set_current_error( Constraint& c, double e ) { c._current_error = e; }
		static double current_error( const Constraint& c ) { return c._current_error; }
		static double simulate_delta( Constraint& c, const std::vector<int>& vars, const std::vector<int>& vals ) { return c.simulate_delta( vars, vals ); }
		static void update( Constraint& c, int var, int value ) { c.update( var, value ); }
		static double cost( const Objective& o ) { return o.cost(); }
	};
}
using Access = ghost::Solver<OracleTag>;

#ifndef REF_MAIN
/* ---- 1. move evaluation by the reference's own constraint code ----------------------------- */

struct RefModel
{
	std::shared_ptr<const DescCopy> desc;
	DescBuilder builder;
	Model model;
	SearchUnitData data;
	algorithms::AdaptiveSearchErrorProjection projection;

	explicit RefModel( std::shared_ptr<const DescCopy> d )
		: desc( d ), builder( d ), model( Access::build( builder ) ), data( model )
	{
		data.initialize_matrix( model ); /* include/search_unit_data.hpp:95-102 */
		projection.initialize_data_structures( data );
	}
};

extern "C" void* ref_model_create( const GhostModelDesc* d )
{
	try { return new RefModel( std::make_shared<const DescCopy>( d ) ); }
	catch( const std::exception& e ) { std::fprintf( stderr, "ref_model_create: %s\n", e.what() ); return nullptr; }
}
extern "C" void ref_model_destroy( void* p ) { delete static_cast<RefModel*>( p ); }

static void load_assignment( RefModel& r, const int32_t* x )
{
	for( int v = 0 ; v < r.desc->V ; ++v ) r.model.variables[v].set_value( x[v] );
	/* SearchUnit::compute_constraints_errors, include/search_unit.hpp:434-447 */
	for( auto& c : r.model.constraints ) Access::set_current_error( *c, Access::error( *c ) );
	r.data.current_sat_error = 0;
	for( auto& c : r.model.constraints ) r.data.current_sat_error += Access::current_error( *c );
	r.projection.compute_variable_errors( r.model.variables, r.model.constraints, r.data );
}

/* Same contract as oracle_eval_moves (NaN = not a candidate). */
extern "C" int ref_eval_moves( void* p, const int32_t* x, int32_t var, double* ctr_error, double* var_error,
                               double* cand_delta, int32_t cand_stride )
{
	RefModel& r = *static_cast<RefModel*>( p );
	try
	{
		load_assignment( r, x );
		Model& model = r.model;
		const int V = r.desc->V;
		for( int c = 0 ; c < static_cast<int>( model.constraints.size() ) ; ++c )
			if( ctr_error ) ctr_error[c] = Access::current_error( *model.constraints[c] );
		for( int v = 0 ; v < V ; ++v )
			if( var_error ) var_error[v] = r.data.error_variables[v];
		if( var < 0 || !cand_delta ) return 0;
		for( int k = 0 ; k < cand_stride ; ++k ) cand_delta[k] = NAN;

		/* the loops of include/search_unit.hpp:831-887, calling the reference's simulate_delta */
		const int variable_to_change = var;
		auto domain_to_explore = model.variables[ variable_to_change ].get_full_domain();
		domain_to_explore.erase( std::find( domain_to_explore.begin(), domain_to_explore.end(), model.variables[ variable_to_change ].get_value() ) );
		std::map<int, std::vector<double>> delta_errors;
		auto& matrix = r.data.matrix_var_ctr;
		if( !model.permutation_problem )
		{
			for( const auto candidate_value : domain_to_explore )
			{
				if( !matrix.at( variable_to_change ).empty() )
					for( const int constraint_id : matrix.at( variable_to_change ) )
						delta_errors[ candidate_value ].push_back( Access::simulate_delta( *model.constraints[ constraint_id ], std::vector<int>{variable_to_change}, std::vector<int>{candidate_value} ) );
				else
					delta_errors[ candidate_value ].push_back( 0.0 );
			}
		}
		else
		{
			for( int variable_id = 0 ; variable_id < V ; ++variable_id )
			{
				auto dom = model.variables[ variable_id ].get_full_domain();
				if( variable_id != variable_to_change
				    && model.variables[ variable_id ].get_value() != model.variables[ variable_to_change ].get_value()
				    && std::find( domain_to_explore.begin(), domain_to_explore.end(), model.variables[ variable_id ].get_value() ) != domain_to_explore.end()
				    && std::find( dom.begin(), dom.end(), model.variables[ variable_to_change ].get_value() ) != dom.end() )
				{
					std::vector<bool> constraint_checked( model.constraints.size(), false );
					int current_value = model.variables[ variable_to_change ].get_value();
					int candidate_value = model.variables[ variable_id ].get_value();
					for( const int constraint_id : matrix.at( variable_to_change ) )
					{
						constraint_checked[ constraint_id ] = true;
						if( model.constraints[ constraint_id ]->has_variable( variable_id ) )
							delta_errors[ variable_id ].push_back( Access::simulate_delta( *model.constraints[ constraint_id ], std::vector<int>{variable_to_change, variable_id}, std::vector<int>{candidate_value, current_value} ) );
						else
							delta_errors[ variable_id ].push_back( Access::simulate_delta( *model.constraints[ constraint_id ], std::vector<int>{variable_to_change}, std::vector<int>{candidate_value} ) );
					}
					for( const int constraint_id : matrix.at( variable_id ) )
						if( !constraint_checked[ constraint_id ] )
							delta_errors[ variable_id ].push_back( Access::simulate_delta( *model.constraints[ constraint_id ], std::vector<int>{variable_id}, std::vector<int>{current_value} ) );
				}
			}
		}
		/* cumulated as AdaptiveSearchValueHeuristic::select_value does (…value_heuristic.cpp:57-58) */
		for( const auto& deltas : delta_errors )
		{
			double sum = std::accumulate( deltas.second.begin(), deltas.second.end(), 0.0 );
			int slot;
			if( model.permutation_problem ) slot = deltas.first;
			else
			{
				/* slot = rank of the value in the sorted domain */
				auto dom = model.variables[ variable_to_change ].get_full_domain();
				std::sort( dom.begin(), dom.end() );
				slot = static_cast<int>( std::lower_bound( dom.begin(), dom.end(), deltas.first ) - dom.begin() );
			}
			if( slot >= 0 && slot < cand_stride ) cand_delta[ slot ] = sum;
		}
		return 0;
	}
	catch( const std::exception& e ) { std::fprintf( stderr, "ref_eval_moves: %s\n", e.what() ); return -1; }
}

/* Apply one move with the reference's update protocol (include/search_unit.hpp:449-541) and return
 * the constraint errors it maintains incrementally: checks update() / _current_error += delta. */
extern "C" int ref_apply_move( void* p, const int32_t* x, int32_t var, int32_t new_value, double* ctr_error_after,
                               int32_t* x_after )
{
	RefModel& r = *static_cast<RefModel*>( p );
	try
	{
		load_assignment( r, x );
		Model& model = r.model;
		auto& matrix = r.data.matrix_var_ctr;
		if( !model.permutation_problem )
		{
			for( const int c : matrix.at( var ) )
			{
				double delta = Access::simulate_delta( *model.constraints[c], std::vector<int>{var}, std::vector<int>{new_value} );
				Access::set_current_error( *model.constraints[c], Access::current_error( *model.constraints[c] ) + delta );
				Access::update( *model.constraints[c], var, new_value );
			}
			model.variables[var].set_value( new_value );
		}
		else
		{
			std::vector<bool> checked( model.constraints.size(), false );
			int current_value = model.variables[var].get_value();
			int next_value = model.variables[new_value].get_value();
			std::vector<std::pair<int,double>> deltas;
			for( const int c : matrix.at( var ) )
			{
				checked[c] = true;
				if( model.constraints[c]->has_variable( new_value ) )
					deltas.emplace_back( c, Access::simulate_delta( *model.constraints[c], std::vector<int>{var, new_value}, std::vector<int>{next_value, current_value} ) );
				else
					deltas.emplace_back( c, Access::simulate_delta( *model.constraints[c], std::vector<int>{var}, std::vector<int>{next_value} ) );
			}
			for( const int c : matrix.at( new_value ) )
				if( !checked[c] )
					deltas.emplace_back( c, Access::simulate_delta( *model.constraints[c], std::vector<int>{new_value}, std::vector<int>{current_value} ) );
			std::fill( checked.begin(), checked.end(), false );
			size_t di = 0;
			for( const int c : matrix.at( var ) )
			{
				checked[c] = true;
				Access::set_current_error( *model.constraints[c], Access::current_error( *model.constraints[c] ) + deltas[di++].second );
				Access::update( *model.constraints[c], var, next_value );
				if( model.constraints[c]->has_variable( new_value ) )
					Access::update( *model.constraints[c], new_value, current_value );
			}
			for( const int c : matrix.at( new_value ) )
				if( !checked[c] )
				{
					Access::set_current_error( *model.constraints[c], Access::current_error( *model.constraints[c] ) + deltas[di++].second );
					Access::update( *model.constraints[c], new_value, current_value );
				}
			model.variables[var].set_value( next_value );
			model.variables[new_value].set_value( current_value );
		}
		for( size_t c = 0 ; c < model.constraints.size() ; ++c )
			ctr_error_after[c] = Access::current_error( *model.constraints[c] );
		for( int v = 0 ; v < r.desc->V ; ++v ) x_after[v] = model.variables[v].get_value();
		return 0;
	}
	catch( const std::exception& e ) { std::fprintf( stderr, "ref_apply_move: %s\n", e.what() ); return -1; }
}

/* Σ error() and user-visible objective value of an assignment: the reference's own checker. */
extern "C" int ref_check( void* p, const int32_t* x, double* sat_error, double* objective )
{
	RefModel& r = *static_cast<RefModel*>( p );
	try
	{
		load_assignment( r, x );
		if( sat_error ) *sat_error = r.data.current_sat_error;
		if( objective )
		{
			double c = Access::cost( *r.model.objective );
			*objective = r.model.objective->is_maximization() ? -c : c;
		}
		return 0;
	}
	catch( const std::exception& e ) { std::fprintf( stderr, "ref_check: %s\n", e.what() ); return -1; }
}

/* ---- 2. the reference solver itself -------------------------------------------------------- */
extern "C" int ref_fast_search( const GhostModelDesc* d, double timeout_us, int parallel_runs, int number_threads,
                                int custom_start, double* final_cost, int32_t* solution, double* elapsed_us )
{
	try
	{
		DescBuilder builder( std::make_shared<const DescCopy>( d ) );
		Solver<DescBuilder> solver( builder );
		Options options;
		options.parallel_runs = parallel_runs != 0;
		if( number_threads > 0 ) options.number_threads = number_threads;
		options.custom_starting_point = custom_start != 0;
		double cost = 0;
		std::vector<int> sol;
		auto t0 = std::chrono::steady_clock::now();
		bool found = solver.fast_search( cost, sol, timeout_us, options );
		std::chrono::duration<double,std::micro> dt = std::chrono::steady_clock::now() - t0;
		if( final_cost ) *final_cost = cost;
		if( solution ) for( size_t i = 0 ; i < sol.size() ; ++i ) solution[i] = sol[i];
		if( elapsed_us ) *elapsed_us = dt.count();
		return found ? 1 : 0;
	}
	catch( const std::exception& e ) { std::fprintf( stderr, "ref_fast_search: %s\n", e.what() ); return -1; }
}

/* ---- 2b. the reference's own complete search (include/solver.hpp:932-987) ------------------------ */
extern "C" int ref_complete_search( const GhostModelDesc* d, int64_t max_solutions, int32_t* solutions, double* costs, int64_t* count )
{
	try
	{
		DescBuilder builder( std::make_shared<const DescCopy>( d ) );
		Solver<DescBuilder> solver( builder );
		Options options;
		std::vector<double> final_costs;
		std::vector<std::vector<int>> final_solutions;
		bool found = solver.complete_search( final_costs, final_solutions, options );
		int64_t n = static_cast<int64_t>( final_solutions.size() );
		if( count ) *count = n;
		int64_t keep = max_solutions > 0 && n > max_solutions ? max_solutions : n;
		for( int64_t i = 0 ; i < keep ; ++i )
		{
			if( solutions ) for( int v = 0 ; v < d->num_variables ; ++v ) solutions[ i * d->num_variables + v ] = final_solutions[i][v];
			if( costs ) costs[i] = final_costs[i];
		}
		return found ? 1 : 0;
	}
	catch( const std::exception& e ) { std::fprintf( stderr, "ref_complete_search: %s\n", e.what() ); return -1; }
}

#else /* REF_MAIN */
/* ---- 3. executable: fast_search on a descriptor file; GHOST_BENCH / GHOST_TRACE decide what the
 *         reference itself prints.  usage: ref_bench desc.txt timeout_us parallel threads [custom_start] */
int main( int argc, char** argv )
{
	if( argc < 5 ) { std::fprintf( stderr, "usage: %s desc.txt timeout_us parallel_runs threads [custom_start [strategy: 0 adaptive search, 1 antidote, 2 culprit projection, 3 hill climbing, 4 random walk]]\n", argv[0] ); return 2; }
	auto desc = std::make_shared<DescCopy>();
	if( !desc->load( argv[1] ) ) { std::fprintf( stderr, "cannot read %s\n", argv[1] ); return 2; }
	DescBuilder builder( desc );
	Solver<DescBuilder> solver( builder );
	Options options;
	options.parallel_runs = std::atoi( argv[3] ) != 0;
	if( std::atoi( argv[4] ) > 0 ) options.number_threads = std::atoi( argv[4] );
	if( argc > 5 ) options.custom_starting_point = std::atoi( argv[5] ) != 0;
	double cost = 0;
	std::vector<int> solution;
	const int strategy = argc > 6 ? std::atoi( argv[6] ) : 0;
	if( strategy >= 1 && strategy <= 4 )
	{
		/* The Antidote heuristic set.  Solver::fast_search never builds it (include/solver.hpp:499-518), so the SearchUnit
		 * is constructed here the way fast_search does for the default set, with the option defaults resolved as at
		 * include/solver.hpp:451-473, and run with its own local_search (and its own GHOST_TRACE output). */
		const int V = desc->V;
		if( options.tabu_time_local_min < 0 ) options.tabu_time_local_min = std::max( std::min( 5, V - 1 ), static_cast<int>( std::ceil( V / 5 ) ) ) + 1;
		if( options.tabu_time_selected < 0 ) options.tabu_time_selected = 0;
		if( options.percent_chance_force_trying_on_plateau < 0 || options.percent_chance_force_trying_on_plateau > 100 ) options.percent_chance_force_trying_on_plateau = 10;
		if( options.reset_threshold < 0 ) options.reset_threshold = options.tabu_time_local_min;
		if( options.restart_threshold < 0 ) options.restart_threshold = V;
		if( options.number_variables_to_reset < 0 ) options.number_variables_to_reset = std::max( 2, static_cast<int>( std::ceil( V * 0.1 ) ) );
		if( options.number_start_samplings < 0 ) options.number_start_samplings = 10;
		/* 1: the three Antidote heuristics over the default projection; 2: the default heuristics over the Culprit projection;
		 * 3 / 4: what Solver::fast_search builds under GHOST_HILL_CLIMBING / GHOST_RANDOM_WALK (include/solver.hpp:502-515) —
		 * 4 only means a random walk in a binary compiled with -DGHOST_RANDOM_WALK (the loop itself changes under that macro,
		 * include/search_unit.hpp:1007-1036): oracle/_ref/ref_trace_rw */
		std::unique_ptr<algorithms::VariableHeuristic> vh;
		std::unique_ptr<algorithms::VariableCandidatesHeuristic> vch;
		std::unique_ptr<algorithms::ValueHeuristic> valh;
		std::unique_ptr<algorithms::ErrorProjection> ep;
		if( strategy == 1 )
		{
			vh = std::make_unique<algorithms::AntidoteSearchVariableHeuristic>();
			vch = std::make_unique<algorithms::AntidoteSearchVariableCandidatesHeuristic>();
			valh = std::make_unique<algorithms::AntidoteSearchValueHeuristic>();
			ep = std::make_unique<algorithms::AdaptiveSearchErrorProjection>();
		}
		else if( strategy == 2 )
		{
			vh = std::make_unique<algorithms::UniformVariableHeuristic>();
			vch = std::make_unique<algorithms::AdaptiveSearchVariableCandidatesHeuristic>();
			valh = std::make_unique<algorithms::AdaptiveSearchValueHeuristic>();
			ep = std::make_unique<algorithms::CulpritSearchErrorProjection>();
		}
		else
		{
#if !defined GHOST_RANDOM_WALK
			if( strategy == 4 ) { std::fprintf( stderr, "strategy 4 needs the binary compiled with -DGHOST_RANDOM_WALK (ref_trace_rw)\n" ); return 2; }
#endif
			vh = std::make_unique<algorithms::UniformVariableHeuristic>();
			vch = std::make_unique<algorithms::AllFreeVariableCandidatesHeuristic>();
			if( strategy == 4 ) valh = std::make_unique<algorithms::RandomWalkValueHeuristic>();
			else valh = std::make_unique<algorithms::AdaptiveSearchValueHeuristic>();
			ep = std::make_unique<algorithms::NullErrorProjection>();
		}
		SearchUnit unit( Access::build( builder ), options, std::move( vh ), std::move( vch ), std::move( valh ), std::move( ep ) );
		unit.local_search( std::atof( argv[2] ) );
		std::cout << "HARNESS antidote best_sat_error=" << unit.data.best_sat_error << " iterations=" << unit.data.search_iterations
		          << " resets=" << unit.data.resets << "\nHARNESS solution:";
		for( int v : unit.final_solution ) std::cout << " " << v;
		std::cout << "\n";
		return 0;
	}
	bool found = solver.fast_search( cost, solution, std::atof( argv[2] ), options );
	std::cout << "HARNESS found=" << found << " cost=" << cost << " threads=" << options.number_threads << "\nHARNESS solution:";
	for( int v : solution ) std::cout << " " << v;
	std::cout << "\n";
	return 0;
}
#endif
