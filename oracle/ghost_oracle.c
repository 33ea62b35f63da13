/*
 * ghost_oracle.c — TEST INFRASTRUCTURE ONLY (see ghost_oracle.h).  Never shipped, never on the
 * product path.
 *
 * Restates, function by function, the reference's hot path (citations are file:line under the
 * reference tree, richoux/GHOST v3.2.0).  Arithmetic is done in double exactly where the
 * reference uses double, so that this file reproduces the reference's numbers, not the GPU's
 * integer shortcut of them.
 */
#include "ghost_oracle.h"

#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------------------------------
 * Philox4x32-10 (Salmon et al., "Parallel random numbers: as easy as 1, 2, 3", SC'11).
 * Not in the reference: replaces thirdparty/randutils.hpp's entropy-seeded mt19937.
 * ------------------------------------------------------------------------------------------ */
void oracle_philox4x32_10(const uint32_t ctr_in[4], const uint32_t key_in[2], uint32_t out[4])
{
	uint32_t c0 = ctr_in[0], c1 = ctr_in[1], c2 = ctr_in[2], c3 = ctr_in[3];
	uint32_t k0 = key_in[0], k1 = key_in[1];
	for (int round = 0; round < 10; ++round) {
		uint64_t p0 = (uint64_t)0xD2511F53u * c0;
		uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
		uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
		uint32_t n1 = (uint32_t)p1;
		uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
		uint32_t n3 = (uint32_t)p0;
		c0 = n0; c1 = n1; c2 = n2; c3 = n3;
		k0 += 0x9E3779B9u;
		k1 += 0xBB67AE85u;
	}
	out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

typedef struct Rng {
	uint64_t seed;
	uint64_t walker;
	uint64_t next_event;
} Rng;

/* word `w` (0..3) of the block (event, slot) of this walker's stream */
static void rng_block(const Rng* r, uint64_t event, uint32_t slot, uint32_t out[4])
{
	uint32_t ctr[4] = { (uint32_t)event, (uint32_t)(event >> 32), slot, (uint32_t)r->walker };
	uint32_t key[2] = { (uint32_t)r->seed, (uint32_t)(r->seed >> 32) ^ (uint32_t)(r->walker >> 32) };
	oracle_philox4x32_10(ctr, key, out);
}
static uint32_t rng_word(const Rng* r, uint64_t event, uint32_t index)
{
	uint32_t blk[4];
	rng_block(r, event, index >> 2, blk);
	return blk[index & 3];
}
/* uniform integer in [0,n): multiply-high (bias <= n/2^32) */
static uint32_t below(uint32_t word, uint32_t n) { return (uint32_t)(((uint64_t)word * n) >> 32); }

/* ------------------------------------------------------------------------------------------ */
struct OracleModel {
	int32_t V, C, permutation;
	int32_t* dom_off;   /* [V+1] */
	int32_t* dom;       /* sorted ascending per variable */
	int32_t* dom_given; /* the same values in the order of the descriptor (complete search) */
	int32_t* start;
	int32_t* kind;
	int32_t* coff;      /* [C+1] */
	int32_t* cvars;
	double*  ccoef;
	double*  cparam;
	/* matrix_var_ctr (include/search_unit_data.hpp:95-102): for every variable the ascending
	 * list of constraints containing it */
	int32_t* v2c_off;   /* [V+1] */
	int32_t* v2c;       /* constraint ids */
	int32_t* v2c_pos;   /* position of the variable inside that constraint's scope */
	/* count tables of the AllDifferent-like constraints */
	int32_t* cnt_off;   /* [C] offset into a per-walker count array, -1 if none */
	int32_t* cnt_base;  /* [C] smallest value counted */
	int32_t* cnt_size;  /* [C] */
	int32_t  cnt_total;
	int32_t  obj_kind;
	double*  obj_coef;  /* dense [V] */
	int32_t  max_candidates;
};

static int pair_lag(const struct OracleModel* m, int c) { return m->kind[c] == GHOST_KIND_ALL_DIFFERENT_LAG_DIFF ? (int)m->cparam[c] : 1; }
static int cmp_i32(const void* a, const void* b)
{
	int32_t x = *(const int32_t*)a, y = *(const int32_t*)b;
	return (x > y) - (x < y);
}

static int is_alldiff_like(int32_t kind) /* kinds that keep a value-count table */
{
	return kind == GHOST_KIND_ALL_DIFFERENT || kind == GHOST_KIND_ALL_DIFFERENT_ABS_DIFF || kind == GHOST_KIND_ALL_EQUAL
	    || kind == GHOST_KIND_ALL_DIFFERENT_OFFSET || kind == GHOST_KIND_ALL_DIFFERENT_LAG_DIFF;
}
/* the two pair-difference kinds: |x[p+1] - x[p]| (ABS_DIFF) and x[p+lag] - x[p] (LAG_DIFF, lag = its parameter) */
static int is_pair_diff(int32_t kind) { return kind == GHOST_KIND_ALL_DIFFERENT_ABS_DIFF || kind == GHOST_KIND_ALL_DIFFERENT_LAG_DIFF; }
static int pair_lag(const struct OracleModel* m, int c);
static int32_t pair_value(int32_t kind, int32_t first, int32_t second) { return kind == GHOST_KIND_ALL_DIFFERENT_ABS_DIFF ? abs(second - first) : second - first; }
static int is_linear(int32_t kind) { return kind >= GHOST_KIND_LINEAR_EQ && kind <= GHOST_KIND_LINEAR_G; }

OracleModel* oracle_model_create(const GhostModelDesc* d)
{
	OracleModel* m = (OracleModel*)calloc(1, sizeof *m);
	int32_t V = d->num_variables, C = d->num_constraints;
	m->V = V; m->C = C; m->permutation = d->permutation_problem;
	int32_t ndom = d->dom_offset[V], nnz = d->ctr_offset[C];
	m->dom_off = (int32_t*)malloc(sizeof(int32_t) * (V + 1));
	m->dom = (int32_t*)malloc(sizeof(int32_t) * (ndom > 0 ? ndom : 1));
	m->start = (int32_t*)malloc(sizeof(int32_t) * (V > 0 ? V : 1));
	memcpy(m->dom_off, d->dom_offset, sizeof(int32_t) * (V + 1));
	memcpy(m->dom, d->dom_values, sizeof(int32_t) * ndom);
	m->dom_given = (int32_t*)malloc(sizeof(int32_t) * (ndom > 0 ? ndom : 1));
	memcpy(m->dom_given, d->dom_values, sizeof(int32_t) * ndom);
	memcpy(m->start, d->start_value, sizeof(int32_t) * V);
	int32_t maxdom = 0;
	for (int v = 0; v < V; ++v) {
		int32_t n = m->dom_off[v + 1] - m->dom_off[v];
		qsort(m->dom + m->dom_off[v], n, sizeof(int32_t), cmp_i32);
		if (n > maxdom) maxdom = n;
	}
	m->max_candidates = m->permutation ? V : maxdom;
	m->kind = (int32_t*)malloc(sizeof(int32_t) * C);
	m->coff = (int32_t*)malloc(sizeof(int32_t) * (C + 1));
	m->cvars = (int32_t*)malloc(sizeof(int32_t) * (nnz > 0 ? nnz : 1));
	m->ccoef = (double*)malloc(sizeof(double) * (nnz > 0 ? nnz : 1));
	m->cparam = (double*)malloc(sizeof(double) * C);
	memcpy(m->kind, d->ctr_kind, sizeof(int32_t) * C);
	memcpy(m->coff, d->ctr_offset, sizeof(int32_t) * (C + 1));
	memcpy(m->cvars, d->ctr_vars, sizeof(int32_t) * nnz);
	for (int i = 0; i < nnz; ++i) m->ccoef[i] = d->ctr_coefs ? d->ctr_coefs[i] : 1.0;
	for (int c = 0; c < C; ++c) m->cparam[c] = d->ctr_param ? d->ctr_param[c] : 0.0;

	/* SearchUnitData::initialize_matrix, include/search_unit_data.hpp:95-102 */
	m->v2c_off = (int32_t*)calloc(V + 2, sizeof(int32_t));
	for (int i = 0; i < nnz; ++i) m->v2c_off[m->cvars[i] + 1]++;
	for (int v = 0; v < V; ++v) m->v2c_off[v + 1] += m->v2c_off[v];
	m->v2c = (int32_t*)malloc(sizeof(int32_t) * (nnz > 0 ? nnz : 1));
	m->v2c_pos = (int32_t*)malloc(sizeof(int32_t) * (nnz > 0 ? nnz : 1));
	int32_t* fill = (int32_t*)calloc(V + 1, sizeof(int32_t));
	for (int c = 0; c < C; ++c)
		for (int i = m->coff[c]; i < m->coff[c + 1]; ++i) {
			int v = m->cvars[i];
			int at = m->v2c_off[v] + fill[v]++;
			m->v2c[at] = c;
			m->v2c_pos[at] = i - m->coff[c];
		}
	free(fill);

	m->cnt_off = (int32_t*)malloc(sizeof(int32_t) * C);
	m->cnt_base = (int32_t*)calloc(C, sizeof(int32_t));
	m->cnt_size = (int32_t*)calloc(C, sizeof(int32_t));
	m->cnt_total = 0;
	for (int c = 0; c < C; ++c) {
		m->cnt_off[c] = -1;
		if (!is_alldiff_like(m->kind[c])) continue;
		int32_t lo = INT32_MAX, hi = INT32_MIN;
		for (int i = m->coff[c]; i < m->coff[c + 1]; ++i) {
			int v = m->cvars[i];
			/* AllDifferentOffset counts the shifted values x + o */
			int32_t off = m->kind[c] == GHOST_KIND_ALL_DIFFERENT_OFFSET ? (int32_t)m->ccoef[i] : 0;
			if (m->dom_off[v + 1] > m->dom_off[v]) {
				int32_t a = m->dom[m->dom_off[v]] + off, b = m->dom[m->dom_off[v + 1] - 1] + off;
				if (a < lo) lo = a;
				if (b > hi) hi = b;
			}
			/* start values may lie outside the domain only if the user built it so; include them */
			if (m->start[v] + off < lo) lo = m->start[v] + off;
			if (m->start[v] + off > hi) hi = m->start[v] + off;
		}
		if (lo > hi) { lo = 0; hi = 0; }
		if (m->kind[c] == GHOST_KIND_ALL_DIFFERENT_ABS_DIFF) { hi = hi - lo; lo = 0; }
		if (m->kind[c] == GHOST_KIND_ALL_DIFFERENT_LAG_DIFF) { hi = hi - lo; lo = -hi; } /* signed differences */
		m->cnt_off[c] = m->cnt_total;
		m->cnt_base[c] = lo;
		m->cnt_size[c] = hi - lo + 1;
		m->cnt_total += m->cnt_size[c];
	}
	m->obj_kind = d->objective_kind;
	m->obj_coef = (double*)calloc(V > 0 ? V : 1, sizeof(double));
	for (int i = 0; i < d->num_objective_terms; ++i) m->obj_coef[d->obj_vars[i]] += d->obj_coefs[i];
	return m;
}

void oracle_model_destroy(OracleModel* m)
{
	if (!m) return;
	free(m->dom_off); free(m->dom); free(m->dom_given); free(m->start); free(m->kind); free(m->coff); free(m->cvars);
	free(m->ccoef); free(m->cparam); free(m->v2c_off); free(m->v2c); free(m->v2c_pos);
	free(m->cnt_off); free(m->cnt_base); free(m->cnt_size); free(m->obj_coef);
	free(m);
}

int32_t oracle_max_candidates(const OracleModel* m) { return m->max_candidates; }

/* ------------------------------------------------------------------------------------------
 * Walker = SearchUnit + SearchUnitData + the mutable caches of its private Model copy.
 * ------------------------------------------------------------------------------------------ */
typedef struct Walker {
	const OracleModel* m;
	GhostSearchOpts opt;
	Rng rng;
	int32_t* x;            /* Variable::_current_value */
	double*  cur_err;      /* Constraint::_current_error */
	double*  cur_sum;      /* LinearEquation::_current_sum */
	int32_t* count;        /* AllDifferent::_count, dense */
	int32_t* tabu;         /* SearchUnitData::tabu_list */
	double*  err_var;      /* SearchUnitData::error_variables */
	double best_sat, best_opt, cur_sat, cur_opt;
	int32_t* best_x;       /* SearchUnit::final_solution */
	int32_t* cand; int32_t ncand; int must_compute;
	GhostCounters k;
	/* scratch */
	int32_t* scratch_a; int32_t* scratch_b; uint8_t* flag; int32_t* sample;
	double* deltas;        /* per-constraint deltas of one candidate */
	/* Antidote strategy: variable_candidates is the VECTOR OF WEIGHTS (projected errors, tabu ones zeroed), not a list of
	 * ids (antidote_search_variable_candidates_heuristic.cpp:39-48); it shrinks by the erase of search_unit.hpp:825-828 */
	double* wvec; int32_t nvec;
	double* cand_sum;      /* cumulated delta of every candidate slot (Antidote value heuristic needs them all) */
	/* Culprit Search projection: CulpritSearchErrorProjection::_error_variables_by_constraints, a dense C x V table in the
	 * reference; row c holds one value per scope entry (culprit[p], p indexing cvars) and ONE common value for every
	 * variable outside the scope (culprit_out[c]) — see culprit_on_constraint */
	double* culprit; double* culprit_out;
} Walker;

static int in_domain(const OracleModel* m, int v, int32_t value)
{
	/* std::find over the domain (include/search_unit.hpp:857-860) */
	int32_t lo = m->dom_off[v], hi = m->dom_off[v + 1];
	while (lo < hi) {
		int32_t mid = (lo + hi) / 2;
		if (m->dom[mid] == value) return 1;
		if (m->dom[mid] < value) lo = mid + 1; else hi = mid;
	}
	return 0;
}

/* AllDifferent::binomial_with_2, src/global_constraints/all_different.cpp:38-47 */
static double binomial_with_2(int value)
{
	if (value <= 1) return 0;
	if (value % 2 == 0) return (double)(value - 1) * (value / 2);
	return (double)value * ((value - 1) / 2);
}

/* LinearEquation*::compute_error: linear_equation_eq.cpp:54-57, _neq.cpp:54-61, _leq.cpp:54-57,
 * _geq.cpp:54-57, _l.cpp:54-61, _g.cpp:54-61 */
static double linear_compute_error(int kind, double sum, double rhs)
{
	switch (kind) {
	case GHOST_KIND_LINEAR_EQ:  return fabs(sum - rhs);
	case GHOST_KIND_LINEAR_NEQ: return rhs == sum ? 1.0 : 0.0;
	case GHOST_KIND_LINEAR_LEQ: return fmax(0.0, sum - rhs);
	case GHOST_KIND_LINEAR_GEQ: return fmax(0.0, rhs - sum);
	case GHOST_KIND_LINEAR_L:   return fmax(0.0, sum - rhs) + (rhs == sum ? 1.0 : 0.0);
	case GHOST_KIND_LINEAR_G:   return fmax(0.0, rhs - sum) + (rhs == sum ? 1.0 : 0.0);
	}
	return 0.0;
}

/* Constraint::error -> required_error (src/constraint.cpp:57-65).  Rebuilds the constraint's
 * cache as the reference's mutable members do. */
static double ctr_error(Walker* w, int c)
{
	const OracleModel* m = w->m;
	int b = m->coff[c], e = m->coff[c + 1];
	switch (m->kind[c]) {
	case GHOST_KIND_ALL_DIFFERENT:
	case GHOST_KIND_ALL_DIFFERENT_OFFSET: {
		/* all_different.cpp:59-75; the extension kind counts x + o instead of x */
		int32_t* cnt = w->count + m->cnt_off[c];
		const int shifted = m->kind[c] == GHOST_KIND_ALL_DIFFERENT_OFFSET;
		memset(cnt, 0, sizeof(int32_t) * m->cnt_size[c]);
		for (int i = b; i < e; ++i) cnt[w->x[m->cvars[i]] + (shifted ? (int32_t)m->ccoef[i] : 0) - m->cnt_base[c]]++;
		double counter = 0;
		for (int t = 0; t < m->cnt_size[c]; ++t)
			if (cnt[t] > 1) counter += binomial_with_2(cnt[t]);
		return counter;
	}
	case GHOST_KIND_ALL_DIFFERENT_ABS_DIFF:
	case GHOST_KIND_ALL_DIFFERENT_LAG_DIFF: {
		/* extensions: same SOFT_ALLDIFF measure on the |x[i+1]-x[i]|, resp. the x[i+lag]-x[i], of the scope */
		int32_t* cnt = w->count + m->cnt_off[c];
		const int lag = pair_lag(m, c);
		memset(cnt, 0, sizeof(int32_t) * m->cnt_size[c]);
		for (int i = b; i + lag < e; ++i) cnt[pair_value(m->kind[c], w->x[m->cvars[i]], w->x[m->cvars[i + lag]]) - m->cnt_base[c]]++;
		double counter = 0;
		for (int t = 0; t < m->cnt_size[c]; ++t)
			if (cnt[t] > 1) counter += binomial_with_2(cnt[t]);
		return counter;
	}
	case GHOST_KIND_ALL_EQUAL: {
		/* all_equal.cpp:47-60: scope size minus the count of the most frequent value */
		int32_t* cnt = w->count + m->cnt_off[c];
		memset(cnt, 0, sizeof(int32_t) * m->cnt_size[c]);
		for (int i = b; i < e; ++i) cnt[w->x[m->cvars[i]] - m->cnt_base[c]]++;
		int32_t mx = 0;
		for (int t = 0; t < m->cnt_size[c]; ++t) if (cnt[t] > mx) mx = cnt[t];
		return (double)(e - b) - mx;
	}
	case GHOST_KIND_FIX_VALUE: {
		/* fix_value.cpp:46-52 */
		double error = 0.;
		int value = (int)m->cparam[c];
		for (int i = b; i < e; ++i) error += abs(w->x[m->cvars[i]] - value);
		return error;
	}
	case GHOST_KIND_PURE_OPTIMIZATION:
		return 0.; /* include/constraint.hpp:320-323 */
	default: {
		/* linear_equation.cpp:50-57 */
		w->cur_sum[c] = 0.0;
		for (int i = b; i < e; ++i) w->cur_sum[c] += m->ccoef[i] * w->x[m->cvars[i]];
		return linear_compute_error(m->kind[c], w->cur_sum[c], m->cparam[c]);
	}
	}
}

/* small multiset of touched count-table entries, to restate all_different.cpp:77-100 without
 * copying the whole table (the reference copies its std::map; only touched keys matter) */
typedef struct Touched { int32_t key[16]; int32_t copy[16]; int n; } Touched;
static int32_t* touched_slot(Touched* t, const int32_t* cnt, int32_t key)
{
	for (int i = 0; i < t->n; ++i) if (t->key[i] == key) return &t->copy[i];
	t->key[t->n] = key; t->copy[t->n] = cnt[key];
	return &t->copy[t->n++];
}

/* Constraint::simulate_delta -> delta_error -> optional_delta_error (src/constraint.cpp:91-96,
 * 67-89).  vars are GLOBAL ids (1 or 2 of them), vals their candidate values. */
/* offset of variable var inside AllDifferentOffset constraint c */
static int32_t scope_offset(const OracleModel* m, int c, int var)
{
	for (int p = m->coff[c]; p < m->coff[c + 1]; ++p) if (m->cvars[p] == var) return (int32_t)m->ccoef[p];
	return 0;
}

static double ctr_delta(const Walker* w, int c, int n, const int32_t* vars, const int32_t* vals)
{
	const OracleModel* m = w->m;
	switch (m->kind[c]) {
	case GHOST_KIND_ALL_DIFFERENT:
	case GHOST_KIND_ALL_DIFFERENT_OFFSET: {
		/* all_different.cpp:77-100; the extension kind on the shifted values */
		const int32_t* cnt = w->count + m->cnt_off[c];
		int32_t base = m->cnt_base[c];
		int32_t shift[2] = { 0, 0 };
		if (m->kind[c] == GHOST_KIND_ALL_DIFFERENT_OFFSET)
			for (int i = 0; i < n && i < 2; ++i) shift[i] = scope_offset(m, c, vars[i]);
		Touched t; t.n = 0;
		double diff = 0.0;
		for (int i = 0; i < n; ++i) {
			*touched_slot(&t, cnt, w->x[vars[i]] + shift[i] - base) -= 1;
			*touched_slot(&t, cnt, vals[i] + shift[i] - base) += 1;
		}
		if (m->kind[c] == GHOST_KIND_ALL_DIFFERENT_OFFSET) {
			/* the extension kind's delta is the true change of the error: every touched key once.  (The stock
			 * loop below visits the (old, new) key of each change, all_different.cpp:88-97, and would count a
			 * key twice when two changes land on the same shifted value — which a swap inside a stock
			 * AllDifferent never produces, its counts do not move at all.) */
			for (int k = 0; k < t.n; ++k) diff += binomial_with_2(t.copy[k]) - binomial_with_2(cnt[t.key[k]]);
			return diff;
		}
		for (int i = 0; i < n; ++i) {
			int32_t ko = w->x[vars[i]] + shift[i] - base, kn = vals[i] + shift[i] - base;
			int32_t co = *touched_slot(&t, cnt, ko), cn = *touched_slot(&t, cnt, kn);
			if (cnt[ko] != co) diff += binomial_with_2(co) - binomial_with_2(cnt[ko]);
			if (cnt[kn] != cn) diff += binomial_with_2(cn) - binomial_with_2(cnt[kn]);
		}
		return diff;
	}
	case GHOST_KIND_ALL_DIFFERENT_ABS_DIFF:
	case GHOST_KIND_ALL_DIFFERENT_LAG_DIFF: {
		/* extensions: remove the differences of every scope pair touching a changed variable,
		 * add the new ones, compare C(count,2) over the touched keys (each key once). */
		const int32_t* cnt = w->count + m->cnt_off[c];
		const int lag = pair_lag(m, c);
		const int32_t kind = m->kind[c], base = m->cnt_base[c];
		int b = m->coff[c], e = m->coff[c + 1];
		int32_t pairs[8]; int np = 0;
		for (int i = 0; i < n; ++i)
			for (int p = b; p < e; ++p)
				if (m->cvars[p] == vars[i]) {
					int cand[2] = { p - lag, p };
					for (int q = 0; q < 2; ++q) {
						int pp = cand[q];
						if (pp < b || pp + lag >= e) continue;
						int seen = 0;
						for (int z = 0; z < np; ++z) if (pairs[z] == pp) seen = 1;
						if (!seen) pairs[np++] = pp;
					}
				}
		Touched t; t.n = 0;
		for (int z = 0; z < np; ++z) {
			int va = m->cvars[pairs[z]], vb = m->cvars[pairs[z] + lag];
			int32_t oa = w->x[va], ob = w->x[vb], na = oa, nb = ob;
			for (int i = 0; i < n; ++i) { if (vars[i] == va) na = vals[i]; if (vars[i] == vb) nb = vals[i]; }
			*touched_slot(&t, cnt, pair_value(kind, oa, ob) - base) -= 1;
			*touched_slot(&t, cnt, pair_value(kind, na, nb) - base) += 1;
		}
		double diff = 0.0;
		for (int i = 0; i < t.n; ++i)
			diff += binomial_with_2(t.copy[i]) - binomial_with_2(cnt[t.key[i]]);
		return diff;
	}
	case GHOST_KIND_ALL_EQUAL: {
		/* all_equal.cpp:62-85: largest count before minus largest count after the changes */
		const int32_t* cnt = w->count + m->cnt_off[c];
		int32_t base = m->cnt_base[c];
		int32_t mx = 0, mx_bis = 0;
		for (int t = 0; t < m->cnt_size[c]; ++t) if (cnt[t] > mx) mx = cnt[t];
		Touched t; t.n = 0;
		for (int i = 0; i < n; ++i) {
			*touched_slot(&t, cnt, w->x[vars[i]] - base) -= 1;
			*touched_slot(&t, cnt, vals[i] - base) += 1;
		}
		for (int k = 0; k < m->cnt_size[c]; ++k) {
			int32_t v = cnt[k];
			for (int i = 0; i < t.n; ++i) if (t.key[i] == k) v = t.copy[i];
			if (v > mx_bis) mx_bis = v;
		}
		return (double)(mx - mx_bis);
	}
	case GHOST_KIND_FIX_VALUE: {
		/* fix_value.cpp:54-64 */
		double diff = 0.;
		int value = (int)m->cparam[c];
		for (int i = 0; i < n; ++i) diff += abs(vals[i] - value) - abs(w->x[vars[i]] - value);
		return diff;
	}
	case GHOST_KIND_PURE_OPTIMIZATION:
		return 0.; /* include/constraint.hpp:325-328 */
	default: {
		/* linear_equation.cpp:59-69; the coefficient of a variable is looked up by its position */
		double sum = w->cur_sum[c];
		for (int i = 0; i < n; ++i) {
			double coef = 0.0;
			for (int p = m->coff[c]; p < m->coff[c + 1]; ++p)
				if (m->cvars[p] == vars[i]) { coef = m->ccoef[p]; break; }
			sum += coef * (vals[i] - w->x[vars[i]]);
		}
		return linear_compute_error(m->kind[c], sum, m->cparam[c]) - w->cur_err[c];
	}
	}
}

/* Constraint::update -> conditional_update_data_structures; must run BEFORE x[var] is overwritten
 * (all_different.cpp:102-110, linear_equation.cpp:71-74; FixValue has none). */
static void ctr_update(Walker* w, int c, int var, int32_t new_value)
{
	const OracleModel* m = w->m;
	if (m->kind[c] == GHOST_KIND_ALL_DIFFERENT || m->kind[c] == GHOST_KIND_ALL_EQUAL || m->kind[c] == GHOST_KIND_ALL_DIFFERENT_OFFSET) { /* all_equal.cpp:87-95 alike */
		int32_t* cnt = w->count + m->cnt_off[c];
		const int32_t shift = m->kind[c] == GHOST_KIND_ALL_DIFFERENT_OFFSET ? scope_offset(m, c, var) : 0;
		cnt[w->x[var] + shift - m->cnt_base[c]] -= 1;
		cnt[new_value + shift - m->cnt_base[c]] += 1;
	} else if (is_linear(m->kind[c])) {
		for (int p = m->coff[c]; p < m->coff[c + 1]; ++p)
			if (m->cvars[p] == var) { w->cur_sum[c] += m->ccoef[p] * (new_value - w->x[var]); break; }
	}
	/* ALL_DIFFERENT_ABS_DIFF: the count table is rebuilt after the value is written, see
	 * abs_diff_rebuild() — its update needs both neighbours' final values. */
}

static int ctr_has_variable(const OracleModel* m, int c, int var)
{
	/* Constraint::has_variable, src/constraint.cpp:116-119 */
	for (int p = m->coff[c]; p < m->coff[c + 1]; ++p) if (m->cvars[p] == var) return 1;
	return 0;
}

/* Objective::cost (src/objective.cpp:53-64) for the linear objective; internal sign. */
static double objective_cost(const Walker* w)
{
	const OracleModel* m = w->m;
	double total = 0.0;
	for (int v = 0; v < m->V; ++v) total += w->x[v] * m->obj_coef[v];
	if (m->obj_kind == GHOST_OBJ_LINEAR_MAXIMIZE) total = -total;
	return total;
}
static int is_optimization(const OracleModel* m) { return m->obj_kind != GHOST_OBJ_NONE; }

/* SearchUnit::compute_constraints_errors, include/search_unit.hpp:434-447 */
static double compute_constraints_errors(Walker* w)
{
	double satisfaction_error = 0.0;
	for (int c = 0; c < w->m->C; ++c) {
		double error = ctr_error(w, c);
		w->cur_err[c] = error;
		satisfaction_error += error;
	}
	return satisfaction_error;
}

/* AdaptiveSearchErrorProjection::compute_variable_errors,
 * src/algorithms/adaptive_search_error_projection_algorithm.cpp:40-49 */
static void culprit_compute_variable_errors(Walker* w);
static void compute_variable_errors(Walker* w)
{
	const OracleModel* m = w->m;
	if (w->opt.strategy == GHOST_STRATEGY_CULPRIT) { culprit_compute_variable_errors(w); return; }
	for (int v = 0; v < m->V; ++v) {
		w->err_var[v] = 0.;
		for (int i = m->v2c_off[v]; i < m->v2c_off[v + 1]; ++i) w->err_var[v] += w->cur_err[m->v2c[i]];
	}
}

/* ---- Culprit Search error projection (GHOST_STRATEGY_CULPRIT) ------------------------------
 * src/algorithms/culprit_search_error_projection_algorithm.cpp.  Forward declaration: ctr_delta is above. */

/* Variable::get_partial_domain(3) (src/variable.cpp:69-136), element 0 and element 2 of what it returns: the domain
 * neighbours of the current value in the order the user gave the domain — except at the two ends, where the window
 * wraps around and the two elements are NOT previous / next (last value: {d[0], d[n-2], d[n-1]} -> d[0] and the value
 * itself; first value: {d[0], d[1], d[n-1]} -> the value itself and d[n-1]); a domain of exactly three values is
 * returned whole, whatever the current value. */
static void partial_domain_ends(const OracleModel* m, int var, int32_t current, int32_t* first, int32_t* third)
{
	const int32_t* d = m->dom_given + m->dom_off[var];
	const int n = m->dom_off[var + 1] - m->dom_off[var];
	int index = 0;
	while (index < n && d[index] != current) ++index;
	if (n <= 3) { *first = d[0]; *third = d[n - 1]; return; }
	const int start = index - 1;
	if (start >= 0) {
		if (index + 2 <= n) { *first = d[start]; *third = d[start + 2]; }
		else { *first = d[0]; *third = d[n - 1]; }          /* {d[0], d[n-2], d[n-1]} */
	} else { *first = d[0]; *third = d[n - 1]; }            /* {d[0], d[1], d[n-1]} */
}

/* CulpritSearchErrorProjection::compute_variable_errors_on_constraint (…:50-110).  The reference fills a vector of V
 * entries: the scope variables get the sum of two simulated deltas, all OTHER variables stay 0; then every entry x
 * becomes max - x (max over all V entries), and the entries are scaled to sum to the constraint's error.  So a variable
 * outside the scope ends with max / sum * error: not zero whenever max > 0. */
static void culprit_on_constraint(Walker* w, int c)
{
	const OracleModel* m = w->m;
	const int b = m->coff[c], e = m->coff[c + 1], outside = m->V - (e - b);
	for (int p = b; p < e; ++p) w->culprit[p] = 0.;
	w->culprit_out[c] = 0.;
	if (!(w->cur_err[c] > 0)) return;
	for (int p = b; p < e; ++p) {
		int32_t var = m->cvars[p];
		const int n = m->dom_off[var + 1] - m->dom_off[var];
		int32_t vars[1] = { var }, vals[1];
		if (n > 2) {
			int32_t first, third;
			partial_domain_ends(m, var, w->x[var], &first, &third);
			vals[0] = first; double a = ctr_delta(w, c, 1, vars, vals);
			vals[0] = third; w->culprit[p] = a + ctr_delta(w, c, 1, vars, vals);
		} else if (n == 2) {
			const int32_t* d = m->dom_given + m->dom_off[var];
			vals[0] = d[0] == w->x[var] ? d[1] : d[0];
			w->culprit[p] = ctr_delta(w, c, 1, vars, vals);
		} else {
			vals[0] = w->x[var];
			w->culprit[p] = ctr_delta(w, c, 1, vars, vals);
		}
	}
	double max = outside > 0 ? 0. : w->culprit[b];
	for (int p = b; p < e; ++p) if (w->culprit[p] > max) max = w->culprit[p];
	double sum = 0.;
	for (int p = b; p < e; ++p) { w->culprit[p] = -w->culprit[p] + max; sum += w->culprit[p]; }
	double out = outside > 0 ? -0. + max : 0.;
	/* std::accumulate walks the V entries in variable order; every term is an integer-valued double here, so the order of
	 * the additions does not matter */
	sum += out * outside;
	for (int p = b; p < e; ++p) w->culprit[p] = w->culprit[p] == 0 ? 0 : (w->culprit[p] / sum) * w->cur_err[c];
	w->culprit_out[c] = (outside > 0 && out != 0) ? (out / sum) * w->cur_err[c] : 0.;
}
/* Row c of the table combined into the projected errors.  mode +1: err[v] = row[v] + err[v] (std::plus, …:122-127, 145-150);
 * mode -1: err[v] = row[v] - err[v] — the "remove" step of update_variable_errors (…:136-141) hands std::transform the row
 * as FIRST and the error vector as SECOND range with std::minus, so it computes row - error, not error - row.  Restated as
 * written: after an update the projected errors are new_row + old_row - old_errors. */
static void culprit_apply(Walker* w, int c, int mode)
{
	const OracleModel* m = w->m;
	const int b = m->coff[c], e = m->coff[c + 1];
	memset(w->flag, 0, m->V);
	for (int p = b; p < e; ++p) {
		const int v = m->cvars[p];
		w->flag[v] = 1;
		w->err_var[v] = mode > 0 ? w->culprit[p] + w->err_var[v] : w->culprit[p] - w->err_var[v];
	}
	for (int v = 0; v < m->V; ++v) if (!w->flag[v]) w->err_var[v] = mode > 0 ? w->culprit_out[c] + w->err_var[v] : w->culprit_out[c] - w->err_var[v];
}
/* CulpritSearchErrorProjection::compute_variable_errors (…:112-129) */
static void culprit_compute_variable_errors(Walker* w)
{
	const OracleModel* m = w->m;
	for (int v = 0; v < m->V; ++v) w->err_var[v] = 0.;
	for (int c = 0; c < m->C; ++c) { culprit_on_constraint(w, c); culprit_apply(w, c, +1); }
}
/* CulpritSearchErrorProjection::update_variable_errors (…:131-151) */
static void culprit_update_variable_errors(Walker* w, int c)
{
	culprit_apply(w, c, -1);
	culprit_on_constraint(w, c);
	culprit_apply(w, c, +1);
}

/* ---- sampling ---------------------------------------------------------------------------- */

/* first `n` elements of a uniform shuffle of 0..V-1 == ordered sample without replacement;
 * drawn by rejection from one RNG event (replaces rng.shuffle, search_unit.hpp:206-208, 239-244) */
static void sample_distinct(Walker* w, int n, int32_t* out)
{
	const OracleModel* m = w->m;
	uint64_t event = w->rng.next_event++;
	memset(w->flag, 0, m->V);
	uint32_t t = 0;
	for (int got = 0; got < n; ++t) {
		uint32_t idx = below(rng_word(&w->rng, event, t), (uint32_t)m->V);
		if (w->flag[idx]) continue;
		w->flag[idx] = 1;
		out[got++] = (int32_t)idx;
	}
}

/* SearchUnit::monte_carlo_sampling, include/search_unit.hpp:201-212 */
static void monte_carlo_sampling(Walker* w, int nb_var)
{
	const OracleModel* m = w->m;
	if (nb_var < 0 || nb_var >= m->V) {
		uint64_t event = w->rng.next_event++;
		for (int v = 0; v < m->V; ++v) {
			uint32_t n = (uint32_t)(m->dom_off[v + 1] - m->dom_off[v]);
			w->x[v] = m->dom[m->dom_off[v] + below(rng_word(&w->rng, event, (uint32_t)v), n)];
		}
		return;
	}
	sample_distinct(w, nb_var, w->sample);
	uint64_t event = w->rng.next_event++;
	for (int i = 0; i < nb_var; ++i) {
		int v = w->sample[i];
		uint32_t n = (uint32_t)(m->dom_off[v + 1] - m->dom_off[v]);
		w->x[v] = m->dom[m->dom_off[v] + below(rng_word(&w->rng, event, (uint32_t)v), n)];
	}
}

static void try_swap(Walker* w, int i, int j)
{
	/* the four conditions of include/search_unit.hpp:224-231 / 247-254 */
	const OracleModel* m = w->m;
	if (i != j && w->x[i] != w->x[j] && in_domain(m, j, w->x[i]) && in_domain(m, i, w->x[j])) {
		int32_t t = w->x[i]; w->x[i] = w->x[j]; w->x[j] = t;
	}
}

/* SearchUnit::random_permutations, include/search_unit.hpp:215-257 */
static void random_permutations(Walker* w, int nb_var)
{
	const OracleModel* m = w->m;
	if (nb_var < 0) {
		uint64_t event = w->rng.next_event++;
		uint32_t blocks = (uint32_t)((m->V + 127) / 128);
		for (int i = 0; i < m->V - 1; ++i)
			for (int j = i + 1; j < m->V; ++j) {
				/* one coin per ordered pair: bit (j mod 128) of block (i, j/128) */
				uint32_t word = rng_word(&w->rng, event, ((uint32_t)i * blocks + ((uint32_t)j >> 7)) * 4u + (((uint32_t)j >> 5) & 3u));
				if ((word >> (j & 31)) & 1u) try_swap(w, i, j);
			}
		return;
	}
	if (nb_var > m->V) nb_var = m->V;
	sample_distinct(w, nb_var, w->scratch_a);
	sample_distinct(w, nb_var, w->scratch_b);
	for (int i = 0; i < nb_var; ++i) try_swap(w, w->scratch_a[i], w->scratch_b[i]);
}

/* SearchUnit::set_initial_configuration, include/search_unit.hpp:153-198.
 * Deviation D1 (documented in DESIGN.md): when a sample lowers data.best_sat_error (180-188) the
 * reference does not copy it to final_solution, so final_solution can lag best_sat_error after a
 * restart; here best_x is updated together with best_sat so that the returned assignment always
 * has the returned error. */
/* Deviation D1.  The reference lowers data.best_sat_error here (search_unit.hpp:180-188)
 * WITHOUT copying the sample into final_solution; initialize_data_structures (285-293) then finds
 * best_sat_error == current_sat_error and does not copy either, so after a restart the assignment a
 * search returns can be an older, worse one than the error it reports (at the first initialisation the
 * copy of 714-717 hides it).  The CUDA walkers and, by default, this restatement keep the best assignment
 * in step with the best error.  oracle_set_faithful_d1(1) switches the restatement to the reference's
 * behaviour so that a test can show the difference (tests/test_distribution.py). */
static int g_faithful_d1 = 0;
void oracle_set_faithful_d1(int on) { g_faithful_d1 = on; }

static void set_initial_configuration(Walker* w, int samplings)
{
	const OracleModel* m = w->m;
	double best_so_far = DBL_MAX, current;
	int32_t* best_values = w->scratch_b + m->V; /* second half of scratch_b */
	if (samplings < 1) samplings = 1;
	int loops = 0;
	do {
		if (m->permutation) random_permutations(w, -1); else monte_carlo_sampling(w, -1);
		current = 0.0;
		for (int c = 0; c < m->C; ++c) current += ctr_error(w, c);
		if (best_so_far > current) {
			best_so_far = current;
			if (w->best_sat > best_so_far) {
				w->best_sat = best_so_far;
				if (!g_faithful_d1) memcpy(w->best_x, w->x, sizeof(int32_t) * m->V); /* D1 */
			}
			memcpy(best_values, w->x, sizeof(int32_t) * m->V);
		}
	} while (++loops < samplings && current > 0.0);
	memcpy(w->x, best_values, sizeof(int32_t) * m->V);
}

/* SearchUnit::initialize_variable_values, include/search_unit.hpp:259-272 */
static void initialize_variable_values(Walker* w)
{
	if (w->opt.custom_starting_point || w->opt.resume_search) {
		if (w->opt.resume_search) w->opt.resume_search = 0;
		memcpy(w->x, w->m->start, sizeof(int32_t) * w->m->V);
	} else
		set_initial_configuration(w, w->opt.number_start_samplings);
}

/* SearchUnit::initialize_data_structures, include/search_unit.hpp:274-317 */
static void initialize_data_structures(Walker* w)
{
	const OracleModel* m = w->m;
	w->must_compute = 1;
	memset(w->tabu, 0, sizeof(int32_t) * m->V);
	for (int c = 0; c < m->C; ++c) w->cur_err[c] = 0.0;
	w->cur_sat = compute_constraints_errors(w);
	if (w->best_sat > w->cur_sat) {
		w->best_sat = w->cur_sat;
		memcpy(w->best_x, w->x, sizeof(int32_t) * m->V);
	}
	if (is_optimization(m)) {
		if (w->cur_sat == 0) {
			w->cur_opt = objective_cost(w);
			if (w->best_opt > w->cur_opt) {
				w->best_opt = w->cur_opt;
				memcpy(w->best_x, w->x, sizeof(int32_t) * m->V);
			}
		} else
			w->cur_opt = DBL_MAX;
	}
	compute_variable_errors(w);
}

/* SearchUnit::reset, include/search_unit.hpp:333-367 */
static void reset(Walker* w)
{
	++w->k.resets;
	if (w->opt.restart_threshold > 0 && (w->k.resets % w->opt.restart_threshold == 0)) {
		++w->k.restarts;
		initialize_variable_values(w);
	} else {
		if (w->m->permutation) random_permutations(w, w->opt.number_variables_to_reset);
		else monte_carlo_sampling(w, w->opt.number_variables_to_reset);
	}
	initialize_data_structures(w);
}

/* AdaptiveSearchVariableCandidatesHeuristic::compute_variable_candidates,
 * src/algorithms/adaptive_search_variable_candidates_heuristic.cpp:38-60 */
static void compute_variable_candidates(Walker* w)
{
	const OracleModel* m = w->m;
	double worst = -1;
	w->ncand = 0;
	if (w->opt.strategy == GHOST_STRATEGY_HILL_CLIMBING || w->opt.strategy == GHOST_STRATEGY_RANDOM_WALK) {
		/* AllFreeVariableCandidatesHeuristic::compute_variable_candidates (all_free_variable_candidates_heuristic.cpp:38-46):
		 * every variable that is not tabu, constrained or not */
		for (int v = 0; v < m->V; ++v) if (w->tabu[v] <= w->k.local_moves) w->cand[w->ncand++] = v;
		return;
	}
	for (int v = 0; v < m->V; ++v)
		if (worst <= w->err_var[v] && w->tabu[v] <= w->k.local_moves
		    && (m->v2c_off[v + 1] > m->v2c_off[v] || (is_optimization(m) && w->cur_sat == 0))) {
			if (worst < w->err_var[v]) {
				w->ncand = 0;
				w->cand[w->ncand++] = v;
				worst = w->err_var[v];
			} else if (worst == w->err_var[v])
				w->cand[w->ncand++] = v;
		}
}

/* The per-constraint deltas of ONE candidate, in the order the reference produces them
 * (include/search_unit.hpp:841-848 and 862-885).  Returns their number; `affected` receives the
 * constraint ids in the same order. */
static int candidate_deltas(const Walker* w, int v, int candidate, double* deltas, int32_t* affected)
{
	const OracleModel* m = w->m;
	int n = 0;
	if (!m->permutation) {
		if (m->v2c_off[v + 1] == m->v2c_off[v]) { deltas[0] = 0.0; if (affected) affected[0] = -1; return 1; }
		for (int i = m->v2c_off[v]; i < m->v2c_off[v + 1]; ++i) {
			int c = m->v2c[i];
			int32_t vars[1] = { v }, vals[1] = { candidate };
			if (affected) affected[n] = c;
			deltas[n++] = ctr_delta(w, c, 1, vars, vals);
		}
		return n;
	}
	int j = candidate;
	int32_t current_value = w->x[v], candidate_value = w->x[j];
	for (int i = m->v2c_off[v]; i < m->v2c_off[v + 1]; ++i) {
		int c = m->v2c[i];
		if (affected) affected[n] = c;
		if (ctr_has_variable(m, c, j)) {
			int32_t vars[2] = { v, j }, vals[2] = { candidate_value, current_value };
			deltas[n++] = ctr_delta(w, c, 2, vars, vals);
		} else {
			int32_t vars[1] = { v }, vals[1] = { candidate_value };
			deltas[n++] = ctr_delta(w, c, 1, vars, vals);
		}
	}
	for (int i = m->v2c_off[j]; i < m->v2c_off[j + 1]; ++i) {
		int c = m->v2c[i];
		if (ctr_has_variable(m, c, v)) continue; /* constraint_checked, search_unit.hpp:883 */
		int32_t vars[1] = { j }, vals[1] = { current_value };
		if (affected) affected[n] = c;
		deltas[n++] = ctr_delta(w, c, 1, vars, vals);
	}
	return n;
}

/* is `k` (value index in the sorted domain / partner variable) a candidate of variable v?
 * include/search_unit.hpp:831-833 and 855-860 */
static int is_candidate(const Walker* w, int v, int k, int32_t* key)
{
	const OracleModel* m = w->m;
	if (!m->permutation) {
		int32_t d = m->dom[m->dom_off[v] + k];
		*key = d;
		return d != w->x[v];
	}
	*key = k;
	return k != v && w->x[k] != w->x[v] && in_domain(m, v, w->x[k]) && in_domain(m, k, w->x[v]);
}
static int num_candidate_slots(const OracleModel* m, int v)
{
	return m->permutation ? m->V : m->dom_off[v + 1] - m->dom_off[v];
}

/* cumulated delta of every candidate + AdaptiveSearchValueHeuristic::select_value's running min
 * and tie list (src/algorithms/adaptive_search_value_heuristic.cpp:43-69).  Candidates are
 * visited in ascending key order like the reference's std::map. */
static int evaluate_candidates(Walker* w, int v, double* cand_sum /* may be NULL */, int32_t* ties,
                               double* min_conflict, int32_t* num_candidates)
{
	int slots = num_candidate_slots(w->m, v), nties = 0, ncandidates = 0;
	*min_conflict = DBL_MAX;
	for (int k = 0; k < slots; ++k) {
		int32_t key;
		if (!is_candidate(w, v, k, &key)) { if (cand_sum) cand_sum[k] = NAN; continue; }
		int n = candidate_deltas(w, v, key, w->deltas, NULL);
		double sum = 0.0; /* std::accumulate(..., 0.0) */
		for (int i = 0; i < n; ++i) sum += w->deltas[i];
		if (cand_sum) cand_sum[k] = sum;
		++ncandidates;
		if (*min_conflict > sum) { nties = 0; ties[nties++] = key; *min_conflict = sum; }
		else if (*min_conflict == sum) ties[nties++] = key;
	}
	if (num_candidates) *num_candidates = ncandidates;
	return nties;
}

/* SearchUnit::update_errors (include/search_unit.hpp:449-513) + the projection update
 * (adaptive_search_error_projection_algorithm.cpp:51-58) */
static void abs_diff_rebuild(Walker* w)
{
	/* extension kind: refresh the difference counts after values were written */
	const OracleModel* m = w->m;
	for (int c = 0; c < m->C; ++c)
		if (is_pair_diff(m->kind[c])) {
			int32_t* cnt = w->count + m->cnt_off[c];
			const int lag = pair_lag(m, c);
			memset(cnt, 0, sizeof(int32_t) * m->cnt_size[c]);
			for (int i = m->coff[c]; i + lag < m->coff[c + 1]; ++i)
				cnt[pair_value(m->kind[c], w->x[m->cvars[i]], w->x[m->cvars[i + lag]]) - m->cnt_base[c]]++;
		}
}

static void update_errors(Walker* w, int v, int new_value)
{
	const OracleModel* m = w->m;
	int32_t* affected = w->scratch_a;
	int n = candidate_deltas(w, v, new_value, w->deltas, affected);
	if (!m->permutation) {
		if (m->v2c_off[v + 1] == m->v2c_off[v]) return;
		for (int i = 0; i < n; ++i) {
			int c = affected[i];
			w->cur_err[c] += w->deltas[i];
			/* the projection is updated BEFORE the constraint's cache and the variable's value (search_unit.hpp:457-465):
			 * Culprit's simulated deltas therefore see the new error with the old sums / counts / assignment */
			if (w->opt.strategy == GHOST_STRATEGY_CULPRIT) culprit_update_variable_errors(w, c);
			else for (int p = m->coff[c]; p < m->coff[c + 1]; ++p) w->err_var[m->cvars[p]] += w->deltas[i];
			ctr_update(w, c, v, new_value);
		}
		return;
	}
	int j = new_value;
	int32_t current_value = w->x[v], next_value = w->x[j];
	for (int i = 0; i < n; ++i) {
		int c = affected[i];
		w->cur_err[c] += w->deltas[i];
		if (w->opt.strategy == GHOST_STRATEGY_CULPRIT) culprit_update_variable_errors(w, c);
		else for (int p = m->coff[c]; p < m->coff[c + 1]; ++p) w->err_var[m->cvars[p]] += w->deltas[i];
		if (ctr_has_variable(m, c, v)) {
			ctr_update(w, c, v, next_value);
			if (ctr_has_variable(m, c, j)) ctr_update(w, c, j, current_value);
		} else
			ctr_update(w, c, j, current_value);
	}
}

/* SearchUnit::local_move, include/search_unit.hpp:516-541 */
static void local_move(Walker* w, int v, int new_value, double min_conflict)
{
	++w->k.local_moves;
	w->cur_sat += min_conflict;
	w->tabu[v] = w->opt.tabu_time_selected + (int32_t)w->k.local_moves;
	w->must_compute = 1;
	update_errors(w, v, new_value);
	if (w->m->permutation) {
		int32_t current_value = w->x[v], next_value = w->x[new_value];
		w->x[v] = next_value;
		w->x[new_value] = current_value;
	} else
		w->x[v] = new_value;
	abs_diff_rebuild(w);
}

/* candidate objective cost with the move applied, include/search_unit.hpp:1063-1094 */
static double candidate_opt_cost(Walker* w, int v, int new_value)
{
	double cost;
	if (w->m->permutation) {
		int32_t a = w->x[v], b = w->x[new_value];
		w->x[v] = b; w->x[new_value] = a;
		cost = objective_cost(w);
		w->x[v] = a; w->x[new_value] = b;
	} else {
		int32_t backup = w->x[v];
		w->x[v] = new_value;
		cost = objective_cost(w);
		w->x[v] = backup;
	}
	return cost;
}

/* Objective::expert_heuristic_value, src/objective.cpp:69-109 (non-permutation COP tie-break) */
static int heuristic_value(Walker* w, int v, const int32_t* possible, int n, uint32_t word)
{
	double min_cost = DBL_MAX;
	int32_t backup = w->x[v];
	int32_t* best = w->sample; int nbest = 0;
	for (int i = 0; i < n; ++i) {
		w->x[v] = possible[i];
		double simulated = objective_cost(w);
		if (min_cost > simulated) { min_cost = simulated; nbest = 0; best[nbest++] = possible[i]; }
		else if (min_cost == simulated) best[nbest++] = possible[i];
	}
	w->x[v] = backup;
	if (nbest > 0) return best[below(word, (uint32_t)nbest)];
	if (n > 0) return possible[below(word, (uint32_t)n)];
	return backup;
}

static Walker* walker_create(const OracleModel* m, const GhostSearchOpts* opts, uint64_t walker_id)
{
	Walker* w = (Walker*)calloc(1, sizeof *w);
	int V = m->V > 0 ? m->V : 1, C = m->C > 0 ? m->C : 1;
	w->m = m;
	if (opts) w->opt = *opts;
	w->rng.seed = opts ? opts->seed : 0; w->rng.walker = walker_id; w->rng.next_event = 0;
	w->x = (int32_t*)calloc(V, sizeof(int32_t));
	w->cur_err = (double*)calloc(C, sizeof(double));
	w->cur_sum = (double*)calloc(C, sizeof(double));
	w->count = (int32_t*)calloc(m->cnt_total > 0 ? m->cnt_total : 1, sizeof(int32_t));
	w->tabu = (int32_t*)calloc(V, sizeof(int32_t));
	w->err_var = (double*)calloc(V, sizeof(double));
	w->best_x = (int32_t*)calloc(V, sizeof(int32_t));
	w->cand = (int32_t*)calloc(V, sizeof(int32_t));
	int big = V > C ? V : C; if (m->max_candidates > big) big = m->max_candidates;
	w->scratch_a = (int32_t*)calloc(2 * big + 2, sizeof(int32_t));
	w->scratch_b = (int32_t*)calloc(2 * big + 2, sizeof(int32_t));
	w->sample = (int32_t*)calloc(big + 2, sizeof(int32_t));
	w->flag = (uint8_t*)calloc(V, 1);
	w->deltas = (double*)calloc(2 * C + 2, sizeof(double));
	w->wvec = (double*)calloc(V, sizeof(double));
	w->culprit = (double*)calloc(m->coff[m->C] > 0 ? m->coff[m->C] : 1, sizeof(double));
	w->culprit_out = (double*)calloc(C, sizeof(double));
	w->cand_sum = (double*)calloc(big + 2, sizeof(double));
	return w;
}
static void walker_destroy(Walker* w)
{
	free(w->x); free(w->cur_err); free(w->cur_sum); free(w->count); free(w->tabu); free(w->err_var);
	free(w->wvec); free(w->cand_sum); free(w->culprit); free(w->culprit_out);
	free(w->best_x); free(w->cand); free(w->scratch_a); free(w->scratch_b); free(w->sample);
	free(w->flag); free(w->deltas); free(w);
}

int oracle_eval_moves(const OracleModel* m, const int32_t* x, int32_t v, double* ctr_error_out,
                      double* var_error_out, double* cand_delta, int32_t* ties, int32_t* num_ties,
                      double* min_conflict)
{
	Walker* w = walker_create(m, NULL, 0);
	memcpy(w->x, x, sizeof(int32_t) * m->V);
	compute_constraints_errors(w);
	compute_variable_errors(w);
	if (ctr_error_out) memcpy(ctr_error_out, w->cur_err, sizeof(double) * m->C);
	if (var_error_out) memcpy(var_error_out, w->err_var, sizeof(double) * m->V);
	if (v >= 0) {
		int32_t* t = ties ? ties : w->scratch_b;
		double mc;
		int n = evaluate_candidates(w, v, cand_delta, t, &mc, NULL);
		if (num_ties) *num_ties = n;
		if (min_conflict) *min_conflict = mc;
	}
	walker_destroy(w);
	return 0;
}

int oracle_check(const OracleModel* m, const int32_t* x, double* sat_error, double* objective)
{
	Walker* w = walker_create(m, NULL, 0);
	memcpy(w->x, x, sizeof(int32_t) * m->V);
	double e = compute_constraints_errors(w);
	if (sat_error) *sat_error = e;
	if (objective) {
		double c = objective_cost(w);
		*objective = m->obj_kind == GHOST_OBJ_LINEAR_MAXIMIZE ? -c : c;
	}
	walker_destroy(w);
	return 0;
}

/* include/solver.hpp:451-473 */
void oracle_resolve_opts(int32_t V, GhostSearchOpts* o)
{
	if (o->tabu_time_local_min < 0) {
		int a = 5 < V - 1 ? 5 : V - 1;
		int b = (int)ceil(V / 5); /* integer division first, as in the reference */
		o->tabu_time_local_min = (a > b ? a : b) + 1;
	}
	if (o->tabu_time_selected < 0) o->tabu_time_selected = 0;
	if (o->percent_chance_force_trying_on_plateau < 0 || o->percent_chance_force_trying_on_plateau > 100)
		o->percent_chance_force_trying_on_plateau = 10;
	if (o->reset_threshold < 0) o->reset_threshold = o->tabu_time_local_min;
	if (o->restart_threshold < 0) o->restart_threshold = V;
	if (o->number_variables_to_reset < 0) {
		int n = (int)ceil(V * 0.1);
		o->number_variables_to_reset = n > 2 ? n : 2;
	}
	if (o->number_start_samplings < 0) o->number_start_samplings = 10;
}

/* SearchUnit::local_search, include/search_unit.hpp:685-1201 */
/* AntidoteSearchVariableCandidatesHeuristic::compute_variable_candidates
 * (src/algorithms/antidote_search_variable_candidates_heuristic.cpp:39-48): the error vector with tabu variables zeroed */
static void antidote_candidates(Walker* w)
{
	const OracleModel* m = w->m;
	for (int v = 0; v < m->V; ++v) w->wvec[v] = w->tabu[v] > w->k.local_moves ? 0.0 : w->err_var[v];
	w->nvec = m->V;
}
/* rng.variate<int, std::discrete_distribution>(weights) restated on one Philox word: index i with probability
 * weights[i] / total — the first i whose running sum exceeds floor(word * total / 2^32).  The weights are
 * integer-valued (errors of integer models); a vector of zeros gives index 0, which is what libstdc++'s
 * discrete_distribution does with NaN probabilities (std::lower_bound finds nothing smaller). */
static int sample_weighted(const double* weights, int n, uint32_t word)
{
	double total = 0.0;
	for (int i = 0; i < n; ++i) total += weights[i];
	if (!(total > 0.0)) return 0;
	uint32_t pick = below(word, (uint32_t)total);
	double run = 0.0;
	for (int i = 0; i < n; ++i) { run += weights[i]; if (run > (double)pick) return i; }
	return n - 1;
}

/* The choices of one iteration taken from the caller instead of the Philox stream (decision replay
 * against the reference's own GHOST_TRACE output, tests/test_decision_replay.py): the variable picked
 * among the candidates (search_unit.hpp:806), the value / swap partner picked among the ties
 * (adaptive_search_value_heuristic.cpp:80, objective.cpp:100-117) and the plateau coin (547). */
typedef struct Forced { int32_t variable, value, force_on_plateau; } Forced;

/* ONE pass of the while body of SearchUnit::local_search (include/search_unit.hpp:731-1189).
 * Returns 0 after an evaluated iteration, 1 after a reset iteration, and with forced choices
 * -1 if the variable is not in the candidate list, -2 if the value is not among the ties (for a
 * non-permutation COP: not among the ties of lowest objective cost).  `ties` receives the tie list
 * (ascending), *nties its length. */
static int walker_iteration(Walker* w, const Forced* forced, OracleTraceRow* row, int32_t* ties, int32_t* nties_out)
{
	const OracleModel* m = w->m;
	const int optimization = is_optimization(m);
	++w->k.search_iterations;
	memset(row, 0, sizeof *row);
	row->iteration = w->k.search_iterations;
	if (nties_out) *nties_out = 0;

	const int antidote = w->opt.strategy == GHOST_STRATEGY_ANTIDOTE;
	const int random_walk = w->opt.strategy == GHOST_STRATEGY_RANDOM_WALK;
	if (w->must_compute) { if (antidote) antidote_candidates(w); else compute_variable_candidates(w); } /* 749-750 */

	int ntabu = 0; /* 761-763 */
	for (int v = 0; v < m->V; ++v) if (w->tabu[v] > w->k.local_moves) ++ntabu;
	if (ntabu >= w->opt.reset_threshold || (antidote ? w->nvec == 0 : w->ncand == 0)) {
		reset(w); /* 769-770 */
		row->variable = -1; row->branch = 0; row->current_sat_error = w->cur_sat;
		return 1;
	}

	/* one RNG event per evaluated iteration: word 0 picks the variable (806), word 1 the
	 * value among ties (adaptive_search_value_heuristic.cpp:80 / objective.cpp:100-101,
	 * 111-117), word 2 the plateau coin (547) */
	uint32_t rnd[4];
	rng_block(&w->rng, w->rng.next_event++, 0, rnd);

	int v;
	if (antidote) {
		/* AntidoteSearchVariableHeuristic::select_variable (antidote_search_variable_heuristic.cpp:39-43): the INDEX
		 * drawn from the weight vector is the variable */
		v = sample_weighted(w->wvec, w->nvec, rnd[0]);
		if (forced) {
			double total = 0.0;
			for (int i = 0; i < w->nvec; ++i) total += w->wvec[i];
			if (forced->variable < 0 || forced->variable >= w->nvec
			    || (total > 0.0 ? !(w->wvec[forced->variable] > 0.0) : forced->variable != 0)) { --w->k.search_iterations; return -1; }
			v = forced->variable;
		}
		/* search_unit.hpp:825-828: std::find looks for the variable's index AS A VALUE of the vector — with Antidote's
		 * weights that erases the first entry whose weight happens to equal the index, if there is one */
		for (int i = 0; i < w->nvec; ++i)
			if (w->wvec[i] == (double)v) { for (int t = i; t + 1 < w->nvec; ++t) w->wvec[t] = w->wvec[t + 1]; --w->nvec; break; }
	} else {
		int pick = (int)below(rnd[0], (uint32_t)w->ncand);
		if (forced) {
			pick = -1;
			for (int i = 0; i < w->ncand; ++i) if (w->cand[i] == forced->variable) pick = i;
			if (pick < 0) { --w->k.search_iterations; return -1; }
		}
		v = w->cand[pick];
		/* erase, keeping the ascending order of the vector (826-828) */
		for (int i = pick; i + 1 < w->ncand; ++i) w->cand[i] = w->cand[i + 1];
		--w->ncand;
	}
	const int cand_empty = antidote ? w->nvec == 0 : w->ncand == 0;

	double min_conflict;
	int32_t ncandidates = 0;
	int nties = evaluate_candidates(w, v, (antidote || random_walk) ? w->cand_sum : NULL, ties, &min_conflict, &ncandidates); /* 831-891 */
	w->k.candidates_evaluated += ncandidates;
	if (nties_out) *nties_out = nties;

	int new_value;
	if (antidote) {
		/* AntidoteSearchValueHeuristic::select_value (antidote_search_value_heuristic.cpp:50-74): candidates in key order,
		 * weight max(0, -cumulated delta); uniform over all candidates when none improves; min_conflict becomes the
		 * CHOSEN candidate's cumulated delta.  (No candidate at all is undefined behaviour there — max_element of an
		 * empty range; here: the variable keeps its value and the iteration counts as a local minimum.) */
		const int slots = num_candidate_slots(m, v);
		int32_t* keys = w->scratch_a; double* wts = w->deltas; /* deltas is free again; it holds 2C+2 doubles: use cand-sized scratch below */
		(void)wts;
		int nc = 0; double maxw = 0.0, total = 0.0;
		for (int k = 0; k < slots; ++k) {
			if (isnan(w->cand_sum[k])) continue;
			int32_t key; is_candidate(w, v, k, &key);
			keys[nc] = key;
			double wt = w->cand_sum[k] >= 0 ? 0.0 : -w->cand_sum[k];
			w->cand_sum[nc] = w->cand_sum[k];       /* compact in place (nc <= k) */
			if (wt > maxw) maxw = wt;
			total += wt;
			++nc;
		}
		if (nc == 0) { new_value = v; min_conflict = DBL_MAX; }
		else {
			int idx;
			if (maxw == 0.0) idx = (int)below(rnd[1], (uint32_t)nc);
			else {
				uint32_t pick = below(rnd[1], (uint32_t)total);
				double run = 0.0; idx = nc - 1;
				for (int i = 0; i < nc; ++i) { double wt = w->cand_sum[i] >= 0 ? 0.0 : -w->cand_sum[i]; run += wt; if (run > (double)pick) { idx = i; break; } }
			}
			if (forced) {
				idx = -1;
				for (int i = 0; i < nc; ++i) if (keys[i] == forced->value) idx = i;
				if (idx < 0) return -2;
				if (maxw > 0.0 && !(w->cand_sum[idx] < 0)) return -2;   /* a weight-0 candidate cannot be drawn */
			}
			new_value = keys[idx];
			min_conflict = w->cand_sum[idx];
		}
	}
	else if (random_walk) {
		/* RandomWalkValueHeuristic::select_value (random_walk_value_heuristic.cpp:47-57): rng.pick over the map of candidates
		 * (ascending key), min_conflict = the picked candidate's cumulated delta.  (An empty map is undefined behaviour there;
		 * here: the variable keeps its value and nothing moves.) */
		const int slots = num_candidate_slots(m, v);
		int nc = 0, idx = -1;
		for (int k = 0; k < slots; ++k) if (!isnan(w->cand_sum[k])) ++nc;
		new_value = v; min_conflict = DBL_MAX;
		if (nc > 0) {
			int want = (int)below(rnd[1], (uint32_t)nc);
			for (int k = 0, seen = 0; k < slots; ++k) {
				if (isnan(w->cand_sum[k])) continue;
				int32_t key; is_candidate(w, v, k, &key);
				if (forced ? key == forced->value : seen == want) { idx = k; new_value = key; break; }
				++seen;
			}
			if (idx < 0) return -2;
			min_conflict = w->cand_sum[idx];
		}
	}
	else if (nties == 0) new_value = v; /* adaptive_search_value_heuristic.cpp:68-69 */
	else if (optimization && !m->permutation) new_value = heuristic_value(w, v, ties, nties, rnd[1]);
	else new_value = ties[below(rnd[1], (uint32_t)nties)];
	if (!antidote && !random_walk && forced && nties > 0) {
		int found = 0;
		for (int i = 0; i < nties; ++i) if (ties[i] == forced->value) found = 1;
		if (found && optimization && !m->permutation) {
			/* Objective::expert_heuristic_value keeps the tied values of lowest cost (objective.cpp:69-109) */
			const int32_t backup = w->x[v];
			double best = DBL_MAX, mine;
			for (int i = 0; i < nties; ++i) { w->x[v] = ties[i]; double c = objective_cost(w); if (c < best) best = c; }
			w->x[v] = forced->value; mine = objective_cost(w);
			w->x[v] = backup;
			if (mine != best) found = 0;
		}
		if (!found) return -2;
		new_value = forced->value;
	}

	row->variable = v; row->new_value = new_value; row->min_conflict = min_conflict;
	row->num_candidates = ncandidates;

	int plateau = 0;
	if (random_walk) { /* search_unit.hpp:1007-1036: the move is made whatever its delta, then the best-so-far bookkeeping */
		if (min_conflict != DBL_MAX) {
			local_move(w, v, new_value, min_conflict);
			if (optimization) w->cur_opt = objective_cost(w);
			row->branch = 1;
		} else { w->must_compute = 0; row->branch = 6; }
	} else
	if (min_conflict < 0.0) { /* 1042-1050 */
		local_move(w, v, new_value, min_conflict);
		if (optimization) w->cur_opt = objective_cost(w);
		row->branch = 1;
	} else if (min_conflict == 0.0) {
		if (optimization && w->opt.enable_optimization_guidance) { /* 1061-1135 */
			double cand_opt = candidate_opt_cost(w, v, new_value);
			if (w->cur_opt > cand_opt) {
				local_move(w, v, new_value, min_conflict);
				w->cur_opt = cand_opt;
				row->branch = 2;
			} else if (w->cur_opt == cand_opt) plateau = 1;
			else {
				/* local_minimum_management, 565-582 */
				w->must_compute = 0;
				if (cand_empty) { w->tabu[v] = w->opt.tabu_time_local_min + (int32_t)w->k.local_moves; ++w->k.local_minimum; row->branch = 5; }
				else row->branch = 6;
			}
		} else plateau = 1; /* 1136-1145 */
	} else { /* 1147-1163 */
		w->must_compute = 0;
		if (cand_empty) { w->tabu[v] = w->opt.tabu_time_local_min + (int32_t)w->k.local_moves; ++w->k.local_minimum; row->branch = 5; }
		else row->branch = 6;
	}
	if (plateau) { /* plateau_management, 545-561 */
		int force = (int)below(rnd[2], 100u) + 1 <= w->opt.percent_chance_force_trying_on_plateau;
		if (forced) force = forced->force_on_plateau;
		if (force) {
			w->tabu[v] = w->opt.tabu_time_local_min + (int32_t)w->k.local_moves;
			w->must_compute = 1;
			++w->k.plateau_force_trying_another_variable;
			row->branch = 4;
		} else {
			local_move(w, v, new_value, 0);
			++w->k.plateau_moves;
			row->branch = 3;
		}
	}

	if (w->best_sat > w->cur_sat) { /* 1166-1176 */
		w->best_sat = w->cur_sat;
		memcpy(w->best_x, w->x, sizeof(int32_t) * m->V);
	} else if (optimization && w->cur_sat == 0.0 && w->best_opt > w->cur_opt) { /* 1177-1188 */
		w->best_opt = w->cur_opt;
		memcpy(w->best_x, w->x, sizeof(int32_t) * m->V);
	}
	row->current_sat_error = w->cur_sat;
	return 0;
}

int oracle_search(const OracleModel* m, const GhostSearchOpts* opts, uint64_t walker_id,
                  int64_t max_iterations, int32_t* best_x, double* best_sat, double* best_opt,
                  GhostCounters* counters, int32_t* final_x,
                  OracleTraceRow* trace, int64_t trace_cap, int64_t* trace_len)
{
	Walker* w = walker_create(m, opts, walker_id);
	int64_t ntrace = 0;
	int optimization = is_optimization(m);
	int32_t* ties = w->scratch_b; /* scratch_b's first half; set_initial_configuration uses the second */

	w->best_sat = DBL_MAX; /* 708-709 */
	w->best_opt = DBL_MAX;
	memcpy(w->x, m->start, sizeof(int32_t) * m->V);
	initialize_variable_values(w);  /* 711 */
	initialize_data_structures(w);  /* 712 */
	memcpy(w->best_x, w->x, sizeof(int32_t) * m->V); /* 714-717 */

	while (w->k.search_iterations < max_iterations
	       && (w->best_sat > 0.0 || (w->best_sat == 0.0 && optimization))) { /* 728-730 */
		OracleTraceRow row;
		walker_iteration(w, NULL, &row, ties, NULL);
		if (trace && ntrace < trace_cap) trace[ntrace] = row;
		++ntrace;
	}

	if (best_x) memcpy(best_x, w->best_x, sizeof(int32_t) * m->V);
	if (final_x) memcpy(final_x, w->x, sizeof(int32_t) * m->V);
	if (best_sat) *best_sat = w->best_sat;
	if (best_opt) *best_opt = w->best_opt;
	if (counters) *counters = w->k;
	if (trace_len) *trace_len = ntrace;
	walker_destroy(w);
	return 0;
}

/* ---- decision replay: the same iteration, its random choices supplied by the caller ------------
 * The reference compiled with GHOST_TRACE prints, per iteration, what it saw and what it chose
 * (include/search_unit.hpp:741-746, 773-779, 808-819, 893-1003, 1042-1164).  A stepper holds one
 * walker; the test feeds it the reference's choices one iteration at a time and compares what the
 * restatement sees (candidate variables, tabu list, current error, ties, branch) with what the
 * reference printed.  After a reset of the reference (whose random perturbation cannot be
 * reproduced: it draws from std::mt19937) the walker is re-synchronised on the printed assignment. */
struct OracleStepper { Walker* w; };

OracleStepper* oracle_stepper_create(const OracleModel* m, const GhostSearchOpts* resolved_opts, const int32_t* x0)
{
	OracleStepper* s = (OracleStepper*)calloc(1, sizeof *s);
	Walker* w = walker_create(m, resolved_opts, 0);
	s->w = w;
	w->best_sat = DBL_MAX; w->best_opt = DBL_MAX; /* 708-709 */
	memcpy(w->x, x0, sizeof(int32_t) * m->V);     /* the reference's assignment after its own initial sampling */
	initialize_data_structures(w);                /* 712 */
	memcpy(w->best_x, w->x, sizeof(int32_t) * m->V);
	return s;
}
void oracle_stepper_destroy(OracleStepper* s) { if (s) { walker_destroy(s->w); free(s); } }

/* the reference went through reset(): same end state, initialize_data_structures() on its assignment (366) */
void oracle_stepper_sync(OracleStepper* s, const int32_t* x)
{
	memcpy(s->w->x, x, sizeof(int32_t) * s->w->m->V);
	initialize_data_structures(s->w);
}

/* What the next iteration starts from: the candidate variables as they are AFTER the recomputation the
 * iteration would do (749-750), the tabu list (entries with tabu_until > local_moves: [variable,
 * until] pairs), counters, and whether the iteration would reset (761-771). */
int oracle_stepper_peek(OracleStepper* s, int32_t* x, int32_t* candidates, int32_t* num_candidates,
                        int32_t* tabu_pairs, int32_t* num_tabu, int64_t* local_moves, double* current_sat_error,
                        int32_t* would_reset)
{
	Walker* w = s->w;
	const OracleModel* m = w->m;
	const int antidote = w->opt.strategy == GHOST_STRATEGY_ANTIDOTE;
	if (w->must_compute) { if (antidote) antidote_candidates(w); else compute_variable_candidates(w); w->must_compute = 0; } /* idempotent: the iteration finds the same list */
	if (x) memcpy(x, w->x, sizeof(int32_t) * m->V);
	if (candidates) memcpy(candidates, w->cand, sizeof(int32_t) * w->ncand);
	if (num_candidates) *num_candidates = w->ncand;
	int nt = 0;
	for (int v = 0; v < m->V; ++v) if (w->tabu[v] > w->k.local_moves) {
		if (tabu_pairs) { tabu_pairs[2 * nt] = v; tabu_pairs[2 * nt + 1] = w->tabu[v]; }
		++nt;
	}
	if (num_tabu) *num_tabu = nt;
	if (local_moves) *local_moves = w->k.local_moves;
	if (current_sat_error) *current_sat_error = w->cur_sat;
	if (would_reset) *would_reset = (nt >= w->opt.reset_threshold || (antidote ? w->nvec == 0 : w->ncand == 0)) ? 1 : 0;
	return 0;
}

/* the projected errors of the walker (SearchUnitData::error_variables) and the constraint errors */
void oracle_stepper_errors(OracleStepper* s, double* var_error, double* ctr_error)
{
	if (var_error) memcpy(var_error, s->w->err_var, sizeof(double) * s->w->m->V);
	if (ctr_error) memcpy(ctr_error, s->w->cur_err, sizeof(double) * s->w->m->C);
}

/* One iteration with the given choices; see walker_iteration for the return value. */
int oracle_stepper_step(OracleStepper* s, int32_t variable, int32_t value, int32_t force_on_plateau,
                        OracleTraceRow* row, int32_t* ties, int32_t* num_ties, GhostCounters* counters)
{
	Forced f; f.variable = variable; f.value = value; f.force_on_plateau = force_on_plateau;
	OracleTraceRow local;
	int rc = walker_iteration(s->w, &f, row ? row : &local, ties ? ties : s->w->scratch_b, num_ties);
	if (counters) *counters = s->w->k;
	return rc;
}

/* ==========================================================================================
 * Solver::complete_search (include/solver.hpp:932-987) restated with its helpers: prefiltering
 * (161-181), ac3_filtering (186-243), has_support (249-295) and the recursive enumeration
 * (300-347).  Domains are per-variable value lists in descriptor order; a "copy" of the domains
 * (the reference passes them by value) is a copy of the size array + the value array.
 * ========================================================================================== */
typedef struct CsDoms { int32_t* val; int32_t* size; } CsDoms; /* val: flat, capacity = original sizes */

typedef struct CsCtx {
	Walker* w;
	const OracleModel* m;
	int32_t ndom;
	int64_t cap, count;
	int32_t* out_x; double* out_cost;
	int truncated;
} CsCtx;

static CsDoms cs_copy(const CsCtx* c, const CsDoms* d)
{
	CsDoms r;
	r.val = (int32_t*)malloc(sizeof(int32_t) * (c->ndom > 0 ? c->ndom : 1));
	r.size = (int32_t*)malloc(sizeof(int32_t) * c->m->V);
	memcpy(r.val, d->val, sizeof(int32_t) * c->ndom);
	memcpy(r.size, d->size, sizeof(int32_t) * c->m->V);
	return r;
}
static void cs_free(CsDoms* d) { free(d->val); free(d->size); }
static void cs_erase(const CsCtx* c, CsDoms* d, int var, int32_t value)
{
	int32_t* v = d->val + c->m->dom_off[var];
	int n = d->size[var], i = 0;
	while (i < n && v[i] != value) ++i;
	if (i == n) return;
	for (; i + 1 < n; ++i) v[i] = v[i + 1]; /* std::vector::erase keeps the order */
	d->size[var] = n - 1;
}

/* solver.hpp:161-181 */
static void cs_prefiltering(CsCtx* c, CsDoms* d)
{
	const OracleModel* m = c->m;
	for (int ctr = 0; ctr < m->C; ++ctr) {
		if (m->coff[ctr + 1] - m->coff[ctr] != 1) continue;
		int index = m->cvars[m->coff[ctr]];
		int n = d->size[index], nrem = 0;
		int32_t* rem = (int32_t*)malloc(sizeof(int32_t) * (n > 0 ? n : 1));
		for (int i = 0; i < n; ++i) {
			int32_t value = d->val[m->dom_off[index] + i];
			c->w->x[index] = value;
			if (ctr_error(c->w, ctr) > 0.0) rem[nrem++] = value;
		}
		for (int i = 0; i < nrem; ++i) cs_erase(c, d, index, rem[i]);
		free(rem);
	}
}

/* solver.hpp:249-295 */
static int cs_has_support(CsCtx* c, int ctr, int variable_id, int index_v, const CsDoms* d)
{
	const OracleModel* m = c->m;
	int b = m->coff[ctr], e = m->coff[ctr + 1];
	int* scope = (int*)malloc(sizeof(int) * (e - b + 1));
	int ns = 0;
	for (int i = b; i < e; ++i) if (m->cvars[i] > index_v && m->cvars[i] != variable_id) scope[ns++] = m->cvars[i];
	if (ns == 0) { free(scope); return ctr_error(c->w, ctr) == 0.0; }
	int* indexes = (int*)calloc(ns + 1, sizeof(int));
	int fake_index = ns, found = 0;
	while (indexes[fake_index] == 0) {
		for (int i = 0; i < fake_index; ++i) c->w->x[scope[i]] = d->val[m->dom_off[scope[i]] + indexes[i]];
		if (ctr_error(c->w, ctr) == 0.0) { found = 1; break; }
		int changed, index = 0;
		do {
			changed = 0;
			++indexes[index];
			if (index < fake_index && indexes[index] >= d->size[scope[index]]) { indexes[index] = 0; changed = 1; ++index; }
		} while (changed);
	}
	free(indexes); free(scope);
	return found;
}

/* solver.hpp:186-243; returns a filtered copy of the domains */
static CsDoms cs_ac3(CsCtx* c, int index_v, const CsDoms* in)
{
	const OracleModel* m = c->m;
	CsDoms d = cs_copy(c, in);
	int cap = 64, head = 0, tail = 0;
	int* qc = (int*)malloc(sizeof(int) * cap); int* qv = (int*)malloc(sizeof(int) * cap);
#define CS_PUSH(cc, vv) do { if (tail == cap) { cap *= 2; qc = (int*)realloc(qc, sizeof(int) * cap); qv = (int*)realloc(qv, sizeof(int) * cap); } qc[tail] = (cc); qv[tail] = (vv); ++tail; } while (0)
	for (int r = m->v2c_off[index_v]; r < m->v2c_off[index_v + 1]; ++r) {
		int ctr = m->v2c[r];
		for (int i = m->coff[ctr]; i < m->coff[ctr + 1]; ++i) if (m->cvars[i] > index_v) CS_PUSH(ctr, m->cvars[i]);
	}
	int32_t* rem = (int32_t*)malloc(sizeof(int32_t) * (c->ndom > 0 ? c->ndom : 1));
	while (head < tail) {
		int ctr = qc[head], variable_id = qv[head];
		++head;
		int nrem = 0, n = d.size[variable_id];
		for (int k = 0; k < n; ++k) {
			int32_t value = d.val[m->dom_off[variable_id] + k];
			c->w->x[variable_id] = value;
			if (!cs_has_support(c, ctr, variable_id, index_v, &d)) {
				rem[nrem++] = value;
				for (int r = m->v2c_off[variable_id]; r < m->v2c_off[variable_id + 1]; ++r) {
					int c_id = m->v2c[r];
					if (c_id == ctr) continue;
					for (int i = m->coff[c_id]; i < m->coff[c_id + 1]; ++i) {
						int v_id = m->cvars[i];
						if (v_id <= index_v || v_id == variable_id) continue;
						int present = 0;
						for (int q = head; q < tail && !present; ++q) present = qc[q] == c_id && qv[q] == v_id;
						if (!present) CS_PUSH(c_id, v_id);
					}
				}
			}
		}
		for (int i = 0; i < nrem; ++i) cs_erase(c, &d, variable_id, rem[i]);
		if (d.size[variable_id] == 0) break; /* once a domain is empty, no need to go further */
	}
#undef CS_PUSH
	free(rem); free(qc); free(qv);
	return d;
}

static void cs_emit(CsCtx* c)
{
	const OracleModel* m = c->m;
	if (c->cap > 0 && c->count >= c->cap) { c->truncated = 1; return; }
	if (c->out_x) memcpy(c->out_x + (size_t)c->count * m->V, c->w->x, sizeof(int32_t) * m->V);
	if (c->out_cost) {
		/* solver.hpp:969-975: objective->cost(), negated back for a maximisation */
		double cost = objective_cost(c->w);
		c->out_cost[c->count] = m->obj_kind == GHOST_OBJ_LINEAR_MAXIMIZE ? -cost : cost;
	}
	++c->count;
}

/* solver.hpp:300-347 */
static void cs_recurse(CsCtx* c, int index_v, const CsDoms* domains)
{
	const OracleModel* m = c->m;
	if (index_v >= m->V) return;
	CsDoms nd;
	if (index_v > 0) {
		nd = cs_ac3(c, index_v, domains);
		for (int v = 0; v < m->V; ++v) if (nd.size[v] == 0) { cs_free(&nd); return; }
	} else nd = cs_copy(c, domains); /* already filtered */
	int next_var = index_v + 1;
	for (int k = 0; k < nd.size[next_var] && !c->truncated; ++k) {
		c->w->x[next_var] = nd.val[m->dom_off[next_var] + k];
		if (next_var == m->V - 1) cs_emit(c);
		else cs_recurse(c, next_var, &nd);
	}
	cs_free(&nd);
}

int oracle_complete_search(const OracleModel* m, int64_t max_solutions, int32_t* solutions, double* costs,
                           int64_t* count, int32_t* truncated)
{
	CsCtx c;
	memset(&c, 0, sizeof c);
	c.m = m; c.w = walker_create(m, NULL, 0);
	c.ndom = m->dom_off[m->V];
	c.cap = max_solutions; c.out_x = solutions; c.out_cost = costs;
	CsDoms d;
	d.val = (int32_t*)malloc(sizeof(int32_t) * (c.ndom > 0 ? c.ndom : 1));
	d.size = (int32_t*)malloc(sizeof(int32_t) * m->V);
	memcpy(d.val, m->dom_given, sizeof(int32_t) * c.ndom);
	for (int v = 0; v < m->V; ++v) d.size[v] = m->dom_off[v + 1] - m->dom_off[v];
	memcpy(c.w->x, m->start, sizeof(int32_t) * m->V);
	cs_prefiltering(&c, &d);
	/* solver.hpp:957-982 */
	for (int k = 0; k < d.size[0] && !c.truncated; ++k) {
		c.w->x[0] = d.val[m->dom_off[0] + k];
		CsDoms nd = cs_ac3(&c, 0, &d);
		int empty = 0;
		for (int v = 0; v < m->V; ++v) if (nd.size[v] == 0) empty = 1;
		if (!empty) {
			if (m->V == 1) cs_emit(&c); /* the reference indexes variable 1 here (undefined for V = 1) */
			else cs_recurse(&c, 0, &nd);
		}
		cs_free(&nd);
	}
	cs_free(&d);
	if (count) *count = c.count;
	if (truncated) *truncated = c.truncated;
	walker_destroy(c.w);
	return 0;
}
