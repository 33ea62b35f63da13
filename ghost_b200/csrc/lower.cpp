// lower.cpp — host-only validation and lowering of a GhostModelDesc (no CUDA calls here, so the
// descriptor checks run on a machine without a GPU).
//
// What it reproduces from the reference:
//  * matrix_var_ctr: for every variable the ascending list of constraints that contain it
//    (include/search_unit_data.hpp:95-102);
//  * each variable's position / coefficient inside a constraint (Constraint::_variables_position,
//    include/constraint.hpp:70; LinearEquation::_coefficients, linear_equation.hpp:46);
//  * the dense count tables that stand for AllDifferent::_count (all_different.hpp:48).
// What it rejects (GHOST_B200_ERR_UNSUPPORTED): user-defined constraints/objectives, non-integer
// coefficients, error magnitudes that do not fit int32, models whose walker state does not fit in
// shared memory.  There is no CPU fallback.
#include "lowered.h"

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <set>

namespace ghost_b200 {

static bool is_linear(int k) { return k >= GHOST_KIND_LINEAR_EQ && k <= GHOST_KIND_LINEAR_G; }
// kinds that keep a value-count table
static bool is_alldiff_like(int k) { return k == GHOST_KIND_ALL_DIFFERENT || k == GHOST_KIND_ALL_DIFFERENT_ABS_DIFF || k == GHOST_KIND_ALL_EQUAL || k == GHOST_KIND_ALL_DIFFERENT_OFFSET || k == GHOST_KIND_ALL_DIFFERENT_LAG_DIFF; }
static bool is_pair_diff(int k) { return k == GHOST_KIND_ALL_DIFFERENT_ABS_DIFF || k == GHOST_KIND_ALL_DIFFERENT_LAG_DIFF; }

static bool as_int(double d, int64_t* out)
{
	if (!(d == d) || std::fabs(d) > 4.0e15) return false;
	double r = std::nearbyint(d);
	if (r != d) return false;
	*out = static_cast<int64_t>(r);
	return true;
}

int lower_model(const GhostModelDesc* d, LoweredModel* out, std::string* why)
{
	auto fail = [&](int code, const std::string& msg) { if (why) *why = msg; return code; };
	if (!d || !out) return fail(GHOST_B200_ERR_INVALID, "null descriptor");
	const int V = d->num_variables, C = d->num_constraints;
	if (V <= 0) return fail(GHOST_B200_ERR_INVALID, "model has no variable");
	if (C <= 0) return fail(GHOST_B200_ERR_INVALID, "model has no constraint (pass one PURE_OPTIMIZATION constraint, as ModelBuilder::build_model does)");
	// size limits of the packed tables, checked before anything is built
	if (V > 65535) return fail(GHOST_B200_ERR_UNSUPPORTED, "more than 65535 variables (scope lists are stored as 16-bit positions)");
	if (C >= GB_ELL_END) return fail(GHOST_B200_ERR_UNSUPPORTED, "too many constraints (20-bit constraint ids in the row tables)");
	if (!d->dom_offset || !d->dom_values || !d->start_value || !d->ctr_kind || !d->ctr_offset || !d->ctr_vars)
		return fail(GHOST_B200_ERR_INVALID, "null array in descriptor");
	if (d->dom_offset[0] != 0 || d->ctr_offset[0] != 0) return fail(GHOST_B200_ERR_INVALID, "offset arrays must start at 0");

	// ---- domains
	std::vector<int32_t> dom_lo(V), dom_size(V), dom_off(V + 1, 0), dom_vals;
	bool contiguous = true, identical = true;
	int max_dom = 0;
	int64_t max_abs_value = 0, min_value = 0, max_value = 0;
	for (int v = 0; v < V; ++v) {
		int b = d->dom_offset[v], e = d->dom_offset[v + 1];
		if (e <= b) return fail(GHOST_B200_ERR_INVALID, "variable " + std::to_string(v) + " has an empty domain");
		std::vector<int32_t> dv(d->dom_values + b, d->dom_values + e);
		if (!std::is_sorted(dv.begin(), dv.end())) std::sort(dv.begin(), dv.end());
		if (std::adjacent_find(dv.begin(), dv.end()) != dv.end())
			return fail(GHOST_B200_ERR_INVALID, "variable " + std::to_string(v) + " has a duplicated domain value");
		if (!std::binary_search(dv.begin(), dv.end(), d->start_value[v]))
			return fail(GHOST_B200_ERR_INVALID, "start value of variable " + std::to_string(v) + " is not in its domain");
		dom_lo[v] = dv.front();
		dom_size[v] = static_cast<int32_t>(dv.size());
		if (static_cast<int64_t>(dv.back()) - dv.front() + 1 != static_cast<int64_t>(dv.size())) contiguous = false;
		dom_off[v + 1] = dom_off[v] + dom_size[v];
		dom_vals.insert(dom_vals.end(), dv.begin(), dv.end());
		max_dom = std::max(max_dom, dom_size[v]);
		min_value = std::min<int64_t>(min_value, dv.front());
		max_value = std::max<int64_t>(max_value, dv.back());
		max_abs_value = std::max<int64_t>(max_abs_value, std::max(std::llabs((long long)dv.front()), std::llabs((long long)dv.back())));
		if (v > 0) {
			if (dom_size[v] != dom_size[0] || !std::equal(dv.begin(), dv.end(), dom_vals.begin())) identical = false;
		}
	}
	if (max_abs_value > (1 << 28)) return fail(GHOST_B200_ERR_UNSUPPORTED, "domain values beyond 2^28 are outside the int32 GPU path");

	// ---- constraints
	const int nnz = d->ctr_offset[C];
	std::vector<int32_t> kind(C), rhs(C, 0), cnt_tab(C, 0), c2v_off(C + 1, 0), c2v(nnz), ent_ctr(nnz), ent_coef(nnz, 0);
	std::vector<int32_t> cnt_owner, cnt_lo(C, 0), cnt_n(C, 0);
	bool has_all_equal = false;
	std::vector<std::vector<std::pair<int, int>>> rows(V); // (constraint, coefficient|position)
	int cnt_total = 0;
	int64_t error_bound = 0, static_error = 0;
	int64_t max_T_bound = 0, max_abs_coef = 0;   // over the linear constraints: |sum - rhs| and |coefficient|
	bool has_abs_diff = false;
	for (int c = 0; c < C; ++c) {
		int k = d->ctr_kind[c];
		int b = d->ctr_offset[c], e = d->ctr_offset[c + 1];
		if (e < b) return fail(GHOST_B200_ERR_INVALID, "constraint offsets not monotone");
		if (!(is_linear(k) || is_alldiff_like(k) || k == GHOST_KIND_FIX_VALUE || k == GHOST_KIND_PURE_OPTIMIZATION))
			return fail(GHOST_B200_ERR_UNSUPPORTED, "constraint " + std::to_string(c) + " is not a stock global constraint "
			            "(user-defined Constraint subclasses are outside the GPU path; there is no CPU fallback)");
		kind[c] = k;
		c2v_off[c + 1] = e;
		std::set<int> seen;
		int64_t lo = std::numeric_limits<int64_t>::max(), hi = std::numeric_limits<int64_t>::min();
		int64_t lin_bound = 0;
		for (int i = b; i < e; ++i) {
			int v = d->ctr_vars[i];
			if (v < 0 || v >= V) return fail(GHOST_B200_ERR_INVALID, "constraint " + std::to_string(c) + " names variable " + std::to_string(v));
			if (!seen.insert(v).second) return fail(GHOST_B200_ERR_INVALID, "constraint " + std::to_string(c) + " lists variable " + std::to_string(v) + " twice");
			c2v[i] = v;
			ent_ctr[i] = c;
			int64_t vlo = dom_lo[v], vhi = dom_vals[dom_off[v + 1] - 1];
			int second = 0;
			if (k == GHOST_KIND_ALL_DIFFERENT_OFFSET) {
				// the count table is indexed by the shifted value x + o
				int64_t off;
				double given = d->ctr_coefs ? d->ctr_coefs[i] : 0.0;
				if (!as_int(given, &off) || std::llabs((long long)off) > (1 << 20))
					return fail(GHOST_B200_ERR_UNSUPPORTED, "constraint " + std::to_string(c) + ": offset " + std::to_string(given) + " is not a (small) integer");
				second = static_cast<int32_t>(off);
				ent_coef[i] = second;
				vlo += off; vhi += off;
			}
			lo = std::min(lo, vlo); hi = std::max(hi, vhi);
			if (is_linear(k)) {
				int64_t a;
				double coef = d->ctr_coefs ? d->ctr_coefs[i] : 1.0;
				if (!as_int(coef, &a) || std::llabs((long long)a) > (1 << 28))
					return fail(GHOST_B200_ERR_UNSUPPORTED, "constraint " + std::to_string(c) + ": coefficient " + std::to_string(coef) +
					            " is not a (small) integer; the GPU path computes in int32");
				ent_coef[i] = static_cast<int32_t>(a);
				second = static_cast<int32_t>(a);
				lin_bound += std::llabs((long long)a) * std::max(std::llabs((long long)vlo), std::llabs((long long)vhi));
				max_abs_coef = std::max<int64_t>(max_abs_coef, std::llabs((long long)a));
			} else if (is_pair_diff(k)) {
				second = i - b; // position in scope
			}
			rows[v].emplace_back(c, second);
		}
		if (is_linear(k) || k == GHOST_KIND_FIX_VALUE) {
			int64_t p;
			double param = d->ctr_param ? d->ctr_param[c] : 0.0;
			if (!as_int(param, &p) || std::llabs((long long)p) > (1LL << 30))
				return fail(GHOST_B200_ERR_UNSUPPORTED, "constraint " + std::to_string(c) + ": rhs/value " + std::to_string(param) + " is not a (small) integer");
			rhs[c] = static_cast<int32_t>(p);
			if (is_linear(k)) {
				if (lin_bound + std::llabs((long long)p) >= (1LL << 30))
					return fail(GHOST_B200_ERR_UNSUPPORTED, "constraint " + std::to_string(c) + ": |sum| can exceed 2^30");
				error_bound += lin_bound + std::llabs((long long)p) + 1;
				max_T_bound = std::max<int64_t>(max_T_bound, lin_bound + std::llabs((long long)p));
			} else {
				for (int i = b; i < e; ++i) {
					int v = c2v[i];
					error_bound += std::max(std::llabs((long long)dom_lo[v] - p), std::llabs((long long)dom_vals[dom_off[v + 1] - 1] - p));
				}
			}
		}
		if (k == GHOST_KIND_ALL_DIFFERENT && d->permutation_problem && (e - b) == V) {
			// A permutation problem only ever swaps values, so an AllDifferent over ALL variables keeps
			// the error its start values give it, and every move's delta on it is 0
			// (all_different.cpp:77-100 with both swapped variables in scope): no cache, no table.
			std::vector<int32_t> vals(d->start_value, d->start_value + V);
			std::sort(vals.begin(), vals.end());
			int64_t err = 0;
			for (size_t i = 0; i < vals.size();) {
				size_t j2 = i;
				while (j2 < vals.size() && vals[j2] == vals[i]) ++j2;
				int64_t n2 = static_cast<int64_t>(j2 - i);
				err += n2 * (n2 - 1) / 2;
				i = j2;
			}
			kind[c] = GB_KIND_STATIC;
			rhs[c] = static_cast<int32_t>(err);
			static_error += err;
			error_bound += err;
		} else if (is_alldiff_like(k)) {
			int64_t n = e - b;
			if (n == 0) { lo = 0; hi = 0; }
			int64_t base = lo, size = hi - lo + 1;
			if (k == GHOST_KIND_ALL_DIFFERENT_ABS_DIFF) { base = 0; size = hi - lo + 1; has_abs_diff = true; }
			if (k == GHOST_KIND_ALL_DIFFERENT_LAG_DIFF) {
				// signed differences x[p+lag] - x[p]; the lag travels in the rhs section
				int64_t lag;
				double param = d->ctr_param ? d->ctr_param[c] : 0.0;
				if (!as_int(param, &lag) || lag < 1 || lag > 65535) return fail(GHOST_B200_ERR_INVALID, "constraint " + std::to_string(c) + ": the lag of AllDifferentLagDiff must be an integer >= 1");
				rhs[c] = static_cast<int32_t>(lag);
				base = -(hi - lo); size = 2 * (hi - lo) + 1; has_abs_diff = true;
			}
			if (size > (1 << 20)) return fail(GHOST_B200_ERR_UNSUPPORTED, "constraint " + std::to_string(c) + ": value range too wide for a dense count table");
			if (n > 65535) return fail(GHOST_B200_ERR_UNSUPPORTED, "AllDifferent scope beyond 65535");
			cnt_tab[c] = static_cast<int32_t>(cnt_total - base);
			cnt_lo[c] = cnt_total; cnt_n[c] = static_cast<int32_t>(size);
			if (k == GHOST_KIND_ALL_EQUAL) { has_all_equal = true; if (n > 32767) return fail(GHOST_B200_ERR_UNSUPPORTED, "AllEqual scope beyond 32767"); }
			for (int64_t t = 0; t < size; ++t) cnt_owner.push_back(c);
			cnt_total += static_cast<int>(size);
			error_bound += k == GHOST_KIND_ALL_EQUAL ? n : n * (n - 1) / 2;
		}
	}
	if (error_bound >= (1LL << 31) - 1)
		return fail(GHOST_B200_ERR_UNSUPPORTED, "total satisfaction error can exceed int32 (bound " + std::to_string(error_bound) + ")");

	// ---- var -> ctr rows (ascending constraint id by construction)
	std::vector<int32_t> v2c_off(V + 1, 0), v2c;
	int max_degree = 0;
	for (int v = 0; v < V; ++v) {
		v2c_off[v + 1] = v2c_off[v] + static_cast<int32_t>(rows[v].size());
		max_degree = std::max<int>(max_degree, static_cast<int>(rows[v].size()));
		for (auto& pr : rows[v]) { v2c.push_back(pr.first); v2c.push_back(pr.second); }
	}
	if (max_degree > 255) return fail(GHOST_B200_ERR_UNSUPPORTED, "a variable appears in more than 255 constraints");

	// ---- objective
	std::vector<int32_t> obj_coef(V, 0);
	int ok = d->objective_kind;
	if (!(ok == GHOST_OBJ_NONE || ok == GHOST_OBJ_LINEAR_MINIMIZE || ok == GHOST_OBJ_LINEAR_MAXIMIZE))
		return fail(GHOST_B200_ERR_UNSUPPORTED, "objective is not a linear objective (user-defined Objective subclasses are outside the GPU path)");
	if (ok != GHOST_OBJ_NONE) {
		int64_t bound = 0;
		for (int i = 0; i < d->num_objective_terms; ++i) {
			int v = d->obj_vars[i];
			int64_t a;
			if (v < 0 || v >= V) return fail(GHOST_B200_ERR_INVALID, "objective names variable " + std::to_string(v));
			if (!as_int(d->obj_coefs[i], &a) || std::llabs((long long)a) > (1 << 28))
				return fail(GHOST_B200_ERR_UNSUPPORTED, "objective coefficient is not a (small) integer");
			obj_coef[v] += static_cast<int32_t>(a);
			bound += std::llabs((long long)a) * std::max(std::llabs((long long)dom_lo[v]), std::llabs((long long)dom_vals[dom_off[v + 1] - 1]));
		}
		// the packed best-key (quality << 24 | walker) keeps 38 bits for a solved COP's cost
		if (bound >= (1LL << 37)) return fail(GHOST_B200_ERR_UNSUPPORTED, "objective magnitude beyond 2^37");
	}

	// ---- blob
	LoweredModel& L = *out;
	L = LoweredModel();
	static std::atomic<uint64_t> next_uid{1};
	L.uid = next_uid.fetch_add(1);
	L.V = V; L.C = C;
	L.desc.dom_off.assign(d->dom_offset, d->dom_offset + V + 1);
	L.desc.dom_val.assign(d->dom_values, d->dom_values + d->dom_offset[V]);
	L.desc.kind.assign(d->ctr_kind, d->ctr_kind + C);
	L.desc.ctr_off = c2v_off; L.desc.ctr_var = c2v; L.desc.ctr_coef = ent_coef;
	L.desc.ctr_param.assign(C, 0);
	for (int c = 0; c < C; ++c) if (is_linear(d->ctr_kind[c]) || d->ctr_kind[c] == GHOST_KIND_FIX_VALUE || d->ctr_kind[c] == GHOST_KIND_ALL_DIFFERENT_LAG_DIFF) L.desc.ctr_param[c] = static_cast<int32_t>(std::llround(d->ctr_param ? d->ctr_param[c] : 0.0));
	L.desc.obj_coef = obj_coef; L.desc.objective = ok;
	L.start.assign(d->start_value, d->start_value + V);
	L.error_bound = error_bound;
	std::vector<int32_t>& blob = L.blob;
	BlobOffsets o{};
	auto put = [&](const std::vector<int32_t>& a) {
		int at = static_cast<int>(blob.size());
		blob.insert(blob.end(), a.begin(), a.end());
		while (blob.size() % 4) blob.push_back(0); // keep every section 16-byte aligned
		return at;
	};
	o.dom_lo = put(dom_lo); o.dom_size = put(dom_size); o.dom_off = put(dom_off); o.dom_vals = put(dom_vals);
	o.start = put(L.start);
	o.v2c_off = put(v2c_off); o.v2c = put(v2c);
	o.kind = put(kind); o.rhs = put(rhs); o.cnt_tab = put(cnt_tab);
	o.c2v_off = put(c2v_off); o.c2v = put(c2v); o.ent_ctr = put(ent_ctr); o.ent_coef = put(ent_coef);
	o.cnt_owner = put(cnt_owner); o.cnt_lo = put(cnt_lo); o.cnt_n = put(cnt_n);
	o.obj_coef = put(obj_coef);
	// ELL rows for the hot loop: one fixed-stride row per variable, class / error weights packed in
	// the entry so that the candidate loop needs no per-constraint global lookups for linear kinds.
	std::vector<std::vector<std::pair<int, int>>> live(V); // rows without the static constraints
	int live_degree = 0;
	bool eq_only = true;
	for (int c = 0; c < C; ++c) if (!(kind[c] == GHOST_KIND_LINEAR_EQ || kind[c] == GB_KIND_STATIC)) eq_only = false;
	for (int v = 0; v < V; ++v) {
		for (auto& pr : rows[v]) if (kind[pr.first] != GB_KIND_STATIC) live[v].push_back(pr);
		live_degree = std::max<int>(live_degree, static_cast<int>(live[v].size()));
	}
	const int stride = std::max(1, live_degree);
	// one row per variable, padded with all-dummy rows to a whole number of 32-variable chunks: the
	// candidate loop of a permutation problem then reads the row of every slot of a chunk unconditionally
	const int vpad_rows = (V + 31) / 32 * 32;
	std::vector<int32_t> ell(static_cast<size_t>(vpad_rows) * stride * 2, 0), has_ctr((V + 31) / 32, 0);
	for (int v = 0; v < vpad_rows; ++v) {
		if (v < V && !rows[v].empty()) has_ctr[v >> 5] |= (1u << (v & 31));
		for (int k = 0; k < stride; ++k) {
			// rows are padded with the dummy constraint C: coefficient 0, error 0, never marked, so the
			// specialised loops need no end-of-row test
			int w0 = C | (GB_CLS_ZERO << 20), w1 = 0;
			if (v < V && k < static_cast<int>(live[v].size())) {
				int c = live[v][k].first, kd = kind[c];
				int cls = GB_CLS_ZERO, wbits = 0;
				switch (kd) {
				case GHOST_KIND_LINEAR_EQ:  cls = GB_CLS_LINEAR; wbits = 3; break; // |t| = max(t,0) + max(-t,0)
				case GHOST_KIND_LINEAR_NEQ: cls = GB_CLS_LINEAR; wbits = 4; break;
				case GHOST_KIND_LINEAR_LEQ: cls = GB_CLS_LINEAR; wbits = 1; break;
				case GHOST_KIND_LINEAR_GEQ: cls = GB_CLS_LINEAR; wbits = 2; break;
				case GHOST_KIND_LINEAR_L:   cls = GB_CLS_LINEAR; wbits = 5; break;
				case GHOST_KIND_LINEAR_G:   cls = GB_CLS_LINEAR; wbits = 6; break;
				case GHOST_KIND_ALL_DIFFERENT: cls = GB_CLS_ALLDIFF; break;
				case GHOST_KIND_ALL_DIFFERENT_OFFSET: cls = GB_CLS_ALLDIFF_OFF; break;
				case GHOST_KIND_FIX_VALUE: cls = GB_CLS_FIX; break;
				case GHOST_KIND_ALL_DIFFERENT_ABS_DIFF: cls = GB_CLS_ABSDIFF; break;
				case GHOST_KIND_ALL_DIFFERENT_LAG_DIFF: cls = GB_CLS_ABSDIFF; break;
				case GHOST_KIND_ALL_EQUAL: cls = GB_CLS_ALLEQ; break;
				default: cls = GB_CLS_ZERO; break;
				}
				int universal = (c2v_off[c + 1] - c2v_off[c] == V) ? 1 : 0;
				w0 = c | (cls << 20) | (wbits << 24) | (universal << 27);
				w1 = live[v][k].second;
			}
			ell[(static_cast<size_t>(v) * stride + k) * 2] = w0;
			ell[(static_cast<size_t>(v) * stride + k) * 2 + 1] = w1;
		}
	}
	o.ell = put(ell);
	// Packed rows of the LinearEq-only paths: 16-bit state offset + 16-bit coefficient per entry, a whole
	// row in one aligned load.  The table (4 B x stride x V) is what the candidate loop streams through
	// L1 every iteration, next to shared-memory carve-outs that leave L1 only a few tens of KB.
	const int stride_eq = stride <= 2 ? stride : (stride + 3) / 4 * 4;
	if ((C + 1) * 8 > 65535) eq_only = false;
	for (size_t i = 0; eq_only && i + 1 < ell.size(); i += 2) if (ell[i + 1] < -32768 || ell[i + 1] > 32767) eq_only = false;
	{
		std::vector<int32_t> ell_eq(eq_only ? static_cast<size_t>(vpad_rows) * stride_eq : 0);
		for (int v = 0; eq_only && v < vpad_rows; ++v)
			for (int k = 0; k < stride_eq; ++k) {
				uint32_t off = static_cast<uint32_t>(C) * 8u, coef = 0;
				if (k < stride) {
					size_t i = (static_cast<size_t>(v) * stride + k) * 2;
					off = static_cast<uint32_t>(ell[i] & 0xFFFFF) * 8u;
					coef = static_cast<uint32_t>(ell[i + 1]) & 0xFFFFu;
				}
				ell_eq[static_cast<size_t>(v) * stride_eq + k] = static_cast<int32_t>(off | (coef << 16));
			}
		o.ell_eq = put(ell_eq);
	}
	// fp32 evaluation of the swap candidates (walker.cuh, evaluate_f32): every term is |T + a*dx| on integers, so
	// fp32 is exact while magnitudes stay below 2^24.  B bounds one term (|T| plus both variables' a*dx); a
	// candidate's sum has at most 4 + 2*4 + 4 terms and is accumulated on top of the constant 1.5*2^23 (its bit
	// pattern is then the integer sum plus a constant: the arg-min needs no conversion), which needs every partial
	// sum inside (-2^22, 2^22).
	bool f32_ok = eq_only && d->permutation_problem && identical && ok == GHOST_OBJ_NONE && stride <= 4 && min_value >= 0 && max_value <= 65535 && V <= 1024;
	if (f32_ok) {
		const int64_t B = max_T_bound + 2 * max_abs_coef * (max_value - min_value);
		if (max_abs_coef > 2048 || 16 * B >= (1LL << 22)) f32_ok = false;
		std::vector<int32_t> sv(d->start_value, d->start_value + V);
		std::sort(sv.begin(), sv.end());
		if (std::adjacent_find(sv.begin(), sv.end()) != sv.end()) f32_ok = false; // the candidates of a variable are then exactly the V - 1 others
		if (std::getenv("GHOST_B200_NO_F32")) f32_ok = false; // test knob: the int32 evaluation of the same model
	}
	{
		auto half_bits = [](int a) -> uint32_t { // IEEE binary16 of an integer |a| <= 2048 (exact)
			if (a == 0) return 0u;
			const uint32_t sign = a < 0 ? 0x8000u : 0u;
			uint32_t mag = static_cast<uint32_t>(a < 0 ? -a : a);
			int e = 0;
			while ((mag >> (e + 1)) != 0) ++e;
			const uint32_t mant = e <= 10 ? (mag << (10 - e)) & 0x3FFu : (mag >> (e - 10)) & 0x3FFu;
			return sign | (static_cast<uint32_t>(e + 15) << 10) | mant;
		};
		// one more all-dummy chunk of rows: the fp32 loop evaluates two chunks per pass
		std::vector<int32_t> ell_f32(f32_ok ? static_cast<size_t>(vpad_rows + 32) * stride_eq : 0);
		for (int v = 0; f32_ok && v < vpad_rows + 32; ++v)
			for (int k = 0; k < stride_eq; ++k) {
				uint32_t off = static_cast<uint32_t>(C) * 8u, coef = 0;
				if (k < stride && v < vpad_rows) {
					size_t i = (static_cast<size_t>(v) * stride + k) * 2;
					off = static_cast<uint32_t>(ell[i] & 0xFFFFF) * 8u;
					coef = half_bits(ell[i + 1]);
				}
				ell_f32[static_cast<size_t>(v) * stride_eq + k] = static_cast<int32_t>((off << 16) | coef);
			}
		o.ell_f32 = put(ell_f32);
	}
	{
		// scope lists as 16-bit scan positions (see scan_pos in walker.cuh)
		const int v128 = V & ~127;
		std::vector<int32_t> packed((c2v.size() + 1) / 2, 0);
		for (size_t i = 0; i < c2v.size(); ++i) {
			int v = c2v[i];
			uint32_t pos = v < v128 ? static_cast<uint32_t>((v & ~127) | ((v & 31) << 2) | ((v >> 5) & 3)) : static_cast<uint32_t>(v);
			packed[i / 2] |= static_cast<int32_t>(pos << ((i & 1) * 16));
		}
		o.c2v_pos = put(packed);
	}
	o.has_ctr = put(has_ctr);
	{
		while (blob.size() % 4) blob.push_back(0);   // 16-byte alignment of the int4 records
		std::vector<int32_t> pd(static_cast<size_t>(C) * 4, 0);
		for (int c = 0; c < C; ++c) {
			if (!is_pair_diff(kind[c])) continue;
			const int b = c2v_off[c], n = c2v_off[c + 1] - b;
			bool identity = true;
			for (int p2 = 0; p2 < n; ++p2) if (c2v[b + p2] != p2) { identity = false; break; }
			const bool absolute = kind[c] == GHOST_KIND_ALL_DIFFERENT_ABS_DIFF;
			pd[4 * c + 0] = b; pd[4 * c + 1] = n; pd[4 * c + 2] = cnt_tab[c];
			pd[4 * c + 3] = static_cast<int32_t>(static_cast<uint32_t>(absolute ? 1 : rhs[c]) | (absolute ? (1u << 30) : 0u) | (identity ? (1u << 31) : 0u));
		}
		o.pd_desc = put(pd);
	}
	bool given_sorted = true;
	{
		std::vector<int32_t> given(d->dom_values, d->dom_values + d->dom_offset[V]);
		given_sorted = given == dom_vals;
		o.dom_given = put(given);
	}
	// ---- structured grid -------------------------------------------------------------------------------------
	// Matrix-shaped models (magic squares, Latin-square sums, ...): with the variables cut into consecutive groups of
	// CW, every variable (g, l) is in U_g, a constraint common to its whole group, and in L_l, a constraint common to
	// position l of every group, both with one common coefficient; few cells carry anything else (the diagonals).
	// The fp32 evaluation then keeps L_l's state in a register of lane l for the whole candidate loop and broadcasts
	// U_g's once per chunk, instead of a row load and a shared-memory gather per candidate and constraint; the few
	// irregular cells are evaluated separately by the lane that owns their chunk's summary (walker.cuh, evaluate_sg).
	int sg_ok = 0, sg_cw = 0, sg_groups = 0, sg_xn = 0, sg_au = 0, sg_al = 0;
	{
		std::vector<int32_t> sg(32 * 4, 0);
		// (groups narrower than half a warp would leave most lanes idle: the plain loop is the better one then)
		for (int cw = 32; f32_ok && !sg_ok && cw >= 16 && !std::getenv("GHOST_B200_NO_SG"); --cw) {
			const int G = (V + cw - 1) / cw;
			if (G > 32 || G < 2) continue;
			// constraints (with coefficient) common to a set of variables
			auto common = [&](const std::vector<int>& vars) {
				std::vector<std::pair<int, int>> acc = live[vars[0]];
				for (size_t i = 1; i < vars.size() && !acc.empty(); ++i) {
					std::vector<std::pair<int, int>> keep;
					for (auto& pr : acc) if (std::find(live[vars[i]].begin(), live[vars[i]].end(), pr) != live[vars[i]].end()) keep.push_back(pr);
					acc.swap(keep);
				}
				return acc;
			};
			std::vector<int> U(G, -1), Lc(cw, -1);
			int au = 0, al = 0;
			bool ok_grid = true;
			for (int g = 0; g < G && ok_grid; ++g) {
				std::vector<int> vars;
				for (int l = 0; l < cw && g * cw + l < V; ++l) vars.push_back(g * cw + l);
				auto acc = common(vars);
				if (g == 0 && !acc.empty()) au = acc[0].second;
				for (auto& pr : acc) if (pr.second == au && U[g] < 0) U[g] = pr.first;
				if (U[g] < 0) ok_grid = false;
			}
			for (int l = 0; l < cw && ok_grid; ++l) {
				std::vector<int> vars;
				for (int g = 0; g < G && g * cw + l < V; ++g) vars.push_back(g * cw + l);
				auto acc = common(vars);
				bool first = true;
				for (auto& pr : acc) {
					if (std::find(U.begin(), U.end(), pr.first) != U.end()) continue; // a group constraint (e.g. one over all variables)
					if (l == 0 && first) { al = pr.second; first = false; }
					if (pr.second == al && Lc[l] < 0) Lc[l] = pr.first;
				}
				if (Lc[l] < 0) ok_grid = false;
			}
			if (!ok_grid || au == 0 || al == 0) continue;
			// the cells with further constraints, per group
			int xn = 0;
			std::vector<std::vector<int>> extra(G);
			for (int v = 0; v < V && ok_grid; ++v) {
				const int g = v / cw, l = v % cw;
				bool has_u = false, has_l = false;
				size_t others = 0;
				for (auto& pr : live[v]) {
					if (pr.first == U[g] && !has_u) has_u = true;
					else if (pr.first == Lc[l] && !has_l) has_l = true;
					else ++others;
				}
				if (!has_u || !has_l) ok_grid = false;
				if (others > 0) extra[g].push_back(l);
			}
			for (int g = 0; g < G; ++g) { if (extra[g].size() > 4) ok_grid = false; xn = std::max<int>(xn, static_cast<int>(extra[g].size())); }
			if (!ok_grid) continue;
			for (int lane = 0; lane < 32; ++lane) {
				sg[lane * 4 + 0] = (lane < G ? U[lane] : C) * 8;
				sg[lane * 4 + 1] = (lane < cw ? Lc[lane] : C) * 8;
				uint32_t skip = 0, packed = 0xFFFFFFFFu;
				for (int g = 0; g < 32; ++g) {
					const bool cell = g < G && lane < cw && g * cw + lane < V;
					const bool irregular = cell && std::find(extra[g].begin(), extra[g].end(), lane) != extra[g].end();
					if (!cell || irregular) skip |= 1u << g;
				}
				if (lane < G) for (size_t e = 0; e < extra[lane].size(); ++e) packed = (packed & ~(0xFFu << (8 * e))) | (static_cast<uint32_t>(extra[lane][e]) << (8 * e));
				sg[lane * 4 + 2] = static_cast<int32_t>(skip);
				sg[lane * 4 + 3] = static_cast<int32_t>(packed);
			}
			sg_ok = 1; sg_cw = cw; sg_groups = G; sg_xn = xn; sg_au = au; sg_al = al;
		}
		o.sg = put(sg);
	}
	o.total = static_cast<int>(blob.size());

	DevModel& m = L.dev;
	m.blob = nullptr;
	m.o = o;
	m.V = V; m.C = C; m.nnz = nnz;
	m.permutation = d->permutation_problem ? 1 : 0;
	m.objective = ok;
	m.contiguous_domains = contiguous ? 1 : 0;
	m.identical_domains = identical ? 1 : 0;
	m.cnt_total = cnt_total;
	m.max_candidates = m.permutation ? V : max_dom;
	m.max_degree = std::max(1, max_degree);
	m.has_abs_diff = has_abs_diff ? 1 : 0;
	{
		int feat = ok != GHOST_OBJ_NONE ? GB_FEAT_OBJ : 0;
		for (int c = 0; c < C; ++c) {
			const int k = d->ctr_kind[c];
			if (is_linear(k) && k != GHOST_KIND_LINEAR_EQ) feat |= GB_FEAT_LINW;
			else if (k == GHOST_KIND_FIX_VALUE) feat |= GB_FEAT_FIX;
			else if (is_pair_diff(k)) feat |= GB_FEAT_PAIRDIFF;
			else if (k == GHOST_KIND_ALL_EQUAL) feat |= GB_FEAT_ALLEQ;
			else if (k == GHOST_KIND_ALL_DIFFERENT_OFFSET) feat |= GB_FEAT_OFFSET;
			else if (k == GHOST_KIND_ALL_DIFFERENT) feat |= GB_FEAT_ALLDIFF;
		}
		m.feat = feat;
	}
	m.has_all_equal = has_all_equal ? 1 : 0;
	m.ell_stride = stride;
	m.ell_eq_stride = stride_eq;
	m.eq_only = eq_only ? 1 : 0;
	m.all_have_ctr = 1;
	for (int v = 0; v < V; ++v) if (rows[v].empty()) m.all_have_ctr = 0;
	m.x16 = (min_value >= 0 && max_value <= 65535) ? 1 : 0;
	m.v128 = V & ~127;
	m.given_sorted = given_sorted ? 1 : 0;
	m.static_error = static_cast<int>(std::min<int64_t>(static_error, INT32_MAX));
	m.f32_ok = f32_ok ? 1 : 0;
	m.sg_ok = sg_ok; m.sg_cw = sg_cw; m.sg_groups = sg_groups; m.sg_xn = sg_xn; m.sg_au = sg_au; m.sg_al = sg_al;
	return GHOST_B200_OK;
}

static int align_up(int x, int a) { return (x + a - 1) / a * a; }

WalkerLayout LoweredModel::layout(int /*threads_per_walker*/, int nreset, int strategy) const
{
	const DevModel& m = dev;
	WalkerLayout w{};
	int at = 0;
	auto take = [&](int bytes) { int r = at; at = align_up(at + bytes, 16); return r; };
	w.nchunks_v = (m.V + 31) / 32;
	w.nchunks_c = (m.max_candidates + 31) / 32;
	const int vpad = w.nchunks_v * 32;
	w.x = take((m.x16 ? 2 : 4) * (vpad + 64));  // + two chunks: the fp32 evaluations read two chunks per pass, the grid form up to 32 slots past a group  // padded: a chunk of swap partners is read without a clamp   // scan without bounds tests: padded, the padding is always tabu
	w.err_var = take(4 * vpad);
	w.tabu = take(2 * vpad);
	w.cs = take(8 * (m.C + 1));  // + the all-zero dummy constraint that pads the ELL rows
	w.cs_em = take(8 * (m.C + 1));
	// (at least 32 entries each, back to back: a one-warp walker keeps these summaries in registers while a launch runs
	//  and the one-warp kernels of the default strategy use the 256 bytes as their cache of 16 Philox blocks, kernels.cuh)
	w.cand_key = take(4 * std::max(w.nchunks_v, 32));
	w.cand_bits = take(4 * std::max(w.nchunks_v, 32));
	w.cnt = take(2 * std::max(1, m.cnt_total));
	if (strategy == GHOST_STRATEGY_CULPRIT) {
		w.errd = take(8 * vpad);
		w.culprit = take(8 * std::max(1, m.nnz));
		w.culprit_out = take(8 * std::max(1, m.C));
		w.cpos = take(2 * vpad);
	}
	w.persist_bytes = at;
	w.counters = take(8 * 8);
	// sampling scratch lives in arrays that are dead while a reset samples: the tabu list (cleared by
	// initialize_data_structures afterwards) and the projected errors (recomputed afterwards)
	w.flags = w.tabu;
	// chunk summaries of the value evaluation; one-warp walkers use the two arrays as ONE staging area of
	// {key, bits} pairs for the variable scan and the fp32 evaluation (hence the common size, 32 + 2 chunks)
	w.val_key = take(4 * (std::max(std::max(w.nchunks_c, w.nchunks_v), 32) + 2));
	w.val_bits = take(4 * (std::max(std::max(w.nchunks_c, w.nchunks_v), 32) + 2));
	w.hoist = take(16 * m.max_degree);
	w.upd = take(4 * 2 * 2 * m.max_degree);
	w.sample = (8 * std::max(1, nreset) <= 4 * m.V) ? w.err_var : take(4 * 2 * std::max(1, nreset));
	w.red = take(4 * 32);   // [0,8) per-warp partials, [8] stop-flag broadcast, [16,32) 64-bit partials
	w.sums = (strategy == GHOST_STRATEGY_ANTIDOTE || strategy == GHOST_STRATEGY_RANDOM_WALK) ? take(4 * 32 * w.nchunks_c) : w.red;
	w.total_bytes = at;
	return w;
}

int LoweredModel::default_threads_per_walker() const
{
	// One warp per walker is the throughput optimum as long as an iteration is a few dozen warp
	// passes (measured on the magic-square 30x30 benchmark: 32 > 64 > 128 threads per walker, the
	// serial part of an iteration is not replicated across warps); wider groups only when one
	// iteration's candidate set (or the variable scan) is thousands of slots.
	int work = std::max(dev.max_candidates, dev.V);
	if (work <= 1024) return 32;
	if (work <= 4096) return 64;
	if (work <= 16384) return 128;
	return 256;
}

int LoweredModel::effective_threads_per_walker(int requested) const
{
	// A one-warp walker keeps its chunk summaries in registers, one chunk per lane: at most 1024
	// variables and 1024 candidate slots.  Larger models asked to run on one warp get two.
	int T = requested > 0 ? requested : default_threads_per_walker();
	if (T == 32 && std::max(dev.max_candidates, dev.V) > 1024) T = 64;
	return T;
}

} // namespace ghost_b200
