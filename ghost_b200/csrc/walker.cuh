// walker.cuh — device code of the local-search walkers (sm_100a).
//
// One walker = one SearchUnit of the reference (include/search_unit.hpp:87) mapped on a group of
// T threads: a warp (T = 32, several walkers per CTA) or a whole CTA (T = 64..512).  The walker's
// assignment, projected errors, tabu list, constraint caches (running sums, AllDifferent count
// tables) and candidate sets live in shared memory; the model (incidence both ways, coefficients,
// domains) is one read-only int32 blob read through the L1/L2 read-only path.
//
// All arithmetic is int32: the stock constraints with integer coefficients produce integer-valued
// doubles in the reference, so results are bit-identical after conversion (lower.cpp rejects
// anything that would not fit).
#pragma once

#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "lowered.h"
#include <math_constants.h>

namespace ghost_b200 {

#define GB_INT_MAX 0x7fffffff
#define GB_I64_MAX 0x7fffffffffffffffLL

// ---------------------------------------------------------------------------------------------
// Philox4x32-10 counter-based RNG (Salmon et al., SC'11).  Replaces the reference's entropy-seeded
// mt19937 (thirdparty/randutils.hpp:605-807).  Stream layout ("RNG protocol"):
//   key = (seed_lo, seed_hi ^ walker_hi), counter = (event_lo, event_hi, slot, walker_lo).
// ---------------------------------------------------------------------------------------------
// Not inlined on purpose (GB_OUTLINE): the generator is used on the hot iteration path and all over the
// reset / sampling paths; one shared copy keeps the instruction working set of an SM — 24 warps in
// different phases of the walk — inside its instruction cache (measured: the inlined copies pushed the
// hot code past it and the icache hit rate from 98% to 77%).
#ifndef GB_OUTLINE
#define GB_OUTLINE __noinline__
#endif
static __device__ GB_OUTLINE uint4 philox4x32_10(uint4 c, uint2 k)
{
	// rolled: 20 more executed instructions per call, 55 fewer in the instruction cache (see GB_OUTLINE above)
#ifndef GB_PHILOX_UNROLLED
#pragma unroll 1
#else
#pragma unroll
#endif
	for (int r = 0; r < 10; ++r) {
		uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
		uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
		c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
		k.x += 0x9E3779B9u;
		k.y += 0xBB67AE85u;
	}
	return c;
}

struct Rng {
	uint64_t seed, walker, next_event;
	__device__ __forceinline__ uint4 block(uint64_t event, uint32_t slot) const
	{
		return philox4x32_10(make_uint4((uint32_t)event, (uint32_t)(event >> 32), slot, (uint32_t)walker),
		                     make_uint2((uint32_t)seed, (uint32_t)(seed >> 32) ^ (uint32_t)(walker >> 32)));
	}
	__device__ __forceinline__ uint32_t word(uint64_t event, uint32_t index) const
	{
		uint4 b = block(event, index >> 2);
		uint32_t q = index & 3u;
		return q == 0 ? b.x : q == 1 ? b.y : q == 2 ? b.z : b.w;
	}
};
__device__ __forceinline__ uint32_t below(uint32_t word, uint32_t n) { return __umulhi(word, n); }

// ---------------------------------------------------------------------------------------------
// Thread group of one walker
// ---------------------------------------------------------------------------------------------
template <int T>
struct Group {
	static constexpr int NW = T / 32;
	__device__ __forceinline__ static void sync()
	{
		if (T == 32) __syncwarp(); else __syncthreads();
	}
	__device__ __forceinline__ static int tid() { return T == 32 ? (threadIdx.x & 31) : threadIdx.x; }
	__device__ __forceinline__ static int warp() { return T == 32 ? 0 : (threadIdx.x >> 5); }

	// Sum over the group; `red` is a private 16-int scratch region that nobody else touches
	// between this call and the next group barrier.
	__device__ __forceinline__ static int sum(int v, int* red)
	{
		v = __reduce_add_sync(0xffffffffu, v);
		if (T == 32) return v;
		if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
		__syncthreads();
		int s = 0;
#pragma unroll
		for (int w = 0; w < NW; ++w) s += red[w];
		return s;
	}
	__device__ __forceinline__ static long long sum64(long long v, int* red)
	{
#pragma unroll
		for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
		if (T == 32) return v;
		long long* r64 = reinterpret_cast<long long*>(red);
		if ((threadIdx.x & 31) == 0) r64[threadIdx.x >> 5] = v;
		__syncthreads();
		long long s = 0;
#pragma unroll
		for (int w = 0; w < NW; ++w) s += r64[w];
		return s;
	}
};

// Storage order of the two arrays the variable scan reads (projected errors, tabu list): inside every
// full block of 128 variables the entry of variable 128b + 32s + l sits at position 128b + 4l + s, so
// that lane l fetches its entry of four consecutive 32-variable chunks with ONE 128-bit (errors) and
// ONE 64-bit (tabu) load.  The tail (< 128 variables) keeps the natural order.  Positions [0, V) are
// exactly the real entries in both parts, so loops over "all entries" need no mapping.
__device__ __forceinline__ int scan_pos(const DevModel& m, int v)
{
	return v < m.v128 ? ((v & ~127) | ((v & 31) << 2) | ((v >> 5) & 3)) : v;
}

// Shared-memory view of one walker; XT = unsigned short when every domain value is in [0, 65535]
// (zero-extending 16-bit loads cost nothing, sign-extending ones an extra instruction), else int
// FEAT: which constraint classes / options the kernel is compiled for (GB_FEAT_*).  The generic kernels carry the code of
// every class on their per-iteration path; a model that uses two of them runs an instantiation that holds only those
// (measured on n-queens 100: instruction-cache hit rate 70 % with everything compiled in).  The device functions read it
// from the view type (feat_on<WS>), so their signatures do not change.
template <class XT, int FEAT = GB_FEAT_ALL>
struct WST {
	typedef XT x_type;
	static constexpr int feat = FEAT;
	XT* x; int* err_var;
	unsigned short* tabu;       // tabu_until - tabu_base, 16 bits (see WalkerRegs::tabu_base)
	int2* cs;                   // per constraint {T = sum - rhs (linear kinds), a_v of the hoisted variable}
	int2* cem;                  // per constraint {current error, mark}
	int* cand_key; unsigned* cand_bits;
	unsigned* cnt32;            // packed 16-bit counters, two per word
	long long* counters;        // GhostCounters, thread 0 only
	unsigned char* flags;
	int* val_key; unsigned* val_bits;
	int4* hoist; int* upd; int* sample; int* red;
	int* sums;                  // Antidote strategy: cumulated delta of every candidate slot
	// Culprit strategy: projected errors (fp64, natural variable order), the table rows, variable -> scope entry scratch
	double* errd; double* culprit; double* culprit_out; unsigned short* cpos;
	__device__ __forceinline__ int cnt(int idx) const { return reinterpret_cast<const unsigned short*>(cnt32)[idx]; }
	__device__ __forceinline__ void cnt_add(int idx, int d) const
	{
		atomicAdd(&cnt32[idx >> 1], (unsigned)d << ((idx & 1) * 16));
	}
	__device__ __forceinline__ int& T_of(int c) const { return reinterpret_cast<int*>(cs + c)[0]; }
	__device__ __forceinline__ int& av_of(int c) const { return reinterpret_cast<int*>(cs + c)[1]; }
	// two arrays of 8-byte stride: {T, a_v} is all the LinearEq evaluation loop gathers; {error, mark} serves the rest
	__device__ __forceinline__ int& err_of(int c) const { return reinterpret_cast<int*>(cem + c)[0]; }
	__device__ __forceinline__ int& mark_of(int c) const { return reinterpret_cast<int*>(cem + c)[1]; }
};

template <class WS> __device__ __forceinline__ constexpr bool feat_on(int f) { return (WS::feat & f) != 0; }
template <class WS>
__device__ __forceinline__ WS make_ws(unsigned char* base_in, const WalkerLayout& L)
{
	// The walker's base is turned into a 32-bit shared-window address and hidden from the optimiser
	// once: otherwise, with the register cap of the hot loop, the compiler re-derives every array
	// pointer from threadIdx (S2R + shifts + multiplies, ~7% of all instructions) instead of keeping
	// one register and adding a constant.
	unsigned sbase = static_cast<unsigned>(__cvta_generic_to_shared(base_in));
	asm volatile("" : "+r"(sbase));
	unsigned char* base = static_cast<unsigned char*>(__cvta_shared_to_generic(sbase));
	WS s;
	s.x = (typename WS::x_type*)(base + L.x); s.err_var = (int*)(base + L.err_var); s.tabu = (unsigned short*)(base + L.tabu);
	s.cs = (int2*)(base + L.cs); s.cem = (int2*)(base + L.cs_em);
	s.cand_key = (int*)(base + L.cand_key); s.cand_bits = (unsigned*)(base + L.cand_bits);
	s.cnt32 = (unsigned*)(base + L.cnt);
	s.counters = (long long*)(base + L.counters); s.flags = base + L.flags;
	s.val_key = (int*)(base + L.val_key); s.val_bits = (unsigned*)(base + L.val_bits);
	s.hoist = (int4*)(base + L.hoist); s.upd = (int*)(base + L.upd); s.sample = (int*)(base + L.sample);
	s.red = (int*)(base + L.red);
	s.sums = (int*)(base + L.sums);
	s.errd = (double*)(base + L.errd); s.culprit = (double*)(base + L.culprit); s.culprit_out = (double*)(base + L.culprit_out);
	s.cpos = (unsigned short*)(base + L.cpos);
	return s;
}

// Model accessors (read-only path)
struct MV {
	const DevModel& m;
	__device__ __forceinline__ MV(const DevModel& mm) : m(mm) {}
	__device__ __forceinline__ int ld(int off, int i) const { return __ldg(m.blob + off + i); }
	__device__ __forceinline__ int dom_lo(int v) const { return ld(m.o.dom_lo, v); }
	__device__ __forceinline__ int dom_size(int v) const { return ld(m.o.dom_size, v); }
	__device__ __forceinline__ int dom_value(int v, int k) const
	{
		return m.contiguous_domains ? dom_lo(v) + k : ld(m.o.dom_vals, ld(m.o.dom_off, v) + k);
	}
	__device__ __forceinline__ bool in_domain(int v, int value) const
	{
		int lo = dom_lo(v), n = dom_size(v);
		if (m.contiguous_domains) return value >= lo && value < lo + n;
		int b = ld(m.o.dom_off, v), e = b + n;
		while (b < e) {
			int mid = (b + e) >> 1, dv = ld(m.o.dom_vals, mid);
			if (dv == value) return true;
			if (dv < value) b = mid + 1; else e = mid;
		}
		return false;
	}
	__device__ __forceinline__ int row_begin(int v) const { return ld(m.o.v2c_off, v); }
	__device__ __forceinline__ int2 row_entry(int e) const { return __ldg(reinterpret_cast<const int2*>(m.blob + m.o.v2c) + e); }
	__device__ __forceinline__ const int2* ell_row(int v) const { return reinterpret_cast<const int2*>(m.blob + m.o.ell) + (size_t)v * m.ell_stride; }
	__device__ __forceinline__ int kind(int c) const { return ld(m.o.kind, c); }
	__device__ __forceinline__ int rhs(int c) const { return ld(m.o.rhs, c); }
	__device__ __forceinline__ int cnt_tab(int c) const { return ld(m.o.cnt_tab, c); }
	__device__ __forceinline__ int scope_begin(int c) const { return ld(m.o.c2v_off, c); }
	__device__ __forceinline__ int scope_var(int i) const { return ld(m.o.c2v, i); }
	// scan_pos() of scope entry i, precomputed (16 bits)
	__device__ __forceinline__ int scope_pos(int i) const { return __ldg(reinterpret_cast<const unsigned short*>(m.blob + m.o.c2v_pos) + i); }
	// packed row entries of the LinearEq-only paths
	__device__ __forceinline__ const unsigned* eq_row(int v) const { return reinterpret_cast<const unsigned*>(m.blob + m.o.ell_eq) + v * m.ell_eq_stride; }
	__device__ __forceinline__ int obj_coef(int v) const { return ld(m.o.obj_coef, v); }
};

// ELL entry decoding (lower.cpp)
__device__ __forceinline__ int ell_ctr(int w0) { return w0 & 0xFFFFF; }
__device__ __forceinline__ int ell_cls(int w0) { return (w0 >> 20) & 0xF; }
__device__ __forceinline__ int ell_w(int w0) { return (w0 >> 24) & 0x7; }
__device__ __forceinline__ bool ell_universal(int w0) { return (w0 >> 27) & 1; }

// LinearEquation*::compute_error on t = sum - rhs
// (reference linear_equation_eq.cpp:54-57, _neq.cpp:54-61, _leq.cpp:54-57, _geq.cpp:54-57,
//  _l.cpp:54-61, _g.cpp:54-61), branch-free: w bit0 counts max(t,0), bit1 max(-t,0), bit2 [t==0].
//  Eq = 3, Leq = 1, Geq = 2, Neq = 4, L = 5, G = 6.
__device__ __forceinline__ int linear_error_w(int w, int t)
{
	int pos = max(t, 0), neg = pos - t;
	int e = (w & 1) ? pos : 0;
	e += (w & 2) ? neg : 0;
	e += ((w & 4) && t == 0) ? 1 : 0;
	return e;
}
// the same for a view type: a model whose linear constraints are all LinearEquationEq needs |t| only
template <class WS>
__device__ __forceinline__ int linear_error_f(int w, int t) { return feat_on<WS>(GB_FEAT_LINW) ? linear_error_w(w, t) : abs(t); }
__device__ __forceinline__ int linear_w_of_kind(int kind)
{
	return kind == GHOST_KIND_LINEAR_EQ ? 3 : kind == GHOST_KIND_LINEAR_LEQ ? 1 : kind == GHOST_KIND_LINEAR_GEQ ? 2
	     : kind == GHOST_KIND_LINEAR_NEQ ? 4 : kind == GHOST_KIND_LINEAR_L ? 5 : 6;
}
__device__ __forceinline__ bool kind_is_linear(int k) { return k >= GHOST_KIND_LINEAR_EQ && k <= GHOST_KIND_LINEAR_G; }
__device__ __forceinline__ int binom2(int n) { return (n * (n - 1)) >> 1; } // all_different.cpp:38-47

// ---------------------------------------------------------------------------------------------
// AllDifferentAbsDiff (extension kind): delta of changing scope position pa to na (and pb to nb
// when pb >= 0): recount the <= 4 neighbouring |differences|.  Kept out of line: only models that
// use the extension pay for it.
// ---------------------------------------------------------------------------------------------
// The two pair-difference kinds share one code path: AllDifferentAbsDiff is (lag 1, absolute values),
// AllDifferentLagDiff (lag = its parameter, signed differences x[p+lag] - x[p]).
__device__ __forceinline__ int pair_value(bool absolute, int first, int second) { return absolute ? abs(second - first) : second - first; }
// Delta of changing scope position pa to na and, when pb >= 0, pb to nb.  At most four pairs (p, p + lag) are touched:
// those starting at pa - lag, pa, pb - lag, pb.  Their old differences are taken out of the counts and the new ones put
// in ONE AT A TIME: removing a value whose count is c lowers sum C(count,2) by c - 1, adding one raises it by c, the
// effective count being the table's minus the removals so far plus the additions so far of that value — all in
// registers, no scratch arrays.  A pair that does not exist (or is the duplicate that arises when |pa - pb| = lag) is
// replaced by "remove 0, add 0", which cancels.
template <class XT>
__device__ __noinline__ int absdiff_delta(const int32_t* blob, int pd_desc_sec, int c2v_sec, const XT* x, const unsigned* cnt32,
                                          int c, int pa, int na, int pb, int nb)
{
	// one 128-bit record per constraint (lower.cpp, pd_desc): scope begin, scope length, count-table base, lag + flags
	const int4 desc = __ldg(reinterpret_cast<const int4*>(blob + pd_desc_sec) + c);
	const int b = desc.x, n = desc.y;
	const unsigned short* cnt = reinterpret_cast<const unsigned short*>(cnt32) + desc.z;
	const int lag = desc.w & 0xFFFF;
	const bool absolute = (desc.w >> 30) & 1, identity = desc.w < 0;   // identity: scope entry p is variable p (no lookup)
	const int32_t* scope = blob + c2v_sec + b;
	int first[4] = { pa - lag, pa, pb - lag, pb };
	if (pb < 0) { first[2] = -1; first[3] = -1; }
	else { if (pb - lag == pa) first[2] = -1; if (pa - lag == pb) first[0] = -1; }
	int ko[4], kn[4];
#pragma unroll
	for (int z = 0; z < 4; ++z) {
		const int p = first[z];
		ko[z] = 0; kn[z] = 0;
		if (p >= 0 && p + lag < n) {
			const int oa = identity ? x[p] : x[__ldg(scope + p)], ob = identity ? x[p + lag] : x[__ldg(scope + p + lag)];
			const int va = (p == pa) ? na : (p == pb ? nb : oa);
			const int vb = (p + lag == pa) ? na : (p + lag == pb ? nb : ob);
			ko[z] = pair_value(absolute, oa, ob);
			kn[z] = pair_value(absolute, va, vb);
		}
	}
	int diff = 0;
#pragma unroll
	for (int i = 0; i < 4; ++i) {
		int ceff = cnt[ko[i]];
#pragma unroll
		for (int j = 0; j < i; ++j) ceff -= ko[j] == ko[i] ? 1 : 0;
		diff -= ceff - 1;
	}
#pragma unroll
	for (int i = 0; i < 4; ++i) {
		int ceff = cnt[kn[i]];
#pragma unroll
		for (int j = 0; j < 4; ++j) ceff -= ko[j] == kn[i] ? 1 : 0;
#pragma unroll
		for (int j = 0; j < i; ++j) ceff += kn[j] == kn[i] ? 1 : 0;
		diff += ceff;
	}
	return diff;
}
template <class WS>
__device__ __forceinline__ int absdiff_delta(const MV& mv, const WS& s, int c, int pa, int na, int pb, int nb)
{
	return absdiff_delta(mv.m.blob, mv.m.o.pd_desc, mv.m.o.c2v, s.x, s.cnt32, c, pa, na, pb, nb);
}

// AllEqual (src/global_constraints/all_equal.cpp:47-95): error = scope size - largest count; the delta of
// a change is (largest count before) - (largest count after).  The constraint's state word keeps the
// largest count (low half) and the largest count once ONE holder of it is set aside (high half:
// equal to it when two values share the lead), so the largest count after moving one variable from
// value a to value d != a is max(cnt[d] + 1, base) with base = the lead unless a held it alone.
__device__ __forceinline__ int alleq_max1(int state) { return state & 0xFFFF; }
__device__ __forceinline__ int alleq_base(int state, int cnt_a)
{
	const int max1 = state & 0xFFFF, max2 = state >> 16;
	return cnt_a == max1 ? max(max2, max1 - 1) : max1;
}
template <class WS>
__device__ __forceinline__ int alleq_single_delta(const WS& s, int state, int tab, int a, int d)
{
	return alleq_max1(state) - max(s.cnt(tab + d) + 1, alleq_base(state, s.cnt(tab + a)));
}
// Recompute the state word and the error of AllEqual constraint c from its count table (one warp).
template <int T, class WS>
__device__ inline void alleq_refresh(const DevModel& m, const WS& s, int c)
{
	if (Group<T>::warp() != 0) return;
	const int lane = threadIdx.x & 31;
	const int lo = __ldg(m.blob + m.o.cnt_lo + c), n = __ldg(m.blob + m.o.cnt_n + c);
	int m1 = 0, k1 = 0, m2 = 0; // largest count, number of values holding it, largest count below it
	for (int t = lane; t < n; t += 32) {
		const int v = s.cnt(lo + t);
		if (v > m1) { m2 = m1; m1 = v; k1 = 1; } else if (v == m1) ++k1; else if (v > m2) m2 = v;
	}
#pragma unroll
	for (int off = 16; off > 0; off >>= 1) {
		const int om1 = __shfl_xor_sync(0xffffffffu, m1, off), ok1 = __shfl_xor_sync(0xffffffffu, k1, off), om2 = __shfl_xor_sync(0xffffffffu, m2, off);
		if (om1 == m1) { k1 += ok1; m2 = max(m2, om2); }
		else if (om1 > m1) { m2 = max(m1, om2); m1 = om1; k1 = ok1; }
		else m2 = max(m2, om1);
	}
	if (lane == 0) {
		const int scope = __ldg(m.blob + m.o.c2v_off + c + 1) - __ldg(m.blob + m.o.c2v_off + c);
		s.T_of(c) = m1 | ((k1 > 1 ? m1 : m2) << 16);
		s.err_of(c) = scope - m1;
	}
}

// AllDifferentOffset, two variables of the scope changing at once (a swap of a permutation problem): the counts
// of up to four shifted values move; the delta is the true change of sum C(count, 2), each distinct key once.
// (For a stock AllDifferent the swap leaves every count alone: delta 0, no call.)
template <class WS>
__device__ __forceinline__ int alldiff_four_key_delta(const WS& s, int k0, int k1, int k2, int k3)
{
	// k0 -1, k1 +1, k2 -1, k3 +1 (indices into the count table)
	const int keys[4] = { k0, k1, k2, k3 };
	int diff = 0;
#pragma unroll
	for (int i = 0; i < 4; ++i) {
		bool seen = false;
		int total = 0;
#pragma unroll
		for (int j = 0; j < 4; ++j) {
			if (keys[j] == keys[i]) { total += (j & 1) ? 1 : -1; if (j < i) seen = true; }
		}
		if (!seen && total != 0) { const int cur = s.cnt(keys[i]); diff += binom2(cur + total) - binom2(cur); }
	}
	return diff;
}

// ---------------------------------------------------------------------------------------------
// Delta of ONE constraint for a change of variable va (row entry second field = ea) to na, and,
// when vb >= 0, of vb (second field eb) to nb.  Restates optional_delta_error of the stock
// constraints (all_different.cpp:77-100, linear_equation.cpp:59-69, fix_value.cpp:54-64) in
// integers.  Used on the update path; the hot evaluation path uses the hoisted forms below and is
// checked against this one by the trajectory tests.
// ---------------------------------------------------------------------------------------------
template <class WS>
__device__ inline int constraint_delta(const MV& mv, const WS& s, int c, int va, int ea, int na, int vb, int eb, int nb)
{
	int kind = mv.kind(c);
	if (feat_on<WS>(GB_FEAT_ALLDIFF) && kind == GHOST_KIND_ALL_DIFFERENT) {
		if (vb >= 0) {
			// two changes; in the search these are always the swap {va,vb} -> {x[vb],x[va]}: no change
			int oa = s.x[va], ob = s.x[vb];
			if (na == ob && nb == oa) return 0;
			// general two-change case, reference formula (counts both (old,new) pairs)
			int tab = mv.cnt_tab(c);
			int keys[4] = { oa, na, ob, nb }, net[4] = { -1, +1, -1, +1 };
			int diff = 0;
			for (int i = 0; i < 4; ++i) {
				int total = 0;
				for (int j = 0; j < 4; ++j) if (keys[j] == keys[i]) total += net[j];
				int cur = s.cnt(tab + keys[i]);
				if (total != 0) diff += binom2(cur + total) - binom2(cur);
			}
			return diff;
		}
		int o = s.x[va];
		if (na == o) return 0;
		int tab = mv.cnt_tab(c);
		return s.cnt(tab + na) - (s.cnt(tab + o) - 1);
	}
	if (kind_is_linear(kind)) {
		int t = s.T_of(c) + ea * (na - s.x[va]);
		if (vb >= 0) t += eb * (nb - s.x[vb]);
		return linear_error_f<WS>(linear_w_of_kind(kind), t) - s.err_of(c);
	}
	if (feat_on<WS>(GB_FEAT_FIX) && kind == GHOST_KIND_FIX_VALUE) {
		int k = mv.rhs(c);
		int d = abs(na - k) - abs(s.x[va] - k);
		if (vb >= 0) d += abs(nb - k) - abs(s.x[vb] - k);
		return d;
	}
	if (feat_on<WS>(GB_FEAT_OFFSET) && kind == GHOST_KIND_ALL_DIFFERENT_OFFSET) {
		// ea / eb = the variables' offsets; the count table is indexed by x + offset
		const int tab = mv.cnt_tab(c), oa = s.x[va];
		if (vb >= 0) return alldiff_four_key_delta(s, tab + ea + oa, tab + ea + na, tab + eb + s.x[vb], tab + eb + nb);
		if (na == oa) return 0;
		return s.cnt(tab + ea + na) - (s.cnt(tab + ea + oa) - 1);
	}
	if (feat_on<WS>(GB_FEAT_PAIRDIFF) && (kind == GHOST_KIND_ALL_DIFFERENT_ABS_DIFF || kind == GHOST_KIND_ALL_DIFFERENT_LAG_DIFF)) return absdiff_delta(mv, s, c, ea, na, vb >= 0 ? eb : -1, nb);
	if (feat_on<WS>(GB_FEAT_ALLEQ) && kind == GHOST_KIND_ALL_EQUAL) {
		const int tab = mv.cnt_tab(c), oa = s.x[va];
		if (vb >= 0) {
			const int ob = s.x[vb];
			if (na == ob && nb == oa) return 0; // a swap inside the scope leaves the counts alone
			// general two-change case (never produced by the search): rescan the table with the changes applied
			const int lo = mv.ld(mv.m.o.cnt_lo, c), n = mv.ld(mv.m.o.cnt_n, c);
			int best = 0;
			for (int t = 0; t < n; ++t) {
				const int key = lo + t - tab;
				best = max(best, s.cnt(lo + t) - (key == oa) - (key == ob) + (key == na) + (key == nb));
			}
			return alleq_max1(s.T_of(c)) - best;
		}
		if (na == oa) return 0;
		return alleq_single_delta(s, s.T_of(c), tab, oa, na);
	}
	return 0; // PURE_OPTIMIZATION
}

// ---------------------------------------------------------------------------------------------
// (Re)compute every constraint error, cache and the projection from the assignment in s.x:
// SearchUnit::compute_constraints_errors (search_unit.hpp:434-447) + Constraint::error's cache
// rebuild (all_different.cpp:59-75, linear_equation.cpp:50-57, fix_value.cpp:46-52) +
// AdaptiveSearchErrorProjection::compute_variable_errors (…projection_algorithm.cpp:40-49).
// Flat over the scope entries with a warp-segmented reduction per constraint.
// Returns the total satisfaction error.  Ends with a group barrier.
// ---------------------------------------------------------------------------------------------
// EQ: every live constraint is a LinearEquationEq (plus, possibly, the static AllDifferent of a
// permutation problem, whose entries carry coefficient 0): no kind lookups, no count tables.
// F32: the linear constraints' T is left as an fp32 number (the fp32 evaluation below reads it as such).
template <int T, bool EQ, bool F32, class WS>
__device__ __forceinline__ int recompute_all_impl(const DevModel& m, const WS& s, int* red, const bool PROJECT)
{
	typedef Group<T> G;
	MV mv(m);
	const int tid = G::tid(), lane = threadIdx.x & 31;
	for (int c = tid; c < m.C; c += T) { s.cs[c] = make_int2(0, 0); s.cem[c] = make_int2(0, 0); }
	if (!EQ) for (int t = tid; t < (m.cnt_total + 1) / 2; t += T) s.cnt32[t] = 0;
	G::sync();
	const int nnz_pad = (m.nnz + 31) & ~31;
	for (int i = tid; i < nnz_pad; i += T) {
		int c = -1, val = 0, kind = -1;
		if (EQ) {
			if (i < m.nnz) { c = mv.ld(m.o.ent_ctr, i); kind = GHOST_KIND_LINEAR_EQ; val = mv.ld(m.o.ent_coef, i) * s.x[mv.scope_var(i)]; }
		} else if (i < m.nnz) {
			c = mv.ld(m.o.ent_ctr, i);
			kind = mv.kind(c);
			int u = mv.scope_var(i);
			int xu = s.x[u];
			if (kind_is_linear(kind)) val = mv.ld(m.o.ent_coef, i) * xu;
			else if (kind == GHOST_KIND_FIX_VALUE) val = abs(xu - mv.rhs(c));
			else if (kind == GHOST_KIND_ALL_DIFFERENT || kind == GHOST_KIND_ALL_EQUAL) s.cnt_add(mv.cnt_tab(c) + xu, 1);
			else if (kind == GHOST_KIND_ALL_DIFFERENT_OFFSET) s.cnt_add(mv.cnt_tab(c) + xu + mv.ld(m.o.ent_coef, i), 1);
			else if (kind == GHOST_KIND_ALL_DIFFERENT_ABS_DIFF || kind == GHOST_KIND_ALL_DIFFERENT_LAG_DIFF) {
				const bool absolute = kind == GHOST_KIND_ALL_DIFFERENT_ABS_DIFF;
				const int lag = absolute ? 1 : mv.rhs(c);
				if (i + lag < mv.scope_begin(c + 1)) s.cnt_add(mv.cnt_tab(c) + pair_value(absolute, xu, s.x[mv.scope_var(i + lag)]), 1);
			}
		}
		// inclusive segmented scan by constraint id
#pragma unroll
		for (int off = 1; off < 32; off <<= 1) {
			int c2 = __shfl_up_sync(0xffffffffu, c, off);
			int v2 = __shfl_up_sync(0xffffffffu, val, off);
			if (lane >= off && c2 == c) val += v2;
		}
		int cnext = __shfl_down_sync(0xffffffffu, c, 1);
		bool tail = (lane == 31) || (cnext != c);
		if (c >= 0 && tail && val != 0) {
			if (kind_is_linear(kind)) atomicAdd(&s.T_of(c), val);
			else if (kind == GHOST_KIND_FIX_VALUE) atomicAdd(&s.err_of(c), val);
		}
	}
	G::sync();
	for (int c = tid; c < m.C; c += T) {
		int kind = mv.kind(c);
		if (kind_is_linear(kind)) {
			int t = s.T_of(c) - mv.rhs(c);
			s.T_of(c) = F32 ? __float_as_int((float)t) : t;
			s.err_of(c) = linear_error_w(linear_w_of_kind(kind), t);
		} else if (kind == GB_KIND_STATIC) s.err_of(c) = mv.rhs(c);
	}
	if (!EQ) for (int t = tid; t < m.cnt_total; t += T) {
		int n = s.cnt(t);
		if (n > 1) {
			const int owner = mv.ld(m.o.cnt_owner, t);
			if (!m.has_all_equal || mv.kind(owner) != GHOST_KIND_ALL_EQUAL) atomicAdd(&s.err_of(owner), binom2(n));
		}
	}
	if (!EQ && m.has_all_equal) {
		G::sync();
		for (int c = 0; c < m.C; ++c) if (mv.kind(c) == GHOST_KIND_ALL_EQUAL) alleq_refresh<T>(m, s, c);
	}
	G::sync();
	int part = 0;
	for (int c = tid; c < m.C; c += T) part += s.err_of(c);
	int total = G::sum(part, red);
	if (PROJECT) for (int v = tid; v < m.V; v += T) {
		int e = 0;
		for (int r = mv.row_begin(v), re = mv.row_begin(v + 1); r < re; ++r) e += s.err_of(mv.row_entry(r).x);
		s.err_var[scan_pos(m, v)] = e;
	}
	G::sync();
	return total;
}
template <int T, bool PROJECT, bool EQ = false, bool F32 = false, class WS>
__device__ __forceinline__ int recompute_all(const DevModel& m, const WS& s, int* red) { return recompute_all_impl<T, EQ, F32>(m, s, red, PROJECT); }

// ---------------------------------------------------------------------------------------------
// Candidate sets are kept as (key, bitmask) per chunk of 32 slots: the set is every set bit of
// the chunks whose key equals the global best key.  pick_kth returns the k-th member in ascending
// slot order — the order of the reference's vectors (variable_candidates in ascending id,
// adaptive_search_variable_candidates_heuristic.cpp:43-57; candidate_values in ascending key,
// adaptive_search_value_heuristic.cpp:60-66).  Executed redundantly by every warp.
// ---------------------------------------------------------------------------------------------
__device__ inline int chunks_count(const int* key, const unsigned* bits, int nchunks, int best)
{
	const int lane = threadIdx.x & 31;
	int n = 0;
	for (int ch = lane; ch < ((nchunks + 31) & ~31); ch += 32) {
		int c = (ch < nchunks && key[ch] == best) ? __popc(bits[ch]) : 0;
		n += __reduce_add_sync(0xffffffffu, c);
	}
	return n;
}
__device__ inline int pick_kth(const int* key, const unsigned* bits, int nchunks, int best, int k)
{
	const int lane = threadIdx.x & 31;
	int result = -1;
	for (int base = 0; base < nchunks; base += 32) {
		int ch = base + lane;
		unsigned b = (ch < nchunks && key[ch] == best) ? bits[ch] : 0u;
		int c = __popc(b);
		int incl = c;
#pragma unroll
		for (int off = 1; off < 32; off <<= 1) {
			int t = __shfl_up_sync(0xffffffffu, incl, off);
			if (lane >= off) incl += t;
		}
		int total = __shfl_sync(0xffffffffu, incl, 31);
		if (k < total) {
			unsigned owner = __ballot_sync(0xffffffffu, incl > k);
			int src = __ffs(owner) - 1;
			int excl = __shfl_sync(0xffffffffu, incl - c, src);
			unsigned wb = __shfl_sync(0xffffffffu, b, src);
			int bit = __fns(wb, 0, k - excl + 1);
			result = (base + src) * 32 + bit;
			break;
		}
		k -= total;
	}
	return result;
}
// min (or max) of the chunk keys; redundantly per warp
__device__ inline int chunks_min(const int* key, int nchunks)
{
	const int lane = threadIdx.x & 31;
	int best = GB_INT_MAX;
	for (int ch = lane; ch < ((nchunks + 31) & ~31); ch += 32) {
		int k = ch < nchunks ? key[ch] : GB_INT_MAX;
		best = min(best, __reduce_min_sync(0xffffffffu, k));
	}
	return best;
}
__device__ inline int chunks_max(const int* key, int nchunks)
{
	const int lane = threadIdx.x & 31;
	int best = -GB_INT_MAX;
	for (int ch = lane; ch < ((nchunks + 31) & ~31); ch += 32) {
		int k = ch < nchunks ? key[ch] : -GB_INT_MAX;
		best = max(best, __reduce_max_sync(0xffffffffu, k));
	}
	return best;
}

// ---------------------------------------------------------------------------------------------
// Chunk summaries of a walker owned by ONE warp (T == 32; at most 32 chunks, i.e. 1024 slots):
// lane ch keeps (key, bits) of chunk ch in two registers instead of shared memory.
// ---------------------------------------------------------------------------------------------
struct ChunkRegs { int key; unsigned bits; };

__device__ __forceinline__ int regs_count(const ChunkRegs& q, int best)
{
	return __reduce_add_sync(0xffffffffu, q.key == best ? __popc(q.bits) : 0);
}
static __device__ GB_OUTLINE int regs_pick_kth_impl(int key, unsigned bits, int best, int k, int n)
{
	const int lane = threadIdx.x & 31;
	const unsigned b = key == best ? bits : 0u;
	if (n == 1) { // a single member (the usual case for the min-conflict value): no prefix sum needed
		const int only = __ffs(__ballot_sync(0xffffffffu, b != 0u)) - 1;
		return only * 32 + __ffs(__shfl_sync(0xffffffffu, b, only)) - 1;
	}
	const int c = __popc(b);
	int incl = c;
#pragma unroll
	for (int off = 1; off < 32; off <<= 1) {
		int t = __shfl_up_sync(0xffffffffu, incl, off);
		if (lane >= off) incl += t;
	}
	const unsigned owner = __ballot_sync(0xffffffffu, incl > k);
	const int src = __ffs(owner) - 1;
	const int excl = __shfl_sync(0xffffffffu, incl - c, src);
	const unsigned wb = __shfl_sync(0xffffffffu, b, src);
	// the (k - excl)-th set bit of wb, by the lanes (lane l: "bit l is set and as many set bits lie below it"); __fns is a
	// software loop of several dozen instructions
	const bool me = ((wb >> lane) & 1u) && __popc(wb & ((1u << lane) - 1u)) == k - excl;
	return src * 32 + __ffs(__ballot_sync(0xffffffffu, me)) - 1;
}
// n = size of the set (known to the caller)
__device__ __forceinline__ int regs_pick_kth(const ChunkRegs& q, int best, int k, int n) { return regs_pick_kth_impl(q.key, q.bits, best, k, n); }

// One chunk of the variable scan: the chunk's largest eligible error and the lanes that hold it.
template <int T, class WS>
__device__ __forceinline__ void commit_scan_chunk(const WS& s, ChunkRegs& q, int ch, int e)
{
	const int c = __reduce_max_sync(0xffffffffu, e);
	const unsigned b = __ballot_sync(0xffffffffu, e == c);   // meaningless when c < 0: such a key never wins
	if (T == 32) { if ((threadIdx.x & 31) == ch) { q.key = c; q.bits = b; } }
	else if ((threadIdx.x & 31) == 0) { s.cand_key[ch] = c; s.cand_bits[ch] = b; }
}

// ---------------------------------------------------------------------------------------------
// AdaptiveSearchVariableCandidatesHeuristic::compute_variable_candidates (…candidates_heuristic.cpp:
// 38-60) fused with the tabu count of the reset test (search_unit.hpp:761-763): one pass over the
// variables.  Leaves the chunk summaries in `q` (T == 32) or in (cand_key, cand_bits); returns ncand,
// writes worst error and tabu count.  Contains group barriers when T > 32.
// Full blocks of 128 variables are read with one 128-bit and one 64-bit load per lane (see scan_pos);
// both arrays are padded to a multiple of 32 entries (padding: always tabu), so no chunk needs a
// bounds test.
// ---------------------------------------------------------------------------------------------
// One pass over the variables: chunk summaries into q / (cand_key, cand_bits), returns this lane's tabu count.
// USE_MASK: variables that appear in no constraint are not candidates (…candidates_heuristic.cpp:45-46) and the
// model has some; a template axis because the predicated form costs every model five instructions per block.
// TQ: tabu variables carry err_var - GB_TABU_OFFSET (see "tabu as an offset" below): the pass reads no tabu entry, counts nothing
template <int T, bool USE_MASK, bool TQ = false, class WS>
__device__ __forceinline__ int scan_pass(const DevModel& m, const WS& s, int nchunks_v, int local_moves, ChunkRegs& q)
{
	typedef Group<T> G;
	MV mv(m);
	const int lane = threadIdx.x & 31;
	const int nblocks = m.v128 >> 7;
	int ntabu = 0; // per lane
	const int4* e4 = reinterpret_cast<const int4*>(s.err_var);
	const uint2* t4 = reinterpret_cast<const uint2*>(s.tabu);
#ifndef GB_UNROLL2   // not unrolled: the loop body has to stay small enough for the per-scheduler instruction cache (measured)
#pragma unroll 1
#endif
	for (int b = G::warp(); b < nblocks; b += G::NW) {
		const int4 e = e4[b * 32 + lane];
		int e0 = e.x, e1 = e.y, e2 = e.z, e3 = e.w;
		if (!TQ) {
			const uint2 t = t4[b * 32 + lane];
			const bool k0 = (int)(t.x & 0xFFFFu) > local_moves, k1 = (int)(t.x >> 16) > local_moves;
			const bool k2 = (int)(t.y & 0xFFFFu) > local_moves, k3 = (int)(t.y >> 16) > local_moves;
			ntabu += (k0 ? 1 : 0) + (k1 ? 1 : 0) + (k2 ? 1 : 0) + (k3 ? 1 : 0);
			e0 = k0 ? -1 : e.x; e1 = k1 ? -1 : e.y; e2 = k2 ? -1 : e.z; e3 = k3 ? -1 : e.w;
		}
		if (USE_MASK) {
			const int4 h = __ldg(reinterpret_cast<const int4*>(m.blob + m.o.has_ctr) + b);
			if (!(((unsigned)h.x >> lane) & 1u)) e0 = -1;
			if (!(((unsigned)h.y >> lane) & 1u)) e1 = -1;
			if (!(((unsigned)h.z >> lane) & 1u)) e2 = -1;
			if (!(((unsigned)h.w >> lane) & 1u)) e3 = -1;
		}
#ifndef GB_SCAN_SPLIT
		if (T == 32) {
#else
		if (false) {
#endif
			// ONE warp reduction for the block's four chunks: a chunk that does not attain the block's largest
			// error gets an empty mask (only chunks whose key equals the global maximum matter); lane 0 stages
			// the four {key, bits} summaries, the owning lanes pick them up after the loop
			const int c = __reduce_max_sync(0xffffffffu, max(max(e0, e1), max(e2, e3)));
			const unsigned b0 = __ballot_sync(0xffffffffu, e0 == c), b1 = __ballot_sync(0xffffffffu, e1 == c);
			const unsigned b2 = __ballot_sync(0xffffffffu, e2 == c), b3 = __ballot_sync(0xffffffffu, e3 == c);
			if (lane == 0) {
				int4* stage = reinterpret_cast<int4*>(s.val_key);
				stage[2 * b] = make_int4(c, (int)b0, c, (int)b1);
				stage[2 * b + 1] = make_int4(c, (int)b2, c, (int)b3);
			}
		} else {
			commit_scan_chunk<T>(s, q, 4 * b + 0, e0);
			commit_scan_chunk<T>(s, q, 4 * b + 1, e1);
			commit_scan_chunk<T>(s, q, 4 * b + 2, e2);
			commit_scan_chunk<T>(s, q, 4 * b + 3, e3);
		}
	}
	for (int ch = 4 * nblocks + G::warp(); ch < nchunks_v; ch += G::NW) { // the tail, natural order
		const int i = ch * 32 + lane;
		int e = s.err_var[i];
		if (!TQ) {
			const bool k = (int)s.tabu[i] > local_moves;
			ntabu += k ? 1 : 0;
			if (k) e = -1;
		}
		if (USE_MASK && !(((unsigned)mv.ld(m.o.has_ctr, ch) >> lane) & 1u)) e = -1;
#ifndef GB_SCAN_SPLIT
		if (T == 32) {
			const int c = __reduce_max_sync(0xffffffffu, e);
			const unsigned bb = __ballot_sync(0xffffffffu, e == c);
			if (lane == 0) reinterpret_cast<int2*>(s.val_key)[ch] = make_int2(c, (int)bb);
		} else
#endif
			commit_scan_chunk<T>(s, q, ch, e);
	}
#ifndef GB_SCAN_SPLIT
	if (T == 32) {
#else
	if (false) {
#endif
		__syncwarp();
		const int2 mine = reinterpret_cast<const int2*>(s.val_key)[lane];   // staged {key, bits}; stale words beyond the last chunk
		q.key = lane < nchunks_v ? mine.x : -1;
		q.bits = lane < nchunks_v ? (unsigned)mine.y : 0u;
	}
	return ntabu;
}

// ---------------------------------------------------------------------------------------------
// Tabu as an offset (the fp32 kernels, whose projected errors lower.cpp bounds below 2^24): while a variable is tabu
// its projected error carries -GB_TABU_OFFSET, so the candidate scan is a plain arg-max that reads no tabu entry — a
// tabu variable's key is negative, which the scan already treats as "not a candidate".  The projection updates are
// additive and never notice.  Which variables carry the offset is a bit mask in registers (lane ch = variables
// 32*ch .. 32*ch+31): marking sets the bit and subtracts the offset, tabu_expire() tests only the marked variables
// (a few dozen at most: the reset threshold bounds them) against the move counter, gives the offset back and clears
// the bit; the population of the mask is the tabu count of the reset test (search_unit.hpp:761-763).  The tabu array
// itself stays as it is (expiry times; rebuilt into a mask when a launch resumes).
// ---------------------------------------------------------------------------------------------
#define GB_TABU_OFFSET (1 << 30)
template <class WS>
__device__ __forceinline__ int tabu_expire(const DevModel& m, const WS& s, int local_moves, unsigned& mask)
{
	const int lane = threadIdx.x & 31;
	unsigned pending = mask;
	while (__any_sync(0xffffffffu, pending != 0u)) {
		if (pending != 0u) {
			const int bit = __ffs(pending) - 1;
			pending &= pending - 1u;
			const int pos = scan_pos(m, lane * 32 + bit);
			if ((int)s.tabu[pos] <= local_moves) { mask &= ~(1u << bit); atomicAdd(&s.err_var[pos], GB_TABU_OFFSET); }
		}
	}
	__syncwarp();
	return __reduce_add_sync(0xffffffffu, __popc(mask));
}
// mark variable v tabu (its tabu entry has just been written by lane 0); tenure 0 marks nothing
template <class WS>
__device__ __forceinline__ void tabu_mark(const DevModel& m, const WS& s, int v, int tenure, unsigned& mask)
{
	if (tenure <= 0) return;
	const int lane = threadIdx.x & 31;
	if (lane == (v >> 5)) mask |= 1u << (v & 31);
	if (lane == 0) atomicSub(&s.err_var[scan_pos(m, v)], GB_TABU_OFFSET);   // (atomic: the projection scatter of a move may follow without a barrier)
}
// a launch resumes: the mask from the tabu entries (the offsets are in the persisted errors already)
template <class WS>
__device__ inline unsigned tabu_mask_rebuild(const DevModel& m, const WS& s, int nchunks_v, int local_moves)
{
	const int lane = threadIdx.x & 31;
	unsigned mine = 0u;
	for (int ch = 0; ch < nchunks_v; ++ch) {
		const int v = ch * 32 + lane;
		const unsigned b = __ballot_sync(0xffffffffu, v < m.V && (int)s.tabu[scan_pos(m, v)] > local_moves);
		if (lane == ch) mine = b;
	}
	return mine;
}

template <int T, bool TQ = false, class WS>
__device__ inline int scan_candidates(const DevModel& m, const WS& s, int nchunks_v, int local_moves,
                                      bool free_vars_ok, int* red, int* worst_out, int* ntabu_out, ChunkRegs& q)
{
	typedef Group<T> G;
	const bool use_mask = !free_vars_ok && !m.all_have_ctr;
	q.key = -1; q.bits = 0u;
	int ntabu = use_mask ? scan_pass<T, true, TQ>(m, s, nchunks_v, local_moves, q) : scan_pass<T, false, TQ>(m, s, nchunks_v, local_moves, q);
	int worst, ncand;
	if (T == 32 && TQ) {
		worst = __reduce_max_sync(0xffffffffu, q.key);
		ncand = regs_count(q, worst);
		if (worst < 0) ncand = 0;
		*worst_out = worst;   // (the tabu count is the caller's: the population of its mask)
		return ncand;
	}
	if (T == 32) {
		ntabu = __reduce_add_sync(0xffffffffu, ntabu);
		worst = __reduce_max_sync(0xffffffffu, q.key);
		ncand = regs_count(q, worst);
	} else {
		ntabu = G::sum(ntabu, red); // barrier inside: the summaries are visible afterwards
		worst = chunks_max(s.cand_key, nchunks_v);
		ncand = worst >= 0 ? chunks_count(s.cand_key, s.cand_bits, nchunks_v, worst) : 0;
	}
	ntabu -= nchunks_v * 32 - m.V; // the padding entries are tabu by construction
	if (worst < 0) ncand = 0;
	*worst_out = worst;
	*ntabu_out = ntabu;
	return ncand;
}

// ---------------------------------------------------------------------------------------------
// Hoist the per-iteration invariants of the selected variable's constraints (these
// are not charged per candidate).  One int4 per hoisted constraint:
//   .x = type (GB_CLS_*) | w << 4 | constraint id << 8
//   linear : .y = a_v, .z = T - a_v*x_v            (error term: linear_error_w(w, .z + .y*d))
//   alldiff: .y = count-table base, .z = cnt[x_v]-1
//   fix    : .y = k
//   absdiff: .y = scope position of v
// The constant part of every candidate's sum (- current errors, - (cnt[x_v]-1), - |x_v-k|) is
// returned in *bias.  In permutation mode mark[c] = entry index + 1; constraints over ALL variables
// (a swap inside them never changes an AllDifferent) are marked GB_MARK_SKIP and not hoisted.
// Returns the number of hoisted entries.  The caller must barrier before reading.
// ---------------------------------------------------------------------------------------------
template <int T, class WS>
__device__ inline int hoist_variable(const DevModel& m, const WS& s, int v, int* bias_out)
{
	MV mv(m);
	const int lane = threadIdx.x & 31;
	const int2* row = mv.ell_row(v);
	const int xv = s.x[v];
	int nh = 0, bias = 0;
	const bool writer = Group<T>::warp() == 0;
	// one lane per row entry (live entries come first in a row, so entry k is hoisted entry k); the count
	// and the bias are warp reductions, every warp of a wider group computes them, warp 0 writes
	for (int base = 0; base < m.ell_stride; base += 32) {
		const int k = base + lane;
		int2 e = make_int2(m.C | (GB_CLS_ZERO << 20), 0);
		if (k < m.ell_stride) e = __ldg(row + k);
		const int c = ell_ctr(e.x);
		const bool valid = c != m.C;
		int contrib = 0;
		if (valid) {
			const int cls = ell_cls(e.x);
			int4 h = make_int4(cls | (ell_w(e.x) << 4) | (c << 8), 0, 0, ell_universal(e.x) ? 1 : 0);   // .w: the scope is every variable
			if (cls == GB_CLS_LINEAR) {
				h.y = e.y; h.z = s.T_of(c) - e.y * xv;
				contrib = -s.err_of(c);
			} else if ((feat_on<WS>(GB_FEAT_ALLDIFF) && cls == GB_CLS_ALLDIFF) || (feat_on<WS>(GB_FEAT_OFFSET) && cls == GB_CLS_ALLDIFF_OFF)) {
				const int tab = mv.cnt_tab(c) + e.y;   // e.y: the variable's offset (0 for a stock AllDifferent)
				const int sub = s.cnt(tab + xv) - 1;
				h.y = tab; h.z = sub;
				contrib = -sub;
			} else if (feat_on<WS>(GB_FEAT_FIX) && cls == GB_CLS_FIX) {
				const int kk = mv.rhs(c);
				h.y = kk;
				contrib = -abs(xv - kk);
			} else if (feat_on<WS>(GB_FEAT_PAIRDIFF) && cls == GB_CLS_ABSDIFF) {
				h.y = e.y;
			} else if (feat_on<WS>(GB_FEAT_ALLEQ) && cls == GB_CLS_ALLEQ) {
				const int tab = mv.cnt_tab(c), state = s.T_of(c);
				h.y = tab; h.z = alleq_base(state, s.cnt(tab + xv));
				contrib = alleq_max1(state);
			}
			if (writer) {
				s.hoist[k] = h;
				if (m.permutation) {
					// mark + the selected variable's coefficient next to the constraint state: the swap
					// partner's loop needs no second lookup for a shared constraint
					s.mark_of(c) = k + 1; s.av_of(c) = e.y;
				}
			}
		}
		nh += __popc(__ballot_sync(0xffffffffu, valid));
		bias += __reduce_add_sync(0xffffffffu, contrib);
	}
	*bias_out = bias;
	return nh;
}
template <int T, class WS>
__device__ inline void unmark_variable(const DevModel& m, const WS& s, int v)
{
	MV mv(m);
	const int2* row = mv.ell_row(v);
	for (int k = Group<T>::tid(); k < m.ell_stride; k += T) {
		int c = ell_ctr(__ldg(row + k).x);
		if (c != m.C) { s.av_of(c) = 0; s.mark_of(c) = 0; }
	}
}

// The same two steps when every live constraint is a LinearEquationEq: the packed row (lower.cpp) is
// read by one lane per entry; live entries come first in a row, so entry k is hoisted entry k.
template <int T, bool PERM, class WS>
__device__ inline int hoist_variable_eq(const DevModel& m, const WS& s, int v, int* bias_out)
{
	MV mv(m);
	const int lane = threadIdx.x & 31;
	const int xv = s.x[v];
	const int nrow = m.ell_eq_stride, dummy = m.C * 8;
	const unsigned* row = mv.eq_row(v);
	const bool writer = Group<T>::warp() == 0;
	int nh = 0, bias = 0;
	for (int base = 0; base < nrow; base += 32) {
		const int k = base + lane;
		const unsigned f = k < nrow ? __ldg(row + k) : (unsigned)dummy;
		const int off = (int)(f & 0xFFFFu), a = (int)f >> 16;
		const bool valid = off != dummy;
		int err = 0;
		if (valid) {
			const int c = off >> 3;
			const int t = s.T_of(c);
			err = s.err_of(c);
			if (writer) {
				s.hoist[k] = make_int4(GB_CLS_LINEAR | (3 << 4) | (c << 8), a, t - a * xv, 0);
				if (PERM) { s.mark_of(c) = k + 1; s.av_of(c) = a; }
			}
		}
		nh += __popc(__ballot_sync(0xffffffffu, valid));
		bias -= __reduce_add_sync(0xffffffffu, err);
	}
	*bias_out = bias;
	return nh;
}
template <int T, class WS>
__device__ inline void unmark_variable_eq(const DevModel& m, const WS& s, int v)
{
	MV mv(m);
	const unsigned* row = mv.eq_row(v);
	for (int k = Group<T>::tid(); k < m.ell_eq_stride; k += T) {
		const int c = (int)(__ldg(row + k) & 0xFFFFu) >> 3;
		s.av_of(c) = 0; s.mark_of(c) = 0; // the dummy entry is all zeros anyway
	}
}

// variable part of hoisted entry h when the selected variable alone takes value d (d != x_v);
// the constant part is in the bias
template <class WS>
__device__ __forceinline__ int hoisted_term(const MV& mv, const WS& s, const int4 h, int d)
{
	int type = h.x & 0xF;
	if (type == GB_CLS_LINEAR) return linear_error_f<WS>((h.x >> 4) & 7, h.z + h.y * d);
	if ((feat_on<WS>(GB_FEAT_ALLDIFF) && type == GB_CLS_ALLDIFF) || (feat_on<WS>(GB_FEAT_OFFSET) && type == GB_CLS_ALLDIFF_OFF)) return s.cnt(h.y + d);
	if (feat_on<WS>(GB_FEAT_FIX) && type == GB_CLS_FIX) return abs(d - h.y);
	if (feat_on<WS>(GB_FEAT_PAIRDIFF) && type == GB_CLS_ABSDIFF) return absdiff_delta(mv, s, h.x >> 8, h.y, d, -1, 0);
	if (feat_on<WS>(GB_FEAT_ALLEQ) && type == GB_CLS_ALLEQ) return -max(s.cnt(h.y + d) + 1, h.z);
	return 0;
}

// Cumulated delta of one candidate (search_unit.hpp:836-887 + std::accumulate of
// adaptive_search_value_heuristic.cpp:57-58).  slot = domain index (non-permutation) or partner
// variable (permutation).  Returns false if the slot is not a candidate.
template <bool PERM, bool EQ, class WS>
__device__ __forceinline__ bool candidate_sum(const DevModel& m, const MV& mv, const WS& s, int v, int xv, int nh, int bias,
                                              int slot, int nslots, int* out)
{
	if (slot >= nslots) return false;
	if (!PERM) {
		int d = mv.dom_value(v, slot);
		if (d == xv) return false;                       // search_unit.hpp:833
		int sum = bias;
		for (int k = 0; k < nh; ++k) sum += hoisted_term(mv, s, s.hoist[k], d);
		*out = sum;                                      // no constraint: 0.0 (search_unit.hpp:847-848)
		return true;
	}
	const int j = slot;
	if (j == v) return false;
	const int xj = s.x[j];
	if (xj == xv) return false;                          // search_unit.hpp:855-860
	if (!m.identical_domains && !(mv.in_domain(v, xj) && mv.in_domain(j, xv))) return false;
	int sum = bias;
	const int dx = xv - xj;
	const int2* row = mv.ell_row(j);
	for (int k = 0; k < nh; ++k) {
		const int4 h = s.hoist[k];
		// a pair-difference constraint over ALL variables holds the partner too: its single-variable term would only be
		// taken out again by the two-variable one below (two recounts out of three saved on a Costas model)
		if (feat_on<WS>(GB_FEAT_PAIRDIFF) && (h.x & 0xF) == GB_CLS_ABSDIFF && h.w) continue;
		sum += hoisted_term(mv, s, h, xj);
	}
	for (int k = 0; k < m.ell_stride; ++k) {
		int2 e = __ldg(row + k);
		int c = ell_ctr(e.x);
		if (c == m.C) break; // padding
		const int4 st = make_int4(s.T_of(c), 0, s.err_of(c), s.mark_of(c));
		int cls = ell_cls(e.x);
		if (st.w == 0) {
			// constraint of the partner only: {j} -> {x_v}  (search_unit.hpp:881-885)
			if (cls == GB_CLS_LINEAR) sum += linear_error_f<WS>(ell_w(e.x), st.x + e.y * dx) - st.z;
			else if ((feat_on<WS>(GB_FEAT_ALLDIFF) && cls == GB_CLS_ALLDIFF) || (feat_on<WS>(GB_FEAT_OFFSET) && cls == GB_CLS_ALLDIFF_OFF)) {
				int tab = mv.cnt_tab(c) + e.y;
				sum += s.cnt(tab + xv) - (s.cnt(tab + xj) - 1);
			} else if (feat_on<WS>(GB_FEAT_FIX) && cls == GB_CLS_FIX) {
				int kk = mv.rhs(c);
				sum += abs(xv - kk) - abs(xj - kk);
			} else if (feat_on<WS>(GB_FEAT_PAIRDIFF) && cls == GB_CLS_ABSDIFF) sum += absdiff_delta(mv, s, c, e.y, xv, -1, 0);
			else if (feat_on<WS>(GB_FEAT_ALLEQ) && cls == GB_CLS_ALLEQ) sum += alleq_single_delta(s, st.x, mv.cnt_tab(c), xj, xv);
		} else {
			// both variables in scope: {v,j} -> {x_j,x_v} (search_unit.hpp:872-874); replace the
			// single-variable delta already added from the hoisted entry
			const int4 h = s.hoist[st.w - 1];
			if (cls == GB_CLS_LINEAR) {
				int w = ell_w(e.x), t1 = h.z + h.y * xj;
				sum += linear_error_f<WS>(w, t1 + e.y * dx) - linear_error_f<WS>(w, t1);
			} else if (feat_on<WS>(GB_FEAT_ALLDIFF) && cls == GB_CLS_ALLDIFF) sum -= s.cnt(h.y + xj) - h.z;
			else if (feat_on<WS>(GB_FEAT_OFFSET) && cls == GB_CLS_ALLDIFF_OFF) {
				// the hoisted single-variable delta out, the swap's own delta in (h.y = table + offset of v)
				const int tabj = mv.cnt_tab(c) + e.y;
				sum += alldiff_four_key_delta(s, h.y + xv, h.y + xj, tabj + xj, tabj + xv) - (s.cnt(h.y + xj) - h.z);
			}
			else if (feat_on<WS>(GB_FEAT_FIX) && cls == GB_CLS_FIX) sum += abs(xv - h.y) - abs(xj - h.y);
			else if (feat_on<WS>(GB_FEAT_PAIRDIFF) && cls == GB_CLS_ABSDIFF) sum += absdiff_delta(mv, s, c, h.y, xj, e.y, xv) - (h.w ? 0 : absdiff_delta(mv, s, c, h.y, xj, -1, 0));
			else if (feat_on<WS>(GB_FEAT_ALLEQ) && cls == GB_CLS_ALLEQ) sum += max(s.cnt(h.y + xj) + 1, h.z) - alleq_max1(st.x); // the swap changes nothing
		}
	}
	*out = sum;
	return true;
}

// One chunk of the value evaluation: the chunk's smallest cumulated delta and the lanes that hold it
// (non-candidates carry INT_MAX; a chunk without candidates gets the key INT_MAX, which never wins).
template <int T, class WS>
__device__ __forceinline__ void commit_value_chunk(const WS& s, ChunkRegs& q, int ch, int sum)
{
	const int cmin = __reduce_min_sync(0xffffffffu, sum);
	const unsigned bits = __ballot_sync(0xffffffffu, sum == cmin);
	if (T == 32) { if ((threadIdx.x & 31) == ch) { q.key = cmin; q.bits = bits; } }
	else if ((threadIdx.x & 31) == 0) { s.val_key[ch] = cmin; s.val_bits[ch] = bits; }
}

// Evaluate every candidate of variable v: chunk minima + tie masks into `q` (T == 32: lane ch keeps
// chunk ch) or into (val_key, val_bits).  If out != nullptr also stores each candidate's sum
// (GHOST_B200_NOT_A_CANDIDATE otherwise).  Returns this lane's number of candidates.
// Specialised evaluation when every live constraint is a LinearEquationEq (linear_equation.cpp:59-69
// with linear_equation_eq.cpp:54-57): |T + a*delta| - error per constraint, no kind dispatch, the
// selected variable's (a, T - a*x_v) pairs held in registers, no end-of-row tests (the ELL rows and
// the assignment are padded to whole chunks), no divergent control flow.
// S = compile-time row length of the padded ELL rows (0: read it at run time)
// FAST = at most four hoisted constraints (all held in registers, absent ones as zeros: ONE loop body for
// every selected variable keeps the hot code small) and, for a permutation problem, one common domain
template <int T, bool PERM, int S, bool FAST, class WS>
__device__ inline int evaluate_candidates_eq_s(const DevModel& m, const WS& s, int v, int nh, int bias, int nchunks_c, int* out, ChunkRegs& q)
{
	typedef Group<T> G;
	MV mv(m);
	const int lane = threadIdx.x & 31;
	int xv = s.x[v];
	asm("" : "+r"(xv));
	const int nslots = PERM ? m.V : mv.dom_size(v);
	int a0 = 0, b0 = 0, a1 = 0, b1 = 0, a2 = 0, b2 = 0, a3 = 0, b3 = 0;
	if (nh > 0) { int4 h = s.hoist[0]; a0 = h.y; b0 = h.z; }
	if (nh > 1) { int4 h = s.hoist[1]; a1 = h.y; b1 = h.z; }
	if (nh > 2) { int4 h = s.hoist[2]; a2 = h.y; b2 = h.z; }
	if (nh > 3) { int4 h = s.hoist[3]; a3 = h.y; b3 = h.z; }
	constexpr int SP = S <= 2 ? S : 4;  // entries per packed row when S is known (one 32/64/128-bit load)
	const unsigned* ell = reinterpret_cast<const unsigned*>(m.blob + m.o.ell_eq);
	const unsigned char* cs_bytes = reinterpret_cast<const unsigned char*>(s.cs);
	int mine = 0;
	q.key = GB_INT_MAX; q.bits = 0u;
	auto load_row = [&](int slot, unsigned* e) {
		if (SP == 1) e[0] = __ldg(ell + slot);
		else if (SP == 2) { const uint2 r = __ldg(reinterpret_cast<const uint2*>(ell) + slot); e[0] = r.x; e[1] = r.y; }
		else { const uint4 r = __ldg(reinterpret_cast<const uint4*>(ell) + slot); e[0] = r.x; e[1] = r.y; e[SP > 2 ? 2 : 0] = r.z; e[SP > 2 ? 3 : 0] = r.w; }
	};
#ifdef GB_PREFETCH
	unsigned en[SP > 0 ? SP : 1];
	if (PERM && S > 0) load_row(G::warp() * 32 + lane, en);
#endif
#ifndef GB_UNROLL2   // not unrolled: the loop body has to stay small enough for the per-scheduler instruction cache (measured)
#pragma unroll 1
#endif
	for (int ch = G::warp(); ch < nchunks_c; ch += G::NW) {
		const int slot = ch * 32 + lane;
		unsigned e[SP > 0 ? SP : 1];
		if (PERM && S > 0) {
			// the partner's row first: this load does not depend on the walker's state
#ifdef GB_PREFETCH
#pragma unroll
			for (int k = 0; k < SP; ++k) e[k] = en[k];
			load_row(min(slot + 32 * G::NW, nchunks_c * 32 - 1), en); // next chunk's row, one pass ahead
#else
			load_row(slot, e);
#endif
		}
		bool ok = slot < nslots;
		int d;
		if (PERM) {
			ok = ok && slot != v;
			d = s.x[slot];                                   // padded: no clamp
			if (!FAST && !m.identical_domains) ok = ok && mv.in_domain(v, d) && mv.in_domain(min(slot, m.V - 1), xv);
		} else {
			d = mv.dom_value(v, min(slot, nslots - 1));
		}
		// keeps the comparison below in 32 bits: narrowed to the 16-bit storage type it would cost a
		// sign extension of both operands per chunk
		asm("" : "+r"(d));
		ok = ok && d != xv;                                  // search_unit.hpp:833 / 855-860
		int sum = bias + abs(b0 + a0 * d) + abs(b1 + a1 * d);
		if (FAST) sum += abs(b2 + a2 * d) + abs(b3 + a3 * d);   // absent entries are 0*d + 0: one loop for every variable
		else if (nh > 2) {
			sum += abs(b2 + a2 * d) + abs(b3 + a3 * d);
			for (int k = 4; k < nh; ++k) { int4 h = s.hoist[k]; sum += abs(h.z + h.y * d); }
		}
		if (PERM) {
			const int dx = xv - d, ndx = d - xv;
			const int nrow = S > 0 ? S : m.ell_eq_stride;
			const unsigned* row = ell + slot * nrow;
#pragma unroll
			for (int k = 0; k < nrow; ++k) {
				const unsigned f = S > 0 ? e[k] : __ldg(row + k); // byte offset of the constraint state | a_j << 16
				const int fy = (int)f >> 16;
				const int2 st = *reinterpret_cast<const int2*>(cs_bytes + (f & 0xFFFFu)); // {T, a_v}
				// Partner's own constraint (a_v = 0; 881-885): |T + a_j*dx| - |T|, |T| being the current
				// error of an Eq constraint.  Shared with v (872-874): the hoisted single-variable term
				// |T - a_v*dx| is replaced by the two-variable one |T - a_v*dx + a_j*dx|.
				const int t1 = st.x + st.y * ndx;
				sum += abs(t1 + fy * dx) - abs(t1);
			}
		}
		mine += ok ? 1 : 0;
		if (!ok) sum = GB_INT_MAX;
		if (out && slot < m.max_candidates) out[slot] = sum;
		commit_value_chunk<T>(s, q, ch, sum);
	}
	return mine;
}

template <int T, bool PERM, class WS>
__device__ inline int evaluate_candidates_eq(const DevModel& m, const WS& s, int v, int nh, int bias, int nchunks_c, int* out, ChunkRegs& q)
{
	if (!PERM) {
		if (nh <= 4) return evaluate_candidates_eq_s<T, PERM, 0, true>(m, s, v, nh, bias, nchunks_c, out, q);
		return evaluate_candidates_eq_s<T, PERM, 0, false>(m, s, v, nh, bias, nchunks_c, out, q);
	}
#ifndef GB_NOFAST
	if (nh <= 4 && m.identical_domains)
#else
	if (false)
#endif
	{
		switch (m.ell_stride) { // warp-uniform: fully unrolled row loops for the usual row lengths
		case 1: return evaluate_candidates_eq_s<T, PERM, 1, true>(m, s, v, nh, bias, nchunks_c, out, q);
		case 2: return evaluate_candidates_eq_s<T, PERM, 2, true>(m, s, v, nh, bias, nchunks_c, out, q);
		case 3: return evaluate_candidates_eq_s<T, PERM, 3, true>(m, s, v, nh, bias, nchunks_c, out, q);
		case 4: return evaluate_candidates_eq_s<T, PERM, 4, true>(m, s, v, nh, bias, nchunks_c, out, q);
		default: break;
		}
	}
	switch (m.ell_stride) {
	case 2: return evaluate_candidates_eq_s<T, PERM, 2, false>(m, s, v, nh, bias, nchunks_c, out, q);
	case 3: return evaluate_candidates_eq_s<T, PERM, 3, false>(m, s, v, nh, bias, nchunks_c, out, q);
	case 4: return evaluate_candidates_eq_s<T, PERM, 4, false>(m, s, v, nh, bias, nchunks_c, out, q);
	default: return evaluate_candidates_eq_s<T, PERM, 0, false>(m, s, v, nh, bias, nchunks_c, out, q);
	}
}

// ---------------------------------------------------------------------------------------------
// The Antidote heuristic set (GhostSearchOpts::strategy = GHOST_STRATEGY_ANTIDOTE; one-warp walkers).
//  * candidates: every variable, weighted by its projected error, tabu ones by 0
//    (antidote_search_variable_candidates_heuristic.cpp:39-48);
//  * variable: drawn in proportion to the weights; the INDEX drawn is the variable (…variable_heuristic.cpp:39-43);
//  * the loop then erases from the weight vector the first entry whose VALUE equals that index
//    (search_unit.hpp:825-828 — written for the default heuristic, whose vector holds variable ids), and while no
//    move happens the shrunken vector is sampled again: the kernel keeps the erased ORIGINAL positions as one bit
//    mask per chunk (lane ch holds chunk ch's, in the register pair that carries the candidate summaries of the
//    default strategy) and maps a drawn position to its index in the shrunken vector by counting the bits before it;
//  * value: drawn in proportion to max(0, -cumulated delta), uniformly over all candidates when none improves;
//    the branch that follows is decided by the CHOSEN candidate's delta (…value_heuristic.cpp:50-74).
// The oracle restates the same with the same Philox words (ghost_oracle.c, sample_weighted).
// ---------------------------------------------------------------------------------------------
// weight of original position id in the current vector, -1 if it is not in it (erased, or padding)
template <class WS>
__device__ __forceinline__ int antidote_weight(const DevModel& m, const WS& s, int id, int lm_rel, unsigned erased_bits)
{
	const int lane = threadIdx.x & 31;
	if (id >= m.V || ((erased_bits >> lane) & 1u)) return -1;
	const int pos = scan_pos(m, id);
	return (int)s.tabu[pos] > lm_rel ? 0 : s.err_var[pos];
}
__device__ __forceinline__ unsigned warp_inclusive_sum(unsigned v)
{
	const int lane = threadIdx.x & 31;
#pragma unroll
	for (int off = 1; off < 32; off <<= 1) { const unsigned t = __shfl_up_sync(0xffffffffu, v, off); if (lane >= off) v += t; }
	return v;
}
// One pass over the vector: tabu count, vector length, and the drawn index (-1: the vector is empty).
template <class WS>
__device__ inline int antidote_select(const DevModel& m, const WS& s, int nchunks_v, int lm_rel, uint32_t word, const ChunkRegs& q,
                                      int* ntabu_out, int* nvec_out)
{
	const int lane = threadIdx.x & 31;
	unsigned my_sum = 0, my_cnt = 0;
	int ntabu = 0;
	for (int ch = 0; ch < nchunks_v; ++ch) {
		const int id = ch * 32 + lane;
		const unsigned er = __shfl_sync(0xffffffffu, q.bits, ch);
		const int w = antidote_weight(m, s, id, lm_rel, er);
		if (id < m.V && (int)s.tabu[scan_pos(m, id)] > lm_rel) ++ntabu;
		const unsigned sum = __reduce_add_sync(0xffffffffu, (unsigned)(w > 0 ? w : 0));
		const unsigned cnt = __popc(__ballot_sync(0xffffffffu, w >= 0));
		if (lane == ch) { my_sum = sum; my_cnt = cnt; }
	}
	const unsigned total = __reduce_add_sync(0xffffffffu, my_sum);
	const int nvec = (int)__reduce_add_sync(0xffffffffu, my_cnt);
	*ntabu_out = __reduce_add_sync(0xffffffffu, ntabu);
	*nvec_out = nvec;
	if (nvec == 0) return -1;
	if (total == 0) return 0;   // a vector of zeros: index 0 (what std::discrete_distribution returns for NaN probabilities)
	const unsigned pick = below(word, total);
	const unsigned incl = warp_inclusive_sum(my_sum), incl_cnt = warp_inclusive_sum(my_cnt);
	const int ch = __ffs(__ballot_sync(0xffffffffu, incl > pick)) - 1;
	const unsigned before = __shfl_sync(0xffffffffu, incl - my_sum, ch), cnt_before = __shfl_sync(0xffffffffu, incl_cnt - my_cnt, ch);
	const unsigned er = __shfl_sync(0xffffffffu, q.bits, ch);
	const int w = antidote_weight(m, s, ch * 32 + lane, lm_rel, er);
	const unsigned incl_w = warp_inclusive_sum((unsigned)(w > 0 ? w : 0));
	const int l = __ffs(__ballot_sync(0xffffffffu, incl_w > pick - before)) - 1;
	const unsigned present = __ballot_sync(0xffffffffu, w >= 0);
	return (int)cnt_before + __popc(present & ((1u << l) - 1u));
}
// search_unit.hpp:825-828 on the weight vector: erase the first entry whose weight equals `value`.  Returns 1 if one was erased.
template <class WS>
__device__ inline int antidote_erase(const DevModel& m, const WS& s, int nchunks_v, int lm_rel, int value, ChunkRegs& q)
{
	const int lane = threadIdx.x & 31;
	for (int ch = 0; ch < nchunks_v; ++ch) {
		const unsigned er = __shfl_sync(0xffffffffu, q.bits, ch);
		const int w = antidote_weight(m, s, ch * 32 + lane, lm_rel, er);
		const unsigned hit = __ballot_sync(0xffffffffu, w >= 0 && w == value);
		if (hit) {
			if (lane == ch) q.bits |= 1u << (__ffs(hit) - 1);
			return 1;
		}
	}
	return 0;
}
// AntidoteSearchValueHeuristic::select_value over the cumulated deltas of every slot (INT_MAX: not a candidate).
// Returns the chosen slot (-1: no candidate) and its delta.
__device__ inline int antidote_value(const int* sums, int nchunks_c, uint32_t word, int* delta_out)
{
	const int lane = threadIdx.x & 31;
	unsigned my_w = 0, my_c = 0;
	for (int ch = 0; ch < nchunks_c; ++ch) {
		const int sum = sums[ch * 32 + lane];
		const bool isc = sum != GB_INT_MAX;
		const unsigned w = __reduce_add_sync(0xffffffffu, (unsigned)(isc && sum < 0 ? -sum : 0));
		const unsigned c = __popc(__ballot_sync(0xffffffffu, isc));
		if (lane == ch) { my_w = w; my_c = c; }
	}
	const unsigned total_w = __reduce_add_sync(0xffffffffu, my_w), nc = __reduce_add_sync(0xffffffffu, my_c);
	if (nc == 0) return -1;
	int slot;
	if (total_w == 0) {   // no candidate improves: uniform over all of them (…value_heuristic.cpp:67-68)
		const unsigned k = below(word, nc);
		const unsigned incl = warp_inclusive_sum(my_c);
		const int ch = __ffs(__ballot_sync(0xffffffffu, incl > k)) - 1;
		const unsigned before = __shfl_sync(0xffffffffu, incl - my_c, ch);
		const unsigned isc = __ballot_sync(0xffffffffu, sums[ch * 32 + lane] != GB_INT_MAX);
		slot = ch * 32 + __fns(isc, 0, (int)(k - before) + 1);
	} else {
		const unsigned pick = below(word, total_w);
		const unsigned incl = warp_inclusive_sum(my_w);
		const int ch = __ffs(__ballot_sync(0xffffffffu, incl > pick)) - 1;
		const unsigned before = __shfl_sync(0xffffffffu, incl - my_w, ch);
		const int sum = sums[ch * 32 + lane];
		const unsigned incl_w = warp_inclusive_sum((unsigned)(sum != GB_INT_MAX && sum < 0 ? -sum : 0));
		slot = ch * 32 + __ffs(__ballot_sync(0xffffffffu, incl_w > pick - before)) - 1;
	}
	*delta_out = sums[slot];
	return slot;
}

// ---------------------------------------------------------------------------------------------
// Culprit Search error projection (GhostSearchOpts::strategy == GHOST_STRATEGY_CULPRIT; one warp per walker).
// Reference: src/algorithms/culprit_search_error_projection_algorithm.cpp; restated in oracle/ghost_oracle.c
// (culprit_on_constraint / culprit_apply / culprit_update_variable_errors), which the decision-replay tests pin
// against the reference's own traces.  Per constraint c the reference keeps a row of V doubles: for a scope
// variable the sum of the error deltas of moving it to the two neighbouring values of its current one in the
// domain AS GIVEN (Variable::get_partial_domain(3)), turned into max - x and scaled so that the row sums to the
// constraint's error; a variable OUTSIDE the scope gets the same treatment of a 0 entry, i.e. max / sum * error.
// Here a row is one double per scope entry (s.culprit, indexed like c2v) plus one for everything outside
// (s.culprit_out[c]).  The projected errors are doubles (s.errd); every operation that rounds (the division, the
// product, the sums of rows into errors) is written with the _rn intrinsics in the reference's order, so the
// values — and with them the arg-max ties of the candidate scan — are the reference's bit for bit.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ double warp_max_f64(double v)
{
#pragma unroll
	for (int off = 16; off > 0; off >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, off));
	return v;
}
__device__ __forceinline__ double warp_sum_f64(double v) // exact here: every term is an integer-valued double far below 2^53
{
#pragma unroll
	for (int off = 16; off > 0; off >>= 1) v = __dadd_rn(v, __shfl_xor_sync(0xffffffffu, v, off));
	return v;
}

// Variable::get_partial_domain(3) around the current value (variable.cpp:60-86), first and last entry of what it
// returns — see partial_domain_ends in the oracle for the two wrap-around cases the reference's iterator arithmetic produces
__device__ inline void culprit_partial_ends(const DevModel& m, const MV& mv, int var, int n, int cur, int* first, int* third)
{
	const int off = mv.ld(m.o.dom_off, var);
	const bool arithmetic = m.given_sorted && m.contiguous_domains;
	const int sec = m.given_sorted ? m.o.dom_vals : m.o.dom_given;
	const int lo = mv.dom_lo(var);
	*first = arithmetic ? lo : mv.ld(sec, off);
	*third = arithmetic ? lo + n - 1 : mv.ld(sec, off + n - 1);
	if (n <= 3) return;
	int index;
	if (arithmetic) index = cur - lo;
	else if (m.given_sorted) {
		int b = 0, e = n;
		while (b < e) { const int mid = (b + e) >> 1; if (mv.ld(sec, off + mid) < cur) b = mid + 1; else e = mid; }
		index = b;
	} else {
		index = 0;
		while (index < n && mv.ld(sec, off + index) != cur) ++index;
	}
	if (index >= 1 && index + 2 <= n) {
		*first = arithmetic ? cur - 1 : mv.ld(sec, off + index - 1);
		*third = arithmetic ? cur + 1 : mv.ld(sec, off + index + 1);
	}
}

// CulpritSearchErrorProjection::compute_variable_errors_on_constraint (…:50-110): row c from the constraint's
// current error, its cache and the assignment.  Lanes over the scope entries; ends with a warp barrier.
template <class WS>
__device__ inline void culprit_on_constraint(const DevModel& m, const WS& s, int c)
{
	MV mv(m);
	const int lane = threadIdx.x & 31;
	const int b = mv.scope_begin(c), e = mv.scope_begin(c + 1), outside = m.V - (e - b);
	const int err = s.err_of(c);
	if (err <= 0) { // uniform over the warp
		for (int p = b + lane; p < e; p += 32) s.culprit[p] = 0.;
		if (lane == 0) s.culprit_out[c] = 0.;
		__syncwarp();
		return;
	}
	const int kind = mv.kind(c);
	const bool by_position = kind == GHOST_KIND_ALL_DIFFERENT_ABS_DIFF || kind == GHOST_KIND_ALL_DIFFERENT_LAG_DIFF;
	double mx = -CUDART_INF;
	for (int p = b + lane; p < e; p += 32) {
		const int var = mv.scope_var(p), n = mv.dom_size(var), cur = s.x[var];
		const int ea = by_position ? p - b : mv.ld(m.o.ent_coef, p);
		int sum;
		if (n > 2) {
			int first, third;
			culprit_partial_ends(m, mv, var, n, cur, &first, &third);
			sum = constraint_delta(mv, s, c, var, ea, first, -1, 0, 0) + constraint_delta(mv, s, c, var, ea, third, -1, 0, 0);
		} else if (n == 2) {
			const int off = mv.ld(m.o.dom_off, var);
			const int d0 = mv.ld(m.o.dom_given, off), d1 = mv.ld(m.o.dom_given, off + 1);
			sum = constraint_delta(mv, s, c, var, ea, d0 == cur ? d1 : d0, -1, 0, 0);
		} else
			sum = constraint_delta(mv, s, c, var, ea, cur, -1, 0, 0);
		const double val = (double)sum;
		s.culprit[p] = val;
		mx = fmax(mx, val);
	}
	mx = warp_max_f64(mx);
	if (outside > 0) mx = fmax(mx, 0.); // the entries of the variables outside the scope are 0 and take part in the maximum
	double part = 0.;
	for (int p = b + lane; p < e; p += 32) { const double t = __dadd_rn(-s.culprit[p], mx); s.culprit[p] = t; part = __dadd_rn(part, t); }
	const double out = outside > 0 ? mx : 0.;
	const double sum = __dadd_rn(warp_sum_f64(part), __dmul_rn(out, (double)outside));
	const double derr = (double)err;
	for (int p = b + lane; p < e; p += 32) { const double t = s.culprit[p]; s.culprit[p] = t == 0. ? 0. : __dmul_rn(__ddiv_rn(t, sum), derr); }
	if (lane == 0) s.culprit_out[c] = (outside > 0 && out != 0.) ? __dmul_rn(__ddiv_rn(out, sum), derr) : 0.;
	__syncwarp();
}

// s.cpos[v] = scope entry of v in constraint c, 0xFFFF outside the scope.  Ends with a warp barrier.
template <class WS>
__device__ inline void culprit_map_scope(const DevModel& m, const WS& s, int c)
{
	MV mv(m);
	const int lane = threadIdx.x & 31;
	const int b = mv.scope_begin(c), e = mv.scope_begin(c + 1);
	for (int v = lane; v < m.V; v += 32) s.cpos[v] = 0xFFFFu;
	__syncwarp();
	for (int p = b + lane; p < e; p += 32) s.cpos[mv.scope_var(p)] = (unsigned short)(p - b);
	__syncwarp();
}
// Row c combined into the projected errors over ALL variables (s.cpos maps c's scope).  ADD: err = row + err
// (std::plus, …:122-127, 145-150); otherwise err = row - err: the "remove" step of update_variable_errors (…:136-141)
// hands std::transform the row as first and the errors as second range with std::minus — restated as written.
template <bool ADD, class WS>
__device__ inline void culprit_apply(const DevModel& m, const WS& s, int c)
{
	MV mv(m);
	const int lane = threadIdx.x & 31;
	const int b = mv.scope_begin(c);
	const double out = s.culprit_out[c];
	for (int v = lane; v < m.V; v += 32) {
		const unsigned q = s.cpos[v];
		const double row = q == 0xFFFFu ? out : s.culprit[b + (int)q];
		s.errd[v] = ADD ? __dadd_rn(row, s.errd[v]) : __dsub_rn(row, s.errd[v]);
	}
	__syncwarp();
}
// CulpritSearchErrorProjection::compute_variable_errors (…:112-129)
template <class WS>
__device__ inline void culprit_compute_variable_errors(const DevModel& m, const WalkerLayout& L, const WS& s)
{
	const int lane = threadIdx.x & 31;
	for (int v = lane; v < L.nchunks_v * 32; v += 32) s.errd[v] = 0.;
	__syncwarp();
	for (int c = 0; c < m.C; ++c) {
		culprit_on_constraint(m, s, c);
		culprit_map_scope(m, s, c);
		culprit_apply<true>(m, s, c);
	}
}
// CulpritSearchErrorProjection::update_variable_errors (…:131-152) for one constraint whose error has just changed
template <class WS>
__device__ inline void culprit_update_variable_errors(const DevModel& m, const WS& s, int c)
{
	culprit_map_scope(m, s, c);
	culprit_apply<false>(m, s, c);
	culprit_on_constraint(m, s, c);
	culprit_apply<true>(m, s, c);
}

// AdaptiveSearchVariableCandidatesHeuristic::compute_variable_candidates over the fp64 projected errors: the
// eligible variables whose error is the largest one and at least -1 (the heuristic's `worst` starts at -1 and the
// Culprit errors can be negative), compared as doubles.  The summaries come back in the default strategy's form
// with key 0 for a chunk that attains the maximum and -1 otherwise, so picking, erasing and persisting a candidate
// set are the default code with cand_max = 0.
template <class WS>
__device__ inline int scan_candidates_culprit(const DevModel& m, const WS& s, int nchunks_v, int local_moves, bool free_vars_ok,
                                              int* ntabu_out, ChunkRegs& q)
{
	MV mv(m);
	const int lane = threadIdx.x & 31;
	const bool use_mask = !free_vars_ok && !m.all_have_ctr;
	double my_key = -CUDART_INF; unsigned my_bits = 0u; int ntabu = 0;
	for (int ch = 0; ch < nchunks_v; ++ch) {
		const int v = ch * 32 + lane;
		const bool k = (int)s.tabu[scan_pos(m, v)] > local_moves;   // padding entries are tabu for ever
		ntabu += k ? 1 : 0;
		double e = k ? -CUDART_INF : s.errd[v];
		if (use_mask && !(((unsigned)mv.ld(m.o.has_ctr, ch) >> lane) & 1u)) e = -CUDART_INF;
		const double c = warp_max_f64(e);
		const unsigned bb = __ballot_sync(0xffffffffu, e == c);
		if (lane == ch) { my_key = c; my_bits = bb; }
	}
	ntabu = __reduce_add_sync(0xffffffffu, ntabu) - (nchunks_v * 32 - m.V);
	const double worst = warp_max_f64(my_key);
	const bool any = worst >= -1.;
	q.key = (any && my_key == worst) ? 0 : -1;
	q.bits = my_bits;
	*ntabu_out = ntabu;
	return any ? regs_count(q, 0) : 0;
}

// ---------------------------------------------------------------------------------------------
// The reference's landscape-analysis builds as strategies (GHOST_STRATEGY_HILL_CLIMBING / _RANDOM_WALK; include/solver.hpp:
// 502-515): AllFreeVariableCandidatesHeuristic — every variable that is not tabu is a candidate, constrained or not
// (all_free_variable_candidates_heuristic.cpp:38-46) — with no error projection; the random walk also draws the value
// uniformly among the candidates (random_walk_value_heuristic.cpp:47-57) and always moves (search_unit.hpp:1007-1036).
// The candidate set comes back in the default strategy's form (key 0 for every chunk), so picking / erasing / persisting
// are the default code with cand_max = 0.
// ---------------------------------------------------------------------------------------------
template <class WS>
__device__ inline int scan_candidates_all_free(const DevModel& m, const WS& s, int nchunks_v, int local_moves, int* ntabu_out, ChunkRegs& q)
{
	const int lane = threadIdx.x & 31;
	unsigned mine = 0u;
	for (int ch = 0; ch < nchunks_v; ++ch) {
		const int v = ch * 32 + lane;
		const unsigned b = __ballot_sync(0xffffffffu, v < m.V && (int)s.tabu[scan_pos(m, v)] <= local_moves);
		if (lane == ch) mine = b;
	}
	q.key = 0; q.bits = mine;
	const int n = __reduce_add_sync(0xffffffffu, (int)__popc(mine));
	*ntabu_out = m.V - n;
	return n;
}
// uniform draw among the candidate slots (sums[slot] != INT_MAX), in ascending slot order; -1 if there is none
__device__ inline int random_walk_value(const int* sums, int nchunks_c, uint32_t word, int* delta_out)
{
	const int lane = threadIdx.x & 31;
	int my_c = 0;
	for (int ch = 0; ch < nchunks_c; ++ch) {
		const unsigned b = __ballot_sync(0xffffffffu, sums[ch * 32 + lane] != GB_INT_MAX);
		if (lane == ch) my_c = __popc(b);
	}
	const int total = __reduce_add_sync(0xffffffffu, my_c);
	if (total == 0) return -1;
	const int k = (int)below(word, (uint32_t)total);
	const int incl = (int)warp_inclusive_sum((unsigned)my_c);
	const int ch = __ffs(__ballot_sync(0xffffffffu, incl > k)) - 1;
	const int before = __shfl_sync(0xffffffffu, incl - my_c, ch);
	const unsigned isc = __ballot_sync(0xffffffffu, sums[ch * 32 + lane] != GB_INT_MAX);
	const bool me = ((isc >> lane) & 1u) && __popc(isc & ((1u << lane) - 1u)) == k - before;
	const int slot = ch * 32 + __ffs(__ballot_sync(0xffffffffu, me)) - 1;
	*delta_out = sums[slot];
	return slot;
}

// ---------------------------------------------------------------------------------------------
// fp32 form of the swap evaluation (DevModel::f32_ok: permutation problem, LinearEquationEq only, one
// common domain, distinct start values, <= 4 live constraints per variable, one warp per walker).
//
// Every term of a candidate's sum is |T + a*dx| on integers that lower.cpp bounds far below 2^24, so
// fp32 computes them exactly, and it computes them on the FMA pipe: t = fma(a, dx, T) and sum += |t|
// (the absolute value is a source modifier of the add) are two instructions where the integer form
// needs IMAD + IABS + IADD, two of them on the ALU pipe that limits this kernel.  The constraint state
// {T, a_v} is kept as two floats; the 16-bit assignment becomes a float by OR-ing it into the mantissa
// of 2^23 (dx is then one exact subtraction of two such numbers); the coefficient of a row entry is
// stored as an IEEE half (one conversion instruction, FMA pipe); and the sum starts at 1.5 * 2^23, so
// that the bit pattern of the finished sum is 0x4B400000 + (integer sum): the arg-min and the tie
// ballot work on those bits directly, no float-to-int conversion per candidate.
// Same results as the integer path bit for bit (the golden and trajectory tests run both).
// ---------------------------------------------------------------------------------------------
#define GB_F32_BASE 12582912.0f
#define GB_F32_BASE_BITS 0x4B400000
template <class WS> __device__ __forceinline__ float& Tf_of(const WS& s, int c) { return reinterpret_cast<float*>(s.cs + c)[0]; }
template <class WS> __device__ __forceinline__ float& avf_of(const WS& s, int c) { return reinterpret_cast<float*>(s.cs + c)[1]; }
__device__ __forceinline__ float half_bits_to_float(unsigned f) { return __half2float(__ushort_as_half((unsigned short)f)); }

// Hoist of the selected variable (see hoist_variable_eq): hoist[k] = {constraint id, a_v, T} with a_v, T as float bits.
template <class WS>
__device__ inline int hoist_variable_f32(const DevModel& m, const WS& s, int v, int* bias_out)
{
	const int lane = threadIdx.x & 31;
	const int nrow = m.ell_eq_stride;
	const unsigned dummy = (unsigned)m.C * 8u;
	const unsigned* row = reinterpret_cast<const unsigned*>(m.blob + m.o.ell_f32) + v * nrow;
	const unsigned f = lane < nrow ? __ldg(row + lane) : (dummy << 16);
	const unsigned off = f >> 16;
	const bool valid = off != dummy;
	int err = 0;
	if (valid) {
		const int c = (int)(off >> 3);
		const float a = half_bits_to_float(f), t = Tf_of(s, c);
		err = s.err_of(c);
		s.hoist[lane] = make_int4(c, __float_as_int(a), __float_as_int(t), 0);
		s.mark_of(c) = lane + 1; avf_of(s, c) = a;
	}
	*bias_out = -__reduce_add_sync(0xffffffffu, err);
	return __popc(__ballot_sync(0xffffffffu, valid));
}
template <class WS>
__device__ inline void unmark_variable_f32(const DevModel& m, const WS& s, int v)
{
	const int lane = threadIdx.x & 31;
	if (lane < m.ell_eq_stride) {
		const int c = (int)(__ldg(reinterpret_cast<const unsigned*>(m.blob + m.o.ell_f32) + v * m.ell_eq_stride + lane) >> 19);
		s.av_of(c) = 0; s.mark_of(c) = 0; // the dummy entry is all zeros anyway
	}
}

// The selected variable's side of a candidate's sum, in registers: (a_v, T) of its hoisted constraints (absent
// ones are 0*dx + 0), x_v as an integer and as 2^23 + x_v, and the start of every sum (bias + 1.5 * 2^23).
// NH = hoisted pairs evaluated per candidate (2 or 4).
template <int NH>
struct HoistF32 { float a0, t0, a1, t1, a2, t2, a3, t3, xvo, base; int xv; };
template <int NH, class WS>
__device__ __forceinline__ HoistF32<NH> load_hoist_f32(const WS& s, int v, int nh, int bias)
{
	HoistF32<NH> h;
	int xv = s.x[v];
	asm("" : "+r"(xv));
	h.xv = xv;
	h.xvo = __int_as_float(0x4B000000 | xv);
	h.a0 = h.t0 = h.a1 = h.t1 = h.a2 = h.t2 = h.a3 = h.t3 = 0.f;
	if (nh > 0) { const int4 e = s.hoist[0]; h.a0 = __int_as_float(e.y); h.t0 = __int_as_float(e.z); }
	if (nh > 1) { const int4 e = s.hoist[1]; h.a1 = __int_as_float(e.y); h.t1 = __int_as_float(e.z); }
	if (NH > 2 && nh > 2) { const int4 e = s.hoist[2]; h.a2 = __int_as_float(e.y); h.t2 = __int_as_float(e.z); }
	if (NH > 2 && nh > 3) { const int4 e = s.hoist[3]; h.a3 = __int_as_float(e.y); h.t3 = __int_as_float(e.z); }
	h.base = (float)bias + GB_F32_BASE;
	return h;
}
// |T - a_v*dx| of the hoisted constraints on top of the base: the selected variable's constraints with x_v replaced by x_j
template <int NH>
__device__ __forceinline__ float hoisted_sum_f32(const HoistF32<NH>& h, float dx)
{
	float sum = h.base + fabsf(fmaf(-h.a0, dx, h.t0));
	sum += fabsf(fmaf(-h.a1, dx, h.t1));
	if (NH > 2) { sum += fabsf(fmaf(-h.a2, dx, h.t2)); sum += fabsf(fmaf(-h.a3, dx, h.t3)); }
	return sum;
}

// Key (bits of 1.5 * 2^23 + sum, or INT_MAX) of the swap with the variable in `slot`, every constraint of the
// partner taken from its packed row.  S = entries per row (a row of three is stored in four words: one 128-bit load).
template <int NH, int S, class WS>
__device__ __forceinline__ int f32_slot_key(const DevModel& m, const WS& s, const HoistF32<NH>& h, int slot, int nslots, int* out)
{
	const unsigned* ell = reinterpret_cast<const unsigned*>(m.blob + m.o.ell_f32);
	const unsigned char* cs_bytes = reinterpret_cast<const unsigned char*>(s.cs);
	constexpr int SW = S == 3 ? 4 : S;
	unsigned e[SW];
	// the partner's row first: this load does not depend on the walker's state
	if (SW == 1) e[0] = __ldg(ell + slot);
	else if (SW == 2) { const uint2 r = __ldg(reinterpret_cast<const uint2*>(ell) + slot); e[0] = r.x; e[SW > 1 ? 1 : 0] = r.y; }
	else { const uint4 r = __ldg(reinterpret_cast<const uint4*>(ell) + slot); e[0] = r.x; e[SW > 1 ? 1 : 0] = r.y; e[SW > 2 ? 2 : 0] = r.z; e[SW > 2 ? 3 : 0] = r.w; }
	int d = s.x[slot];                                   // padded: no clamp
	asm("" : "+r"(d));
	const float dx = h.xvo - __int_as_float(0x4B000000 | d);   // x_v - x_j, exact
	float sum = hoisted_sum_f32(h, dx);
	float part = 0.f;   // a second accumulator: the two chains overlap (any association is exact here)
#pragma unroll
	for (int k = 0; k < S; ++k) {
		const unsigned f = e[k];                         // state offset << 16 | a_j as a half
		const float2 st = *reinterpret_cast<const float2*>(cs_bytes + (f >> 16)); // {T, a_v}
		const float aj = half_bits_to_float(f);
		// Partner's own constraint (a_v = 0; search_unit.hpp:881-885): |T + a_j*dx| - |T|.  Shared with v
		// (872-874): the hoisted single-variable term |T - a_v*dx| is replaced by |T - a_v*dx + a_j*dx|.
		const float u1 = fmaf(-st.y, dx, st.x);
		const float u2 = fmaf(aj, dx, u1);
		part += fabsf(u2);
		part -= fabsf(u1);
	}
	sum += part;
	const bool ok = d != h.xv && slot < nslots;          // search_unit.hpp:855-860 (the values are pairwise distinct)
	const int key = ok ? __float_as_int(sum) : GB_INT_MAX;
	if (out && slot < m.max_candidates) out[slot] = ok ? key - GB_F32_BASE_BITS : GB_INT_MAX;
	return key;
}

// Two chunks of 32 partners per loop pass: two independent dependency chains in flight, ONE warp reduction for
// both (the pair's minimum; a chunk that does not attain it gets an empty tie mask, which is all the pick needs:
// only chunks whose key equals the global minimum matter), and the pair's two summaries leave in one 128-bit
// store by lane 0 into the (val_key, val_bits) region, read back as {key, bits}[chunk] after the loop by the
// lane that owns the chunk.  The packed rows and the assignment are padded so that the second chunk of the last
// pair needs no clamp.
__device__ __forceinline__ void commit_pair(int4* summaries, int ch, int ka, int kb)
{
	const int cmin = __reduce_min_sync(0xffffffffu, min(ka, kb));
	const unsigned ba = __ballot_sync(0xffffffffu, ka == cmin), bb = __ballot_sync(0xffffffffu, kb == cmin);
	if ((threadIdx.x & 31) == 0) summaries[ch >> 1] = make_int4(cmin, (int)ba, cmin, (int)bb);
}
__device__ __forceinline__ void commit_quad(int4* summaries, int ch, int ka, int kb, int kc, int kd)
{
	const int cmin = __reduce_min_sync(0xffffffffu, min(min(ka, kb), min(kc, kd)));
	const unsigned ba = __ballot_sync(0xffffffffu, ka == cmin), bb = __ballot_sync(0xffffffffu, kb == cmin);
	const unsigned bc = __ballot_sync(0xffffffffu, kc == cmin), bd = __ballot_sync(0xffffffffu, kd == cmin);
	if ((threadIdx.x & 31) == 0) {
		summaries[ch >> 1] = make_int4(cmin, (int)ba, cmin, (int)bb);
		summaries[(ch >> 1) + 1] = make_int4(cmin, (int)bc, cmin, (int)bd);
	}
}
__device__ __forceinline__ void fetch_summary(const int4* summaries, int nchunks, ChunkRegs& q)
{
	const int lane = threadIdx.x & 31;
	__syncwarp();
	const int2 mine = reinterpret_cast<const int2*>(summaries)[lane];   // lanes beyond the last chunk read stale words: masked
	q.key = lane < nchunks ? mine.x : GB_INT_MAX;
	q.bits = lane < nchunks ? (unsigned)mine.y : 0u;
}

template <int NH, int S, class WS>
__device__ inline int evaluate_f32(const DevModel& m, const WS& s, int v, int nh, int bias, int nchunks_c, int* out, ChunkRegs& q)
{
	const int lane = threadIdx.x & 31;
	const HoistF32<NH> h = load_hoist_f32<NH>(s, v, nh, bias);
	const int nslots = m.V;
	int4* summaries = reinterpret_cast<int4*>(s.val_key);   // {key, bits} per chunk, two chunks per int4
#ifndef GB_F32_SINGLE
#pragma unroll 1
	for (int ch = 0; ch < nchunks_c; ch += 2) {
		const int slot = ch * 32 + lane;
		const int ka = f32_slot_key<NH, S>(m, s, h, slot, nslots, out), kb = f32_slot_key<NH, S>(m, s, h, slot + 32, nslots, out);
		commit_pair(summaries, ch, ka, kb);
	}
#else
#pragma unroll 1
	for (int ch = 0; ch < nchunks_c; ++ch) {
		const int ka = f32_slot_key<NH, S>(m, s, h, ch * 32 + lane, nslots, out);
		const int cmin = __reduce_min_sync(0xffffffffu, ka);
		const unsigned ba = __ballot_sync(0xffffffffu, ka == cmin);
		if (lane == 0) reinterpret_cast<int2*>(summaries)[ch] = make_int2(cmin, (int)ba);
	}
#endif
	fetch_summary(summaries, nchunks_c, q);
	// this lane's candidates: its slots below V, minus the selected variable itself
	return ((m.V - lane + 31) >> 5) - (lane == (v & 31) ? 1 : 0);
}

// Structured grid (DevModel::sg_ok, lower.cpp): chunk g of the candidate loop is group g of the variables (sg_cw of
// them, lanes beyond idle), so the group constraint U_g is the same for the whole chunk — its {T, a_v} sits in lane
// g and is broadcast with two shuffles — and the position constraint L_l is the same for lane l in every chunk —
// its {T, a_v} stays in two registers for the whole loop.  No row load, no gather: per candidate the x load, the
// hoisted terms and two constraint terms with register operands.  Cells that carry more constraints (the diagonals
// of a magic square) are skipped here and evaluated afterwards in full (f32_slot_key) by the lane that owns their
// chunk's summary, which merges them in place.  Tie masks are in (group, lane) coordinates: sg_slot_to_var().
__device__ __forceinline__ int sg_slot_to_var(const DevModel& m, int slot) { return (slot >> 5) * m.sg_cw + (slot & 31); }
template <int NH, int S, class WS>
__device__ inline int evaluate_sg(const DevModel& m, const WS& s, int v, int nh, int bias, int* out, ChunkRegs& q)
{
	const int lane = threadIdx.x & 31;
	const HoistF32<NH> h = load_hoist_f32<NH>(s, v, nh, bias);
	const unsigned char* cs_bytes = reinterpret_cast<const unsigned char*>(s.cs);
	const int4 row = __ldg(reinterpret_cast<const int4*>(m.blob + m.o.sg) + lane);
	const float2 stU = *reinterpret_cast<const float2*>(cs_bytes + row.x);   // {T, a_v} of group `lane`'s constraint
	const float2 stL = *reinterpret_cast<const float2*>(cs_bytes + row.y);   // {T, a_v} of position `lane`'s constraint
	const float au = (float)m.sg_au, al = (float)m.sg_al;
	const unsigned skip = (unsigned)row.z;
	const int cw = m.sg_cw, G = m.sg_groups;
	int4* summaries = reinterpret_cast<int4*>(s.val_key);
	// the loop walks the groups in order: a running pointer into the assignment, and the skip mask bit-reversed so that the
	// group's bit is the sign bit (one compare, chained with the value test, instead of shift + and + compare)
	const typename WS::x_type* px = s.x + lane;
	unsigned rm = __brev(skip);
	auto cell_key = [&](int ch) -> int {
		const int u = ch * cw + lane;
		int d = *px;                                         // padded by two chunks: no clamp
		px += cw;
		asm("" : "+r"(d));
		const float dx = h.xvo - __int_as_float(0x4B000000 | d);
		const float TU = __shfl_sync(0xffffffffu, stU.x, ch), aU = __shfl_sync(0xffffffffu, stU.y, ch);
		float sum = hoisted_sum_f32(h, dx);
		// as in f32_slot_key: |T - a_v*dx + a_j*dx| - |T - a_v*dx|, a_v = 0 unless the selected variable is in the constraint
		const float u1 = fmaf(-aU, dx, TU), u2 = fmaf(au, dx, u1);
		float part = fabsf(u2) - fabsf(u1);
		const float w1 = fmaf(-stL.y, dx, stL.x), w2 = fmaf(al, dx, w1);
		part += fabsf(w2) - fabsf(w1);
		sum += part;
		const bool cell = (int)rm >= 0;
		const unsigned rm_now = rm;
		rm <<= 1;
		const bool ok = cell & (d != h.xv);
#ifndef GB_SG_NO_CHAIN   // predicate chain spelled out: the compiler emits two selects otherwise
		int key;
		asm("{ .reg .pred p; setp.ge.s32 p, %2, 0; setp.ne.and.s32 p, %3, %4, p; selp.b32 %0, %1, 0x7fffffff, p; }"
		    : "=r"(key) : "r"(__float_as_int(sum)), "r"((int)rm_now), "r"(d), "r"(h.xv));
#else
		const int key = ok ? __float_as_int(sum) : GB_INT_MAX;
#endif
		if (out && cell) out[u] = ok ? key - GB_F32_BASE_BITS : GB_INT_MAX;
		return key;
	};
#ifdef GB_SG_QUAD
#pragma unroll 1
	for (int ch = 0; ch < G; ch += 4) {
		const int ka = cell_key(ch), kb = cell_key(ch + 1), kc = cell_key(ch + 2), kd = cell_key(ch + 3);
		commit_quad(summaries, ch, ka, kb, kc, kd);
	}
#elif !defined(GB_SG_SINGLE)   // two groups per pass with ONE reduction (measured faster once the per-iteration path had shrunk)
#ifdef GB_SG_PAIR_U2
#pragma unroll 2
#else
#pragma unroll 1
#endif
	for (int ch = 0; ch < G; ch += 2) {
		const int ka = cell_key(ch), kb = cell_key(ch + 1);
		commit_pair(summaries, ch, ka, kb);
	}
#else
#ifdef GB_SG_UNROLL2
#pragma unroll 2
#else
#pragma unroll 1
#endif
	for (int ch = 0; ch < G; ++ch) {
		const int ka = cell_key(ch);
		const int cmin = __reduce_min_sync(0xffffffffu, ka);
		const unsigned ba = __ballot_sync(0xffffffffu, ka == cmin);
		if (lane == 0) reinterpret_cast<int2*>(summaries)[ch] = make_int2(cmin, (int)ba);
	}
#endif
	fetch_summary(summaries, G, q);
	// the irregular cells of group `lane`, one per pass
	for (int e = 0; e < m.sg_xn; ++e) {
		const unsigned le = ((unsigned)row.w >> (8 * e)) & 0xFFu;
		const bool has = le != 0xFFu;
		const int u = has ? lane * cw + (int)le : 0;
		const int key = f32_slot_key<NH, S>(m, s, h, u, has ? m.V : 0, has ? out : nullptr);
		if (key < q.key) { q.key = key; q.bits = 1u << le; }
		else if (key == q.key && has) q.bits |= 1u << le;
	}
	return lane == 0 ? m.V - 1 : 0;
}

// FS: what the kernel was instantiated for — 0: decided here at run time; 1..4: the plain fp32 loop for packed rows
// of that length; 11..14: the structured-grid loop (rows of length FS - 10 for its irregular cells).  One
// instantiation of the search kernel per case keeps the other loop bodies out of its code: the per-iteration path of
// 24 warps in different phases has to stay inside the SM's instruction cache ("Instruction cache").
template <int FS, class WS>
__device__ inline int evaluate_candidates_f32(const DevModel& m, const WS& s, int v, int nh, int bias, int nchunks_c, int* out, ChunkRegs& q)
{
#ifdef GB_SG_TWO_NH
	if (FS > 10 && nh <= 2) return evaluate_sg<2, (FS > 10 ? FS - 10 : 1)>(m, s, v, nh, bias, out, q);
#endif
	if (FS > 10) return evaluate_sg<4, (FS > 10 ? FS - 10 : 1)>(m, s, v, nh, bias, out, q);
	if (FS > 0) {
#ifdef GB_F32_TWO_NH
		if (nh <= 2) return evaluate_f32<2, ((FS > 0 && FS < 10) ? FS : 1)>(m, s, v, nh, bias, nchunks_c, out, q);
#endif
		return evaluate_f32<4, ((FS > 0 && FS < 10) ? FS : 1)>(m, s, v, nh, bias, nchunks_c, out, q);
	}
	if (m.sg_ok) {
		switch (m.ell_stride) { // warp-uniform
		case 1: return evaluate_sg<4, 1>(m, s, v, nh, bias, out, q);
		case 2: return evaluate_sg<4, 2>(m, s, v, nh, bias, out, q);
		case 3: return evaluate_sg<4, 3>(m, s, v, nh, bias, out, q);
		default: return evaluate_sg<4, 4>(m, s, v, nh, bias, out, q);
		}
	}
	switch (m.ell_stride) {
	case 1: return evaluate_f32<4, 1>(m, s, v, nh, bias, nchunks_c, out, q);
	case 2: return evaluate_f32<4, 2>(m, s, v, nh, bias, nchunks_c, out, q);
	case 3: return evaluate_f32<4, 3>(m, s, v, nh, bias, nchunks_c, out, q);
	default: return evaluate_f32<4, 4>(m, s, v, nh, bias, nchunks_c, out, q);
	}
}

// The update part of a swap (same contract as apply_move_eq), constraint state in fp32.
template <class WS>
__device__ inline void apply_move_f32(const DevModel& m, const WS& s, int v, int j, int nh)
{
	MV mv(m);
	const int lane = threadIdx.x & 31;
	const int xv = s.x[v], xj = s.x[j];
	const float dxf = (float)(xv - xj);
	const int nrow = m.ell_eq_stride;
	const unsigned* rowj = reinterpret_cast<const unsigned*>(m.blob + m.o.ell_f32) + j * nrow;
	int myd = 0, myb = 0, mye = 0;
	if (lane < nh + nrow) {
		int c = -1;
		float told = 0.f, tnew = 0.f;
		if (lane < nh) {
			const int4 h = s.hoist[lane];
			c = h.x;
			told = Tf_of(s, c);
			tnew = fmaf(-__int_as_float(h.y), dxf, told);
			for (int e = 0; e < nrow; ++e) {
				const unsigned f = __ldg(rowj + e);
				if ((int)(f >> 19) == c) { tnew = fmaf(half_bits_to_float(f), dxf, tnew); break; } // the partner is in this scope too
			}
		} else {
			const unsigned f = __ldg(rowj + (lane - nh));
			const int cc = (int)(f >> 19);
			if (cc != m.C && s.mark_of(cc) == 0) { // not padding, not shared with v (handled above)
				c = cc;
				told = Tf_of(s, c);
				tnew = fmaf(half_bits_to_float(f), dxf, told);
			}
		}
		if (c >= 0) {
			myd = __float2int_rn(fabsf(tnew) - fabsf(told));
			Tf_of(s, c) = tnew; s.err_of(c) = __float2int_rn(fabsf(tnew));
			if (myd != 0) { myb = mv.scope_begin(c); mye = mv.scope_begin(c + 1); }
		}
	}
	// only the constraints whose error changed (padding entries, shared ones and zero deltas drop out here)
	unsigned todo = __ballot_sync(0xffffffffu, myd != 0);
	while (todo) {
		const int k = __ffs(todo) - 1;
		todo &= todo - 1;
		const int d = __shfl_sync(0xffffffffu, myd, k), b = __shfl_sync(0xffffffffu, myb, k), e = __shfl_sync(0xffffffffu, mye, k);
		for (int i = b + lane; i < e; i += 32) atomicAdd(&s.err_var[mv.scope_pos(i)], d);
	}
	__syncwarp();
	if (lane == 0) { s.x[v] = xj; s.x[j] = xv; }
	__syncwarp();
}

template <int T, bool PERM, bool EQ, bool F32 = false, int FS = 0, class WS>
__device__ inline int evaluate_candidates(const DevModel& m, const WS& s, int v, int nh, int bias, int nchunks_c, int* out, ChunkRegs& q)
{
	if (F32) return evaluate_candidates_f32<FS>(m, s, v, nh, bias, nchunks_c, out, q);
	if (EQ) return evaluate_candidates_eq<T, PERM>(m, s, v, nh, bias, nchunks_c, out, q);
	typedef Group<T> G;
	MV mv(m);
	const int lane = threadIdx.x & 31;
	const int xv = s.x[v];
	const int nslots = PERM ? m.V : mv.dom_size(v);
	int mine = 0;
	q.key = GB_INT_MAX; q.bits = 0u;
	if (!PERM && T == 32 && nslots <= 4 && nh > 2) {
		// A handful of candidates over many constraints (binary domains: the knapsack rows): candidates across lanes
		// would leave one lane summing nh terms while 31 wait ("low candidate counts").  Here the lanes
		// take the CONSTRAINTS of one candidate, a warp reduction adds them up, and lane k keeps candidate k's sum.
		int sum = GB_INT_MAX;
		for (int k = 0; k < nslots; ++k) {                       // warp-uniform
			const int d = mv.dom_value(v, k);
			int total = GB_INT_MAX;
			if (d != xv) {                                       // search_unit.hpp:833
				int part = 0;
				for (int t = lane; t < nh; t += 32) part += hoisted_term(mv, s, s.hoist[t], d);
				total = __reduce_add_sync(0xffffffffu, part) + bias;
				if (lane == k) { sum = total; mine = 1; }
			}
			if (out && lane == 0) out[k] = total;
		}
		if (out) for (int k = nslots + lane; k < m.max_candidates; k += 32) out[k] = GB_INT_MAX;
		commit_value_chunk<T>(s, q, 0, sum);
		return mine;
	}
	for (int ch = G::warp(); ch < nchunks_c; ch += G::NW) {
		int slot = ch * 32 + lane;
		int sum = GB_INT_MAX;
		bool ok = candidate_sum<PERM, EQ>(m, mv, s, v, xv, nh, bias, slot, nslots, &sum);
		mine += ok ? 1 : 0;
		if (!ok) sum = GB_INT_MAX;
		if (out && slot < m.max_candidates) out[slot] = sum;
		commit_value_chunk<T>(s, q, ch, sum);
	}
	return mine;
}

// ---------------------------------------------------------------------------------------------
// The update part of a move when every live constraint is a LinearEquationEq (same contract as
// apply_move below): the affected constraints are the hoisted ones of v plus, for a swap, the
// partner's own row entries; their new state is T' = T + a*delta and error |T'| — no kind lookups,
// no CSR walks.  The projection scatter first gathers each affected constraint's scope bounds in one
// lane each, then broadcasts them, so that the dependent loads of different constraints overlap.
// ---------------------------------------------------------------------------------------------
template <int T, bool PERM, class WS>
__device__ inline void apply_move_eq(const DevModel& m, const WS& s, int v, int new_value, int nh)
{
	typedef Group<T> G;
	MV mv(m);
	const int tid = G::tid(), lane = threadIdx.x & 31;
	const int xv = s.x[v];
	int j = 0, xj = 0;
	if (PERM) { j = new_value; xj = s.x[j]; }
	const int nv_new = PERM ? xj : new_value;
	const int nrow = m.ell_eq_stride;
	const unsigned* rowj = mv.eq_row(j);
	const int naff = nh + (PERM ? nrow : 0);
	// new state of affected constraint k (hoisted ones first, then the partner's row); returns its id
	// (-1: nothing to do) and the change of its error
	auto update = [&](int k, int* delta_out) -> int {
		int c = -1, delta = 0;
		if (k < nh) {
			const int4 h = s.hoist[k];
			c = h.x >> 8;
			const int told = s.T_of(c);
			int tnew = told + h.y * (nv_new - xv);
			if (PERM)
				for (int e = 0; e < nrow; ++e) {
					const unsigned f = __ldg(rowj + e);
					if ((int)(f & 0xFFFFu) == c * 8) { tnew += ((int)f >> 16) * (xv - xj); break; } // the partner is in this scope too
				}
			delta = abs(tnew) - abs(told);
			s.T_of(c) = tnew; s.err_of(c) = abs(tnew);
		} else {
			const unsigned f = __ldg(rowj + (k - nh));
			const int cc = (int)(f & 0xFFFFu) >> 3;
			if (cc != m.C && s.mark_of(cc) == 0) { // not padding, not shared with v (handled above)
				c = cc;
				const int told = s.T_of(c), tnew = told + ((int)f >> 16) * (xv - xj);
				delta = abs(tnew) - abs(told);
				s.T_of(c) = tnew; s.err_of(c) = abs(tnew);
			}
		}
		*delta_out = delta;
		return c;
	};
	if (T == 32 && naff <= 32) {
		// one warp: everything stays in registers
		int myd = 0, myb = 0, mye = 0;
		if (lane < naff) {
			const int c = update(lane, &myd);
			if (c >= 0 && myd != 0) { myb = mv.scope_begin(c); mye = mv.scope_begin(c + 1); }
		}
		// only the constraints whose error changed (padding entries, shared ones and zero deltas drop out here)
		unsigned todo = __ballot_sync(0xffffffffu, myd != 0);
		while (todo) {
			const int k = __ffs(todo) - 1;
			todo &= todo - 1;
			const int d = __shfl_sync(0xffffffffu, myd, k), b = __shfl_sync(0xffffffffu, myb, k), e = __shfl_sync(0xffffffffu, mye, k);
			for (int i = b + lane; i < e; i += 32) atomicAdd(&s.err_var[mv.scope_pos(i)], d);
		}
		__syncwarp();
		if (lane == 0) {
			s.x[v] = nv_new;
			if (PERM) s.x[j] = xv;
		}
		__syncwarp();
		return;
	}
	for (int k = tid; k < naff; k += T) {
		int delta;
		const int c = update(k, &delta);
		s.upd[2 * k] = c; s.upd[2 * k + 1] = delta;
	}
	G::sync();
	if (naff <= 32) {
		int myd = 0, myb = 0, mye = 0;
		if (lane < naff) {
			const int c = s.upd[2 * lane];
			myd = s.upd[2 * lane + 1];
			if (c >= 0 && myd != 0) { myb = mv.scope_begin(c); mye = mv.scope_begin(c + 1); }
		}
		for (int k = 0; k < naff; ++k) {
			const int d = __shfl_sync(0xffffffffu, myd, k), b = __shfl_sync(0xffffffffu, myb, k), e = __shfl_sync(0xffffffffu, mye, k);
			if (d == 0) continue;
			for (int i = b + tid; i < e; i += T) atomicAdd(&s.err_var[mv.scope_pos(i)], d);
		}
	} else {
		for (int k = 0; k < naff; ++k) {
			const int c = s.upd[2 * k], d = s.upd[2 * k + 1];
			if (c < 0 || d == 0) continue;
			for (int i = mv.scope_begin(c) + tid, e = mv.scope_begin(c + 1); i < e; i += T) atomicAdd(&s.err_var[mv.scope_pos(i)], d);
		}
	}
	if (tid == 0) {
		s.x[v] = nv_new;
		if (PERM) s.x[j] = xv;
	}
	G::sync();
}

// ---------------------------------------------------------------------------------------------
// SearchUnit::local_move's update part (search_unit.hpp:449-541): constraint errors, caches,
// projection, then the value(s).  One thread per affected constraint computes its delta with the
// generic constraint_delta(); the projection update (…projection_algorithm.cpp:51-58) is a
// shared-memory scatter.  Ends with a barrier.  `new_value` is the swap partner in permutation
// mode, the value otherwise.
// ---------------------------------------------------------------------------------------------
// CULPRIT: the Culprit projection is updated per affected constraint, in the reference's order, from the constraint's
// NEW error with its OLD cache and the OLD assignment (search_unit.hpp:457-465 updates the projection before the
// constraint): the deltas are computed first, the caches are touched in a second pass after the projection.
template <int T, bool CULPRIT = false, class WS>
__device__ inline void apply_move(const DevModel& m, const WS& s, int v, int new_value)
{
	typedef Group<T> G;
	MV mv(m);
	const int tid = G::tid();
	const int rbv = mv.row_begin(v), degv = mv.row_begin(v + 1) - rbv;
	const int xv = s.x[v];
	int j = -1, xj = 0, rbj = 0, degj = 0;
	if (m.permutation) { j = new_value; xj = s.x[j]; rbj = mv.row_begin(j); degj = mv.row_begin(j + 1) - rbj; }
	const int nv_new = m.permutation ? xj : new_value;
	const int naff = degv + degj;
	// the cache of a constraint of v's row / of the partner's row follows the move (FORCE: the Culprit second pass)
	auto first_row_cache = [&](int c, int ey, bool both, int eb, bool force) {
		if (CULPRIT && !force) return;
		const int kind = mv.kind(c);
		if (feat_on<WS>(GB_FEAT_ALLDIFF | GB_FEAT_ALLEQ) && (kind == GHOST_KIND_ALL_DIFFERENT || kind == GHOST_KIND_ALL_EQUAL)) {
			if (!both && nv_new != xv) { int tab = mv.cnt_tab(c); s.cnt_add(tab + xv, -1); s.cnt_add(tab + nv_new, +1); }
		} else if (feat_on<WS>(GB_FEAT_OFFSET) && kind == GHOST_KIND_ALL_DIFFERENT_OFFSET) {
			const int tab = mv.cnt_tab(c);
			if (nv_new != xv) { s.cnt_add(tab + ey + xv, -1); s.cnt_add(tab + ey + nv_new, +1); }
			if (both && xj != xv) { s.cnt_add(tab + eb + xj, -1); s.cnt_add(tab + eb + xv, +1); } // the partner's shifted value moves too
		} else if (kind_is_linear(kind)) {
			s.T_of(c) += ey * (nv_new - xv) + (both ? eb * (xv - xj) : 0);
		}
	};
	auto second_row_cache = [&](int c, int ey, bool force) {
		if (CULPRIT && !force) return;
		const int kind = mv.kind(c);
		if (feat_on<WS>(GB_FEAT_ALLDIFF | GB_FEAT_ALLEQ) && (kind == GHOST_KIND_ALL_DIFFERENT || kind == GHOST_KIND_ALL_EQUAL)) {
			if (xv != xj) { int tab = mv.cnt_tab(c); s.cnt_add(tab + xj, -1); s.cnt_add(tab + xv, +1); }
		} else if (feat_on<WS>(GB_FEAT_OFFSET) && kind == GHOST_KIND_ALL_DIFFERENT_OFFSET) {
			if (xv != xj) { int tab = mv.cnt_tab(c) + ey; s.cnt_add(tab + xj, -1); s.cnt_add(tab + xv, +1); }
		} else if (kind_is_linear(kind)) s.T_of(c) += ey * (xv - xj);
	};
	for (int k = tid; k < naff; k += T) {
		int c, delta;
		if (k < degv) {
			int2 e = mv.row_entry(rbv + k);
			c = e.x;
			int eb = 0; bool both = false;
			if (m.permutation) {
				for (int r = rbj; r < rbj + degj; ++r) { int2 f = mv.row_entry(r); if (f.x == c) { both = true; eb = f.y; break; } }
			}
			delta = constraint_delta(mv, s, c, v, e.y, nv_new, both ? j : -1, eb, xv);
			first_row_cache(c, e.y, both, eb, false);
			s.err_of(c) += delta;
		} else {
			int2 e = mv.row_entry(rbj + (k - degv));
			c = e.x;
			bool shared = s.mark_of(c) != 0; // shared with v: handled above
			if (CULPRIT && !shared) // (the marks leave out the constraints whose delta is always 0; here every constraint must count once)
				for (int r = rbv; r < rbv + degv; ++r) if (mv.row_entry(r).x == c) { shared = true; break; }
			if (shared) { s.upd[2 * k] = -1; continue; }
			delta = constraint_delta(mv, s, c, j, e.y, xv, -1, 0, 0);
			second_row_cache(c, e.y, false);
			s.err_of(c) += delta;
		}
		s.upd[2 * k] = c; s.upd[2 * k + 1] = delta;
	}
	G::sync();
	if (CULPRIT) {
		for (int k = 0; k < naff; ++k) {
			const int c = s.upd[2 * k];
			if (c >= 0) culprit_update_variable_errors(m, s, c);
		}
		for (int k = tid; k < naff; k += T) {
			if (s.upd[2 * k] < 0) continue;
			if (k < degv) {
				const int2 e = mv.row_entry(rbv + k);
				int eb = 0; bool both = false;
				if (m.permutation) {
					for (int r = rbj; r < rbj + degj; ++r) { int2 f = mv.row_entry(r); if (f.x == e.x) { both = true; eb = f.y; break; } }
				}
				first_row_cache(e.x, e.y, both, eb, true);
			} else {
				const int2 e = mv.row_entry(rbj + (k - degv));
				second_row_cache(e.x, e.y, true);
			}
		}
		G::sync();
	}
	// projection update (…projection_algorithm.cpp:51-58).  A constraint with a short scope (the binary
	// and ternary links of the n-queens / Costas models) is scattered by ONE thread, all of them at once;
	// a long scope (AllDifferent, knapsack rows) by the whole group.
	for (int base = 0; base < naff; base += T) {
		const int k = base + tid;
		int c = -1, delta = 0, b = 0, e = 0;
		if (k < naff) { c = s.upd[2 * k]; delta = s.upd[2 * k + 1]; }
		const bool live = c >= 0 && delta != 0;
		if (live) { b = mv.scope_begin(c); e = mv.scope_begin(c + 1); }
		const bool small = live && e - b <= 8;
		if (small) for (int i = b; i < e; ++i) atomicAdd(&s.err_var[mv.scope_pos(i)], delta);
		if (live && !small) s.upd[2 * k + 1] = delta; else if (k < naff) s.upd[2 * k + 1] = 0; // what is left for the group pass
	}
	G::sync();
	for (int k = 0; k < naff; ++k) {
		int c = s.upd[2 * k], delta = s.upd[2 * k + 1];
		if (c < 0 || delta == 0) continue;
		int b = mv.scope_begin(c), e = mv.scope_begin(c + 1);
		for (int i = b + tid; i < e; i += T) atomicAdd(&s.err_var[mv.scope_pos(i)], delta);
	}
	if (feat_on<WS>(GB_FEAT_ALLEQ) && m.has_all_equal) {
		// the count tables changed: new lead counts of the AllEqual constraints among the affected ones
		for (int k = 0; k < naff; ++k) {
			const int c = s.upd[2 * k];
			if (c >= 0 && mv.kind(c) == GHOST_KIND_ALL_EQUAL) alleq_refresh<T>(m, s, c);
		}
	}
	if (tid == 0) {
		s.x[v] = nv_new;
		if (m.permutation) s.x[j] = xv;
	}
	G::sync();
	if (feat_on<WS>(GB_FEAT_PAIRDIFF) && m.has_abs_diff) {
		// refresh the |difference| tables of the extension kind from the new assignment
		for (int k = tid; k < naff; k += T) {
			int c = s.upd[2 * k];
			if (c < 0) continue;
			const int kd = mv.kind(c);
			if (kd != GHOST_KIND_ALL_DIFFERENT_ABS_DIFF && kd != GHOST_KIND_ALL_DIFFERENT_LAG_DIFF) continue;
			const bool absolute = kd == GHOST_KIND_ALL_DIFFERENT_ABS_DIFF;
			const int lag = absolute ? 1 : mv.rhs(c);
			int b = mv.scope_begin(c), n = mv.scope_begin(c + 1) - b, tab = mv.cnt_tab(c);
			int2 e = (k < degv) ? mv.row_entry(rbv + k) : mv.row_entry(rbj + (k - degv));
			int pa = e.y, pb = -1;
			if (k < degv && m.permutation)
				for (int r = rbj; r < rbj + degj; ++r) { int2 f = mv.row_entry(r); if (f.x == c) { pb = f.y; break; } }
			int pairs[4]; int np = 0;
			auto add_pair = [&](int p) { if (p < 0 || p + lag >= n) return; for (int z = 0; z < np; ++z) if (pairs[z] == p) return; pairs[np++] = p; };
			add_pair(pa - lag); add_pair(pa);
			if (pb >= 0) { add_pair(pb - lag); add_pair(pb); }
			int va = (k < degv) ? v : j, vb2 = (k < degv) ? j : -1;
			for (int z = 0; z < np; ++z) {
				int ua = mv.scope_var(b + pairs[z]), ub = mv.scope_var(b + pairs[z] + lag);
				int na = s.x[ua], nb = s.x[ub];
				// old values: undo the write
				int oa = (ua == va) ? (k < degv ? xv : xj) : ((ua == vb2 && pb >= 0) ? xj : na);
				int ob = (ub == va) ? (k < degv ? xv : xj) : ((ub == vb2 && pb >= 0) ? xj : nb);
				s.cnt_add(tab + pair_value(absolute, oa, ob), -1);
				s.cnt_add(tab + pair_value(absolute, na, nb), +1);
			}
		}
		G::sync();
	}
}

} // namespace ghost_b200
