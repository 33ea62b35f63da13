// Host C++ API checks (no gtest in the image: plain CHECK macro).  `cpu` mode needs no GPU;
// `gpu` mode runs Solver::fast_search through the C ABI on the device.
#include <cmath>
#include <cstring>
#include <iostream>
#include <set>
#include <vector>

#include <ghost/solver.hpp>

static int failures = 0;
#define CHECK(cond) do { if (!(cond)) { std::cerr << __FILE__ << ":" << __LINE__ << ": CHECK failed: " #cond "\n"; ++failures; } } while (0)
#define CHECK_THROWS(expr) do { bool thrown = false; try { expr; } catch (...) { thrown = true; } if (!thrown) { std::cerr << __FILE__ << ":" << __LINE__ << ": no throw: " #expr "\n"; ++failures; } } while (0)

using ghost::Variable;
typedef std::vector<int> V;

// The expectations of the reference's only test file (tests/src/test_variable.cpp:38-195).
static void test_variable()
{
	Variable c1(V{1, 9, 5, 7, 3}, 3, "var_ctor1"), c1d(V{1, 9, 5, 7, 3}), c2(10, 6, 4, "var_ctor2"), c2d(10, 6),
		c3(V{2, 8, 6, 4, 0}, "var_ctor3"), c4(20, 3, "var_ctor4");
	CHECK((c1.get_full_domain() == V{1, 9, 5, 7, 3}));
	CHECK((c2.get_full_domain() == V{10, 11, 12, 13, 14, 15}));
	CHECK((c4.get_full_domain() == V{20, 21, 22}));
	CHECK((c1.get_partial_domain(3) == V{5, 7, 3}));
	CHECK((c1.get_partial_domain(5) == V{1, 9, 5, 7, 3}));
	CHECK((c1.get_partial_domain(10) == V{1, 9, 5, 7, 3}));
	CHECK(c1.get_partial_domain(0).empty() && c1.get_partial_domain(-3).empty());
	CHECK((c1d.get_partial_domain(3) == V{1, 9, 3}));
	CHECK((c2.get_partial_domain(3) == V{13, 14, 15}));
	CHECK((c2.get_partial_domain(5) == V{10, 12, 13, 14, 15}));
	CHECK((c2d.get_partial_domain(3) == V{10, 11, 15}));
	CHECK((c3.get_partial_domain(1) == V{2}));
	CHECK((c3.get_partial_domain(2) == V{2, 0}));
	CHECK((c3.get_partial_domain(3) == V{2, 8, 0}));
	CHECK((c3.get_partial_domain(4) == V{2, 8, 4, 0}));
	CHECK((c4.get_partial_domain(2) == V{20, 22}));
	CHECK((c4.get_partial_domain(4) == V{20, 21, 22}));
	CHECK(c1.get_value() == 7 && c1d.get_value() == 1 && c2.get_value() == 14 && c2d.get_value() == 10 && c3.get_value() == 2 && c4.get_value() == 20);
	c1.set_value(5); c1d.set_value(5); c2.set_value(11); c2d.set_value(13); c3.set_value(6); c4.set_value(21);
	CHECK((c1.get_partial_domain(3) == V{9, 5, 7}));
	CHECK((c2.get_partial_domain(3) == V{10, 11, 12}));
	CHECK((c2d.get_partial_domain(3) == V{12, 13, 14}));
	CHECK((c3.get_partial_domain(3) == V{8, 6, 4}));
	CHECK((c4.get_partial_domain(1) == V{21}));
	CHECK_THROWS(c1.set_value(2)); CHECK_THROWS(c1.set_value(-5)); CHECK_THROWS(c1.set_value(500));
	CHECK_THROWS(c2.set_value(9)); CHECK_THROWS(c2.set_value(16)); CHECK_THROWS(c3.set_value(3)); CHECK_THROWS(c4.set_value(23));
	CHECK(c1.get_domain_size() == 5 && c2.get_domain_size() == 6 && c4.get_domain_size() == 3);
	CHECK(c1.get_domain_min_value() == 1 && c2.get_domain_min_value() == 10 && c3.get_domain_min_value() == 0);
	CHECK(c1.get_domain_max_value() == 9 && c2.get_domain_max_value() == 15 && c3.get_domain_max_value() == 8 && c4.get_domain_max_value() == 22);
	CHECK(c1.get_name() == "var_ctor1" && c1d.get_name().empty() && c4.get_name() == "var_ctor4");
}

class QueensBuilder : public ghost::ModelBuilder
{
public:
	int n;
	QueensBuilder(int n) : n(n) {}
	void declare_variables() override
	{
		create_n_variables(n, 0, n);          // q_i
		create_n_variables(n, 0, 2 * n - 1);  // d1_i = q_i + i
		create_n_variables(n, -(n - 1), 2 * n - 1, n - 1); // d2_i = q_i - i
	}
	void declare_constraints() override
	{
		using namespace ghost::global_constraints;
		V q, d1, d2;
		for (int i = 0; i < n; ++i) { q.push_back(i); d1.push_back(n + i); d2.push_back(2 * n + i); }
		constraints.emplace_back(std::make_shared<AllDifferent>(q));
		constraints.emplace_back(std::make_shared<AllDifferent>(d1));
		constraints.emplace_back(std::make_shared<AllDifferent>(d2));
		for (int i = 0; i < n; ++i) constraints.emplace_back(std::make_shared<LinearEquationEq>(V{i, n + i}, -i, std::vector<double>{1, -1}));
		for (int i = 0; i < n; ++i) constraints.emplace_back(std::make_shared<LinearEquationEq>(V{i, 2 * n + i}, i, std::vector<double>{1, -1}));
	}
};

// a user-defined constraint: fine on the host, refused by the GPU path
class MyConstraint : public ghost::Constraint
{
	double required_error(const std::vector<Variable*>& v) const override { return v[0]->get_value() == v[1]->get_value() ? 1.0 : 0.0; }
public:
	MyConstraint(const V& ids) : ghost::Constraint(ids) {}
};
class UserBuilder : public ghost::ModelBuilder
{
public:
	void declare_variables() override { create_n_variables(3, 0, 3); }
	void declare_constraints() override { constraints.emplace_back(std::make_shared<MyConstraint>(V{0, 1})); }
};

class TinyBuilder : public ghost::ModelBuilder
{
public:
	void declare_variables() override { create_n_variables(3, 1, 3); }
	void declare_constraints() override
	{
		using namespace ghost::global_constraints;
		constraints.emplace_back(std::make_shared<AllDifferent>(variables));
		constraints.emplace_back(std::make_shared<LinearEquationLeq>(V{0, 1}, 3));
	}
	void declare_objective() override
	{
		objective = std::make_shared<ghost::global_objectives::LinearObjective>(variables, std::vector<double>{1, 10, 100}, true);
	}
};

static void test_complete_search()
{
	// 4-queens has exactly 2 solutions; enumeration is in variable-then-value order
	QueensBuilder qb(4);
	ghost::Solver<QueensBuilder> s(qb);
	std::vector<double> costs; std::vector<V> sols;
	CHECK(s.complete_search(costs, sols));
	CHECK(sols.size() == 2);
	if (sols.size() == 2) { CHECK((V(sols[0].begin(), sols[0].begin() + 4) == V{1, 3, 0, 2})); CHECK((V(sols[1].begin(), sols[1].begin() + 4) == V{2, 0, 3, 1})); }
	// permutations of {1,2,3} with x0 + x1 <= 3: (1,2,3) and (2,1,3); objective x0 + 10 x1 + 100 x2
	TinyBuilder tb;
	ghost::Solver<TinyBuilder> t(tb);
	costs.clear(); sols.clear();
	CHECK(t.complete_search(costs, sols));
	CHECK(sols.size() == 2 && costs.size() == 2);
	if (sols.size() == 2) { CHECK((sols[0] == V{1, 2, 3}) && costs[0] == 321); CHECK((sols[1] == V{2, 1, 3}) && costs[1] == 312); }
}

static void test_user_constraint_rejected()
{
	UserBuilder ub;
	ghost::Solver<UserBuilder> s(ub);
	double cost; V sol;
	bool thrown = false;
	try { s.fast_search(cost, sol, 1000.0); }
	catch (const std::runtime_error& e) { thrown = std::strstr(e.what(), "no CPU fallback") != nullptr; }
	CHECK(thrown);
}

static void test_gpu_searches()
{
	{
		QueensBuilder qb(8);
		ghost::Solver<QueensBuilder> s(qb);
		ghost::Options o; o.seed = 7; o.number_walkers = 2048;
		double cost = -1; V sol;
		bool found = s.fast_search(cost, sol, 20e6, o);
		CHECK(found && cost == 0.0 && sol.size() == 24);
		std::set<int> cols(sol.begin(), sol.begin() + 8), d1, d2;
		for (int i = 0; i < 8; ++i) { d1.insert(sol[i] + i); d2.insert(sol[i] - i); }
		CHECK(cols.size() == 8 && d1.size() == 8 && d2.size() == 8);
		CHECK(s.get_statistics().total.candidates_evaluated > 0 && s.get_statistics().walkers == 2048);
		CHECK(s.get_variables()[3].get_value() == sol[3]);
	}
	{
		TinyBuilder tb;
		ghost::Solver<TinyBuilder> s(tb);
		ghost::Options o; o.seed = 3; o.number_walkers = 256;
		double cost = 0; V sol;
		bool found = s.fast_search(cost, sol, 0.2e6, o);
		CHECK(found && cost == 321 && (sol == V{1, 2, 3}));   // the better of the two feasible assignments, sign restored
	}
}

int main(int argc, char** argv)
{
	const bool gpu = argc > 1 && std::strcmp(argv[1], "gpu") == 0;
	test_variable();
	test_user_constraint_rejected();
	if (gpu) { test_gpu_searches(); test_complete_search(); } // complete_search enumerates on the device
	std::cout << (failures ? "FAILED" : "OK") << " (" << failures << " failures)\n";
	return failures ? 1 : 0;
}
