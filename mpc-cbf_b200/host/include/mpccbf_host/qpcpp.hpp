// Host mirror of the reference's solver-agnostic QP container and solver interface, backed by the
// B200 dense QP kernel through the C ABI (mpccbf_solve_qp_dense).
//
// Same public names and call semantics as
//   qpcpp::Variable / LinearConstraint / CostFunction / Problem   qpcpp/include/qpcpp/Problem.h:16-196
//   qpcpp::SolveStatus / Solver<T>                                qpcpp/include/qpcpp/solvers/Solver.h:13-37
// with qpcpp::B200Solver<T> taking the place of qpcpp::CPLEXSolver<T>
// (qpcpp/include/qpcpp/solvers/CPLEX.h; CPLEX.cpp:35-178). The storage is NOT the reference's
// pointer-keyed hash maps: variables and rows carry dense indices so that a problem flattens
// into the contiguous P, q, A, l, u, lb, ub arrays the device kernel consumes without any hashing,
// and so that many problems of one shape can be solved in ONE batched launch (solveBatch).
//
// Semantics kept: objective is sum_{i<=j} c_ij x_i x_j + sum c_i x_i + const with (i,j) and (j,i)
// accumulating into one coefficient (Problem.cpp:101-116); default bounds are
// numeric_limits::lowest()/max() meaning "free" (|bound| >= 1e20 is infinite, like CPLEX);
// misuse (a variable of another problem) throws std::runtime_error; the solution is written to
// Variable::solution_value only for OPTIMAL / FEASIBLE (CPLEX.cpp:148-156); resetProblem() clears
// cost and rows but keeps the variables (Problem.cpp:267).
#pragma once
#include <cstdint>
#include <deque>
#include <limits>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <utility>
#include <vector>

#include "../../../../include/mpccbf.h"

namespace qpcpp {

enum class SolveStatus { OPTIMAL, FEASIBLE, UNBOUNDED, INFEASIBLE, ERROR, UNKNOWN, INFEASIBLEORUNBOUNDED };

inline std::string SolveStatusToStr(SolveStatus s) {
    static const char* names[] = {"OPTIMAL", "FEASIBLE", "UNBOUNDED", "INFEASIBLE", "ERROR", "UNKNOWN",
                                  "INFEASIBLEORUNBOUNDED"};
    const int i = static_cast<int>(s);
    return (i >= 0 && i < 7) ? names[i] : "";
}

template <typename T>
class Problem;

template <typename T>
class Variable {
  public:
    const Problem<T>& context_problem() const { return *owner_; }
    T min() const { return min_; }
    T max() const { return max_; }
    void set_min(T v) { min_ = v; }
    void set_max(T v) { max_ = v; }
    void set_solution_value(T v) { solution_ = v; }
    T solution_value() const { return solution_; }
    std::size_t index() const { return index_; }  // dense column of this variable (creation order)

  private:
    Variable(const Problem<T>* owner, std::size_t index, T mn, T mx)
        : owner_(owner), index_(index), min_(mn), max_(mx), solution_(T(0)) {}
    const Problem<T>* owner_;
    std::size_t index_;
    T min_, max_, solution_;
    friend class Problem<T>;
};

template <typename T>
class LinearConstraint {
  public:
    void setCoefficient(const Variable<T>* v, T coefficient) {
        const std::size_t j = owner_->column_of(v, "LinearConstraint::setCoefficient");
        for (auto& e : entries_)
            if (e.first == j) {
                e.second = coefficient;
                return;
            }
        entries_.emplace_back(j, coefficient);
    }
    T getCoefficient(const Variable<T>* v) const {
        const std::size_t j = owner_->column_of(v, "LinearConstraint::getCoefficient");
        for (const auto& e : entries_)
            if (e.first == j) return e.second;
        return T(0);
    }
    T min() const { return min_; }
    T max() const { return max_; }
    const std::vector<std::pair<std::size_t, T>>& entries() const { return entries_; }

  private:
    LinearConstraint(const Problem<T>* owner, T mn, T mx) : owner_(owner), min_(mn), max_(mx) {}
    const Problem<T>* owner_;
    T min_, max_;
    std::vector<std::pair<std::size_t, T>> entries_;
    friend class Problem<T>;
};

template <typename T>
class CostFunction {
  public:
    void addQuadraticTerm(const Variable<T>* a, const Variable<T>* b, T coefficient) {
        quad_[key(owner_->column_of(a, "CostFunction::addQuadraticTerm"),
                  owner_->column_of(b, "CostFunction::addQuadraticTerm"))] += coefficient;
    }
    T getQuadraticCoefficient(const Variable<T>* a, const Variable<T>* b) const {
        auto it = quad_.find(key(owner_->column_of(a, "CostFunction::getQuadraticCoefficient"),
                                 owner_->column_of(b, "CostFunction::getQuadraticCoefficient")));
        return it == quad_.end() ? T(0) : it->second;
    }
    void addLinearTerm(const Variable<T>* v, T coefficient) {
        const std::size_t j = owner_->column_of(v, "CostFunction::addLinearTerm");
        if (lin_.size() <= j) lin_.resize(j + 1, T(0));
        lin_[j] += coefficient;
    }
    T getLinearCoefficient(const Variable<T>* v) const {
        const std::size_t j = owner_->column_of(v, "CostFunction::getLinearCoefficient");
        return j < lin_.size() ? lin_[j] : T(0);
    }
    void add_constant(T c) { constant_ += c; }
    T constant() const { return constant_; }
    void setZero() {
        quad_.clear();
        lin_.clear();
        constant_ = T(0);
    }

  private:
    explicit CostFunction(const Problem<T>* owner) : owner_(owner), constant_(T(0)) {}
    static std::uint64_t key(std::size_t i, std::size_t j) {  // unordered pair -> (min, max)
        if (i > j) std::swap(i, j);
        return (static_cast<std::uint64_t>(i) << 32) | static_cast<std::uint64_t>(j);
    }
    const Problem<T>* owner_;
    std::unordered_map<std::uint64_t, T> quad_;
    std::vector<T> lin_;
    T constant_;
    friend class Problem<T>;
};

template <typename T>
class Problem {
  public:
    Problem() : cost_(this) {}
    Problem(const Problem&) = delete;
    Problem& operator=(const Problem&) = delete;

    std::size_t numVariables() const { return vars_.size(); }
    Variable<T>* addVariable(T mn = std::numeric_limits<T>::lowest(), T mx = std::numeric_limits<T>::max()) {
        vars_.push_back(Variable<T>(this, vars_.size(), mn, mx));
        return &vars_.back();
    }
    bool hasVariable(const Variable<T>* v) const {
        return v != nullptr && v->owner_ == this && v->index_ < vars_.size() && &vars_[v->index_] == v;
    }
    std::size_t numLinearConstraints() const { return rows_.size(); }
    LinearConstraint<T>* addLinearConstraint(T mn = std::numeric_limits<T>::lowest(),
                                             T mx = std::numeric_limits<T>::max()) {
        rows_.push_back(LinearConstraint<T>(this, mn, mx));
        return &rows_.back();
    }
    const std::deque<Variable<T>>& variables() const { return vars_; }
    std::deque<Variable<T>>& mutable_variables() { return vars_; }
    const std::deque<LinearConstraint<T>>& linear_constraints() const { return rows_; }
    CostFunction<T>* cost_function() { return &cost_; }
    const CostFunction<T>* cost_function() const { return &cost_; }
    void clearLinearConstraints() { rows_.clear(); }
    void setCostFunctionToZero() { cost_.setZero(); }
    void resetProblem() {
        clearLinearConstraints();
        setCostFunctionToZero();
    }

    // Dense row-major arrays in the layout mpccbf_solve_qp_dense takes (reference objective
    // convention x'Px + q'x: the merged coefficient c_ij, i <= j, is placed at P[i][j]).
    void flatten(std::vector<double>& P, std::vector<double>& q, std::vector<double>& A, std::vector<double>& l,
                 std::vector<double>& u, std::vector<double>& lb, std::vector<double>& ub) const {
        const std::size_t n = vars_.size(), m = rows_.size();
        P.assign(n * n, 0.0);
        q.assign(n, 0.0);
        A.assign(m * n, 0.0);
        l.resize(m);
        u.resize(m);
        lb.resize(n);
        ub.resize(n);
        for (const auto& kv : cost_.quad_) {
            const std::size_t i = static_cast<std::size_t>(kv.first >> 32), j = static_cast<std::size_t>(kv.first & 0xffffffffu);
            P[i * n + j] = static_cast<double>(kv.second);
        }
        for (std::size_t j = 0; j < cost_.lin_.size(); ++j) q[j] = static_cast<double>(cost_.lin_[j]);
        for (std::size_t r = 0; r < m; ++r) {
            for (const auto& e : rows_[r].entries_) A[r * n + e.first] = static_cast<double>(e.second);
            l[r] = static_cast<double>(rows_[r].min_);
            u[r] = static_cast<double>(rows_[r].max_);
        }
        for (std::size_t j = 0; j < n; ++j) {
            lb[j] = static_cast<double>(vars_[j].min_);
            ub[j] = static_cast<double>(vars_[j].max_);
        }
    }

  private:
    std::size_t column_of(const Variable<T>* v, const char* who) const {
        if (!hasVariable(v)) throw std::runtime_error(std::string(who) + ": variable does not exist in the problem");
        return v->index_;
    }
    std::deque<Variable<T>> vars_;       // deque: stable addresses, creation order = column order
    std::deque<LinearConstraint<T>> rows_;
    CostFunction<T> cost_;
    friend class LinearConstraint<T>;
    friend class CostFunction<T>;
};

template <typename T>
class Solver {
  public:
    virtual ~Solver() = default;
    virtual SolveStatus solve(Problem<T>& problem) = 0;
};

// The B200 backend: one CTA per QP (ADMM + active-set polish). Owns (or borrows) a device handle.
template <typename T>
class B200Solver : public Solver<T> {
  public:
    explicit B200Solver(int device = 0) : owned_(true) {
        mpccbf_config cfg;
        mpccbf_config_default(&cfg);
        cfg.device = device;
        cfg.max_scenarios = 1;
        if (mpccbf_create(&cfg, &h_) != MPCCBF_OK) throw std::runtime_error(std::string("B200Solver: ") + mpccbf_last_error());
    }
    explicit B200Solver(mpccbf_handle* borrowed) : h_(borrowed), owned_(false) {}
    B200Solver(const B200Solver&) = delete;
    B200Solver& operator=(const B200Solver&) = delete;
    ~B200Solver() override {
        if (owned_ && h_) mpccbf_destroy(h_);
    }

    SolveStatus solve(Problem<T>& problem) override {
        std::vector<Problem<T>*> one{&problem};
        return solveBatch(one)[0];
    }

    // All problems must have the same number of variables and rows; one kernel launch.
    std::vector<SolveStatus> solveBatch(const std::vector<Problem<T>*>& problems) {
        std::vector<SolveStatus> out(problems.size(), SolveStatus::ERROR);
        if (problems.empty()) return out;
        const std::size_t n = problems[0]->numVariables(), m = problems[0]->numLinearConstraints(), B = problems.size();
        std::vector<double> P(B * n * n), q(B * n), A(B * m * n), l(B * m), u(B * m), lb(B * n), ub(B * n), x(B * n), obj(B);
        std::vector<int32_t> st(B), it(B);
        std::vector<double> p1, q1, a1, l1, u1, lb1, ub1;
        for (std::size_t b = 0; b < B; ++b) {
            if (problems[b]->numVariables() != n || problems[b]->numLinearConstraints() != m)
                throw std::invalid_argument("B200Solver::solveBatch: problems must share one shape");
            problems[b]->flatten(p1, q1, a1, l1, u1, lb1, ub1);
            std::copy(p1.begin(), p1.end(), P.begin() + b * n * n);
            std::copy(q1.begin(), q1.end(), q.begin() + b * n);
            std::copy(a1.begin(), a1.end(), A.begin() + b * m * n);
            std::copy(l1.begin(), l1.end(), l.begin() + b * m);
            std::copy(u1.begin(), u1.end(), u.begin() + b * m);
            std::copy(lb1.begin(), lb1.end(), lb.begin() + b * n);
            std::copy(ub1.begin(), ub1.end(), ub.begin() + b * n);
        }
        const int rc = mpccbf_solve_qp_dense(h_, static_cast<int32_t>(B), static_cast<int32_t>(n), static_cast<int32_t>(m),
                                             P.data(), q.data(), A.data(), l.data(), u.data(), lb.data(), ub.data(),
                                             x.data(), obj.data(), st.data(), it.data());
        if (rc != MPCCBF_OK) throw std::runtime_error(std::string("B200Solver: ") + mpccbf_last_error());
        for (std::size_t b = 0; b < B; ++b) {
            out[b] = static_cast<SolveStatus>(st[b]);
            if (out[b] == SolveStatus::OPTIMAL || out[b] == SolveStatus::FEASIBLE) {
                auto& vars = problems[b]->mutable_variables();
                for (std::size_t j = 0; j < n; ++j) vars[j].set_solution_value(static_cast<T>(x[b * n + j]));
            }
        }
        last_objective_.assign(obj.begin(), obj.end());
        last_iterations_.assign(it.begin(), it.end());
        return out;
    }
    const std::vector<double>& last_objectives() const { return last_objective_; }  // x'Px + q'x (no constant)
    const std::vector<int32_t>& last_iterations() const { return last_iterations_; }

  private:
    mpccbf_handle* h_ = nullptr;
    bool owned_;
    std::vector<double> last_objective_;
    std::vector<int32_t> last_iterations_;
};

}  // namespace qpcpp
