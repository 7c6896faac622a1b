// ORACLE — TEST INFRASTRUCTURE ONLY (see vfem_oracle.hpp).
// Restatement of TopologyOptimization{Filter,Constraint,Objective,Problem}.hh and
// OptimalityCriterion.hh.
#pragma once
#include "vfem_oracle_mg.hpp"

namespace orc {

struct Filter {
    std::vector<size_t> dims;
    virtual ~Filter() {}
    virtual void setGridDimensions(const std::vector<size_t> &d) { dims = d; }
    virtual void apply(const std::vector<Real> &in, std::vector<Real> &out) = 0;
    virtual void backprop(const std::vector<Real> &in, const std::vector<Real> &vars, std::vector<Real> &out) const = 0;
};

// TopologyOptimizationFilter.hh:55-79
struct ProjectionFilter : Filter {
    Real beta = 1.0;
    void apply(const std::vector<Real> &in, std::vector<Real> &out) override {
        out.resize(in.size());
        for (size_t i = 0; i < in.size(); ++i)
            out[i] = 0.5 * (std::tanh(0.5 * beta) + std::tanh(beta * (in[i] - 0.5))) / std::tanh(0.5 * beta);
    }
    void backprop(const std::vector<Real> &in, const std::vector<Real> &vars, std::vector<Real> &out) const override {
        out.resize(in.size());
        for (size_t i = 0; i < in.size(); ++i) {
            Real t = std::tanh(beta * (vars[i] - 0.5));
            out[i] = in[i] * 0.5 * beta * (1.0 - t * t) / std::tanh(0.5 * beta);
        }
    }
};

// TopologyOptimizationFilter.hh:105-162; stencil = cubic box clipped to the grid (NDVector.hh:235-262)
struct SmoothingFilter : Filter {
    size_t radius = 1;
    template <class F>
    void visitStencil(size_t i, F &&f) const {
        const size_t D = dims.size();
        std::vector<long long> c(D), lo(D), hi(D), cur(D);
        size_t r = i;
        for (size_t d = D; d-- > 0;) { c[d] = (long long)(r % dims[d]); r /= dims[d]; }
        size_t cnt = 1;
        for (size_t d = 0; d < D; ++d) { lo[d] = c[d] - (long long)radius; hi[d] = c[d] + (long long)radius; cnt *= 2 * radius + 1; }
        cur = lo;
        for (size_t k = 0; k < cnt; ++k) {
            bool in = true; size_t flat = 0;
            for (size_t d = 0; d < D; ++d) { if (cur[d] < 0 || cur[d] >= (long long)dims[d]) in = false; flat = flat * dims[d] + size_t(cur[d] < 0 ? 0 : cur[d]); }
            if (in) f(flat);
            for (size_t d = 0; d < D; ++d) { if (++cur[d] <= hi[d]) break; cur[d] = lo[d]; } // axis-0-fastest counter
        }
    }
    void apply(const std::vector<Real> &in, std::vector<Real> &out) override {
        out.assign(in.size(), 0.0);
#pragma omp parallel for schedule(static)
        for (long long i = 0; i < (long long)in.size(); ++i) {
            size_t cnt = 0; Real s = 0;
            visitStencil(size_t(i), [&](size_t k) { ++cnt; s += in[k]; });
            out[i] = s * (1.0 / Real(cnt)); // weights 1/numInStencil (Filter.hh:143-146)
        }
    }
    void backprop(const std::vector<Real> &in, const std::vector<Real> &, std::vector<Real> &out) const override {
        // transpose: out[k] = sum_{i : k in stencil(i)} in[i]/cnt_i; the box stencil is symmetric (k in S(i) <=> i in S(k))
        std::vector<Real> scaled(in.size());
#pragma omp parallel for schedule(static)
        for (long long i = 0; i < (long long)in.size(); ++i) {
            size_t cnt = 0; visitStencil(size_t(i), [&](size_t) { ++cnt; });
            scaled[i] = in[i] * (1.0 / Real(cnt));
        }
        out.assign(in.size(), 0.0);
#pragma omp parallel for schedule(static)
        for (long long k = 0; k < (long long)in.size(); ++k) {
            Real s = 0; visitStencil(size_t(k), [&](size_t i) { s += scaled[i]; });
            out[k] = s;
        }
    }
};

// TopologyOptimizationConstraint.hh:21-35
struct TotalVolumeConstraint {
    Real volumeFraction;
    explicit TotalVolumeConstraint(Real v) : volumeFraction(v) {}
    Real evaluate(const std::vector<Real> &v) const {
        Real s = 0; for (Real a : v) s += a;
        return 1.0 - (s / Real(v.size())) / volumeFraction;
    }
    void backprop(const std::vector<Real> &v, std::vector<Real> &out) const { out.assign(v.size(), -1.0 / (volumeFraction * Real(v.size()))); }
};

// TopologyOptimizationObjective.hh:24-102
template <int N>
struct ComplianceObjective {
    Sim<N> &sim; std::shared_ptr<MG<N>> mg; // mg == nullptr -> direct solve
    std::vector<Real> u, f;
    size_t cgIter = 100; Real tol = 1e-5; size_t mgIterations = 1, mgSmoothingIterations = 2;
    bool fullMultigrid = true, zeroInit = false;
    size_t lastIters = 0;
    std::function<void(size_t, Real)> residual_cb;
    ComplianceObjective(Sim<N> &s, std::shared_ptr<MG<N>> m) : sim(s), mg(m) {
        f.resize(sim.nn * N); sim.buildLoadVector(f.data());
        u.assign(sim.nn * N, 0.0);
        updateCache(sim.rho);
    }
    Real compliance() const { Real s = 0; for (size_t i = 0; i < u.size(); ++i) s += f[i] * u[i]; return 0.5 * s; }
    void gradient(std::vector<Real> &g) const { g.resize(sim.ne); sim.complianceGradient(u.data(), g.data()); }
    void updateCache(const std::vector<Real> &xPhys) {
        if (xPhys.size() != sim.ne) throw std::runtime_error("size mismatch");
        sim.rho = xPhys; sim.factorUpToDate = false;
        if (!mg) { sim.solveDirect(f.data(), u.data()); return; }
        typename MG<N>::PcgCb cb;
        if (residual_cb) cb = [&](size_t it, const Real *, const Real *r) {
            Real s = 0; for (size_t i = 0; i < u.size(); ++i) s += r[i] * r[i]; residual_cb(it, std::sqrt(s)); };
        if (zeroInit) std::fill(u.begin(), u.end(), 0.0);
        std::vector<Real> out(u.size());
        lastIters = mg->pcg(u.data(), f.data(), cgIter, tol, cb, mgIterations, mgSmoothingIterations, fullMultigrid, out.data());
        u.swap(out);
    }
};

// TopologyOptimizationProblem.hh:17-204
template <int N>
struct TOProblem {
    Sim<N> &sim; std::shared_ptr<ComplianceObjective<N>> objective;
    std::vector<std::shared_ptr<TotalVolumeConstraint>> constraints;
    std::vector<std::shared_ptr<Filter>> filters;
    std::vector<std::vector<Real>> cached; bool varsAreSet = false;
    TOProblem(Sim<N> &s, std::shared_ptr<ComplianceObjective<N>> o, std::vector<std::shared_ptr<TotalVolumeConstraint>> c,
              std::vector<std::shared_ptr<Filter>> f) : sim(s), objective(o), constraints(c), filters(f) {
        cached.assign(filters.size() + 1, std::vector<Real>(sim.ne, 0.0));
        std::vector<size_t> dims(sim.E, sim.E + N);
        for (auto &fl : filters) fl->setGridDimensions(dims);
    }
    size_t numVars() const { return sim.ne; }
    bool setVars(const Real *xin, bool forceUpdate = false) {
        std::vector<Real> &xDesign = cached[0];
        if (!forceUpdate && varsAreSet) {
            Real s = 0; for (size_t i = 0; i < sim.ne; ++i) { Real d = xin[i] - xDesign[i]; s += d * d; }
            if (std::sqrt(s) < 1e-16) return false;
        }
        std::copy(xin, xin + sim.ne, xDesign.begin());
        for (size_t i = 0; i < filters.size(); ++i) filters[i]->apply(cached[i], cached[i + 1]);
        objective->updateCache(cached.back());
        varsAreSet = true;
        return true;
    }
    Real evaluateOCConstraintAtVars(std::vector<Real> &xv, std::vector<Real> &scratch) const {
        if (constraints.size() != 1) throw std::runtime_error("Applicable only for a topology optimization with a single (volume) constraint");
        for (const auto &fl : filters) { fl->apply(xv, scratch); xv.swap(scratch); }
        return constraints[0]->evaluate(xv);
    }
    Real evaluateObjective() const { if (!varsAreSet) throw std::runtime_error("Must call setVars first!"); return objective->compliance(); }
    void evaluateObjectiveGradient(std::vector<Real> &grad) const {
        if (!varsAreSet) throw std::runtime_error("Must call setVars first!");
        size_t nf = filters.size();
        objective->gradient(grad);
        std::vector<Real> old(sim.ne);
        for (size_t i = 0; i < nf; ++i) { grad.swap(old); filters[nf - 1 - i]->backprop(old, cached[nf - 1 - i], grad); }
    }
    void evaluateConstraints(std::vector<Real> &c) const {
        if (!varsAreSet) throw std::runtime_error("Must call setVars first!");
        c.resize(constraints.size());
        for (size_t i = 0; i < constraints.size(); ++i) c[i] = constraints[i]->evaluate(cached.back());
    }
    void evaluateConstraintJacobianRow(size_t n, std::vector<Real> &row) const {
        if (!varsAreSet) throw std::runtime_error("Must call setVars first!");
        size_t nf = filters.size();
        std::vector<Real> d(sim.ne), dOld(sim.ne);
        constraints[n]->backprop(cached.back(), dOld);
        for (size_t i = 0; i < nf; ++i) { filters[nf - 1 - i]->backprop(dOld, cached[nf - 1 - i], d); dOld.swap(d); }
        row = dOld;
    }
};

// OptimalityCriterion.hh:32-81
template <int N>
struct OCOptimizer {
    TOProblem<N> &p; Real lambda_min = 1, lambda_max = 2; Real lastLambda = 0; int lastBisections = 0;
    explicit OCOptimizer(TOProblem<N> &prob) : p(prob) {}
    void step(Real m = 0.2, Real ctol = 1e-6) {
        std::vector<Real> dJ, dc; p.evaluateObjectiveGradient(dJ); p.evaluateConstraintJacobianRow(0, dc);
        const std::vector<Real> x0 = p.cached[0];
        const size_t n = x0.size();
        std::vector<Real> xs(n), scratch(n);
        auto stepped = [&](Real lambda, std::vector<Real> &out) {
            out.resize(n);
            for (size_t i = 0; i < n; ++i) {
                Real v = x0[i] * std::sqrt(dJ[i] / (dc[i] * lambda));
                v = std::max(v, x0[i] - m); v = std::max(v, 0.0); v = std::min(v, x0[i] + m); v = std::min(v, 1.0);
                out[i] = v;
            }
        };
        auto ceval = [&](Real lambda) { stepped(lambda, xs); ++lastBisections; return p.evaluateOCConstraintAtVars(xs, scratch); };
        lastBisections = 0;
        while (ceval(lambda_min) > 0) { lambda_max = lambda_min; lambda_min /= 2; }
        while (ceval(lambda_max) < 0) { lambda_min = lambda_max; lambda_max *= 2; }
        Real mid = 0.5 * (lambda_min + lambda_max);
        Real vol = ceval(mid);
        while (std::abs(vol) > ctol) {
            if (vol < 0) lambda_min = mid;
            if (vol > 0) lambda_max = mid;
            mid = 0.5 * (lambda_min + lambda_max);
            vol = ceval(mid);
        }
        stepped(mid, xs);
        p.setVars(xs.data());
        lastLambda = mid;
    }
};

} // namespace orc
