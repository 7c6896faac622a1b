// ndr_b200 — density pipeline of a topology-optimisation iteration, device resident:
// filters (TopologyOptimizationFilter.hh:55-162), volume constraint (…Constraint.hh:21-35),
// compliance objective (…Objective.hh:24-102), problem (…Problem.hh:17-204) and the
// optimality-criterion step (OptimalityCriterion.hh:32-81).
#pragma once
#include "mg.h"

namespace vfem {

enum FilterType { FILTER_SMOOTHING = 0, FILTER_PROJECTION = 1, FILTER_CALLBACK = 2 };
// host callback filter (PythonFilter, Filter.hh:81-103): out = apply(in) / out = backprop(in, vars); host arrays of length n
typedef void (*FilterCallback)(void *ctx, int backprop, const Real *in, const Real *vars, Real *out, size_t n);
struct FilterDesc {
    int type = FILTER_SMOOTHING;
    Real param = 1.0;            // smoothing radius (truncated to an integer, Filter.hh:159) or projection beta
    FilterCallback cb = nullptr;
    void *ctx = nullptr;
};

struct ObjectiveBase { int dim = 0; virtual ~ObjectiveBase() {} };
struct ProblemBase { int dim = 0; virtual ~ProblemBase() {} };

template <int N>
struct Objective : ObjectiveBase {
    std::shared_ptr<Sim<N>> sim;
    std::shared_ptr<MG<N>> mg;               // null: ComplianceObjective, the direct solve (Objective.hh:25-68)
    DevBuf<Real> u, f;
    Reducer red;                             // reductions of the density pipeline (slots 5..7)
    size_t cgIter = 100; Real tol = 1e-5; size_t mgIterations = 1, mgSmoothingIterations = 2; // Objective.hh:96-99
    bool fullMultigrid = true, zeroInit = false;
    void (*residual_cb)(void *, size_t, Real) = nullptr; void *residual_ctx = nullptr;
    size_t lastIterations = 0; Real lastResidual = 0;

    explicit Objective(std::shared_ptr<MG<N>> mg_);    // MultigridComplianceObjective (Objective.hh:70-102)
    explicit Objective(std::shared_ptr<Sim<N>> sim_);  // ComplianceObjective(simulator): every updateCache is a direct solve
    void updateCache(const Real *xPhysDev);  // Objective.hh:85-93
    Real compliance();                       // Objective.hh:39-41
    void gradient(Real *gDev) { sim->complianceGradient(u.p, gDev); } // Objective.hh:43-45
};

template <int N>
struct Problem : ProblemBase {
    std::shared_ptr<Sim<N>> sim;
    std::shared_ptr<Objective<N>> obj;
    std::vector<Real> volumeFractions;       // one TotalVolumeConstraint each
    std::vector<FilterDesc> filters;
    std::vector<DevBuf<Real>> cached;        // m_cachedVars: [0] design ... [nf] physical
    DevBuf<Real> s0, s1, s2, s3;             // scratch
    bool varsAreSet = false;
    // OC state (OptimalityCriterion.hh:78)
    Real lambdaMin = 1.0, lambdaMax = 2.0, lastLambda = 0.0; int lastBisections = 0;
    std::vector<Real> hostA, hostB, hostC;   // staging for callback filters

    // Slab decomposition ("rho/E ghost element layer"): the variables of a rank are its LOCAL element layers - `ghost`
    // layers below that mirror the lower neighbour's top layers, then the owned ones.  Pointwise passes run on all of them
    // (consistent in -> consistent out); the box filter fetches `radius` layers from the rank above into the tail every buffer
    // carries, computes the owned layers with the GLOBAL clipping, and refreshes the ghost layers from the owner; reductions
    // run over the owned layers + one all-reduce.  Single device: ghost = 0, no tail, no exchange.
    bool slab() const { return obj->mg && obj->mg->slab(); }
    size_t planeElems() const { size_t p = 1; for (int d = 1; d < N; ++d) p *= size_t(sim->g.E[d]); return p; }
    int ghostLayers() const { return slab() ? obj->mg->lv[0].g.own0 : 0; }
    int ghostSpan() const { return slab() ? 1 << (int(obj->mg->numLevels()) - 1) : 0; }     // layers a rank mirrors from below (ranks > 0)
    size_t ownedOffset() const { return size_t(ghostLayers()) * planeElems(); }
    size_t ownedCount() const { return numVars() - ownedOffset(); }
    size_t globalVars() const { return size_t(sim->globalE0) * planeElems(); }
    void refreshGhostBelow(Real *v);            // ghost layers <- the lower neighbour's top owned layers
    void fetchAbove(Real *v, int layers);       // tail of v <- the upper neighbour's first owned layers
    Real reduceOwned(int slot);                 // all-reduce + read back of a per-rank partial sum

    Problem(std::shared_ptr<Sim<N>> s, std::shared_ptr<Objective<N>> o, const std::vector<Real> &volFracs, const std::vector<FilterDesc> &fl);
    size_t numVars() const { return sim->numElements(); }
    void applyFilter(const FilterDesc &f, const Real *in, Real *out);
    void backpropFilter(const FilterDesc &f, const Real *in, const Real *vars, Real *out);
    void slabSmoothing(const FilterDesc &f, bool backprop, Real *in, Real *out);
    bool setVars(const Real *x, MemSpace ms, bool forceUpdate);          // Problem.hh:42-65
    Real evaluateObjective();                                            // Problem.hh:88-94
    void evaluateObjectiveGradient(Real *gradDev);                       // Problem.hh:98-113
    Real evaluateConstraint(size_t i);                                   // Problem.hh:117-127
    void evaluateConstraintJacobianRow(size_t i, Real *rowDev);          // Problem.hh:131-150
    Real ocConstraintAt(Real lambda, Real m, const Real *dJ, const Real *dc, Real *xOut); // OC.hh:16-30 + Problem.hh:72-84
    void ocStep(Real m, Real ctol);                                      // OC.hh:41-75
    Real meanOf(const Real *v);
};

// standalone filter application on a grid (pyVoxelFEM.applyFilter)
void applyFilterStandalone(int ndim, const size_t *dims, const FilterDesc &f, const Real *xHost, Real *outHost);

} // namespace vfem
