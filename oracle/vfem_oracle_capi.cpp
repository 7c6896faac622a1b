// ORACLE — TEST INFRASTRUCTURE ONLY (see vfem_oracle.hpp).  C entry points for ctypes.
#include "vfem_oracle_topopt.hpp"
#ifdef _OPENMP
#include <omp.h>
#endif

using namespace orc;

namespace {
thread_local std::string g_err;

struct CtxBase { int dim; virtual ~CtxBase() {} };
template <int N>
struct Ctx : CtxBase {
    std::shared_ptr<Sim<N>> sim;
    std::shared_ptr<MG<N>> mg;
    std::shared_ptr<ComplianceObjective<N>> obj;
    std::shared_ptr<TOProblem<N>> top;
    std::shared_ptr<OCOptimizer<N>> oc;
};
} // namespace

#define ORC_TRY try {
#define ORC_CATCH } catch (const std::exception &e) { g_err = e.what(); return 1; } return 0;
#define DISPATCH(h, BODY) do { CtxBase *cb_ = (CtxBase *)(h); \
    if (cb_->dim == 2) { auto &c = *static_cast<Ctx<2> *>(cb_); constexpr int N = 2; (void)N; BODY } \
    else { auto &c = *static_cast<Ctx<3> *>(cb_); constexpr int N = 3; (void)N; BODY } } while (0)

extern "C" {

const char *orc_last_error() { return g_err.c_str(); }
void orc_set_num_threads(int n) {
#ifdef _OPENMP
    omp_set_num_threads(n);
#else
    (void)n;
#endif
}
int orc_max_threads() {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

int orc_sim_create(int dim, const double *bbmin, const double *bbmax, const size_t *elems, void **out) {
    ORC_TRY
    if (dim == 2) { auto *c = new Ctx<2>(); c->dim = 2; c->sim = std::make_shared<Sim<2>>(bbmin, bbmax, elems); *out = c; }
    else if (dim == 3) { auto *c = new Ctx<3>(); c->dim = 3; c->sim = std::make_shared<Sim<3>>(bbmin, bbmax, elems); *out = c; }
    else throw std::runtime_error("No template instantiation matching degreesPerDimension!");
    ORC_CATCH
}
void orc_sim_destroy(void *h) { delete (CtxBase *)h; }
size_t orc_num_nodes(void *h) { size_t r = 0; DISPATCH(h, r = c.sim->nn;); return r; }
size_t orc_num_elements(void *h) { size_t r = 0; DISPATCH(h, r = c.sim->ne;); return r; }
int orc_set_isotropic(void *h, double E, double nu) { ORC_TRY DISPATCH(h, c.sim->setIsotropic(E, nu);); ORC_CATCH }
int orc_set_simp(void *h, double E0, double Emin, double gamma) { ORC_TRY DISPATCH(h, c.sim->E0 = E0; c.sim->Emin = Emin; c.sim->gamma = gamma;); ORC_CATCH }
int orc_add_force(void *h, const double *lo, const double *hi, const double *val) { ORC_TRY DISPATCH(h, c.sim->addForceRegion(lo, hi, val);); ORC_CATCH }
int orc_add_dirichlet(void *h, const double *lo, const double *hi, const uint8_t *mask, const double *val) {
    ORC_TRY DISPATCH(h, c.sim->addDirichletRegion(lo, hi, mask, val);); ORC_CATCH }
int orc_set_densities(void *h, const double *rho) { ORC_TRY DISPATCH(h, std::copy(rho, rho + c.sim->ne, c.sim->rho.begin()); c.sim->factorUpToDate = false;); ORC_CATCH }
int orc_get_densities(void *h, double *rho) { ORC_TRY DISPATCH(h, std::copy(c.sim->rho.begin(), c.sim->rho.end(), rho);); ORC_CATCH }
int orc_get_k0(void *h, double *out) { ORC_TRY DISPATCH(h, std::copy(c.sim->K0, c.sim->K0 + Sim<N>::KS * Sim<N>::KS, out);); ORC_CATCH }
int orc_load_vector(void *h, double *f) { ORC_TRY DISPATCH(h, c.sim->buildLoadVector(f);); ORC_CATCH }
int orc_apply_k(void *h, const double *u, double *out) { ORC_TRY DISPATCH(h, c.sim->applyK(u, out);); ORC_CATCH }
int orc_solve_direct(void *h, const double *f, double *x) { ORC_TRY DISPATCH(h, c.sim->solveDirect(f, x);); ORC_CATCH }
int orc_compliance_gradient(void *h, const double *u, double *g) { ORC_TRY DISPATCH(h, c.sim->complianceGradient(u, g);); ORC_CATCH }
int orc_node_position(void *h, size_t ni, double *p) { ORC_TRY DISPATCH(h, c.sim->nodePosition(ni, p);); ORC_CATCH }

// ---- multigrid ----
int orc_mg_create(void *h, size_t levels) { ORC_TRY DISPATCH(h, c.mg = std::make_shared<MG<N>>(c.sim, levels);); ORC_CATCH }
int orc_mg_set_flags(void *h, int symGS, int buildBlock, int buildFinest) {
    ORC_TRY DISPATCH(h, c.mg->symmetricGS = symGS; c.mg->buildBlockStiffnessMatrices = buildBlock; c.mg->buildFinestBlockStiffnessMatrix = buildFinest;); ORC_CATCH }
size_t orc_mg_level_nodes(void *h, size_t l) { size_t r = 0; DISPATCH(h, r = c.mg->sims.at(l)->nn;); return r; }
size_t orc_mg_level_elements(void *h, size_t l) { size_t r = 0; DISPATCH(h, r = c.mg->sims.at(l)->ne;); return r; }
int orc_mg_level_dmask(void *h, size_t l, uint8_t *out) { ORC_TRY DISPATCH(h, auto &s = *c.mg->sims.at(l); std::copy(s.dmask.begin(), s.dmask.end(), out);); ORC_CATCH }
int orc_mg_level_dvals(void *h, size_t l, double *out) { ORC_TRY DISPATCH(h, auto &s = *c.mg->sims.at(l); std::copy(s.dval.begin(), s.dval.end(), out);); ORC_CATCH }
int orc_mg_update_stiffness(void *h) { ORC_TRY DISPATCH(h, c.mg->updateElementStiffnessMatrices();); ORC_CATCH }
int orc_mg_update_blockks(void *h) { ORC_TRY DISPATCH(h, c.mg->updateBlockKs();); ORC_CATCH }
int orc_mg_level_ke(void *h, size_t l, size_t e, double *out) {
    ORC_TRY DISPATCH(h, auto &s = *c.mg->sims.at(l); if (!s.hasCachedKe()) throw std::runtime_error("no cached Ke");
                     std::copy(s.elementKe(e), s.elementKe(e) + Sim<N>::KS * Sim<N>::KS, out);); ORC_CATCH }
int orc_mg_level_stencil(void *h, size_t l, double *out) {
    ORC_TRY DISPATCH(h, auto &s = *c.mg->sims.at(l); if (!s.hasBlockK) throw std::runtime_error("no blockK");
                     std::copy(s.stencil.begin(), s.stencil.end(), out);); ORC_CATCH }
int orc_mg_apply_k(void *h, size_t l, const double *u, double *out) { ORC_TRY DISPATCH(h, c.mg->applyK(l, u, out);); ORC_CATCH }
int orc_mg_residual(void *h, size_t l, const double *u, const double *b, double *out) { ORC_TRY DISPATCH(h, c.mg->computeResidual(l, u, b, out);); ORC_CATCH }
int orc_mg_smooth_colored(void *h, size_t l, double *u, const double *b, int forward) { ORC_TRY DISPATCH(h, c.mg->smoothingMulticoloredGS(l, u, b, forward != 0);); ORC_CATCH }
int orc_mg_smooth_serial(void *h, size_t l, double *u, const double *b, int forward) { ORC_TRY DISPATCH(h, c.mg->smoothingSerial(l, u, b, forward != 0);); ORC_CATCH }
int orc_mg_restrict(void *h, size_t lf, const double *vf, double *outc) { ORC_TRY DISPATCH(h, c.mg->restriction(lf, vf, outc);); ORC_CATCH }
int orc_mg_interpolate(void *h, size_t lf, const double *vc, double *outf, int accum) {
    ORC_TRY DISPATCH(h, if (accum) c.mg->template interpolate<true>(lf, vc, outf); else c.mg->template interpolate<false>(lf, vc, outf);); ORC_CATCH }
int orc_mg_zero_dirichlet(void *h, size_t l, double *u) { ORC_TRY DISPATCH(h, c.mg->zeroOutDirichletComponents(l, u);); ORC_CATCH }
int orc_mg_color_order(void *h, int *order) { ORC_TRY DISPATCH(h, c.mg->debugMulticolorVisit(order);); ORC_CATCH }
int orc_mg_solve(void *h, const double *u, const double *f, size_t numSteps, size_t nu, int stiffnessUpdated, int zeroDirichlet, int fmg, double *out) {
    ORC_TRY DISPATCH(h, c.mg->solve(u, f, numSteps, nu, stiffnessUpdated != 0, zeroDirichlet != 0, nullptr, fmg != 0, out);); ORC_CATCH }
int orc_mg_debug_xb(void *h, size_t l, double *x, double *b) {
    ORC_TRY DISPATCH(h, std::copy(c.mg->x.at(l).begin(), c.mg->x.at(l).end(), x); std::copy(c.mg->b.at(l).begin(), c.mg->b.at(l).end(), b);); ORC_CATCH }
// resid_hist (capacity hist_cap) receives ||r|| after every iteration
int orc_mg_pcg(void *h, const double *u, const double *b, size_t maxIter, double tol, size_t mgIt, size_t nu, int fmg,
               double *out, size_t *iters, double *resid_hist, size_t hist_cap) {
    ORC_TRY DISPATCH(h,
        size_t n = c.sim->nn * N;
        typename MG<N>::PcgCb cb;
        if (resid_hist) cb = [&](size_t it, const Real *, const Real *r) {
            Real s = 0; for (size_t i = 0; i < n; ++i) s += r[i] * r[i];
            if (it - 1 < hist_cap) resid_hist[it - 1] = std::sqrt(s); };
        size_t k = c.mg->pcg(u, b, maxIter, tol, cb, mgIt, nu, fmg != 0, out);
        if (iters) *iters = k;);
    ORC_CATCH }

// ---- topology optimisation ----
// filter_types: 0 = Smoothing(param = radius), 1 = Projection(param = beta)
int orc_top_create(void *h, int use_mg, double volfrac, int nfilters, const int *ftypes, const double *fparams) {
    ORC_TRY DISPATCH(h,
        c.obj = std::make_shared<ComplianceObjective<N>>(*c.sim, use_mg ? c.mg : nullptr);
        std::vector<std::shared_ptr<Filter>> fl;
        for (int i = 0; i < nfilters; ++i) {
            if (ftypes[i] == 0) { auto f = std::make_shared<SmoothingFilter>(); f->radius = size_t(fparams[i]); fl.push_back(f); }
            else { auto f = std::make_shared<ProjectionFilter>(); f->beta = fparams[i]; fl.push_back(f); }
        }
        std::vector<std::shared_ptr<TotalVolumeConstraint>> cs{std::make_shared<TotalVolumeConstraint>(volfrac)};
        c.top = std::make_shared<TOProblem<N>>(*c.sim, c.obj, cs, fl);
        c.oc = std::make_shared<OCOptimizer<N>>(*c.top););
    ORC_CATCH }
int orc_top_set_solver(void *h, size_t cgIter, double tol, size_t mgIt, size_t nu, int fmg, int zeroInit) {
    ORC_TRY DISPATCH(h, c.obj->cgIter = cgIter; c.obj->tol = tol; c.obj->mgIterations = mgIt; c.obj->mgSmoothingIterations = nu;
                     c.obj->fullMultigrid = fmg != 0; c.obj->zeroInit = zeroInit != 0;); ORC_CATCH }
int orc_top_set_vars(void *h, const double *x, int force, int *changed) { ORC_TRY DISPATCH(h, bool ch = c.top->setVars(x, force != 0); if (changed) *changed = ch;); ORC_CATCH }
int orc_top_objective(void *h, double *J) { ORC_TRY DISPATCH(h, *J = c.top->evaluateObjective();); ORC_CATCH }
int orc_top_gradient(void *h, double *g) { ORC_TRY DISPATCH(h, std::vector<Real> v; c.top->evaluateObjectiveGradient(v); std::copy(v.begin(), v.end(), g);); ORC_CATCH }
int orc_top_constraint(void *h, double *cv) { ORC_TRY DISPATCH(h, std::vector<Real> v; c.top->evaluateConstraints(v); *cv = v[0];); ORC_CATCH }
int orc_top_constraint_jac(void *h, double *row) { ORC_TRY DISPATCH(h, std::vector<Real> v; c.top->evaluateConstraintJacobianRow(0, v); std::copy(v.begin(), v.end(), row);); ORC_CATCH }
int orc_top_get_u(void *h, double *u) { ORC_TRY DISPATCH(h, std::copy(c.obj->u.begin(), c.obj->u.end(), u);); ORC_CATCH }
int orc_top_get_vars(void *h, int which, double *out) { ORC_TRY DISPATCH(h, auto &v = which < 0 ? c.top->cached.back() : c.top->cached.at(which); std::copy(v.begin(), v.end(), out);); ORC_CATCH }
size_t orc_top_last_iters(void *h) { size_t r = 0; DISPATCH(h, r = c.obj->lastIters;); return r; }
int orc_oc_step(void *h, double m, double ctol, double *lambda, int *nbisect) {
    ORC_TRY DISPATCH(h, c.oc->step(m, ctol); if (lambda) *lambda = c.oc->lastLambda; if (nbisect) *nbisect = c.oc->lastBisections;); ORC_CATCH }

// standalone filter application (pyVoxelFEM.applyFilter, VoxelFEM.cc:275-281)
int orc_filter(int type, double param, const size_t *dims, int ndim, const double *x, const double *vars, int backprop, double *out) {
    ORC_TRY
    std::vector<size_t> d(dims, dims + ndim); size_t n = 1; for (size_t v : d) n *= v;
    std::shared_ptr<Filter> f;
    if (type == 0) { auto s = std::make_shared<SmoothingFilter>(); s->radius = size_t(param); f = s; }
    else { auto p = std::make_shared<ProjectionFilter>(); p->beta = param; f = p; }
    f->setGridDimensions(d);
    std::vector<Real> in(x, x + n), o, v;
    if (backprop) { v.assign(vars, vars + n); f->backprop(in, v, o); } else f->apply(in, o);
    std::copy(o.begin(), o.end(), out);
    ORC_CATCH
}

// tensor-product Gauss quadrature of a monomial (TensorProductQuadrature.hh): rule exact to ruleDegree[d] per axis
int orc_gauss_monomial(int ndim, const int *ruleDegree, const int *expo, double *out) {
    ORC_TRY
    Real result = 1;
    for (int d = 0; d < ndim; ++d) {
        std::vector<Real> x, w; gaussRule(gaussPointsForDegree(ruleDegree[d]), x, w);
        Real s = 0; for (size_t i = 0; i < x.size(); ++i) s += w[i] * std::pow(x[i], expo[d]);
        result *= s;
    }
    *out = result;
    ORC_CATCH
}

} // extern "C"
