// ndr_b200 — Objective / Problem / OC member functions (included by topopt.cu inside namespace vfem).

template <int N>
Objective<N>::Objective(std::shared_ptr<MG<N>> mg_) : sim(mg_->sim), mg(mg_) {
    dim = N;
    const size_t nd = sim->numDofs();
    std::vector<Real> fh(nd);
    sim->buildLoadVectorHost(fh.data());
    f.upload(fh, sim->stream);
    u.alloc(nd); u.zero(sim->stream);          // Objective.hh:80
    red.init(sim->stream);
    updateCache(sim->d_rho.p);                  // Objective.hh:81: a full solve with the default knobs
}

template <int N>
Objective<N>::Objective(std::shared_ptr<Sim<N>> sim_) : sim(sim_) {
    dim = N;
    const size_t nd = sim->numDofs();
    std::vector<Real> fh(nd);
    sim->buildLoadVectorHost(fh.data());       // Objective.hh:32
    f.upload(fh, sim->stream);
    u.alloc(nd); u.zero(sim->stream);
    red.init(sim->stream);
    updateCache(sim->d_rho.p);                  // Objective.hh:33
}

template <int N>
void Objective<N>::updateCache(const Real *xPhysDev) {
    if (xPhysDev != sim->d_rho.p) sim->setDensities(xPhysDev, MS_DEVICE);
    else sim->touchStiffness();
    if (!mg) { sim->solveDirect(f.p, u.p); lastIterations = 0; lastResidual = 0; return; } // Objective.hh:48-51
    if (zeroInit) u.zero(sim->stream);
    typename MG<N>::PcgCb cb;
    if (residual_cb) cb = [this](size_t it, const Real *, const Real *, Real rnorm) { residual_cb(residual_ctx, it, rnorm); };
    lastIterations = mg->pcg(u.p, f.p, cgIter, tol, cb, mgIterations, mgSmoothingIterations, fullMultigrid, u.p, &lastResidual);
}

template <int N>
Real Objective<N>::compliance() {
    if (mg) return 0.5 * mg->dot(f.p, u.p, sim->numDofs());
    const size_t n = sim->numDofs();
    dotKernel<<<reduceBlocks(n), kRedBlock, 0, sim->stream>>>(f.p, u.p, n, red.rw, red.scal.p + 5);
    VFEM_CHECK_LAUNCH(); countLaunch();
    return 0.5 * red.fetch(5, sim->stream);
}

template <int N>
Problem<N>::Problem(std::shared_ptr<Sim<N>> s, std::shared_ptr<Objective<N>> o, const std::vector<Real> &volFracs,
                    const std::vector<FilterDesc> &fl)
    : sim(s), obj(o), volumeFractions(volFracs), filters(fl) {
    dim = N;
    const size_t ne = sim->numElements();
    cached.resize(filters.size() + 1);
    for (auto &c : cached) { c.alloc(ne); c.zero(sim->stream); }
    s0.alloc(ne); s1.alloc(ne); s2.alloc(ne); s3.alloc(ne);
}

template <int N>
void Problem<N>::applyFilter(const FilterDesc &f, const Real *in, Real *out) {
    launchFilter(dimsOf(*sim), f, false, in, nullptr, out, numVars(), sim->stream, &hostA, &hostB, &hostC);
}
template <int N>
void Problem<N>::backpropFilter(const FilterDesc &f, const Real *in, const Real *vars, Real *out) {
    launchFilter(dimsOf(*sim), f, true, in, vars, out, numVars(), sim->stream, &hostA, &hostB, &hostC);
}

template <int N>
bool Problem<N>::setVars(const Real *x, MemSpace ms, bool forceUpdate) {
    const size_t ne = numVars();
    Reducer &red = obj->red;
    const Real *xd = x;
    if (ms == MS_HOST) { // stage the incoming design variables
        VFEM_CUDA(cudaMemcpyAsync(s0.p, x, ne * sizeof(Real), cudaMemcpyHostToDevice, sim->stream));
        xd = s0.p;
    }
    if (!forceUpdate && varsAreSet) { // Problem.hh:50-51
        diffSqKernel<<<reduceBlocks(ne), kRedBlock, 0, sim->stream>>>(xd, cached[0].p, ne, red.rw, red.scal.p + 6);
        VFEM_CHECK_LAUNCH(); countLaunch();
        if (std::sqrt(red.fetch(6, sim->stream)) < 1e-16) return false;
    }
    if (xd != cached[0].p) VFEM_CUDA(cudaMemcpyAsync(cached[0].p, xd, ne * sizeof(Real), cudaMemcpyDeviceToDevice, sim->stream));
    for (size_t i = 0; i < filters.size(); ++i) applyFilter(filters[i], cached[i].p, cached[i + 1].p);
    obj->updateCache(cached.back().p);
    varsAreSet = true;
    return true;
}

template <int N>
Real Problem<N>::evaluateObjective() {
    if (!varsAreSet) throw std::runtime_error("Must call setVars first!");
    return obj->compliance();
}

template <int N>
void Problem<N>::evaluateObjectiveGradient(Real *gradDev) {
    if (!varsAreSet) throw std::runtime_error("Must call setVars first!");
    const size_t nf = filters.size();
    Real *cur = (nf % 2 == 0) ? gradDev : s3.p, *other = (nf % 2 == 0) ? s3.p : gradDev;
    obj->gradient(cur);
    for (size_t i = 0; i < nf; ++i) { // reverse filter chain, each evaluated at its own input vars (Problem.hh:108-111)
        backpropFilter(filters[nf - 1 - i], cur, cached[nf - 1 - i].p, other);
        std::swap(cur, other);
    }
}

template <int N>
Real Problem<N>::meanOf(const Real *v) {
    Reducer &red = obj->red;
    const size_t ne = numVars();
    sumKernel<<<reduceBlocks(ne), kRedBlock, 0, sim->stream>>>(v, ne, red.rw, red.scal.p + 7);
    VFEM_CHECK_LAUNCH(); countLaunch();
    return red.fetch(7, sim->stream) / Real(ne);
}

template <int N>
Real Problem<N>::evaluateConstraint(size_t i) { // 1 - mean(xPhys)/V  (Constraint.hh:25-27)
    if (!varsAreSet) throw std::runtime_error("Must call setVars first!");
    return 1.0 - meanOf(cached.back().p) / volumeFractions.at(i);
}

template <int N>
void Problem<N>::evaluateConstraintJacobianRow(size_t ci, Real *rowDev) {
    if (!varsAreSet) throw std::runtime_error("Must call setVars first!");
    const size_t nf = filters.size(), ne = numVars();
    Real *cur = (nf % 2 == 0) ? rowDev : s3.p, *other = (nf % 2 == 0) ? s3.p : rowDev;
    fillKernel<<<eblocks(ne), kRedBlock, 0, sim->stream>>>(cur, -1.0 / (volumeFractions.at(ci) * Real(ne)), ne); // Constraint.hh:29-31
    VFEM_CHECK_LAUNCH(); countLaunch();
    for (size_t i = 0; i < nf; ++i) {
        backpropFilter(filters[nf - 1 - i], cur, cached[nf - 1 - i].p, other);
        std::swap(cur, other);
    }
}

// constraint value of the OC candidate x(lambda) after the full filter chain; leaves x(lambda) in xOut
template <int N>
Real Problem<N>::ocConstraintAt(Real lambda, Real m, const Real *dJ, const Real *dc, Real *xOut) {
    const size_t ne = numVars();
    ocCandidateKernel<<<eblocks(ne), kRedBlock, 0, sim->stream>>>(cached[0].p, dJ, dc, lambda, m, xOut, ne);
    VFEM_CHECK_LAUNCH(); countLaunch();
    ++lastBisections;
    const Real *cur = xOut;
    Real *pp[2] = {s2.p, s3.p};
    for (size_t i = 0; i < filters.size(); ++i) { applyFilter(filters[i], cur, pp[i & 1]); cur = pp[i & 1]; }
    return 1.0 - meanOf(cur) / volumeFractions[0];
}

template <int N>
void Problem<N>::ocStep(Real m, Real ctol) {
    if (volumeFractions.size() != 1) throw std::runtime_error("Applicable only for a topology optimization with a single (volume) constraint");
    const size_t ne = numVars();
    DevBuf<Real> dJ(ne), dc(ne);
    evaluateObjectiveGradient(dJ.p);
    evaluateConstraintJacobianRow(0, dc.p);
    lastBisections = 0;
    Real *xs = s1.p;
    auto ceval = [&](Real lambda) { return ocConstraintAt(lambda, m, dJ.p, dc.p, xs); };
    while (ceval(lambdaMin) > 0) { lambdaMax = lambdaMin; lambdaMin /= 2; }   // OC.hh:60
    while (ceval(lambdaMax) < 0) { lambdaMin = lambdaMax; lambdaMax *= 2; }   // OC.hh:61
    Real mid = 0.5 * (lambdaMin + lambdaMax);
    Real vol = ceval(mid);
    while (std::abs(vol) > ctol) {                                            // OC.hh:64-69
        if (vol < 0) lambdaMin = mid;
        if (vol > 0) lambdaMax = mid;
        mid = 0.5 * (lambdaMin + lambdaMax);
        vol = ceval(mid);
    }
    // xs holds x(mid) from the last evaluation
    setVars(xs, MS_DEVICE, false);
    lastLambda = mid;
}
