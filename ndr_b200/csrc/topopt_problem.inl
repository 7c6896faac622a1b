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
    const size_t tail = size_t(ghostSpan()) * planeElems(); // room for the layers fetched from the rank above (radius <= ghost span)
    cached.resize(filters.size() + 1);
    for (auto &c : cached) { c.alloc(ne + tail); c.zero(sim->stream); }
    s0.alloc(ne + tail); s1.alloc(ne + tail); s2.alloc(ne + tail); s3.alloc(ne + tail);
    if (slab())
        for (const FilterDesc &f : filters)
            if (f.type == FILTER_CALLBACK) throw std::runtime_error("callback filters are not available on a slab-decomposed problem");
}

template <int N>
void Problem<N>::refreshGhostBelow(Real *v) {
    if (!slab()) return;
    const size_t cnt = size_t(ghostSpan()) * planeElems();
    obj->mg->comm->shift(v + numVars() - cnt, v, v, v, cnt, true, false, sim->stream);
}
template <int N>
void Problem<N>::fetchAbove(Real *v, int layers) {
    if (!slab() || layers <= 0) return;
    const size_t cnt = size_t(layers) * planeElems();
    obj->mg->comm->shift(v, v, v + ownedOffset(), v + numVars(), cnt, false, true, sim->stream);
}
template <int N>
Real Problem<N>::reduceOwned(int slot) {
    Reducer &red = obj->red;
    if (slab()) obj->mg->comm->allreduceSum(red.scal.p + slot, 1, sim->stream);
    return red.fetch(slot, sim->stream);
}

template <int N>
void Problem<N>::applyFilter(const FilterDesc &f, const Real *in, Real *out) {
    if (slab() && f.type == FILTER_SMOOTHING) { slabSmoothing(f, false, const_cast<Real *>(in), out); return; }
    launchFilter(dimsOf(*sim), f, false, in, nullptr, out, numVars(), sim->stream, &hostA, &hostB, &hostC);
}
template <int N>
void Problem<N>::backpropFilter(const FilterDesc &f, const Real *in, const Real *vars, Real *out) {
    if (slab() && f.type == FILTER_SMOOTHING) { slabSmoothing(f, true, const_cast<Real *>(in), out); return; }
    launchFilter(dimsOf(*sim), f, true, in, vars, out, numVars(), sim->stream, &hostA, &hostB, &hostC);
}
// box filter (or its transpose) on a slab: `in` is one of the problem's own buffers (consistent ghost layers, room for the tail)
template <int N>
void Problem<N>::slabSmoothing(const FilterDesc &f, bool backprop, Real *in, Real *out) {
    const int r = int(size_t(f.param));
    if (r > ghostSpan()) throw std::runtime_error("slab-decomposed smoothing filter: the radius must not exceed 2^numCoarseningLevels element layers");
    fetchAbove(in, r);
    Dims D = dimsOf(*sim);
    D.E[0] = sim->globalE0;
    launchFilter(D, f, backprop, in, nullptr, out + ownedOffset(), ownedCount(), sim->stream, &hostA, &hostB, &hostC, sim->layerLo,
                 sim->layerLo + ghostLayers());
    refreshGhostBelow(out);
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
        diffSqKernel<<<reduceBlocks(ownedCount()), kRedBlock, 0, sim->stream>>>(xd + ownedOffset(), cached[0].p + ownedOffset(), ownedCount(), red.rw, red.scal.p + 6);
        VFEM_CHECK_LAUNCH(); countLaunch();
        if (std::sqrt(reduceOwned(6)) < 1e-16) return false;
    }
    if (xd != cached[0].p) VFEM_CUDA(cudaMemcpyAsync(cached[0].p, xd, ne * sizeof(Real), cudaMemcpyDeviceToDevice, sim->stream));
    refreshGhostBelow(cached[0].p); // the caller's ghost layers are not trusted
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
    Real *cur = s2.p, *other = s3.p; // the problem's own buffers: a slab filter needs their tail
    obj->gradient(cur);
    refreshGhostBelow(cur); // ghost elements see stale displacement planes: their sensitivities come from the owner
    for (size_t i = 0; i < nf; ++i) { // reverse filter chain, each evaluated at its own input vars (Problem.hh:108-111)
        backpropFilter(filters[nf - 1 - i], cur, cached[nf - 1 - i].p, other);
        std::swap(cur, other);
    }
    VFEM_CUDA(cudaMemcpyAsync(gradDev, cur, numVars() * sizeof(Real), cudaMemcpyDeviceToDevice, sim->stream));
}

template <int N>
Real Problem<N>::meanOf(const Real *v) {
    Reducer &red = obj->red;
    const size_t ne = ownedCount();
    sumKernel<<<reduceBlocks(ne), kRedBlock, 0, sim->stream>>>(v + ownedOffset(), ne, red.rw, red.scal.p + 7);
    VFEM_CHECK_LAUNCH(); countLaunch();
    return reduceOwned(7) / Real(slab() ? globalVars() : ne);
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
    Real *cur = s2.p, *other = s3.p;
    fillKernel<<<eblocks(ne), kRedBlock, 0, sim->stream>>>(cur, -1.0 / (volumeFractions.at(ci) * Real(slab() ? globalVars() : ne)), ne); // Constraint.hh:29-31
    VFEM_CHECK_LAUNCH(); countLaunch();
    for (size_t i = 0; i < nf; ++i) {
        backpropFilter(filters[nf - 1 - i], cur, cached[nf - 1 - i].p, other);
        std::swap(cur, other);
    }
    VFEM_CUDA(cudaMemcpyAsync(rowDev, cur, ne * sizeof(Real), cudaMemcpyDeviceToDevice, sim->stream));
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
