// ndr_b200 — V-cycle, full multigrid and MG-preconditioned CG (included by mg.cu inside namespace vfem).

// vcycle (MG.hh:516-553)
template <int N>
void MG<N>::vcycle(size_t l, size_t nu, bool residualSystem) {
    const size_t coarsest = lv.size() - 1;
    Level &L = lv[l];
    if (l == coarsest) { coarseSolve(L.b.p, L.x.p); return; }
    enforceDirichlet(l, L.x.p, residualSystem);
    smooth(l, L.x.p, L.b.p, true, nu);
    computeResidual(l, L.x.p, L.b.p, L.r.p);
    restriction(l, L.r.p, lv[l + 1].b.p);
    lv[l + 1].x.zero(stream);
    vcycle(l + 1, nu, true);
    interpolate(l, lv[l + 1].x.p, L.x.p, true);
    smooth(l, L.x.p, L.b.p, !symmetricGS, nu);
}

// fullMultigrid (MG.hh:486-508)
template <int N>
void MG<N>::fullMultigrid(size_t l, size_t nu, bool residualSystem) {
    const size_t coarsest = lv.size() - 1;
    Level &L = lv[l];
    if (l == coarsest) { coarseSolve(L.b.p, L.x.p); return; }
    restriction(l, L.b.p, lv[l + 1].b.p);
    fullMultigrid(l + 1, nu, residualSystem);
    interpolate(l, lv[l + 1].x.p, L.x.p, false);
    vcycle(l, nu, residualSystem);
}

// solve (MG.hh:446-472).  u may be nullptr (= zero initial guess); out may alias lv[0].x.
template <int N>
void MG<N>::solve(const Real *u, const Real *f, size_t numSteps, size_t nu, bool stiffnessUpdated, bool zeroDirichlet,
                  const SolveCb &cb, bool fmg, Real *out) {
    const size_t bytes = levelDofs(0) * sizeof(Real);
    syncMasks();
    if (!stiffnessUpdated) updateElementStiffnessMatrices(true);
    if (numSteps == 0) {
        if (u && out != u) VFEM_CUDA(cudaMemcpyAsync(out, u, bytes, cudaMemcpyDeviceToDevice, stream));
        if (!u) VFEM_CUDA(cudaMemsetAsync(out, 0, bytes, stream));
        return;
    }
    Level &L0 = lv[0];
    if (u) { if (u != L0.x.p) VFEM_CUDA(cudaMemcpyAsync(L0.x.p, u, bytes, cudaMemcpyDeviceToDevice, stream)); }
    else L0.x.zero(stream);
    if (f != L0.b.p) VFEM_CUDA(cudaMemcpyAsync(L0.b.p, f, bytes, cudaMemcpyDeviceToDevice, stream));
    size_t first = 0;
    if (fmg) { fullMultigrid(0, nu, zeroDirichlet); if (cb) cb(0, L0.x.p); first = 1; }
    for (size_t i = first; i < numSteps; ++i) { vcycle(0, nu, zeroDirichlet); if (cb) cb(i, L0.x.p); }
    if (out != L0.x.p) VFEM_CUDA(cudaMemcpyAsync(out, L0.x.p, bytes, cudaMemcpyDeviceToDevice, stream));
}

// applyPreconditionerInv for nu > 0 (MG.hh:476-479): solve(0, f, steps, nu, true, true, null, fmg) into lv[0].x
template <int N>
void MG<N>::precondition(const Real *f, size_t steps, size_t nu, bool fmg) {
    if (!useGraphs) { solve(nullptr, f, steps, nu, true, true, nullptr, fmg, lv[0].x.p); return; }
    PrecondKey key; key.f = f; key.steps = steps; key.nu = nu; key.fmg = fmg; key.sym = symmetricGS; key.P = coarseP; key.planes = coarsePlanes;
    key.k0Version = sim->k0Version; key.applyVariant = sim->applyVariant; key.gsVariant = gsVariant;
    if (!pgExec || !(key == pgKey)) {
        if (pgExec) { cudaGraphExecDestroy(pgExec); pgExec = nullptr; }
        if (!captureStream) VFEM_CUDA(cudaStreamCreateWithFlags(&captureStream, cudaStreamNonBlocking));
        syncMasks();
        sim->moduli();
        VFEM_CUDA(cudaStreamSynchronize(stream));
        const cudaStream_t saved = stream, savedSim = sim->stream;
        const unsigned long long before = g_launchCount;
        stream = captureStream; sim->stream = captureStream;
        cudaGraph_t graph = nullptr;
        VFEM_CUDA(cudaStreamBeginCapture(captureStream, cudaStreamCaptureModeRelaxed));
        try {
            solve(nullptr, f, steps, nu, true, true, nullptr, fmg, lv[0].x.p);
        } catch (...) {
            cudaStreamEndCapture(captureStream, &graph);
            if (graph) cudaGraphDestroy(graph);
            stream = saved; sim->stream = savedSim;
            throw;
        }
        stream = saved; sim->stream = savedSim;
        VFEM_CUDA(cudaStreamEndCapture(captureStream, &graph));
        pgLaunches = g_launchCount - before;
        g_launchCount = before; // nothing ran yet
        VFEM_CUDA(cudaGraphInstantiate(&pgExec, graph, 0));
        cudaGraphDestroy(graph);
        pgKey = key;
    }
    VFEM_CUDA(cudaGraphLaunch(pgExec, stream));
    countLaunch(pgLaunches);
}

// deterministic dot product over the OWNED dofs of two level-0 vectors (+ all-reduce over the slabs), synchronous
template <int N>
Real MG<N>::dot(const Real *a, const Real *b, size_t n) {
    if (slab()) { const size_t off = ownedOffset(0); a += off; b += off; n = ownedDofs(0); }
    dotKernel<<<reduceBlocks(n), kRedBlock, 0, stream>>>(a, b, n, rw, d_scal.p + 5);
    VFEM_CHECK_LAUNCH(); countLaunch();
    allreduce(d_scal.p + 5, 1);
    VFEM_CUDA(cudaMemcpyAsync(h_scal + 5, d_scal.p + 5, sizeof(Real), cudaMemcpyDeviceToHost, stream));
    VFEM_CUDA(cudaStreamSynchronize(stream));
    return h_scal[5];
}

// preconditionedConjugateGradient (MG.hh:679-732).  All vectors stay in HBM; per iteration the host
// reads back one scalar (||r||^2) for the stopping test.  alpha and beta are formed on the device
// from the fused reductions, so there is no other host round trip.
// Slab mode: u, b, out are the rank-local arrays (ghost planes included; u must be consistent, out is); vector
// updates and reductions run over the owned planes, each reduction is followed by a one-scalar all-reduce, and the
// search direction gets its ghost planes refreshed before K d.
template <int N>
size_t MG<N>::pcg(const Real *u, const Real *b, size_t maxIter, Real tol, const PcgCb &cb, size_t mgIterations, size_t nu,
                  bool fmg, Real *out, Real *finalResidualNorm) {
    const size_t nAll = levelDofs(0), bytes = nAll * sizeof(Real);
    const size_t off = slab() ? ownedOffset(0) : 0, n = slab() ? ownedDofs(0) : nAll;
    const int rb = reduceBlocks(n);
    syncMasks();
    Real *x = p_x.p, *r = p_r.p, *d = p_d.p, *Ad = p_Ad.p;
    Real *sc = d_scal.p;
    VFEM_CUDA(cudaMemcpyAsync(x, u, bytes, cudaMemcpyDeviceToDevice, stream));
    if (b != p_b.p) VFEM_CUDA(cudaMemcpyAsync(p_b.p, b, bytes, cudaMemcpyDeviceToDevice, stream));
    b = p_b.p;
    enforceDirichlet(0, x, false);
    updateElementStiffnessMatrices(true); // MG.hh:688-689 (+ updateBlockKs)
    dotKernel<<<rb, kRedBlock, 0, stream>>>(b + off, b + off, n, rw, sc + 4); VFEM_CHECK_LAUNCH(); countLaunch();
    computeResidual(0, x, b, r);
    dotKernel<<<rb, kRedBlock, 0, stream>>>(r + off, r + off, n, rw, sc + 3); VFEM_CHECK_LAUNCH(); countLaunch();
    allreduce(sc + 3, 2);
    VFEM_CUDA(cudaMemcpyAsync(h_scal + 3, sc + 3, 2 * sizeof(Real), cudaMemcpyDeviceToHost, stream));
    VFEM_CUDA(cudaStreamSynchronize(stream));
    const Real bnsq = h_scal[4];
    Real rr = h_scal[3];
    size_t i = 0; // the reference leaves `i` uninitialised (MG.hh:710); 0 is the intended value
    bool first = true;
    if (slab() && first) VFEM_CUDA(cudaMemsetAsync(d, 0, bytes, stream)); // ghost planes of d are read before they are first exchanged
    while ((i++ < maxIter) && (rr > tol * tol * bnsq)) {
        Real *s;
        if (nu == 0) { // applyPreconditionerInv returns r unchanged (MG.hh:477)
            if (p_s.n != nAll) p_s.alloc(nAll);
            VFEM_CUDA(cudaMemcpyAsync(p_s.p, r, bytes, cudaMemcpyDeviceToDevice, stream));
            s = p_s.p;
        } else {
            precondition(r, mgIterations, nu, fmg);
            s = lv[0].x.p;
        }
        copyScalarKernel<<<1, 1, 0, stream>>>(sc + 1, sc + 0); VFEM_CHECK_LAUNCH(); countLaunch();
        maskDotKernel<N><<<rb, kRedBlock, 0, stream>>>(s + off, r + off, maskOf(*this, 0) + off / N, n, rw, sc + 0); VFEM_CHECK_LAUNCH(); countLaunch();
        allreduce(sc + 0, 1);
        updateDirectionKernel<<<rb, kRedBlock, 0, stream>>>(d + off, s + off, n, sc + 0, sc + 1, first ? 1 : 0); VFEM_CHECK_LAUNCH(); countLaunch();
        first = false;
        exchangeApplyFine<2>(d, nullptr, Ad);
        dotKernel<<<rb, kRedBlock, 0, stream>>>(d + off, Ad + off, n, rw, sc + 2); VFEM_CHECK_LAUNCH(); countLaunch();
        allreduce(sc + 2, 1);
        updateSolutionKernel<<<rb, kRedBlock, 0, stream>>>(x + off, r + off, d + off, Ad + off, n, sc + 0, sc + 2, rw, sc + 3); VFEM_CHECK_LAUNCH(); countLaunch();
        allreduce(sc + 3, 1);
        VFEM_CUDA(cudaMemcpyAsync(h_scal + 3, sc + 3, sizeof(Real), cudaMemcpyDeviceToHost, stream));
        VFEM_CUDA(cudaStreamSynchronize(stream));
        rr = h_scal[3];
        if (cb) cb(i, x, r, std::sqrt(rr));
    }
    exchange(0, x, XUP | XDOWN);
    if (out != x) VFEM_CUDA(cudaMemcpyAsync(out, x, bytes, cudaMemcpyDeviceToDevice, stream));
    if (finalResidualNorm) *finalResidualNorm = std::sqrt(rr);
    return i - 1;
}
