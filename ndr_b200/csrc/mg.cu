// ndr_b200 — MultigridSolver host logic: hierarchy setup, operator refresh, launches.
#include "mg.h"
#include "kernels.cuh"
#include "kernels_mg.cuh"
#include "gs_tile.cuh"
#include "block_tri.h"

namespace vfem {

template <int N>
MG<N>::MG(std::shared_ptr<Sim<N>> fine, size_t numCoarseningLevels, std::shared_ptr<SlabComm> comm_) : sim(fine), comm(comm_) {
    dim = N;
    const int Lc = int(numCoarseningLevels);
    if (slab()) {
        const int w = 1 << Lc, r = comm->rank, W = comm->world;
        const int ghostLo = r > 0 ? w : 0;
        if (fine->globalE0 % w || fine->layerLo % w || fine->g.E[0] % w || fine->g.E[0] <= ghostLo)
            throw std::runtime_error("slab solve: the global grid and every slab boundary must be multiples of 2^numCoarseningLevels layers");
        if ((r == 0 && fine->layerLo != 0) || (r == W - 1) != (fine->layerLo + fine->g.E[0] == fine->globalE0))
            throw std::runtime_error("slab solve: rank 0 must hold layer 0 and the last rank the last layer");
        useGraphs = false;
    }
    stream = fine->stream;
    useGraphs = getenv("VFEM_NO_GRAPH") == nullptr && (!slab() || comm->graphSafe());
    if (const char *v = getenv("VFEM_GS_VARIANT")) gsVariant = atoi(v);
    fusedSweeps = getenv("VFEM_NO_FUSED_SWEEPS") == nullptr;
    int E[N];
    for (int d = 0; d < N; ++d) E[d] = fine->g.E[d];
    lv.resize(numCoarseningLevels + 1);
    for (size_t l = 0; l <= numCoarseningLevels; ++l) {
        Level &L = lv[l];
        if (l == 0) { L.g = fine->g; L.dmask = fine->dmask; }
        else {
            if (slab() && l >= 2) exchangeHostMask(l - 1); // the coarsening below reads the ghost planes of the finer mask (level 0's come from node positions)
            for (int d = 0; d < N; ++d) {
                if (E[d] % 2 == 1)
                    throw std::runtime_error("Grid size currently must be divisible by 2^numCoarseningLevels (nonuniform coarsening not yet implemented)"); // MG.hh:36-37
                E[d] /= 2;
            }
            L.g = makeGrid<N>(E);
            L.dmask.assign(L.g.nn, 0);
            // Dirichlet coarsening (MG.hh:43-84): a fine constrained node lying on a coarse vertex /
            // edge / face constrains, with the same component mask, every coarse node of that entity.
            const Level &F = lv[l - 1];
            for (int nf = 0; nf < F.g.nn; ++nf) {
                const uint8_t m = F.dmask[nf];
                if (!m) continue;
                int fi[N], r = nf, lo[N], cnt[N]; bool interior = true;
                for (int d = N - 1; d >= 0; --d) { fi[d] = r % F.g.n[d]; r /= F.g.n[d]; }
                for (int d = 0; d < N; ++d) { lo[d] = fi[d] / 2; cnt[d] = (fi[d] & 1) ? 2 : 1; if (cnt[d] == 1) interior = false; }
                if (interior) throw std::runtime_error("Dirichlet constraints on internal nodes are not supported"); // MG.hh:70
                for (int j = 0; j < NPE; ++j) {
                    int cn = 0; bool use = true;
                    for (int d = 0; d < N; ++d) { int bit = (j >> (N - 1 - d)) & 1; if (bit >= cnt[d]) use = false; cn += (lo[d] + bit) * L.g.ninc[d]; }
                    if (use) L.dmask[cn] |= m;
                }
            }
            L.S.alloc(size_t(L.g.nn) * NST * N * N);
            L.d_dval.alloc(size_t(L.g.nn) * N); L.d_dval.zero(stream);
        }
        if (slab()) { // owned node planes of this level (see mg.h)
            L.g.own0 = comm->rank > 0 ? (1 << (Lc - int(l))) : 0;
            L.g.own1 = comm->rank + 1 < comm->world ? L.g.n[0] - 2 : L.g.n[0] - 1;
        }
        L.x.alloc(size_t(L.g.nn) * N); L.b.alloc(size_t(L.g.nn) * N); L.r.alloc(size_t(L.g.nn) * N);
        L.x.zero(stream); L.b.zero(stream);
    }
    const size_t nd = levelDofs(0);
    p_x.alloc(nd); p_r.alloc(nd); p_d.alloc(nd); p_Ad.alloc(nd); p_b.alloc(nd);
    d_scal.alloc(8); d_scal.zero(stream);
    d_partials.alloc(size_t(kRedMaxBlocks) * 4); d_ticket.alloc(1); d_ticket.zero(stream);
    rw.partials = d_partials.p; rw.ticket = d_ticket.p;
    VFEM_CUDA(cudaMallocHost((void **)&h_scal, 8 * sizeof(Real)));
    d_C.alloc(size_t(NPE) * KS * KS);
    bt.reset(new BlockTri<N>());
    gCoarse = lv.back().g;
    if (slab()) {
        sim->computeLo = lv[0].g.own0 > 1 ? lv[0].g.own0 - 1 : 0;
        // global coarsest grid + its Dirichlet mask (owned rows summed over the ranks)
        int Eg[N];
        for (int d = 0; d < N; ++d) Eg[d] = lv.back().g.E[d];
        Eg[0] = fine->globalE0 >> Lc;
        gCoarse = makeGrid<N>(Eg);
        coarsePlaneOffset = fine->layerLo >> Lc;
        const Level &C = lv.back();
        const size_t pn = size_t(C.g.ninc[0]);
        std::vector<uint8_t> mg(size_t(gCoarse.nn), 0);
        for (int p = C.g.own0; p <= C.g.own1; ++p)
            for (size_t i = 0; i < pn; ++i) mg[size_t(coarsePlaneOffset + p) * pn + i] = C.dmask[size_t(p) * pn + i];
        dmaskG.upload(mg, stream);
        comm->allreduceSumBytes(dmaskG.p, dmaskG.n, stream);
        std::vector<uint8_t> full = dmaskG.toHost(stream);
        // the local coarsest mask (ghost planes included) is a window of the global one
        Level &Cm = lv.back();
        for (int p = 0; p < Cm.g.n[0]; ++p)
            for (size_t i = 0; i < pn; ++i) Cm.dmask[size_t(p) * pn + i] = full[size_t(coarsePlaneOffset + p) * pn + i];
        Sg.alloc(size_t(gCoarse.nn) * NST * N * N); bg.alloc(size_t(gCoarse.nn) * N); xg.alloc(size_t(gCoarse.nn) * N);
    }
    for (size_t l = 1; l <= numCoarseningLevels; ++l) lv[l].d_dmask.upload(lv[l].dmask, stream);
    VFEM_CUDA(cudaStreamSynchronize(stream));
}

// ghost planes of the level-l host Dirichlet mask <- the owners' planes (masks are static: done once, through the
// device transport with the mask widened to doubles)
template <int N>
void MG<N>::exchangeHostMask(size_t l) {
    Level &L = lv[l];
    std::vector<uint8_t> &m = L.dmask;
    const size_t pn = size_t(L.g.ninc[0]);
    std::vector<Real> w(m.begin(), m.end());
    DevBuf<Real> d; d.upload(w, stream);
    const int o0 = L.g.own0, o1 = L.g.own1;
    comm->shift(d.p + size_t(o1) * pn, o0 > 0 ? d.p + size_t(o0 - 1) * pn : d.p, d.p + size_t(o0) * pn,
                o1 + 1 < L.g.n[0] ? d.p + size_t(o1 + 1) * pn : d.p, pn, true, true, stream);
    d.download(w.data(), stream);
    for (size_t i = 0; i < w.size(); ++i) m[i] = uint8_t(w[i]);
}

template <int N>
void MG<N>::exchange(size_t l, Real *v, int dirs) {
    if (!slab()) return;
    const Level &L = lv[l];
    const size_t pd = size_t(L.g.ninc[0]) * N;
    const int o0 = L.g.own0, o1 = L.g.own1;
    comm->shift(v + size_t(o1) * pd, o0 > 0 ? v + size_t(o0 - 1) * pd : v, v + size_t(o0) * pd, o1 + 1 < L.g.n[0] ? v + size_t(o1 + 1) * pd : v, pd,
                (dirs & XUP) != 0, (dirs & XDOWN) != 0, stream);
}

template <int N>
MG<N>::~MG() {
    if (pgExec) cudaGraphExecDestroy(pgExec);
    if (captureStream) cudaStreamDestroy(captureStream);
    if (h_scal) cudaFreeHost(h_scal);
    for (auto &kv : haloGraphs) if (kv.second) cudaGraphExecDestroy(kv.second);
    if (haloStream) cudaStreamDestroy(haloStream);
    if (haloFork) cudaEventDestroy(haloFork);
    if (haloJoin) cudaEventDestroy(haloJoin);
    if (haloStream2) cudaStreamDestroy(haloStream2);
    if (haloShifted) cudaEventDestroy(haloShifted);
    if (haloJoin2) cudaEventDestroy(haloJoin2);
}

template <int N>
void MG<N>::syncMasks() {
    sim->syncBC();
}

// C_fi = I_fi^T K0 I_fi with I_fi[(i,c),(J,d)] = phi_J(xi_i) delta_cd, xi_i the position of node i of
// child fi inside the parent element (MG.hh:559-583, 644-648; child bit d <-> axis d).
template <int N>
static void coarsenedReferenceStiffness(const K0Param<N> &K0, std::vector<Real> &C) {
    constexpr int NPE = 1 << N, KS = N * NPE;
    C.assign(size_t(NPE) * KS * KS, 0.0);
    for (int fi = 0; fi < NPE; ++fi) {
        Real phi[NPE][NPE]; // [fine node i][coarse node J]
        for (int i = 0; i < NPE; ++i)
            for (int J = 0; J < NPE; ++J) {
                Real v = 1.0;
                for (int d = 0; d < N; ++d) {
                    Real xi = 0.5 * ((i >> (N - 1 - d)) & 1) + 0.5 * ((fi >> d) & 1);
                    v *= ((J >> (N - 1 - d)) & 1) ? xi : 1.0 - xi;
                }
                phi[i][J] = v;
            }
        Real *Cf = &C[size_t(fi) * KS * KS];
        for (int J = 0; J < NPE; ++J)
            for (int J2 = 0; J2 < NPE; ++J2)
                for (int c = 0; c < N; ++c)
                    for (int d = 0; d < N; ++d) {
                        Real s = 0.0;
                        for (int i = 0; i < NPE; ++i)
                            for (int i2 = 0; i2 < NPE; ++i2) s += phi[i][J] * K0.k[(i * N + c) * KS + i2 * N + d] * phi[i2][J2];
                        Cf[(J * N + c) * KS + J2 * N + d] = s;
                    }
    }
}

template <int N>
void MG<N>::updateElementStiffnessMatrices(bool force) {
    if (!force && builtVersion == sim->densityVersion) return;
    const size_t L = lv.size() - 1;
    if (L == 0) throw std::runtime_error("multigridSolver: at least one coarsening level is required on the device path");
    const Real *Ey = sim->moduli();
    std::vector<Real> C; coarsenedReferenceStiffness<N>(sim->K0, C);
    d_C.upload(C, stream);
    {
        long long tot = (long long)lv[1].g.nn * NST;
        buildLevel1StencilKernel<N><<<unsigned((tot + kBlock - 1) / kBlock), kBlock, 0, stream>>>(lv[0].g, lv[1].g, Ey, d_C.p, lv[1].S.p);
        VFEM_CHECK_LAUNCH(); countLaunch();
    }
    for (size_t l = 2; l <= L; ++l) {
        long long tot = (long long)lv[l].g.nn * NST;
        static const bool threadForm = getenv("VFEM_GALERKIN_THREAD") != nullptr; // the round-1 kernel (thread per output block), for comparison
        if (threadForm) galerkinStencilKernel<N><<<unsigned((tot + kBlock - 1) / kBlock), kBlock, 0, stream>>>(lv[l - 1].g, lv[l].g, lv[l - 1].S.p, lv[l].S.p);
        else {
            constexpr size_t smem = size_t(NST) * N * NST * N * sizeof(Real);
            static bool attr = false;
            if (!attr) { VFEM_CUDA(cudaFuncSetAttribute(galerkinStencilCtaKernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem))); attr = true; }
            if (!galStart.p) { GalerkinTable<N> t; galStart.upload(t.start, stream); galOff.upload(t.off, stream); galW.upload(t.w, stream); VFEM_CUDA(cudaStreamSynchronize(stream)); }
            galerkinStencilCtaKernel<N><<<unsigned(lv[l].g.nn), 256, smem, stream>>>(lv[l - 1].g, lv[l].g, lv[l - 1].S.p, lv[l].S.p, galStart.p, galOff.p, galW.p);
        }
        VFEM_CHECK_LAUNCH(); countLaunch();
    }
    // coarsest level: block-tridiagonal LDL^T of the Dirichlet-reduced operator (coarse_solve.cuh).  Slab mode: the
    // owned rows of every rank are summed into the global operator and every rank factorises it (redundantly).
    const Grid<N> &cgrid = gCoarse;
    const Real *cS = lv[L].S.p;
    const uint8_t *cmask = lv[L].d_dmask.p;
    if (slab()) {
        const Level &Ll = lv[L];
        const size_t rowLen = size_t(Ll.g.ninc[0]) * NST * N * N;
        Sg.zero(stream);
        VFEM_CUDA(cudaMemcpyAsync(Sg.p + size_t(coarsePlaneOffset + Ll.g.own0) * rowLen, Ll.S.p + size_t(Ll.g.own0) * rowLen,
                                  size_t(Ll.g.own1 - Ll.g.own0 + 1) * rowLen * sizeof(Real), cudaMemcpyDeviceToDevice, stream));
        comm->allreduceSum(Sg.p, Sg.n, stream);
        cS = Sg.p; cmask = dmaskG.p;
    }
    bt->factor(cgrid, cS, cmask, stream);
    coarseP = bt->P; coarsePlanes = bt->planes;
    bt->check(stream);
    builtVersion = sim->densityVersion;
}

template <int N>
void MG<N>::coarseSolve(const Real *b, Real *x) {
    const Level &Lc = lv.back();
    Real *xl = x;
    const uint8_t *cmask = Lc.d_dmask.p;
    if (slab()) { // gather the owned planes of b into the global right-hand side; every rank solves the same system
        const size_t pd = size_t(Lc.g.ninc[0]) * N;
        bg.zero(stream);
        VFEM_CUDA(cudaMemcpyAsync(bg.p + size_t(coarsePlaneOffset + Lc.g.own0) * pd, b + size_t(Lc.g.own0) * pd,
                                  ownedDofs(lv.size() - 1) * sizeof(Real), cudaMemcpyDeviceToDevice, stream));
        comm->allreduceSum(bg.p, bg.n, stream);
        b = bg.p; x = xg.p; cmask = dmaskG.p;
    }
    bt->solve(b, x, cmask, stream);
    if (slab()) // the local window (ghost planes included) of the global solution
        VFEM_CUDA(cudaMemcpyAsync(xl, xg.p + size_t(coarsePlaneOffset) * Lc.g.ninc[0] * N, size_t(Lc.g.nn) * N * sizeof(Real),
                                  cudaMemcpyDeviceToDevice, stream));
}

// ---- per-level operators ------------------------------------------------------------------------------
template <int N>
static inline const uint8_t *maskOf(MG<N> &mg, size_t l) { return l == 0 ? mg.sim->d_dmask.p : mg.lv[l].d_dmask.p; }
template <int N>
static inline const Real *dvalOf(MG<N> &mg, size_t l) { return l == 0 ? mg.sim->d_dval.p : mg.lv[l].d_dval.p; }

template <int N>
template <int MODE>
void MG<N>::applyMode(size_t l, const Real *u, const Real *b, Real *out) {
    const Level &L = lv.at(l);
    const unsigned blocks = unsigned((L.g.nn + kBlock - 1) / kBlock);
    if (l == 0) { sim->template applyFine<MODE>(u, b, maskOf(*this, 0), out); return; }
    (void)blocks;
    applyStencilWarpKernel<N, MODE><<<warpGridBlocks(L.g.nn), kWarpBlock, 0, stream>>>(L.g, L.S.p, u, b, maskOf(*this, l), out);
    VFEM_CHECK_LAUNCH(); countLaunch();
}

// Slab mode, fine level: out = op(u) with the ghost planes of u refreshed on the way.  The element layers that touch no
// ghost plane are applied while the one-plane shifts cross NVLink on haloStream; the two layer pairs around the first and
// last owned plane follow once the planes have landed.  Views store complete rows only (edgeSkip), so the three launches
// write disjoint planes.
template <int N>
template <int MODE>
void MG<N>::exchangeApplyFine(Real *u, const Real *b, Real *out) {
    const Grid<N> &g = lv[0].g;
    const int o0 = g.own0, o1 = g.own1;
    if (!slab() || o1 - o0 < 3 || getenv("VFEM_NO_HALO_OVERLAP")) {
        exchange(0, u, XUP | XDOWN);
        applyMode<MODE>(0, u, b, out);
        return;
    }
    sim->moduli(); // a stale E(rho) is refreshed on the solver stream BEFORE the fork
    syncMasks();
    // The overlapped form is 1 NCCL group + 3 kernel launches + 4 event operations on two streams: launch-bound next to a
    // 20 us kernel.  With VFEM_HALO_GRAPH=1 its SECOND call with the same arrays captures it into a CUDA graph (fork / join
    // included) and later calls replay it.  Measured on 2 x B200 (gpurun r2c): the replay is SLOWER than the plain launches
    // (37.8 us against 28.1 us per halo'd apply of 128x64x64 layers per rank), so it is opt-in.
    if (useGraphs && comm->graphSafe() && comm->sideStreamSafe() && getenv("VFEM_HALO_GRAPH") && !getenv("VFEM_TRACE_HALO")) {
        HaloGraphKey key{u, b, out, MODE, sim->k0Version, sim->applyVariant};
        auto it = haloGraphs.find(key);
        if (it == haloGraphs.end()) {
            if (haloGraphs.size() >= 64) { for (auto &kv : haloGraphs) if (kv.second) cudaGraphExecDestroy(kv.second); haloGraphs.clear(); }
            haloGraphs.emplace(key, nullptr);   // first call: run it plainly (also sets function attributes, plans the views)
        } else {
            if (!it->second) {
                if (!captureStream) VFEM_CUDA(cudaStreamCreateWithFlags(&captureStream, cudaStreamNonBlocking));
                VFEM_CUDA(cudaStreamSynchronize(stream));
                const cudaStream_t saved = stream, savedSim = sim->stream;
                const unsigned long long before = g_launchCount;
                stream = captureStream; sim->stream = captureStream;
                cudaGraph_t graph = nullptr;
                VFEM_CUDA(cudaStreamBeginCapture(captureStream, cudaStreamCaptureModeRelaxed));
                try { exchangeApplyFineBody<MODE>(u, b, out); }
                catch (...) { cudaStreamEndCapture(captureStream, &graph); if (graph) cudaGraphDestroy(graph); stream = saved; sim->stream = savedSim; throw; }
                stream = saved; sim->stream = savedSim;
                VFEM_CUDA(cudaStreamEndCapture(captureStream, &graph));
                haloGraphLaunches = g_launchCount - before;
                g_launchCount = before;
                VFEM_CUDA(cudaGraphInstantiate(&it->second, graph, 0));
                cudaGraphDestroy(graph);
            }
            VFEM_CUDA(cudaGraphLaunch(it->second, stream));
            countLaunch(haloGraphLaunches);
            return;
        }
    }
    exchangeApplyFineBody<MODE>(u, b, out);
}

template <int N>
template <int MODE>
void MG<N>::exchangeApplyFineBody(Real *u, const Real *b, Real *out) {
    const Grid<N> &g = lv[0].g;
    const int o0 = g.own0, o1 = g.own1, E0 = g.E[0];
    const bool below = o0 > 0, above = o1 + 1 < g.n[0];
    cudaStream_t hs = stream;
    if (comm->sideStreamSafe()) {
        if (!haloStream) {
            VFEM_CUDA(cudaStreamCreateWithFlags(&haloStream, cudaStreamNonBlocking));
            VFEM_CUDA(cudaEventCreateWithFlags(&haloFork, cudaEventDisableTiming));
            VFEM_CUDA(cudaEventCreateWithFlags(&haloJoin, cudaEventDisableTiming));
        }
        hs = haloStream;
        VFEM_CUDA(cudaEventRecord(haloFork, stream));
        VFEM_CUDA(cudaStreamWaitEvent(hs, haloFork, 0));
    }
    const size_t pd = size_t(g.ninc[0]) * N;
    // VFEM_TRACE_HALO=1 (diagnosis of a stalled exchange): synchronise after every part and say which one was reached
    static const bool traceHalo = getenv("VFEM_TRACE_HALO") != nullptr;
    static unsigned long long traceCall = 0;
    auto mark = [&](const char *what, cudaStream_t s) {
        if (!traceHalo) return;
        fprintf(stderr, "[halo rank %d call %llu] %s issued\n", comm->rank, traceCall, what); fflush(stderr);
        VFEM_CUDA(cudaStreamSynchronize(s));
        fprintf(stderr, "[halo rank %d call %llu] %s done\n", comm->rank, traceCall, what); fflush(stderr);
    };
    if (traceHalo) ++traceCall;
    comm->shift(u + size_t(o1) * pd, below ? u + size_t(o0 - 1) * pd : u, u + size_t(o0) * pd, above ? u + size_t(o1 + 1) * pd : u, pd, true, true, hs);
    mark("shift", hs);
    const uint8_t *mask = maskOf(*this, 0);
    sim->template applyFineRange<MODE>(u, b, mask, out, below ? o0 : 0, above ? o1 : E0, stream, (below ? 1 : 0) | (above ? 2 : 0));
    mark("interior apply", stream);
    // the thin boundary views follow the shift on ITS stream: they share the SMs with the tail of the interior launch
    // a middle rank has two of them: the second one runs beside the first on a stream of its own (they write disjoint planes)
    static const bool serialViews = getenv("VFEM_HALO_SERIAL_VIEWS") != nullptr;
    const bool split = below && above && hs != stream && !serialViews && !traceHalo;
    if (split) {
        if (!haloStream2) {
            VFEM_CUDA(cudaStreamCreateWithFlags(&haloStream2, cudaStreamNonBlocking));
            VFEM_CUDA(cudaEventCreateWithFlags(&haloShifted, cudaEventDisableTiming));
            VFEM_CUDA(cudaEventCreateWithFlags(&haloJoin2, cudaEventDisableTiming));
        }
        VFEM_CUDA(cudaEventRecord(haloShifted, hs));
        VFEM_CUDA(cudaStreamWaitEvent(haloStream2, haloShifted, 0));
    }
    if (below) sim->template applyFineRange<MODE>(u, b, mask, out, o0 - 1, o0 + 1, hs, 3);
    if (above) sim->template applyFineRange<MODE>(u, b, mask, out, o1 - 1, o1 + 1, split ? haloStream2 : hs, 3);
    if (split) {
        VFEM_CUDA(cudaEventRecord(haloJoin2, haloStream2));
        VFEM_CUDA(cudaStreamWaitEvent(stream, haloJoin2, 0));
    }
    mark("boundary views", hs);
    if (hs != stream) {
        VFEM_CUDA(cudaEventRecord(haloJoin, hs));
        VFEM_CUDA(cudaStreamWaitEvent(stream, haloJoin, 0));
    }
}

template <int N> void MG<N>::applyK(size_t l, const Real *u, Real *out) { applyMode<0>(l, u, nullptr, out); }
template <int N> void MG<N>::computeResidual(size_t l, const Real *u, const Real *b, Real *r) { applyMode<1>(l, u, b, r); }
template <int N> void MG<N>::applyKMasked(size_t l, const Real *u, Real *out) { applyMode<2>(l, u, nullptr, out); }

// 3-D fine level: shared-memory tile kernel (gs_tile.cuh)
template <class Cfg>
static inline void launchGsFineTileCfg(const Grid<3> &g, const K0Param<3> &K, const Real *Ey, Real *u, const Real *b, const uint8_t *dmask,
                                       int colour, int forward, cudaStream_t stream) {
    static bool once = (cudaFuncSetAttribute(gsFineTileKernel<Cfg>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(Cfg::SMEM)), true);
    (void)once;
    const GsTilePlan p = planGsTile<Cfg>(g, colour);
    gsFineTileKernel<Cfg><<<p.grid, Cfg::NT, Cfg::SMEM, stream>>>(g, K, Ey, u, b, dmask, colour, forward, p.tz);
}
static inline bool launchGsFineTile(const Grid<3> &g, const K0Param<3> &K, const Real *Ey, Real *u, const Real *b, const uint8_t *dmask,
                                    int colour, int forward, cudaStream_t stream, int variant) {
    if (reinterpret_cast<size_t>(u) & 7) return false;
    switch (variant) { // lattice sites per CTA along axes 0, 1 (experiment knob VFEM_GS_VARIANT; 1 = default)
    case 2: launchGsFineTileCfg<GsTileCfg<1, 2>>(g, K, Ey, u, b, dmask, colour, forward, stream); break;
    case 3: launchGsFineTileCfg<GsTileCfg<2, 4>>(g, K, Ey, u, b, dmask, colour, forward, stream); break;
    case 4: launchGsFineTileCfg<GsTileCfg<1, 4>>(g, K, Ey, u, b, dmask, colour, forward, stream); break;
    default: launchGsFineTileCfg<GsTileCfg<2, 2>>(g, K, Ey, u, b, dmask, colour, forward, stream); break;
    }
    return true;
}
static inline bool launchGsFineTile(const Grid<2> &, const K0Param<2> &, const Real *, Real *, const Real *, const uint8_t *, int, int, cudaStream_t, int) { return false; }

// Halo traffic of a coloured sweep (slab mode): the colours whose axis-0 parity is even (the first half of the forward
// order, MG.hh:292-294) only read odd ghost planes and vice versa, and own0 is even / own1 odd on interior faces, so a
// sweep needs ONE one-directional plane shift after each half: even half -> first owned plane goes down, odd half ->
// last owned plane goes up.
template <int N>
void MG<N>::smoothColoured(size_t l, Real *u, const Real *b, bool forward) {
    const Level &L = lv.at(l);
    const Real *Ey = (l == 0) ? sim->moduli() : nullptr;
    for (int i = 0; i < NPE; ++i) {
        const int colour = forward ? i : NPE - 1 - i; // MG.hh:318-320
        const int cnt = colourCount<N>(L.g, colour);
        if (cnt > 0) {
            const unsigned blocks = unsigned((cnt + kBlock - 1) / kBlock);
            if (l == 0) { // gsVariant 0: one thread per node gathering from global memory (the 2-D path and the fallback)
                if (!(N == 3 && gsVariant != 0 && launchGsFineTile(L.g, sim->K0, Ey, u, b, maskOf(*this, 0), colour, forward ? 1 : 0, stream, gsVariant == 1 && cnt < 60000 ? 3 : gsVariant)))
                    gsFineKernel<N, 4><<<blocks, kBlock, 0, stream>>>(L.g, sim->K0, Ey, u, b, maskOf(*this, 0), colour, forward ? 1 : 0);
            }
            else gsStencilWarpKernel<N><<<warpGridBlocks(cnt), kWarpBlock, 0, stream>>>(L.g, L.S.p, u, b, maskOf(*this, l), colour, forward ? 1 : 0, cnt);
            VFEM_CHECK_LAUNCH(); countLaunch();
        }
        if (slab() && (i + 1) % (NPE / 2) == 0) exchange(l, u, (colour >> (N - 1)) & 1 ? XUP : XDOWN);
    }
}

template <int N>
void MG<N>::smoothSerial(size_t l, Real *u, const Real *b, bool forward) {
    if (slab()) throw std::runtime_error("serial Gauss-Seidel is not available on a slab-decomposed grid");
    const Level &L = lv.at(l);
    if (l + 1 == lv.size()) throw std::runtime_error("smoothing: the coarsest level is solved exactly (no operator rows are kept for it)");
    gsSerialKernel<N><<<1, 32, 0, stream>>>(L.g, sim->K0, l == 0 ? sim->moduli() : nullptr, l == 0 ? nullptr : L.S.p, u, b, maskOf(*this, l), forward ? 1 : 0);
    VFEM_CHECK_LAUNCH(); countLaunch();
}

template <int N>
void MG<N>::smooth(size_t l, Real *u, const Real *b, bool forward, size_t sweeps) {
    if (l == 0 || !fusedSweeps || sweeps == 0) { for (size_t i = 0; i < sweeps; ++i) smoothColoured(l, u, b, forward); return; }
    const Level &L = lv.at(l);
    if (sweepGridBlocks == 0) {
        int dev = 0, sms = 148, per = 1;
        VFEM_CUDA(cudaGetDevice(&dev));
        VFEM_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        VFEM_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per, gsStencilSweepKernel<N>, kWarpBlock, 0));
        if (const char *v = getenv("VFEM_SWEEP_BPS")) per = std::min(per, std::max(1, atoi(v)));
        sweepGridBlocks = sms * (per > 0 ? per : 1);
    }
    long long most = 1;
    for (int c = 0; c < NPE; ++c) most = std::max<long long>(most, colourCount<N>(L.g, c));
    const long long need = most * 32 / kWarpBlock + 1;
    const unsigned blocks = unsigned(need < sweepGridBlocks ? need : sweepGridBlocks);
    Grid<N> g = L.g; const Real *S = L.S.p; const uint8_t *m = maskOf(*this, l);
    int fw = forward ? 1 : 0;
    // one launch per run of phases between two halo exchanges (the whole sequence on a single device)
    const int total = int(sweeps) * NPE, step = slab() ? NPE / 2 : total;
    for (int q0 = 0; q0 < total; q0 += step) {
        int q1 = q0 + step;
        void *args[] = {&g, &S, &u, &b, &m, &fw, &q0, &q1};
        VFEM_CUDA(cudaLaunchCooperativeKernel((const void *)gsStencilSweepKernel<N>, dim3(blocks), dim3(kWarpBlock), args, 0, stream));
        countLaunch();
        if (slab()) {
            const int colour = forward ? ((q1 - 1) % NPE) : (NPE - 1 - (q1 - 1) % NPE);
            exchange(l, u, (colour >> (N - 1)) & 1 ? XUP : XDOWN);
        }
    }
}

template <int N>
void MG<N>::restriction(size_t lf, const Real *vf, Real *bc) {
    const Level &F = lv.at(lf), &C = lv.at(lf + 1);
    exchange(lf, const_cast<Real *>(vf), XUP); // the lowest owned coarse plane gathers from the fine ghost plane below it
    restrictKernel<N><<<(C.g.nn + kBlock - 1) / kBlock, kBlock, 0, stream>>>(F.g, C.g, vf, bc);
    VFEM_CHECK_LAUNCH(); countLaunch();
}

template <int N>
void MG<N>::interpolate(size_t lf, const Real *xc, Real *xf, bool accumulate) {
    const Level &F = lv.at(lf), &C = lv.at(lf + 1);
    const unsigned blocks = unsigned((F.g.nn + kBlock - 1) / kBlock);
    if (accumulate) prolongKernel<N, true><<<blocks, kBlock, 0, stream>>>(F.g, C.g, xc, xf);
    else prolongKernel<N, false><<<blocks, kBlock, 0, stream>>>(F.g, C.g, xc, xf);
    VFEM_CHECK_LAUNCH(); countLaunch();
}

template <int N>
void MG<N>::enforceDirichlet(size_t l, Real *u, bool zero) {
    const size_t n = levelDofs(l);
    enforceDirichletKernel<N><<<reduceBlocks(n), kRedBlock, 0, stream>>>(u, maskOf(*this, l), dvalOf(*this, l), n, zero ? 1 : 0);
    VFEM_CHECK_LAUNCH(); countLaunch();
}
template <int N> void MG<N>::zeroOutDirichlet(size_t l, Real *u) { enforceDirichlet(l, u, true); }

#include "mg_cycles.inl"

template struct MG<2>;
template struct MG<3>;
template void MG<2>::exchangeApplyFine<0>(Real *, const Real *, Real *);
template void MG<3>::exchangeApplyFine<0>(Real *, const Real *, Real *);

} // namespace vfem
