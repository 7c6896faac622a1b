// ORACLE — TEST INFRASTRUCTURE ONLY (see vfem_oracle.hpp).
// Restatement of VoxelFEM/MultigridSolver.hh (MG.hh): hierarchy, Dirichlet coarsening,
// multicoloured Gauss-Seidel, trilinear transfer, Galerkin coarsening, V-cycle, FMG, PCG.
#pragma once
#include "vfem_oracle_solve.hpp"

namespace orc {

template <int N>
struct MG {
    typedef Sim<N> S;
    static constexpr int NPE = S::NPE, KS = S::KS, NST = S::NST;
    std::vector<std::shared_ptr<S>> sims;
    std::vector<std::vector<Real>> x, b;
    bool symmetricGS = true;               // MG.hh:758
    bool buildBlockStiffnessMatrices = true;      // MG.hh:746
    bool buildFinestBlockStiffnessMatrix = false; // MG.hh:750

    // MG.hh:22-90
    MG(std::shared_ptr<S> fine, size_t numCoarseningLevels) {
        size_t elems[N]; for (int d = 0; d < N; ++d) elems[d] = fine->E[d];
        for (size_t l = 0; l <= numCoarseningLevels; ++l) {
            std::shared_ptr<S> tps = fine;
            if (l > 0) {
                for (int d = 0; d < N; ++d) {
                    if (elems[d] % 2 == 1)
                        throw std::runtime_error("Grid size currently must be divisible by 2^numCoarseningLevels (nonuniform coarsening not yet implemented)");
                    elems[d] /= 2;
                }
                tps = std::make_shared<S>(fine->mn, fine->mx, elems);
                tps->lambda = fine->lambda; tps->mu = fine->mu; tps->updateK0();
                const S &finer = *sims.back();
                S &coarser = *tps;
                for (size_t nfi = 0; nfi < finer.nn; ++nfi) {
                    const uint8_t *fm = &finer.dmask[nfi * N];
                    bool any = false; for (int c = 0; c < N; ++c) any |= fm[c] != 0;
                    if (!any) continue;
                    // containing coarse element and position inside it (TPS.hh:1178-1216 with the
                    // float division replaced by the equivalent integer arithmetic on node indices)
                    size_t fi[N], ce[N]; int onb[N]; finer.nodeIndex(nfi, fi);
                    bool allInterior = true;
                    for (int d = 0; d < N; ++d) {
                        ce[d] = std::min(fi[d] / 2, coarser.E[d] - 1);
                        size_t rel = fi[d] - 2 * ce[d]; // 0, 1 (=0.5), 2 (=1.0)
                        onb[d] = rel == 0 ? 0 : (rel == 2 ? 1 : -1);
                        if (onb[d] >= 0) allInterior = false;
                    }
                    if (allInterior) throw std::runtime_error("Dirichlet constraints on internal nodes are not supported");
                    size_t cbase = coarser.flatNode(ce);
                    Real zero[N]; for (int c = 0; c < N; ++c) zero[c] = 0;
                    for (int j = 0; j < NPE; ++j) {
                        bool match = true;
                        for (int d = 0; d < N; ++d) { int bit = (j >> (N - 1 - d)) & 1; if (onb[d] != -1 && onb[d] != bit) match = false; }
                        if (match) coarser.setDirichlet(cbase + coarser.refNodes[j], fm, zero);
                    }
                }
            }
            x.emplace_back(tps->nn * N, 0.0);
            b.emplace_back(tps->nn * N, 0.0);
            sims.push_back(tps);
        }
    }
    size_t numLevels() const { return sims.size(); }

    // ---- transfers (MG.hh:104-180): trilinear weights, products of {1, 1/2} ----
    template <bool ACCUM>
    void interpolate(size_t lf, const Real *vc, Real *outf) const {
        const S &fs = *sims[lf], &cs = *sims[lf + 1];
#pragma omp parallel for schedule(static)
        for (long long t = 0; t < (long long)fs.nn; ++t) {
            size_t fi[N]; fs.nodeIndex(size_t(t), fi);
            size_t lo[N]; int odd[N];
            for (int d = 0; d < N; ++d) { lo[d] = fi[d] / 2; odd[d] = fi[d] & 1; }
            Real acc[N]; for (int c = 0; c < N; ++c) acc[c] = 0;
            for (int j = 0; j < NPE; ++j) {
                Real w = 1; size_t ci[N]; bool skip = false;
                for (int d = 0; d < N; ++d) {
                    int bit = (j >> (N - 1 - d)) & 1;
                    if (odd[d]) { w *= 0.5; ci[d] = lo[d] + bit; }
                    else { if (bit) skip = true; ci[d] = lo[d]; }
                }
                if (skip) continue;
                size_t cn = cs.flatNode(ci);
                for (int c = 0; c < N; ++c) acc[c] += w * vc[cn * N + c];
            }
            for (int c = 0; c < N; ++c) { if (ACCUM) outf[size_t(t) * N + c] += acc[c]; else outf[size_t(t) * N + c] = acc[c]; }
        }
    }
    // result_c = P^T values_f, written as a gather over the 3^N fine nodes around each coarse node
    void restriction(size_t lf, const Real *vf, Real *outc) const {
        const S &fs = *sims[lf], &cs = *sims[lf + 1];
#pragma omp parallel for schedule(static)
        for (long long t = 0; t < (long long)cs.nn; ++t) {
            size_t ci[N]; cs.nodeIndex(size_t(t), ci);
            Real acc[N]; for (int c = 0; c < N; ++c) acc[c] = 0;
            for (int k = 0; k < NST; ++k) {
                int tt = k; Real w = 1; size_t fi[N]; bool ok = true;
                for (int d = N - 1; d >= 0; --d) {
                    int dl = tt % 3 - 1; tt /= 3;
                    long long v = 2 * (long long)ci[d] + dl;
                    if (v < 0 || v >= (long long)fs.nper[d]) { ok = false; }
                    fi[d] = size_t(v); if (dl != 0) w *= 0.5;
                }
                if (!ok) continue;
                size_t fn = fs.flatNode(fi);
                for (int c = 0; c < N; ++c) acc[c] += w * vf[fn * N + c];
            }
            for (int c = 0; c < N; ++c) outc[size_t(t) * N + c] = acc[c];
        }
    }

    // ---- Gauss-Seidel node update (MG.hh:193-265) ----
    void smoothNode(const S &sim, const size_t *gidx, Real *u, const Real *bb, bool forward) const {
        const size_t n = sim.flatNode(gidx);
        Real t[N], M[N * N];
        for (int c = 0; c < N; ++c) t[c] = bb[n * N + c];
        for (int i = 0; i < N * N; ++i) M[i] = 0;
        if (!sim.hasBlockK) {
            const bool cached = sim.hasCachedKe();
            // visitIncidentElements (TPS.hh:1233-1281)
            for (int i = 0; i < NPE; ++i) {
                size_t eidx[N]; int loc = 0; bool valid = true;
                for (int d = 0; d < N; ++d) {
                    size_t elem = gidx[d]; int localNode = 0;
                    if (elem == sim.E[d]) { elem = sim.E[d] - 1; localNode = 1; }
                    if ((1 << d) & i) { eidx[d] = elem; loc |= localNode << (N - 1 - d); }
                    else {
                        if (localNode != 0 || elem == 0) { valid = false; break; }
                        eidx[d] = elem - 1; loc |= 1 << (N - 1 - d);
                    }
                }
                if (!valid) continue;
                size_t ek = sim.flatElem(eidx), base = sim.flatNode(eidx);
                Real ul[KS];
                for (int m = 0; m < NPE; ++m) for (int c = 0; c < N; ++c) ul[m * N + c] = u[(base + sim.refNodes[m]) * N + c];
                if (!cached) {
                    Real Ey = sim.youngModulus(ek);
                    for (int c = 0; c < N; ++c) {
                        Real s = 0;
                        for (int j = 0; j < KS; ++j) s += sim.K0[j * KS + (loc * N + c)] * ul[j]; // column access (MG.hh:216)
                        t[c] -= Ey * s;
                        for (int c2 = 0; c2 < N; ++c2) M[c * N + c2] += Ey * sim.K0[(loc * N + c) * KS + (loc * N + c2)];
                    }
                } else {
                    const Real *Ke = sim.elementKe(ek);
                    for (int c = 0; c < N; ++c) {
                        Real s = 0;
                        for (int j = 0; j < KS; ++j) s += Ke[(loc * N + c) * KS + j] * ul[j];
                        t[c] -= s;
                        for (int c2 = 0; c2 < N; ++c2) M[c * N + c2] += Ke[(loc * N + c) * KS + (loc * N + c2)];
                    }
                }
            }
        } else {
            for (int k = 0; k < NST; ++k) {
                size_t m = sim.neighbour(gidx, k); if (m == SIZE_MAX) continue;
                const Real *blk = &sim.stencil[(n * NST + k) * N * N];
                for (int a = 0; a < N; ++a) for (int c2 = 0; c2 < N; ++c2) t[a] -= blk[a * N + c2] * u[m * N + c2];
                if (m == n) for (int i = 0; i < N * N; ++i) M[i] = blk[i];
            }
        }
        Real du[N]; for (int c = 0; c < N; ++c) du[c] = 0;
        auto upd = [&](int i) {
            Real s = 0; for (int j = 0; j < N; ++j) s += M[i * N + j] * du[j];
            du[i] = (t[i] - s) * (Real(!sim.dmask[n * N + i]) / M[i * N + i]);
        };
        if (forward) for (int i = 0; i < N; ++i) upd(i);
        else for (int i = N - 1; i >= 0; --i) upd(i);
        for (int c = 0; c < N; ++c) u[n * N + c] += du[c];
    }
    // serial lexicographic sweep (MG.hh:269-282; the python `smoothing` binding)
    void smoothingSerial(size_t l, Real *u, const Real *bb, bool forward = true) const {
        const S &sim = *sims[l];
        for (size_t n_ = 0; n_ < sim.nn; ++n_) {
            size_t n = forward ? n_ : sim.nn - (n_ + 1);
            size_t idx[N]; sim.nodeIndex(n, idx);
            smoothNode(sim, idx, u, bb, forward);
        }
    }
    // MG.hh:284-340
    void smoothingMulticoloredGS(size_t l, Real *u, const Real *bb, bool forward = true) const {
        const S &sim = *sims[l];
        for (int i = 0; i < NPE; ++i) {
            int lni = forward ? i : (NPE - i - 1);
            size_t cnt[N], tot = 1; int par[N];
            for (int d = 0; d < N; ++d) { par[d] = (lni >> (N - 1 - d)) & 1; cnt[d] = (sim.nper[d] - 1 - par[d]) / 2 + 1; tot *= cnt[d]; }
#pragma omp parallel for schedule(static)
            for (long long t = 0; t < (long long)tot; ++t) {
                size_t idx[N], r = size_t(t);
                for (int d = N - 1; d >= 0; --d) { idx[d] = par[d] + 2 * (r % cnt[d]); r /= cnt[d]; }
                smoothNode(sim, idx, u, bb, forward);
            }
        }
    }
    void debugMulticolorVisit(int *order) const { // MG.hh:328-334
        const S &sim = *sims[0]; int ctr = 0;
        for (int lni = 0; lni < NPE; ++lni) {
            size_t cnt[N], tot = 1; int par[N];
            for (int d = 0; d < N; ++d) { par[d] = (lni >> (N - 1 - d)) & 1; cnt[d] = (sim.nper[d] - 1 - par[d]) / 2 + 1; tot *= cnt[d]; }
            for (size_t t = 0; t < tot; ++t) {
                size_t idx[N], r = t;
                for (int d = N - 1; d >= 0; --d) { idx[d] = par[d] + 2 * (r % cnt[d]); r /= cnt[d]; }
                order[sim.flatNode(idx)] = ctr++;
            }
        }
    }

    void applyK(size_t l, const Real *u, Real *out) { // MG.hh:353-358
        const S &sim = *sims.at(l);
        if (l > 0 && !sim.hasCachedKe()) updateElementStiffnessMatrices();
        if (sim.hasBlockK) sim.applyBlockK(u, out); else sim.applyK(u, out);
    }
    void zeroOutDirichletComponents(size_t l, Real *u) const { // MG.hh:364-378
        const S &sim = *sims[l];
        for (size_t i = 0; i < sim.nn * N; ++i) if (sim.dmask[i]) u[i] = 0;
    }
    void enforceDirichletConditions(size_t l, Real *u, bool zero) const { // MG.hh:386-398
        const S &sim = *sims[l];
        for (size_t i = 0; i < sim.nn * N; ++i) if (sim.dmask[i]) u[i] = zero ? 0.0 : sim.dval[i];
    }
    void computeResidual(size_t l, const Real *u, const Real *bb, Real *r) { // MG.hh:401-413
        const S &sim = *sims[l];
        applyK(l, u, r);
        for (size_t i = 0; i < sim.nn * N; ++i) r[i] = bb[i] - r[i];
        zeroOutDirichletComponents(l, r);
    }

    // ---- Galerkin coarsening (MG.hh:559-669) ----
    static void compressedInterpolation(Real phis[NPE][NPE * NPE]) { // phis[fi][fine_n*NPE + coarse_n]
        for (int fn = 0; fn < NPE; ++fn)
            for (int fi = 0; fi < NPE; ++fi) {
                Real xi[N];
                for (int d = 0; d < N; ++d) xi[d] = 0.5 * ((fn >> (N - 1 - d)) & 1) + ((fi & (1 << d)) ? 0.5 : 0.0);
                for (int cn = 0; cn < NPE; ++cn) {
                    Real v = 1;
                    for (int d = 0; d < N; ++d) v *= ((cn >> (N - 1 - d)) & 1) ? xi[d] : (1.0 - xi[d]);
                    phis[fi][fn * NPE + cn] = v;
                }
            }
    }
    static void accumulateCoarsened(const Real *phi, const Real *Kef, Real *Kec) { // Kec += I^T Kef I
        Real tmp[KS * KS]; // tmp = Kef * I : tmp[(i,c),(J,d)] = sum_i' Kef[(i,c),(i',d)] phi(i',J)
        for (int r = 0; r < KS; ++r)
            for (int J = 0; J < NPE; ++J)
                for (int d = 0; d < N; ++d) {
                    Real s = 0;
                    for (int ip = 0; ip < NPE; ++ip) s += Kef[r * KS + ip * N + d] * phi[ip * NPE + J];
                    tmp[r * KS + J * N + d] = s;
                }
        for (int J = 0; J < NPE; ++J)
            for (int c = 0; c < N; ++c)
                for (int col = 0; col < KS; ++col) {
                    Real s = 0;
                    for (int i = 0; i < NPE; ++i) s += phi[i * NPE + J] * tmp[(i * N + c) * KS + col];
                    Kec[(J * N + c) * KS + col] += s;
                }
    }
    void buildPESCoarse(size_t lc) {
        const S &finer = *sims[lc - 1]; S &coarser = *sims[lc];
        Real phis[NPE][NPE * NPE]; compressedInterpolation(phis);
        std::vector<Real> KeVec(coarser.ne * size_t(KS) * KS, 0.0);
        const bool finerCached = finer.hasCachedKe();
        std::vector<Real> coarsenedK0(size_t(NPE) * KS * KS, 0.0);
        if (!finerCached) for (int fi = 0; fi < NPE; ++fi) accumulateCoarsened(phis[fi], finer.K0, &coarsenedK0[size_t(fi) * KS * KS]);
#pragma omp parallel for schedule(static)
        for (long long ec = 0; ec < (long long)coarser.ne; ++ec) {
            size_t cidx[N]; coarser.elemIndex(size_t(ec), cidx);
            Real *Kec = &KeVec[size_t(ec) * KS * KS];
            for (int fi = 0; fi < NPE; ++fi) {
                size_t fidx[N]; bool ok = true;
                for (int d = 0; d < N; ++d) { fidx[d] = 2 * cidx[d] + ((fi & (1 << d)) ? 1 : 0); if (fidx[d] >= finer.E[d]) ok = false; }
                if (!ok) continue;
                size_t ef = finer.flatElem(fidx);
                if (!finerCached) {
                    Real Ey = finer.youngModulus(ef);
                    const Real *c0 = &coarsenedK0[size_t(fi) * KS * KS];
                    for (int i = 0; i < KS * KS; ++i) Kec[i] += Ey * c0[i];
                } else accumulateCoarsened(phis[fi], finer.elementKe(ef), Kec);
            }
        }
        coarser.KeCache = std::move(KeVec);
        coarser.factorUpToDate = false;
    }
    void updateElementStiffnessMatrices() { for (size_t l = 1; l < sims.size(); ++l) buildPESCoarse(l); } // MG.hh:415-425
    void updateBlockKs() { // MG.hh:427-441
        const size_t numLevels = sims.size() - 1;
        // NOTE: like the reference, there is no early return here, so block matrices for levels
        // 1..L-1 are (re)built even when buildBlockStiffnessMatrices is false.
        if (!buildBlockStiffnessMatrices) { for (size_t l = 0; l < numLevels; ++l) sims[l]->clearBlockK(); }
        if (!buildFinestBlockStiffnessMatrix) sims[0]->clearBlockK();
        for (size_t l = buildFinestBlockStiffnessMatrix ? 0 : 1; l < numLevels; ++l) sims[l]->updateBlockK();
    }

    // ---- cycles (MG.hh:446-553) ----
    void vcycle(size_t l, size_t nu, bool residualSystem) {
        const size_t coarsest = sims.size() - 1;
        if (l == coarsest) { sims[l]->solveDirect(b[l].data(), x[l].data()); return; }
        enforceDirichletConditions(l, x[l].data(), residualSystem);
        for (size_t i = 0; i < nu; ++i) smoothingMulticoloredGS(l, x[l].data(), b[l].data(), true);
        std::vector<Real> r(sims[l]->nn * N);
        computeResidual(l, x[l].data(), b[l].data(), r.data());
        restriction(l, r.data(), b[l + 1].data());
        std::fill(x[l + 1].begin(), x[l + 1].end(), 0.0);
        vcycle(l + 1, nu, true);
        interpolate<true>(l, x[l + 1].data(), x[l].data());
        for (size_t i = 0; i < nu; ++i) smoothingMulticoloredGS(l, x[l].data(), b[l].data(), !symmetricGS);
    }
    void fullMultigrid(size_t l, size_t nu, bool residualSystem) {
        const size_t coarsest = sims.size() - 1;
        if (l == coarsest) { sims[l]->solveDirect(b[l].data(), x[l].data()); return; }
        restriction(l, b[l].data(), b[l + 1].data());
        fullMultigrid(l + 1, nu, residualSystem);
        interpolate<false>(l, x[l + 1].data(), x[l].data());
        vcycle(l, nu, residualSystem);
    }
    typedef std::function<void(size_t, const Real *)> SolveCb;
    void solve(const Real *u, const Real *f, size_t numSteps, size_t nu, bool stiffnessUpdated, bool zeroDirichlet,
               const SolveCb &cb, bool fmg, Real *out) {
        const size_t n = sims[0]->nn * N;
        if (!stiffnessUpdated) updateElementStiffnessMatrices();
        if (numSteps == 0) { std::copy(u, u + n, out); return; }
        std::copy(u, u + n, x[0].begin()); std::copy(f, f + n, b[0].begin());
        if (fmg) {
            fullMultigrid(0, nu, zeroDirichlet);
            if (cb) cb(0, x[0].data());
            for (size_t i = 1; i < numSteps; ++i) { vcycle(0, nu, zeroDirichlet); if (cb) cb(i, x[0].data()); }
        } else
            for (size_t i = 0; i < numSteps; ++i) { vcycle(0, nu, zeroDirichlet); if (cb) cb(i, x[0].data()); }
        std::copy(x[0].begin(), x[0].end(), out);
    }
    void applyPreconditionerInv(const Real *r, size_t numSteps, size_t nu, bool fmg, Real *s) { // MG.hh:476-479
        const size_t n = sims[0]->nn * N;
        if (nu == 0) { std::copy(r, r + n, s); return; }
        std::vector<Real> zero(n, 0.0);
        solve(zero.data(), r, numSteps, nu, true, true, nullptr, fmg, s);
    }

    // ---- MG-preconditioned CG (MG.hh:679-732); returns the iteration count ----
    typedef std::function<void(size_t, const Real *, const Real *)> PcgCb;
    size_t pcg(const Real *u, const Real *bvec, size_t maxIter, Real tol, const PcgCb &cb, size_t mgIterations,
               size_t nu, bool fmg, Real *xout) {
        const size_t n = sims[0]->nn * N;
        std::vector<Real> xv(u, u + n), r(n), s(n), d, Ad(n);
        enforceDirichletConditions(0, xv.data(), false);
        updateElementStiffnessMatrices();
        updateBlockKs();
        Real bnsq = 0; for (size_t i = 0; i < n; ++i) bnsq += bvec[i] * bvec[i];
        computeResidual(0, xv.data(), bvec, r.data());
        Real rMr = 0;
        size_t i = 0; // the reference leaves `i` uninitialised (MG.hh:710); treated as 0
        auto rsq = [&]() { Real v = 0; for (size_t k = 0; k < n; ++k) v += r[k] * r[k]; return v; };
        while ((i++ < maxIter) && (rsq() > tol * tol * bnsq)) {
            applyPreconditionerInv(r.data(), mgIterations, nu, fmg, s.data());
            zeroOutDirichletComponents(0, s.data());
            Real rMrOld = rMr; rMr = 0;
            for (size_t k = 0; k < n; ++k) rMr += r[k] * s[k];
            if (!d.empty()) { Real beta = rMr / rMrOld; for (size_t k = 0; k < n; ++k) d[k] = s[k] + beta * d[k]; }
            else d = s;
            applyK(0, d.data(), Ad.data());
            zeroOutDirichletComponents(0, Ad.data());
            Real dAd = 0; for (size_t k = 0; k < n; ++k) dAd += d[k] * Ad[k];
            Real alpha = rMr / dAd;
            for (size_t k = 0; k < n; ++k) { xv[k] += alpha * d[k]; r[k] -= alpha * Ad[k]; }
            if (cb) cb(i, xv.data(), r.data());
        }
        std::copy(xv.begin(), xv.end(), xout);
        return i - 1;
    }
};

} // namespace orc
