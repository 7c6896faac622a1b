// ORACLE — TEST INFRASTRUCTURE ONLY.  Not part of the shipped product.
//
// CPU restatement (C++17, no Eigen/TBB/CHOLMOD) of the reference algorithm of
// Nikronic/ndr's VoxelFEM elasticity solve.  Only tests/, __graft_entry__.smoke()
// and bench.py's cpu_baseline / --impl reference legs may use it, and only as
// the checker / CPU baseline; the product (ndr_b200/) never links or calls it.
//
// Parity pinning: the reference binary cannot be built here (Eigen/TBB/
// SuiteSparse absent), so this restatement is pinned against the reference's own
// golden numbers: quadrature tables (VoxelFEM/tests/tp_quadrature_*var_test.inl)
// and the OC compliance trajectories in logs/slurm/gt/*.log (see
// tests/test_oracle_golden.py and tests/golden/).
//
// Every function cites the reference file:line it follows
// (TPS.hh = VoxelFEM/TensorProductSimulator.hh, MG.hh = VoxelFEM/MultigridSolver.hh).
#pragma once
#include <algorithm>
#include <array>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <functional>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace orc {
typedef double Real;

// ---------------------------------------------------------------------------
// 1-D Gauss rules on [0,1] (TensorProductQuadrature.hh:118-173).  A rule that
// must integrate degree D exactly uses ceil((D+1)/2) points.
// ---------------------------------------------------------------------------
inline void gaussRule(int npts, std::vector<Real> &x, std::vector<Real> &w) {
    x.clear(); w.clear();
    switch (npts) {
    case 1: x = {0.5}; w = {1.0}; break;
    case 2: x = {0.21132486540518711775, 0.78867513459481288225}; w = {0.5, 0.5}; break;
    case 3: x = {0.5, 0.11270166537925831148, 0.88729833462074168852};
            w = {8 / 18., 5 / 18., 5 / 18.}; break;
    case 4: x = {0.33000947820757186760, 0.66999052179242813240, 0.93056815579702628761, 0.06943184420297371239};
            w = {0.326072577431273071, 0.326072577431273071, 0.1739274225687269287, 0.1739274225687269287}; break;
    case 5: x = {0.5, 0.23076534494715845448, 0.76923465505284154552, 0.04691007703066800360, 0.95308992296933199640};
            w = {64 / 225., 0.239314335249683234, 0.239314335249683234, 0.1184634425280945438, 0.1184634425280945438}; break;
    default: throw std::runtime_error("gaussRule: unsupported point count");
    }
}
inline int gaussPointsForDegree(int degree) { return degree / 2 + 1; }

// ---------------------------------------------------------------------------
// Regular grid simulator, Q1 elements (TPS.hh:219-1419)
// ---------------------------------------------------------------------------
template <int N>
struct Sim {
    static constexpr int NPE = 1 << N;   // nodes per element
    static constexpr int KS = N * NPE;   // element stiffness size (24 in 3-D, 8 in 2-D)
    static constexpr int NST = (N == 3) ? 27 : 9; // nodes in the 3^N neighbourhood

    size_t E[N], nper[N], ninc[N], einc[N];
    size_t nn = 0, ne = 0;
    Real mn[N], mx[N], h[N], spacing[N];
    size_t refNodes[NPE];
    Real lambda = 0.0, mu = 0.5; // default ETensor(1, 0) (TPS.hh:1379)
    Real K0[KS * KS];
    std::vector<uint8_t> dmask;   // nn*N Dirichlet component flags
    std::vector<Real> dval;       // nn*N Dirichlet values
    std::vector<Real> force;      // nn*N
    std::vector<uint8_t> hasForce;
    std::vector<Real> rho;
    Real E0 = 1, Emin = 1e-9, gamma = 3; // TPS.hh:1392-1394

    std::vector<Real> KeCache;    // ne*KS*KS when cached (coarse levels)
    std::vector<Real> stencil;    // nn*NST*N*N assembled block rows ("blockK", TPS.hh:649-720)
    bool hasBlockK = false;

    // banded Cholesky factor for the direct solve (stands in for CHOLMOD, TPS.hh:834-865)
    mutable std::vector<Real> band;
    mutable size_t bandw = 0;
    mutable bool factorUpToDate = false;

    // TPS.hh:252-316
    Sim(const Real *bbmin, const Real *bbmax, const size_t *elems) {
        nn = 1; ne = 1;
        for (int d = 0; d < N; ++d) {
            E[d] = elems[d]; nper[d] = elems[d] + 1;
            mn[d] = bbmin[d]; mx[d] = bbmax[d];
            nn *= nper[d]; ne *= E[d];
        }
        ninc[N - 1] = 1; einc[N - 1] = 1;
        for (int d = N - 2; d >= 0; --d) { ninc[d] = ninc[d + 1] * nper[d + 1]; einc[d] = einc[d + 1] * E[d + 1]; }
        for (int d = 0; d < N; ++d) {
            Real dim = mx[d] - mn[d];
            spacing[d] = dim / (Real(nper[d]) - 1.0);  // TPS.hh:275
            h[d] = dim / Real(E[d]);                   // TPS.hh:287
        }
        for (int j = 0; j < NPE; ++j) {  // local node j, first axis slowest (NDArray.hh:100-105)
            size_t off = 0;
            for (int d = 0; d < N; ++d) off += size_t((j >> (N - 1 - d)) & 1) * ninc[d];
            refNodes[j] = off;
        }
        dmask.assign(nn * N, 0); dval.assign(nn * N, 0.0);
        force.assign(nn * N, 0.0); hasForce.assign(nn, 0);
        rho.assign(ne, 0.0); // NDVector<Real> value-initialises
        updateK0();
    }

    void nodeIndex(size_t ni, size_t *idx) const { for (int d = N - 1; d >= 0; --d) { idx[d] = ni % nper[d]; ni /= nper[d]; } }
    void elemIndex(size_t ei, size_t *idx) const { for (int d = N - 1; d >= 0; --d) { idx[d] = ei % E[d]; ei /= E[d]; } }
    size_t flatNode(const size_t *idx) const { size_t r = 0; for (int d = 0; d < N; ++d) r += idx[d] * ninc[d]; return r; }
    size_t flatElem(const size_t *idx) const { size_t r = 0; for (int d = 0; d < N; ++d) r += idx[d] * einc[d]; return r; }
    // TPS.hh:998-1012
    size_t firstNodeOfElement(size_t ei) const { size_t idx[N]; elemIndex(ei, idx); return flatNode(idx); }
    void nodePosition(size_t ni, Real *p) const {  // TPS.hh:274-279
        size_t idx[N]; nodeIndex(ni, idx);
        for (int d = 0; d < N; ++d) p[d] = mn[d] + Real(idx[d]) * spacing[d];
    }

    // ElasticityTensor.hh:100-115 (plane stress in 2-D)
    void setIsotropic(Real young, Real nu) {
        lambda = (nu * young) / ((1.0 + nu) * (1.0 - 2.0 * nu));
        mu = young / (2.0 + 2.0 * nu);
        if (N == 2) lambda = (nu * young) / (1.0 - nu * nu);
        updateK0();
    }

    // K0 = vol * int eps_i : C : eps_j, 2-point Gauss per axis for Q1
    // (TPS.hh:127-140,1358-1366; TensorProductPolynomialInterpolant.hh:161-264)
    void updateK0() {
        std::vector<Real> gx, gw; gaussRule(gaussPointsForDegree(2), gx, gw);
        const int ng = int(gx.size());
        int ngp = 1; for (int d = 0; d < N; ++d) ngp *= ng;
        Real vol = 1; for (int d = 0; d < N; ++d) vol *= h[d];
        // strains at Gauss points: eps[basis][gp][N*N]
        std::vector<Real> eps(size_t(KS) * ngp * N * N, 0.0), wgp(ngp);
        for (int g = 0; g < ngp; ++g) {
            Real p[N]; Real w = 1; int t = g;
            for (int d = N - 1; d >= 0; --d) { p[d] = gx[t % ng]; w *= gw[t % ng]; t /= ng; }
            wgp[g] = w;
            for (int a = 0; a < NPE; ++a) {
                Real grad[N];
                for (int i = 0; i < N; ++i) {
                    Real v = 1;
                    for (int d = 0; d < N; ++d) {
                        int bit = (a >> (N - 1 - d)) & 1;
                        if (d == i) v *= (bit ? 1.0 : -1.0) / h[d];
                        else v *= bit ? p[d] : (1.0 - p[d]);
                    }
                    grad[i] = v;
                }
                for (int c = 0; c < N; ++c) {
                    Real *e = &eps[(size_t(a * N + c) * ngp + g) * N * N];
                    for (int i = 0; i < N; ++i) { e[c * N + i] = 0.5 * grad[i]; e[i * N + c] = 0.5 * grad[i]; }
                    e[c * N + c] = grad[c];
                }
            }
        }
        for (int i = 0; i < KS; ++i)
            for (int j = i; j < KS; ++j) {
                Real acc = 0;
                for (int g = 0; g < ngp; ++g) {
                    const Real *ei = &eps[(size_t(i) * ngp + g) * N * N];
                    const Real *ej = &eps[(size_t(j) * ngp + g) * N * N];
                    Real tr = 0; for (int d = 0; d < N; ++d) tr += ej[d * N + d];
                    Real s = 0;
                    for (int a = 0; a < N; ++a)
                        for (int b = 0; b < N; ++b) {
                            Real sig = 2.0 * mu * ej[a * N + b] + (a == b ? lambda * tr : 0.0);
                            s += ei[a * N + b] * sig;
                        }
                    acc += wgp[g] * s;
                }
                K0[i * KS + j] = acc * vol;
                K0[j * KS + i] = K0[i * KS + j];
            }
        factorUpToDate = false;
    }

    // SIMP (TPS.hh:725-727)
    Real youngModulus(size_t ei) const { return Emin + std::pow(rho[ei], gamma) * (E0 - Emin); }

    bool hasCachedKe() const { return KeCache.size() == ne * size_t(KS) * KS; }
    const Real *elementKe(size_t ei) const { return &KeCache[ei * size_t(KS) * KS]; }

    // ---- boundary conditions (TPS.hh:358-409, Node::setDirichlet TPS.hh:49-63) ----
    bool inBox(const Real *p, const Real *lo, const Real *hi) const {
        for (int d = 0; d < N; ++d) if (!(p[d] >= lo[d] && p[d] <= hi[d])) return false; // Geometry.hh:276-279
        return true;
    }
    void addForceRegion(const Real *lo, const Real *hi, const Real *value) {
        std::vector<size_t> in;
        Real p[N];
        for (size_t ni = 0; ni < nn; ++ni) { nodePosition(ni, p); if (inBox(p, lo, hi)) in.push_back(ni); }
        if (in.empty()) throw std::runtime_error("Force constraint region unmatched");
        for (size_t ni : in) {
            hasForce[ni] = 1;
            for (int d = 0; d < N; ++d) force[ni * N + d] = value[d] / Real(in.size()); // overwrite (TPS.hh:387)
        }
    }
    void setDirichlet(size_t ni, const uint8_t *mask, const Real *val) {
        for (int c = 0; c < N; ++c) {
            if (!mask[c]) continue;
            if (!dmask[ni * N + c]) { dmask[ni * N + c] = 1; dval[ni * N + c] = val[c]; }
            else if (std::abs(dval[ni * N + c] - val[c]) > 1e-10) throw std::runtime_error("Conflicting dirichlet displacements.");
        }
    }
    void addDirichletRegion(const Real *lo, const Real *hi, const uint8_t *mask, const Real *value) {
        size_t cnt = 0; Real p[N];
        for (size_t ni = 0; ni < nn; ++ni) { nodePosition(ni, p); if (inBox(p, lo, hi)) { ++cnt; setDirichlet(ni, mask, value); } }
        if (cnt == 0) throw std::runtime_error("Dirichlet region unmatched");
        factorUpToDate = false;
    }
    void buildLoadVector(Real *f) const { // TPS.hh:893-901
        for (size_t i = 0; i < nn * N; ++i) f[i] = 0;
        for (size_t ni = 0; ni < nn; ++ni) if (hasForce[ni]) for (int d = 0; d < N; ++d) f[ni * N + d] += force[ni * N + d];
    }

    // ---- K u (TPS.hh:905-952).  Elements are visited in 2^N parity colours so the
    // OpenMP scatter is race-free and deterministic (the reference uses per-thread
    // accumulators; summation order is roundoff-only). ----
    void applyK(const Real *u, Real *out) const {
        for (size_t i = 0; i < nn * N; ++i) out[i] = 0;
        const bool cached = hasCachedKe();
        for (int color = 0; color < NPE; ++color) {
            size_t cnt[N], tot = 1;
            for (int d = 0; d < N; ++d) { int par = (color >> (N - 1 - d)) & 1; cnt[d] = (E[d] + 1 - par) / 2; tot *= cnt[d]; }
#pragma omp parallel for schedule(static)
            for (long long t = 0; t < (long long)tot; ++t) {
                size_t idx[N]; size_t r = size_t(t);
                for (int d = N - 1; d >= 0; --d) { int par = (color >> (N - 1 - d)) & 1; idx[d] = 2 * (r % cnt[d]) + par; r /= cnt[d]; }
                size_t ei = flatElem(idx), base = flatNode(idx);
                Real ul[KS], kl[KS];
                for (int m = 0; m < NPE; ++m) for (int c = 0; c < N; ++c) ul[m * N + c] = u[(base + refNodes[m]) * N + c];
                const Real *Ke = cached ? elementKe(ei) : K0;
                for (int i = 0; i < KS; ++i) { Real s = 0; for (int j = 0; j < KS; ++j) s += Ke[i * KS + j] * ul[j]; kl[i] = s; }
                if (!cached) { Real Ey = youngModulus(ei); for (int i = 0; i < KS; ++i) kl[i] *= Ey; }
                for (int m = 0; m < NPE; ++m) for (int c = 0; c < N; ++c) out[(base + refNodes[m]) * N + c] += kl[m * N + c];
            }
        }
    }

    // Assembled block-row operator (TPS.hh:649-720): stencil[n][k] = K_{n, n+off(k)}.
    static int stencilSlot(const int *delta) { int k = 0; for (int d = 0; d < N; ++d) k = k * 3 + (delta[d] + 1); return k; }
    void updateBlockK() {
        stencil.assign(nn * size_t(NST) * N * N, 0.0);
        const bool cached = hasCachedKe();
        std::vector<Real> tmp(size_t(KS) * KS);
        for (size_t ei = 0; ei < ne; ++ei) {
            size_t base = firstNodeOfElement(ei);
            const Real *Ke;
            if (cached) Ke = elementKe(ei);
            else { Real Ey = youngModulus(ei); for (int i = 0; i < KS * KS; ++i) tmp[i] = Ey * K0[i]; Ke = tmp.data(); }
            for (int i = 0; i < NPE; ++i)
                for (int j = 0; j < NPE; ++j) {
                    int delta[N];
                    for (int d = 0; d < N; ++d) delta[d] = ((j >> (N - 1 - d)) & 1) - ((i >> (N - 1 - d)) & 1);
                    Real *blk = &stencil[((base + refNodes[i]) * NST + stencilSlot(delta)) * N * N];
                    for (int a = 0; a < N; ++a) for (int b = 0; b < N; ++b) blk[a * N + b] += Ke[(i * N + a) * KS + (j * N + b)];
                }
        }
        hasBlockK = true;
    }
    void clearBlockK() { hasBlockK = false; stencil.clear(); }
    // neighbour flat index or SIZE_MAX if outside the grid
    size_t neighbour(const size_t *idx, int k) const {
        size_t r = 0; int t = k;
        int delta[N];
        for (int d = N - 1; d >= 0; --d) { delta[d] = t % 3 - 1; t /= 3; }
        for (int d = 0; d < N; ++d) {
            long long v = (long long)idx[d] + delta[d];
            if (v < 0 || v >= (long long)nper[d]) return SIZE_MAX;
            r += size_t(v) * ninc[d];
        }
        return r;
    }
    void applyBlockK(const Real *u, Real *out) const { // TPS.hh:960-964
#pragma omp parallel for schedule(static)
        for (long long n = 0; n < (long long)nn; ++n) {
            size_t idx[N]; nodeIndex(size_t(n), idx);
            Real acc[N]; for (int a = 0; a < N; ++a) acc[a] = 0;
            for (int k = 0; k < NST; ++k) {
                size_t m = neighbour(idx, k); if (m == SIZE_MAX) continue;
                const Real *blk = &stencil[(size_t(n) * NST + k) * N * N];
                for (int a = 0; a < N; ++a) for (int b = 0; b < N; ++b) acc[a] += blk[a * N + b] * u[m * N + b];
            }
            for (int a = 0; a < N; ++a) out[size_t(n) * N + a] = acc[a];
        }
    }

    // compliance gradient (TPS.hh:730-751)
    void complianceGradient(const Real *u, Real *g) const {
#pragma omp parallel for schedule(static)
        for (long long e = 0; e < (long long)ne; ++e) {
            size_t base = firstNodeOfElement(size_t(e));
            Real ul[KS];
            for (int m = 0; m < NPE; ++m) for (int c = 0; c < N; ++c) ul[m * N + c] = u[(base + refNodes[m]) * N + c];
            Real q = 0;
            for (int i = 0; i < KS; ++i) { Real s = 0; for (int j = 0; j < KS; ++j) s += K0[i * KS + j] * ul[j]; q += ul[i] * s; }
            g[e] = -0.5 * gamma * std::pow(rho[e], gamma - 1.0) * (E0 - Emin) * q;
        }
    }

    void solveDirect(const Real *f, Real *x) const; // defined in vfem_oracle_solve.hpp
};

} // namespace orc
