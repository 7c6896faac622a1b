// ORACLE — TEST INFRASTRUCTURE ONLY (see vfem_oracle.hpp).
// Direct solve of the Dirichlet-reduced system (TPS.hh:834-887).  The reference
// factors with CHOLMOD (SuiteSparse 4.5.6, absent here); the solution of an SPD
// system is unique, so a banded Cholesky reproduces it to roundoff.  Fixed dofs are
// kept as identity rows/columns, which is algebraically the row/column removal of
// TPS.hh:849 followed by extractFullSolution (TPS.hh:878-887).
#pragma once
#include "vfem_oracle.hpp"

namespace orc {

template <int N>
void Sim<N>::solveDirect(const Real *f, Real *x) const {
    const size_t n = nn * N;
    for (size_t i = 0; i < n; ++i)
        if (dmask[i] && dval[i] != 0) throw std::runtime_error("Nonzero Dirichlet constraints currently unsupported"); // TPS.hh:840
    size_t span = 0; for (int d = 0; d < N; ++d) span += ninc[d];
    const size_t bw = N * span + (N - 1);
    const size_t W = bw + 1;
    if (!factorUpToDate || bandw != bw || band.size() != n * W) {
        bandw = bw;
        band.assign(n * W, 0.0);
        auto at = [&](size_t i, size_t j) -> Real & { return band[i * W + (j + bw - i)]; }; // j <= i
        const bool cached = hasCachedKe();
        // assembly (TPS.hh:589-625), lower triangle
        for (size_t ei = 0; ei < ne; ++ei) {
            size_t base = firstNodeOfElement(ei);
            Real Ey = cached ? 1.0 : youngModulus(ei);
            const Real *Ke = cached ? elementKe(ei) : K0;
            for (int a = 0; a < NPE; ++a)
                for (int b = 0; b < NPE; ++b)
                    for (int ca = 0; ca < N; ++ca)
                        for (int cb = 0; cb < N; ++cb) {
                            size_t i = (base + refNodes[a]) * N + ca, j = (base + refNodes[b]) * N + cb;
                            if (j > i) continue;
                            at(i, j) += Ey * Ke[(a * N + ca) * KS + (b * N + cb)];
                        }
        }
        // Dirichlet rows/cols -> identity
        for (size_t i = 0; i < n; ++i) {
            size_t lo = i > bw ? i - bw : 0;
            if (dmask[i]) { for (size_t j = lo; j < i; ++j) at(i, j) = 0; at(i, i) = 1; }
            else for (size_t j = lo; j < i; ++j) if (dmask[j]) at(i, j) = 0;
        }
        // banded Cholesky, row by row
        for (size_t i = 0; i < n; ++i) {
            size_t lo = i > bw ? i - bw : 0;
            Real *Li = &band[i * W + bw - i]; // Li[k] = L(i,k)
            for (size_t j = lo; j <= i; ++j) {
                const Real *Lj = &band[j * W + bw - j];
                size_t k0 = std::max(lo, j > bw ? j - bw : size_t(0));
                Real s = Li[j];
                Real acc = 0;
                for (size_t k = k0; k < j; ++k) acc += Li[k] * Lj[k];
                s -= acc;
                if (j < i) Li[j] = s / Lj[j];
                else {
                    if (!(s > 0)) throw std::runtime_error("Cholesky: matrix not positive definite");
                    Li[i] = std::sqrt(s);
                }
            }
        }
        factorUpToDate = true;
    }
    // forward / backward substitution
    std::vector<Real> y(n);
    for (size_t i = 0; i < n; ++i) {
        size_t lo = i > bw ? i - bw : 0;
        const Real *Li = &band[i * W + bw - i];
        Real s = dmask[i] ? 0.0 : f[i];
        for (size_t k = lo; k < i; ++k) s -= Li[k] * y[k];
        y[i] = s / Li[i];
    }
    for (size_t ii = n; ii-- > 0;) {
        Real s = y[ii] / band[ii * W + bw];
        y[ii] = s;
        size_t lo = ii > bw ? ii - bw : 0;
        const Real *Li = &band[ii * W + bw - ii];
        for (size_t k = lo; k < ii; ++k) y[k] -= Li[k] * s;
    }
    for (size_t i = 0; i < n; ++i) x[i] = dmask[i] ? 0.0 : y[i];
}

} // namespace orc
