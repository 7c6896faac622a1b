// ndr_b200 — TensorProductSimulator host logic + fine-level launches.
#include "sim.h"
#include "kernels.cuh"
#include "kernels_vec.cuh"
#include "block_tri.h"
#include "apply_march_tma.cuh"

namespace vfem {

unsigned long long g_launchCount = 0;

// ---- direct solve (TPS::solve, TPS.hh:834-865) ----------------------------------------------------------------------
// Fine-level operator assembled as a 3^N-point block stencil: S[n][k] = sum over the elements e containing both n and its
// neighbour n + delta(k) of E_e K0[(loc_e(n), :), (loc_e(n + delta), :)]  (what TPS::m_assembleStiffnessMatrix, TPS.hh:589-625,
// builds as a CSC matrix).  One thread per (node, slot).
template <int N>
static __global__ void __launch_bounds__(kBlock) buildFineStencilKernel(Grid<N> g, K0Param<N> K, const Real *__restrict__ Ey, Real *__restrict__ S) {
    constexpr int NPE = 1 << N, KS = N * NPE, NST = ipow3(N);
    const long long t = (long long)blockIdx.x * kBlock + threadIdx.x;
    if (t >= (long long)g.nn * NST) return;
    const int n = int(t / NST), k = int(t % NST);
    int idx[N], dl[N];
    unflattenNode<N>(g, n, idx);
    { int tt = k; for (int d = N - 1; d >= 0; --d) { dl[d] = tt % 3 - 1; tt /= 3; } }
    Real blk[N * N];
#pragma unroll
    for (int i = 0; i < N * N; ++i) blk[i] = 0.0;
    for (int s = 0; s < NPE; ++s) {
        int ei = 0, locI = 0, locJ = 0; bool ok = true;
#pragma unroll
        for (int d = 0; d < N; ++d) {
            const int sd = (s >> (N - 1 - d)) & 1, e = idx[d] - 1 + sd;
            ok = ok && e >= 0 && e < g.E[d];
            const int li = 1 - sd, lj = li + dl[d];
            ok = ok && lj >= 0 && lj <= 1;
            ei += e * g.einc[d]; locI |= li << (N - 1 - d); locJ |= lj << (N - 1 - d);
        }
        if (!ok) continue;
        const Real Ee = Ey[ei];
#pragma unroll
        for (int a = 0; a < N; ++a)
#pragma unroll
            for (int b = 0; b < N; ++b) blk[a * N + b] = fma(Ee, K.k[(locI * N + a) * KS + locJ * N + b], blk[a * N + b]);
    }
#pragma unroll
    for (int i = 0; i < N * N; ++i) S[stencilIndex<N>(n, k, i)] = blk[i];
}

template <int N>
struct DirectSolver {
    DevBuf<Real> S, r, dx;
    BlockTri<N> bt;
    unsigned long long densityVersion = 0, k0Version = 0, bcVersion = 0;
};

template <int N>
Sim<N>::~Sim() {
    for (cudaStream_t st : {h2dStream, d2hStream, hostComputeStream}) if (st) cudaStreamDestroy(st);
    for (int k = 0; k < kHostChunks; ++k) for (cudaEvent_t e : {evIn[k], evKernel[k], evOut[k]}) if (e) cudaEventDestroy(e);
}

template <int N>
void Sim<N>::solveDirect(const Real *f, Real *u, int refinements) {
    if (globalE0 != g.E[0]) throw std::runtime_error("direct solve is not available on a slab of a larger grid");
    for (Real v : dval) if (v != 0.0) throw std::runtime_error("Non-zero Dirichlet conditions are not supported by the direct solver"); // TPS.hh:846-848 (the reference throws too)
    syncBC();
    if (!direct) direct.reset(new DirectSolver<N>());
    DirectSolver<N> &ds = *direct;
    const size_t nd = numDofs();
    const Real *Ey = moduli();
    if (ds.densityVersion != densityVersion || ds.k0Version != k0Version || ds.bcVersion != bcVersion) { // numeric factorisation (TPS.hh:854-858)
        ds.S.alloc(size_t(g.nn) * NST * N * N); ds.r.alloc(nd); ds.dx.alloc(nd);
        const long long tot = (long long)g.nn * NST;
        buildFineStencilKernel<N><<<unsigned((tot + kBlock - 1) / kBlock), kBlock, 0, stream>>>(g, K0, Ey, ds.S.p);
        VFEM_CHECK_LAUNCH(); countLaunch();
        ds.bt.factor(g, ds.S.p, d_dmask.p, stream);
        ds.bt.check(stream);
        ds.densityVersion = densityVersion; ds.k0Version = k0Version; ds.bcVersion = bcVersion;
    }
    // the chain kernel masks its right-hand side and its result: the solution is 0 on the constrained dofs (TPS.hh:869-887)
    ds.bt.solve(f, u, d_dmask.p, stream);
    for (int it = 0; it < refinements; ++it) { // r = mask(f - K u) matrix-free, u += A^-1 r
        applyFine<1>(u, f, d_dmask.p, ds.r.p);
        ds.bt.solve(ds.r.p, ds.dx.p, d_dmask.p, stream);
        axpyKernel<<<reduceBlocks(nd), kRedBlock, 0, stream>>>(u, ds.dx.p, 1.0, nd);
        VFEM_CHECK_LAUNCH(); countLaunch();
    }
}

template <int N>
Sim<N>::Sim(const Real *bbmin, const Real *bbmax, const size_t *elems, size_t globalE0_, size_t layerLo_) {
    dim = N;
    if (globalE0_ != 0 && layerLo_ + elems[0] > globalE0_) throw std::runtime_error("slab layers exceed the global grid");
    globalE0 = int(globalE0_ ? globalE0_ : elems[0]); layerLo = int(globalE0_ ? layerLo_ : 0);
    int E[N];
    for (int d = 0; d < N; ++d) {
        if (elems[d] == 0 || elems[d] > (1u << 20)) throw std::runtime_error("invalid grid size");
        E[d] = int(elems[d]); mn[d] = bbmin[d]; mx[d] = bbmax[d];
    }
    g = makeGrid<N>(E);
    for (int d = 0; d < N; ++d) {
        Real len = mx[d] - mn[d];
        const int Eg = (d == 0) ? globalE0 : g.E[d];
        spacing[d] = len / (Real(Eg + 1) - 1.0); // TPS.hh:275
        h[d] = len / Real(Eg);                   // TPS.hh:287
    }
    dmask.assign(g.nn, 0); dval.assign(size_t(g.nn) * N, 0.0);
    force.assign(size_t(g.nn) * N, 0.0); hasForce.assign(g.nn, 0);
    if (const char *v = getenv("VFEM_APPLY_VARIANT")) applyVariant = atoi(v); // experiment knob (bench / profiles)
    d_rho.alloc(g.ne); d_rho.zero(stream); // NDVector value-initialises densities to 0
    d_Ey.alloc(g.ne);
    if (N == 3) {
        int dev = 0, sms = 148;
        if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        marchPlan = planMarch(reinterpret_cast<const Grid<3> &>(g), sms > 0 ? sms : 148);
    }
    updateK0();
}

template <int N>
void Sim<N>::setIsotropic(Real young, Real nu) {
    lambda = (N == 2) ? (nu * young) / (1.0 - nu * nu)                      // plane stress
                      : (nu * young) / ((1.0 + nu) * (1.0 - 2.0 * nu));
    mu = young / (2.0 + 2.0 * nu);
    updateK0();
}

// K0[(p,a),(q,b)] = vol * sum_gp w * [ lambda d_a phi_p d_b phi_q + mu d_b phi_p d_a phi_q
//                                     + mu delta_ab grad phi_p . grad phi_q ]
// which is eps(phi_p e_a) : C : eps(phi_q e_b) for isotropic C, integrated with the 2-point
// Gauss rule per axis (TensorProductQuadrature.hh:123-132); gradients scaled by 1/h_d.
template <int N>
void Sim<N>::updateK0() {
    const Real gp[2] = {0.5 - 0.5 / std::sqrt(3.0), 0.5 + 0.5 / std::sqrt(3.0)};
    Real vol = 1.0;
    for (int d = 0; d < N; ++d) vol *= h[d];
    for (int i = 0; i < KS * KS; ++i) K0.k[i] = 0.0;
    for (int q = 0; q < (1 << N); ++q) { // quadrature point
        Real xi[N], w = 1.0;
        for (int d = 0; d < N; ++d) { xi[d] = gp[(q >> (N - 1 - d)) & 1]; w *= 0.5; }
        Real grad[NPE][N];
        for (int p = 0; p < NPE; ++p)
            for (int a = 0; a < N; ++a) {
                Real v = 1.0;
                for (int d = 0; d < N; ++d) {
                    int bit = (p >> (N - 1 - d)) & 1;
                    if (d == a) v *= (bit ? 1.0 : -1.0) / h[d];
                    else v *= bit ? xi[d] : 1.0 - xi[d];
                }
                grad[p][a] = v;
            }
        for (int p = 0; p < NPE; ++p)
            for (int r = 0; r < NPE; ++r) {
                Real gg = 0.0;
                for (int k = 0; k < N; ++k) gg += grad[p][k] * grad[r][k];
                for (int a = 0; a < N; ++a)
                    for (int b = 0; b < N; ++b) {
                        Real v = lambda * grad[p][a] * grad[r][b] + mu * grad[p][b] * grad[r][a] + (a == b ? mu * gg : 0.0);
                        K0.k[(p * N + a) * KS + r * N + b] += w * vol * v;
                    }
            }
    }
    // exact symmetry (the reference mirrors the upper triangle, TPS.hh:136-138)
    for (int i = 0; i < KS; ++i)
        for (int j = i + 1; j < KS; ++j) K0.k[j * KS + i] = K0.k[i * KS + j];
    if (N == 3) khatStructure = computeKhat(K0.k, Khat);
    ++k0Version;
    touchStiffness();
}

template <int N>
static bool inRegion(const Real *p, const Real *lo, const Real *hi) { // inclusive, Geometry.hh:276-279
    for (int d = 0; d < N; ++d) if (!(p[d] >= lo[d] && p[d] <= hi[d])) return false;
    return true;
}

template <int N>
size_t Sim<N>::countRegionNodes(const Real *lo, const Real *hi, size_t planeLo, size_t planeHi) const {
    size_t cnt = 0; Real p[N];
    const size_t pn = size_t(g.ninc[0]);
    for (size_t ni = planeLo * pn; ni < (planeHi + 1) * pn && ni < size_t(g.nn); ++ni) { nodePosition(ni, p); if (inRegion<N>(p, lo, hi)) ++cnt; }
    return cnt;
}

template <int N>
void Sim<N>::addForceRegion(const Real *lo, const Real *hi, const Real *value, size_t divisor) {
    std::vector<size_t> hit;
    Real p[N];
    for (size_t ni = 0; ni < size_t(g.nn); ++ni) { nodePosition(ni, p); if (inRegion<N>(p, lo, hi)) hit.push_back(ni); }
    if (hit.empty() && divisor == 0) throw std::runtime_error("Force constraint region unmatched");
    const Real cnt = Real(divisor ? divisor : hit.size());
    for (size_t ni : hit) {
        hasForce[ni] = 1;
        for (int d = 0; d < N; ++d) force[ni * N + d] = value[d] / cnt; // overwrite, TPS.hh:385-388
    }
}

template <int N>
void Sim<N>::setDirichlet(size_t ni, const uint8_t *compMask, const Real *value) {
    for (int c = 0; c < N; ++c) {
        if (!compMask[c]) continue;
        if (!((dmask[ni] >> c) & 1)) { dmask[ni] |= uint8_t(1u << c); dval[ni * N + c] = value[c]; }
        else if (std::abs(dval[ni * N + c] - value[c]) > 1e-10) throw std::runtime_error("Conflicting dirichlet displacements.");
    }
    bcDirty = true; ++bcVersion;
}

template <int N>
void Sim<N>::addDirichletRegion(const Real *lo, const Real *hi, const uint8_t *compMask, const Real *value) {
    size_t cnt = 0; Real p[N];
    for (size_t ni = 0; ni < size_t(g.nn); ++ni) { nodePosition(ni, p); if (inRegion<N>(p, lo, hi)) { ++cnt; setDirichlet(ni, compMask, value); } }
    if (cnt == 0 && globalE0 == g.E[0]) throw std::runtime_error("Dirichlet region unmatched"); // a slab may legitimately miss the region
}

template <int N>
void Sim<N>::buildLoadVectorHost(Real *f) const {
    const size_t nd = numDofs();
    for (size_t i = 0; i < nd; ++i) f[i] = 0.0;
    for (size_t ni = 0; ni < size_t(g.nn); ++ni)
        if (hasForce[ni]) for (int d = 0; d < N; ++d) f[ni * N + d] += force[ni * N + d];
}

template <int N>
void Sim<N>::syncBC() {
    if (!bcDirty && d_dmask.n == size_t(g.nn)) return;
    d_dmask.upload(dmask, stream);
    d_dval.upload(dval, stream);
    VFEM_CUDA(cudaStreamSynchronize(stream));
    bcDirty = false;
}

template <int N>
void Sim<N>::setDensities(const Real *rho, MemSpace ms) {
    VFEM_CUDA(cudaMemcpyAsync(d_rho.p, rho, size_t(g.ne) * sizeof(Real), ms == MS_HOST ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToDevice, stream));
    if (ms == MS_HOST) VFEM_CUDA(cudaStreamSynchronize(stream));
    touchStiffness();
}

template <int N>
void Sim<N>::setUniformDensities(Real v) {
    if (v < 0 || v > 1) throw std::runtime_error("Density value must be in [0, 1]"); // TPS.hh:457
    std::vector<Real> hrho(g.ne, v);
    setDensities(hrho.data(), MS_HOST);
}

template <int N>
void Sim<N>::setElementDensity(size_t ei, Real v) {
    if (v < 0 || v > 1) throw std::runtime_error("Density value must be in [0, 1]"); // TPS.hh:464
    if (ei >= size_t(g.ne)) throw std::runtime_error("element index out of range");
    VFEM_CUDA(cudaMemcpyAsync(d_rho.p + ei, &v, sizeof(Real), cudaMemcpyHostToDevice, stream));
    VFEM_CUDA(cudaStreamSynchronize(stream));
    touchStiffness();
}

template <int N>
Real Sim<N>::elementDensity(size_t ei) const {
    if (ei >= size_t(g.ne)) throw std::runtime_error("element index out of range");
    Real v;
    VFEM_CUDA(cudaMemcpyAsync(&v, d_rho.p + ei, sizeof(Real), cudaMemcpyDeviceToHost, stream));
    VFEM_CUDA(cudaStreamSynchronize(stream));
    return v;
}

template <int N>
void Sim<N>::getDensities(Real *out, MemSpace ms) const {
    VFEM_CUDA(cudaMemcpyAsync(out, d_rho.p, size_t(g.ne) * sizeof(Real), ms == MS_HOST ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice, stream));
    VFEM_CUDA(cudaStreamSynchronize(stream));
}

template <int N>
const Real *Sim<N>::moduli() {
    if (EyDirty) {
        int blocks = reduceBlocks(g.ne);
        simpKernel<<<blocks, kRedBlock, 0, stream>>>(d_rho.p, d_Ey.p, size_t(g.ne), E0, Emin, gamma);
        VFEM_CHECK_LAUNCH(); countLaunch();
        EyDirty = false;
    }
    return d_Ey.p;
}

// K u restricted to the element layers [layerLo, layerHi) of the local grid, by viewing the grid from layer layerLo on
// (the kernels do not depend on the absolute plane index).  Node planes layerLo+1 .. layerHi-1 (and the grid's own
// boundary planes) receive complete rows, the two edge planes of the view partial ones.
template <int N>
template <int MODE>
void Sim<N>::applyFineRange(const Real *u, const Real *b, const uint8_t *mask, Real *out, int layerLo_, int layerHi_, cudaStream_t st,
                            int edgeSkip) {
    const Real *Ey = moduli();
    Grid<N> gv = g;
    const MarchPlan *plan = &marchPlan;
    if (layerLo_ > 0 || layerHi_ < g.E[0]) {
        int E[N];
        for (int d = 0; d < N; ++d) E[d] = g.E[d];
        E[0] = layerHi_ - layerLo_;
        gv = makeGrid<N>(E);
        const size_t po = size_t(layerLo_) * g.ninc[0];
        u += po * N; out += po * N; Ey += size_t(layerLo_) * g.einc[0];
        if (b) b += po * N;
        if (mask) mask += po;
        if (N == 3) {
            auto it = viewPlans.find(E[0]);
            if (it == viewPlans.end()) {
                int dev = 0, sms = 148;
                if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
                it = viewPlans.emplace(E[0], planMarch(reinterpret_cast<const Grid<3> &>(gv), sms > 0 ? sms : 148)).first;
            }
            plan = &it->second;
        }
    }
    if (useMarch() && applyVariant != 4)   // ring-fed marching kernel (apply_march_tma.cuh)
        launchApplyMarchTma<MODE>(reinterpret_cast<const Grid<3> &>(gv), Khat, khatStructure == 2 && applyVariant != 2, *plan, Ey, u, b, mask, out, st,
                                  edgeSkip);
    else if (useMarch())                   // variant 4: first-generation marching kernel (per-thread global loads, CTA barrier per plane)
        launchApplyMarch<MODE>(reinterpret_cast<const Grid<3> &>(gv), Khat, khatStructure == 2, *plan, Ey, u, b, mask, out, st, edgeSkip);
    else
        applyFineKernel<N, MODE><<<(gv.nn + kBlock - 1) / kBlock, kBlock, 0, st>>>(gv, K0, Ey, u, b, mask, out, edgeSkip);
    VFEM_CHECK_LAUNCH(); countLaunch();
}

// slab solver: the element layers below computeLo only feed ghost planes nobody reads - skipped
template <int N>
template <int MODE>
void Sim<N>::applyFine(const Real *u, const Real *b, const uint8_t *mask, Real *out) {
    applyFineRange<MODE>(u, b, mask, out, computeLo, g.E[0], stream);
}

// Reference-facing call with HOST arrays (applyK(numpy) -> numpy): instead of copy-in / kernel / copy-out, the grid is cut
// into chunks of node planes along axis 0 and the three stages are pipelined over two copy streams and the compute stream,
// so host->device and device->host traffic overlap (PCIe is full duplex) and the kernel time disappears behind them.
// Chunk k owns the node planes [a_k, b_k); its kernel covers the element layers [a_k - 1, b_k) and stores the complete
// rows only (edgeSkip drops the partial planes a_k - 1 and b_k), so the chunks write disjoint planes.
template <int N>
void Sim<N>::applyKHost(const Real *hu, Real *hout) {
    const size_t nd = numDofs(), pd = size_t(g.ninc[0]) * N;
    if (hostIn.n != nd) { hostIn.alloc(nd); hostOut.alloc(nd); }
    if (!h2dStream) {
        if (const char *v = getenv("VFEM_HOST_CHUNKS")) hostChunks = std::max(1, atoi(v));
        VFEM_CUDA(cudaStreamCreateWithFlags(&h2dStream, cudaStreamNonBlocking));
        VFEM_CUDA(cudaStreamCreateWithFlags(&d2hStream, cudaStreamNonBlocking));
        VFEM_CUDA(cudaStreamCreateWithFlags(&hostComputeStream, cudaStreamNonBlocking));
        for (int k = 0; k < kHostChunks; ++k) {
            VFEM_CUDA(cudaEventCreateWithFlags(&evIn[k], cudaEventDisableTiming));
            VFEM_CUDA(cudaEventCreateWithFlags(&evKernel[k], cudaEventDisableTiming));
            VFEM_CUDA(cudaEventCreateWithFlags(&evOut[k], cudaEventDisableTiming));
        }
    }
    moduli(); // refresh E(rho) on the main stream if stale, then make the side streams see it
    VFEM_CUDA(cudaStreamSynchronize(stream));
    const int E0 = g.E[0], planes = g.n[0];
    const int C = std::max(1, std::min(std::min(hostChunks, kHostChunks), E0 / 4));
    auto bound = [&](int k) { return k >= C ? planes : int((long long)E0 * k / C); }; // chunk k owns node planes [bound(k), bound(k+1))
    for (int k = 0; k < C; ++k) { // inputs: chunk k's kernel needs the node planes up to bound(k+1) inclusive
        const int p0 = k == 0 ? 0 : std::min(bound(k) + 1, planes), p1 = std::min(bound(k + 1) + 1, planes);
        if (p1 > p0) VFEM_CUDA(cudaMemcpyAsync(hostIn.p + size_t(p0) * pd, hu + size_t(p0) * pd, size_t(p1 - p0) * pd * sizeof(Real), cudaMemcpyHostToDevice, h2dStream));
        VFEM_CUDA(cudaEventRecord(evIn[k], h2dStream));
    }
    for (int k = 0; k < C; ++k) {
        const int a = bound(k), bnd = bound(k + 1);
        VFEM_CUDA(cudaStreamWaitEvent(hostComputeStream, evIn[k], 0));
        applyFineRange<0>(hostIn.p, nullptr, nullptr, hostOut.p, std::max(a - 1, 0), std::min(bnd, E0), hostComputeStream,
                          (a > 0 ? 1 : 0) | (bnd < E0 ? 2 : 0));
        VFEM_CUDA(cudaEventRecord(evKernel[k], hostComputeStream));
        VFEM_CUDA(cudaStreamWaitEvent(d2hStream, evKernel[k], 0));
        VFEM_CUDA(cudaMemcpyAsync(hout + size_t(a) * pd, hostOut.p + size_t(a) * pd, size_t(bnd - a) * pd * sizeof(Real), cudaMemcpyDeviceToHost, d2hStream));
        VFEM_CUDA(cudaEventRecord(evOut[k], d2hStream));
    }
    VFEM_CUDA(cudaStreamSynchronize(d2hStream));
}

template <int N>
void Sim<N>::applyK(const Real *u, Real *out) { applyFine<0>(u, nullptr, nullptr, out); }

// g_e = -1/2 gamma rho_e^(gamma-1) (E0 - Emin) u_e^T K0 u_e     (TPS.hh:730-751)
template <int N>
__global__ void __launch_bounds__(kBlock) complianceGradientKernel(Grid<N> g, K0Param<N> K, const Real *__restrict__ rho,
                                                                    const Real *__restrict__ u, Real *__restrict__ grad, Real gamma,
                                                                    Real dE) {
    constexpr int NPE = 1 << N, KS = N * NPE;
    int e = blockIdx.x * kBlock + threadIdx.x;
    if (e >= g.ne) return;
    int base = 0, r = e;
#pragma unroll
    for (int d = N - 1; d > 0; --d) { int q = r / g.E[d]; base += (r - q * g.E[d]) * g.ninc[d]; r = q; }
    base += r * g.ninc[0];
    Real ul[KS];
#pragma unroll
    for (int m = 0; m < NPE; ++m) {
        int off = 0;
#pragma unroll
        for (int d = 0; d < N; ++d) off += ((m >> (N - 1 - d)) & 1) * g.ninc[d];
#pragma unroll
        for (int c = 0; c < N; ++c) ul[m * N + c] = u[size_t(base + off) * N + c];
    }
    Real q = 0.0;
#pragma unroll
    for (int i = 0; i < KS; ++i) {
        Real s = 0.0;
#pragma unroll
        for (int j = 0; j < KS; ++j) s = fma(K.k[i * KS + j], ul[j], s);
        q = fma(ul[i], s, q);
    }
    grad[e] = -0.5 * gamma * pow(rho[e], gamma - 1.0) * dE * q;
}

template <int N>
void Sim<N>::complianceGradient(const Real *u, Real *grad) {
    complianceGradientKernel<N><<<(g.ne + kBlock - 1) / kBlock, kBlock, 0, stream>>>(g, K0, d_rho.p, u, grad, gamma, E0 - Emin);
    VFEM_CHECK_LAUNCH(); countLaunch();
}

template struct Sim<2>;
template struct Sim<3>;
#define VFEM_INST(N, M) template void Sim<N>::applyFine<M>(const Real *, const Real *, const uint8_t *, Real *);
VFEM_INST(2, 0) VFEM_INST(2, 1) VFEM_INST(2, 2) VFEM_INST(3, 0) VFEM_INST(3, 1) VFEM_INST(3, 2)
#undef VFEM_INST
#define VFEM_INST(N, M) template void Sim<N>::applyFineRange<M>(const Real *, const Real *, const uint8_t *, Real *, int, int, cudaStream_t, int);
VFEM_INST(2, 0) VFEM_INST(2, 1) VFEM_INST(2, 2) VFEM_INST(3, 0) VFEM_INST(3, 1) VFEM_INST(3, 2)
#undef VFEM_INST

} // namespace vfem
