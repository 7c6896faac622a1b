// ndr_b200 — MultigridSolver host logic: hierarchy setup, operator refresh, launches.
#include "mg.h"
#include "kernels.cuh"
#include "kernels_mg.cuh"
#include "coarse_solve.cuh"
#include <cublas_v2.h>
#include <cusolverDn.h>

namespace vfem {

static const size_t kMaxCoarseBytes = size_t(32) << 30; // block factors of the coarsest level

#define VFEM_CUSOLVER(call)                                                                              \
    do {                                                                                                 \
        cusolverStatus_t s_ = (call);                                                                    \
        if (s_ != CUSOLVER_STATUS_SUCCESS)                                                               \
            throw std::runtime_error("cuSOLVER error " + std::to_string(int(s_)) + " at " + __FILE__ + ":" + std::to_string(__LINE__)); \
    } while (0)

template <int N>
MG<N>::MG(std::shared_ptr<Sim<N>> fine, size_t numCoarseningLevels) : sim(fine) {
    dim = N;
    stream = fine->stream;
    useGraphs = getenv("VFEM_NO_GRAPH") == nullptr;
    if (const char *v = getenv("VFEM_GS_VARIANT")) gsVariant = atoi(v);
    fusedSweeps = getenv("VFEM_NO_FUSED_SWEEPS") == nullptr;
    int E[N];
    for (int d = 0; d < N; ++d) E[d] = fine->g.E[d];
    lv.resize(numCoarseningLevels + 1);
    for (size_t l = 0; l <= numCoarseningLevels; ++l) {
        Level &L = lv[l];
        if (l == 0) { L.g = fine->g; L.dmask = fine->dmask; }
        else {
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
            L.d_dmask.upload(L.dmask, stream);
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
    d_info.alloc(1);
    VFEM_CUDA(cudaStreamSynchronize(stream));
}

template <int N>
MG<N>::~MG() {
    if (pgExec) cudaGraphExecDestroy(pgExec);
    if (captureStream) cudaStreamDestroy(captureStream);
    if (h_scal) cudaFreeHost(h_scal);
    if (cusolverHandle) cusolverDnDestroy((cusolverDnHandle_t)cusolverHandle);
    if (cublasHandle) cublasDestroy((cublasHandle_t)cublasHandle);
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
        galerkinStencilKernel<N><<<unsigned((tot + kBlock - 1) / kBlock), kBlock, 0, stream>>>(lv[l - 1].g, lv[l].g, lv[l - 1].S.p, lv[l].S.p);
        VFEM_CHECK_LAUNCH(); countLaunch();
    }
    // coarsest level: block-tridiagonal Cholesky of the Dirichlet-reduced operator (coarse_solve.cuh)
    const Level &Lc = lv[L];
    const size_t nd = size_t(Lc.g.nn) * N;
    const int P = Lc.g.ninc[0] * N, planes = Lc.g.n[0];
    const size_t PP = size_t(P) * P;
    if (PP * size_t(6 * planes) * sizeof(Real) > kMaxCoarseBytes)
        throw std::runtime_error("coarsest multigrid level has " + std::to_string(nd) + " dofs in planes of " + std::to_string(P) +
                                 "; its block factorisation would need " + std::to_string(PP * size_t(6 * planes) * sizeof(Real) >> 20) +
                                 " MB (use more coarsening levels)");
    coarseDofs = int(nd); coarseP = P; coarsePlanes = planes;
    const size_t offBlocks = size_t(planes > 1 ? planes - 1 : 1);
    d_blkD.alloc(PP * planes); d_blkW.alloc(PP * planes); d_blkLo.alloc(PP * offBlocks);
    d_Sinv.alloc(PP * planes); d_Trm.alloc(PP * offBlocks); d_Tcm.alloc(PP * offBlocks); d_z.alloc(nd);
    d_blkD.zero(stream); d_blkLo.zero(stream);
    {
        long long tot = (long long)Lc.g.nn * NST;
        assembleBlockTriKernel<N><<<unsigned((tot + 127) / 128), 128, 0, stream>>>(Lc.g, Lc.S.p, Lc.d_dmask.p, d_blkD.p, d_blkLo.p, P);
        VFEM_CHECK_LAUNCH(); countLaunch();
        setIdentityBlocksKernel<<<296, 256, 0, stream>>>(d_blkW.p, P, planes);
        VFEM_CHECK_LAUNCH(); countLaunch();
    }
    if (!cusolverHandle) {
        cusolverDnHandle_t hnd; VFEM_CUSOLVER(cusolverDnCreate(&hnd)); cusolverHandle = hnd;
        cublasHandle_t bh; if (cublasCreate(&bh) != CUBLAS_STATUS_SUCCESS) throw std::runtime_error("cublasCreate failed"); cublasHandle = bh;
    }
    cusolverDnHandle_t hnd = (cusolverDnHandle_t)cusolverHandle;
    cublasHandle_t bh = (cublasHandle_t)cublasHandle;
    VFEM_CUSOLVER(cusolverDnSetStream(hnd, stream));
    if (cublasSetStream(bh, stream) != CUBLAS_STATUS_SUCCESS) throw std::runtime_error("cublasSetStream failed");
    int lw = 0;
    VFEM_CUSOLVER(cusolverDnDpotrf_bufferSize(hnd, CUBLAS_FILL_MODE_LOWER, P, d_blkD.p, P, &lw));
    if (d_solverWork.n < size_t(lw)) d_solverWork.alloc(size_t(lw));
    if (d_info.n < size_t(planes)) d_info.alloc(size_t(planes));
    const Real one = 1.0, mone = -1.0, zero = 0.0;
    auto gemm = [&](cublasOperation_t ta, cublasOperation_t tb, const Real *alpha, const Real *A, const Real *B, const Real *beta, Real *Cm) {
        if (cublasDgemm(bh, ta, tb, P, P, P, alpha, A, P, B, P, beta, Cm, P) != CUBLAS_STATUS_SUCCESS) throw std::runtime_error("cublasDgemm failed");
        countLaunch();
    };
    for (int p = 0; p < planes; ++p) {
        Real *Dp = d_blkD.p + PP * p, *Wp = d_blkW.p + PP * p, *Sp = d_Sinv.p + PP * p;
        VFEM_CUSOLVER(cusolverDnDpotrf(hnd, CUBLAS_FILL_MODE_LOWER, P, Dp, P, d_solverWork.p, lw, d_info.p + p));
        // W_p = C_p^{-1} (solve C_p X = I), S_p^{-1} = W_p^T W_p
        if (cublasDtrsm(bh, CUBLAS_SIDE_LEFT, CUBLAS_FILL_MODE_LOWER, CUBLAS_OP_N, CUBLAS_DIAG_NON_UNIT, P, P, &one, Dp, P, Wp, P) != CUBLAS_STATUS_SUCCESS)
            throw std::runtime_error("cublasDtrsm failed");
        countLaunch(2);
        gemm(CUBLAS_OP_T, CUBLAS_OP_N, &one, Wp, Wp, &zero, Sp);
        if (p + 1 < planes) {
            const Real *Ap = d_blkLo.p + PP * p; // A_{p+1,p}
            Real *Trm = d_Trm.p + PP * p, *Tcm = d_Tcm.p + PP * p, *Dn = d_blkD.p + PP * (p + 1);
            gemm(CUBLAS_OP_N, CUBLAS_OP_N, &one, Ap, Sp, &zero, Tcm);  // T_{p+1} = A_{p+1,p} S_p^{-1}
            gemm(CUBLAS_OP_N, CUBLAS_OP_T, &one, Sp, Ap, &zero, Trm);  // T_{p+1}^T (column-major) = T_{p+1} row-major
            gemm(CUBLAS_OP_N, CUBLAS_OP_T, &mone, Tcm, Ap, &one, Dn);  // S_{p+1} = A_{p+1,p+1} - T_{p+1} A_{p+1,p}^T
        }
    }
    std::vector<int> info(planes, 0);
    VFEM_CUDA(cudaMemcpyAsync(info.data(), d_info.p, sizeof(int) * planes, cudaMemcpyDeviceToHost, stream));
    VFEM_CUDA(cudaStreamSynchronize(stream));
    for (int p = 0; p < planes; ++p)
        if (info[p] != 0) throw std::runtime_error("Cholesky factorization failed: coarsest-level matrix not positive definite (plane " + std::to_string(p) + ", info " + std::to_string(info[p]) + ")");
    builtVersion = sim->densityVersion;
}

template <int N>
void MG<N>::coarseSolve(const Real *b, Real *x) {
    const Level &Lc = lv.back();
    const size_t smem = size_t(2) * coarseP * sizeof(Real);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(kChainCluster); cfg.blockDim = dim3(kChainThreads); cfg.dynamicSmemBytes = smem; cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = kChainCluster; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    const bool pf = coarseP <= 256 && coarseP <= 2 * kChainCluster * (kChainThreads / 32);
    auto kern = pf ? blockTriChainKernel<N, true> : blockTriChainKernel<N, false>;
    if (smem > 48 * 1024) VFEM_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    VFEM_CUDA(cudaLaunchKernelEx(&cfg, kern, coarseP, coarsePlanes, (const Real *)d_Trm.p, (const Real *)d_Tcm.p,
                                 (const Real *)d_Sinv.p, b, (const uint8_t *)Lc.d_dmask.p, x, d_z.p));
    countLaunch();
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

template <int N> void MG<N>::applyK(size_t l, const Real *u, Real *out) { applyMode<0>(l, u, nullptr, out); }
template <int N> void MG<N>::computeResidual(size_t l, const Real *u, const Real *b, Real *r) { applyMode<1>(l, u, b, r); }
template <int N> void MG<N>::applyKMasked(size_t l, const Real *u, Real *out) { applyMode<2>(l, u, nullptr, out); }

template <int N>
void MG<N>::smoothColoured(size_t l, Real *u, const Real *b, bool forward) {
    const Level &L = lv.at(l);
    const Real *Ey = (l == 0) ? sim->moduli() : nullptr;
    for (int i = 0; i < NPE; ++i) {
        const int colour = forward ? i : NPE - 1 - i; // MG.hh:318-320
        const int cnt = colourCount<N>(L.g, colour);
        const unsigned blocks = unsigned((cnt + kBlock - 1) / kBlock);
        if (l == 0) {
            if (gsVariant == 4) gsFineKernel<N, 4><<<blocks, kBlock, 0, stream>>>(L.g, sim->K0, Ey, u, b, maskOf(*this, 0), colour, forward ? 1 : 0);
            else if (gsVariant == 3) gsFineKernel<N, 3><<<blocks, kBlock, 0, stream>>>(L.g, sim->K0, Ey, u, b, maskOf(*this, 0), colour, forward ? 1 : 0);
            else gsFineKernel<N, 2><<<blocks, kBlock, 0, stream>>>(L.g, sim->K0, Ey, u, b, maskOf(*this, 0), colour, forward ? 1 : 0);
        }
        else gsStencilWarpKernel<N><<<warpGridBlocks(cnt), kWarpBlock, 0, stream>>>(L.g, L.S.p, u, b, maskOf(*this, l), colour, forward ? 1 : 0, cnt);
        VFEM_CHECK_LAUNCH(); countLaunch();
    }
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
        sweepGridBlocks = sms * (per > 0 ? per : 1);
    }
    const long long need = (long long)(colourCount<N>(L.g, 0)) * 32 / kWarpBlock + 1;
    const unsigned blocks = unsigned(need < sweepGridBlocks ? need : sweepGridBlocks);
    Grid<N> g = L.g; const Real *S = L.S.p; const uint8_t *m = maskOf(*this, l);
    int fw = forward ? 1 : 0, q0 = 0, q1 = int(sweeps) * NPE;
    void *args[] = {&g, &S, &u, &b, &m, &fw, &q0, &q1};
    VFEM_CUDA(cudaLaunchCooperativeKernel((const void *)gsStencilSweepKernel<N>, dim3(blocks), dim3(kWarpBlock), args, 0, stream));
    countLaunch();
}

template <int N>
void MG<N>::restriction(size_t lf, const Real *vf, Real *bc) {
    const Level &F = lv.at(lf), &C = lv.at(lf + 1);
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

} // namespace vfem
