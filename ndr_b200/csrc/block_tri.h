// ndr_b200 — exact solver for a grid operator given as an assembled 3^N-point block stencil: block-tridiagonal LDL^T over
// the node planes of axis 0 (block_tri_factor.cuh) + the two-ended chain solve (coarse_solve.cuh).
// Stands where the reference calls CHOLMOD (TensorProductSimulator::solve, TPS.hh:834-865): MG's coarsest level and the
// direct solve behind ComplianceObjective.  Include from .cu files only.
#pragma once
#include "block_tri_factor.cuh"

namespace vfem {

template <int N>
struct BlockTri {
    static constexpr int NST = ipow3(N);
    DevBuf<Real> D, Lo, Trm, Tcm, z;   // S_p^{-1} (in place of A_pp), A_{p+1,p}, T blocks (row- / column-major), chain scratch
    DevBuf<int> info;
    int P = 0, planes = 0, mid = 0;
    int chainCluster = 0, chainGridBlocks = 0, factorBlocksMax = 0;
    static constexpr size_t kMaxBytes = size_t(48) << 30;

    size_t bytesFor(int P_, int planes_) const { return size_t(P_) * P_ * size_t(4 * planes_) * sizeof(Real); }

    // S: stencil of grid g (node-major, stencilIndex), dmask: Dirichlet bits per node (identity rows / columns).
    // Enqueues the whole factorisation on `stream` (2 launches) and returns; check() reports a failed pivot.
    void factor(const Grid<N> &g, const Real *S, const uint8_t *dmask, cudaStream_t stream) {
        P = g.ninc[0] * N; planes = g.n[0]; mid = planes / 2;
        const size_t PP = size_t(P) * P;
        if (bytesFor(P, planes) > kMaxBytes)
            throw std::runtime_error("exact solve: " + std::to_string(size_t(g.nn) * N) + " dofs in planes of " + std::to_string(P) +
                                     " need " + std::to_string(bytesFor(P, planes) >> 20) + " MB of block factors (use more coarsening levels)");
        const size_t off = size_t(planes > 1 ? planes - 1 : 1);
        D.alloc(PP * planes); Lo.alloc(PP * off); Trm.alloc(PP * off); Tcm.alloc(PP * off); z.alloc(size_t(planes) * P);
        info.alloc(1);
        D.zero(stream); Lo.zero(stream); info.zero(stream);
        const long long tot = (long long)g.nn * NST;
        assembleBlockTriKernel<N><<<unsigned((tot + 127) / 128), 128, 0, stream>>>(g, S, dmask, D.p, Lo.p, P);
        VFEM_CHECK_LAUNCH(); countLaunch();
        if (factorBlocksMax == 0) {
            int dev = 0, sms = 148, per = 1;
            VFEM_CUDA(cudaGetDevice(&dev));
            VFEM_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
            VFEM_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per, blockTriFactorKernel<N>, kFactorThreads, 0));
            if (per < 1) throw std::runtime_error("blockTriFactorKernel does not fit on this device");
            factorBlocksMax = sms; // one CTA per SM: the kernel is bound by its grid barriers, not by occupancy
        }
        const int nbk = (P + kFB - 1) / kFB;
        int blocks = std::min(factorBlocksMax, std::max(16, nbk * nbk));
        if (const char *v = getenv("VFEM_FACTOR_BLOCKS")) blocks = std::max(1, std::min(factorBlocksMax, atoi(v)));
        FactorParams<N> fp;
        fp.P = P; fp.planes = planes; fp.mid = mid;
        for (int d = 1; d < N; ++d) fp.pn[d - 1] = g.n[d];
        Real *Dp = D.p, *trm = Trm.p, *tcm = Tcm.p; const Real *lo = Lo.p; int *inf = info.p;
        void *args[] = {&fp, &Dp, &lo, &trm, &tcm, &inf};
        VFEM_CUDA(cudaLaunchCooperativeKernel((const void *)blockTriFactorKernel<N>, dim3(blocks), dim3(kFactorThreads), args, 0, stream));
        countLaunch();
    }

    // synchronises `stream`; throws when a pivot was not positive (the reference's CHOLMOD reports "not positive definite")
    void check(cudaStream_t stream) {
        int h = 0;
        VFEM_CUDA(cudaMemcpyAsync(&h, info.p, sizeof(int), cudaMemcpyDeviceToHost, stream));
        VFEM_CUDA(cudaStreamSynchronize(stream));
        if (h != 0) throw std::runtime_error("Cholesky factorization failed: matrix not positive definite (plane " + std::to_string(h - 1) + ")");
    }

    // x = A^{-1} b on the free dofs, 0 on the constrained ones; one launch
    void solve(const Real *b, Real *x, const uint8_t *dmask, cudaStream_t stream) {
        const bool small = P <= 256 && 2 * P <= kChainSlots * kChainCluster * (kChainThreads / 32);
        ChainIdx ci{planes, mid};
        if (small) { // one thread-block cluster, plane vectors in distributed shared memory
            const size_t smem = size_t(4) * P * sizeof(Real);
            cudaLaunchConfig_t cfg = {};
            cudaLaunchAttribute attr[1];
            attr[0].id = cudaLaunchAttributeClusterDimension;
            attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
            cfg.blockDim = dim3(kChainThreads); cfg.dynamicSmemBytes = smem; cfg.stream = stream;
            cfg.attrs = attr; cfg.numAttrs = 1;
            if (chainCluster == 0) { // 16 CTAs (non-portable size) halve the matrix rows per SM and step; 8 where that does not fit
                chainCluster = kChainCluster;
                const char *v = getenv("VFEM_CHAIN_CLUSTER");
                const int want = v ? atoi(v) : 16;
                if (want > kChainCluster && want <= 16 &&
                    cudaFuncSetAttribute(blockTriChainKernel<N, 2>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) == cudaSuccess &&
                    cudaFuncSetAttribute(blockTriChainKernel<N, kChainSlots>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) == cudaSuccess) {
                    cfg.gridDim = dim3(want); attr[0].val.clusterDim.x = want;
                    int n = 0;
                    if (cudaOccupancyMaxActiveClusters(&n, blockTriChainKernel<N, 2>, &cfg) == cudaSuccess && n >= 1) chainCluster = want;
                } else if (want >= 1 && want < kChainCluster && 2 * P <= kChainSlots * want * (kChainThreads / 32)) chainCluster = want;
                (void)cudaGetLastError();
            }
            cfg.gridDim = dim3(chainCluster); attr[0].val.clusterDim.x = chainCluster;
            if (2 * P <= 2 * chainCluster * (kChainThreads / 32)) // a warp per 2 rows: the instantiation without register spills
                VFEM_CUDA(cudaLaunchKernelEx(&cfg, blockTriChainKernel<N, 2>, P, ci, (const Real *)Trm.p, (const Real *)Tcm.p, (const Real *)D.p, b, dmask, x, z.p));
            else
                VFEM_CUDA(cudaLaunchKernelEx(&cfg, blockTriChainKernel<N, kChainSlots>, P, ci, (const Real *)Trm.p, (const Real *)Tcm.p, (const Real *)D.p, b, dmask, x, z.p));
        } else { // large planes: the whole device, one grid barrier per chain step
            if (chainGridBlocks == 0) {
                int dev = 0, sms = 148;
                VFEM_CUDA(cudaGetDevice(&dev));
                VFEM_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
                chainGridBlocks = sms;
            }
            int Pv = P;
            const Real *trm = Trm.p, *tcm = Tcm.p, *sinv = D.p; Real *zp = z.p;
            void *args[] = {&Pv, &ci, &trm, &tcm, &sinv, &b, &dmask, &x, &zp};
            const size_t smem = size_t(2) * P * sizeof(Real);
            if (smem > 48 * 1024) VFEM_CUDA(cudaFuncSetAttribute(blockTriChainGridKernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
            // a forward / backward step has 2 P rows (two chains): enough warps for one row each where the CTA size allows it
            int threads = kChainGridThreads;
            while (threads < kChainGridMaxThreads && chainGridBlocks * (threads / 32) < 2 * P) threads += 128;
            if (const char *v = getenv("VFEM_CHAIN_GRID_THREADS")) threads = std::max(64, std::min(kChainGridMaxThreads, atoi(v) / 32 * 32));
            VFEM_CUDA(cudaLaunchCooperativeKernel((const void *)blockTriChainGridKernel<N>, dim3(chainGridBlocks), dim3(threads), args, smem, stream));
        }
        countLaunch();
    }
};

} // namespace vfem
