// ndr_b200 — geometric multigrid + MG-preconditioned CG, device resident.
// Mirrors MultigridSolver<1,...> (VoxelFEM/MultigridSolver.hh:11-759).
//
// Design (not a port): level 0 is matrix-free on the SIMP moduli; every coarser level stores its
// Galerkin operator P^T K P as an assembled 3^N-point block stencil in SoA layout.  The reference
// keeps a dense 24x24 Ke per coarse element (4.6 KB each) AND a block-CSC matrix per level; the
// assembled stencil is the same linear operator, so smoothing/apply/restriction results agree to
// roundoff.  The coarsest level is solved exactly through a block-tridiagonal LDL^T factorisation over
// its node planes (hand-written cooperative kernel, block_tri_factor.cuh, refreshed once per density update) whose
// solves are one thread-block-cluster launch each (coarse_solve.cuh), where the reference calls CHOLMOD.
#pragma once
#include "sim.h"
#include "kernels_vec.cuh"
#include "comm.h"
#include <functional>
#include <map>
#include <tuple>

namespace vfem {

template <int N> struct BlockTri;

struct MGBase {
    int dim = 0;
    virtual ~MGBase() {}
};

template <int N>
struct MG : MGBase {
    static constexpr int NPE = 1 << N, KS = N * NPE, NST = ipow3(N);
    struct Level {
        Grid<N> g;
        std::vector<uint8_t> dmask; // host mirror (coarse levels: MG.hh:43-84)
        DevBuf<uint8_t> d_dmask;
        DevBuf<Real> d_dval;        // zeros on coarse levels
        DevBuf<Real> x, b, r;       // work vectors m_x, m_b (MG.hh:755-756) + residual scratch
        DevBuf<Real> S;             // block stencil, levels >= 1
    };
    std::shared_ptr<Sim<N>> sim;
    std::vector<Level> lv;
    DevBuf<Real> d_C;               // 2^N coarsened reference stiffness matrices (MG.hh:644-648)
    DevBuf<int> galStart, galOff;   // admissible (fine row, fine column) pairs of the Galerkin product per stencil slot (kernels_mg.cuh)
    DevBuf<Real> galW;
    // coarsest level: block-tridiagonal LDL^T over the node planes of axis 0 (block_tri.h), refreshed once per density update
    std::unique_ptr<BlockTri<N>> bt;
    int coarseP = 0, coarsePlanes = 0;
    cudaStream_t haloStream = nullptr; // slab mode: the fine-level halo shift runs here, behind the interior K-apply
    cudaEvent_t haloFork = nullptr, haloJoin = nullptr;
    cudaStream_t haloStream2 = nullptr; // the second boundary view of a middle rank runs beside the first one
    cudaEvent_t haloShifted = nullptr, haloJoin2 = nullptr;
    unsigned long long builtVersion = 0;
    bool symmetricGS = true;                       // MG.hh:758
    bool buildBlockStiffnessMatrices = true;       // MG.hh:746 (kept for API parity; no effect on results)
    bool buildFinestBlockStiffnessMatrix = false;  // MG.hh:750
    cudaStream_t stream = 0;
    // The preconditioner application (one FMG / V-cycle pass: ~200 small launches) is launch-bound, so it is captured
    // once into a CUDA graph on a private stream and replayed per PCG iteration.
    struct PrecondKey {
        const Real *f = nullptr; size_t steps = 0, nu = 0; bool fmg = false, sym = true; int P = 0, planes = 0;
        unsigned long long k0Version = 0; int applyVariant = 0, gsVariant = 0; // K0 / Khat and the kernel choice are baked into the captured launches
        bool operator==(const PrecondKey &o) const {
            return f == o.f && steps == o.steps && nu == o.nu && fmg == o.fmg && sym == o.sym && P == o.P && planes == o.planes &&
                   k0Version == o.k0Version && applyVariant == o.applyVariant && gsVariant == o.gsVariant;
        }
    };
    bool useGraphs = true, fusedSweeps = true;
    int gsVariant = 1;              // fine-level GS, 3-D: 1 = shared-memory tile kernel (gs_tile.cuh; 2 x 4 site tiles below 60 k nodes per colour, 2 x 2 above), 2..4 = forced tile shape, 0 = per-node gather from global memory
    cudaStream_t captureStream = nullptr;
    cudaGraphExec_t pgExec = nullptr;
    PrecondKey pgKey;
    unsigned long long pgLaunches = 0;
    void precondition(const Real *f, size_t steps, size_t nu, bool fmg); // lv[0].x = M^{-1} f (MG.hh:476-479)

    // PCG state
    DevBuf<Real> p_x, p_r, p_d, p_Ad, p_s, p_b;
    DevBuf<Real> d_scal;            // [0] rMr [1] rMrOld [2] dAd [3] rr [4] bb [5..7] scratch
    DevBuf<Real> d_partials; DevBuf<unsigned int> d_ticket;
    Real *h_scal = nullptr;         // pinned mirror of d_scal
    ReduceWork rw;

    // ---- slab decomposition -------------------------------------------------------------------------
    // Rank r owns the coarsest-level element layers [c0, c1) of axis 0 and, on every level l, the node planes
    // [c0, c1) * 2^(L-l) (the last rank also the final plane).  Its local grids hold the coarsest layers [c0-1, c1)
    // refined 2^(L-l) times, so that local fine = 2 x local coarse on every level and the transfer / Galerkin kernels
    // run unchanged: 2^(L-l) ghost planes below the owned range (only the nearest one is ever read) and one ghost
    // plane above it.  Vectors are "consistent" when those two ghost planes hold the neighbours' values; stencil-type
    // operators need consistent inputs and leave ghost outputs undefined, exchange() restores consistency.
    // The coarsest level is assembled on every rank (all-reduce of the owned rows) and solved redundantly.
    std::shared_ptr<SlabComm> comm;  // null: single device
    int coarsePlaneOffset = 0;       // global coarsest-level plane index of local plane 0
    Grid<N> gCoarse;                 // global coarsest grid (== lv.back().g without a communicator)
    DevBuf<Real> Sg, bg, xg;         // global coarsest operator / right-hand side / solution (slab mode)
    DevBuf<uint8_t> dmaskG;
    enum { XUP = 1, XDOWN = 2 };
    bool slab() const { return comm && comm->world > 1; }
    void exchange(size_t l, Real *v, int dirs);                 // halo planes of a level-l vector
    // fine level: refresh the ghost planes of u AND apply the operator, the shift hidden behind the interior layers
    template <int MODE> void exchangeApplyFine(Real *u, const Real *b, Real *out);
    template <int MODE> void exchangeApplyFineBody(Real *u, const Real *b, Real *out);
    struct HaloGraphKey {
        const Real *u, *b, *out; int mode; unsigned long long k0Version; int applyVariant;
        bool operator<(const HaloGraphKey &o) const {
            return std::tie(u, b, out, mode, k0Version, applyVariant) < std::tie(o.u, o.b, o.out, o.mode, o.k0Version, o.applyVariant);
        }
    };
    std::map<HaloGraphKey, cudaGraphExec_t> haloGraphs;   // captured halo-shift + apply, per array triple
    unsigned long long haloGraphLaunches = 0;
    void exchangeHostMask(size_t l);
    void allreduce(Real *devScalars, size_t n) { if (slab()) comm->allreduceSum(devScalars, n, stream); }
    size_t ownedOffset(size_t l) const { return size_t(lv[l].g.own0) * lv[l].g.ninc[0] * N; }
    size_t ownedDofs(size_t l) const { return size_t(lv[l].g.own1 - lv[l].g.own0 + 1) * lv[l].g.ninc[0] * N; }

    MG(std::shared_ptr<Sim<N>> fine, size_t numCoarseningLevels, std::shared_ptr<SlabComm> comm = nullptr);
    ~MG();
    size_t numLevels() const { return lv.size(); }
    size_t levelDofs(size_t l) const { return size_t(lv.at(l).g.nn) * N; }
    void syncMasks();

    // operator hierarchy (MG.hh:415-441)
    void updateElementStiffnessMatrices(bool force = true);
    // device-pointer operators
    template <int MODE> void applyMode(size_t l, const Real *u, const Real *b, Real *out);
    void applyK(size_t l, const Real *u, Real *out);                                   // MG.hh:353-358
    void computeResidual(size_t l, const Real *u, const Real *b, Real *r);             // MG.hh:401-413
    void applyKMasked(size_t l, const Real *u, Real *out);
    void smoothColoured(size_t l, Real *u, const Real *b, bool forward);               // MG.hh:336-340
    void smoothSerial(size_t l, Real *u, const Real *b, bool forward);                 // MG.hh:269-282 (lexicographic; debug helper)
    void smooth(size_t l, Real *u, const Real *b, bool forward, size_t sweeps);        // `sweeps` x smoothColoured, fused on the assembled levels
    int sweepGridBlocks = 0;                             // co-resident blocks of gsStencilSweepKernel (cooperative launch)
    void restriction(size_t lf, const Real *vf, Real *bc);                             // MG.hh:145-180
    void interpolate(size_t lf, const Real *xc, Real *xf, bool accumulate);            // MG.hh:115-141
    void zeroOutDirichlet(size_t l, Real *u);                                          // MG.hh:364-378
    void enforceDirichlet(size_t l, Real *u, bool zero);                               // MG.hh:386-398
    void coarseSolve(const Real *b, Real *x);                                          // TPS.hh:834-865
    // cycles (MG.hh:446-553)
    void vcycle(size_t l, size_t nu, bool residualSystem);
    void fullMultigrid(size_t l, size_t nu, bool residualSystem);
    typedef std::function<void(size_t, const Real * /*device x*/)> SolveCb;
    void solve(const Real *u, const Real *f, size_t numSteps, size_t nu, bool stiffnessUpdated, bool zeroDirichlet,
               const SolveCb &cb, bool fmg, Real *out);
    // MG.hh:679-732.  u, b, out: device pointers.  cb(it, x_dev, r_dev, ||r||).  Returns iterations done.
    typedef std::function<void(size_t, const Real *, const Real *, Real)> PcgCb;
    size_t pcg(const Real *u, const Real *b, size_t maxIter, Real tol, const PcgCb &cb, size_t mgIterations, size_t nu,
               bool fmg, Real *out, Real *finalResidualNorm = nullptr);
    Real dot(const Real *a, const Real *b, size_t n); // deterministic device dot, synchronous
};

} // namespace vfem
