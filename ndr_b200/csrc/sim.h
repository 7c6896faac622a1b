// ndr_b200 — host side of the regular-grid elasticity simulator.
// Mirrors the public surface of the reference's TensorProductSimulator<1,...> for the solve path
// (VoxelFEM/TensorProductSimulator.hh:219-1419): grid + indexing, material -> K0, boundary
// conditions, SIMP moduli, K u, compliance gradient.  All field data lives in HBM.
#pragma once
#include "common.h"
#include "apply_march.cuh"
#include <cmath>
#include <map>
#include <memory>

namespace vfem {

template <int N> struct DirectSolver; // sim.cu: assembled fine operator + block-tridiagonal factor (TPS::solve)

struct SimBase {
    int dim = 0;
    cudaStream_t stream = 0;
    virtual ~SimBase() {}
};

template <int N>
struct Sim : SimBase {
    static constexpr int NPE = 1 << N, KS = N * NPE, NST = ipow3(N);

    Grid<N> g;
    Real mn[N], mx[N], h[N], spacing[N];
    Real lambda = 0.0, mu = 0.5;       // default ETensor(1, 0), TPS.hh:1379
    K0Param<N> K0;                     // reference element stiffness for unit modulus
    Real E0 = 1.0, Emin = 1e-9, gamma = 3.0; // TPS.hh:1392-1394
    // 3-D only: Walsh-Hadamard block form of K0 for the marching apply (apply_march.cuh)
    KhatParam Khat;
    int khatStructure = 0;             // 0: not block diagonal -> gather kernel; 1: dense blocks; 2: isotropic pattern
    MarchPlan marchPlan;
    int applyVariant = 0;              // 0 = auto (ring-fed marching kernel when K0 allows it), 1 = per-node gather kernel, 2 = marching kernel with dense 3x3 blocks, 4 = first-generation marching kernel
    bool useMarch() const { return N == 3 && khatStructure > 0 && applyVariant != 1; }

    // boundary data (host mirrors; device copies refreshed lazily)
    std::vector<uint8_t> dmask;        // per node, bit c = component c fixed
    std::vector<Real> dval;            // nn*N
    std::vector<Real> force;           // nn*N
    std::vector<uint8_t> hasForce;     // nn
    bool bcDirty = true;
    unsigned long long bcVersion = 1;  // bumped whenever the Dirichlet mask changes (the direct factorisation depends on it)
    DevBuf<uint8_t> d_dmask;
    DevBuf<Real> d_dval;

    // densities and moduli (device is authoritative)
    DevBuf<Real> d_rho, d_Ey;
    bool EyDirty = true;
    unsigned long long densityVersion = 1; // bumped whenever K(rho) changes
    unsigned long long k0Version = 1;      // bumped whenever K0 / Khat change (they are baked into captured graphs as kernel parameters)

    // Slab of a larger grid: this simulator holds the element layers [layerLo, layerLo + g.E[0]) of a grid
    // with globalE0 layers along axis 0.  mn/mx/spacing/h describe the GLOBAL domain, so node positions (and with them
    // the boundary-condition regions) and K0 are bit-identical to the undecomposed grid's.
    int layerLo = 0, globalE0 = 0;
    int computeLo = 0;                        // first local element layer K u is evaluated on (set by the slab solver)
    std::map<int, MarchPlan> viewPlans;       // march plans of sub-grid views, keyed by their layer count
    // pipelined host-array apply (applyKHost)
    static constexpr int kHostChunks = 16; // upper bound; hostChunks are used
    int hostChunks = 4;
    DevBuf<Real> hostIn, hostOut;
    cudaStream_t h2dStream = nullptr, d2hStream = nullptr, hostComputeStream = nullptr;
    cudaEvent_t evIn[kHostChunks] = {}, evKernel[kHostChunks] = {}, evOut[kHostChunks] = {};

    // elems: local element counts; globalE0 = 0 means "not a slab" (the local grid is the whole grid)
    Sim(const Real *bbmin, const Real *bbmax, const size_t *elems, size_t globalE0 = 0, size_t layerLo = 0);
    ~Sim() override;

    // TPS.hh:274-279
    void nodePosition(size_t ni, Real *p) const {
        for (int d = N - 1; d >= 0; --d) {
            size_t i = ni % size_t(g.n[d]); ni /= size_t(g.n[d]);
            if (d == 0) i += size_t(layerLo);
            p[d] = mn[d] + Real(i) * spacing[d];
        }
    }
    size_t numNodes() const { return size_t(g.nn); }
    size_t numElements() const { return size_t(g.ne); }
    size_t numDofs() const { return size_t(g.nn) * N; }

    void setIsotropic(Real young, Real nu);   // ElasticityTensor.hh:100-115
    void updateK0();                          // TPS.hh:127-140,1358-1366
    void touchStiffness() { EyDirty = true; ++densityVersion; }

    // boundary conditions (TPS.hh:358-409, Node::setDirichlet TPS.hh:49-63)
    // divisor: number of nodes the force is split over; 0 = the nodes of this grid inside the region (TPS.hh:385-388)
    void addForceRegion(const Real *lo, const Real *hi, const Real *value, size_t divisor = 0);
    size_t countRegionNodes(const Real *lo, const Real *hi, size_t planeLo, size_t planeHi) const; // local planes [planeLo, planeHi]
    void addDirichletRegion(const Real *lo, const Real *hi, const uint8_t *compMask, const Real *value);
    void setDirichlet(size_t ni, const uint8_t *compMask, const Real *value);
    void buildLoadVectorHost(Real *f) const;  // TPS.hh:893-901
    void syncBC();

    // densities
    void setDensities(const Real *rho, MemSpace ms);
    void setUniformDensities(Real v);         // TPS.hh:456-461
    void setElementDensity(size_t ei, Real v);// TPS.hh:463-468
    Real elementDensity(size_t ei) const;
    void getDensities(Real *out, MemSpace ms) const;
    const Real *moduli();                     // refreshes E(rho) when stale

    // operators on device pointers
    void applyK(const Real *u, Real *out);    // TPS.hh:905-952
    // MODE 0: K u; 1: mask(b - K u); 2: mask(K u)   (fine level; MG.hh:401-413, 719-720)
    template <int MODE> void applyFine(const Real *u, const Real *b, const uint8_t *mask, Real *out);
    template <int MODE> void applyFineRange(const Real *u, const Real *b, const uint8_t *mask, Real *out, int layerLo, int layerHi, cudaStream_t st,
                                               int edgeSkip = 0); // edgeSkip bit 0 / 1: do not store the view's first / last node plane
    void applyKHost(const Real *hu, Real *hout);  // host arrays in / out, copies pipelined with the kernel
    void complianceGradient(const Real *u, Real *g); // TPS.hh:730-751
    // u = K(rho)^{-1} f with the Dirichlet rows / columns removed (TPS::solve, TPS.hh:834-865); device pointers.
    // The numeric factorisation is redone only when the densities, the material or the Dirichlet mask changed
    // (m_numericFactorizationUpToDate).  refinements: steps of iterative refinement with the matrix-free residual.
    void solveDirect(const Real *f, Real *u, int refinements = 2);
    std::unique_ptr<DirectSolver<N>> direct;
};

} // namespace vfem
