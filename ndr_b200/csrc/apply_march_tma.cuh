// ndr_b200 — element-marching K(rho) u, second generation: the algorithm of apply_march.cuh (same arithmetic, same
// summation order, bit-identical results) with the two latency chains of that kernel taken off the per-plane critical
// path (profiles/r01r_apply_full_march.txt: 18 % active warps, 3x latency exposure):
//
//  * INPUT: the node planes are not loaded by the threads any more.  Warp 0 streams them into a 4-stage shared-memory ring
//    with bulk asynchronous copies (cp.async.bulk -> UBLKCP, one copy per node row of the tile, completion counted on an
//    mbarrier per stage), three planes ahead of the march.  A thread reads its four corner nodes of the next plane from the
//    ring (conflict-free 64-bit LDS at a 24-byte lane stride), which also removes the forward z shuffles and the special
//    case of the lanes whose z+1 neighbour sits in another warp.
//  * EXCHANGE: the y / cross-warp z exchange of the backward butterflies is split-phase.  A thread publishes its part of
//    plane x, ARRIVES on an mbarrier and goes straight on to the forward half of layer x+1; it WAITS for the barrier only
//    after that arithmetic, finishes plane x and stores it.  The CTA-wide __syncthreads per plane (barrier 15 % + short
//    scoreboard 11 % of the stall samples) becomes a wait that has a whole layer of FP64 work to hide behind.
// Rows start at 24 * node bytes, bulk copies need 16-byte alignment: see async_copy.cuh and rowParity below.
#pragma once
#include "apply_march.cuh"
#include "async_copy.cuh"

namespace vfem {

constexpr int kMarchStages = 4; // planes in the ring: one being read, three in flight

template <int RY, int ZT>
struct MarchTmaCfg {
    static constexpr int NT = MarchCfg<RY, ZT>::NT;
    static constexpr int XS = MarchCfg<RY, ZT>::XS;
    static constexpr int ROWS = RY + 1, NODES = ZT + 1;
    static constexpr int PITCH = ((3 * NODES + 4) + 1) / 2 * 2;  // doubles per staged row (even: rows stay 16-byte aligned)
    static constexpr size_t STAGE = size_t(ROWS) * PITCH;        // doubles per stage
    static constexpr size_t SMEM = kMarchStages * STAGE * sizeof(Real);
};

template <int MODE, bool SPARSE, int RY, int ZT>
__global__ void __launch_bounds__(MarchCfg<RY, ZT>::NT, MarchCfg<RY, ZT>::MINB)
applyMarchTmaKernel(Grid<3> g, KhatParam kh, const Real *__restrict__ Ey, const Real *__restrict__ u, const Real *__restrict__ b,
                    const uint8_t *__restrict__ dmask, Real *__restrict__ out, int XC, int edgeSkip) {
    using Cfg = MarchTmaCfg<RY, ZT>;
    constexpr int XS = Cfg::XS, NT = Cfg::NT, PITCH = Cfg::PITCH;
    extern __shared__ __align__(16) unsigned char marchRaw[];
    Real *ring = reinterpret_cast<Real *>(marchRaw);      // [stage][row][PITCH]
    __shared__ Real sYx[2][RY][3][ZT];   // row ry -> slot ry+1 : y-stage part for node row y+1
    __shared__ Real sX[2][RY][XS][6];    // zh of a thread whose z+1 neighbour sits in the next warp
    __shared__ Real sYsp[2][RY][3];      // y exchange of the special last-z node column
    __shared__ __align__(8) unsigned long long full[kMarchStages], exch[2];

    const int nx = g.n[0], ny = g.n[1], nz = g.n[2], Ex = g.E[0];
    const int x0 = blockIdx.z * (XC - 1);
    const int x1 = min(x0 + XC, Ex);
    const int z0 = blockIdx.x * (ZT - 1), y0 = blockIdx.y * (RY - 1);
    unsigned flags, eOff, nodeBase, sIdx; // sIdx = ry * 3 * ZT + tx : element index into sYx[buf]
    int oA, oA1, oB, oB1;                  // ring offsets (doubles, parity shift excluded) of nodes (y,z) (y,z+1) (y+1,z) (y+1,z+1)
    int parA, parB;                        // row parity of rows y / y+1 at plane 0 (flips with the plane when ny*nz is odd)
    const long long baseOdd = (long long)(reinterpret_cast<size_t>(u) >> 3) & 1;
    {
        const int t = threadIdx.x, lane = t & 31;
        int ry = t / ZT;
        const int tx = t - ry * ZT;
        const bool live = ry < RY;
        if (!live) ry = RY - 1;
        const int y = y0 + ry, z = z0 + tx;
        const bool nextInWarp = (lane < 31) && (tx + 1 < ZT); // z+1 neighbour is lane+1 of this warp
        const bool prevInWarp = (lane > 0) && (tx > 0);
        const bool spz = live && (tx == ZT - 1) && (z + 1 == nz - 1);
        const bool top = live && (y0 + RY == ny - 1) && (ry == RY - 1); // this thread row also owns node row ny-1
        const bool ownY = (ry > 0 || blockIdx.y == 0), ownZ = (tx > 0 || blockIdx.x == 0);
        flags = (live ? MF_LIVE : 0u) | ((live && y < ny - 1 && z < nz - 1) ? MF_COL : 0u) |
                ((live && !nextInWarp && tx + 1 < ZT) ? MF_XSEND : 0u) | ((live && !prevInWarp && tx > 0) ? MF_XRECV : 0u) |
                ((live && ry > 0 && tx > 0 && (((t - ZT) & 31) == 0)) ? MF_XRECVB : 0u) | (spz ? MF_SPZ : 0u) |
                ((live && y < ny && z < nz && ownY && ownZ) ? MF_STORE : 0u) | ((live && ry + 1 < RY) ? MF_YSEND : 0u) |
                ((live && ry > 0) ? MF_YRECV : 0u) | (prevInWarp ? MF_PREV : 0u) | ((spz && y < ny && ownY) ? MF_SPSTORE : 0u) |
                (top ? MF_TOP : 0u) | ((top && z < nz && ownZ) ? MF_TOPSTORE : 0u) |
                ((blockIdx.z == 0 && !(edgeSkip & 1)) ? MF_CHUNK0 : 0u);
        // out-of-grid corner nodes are clamped into the staged (in-grid) part of the tile: they only ever feed element columns
        // whose modulus is forced to 0 (MF_COL clear), so any finite value will do
        const int rowsIn = min(RY + 1, ny - y0), nodesIn = min(ZT + 1, nz - z0); // staged rows / nodes per row
        const int ra = min(ry, rowsIn - 1), rb = min(ry + 1, rowsIn - 1), ca = min(tx, nodesIn - 1), cb = min(tx + 1, nodesIn - 1);
        oA = ra * PITCH + 2 + 3 * ca; oA1 = ra * PITCH + 2 + 3 * cb; oB = rb * PITCH + 2 + 3 * ca; oB1 = rb * PITCH + 2 + 3 * cb;
        parA = int((baseOdd + (long long)(y0 + ra) * nz + z0) & 1); parB = int((baseOdd + (long long)(y0 + rb) * nz + z0) & 1);
        const int yc = min(y, ny - 1), zc = min(z, nz - 1);
        eOff = unsigned(min(y, g.E[1] - 1) * g.E[2] + min(z, g.E[2] - 1));
        nodeBase = unsigned(yc * nz + zc);
        sIdx = unsigned(ry * 3 * ZT + tx);
    }
    asm volatile("" : "+r"(flags), "+r"(sIdx));
    const Real prevMask = (flags & MF_PREV) ? 1.0 : 0.0;
    const unsigned planeStride = unsigned(ny) * unsigned(nz) * 3u, planeNodes = unsigned(ny) * unsigned(nz);
    const unsigned eStride = unsigned(g.E[1]) * unsigned(g.E[2]);
    const int flipX = (ny * nz) & 1;

    if (threadIdx.x == 0) {
#pragma unroll
        for (int s = 0; s < kMarchStages; ++s) mbarInit(&full[s], RY + 1);
        mbarInit(&exch[0], NT); mbarInit(&exch[1], NT);
    }
    __syncthreads();

    // ---- producer: lanes 0..RY of warp 0, one node row of the tile each ----
    const int nPlanes = x1 - x0 + 1;                    // planes x0 .. x1 of this chunk (local index 0 .. nPlanes-1)
    auto issuePlane = [&](int pl) {                     // pl: local plane index; called by threads 0..RY
        const int r = threadIdx.x, y = y0 + r, s = pl & (kMarchStages - 1);
        unsigned bytes = 0;
        if (pl < nPlanes && y < ny) {
            const int nodes = min(ZT + 1, nz - z0);
            const long long node0 = ((long long)(x0 + pl) * ny + y) * nz + z0, src = node0 * 24;
            const long long ubase = (long long)reinterpret_cast<size_t>(u), total = (long long)g.nn * 24;
            const int delta = int((ubase + src) & 8);
            Real *dst = ring + size_t(s) * Cfg::STAGE + r * PITCH + 2 + int((baseOdd + node0) & 1) - (delta >> 3);
            const int need = nodes * 24 + delta;
            int size = (need + 15) & ~15;
            if (src - delta + size > total) { // the field ends inside the last 16-byte chunk: that double is copied by hand
                size = need & ~15;
                for (int i = size >> 3; i < (need >> 3); ++i) dst[i] = reinterpret_cast<const Real *>(reinterpret_cast<const char *>(u) + src - delta)[i];
            }
            if (size > 0) { bulkLoad(dst, reinterpret_cast<const char *>(u) + src - delta, unsigned(size), &full[s]); bytes = unsigned(size); }
        }
        mbarArriveExpectTx(&full[s], bytes);
    };
    const bool producer = threadIdx.x <= RY;
    if (producer)
        for (int pl = 0; pl < kMarchStages; ++pl) issuePlane(pl);

    // faces of local plane pl from the ring
    auto faces = [&](int pl, Real (&F)[4][3]) {
        mbarWait(&full[pl & (kMarchStages - 1)], unsigned(pl / kMarchStages) & 1u);
        const Real *st = ring + size_t(pl & (kMarchStages - 1)) * Cfg::STAGE;
        const int fl = flipX & (x0 + pl);
        const Real *pA = st + (parA ^ fl), *pB = st + (parB ^ fl);
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const Real a = pA[oA + c], an = pA[oA1 + c], bq = pB[oB + c], bn = pB[oB1 + c];
            const Real zs0 = a + an, zd0 = a - an, zs1 = bq + bn, zd1 = bq - bn;
            F[0][c] = zs0 + zs1; F[2][c] = zs0 - zs1; F[1][c] = zd0 + zd1; F[3][c] = zd0 - zd1;
        }
    };

    Real FA[4][3], FB[4][3], Gprev[4][3], Enext;
#pragma unroll
    for (int k = 0; k < 4; ++k)
#pragma unroll
        for (int c = 0; c < 3; ++c) Gprev[k][c] = 0.0;
    const Real *eP = Ey + size_t(x0) * eStride + eOff;
    Enext = *eP;
    faces(0, FA);

    // what a plane needs after the exchange to be finished and stored (carried across the next layer's forward half)
    Real res[3], yhh[3], zh[2][3];
#pragma unroll
    for (int c = 0; c < 3; ++c) { res[c] = 0.0; yhh[c] = 0.0; zh[0][c] = 0.0; zh[1][c] = 0.0; }

    // finish local plane `it` (global plane x0 + it): wait for everybody's exchange data, add it, store
    auto finish = [&](int it) {
        const int buf = it & 1, x = x0 + it;
        mbarWait(&exch[buf], unsigned(it >> 1) & 1u);
        const Real *sYb = &sYx[buf][0][0][0] + sIdx;
        const Real *sXb = &sX[buf][0][0][0];
        if (flags & MF_YRECV) {
#pragma unroll
            for (int c = 0; c < 3; ++c) res[c] += sYb[c * ZT];
        }
        const Real *sXr = sXb;
        if (flags & (MF_XRECV | MF_XRECVB)) {
            const int ry = sIdx / (3 * ZT), tx = sIdx - ry * 3 * ZT;
            sXr = sXb + (ry * XS + (tx >> 5)) * 6;
        }
        if (flags & MF_XRECV) {
#pragma unroll
            for (int c = 0; c < 3; ++c) res[c] += sXr[c] + sXr[3 + c];
        }
        if (flags & MF_XRECVB) {
#pragma unroll
            for (int c = 0; c < 3; ++c) res[c] += sXr[c - XS * 6] - sXr[3 + c - XS * 6];
        }
        if (it > 0 || (flags & MF_CHUNK0)) {
            const unsigned n = unsigned(x) * planeNodes + nodeBase;
            if (flags & MF_STORE) marchStore<MODE>(n, res, b, dmask, out);
            if (flags & (MF_SPSTORE | MF_TOP)) { // rare paths: last node column / last node row of the grid
                if (flags & MF_SPSTORE) {
                    Real v[3];
#pragma unroll
                    for (int c = 0; c < 3; ++c) v[c] = zh[0][c] + zh[1][c] + ((flags & MF_YRECV) ? sYsp[buf][sIdx / (3 * ZT)][c] : 0.0);
                    marchStore<MODE>(n + 1, v, b, dmask, out);
                }
                if (flags & MF_TOPSTORE) {
                    Real v[3];
#pragma unroll
                    for (int c = 0; c < 3; ++c) v[c] = yhh[c];
                    if (flags & MF_XRECV) {
#pragma unroll
                        for (int c = 0; c < 3; ++c) v[c] += sXr[c] - sXr[3 + c];
                    }
                    marchStore<MODE>(n + nz, v, b, dmask, out);
                }
                if ((flags & (MF_TOP | MF_SPZ)) == (MF_TOP | MF_SPZ)) {
                    Real v[3];
#pragma unroll
                    for (int c = 0; c < 3; ++c) v[c] = zh[0][c] - zh[1][c];
                    marchStore<MODE>(n + nz + 1, v, b, dmask, out);
                }
            }
        }
    };

    // LAYER: the step advances through element layer x0 + it (always, except the closing step that only flushes plane Ex)
    auto step = [&](auto hasLayer, const int it, Real (&Fold)[4][3], Real (&Fnew)[4][3]) {
        constexpr bool LAYER = decltype(hasLayer)::value;
        const int x = x0 + it, buf = it & 1;
        Real Gtot[4][3];
        if (LAYER) {
            faces(it + 1, Fnew);
            const Real Ee = (flags & MF_COL) ? Enext : 0.0;
            if (x + 2 < nx) { eP += eStride; Enext = *eP; } // the next layer's modulus, one step ahead
            marchLayer<SPARSE>(kh, Fold, Fnew, Ee, Gprev, Gtot);
        } else {
#pragma unroll
            for (int k = 0; k < 4; ++k)
#pragma unroll
                for (int c = 0; c < 3; ++c) Gtot[k][c] = Gprev[k][c];
        }
        // the previous plane's exchange has had this layer's forward half to complete
        if (it > 0) finish(it - 1);
        // its stage of the ring (plane it, read in the previous step) is free now: refill it three planes ahead
        if (producer && it > 0 && it - 1 + kMarchStages < nPlanes) issuePlane(it - 1 + kMarchStages);
        // ---- backward, plane x.  z stage (index k = my*2 + mz), then the y stage's send half ----
        Real Ye[2][3];
#pragma unroll
        for (int my = 0; my < 2; ++my)
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                const Real zl = Gtot[my * 2][c] + Gtot[my * 2 + 1][c];
                zh[my][c] = Gtot[my * 2][c] - Gtot[my * 2 + 1][c];
                const Real r = __shfl_up_sync(0xffffffffu, zh[my][c], 1);
                Ye[my][c] = fma(r, prevMask, zl);
            }
        Real *sYb = &sYx[buf][0][0][0] + sIdx;
        Real *sXb = &sX[buf][0][0][0];
        if (flags & MF_XSEND) {
            const int ry = sIdx / (3 * ZT), tx = sIdx - ry * 3 * ZT;
#pragma unroll
            for (int my = 0; my < 2; ++my)
#pragma unroll
                for (int c = 0; c < 3; ++c) sXb[(ry * XS + ((tx + 1) >> 5)) * 6 + my * 3 + c] = zh[my][c];
        }
#pragma unroll
        for (int c = 0; c < 3; ++c) { res[c] = Ye[0][c] + Ye[1][c]; yhh[c] = Ye[0][c] - Ye[1][c]; }
        if (flags & MF_YSEND) {
#pragma unroll
            for (int c = 0; c < 3; ++c) sYb[(3 + c) * ZT] = yhh[c];
        }
        if ((flags & (MF_SPZ | MF_YSEND)) == (MF_SPZ | MF_YSEND)) {
#pragma unroll
            for (int c = 0; c < 3; ++c) sYsp[buf][sIdx / (3 * ZT) + 1][c] = zh[0][c] - zh[1][c];
        }
        mbarArrive(&exch[buf]);
    };

    const int nIter = (x1 - x0) + ((x1 == Ex && !(edgeSkip & 2)) ? 1 : 0); // edgeSkip: the grid is a view whose first / last plane is partial
    const int nLayers = x1 - x0;
    int it = 0;
    for (; it + 1 < nLayers; it += 2) { // two steps per trip so the face registers ping-pong without copies
        step(std::true_type{}, it, FA, FB);
        step(std::true_type{}, it + 1, FB, FA);
    }
    if (it < nLayers) { // odd layer count
        step(std::true_type{}, it, FA, FB);
        if (nIter > nLayers) step(std::false_type{}, it + 1, FB, FA);
    } else if (nIter > nLayers) step(std::false_type{}, it, FA, FB);
    finish(nIter - 1);
}

template <int MODE>
inline void launchApplyMarchTma(const Grid<3> &g, const KhatParam &kh, bool sparse, const MarchPlan &pl, const Real *Ey, const Real *u,
                                const Real *b, const uint8_t *dmask, Real *out, cudaStream_t stream, int edgeSkip = 0) {
#define VFEM_MARCH_TMA_CASE(S, RY, ZT)                                                                                         \
    case S: {                                                                                                                  \
        constexpr size_t smem = MarchTmaCfg<RY, ZT>::SMEM;                                                                     \
        if (sparse) {                                                                                                          \
            static bool once = (cudaFuncSetAttribute(applyMarchTmaKernel<MODE, true, RY, ZT>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)), true); \
            (void)once;                                                                                                        \
            applyMarchTmaKernel<MODE, true, RY, ZT><<<pl.grid, MarchCfg<RY, ZT>::NT, smem, stream>>>(g, kh, Ey, u, b, dmask, out, pl.XC, edgeSkip); \
        } else {                                                                                                               \
            static bool once = (cudaFuncSetAttribute(applyMarchTmaKernel<MODE, false, RY, ZT>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)), true); \
            (void)once;                                                                                                        \
            applyMarchTmaKernel<MODE, false, RY, ZT><<<pl.grid, MarchCfg<RY, ZT>::NT, smem, stream>>>(g, kh, Ey, u, b, dmask, out, pl.XC, edgeSkip); \
        }                                                                                                                      \
    } break;
    switch (pl.cfg) {
        VFEM_MARCH_TMA_CASE(0, 8, 64)
        VFEM_MARCH_TMA_CASE(1, 16, 32)
        VFEM_MARCH_TMA_CASE(2, 15, 33)
        VFEM_MARCH_TMA_CASE(3, 7, 65)
        VFEM_MARCH_TMA_CASE(4, 6, 64)
        VFEM_MARCH_TMA_CASE(5, 4, 64)
        VFEM_MARCH_TMA_CASE(6, 8, 32)
        VFEM_MARCH_TMA_CASE(7, 8, 33)
        VFEM_MARCH_TMA_CASE(8, 5, 64)
        VFEM_MARCH_TMA_CASE(9, 4, 33)
        VFEM_MARCH_TMA_CASE(10, 6, 65)
        VFEM_MARCH_TMA_CASE(11, 5, 33)
        VFEM_MARCH_TMA_CASE(12, 11, 33)
    }
#undef VFEM_MARCH_TMA_CASE
}

} // namespace vfem
