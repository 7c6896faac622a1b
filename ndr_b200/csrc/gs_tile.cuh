// ndr_b200 — fine-level (matrix-free) coloured Gauss-Seidel through a shared-memory tile (3-D Q1, FP64, sm_100a).
//
// Reference semantics: MultigridSolver::m_smoothNode / visitNodesMulticolored, MG.hh:193-340 (one colour phase;
// colour = parity vector of the node index, same-colour nodes are 2 nodes apart).
//
// One thread per node of the active colour, as gsFineKernel (kernels.cuh), but the 27-node neighbourhoods are not
// gathered from global memory: on the stride-2 colour lattice every warp-wide gather touches 12-13 lines for 256 useful
// bytes and the sweep is bound by L1 wavefronts (profiles/r01s, r01v).  Here a CTA owns a box of TX x TY x TZ lattice
// sites; the (2TX+1) x (2TY+1) node rows it reads are contiguous in memory (axis 2 fastest, xyz interleaved), so warp 0
// fetches each row with ONE bulk asynchronous copy (cp.async.bulk -> UBLKCP, completion counted on an mbarrier) and the
// threads read their neighbourhoods from shared memory.  cp.async.bulk wants 16-byte aligned addresses and sizes while a
// row starts at 24 * node bytes: a copy starts at the 16-byte boundary at or below its first node and lands at a row
// offset of matching parity (`rowParity`), which the readers re-derive from the row's node index.  Out-of-grid
// neighbours are never read from the tile (select against the index test), so its uninitialised parts are harmless.
// Update rule, colour order and in-node substitution are the reference's; results are bit-identical to gsFineKernel.
#pragma once
#include "kernels.cuh"
#include "async_copy.cuh"

namespace vfem {

constexpr int kGsTileTZMax = 34;                  // lattice sites per CTA along axis 2 (at most)
template <int TX_, int TY_>
struct GsTileCfg {
    static constexpr int TX = TX_, TY = TY_;      // lattice sites per CTA along axes 0 and 1
    static constexpr int NT = (TX * TY * kGsTileTZMax + 31) / 32 * 32;   // one thread per lattice site
    static constexpr int BX = 2 * TX + 1, BY = 2 * TY + 1;
    static constexpr int PITCH = 3 * (2 * kGsTileTZMax + 1) + 7;         // doubles per staged row (even: rows stay 16-byte aligned)
    static constexpr size_t SMEM = size_t(BX) * BY * PITCH * sizeof(Real);
    static constexpr int MINB = (NT <= 96) ? 5 : (NT <= 160) ? 3 : 2;    // resident CTAs per SM the register budget is sized for
    static_assert(PITCH % 2 == 0, "row pitch must keep 16-byte alignment");
};

struct GsTilePlan { dim3 grid; int tz; };

// tiles of the colour lattice (restricted to the owned planes of axis 0)
template <class Cfg>
inline GsTilePlan planGsTile(const Grid<3> &g, int colour) {
    const int cx = colourPlaneCount<3>(g, colour);
    const int cy = (g.n[1] - 1 - ((colour >> 1) & 1)) / 2 + 1, cz = (g.n[2] - 1 - (colour & 1)) / 2 + 1;
    GsTilePlan p;
    const int ntz = (cz + kGsTileTZMax - 1) / kGsTileTZMax;
    p.tz = (cz + ntz - 1) / ntz;
    p.grid = dim3(unsigned(ntz), unsigned((cy + Cfg::TY - 1) / Cfg::TY), unsigned((cx + Cfg::TX - 1) / Cfg::TX));
    return p;
}

template <class Cfg>
__global__ void __launch_bounds__(Cfg::NT, Cfg::MINB)
gsFineTileKernel(Grid<3> g, K0Param<3> K, const Real *__restrict__ Ey, Real *u, const Real *__restrict__ b,
                 const uint8_t *__restrict__ dmask, int colour, int forward, int tz) {
    constexpr int TX = Cfg::TX, TY = Cfg::TY, BX = Cfg::BX, BY = Cfg::BY, RP = Cfg::PITCH;
    extern __shared__ __align__(16) unsigned char gsTileRaw[];
    Real *box = reinterpret_cast<Real *>(gsTileRaw);
    __shared__ __align__(8) unsigned long long bar;

    const int nx = g.n[0], ny = g.n[1], nz = g.n[2];
    const int py = (colour >> 1) & 1, pz = colour & 1;
    const int cntX = colourPlaneCount<3>(g, colour), cntY = (ny - 1 - py) / 2 + 1, cntZ = (nz - 1 - pz) / 2 + 1;
    const int lx0 = blockIdx.z * TX, ly0 = blockIdx.y * TY, lz0 = blockIdx.x * tz;
    const int tzc = min(tz, cntZ - lz0);                 // lattice sites of this tile along axis 2
    const int x0 = colourFirstPlane<3>(g, colour) + 2 * lx0, y0 = py + 2 * ly0, z0 = pz + 2 * lz0; // first node of the tile
    const int zLo = z0 - 1, BZ = 2 * tzc + 1;            // node window [zLo, zLo + BZ) of every staged row

    const long long baseOdd = (long long)(reinterpret_cast<size_t>(u) >> 3) & 1;
    if (threadIdx.x == 0) mbarInit(&bar, 32);
    __syncthreads();
    if (threadIdx.x < 32) { // warp 0 stages the rows; every lane arrives once with the bytes of the copies it issued
        unsigned bytes = 0;
        const int za = max(zLo, 0), zb = min(zLo + BZ, nz);
        const long long total = (long long)g.nn * 24;
        const long long ubase = (long long)reinterpret_cast<size_t>(u);
        for (int r = threadIdx.x; r < BX * BY; r += 32) {
            const int bx = r / BY, by = r - bx * BY;
            const int x = x0 - 1 + bx, y = y0 - 1 + by;
            if (x < 0 || x >= nx || y < 0 || y >= ny) continue;
            const long long row0 = ((long long)x * ny + y) * nz;
            const long long src = (row0 + za) * 24;
            const int delta = int((ubase + src) & 8);               // u itself may be only 8-byte aligned (a view into a larger field)
            const int orow = 2 + int((baseOdd + row0 + z0 + 1) & 1); // rowParity: see the readers below
            Real *dst = box + r * RP + orow + 3 * (za - zLo) - (delta >> 3);
            const int need = (zb - za) * 24 + delta;
            int size = (need + 15) & ~15;
            if (src - delta + size > total) { // the array ends inside the last 16-byte chunk: copy that double by hand
                size = need & ~15;
                const int first = size >> 3; // doubles covered by the bulk copy
                for (int i = first; i < (need >> 3); ++i) dst[i] = reinterpret_cast<const Real *>(reinterpret_cast<const char *>(u) + src - delta)[i];
            }
            if (size > 0) {
                bulkLoad(dst, reinterpret_cast<const char *>(u) + src - delta, unsigned(size), &bar);
                bytes += unsigned(size);
            }
        }
        mbarArriveExpectTx(&bar, bytes);
    }

    // this thread's lattice site
    const int t = threadIdx.x;
    const int lz = t % tzc, q = t / tzc, ly = q % TY, lx = q / TY;
    const int x = x0 + 2 * lx, y = y0 + 2 * ly, z = z0 + 2 * lz;
    const bool active = lx < TX && lx0 + lx < cntX && ly0 + ly < cntY;
    const int n = active ? (x * ny + y) * nz + z : 0;
    // loads that do not depend on the tile are issued before the wait
    Real bn[3] = {0.0, 0.0, 0.0};
    unsigned mask = 0;
    if (active) {
#pragma unroll
        for (int c = 0; c < 3; ++c) bn[c] = b[size_t(n) * 3 + c];
        mask = dmask[n];
    }
    Real Ee[8]; // moduli of the 8 incident elements (0 outside the grid)
#pragma unroll
    for (int s = 0; s < 8; ++s) {
        const int e0 = x - 1 + ((s >> 2) & 1), e1 = y - 1 + ((s >> 1) & 1), e2 = z - 1 + (s & 1);
        const bool ok = active && e0 >= 0 && e0 < g.E[0] && e1 >= 0 && e1 < g.E[1] && e2 >= 0 && e2 < g.E[2];
        Ee[s] = ok ? Ey[e0 * g.einc[0] + e1 * g.einc[1] + e2] : 0.0;
    }
    mbarWait(&bar, 0);
    if (!active) return;

    // row (x + dx, y + dy) was staged at offset 2 + ((row0 + z0 + 1) & 1): the parity that makes its first node's
    // shared-memory address congruent (mod 16) to its global address
    const int p0 = int((baseOdd + ((long long)x * ny + y) * nz + z0 + 1) & 1), fx = (ny * nz) & 1, fy = nz & 1;
    // neighbour-major accumulation: node k's three values are read from the tile once and folded into the running rows of
    // every incident element that contains it (8 elements x 3 rows of accumulators instead of the whole 27-node neighbourhood
    // in registers).  Per element the terms arrive in the same order as in fineNodeRowCompute, so the sums are bit-identical.
    Real re[8][3], uc[3] = {0.0, 0.0, 0.0};
#pragma unroll
    for (int s = 0; s < 8; ++s) re[s][0] = re[s][1] = re[s][2] = 0.0;
#pragma unroll
    for (int r = 0; r < 9; ++r) {
        const int dx = r / 3 - 1, dy = r % 3 - 1;
        // the three z-neighbours of row (dx, dy) are 9 contiguous doubles starting at an index of the row's parity (the same for
        // every thread of the CTA: x, y advance in steps of 2): four aligned 16-byte loads + one 8-byte load instead of nine
        // 8-byte loads that conflict 2-way at the 48-byte lane stride
        const int par = p0 ^ (dx != 0 ? fx : 0) ^ (dy != 0 ? fy : 0);
        const Real *row = box + ((2 * lx + 1 + dx) * BY + (2 * ly + 1 + dy)) * RP + 2 + 6 * lz; // even index: 16-byte aligned
        Real w[9];
        if (par) {
            w[0] = row[1];
#pragma unroll
            for (int i = 0; i < 4; ++i) { const double2 t2 = *reinterpret_cast<const double2 *>(row + 2 + 2 * i); w[1 + 2 * i] = t2.x; w[2 + 2 * i] = t2.y; }
        } else {
#pragma unroll
            for (int i = 0; i < 4; ++i) { const double2 t2 = *reinterpret_cast<const double2 *>(row + 2 * i); w[2 * i] = t2.x; w[2 * i + 1] = t2.y; }
            w[8] = row[8];
        }
        const bool okr = (x + dx >= 0) && (x + dx < nx) && (y + dy >= 0) && (y + dy < ny);
#pragma unroll
        for (int dzi = 0; dzi < 3; ++dzi) {
            const int dz = dzi - 1, k = r * 3 + dzi;
            const bool ok = okr && (z + dz >= 0) && (z + dz < nz);
            Real v[3];
#pragma unroll
            for (int c = 0; c < 3; ++c) v[c] = ok ? w[dzi * 3 + c] : 0.0;
            if (k == 13) { uc[0] = v[0]; uc[1] = v[1]; uc[2] = v[2]; }
#pragma unroll
            for (int s = 0; s < 8; ++s) { // incident element s holds this node at offset sd - 1 + qd = dd  =>  qd = dd + 1 - sd in {0, 1}
                const int s0 = (s >> 2) & 1, s1 = (s >> 1) & 1, s2 = s & 1;
                const int q0 = dx + 1 - s0, q1 = dy + 1 - s1, q2 = dz + 1 - s2;
                if (q0 < 0 || q0 > 1 || q1 < 0 || q1 > 1 || q2 < 0 || q2 > 1) continue;
                const int q = q0 * 4 + q1 * 2 + q2, loc = ((1 - s0) << 2) | ((1 - s1) << 1) | (1 - s2);
#pragma unroll
                for (int bb = 0; bb < 3; ++bb)
#pragma unroll
                    for (int c = 0; c < 3; ++c) re[s][c] = fma(K.k[(loc * 3 + c) * 24 + q * 3 + bb], v[bb], re[s][c]);
            }
        }
    }
    Real acc[3] = {0.0, 0.0, 0.0}, M[9];
#pragma unroll
    for (int i = 0; i < 9; ++i) M[i] = 0.0;
#pragma unroll
    for (int s = 0; s < 8; ++s) {
        const int loc = ((1 - ((s >> 2) & 1)) << 2) | ((1 - ((s >> 1) & 1)) << 1) | (1 - (s & 1));
#pragma unroll
        for (int c = 0; c < 3; ++c) acc[c] = fma(Ee[s], re[s][c], acc[c]);
#pragma unroll
        for (int c = 0; c < 3; ++c)
#pragma unroll
            for (int c2 = 0; c2 < 3; ++c2) M[c * 3 + c2] = fma(Ee[s], K.k[(loc * 3 + c) * 24 + loc * 3 + c2], M[c * 3 + c2]);
    }
    Real tt[3], du[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) tt[c] = bn[c] - acc[c];
    if (forward) gsNodeSubst<3, true>(tt, M, mask, du); else gsNodeSubst<3, false>(tt, M, mask, du);
#pragma unroll
    for (int c = 0; c < 3; ++c) u[size_t(n) * 3 + c] = uc[c] + du[c];
}

} // namespace vfem
