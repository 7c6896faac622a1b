// ndr_b200 — element-marching Walsh-Hadamard form of the fine-level matrix-free K(rho) u (3-D Q1, FP64, sm_100a).
//
// Reference semantics: TensorProductSimulator::applyK, matrix-free branch (TPS.hh:905-952):
//     out = sum_e E(rho_e) * scatter(K0 * gather(u)).
// On B200 the textbook form (576 DFMA per node) is bound by the FP64 pipe (64 DFMA/clk/SM) at ~27 % of
// the HBM roofline, so this kernel reduces the arithmetic instead of the traffic:
//
//  * K0 of a box element with an axis-aligned isotropic C commutes with the reflections of the element,
//    so in the per-axis sum/difference (Walsh-Hadamard) basis  uh[m][c] = sum_j (-1)^(m.j) u[j][c]
//    it splits into 8 blocks of 3x3: block g couples (m = g^e_c, c), c = 0..2  (e_0=4, e_1=2, e_2=1).
//    K0 = W^T Khat W with Khat = W K0 W^T / 64.  Rigid-body modes and a further symmetry leave only 45
//    non-zero block entries (`khatEntryNonZero`); the host verifies both facts numerically for the K0 at
//    hand and otherwise routes to the dense gather kernel (kernels.cuh).
//  * the butterflies are per axis, so their first stages are shared between neighbouring elements: one
//    thread owns one (y,z) element column and marches along axis 0.  Per step it loads its node of the
//    next plane (+ the row above), forms the plane's face quantities F[my][mz][c], combines them with the
//    previous plane's (kept in registers) into uh, applies the blocks, and runs the transposed butterflies
//    back: along x in registers (the contribution to the next plane, scaled by E_e, is carried in Gprev),
//    along z with warp shuffles, along y through shared memory (one barrier per step, double buffered).
//  * tiles of RY rows x ZT lanes overlap their low-side neighbour by one element column (the ghost row /
//    lane recomputes that column); chunks of XC layers overlap by one layer along the march.
// ~160 FP64 instructions per element instead of 576; u, E read once, K u written once; no atomics, fixed
// summation order => bit-reproducible.  tools/march_model.py is a numpy model of exactly this algorithm.
#pragma once
#include "common.h"
#include <algorithm>
#include <cmath>

namespace vfem {

struct KhatParam { Real b[8][3][3]; }; // b[g][c][c2] = Khat[(g^e_c, c), (g^e_c2, c2)]

__host__ __device__ constexpr int khatAxisBit(int c) { return 4 >> c; }
__host__ __device__ constexpr int khatPopc(int g) { return (g & 1) + ((g >> 1) & 1) + ((g >> 2) & 1); }
// sparsity pattern of the blocks for an isotropic box element
__host__ __device__ constexpr bool khatEntryNonZero(int g, int c, int c2) {
    const int pc = khatPopc(g);
    if (pc == 0 || pc == 3) return true;
    if (pc == 1) { // the component whose mode index is 0 (rigid translation) drops out
        return (g != khatAxisBit(c)) && (g != khatAxisBit(c2));
    }
    // pc == 2: the component of the missing axis decouples
    const int k = (g == 3) ? 0 : (g == 5) ? 1 : 2;
    return (c == k && c2 == k) || (c != k && c2 != k);
}

template <int RY, int ZT>
struct MarchCfg {
    static constexpr int NT = ((RY * ZT + 31) / 32) * 32;
    static constexpr int XS = ZT / 32 + 2; // cross-warp slots per row
};

// MODE 0: out = K u;  MODE 1: out = mask(b - K u);  MODE 2: out = mask(K u)
template <int MODE>
__device__ __forceinline__ void marchStore(const Grid<3> &g, int x, int y, int z, const Real *v, const Real *__restrict__ b,
                                           const uint8_t *__restrict__ dmask, Real *__restrict__ out) {
    const size_t n = (size_t(x) * g.n[1] + y) * g.n[2] + z;
    unsigned m = 0;
    if (MODE != 0) m = dmask[n];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        Real r = v[c];
        if (MODE == 1) r = b[n * 3 + c] - r;
        if (MODE != 0 && ((m >> c) & 1u)) r = 0.0;
        out[n * 3 + c] = r;
    }
}

template <int MODE, bool SPARSE, int RY, int ZT>
__global__ void __launch_bounds__(MarchCfg<RY, ZT>::NT, 1)
applyMarchKernel(Grid<3> g, KhatParam kh, const Real *__restrict__ Ey, const Real *__restrict__ u, const Real *__restrict__ b,
                 const uint8_t *__restrict__ dmask, Real *__restrict__ out, int XC) {
    constexpr int XS = MarchCfg<RY, ZT>::XS;
    __shared__ Real sYx[2][RY][3][ZT];   // row ry -> slot ry+1 : y-stage part for node row y+1
    __shared__ Real sX[2][RY][XS][6];    // zh of a lane-31 thread for the lane-0 thread of the next warp
    __shared__ Real sYsp[2][RY][3];      // y exchange of the special last-z node column

    const int t = threadIdx.x;
    const int ry = t / ZT, tx = t - ry * ZT, lane = t & 31;
    const bool live = ry < RY;
    const int nx = g.n[0], ny = g.n[1], nz = g.n[2];
    const int Ex = g.E[0];
    const int z0 = blockIdx.x * (ZT - 1), y0 = blockIdx.y * (RY - 1), x0 = blockIdx.z * (XC - 1);
    const int x1 = min(x0 + XC, Ex);
    const int y = y0 + ry, z = z0 + tx;
    const bool nodeA = live && y < ny && z < nz;          // (y, z)
    const bool nodeB = live && y + 1 < ny && z < nz;      // (y+1, z)
    const bool colOk = live && y < ny - 1 && z < nz - 1;
    const bool nextInWarp = (lane < 31) && (tx + 1 < ZT); // z+1 neighbour is lane+1 of this warp
    const bool prevInWarp = (lane > 0) && (tx > 0);
    const bool edgeLoad = live && !nextInWarp;            // loads node z+1 itself
    const bool xSend = live && !nextInWarp && (tx + 1 < ZT);
    const bool xRecv = live && !prevInWarp && (tx > 0);
    const bool xRecvBelow = live && ry > 0 && tx > 0 && (((t - ZT) & 31) == 0); // thread (ry-1, tx) is an xRecv thread
    const bool spz = live && (tx == ZT - 1) && (z + 1 == nz - 1);
    const bool spRow = (y0 + RY == ny - 1);               // thread row RY-1 also owns node row ny-1
    const bool ownYZ = (ry > 0 || blockIdx.y == 0) && (tx > 0 || blockIdx.x == 0);
    const size_t planeStride = size_t(ny) * nz * 3;
    const Real *uA = u + (size_t(y) * nz + z) * 3;        // + p * planeStride
    const Real *eP = Ey + size_t(y) * g.E[2] + z;         // + x * E1*E2
    const size_t eStride = size_t(g.E[1]) * g.E[2];

    // raw node values of the plane being prefetched
    Real pa[3], pb[3];
    auto loadRaw = [&](int p) {
        const Real *q = uA + size_t(p) * planeStride;
#pragma unroll
        for (int c = 0; c < 3; ++c) { pa[c] = nodeA ? q[c] : 0.0; pb[c] = nodeB ? q[size_t(nz) * 3 + c] : 0.0; }
    };
    // face quantities of plane p from the raw values: F[my*2+mz][c]
    auto faces = [&](int p, Real (&F)[4][3]) {
        Real an[3], bn[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) { an[c] = __shfl_down_sync(0xffffffffu, pa[c], 1); bn[c] = __shfl_down_sync(0xffffffffu, pb[c], 1); }
        if (edgeLoad) {
            const Real *q = uA + size_t(p) * planeStride + 3;
            const bool okA = y < ny && z + 1 < nz, okB = y + 1 < ny && z + 1 < nz;
#pragma unroll
            for (int c = 0; c < 3; ++c) { an[c] = okA ? q[c] : 0.0; bn[c] = okB ? q[size_t(nz) * 3 + c] : 0.0; }
        }
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const Real zs0 = pa[c] + an[c], zd0 = pa[c] - an[c], zs1 = pb[c] + bn[c], zd1 = pb[c] - bn[c];
            F[0][c] = zs0 + zs1; F[2][c] = zs0 - zs1; F[1][c] = zd0 + zd1; F[3][c] = zd0 - zd1;
        }
    };

    Real Fprev[4][3], Gprev[4][3];
#pragma unroll
    for (int k = 0; k < 4; ++k)
#pragma unroll
        for (int c = 0; c < 3; ++c) Gprev[k][c] = 0.0;
    loadRaw(x0);
    faces(x0, Fprev);
    if (x0 + 1 < nx) loadRaw(x0 + 1);
    Real Enext = (colOk && x0 < Ex) ? eP[size_t(x0) * eStride] : 0.0;

    const int nIter = (x1 - x0) + (x1 == Ex ? 1 : 0);
    for (int it = 0; it < nIter; ++it) {
        const int x = x0 + it, buf = it & 1;
        Real Gtot[4][3];
        if (x < Ex) {
            Real Fnew[4][3];
            faces(x + 1, Fnew);
            const Real Ee = Enext;
            // prefetch the next step's inputs before the arithmetic of this one
            if (x + 2 < nx && it + 1 < nIter) loadRaw(x + 2);
            Enext = (colOk && x + 1 < Ex) ? eP[size_t(x + 1) * eStride] : 0.0;
            // x stage: uh[mx*4 + k][c]
            Real uh[8][3], yh[8][3];
#pragma unroll
            for (int k = 0; k < 4; ++k)
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    if (!(SPARSE && k == 0)) uh[k][c] = Fprev[k][c] + Fnew[k][c];
                    uh[4 + k][c] = Fprev[k][c] - Fnew[k][c];
                    Fprev[k][c] = Fnew[k][c];
                }
            // 3x3 blocks
#pragma unroll
            for (int gi = 0; gi < 8; ++gi)
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    const int m = gi ^ khatAxisBit(c);
                    bool first = true;
                    Real acc = 0.0;
#pragma unroll
                    for (int c2 = 0; c2 < 3; ++c2) {
                        if (SPARSE && !khatEntryNonZero(gi, c, c2)) continue;
                        const int m2 = gi ^ khatAxisBit(c2);
                        acc = first ? kh.b[gi][c][c2] * uh[m2][c2] : fma(kh.b[gi][c][c2], uh[m2][c2], acc);
                        first = false;
                    }
                    yh[m][c] = acc;
                }
            // inverse x stage + modulus: plane x gets Gprev + E (yh0 + yh1); plane x+1 gets E (yh0 - yh1)
            const Real nEe = -Ee;
#pragma unroll
            for (int k = 0; k < 4; ++k)
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    if (SPARSE && k == 0) { // yh[0][c] == 0
                        Gtot[k][c] = fma(Ee, yh[4][c], Gprev[k][c]);
                        Gprev[k][c] = nEe * yh[4][c];
                    } else {
                        Gtot[k][c] = fma(Ee, yh[k][c] + yh[4 + k][c], Gprev[k][c]);
                        Gprev[k][c] = Ee * (yh[k][c] - yh[4 + k][c]);
                    }
                }
        } else {
#pragma unroll
            for (int k = 0; k < 4; ++k)
#pragma unroll
                for (int c = 0; c < 3; ++c) Gtot[k][c] = Gprev[k][c];
        }
        // ---- backward, plane x.  z stage: index k = my*2 + mz ----
        Real zl[2][3], zh[2][3], Ye[2][3];
#pragma unroll
        for (int my = 0; my < 2; ++my)
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                zl[my][c] = Gtot[my * 2][c] + Gtot[my * 2 + 1][c];
                zh[my][c] = Gtot[my * 2][c] - Gtot[my * 2 + 1][c];
                Real r = __shfl_up_sync(0xffffffffu, zh[my][c], 1);
                if (!prevInWarp) r = 0.0;
                Ye[my][c] = zl[my][c] + r;
            }
        if (xSend) {
#pragma unroll
            for (int my = 0; my < 2; ++my)
#pragma unroll
                for (int c = 0; c < 3; ++c) sX[buf][ry][(tx + 1) >> 5][my * 3 + c] = zh[my][c];
        }
        Real yl[3], yhh[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) { yl[c] = Ye[0][c] + Ye[1][c]; yhh[c] = Ye[0][c] - Ye[1][c]; }
        if (live && ry + 1 < RY) {
#pragma unroll
            for (int c = 0; c < 3; ++c) sYx[buf][ry + 1][c][tx] = yhh[c];
        }
        Real spl[3], sph[3];
        if (spz) {
#pragma unroll
            for (int c = 0; c < 3; ++c) { spl[c] = zh[0][c] + zh[1][c]; sph[c] = zh[0][c] - zh[1][c]; }
            if (ry + 1 < RY) {
#pragma unroll
                for (int c = 0; c < 3; ++c) sYsp[buf][ry + 1][c] = sph[c];
            }
        }
        __syncthreads();
        const bool storePlane = (it > 0) || (blockIdx.z == 0);
        Real res[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) res[c] = yl[c] + ((live && ry > 0) ? sYx[buf][ry][c][tx] : 0.0);
        if (xRecv) { // the part of the previous warp's last column that the shuffle could not deliver
            const int s = tx >> 5;
#pragma unroll
            for (int c = 0; c < 3; ++c) res[c] += sX[buf][ry][s][c] + sX[buf][ry][s][3 + c];
        }
        if (xRecvBelow) { // ... and the same omission in what the thread of the row below sent up
            const int s = tx >> 5;
#pragma unroll
            for (int c = 0; c < 3; ++c) res[c] += sX[buf][ry - 1][s][c] - sX[buf][ry - 1][s][3 + c];
        }
        if (storePlane) {
            if (nodeA && ownYZ) marchStore<MODE>(g, x, y, z, res, b, dmask, out);
            if (spz && y < ny && (ry > 0 || blockIdx.y == 0)) {
                Real v[3];
#pragma unroll
                for (int c = 0; c < 3; ++c) v[c] = spl[c] + (ry > 0 ? sYsp[buf][ry][c] : 0.0);
                marchStore<MODE>(g, x, y, z + 1, v, b, dmask, out);
            }
            if (spRow && live && ry == RY - 1) { // node row ny-1 = y+1: only this element row touches it
                if (z < nz && (tx > 0 || blockIdx.x == 0)) {
                    Real v[3];
#pragma unroll
                    for (int c = 0; c < 3; ++c) v[c] = yhh[c];
                    if (xRecv) {
                        const int s = tx >> 5;
#pragma unroll
                        for (int c = 0; c < 3; ++c) v[c] += sX[buf][ry][s][c] - sX[buf][ry][s][3 + c];
                    }
                    marchStore<MODE>(g, x, y + 1, z, v, b, dmask, out);
                }
                if (spz) marchStore<MODE>(g, x, y + 1, z + 1, sph, b, dmask, out);
            }
        }
    }
}

// tiles of t columns overlapping by one: 1 + ceil(max(0, ne - t) / (t - 1))
inline int marchTiles(int ne, int t) { return 1 + (ne > t ? (ne - t + t - 2) / (t - 1) : 0); }

// Khat = W K0 W^T / 64 in block order.  structure: 0 = K0 is not block diagonal in the twisted Hadamard
// basis (use the gather kernel), 1 = block diagonal (dense 3x3 blocks), 2 = block diagonal with the
// isotropic sparsity pattern of khatEntryNonZero.
inline int computeKhat(const Real *K0 /*24x24*/, KhatParam &kh) {
    static Real W[24][24];
    for (int m = 0; m < 8; ++m)
        for (int j = 0; j < 8; ++j) {
            const int pc = khatPopc(m & j);
            for (int c = 0; c < 3; ++c)
                for (int c2 = 0; c2 < 3; ++c2) W[m * 3 + c][j * 3 + c2] = (c == c2) ? ((pc & 1) ? -1.0 : 1.0) : 0.0;
        }
    static Real T[24][24], Kh[24][24];
    Real kmax = 0.0;
    for (int a = 0; a < 24; ++a)
        for (int j = 0; j < 24; ++j) {
            Real s = 0.0;
            for (int i = 0; i < 24; ++i) s += W[a][i] * K0[i * 24 + j];
            T[a][j] = s;
        }
    for (int a = 0; a < 24; ++a)
        for (int bb = 0; bb < 24; ++bb) {
            Real s = 0.0;
            for (int j = 0; j < 24; ++j) s += T[a][j] * W[bb][j];
            Kh[a][bb] = s / 64.0;
            if (std::abs(Kh[a][bb]) > kmax) kmax = std::abs(Kh[a][bb]);
        }
    bool used[24][24] = {};
    int structure = 2;
    for (int gi = 0; gi < 8; ++gi)
        for (int c = 0; c < 3; ++c)
            for (int c2 = 0; c2 < 3; ++c2) {
                const int a = (gi ^ khatAxisBit(c)) * 3 + c, bb = (gi ^ khatAxisBit(c2)) * 3 + c2;
                kh.b[gi][c][c2] = Kh[a][bb]; used[a][bb] = true;
                if (!khatEntryNonZero(gi, c, c2) && std::abs(Kh[a][bb]) > 1e-14 * kmax) structure = 1;
            }
    for (int a = 0; a < 24; ++a)
        for (int bb = 0; bb < 24; ++bb)
            if (!used[a][bb] && std::abs(Kh[a][bb]) > 1e-14 * kmax) return 0;
    return structure;
}

struct MarchPlan { int cfg = 0, XC = 2; dim3 grid; };

// picks the tile shape with the fewest thread-steps, then the chunk length that fills the SMs (1 CTA / SM)
inline MarchPlan planMarch(const Grid<3> &g, int numSMs) {
    static const int shapes[4][2] = {{8, 64}, {16, 32}, {15, 33}, {7, 65}};
    MarchPlan best; double bestCost = 1e300;
    for (int s = 0; s < 4; ++s) {
        const int RY = shapes[s][0], ZT = shapes[s][1], NT = ((RY * ZT + 31) / 32) * 32;
        const int ty = marchTiles(g.E[1], RY), tz = marchTiles(g.E[2], ZT);
        for (int XC = 2; XC <= (g.E[0] > 2 ? g.E[0] : 2); ++XC) {
            const int nch = marchTiles(g.E[0], XC);
            const long long ctas = (long long)ty * tz * nch;
            const long long waves = (ctas + numSMs - 1) / numSMs;
            const double cost = double(waves) * (std::min(XC, g.E[0]) + 2.5) * NT;
            if (cost < bestCost) { bestCost = cost; best.cfg = s; best.XC = XC; best.grid = dim3(tz, ty, nch); }
        }
    }
    return best;
}

template <int MODE>
inline void launchApplyMarch(const Grid<3> &g, const KhatParam &kh, bool sparse, const MarchPlan &pl, const Real *Ey, const Real *u,
                             const Real *b, const uint8_t *dmask, Real *out, cudaStream_t stream) {
#define VFEM_MARCH_CASE(S, RY, ZT)                                                                                             \
    case S:                                                                                                                    \
        if (sparse) applyMarchKernel<MODE, true, RY, ZT><<<pl.grid, MarchCfg<RY, ZT>::NT, 0, stream>>>(g, kh, Ey, u, b, dmask, out, pl.XC); \
        else applyMarchKernel<MODE, false, RY, ZT><<<pl.grid, MarchCfg<RY, ZT>::NT, 0, stream>>>(g, kh, Ey, u, b, dmask, out, pl.XC);      \
        break;
    switch (pl.cfg) {
        VFEM_MARCH_CASE(0, 8, 64)
        VFEM_MARCH_CASE(1, 16, 32)
        VFEM_MARCH_CASE(2, 15, 33)
        VFEM_MARCH_CASE(3, 7, 65)
    }
#undef VFEM_MARCH_CASE
}

} // namespace vfem
