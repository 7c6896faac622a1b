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
#include <cstdlib>
#include <type_traits>

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
    static constexpr int MINB = (512 / NT) > 0 ? (512 / NT) : 1; // resident CTAs per SM the register budget is sized for
    static constexpr int XS = ZT / 32 + 2; // cross-warp slots per row
    // running 64-bit pointers (8 more registers, ~50 fewer integer instructions per step) where the register budget has room
    // registers per thread with MINB CTAs resident: the register file is 4 x 16384 (one bank per scheduler) and the warps are
    // dealt round-robin, so 13-16 warps per SM stop at 128 registers and 9-12 at 168 (what __launch_bounds__ arrives at)
    static constexpr int REGS_FIT = 16384 / (((NT / 32) * MINB + 3) / 4) / 32 / 8 * 8;
    static constexpr bool PTR = REGS_FIT >= 160;
};

// per-thread role bits (kept in one register; tested with single-bit predicates)
enum : unsigned {
    MF_LIVE = 1u, MF_COL = 2u, MF_EDGE = 4u, MF_XSEND = 8u, MF_XRECV = 16u, MF_XRECVB = 32u, MF_SPZ = 64u, MF_STORE = 128u,
    MF_YSEND = 256u, MF_YRECV = 512u, MF_PREV = 1024u, MF_SPSTORE = 2048u, MF_TOP = 4096u, MF_TOPSTORE = 8192u, MF_DZ = 16384u, MF_CHUNK0 = 32768u
};

// MODE 0: out = K u;  MODE 1: out = mask(b - K u);  MODE 2: out = mask(K u).   n = flat node index
template <int MODE>
__device__ __forceinline__ void marchStore(unsigned n, const Real *v, const Real *__restrict__ b, const uint8_t *__restrict__ dmask,
                                           Real *__restrict__ out) {
    Real *o = out + size_t(n) * 3;
    if (MODE == 0) { o[0] = v[0]; o[1] = v[1]; o[2] = v[2]; return; }
    const unsigned m = dmask[n];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        Real r = v[c];
        if (MODE == 1) r = b[size_t(n) * 3 + c] - r;
        if ((m >> c) & 1u) r = 0.0;
        o[c] = r;
    }
}

// Face quantities F[my*2+mz][c] of one node plane from the raw values of rows y (a) and y+1 (b):
// z stage by shuffle (threads whose z+1 neighbour is not lane+1 load that node themselves), then y stage.
__device__ __forceinline__ void marchFaces(const Real (&pa)[3], const Real (&pb)[3], unsigned flags, const Real *__restrict__ qa,
                                           const Real *__restrict__ qb, Real (&F)[4][3]) {
    Real an[3], bn[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) { an[c] = __shfl_down_sync(0xffffffffu, pa[c], 1); bn[c] = __shfl_down_sync(0xffffffffu, pb[c], 1); }
    if (flags & MF_EDGE) { // node z+1 (clamped into the grid) of rows y and y+1
        const int d = (flags & MF_DZ) ? 3 : 0;
#pragma unroll
        for (int c = 0; c < 3; ++c) { an[c] = qa[d + c]; bn[c] = qb[d + c]; }
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const Real zs0 = pa[c] + an[c], zd0 = pa[c] - an[c], zs1 = pb[c] + bn[c], zd1 = pb[c] - bn[c];
        F[0][c] = zs0 + zs1; F[2][c] = zs0 - zs1; F[1][c] = zd0 + zd1; F[3][c] = zd0 - zd1;
    }
}

// One element layer: uh = Fold +- Fnew, yh = blocks * uh, Gtot = Gprev + E (yh0 + yh1), Gprev = E (yh0 - yh1).
template <bool SPARSE>
__device__ __forceinline__ void marchLayer(const KhatParam &kh, const Real (&Fold)[4][3], const Real (&Fnew)[4][3], Real Ee,
                                           Real (&Gprev)[4][3], Real (&Gtot)[4][3]) {
    Real uh[8][3], yh[8][3];
#pragma unroll
    for (int k = 0; k < 4; ++k)
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            if (!(SPARSE && k == 0)) uh[k][c] = Fold[k][c] + Fnew[k][c];
            uh[4 + k][c] = Fold[k][c] - Fnew[k][c];
        }
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
}

template <int MODE, bool SPARSE, int RY, int ZT>
__global__ void __launch_bounds__(MarchCfg<RY, ZT>::NT, MarchCfg<RY, ZT>::MINB)
applyMarchKernel(Grid<3> g, KhatParam kh, const Real *__restrict__ Ey, const Real *__restrict__ u, const Real *__restrict__ b,
                 const uint8_t *__restrict__ dmask, Real *__restrict__ out, int XC, int edgeSkip) {
    constexpr int XS = MarchCfg<RY, ZT>::XS;
    constexpr bool PTR = MarchCfg<RY, ZT>::PTR;
    __shared__ Real sYx[2][RY][3][ZT];   // row ry -> slot ry+1 : y-stage part for node row y+1
    __shared__ Real sX[2][RY][XS][6];    // zh of a thread whose z+1 neighbour sits in the next warp
    __shared__ Real sYsp[2][RY][3];      // y exchange of the special last-z node column

    const int nx = g.n[0], ny = g.n[1], nz = g.n[2], Ex = g.E[0];
    const int x0 = blockIdx.z * (XC - 1);
    const int x1 = min(x0 + XC, Ex);
    unsigned flags, offA, offB, eOff, nodeBase, sIdx; // sIdx = ry * 3 * ZT + tx : element index into sYx[buf]
    {
        const int t = threadIdx.x, lane = t & 31;
        int ry = t / ZT;
        const int tx = t - ry * ZT;
        const bool live = ry < RY;
        if (!live) ry = RY - 1;
        const int z0 = blockIdx.x * (ZT - 1), y0 = blockIdx.y * (RY - 1);
        const int y = y0 + ry, z = z0 + tx;
        const bool nextInWarp = (lane < 31) && (tx + 1 < ZT); // z+1 neighbour is lane+1 of this warp
        const bool prevInWarp = (lane > 0) && (tx > 0);
        const bool spz = live && (tx == ZT - 1) && (z + 1 == nz - 1);
        const bool top = live && (y0 + RY == ny - 1) && (ry == RY - 1); // this thread row also owns node row ny-1
        const bool ownY = (ry > 0 || blockIdx.y == 0), ownZ = (tx > 0 || blockIdx.x == 0);
        flags = (live ? MF_LIVE : 0u) | ((live && y < ny - 1 && z < nz - 1) ? MF_COL : 0u) | ((live && !nextInWarp) ? MF_EDGE : 0u) |
                ((live && !nextInWarp && tx + 1 < ZT) ? MF_XSEND : 0u) | ((live && !prevInWarp && tx > 0) ? MF_XRECV : 0u) |
                ((live && ry > 0 && tx > 0 && (((t - ZT) & 31) == 0)) ? MF_XRECVB : 0u) | (spz ? MF_SPZ : 0u) |
                ((live && y < ny && z < nz && ownY && ownZ) ? MF_STORE : 0u) | ((live && ry + 1 < RY) ? MF_YSEND : 0u) |
                ((live && ry > 0) ? MF_YRECV : 0u) | (prevInWarp ? MF_PREV : 0u) | ((spz && y < ny && ownY) ? MF_SPSTORE : 0u) |
                (top ? MF_TOP : 0u) | ((top && z < nz && ownZ) ? MF_TOPSTORE : 0u) | ((z + 1 < nz) ? MF_DZ : 0u) |
                ((blockIdx.z == 0 && !(edgeSkip & 1)) ? MF_CHUNK0 : 0u);
        // addresses are clamped into the grid instead of predicating the loads: an out-of-grid node only ever
        // feeds element columns whose modulus is forced to 0 (MF_COL clear), so its (finite) value is irrelevant
        const int yc = min(y, ny - 1), yb = min(y + 1, ny - 1), zc = min(z, nz - 1);
        offA = unsigned(yc * nz + zc) * 3u; offB = unsigned(yb * nz + zc) * 3u;
        eOff = unsigned(min(y, g.E[1] - 1) * g.E[2] + min(z, g.E[2] - 1));
        nodeBase = unsigned(yc * nz + zc);
        sIdx = unsigned(ry * 3 * ZT + tx);
    }
    // opaque to the optimiser: otherwise it re-derives single role bits from threadIdx / blockIdx inside the march loop
    // (~25 integer instructions per step) instead of testing the register
    asm volatile("" : "+r"(flags), "+r"(sIdx));
    const Real prevMask = (flags & MF_PREV) ? 1.0 : 0.0;
    const unsigned planeStride = unsigned(ny) * unsigned(nz) * 3u, planeNodes = unsigned(ny) * unsigned(nz);
    const unsigned eStride = unsigned(g.E[1]) * unsigned(g.E[2]);

    Real pa[3], pb[3], FA[4][3], FB[4][3], Gprev[4][3], Enext;
#pragma unroll
    for (int k = 0; k < 4; ++k)
#pragma unroll
        for (int c = 0; c < 3; ++c) Gprev[k][c] = 0.0;
    // running pointers, advanced by one plane / layer per step (no per-step 64-bit index arithmetic):
    // uA / uB -> this thread's node of rows y / y+1 in the plane the NEXT marchFaces consumes; eP -> the modulus in Enext;
    // oP -> this thread's output node in the plane being stored
    const Real *uA = u + size_t(x0) * planeStride + offA, *uB = u + size_t(x0) * planeStride + offB;
    const Real *eP = Ey + size_t(x0) * eStride + eOff;
    Real *oP = out + (size_t(x0) * planeNodes + nodeBase) * 3;
    {
#pragma unroll
        for (int c = 0; c < 3; ++c) { pa[c] = uA[c]; pb[c] = uB[c]; }
        marchFaces(pa, pb, flags, uA, uB, FA);
        uA += planeStride; uB += planeStride; // x0 + 1 <= Ex is always a valid plane
#pragma unroll
        for (int c = 0; c < 3; ++c) { pa[c] = uA[c]; pb[c] = uB[c]; }
        Enext = *eP;
    }

    // LAYER: the step advances through element layer x (always, except the closing step that only flushes plane Ex)
    auto step = [&](auto hasLayer, const int it, Real (&Fold)[4][3], Real (&Fnew)[4][3]) {
        constexpr bool LAYER = decltype(hasLayer)::value;
        const int x = x0 + it, buf = it & 1;
        Real Gtot[4][3];
        if (LAYER) {
            if (!PTR) { uA = u + size_t(x + 1) * planeStride + offA; uB = u + size_t(x + 1) * planeStride + offB; }
            marchFaces(pa, pb, flags, uA, uB, Fnew); // uA / uB: plane x + 1
            const Real Ee = (flags & MF_COL) ? Enext : 0.0;
            if (x + 2 < nx) { // prefetch the next step's inputs before the arithmetic of this one
                if (PTR) { uA += planeStride; uB += planeStride; eP += eStride; }
                else { uA += planeStride; uB += planeStride; eP = Ey + size_t(x + 1) * eStride + eOff; }
#pragma unroll
                for (int c = 0; c < 3; ++c) { pa[c] = uA[c]; pb[c] = uB[c]; }
                Enext = *eP;
            }
            marchLayer<SPARSE>(kh, Fold, Fnew, Ee, Gprev, Gtot);
        } else {
#pragma unroll
            for (int k = 0; k < 4; ++k)
#pragma unroll
                for (int c = 0; c < 3; ++c) Gtot[k][c] = Gprev[k][c];
        }
        // ---- backward, plane x.  z stage (index k = my*2 + mz), then y stage ----
        Real zh[2][3], Ye[2][3];
#pragma unroll
        for (int my = 0; my < 2; ++my)
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                const Real zl = Gtot[my * 2][c] + Gtot[my * 2 + 1][c];
                zh[my][c] = Gtot[my * 2][c] - Gtot[my * 2 + 1][c];
                const Real r = __shfl_up_sync(0xffffffffu, zh[my][c], 1);
                Ye[my][c] = fma(r, prevMask, zl); // prevMask = 0 where lane-1 is not the z-1 neighbour (r is then this lane's own, finite, value)
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
        Real res[3], yhh[3];
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
        __syncthreads();
        if (flags & MF_YRECV) {
#pragma unroll
            for (int c = 0; c < 3; ++c) res[c] += sYb[c * ZT];
        }
        const Real *sXr = sXb; // -> sX[buf][ry][tx >> 5][0], only formed on the rare paths
        if (flags & (MF_XRECV | MF_XRECVB)) {
            const int ry = sIdx / (3 * ZT), tx = sIdx - ry * 3 * ZT;
            sXr = sXb + (ry * XS + (tx >> 5)) * 6;
        }
        if (flags & MF_XRECV) { // the part of the previous warp's last column that the shuffle could not deliver
#pragma unroll
            for (int c = 0; c < 3; ++c) res[c] += sXr[c] + sXr[3 + c];
        }
        if (flags & MF_XRECVB) { // ... and the same omission in what the thread of the row below sent up
#pragma unroll
            for (int c = 0; c < 3; ++c) res[c] += sXr[c - XS * 6] - sXr[3 + c - XS * 6];
        }
        if (it > 0 || (flags & MF_CHUNK0)) {
            const unsigned n = unsigned(x) * planeNodes + nodeBase;
            if (flags & MF_STORE) {
                if (MODE == 0 && PTR) { oP[0] = res[0]; oP[1] = res[1]; oP[2] = res[2]; }
                else marchStore<MODE>(n, res, b, dmask, out);
            }
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
        if (PTR) oP += planeStride;
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
}

// tiles of t columns overlapping by one: 1 + ceil(max(0, ne - t) / (t - 1))
inline int marchTiles(int ne, int t) { return 1 + (ne > t ? (ne - t + t - 2) / (t - 1) : 0); }

// Khat = W K0 W^T / 64 in block order.  structure: 0 = K0 is not block diagonal in the twisted Hadamard
// basis (use the gather kernel), 1 = block diagonal (dense 3x3 blocks), 2 = block diagonal with the
// isotropic sparsity pattern of khatEntryNonZero.
inline int computeKhat(const Real *K0 /*24x24*/, KhatParam &kh) {
    Real W[24][24]; // locals: simulators are built concurrently by the slab tests' rank threads
    for (int m = 0; m < 8; ++m)
        for (int j = 0; j < 8; ++j) {
            const int pc = khatPopc(m & j);
            for (int c = 0; c < 3; ++c)
                for (int c2 = 0; c2 < 3; ++c2) W[m * 3 + c][j * 3 + c2] = (c == c2) ? ((pc & 1) ? -1.0 : 1.0) : 0.0;
        }
    Real T[24][24], Kh[24][24];
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

constexpr int kMarchShapes = 13;
struct MarchPlan { int cfg = 0, XC = 2; dim3 grid; };

// Picks tile shape and chunk length from a measured cost model: a CTA advances one plane per step, and a step takes
// stepUs[shape] on B200 whatever the chunk length (the march is latency-bound: profiles/r01d_march_shape_sweep.txt), so
// cost = waves * (steps per chunk + pipeline fill) * stepUs, with waves = CTAs / (SMs * resident CTAs per SM).
inline MarchPlan planMarch(const Grid<3> &g, int numSMs) {
    static const int shapes[kMarchShapes][2] = {{8, 64}, {16, 32}, {15, 33}, {7, 65}, {6, 64}, {4, 64}, {8, 32}, {8, 33}, {5, 64}, {4, 33}, {6, 65}, {5, 33}, {11, 33}};
    static const double stepUs[kMarchShapes] = {1.78, 1.90, 1.98, 1.69, 1.195, 1.68, 1.64, 1.06, 1.10, 1.66, 1.58, 1.11, 1.20}; // profiles/r01n_march_loop_sass.txt, r01q
    MarchPlan best; double bestCost = 1e300;
    const char *fs = getenv("VFEM_MARCH_SHAPE"), *fx = getenv("VFEM_MARCH_XC"); // experiment overrides
    for (int s = 0; s < kMarchShapes; ++s) {
        if (fs) s = atoi(fs);
        const int RY = shapes[s][0], ZT = shapes[s][1], NT = ((RY * ZT + 31) / 32) * 32, MINB = 512 / NT > 0 ? 512 / NT : 1;
        const int ty = marchTiles(g.E[1], RY), tz = marchTiles(g.E[2], ZT);
        for (int XC = 2; XC <= (g.E[0] > 2 ? g.E[0] : 2); ++XC) {
            if (fx) XC = std::max(2, atoi(fx));
            const int nch = marchTiles(g.E[0], XC);
            const long long ctas = (long long)ty * tz * nch;
            const long long waves = (ctas + (long long)numSMs * MINB - 1) / ((long long)numSMs * MINB);
            const double cost = double(waves) * (std::min(XC, g.E[0]) + 2.5) * stepUs[s];
            if (cost < bestCost) { bestCost = cost; best.cfg = s; best.XC = XC; best.grid = dim3(tz, ty, nch); }
            if (fx) break;
        }
        if (fs) break;
    }
    return best;
}

template <int MODE>
inline void launchApplyMarch(const Grid<3> &g, const KhatParam &kh, bool sparse, const MarchPlan &pl, const Real *Ey, const Real *u,
                             const Real *b, const uint8_t *dmask, Real *out, cudaStream_t stream, int edgeSkip = 0) {
#define VFEM_MARCH_CASE(S, RY, ZT)                                                                                             \
    case S:                                                                                                                    \
        if (sparse) applyMarchKernel<MODE, true, RY, ZT><<<pl.grid, MarchCfg<RY, ZT>::NT, 0, stream>>>(g, kh, Ey, u, b, dmask, out, pl.XC, edgeSkip); \
        else applyMarchKernel<MODE, false, RY, ZT><<<pl.grid, MarchCfg<RY, ZT>::NT, 0, stream>>>(g, kh, Ey, u, b, dmask, out, pl.XC, edgeSkip);      \
        break;
    switch (pl.cfg) {
        VFEM_MARCH_CASE(0, 8, 64)
        VFEM_MARCH_CASE(1, 16, 32)
        VFEM_MARCH_CASE(2, 15, 33)
        VFEM_MARCH_CASE(3, 7, 65)
        VFEM_MARCH_CASE(4, 6, 64)
        VFEM_MARCH_CASE(5, 4, 64)
        VFEM_MARCH_CASE(6, 8, 32)
        VFEM_MARCH_CASE(7, 8, 33)
        VFEM_MARCH_CASE(8, 5, 64)
        VFEM_MARCH_CASE(9, 4, 33)
        VFEM_MARCH_CASE(10, 6, 65)
        VFEM_MARCH_CASE(11, 5, 33)
        VFEM_MARCH_CASE(12, 11, 33)
    }
#undef VFEM_MARCH_CASE
}

} // namespace vfem
