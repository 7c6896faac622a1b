// ndr_b200 — CUDA kernels of the elasticity solve (sm_100a, FP64).
//
// Every kernel here is a gather: one thread owns one output node (or one stencil block) and
// reads its 3^N neighbourhood, so no atomics are needed and results are deterministic.
// Reference semantics (cited per kernel): TPS.hh = VoxelFEM/TensorProductSimulator.hh,
// MG.hh = VoxelFEM/MultigridSolver.hh.
#pragma once
#include "common.h"
#include <cooperative_groups.h>

namespace vfem {

constexpr int kBlock = 128;

// ---- index helpers --------------------------------------------------------------------------
template <int N>
__device__ __forceinline__ void unflattenNode(const Grid<N> &g, int n, int *idx) {
#pragma unroll
    for (int d = N - 1; d > 0; --d) { int q = n / g.n[d]; idx[d] = n - q * g.n[d]; n = q; }
    idx[0] = n;
}
template <int N>
__device__ __forceinline__ int flattenNode(const Grid<N> &g, const int *idx) {
    int r = 0;
#pragma unroll
    for (int d = 0; d < N; ++d) r += idx[d] * g.ninc[d];
    return r;
}
// thread t of colour lattice `par` (idx_d = par_d + 2*k_d), restricted to the owned planes [own0, own1] of axis 0;
// returns false when t is out of range
template <int N>
__host__ __device__ __forceinline__ int colourFirstPlane(const Grid<N> &g, int colour) {
    const int par = (colour >> (N - 1)) & 1;
    return g.own0 + ((g.own0 ^ par) & 1);
}
template <int N>
__host__ __device__ __forceinline__ int colourPlaneCount(const Grid<N> &g, int colour) {
    const int first = colourFirstPlane<N>(g, colour);
    return first <= g.own1 ? (g.own1 - first) / 2 + 1 : 0;
}
template <int N>
__device__ __forceinline__ bool colourNode(const Grid<N> &g, int t, int colour, int *idx) {
    int cnt[N];
#pragma unroll
    for (int d = 1; d < N; ++d) { int par = (colour >> (N - 1 - d)) & 1; cnt[d] = (g.n[d] - 1 - par) / 2 + 1; }
#pragma unroll
    for (int d = N - 1; d > 0; --d) { int q = t / cnt[d]; idx[d] = ((colour >> (N - 1 - d)) & 1) + 2 * (t - q * cnt[d]); t = q; }
    idx[0] = colourFirstPlane<N>(g, colour) + 2 * t;
    return t < colourPlaneCount<N>(g, colour);
}
template <int N>
__host__ __device__ inline int colourCount(const Grid<N> &g, int colour) {
    long long tot = colourPlaneCount<N>(g, colour);
    for (int d = 1; d < N; ++d) { int par = (colour >> (N - 1 - d)) & 1; tot *= (g.n[d] - 1 - par) / 2 + 1; }
    return int(tot);
}

// ---- fine level, matrix-free: row n of K(rho) u ------------------------------------------------
// acc[c] += sum_{e incident to n} E_e * K0[(loc_e(n), c), :] . u_e      (TPS.hh:933-951 gathered per
// node; the same row is what MG.hh:205-222 subtracts from b in the smoother).
// M (optional) = sum_e E_e * K0[(loc,:),(loc,:)], the node's diagonal block (MG.hh:223-225).
// Out-of-grid neighbours read 0 and out-of-grid elements get modulus 0: branch-free, and
// adding exact zeros leaves the result unchanged.
// row of K(rho) u at one node from its gathered 3^N neighbourhood un[k][c] (zeros outside the grid)
template <int N, bool WITH_M>
__device__ __forceinline__ void fineNodeRowCompute(const Grid<N> &g, const K0Param<N> &K, const Real *__restrict__ Ey, const int *idx,
                                                   const Real (&un)[ipow3(N)][N], Real *acc, Real *M) {
    constexpr int NPE = 1 << N, KS = N * NPE;
#pragma unroll
    for (int s = 0; s < NPE; ++s) {
        int ei = 0, loc = 0; bool ok = true;
#pragma unroll
        for (int d = 0; d < N; ++d) {
            int sd = (s >> (N - 1 - d)) & 1;
            int e = idx[d] - 1 + sd; ok = ok && (e >= 0) && (e < g.E[d]);
            ei += e * g.einc[d];
            loc |= (1 - sd) << (N - 1 - d);
        }
        const Real Ee = ok ? Ey[ei] : 0.0;
        Real r[N];
#pragma unroll
        for (int c = 0; c < N; ++c) r[c] = 0.0;
#pragma unroll
        for (int q = 0; q < NPE; ++q) {
            int k = 0; // neighbourhood slot of local node q of incident element s
#pragma unroll
            for (int d = 0; d < N; ++d) {
                int sd = (s >> (N - 1 - d)) & 1, qd = (q >> (N - 1 - d)) & 1;
                k = k * 3 + (sd - 1 + qd + 1);
            }
#pragma unroll
            for (int b = 0; b < N; ++b)
#pragma unroll
                for (int c = 0; c < N; ++c) r[c] = fma(K.k[(loc * N + c) * KS + q * N + b], un[k][b], r[c]);
        }
#pragma unroll
        for (int c = 0; c < N; ++c) acc[c] = fma(Ee, r[c], acc[c]);
        if (WITH_M) {
#pragma unroll
            for (int c = 0; c < N; ++c)
#pragma unroll
                for (int c2 = 0; c2 < N; ++c2) M[c * N + c2] = fma(Ee, K.k[(loc * N + c) * KS + loc * N + c2], M[c * N + c2]);
        }
    }
}

template <int N, bool WITH_M, bool LDCG = false>
__device__ __forceinline__ void fineNodeRow(const Grid<N> &g, const K0Param<N> &K, const Real *__restrict__ Ey,
                                            const Real *u, const int *idx, Real *acc, Real *M) {
    constexpr int NST = ipow3(N);
    Real un[NST][N];
#pragma unroll
    for (int k = 0; k < NST; ++k) {
        int t = k, off = 0; bool ok = true;
#pragma unroll
        for (int d = N - 1; d >= 0; --d) {
            int dl = t % 3 - 1; t /= 3;
            int v = idx[d] + dl; ok = ok && (v >= 0) && (v < g.n[d]);
            off += dl * g.ninc[d];
        }
        int base = flattenNode<N>(g, idx) + off;
#pragma unroll
        for (int c = 0; c < N; ++c) un[k][c] = ok ? (LDCG ? __ldcg(u + size_t(base) * N + c) : u[size_t(base) * N + c]) : 0.0;
    }
    fineNodeRowCompute<N, WITH_M>(g, K, Ey, idx, un, acc, M);
}

// ---- coarse levels: assembled 3^N-point block stencil ---------------------------------------------
// Node-major, row after row: S[((n*N + a)*3^N + k)*N + b] = K_l[(n,a), (n + delta(k), b)] (zero for out-of-grid
// neighbours), so the 3^N * N * N coefficients of one node are contiguous and a warp reads them fully coalesced.

// The reference stores the same operator as a block-CSC matrix (TPS.hh:649-720) and applies it
// with applyTransposeParallel (SparseMatrices.hh:1519-1568).

// Warp-per-node rows of the assembled operator.  Lane l owns the row entries e = l, l+32, ... (< 3^N*N) of every
// matrix row; entry e multiplies u[n + delta(e / N)][e % N], an address offset that depends on the lane only, so it is
// computed once per warp (StencilLane) and reused for all nodes and all N rows.  Out-of-grid neighbours have zero
// coefficients (buildLevel1StencilKernel / galerkinStencilKernel), so their addresses are clamped instead of tested.
// The N partial row sums are combined with a fixed shuffle tree (deterministic); all lanes return the full sums.
template <int N>
struct StencilLane {
    static constexpr int ROW = ipow3(N) * N, J = (ROW + 31) / 32;
    int uoff[J];   // dof offset of the neighbour value relative to n*N
    bool valid[J];
    __device__ __forceinline__ void init(const Grid<N> &g, int lane) {
#pragma unroll
        for (int j = 0; j < J; ++j) {
            const int e = lane + 32 * j;
            valid[j] = e < ROW;
            const int k = valid[j] ? e / N : 0, b = valid[j] ? e % N : 0;
            int t = k, off = 0;
#pragma unroll
            for (int d = N - 1; d >= 0; --d) { off += (t % 3 - 1) * g.ninc[d]; t /= 3; }
            uoff[j] = off * N + b;
        }
    }
};

template <int N, bool LDCG = false>
__device__ __forceinline__ void stencilWarpRow(const StencilLane<N> &sl, int ndof, const Real *__restrict__ S, const Real *u, int n,
                                               int lane, Real *acc) {
    constexpr int ROW = StencilLane<N>::ROW, J = StencilLane<N>::J;
    const Real *Sn = S + size_t(n) * N * ROW + lane;
    Real uv[J];
#pragma unroll
    for (int j = 0; j < J; ++j) {
        const int a = min(max(n * N + sl.uoff[j], 0), ndof - 1);
        uv[j] = sl.valid[j] ? (LDCG ? __ldcg(u + a) : u[a]) : 0.0;
    }
#pragma unroll
    for (int a = 0; a < N; ++a) {
        Real s = 0.0;
#pragma unroll
        for (int j = 0; j < J; ++j)
            if (sl.valid[j]) s = fma(Sn[a * ROW + 32 * j], uv[j], s);
        acc[a] = s;
    }
#pragma unroll
    for (int a = 0; a < N; ++a)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc[a] += __shfl_xor_sync(0xffffffffu, acc[a], o);
}

// Gauss-Seidel update of one node (MG.hh:248-264): t = b_n - (K u)_n, then one component-wise
// forward/backward substitution through the diagonal block M; constrained components get 0.
template <int N, bool FWD>
__device__ __forceinline__ void gsNodeSubst(const Real *t, const Real *M, unsigned mask, Real *du) {
#pragma unroll
    for (int c = 0; c < N; ++c) du[c] = 0.0;
#pragma unroll
    for (int ii = 0; ii < N; ++ii) {
        constexpr int dummy = 0; (void)dummy;
        const int i = FWD ? ii : (N - 1 - ii);
        Real s = 0.0;
#pragma unroll
        for (int j = 0; j < N; ++j) s = fma(M[i * N + j], du[j], s);
        const Real inv = ((mask >> i) & 1u) ? 0.0 : 1.0 / M[i * N + i];
        du[i] = (t[i] - s) * inv;
    }
}
template <int N>
__device__ __forceinline__ void gsNodeUpdate(const Real *t, const Real *M, unsigned mask, bool forward, Real *__restrict__ u, int n) {
    Real du[N];
    if (forward) gsNodeSubst<N, true>(t, M, mask, du);
    else gsNodeSubst<N, false>(t, M, mask, du);
#pragma unroll
    for (int c = 0; c < N; ++c) u[size_t(n) * N + c] += du[c];
}

// ---- K u / residual kernels ---------------------------------------------------------------------
// MODE 0: out = K u (TPS.hh:905-952, MG.hh:353-358)
// MODE 1: out = mask(b - K u)     (MG::computeResidual, MG.hh:401-413)
// MODE 2: out = mask(K u)         (PCG's A d, MG.hh:719-720)
template <int N, int MODE>
__global__ void __launch_bounds__(kBlock) applyFineKernel(Grid<N> g, K0Param<N> K, const Real *__restrict__ Ey,
                                                           const Real *__restrict__ u, const Real *__restrict__ b,
                                                           const uint8_t *__restrict__ dmask, Real *__restrict__ out, int edgeSkip) {
    int n = blockIdx.x * kBlock + threadIdx.x;
    if (n >= g.nn) return;
    int idx[N]; unflattenNode<N>(g, n, idx);
    // the grid is a view of a larger one: rows of its first (bit 0) / last (bit 1) plane would be partial - not stored
    if (((edgeSkip & 1) && idx[0] == 0) || ((edgeSkip & 2) && idx[0] == g.E[0])) return;
    Real acc[N];
#pragma unroll
    for (int c = 0; c < N; ++c) acc[c] = 0.0;
    fineNodeRow<N, false>(g, K, Ey, u, idx, acc, nullptr);
    unsigned m = (MODE != 0) ? dmask[n] : 0u;
#pragma unroll
    for (int c = 0; c < N; ++c) {
        Real v = (MODE == 1) ? b[size_t(n) * N + c] - acc[c] : acc[c];
        if ((m >> c) & 1u) v = 0.0;
        out[size_t(n) * N + c] = v;
    }
}

// ---- multicoloured Gauss-Seidel: one colour phase (MG.hh:284-340) --------------------------------
template <int N, int MINB>
__global__ void __launch_bounds__(kBlock, MINB) gsFineKernel(Grid<N> g, K0Param<N> K, const Real *__restrict__ Ey, Real *__restrict__ u,
                                                        const Real *__restrict__ b, const uint8_t *__restrict__ dmask, int colour,
                                                        int forward) {
    int t = blockIdx.x * kBlock + threadIdx.x;
    int idx[N];
    if (!colourNode<N>(g, t, colour, idx)) return;
    int n = flattenNode<N>(g, idx);
    Real acc[N], M[N * N];
#pragma unroll
    for (int c = 0; c < N; ++c) acc[c] = 0.0;
#pragma unroll
    for (int i = 0; i < N * N; ++i) M[i] = 0.0;
    fineNodeRow<N, true>(g, K, Ey, u, idx, acc, M);
    Real tt[N];
#pragma unroll
    for (int c = 0; c < N; ++c) tt[c] = b[size_t(n) * N + c] - acc[c];
    gsNodeUpdate<N>(tt, M, dmask[n], forward != 0, u, n);
}


// ---- serial lexicographic Gauss-Seidel (MG::smoothing, MG.hh:269-282; the debug helper behind pyVoxelFEM's
// MultigridSolver.smoothing, VoxelFEM.cc:99-104): ONE warp walks the nodes in index order (reverse order for a
// backward sweep).  Not on the solve path - the cycles use the multicoloured sweep.
template <int N>
__global__ void __launch_bounds__(32) gsSerialKernel(Grid<N> g, K0Param<N> K, const Real *__restrict__ Ey, const Real *__restrict__ S, Real *u,
                                                     const Real *__restrict__ b, const uint8_t *__restrict__ dmask, int forward) {
    const int lane = threadIdx.x;
    StencilLane<N> sl; sl.init(g, lane);
    for (int n_ = 0; n_ < g.nn; ++n_) {
        const int n = forward ? n_ : g.nn - 1 - n_;
        Real acc[N], M[N * N];
        if (S) { // assembled level: the warp forms the row sums, the diagonal block is stencil slot 3^N / 2
            stencilWarpRow<N, true>(sl, g.nn * N, S, u, n, lane, acc);
#pragma unroll
            for (int i = 0; i < N * N; ++i) M[i] = S[stencilIndex<N>(n, ipow3(N) / 2, i)];
        } else if (lane == 0) {
            int idx[N]; unflattenNode<N>(g, n, idx);
#pragma unroll
            for (int c = 0; c < N; ++c) acc[c] = 0.0;
#pragma unroll
            for (int i = 0; i < N * N; ++i) M[i] = 0.0;
            fineNodeRow<N, true, true>(g, K, Ey, u, idx, acc, M);
        }
        if (lane == 0) {
            Real tt[N];
#pragma unroll
            for (int c = 0; c < N; ++c) tt[c] = b[size_t(n) * N + c] - acc[c];
            gsNodeUpdate<N>(tt, M, dmask[n], forward != 0, u, n);
            __threadfence();
        }
        __syncwarp();
    }
}

// ---- warp-per-node kernels for the assembled levels: persistent warps striding over the nodes ----
constexpr int kWarpBlock = 256; // 8 warps per block
inline unsigned warpGridBlocks(long long nodes) {
    const long long need = (nodes * 32 + kWarpBlock - 1) / kWarpBlock, cap = 148 * 8;
    return unsigned(need < cap ? (need > 0 ? need : 1) : cap);
}

template <int N, int MODE>
__global__ void __launch_bounds__(kWarpBlock) applyStencilWarpKernel(Grid<N> g, const Real *__restrict__ S, const Real *__restrict__ u,
                                                                      const Real *__restrict__ b, const uint8_t *__restrict__ dmask,
                                                                      Real *__restrict__ out) {
    const int lane = threadIdx.x & 31, nwarps = (gridDim.x * kWarpBlock) >> 5;
    StencilLane<N> sl; sl.init(g, lane);
    for (int n = (blockIdx.x * kWarpBlock + threadIdx.x) >> 5; n < g.nn; n += nwarps) {
        Real acc[N];
        stencilWarpRow<N>(sl, g.nn * N, S, u, n, lane, acc);
        if (lane < N) {
            const unsigned m = (MODE != 0) ? dmask[n] : 0u;
            Real v = acc[0];
#pragma unroll
            for (int c = 1; c < N; ++c) if (lane == c) v = acc[c];
            if (MODE == 1) v = b[size_t(n) * N + lane] - v;
            if ((m >> lane) & 1u) v = 0.0;
            out[size_t(n) * N + lane] = v;
        }
    }
}

template <int N>
__global__ void __launch_bounds__(kWarpBlock) gsStencilWarpKernel(Grid<N> g, const Real *__restrict__ S, Real *__restrict__ u,
                                                                   const Real *__restrict__ b, const uint8_t *__restrict__ dmask, int colour,
                                                                   int forward, int count) {
    constexpr int NST = ipow3(N), NN = N * N;
    const int lane = threadIdx.x & 31, nwarps = (gridDim.x * kWarpBlock) >> 5;
    StencilLane<N> sl; sl.init(g, lane);
    for (int t = (blockIdx.x * kWarpBlock + threadIdx.x) >> 5; t < count; t += nwarps) {
        int idx[N];
        colourNode<N>(g, t, colour, idx);
        const int n = flattenNode<N>(g, idx);
        // the diagonal block and right-hand side do not depend on the row sums: fetch them before the reduction
        Real M[NN], tt[N];
        unsigned mask = 0;
        if (lane == 0) {
#pragma unroll
            for (int i = 0; i < NN; ++i) M[i] = S[stencilIndex<N>(n, NST / 2, i)];
#pragma unroll
            for (int c = 0; c < N; ++c) tt[c] = b[size_t(n) * N + c];
            mask = dmask[n];
        }
        Real acc[N];
        stencilWarpRow<N>(sl, g.nn * N, S, u, n, lane, acc);
        if (lane == 0) {
#pragma unroll
            for (int c = 0; c < N; ++c) tt[c] -= acc[c];
            gsNodeUpdate<N>(tt, M, mask, forward != 0, u, n);
        }
    }
}

// All colour phases of one or more Gauss-Seidel sweeps in ONE cooperative launch: phase q of the sequence visits colour
// (forward ? q % 2^N : 2^N-1 - q % 2^N) and a grid-wide barrier separates the phases (the per-colour launches of the
// small assembled levels are otherwise pure launch latency).  u is read with ld.global.cg: other SMs rewrite it between
// the phases.  [q0, q1) selects the part of the sequence to run (the slab solver splits a sweep at its halo exchanges).
template <int N>
__global__ void __launch_bounds__(kWarpBlock, 2) gsStencilSweepKernel(Grid<N> g, const Real *__restrict__ S, Real *u, const Real *__restrict__ b,
                                                                    const uint8_t *__restrict__ dmask, int forward, int q0, int q1) {
    constexpr int NST = ipow3(N), NN = N * N, NPE = 1 << N;
    namespace cg = cooperative_groups;
    cg::grid_group grid = cg::this_grid();
    const int lane = threadIdx.x & 31, nwarps = (gridDim.x * kWarpBlock) >> 5;
    StencilLane<N> sl; sl.init(g, lane);
    for (int q = q0; q < q1; ++q) {
        const int colour = forward ? (q % NPE) : (NPE - 1 - q % NPE);
        const int count = colourCount<N>(g, colour);
        // software pipeline over the warp's nodes of this colour: the loads of node t + nwarps are issued before the
        // reduction / substitution of node t (same-colour nodes are not neighbours, so u read ahead inside a phase is current)
        constexpr int ROW = StencilLane<N>::ROW, J = StencilLane<N>::J;
        struct Slot { Real Sv[N][J], uv[J], bv; unsigned mask; int n; }; // bv: lanes 0..N-1 hold b[n][lane]; lane 0 holds the mask
        auto load = [&](int t, Slot &L) {
            int idx[N];
            colourNode<N>(g, t, colour, idx);
            L.n = flattenNode<N>(g, idx);
            const Real *Sn = S + size_t(L.n) * N * ROW + lane;
#pragma unroll
            for (int j = 0; j < J; ++j) {
                const int a = min(max(L.n * N + sl.uoff[j], 0), g.nn * N - 1);
                L.uv[j] = sl.valid[j] ? __ldcg(u + a) : 0.0;
#pragma unroll
                for (int r = 0; r < N; ++r) L.Sv[r][j] = sl.valid[j] ? Sn[r * ROW + 32 * j] : 0.0;
            }
            L.bv = (lane < N) ? b[size_t(L.n) * N + lane] : 0.0;
            L.mask = (lane == 0) ? unsigned(dmask[L.n]) : 0u;
        };
        int t = (blockIdx.x * kWarpBlock + threadIdx.x) >> 5;
        Slot cur;
        if (t < count) load(t, cur);
        while (t < count) {
            const int tn = t + nwarps;
            Slot nxt;
            if (tn < count) load(tn, nxt);
            Real acc[N];
#pragma unroll
            for (int r = 0; r < N; ++r) {
                Real s = 0.0;
#pragma unroll
                for (int j = 0; j < J; ++j)
                    if (sl.valid[j]) s = fma(cur.Sv[r][j], cur.uv[j], s);
                acc[r] = s;
            }
#pragma unroll
            for (int r = 0; r < N; ++r)
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) acc[r] += __shfl_xor_sync(0xffffffffu, acc[r], o);
            // the diagonal block is part of the rows the lanes already hold (entries (NST/2)*N + c of each row)
            Real M[NN], tt[N];
#pragma unroll
            for (int r = 0; r < N; ++r)
#pragma unroll
                for (int c = 0; c < N; ++c) {
                    constexpr int e0 = (NST / 2) * N;
                    const int e = e0 + c;
                    M[r * N + c] = __shfl_sync(0xffffffffu, cur.Sv[r][e / 32], e % 32);
                }
#pragma unroll
            for (int c = 0; c < N; ++c) tt[c] = __shfl_sync(0xffffffffu, cur.bv, c) - acc[c];
            if (lane == 0) {
                Real du[N];
                if (forward) gsNodeSubst<N, true>(tt, M, cur.mask, du); else gsNodeSubst<N, false>(tt, M, cur.mask, du);
#pragma unroll
                for (int c = 0; c < N; ++c) u[size_t(cur.n) * N + c] = __ldcg(u + size_t(cur.n) * N + c) + du[c];
            }
            if (tn < count) cur = nxt;
            t = tn;
        }
        if (q + 1 < q1) grid.sync();
    }
}

} // namespace vfem
