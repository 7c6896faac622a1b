// ndr_b200 — multigrid transfer operators, Galerkin coarsening and the coarsest-level dense system.
#pragma once
#include "kernels.cuh"

namespace vfem {

// ---- prolongation (MG.hh:115-141): x_f (+)= P x_c, trilinear weights = products of {1, 1/2} -----
template <int N, bool ACCUM>
static __global__ void __launch_bounds__(kBlock) prolongKernel(Grid<N> gf, Grid<N> gc, const Real *__restrict__ xc, Real *__restrict__ xf) {
    constexpr int NPE = 1 << N;
    int n = blockIdx.x * kBlock + threadIdx.x;
    if (n >= gf.nn) return;
    int fi[N]; unflattenNode<N>(gf, n, fi);
    Real acc[N];
#pragma unroll
    for (int c = 0; c < N; ++c) acc[c] = 0.0;
#pragma unroll
    for (int j = 0; j < NPE; ++j) {
        Real w = 1.0; int cn = 0; bool skip = false;
#pragma unroll
        for (int d = 0; d < N; ++d) {
            int bit = (j >> (N - 1 - d)) & 1, lo = fi[d] >> 1, odd = fi[d] & 1;
            if (odd) { w *= 0.5; cn += (lo + bit) * gc.ninc[d]; }
            else { skip = skip || bit; cn += lo * gc.ninc[d]; }
        }
        if (skip) continue;
#pragma unroll
        for (int c = 0; c < N; ++c) acc[c] = fma(w, xc[size_t(cn) * N + c], acc[c]);
    }
#pragma unroll
    for (int c = 0; c < N; ++c) {
        if (ACCUM) xf[size_t(n) * N + c] += acc[c];
        else xf[size_t(n) * N + c] = acc[c];
    }
}

// ---- restriction (MG.hh:145-180): b_c = P^T v_f, gathered over the 3^N fine nodes around a coarse node
template <int N>
static __global__ void __launch_bounds__(kBlock) restrictKernel(Grid<N> gf, Grid<N> gc, const Real *__restrict__ vf, Real *__restrict__ bc) {
    constexpr int NST = ipow3(N);
    int n = blockIdx.x * kBlock + threadIdx.x;
    if (n >= gc.nn) return;
    int ci[N]; unflattenNode<N>(gc, n, ci);
    Real acc[N];
#pragma unroll
    for (int c = 0; c < N; ++c) acc[c] = 0.0;
#pragma unroll
    for (int k = 0; k < NST; ++k) {
        int t = k, fn = 0; Real w = 1.0; bool ok = true;
#pragma unroll
        for (int d = N - 1; d >= 0; --d) {
            int dl = t % 3 - 1; t /= 3;
            int v = 2 * ci[d] + dl; ok = ok && (v >= 0) && (v < gf.n[d]);
            fn += v * gf.ninc[d];
            if (dl != 0) w *= 0.5;
        }
        if (!ok) continue;
#pragma unroll
        for (int c = 0; c < N; ++c) acc[c] = fma(w, vf[size_t(fn) * N + c], acc[c]);
    }
#pragma unroll
    for (int c = 0; c < N; ++c) bc[size_t(n) * N + c] = acc[c];
}

// ---- level-1 operator straight from the fine moduli (MG.hh:644-656 + TPS.hh:649-720) --------------
// Ke_c = sum_{children fi} E(rho_child) * C_fi with the 2^N constant matrices C_fi = I_fi^T K0 I_fi
// (child bit d <-> axis d, MG.hh:589-600), scattered into the block stencil.  One thread per
// (coarse node, stencil slot) gathers its block; C is [fi][KS*KS] row-major in global memory.
template <int N>
static __global__ void __launch_bounds__(kBlock) buildLevel1StencilKernel(Grid<N> gf, Grid<N> gc, const Real *__restrict__ Ey,
                                                                    const Real *__restrict__ C, Real *__restrict__ S) {
    constexpr int NPE = 1 << N, KS = N * NPE, NST = ipow3(N);
    long long t = (long long)blockIdx.x * kBlock + threadIdx.x;
    if (t >= (long long)gc.nn * NST) return;
    const int n = int(t / NST), k = int(t % NST);
    int ci[N]; unflattenNode<N>(gc, n, ci);
    int dl[N];
    { int tt = k; for (int d = N - 1; d >= 0; --d) { dl[d] = tt % 3 - 1; tt /= 3; } }
    Real blk[N * N];
#pragma unroll
    for (int i = 0; i < N * N; ++i) blk[i] = 0.0;
    for (int s = 0; s < NPE; ++s) {
        int e[N], locI = 0, locJ = 0; bool ok = true;
#pragma unroll
        for (int d = 0; d < N; ++d) {
            int sd = (s >> (N - 1 - d)) & 1;
            e[d] = ci[d] - 1 + sd; ok = ok && (e[d] >= 0) && (e[d] < gc.E[d]);
            int li = 1 - sd, lj = li + dl[d]; ok = ok && (lj >= 0) && (lj <= 1);
            locI |= li << (N - 1 - d); locJ |= lj << (N - 1 - d);
        }
        if (!ok) continue;
        for (int fi = 0; fi < NPE; ++fi) {
            int ef = 0;
#pragma unroll
            for (int d = 0; d < N; ++d) ef += (2 * e[d] + ((fi >> d) & 1)) * gf.einc[d];
            const Real Ee = Ey[ef];
            const Real *Cb = C + size_t(fi) * KS * KS + (locI * N) * KS + locJ * N;
#pragma unroll
            for (int a = 0; a < N; ++a)
#pragma unroll
                for (int b = 0; b < N; ++b) blk[a * N + b] = fma(Ee, Cb[a * KS + b], blk[a * N + b]);
        }
    }
#pragma unroll
    for (int i = 0; i < N * N; ++i) S[stencilIndex<N>(n, k, i)] = blk[i];
}

// ---- Galerkin coarsening of a block stencil: S_c = P^T S_f P (MG.hh:659-665 assembled) ------------
template <int N>
static __global__ void __launch_bounds__(kBlock) galerkinStencilKernel(Grid<N> gf, Grid<N> gc, const Real *__restrict__ Sf, Real *__restrict__ Sc) {
    constexpr int NST = ipow3(N);
    long long t = (long long)blockIdx.x * kBlock + threadIdx.x;
    if (t >= (long long)gc.nn * NST) return;
    const int n = int(t / NST), k = int(t % NST);
    int ci[N]; unflattenNode<N>(gc, n, ci);
    int D[N]; bool Jok = true;
    { int tt = k; for (int d = N - 1; d >= 0; --d) { D[d] = tt % 3 - 1; tt /= 3; int v = ci[d] + D[d]; Jok = Jok && v >= 0 && v < gc.n[d]; } }
    Real blk[N * N];
#pragma unroll
    for (int i = 0; i < N * N; ++i) blk[i] = 0.0;
    if (Jok) {
        for (int ki = 0; ki < NST; ++ki) {
            int di[N], fn = 0; Real wi = 1.0; bool ok = true;
            { int tt = ki; for (int d = N - 1; d >= 0; --d) { di[d] = tt % 3 - 1; tt /= 3; } }
#pragma unroll
            for (int d = 0; d < N; ++d) {
                int v = 2 * ci[d] + di[d]; ok = ok && v >= 0 && v < gf.n[d];
                fn += v * gf.ninc[d]; if (di[d]) wi *= 0.5;
            }
            if (!ok) continue;
            for (int kj = 0; kj < NST; ++kj) {
                int dj[N]; Real w = wi; bool okj = true;
                { int tt = kj; for (int d = N - 1; d >= 0; --d) { dj[d] = tt % 3 - 1; tt /= 3; } }
#pragma unroll
                for (int d = 0; d < N; ++d) {
                    int o = di[d] + dj[d] - 2 * D[d]; okj = okj && o >= -1 && o <= 1;
                    int v = 2 * ci[d] + di[d] + dj[d]; okj = okj && v >= 0 && v < gf.n[d];
                    if (o) w *= 0.5;
                }
                if (!okj) continue;
#pragma unroll
                for (int i = 0; i < N * N; ++i) blk[i] = fma(w, Sf[stencilIndex<N>(fn, kj, i)], blk[i]);
            }
        }
    }
#pragma unroll
    for (int i = 0; i < N * N; ++i) Sc[stencilIndex<N>(n, k, i)] = blk[i];
}

// Same product, one CTA per coarse node: the 3^N fine rows around 2 ci (3^N x N x 3^N x N doubles, 52 KB in 3-D) are staged
// in shared memory with coalesced loads - every fine row is read by at most 2^N CTAs instead of by up to 27 x 27 threads in
// scattered 24-byte pieces - and each warp adds up whole output blocks over the admissible (fine row, fine column) pairs
// only: per axis the pairs (di, dj) with di + dj - 2 D in [-1, 1], 7 for D = 0 and 3 for D = +-1, weight
// (di ? 1/2 : 1) (o ? 1/2 : 1).  Out-of-grid fine rows are staged as zeros; out-of-grid fine columns are
// zero entries of S_f already (every kernel that writes a stencil writes all of its slots).  Measured (gpurun r2m): operator
// refresh 2.82 ms against 3.42 ms with the thread-per-block form above at 128x64x64, 25.2 against 27.4 ms at 256x128x128.
// The admissible pairs of every output slot k are tabulated once on the host (GalerkinTable): start[k] .. start[k+1] index
// (offset of the fine block in the staged rows, weight); 13^N entries in all.
template <int N>
struct GalerkinTable {
    static constexpr int NST = ipow3(N), ROWLEN = N * NST * N;
    std::vector<int> start, off;
    std::vector<Real> w;
    GalerkinTable() {
        const int di[13] = {0, -1, -1, -1, -1, 0, 0, 0, 1, 1, 0, 1, 1}, dj[13] = {-1, 0, -1, 0, 1, -1, 0, 1, -1, 0, 1, 0, 1}; // 0..2: D = -1, 3..9: D = 0, 10..12: D = +1
        start.push_back(0);
        for (int k = 0; k < NST; ++k) {
            int D[N], cnt[N], base[N], total = 1;
            { int tt = k; for (int d = N - 1; d >= 0; --d) { D[d] = tt % 3 - 1; tt /= 3; cnt[d] = D[d] == 0 ? 7 : 3; base[d] = D[d] == 0 ? 3 : (D[d] > 0 ? 10 : 0); total *= cnt[d]; } }
            for (int q = 0; q < total; ++q) {
                int rem = q, ki = 0, kj = 0, p3 = 1; Real wt = 1.0;
                for (int d = N - 1; d >= 0; --d) { // axis N-1 is the least significant digit of a stencil slot
                    const int x = base[d] + rem % cnt[d]; rem /= cnt[d];
                    const int o = di[x] + dj[x] - 2 * D[d];
                    ki += (di[x] + 1) * p3; kj += (dj[x] + 1) * p3; p3 *= 3; wt *= (di[x] ? 0.5 : 1.0) * (o ? 0.5 : 1.0);
                }
                off.push_back(ki * ROWLEN + kj * N); w.push_back(wt);
            }
            start.push_back(int(off.size()));
        }
    }
};
template <int N>
static __global__ void __launch_bounds__(256) galerkinStencilCtaKernel(Grid<N> gf, Grid<N> gc, const Real *__restrict__ Sf, Real *__restrict__ Sc,
                                                                      const int *__restrict__ tStart, const int *__restrict__ tOff,
                                                                      const Real *__restrict__ tW) {
    constexpr int NST = ipow3(N), ROWLEN = N * NST * N, NN = N * N; // ROWLEN: doubles per node
    extern __shared__ __align__(16) Real galSm[];                    // [NST][ROWLEN]
    const int n = blockIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    int ci[N]; unflattenNode<N>(gc, n, ci);
    for (int ki = warp; ki < NST; ki += nwarps) { // a warp stages whole fine rows: the node arithmetic is per row, not per double
        int tt = ki, fn = 0; bool ok = true;
#pragma unroll
        for (int d = N - 1; d >= 0; --d) { const int di = tt % 3 - 1; tt /= 3; const int v = 2 * ci[d] + di; ok = ok && v >= 0 && v < gf.n[d]; fn += v * gf.ninc[d]; }
        const Real *src = Sf + size_t(ok ? fn : 0) * ROWLEN;
        for (int r = lane; r < ROWLEN; r += 32) galSm[ki * ROWLEN + r] = ok ? src[r] : 0.0;
    }
    __syncthreads();
    // warp w forms the output blocks k = w, w + 8, ...; its lanes share the admissible pairs of that block (3^a 7^b of them,
    // a + b = N), each lane accumulating all N x N entries, and a shuffle butterfly adds the lanes up (fixed order)
    for (int k = warp; k < NST; k += nwarps) {
        bool Jok = true;
        { int tt = k;
#pragma unroll
          for (int d = N - 1; d >= 0; --d) { const int D = tt % 3 - 1; tt /= 3; const int v = ci[d] + D; Jok = Jok && v >= 0 && v < gc.n[d]; } }
        Real acc[NN];
#pragma unroll
        for (int i = 0; i < NN; ++i) acc[i] = 0.0;
        if (Jok) {
            const int q1 = tStart[k + 1];
            for (int q = tStart[k] + lane; q < q1; q += 32) {
                const Real w = tW[q];
                const Real *blk = galSm + tOff[q];
#pragma unroll
                for (int a = 0; a < N; ++a)
#pragma unroll
                    for (int b = 0; b < N; ++b) acc[a * N + b] = fma(w, blk[a * NST * N + b], acc[a * N + b]);
            }
        }
#pragma unroll
        for (int i = 0; i < NN; ++i)
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) acc[i] += __shfl_xor_sync(0xffffffffu, acc[i], o);
        Real mine = 0.0;
#pragma unroll
        for (int i = 0; i < NN; ++i) if (lane == i) mine = acc[i];
        if (lane < NN) Sc[size_t(n) * ROWLEN + ((lane / N) * NST + k) * N + lane % N] = mine;
    }
}

// ---- coarsest level: dense Dirichlet-reduced matrix (TPS.hh:589-625,849: rows/cols of fixed dofs
// removed; here kept as identity rows/cols, which yields the same solution on the free dofs) -----------
template <int N>
static __global__ void __launch_bounds__(kBlock) assembleDenseKernel(Grid<N> g, const Real *__restrict__ S, const uint8_t *__restrict__ dmask,
                                                               Real *__restrict__ A) {
    constexpr int NST = ipow3(N);
    long long t = (long long)blockIdx.x * kBlock + threadIdx.x;
    if (t >= (long long)g.nn * NST) return;
    const int n = int(t / NST), k = int(t % NST);
    int idx[N]; unflattenNode<N>(g, n, idx);
    int off = 0; bool ok = true;
    { int tt = k; for (int d = N - 1; d >= 0; --d) { int dl = tt % 3 - 1; tt /= 3; int v = idx[d] + dl; ok = ok && v >= 0 && v < g.n[d]; off += dl * g.ninc[d]; } }
    if (!ok) return;
    const int m = n + off;
    const size_t nd = size_t(g.nn) * N;
    const unsigned mn = dmask[n], mm = dmask[m];
#pragma unroll
    for (int a = 0; a < N; ++a)
#pragma unroll
        for (int b = 0; b < N; ++b) {
            size_t row = size_t(n) * N + a, col = size_t(m) * N + b;
            Real v = S[stencilIndex<N>(n, k, a * N + b)];
            if (((mn >> a) & 1u) || ((mm >> b) & 1u)) v = (row == col) ? 1.0 : 0.0;
            A[row * nd + col] = v;
        }
}

// copy the triangle cuSOLVER wrote (column-major lower == row-major upper) onto the other one
static __global__ void symmetrizeKernel(Real *__restrict__ A, int nd) {
    int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y * blockDim.y + threadIdx.y;
    if (i < nd && j < nd && j < i) A[size_t(i) * nd + j] = A[size_t(j) * nd + i];
}

// x = mask(Ainv * mask(b)): one warp per row, fixed summation order (deterministic)
template <int N>
static __global__ void denseSolveKernel(const Real *__restrict__ Ainv, const Real *__restrict__ b, const uint8_t *__restrict__ dmask,
                                 Real *__restrict__ x, int nd) {
    int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (row >= nd) return;
    const Real *Ar = Ainv + size_t(row) * nd;
    Real s = 0.0;
    for (int j = lane; j < nd; j += 32) {
        Real bj = ((dmask[j / N] >> (j % N)) & 1u) ? 0.0 : b[j];
        s = fma(Ar[j], bj, s);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) x[row] = ((dmask[row / N] >> (row % N)) & 1u) ? 0.0 : s;
}

} // namespace vfem
