// ndr_b200 — exact solve of the coarsest multigrid level (reference: TensorProductSimulator::solve, TPS.hh:834-865,
// which removes the Dirichlet rows/columns and calls CHOLMOD).
//
// With nodes numbered axis-0-slowest the coarsest operator is block tridiagonal over the node planes of axis 0
// (a node only couples to the planes x-1, x, x+1): P = N * nodesPerPlane dofs per block.  Once per density update the
// host runs a block Cholesky  A = L L^T,  L block lower bidiagonal with diagonal blocks C_p and sub-diagonal blocks
// L_p = A_{p+1,p} C_p^{-T}  (cuSOLVER potrf + cuBLAS trsm/syrk on P x P blocks) and forms W_p = C_p^{-1} explicitly, so that
// every solve is a chain of dense matrix-vector products done by ONE kernel launch (blockTriSolveKernel):
//     forward   y_p = W_p (b_p - L_{p-1} y_{p-1}),      backward   x_p = W_p^T (y_p - L_p^T x_{p+1}).
// Constrained dofs keep an identity row/column and a zero right-hand side, which gives the reduced system's solution
// on the free dofs and 0 on the fixed ones.  All blocks are column-major (cuBLAS convention), leading dimension P.
#pragma once
#include "common.h"

namespace vfem {

// D[p] (diagonal blocks, lower triangle referenced) and Lo[p] = A_{p+1,p} from the assembled stencil; both zeroed before.
template <int N>
static __global__ void __launch_bounds__(128) assembleBlockTriKernel(Grid<N> g, const Real *__restrict__ S, const uint8_t *__restrict__ dmask,
                                                                     Real *__restrict__ D, Real *__restrict__ Lo, int P) {
    constexpr int NST = ipow3(N);
    long long t = (long long)blockIdx.x * 128 + threadIdx.x;
    if (t >= (long long)g.nn * NST) return;
    const int n = int(t / NST), k = int(t % NST);
    int idx[N];
    { int r = n; for (int d = N - 1; d > 0; --d) { int q = r / g.n[d]; idx[d] = r - q * g.n[d]; r = q; } idx[0] = r; }
    int off = 0, dl0 = 0; bool ok = true;
    { int tt = k; for (int d = N - 1; d >= 0; --d) { int dl = tt % 3 - 1; tt /= 3; int v = idx[d] + dl; ok = ok && v >= 0 && v < g.n[d]; off += dl * g.ninc[d]; if (d == 0) dl0 = dl; } }
    if (!ok || dl0 > 0) return; // the upper blocks are the transposes of the lower ones
    const int m = n + off, npp = g.ninc[0];
    const int p = idx[0], rloc = (n - p * npp) * N, cloc = (m - (p + dl0) * npp) * N;
    Real *B = (dl0 == 0) ? D + size_t(p) * P * P : Lo + size_t(p - 1) * P * P;
    const unsigned mn = dmask[n], mm = dmask[m];
#pragma unroll
    for (int a = 0; a < N; ++a)
#pragma unroll
        for (int b = 0; b < N; ++b) {
            Real v = S[stencilIndex<N>(n, k, a * N + b)];
            if (((mn >> a) & 1u) || ((mm >> b) & 1u)) v = (dl0 == 0 && rloc + a == cloc + b) ? 1.0 : 0.0;
            B[size_t(rloc + a) + size_t(cloc + b) * P] = v;
        }
}

// W <- identity (right-hand side of the triangular solve that forms C^{-1})
static __global__ void setIdentityBlocksKernel(Real *__restrict__ W, int P, int nblocks) {
    const size_t tot = size_t(P) * P * nblocks;
    for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < tot; i += size_t(gridDim.x) * blockDim.x) {
        const size_t e = i % (size_t(P) * P);
        W[i] = (e % P == e / P) ? 1.0 : 0.0;
    }
}

constexpr int kTriThreads = 1024;

// One CTA walks the planes; see the header comment.  W: C_p^{-1} (lower triangular, col-major), Lo: L_p.
// dynamic shared memory: (3 + groups) * P doubles.
template <int N>
static __global__ void __launch_bounds__(kTriThreads) blockTriSolveKernel(int P, int planes, const Real *__restrict__ W, const Real *__restrict__ Lo,
                                                                          const Real *__restrict__ b, const uint8_t *__restrict__ dmask,
                                                                          Real *__restrict__ x, int groups) {
    extern __shared__ Real sm[];
    Real *cur = sm, *prev = sm + P, *tmp = sm + 2 * P, *part = sm + 3 * P; // part[groups][P]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = kTriThreads / 32;
    const int rowsPerPass = kTriThreads / groups; // threads per column group
    const int gq = tid / rowsPerPass, r0 = tid - gq * rowsPerPass;
    const size_t PP = size_t(P) * P;
    // ---- forward: y_p = W_p (b_p - L_{p-1} y_{p-1}) ----
    for (int p = 0; p < planes; ++p) {
        const Real *Lp = (p > 0) ? Lo + size_t(p - 1) * PP : Lo, *Wp = W + size_t(p) * PP;
        // tmp = b_p (masked) - L_{p-1} prev : column groups -> partial sums, fixed-order reduction
        if (p > 0 && gq < groups) {
            for (int i = r0; i < P; i += rowsPerPass) {
                Real s = 0.0;
                for (int j = gq; j < P; j += groups) s = fma(Lp[size_t(i) + size_t(j) * P], prev[j], s);
                part[gq * P + i] = s;
            }
        }
        __syncthreads();
        for (int i = tid; i < P; i += kTriThreads) {
            const int dof = p * P + i;
            Real v = ((dmask[dof / N] >> (dof % N)) & 1u) ? 0.0 : b[dof];
            if (p > 0) for (int q = 0; q < groups; ++q) v -= part[q * P + i];
            tmp[i] = v;
        }
        __syncthreads();
        if (gq < groups) {
            for (int i = r0; i < P; i += rowsPerPass) {
                Real s = 0.0;
                for (int j = gq; j <= i; j += groups) s = fma(Wp[size_t(i) + size_t(j) * P], tmp[j], s);
                part[gq * P + i] = s;
            }
        }
        __syncthreads();
        for (int i = tid; i < P; i += kTriThreads) {
            Real v = 0.0;
            for (int q = 0; q < groups; ++q) v += part[q * P + i];
            cur[i] = v; x[p * P + i] = v; // y_p parked in x
        }
        __syncthreads();
        Real *sw = cur; cur = prev; prev = sw; // prev = y_p
    }
    // ---- backward: x_p = W_p^T (y_p - L_p^T x_{p+1});  prev holds x_{p+1} ----
    for (int p = planes - 1; p >= 0; --p) {
        const Real *Lp = Lo + size_t(p) * PP, *Wp = W + size_t(p) * PP;
        for (int j = warp; j < P; j += nwarps) { // tmp_j = y_j - column_j(L_p) . x_{p+1}
            Real s = 0.0;
            if (p < planes - 1) for (int i = lane; i < P; i += 32) s = fma(Lp[size_t(i) + size_t(j) * P], prev[i], s);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            if (lane == 0) tmp[j] = x[p * P + j] - s;
        }
        __syncthreads();
        for (int j = warp; j < P; j += nwarps) { // x_j = column_j(W_p)[j:] . tmp[j:]
            Real s = 0.0;
            for (int i = j + lane; i < P; i += 32) s = fma(Wp[size_t(i) + size_t(j) * P], tmp[i], s);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            if (lane == 0) {
                const int dof = p * P + j;
                const Real v = ((dmask[dof / N] >> (dof % N)) & 1u) ? 0.0 : s;
                cur[j] = v; x[dof] = v;
            }
        }
        __syncthreads();
        Real *sw = cur; cur = prev; prev = sw; // prev = x_p
    }
}

} // namespace vfem
