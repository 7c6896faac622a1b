// ndr_b200 — exact solve of the coarsest multigrid level (reference: TensorProductSimulator::solve, TPS.hh:834-865,
// which removes the Dirichlet rows/columns and calls CHOLMOD).
//
// With nodes numbered axis-0-slowest the coarsest operator is block tridiagonal over the node planes of axis 0
// (a node only couples to the planes x-1, x, x+1): P = N * nodesPerPlane dofs per block.  Once per density update the
// host runs the block LDL^T recurrence  S_0 = A_00,  T_p = A_{p,p-1} S_{p-1}^{-1},  S_p = A_pp - T_p A_{p,p-1}^T
// (cuSOLVER potrf + cuBLAS trsm/gemm on P x P blocks) and stores S_p^{-1}, T_p explicitly, so that a solve is
//     forward    z_p = b_p - T_p z_{p-1}            (a chain of planes-1 dependent matrix-vector products)
//     diagonal   v_p = S_p^{-1} z_p                 (independent over the planes)
//     backward   x_p = v_p - T_{p+1}^T x_{p+1}      (a chain again)
// done by ONE launch of ONE thread-block cluster (blockTriChainKernel): the rows of each product are spread over the
// warps of the cluster, every CTA keeps the current plane vector in its shared memory, and a finished entry is written
// into all CTAs' copies through distributed shared memory, so a chain step costs one cluster barrier instead of a
// kernel launch or a grid-wide barrier.  Constrained dofs keep an identity row/column and a zero right-hand side,
// which gives the reduced system's solution on the free dofs and 0 on the fixed ones.
// Blocks handed to cuBLAS are column-major with leading dimension P; Trm/Sinv are read row-wise, Tcm column-wise.
#pragma once
#include "common.h"
#include <cooperative_groups.h>

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

constexpr int kChainThreads = 512; // per CTA
constexpr int kChainCluster = 8;   // portable cluster size

// One cluster walks the planes; see the header comment.  Trm[p-1]: T_p row-major, Tcm[p-1]: T_p column-major (p = 1..planes-1),
// Sinv[p]: S_p^{-1}.  x doubles as the scratch for v; z: planes*P scratch.  Dynamic shared memory: 2*P doubles.
// Requires P <= 256 (at most 2 rows per warp, 8 columns per lane): the matrix rows of the NEXT chain step are fetched into
// registers before the cluster barrier of the current one, so the L2/HBM latency of the (vector-independent) matrix
// overlaps the barrier.  Larger planes use blockTriChainGridKernel below.
template <int N>
static __global__ void __launch_bounds__(kChainThreads) blockTriChainKernel(int P, int planes, const Real *__restrict__ Trm, const Real *__restrict__ Tcm,
                                                                            const Real *__restrict__ Sinv, const Real *__restrict__ b,
                                                                            const uint8_t *__restrict__ dmask, Real *x, Real *z) {
    namespace cg = cooperative_groups;
    cg::cluster_group cl = cg::this_cluster();
    extern __shared__ Real sm[];
    Real *cur = sm, *nxt = sm + P;
    const int tid = threadIdx.x, lane = tid & 31, nw = kChainThreads / 32;
    const int ncta = int(cl.num_blocks()), rank = int(cl.block_rank());
    const int gw = rank * nw + (tid >> 5), tw = ncta * nw;
    const size_t PP = size_t(P) * P;
    auto masked = [&](int dof, Real v) { return ((dmask[dof / N] >> (dof % N)) & 1u) ? 0.0 : v; };
    // every entry is published to all CTAs' copies of the next plane vector (lanes 0..ncta-1 each serve one CTA)
    auto publish = [&](int i, Real v) {
        if (lane < ncta) cl.map_shared_rank(nxt, lane)[i] = v;
    };
    constexpr int JP = 8;
    Real t0[JP], t1[JP];
    const bool has0 = gw < P, has1 = gw + tw < P;
    auto fetch = [&](const Real *T) { // rows gw and gw + tw of T
#pragma unroll
        for (int j = 0; j < JP; ++j) {
            const int c = lane + 32 * j;
            t0[j] = (has0 && c < P) ? T[size_t(gw) * P + c] : 0.0;
            t1[j] = (has1 && c < P) ? T[size_t(gw + tw) * P + c] : 0.0;
        }
    };
    auto dots = [&](Real &s0, Real &s1) {
        Real a0 = 0.0, a1 = 0.0;
#pragma unroll
        for (int j = 0; j < JP; ++j) {
            const int c = lane + 32 * j;
            const Real vj = c < P ? cur[c] : 0.0;
            a0 = fma(t0[j], vj, a0); a1 = fma(t1[j], vj, a1);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { a0 += __shfl_xor_sync(0xffffffffu, a0, o); a1 += __shfl_xor_sync(0xffffffffu, a1, o); }
        s0 = a0; s1 = a1;
    };
    // one chain step: out_p[i] = masked(rhs[p*P+i] - row_i(T) . cur), published to every CTA and stored to `store`
    auto chainStep = [&](const Real *rhs, Real *store, int p, const Real *Tnext) {
        Real s0, s1;
        dots(s0, s1);
        if (has0) { const Real v = masked(p * P + gw, rhs[p * P + gw] - s0); publish(gw, v); if (lane == 0) store[p * P + gw] = v; }
        if (has1) { const Real v = masked(p * P + gw + tw, rhs[p * P + gw + tw] - s1); publish(gw + tw, v); if (lane == 0) store[p * P + gw + tw] = v; }
        if (Tnext) fetch(Tnext);
        cl.sync();
        Real *sw = cur; cur = nxt; nxt = sw;
    };
    // ---- forward: z_0 = b_0,  z_p = b_p - T_p z_{p-1} ----
    if (planes > 1) fetch(Trm);
    for (int i = tid; i < P; i += kChainThreads) { const Real v = masked(i, b[i]); cur[i] = v; if (rank == 0) z[i] = v; }
    cl.sync();
    for (int p = 1; p < planes; ++p)
        chainStep(b, z, p, p + 1 < planes ? Trm + size_t(p) * PP : nullptr);
    __threadfence();
    cl.sync();
    // ---- diagonal: v_p = S_p^{-1} z_p, all planes at once (v parked in x) ----
    if (planes > 1) fetch(Tcm + size_t(planes - 2) * PP);
    for (int row = gw; row < planes * P; row += tw) {
        const int p = row / P, i = row - p * P;
        const Real *Sr = Sinv + size_t(p) * PP + size_t(i) * P, *zp = z + size_t(p) * P;
        Real a = 0.0;
        for (int j = lane; j < P; j += 32) a = fma(Sr[j], zp[j], a);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
        if (lane == 0) x[row] = masked(row, a);
    }
    __threadfence();
    cl.sync();
    // ---- backward: x_last = v_last,  x_p = v_p - T_{p+1}^T x_{p+1}  (column i of T_{p+1} is contiguous in Tcm) ----
    for (int i = tid; i < P; i += kChainThreads) cur[i] = x[size_t(planes - 1) * P + i];
    __syncthreads();
    for (int p = planes - 2; p >= 0; --p)
        chainStep(x, x, p, p > 0 ? Tcm + size_t(p - 1) * PP : nullptr);
}

// Large planes (P > 256): the same three phases on the WHOLE device.  One cooperative launch; a warp owns a matrix row,
// the plane vectors live in global memory and are staged into every CTA's shared memory at the start of a step (read
// with ld.global.cg: other CTAs wrote them in the previous step); a grid barrier per chain step.  A lane issues all its
// loads of a row (up to 32) before the first multiply, so a step costs about one memory round trip, and the rows of the
// next step are prefetched into L2 before the barrier.  Dynamic shared memory: P doubles.
constexpr int kChainGridThreads = 256;
template <int N>
static __global__ void __launch_bounds__(kChainGridThreads) blockTriChainGridKernel(int P, int planes, const Real *__restrict__ Trm,
                                                                                const Real *__restrict__ Tcm, const Real *__restrict__ Sinv,
                                                                                const Real *__restrict__ b, const uint8_t *__restrict__ dmask,
                                                                                Real *x, Real *z) {
    namespace cg = cooperative_groups;
    cg::grid_group grid = cg::this_grid();
    extern __shared__ Real sv[];
    const int lane = threadIdx.x & 31;
    const int gw = (blockIdx.x * kChainGridThreads + threadIdx.x) >> 5, tw = (gridDim.x * kChainGridThreads) >> 5;
    const size_t PP = size_t(P) * P;
    auto masked = [&](int dof, Real v) { return ((dmask[dof / N] >> (dof % N)) & 1u) ? 0.0 : v; };
    auto stage = [&](const Real *v) { // plane vector -> shared memory (block-wide)
        __syncthreads();
        for (int i = threadIdx.x; i < P; i += kChainGridThreads) sv[i] = __ldcg(v + i);
        __syncthreads();
    };
    auto rowDot = [&](const Real *__restrict__ row) { // row . sv, all lanes return the sum
        Real a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
        for (int base = 0; base < P; base += 1024) {
            Real rv[32];
#pragma unroll
            for (int k = 0; k < 32; ++k) { const int j = base + lane + 32 * k; rv[k] = j < P ? row[j] : 0.0; }
#pragma unroll
            for (int k = 0; k < 32; k += 4) {
                const int j = base + lane + 32 * k;
                a0 = fma(rv[k], j < P ? sv[j] : 0.0, a0); a1 = fma(rv[k + 1], j + 32 < P ? sv[j + 32] : 0.0, a1);
                a2 = fma(rv[k + 2], j + 64 < P ? sv[j + 64] : 0.0, a2); a3 = fma(rv[k + 3], j + 96 < P ? sv[j + 96] : 0.0, a3);
            }
        }
        Real a = (a0 + a1) + (a2 + a3);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
        return a;
    };
    // the matrices do not depend on the chain: pull the rows of the NEXT step into L2 before waiting at the barrier
    auto prefetchRows = [&](const Real *T) {
        for (int i = gw; i < P; i += tw)
            for (int j = lane * 16; j < P; j += 512) asm volatile("prefetch.global.L2 [%0];" ::"l"(T + size_t(i) * P + j));
    };
    for (int i = blockIdx.x * kChainGridThreads + threadIdx.x; i < P; i += gridDim.x * kChainGridThreads) z[i] = masked(i, b[i]);
    if (planes > 1) prefetchRows(Trm);
    grid.sync();
    for (int p = 1; p < planes; ++p) { // forward
        stage(z + size_t(p - 1) * P);
        for (int i = gw; i < P; i += tw) {
            const Real s = rowDot(Trm + size_t(p - 1) * PP + size_t(i) * P);
            if (lane == 0) z[p * P + i] = masked(p * P + i, b[p * P + i] - s);
        }
        if (p + 1 < planes) prefetchRows(Trm + size_t(p) * PP);
        grid.sync();
    }
    for (int p = 0; p < planes; ++p) { // diagonal: no dependency between the planes, only block-level staging
        stage(z + size_t(p) * P);
        for (int i = gw; i < P; i += tw) {
            const Real s = rowDot(Sinv + size_t(p) * PP + size_t(i) * P);
            if (lane == 0) x[p * P + i] = masked(p * P + i, s);
        }
    }
    if (planes > 1) prefetchRows(Tcm + size_t(planes - 2) * PP);
    grid.sync();
    for (int p = planes - 2; p >= 0; --p) { // backward
        stage(x + size_t(p + 1) * P);
        for (int i = gw; i < P; i += tw) {
            const Real s = rowDot(Tcm + size_t(p) * PP + size_t(i) * P);
            if (lane == 0) x[p * P + i] = masked(p * P + i, __ldcg(x + p * P + i) - s);
        }
        if (p > 0) prefetchRows(Tcm + size_t(p - 1) * PP);
        grid.sync();
    }
}

} // namespace vfem
