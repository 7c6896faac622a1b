// ndr_b200 — exact solve of the coarsest multigrid level (reference: TensorProductSimulator::solve, TPS.hh:834-865,
// which removes the Dirichlet rows/columns and calls CHOLMOD).
//
// With nodes numbered axis-0-slowest the coarsest operator is block tridiagonal over the node planes of axis 0
// (a node only couples to the planes x-1, x, x+1): P = N * nodesPerPlane dofs per block.  Once per density update the
// host runs the block LDL^T recurrence  S_0 = A_00,  T_p = A_{p,p-1} S_{p-1}^{-1},  S_p = A_pp - T_p A_{p,p-1}^T
// (cuSOLVER potrf + cuBLAS trsm/gemm on P x P blocks) and stores S_p^{-1}, T_p explicitly, so that a solve is
//     forward    z_p = b_p - T_p z_{p-1}            (a chain of dependent matrix-vector products)
//     diagonal   v_p = S_p^{-1} z_p                 (independent over the planes)
//     backward   x_p = v_p - T_{p+1}^T x_{p+1}      (a chain again).
// Both the factorisation and the solve run the recurrence from BOTH ends of the plane sequence towards the middle
// plane (two independent chains, see ChainIdx), which halves the number of dependent steps.
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
constexpr int kChainSlots = 4;     // matrix rows per warp and step at most: 2 chains x P rows <= kChainSlots * 128 warps (8 CTAs);
                                   // the kernel is instantiated with SLOTS = 2 as well: a 16-CTA cluster has a warp per 2 rows up to P = 256, and
                                   // 4 x 8 prefetched doubles per lane do not fit the 128 registers of a 512-thread CTA (248 bytes of spills)

// Index helpers of the two-ended ("burn at both ends") elimination shared by both kernels and the host factorisation.
// Planes 0..mid-1 are eliminated downwards (chain 0), planes-1..mid+1 upwards (chain 1), plane `mid` last.
// Block q of Trm / Tcm (row- / column-major) holds  T_{q+1} = A_{q+1,q} S_q^{-1}   for q <  mid  (chain 0)
//                                              and  U_q     = A_{q+1,q}^T S_{q+1}^{-1} for q >= mid (chain 1).
struct ChainIdx {
    int planes, mid;
    __host__ __device__ int len(int c) const { return c == 0 ? mid : planes - 1 - mid; }            // planes of chain c
    __host__ __device__ int plane(int c, int s) const { return c == 0 ? s : planes - 1 - s; }          // s-th plane of chain c
    __host__ __device__ int fwdBlock(int c, int s) const { return c == 0 ? s - 1 : planes - 1 - s; }   // z_s = b_s - M z_{s-1}
    __host__ __device__ int midBlock(int c) const { return c == 0 ? mid - 1 : mid; }
    __host__ __device__ int bwdPlane(int c, int s) const { return c == 0 ? mid - s : mid + s; }        // s = 1..len(c)
    __host__ __device__ int bwdBlock(int c, int s) const { return c == 0 ? mid - s : mid + s - 1; }
    __host__ __device__ int steps() const { int a = len(0), b2 = len(1); return a > b2 ? a : b2; }
};

// One cluster walks the planes; see the header comment.  x doubles as the scratch for v; z: planes*P scratch.
// Dynamic shared memory: 4*P doubles (current / next vector of both chains).  Requires 2 P <= 512 rows per step.
// The matrix rows of the NEXT step are fetched into registers before the cluster barrier of the current one, so the
// L2/HBM latency of the (vector-independent) matrices overlaps the barrier.  Larger planes: blockTriChainGridKernel.
template <int N, int SLOTS>
static __global__ void __launch_bounds__(kChainThreads) blockTriChainKernel(int P, ChainIdx ci, const Real *__restrict__ Trm, const Real *__restrict__ Tcm,
                                                                            const Real *__restrict__ Sinv, const Real *__restrict__ b,
                                                                            const uint8_t *__restrict__ dmask, Real *x, Real *z) {
    namespace cg = cooperative_groups;
    cg::cluster_group cl = cg::this_cluster();
    extern __shared__ Real sm[];
    Real *cur = sm, *nxt = sm + 2 * P; // [chain][P]
    const int tid = threadIdx.x, lane = tid & 31, nw = kChainThreads / 32;
    const int ncta = int(cl.num_blocks()), rank = int(cl.block_rank());
    const int gw = rank * nw + (tid >> 5), tw = ncta * nw;
    const int planes = ci.planes, mid = ci.mid;
    const size_t PP = size_t(P) * P;
    auto masked = [&](int dof, Real v) { return ((dmask[dof / N] >> (dof % N)) & 1u) ? 0.0 : v; };
    // every entry is published to all CTAs' copies of the next vectors (lanes 0..ncta-1 each serve one CTA)
    auto publish = [&](int e, Real v) {
        if (lane < ncta) cl.map_shared_rank(nxt, lane)[e] = v;
    };
    constexpr int JP = 8;
    Real t[SLOTS][JP];
    int sc[SLOTS], si[SLOTS]; // chain and row of this warp's slots (-1: none)
#pragma unroll
    for (int k = 0; k < SLOTS; ++k) { const int r = gw + k * tw; sc[k] = r < 2 * P ? r / P : -1; si[k] = r < 2 * P ? r - (r / P) * P : 0; }
    // phase 0: forward step s, 1: middle, 2: backward step s  ->  matrix block of chain c, or -1 when the chain is idle
    auto blockOf = [&](int phase, int s, int c) {
        if (phase == 0) return s <= ci.len(c) - 1 ? ci.fwdBlock(c, s) : -1;
        if (phase == 1) return ci.len(c) >= 1 ? ci.midBlock(c) : -1;
        return s <= ci.len(c) ? ci.bwdBlock(c, s) : -1;
    };
    auto fetch = [&](int phase, int s) {
#pragma unroll
        for (int k = 0; k < SLOTS; ++k) {
            const int q = sc[k] >= 0 ? blockOf(phase, s, sc[k]) : -1;
            const Real *row = (phase == 2 ? Tcm : Trm) + size_t(q < 0 ? 0 : q) * PP + size_t(si[k]) * P;
#pragma unroll
            for (int j = 0; j < JP; ++j) { const int c = lane + 32 * j; t[k][j] = (q >= 0 && c < P) ? row[c] : 0.0; }
        }
    };
    auto dots = [&](Real (&s)[SLOTS]) {
#pragma unroll
        for (int k = 0; k < SLOTS; ++k) {
            const Real *v = cur + (sc[k] > 0 ? P : 0);
            Real a = 0.0;
#pragma unroll
            for (int j = 0; j < JP; ++j) { const int c = lane + 32 * j; a = fma(t[k][j], c < P ? v[c] : 0.0, a); }
            s[k] = a;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1)
#pragma unroll
            for (int k = 0; k < SLOTS; ++k) s[k] += __shfl_xor_sync(0xffffffffu, s[k], o);
    };
    // one step of both chains; the registers already hold its matrix rows, (nphase, ns) names the step to prefetch
    auto chainStep = [&](int phase, int s, int nphase, int ns) {
        Real d[SLOTS];
        dots(d);
#pragma unroll
        for (int k = 0; k < SLOTS; ++k) {
            if (sc[k] < 0) continue;
            const int c = sc[k], i = si[k], e = c * P + i;
            if (blockOf(phase, s, c) < 0) { publish(e, phase == 1 ? 0.0 : cur[e]); continue; } // idle chain: keep its vector
            if (phase == 1) { publish(e, d[k]); continue; }                                     // partial sums of the middle plane
            const int pl = phase == 0 ? ci.plane(c, s) : ci.bwdPlane(c, s);
            const Real rhs = phase == 0 ? b[pl * P + i] : x[pl * P + i];
            const Real v = masked(pl * P + i, rhs - d[k]);
            publish(e, v);
            if (lane == 0) (phase == 0 ? z : x)[pl * P + i] = v;
        }
        if (nphase >= 0) fetch(nphase, ns);
        cl.sync();
        Real *sw = cur; cur = nxt; nxt = sw;
    };
    // ---- forward: both chains start from their end planes ----
    const int S = ci.steps();
    fetch(S > 1 ? 0 : 1, 1);
    for (int i = tid; i < P; i += kChainThreads) {
        const Real v0 = masked(i, b[i]), v1 = masked((planes - 1) * P + i, b[(planes - 1) * P + i]);
        cur[i] = v0; cur[P + i] = v1;
        if (rank == 0) { if (ci.len(0) >= 1) z[i] = v0; if (ci.len(1) >= 1) z[(planes - 1) * P + i] = v1; }
    }
    cl.sync();
    for (int s = 1; s <= S - 1; ++s) chainStep(0, s, s + 1 <= S - 1 ? 0 : 1, s + 1);
    // ---- middle plane: z_mid = b_mid - T_mid z_{mid-1} - U_mid w_{mid+1} ----
    chainStep(1, 0, -1, 0);
    for (int i = tid; i < P; i += kChainThreads) {
        const Real v = masked(mid * P + i, b[mid * P + i] - cur[i] - cur[P + i]);
        if (rank == 0) z[mid * P + i] = v;
    }
    __threadfence();
    cl.sync();
    // ---- diagonal: v_p = S_p^{-1} z_p, all planes at once (v parked in x) ----
    // (4 rows per warp at a time, all their loads issued before the first multiply: the phase is L2-latency bound)
    for (int row0 = gw; row0 < planes * P; row0 += kChainSlots * tw) {
        Real sv4[kChainSlots][JP];
#pragma unroll
        for (int k = 0; k < kChainSlots; ++k) {
            const int row = row0 + k * tw;
            const bool ok = row < planes * P;
            const Real *Sr = Sinv + size_t(ok ? row : 0) * P; // row (p, i) of the stacked S^{-1} blocks: (p*P + i)*P
#pragma unroll
            for (int j = 0; j < JP; ++j) { const int c = lane + 32 * j; sv4[k][j] = (ok && c < P) ? Sr[c] : 0.0; }
        }
        Real acc[kChainSlots];
#pragma unroll
        for (int k = 0; k < kChainSlots; ++k) {
            const int row = row0 + k * tw, p = row < planes * P ? row / P : 0;
            const Real *zp = z + size_t(p) * P;
            Real a = 0.0;
#pragma unroll
            for (int j = 0; j < JP; ++j) { const int c = lane + 32 * j; a = fma(sv4[k][j], c < P ? zp[c] : 0.0, a); }
            acc[k] = a;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1)
#pragma unroll
            for (int k = 0; k < kChainSlots; ++k) acc[k] += __shfl_xor_sync(0xffffffffu, acc[k], o);
#pragma unroll
        for (int k = 0; k < kChainSlots; ++k) {
            const int row = row0 + k * tw;
            if (lane == 0 && row < planes * P) x[row] = masked(row, acc[k]);
        }
    }
    if (S >= 1) fetch(2, 1);
    __threadfence();
    cl.sync();
    // ---- backward: outwards from x_mid = v_mid ----
    for (int i = tid; i < P; i += kChainThreads) { const Real v = x[size_t(mid) * P + i]; cur[i] = v; cur[P + i] = v; }
    __syncthreads();
    for (int s = 1; s <= S; ++s) chainStep(2, s, s + 1 <= S ? 2 : -1, s + 1);
}

// Large planes (P > 256): the same three phases on the WHOLE device.  One cooperative launch; a warp owns a matrix row,
// the plane vectors live in global memory and are staged into every CTA's shared memory at the start of a step (read
// with ld.global.cg: other CTAs wrote them in the previous step); a grid barrier per chain step.  A lane issues all its
// loads of a row (up to 32) before the first multiply, so a step costs about one memory round trip, and the rows of the
// next step are prefetched into L2 before the barrier.  Dynamic shared memory: P doubles.
constexpr int kChainGridThreads = 256, kChainGridMaxThreads = 384; // default / upper bound of the CTA size (block_tri.h picks it: a warp per row of a step where that fits)
template <int N>
static __global__ void __launch_bounds__(kChainGridMaxThreads) blockTriChainGridKernel(int P, ChainIdx ci, const Real *__restrict__ Trm,
                                                                                const Real *__restrict__ Tcm, const Real *__restrict__ Sinv,
                                                                                const Real *__restrict__ b, const uint8_t *__restrict__ dmask,
                                                                                Real *x, Real *z) {
    namespace cg = cooperative_groups;
    cg::grid_group grid = cg::this_grid();
    extern __shared__ Real sv[]; // [chain][P]
    const int lane = threadIdx.x & 31, gtid = blockIdx.x * blockDim.x + threadIdx.x, gthreads = gridDim.x * blockDim.x;
    const int gw = gtid >> 5, tw = gthreads >> 5;
    const int planes = ci.planes, mid = ci.mid, S = ci.steps();
    const size_t PP = size_t(P) * P;
    auto masked = [&](int dof, Real v) { return ((dmask[dof / N] >> (dof % N)) & 1u) ? 0.0 : v; };
    auto stage = [&](const Real *v0, const Real *v1) { // plane vectors of both chains -> shared memory (block-wide)
        __syncthreads();
        for (int i = threadIdx.x; i < P; i += blockDim.x) { sv[i] = v0 ? __ldcg(v0 + i) : 0.0; sv[P + i] = v1 ? __ldcg(v1 + i) : 0.0; }
        __syncthreads();
    };
    auto rowDot = [&](const Real *__restrict__ row, const Real *v) { // row . v (v in shared memory), all lanes return the sum
        Real a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
        for (int base = 0; base < P; base += 1024) {
            Real rv[32];
#pragma unroll
            for (int k = 0; k < 32; ++k) { const int j = base + lane + 32 * k; rv[k] = j < P ? row[j] : 0.0; }
#pragma unroll
            for (int k = 0; k < 32; k += 4) {
                const int j = base + lane + 32 * k;
                a0 = fma(rv[k], j < P ? v[j] : 0.0, a0); a1 = fma(rv[k + 1], j + 32 < P ? v[j + 32] : 0.0, a1);
                a2 = fma(rv[k + 2], j + 64 < P ? v[j + 64] : 0.0, a2); a3 = fma(rv[k + 3], j + 96 < P ? v[j + 96] : 0.0, a3);
            }
        }
        Real a = (a0 + a1) + (a2 + a3);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
        return a;
    };
    // the matrices do not depend on the chains: pull the rows of the NEXT step into L2 before waiting at the barrier
    auto prefetchStep = [&](const Real *T, int q0, int q1) {
        for (int r = gw; r < 2 * P; r += tw) {
            const int c = r / P, i = r - c * P, q = c == 0 ? q0 : q1;
            if (q < 0) continue;
            for (int j = lane * 16; j < P; j += 512) asm volatile("prefetch.global.L2 [%0];" ::"l"(T + size_t(q) * PP + size_t(i) * P + j));
        }
    };
    auto fwdBlock = [&](int c, int s) { return s <= ci.len(c) - 1 ? ci.fwdBlock(c, s) : -1; };
    auto midBlock = [&](int c) { return ci.len(c) >= 1 ? ci.midBlock(c) : -1; };
    auto bwdBlock = [&](int c, int s) { return s <= ci.len(c) ? ci.bwdBlock(c, s) : -1; };
    // ---- forward ----
    for (int i = gtid; i < P; i += gthreads) {
        if (ci.len(0) >= 1) z[i] = masked(i, b[i]);
        if (ci.len(1) >= 1) z[(planes - 1) * P + i] = masked((planes - 1) * P + i, b[(planes - 1) * P + i]);
    }
    if (S > 1) prefetchStep(Trm, fwdBlock(0, 1), fwdBlock(1, 1)); else prefetchStep(Trm, midBlock(0), midBlock(1));
    grid.sync();
    for (int s = 1; s <= S - 1; ++s) {
        const int q0 = fwdBlock(0, s), q1 = fwdBlock(1, s);
        stage(q0 >= 0 ? z + size_t(ci.plane(0, s - 1)) * P : nullptr, q1 >= 0 ? z + size_t(ci.plane(1, s - 1)) * P : nullptr);
        for (int r = gw; r < 2 * P; r += tw) {
            const int c = r / P, i = r - c * P, q = c == 0 ? q0 : q1;
            if (q < 0) continue;
            const int pl = ci.plane(c, s);
            const Real d = rowDot(Trm + size_t(q) * PP + size_t(i) * P, sv + c * P);
            if (lane == 0) z[pl * P + i] = masked(pl * P + i, b[pl * P + i] - d);
        }
        if (s + 1 <= S - 1) prefetchStep(Trm, fwdBlock(0, s + 1), fwdBlock(1, s + 1)); else prefetchStep(Trm, midBlock(0), midBlock(1));
        grid.sync();
    }
    { // ---- middle plane: partial products parked in x[0 .. 2P) (x is not in use yet) ----
        const int q0 = midBlock(0), q1 = midBlock(1);
        stage(q0 >= 0 ? z + size_t(mid - 1) * P : nullptr, q1 >= 0 ? z + size_t(mid + 1) * P : nullptr);
        for (int r = gw; r < 2 * P; r += tw) {
            const int c = r / P, i = r - c * P, q = c == 0 ? q0 : q1;
            if (q < 0) continue;
            const Real d = rowDot(Trm + size_t(q) * PP + size_t(i) * P, sv + c * P);
            if (lane == 0) x[c * P + i] = d;
        }
        grid.sync();
        for (int i = gtid; i < P; i += gthreads)
            z[mid * P + i] = masked(mid * P + i, b[mid * P + i] - (q0 >= 0 ? __ldcg(x + i) : 0.0) - (q1 >= 0 ? __ldcg(x + P + i) : 0.0));
        grid.sync();
    }
    for (int p = 0; p < planes; ++p) { // ---- diagonal: no dependency between the planes, only block-level staging ----
        stage(z + size_t(p) * P, nullptr);
        for (int i = gw; i < P; i += tw) {
            const Real d = rowDot(Sinv + size_t(p) * PP + size_t(i) * P, sv);
            if (lane == 0) x[p * P + i] = masked(p * P + i, d);
        }
    }
    if (S >= 1) prefetchStep(Tcm, bwdBlock(0, 1), bwdBlock(1, 1));
    grid.sync();
    for (int s = 1; s <= S; ++s) { // ---- backward, outwards from the middle plane ----
        const int q0 = bwdBlock(0, s), q1 = bwdBlock(1, s);
        stage(q0 >= 0 ? x + size_t(ci.bwdPlane(0, s - 1)) * P : nullptr, q1 >= 0 ? x + size_t(ci.bwdPlane(1, s - 1)) * P : nullptr);
        for (int r = gw; r < 2 * P; r += tw) {
            const int c = r / P, i = r - c * P, q = c == 0 ? q0 : q1;
            if (q < 0) continue;
            const int pl = ci.bwdPlane(c, s);
            const Real d = rowDot(Tcm + size_t(q) * PP + size_t(i) * P, sv + c * P);
            if (lane == 0) x[pl * P + i] = masked(pl * P + i, __ldcg(x + pl * P + i) - d);
        }
        if (s + 1 <= S) prefetchStep(Tcm, bwdBlock(0, s + 1), bwdBlock(1, s + 1));
        grid.sync();
    }
}

} // namespace vfem
