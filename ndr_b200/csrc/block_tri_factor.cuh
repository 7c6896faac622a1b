// ndr_b200 — hand-written block LDL^T factorisation of a block-tridiagonal SPD operator (sm_100a, FP64).
//
// Replaces the reference's CHOLMOD numeric factorisation (TPS.hh:834-865, once per density update,
// m_numericFactorizationUpToDate) for the exact solves of this library: the coarsest multigrid level and the direct
// solve of ComplianceObjective.  No cuSOLVER / cuBLAS: ONE cooperative launch runs the whole recurrence
//     S_0 = A_00,   T_{p+1} = A_{p+1,p} S_p^{-1},   S_{p+1} = A_{p+1,p+1} - T_{p+1} A_{p+1,p}^T
// from BOTH ends of the plane sequence (ChainIdx, coarse_solve.cuh): the first half of the grid's CTAs eliminates the
// planes 0..mid-1 downwards, the second half planes-1..mid+1 upwards, both meet at plane `mid`.
//  * S_p^{-1} is formed IN PLACE by a blocked Gauss-Jordan inversion without pivoting (S_p is SPD; constrained dofs are
//    identity rows).  Block size 32: every CTA inverts the 32x32 pivot block in shared memory itself (no broadcast
//    step), then updates its share of the trailing 32x32 tiles  C_ij -= (F_i D^{-1}) R_j ; a grid barrier; the pivot
//    row / column panels ; a grid barrier.  2 P^3 flops per plane, 2 ceil(P/32) grid barriers.
//  * A_{p+1,p} couples a node only to the 3^(N-1) nodes around it in the neighbouring plane, so T = A S^{-1} and the
//    Schur update T A^T are SPARSE-times-dense products (<= 3^(N-1) N terms per entry) instead of P^3 GEMMs.
// Outputs are the arrays the chain-solve kernels (coarse_solve.cuh) read: S_p^{-1} (in D), T row- and column-major.
#pragma once
#include "coarse_solve.cuh"

namespace vfem {

constexpr int kFB = 32;             // Gauss-Jordan block size
constexpr int kFactorThreads = 256; // 32 rows x 8 column groups of 4

template <int N>
struct FactorParams {
    int P, planes, mid;
    int pn[N > 1 ? N - 1 : 1];      // nodes per in-plane axis (axes 1..N-1 of the grid)
};

constexpr int planeNeighbourMax(int N) { return ipow3(N - 1) * N; }

// in-plane neighbours of dof i = (node a, component): the dofs of the nodes within +-1 of a; slot s of the fixed-size list
// is -1 where the neighbour lies outside the plane
template <int N>
__device__ __forceinline__ void planeNeighbours(const FactorParams<N> &fp, int i, int (&nb)[planeNeighbourMax(N)]) {
    const int a = i / N;
    if (N == 2) {
#pragma unroll
        for (int d = -1; d <= 1; ++d) {
            const int y = a + d;
            const bool ok = y >= 0 && y < fp.pn[0];
#pragma unroll
            for (int c = 0; c < N; ++c) nb[(d + 1) * N + c] = ok ? y * N + c : -1;
        }
    } else {
        const int nz = fp.pn[N - 2], y = a / nz, z = a - y * nz;
#pragma unroll
        for (int dy = -1; dy <= 1; ++dy)
#pragma unroll
            for (int dz = -1; dz <= 1; ++dz) {
                const int yy = y + dy, zz = z + dz;
                const bool ok = yy >= 0 && yy < fp.pn[0] && zz >= 0 && zz < nz;
#pragma unroll
                for (int c = 0; c < N; ++c) nb[((dy + 1) * 3 + dz + 1) * N + c] = ok ? (yy * nz + zz) * N + c : -1;
            }
    }
}

// In-place inverse of the SPD matrix M (P x P, leading dimension P; nullptr = this team idles but keeps the barrier
// count) by the team's CTAs [0, ncb); every CTA of the GRID calls this with the same P.
__device__ __forceinline__ void gaussJordanInvert(Real *M, int P, int cb, int ncb, int *info, int plane) {
    namespace cg = cooperative_groups;
    cg::grid_group grid = cg::this_grid();
    __shared__ Real sD[kFB][kFB + 1], sG[kFB][kFB + 1], sF[kFB][kFB + 1], sR[kFB][kFB + 1];
    const int tid = threadIdx.x, tr = tid & 31, tc = tid >> 5; // row, column group (columns 4 tc .. 4 tc + 3)
    const int nbk = (P + kFB - 1) / kFB;
    auto loadTile = [&](Real (&s)[kFB][kFB + 1], int r0, int c0, bool identityPad) {
        for (int e = tid; e < kFB * kFB; e += kFactorThreads) {
            const int r = e & 31, c = e >> 5;
            const bool in = r0 + r < P && c0 + c < P;
            s[r][c] = in ? __ldcg(M + size_t(r0 + r) + size_t(c0 + c) * P) : ((identityPad && r == c) ? 1.0 : 0.0);
        }
    };
    for (int kb = 0; kb < nbk; ++kb) {
        const int k0 = kb * kFB;
        if (M) {
            // ---- pivot block -> its inverse, in shared memory (unblocked Gauss-Jordan, every CTA for itself) ----
            __syncthreads();
            loadTile(sD, k0, k0, true);
            __syncthreads();
            for (int k = 0; k < kFB; ++k) {
                const Real d = sD[k][k];
                if (!(d > 0.0) && tid == 0 && cb == 0) atomicCAS(info, 0, plane + 1);
                const Real piv = 1.0 / d;
                Real nv[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int c = 4 * tc + q;
                    const Real f = sD[tr][k], rk = sD[k][c];
                    if (tr == k) nv[q] = (c == k) ? piv : rk * piv;
                    else nv[q] = (c == k) ? -f * piv : fma(-f * piv, rk, sD[tr][c]);
                }
                __syncthreads();
#pragma unroll
                for (int q = 0; q < 4; ++q) sD[tr][4 * tc + q] = nv[q];
                __syncthreads();
            }
            // ---- trailing tiles (i, j), i != kb, j != kb:  C -= (F_i D^-1) R_j ----
            const int nt = (nbk - 1) * (nbk - 1);
            const int t0 = int((long long)cb * nt / ncb), t1 = int((long long)(cb + 1) * nt / ncb);
            // the R tile and the C entries of the NEXT tile are fetched into registers before the arithmetic of the current
            // one (a tile is 3 dependent L2 round trips otherwise: the loop is latency-, not FP64-bound)
            auto tileOf = [&](int t, int &i, int &j) {
                const int ii = t / (nbk - 1), jj = t - ii * (nbk - 1);
                i = ii + (ii >= kb); j = jj + (jj >= kb);
            };
            auto fetchR = [&](int j, Real (&rr)[4]) {
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int e = tid + q * kFactorThreads, r = k0 + (e & 31), c = j * kFB + (e >> 5);
                    rr[q] = (r < P && c < P) ? __ldcg(M + size_t(r) + size_t(c) * P) : 0.0;
                }
            };
            auto fetchC = [&](int i, int j, Real (&cv)[4]) {
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int r = i * kFB + tr, c = j * kFB + 4 * tc + q;
                    cv[q] = (r < P && c < P) ? __ldcg(M + size_t(r) + size_t(c) * P) : 0.0;
                }
            };
            int curI = -1, i = 0, j = 0;
            Real rr[4], cn[4];
            if (t0 < t1) { tileOf(t0, i, j); fetchR(j, rr); fetchC(i, j, cn); }
            for (int t = t0; t < t1; ++t) {
                tileOf(t, i, j);
                __syncthreads();
                if (i != curI) {
                    loadTile(sF, i * kFB, k0, false);
                    __syncthreads();
                    Real g[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll 8
                    for (int k = 0; k < kFB; ++k) {
                        const Real f = sF[tr][k];
#pragma unroll
                        for (int q = 0; q < 4; ++q) g[q] = fma(f, sD[k][4 * tc + q], g[q]);
                    }
#pragma unroll
                    for (int q = 0; q < 4; ++q) sG[tr][4 * tc + q] = g[q];
                    curI = i;
                }
                Real cv[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) { const int e = tid + q * kFactorThreads; sR[e & 31][e >> 5] = rr[q]; cv[q] = cn[q]; }
                if (t + 1 < t1) { int i2, j2; tileOf(t + 1, i2, j2); fetchR(j2, rr); fetchC(i2, j2, cn); }
                __syncthreads();
                const int r = i * kFB + tr;
                Real acc[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll 8
                for (int k = 0; k < kFB; ++k) {
                    const Real gk = sG[tr][k];
#pragma unroll
                    for (int q = 0; q < 4; ++q) acc[q] = fma(gk, sR[k][4 * tc + q], acc[q]);
                }
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int c = j * kFB + 4 * tc + q;
                    if (r < P && c < P) M[size_t(r) + size_t(c) * P] = cv[q] - acc[q];
                }
            }
        }
        grid.sync();
        if (M) {
            // ---- pivot row panel  D^-1 R_j, pivot column panel  -F_i D^-1, pivot block  D^-1 ----
            const int nt = 2 * (nbk - 1) + 1;
            for (int t = cb; t < nt; t += ncb) {
                __syncthreads();
                if (t == nt - 1) {
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const int r = k0 + tr, c = k0 + 4 * tc + q;
                        if (r < P && c < P) M[size_t(r) + size_t(c) * P] = sD[tr][4 * tc + q];
                    }
                    continue;
                }
                const bool rowPanel = t < nbk - 1;
                const int o = rowPanel ? t : t - (nbk - 1), blk = o + (o >= kb);
                loadTile(sR, rowPanel ? k0 : blk * kFB, rowPanel ? blk * kFB : k0, false);
                __syncthreads();
                Real acc[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll 8
                for (int k = 0; k < kFB; ++k) {
#pragma unroll
                    for (int q = 0; q < 4; ++q)
                        acc[q] = rowPanel ? fma(sD[tr][k], sR[k][4 * tc + q], acc[q]) : fma(sR[tr][k], sD[k][4 * tc + q], acc[q]);
                }
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int r = (rowPanel ? k0 : blk * kFB) + tr, c = (rowPanel ? blk * kFB : k0) + 4 * tc + q;
                    if (r < P && c < P) M[size_t(r) + size_t(c) * P] = rowPanel ? acc[q] : -acc[q];
                }
            }
        }
        grid.sync();
    }
}

// D: planes blocks (A_pp on entry, S_p^{-1} on exit), Lo: A_{p+1,p}, both column-major with leading dimension P;
// Trm / Tcm: block q as ChainIdx describes.  info: 0, or 1 + the plane whose pivot was not positive.
template <int N>
static __global__ void __launch_bounds__(kFactorThreads) blockTriFactorKernel(FactorParams<N> fp, Real *D, const Real *__restrict__ Lo,
                                                                              Real *Trm, Real *Tcm, int *info) {
    namespace cg = cooperative_groups;
    cg::grid_group grid = cg::this_grid();
    const int P = fp.P, planes = fp.planes, mid = fp.mid;
    const size_t PP = size_t(P) * P;
    const ChainIdx ci{planes, mid};
    const int G = gridDim.x, Gh = G / 2;
    // teams: chain 0 = CTAs [0, Gh), chain 1 = [Gh, 2 Gh); a single-CTA grid (or the odd CTA) belongs to chain 0
    const int c = (G >= 2 && int(blockIdx.x) >= Gh && int(blockIdx.x) < 2 * Gh) ? 1 : 0;
    const bool spare = G >= 2 && int(blockIdx.x) >= 2 * Gh;
    const int ncb = G >= 2 ? Gh : 1, cb = G >= 2 ? int(blockIdx.x) - c * Gh : 0;
    constexpr int NB = planeNeighbourMax(N);
    const int lane = threadIdx.x & 31, tw = cb * (kFactorThreads / 32) + (threadIdx.x >> 5), tnw = ncb * (kFactorThreads / 32); // warp of the team
    const int S = ci.steps();
    const bool serial = G < 2; // one team walks both chains one after the other
    for (int pass = 0; pass < (serial ? 2 : 1); ++pass) {
        const int cc = serial ? pass : c;
        for (int s = 0; s < (serial ? ci.len(cc) : S); ++s) {
            const bool active = !spare && s < ci.len(cc);
            const int p = ci.plane(cc, s);
            gaussJordanInvert(active ? D + PP * p : nullptr, P, cb, ncb, info, p);
            // ---- T = A S^-1 (chain 0: T_{p+1} = A_{p+1,p} S_p^-1, block p; chain 1: U_{p-1} = A_{p,p-1}^T S_p^-1, block p-1) ----
            const int q = cc == 0 ? p : p - 1;
            if (active) { // one warp per row i of T: its <= 3^(N-1) N coefficients of A first, then lanes over the columns
                const Real *A = Lo + PP * q, *Si = D + PP * p;
                Real *tr = Trm + PP * q, *tc = Tcm + PP * q;
                for (int i = tw; i < P; i += tnw) {
                    int nb[NB]; Real av[NB];
                    planeNeighbours<N>(fp, i, nb);
#pragma unroll
                    for (int m = 0; m < NB; ++m) av[m] = nb[m] < 0 ? 0.0 : (cc == 0 ? A[size_t(i) + size_t(nb[m]) * P] : A[size_t(nb[m]) + size_t(i) * P]);
                    for (int k = lane; k < P; k += 32) {
                        Real sv[NB];
#pragma unroll
                        for (int m = 0; m < NB; ++m) sv[m] = nb[m] < 0 ? 0.0 : __ldcg(Si + size_t(k) + size_t(nb[m]) * P); // S^-1 is symmetric: (j,k) read as (k,j)
                        Real acc = 0.0;
#pragma unroll
                        for (int m = 0; m < NB; ++m) acc = fma(av[m], sv[m], acc);
                        tr[size_t(i) * P + k] = acc;
                        tc[size_t(i) + size_t(k) * P] = acc;
                    }
                }
            }
            grid.sync();
            // ---- Schur complement of the next plane of the chain; both chains end on plane `mid`: chain 1 goes second ----
            const int target = cc == 0 ? p + 1 : p - 1;
            for (int turn = 0; turn < 2; ++turn) {
                const bool clash = !serial && s == ci.len(0) - 1 && s == ci.len(1) - 1; // both targets are `mid` in this step
                if (turn == 1 && !clash) break;
                if (active && (clash ? turn == cc : turn == 0)) { // one warp per column k of the target block, lanes over its rows
                    const Real *A = Lo + PP * q, *T = Tcm + PP * q;
                    Real *Dn = D + PP * target;
                    for (int k = tw; k < P; k += tnw) {
                        int nb[NB]; Real av[NB];
                        planeNeighbours<N>(fp, k, nb);
#pragma unroll
                        for (int m = 0; m < NB; ++m) av[m] = nb[m] < 0 ? 0.0 : (cc == 0 ? A[size_t(k) + size_t(nb[m]) * P] : A[size_t(nb[m]) + size_t(k) * P]);
                        for (int i = lane; i < P; i += 32) {
                            Real tv[NB];
#pragma unroll
                            for (int m = 0; m < NB; ++m) tv[m] = nb[m] < 0 ? 0.0 : __ldcg(T + size_t(i) + size_t(nb[m]) * P);
                            Real *d = Dn + size_t(i) + size_t(k) * P;
                            Real acc = __ldcg(d);
#pragma unroll
                            for (int m = 0; m < NB; ++m) acc = fma(-tv[m], av[m], acc);
                            *d = acc;
                        }
                    }
                }
                grid.sync();
            }
        }
    }
    gaussJordanInvert(spare ? nullptr : (c == 0 ? D + PP * mid : nullptr), P, cb, ncb, info, mid);
}

} // namespace vfem
