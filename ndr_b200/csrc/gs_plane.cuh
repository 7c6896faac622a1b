// ndr_b200 — fine-level coloured Gauss-Seidel, plane-fused form (3-D Q1, FP64, sm_100a): TWO launches per sweep instead
// of eight.
//
// Reference semantics: MultigridSolver::smoothingMulticoloredGS, MG.hh:284-340.  The colour order (MG.hh:292-294) visits the
// four colours of the EVEN node planes of axis 0 first, then the four colours of the odd planes.  A node only couples to the
// planes x-1, x, x+1, so within one half
//   * different planes of that parity are independent of each other (their neighbours x +- 1 have the other parity and
//     are not touched in this half), and
//   * the four colours of a plane only depend on each other THROUGH THAT PLANE.
// One thread-block cluster therefore owns one node plane for a whole half sweep: it stages the planes x-1, x, x+1 in
// shared memory once (row-wise cp.async.bulk copies, as gs_tile.cuh), runs the four colour phases on them with a cluster
// barrier in between - updated values are written to the staged plane, to global memory and, for the rows at a strip
// boundary, into the neighbour CTA's ghost row through distributed shared memory - and never goes back to HBM/L2 for u.
// The CTAs of a cluster split the plane into strips of rows.  Per node the arithmetic and its order are those of
// gsFineKernel / gs_tile.cuh: results are bit-identical (any schedule that respects the colour order gives the same sweep).
// The per-colour tile kernel (gs_tile.cuh) remains the path for planes that do not fit a cluster's shared memory (nz > 129).
#pragma once
#include "kernels.cuh"
#include "async_copy.cuh"
#include <cooperative_groups.h>

namespace vfem {

constexpr int kGsPlaneThreads = 320;
constexpr size_t kGsPlaneSmemMax = 220 * 1024;
constexpr int kGsPlanePad = 8; // doubles in front of the staged box: the z-1 read of a z = 0 node stays inside shared memory

struct GsPlanePlan { int cy = 0, h = 0, rows = 0, pitch = 0; size_t smem = 0; };

// strips of h rows over cy CTAs (cy in {1,2,4,8}); cy = 0: the plane does not fit
inline GsPlanePlan planGsPlane(const Grid<3> &g, int forceCy = 0) {
    GsPlanePlan p;
    const int ny = g.n[1], nz = g.n[2];
    p.pitch = (3 * nz + 6 + 1) / 2 * 2;
    for (int cy : {1, 2, 4, 8}) {
        if (forceCy && cy != forceCy) continue;
        const int h = (ny + cy - 1) / cy;
        const size_t smem = (size_t(3) * (h + 2) * p.pitch + kGsPlanePad) * sizeof(Real);
        if (smem > kGsPlaneSmemMax) continue;
        if (!forceCy && h > 17 && cy < 8) continue; // prefer strips of <= 17 rows: enough CTAs to fill the device
        p.cy = cy; p.h = h; p.rows = h + 2; p.smem = smem;
        break;
    }
    return p;
}

__global__ void __launch_bounds__(kGsPlaneThreads, 1)
gsFinePlaneKernel(Grid<3> g, K0Param<3> K, const Real *__restrict__ Ey, Real *u, const Real *__restrict__ b, const uint8_t *__restrict__ dmask,
                  int px, int forward, int h, int rows, int RP) {
    namespace cg = cooperative_groups;
    cg::cluster_group cl = cg::this_cluster();
    extern __shared__ __align__(16) unsigned char gsPlaneRaw[];
    Real *box = reinterpret_cast<Real *>(gsPlaneRaw) + kGsPlanePad; // [plane 0..2][row][RP]
    __shared__ __align__(8) unsigned long long bar;

    const int nx = g.n[0], ny = g.n[1], nz = g.n[2];
    const int CY = int(cl.num_blocks()), r = int(cl.block_rank());
    const int slot = blockIdx.x / CY;
    const int x = colourFirstPlane<3>(g, px << 2) + 2 * slot;         // the node plane of this cluster
    const int yOwn0 = r * h, yOwn1 = min((r + 1) * h, ny) - 1;         // owned rows (empty when yOwn0 > yOwn1)
    const int yLo = yOwn0 - 1;                                          // global row of staged row 0 (may be -1: not staged)
    const long long baseOdd = (long long)(reinterpret_cast<size_t>(u) >> 3) & 1;
    const size_t planeSz = size_t(rows) * RP;

    if (threadIdx.x == 0) mbarInit(&bar, 32);
    __syncthreads();
    if (threadIdx.x < 32) { // warp 0 stages the rows of the three planes; every lane arrives once with the bytes it asked for
        unsigned bytes = 0;
        const long long total = (long long)g.nn * 24, ubase = (long long)reinterpret_cast<size_t>(u);
        for (int q = threadIdx.x; q < 3 * rows; q += 32) {
            const int pl = q / rows, lr = q - pl * rows;
            const int xx = x + pl - 1, yy = yLo + lr;
            if (xx < 0 || xx >= nx || yy < 0 || yy >= ny || yy > yOwn1 + 1 || yOwn0 > yOwn1) continue;
            const long long row0 = ((long long)xx * ny + yy) * nz, src = row0 * 24;
            const int delta = int((ubase + src) & 8);
            Real *dst = box + size_t(pl) * planeSz + size_t(lr) * RP + 2; // 16-byte aligned; node 0 lands at dst + delta / 8
            const int need = nz * 24 + delta;
            int size = (need + 15) & ~15;
            if (src - delta + size > total) {
                size = need & ~15;
                for (int i = size >> 3; i < (need >> 3); ++i) dst[i] = reinterpret_cast<const Real *>(reinterpret_cast<const char *>(u) + src - delta)[i];
            }
            if (size > 0) { bulkLoad(dst, reinterpret_cast<const char *>(u) + src - delta, unsigned(size), &bar); bytes += unsigned(size); }
        }
        mbarArriveExpectTx(&bar, bytes);
    }
    cl.sync(); // every CTA of the cluster is running: its shared memory may be written remotely from here on
    mbarWait(&bar, 0);

    const int fxp = (ny * nz) & 1, fyp = nz & 1; // parity flips of a row's 8-byte shift with x and y
    for (int ci = 0; ci < 4; ++ci) {
        const int cyz = forward ? ci : 3 - ci;   // MG.hh:318-320: reversed colour order for the backward sweep
        const int py = (cyz >> 1) & 1, pz = cyz & 1;
        const int yFirst = yOwn0 + ((yOwn0 ^ py) & 1);
        const int cntY = yFirst <= yOwn1 ? (yOwn1 - yFirst) / 2 + 1 : 0, cntZ = (nz - 1 - pz) / 2 + 1;
        for (int t = threadIdx.x; t < cntY * cntZ; t += kGsPlaneThreads) {
            const int iy = t / cntZ, iz = t - iy * cntZ;
            const int y = yFirst + 2 * iy, z = pz + 2 * iz, ly = y - yLo;
            const int n = (x * ny + y) * nz + z;
            Real Ee[8];
#pragma unroll
            for (int s = 0; s < 8; ++s) {
                const int e0 = x - 1 + ((s >> 2) & 1), e1 = y - 1 + ((s >> 1) & 1), e2 = z - 1 + (s & 1);
                const bool ok = e0 >= 0 && e0 < g.E[0] && e1 >= 0 && e1 < g.E[1] && e2 >= 0 && e2 < g.E[2];
                Ee[s] = ok ? Ey[e0 * g.einc[0] + e1 * g.einc[1] + e2] : 0.0;
            }
            Real bn[3];
#pragma unroll
            for (int c = 0; c < 3; ++c) bn[c] = b[size_t(n) * 3 + c];
            const unsigned mask = dmask[n];
            // staged row (x + dx, y + dy): node z', component c at index 2 + par + 3 z' + c, par = 8-byte shift of that row
            const int p0 = int((baseOdd + ((long long)x * ny + y) * nz) & 1);
            Real re[8][3], uc[3] = {0.0, 0.0, 0.0};
#pragma unroll
            for (int s = 0; s < 8; ++s) re[s][0] = re[s][1] = re[s][2] = 0.0;
#pragma unroll
            for (int rr = 0; rr < 9; ++rr) {
                const int dx = rr / 3 - 1, dy = rr % 3 - 1;
                const int par = p0 ^ (dx != 0 ? fxp : 0) ^ (dy != 0 ? fyp : 0);
                const int s0 = 2 + par + 3 * (z - 1), e = s0 & 1;                 // first of the 9 doubles of nodes z-1, z, z+1; its parity
                const Real *base = box + size_t(dx + 1) * planeSz + size_t(ly + dy) * RP + (s0 - e); // even index: 16-byte aligned
                Real w[9];
                if (e) {
                    w[0] = base[1];
#pragma unroll
                    for (int i = 0; i < 4; ++i) { const double2 t2 = *reinterpret_cast<const double2 *>(base + 2 + 2 * i); w[1 + 2 * i] = t2.x; w[2 + 2 * i] = t2.y; }
                } else {
#pragma unroll
                    for (int i = 0; i < 4; ++i) { const double2 t2 = *reinterpret_cast<const double2 *>(base + 2 * i); w[2 * i] = t2.x; w[2 * i + 1] = t2.y; }
                    w[8] = base[8];
                }
                const bool okr = (x + dx >= 0) && (x + dx < nx) && (y + dy >= 0) && (y + dy < ny);
#pragma unroll
                for (int dzi = 0; dzi < 3; ++dzi) {
                    const int dz = dzi - 1, k = rr * 3 + dzi;
                    const bool ok = okr && (z + dz >= 0) && (z + dz < nz);
                    Real v[3];
#pragma unroll
                    for (int c = 0; c < 3; ++c) v[c] = ok ? w[dzi * 3 + c] : 0.0;
                    if (k == 13) { uc[0] = v[0]; uc[1] = v[1]; uc[2] = v[2]; }
#pragma unroll
                    for (int s = 0; s < 8; ++s) {
                        const int sx = (s >> 2) & 1, sy = (s >> 1) & 1, sz = s & 1;
                        const int q0 = dx + 1 - sx, q1 = dy + 1 - sy, q2 = dz + 1 - sz;
                        if (q0 < 0 || q0 > 1 || q1 < 0 || q1 > 1 || q2 < 0 || q2 > 1) continue;
                        const int q = q0 * 4 + q1 * 2 + q2, loc = ((1 - sx) << 2) | ((1 - sy) << 1) | (1 - sz);
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
            // new value -> global memory, the staged plane, and the ghost row of the strip next door
            const size_t own = size_t(1) * planeSz + size_t(ly) * RP + 2 + p0 + 3 * z;
            Real *gup = (y == yOwn0 && r > 0) ? cl.map_shared_rank(box, r - 1) + planeSz + size_t(h + 1) * RP + 2 + p0 + 3 * z : nullptr;
            Real *gdn = (y == yOwn1 && r + 1 < CY && yOwn1 + 1 < ny) ? cl.map_shared_rank(box, r + 1) + planeSz + 2 + p0 + 3 * z : nullptr;
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                const Real nv = uc[c] + du[c];
                u[size_t(n) * 3 + c] = nv;
                box[own + c] = nv;
                if (gup) gup[c] = nv;
                if (gdn) gdn[c] = nv;
            }
        }
        cl.sync(); // the colour's updates (own strip and the pushed ghost rows) are visible to the whole cluster
    }
}

// ---- assembled levels (l >= 1): the same plane-fused half sweep on the 27-point block stencil -------------------------
// Replaces the cooperative all-colour kernel (gsStencilSweepKernel: 8 grid barriers per sweep, 2-3.3 TB/s).  A warp owns a
// node: its 3 x 81 stencil coefficients stream from global memory (coalesced; two nodes per warp in flight), the 27
// neighbour values come from the staged planes.  The staged box is zero-filled first: out-of-grid neighbours carry zero
// coefficients (buildLevel1StencilKernel / galerkinStencilKernel) and must meet finite values.
constexpr int kGsStPlaneThreads = 384;

__global__ void __launch_bounds__(kGsStPlaneThreads, 1)
gsStencilPlaneKernel(Grid<3> g, const Real *__restrict__ S, Real *u, const Real *__restrict__ b, const uint8_t *__restrict__ dmask, int px,
                     int forward, int h, int rows, int RP) {
    namespace cg = cooperative_groups;
    cg::cluster_group cl = cg::this_cluster();
    extern __shared__ __align__(16) unsigned char gsPlaneRaw[];
    Real *raw = reinterpret_cast<Real *>(gsPlaneRaw), *box = raw + kGsPlanePad;
    __shared__ __align__(8) unsigned long long bar;
    constexpr int ROW = 81, J = 3;

    const int nx = g.n[0], ny = g.n[1], nz = g.n[2];
    const int CY = int(cl.num_blocks()), r = int(cl.block_rank());
    const int slot = blockIdx.x / CY;
    const int x = colourFirstPlane<3>(g, px << 2) + 2 * slot;
    const int yOwn0 = r * h, yOwn1 = min((r + 1) * h, ny) - 1;
    const int yLo = yOwn0 - 1;
    const long long baseOdd = (long long)(reinterpret_cast<size_t>(u) >> 3) & 1;
    const int planeSz = rows * RP;

    { // zero the box (generic proxy), then hand it to the async proxy
        double2 *z2 = reinterpret_cast<double2 *>(raw);
        const int n2 = (3 * planeSz + kGsPlanePad) / 2;
        for (int i = threadIdx.x; i < n2; i += kGsStPlaneThreads) z2[i] = make_double2(0.0, 0.0);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    if (threadIdx.x == 0) mbarInit(&bar, 32);
    __syncthreads();
    if (threadIdx.x < 32) {
        unsigned bytes = 0;
        const long long total = (long long)g.nn * 24, ubase = (long long)reinterpret_cast<size_t>(u);
        for (int q = threadIdx.x; q < 3 * rows; q += 32) {
            const int pl = q / rows, lr = q - pl * rows;
            const int xx = x + pl - 1, yy = yLo + lr;
            if (xx < 0 || xx >= nx || yy < 0 || yy >= ny || yy > yOwn1 + 1 || yOwn0 > yOwn1) continue;
            const long long row0 = ((long long)xx * ny + yy) * nz;
            long long src = row0 * 24;
            const int delta = int((ubase + src) & 8);
            Real *dst = box + size_t(pl) * planeSz + size_t(lr) * RP + 2;
            int need = nz * 24 + delta;
            src -= delta;
            if (src < 0) { dst[1] = u[0]; src += 16; dst += 2; need -= 16; }       // never read in front of the field: its first double by hand
            int size = (need + 15) & ~15;
            if (src + size > total) {                                              // ... nor behind it
                size = need & ~15;
                for (int i = size >> 3; i < (need >> 3); ++i) dst[i] = reinterpret_cast<const Real *>(reinterpret_cast<const char *>(u) + src)[i];
            }
            if (size > 0) { bulkLoad(dst, reinterpret_cast<const char *>(u) + src, unsigned(size), &bar); bytes += unsigned(size); }
        }
        mbarArriveExpectTx(&bar, bytes);
    }
    cl.sync();
    mbarWait(&bar, 0);

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = kGsStPlaneThreads / 32;
    const int fxp = (ny * nz) & 1, fyp = nz & 1;
    int soff[J], sflip[J]; bool valid[J];
#pragma unroll
    for (int j = 0; j < J; ++j) { // entry e = lane + 32 j of a stencil row multiplies u[n + delta(e / 3)][e % 3]
        const int e = lane + 32 * j;
        valid[j] = e < ROW;
        const int k = valid[j] ? e / 3 : 13, bc = valid[j] ? e % 3 : 0;
        const int dx = k / 9 - 1, dy = (k / 3) % 3 - 1, dz = k % 3 - 1;
        soff[j] = dx * planeSz + dy * RP + 3 * dz + bc;   // relative to the node itself in the staged centre plane
        sflip[j] = (dx != 0 ? fxp : 0) ^ (dy != 0 ? fyp : 0);
    }
    struct Slot { Real Sv[3][J], uv[J], bv; unsigned mask; int n, own, p0, y, z; bool on; };
    for (int ci = 0; ci < 4; ++ci) {
        const int cyz = forward ? ci : 3 - ci;
        const int py = (cyz >> 1) & 1, pz = cyz & 1;
        const int yFirst = yOwn0 + ((yOwn0 ^ py) & 1);
        const int cntY = yFirst <= yOwn1 ? (yOwn1 - yFirst) / 2 + 1 : 0, cntZ = (nz - 1 - pz) / 2 + 1, cnt = cntY * cntZ;
        auto load = [&](int t, Slot &L) {
            L.on = t < cnt;
            if (!L.on) return;
            const int iy = t / cntZ, iz = t - iy * cntZ;
            L.y = yFirst + 2 * iy; L.z = pz + 2 * iz;
            L.n = (x * ny + L.y) * nz + L.z;
            L.p0 = int((baseOdd + ((long long)x * ny + L.y) * nz) & 1);
            L.own = planeSz + (L.y - yLo) * RP + 2 + 3 * L.z;          // + p0: this node in the staged centre plane
            const Real *Sn = S + size_t(L.n) * 3 * ROW + lane;
#pragma unroll
            for (int j = 0; j < J; ++j) {
                L.uv[j] = valid[j] ? box[L.own + soff[j] + (L.p0 ^ sflip[j])] : 0.0;
#pragma unroll
                for (int a = 0; a < 3; ++a) L.Sv[a][j] = valid[j] ? Sn[a * ROW + 32 * j] : 0.0;
            }
            L.bv = (lane < 3) ? b[size_t(L.n) * 3 + lane] : 0.0;
            L.mask = (lane == 0) ? unsigned(dmask[L.n]) : 0u;
        };
        auto update = [&](const Slot &L) {
            if (!L.on) return;
            Real acc[3];
#pragma unroll
            for (int a = 0; a < 3; ++a) {
                Real s = 0.0;
#pragma unroll
                for (int j = 0; j < J; ++j)
                    if (valid[j]) s = fma(L.Sv[a][j], L.uv[j], s);
                acc[a] = s;
            }
#pragma unroll
            for (int a = 0; a < 3; ++a)
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) acc[a] += __shfl_xor_sync(0xffffffffu, acc[a], o);
            Real M[9], tt[3], uo[3];
#pragma unroll
            for (int a = 0; a < 3; ++a)
#pragma unroll
                for (int c = 0; c < 3; ++c) { constexpr int e0 = 13 * 3; const int e = e0 + c; M[a * 3 + c] = __shfl_sync(0xffffffffu, L.Sv[a][e / 32], e % 32); }
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                constexpr int e0 = 13 * 3; const int e = e0 + c;
                tt[c] = __shfl_sync(0xffffffffu, L.bv, c) - acc[c];
                uo[c] = __shfl_sync(0xffffffffu, L.uv[e / 32], e % 32);
            }
            if (lane == 0) {
                Real du[3];
                if (forward) gsNodeSubst<3, true>(tt, M, L.mask, du); else gsNodeSubst<3, false>(tt, M, L.mask, du);
                const int o = L.own + L.p0;
                Real *gup = (L.y == yOwn0 && r > 0) ? cl.map_shared_rank(box, r - 1) + planeSz + (h + 1) * RP + 2 + L.p0 + 3 * L.z : nullptr;
                Real *gdn = (L.y == yOwn1 && r + 1 < CY && yOwn1 + 1 < ny) ? cl.map_shared_rank(box, r + 1) + planeSz + 2 + L.p0 + 3 * L.z : nullptr;
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    const Real nv = uo[c] + du[c];
                    u[size_t(L.n) * 3 + c] = nv;
                    box[o + c] = nv;
                    if (gup) gup[c] = nv;
                    if (gdn) gdn[c] = nv;
                }
            }
        };
        // two nodes per warp in flight, the next pair's coefficients requested before the current pair is reduced
        Slot A, B, A2, B2;
        int t = warp;
        load(t, A); load(t + nw, B);
        while (t < cnt) {
            const int tn = t + 2 * nw;
            load(tn, A2); load(tn + nw, B2);
            update(A); update(B);
            A = A2; B = B2;
            t = tn;
        }
        cl.sync();
    }
}

} // namespace vfem
