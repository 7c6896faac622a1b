// ndr_b200 — filters, objective, problem and OC step on the device.
#include "topopt.h"
#include "kernels_vec.cuh"
#include <cmath>

namespace vfem {

struct Dims { int nd; int E[3]; };

// ---- SmoothingFilter (Filter.hh:105-162): row i has weight 1/k_i on every in-grid voxel of the cube
// |delta_d| <= r (NDVector.hh:235-262).  BACKPROP applies the transposed matrix. ------------------------
// Slab form: D holds the GLOBAL dimensions (the clipping of the box and the counts are those of the undecomposed grid); `in`
// starts at global layer inLo of axis 0 and must hold every in-grid layer within r of the outputs, `out` starts at global layer
// outLo and has n entries.  The undecomposed call is inLo = outLo = 0, n = all elements.
template <bool BACKPROP>
__global__ void __launch_bounds__(kRedBlock) smoothingKernel(Dims D, int r, const Real *__restrict__ in, Real *__restrict__ out, size_t n,
                                                             int inLo, int outLo) {
    size_t i = size_t(blockIdx.x) * kRedBlock + threadIdx.x;
    if (i >= n) return;
    // grids with fewer than 3 axes are padded with trailing unit axes (the clipped box is then 1 wide there)
    int c[3], E[3] = {D.E[0], D.E[1], D.E[2]};
    { size_t t = i; for (int d = 2; d > 0; --d) { c[d] = int(t % size_t(E[d])); t /= size_t(E[d]); } c[0] = int(t) + outLo; }
    in -= size_t(inLo) * E[1] * E[2];
    const int rr[3] = {r, r, r};
    int lo[3], hi[3];
    for (int d = 0; d < 3; ++d) { lo[d] = max(c[d] - rr[d], 0); hi[d] = min(c[d] + rr[d], E[d] - 1); }
    Real s = 0.0;
    if (!BACKPROP) {
        for (int a = lo[0]; a <= hi[0]; ++a)
            for (int b = lo[1]; b <= hi[1]; ++b)
                for (int k = lo[2]; k <= hi[2]; ++k) s += in[(size_t(a) * E[1] + b) * E[2] + k];
        const int cnt = (hi[0] - lo[0] + 1) * (hi[1] - lo[1] + 1) * (hi[2] - lo[2] + 1);
        out[i] = s * (1.0 / Real(cnt));
    } else {
        for (int a = lo[0]; a <= hi[0]; ++a) {
            const int ca = min(a + rr[0], E[0] - 1) - max(a - rr[0], 0) + 1;
            for (int b = lo[1]; b <= hi[1]; ++b) {
                const int cb = min(b + rr[1], E[1] - 1) - max(b - rr[1], 0) + 1;
                for (int k = lo[2]; k <= hi[2]; ++k) {
                    const int ck = min(k + rr[2], E[2] - 1) - max(k - rr[2], 0) + 1;
                    s += in[(size_t(a) * E[1] + b) * E[2] + k] * (1.0 / Real(ca * cb * ck));
                }
            }
        }
        out[i] = s;
    }
}

// ---- ProjectionFilter (Filter.hh:55-79) -----------------------------------------------------------------
__global__ void __launch_bounds__(kRedBlock) projectionKernel(Real beta, const Real *__restrict__ in, Real *__restrict__ out, size_t n) {
    size_t i = size_t(blockIdx.x) * kRedBlock + threadIdx.x;
    if (i >= n) return;
    const Real th = tanh(0.5 * beta);
    out[i] = 0.5 * (th + tanh(beta * (in[i] - 0.5))) / th;
}
__global__ void __launch_bounds__(kRedBlock) projectionBackpropKernel(Real beta, const Real *__restrict__ in, const Real *__restrict__ vars,
                                                                       Real *__restrict__ out, size_t n) {
    size_t i = size_t(blockIdx.x) * kRedBlock + threadIdx.x;
    if (i >= n) return;
    const Real t = tanh(beta * (vars[i] - 0.5));
    out[i] = in[i] * 0.5 * beta * (1.0 - t * t) / tanh(0.5 * beta);
}

// x(lambda) = min(1, x0+m, max(0, x0-m, x0 sqrt(dJ / (dc lambda))))    (OptimalityCriterion.hh:21-26)
__global__ void __launch_bounds__(kRedBlock) ocCandidateKernel(const Real *__restrict__ x0, const Real *__restrict__ dJ, const Real *__restrict__ dc,
                                                                Real lambda, Real m, Real *__restrict__ out, size_t n) {
    size_t i = size_t(blockIdx.x) * kRedBlock + threadIdx.x;
    if (i >= n) return;
    const Real x = x0[i];
    Real v = x * sqrt(dJ[i] / (dc[i] * lambda));
    v = fmax(v, x - m); v = fmax(v, 0.0); v = fmin(v, x + m); v = fmin(v, 1.0);
    out[i] = v;
}

__global__ void __launch_bounds__(kRedBlock) fillKernel(Real *__restrict__ out, Real v, size_t n) {
    size_t i = size_t(blockIdx.x) * kRedBlock + threadIdx.x;
    if (i < n) out[i] = v;
}

// out[0] = sum (a-b)^2
__global__ void __launch_bounds__(kRedBlock) diffSqKernel(const Real *__restrict__ a, const Real *__restrict__ b, size_t n, ReduceWork w, Real *out) {
    Real v[1] = {0.0};
    for (size_t i = size_t(blockIdx.x) * kRedBlock + threadIdx.x; i < n; i += size_t(gridDim.x) * kRedBlock) { Real d = a[i] - b[i]; v[0] = fma(d, d, v[0]); }
    blockReduceAndFinish<1>(v, w.partials, w.ticket, out);
}

static inline unsigned eblocks(size_t n) { return unsigned((n + kRedBlock - 1) / kRedBlock); }

static void launchFilter(const Dims &D, const FilterDesc &f, bool backprop, const Real *in, const Real *vars, Real *out, size_t n,
                         cudaStream_t stream, std::vector<Real> *hA, std::vector<Real> *hB, std::vector<Real> *hC, int inLo = 0, int outLo = 0) {
    if (f.type == FILTER_SMOOTHING) {
        const int r = int(size_t(f.param));
        if (backprop) smoothingKernel<true><<<eblocks(n), kRedBlock, 0, stream>>>(D, r, in, out, n, inLo, outLo);
        else smoothingKernel<false><<<eblocks(n), kRedBlock, 0, stream>>>(D, r, in, out, n, inLo, outLo);
        VFEM_CHECK_LAUNCH(); countLaunch();
    } else if (f.type == FILTER_PROJECTION) {
        if (backprop) projectionBackpropKernel<<<eblocks(n), kRedBlock, 0, stream>>>(f.param, in, vars, out, n);
        else projectionKernel<<<eblocks(n), kRedBlock, 0, stream>>>(f.param, in, out, n);
        VFEM_CHECK_LAUNCH(); countLaunch();
    } else if (f.type == FILTER_CALLBACK) {
        if (!f.cb) throw std::runtime_error("callback filter without callback");
        hA->resize(n); hB->resize(n); hC->resize(n);
        VFEM_CUDA(cudaMemcpyAsync(hA->data(), in, n * sizeof(Real), cudaMemcpyDeviceToHost, stream));
        if (backprop) VFEM_CUDA(cudaMemcpyAsync(hB->data(), vars, n * sizeof(Real), cudaMemcpyDeviceToHost, stream));
        VFEM_CUDA(cudaStreamSynchronize(stream));
        f.cb(f.ctx, backprop ? 1 : 0, hA->data(), backprop ? hB->data() : nullptr, hC->data(), n);
        VFEM_CUDA(cudaMemcpyAsync(out, hC->data(), n * sizeof(Real), cudaMemcpyHostToDevice, stream));
        VFEM_CUDA(cudaStreamSynchronize(stream));
    } else throw std::runtime_error("unknown filter type");
}

template <int N>
static Dims dimsOf(const Sim<N> &s) { Dims D; D.nd = N; for (int d = 0; d < 3; ++d) D.E[d] = d < N ? s.g.E[d] : 1; return D; }

void applyFilterStandalone(int ndim, const size_t *dims, const FilterDesc &f, const Real *xHost, Real *outHost) {
    if (ndim < 1 || ndim > 3) throw std::runtime_error("applyFilter: 1 to 3 grid dimensions supported");
    Dims D; D.nd = ndim; size_t n = 1;
    for (int d = 0; d < 3; ++d) { D.E[d] = d < ndim ? int(dims[d]) : 1; n *= size_t(D.E[d]); }
    DevBuf<Real> in(n), out(n);
    in.upload(xHost, n);
    std::vector<Real> a, b, c;
    launchFilter(D, f, false, in.p, nullptr, out.p, n, 0, &a, &b, &c);
    out.download(outHost);
}

#include "topopt_problem.inl"

template struct Objective<2>;
template struct Objective<3>;
template struct Problem<2>;
template struct Problem<3>;

} // namespace vfem
