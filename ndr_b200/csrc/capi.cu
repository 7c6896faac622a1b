// ndr_b200 — C ABI (include/vfem_b200.h) over the C++ host classes.
#include "../../include/vfem_b200.h"
#include "topopt.h"

using namespace vfem;

namespace {
thread_local std::string g_err;

struct TpsH { int dim; std::shared_ptr<Sim<2>> s2; std::shared_ptr<Sim<3>> s3; };
struct MgH { int dim; std::shared_ptr<MG<2>> m2; std::shared_ptr<MG<3>> m3; };
struct ObjH { int dim; std::shared_ptr<Objective<2>> o2; std::shared_ptr<Objective<3>> o3; };
struct CommH { std::shared_ptr<SlabComm> c; };
struct TopH { int dim; std::shared_ptr<Problem<2>> p2; std::shared_ptr<Problem<3>> p3; };

template <int N> struct Pick;
template <> struct Pick<2> {
    static std::shared_ptr<Sim<2>> &sim(TpsH *h) { return h->s2; }
    static std::shared_ptr<MG<2>> &mg(MgH *h) { return h->m2; }
    static std::shared_ptr<Objective<2>> &obj(ObjH *h) { return h->o2; }
    static std::shared_ptr<Problem<2>> &top(TopH *h) { return h->p2; }
};
template <> struct Pick<3> {
    static std::shared_ptr<Sim<3>> &sim(TpsH *h) { return h->s3; }
    static std::shared_ptr<MG<3>> &mg(MgH *h) { return h->m3; }
    static std::shared_ptr<Objective<3>> &obj(ObjH *h) { return h->o3; }
    static std::shared_ptr<Problem<3>> &top(TopH *h) { return h->p3; }
};

// Stages caller arrays: host arrays get a temporary device copy (and a copy back for outputs).
struct Staged {
    cudaStream_t stream;
    std::vector<std::pair<std::pair<double *, double *>, size_t>> outs; // (host, dev), count
    std::vector<double *> temps;
    explicit Staged(cudaStream_t s) : stream(s) {}
    const double *in(const double *p, size_t n, int ms) {
        if (ms == VFEM_DEVICE || !p) return p;
        double *d; VFEM_CUDA(cudaMalloc((void **)&d, n * sizeof(double))); temps.push_back(d);
        VFEM_CUDA(cudaMemcpyAsync(d, p, n * sizeof(double), cudaMemcpyHostToDevice, stream));
        return d;
    }
    double *out(double *p, size_t n, int ms, bool alsoIn = false) {
        if (ms == VFEM_DEVICE || !p) return p;
        double *d; VFEM_CUDA(cudaMalloc((void **)&d, n * sizeof(double))); temps.push_back(d);
        if (alsoIn) VFEM_CUDA(cudaMemcpyAsync(d, p, n * sizeof(double), cudaMemcpyHostToDevice, stream));
        outs.push_back({{p, d}, n});
        return d;
    }
    void finish() {
        for (auto &o : outs) VFEM_CUDA(cudaMemcpyAsync(o.first.first, o.first.second, o.second * sizeof(double), cudaMemcpyDeviceToHost, stream));
        // device-resident calls stay asynchronous on the library stream (the caller synchronises once, like any
        // CUDA library); calls with host arrays return only when the results have landed
        if (!outs.empty() || !temps.empty()) VFEM_CUDA(cudaStreamSynchronize(stream));
        outs.clear();
    }
    ~Staged() { for (double *t : temps) cudaFree(t); }
};
} // namespace

#define API_TRY try {
#define API_CATCH } catch (const std::exception &e) { g_err = e.what(); return 1; } catch (...) { g_err = "unknown error"; return 1; } return 0;
#define WITH_SIM(h, BODY) do { TpsH *hh_ = (TpsH *)(h); if (hh_->dim == 2) { constexpr int N = 2; auto &s = *hh_->s2; (void)N; BODY } else { constexpr int N = 3; auto &s = *hh_->s3; (void)N; BODY } } while (0)
#define WITH_MG(h, BODY) do { MgH *hh_ = (MgH *)(h); if (hh_->dim == 2) { constexpr int N = 2; auto &mg = *hh_->m2; (void)N; BODY } else { constexpr int N = 3; auto &mg = *hh_->m3; (void)N; BODY } } while (0)
#define WITH_OBJ(h, BODY) do { ObjH *hh_ = (ObjH *)(h); if (hh_->dim == 2) { constexpr int N = 2; auto &ob = *hh_->o2; (void)N; BODY } else { constexpr int N = 3; auto &ob = *hh_->o3; (void)N; BODY } } while (0)
#define WITH_TOP(h, BODY) do { TopH *hh_ = (TopH *)(h); if (hh_->dim == 2) { constexpr int N = 2; auto &tp = *hh_->p2; (void)N; BODY } else { constexpr int N = 3; auto &tp = *hh_->p3; (void)N; BODY } } while (0)

namespace vfem {
void setLastError(const std::string &msg) { g_err = msg; } // for entry points defined in other files (no_alloc.cu)
}

extern "C" {

const char *vfem_last_error(void) { return g_err.c_str(); }
int vfem_device_count(int *count) { API_TRY VFEM_CUDA(cudaGetDeviceCount(count)); API_CATCH }
int vfem_set_device(int device) { API_TRY VFEM_CUDA(cudaSetDevice(device)); API_CATCH }
int vfem_device_synchronize(void) { API_TRY VFEM_CUDA(cudaDeviceSynchronize()); API_CATCH }
unsigned long long vfem_launch_count(void) { return g_launchCount; }

// ---- TensorProductSimulator -------------------------------------------------------------------------
int vfem_tps_create(int dim, const size_t *degrees, const double *bbmin, const double *bbmax, const size_t *elems, vfem_tps_t *out) {
    API_TRY
    bool q1 = (dim == 2 || dim == 3);
    for (int d = 0; q1 && d < dim; ++d) q1 = degrees[d] == 1;
    if (!q1) throw std::runtime_error("No template instantiation matching degreesPerDimension!"); // VoxelFEM.cc:238
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        throw std::runtime_error("libvfem_b200: no CUDA device available (this library has no CPU fallback)");
    TpsH *h = new TpsH();
    h->dim = dim;
    try {
        if (dim == 2) h->s2 = std::make_shared<Sim<2>>(bbmin, bbmax, elems);
        else h->s3 = std::make_shared<Sim<3>>(bbmin, bbmax, elems);
    } catch (...) { delete h; throw; }
    *out = h;
    API_CATCH
}
int vfem_tps_create_slab(int dim, const size_t *degrees, const double *bbmin, const double *bbmax, const size_t *elems_global,
                         size_t layer_lo, size_t layer_hi, vfem_tps_t *out) {
    API_TRY
    bool q1 = (dim == 2 || dim == 3);
    for (int d = 0; q1 && d < dim; ++d) q1 = degrees[d] == 1;
    if (!q1) throw std::runtime_error("No template instantiation matching degreesPerDimension!");
    if (layer_hi <= layer_lo || layer_hi > elems_global[0]) throw std::runtime_error("invalid slab layer range");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        throw std::runtime_error("libvfem_b200: no CUDA device available (this library has no CPU fallback)");
    size_t el[3] = {layer_hi - layer_lo, elems_global[1], dim == 3 ? elems_global[2] : 0};
    TpsH *h = new TpsH();
    h->dim = dim;
    try {
        if (dim == 2) h->s2 = std::make_shared<Sim<2>>(bbmin, bbmax, el, elems_global[0], layer_lo);
        else h->s3 = std::make_shared<Sim<3>>(bbmin, bbmax, el, elems_global[0], layer_lo);
    } catch (...) { delete h; throw; }
    *out = h;
    API_CATCH
}
void vfem_tps_destroy(vfem_tps_t h) { delete (TpsH *)h; }
int vfem_tps_dim(vfem_tps_t h) { return ((TpsH *)h)->dim; }
size_t vfem_tps_num_nodes(vfem_tps_t h) { size_t r = 0; WITH_SIM(h, r = s.numNodes();); return r; }
size_t vfem_tps_num_elements(vfem_tps_t h) { size_t r = 0; WITH_SIM(h, r = s.numElements();); return r; }
int vfem_tps_set_isotropic(vfem_tps_t h, double young, double poisson) { API_TRY WITH_SIM(h, s.setIsotropic(young, poisson);); API_CATCH }
int vfem_tps_get_lame(vfem_tps_t h, double *lambda, double *mu) { API_TRY WITH_SIM(h, *lambda = s.lambda; *mu = s.mu;); API_CATCH }
int vfem_tps_set_simp(vfem_tps_t h, double E_0, double E_min, double gamma) {
    API_TRY WITH_SIM(h, s.E0 = E_0; s.Emin = E_min; s.gamma = gamma; s.touchStiffness();); API_CATCH }
int vfem_tps_get_simp(vfem_tps_t h, double *E_0, double *E_min, double *gamma) { API_TRY WITH_SIM(h, *E_0 = s.E0; *E_min = s.Emin; *gamma = s.gamma;); API_CATCH }
int vfem_tps_add_force_region(vfem_tps_t h, const double *lo, const double *hi, const double *value) { API_TRY WITH_SIM(h, s.addForceRegion(lo, hi, value);); API_CATCH }
int vfem_tps_add_force_region_split(vfem_tps_t h, const double *lo, const double *hi, const double *value, size_t divisor) {
    API_TRY WITH_SIM(h, s.addForceRegion(lo, hi, value, divisor);); API_CATCH }
int vfem_tps_count_region_nodes(vfem_tps_t h, const double *lo, const double *hi, size_t plane_lo, size_t plane_hi, size_t *count) {
    API_TRY WITH_SIM(h, *count = s.countRegionNodes(lo, hi, plane_lo, plane_hi);); API_CATCH }
int vfem_tps_add_dirichlet_region(vfem_tps_t h, const double *lo, const double *hi, const uint8_t *comp_mask, const double *value) {
    API_TRY WITH_SIM(h, s.addDirichletRegion(lo, hi, comp_mask, value);); API_CATCH }
int vfem_tps_get_dirichlet(vfem_tps_t h, uint8_t *mask, double *values) {
    API_TRY WITH_SIM(h,
        for (size_t n = 0; n < s.numNodes(); ++n) for (int c = 0; c < N; ++c) {
            if (mask) mask[n * N + c] = (s.dmask[n] >> c) & 1;
            if (values) values[n * N + c] = s.dval[n * N + c];
        });
    API_CATCH }
int vfem_tps_set_dirichlet(vfem_tps_t h, const uint8_t *mask, const double *values) {
    API_TRY WITH_SIM(h,
        for (size_t n = 0; n < s.numNodes(); ++n) {
            if (mask) { uint8_t m = 0; for (int c = 0; c < N; ++c) if (mask[n * N + c]) m |= uint8_t(1u << c); s.dmask[n] = m; }
            if (values) for (int c = 0; c < N; ++c) s.dval[n * N + c] = values[n * N + c];
        }
        s.bcDirty = true; ++s.bcVersion;);
    API_CATCH }
int vfem_tps_get_force_mask(vfem_tps_t h, uint8_t *mask) { API_TRY WITH_SIM(h, std::copy(s.hasForce.begin(), s.hasForce.end(), mask);); API_CATCH }
int vfem_tps_build_load_vector(vfem_tps_t h, double *f, int ms) {
    API_TRY WITH_SIM(h,
        if (ms == VFEM_HOST) s.buildLoadVectorHost(f);
        else { std::vector<double> t(s.numDofs()); s.buildLoadVectorHost(t.data());
               VFEM_CUDA(cudaMemcpy(f, t.data(), t.size() * sizeof(double), cudaMemcpyHostToDevice)); });
    API_CATCH }
int vfem_tps_node_position(vfem_tps_t h, size_t ni, double *p) { API_TRY WITH_SIM(h, if (ni >= s.numNodes()) throw std::runtime_error("node index out of range"); s.nodePosition(ni, p);); API_CATCH }
int vfem_tps_get_k0(vfem_tps_t h, double *k0) { API_TRY WITH_SIM(h, std::copy(s.K0.k, s.K0.k + Sim<N>::KS * Sim<N>::KS, k0);); API_CATCH }
int vfem_tps_set_densities(vfem_tps_t h, const double *rho, int ms) { API_TRY WITH_SIM(h, s.setDensities(rho, MemSpace(ms));); API_CATCH }
int vfem_tps_set_uniform_densities(vfem_tps_t h, double v) { API_TRY WITH_SIM(h, s.setUniformDensities(v);); API_CATCH }
int vfem_tps_set_element_density(vfem_tps_t h, size_t ei, double v) { API_TRY WITH_SIM(h, s.setElementDensity(ei, v);); API_CATCH }
int vfem_tps_element_density(vfem_tps_t h, size_t ei, double *v) { API_TRY WITH_SIM(h, *v = s.elementDensity(ei);); API_CATCH }
int vfem_tps_get_densities(vfem_tps_t h, double *rho, int ms) { API_TRY WITH_SIM(h, s.getDensities(rho, MemSpace(ms));); API_CATCH }
int vfem_tps_apply_k(vfem_tps_t h, const double *u, double *out, int ms) {
    API_TRY WITH_SIM(h, if (ms == VFEM_HOST) { s.applyKHost(u, out); } else { s.applyK(u, out); }); API_CATCH }
int vfem_tps_solve(vfem_tps_t h, const double *f, double *u, int ms) {
    API_TRY WITH_SIM(h, Staged st(s.stream); const double *df = st.in(f, s.numDofs(), ms); double *du = st.out(u, s.numDofs(), ms);
                     s.solveDirect(df, du); st.finish();); API_CATCH }
int vfem_tps_apply_k_ring(vfem_tps_t h, const double *const *us, double *const *outs, size_t nsets, size_t start, size_t count) {
    API_TRY WITH_SIM(h, for (size_t k = 0; k < count; ++k) s.applyK(us[(start + k) % nsets], outs[(start + k) % nsets]);); API_CATCH }
int vfem_tps_set_apply_variant(vfem_tps_t h, int variant) { API_TRY WITH_SIM(h, s.applyVariant = variant;); API_CATCH }
int vfem_tps_apply_info(vfem_tps_t h, int *info) {
    API_TRY WITH_SIM(h, info[0] = s.useMarch() ? 1 : 0; info[1] = s.khatStructure; info[2] = s.marchPlan.cfg; info[3] = s.marchPlan.XC;
                     info[4] = int(s.marchPlan.grid.x); info[5] = int(s.marchPlan.grid.y); info[6] = int(s.marchPlan.grid.z);); API_CATCH }
int vfem_tps_compliance_gradient(vfem_tps_t h, const double *u, double *g, int ms) {
    API_TRY WITH_SIM(h, Staged st(s.stream); const double *du = st.in(u, s.numDofs(), ms); double *dg = st.out(g, s.numElements(), ms);
                     s.complianceGradient(du, dg); st.finish();); API_CATCH }

#include "capi_mg.inl"

} // extern "C"
