// ndr_b200 — C ABI, multigrid / objective / problem entry points (included by capi.cu inside extern "C").

int vfem_mg_create(vfem_tps_t tps, size_t levels, vfem_mg_t *out) {
    API_TRY
    TpsH *t = (TpsH *)tps;
    MgH *h = new MgH(); h->dim = t->dim;
    try {
        if (t->dim == 2) h->m2 = std::make_shared<MG<2>>(t->s2, levels);
        else h->m3 = std::make_shared<MG<3>>(t->s3, levels);
    } catch (...) { delete h; throw; }
    *out = h;
    API_CATCH
}
// ---- slab decomposition -------------------------------------------------------------------------------------
int vfem_comm_nccl_unique_id(uint8_t *id128) { API_TRY newNcclUniqueId(id128); API_CATCH }
int vfem_comm_create_nccl(const uint8_t *id128, int rank, int world, vfem_comm_t *out) {
    API_TRY CommH *h = new CommH(); try { h->c = makeNcclComm(id128, rank, world); } catch (...) { delete h; throw; } *out = h; API_CATCH }
int vfem_comm_create_threads(int world, vfem_comm_t *out_world) {
    API_TRY auto cs = makeThreadComms(world); for (int r = 0; r < world; ++r) { CommH *h = new CommH(); h->c = cs[r]; out_world[r] = h; } API_CATCH }
void vfem_comm_destroy(vfem_comm_t h) { delete (CommH *)h; }
int vfem_mg_create_slab(vfem_tps_t tps, size_t levels, vfem_comm_t comm, vfem_mg_t *out) {
    API_TRY
    TpsH *t = (TpsH *)tps;
    MgH *h = new MgH(); h->dim = t->dim;
    std::shared_ptr<SlabComm> c = comm ? ((CommH *)comm)->c : nullptr;
    try {
        if (t->dim == 2) h->m2 = std::make_shared<MG<2>>(t->s2, levels, c);
        else h->m3 = std::make_shared<MG<3>>(t->s3, levels, c);
    } catch (...) { delete h; throw; }
    *out = h;
    API_CATCH
}
int vfem_mg_owned_planes(vfem_mg_t h, size_t l, size_t *first, size_t *last) {
    API_TRY WITH_MG(h, *first = size_t(mg.lv.at(l).g.own0); *last = size_t(mg.lv.at(l).g.own1);); API_CATCH }
int vfem_mg_exchange(vfem_mg_t h, size_t l, double *v, int ms) {
    API_TRY WITH_MG(h, if (ms != VFEM_DEVICE) throw std::runtime_error("vfem_mg_exchange needs a device array");
                    mg.exchange(l, v, MG<N>::XUP | MG<N>::XDOWN);); API_CATCH }
int vfem_mg_apply_k_halo(vfem_mg_t h, double *u, double *out, int ms) {
    API_TRY WITH_MG(h, if (ms != VFEM_DEVICE) throw std::runtime_error("vfem_mg_apply_k_halo needs device arrays");
                    mg.template exchangeApplyFine<0>(u, nullptr, out);); API_CATCH }
int vfem_mg_apply_k_halo_ring(vfem_mg_t h, double *const *us, double *const *outs, size_t nsets, size_t start, size_t count) {
    // at most 16 halo'd applies (each: an NCCL group + kernels on two streams) are kept in flight: two ranks that queue thousands of
    // them without ever yielding were seen to stall each other (r02c); the wait costs ~1 % of a 16-apply burst
    API_TRY WITH_MG(h, for (size_t k = 0; k < count; ++k) {
                        mg.template exchangeApplyFine<0>(us[(start + k) % nsets], nullptr, outs[(start + k) % nsets]);
                        if (mg.slab() && (k + 1) % 16 == 0 && k + 1 < count) VFEM_CUDA(cudaStreamSynchronize(mg.stream));
                    }); API_CATCH }
int vfem_mg_dot(vfem_mg_t h, const double *a, const double *b, int ms, double *result) {
    API_TRY WITH_MG(h, size_t n = mg.levelDofs(0); Staged st(mg.stream); const double *da = st.in(a, n, ms); const double *db = st.in(b, n, ms);
                    *result = mg.dot(da, db, n);); API_CATCH }
void vfem_mg_destroy(vfem_mg_t h) { delete (MgH *)h; }
size_t vfem_mg_num_levels(vfem_mg_t h) { size_t r = 0; WITH_MG(h, r = mg.numLevels();); return r; }
size_t vfem_mg_level_nodes(vfem_mg_t h, size_t l) { size_t r = 0; WITH_MG(h, r = l < mg.lv.size() ? size_t(mg.lv[l].g.nn) : 0;); return r; }
int vfem_mg_level_grid(vfem_mg_t h, size_t l, size_t *elems) { API_TRY WITH_MG(h, for (int d = 0; d < N; ++d) elems[d] = size_t(mg.lv.at(l).g.E[d]);); API_CATCH }
int vfem_mg_level_dirichlet_mask(vfem_mg_t h, size_t l, uint8_t *mask) {
    API_TRY WITH_MG(h,
        const std::vector<uint8_t> &m = (l == 0) ? mg.sim->dmask : mg.lv.at(l).dmask;
        for (size_t n = 0; n < m.size(); ++n) for (int c = 0; c < N; ++c) mask[n * N + c] = (m[n] >> c) & 1;);
    API_CATCH }
int vfem_mg_set_symmetric_gs(vfem_mg_t h, int symmetric) { API_TRY WITH_MG(h, mg.symmetricGS = symmetric != 0;); API_CATCH }
int vfem_mg_set_flags(vfem_mg_t h, int build_block, int build_finest) {
    API_TRY WITH_MG(h, mg.buildBlockStiffnessMatrices = build_block != 0; mg.buildFinestBlockStiffnessMatrix = build_finest != 0;); API_CATCH }
int vfem_mg_update_stiffness(vfem_mg_t h) { API_TRY WITH_MG(h, mg.syncMasks(); mg.updateElementStiffnessMatrices(true); VFEM_CUDA(cudaStreamSynchronize(mg.stream));); API_CATCH }

// debug-hook operators of levels >= 1 need the hierarchy: refresh it lazily like the reference's applyK(l) (MG.hh:353-358)
#define MG_READY(l) do { mg.syncMasks(); if ((l) > 0) mg.updateElementStiffnessMatrices(false); } while (0)

int vfem_mg_apply_k(vfem_mg_t h, size_t l, const double *u, double *out, int ms) {
    API_TRY WITH_MG(h, MG_READY(l); size_t n = mg.levelDofs(l); Staged st(mg.stream);
                    const double *du = st.in(u, n, ms); double *dout = st.out(out, n, ms); mg.applyK(l, du, dout); st.finish();); API_CATCH }
int vfem_mg_compute_residual(vfem_mg_t h, size_t l, const double *u, const double *b, double *out, int ms) {
    API_TRY WITH_MG(h, MG_READY(l); size_t n = mg.levelDofs(l); Staged st(mg.stream);
                    const double *du = st.in(u, n, ms); const double *db = st.in(b, n, ms); double *dout = st.out(out, n, ms);
                    mg.computeResidual(l, du, db, dout); st.finish();); API_CATCH }
int vfem_mg_zero_dirichlet(vfem_mg_t h, size_t l, double *u, int ms) {
    API_TRY WITH_MG(h, mg.syncMasks(); size_t n = mg.levelDofs(l); Staged st(mg.stream); double *du = st.out(u, n, ms, true);
                    mg.zeroOutDirichlet(l, du); st.finish();); API_CATCH }
int vfem_mg_smooth_serial(vfem_mg_t h, size_t l, double *u, const double *b, int forward, int ms) {
    API_TRY WITH_MG(h, MG_READY(l); size_t n = mg.levelDofs(l); Staged st(mg.stream);
                    double *du = st.out(u, n, ms, true); const double *db = st.in(b, n, ms);
                    mg.smoothSerial(l, du, db, forward != 0); st.finish();); API_CATCH }
int vfem_mg_smooth_colored(vfem_mg_t h, size_t l, double *u, const double *b, int forward, int ms) {
    API_TRY WITH_MG(h, MG_READY(l); size_t n = mg.levelDofs(l); Staged st(mg.stream);
                    double *du = st.out(u, n, ms, true); const double *db = st.in(b, n, ms);
                    mg.smoothColoured(l, du, db, forward != 0); st.finish();); API_CATCH }
int vfem_mg_restrict(vfem_mg_t h, size_t lf, const double *vf, double *outc, int ms) {
    API_TRY WITH_MG(h, Staged st(mg.stream); const double *dv = st.in(vf, mg.levelDofs(lf), ms); double *dc = st.out(outc, mg.levelDofs(lf + 1), ms);
                    mg.restriction(lf, dv, dc); st.finish();); API_CATCH }
int vfem_mg_interpolate(vfem_mg_t h, size_t lf, const double *xc, double *xf, int accumulate, int ms) {
    API_TRY WITH_MG(h, Staged st(mg.stream); const double *dc = st.in(xc, mg.levelDofs(lf + 1), ms); double *df = st.out(xf, mg.levelDofs(lf), ms, accumulate != 0);
                    mg.interpolate(lf, dc, df, accumulate != 0); st.finish();); API_CATCH }
int vfem_mg_level_stencil(vfem_mg_t h, size_t l, double *out) {
    API_TRY WITH_MG(h, if (l == 0 || l >= mg.lv.size()) throw std::runtime_error("level has no assembled operator");
                    MG_READY(l);
                    { std::vector<double> h = mg.lv[l].S.toHost(mg.stream); const int nn = mg.lv[l].g.nn; constexpr int NST = ipow3(N);
                      for (int n = 0; n < nn; ++n) for (int k = 0; k < NST; ++k) for (int i = 0; i < N * N; ++i)
                          out[size_t(k * N * N + i) * nn + n] = h[stencilIndex<N>(n, k, i)]; }); API_CATCH }
int vfem_mg_debug_get(vfem_mg_t h, size_t l, double *x, double *b) {
    API_TRY WITH_MG(h, if (x) mg.lv.at(l).x.download(x, mg.stream); if (b) mg.lv.at(l).b.download(b, mg.stream);); API_CATCH }

int vfem_mg_time_op(vfem_mg_t h, int op, size_t l, int reps, double *ms_per_call) {
    API_TRY WITH_MG(h,
        mg.syncMasks(); mg.updateElementStiffnessMatrices(false);
        auto &L = mg.lv.at(l);
        auto run = [&]() {
            switch (op) {
            case 0: mg.smooth(l, L.x.p, L.b.p, true, 1); break;
            case 1: mg.computeResidual(l, L.x.p, L.b.p, L.r.p); break;
            case 2: mg.restriction(l, L.r.p, mg.lv.at(l + 1).b.p); break;
            case 3: mg.interpolate(l, mg.lv.at(l + 1).x.p, L.x.p, true); break;
            case 4: mg.coarseSolve(mg.lv.back().b.p, mg.lv.back().x.p); break;
            case 5: mg.updateElementStiffnessMatrices(true); break;
            case 6: mg.precondition(mg.p_r.p, 1, 2, true); break;
            case 7: mg.applyKMasked(l, L.x.p, L.r.p); break;
            default: throw std::runtime_error("vfem_mg_time_op: unknown op");
            }
        };
        run();
        cudaEvent_t e0; cudaEvent_t e1; VFEM_CUDA(cudaEventCreate(&e0)); VFEM_CUDA(cudaEventCreate(&e1));
        VFEM_CUDA(cudaStreamSynchronize(mg.stream));
        VFEM_CUDA(cudaEventRecord(e0, mg.stream));
        for (int r = 0; r < reps; ++r) run();
        VFEM_CUDA(cudaEventRecord(e1, mg.stream));
        VFEM_CUDA(cudaEventSynchronize(e1));
        float ms = 0.f; VFEM_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        cudaEventDestroy(e0); cudaEventDestroy(e1);
        *ms_per_call = double(ms) / reps;);
    API_CATCH }

int vfem_mg_solve(vfem_mg_t h, const double *u, const double *f, size_t num_steps, size_t nu, int stiffness_updated, int zero_dirichlet,
                  vfem_solve_cb cb, void *cb_ctx, int fmg, double *out, int ms) {
    API_TRY WITH_MG(h,
        size_t n = mg.levelDofs(0); Staged st(mg.stream);
        const double *du = st.in(u, n, ms); const double *df = st.in(f, n, ms); double *dout = st.out(out, n, ms);
        std::vector<double> hx;
        typename MG<N>::SolveCb scb;
        if (cb) scb = [&](size_t it, const double *xdev) {
            hx.resize(n);
            VFEM_CUDA(cudaMemcpyAsync(hx.data(), xdev, n * sizeof(double), cudaMemcpyDeviceToHost, mg.stream));
            VFEM_CUDA(cudaStreamSynchronize(mg.stream));
            cb(cb_ctx, it, hx.data()); };
        mg.solve(du, df, num_steps, nu, stiffness_updated != 0, zero_dirichlet != 0, scb, fmg != 0, dout);
        st.finish(););
    API_CATCH }

int vfem_mg_pcg(vfem_mg_t h, const double *u, const double *b, size_t max_iter, double tol, vfem_pcg_cb cb, void *cb_ctx, int want_vectors,
                size_t mg_iterations, size_t nu, int fmg, double *out, int ms, size_t *iterations, double *residual_norm) {
    API_TRY WITH_MG(h,
        size_t n = mg.levelDofs(0); Staged st(mg.stream);
        const double *du = st.in(u, n, ms); const double *db = st.in(b, n, ms); double *dout = st.out(out, n, ms);
        std::vector<double> hx; std::vector<double> hr;
        typename MG<N>::PcgCb pcb;
        if (cb) pcb = [&](size_t it, const double *xdev, const double *rdev, double rn) {
            if (want_vectors) { // it_callback(i, x, r) hands host arrays to the caller (MG.hh:726-729)
                hx.resize(n); hr.resize(n);
                VFEM_CUDA(cudaMemcpyAsync(hx.data(), xdev, n * sizeof(double), cudaMemcpyDeviceToHost, mg.stream));
                VFEM_CUDA(cudaMemcpyAsync(hr.data(), rdev, n * sizeof(double), cudaMemcpyDeviceToHost, mg.stream));
                VFEM_CUDA(cudaStreamSynchronize(mg.stream));
                cb(cb_ctx, it, hx.data(), hr.data(), rn);
            } else cb(cb_ctx, it, nullptr, nullptr, rn); };
        double rn = 0;
        size_t its = mg.pcg(du, db, max_iter, tol, pcb, mg_iterations, nu, fmg != 0, dout, &rn);
        if (iterations) *iterations = its;
        if (residual_norm) *residual_norm = rn;
        st.finish(););
    API_CATCH }

// ---- objective ----------------------------------------------------------------------------------------------
int vfem_obj_create(vfem_mg_t mgh, vfem_obj_t *out) {
    API_TRY
    MgH *m = (MgH *)mgh;
    ObjH *h = new ObjH(); h->dim = m->dim;
    try {
        if (m->dim == 2) h->o2 = std::make_shared<Objective<2>>(m->m2);
        else h->o3 = std::make_shared<Objective<3>>(m->m3);
    } catch (...) { delete h; throw; }
    *out = h;
    API_CATCH
}
int vfem_obj_create_direct(vfem_tps_t tps, vfem_obj_t *out) {
    API_TRY
    TpsH *t = (TpsH *)tps;
    ObjH *h = new ObjH(); h->dim = t->dim;
    try {
        if (t->dim == 2) h->o2 = std::make_shared<Objective<2>>(t->s2);
        else h->o3 = std::make_shared<Objective<3>>(t->s3);
    } catch (...) { delete h; throw; }
    *out = h;
    API_CATCH
}
void vfem_obj_destroy(vfem_obj_t h) { delete (ObjH *)h; }
int vfem_obj_set_params(vfem_obj_t h, size_t cg_iter, double tol, size_t mg_it, size_t nu, int fmg, int zero_init) {
    API_TRY WITH_OBJ(h, ob.cgIter = cg_iter; ob.tol = tol; ob.mgIterations = mg_it; ob.mgSmoothingIterations = nu; ob.fullMultigrid = fmg != 0; ob.zeroInit = zero_init != 0;); API_CATCH }
int vfem_obj_get_params(vfem_obj_t h, size_t *cg_iter, double *tol, size_t *mg_it, size_t *nu, int *fmg, int *zero_init) {
    API_TRY WITH_OBJ(h, *cg_iter = ob.cgIter; *tol = ob.tol; *mg_it = ob.mgIterations; *nu = ob.mgSmoothingIterations; *fmg = ob.fullMultigrid; *zero_init = ob.zeroInit;); API_CATCH }
int vfem_obj_set_residual_cb(vfem_obj_t h, vfem_residual_cb cb, void *ctx) { API_TRY WITH_OBJ(h, ob.residual_cb = cb; ob.residual_ctx = ctx;); API_CATCH }
int vfem_obj_compliance(vfem_obj_t h, double *J) { API_TRY WITH_OBJ(h, *J = ob.compliance();); API_CATCH }
int vfem_obj_gradient(vfem_obj_t h, double *g, int ms) {
    API_TRY WITH_OBJ(h, Staged st(ob.sim->stream); double *dg = st.out(g, ob.sim->numElements(), ms); ob.gradient(dg); st.finish();); API_CATCH }
static int copyOut(const double *dev, double *dst, size_t n, int ms) {
    API_TRY VFEM_CUDA(cudaMemcpy(dst, dev, n * sizeof(double), ms == VFEM_HOST ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice)); API_CATCH }
int vfem_obj_get_u(vfem_obj_t h, double *u, int ms) { int rc = 0; WITH_OBJ(h, rc = copyOut(ob.u.p, u, ob.u.n, ms);); return rc; }
int vfem_obj_get_f(vfem_obj_t h, double *f, int ms) { int rc = 0; WITH_OBJ(h, rc = copyOut(ob.f.p, f, ob.f.n, ms);); return rc; }
int vfem_obj_update_cache(vfem_obj_t h, const double *x, int ms) {
    API_TRY WITH_OBJ(h, Staged st(ob.sim->stream); const double *dx = st.in(x, ob.sim->numElements(), ms); ob.updateCache(dx); st.finish();); API_CATCH }
int vfem_obj_last_solve(vfem_obj_t h, size_t *iterations, double *rn) { API_TRY WITH_OBJ(h, if (iterations) *iterations = ob.lastIterations; if (rn) *rn = ob.lastResidual;); API_CATCH }

// ---- problem + OC ----------------------------------------------------------------------------------------------
int vfem_top_create(vfem_tps_t tps, vfem_obj_t objective, int nc, const double *vf, int nf, const int *ftypes, const double *fparams,
                    const vfem_filter_cb *fcbs, void *const *fctxs, vfem_top_t *out) {
    API_TRY
    TpsH *t = (TpsH *)tps; ObjH *o = (ObjH *)objective;
    if (t->dim != o->dim) throw std::runtime_error("simulator/objective dimension mismatch");
    std::vector<double> vfs(vf, vf + nc);
    std::vector<FilterDesc> fl(nf);
    for (int i = 0; i < nf; ++i) { fl[i].type = ftypes[i]; fl[i].param = fparams[i]; fl[i].cb = fcbs ? fcbs[i] : nullptr; fl[i].ctx = fctxs ? fctxs[i] : nullptr; }
    TopH *h = new TopH(); h->dim = t->dim;
    try {
        if (t->dim == 2) h->p2 = std::make_shared<Problem<2>>(t->s2, o->o2, vfs, fl);
        else h->p3 = std::make_shared<Problem<3>>(t->s3, o->o3, vfs, fl);
    } catch (...) { delete h; throw; }
    *out = h;
    API_CATCH
}
void vfem_top_destroy(vfem_top_t h) { delete (TopH *)h; }
size_t vfem_top_num_vars(vfem_top_t h) { size_t r = 0; WITH_TOP(h, r = tp.numVars();); return r; }
int vfem_top_set_filter_param(vfem_top_t h, int filter, double param) { API_TRY WITH_TOP(h, tp.filters.at(filter).param = param;); API_CATCH }
int vfem_top_set_vars(vfem_top_t h, const double *x, int ms, int force, int *changed) {
    API_TRY WITH_TOP(h, bool ch = tp.setVars(x, MemSpace(ms), force != 0); if (changed) *changed = ch ? 1 : 0;); API_CATCH }
int vfem_top_get_vars(vfem_top_t h, int which, double *out, int ms) {
    int rc = 0; WITH_TOP(h, if (which >= int(tp.cached.size())) { g_err = "filter stage out of range"; return 1; }
                         auto &c = which < 0 ? tp.cached.back() : tp.cached[which]; rc = copyOut(c.p, out, tp.numVars(), ms);); return rc; }
int vfem_top_objective(vfem_top_t h, double *J) { API_TRY WITH_TOP(h, *J = tp.evaluateObjective();); API_CATCH }
int vfem_top_objective_gradient(vfem_top_t h, double *g, int ms) {
    API_TRY WITH_TOP(h, Staged st(tp.sim->stream); double *dg = st.out(g, tp.numVars(), ms); tp.evaluateObjectiveGradient(dg); st.finish();); API_CATCH }
int vfem_top_constraints(vfem_top_t h, double *c) { API_TRY WITH_TOP(h, for (size_t i = 0; i < tp.volumeFractions.size(); ++i) c[i] = tp.evaluateConstraint(i);); API_CATCH }
int vfem_top_constraints_jacobian(vfem_top_t h, double *jac, int ms) {
    API_TRY WITH_TOP(h, size_t ne = tp.numVars(); Staged st(tp.sim->stream); double *dj = st.out(jac, ne * tp.volumeFractions.size(), ms);
                     for (size_t i = 0; i < tp.volumeFractions.size(); ++i) tp.evaluateConstraintJacobianRow(i, dj + i * ne); st.finish();); API_CATCH }
int vfem_oc_reset(vfem_top_t h) { API_TRY WITH_TOP(h, tp.lambdaMin = 1.0; tp.lambdaMax = 2.0;); API_CATCH }
int vfem_oc_step(vfem_top_t h, double m, double ctol, double *lambda, int *nevals) {
    API_TRY WITH_TOP(h, tp.ocStep(m, ctol); if (lambda) *lambda = tp.lastLambda; if (nevals) *nevals = tp.lastBisections;); API_CATCH }

int vfem_apply_filter(int type, double param, int ndim, const size_t *dims, const double *x, double *out) {
    API_TRY FilterDesc f; f.type = type; f.param = param; applyFilterStandalone(ndim, dims, f, x, out); API_CATCH }
