/* vfem_b200.h — C ABI of libvfem_b200.so, the B200-native (sm_100a) replacement for the elasticity
 * solve of Nikronic/ndr's VoxelFEM.  This is the drop-in boundary: every entry point names the
 * reference interface it replaces (VoxelFEM/python_bindings/VoxelFEM.cc = "VoxelFEM.cc",
 * TensorProductSimulator.hh = "TPS.hh", MultigridSolver.hh = "MG.hh").
 *
 * Conventions
 *  - every call returns 0 on success, non-zero on failure; vfem_last_error() gives the message of the
 *    last failure on the calling thread (the reference throws std::runtime_error with the same text);
 *  - handles are opaque; the caller owns all array buffers, the library owns the handles;
 *  - vector fields are (numNodes, N) row-major float64 (TPS.hh:227), densities / gradients are
 *    (numElements,) float64 in NDVector order (first axis slowest, NDVector.hh:279-293);
 *  - `memspace` selects where the caller's arrays live: VFEM_HOST (copied through) or VFEM_DEVICE
 *    (CUDA pointers on the library's device, used in place — the DLPack zero-copy path);
 *  - one host thread drives one problem (same contract as the reference, which holds the GIL).
 */
#ifndef VFEM_B200_H
#define VFEM_B200_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define VFEM_HOST 0
#define VFEM_DEVICE 1

typedef void *vfem_tps_t;     /* TensorProductSimulator<1,..>          (VoxelFEM.cc:48-91)   */
typedef void *vfem_mg_t;      /* MultigridSolver<1,..>                 (VoxelFEM.cc:94-131)  */
typedef void *vfem_obj_t;     /* MultigridComplianceObjective          (VoxelFEM.cc:196-207) */
typedef void *vfem_top_t;     /* TopologyOptimizationProblem           (VoxelFEM.cc:170-183) */
typedef void *vfem_comm_t;    /* slab transport: no reference counterpart (the reference is single-node TBB)  */

const char *vfem_last_error(void);
int vfem_device_count(int *count);
int vfem_set_device(int device);
int vfem_device_synchronize(void);
unsigned long long vfem_launch_count(void);       /* kernels launched by this library so far */

/* ---- TensorProductSimulator ------------------------------------------------------------------ */
/* factory pyVoxelFEM.TensorProductSimulator(degrees, bbox, elems) (VoxelFEM.cc:234-240); only Q1
 * (all degrees 1) in 2-D and 3-D is instantiated, as in the reference (VoxelFEM.cc:226-229). */
int vfem_tps_create(int dim, const size_t *degrees, const double *bbox_min, const double *bbox_max, const size_t *elems, vfem_tps_t *out);
void vfem_tps_destroy(vfem_tps_t h);
int vfem_tps_dim(vfem_tps_t h);
size_t vfem_tps_num_nodes(vfem_tps_t h);            /* numNodes      */
size_t vfem_tps_num_elements(vfem_tps_t h);         /* numElements   */
int vfem_tps_set_isotropic(vfem_tps_t h, double young, double poisson); /* readMaterial -> ETensor::setIsotropic (TPS.hh:326-330) */
int vfem_tps_get_lame(vfem_tps_t h, double *lambda, double *mu);
int vfem_tps_set_simp(vfem_tps_t h, double E_0, double E_min, double gamma); /* E_0 / E_min / gamma setters (TPS.hh:1170-1172) */
int vfem_tps_get_simp(vfem_tps_t h, double *E_0, double *E_min, double *gamma);
int vfem_tps_add_force_region(vfem_tps_t h, const double *lo, const double *hi, const double *value);       /* applyDisplacementsAndLoads (TPS.hh:381-389) */
int vfem_tps_add_dirichlet_region(vfem_tps_t h, const double *lo, const double *hi, const uint8_t *comp_mask, const double *value); /* TPS.hh:391-406 */
int vfem_tps_get_dirichlet(vfem_tps_t h, uint8_t *mask_nnxN, double *values_nnxN);   /* dirichletMask / dirichletValues getters */
int vfem_tps_set_dirichlet(vfem_tps_t h, const uint8_t *mask_nnxN, const double *values_nnxN); /* setters (either may be NULL) */
int vfem_tps_get_force_mask(vfem_tps_t h, uint8_t *mask_nn);                         /* getForceMask */
int vfem_tps_build_load_vector(vfem_tps_t h, double *f, int memspace);              /* buildLoadVector (TPS.hh:893-901) */
int vfem_tps_node_position(vfem_tps_t h, size_t ni, double *p);                     /* nodePosition */
int vfem_tps_get_k0(vfem_tps_t h, double *k0);                                      /* fullDensityElementStiffnessMatrix */
int vfem_tps_set_densities(vfem_tps_t h, const double *rho, int memspace);          /* setElementDensities (TPS.hh:1157-1161) */
int vfem_tps_set_uniform_densities(vfem_tps_t h, double v);                         /* setUniformDensities */
int vfem_tps_set_element_density(vfem_tps_t h, size_t ei, double v);                /* setElementDensity */
int vfem_tps_element_density(vfem_tps_t h, size_t ei, double *v);                   /* elementDensity */
int vfem_tps_get_densities(vfem_tps_t h, double *rho, int memspace);                /* getDensities */
int vfem_tps_apply_k(vfem_tps_t h, const double *u, double *out, int memspace);     /* applyK (TPS.hh:905-952) */
int vfem_tps_solve(vfem_tps_t h, const double *f, double *u, int memspace);          /* solve(f) (TPS.hh:834-865): Dirichlet rows/cols removed, block LDL^T refreshed once per density update */
/* `count` applyK calls issued back to back by one host call (no reference counterpart; a caller that applies K to many
 * fields, and bench.py's timed loop): call k reads us[(start + k) % nsets] and writes outs[(start + k) % nsets]; device arrays */
int vfem_tps_apply_k_ring(vfem_tps_t h, const double *const *us, double *const *outs, size_t nsets, size_t start, size_t count);
/* kernel selection for the fine-level apply (no reference counterpart): variant 0 = auto (element-marching
 * Walsh-Hadamard kernel when K0 has the required symmetry), 1 = per-node gather kernel, 2 = marching kernel
 * with dense 3x3 blocks (the variant used when K0 lacks the isotropic zero pattern).  info[7] =
 * {march kernel in use, K0 structure (0 none / 1 dense blocks / 2 isotropic pattern), tile shape id, layers per
 * chunk, grid x, grid y, grid z} */
int vfem_tps_set_apply_variant(vfem_tps_t h, int variant);
int vfem_tps_apply_info(vfem_tps_t h, int *info);
int vfem_tps_compliance_gradient(vfem_tps_t h, const double *u, double *g, int memspace); /* complianceGradient (TPS.hh:730-751) */

/* ---- MultigridSolver ------------------------------------------------------------------------- */
int vfem_mg_create(vfem_tps_t tps, size_t num_coarsening_levels, vfem_mg_t *out);   /* tps.multigridSolver(L) (VoxelFEM.cc:61) */
void vfem_mg_destroy(vfem_mg_t h);
size_t vfem_mg_num_levels(vfem_mg_t h);
size_t vfem_mg_level_nodes(vfem_mg_t h, size_t l);                                  /* getSimulator(l).numNodes() */
int vfem_mg_level_grid(vfem_mg_t h, size_t l, size_t *elems);                       /* getSimulator(l) grid */
int vfem_mg_level_dirichlet_mask(vfem_mg_t h, size_t l, uint8_t *mask_nnxN);        /* getSimulator(l).dirichletMask */
int vfem_mg_set_symmetric_gs(vfem_mg_t h, int symmetric);                           /* setSymmetricGaussSeidel */
int vfem_mg_set_flags(vfem_mg_t h, int build_block, int build_finest_block);        /* buildBlockStiffnessMatrices / buildFinestBlockStiffnessMatrix */
int vfem_mg_update_stiffness(vfem_mg_t h);                                          /* updateElementStiffnessMatrices + updateBlockKs */
int vfem_mg_apply_k(vfem_mg_t h, size_t l, const double *u, double *out, int memspace);              /* applyK(l,u) */
int vfem_mg_compute_residual(vfem_mg_t h, size_t l, const double *u, const double *b, double *out, int memspace); /* computeResidual */
int vfem_mg_zero_dirichlet(vfem_mg_t h, size_t l, double *u, int memspace);                          /* zeroOutDirichletComponents */
int vfem_mg_smooth_colored(vfem_mg_t h, size_t l, double *u, const double *b, int forward, int memspace); /* smoothingMulticoloredGS (MG.hh:336-340) */
int vfem_mg_smooth_serial(vfem_mg_t h, size_t l, double *u, const double *b, int forward, int memspace);  /* smoothing(l,u,b): serial lexicographic GS (MG.hh:269-282, VoxelFEM.cc:99-104; debug helper, one warp) */
int vfem_mg_restrict(vfem_mg_t h, size_t l_fine, const double *v_fine, double *out_coarse, int memspace);   /* restriction (MG.hh:145-180) */
int vfem_mg_interpolate(vfem_mg_t h, size_t l_fine, const double *x_coarse, double *x_fine, int accumulate, int memspace); /* MG.hh:115-141 */
int vfem_mg_level_stencil(vfem_mg_t h, size_t l, double *out);                      /* debug: assembled operator of level l >= 1, (3^N, N, N, nn) */
int vfem_mg_debug_get(vfem_mg_t h, size_t l, double *x, double *b);                 /* debug_get_x / debug_get_b */
/* profiling hook (no reference counterpart): average device time of one operator of level l over `reps` back-to-back calls.
 * op: 0 one forward GS sweep, 1 residual, 2 restriction l->l+1, 3 prolongation l+1->l, 4 coarsest solve, 5 operator refresh,
 * 6 one FMG preconditioner pass (2 sweeps), 7 masked K u */
int vfem_mg_time_op(vfem_mg_t h, int op, size_t l, int reps, double *ms_per_call);
typedef void (*vfem_solve_cb)(void *ctx, size_t it, const double *x_host);
int vfem_mg_solve(vfem_mg_t h, const double *u, const double *f, size_t num_steps, size_t num_smoothing_steps, int stiffness_updated,
                  int zero_dirichlet, vfem_solve_cb cb, void *cb_ctx, int full_multigrid, double *out, int memspace); /* solve (MG.hh:446-472) */
/* it_callback(i, x, r) of the reference; x/r are host copies made only when want_vectors != 0 */
typedef void (*vfem_pcg_cb)(void *ctx, size_t it, const double *x_host, const double *r_host, double r_norm);
int vfem_mg_pcg(vfem_mg_t h, const double *u, const double *b, size_t max_iter, double tol, vfem_pcg_cb cb, void *cb_ctx, int want_vectors,
                size_t mg_iterations, size_t mg_smoothing_iterations, int full_multigrid, double *out, int memspace,
                size_t *iterations, double *residual_norm);                         /* preconditionedConjugateGradient (MG.hh:679-732) */

/* ---- slab decomposition along storage axis 0 (no reference counterpart) -----------------------------
 * One rank per GPU.  A rank's simulator holds the element layers [layer_lo, layer_hi) of the global grid; node positions,
 * boundary-condition regions and K0 are those of the global grid (TPS.hh:274-287 evaluated with global indices).
 * Layout of a rank for L coarsening levels: its owned coarsest-level layers [c0, c1) plus ONE coarsest layer below
 * (ranks > 0), i.e. layer_lo = (c0 - [rank > 0]) * 2^L, layer_hi = c1 * 2^L; see ndr_b200/slab.py: slab_layers().
 * Vectors passed to the vfem_mg_* calls of a slab solver are the rank-local (numNodes_local, N) arrays. */
int vfem_tps_create_slab(int dim, const size_t *degrees, const double *bbox_min, const double *bbox_max, const size_t *elems_global,
                         size_t layer_lo, size_t layer_hi, vfem_tps_t *out);
int vfem_tps_count_region_nodes(vfem_tps_t h, const double *lo, const double *hi, size_t plane_lo, size_t plane_hi, size_t *count); /* nodes of the local planes [plane_lo, plane_hi] inside a box */
int vfem_tps_add_force_region_split(vfem_tps_t h, const double *lo, const double *hi, const double *value, size_t divisor); /* TPS.hh:381-389 with the GLOBAL node count of the region as divisor */
int vfem_comm_nccl_unique_id(uint8_t *id128);                                       /* rank 0; broadcast the 128 bytes to the other ranks */
int vfem_comm_create_nccl(const uint8_t *id128, int rank, int world, vfem_comm_t *out); /* collective: ncclCommInitRank on the current device */
int vfem_comm_create_threads(int world, vfem_comm_t *out_world);                    /* `world` ranks as host threads on ONE device (single-GPU tests of the slab algorithm) */
void vfem_comm_destroy(vfem_comm_t h);
int vfem_mg_create_slab(vfem_tps_t tps_local, size_t num_coarsening_levels, vfem_comm_t comm, vfem_mg_t *out); /* collective */
int vfem_mg_owned_planes(vfem_mg_t h, size_t l, size_t *first, size_t *last);       /* owned node planes (inclusive) of level l, local indices */
int vfem_mg_exchange(vfem_mg_t h, size_t l, double *v_device, int memspace);        /* refresh the ghost planes of a level-l vector (collective) */
int vfem_mg_apply_k_halo(vfem_mg_t h, double *u_device, double *out_device, int memspace); /* ghost planes of u refreshed + out = K u on the local grid (owned planes valid); the shift overlaps the interior layers (collective) */
int vfem_mg_apply_k_halo_ring(vfem_mg_t h, double *const *us, double *const *outs, size_t nsets, size_t start, size_t count); /* `count` vfem_mg_apply_k_halo calls over a ring of device arrays, issued by one host call (collective) */
int vfem_mg_dot(vfem_mg_t h, const double *a, const double *b, int memspace, double *result); /* dot product over the owned dofs of all ranks (collective) */

/* ---- MultigridComplianceObjective --------------------------------------------------------------- */
int vfem_obj_create(vfem_mg_t mg, vfem_obj_t *out);                                 /* ctor solves once (Objective.hh:78-82) */
int vfem_obj_create_direct(vfem_tps_t tps, vfem_obj_t *out);                        /* pyVoxelFEM.ComplianceObjective(simulator) (VoxelFEM.cc:160-163, Objective.hh:25-68): direct solve per update */
void vfem_obj_destroy(vfem_obj_t h);
int vfem_obj_set_params(vfem_obj_t h, size_t cg_iter, double tol, size_t mg_iterations, size_t mg_smoothing_iterations, int full_multigrid, int zero_init);
int vfem_obj_get_params(vfem_obj_t h, size_t *cg_iter, double *tol, size_t *mg_iterations, size_t *mg_smoothing_iterations, int *full_multigrid, int *zero_init);
typedef void (*vfem_residual_cb)(void *ctx, size_t it, double r_norm);
int vfem_obj_set_residual_cb(vfem_obj_t h, vfem_residual_cb cb, void *ctx);         /* residual_cb */
int vfem_obj_compliance(vfem_obj_t h, double *J);                                   /* compliance() = 1/2 f.u */
int vfem_obj_gradient(vfem_obj_t h, double *g, int memspace);                       /* gradient() */
int vfem_obj_get_u(vfem_obj_t h, double *u, int memspace);                          /* u() */
int vfem_obj_get_f(vfem_obj_t h, double *f, int memspace);                          /* f() */
int vfem_obj_update_cache(vfem_obj_t h, const double *x_phys, int memspace);        /* updateCache(xPhys) */
int vfem_obj_last_solve(vfem_obj_t h, size_t *iterations, double *residual_norm);

/* ---- TopologyOptimizationProblem + OCOptimizer ----------------------------------------------------- */
#define VFEM_FILTER_SMOOTHING 0   /* SmoothingFilter{radius}   */
#define VFEM_FILTER_PROJECTION 1  /* ProjectionFilter{beta}    */
#define VFEM_FILTER_CALLBACK 2    /* PythonFilter{apply_cb, backprop_cb} */
typedef void (*vfem_filter_cb)(void *ctx, int backprop, const double *in, const double *vars, double *out, size_t n);
int vfem_top_create(vfem_tps_t tps, vfem_obj_t objective, int num_constraints, const double *volume_fractions, int num_filters,
                    const int *filter_types, const double *filter_params, const vfem_filter_cb *filter_cbs, void *const *filter_ctxs,
                    vfem_top_t *out);                                               /* VoxelFEM.cc:156-159 */
void vfem_top_destroy(vfem_top_t h);
size_t vfem_top_num_vars(vfem_top_t h);
int vfem_top_set_filter_param(vfem_top_t h, int filter, double param);
int vfem_top_set_vars(vfem_top_t h, const double *x, int memspace, int force_update, int *changed); /* setVars (Problem.hh:42-65) */
int vfem_top_get_vars(vfem_top_t h, int which, double *out, int memspace);          /* getVars (which=0) / getDensities (which=-1) */
int vfem_top_objective(vfem_top_t h, double *J);                                    /* evaluateObjective */
int vfem_top_objective_gradient(vfem_top_t h, double *g, int memspace);             /* evaluateObjectiveGradient */
int vfem_top_constraints(vfem_top_t h, double *c);                                  /* evaluateConstraints */
int vfem_top_constraints_jacobian(vfem_top_t h, double *jac, int memspace);         /* evaluateConstraintsJacobian, (numConstraints, numVars) */
int vfem_oc_reset(vfem_top_t h);                                                    /* OCOptimizer(problem): bracket [1,2] */
int vfem_oc_step(vfem_top_t h, double m, double ctol, double *lambda, int *num_evals); /* OCOptimizer.step (OptimalityCriterion.hh:41-75) */

/* pyVoxelFEM.applyFilter(filter, x) (VoxelFEM.cc:275-281); x has prod(dims) entries, host memory */
int vfem_apply_filter(int filter_type, double param, int ndim, const size_t *dims, const double *x, double *out);

/* ---- allocation-free entry point ---------------------------------------------------------------------------------- */
/* NO counterpart in the reference: added for callers that must not trigger
   cudaMalloc (CUDA-graph capture, serving loops).  result_device[0] = a . b for two device arrays of n doubles, on the
   legacy default stream, asynchronously; the reduction scratch is a static __device__ array of 64.25 KB per GPU
   (ndr_b200/csrc/no_alloc.cu), so the call allocates nothing and n is limited to vfem_dot_no_alloc_max_n() = 33 554 432
   entries (larger n: error).  All calls share that array: calls that overlap on the GPU race. */
int vfem_dot_no_alloc(const double *a_device, const double *b_device, size_t n, double *result_device);
size_t vfem_dot_no_alloc_max_n(void);

#ifdef __cplusplus
}
#endif
#endif
