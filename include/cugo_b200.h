/*
 * cugo_b200.h -- C ABI of the B200-native bundle-adjustment engine.
 *
 * This is the drop-in boundary for the reference's hot path: the seam between
 * cugo::BlockSolver (src/block_solver.h:46-173) and the free functions of
 * namespace cugo::gpu (src/cuda/cuda_block_solver.h:55-256).  Every entry point
 * names the reference interface it replaces.  Plain pointers and sizes only; all
 * pointers are HOST pointers unless stated; every function returns 0 on success
 * or a negative CUGO_B200_E* code (cugo_b200_last_error() gives the text).  The
 * library never throws across this boundary and has no CPU fallback: without a
 * CUDA device cugo_b200_create() fails with CUGO_B200_ENODEVICE.
 *
 * Data contract (same as the reference):
 *   pose      7 doubles  qx qy qz qw tx ty tz   (world -> camera, Se3D src/fixed_vector.h:393-471)
 *   landmark  3 doubles
 *   vertices are passed in INDEX order: non-fixed first (0..active-1), then fixed
 *             (VertexSet::generateEstimateData, src/optimisable_graph.hpp:84-126)
 *   edge      (iP, iL) vertex indices, measurement (2 mono / 3 stereo doubles),
 *             information omega (per edge or one per set), camera fx fy cx cy bf
 *             (per edge or one per set)  -- EdgeSet::init, src/optimisable_graph.hpp:474-572
 *   blocks    column-major, full: Hpp 6x6, Hll 3x3, Hpl 6x3; x = [xp (6P) | xl (3L)]
 */
#ifndef CUGO_B200_H
#define CUGO_B200_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct cugo_b200_solver cugo_b200_solver;

enum {
    CUGO_B200_OK = 0,
    CUGO_B200_EINVAL = -1,      /* bad argument */
    CUGO_B200_ENODEVICE = -2,   /* no usable CUDA device (there is no CPU path) */
    CUGO_B200_ECUDA = -3,       /* CUDA runtime error */
    CUGO_B200_ESTATE = -4,      /* call order violated (e.g. solve before build_system) */
    CUGO_B200_EUNSUPPORTED = -5 /* outside what the engine covers (e.g. ICP edge sets on several GPUs) */
};

/* robust kernels: reference RobustKernelType {None, Cauchy, Tukey} (src/robust_kernel.h:12-17);
 * Huber (g2o RobustKernelHuber) is added because BASELINE.json's configs name it. */
enum { CUGO_B200_RK_NONE = 0, CUGO_B200_RK_CAUCHY = 1, CUGO_B200_RK_TUKEY = 2, CUGO_B200_RK_HUBER = 3 };

/* structure arrays readable through cugo_b200_get_structure (int32) */
enum {
    CUGO_B200_HPL_COLPTR = 0, /* [L+1]   d_Hpl_.outerIndices   gpu::buildHplStructure cuda_block_solver.cu:1571-1604 */
    CUGO_B200_HPL_ROWIND = 1, /* [nHpl]  d_Hpl_.innerIndices */
    CUGO_B200_HSC_ROWPTR = 2, /* [P+1]   Hsc_.outerIndices     sparse_block_matrix.cpp:63-156 */
    CUGO_B200_HSC_COLIND = 3, /* [nblk]  Hsc_.innerIndices */
    CUGO_B200_CSR_ROWPTR = 4, /* [6P+1]  Hsc_.rowPtr()         sparse_block_matrix.cpp:158-262 */
    CUGO_B200_CSR_COLIND = 5, /* [nnzSymm] Hsc_.colInd() */
    CUGO_B200_BSR2CSR = 6,    /* [nnzSymm] Hsc_.BSR2CSR() */
    CUGO_B200_MULIDS = 7,     /* [3 nmul] d_HscMulBlockIds_, sorted by (i,j)  cuda_block_solver.cu:1606-1634 */
    CUGO_B200_EDGE2HPL = 8,   /* [E] d_edge2Hpl_ in input edge order (-1: no block) */
    CUGO_B200_PERM = 9,       /* [P] fill-reducing block ordering, new -> old (ours; the reference's is METIS) */
    /* the rest of the symbolic analysis (no reference counterpart that is visible: it lives inside cuSOLVER's
     * csrcholInfo_t, src/cholesky.hpp:94-101); compared bit for bit with the host run of csrc/symbolic.h */
    CUGO_B200_SN_START = 10,  /* [nsn+1] first permuted column of every supernode */
    CUGO_B200_SN_ROWPTR = 11, /* [nsn+1] */
    CUGO_B200_SN_ROWS = 12,   /* [nrows] row structure (permuted block rows below the supernode), ascending */
    CUGO_B200_SN_PARENT = 13, /* [nsn] assembly tree */
    CUGO_B200_SN_LEVEL = 14,  /* [nsn] */
    CUGO_B200_LEVEL_SN = 15,  /* [nsn] supernodes grouped by level (the launch schedule) */
    CUGO_B200_CHILD_IDX = 16, /* [non-roots] children grouped by parent, ascending */
    CUGO_B200_INV_MAP = 17,   /* [ninv] inverse relative indices child -> parent front */
    CUGO_B200_ASM_BLK = 18, CUGO_B200_ASM_ROW = 19, CUGO_B200_ASM_COL = 20   /* [nblk] Hsc block -> front scatter */
};

/* value arrays readable through cugo_b200_get_values (double) */
enum {
    CUGO_B200_HPP = 0, CUGO_B200_HLL = 1, CUGO_B200_HPL = 2, CUGO_B200_B = 3, CUGO_B200_X = 4,
    CUGO_B200_BSC = 5, CUGO_B200_HSC = 6, CUGO_B200_INVHLL = 7, CUGO_B200_CHI_PER_EDGE = 8
};

/* -- lifetime: replaces CudaDevice (src/cuda_device.cpp:15-42) + BlockSolver ctor. `device` is the
 * CUDA ordinal chosen by the caller (one solver = one device + one stream).
 * use_float32 != 0 selects the counterpart of the reference's USE_FLOAT32 build (include/scalar.h; that
 * build itself does not compile): the Hpl blocks -- 18 values per edge, the bulk of the bytes
 * every bandwidth-bound stage moves -- are STORED in single precision; estimates, measurements, residuals,
 * Jacobians, all accumulation, the Schur complement and the factorisation stay FP64 (as the reference keeps
 * vertices and measurements double under USE_FLOAT32).  chi2 per iteration stays within 1e-4 relative of the
 * FP64 run.  cugo_b200_get_values(CUGO_B200_HPL) returns the stored blocks widened to double. */
int cugo_b200_create(int device, int use_float32, cugo_b200_solver** out);
int cugo_b200_destroy(cugo_b200_solver* s);
const char* cugo_b200_last_error(const cugo_b200_solver* s);

/* -- BlockSolver::initialize (src/block_solver.cpp:21-137): vertices ... */
int cugo_b200_set_vertices(cugo_b200_solver* s, int nP_all, int nP_active, const double* poses7,
                           int nL_all, int nL_active, const double* landmarks3);
/* ... and edge sets (EdgeSet::init / mapDevice, src/optimisable_graph.hpp:474-601).  dim: 2 =
 * MonoEdgeSet, 3 = StereoEdgeSet (include/ba_types.h:34-169), 4 = DepthEdgeSet (below: 3 measurement columns).  nomega / ncam are 1 or n.  Only
 * active edges (not both endpoints fixed) may be passed.  Returns the set id in *set_id. */
int cugo_b200_clear_edge_sets(cugo_b200_solver* s);
int cugo_b200_add_edge_set(cugo_b200_solver* s, int dim, int n, const int* iP, const int* iL,
                           const double* measurements, const double* omegas, int nomega,
                           const double* cameras5, int ncam, int robust_type, double robust_delta,
                           int* set_id);
/* -- the reference's other edge types.
 * DepthEdgeSet (include/ba_types.h:171-234): measurement (u, v, 1/z), three rows.  Same arguments as
 * cugo_b200_add_edge_set (equivalent to dim = 4 there).  The reference pairs the residual meas - proj with the
 * stereo Jacobian (src/cuda/cuda_block_solver.cu:2341-2343 with ba_types.h:217), which turns every step into an
 * ascent direction; here the residual is proj - meas like the other edges and the third Jacobian row is that of 1/Z. */
int cugo_b200_add_depth_edge_set(cugo_b200_solver* s, int n, const int* iP, const int* iL,
                                 const double* measurements, const double* omegas, int nomega,
                                 const double* cameras5, int ncam, int robust_type, double robust_delta,
                                 int* set_id);
/* PlaneEdgeSet / LineEdgeSet (include/icp_types.h:19-107): pose-only edges with one scalar residual, added to Hpp / bp of
 * their pose.  matches7 / matches10 are the reference's match structs as they lie in memory (src/measurements.h:13-101):
 *   plane: normal[3], originDistance, pointP[3]          e = normal . (T pointP) - originDistance
 *   line : a[3], b[3], length (= |a - b| > 0), pointP[3] e = |(T pointP - a) x (T pointP - b)| / length
 * with T pointP = R(q) pointP + t.  Edges on fixed poses are rejected (inactive).  The ids returned for these sets start
 * at CUGO_B200_ICP_SET_BASE; they are accepted by cugo_b200_set_robust_kernel only (the reference's ICP kernels know
 * no outlier mask).  One GPU.  A graph whose landmarks are all fixed (or that has none) is solved without the Schur
 * complement: Hpp is block diagonal there (reference: DenseLinearSolver over the whole 6P x 6P matrix,
 * src/block_solver.cpp:226-234,344-354, src/cholesky.hpp:314-516). */
#define CUGO_B200_ICP_SET_BASE 1024
int cugo_b200_add_plane_edge_set(cugo_b200_solver* s, int n, const int* iP, const double* matches7,
                                 const double* omegas, int nomega, int robust_type, double robust_delta, int* set_id);
int cugo_b200_add_line_edge_set(cugo_b200_solver* s, int n, const int* iP, const double* matches10,
                                const double* omegas, int nomega, int robust_type, double robust_delta, int* set_id);
/* EdgeSet::setRobustKernel (src/optimisable_graph.hpp:401-406); per set here, process-global there */
int cugo_b200_set_robust_kernel(cugo_b200_solver* s, int set_id, int robust_type, double robust_delta);
/* caller-supplied outlier mask of a set (reference d_outliers / BaseEdge::inactivate, src/optimisable_graph.h:341-408);
 * changing it invalidates the structure: masked edges own no Hpl block after the rebuild, as in the reference's
 * buildHplBlockPos (src/optimisable_graph.hpp:642-661) */
int cugo_b200_set_outliers(cugo_b200_solver* s, int set_id, const int* mask);
/* -- outlier gating on the device.  EdgeSet::setOutlierThreshold (src/optimisable_graph.hpp:415-420): with a
 * threshold > 0, the end of every cugo_b200_optimize() flags the set's edges whose robustified chi2 at the last
 * evaluated estimates exceeds it -- BlockSolver::updateEdges (src/block_solver.cpp:423-429) ->
 * EdgeSet::updateEdges (src/optimisable_graph.hpp:603-640) -> computeOutliersKernel
 * (cuda_block_solver.cu:1135-1149), called from src/cuda_graph_optimisation.cpp:150-151.  Flagged edges are
 * skipped by every later stage.  Unlike the reference (mask to the host, edges inactivated, structure rebuilt)
 * the structure, its symbolic analysis and the trial graph are KEPT: a flagged edge's Hpl block stays in the
 * pattern as a zero block.  cugo_b200_update_edges is the staged call (uses the estimates of the last
 * cugo_b200_compute_errors). */
int cugo_b200_set_outlier_threshold(cugo_b200_solver* s, int set_id, double threshold);
int cugo_b200_update_edges(cugo_b200_solver* s);
/* EdgeSet::getOutlierCount / getInlierCount (src/optimisable_graph.hpp:664-675) and the mask the reference
 * downloads in updateEdges (edgeOutliers_): mask[n] in the set's edge order (may be NULL) */
int cugo_b200_get_outliers(cugo_b200_solver* s, int set_id, int* mask, int* n_outliers, int* n_inliers);
/* re-upload estimates only (a second initialize() on an unchanged graph keeps the structure,
 * src/block_solver.cpp:161-170) */
int cugo_b200_set_estimates(cugo_b200_solver* s, const double* poses7, const double* landmarks3);

/* -- multi-GPU, one process per GPU (no counterpart in the reference, which is single-GPU): each rank holds
 * ALL poses and a shard of the landmarks with all their edges (landmark indices local to the shard).  Per LM
 * trial the ranks' partial reduced camera systems [Hsc | bsc] are summed with one ncclAllReduce on the
 * solver's stream, every rank factorises redundantly and back-substitutes its own landmarks.  Rank 0 calls
 * cugo_b200_comm_unique_id and the host distributes the 128 bytes (e.g. torch.distributed.broadcast); every
 * rank then calls cugo_b200_comm_init before cugo_b200_build_structure.  With nranks > 1 only the
 * device-resident cugo_b200_optimize path is available (the staged calls return CUGO_B200_ESTATE). */
int cugo_b200_comm_unique_id(char id[128]);
int cugo_b200_comm_init(cugo_b200_solver* s, int nranks, int rank, const char id[128]);

/* -- BlockSolver::buildStructure (src/block_solver.cpp:139-248) incl. symbolic analysis */
int cugo_b200_build_structure(cugo_b200_solver* s);

/* -- stage API, one call per BlockSolver method (src/block_solver.h:54-173) */
int cugo_b200_compute_errors(cugo_b200_solver* s, double* chi2);              /* computeErrors   :250-272 */
int cugo_b200_build_system(cugo_b200_solver* s);                              /* buildSystem     :274-307 */
int cugo_b200_max_diagonal(cugo_b200_solver* s, double* max_diag);            /* maxDiagonal     :309-320 */
int cugo_b200_set_lambda(cugo_b200_solver* s, double lambda);                 /* setLambda       :322-329 */
int cugo_b200_restore_diagonal(cugo_b200_solver* s);                          /* restoreDiagonal :331-338 */
int cugo_b200_solve(cugo_b200_solver* s, int* ok);                            /* solve           :340-388 */
int cugo_b200_update(cugo_b200_solver* s);                                    /* update          :390-413 */
int cugo_b200_compute_scale(cugo_b200_solver* s, double lambda, double* scale); /* computeScale  :415-421 */
int cugo_b200_push(cugo_b200_solver* s);                                      /* push            :431-434 */
int cugo_b200_pop(cugo_b200_solver* s);                                       /* pop             :436-439 */

/* -- CudaGraphOptimisationImpl::optimize (src/cuda_graph_optimisation.cpp:48-154) with the LM
 * control on the device.  chi2 / lambda / trials receive one entry per outer iteration
 * (BatchStatistics semantics: chi2 after that iteration); capacity entries each (may be NULL). */
int cugo_b200_optimize(cugo_b200_solver* s, int niterations, int capacity, double* chi2,
                       double* lambda, int* trials, int* nstats);

/* -- BlockSolver::finalize / VertexSet::finalise (src/optimisable_graph.hpp:137-154) */
int cugo_b200_get_estimates(cugo_b200_solver* s, double* poses7, double* landmarks3);

/* -- introspection for parity tests */
int cugo_b200_get_structure(cugo_b200_solver* s, int which, int* out, long capacity, long* n);
int cugo_b200_get_values(cugo_b200_solver* s, int which, double* out, long capacity, long* n);
/* sizes: P_active L_active E nHpl nblk nmul nnzSymm nsupernodes nlevels kernels_per_trial */
int cugo_b200_get_sizes(cugo_b200_solver* s, long sizes[10]);
/* device time of the last optimize() in ms (CUDA events on the solver's stream) and the number of
 * kernels it launched */
int cugo_b200_get_last_timing(cugo_b200_solver* s, double* optimize_ms, long* kernels_launched);
/* per-kernel device times (ms, CUDA events on the solver's stream) of ONE LM trial run outside the
 * CUDA graph: names is a capacity x name_len char matrix.  Leaves the estimates advanced by that trial. */
int cugo_b200_profile_trial(cugo_b200_solver* s, int capacity, char* names, int name_len, double* ms, int* n);
/* new measurements / information for an unchanged graph: the reference's second initialize() re-uploads
 * the edge arena but keeps the structure (src/block_solver.cpp:161-170, EdgeSet::mapDevice
 * src/optimisable_graph.hpp:574-601) */
int cugo_b200_update_measurements(cugo_b200_solver* s, int set_id, const double* measurements,
                                  const double* omegas, int nomega);
/* algorithmic flops of one numeric factorisation under the ordering in use: panel
 * factorisation (Cholesky of the diagonal blocks + rows below, forward substitution included), symmetric
 * updates, backward substitution, and the number of doubles stored in L */
int cugo_b200_get_factor_stats(cugo_b200_solver* s, double stats[4]);
/* diagnostics: phase timestamps (SM clock) the factorisation kernels record per tree level -- only in a diagnostic build
 * of the library (make DIAG=1, run with CUGO_B200_DEBUG=1); the shipped build returns CUGO_B200_EUNSUPPORTED */
int cugo_b200_debug_counters(long long* out, int n);
/* scheduling variants of the LM trial, for A/B measurements and the parity test that checks every variant: name one of
 * "pdl" (programmatic dependent launch, else plain stream order), "fuse_build" (both passes of the system build in
 * one launch), "panel_dc" (decoupled diagonal chain in the panel kernel), "zinv" / "bwd_split" (backward solves with the
 * inverse of L11 / shared by several CTAs), "parallel_symbolic" (0: the sequential statement of the analysis,
 * csrc/symbolic.h, run by one device thread).  All default to 1; changing one rebuilds the structure and the graph. */
int cugo_b200_set_option(cugo_b200_solver* s, const char* name, int value);
/* the solver's CUDA stream (cudaStream_t) for callers that time or chain work on it */
void* cugo_b200_stream(cugo_b200_solver* s);

/* -- allocation-free chi2 of ONE edge set.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same per-edge arithmetic as cugo_b200_compute_errors (sum over the set of
 * rho(omega |e|^2)), but every buffer is the caller's and ALL pointers except camera5 / errbuf are DEVICE
 * pointers in the caller's own layout (poses 7 doubles each, measurements [n x dim] interleaved, vertex
 * indices as for cugo_b200_add_edge_set).  The per-CTA partial sums live in a static __device__ array inside the
 * library (64.25 KB, held by every GPU that loads the library until exit) instead of a cudaMalloc'ed buffer, which
 * bounds the set: n <= CUGO_B200_NOALLOC_MAX_EDGES, larger sets are rejected with CUGO_B200_EINVAL (split them).
 * The call makes no cudaMalloc / cudaFree / allocating library call; it enqueues one 256-byte memset and one kernel
 * on `stream` (a cudaStream_t, NULL = default stream) and returns without synchronising: *d_chi2 (one double, device)
 * is valid once the stream has reached that point.  All calls on a device share the one scratch array: calls that
 * may OVERLAP on the GPU (different streams, no ordering between them) race -- serialise them.
 * errbuf (host, may be NULL) receives the error text. */
#define CUGO_B200_NOALLOC_MAX_EDGES (8 * 1024 * 1024)
int cugo_b200_chi2_noalloc(int device, void* stream, int dim, int n, const int* d_iP, const int* d_iL,
                           const double* d_measurements, const double* d_omegas, int nomega,
                           const double camera5[5], int robust_type, double robust_delta,
                           const double* d_poses7, const double* d_landmarks3, double* d_chi2,
                           char* errbuf, int errbuf_len);

#ifdef __cplusplus
}
#endif
#endif /* CUGO_B200_H */
