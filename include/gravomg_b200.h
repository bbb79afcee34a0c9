/* gravomg_b200 -- C ABI of the B200-native Gravo MG hierarchy-construction path.
 *
 * This is the drop-in boundary: plain pointers and sizes, no C++/torch types. Each entry point
 * replaces one reference function (paths relative to the reference repository):
 *
 *   gmg_to_edge_distance_matrix        GravoMG::toEdgeDistanceMatrix        include/gravomg/utility.h:22,  src/utility.cpp:50-56
 *   gmg_extract_edges                  GravoMG::extractEdges                include/gravomg/utility.h:24,  src/utility.cpp:58-71
 *   gmg_average_edge_length            GravoMG::averageEdgeLength           include/gravomg/multigrid.h:36, src/multigrid.cpp:127-133
 *   gmg_fast_disc_sample               GravoMG::fastDiscSample              include/gravomg/sampling.h:14-18, src/sampling.cpp:7-53
 *   gmg_assign_parents                 GravoMG::assignParents               include/gravomg/multigrid.h:30-34, src/multigrid.cpp:77-125
 *   gmg_extract_coarse_edges           GravoMG::extractCoarseEdges          include/gravomg/multigrid.h:38-43, src/multigrid.cpp:135-169
 *   gmg_coarse_from_mean_of_fine_children  GravoMG::coarseFromMeanOfFineChildren  multigrid.h:45-50, src/multigrid.cpp:171-207
 *   gmg_construct_voronoi_triangles    GravoMG::constructVoronoiTriangles   include/gravomg/multigrid.h:52-55, src/multigrid.cpp:209-263
 *   gmg_construct_prolongation         GravoMG::constructProlongation       include/gravomg/multigrid.h:57-63, src/multigrid.cpp:265-498
 *   gmg_projected_points               GravoMG::projectedPoints             include/gravomg/multigrid.h:65,  src/multigrid.cpp:500-510
 *   gmg_hierarchy_*                    [not in reference] the per-level call order of test/main.cpp:68-121,
 *                                      looped over levels with everything resident in HBM, plus the Galerkin
 *                                      operator, SpMV and multicolour Gauss-Seidel BASELINE.json asks for.
 *
 * Data conventions (those of the reference's Eigen types, include/gravomg/utility.h:10-18):
 *   points     column-major n x 3 doubles with leading dimension ld >= n  (Eigen::MatrixXd::data(), ld = rows())
 *   edge matrix  compressed column-major: int32 colptr[n+1], int32 rowidx[nnz] ascending per column,
 *              f64 val[nnz]   (Eigen::SparseMatrix<double>: outerIndexPtr / innerIndexPtr / valuePtr)
 *   P          compressed row-major: int32 rowptr[n+1], int32 colidx[nnz] ascending per row, f64 val[nnz]
 *   std::vector<Eigen::Index>  -> int64 pointer + length
 * All pointers in the per-function entry points are HOST pointers; the library copies to and from
 * the device inside the call. gmg_hierarchy_* keeps everything on the device between stages.
 *
 * Variable-size outputs use a two-call convention: the compute call returns the sizes and keeps
 * the result in the context; gmg_fetch_* copies it into caller-allocated arrays.
 *
 * Every function returns a gmg_status; gmg_last_error() gives the text. There is no CPU fallback:
 * without a usable CUDA device gmg_create fails with GMG_ERR_CUDA.
 */
#ifndef GRAVOMG_B200_H
#define GRAVOMG_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct gmg_ctx gmg_ctx;
typedef struct gmg_hierarchy gmg_hierarchy;

typedef enum {
    GMG_OK = 0,
    GMG_ERR_CUDA = 1,            /* a CUDA runtime call or kernel failed */
    GMG_ERR_INVALID = 2,         /* malformed input the reference would treat as undefined behaviour */
    GMG_ERR_FALLBACK_RATIO = 3,  /* mirrors assert(fallbackCount / n < 0.5), src/multigrid.cpp:488 */
    GMG_ERR_ZERO_LENGTH_EDGE = 4,/* an edge that does not increase the path length: Dijkstra tie order is dynamic */
    GMG_ERR_ASYMMETRIC = 5,      /* the sparsity pattern is not symmetric; the parallel formulation needs it */
    GMG_ERR_STATE = 6,           /* fetch without a matching compute call, or bad handle */
    GMG_ERR_NCCL = 7
} gmg_status;

/* Weighting of include/gravomg/multigrid.h:12-16 */
enum { GMG_BARYCENTRIC = 0, GMG_UNIFORM = 1, GMG_INVDIST = 2 };
/* NESTED mode [the reference's todo at src/multigrid.cpp:291-292], OR-ed into a weighting: the coarse points of a level
 * are the positions of its samples (gmg_coarse_from_samples; gmg_hierarchy_build / the sharded build do this themselves)
 * instead of the means of their children, and a fine point that coincides with its parent's coarse point -- every
 * sample, then -- gets the row (parent, 1.0), so that P restricted to the samples is the identity. */
enum { GMG_NESTED = 0x100 };
/* Sampler modes: the same greedy rule (src/sampling.cpp:22-50) under three precedence orders.
 *   EXACT     index order: the reference's sample set bit-for-bit.
 *   PARALLEL  the index-order sweep restarted in 2 contiguous super-tiles that are swept concurrently (their
 *             flow tiles are interleaved in the order; the seed is not used): a different set, within 2 % of the
 *             reference on sample count and coverage radius (the north star's statistical criterion) also on
 *             index-coherent grids, where the reference's row-by-row sweep packs much more tightly than any
 *             order-independent rule; half the index order's dependency depth. Below 2 M
 *             vertices (where a second sweep start would cost more than 2 % of the samples) it is the index order.
 *   HASHED    every vertex hashed: dependency depth O(log n); what the row-sharded build runs for both
 *             non-exact modes. Statistically equal to the reference on unordered inputs (point clouds,
 *             subdivision orders), 28 % fewer samples on a row-major grid. */
enum { GMG_SAMPLER_EXACT = 0, GMG_SAMPLER_PARALLEL = 1, GMG_SAMPLER_HASHED = 2 };

int gmg_create(int device, void* cuda_stream, gmg_ctx** out);
void gmg_destroy(gmg_ctx* ctx);
/* The context parks released device buffers for reuse by later calls (a build repeats the same sizes).
 * gmg_trim hands them back to the driver's pool; *cached_bytes (optional) = how much was parked. */
int gmg_trim(gmg_ctx* ctx, int64_t* cached_bytes);
const char* gmg_last_error(const gmg_ctx* ctx);
/* milliseconds spent in kernels of the most recent compute call, measured with CUDA events on the
 * context's stream (excludes host<->device copies) */
double gmg_last_kernel_ms(const gmg_ctx* ctx);
/* number of kernel launches issued by this context since creation */
int64_t gmg_launch_count(const gmg_ctx* ctx);

int gmg_to_edge_distance_matrix(gmg_ctx* ctx, int64_t n, const int32_t* colptr, const int32_t* rowidx,
                                const double* points, int64_t ld, double* out_val);

/* out_edges: column-major nnz x 2 int32 (column 0 = col, column 1 = row), out_values: nnz */
int gmg_extract_edges(gmg_ctx* ctx, int64_t n, const int32_t* colptr, const int32_t* rowidx, const double* val,
                      int32_t* out_edges, double* out_values);

/* sum of |p_j - p_i| over all rows of the edge list, divided by (n_edges - n). The summation order
 * is a fixed pairwise tree (deterministic), not libstdc++'s transform_reduce order. */
int gmg_average_edge_length(gmg_ctx* ctx, int64_t n, const double* points, int64_t ld, int64_t n_edges,
                            const int32_t* edges, double* out_mean);

/* averageEdgeLength for callers that cannot allocate [no counterpart in the reference]:
 * code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc. Everything is a
 * DEVICE pointer on ctx's device: d_points n x 3 column-major packed (x | y | z, leading dimension n), d_edges
 * n_edges x 2 column-major, d_out_mean one double. The call allocates nothing (no cudaMalloc / cudaFree / pool): its
 * scratch -- one partial sum per 2048 edge rows -- is a fixed 512 KiB __device__ array inside the library, which every
 * GPU that loads the library holds until exit. LIMIT: n_edges <= 134,217,728, more is GMG_ERR_INVALID. The work is
 * queued on ctx's stream and not waited for. All calls on a device share the scratch: calls that overlap on the GPU
 * (other streams, other contexts of the same device) race and must be serialised by the caller; consecutive calls
 * need no clearing. Same summation tree as gmg_average_edge_length: identical bits. (csrc/k_noalloc.cu) */
int gmg_average_edge_length_noalloc(gmg_ctx* ctx, int64_t n, const double* d_points, int64_t n_edges,
                                    const int32_t* d_edges, double* d_out_mean);

/* selection_out has room for n entries, ascending; *m_out receives the count */
int gmg_fast_disc_sample(gmg_ctx* ctx, int64_t n, const double* points, int64_t ld, const int32_t* colptr,
                         const int32_t* rowidx, double radius, int mode, uint64_t seed, int64_t* selection_out,
                         int64_t* m_out);

/* The bounds the reference marks as a todo (src/sampling.cpp:11-12: min_selection = 0, max_selection = max) [semantics
 * not in reference]: a radius that yields fewer than min_selection samples is halved and the sampling repeated until
 * there are enough (or every vertex is selected); more than max_selection samples are cut to the first max_selection
 * of the ascending selection (where the reference's loop would have stopped selecting). *radius_used (optional)
 * receives the radius of the run that was kept. */
int gmg_fast_disc_sample_bounded(gmg_ctx* ctx, int64_t n, const double* points, int64_t ld, const int32_t* colptr,
                                 const int32_t* rowidx, double radius, int64_t min_selection, int64_t max_selection, int mode,
                                 uint64_t seed, int64_t* selection_out, int64_t* m_out, double* radius_used);

int gmg_assign_parents(gmg_ctx* ctx, int64_t n, const double* points, int64_t ld, const int32_t* colptr,
                       const int32_t* rowidx, int64_t m, const int64_t* samples, int64_t* parents_out);

/* out_colptr has m+1 entries; then gmg_fetch_sparse(ctx, rowidx[nnz], val[nnz]) */
int gmg_extract_coarse_edges(gmg_ctx* ctx, int64_t n, const int32_t* colptr, const int32_t* rowidx,
                             const double* val, int64_t m, const int64_t* parents, int32_t* out_colptr,
                             int64_t* nnz_out);
int gmg_fetch_sparse(gmg_ctx* ctx, int32_t* out_inner, double* out_val);

int gmg_coarse_from_mean_of_fine_children(gmg_ctx* ctx, int64_t n, const double* points, int64_t ld,
                                          const int32_t* colptr, const int32_t* rowidx, const int64_t* parents,
                                          int64_t m, double* out_points, int64_t out_ld);

/* out_fan_ptr has m+1 entries; then gmg_fetch_triangles(ctx, tri[3T] row-major int64, normal[3T]
 * row-major f64, fan[3T] int64) */
int gmg_construct_voronoi_triangles(gmg_ctx* ctx, int64_t m, const double* points, int64_t ld,
                                    const int32_t* colptr, const int32_t* rowidx, const double* val,
                                    int64_t* T_out, int64_t* out_fan_ptr);
int gmg_fetch_triangles(gmg_ctx* ctx, int64_t* out_tri, double* out_normal, int64_t* out_fan);

/* counters the reference computes and drops (src/multigrid.cpp:281-283) */
typedef struct {
    int64_t no_triangle_found;   /* notrisfound */
    int64_t edges_found;         /* edgesfound */
    int64_t fallback_count;      /* fallbackCount */
} gmg_prolongation_stats;

/* out_rowptr has n+1 entries; then gmg_fetch_sparse(ctx, colidx[nnz], val[nnz]) */
int gmg_construct_prolongation(gmg_ctx* ctx, int64_t n, const double* fine_points, int64_t fine_ld, int64_t m,
                               const double* coarse_points, int64_t coarse_ld, const int32_t* c_colptr,
                               const int32_t* c_rowidx, const double* c_val, const int64_t* parents, int weighting,
                               int32_t* out_rowptr, int64_t* nnz_out, gmg_prolongation_stats* stats_out);

/* nested mode: out (column-major m x 3, leading dimension out_ld) = the positions of the samples */
int gmg_coarse_from_samples(gmg_ctx* ctx, int64_t n, const double* points, int64_t ld, int64_t m, const int64_t* samples,
                            double* out_points, int64_t out_ld);

int gmg_projected_points(gmg_ctx* ctx, int64_t n, int64_t m, const int32_t* rowptr, const int32_t* colidx,
                         const double* val, const double* coarse_points, int64_t coarse_ld, double* out_points,
                         int64_t out_ld);

/* ------------------------------------------------------------------------------------------------
 * Resident hierarchy builder [not in reference]: the per-level call order of test/main.cpp:68-121
 * (toEdgeDistanceMatrix once, then per level: averageEdgeLength -> fastDiscSample -> assignParents ->
 * extractCoarseEdges -> coarseFromMeanOfFineChildren -> constructVoronoiTriangles ->
 * constructProlongation), feeding (coarse_points, coarse_edge_matrix) back in as the next level's
 * fine inputs. Everything stays in HBM between stages and levels.
 * ------------------------------------------------------------------------------------------------ */
typedef struct gmg_input gmg_input;

/* Copy a level-0 problem to the device. A_val may be NULL when no Galerkin operator is wanted; the
 * pattern of A (colptr/rowidx) IS the level-0 graph, as in test/main.cpp:68-69. */
int gmg_input_create(gmg_ctx* ctx, int64_t n, const double* points, int64_t ld, const int32_t* colptr,
                     const int32_t* rowidx, const double* A_val, gmg_input** out);
void gmg_input_destroy(gmg_input* in);
/* Row-sharded upload [not in reference]: an input whose arrays the caller fills ON THE DEVICE -- each rank
 * copies its own slice from the host and the slices are all-gathered over NVLink, instead of every rank
 * pulling the whole mesh over PCIe (gravo-mg_b200/sharded.py upload_sharded). gmg_input_device_ptr: 0 points
 * (f64[3n]: x | y | z), 1 colptr (int32[n+1]), 2 rowidx (int32[nnz]), 3 A values (f64[nnz]).
 * gmg_input_commit validates the completed arrays (same checks as gmg_input_create). */
/* A values for an input created with A_val = NULL, copied on the context's copy stream while the compute stream goes
 * on (a build needs them only at its first Galerkin product). A_val must stay valid (and should be pinned) until a call
 * that uses the values has returned, or gmg_copy_sync has. */
int gmg_input_set_values_async(gmg_ctx* ctx, gmg_input* in, const double* A_val);
int gmg_input_create_device(gmg_ctx* ctx, int64_t n, int64_t nnz, int with_a, gmg_input** out);
int gmg_input_device_ptr(gmg_input* in, int which, void** ptr);
int gmg_input_commit(gmg_ctx* ctx, gmg_input* in);

/* Level-0 problems generated on the device [not in reference; the step before the hot path]: the noisy
 * torus grid of configs C3/C5 (nu x nv vertices, vertex (i,j) = i*nv + j, quads split on the same
 * diagonal, positions displaced by amplitude * h * splitmix64 noise; pattern = adjacency + diagonal).
 * gmg_input_fetch copies an array back to the host: 0 points (f64[3n], column-major n x 3),
 * 1 colptr (int32[n+1]), 2 rowidx (int32[nnz]), 3 A values (f64[nnz]).
 * gmg_input_torus_operator adds the system matrix of C3/C5 to such an input: cotangent Laplacian +
 * mass_shift * lumped mass on the same pattern. */
int gmg_generate_torus(gmg_ctx* ctx, int32_t nu, int32_t nv, double R, double r, double amplitude, uint64_t seed,
                       gmg_input** out);
int gmg_input_torus_operator(gmg_ctx* ctx, gmg_input* in, int32_t nu, int32_t nv, double mass_shift);
int gmg_input_fetch(gmg_ctx* ctx, const gmg_input* in, int which, void* host_out);
int gmg_input_size(const gmg_input* in, int64_t* n, int64_t* nnz);

/* Level-0 graphs BUILT on the device from raw geometry [not in reference; SURVEY.md 8(f) rank 2: what the reference's
 * driver gets from an external buildPointCloudLaplacian / mesh loader, test/main.cpp:68]. points: column-major n x 3,
 * leading dimension ld, on the host.
 *   gmg_input_from_knn   exact k nearest neighbours (1 <= k <= 32, ties by index) by grid hashing, union-symmetrised,
 *                        plus the diagonal: the graph of config C4. No A values.
 *   gmg_input_from_mesh  faces: nf x 3 int32 row-major. Pattern = mesh adjacency + diagonal, A = cotangent Laplacian
 *                        + mass_shift * lumped mass (the system matrix of C1/C2/C3).
 *   gmg_morton_order     perm_out[r] = index of the point of rank r along the Morton curve of the bounding cube
 *                        (21 bits per axis, ties by index): the reordering that makes contiguous row blocks
 *                        spatially compact before a row-sharded build. */
int gmg_input_from_knn(gmg_ctx* ctx, int64_t n, const double* points, int64_t ld, int32_t k, gmg_input** out);
int gmg_input_from_mesh(gmg_ctx* ctx, int64_t n, const double* points, int64_t ld, int64_t nf, const int32_t* faces,
                        double mass_shift, gmg_input** out);
int gmg_morton_order(gmg_ctx* ctx, int64_t n, const double* points, int64_t ld, int32_t* perm_out);

typedef struct {
    int32_t levels;          /* levels including the finest; levels-1 prolongations are built */
    double ratio;            /* radius = cbrt(ratio) * averageEdgeLength  (test/main.cpp:74) */
    int32_t weighting;       /* GMG_BARYCENTRIC / GMG_UNIFORM / GMG_INVDIST */
    int32_t sampler_mode;    /* GMG_SAMPLER_EXACT / GMG_SAMPLER_PARALLEL / GMG_SAMPLER_HASHED */
    uint64_t seed;           /* priority hash seed of the parallel sampler */
    int32_t radius_policy;   /* 0: averageEdgeLength verbatim, divisor nnz - n on every level (src/multigrid.cpp:132);
                                1: divisor = number of off-diagonal entries */
    const double* radius_override; /* optional host array [levels-1]; entries > 0 replace the computed radius */
    int32_t build_galerkin;  /* 1: also A_c = P^T A P on every level (needs A_val) */
    /* optional: called on the building thread as soon as a level is complete on the device, with the
     * hierarchy under construction (valid for gmg_hierarchy_level_info / gmg_hierarchy_fetch_async of levels
     * <= level only). Lets a finished level leave the device while the next one builds. */
    void (*on_level)(void* user, gmg_hierarchy* h, int level);
    void* user;
} gmg_hierarchy_options;
/* With build_galerkin, on_level is called TWICE per level: first with (level | GMG_LEVEL_P_READY) as soon as P is
 * complete (its three arrays may be fetched; A_c does not exist yet), then with the plain level once A_c is. */
enum { GMG_LEVEL_P_READY = 0x40000000 };

typedef struct {
    int64_t n, nnz;          /* fine side of this level */
    int64_t m, nnz_c;        /* coarse side */
    int64_t T;               /* Voronoi-neighbour triangles of the coarse graph */
    int64_t nnz_p;           /* stored entries of P */
    int64_t nnz_ac;          /* stored entries of A_c (0 unless build_galerkin) */
    double radius;
    int32_t sampler_rounds, parent_rounds;
    gmg_prolongation_stats stats;
    /* device milliseconds per stage (CUDA events on the context stream):
       0 edge lengths+radius, 1 sampling, 2 parents, 3 cluster CSR, 4 coarse edges, 5 coarse points,
       6 triangle fans, 7 prolongation rows, 8 Galerkin */
    double stage_ms[9];
    /* stage 8 split: transpose R = P^T; first product T = P^T A (sizing pass + values); second product A_c = T P
     * (sizing pass + values + compaction). 0 unless build_galerkin */
    double galerkin_ms[3];
} gmg_level_info;

int gmg_hierarchy_build(gmg_ctx* ctx, const gmg_input* in, const gmg_hierarchy_options* opt, gmg_hierarchy** out);
void gmg_hierarchy_destroy(gmg_hierarchy* h);
int gmg_hierarchy_num_levels(const gmg_hierarchy* h);                 /* number of prolongations built */
int gmg_hierarchy_level_info(const gmg_hierarchy* h, int level, gmg_level_info* out);

/* Copy one array of level `level` to the host. Sizes follow gmg_level_info. */
typedef enum {
    GMG_H_SAMPLES = 0,        /* int64[m]      fastDiscSample result */
    GMG_H_PARENTS = 1,        /* int64[n]      assignParents result */
    GMG_H_COARSE_COLPTR = 2,  /* int32[m+1] */
    GMG_H_COARSE_ROWIDX = 3,  /* int32[nnz_c] */
    GMG_H_COARSE_VAL = 4,     /* f64[nnz_c] */
    GMG_H_COARSE_POINTS = 5,  /* f64[3m], column-major, ld = m */
    GMG_H_TRIANGLES = 6,      /* int64[3T] row-major */
    GMG_H_NORMALS = 7,        /* f64[3T] row-major */
    GMG_H_FAN_PTR = 8,        /* int64[m+1] */
    GMG_H_FAN = 9,            /* int64[3T] */
    GMG_H_P_ROWPTR = 10,      /* int32[n+1] */
    GMG_H_P_COLIDX = 11,      /* int32[nnz_p] */
    GMG_H_P_VAL = 12,         /* f64[nnz_p] */
    GMG_H_AC_COLPTR = 13,     /* int32[m+1]   A_c = P^T A P, compressed column-major like A */
    GMG_H_AC_ROWIDX = 14,     /* int32[nnz_ac] ascending per column, structural (un-pruned) pattern */
    GMG_H_AC_VAL = 15,        /* f64[nnz_ac] */
    GMG_H_FINE_VAL = 16,      /* f64[nnz] level-0 only: the toEdgeDistanceMatrix values */
    GMG_H_PROJECTED = 17      /* f64[3n], column-major, ld = n: projectedPoints(P, coarse_points) = P X_c
                                 (src/multigrid.cpp:500-510, test/main.cpp:147), computed on the device from the
                                 resident P and coarse points when it is fetched */
} gmg_hierarchy_array;
int gmg_hierarchy_fetch(gmg_ctx* ctx, const gmg_hierarchy* h, int level, int which, void* out);
/* The same copy on the context's COPY stream, not waited for (arrays without an int64 conversion only: the
 * coarse graph, points, normals, P, A_c, fine values). Call it from gmg_hierarchy_options.on_level so a
 * finished level leaves the device while the next one builds; `out` should be pinned host memory.
 * gmg_copy_sync waits for every such copy. */
int gmg_hierarchy_fetch_async(gmg_ctx* ctx, const gmg_hierarchy* h, int level, int which, void* out);
int gmg_copy_sync(gmg_ctx* ctx);

/* ------------------------------------------------------------------------------------------------
 * SpMV, multicolour Gauss-Seidel and the V-cycle that validates a hierarchy [not in reference;
 * BASELINE.json north star]. Matrices are compressed by outer index; gmg_spmv computes
 * y[i] = sum_k val[k] * x[inner[k]] over outer vector i (A x for row-major storage, A^T x for
 * column-major storage -- the same thing for the symmetric operators used here).
 * ------------------------------------------------------------------------------------------------ */
int gmg_spmv(gmg_ctx* ctx, int64_t n, int64_t n_inner, const int32_t* outer, const int32_t* inner, const double* val,
             const double* x, double* y);
/* `sweeps` multicolour Gauss-Seidel sweeps on A x = b, x updated in place. colours_out (optional,
 * int32[n]) receives the Jones-Plassmann colouring used; the sweep equals sequential Gauss-Seidel in
 * (colour, row) order. */
int gmg_gauss_seidel(gmg_ctx* ctx, int64_t n, const int32_t* outer, const int32_t* inner, const double* val,
                     const double* b, double* x, int sweeps, int32_t* colours_out, int32_t* num_colours_out);

typedef struct gmg_solver gmg_solver;
/* Needs a hierarchy built with build_galerkin = 1; `in` and `h` must outlive the solver. */
int gmg_solver_create(gmg_ctx* ctx, const gmg_input* in, const gmg_hierarchy* h, gmg_solver** out);
void gmg_solver_destroy(gmg_solver* s);
/* `cycles` V-cycles on A x = b from the x passed in: pre/post multicolour Gauss-Seidel sweeps per
 * level (post in reverse colour order), restriction r_c = P^T r, prolongation x += P e,
 * coarse_sweeps sweeps on the coarsest operator. residual_out (optional) has cycles+1 entries:
 * ||b - A x||_2 before the first and after every cycle. */
int gmg_solver_vcycle(gmg_ctx* ctx, gmg_solver* s, const double* b, double* x, int cycles, int pre, int post,
                      int coarse_sweeps, double* residual_out);
/* Conjugate gradients on A x = b preconditioned by one SYMMETRIC V(sweeps, sweeps) cycle from a zero guess (post
 * sweeps in reverse colour order; the coarsest level sweeps forward then backward). Stops when
 * ||b - A x||_2 <= rel_tol * ||b||_2 or after max_iter iterations; residual_out (optional, max_iter + 1
 * entries) receives the residual norm before the first and after every iteration. */
int gmg_solver_pcg(gmg_ctx* ctx, gmg_solver* s, const double* b, double* x, double rel_tol, int max_iter, int sweeps,
                   int coarse_sweeps, int* iterations_out, double* residual_out);
/* device milliseconds of the last gmg_spmv / gmg_solver_vcycle / gmg_solver_pcg: gmg_last_kernel_ms */

/* ------------------------------------------------------------------------------------------------
 * Row-sharded hierarchy build [not in reference; one process per GPU]. Fine vertices are split into
 * contiguous ranges, one per rank. Every rank holds the whole level graph; the work and the large
 * products (coverer lists, relaxation, coarse rows, rows of P) are computed for the owned range only.
 * A rank writes only the entries it owns of a replicated per-vertex buffer, and the HOST all-gathers
 * the owned slices after the phases named below (NCCL through torch.distributed in
 * gravo-mg_b200/sharded.py; an in-process loop-back in the single-GPU tests). Results equal the
 * single-GPU build bit for bit; P comes back sharded by rows.
 *
 * Driver loop per level (G = all-gather of that buffer over the ranks):
 *   LEVEL_BEGIN (io[0] = 0: no level left) ; G(LEN_PARTIAL) ; RADIUS ; SAMPLER_BUILD ; G(STATE) ;
 *   repeat { SAMPLER_CHECK -> io[0] = undecided vertices; stop at 0 ; SAMPLER_ROUND ; G(STATE) } ;
 *   SAMPLES -> io[0] = m ; repeat { PARENTS_PULL ; G(DIST) ; G(CHANGED) ; PARENTS_CHECK -> io[0] =
 *   improved vertices; stop at 0 } ; PARENTS_PRED ; G(PRED) ; PARENTS_RESOLVE ; G(PARENTS) ;
 *   COARSE_ROWS ; G(COARSE_COUNT) ; COARSE_PACK ; G(ROWS_ID) ; G(ROWS_VAL) ; COARSE_ASSEMBLE ;
 *   PROLONG ; [Galerkin, below] ; LEVEL_STATUS ; G(STATUS) ; LEVEL_END -> io = {m, nnz_c, T, nnz_p of the owned rows,
 *   nnz_ac}. LEVEL_STATUS puts the rank's deferred error flags and counters (zero-length edge, missing offer, coarse
 *   overflow / asymmetry, degenerate fan, fallback rows, owned rows) into its slice of STATUS; LEVEL_END evaluates
 *   the checks on the values of ALL ranks -- flags OR-ed, the fallback ratio of src/multigrid.cpp:488 over the whole
 *   level -- so that every rank raises the same error together (a rank that left the loop alone would leave the
 *   others waiting in the next all-gather) and the error behaviour equals the single-GPU build's.
 * With build_galerkin (A_c = P^T A P on every level, replicated because it is the next level's A):
 *   P_COUNT ; G(P_COUNT) ; P_PACK ; G(P_COL) ; G(P_VAL) ; GALERKIN_ROWS (transpose the owned coarse
 *   columns of P, two-pass SpGEMM on the owned outer vectors) ; G(COARSE_COUNT) ; GALERKIN_PACK ;
 *   G(ROWS_ID) ; G(ROWS_VAL) ; GALERKIN_ASSEMBLE.
 * ------------------------------------------------------------------------------------------------ */
typedef struct gmg_shard gmg_shard;
enum { GMG_PH_LEVEL_BEGIN = 0, GMG_PH_SAMPLER_BUILD, GMG_PH_SAMPLER_ROUND, GMG_PH_SAMPLER_CHECK, GMG_PH_SAMPLES,
       GMG_PH_PARENTS_PULL, GMG_PH_PARENTS_CHECK, GMG_PH_PARENTS_PRED, GMG_PH_PARENTS_RESOLVE, GMG_PH_COARSE_ROWS,
       GMG_PH_COARSE_PACK, GMG_PH_COARSE_ASSEMBLE, GMG_PH_PROLONG, GMG_PH_LEVEL_END, GMG_PH_RADIUS,
       GMG_PH_P_COUNT, GMG_PH_P_PACK, GMG_PH_GALERKIN_ROWS, GMG_PH_GALERKIN_PACK, GMG_PH_GALERKIN_ASSEMBLE,
       GMG_PH_LEVEL_STATUS };
enum { GMG_XB_STATE = 0, GMG_XB_CHANGED, GMG_XB_DIST, GMG_XB_PRED, GMG_XB_PARENTS, GMG_XB_COARSE_COUNT, GMG_XB_ROWS_ID,
       GMG_XB_ROWS_VAL, GMG_XB_LEN_PARTIAL, GMG_XB_P_COUNT, GMG_XB_P_COL, GMG_XB_P_VAL, GMG_XB_STATUS };
/* `in` is the WHOLE level-0 problem, resident on this rank's device (the same on every rank); it must
 * outlive the shard */
int gmg_shard_create(gmg_ctx* ctx, int rank, int world, const gmg_input* in, const gmg_hierarchy_options* opt,
                     gmg_shard** out);
void gmg_shard_destroy(gmg_shard* s);
/* io: host array of at least 8 int64 */
int gmg_shard_phase(gmg_ctx* ctx, gmg_shard* s, int phase, int64_t* io);
/* exchange buffer: device pointer, element size, slice length in elements; the buffer holds
 * slice * world elements and rank r owns [r * slice, (r + 1) * slice) */
int gmg_shard_buffer(gmg_shard* s, int which, void** ptr, int32_t* elem_bytes, int64_t* slice);
/* ROWS_ID, ROWS_VAL, P_COL, P_VAL have slices of different sizes (packed rows): the buffer is the final
 * row-major array itself and rank r owns elements [offsets[r], offsets[r + 1]) (world + 1 entries). */
int gmg_shard_buffer_v(gmg_shard* s, int which, void** ptr, int32_t* elem_bytes, int64_t* offsets);
/* The same driver loop in ONE call, exchanges as ncclAllGather on the context's stream (csrc/shard_nccl.cu). NCCL is
 * opened at run time (libnccl.so.2), not linked. The communicator is the library's own: rank 0 asks for a unique id
 * (128 bytes), the caller hands it to every rank by any means, every rank calls gmg_nccl_init.
 * stats (optional, 2 entries): exchanges issued, bytes gathered (summed over the ranks' slices). */
typedef struct gmg_nccl gmg_nccl;
int gmg_nccl_unique_id(void* out128);
int gmg_nccl_init(gmg_ctx* ctx, const void* id128, int rank, int world, gmg_nccl** out);
void gmg_nccl_destroy(gmg_nccl* comm);
int gmg_shard_build_nccl(gmg_ctx* ctx, gmg_shard* s, gmg_nccl* comm, int build_galerkin, int64_t* stats);
int gmg_shard_num_levels(const gmg_shard* s);
int gmg_shard_level_info(const gmg_shard* s, int level, gmg_level_info* out, int64_t* row_lo, int64_t* row_hi);
/* array ids of gmg_hierarchy_array; replicated arrays are complete on every rank, the P arrays hold
 * the owned rows [row_lo, row_hi) (rowptr: row_hi - row_lo + 1 entries starting at 0) */
int gmg_shard_fetch(gmg_ctx* ctx, const gmg_shard* s, int level, int which, void* out);

/* ------------------------------------------------------------------------------------------------
 * Row-sharded smoothing and V-cycle on a shard built with build_galerkin [not in reference]. Every rank
 * keeps full-length vectors per level (levels = prolongations + 1); only the owned entries [lo, hi) and the
 * ghost entries its rows read are valid. An op acts on the owned rows; between ops the HOST exchanges
 * ghost entries (packed all-to-all over NCCL in gravo-mg_b200/sharded.py), guided by the halo masks.
 *   First replicated level (small levels, at least the coarsest): all-gather b, VCYCLE_ALL with
 *   arg = pre | post << 8 | coarse_sweeps << 16 (every rank finishes the cycle whole; no halo exchange).
 *   V-cycle on level l:  for each colour k { GS_COLOUR(k) ; exchange x over the ghosts of colour k } (pre
 *   sweeps) ; RESIDUAL ; exchange r over pattern R ; RESTRICT ; ZERO_X(l+1) ; recurse ; exchange x(l+1) over
 *   pattern P ; PROLONG ; exchange x over pattern A ; post sweeps in reverse colour order.
 * ------------------------------------------------------------------------------------------------ */
typedef struct gmg_shard_solver gmg_shard_solver;
enum { GMG_SV_X = 0, GMG_SV_B, GMG_SV_R, GMG_SV_COLOUR, GMG_SV_MASK };
enum { GMG_HP_A = 0, GMG_HP_R, GMG_HP_P };
enum { GMG_SO_GS_COLOUR = 0, GMG_SO_RESIDUAL, GMG_SO_RESTRICT, GMG_SO_PROLONG, GMG_SO_ZERO_X, GMG_SO_SUMSQ_R,
       GMG_SO_GS_ALL, GMG_SO_VCYCLE_ALL };
/* `shard` and `in` must outlive the solver */
/* replicate_below: levels >= 1 with at most this many vertices (the build keeps P whole up to 4 M), and
 * every coarser one, run whole on every rank: no halo exchange below that level. The coarsest always does. */
int gmg_shard_solver_create(gmg_ctx* ctx, const gmg_shard* shard, const gmg_input* in, int64_t replicate_below,
                            gmg_shard_solver** out);
void gmg_shard_solver_destroy(gmg_shard_solver* s);
int gmg_shard_solver_num_levels(const gmg_shard_solver* s);
/* io (>= 8 int64) = {n, lo, hi, slice length of the level's split (owner of vertex j = j / slice), colours,
 * 1 if the level is replicated} */
int gmg_shard_solver_level(const gmg_shard_solver* s, int level, int64_t* io);
/* device pointer of a level's vector: X, B, R f64[max(n, slice * world)]; COLOUR int32[n]; MASK uint8[same as X] */
int gmg_shard_solver_vector(gmg_shard_solver* s, int level, int which, void** ptr);
/* Fill the MASK vector with a halo pattern: 1 for every entry outside the owned range that the owned rows of
 * the operator read. GMG_HP_A: x of `level` read by A; GMG_HP_R: r of `level` read by the restriction to
 * level + 1; GMG_HP_P: x of level + 1 read by the prolongation (written to the mask OF level + 1).
 * *space = number of entries of that mask. */
int gmg_shard_solver_halo_mask(gmg_ctx* ctx, gmg_shard_solver* s, int level, int pattern, int64_t* space);
/* One operator on the owned rows. GS_COLOUR: arg = colour. RESIDUAL: r = b - A x. RESTRICT: b(level+1) = R r.
 * PROLONG: x += P x(level+1). ZERO_X: x = 0 (all entries). SUMSQ_R: *out = sum over owned rows of r^2.
 * GS_ALL: arg sweeps over ALL rows (the coarsest level, after the host has all-gathered b: X, B and R hold
 * slice * world entries, rank r owning slice r, so an in-place all-gather of the owned slices completes them).
 * VCYCLE_ALL (replicated levels only): the rest of the V-cycle from this level down on ALL rows, after the host
 * has all-gathered b; arg = pre | post << 8 | coarse_sweeps << 16. */
int gmg_shard_solver_op(gmg_ctx* ctx, gmg_shard_solver* s, int level, int op, int arg, double* out);

#ifdef __cplusplus
}
#endif
#endif /* GRAVOMG_B200_H */
