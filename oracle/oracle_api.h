/* TEST INFRASTRUCTURE ONLY -- never linked into, imported by or called from the product path.
 *
 * One raw-array C interface, implemented twice:
 *   ref_*  : oracle/ref_driver.cpp -- the UNMODIFIED reference sources
 *            (/root/reference/src/{multigrid,sampling,utility}.cpp) compiled against
 *            oracle/eigen_shim and called through their own GravoMG:: entry points.
 *            Built into oracle/_ref/libgravomg_ref.so.  "reference source on shim".
 *   orc_*  : oracle/gravomg_oracle.c -- a plain-C restatement of the same algorithms, every
 *            function citing the reference file:line it follows. Built into
 *            oracle/liboracle.so.
 * tests/ pin orc_* against ref_* and against tests/golden/, then use orc_* to check the CUDA path.
 *
 * Conventions (same as the reference's Eigen types, include/gravomg/utility.h:10-18):
 *   points   : column-major n x 3 doubles, i.e. x[0..n) y[0..n) z[0..n)        (PointMatrix)
 *   edge mat : compressed column-major, int32 colptr[n+1], int32 rowidx[nnz], f64 val[nnz]
 *   P        : compressed row-major,    int32 rowptr[n+1], int32 colidx[nnz], f64 val[nnz]
 *   indices crossing as std::vector<Index> are int64
 * Variable-size outputs are malloc'ed by the callee and released with PFX_free().
 */
#ifndef GRAVOMG_ORACLE_API_H
#define GRAVOMG_ORACLE_API_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#ifndef ORACLE_PFX
#error "define ORACLE_PFX(name) before including oracle_api.h"
#endif

void ORACLE_PFX(free)(void* p);

/* src/utility.cpp:50-56 */
void ORACLE_PFX(to_edge_distance_matrix)(int64_t n, const int32_t* colptr, const int32_t* rowidx,
                                         const double* points, double* out_val);

/* src/utility.cpp:58-71. out_edges is column-major nnz x 2 (col then row), like Eigen::MatrixXi */
void ORACLE_PFX(extract_edges)(int64_t n, const int32_t* colptr, const int32_t* rowidx, const double* val,
                               int32_t* out_edges, double* out_values);

/* src/multigrid.cpp:127-133 */
double ORACLE_PFX(average_edge_length)(int64_t n, const double* points, int64_t n_edges, const int32_t* edges);

/* src/sampling.cpp:7-53. selection_out has room for n entries; returns m. */
int64_t ORACLE_PFX(fast_disc_sample)(int64_t n, const double* points, const int32_t* colptr,
                                     const int32_t* rowidx, double radius, int64_t* selection_out);

/* src/multigrid.cpp:77-125 */
void ORACLE_PFX(assign_parents)(int64_t n, const double* points, const int32_t* colptr, const int32_t* rowidx,
                                int64_t m, const int64_t* samples, int64_t* parents_out);

/* src/multigrid.cpp:135-169. out_colptr has m+1 entries; *out_rowidx / *out_val are malloc'ed. */
void ORACLE_PFX(extract_coarse_edges)(int64_t n, const int32_t* colptr, const int32_t* rowidx, const double* val,
                                      int64_t m, const int64_t* parents, int32_t* out_colptr,
                                      int32_t** out_rowidx, double** out_val);

/* src/multigrid.cpp:171-207 */
void ORACLE_PFX(coarse_from_mean_of_fine_children)(int64_t n, const double* points, const int32_t* colptr,
                                                   const int32_t* rowidx, const int64_t* parents, int64_t m,
                                                   double* out_points);

/* src/multigrid.cpp:209-263. Returns T. *out_tri = int64[3T] (row-major triples), *out_normal =
 * f64[3T] (row-major), out_fan_ptr = int64[m+1], *out_fan = int64[3T]. */
int64_t ORACLE_PFX(construct_voronoi_triangles)(int64_t m, const double* points, const int32_t* colptr,
                                                const int32_t* rowidx, const double* val, int64_t** out_tri,
                                                double** out_normal, int64_t* out_fan_ptr, int64_t** out_fan);

/* src/multigrid.cpp:265-498. weighting: 0 barycentric, 1 uniform, 2 inverse distance.
 * out_rowptr has n+1 entries. Returns 0, or 1 if the reference's assert (:488) would fire. */
int ORACLE_PFX(construct_prolongation)(int64_t n, const double* fine_points, int64_t m, const double* coarse_points,
                                       const int32_t* c_colptr, const int32_t* c_rowidx, const double* c_val,
                                       const int64_t* parents, int weighting, int32_t* out_rowptr,
                                       int32_t** out_colidx, double** out_val);

/* src/multigrid.cpp:500-510 */
void ORACLE_PFX(projected_points)(int64_t n, int64_t m, const int32_t* rowptr, const int32_t* colidx,
                                  const double* val, const double* coarse_points, double* out_points);

#ifdef __cplusplus
}
#endif
#endif
