// Host-side stage functions over DEVICE arrays. The C ABI entry points (capi.cu) and the resident
// hierarchy builder (hierarchy.cu) are both thin sequences of these calls.
#pragma once

#include <cstring>

#include "common.cuh"

namespace gmg {

// ---- k_basic.cu -------------------------------------------------------------------------------
// col_mask (optional, n bytes): compute only the flagged columns (row-sharded build)
void edge_lengths(gmg_ctx* c, int32_t n, const int32_t* colptr, const int32_t* rowidx, const double* pts,
                  double* out_len, const uint8_t* col_mask = nullptr);
void extract_edges_dev(gmg_ctx* c, int32_t n, const int32_t* colptr, const int32_t* rowidx, const double* val,
                       int64_t nnz, int32_t* out_edges, double* out_values);
void average_edge_length_dev(gmg_ctx* c, int32_t n, const double* pts, int64_t n_edges, const int32_t* edges,
                             double* d_out);
// mean stored length = (sum of all entries) / denom, with a summation tree defined on groups of 32 columns
// so that the row-sharded build (owned groups per rank, partials all-gathered) reproduces it bit for bit
void mean_length_dev(gmg_ctx* c, int32_t n, int64_t nnz, const int32_t* colptr, const double* len, double denom,
                     double* d_out);
void length_group_partials(gmg_ctx* c, int32_t n, int32_t g_lo, int32_t g_hi, const int32_t* colptr, const double* len,
                           double* partial);
void mean_from_group_partials(gmg_ctx* c, int64_t groups, const double* partial, double denom, double* d_out);
void check_pattern(gmg_ctx* c, int32_t n, const int32_t* colptr, const int32_t* rowidx, bool need_symmetric);

// Device status block that a stage writes INSTEAD of synchronising with the host; the hierarchy builder
// fetches it once per level. A stage called without one (the per-function C ABI) uses a private block
// and checks it before returning, which is the old behaviour.
struct StageStatus {
    enum : int { kZeroEdge = 0, kNoOffer = 1, kParentRounds = 2, kCoarseHardOverflow = 3, kCoarseAsym = 4,
                 kProlongFallback = 5 /* rows handed to the fallback kernel */,
                 kProlongStats = 8 /* 4 x uint64 */, kProlongNnz = 16 /* int64 */, kWords = 32 };
    DevBuf<int32_t> dev;
    int32_t host[kWords] = {0};
    void begin(gmg_ctx* c) { if (!dev.p) dev.alloc(c, kWords); dev.zero(); }
    void fetch(gmg_ctx* c) { dev.download(host, kWords); sync(c); }
    int64_t host_i64(int word) const { int64_t v; memcpy(&v, host + word, sizeof v); return v; }
    void raise_parents() const;
    void raise_coarse() const;
    void raise_prolong(int64_t n) const;
};

// ---- k_sampler.cu -----------------------------------------------------------------------------
// Greedy disc sampling in priority order (index order = the reference's set). `len` holds the
// Euclidean length of every stored entry. Returns m; selection receives ascending fine indices.
struct SamplerStats { int rounds; };
// `path_bound` = nnz + sum over columns of deg^2 when the caller already knows it (saves a host round
// trip), or < 0 to have it computed here.
int32_t disc_sample(gmg_ctx* c, int32_t n, const int32_t* colptr, const int32_t* rowidx, const double* len,
                    double radius, int mode, uint64_t seed, DevBuf<int32_t>& selection, SamplerStats* stats,
                    int64_t path_bound = -1);
// *d_out (device) += sum over columns of deg^2
void degree_square_sum(gmg_ctx* c, int32_t n, const int32_t* colptr, unsigned long long* d_out);

// Row-sharded variant: rank-local lists of preceding coverers for the owned vertices [lo, hi); the
// uint8 state array (0 undecided, 1 in, 2 out) is replicated and all-gathered by the caller between rounds.
struct ShardSampler {
    int32_t lo = 0, hi = 0;
    DevBuf<long long> lptr;
    DevBuf<int32_t> lcnt, lst;
};
void shard_sampler_build(gmg_ctx* c, int32_t n, int32_t lo, int32_t hi, const int32_t* colptr, const int32_t* rowidx,
                         const double* len, double radius, int mode, uint64_t seed, uint8_t* state8, ShardSampler& out);
void shard_sampler_round(gmg_ctx* c, const ShardSampler& s, uint8_t* state8);
int64_t count_state(gmg_ctx* c, int64_t n, const uint8_t* state8, int what);
int32_t compact_selected(gmg_ctx* c, int32_t n, const uint8_t* state8, DevBuf<int32_t>& selection);

// ---- k_parents.cu -----------------------------------------------------------------------------
struct ParentStats { int rounds; };
void assign_parents_dev(gmg_ctx* c, int32_t n, const int32_t* colptr, const int32_t* rowidx, const double* len,
                        int32_t m, const int32_t* samples, int32_t* parents, ParentStats* stats,
                        StageStatus* status = nullptr);

// Row-sharded variant: pull relaxation over the owned vertices [lo, hi); dist (uint64 bit patterns of
// non-negative doubles), the "improved last round" bytes and srcid are replicated arrays.
void shard_parents_init(gmg_ctx* c, int32_t n, int32_t m, const int32_t* samples, unsigned long long* dist,
                        uint8_t* changed, int32_t* srcid);
void shard_parents_pull(gmg_ctx* c, int32_t lo, int32_t hi, const int32_t* colptr, const int32_t* rowidx,
                        const double* len, unsigned long long* dist, const uint8_t* changed_in, uint8_t* changed_out);
void shard_parents_pred(gmg_ctx* c, int32_t n, int32_t lo, int32_t hi, const int32_t* colptr, const int32_t* rowidx,
                        const double* len, const unsigned long long* dist, const int32_t* srcid, int32_t* pred,
                        int32_t* d_flags);
void shard_parents_resolve(gmg_ctx* c, int32_t n, int32_t lo, int32_t hi, const int32_t* pred, const int32_t* srcid,
                           int32_t* parents);

// ---- k_cluster.cu -----------------------------------------------------------------------------
struct Clusters {
    DevBuf<int32_t> ptr;    // m+1
    DevBuf<int32_t> child;  // n, ascending within each cluster
};
void build_clusters(gmg_ctx* c, int32_t n, int32_t m, const int32_t* parents, Clusters& out);

struct DeviceCsc {
    int32_t n = 0;
    int64_t nnz = 0;
    DevBuf<int32_t> colptr, rowidx;
    DevBuf<double> val;
};
void coarse_edges(gmg_ctx* c, int32_t n, int64_t nnz, const int32_t* colptr, const int32_t* rowidx, const double* val,
                  int32_t m, const int32_t* parents, const Clusters& cl, DeviceCsc& out, StageStatus* status = nullptr);
// The two halves of coarse_edges, used separately by the row-sharded build: rows of the owned clusters
// [p_lo, p_hi), then (after the ranks have exchanged their rows) the column-major matrix.
struct CoarseRows {
    int32_t p_lo = 0, p_hi = 0;
    int64_t bound = -1;          // candidates in the owned rows; < 0: compute (one host round trip)
    DevBuf<int32_t> boff;        // p_hi - p_lo + 1 scratch offsets
    DevBuf<int32_t> sQ;          // neighbour clusters, ascending per row
    DevBuf<double> st;           // min-reduced value per neighbour
};
void coarse_rows_range(gmg_ctx* c, int32_t n, const int32_t* colptr, const int32_t* rowidx, const double* val,
                       int32_t p_lo, int32_t p_hi, const int32_t* parents, const Clusters& cl, int32_t* cnt,
                       CoarseRows& out, int32_t* d_hard_overflow);
void coarse_assemble(gmg_ctx* c, int32_t m, const int32_t* boff, const int32_t* cnt, const int32_t* sQ, const double* st,
                     DeviceCsc& out, int32_t* d_asym);
void sample_points(gmg_ctx* c, int32_t n, const double* pts, int32_t m, const int32_t* samples, double* out_pts);   // nested mode
void coarse_points(gmg_ctx* c, int32_t n, const double* pts, const int32_t* colptr, const int32_t* rowidx, int32_t m,
                   const Clusters& cl, double* out_pts);

// ---- k_triangles.cu ---------------------------------------------------------------------------
struct Triangles {
    int64_t T = 0;
    DevBuf<int32_t> tri;      // 3T, row-major (v0,v1,v2) ascending triples in lexicographic order
    DevBuf<double> normal;    // 3T, row-major
    DevBuf<int32_t> fan_ptr;  // m+1
    DevBuf<int32_t> fan_tri;  // 3T triangle ids, ascending per vertex
    DevBuf<int32_t> fan_o1;   // 3T: second vertex after rotating the owner to the front
    DevBuf<int32_t> fan_o2;   // 3T: third vertex after rotation
};
// v0_mask (optional, m bytes): enumerate only the triangles whose smallest vertex is flagged. Ids then
// number the enumerated subset in the same lexicographic order, and only vertices whose whole closed
// neighbourhood is flagged have complete fans (the row-sharded build flags what its own rows of P need).
void voronoi_triangles(gmg_ctx* c, int32_t m, int64_t nnz, const double* pts, const int32_t* colptr,
                       const int32_t* rowidx, const double* val, Triangles& out, const uint8_t* v0_mask = nullptr);

// ---- k_prolong.cu -----------------------------------------------------------------------------
struct DeviceCsr {
    int32_t rows = 0, cols = 0;
    int64_t nnz = 0;
    DevBuf<int32_t> rowptr, colidx;
    DevBuf<double> val;
};
void prolongation(gmg_ctx* c, int32_t n, const double* fine_pts, int32_t m, const double* coarse_pts,
                  const int32_t* c_colptr, const int32_t* c_rowidx, const int32_t* parents, const Triangles& tr,
                  int weighting, DeviceCsr& P, gmg_prolongation_stats* stats, StageStatus* status = nullptr,
                  int32_t f_lo = 0, int32_t f_hi = -1);   // rows [f_lo, f_hi) only (default: all); P.rows = f_hi - f_lo
// With a StageStatus, P.nnz and the counters are only known after status->fetch(): this fills them in.
void finish_prolongation(const StageStatus& status, DeviceCsr& P, gmg_prolongation_stats* stats);
void projected_points_dev(gmg_ctx* c, int32_t n, int32_t m, const int32_t* rowptr, const int32_t* colidx,
                          const double* val, const double* coarse_pts, double* out_pts);

// ---- k_galerkin.cu ----------------------------------------------------------------------------
// A_c = P^T A P as a two-pass (symbolic, numeric) SpGEMM. A and A_c are compressed column-major;
// the structural pattern is kept (no numerical pruning), inner indices ascending.
// With [c_lo, c_hi) only those outer vectors are formed (Ac.n = c_hi - c_lo, indexed from c_lo): the
// row-sharded build, where P is the whole operator and every rank forms the coarse rows it owns.
// marks (optional): 4 events recorded before the transpose, after it, after the symbolic and after the numeric pass
void galerkin(gmg_ctx* c, int32_t n, const int32_t* a_colptr, const int32_t* a_rowidx, const double* a_val,
              const DeviceCsr& P, DeviceCsc& Ac, int32_t c_lo = 0, int32_t c_hi = -1, DeviceCsr* keep_R = nullptr,
              cudaEvent_t* marks = nullptr);
// R = P^T in CSR (rows = coarse columns [c_lo, c_hi) of P), fine indices ascending per row
// sort_rows = false leaves every row in arrival order (the Galerkin kernel sorts what it stages)
void transpose_csr(gmg_ctx* c, const DeviceCsr& P, DeviceCsr& R, int32_t c_lo = 0, int32_t c_hi = -1, bool sort_rows = true);

// ---- k_solver.cu ------------------------------------------------------------------------------
struct Colouring {
    int num_colours = 0;
    DevBuf<int32_t> colour;        // n
    DevBuf<int32_t> rows;          // n, rows bucketed by colour, ascending inside a bucket
    std::vector<int32_t> offsets;  // num_colours + 1 (host)
};
// y = alpha * (M x) + b over outer vectors (b may be null)
void spmv(gmg_ctx* c, int32_t n, const int32_t* outer, const int32_t* inner, const double* val, const double* x,
          double* y, double alpha, const double* b);
void axpy(gmg_ctx* c, int32_t n, double a, const double* x, double* y);
void colour_graph(gmg_ctx* c, int32_t n, const int32_t* outer, const int32_t* inner, Colouring& out);
void gauss_seidel(gmg_ctx* c, const Colouring& col, const int32_t* outer, const int32_t* inner, const double* val,
                  const double* b, double* x, int sweeps, bool reverse);
// one colour's worth: x[i] = (b[i] - sum_{j != i} a_ij x[j]) / a_ii for the listed (mutually independent) rows
void gauss_seidel_rows(gmg_ctx* c, const int32_t* rows, int32_t count, const int32_t* outer, const int32_t* inner,
                       const double* val, const double* b, double* x);
double norm2(gmg_ctx* c, int32_t n, const double* x);
double dot(gmg_ctx* c, int32_t n, const double* x, const double* y);
void xpby(gmg_ctx* c, int32_t n, const double* x, double b, double* y);     // y = x + b * y

// index conversions at the ABI boundary (std::vector<Eigen::Index> is int64, device indices int32)
void narrow_i64(gmg_ctx* c, const int64_t* in, int32_t* out, int64_t n, int64_t upper, int32_t* d_flag);
void widen_i32(gmg_ctx* c, const int32_t* in, int64_t* out, int64_t n);

}  // namespace gmg
