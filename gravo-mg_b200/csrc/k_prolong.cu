// Rows of the prolongation operator P (GravoMG::constructProlongation, reference
// src/multigrid.cpp:265-498; inTriangle :18-55; inverseDistanceWeights :63-75) and
// GravoMG::projectedPoints (:500-510).
//
// One thread per fine vertex f with parent c:
//   degree(c) == 0 : (c, 1)                                                        (:295-299)
//   degree(c) == 1 : clamp-projected weights along the only edge                  (:301-334)
//   degree(c) >= 2 : walk c's fan in ascending triangle id and take the FIRST triangle whose three
//                    barycentrics are >= 0 (:374-380 breaks on the first hit); else the edge
//                    fallback; else the three-nearest fallback.
// The reference threads a std::map<Index,float> through the fan walk (:37-48, :414-421). Only this
// is ever consumed from it: the SMALLEST key that (a) was never flagged -1 by any fan triangle and
// (b) whose first-seen value, a float-narrowed norm that depends on (f, c, key) only, is finite.
// Flags are sticky, and when no triangle matched every fan triangle was visited, so the map is
// replaced by per-triangle flag bits (fans of <= 32 triangles, in a register) or by recomputation.
// Rows are emitted like setFromTriplets leaves them: duplicates summed in order, columns
// ascending, explicit zeros kept.
// Algorithmic bytes: 32n + 12nnzP + 32m + 4nnz_c + 48T.
#include "stages.h"

#include <algorithm>

namespace gmg {

struct ProlongArgs {
    int32_t f_lo, f_hi;    // fine rows built by this call; row_cnt/row_col/row_val are indexed by f - f_lo
    int32_t n, m;
    const double* fp;
    const double* cp;
    const int32_t* c_colptr;
    const int32_t* c_rowidx;
    const int32_t* parents;
    const int32_t* fan_ptr;
    const int32_t* fan_tri;
    const int32_t* fan_o1;
    const int32_t* fan_o2;
    const double* normal;
    int weighting;
    int nested;            // GMG_NESTED: a fine point that coincides with its parent's coarse point gets the row (parent, 1.0)
    int32_t* row_cnt;      // n (+1 for the scan)
    int32_t* row_col;      // 3n scratch
    double* row_val;       // 3n scratch
    unsigned long long* stats;  // [0] no triangle, [1] edges found, [2] fallback, [3] invalid input
};

// inTriangle (:18-55) for the triangle (c, k1, k2) already rotated so that c comes first.
// Returns true when all three barycentrics are >= 0 and |dist| < DBL_MAX (:50-52, :374).
__device__ __forceinline__ bool in_triangle(V3 p, V3 v1, V3 v2, V3 v3, V3 nrm, double* bary, bool& bad1, bool& bad2) {
    const double dist = dot(p - v1, nrm);                                     // :29
    const V3 pproj = p - nrm * dist;                                          // :30
    const double dbl_area = dot(cross(v2 - v1, v3 - v1), nrm);                // :32
    bary[0] = dot(cross(v3 - v2, pproj - v2), nrm) / dbl_area;                // :33
    bary[1] = dot(cross(v1 - v3, pproj - v3), nrm) / dbl_area;                // :34
    bary[2] = 1. - bary[0] - bary[1];                                         // :35
    bad1 = bary[0] < 0. || bary[1] < 0.;                                      // :43
    bad2 = bary[0] < 0. || bary[2] < 0.;                                      // :46
    return bary[0] >= 0. && bary[1] >= 0. && bary[2] >= 0. && fabs(dist) < DBL_MAX;
}

// value first stored for key k in the reference's map (:37-42), narrowed to float; usable iff >= 0
// and < DBL_MAX once widened (:415)
__device__ __forceinline__ bool edge_value_usable(V3 p, V3 pc, V3 pk) {
    const V3 v1p = p - pc, e = pk - pc;
    const float d = float(norm(v1p - e * dot(v1p, e)));
    return double(d) >= 0. && double(d) < DBL_MAX;
}

// :309-315 and :424-429
__device__ __forceinline__ double edge_weight(V3 p, V3 pc, V3 pq) {
    const V3 e = pq - pc;
    double len = norm(e);
    len = (len < 1e-8) ? 1e-8 : len;
    double w = dot(p - pc, normalized(e)) / len;
    if (w < 0.0) w = 0.0;
    if (1.0 < w) w = 1.0;
    return w;
}

// :63-75
__device__ __forceinline__ void inverse_distance_weights(const double* __restrict__ cp, int32_t m, V3 p,
                                                         const int32_t* ids, int count, double* w) {
    double sum = 0.0;
    for (int j = 0; j < count; ++j) {
        const double d = norm(p - ld3(cp, m, ids[j]));
        w[j] = 1.0 / ((1e-8 < d) ? d : 1e-8);
        sum += w[j];
    }
    for (int j = 0; j < count; ++j) w[j] = w[j] / sum;
}

__device__ __forceinline__ void emit_row(const ProlongArgs& a, int32_t f, const int32_t* ids, const double* w, int count) {
    int32_t c[3]; double v[3]; int k = 0;
    for (int i = 0; i < count; ++i) {
        int j = 0;
        for (; j < k; ++j) if (c[j] == ids[i]) break;
        if (j < k) v[j] = v[j] + w[i]; else { c[k] = ids[i]; v[k] = w[i]; ++k; }
    }
    if (k > 1 && c[1] < c[0]) { int32_t tc = c[0]; c[0] = c[1]; c[1] = tc; double tv = v[0]; v[0] = v[1]; v[1] = tv; }
    if (k > 2) {
        if (c[2] < c[1]) { int32_t tc = c[1]; c[1] = c[2]; c[2] = tc; double tv = v[1]; v[1] = v[2]; v[2] = tv; }
        if (c[1] < c[0]) { int32_t tc = c[0]; c[0] = c[1]; c[1] = tc; double tv = v[0]; v[0] = v[1]; v[1] = tv; }
    }
    const int64_t r = f - a.f_lo;
    a.row_cnt[r] = k;
    for (int i = 0; i < k; ++i) { a.row_col[3 * r + i] = c[i]; a.row_val[3 * r + i] = v[i]; }
}

// Common case, one thread per fine vertex: degree-0 / degree-1 parents and the first-hit walk over the parent's fan.
// A vertex that lies in none of the fan's triangles (a few per cent) is appended to `fallback` and finished by
// prolong_fallback_kernel: keeping the rare edge / three-nearest fallbacks out of this kernel keeps its registers
// and code size down.
__global__ void __launch_bounds__(kBlock, 4) prolong_rows_kernel(ProlongArgs a, int32_t* __restrict__ fallback,
                                                                 int32_t* __restrict__ fallback_count) {
    const int64_t gid = int64_t(a.f_lo) + int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (gid >= a.f_hi) return;
    const int32_t f = int32_t(gid);
    const int32_t c = __ldg(a.parents + f);
    const V3 p = ld3(a.fp, a.n, f), pc = ld3(a.cp, a.m, c);
    const int32_t cb = __ldg(a.c_colptr + c), ce = __ldg(a.c_colptr + c + 1);
    int32_t ids[3]; double w[3];
    // "Exact matches get a weight of 1.0 (if in nested mode)" -- the reference's todo at :291-292
    if (a.nested && p.x == pc.x && p.y == pc.y && p.z == pc.z) {
        ids[0] = c; w[0] = 1.0; emit_row(a, f, ids, w, 1);
        return;
    }
    if (cb == ce) {                                                                       // :295-299
        ids[0] = c; w[0] = 1.0; emit_row(a, f, ids, w, 1);
        return;
    }
    if (ce - cb == 1) {                                                                   // :301-334
        const int32_t q = __ldg(a.c_rowidx + cb);
        const double wq = edge_weight(p, pc, ld3(a.cp, a.m, q));
        ids[0] = c; ids[1] = q;
        if (a.weighting == GMG_BARYCENTRIC) { w[0] = 1 - wq; w[1] = wq; }
        else if (a.weighting == GMG_UNIFORM) { w[0] = w[1] = 1. / 2; }
        else inverse_distance_weights(a.cp, a.m, p, ids, 2, w);
        emit_row(a, f, ids, w, 2);
        return;
    }
    const int32_t fb = __ldg(a.fan_ptr + c), fe = __ldg(a.fan_ptr + c + 1);
    for (int32_t i = fb; i < fe; ++i) {                                                   // :356
        const int32_t k1 = __ldg(a.fan_o1 + i), k2 = __ldg(a.fan_o2 + i);
        const int64_t t = __ldg(a.fan_tri + i);
        const V3 nn{__ldg(a.normal + 3 * t), __ldg(a.normal + 3 * t + 1), __ldg(a.normal + 3 * t + 2)};
        double bary[3]; bool b1, b2;
        if (in_triangle(p, pc, ld3(a.cp, a.m, k1), ld3(a.cp, a.m, k2), nn, bary, b1, b2)) {   // :374-380
            ids[0] = c; ids[1] = k1; ids[2] = k2;
            if (a.weighting == GMG_BARYCENTRIC) { w[0] = bary[0]; w[1] = bary[1]; w[2] = bary[2]; }
            else if (a.weighting == GMG_UNIFORM) { w[0] = w[1] = w[2] = 1. / 3; }
            else inverse_distance_weights(a.cp, a.m, p, ids, 3, w);
            emit_row(a, f, ids, w, 3);
            return;
        }
    }
    warp_count(a.stats + 0);
    fallback[warp_append(fallback_count)] = f;
}

// The vertices the walk above found no triangle for (grid-stride over the list; its length is read on the device).
__global__ void __launch_bounds__(kBlock) prolong_fallback_kernel(ProlongArgs a, const int32_t* __restrict__ fallback,
                                                                  const int32_t* __restrict__ fallback_count) {
    const int32_t count = *fallback_count;
    for (int64_t r = int64_t(blockIdx.x) * kBlock + threadIdx.x; r < count; r += int64_t(gridDim.x) * kBlock) {
        const int32_t f = fallback[r];
        const int32_t c = __ldg(a.parents + f);
        const V3 p = ld3(a.fp, a.n, f), pc = ld3(a.cp, a.m, c);
        const int32_t cb = __ldg(a.c_colptr + c), ce = __ldg(a.c_colptr + c + 1);
        const int32_t fb = __ldg(a.fan_ptr + c), fe = __ldg(a.fan_ptr + c + 1);
        const int32_t F = fe - fb;
        int32_t ids[3]; double w[3];
        // the edge-side flags of the walk that failed (2 bits per fan triangle, kept when F <= 32)
        unsigned long long bits = 0ull;
        for (int32_t i = 0; i < F && i < 32; ++i) {
            const int64_t t = __ldg(a.fan_tri + fb + i);
            const V3 nn{__ldg(a.normal + 3 * t), __ldg(a.normal + 3 * t + 1), __ldg(a.normal + 3 * t + 2)};
            double bary[3]; bool b1, b2;
            in_triangle(p, pc, ld3(a.cp, a.m, __ldg(a.fan_o1 + fb + i)), ld3(a.cp, a.m, __ldg(a.fan_o2 + fb + i)), nn, bary, b1, b2);
            bits |= (unsigned long long)(b1 ? 1 : 0) << (2 * i) | (unsigned long long)(b2 ? 1 : 0) << (2 * i + 1);
        }
        // First fallback (:407-448): smallest never-flagged key with a usable first-seen value.
        int32_t best = 0x7fffffff;
        for (int32_t i = 0; i < F; ++i) {
            for (int s = 0; s < 2; ++s) {
                const int32_t k = __ldg((s ? a.fan_o2 : a.fan_o1) + fb + i);
                if (k >= best) continue;
                bool bad = false;
                for (int32_t j = 0; j < F && !bad; ++j) {
                    const int32_t j1 = __ldg(a.fan_o1 + fb + j), j2 = __ldg(a.fan_o2 + fb + j);
                    if (j1 != k && j2 != k) continue;
                    bool b1, b2;
                    if (F <= 32) { b1 = (bits >> (2 * j)) & 1; b2 = (bits >> (2 * j + 1)) & 1; }
                    else {
                        const int64_t t = __ldg(a.fan_tri + fb + j);
                        const V3 nn{__ldg(a.normal + 3 * t), __ldg(a.normal + 3 * t + 1), __ldg(a.normal + 3 * t + 2)};
                        double bary[3];
                        in_triangle(p, pc, ld3(a.cp, a.m, j1), ld3(a.cp, a.m, j2), nn, bary, b1, b2);
                    }
                    // the map is keyed by vertex: tri[1] and tri[2] may name the same key (:43-48)
                    bad = (j1 == k && b1) || (j2 == k && b2);
                }
                if (!bad && edge_value_usable(p, pc, ld3(a.cp, a.m, k))) best = k;
            }
        }
        if (best != 0x7fffffff) {                                                         // :422-448
            warp_count(a.stats + 1);
            const double w2 = edge_weight(p, pc, ld3(a.cp, a.m, best));
            ids[0] = c; ids[1] = best;
            if (a.weighting == GMG_BARYCENTRIC) { w[0] = 1. - w2; w[1] = w2; }
            else if (a.weighting == GMG_UNIFORM) { w[0] = w[1] = 1. / 2; }
            else inverse_distance_weights(a.cp, a.m, p, ids, 2, w);
            emit_row(a, f, ids, w, 2);
            continue;
        }
        // Second fallback (:449-483): the two nearest coarse neighbours of c by (distance, index), plus c.
        warp_count(a.stats + 2);
        double d1 = 0, d2 = 0; int32_t i1 = -1, i2 = -1;
        for (int32_t k = cb; k < ce; ++k) {
            const int32_t nb = __ldg(a.c_rowidx + k);
            if (nb == c) continue;                                                        // :462
            const double d = norm(p - ld3(a.cp, a.m, nb));
            // std::pair operator< : (d, nb) < (d1, i1)
            if (i1 < 0 || d < d1 || (!(d1 < d) && nb < i1)) { d2 = d1; i2 = i1; d1 = d; i1 = nb; }
            else if (i2 < 0 || d < d2 || (!(d2 < d) && nb < i2)) { d2 = d; i2 = nb; }
        }
        if (i2 < 0) {  // fewer than two non-self neighbours: out-of-range read in the reference (:472)
            warp_count(a.stats + 3);
            ids[0] = c; w[0] = 1.0; emit_row(a, f, ids, w, 1);
            continue;
        }
        ids[0] = c; ids[1] = i1; ids[2] = i2;
        inverse_distance_weights(a.cp, a.m, p, ids, 3, w);                                // :478
        emit_row(a, f, ids, w, 3);
    }
}

__global__ void __launch_bounds__(kBlock) prolong_compact_kernel(int32_t n, const int32_t* __restrict__ rowptr,
                                                                 const int32_t* __restrict__ row_col,
                                                                 const double* __restrict__ row_val,
                                                                 int32_t* __restrict__ colidx, double* __restrict__ val) {
    const int64_t f = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (f >= n) return;
    const int32_t b = rowptr[f], k = rowptr[f + 1] - b;
    for (int i = 0; i < k; ++i) { colidx[b + i] = row_col[3 * f + i]; val[b + i] = row_val[3 * f + i]; }
}

void prolongation(gmg_ctx* c, int32_t n, const double* fine_pts, int32_t m, const double* coarse_pts,
                  const int32_t* c_colptr, const int32_t* c_rowidx, const int32_t* parents, const Triangles& tr,
                  int weighting, DeviceCsr& P, gmg_prolongation_stats* stats, StageStatus* status, int32_t f_lo,
                  int32_t f_hi) {
    const int nested = (weighting & GMG_NESTED) ? 1 : 0;
    weighting &= ~GMG_NESTED;
    GMG_REQUIRE(weighting >= 0 && weighting <= 2, GMG_ERR_INVALID, "unknown weighting scheme");
    if (f_hi < 0) { f_lo = 0; f_hi = n; }
    const int32_t n_all = n;                       // fine vertices of the level (stride of the point arrays)
    n = f_hi - f_lo;                               // rows built here
    StageStatus local;
    StageStatus* st = status ? status : &local;
    if (!status) local.begin(c);
    P.rows = n; P.cols = m;
    P.rowptr.alloc(c, int64_t(n) + 1);
    DevBuf<int32_t> row_col(c, 3 * int64_t(n));
    DevBuf<double> row_val(c, 3 * int64_t(n));
    unsigned long long* dstats = reinterpret_cast<unsigned long long*>(st->dev.p + StageStatus::kProlongStats);
    int64_t* total = reinterpret_cast<int64_t*>(st->dev.p + StageStatus::kProlongNnz);
    P.rowptr.zero();
    ProlongArgs a{f_lo, f_hi, n_all, m, fine_pts, coarse_pts, c_colptr, c_rowidx, parents, tr.fan_ptr.p, tr.fan_tri.p, tr.fan_o1.p,
                  tr.fan_o2.p, tr.normal.p, weighting, nested, P.rowptr.p, row_col.p, row_val.p, dstats};
    if (n > 0) {
        DevBuf<int32_t> fallback(c, n);
        int32_t* fallback_count = st->dev.p + StageStatus::kProlongFallback;      // zeroed with the status block
        GMG_LAUNCH(c, prolong_rows_kernel, div_up(n, kBlock), kBlock, 0, a, fallback.p, fallback_count);
        const int grid = std::min(div_up(n, kBlock), 2 * c->sm_count);
        GMG_LAUNCH(c, prolong_fallback_kernel, grid, kBlock, 0, a, fallback.p, fallback_count);
    }
    exclusive_scan_i32(c, P.rowptr.p, P.rowptr.p, int64_t(n) + 1, total);
    // rows hold at most three entries: the arrays are sized 3n, so the host need not know nnz yet
    P.colidx.alloc(c, 3 * int64_t(n));
    P.val.alloc(c, 3 * int64_t(n));
    if (n > 0) GMG_LAUNCH(c, prolong_compact_kernel, div_up(n, kBlock), kBlock, 0, n, P.rowptr.p, row_col.p, row_val.p,
                          P.colidx.p, P.val.p);
    if (status) return;
    local.fetch(c);
    finish_prolongation(local, P, stats);
    local.raise_prolong(n);
}

void finish_prolongation(const StageStatus& status, DeviceCsr& P, gmg_prolongation_stats* stats) {
    P.nnz = status.host_i64(StageStatus::kProlongNnz);
    if (stats) {
        stats->no_triangle_found = status.host_i64(StageStatus::kProlongStats);
        stats->edges_found = status.host_i64(StageStatus::kProlongStats + 2);
        stats->fallback_count = status.host_i64(StageStatus::kProlongStats + 4);
    }
}

void StageStatus::raise_prolong(int64_t n) const {
    GMG_REQUIRE(host_i64(kProlongStats + 6) == 0, GMG_ERR_INVALID,
                "constructProlongation: a coarse point with degree >= 2 has fewer than two non-self neighbours");
    GMG_REQUIRE(n == 0 || float(host_i64(kProlongStats + 4)) / float(n) < 0.5f, GMG_ERR_FALLBACK_RATIO,
                "constructProlongation: more than half of the fine points fell back to three-nearest weights");
}

// projectedPoints (:500-510): row f = sum over stored entries in ascending column of coarse * weight
__global__ void __launch_bounds__(kBlock) projected_points_kernel(int32_t n, int32_t m, const int32_t* __restrict__ rowptr,
                                                                  const int32_t* __restrict__ colidx,
                                                                  const double* __restrict__ val,
                                                                  const double* __restrict__ cp, double* __restrict__ out) {
    const int64_t f = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (f >= n) return;
    double x = 0.0, y = 0.0, z = 0.0;
    for (int32_t k = rowptr[f]; k < rowptr[f + 1]; ++k) {
        const int32_t col = __ldg(colidx + k);
        const double v = __ldg(val + k);
        x = x + __ldg(cp + col) * v; y = y + __ldg(cp + m + col) * v; z = z + __ldg(cp + 2 * int64_t(m) + col) * v;
    }
    out[f] = x; out[n + f] = y; out[2 * int64_t(n) + f] = z;
}

void projected_points_dev(gmg_ctx* c, int32_t n, int32_t m, const int32_t* rowptr, const int32_t* colidx,
                          const double* val, const double* coarse_pts, double* out_pts) {
    if (n > 0) GMG_LAUNCH(c, projected_points_kernel, div_up(n, kBlock), kBlock, 0, n, m, rowptr, colidx, val, coarse_pts, out_pts);
}

}  // namespace gmg
