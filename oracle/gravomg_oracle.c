/* TEST INFRASTRUCTURE ONLY -- never linked into, imported by or called from the product path.
 *
 * Plain-C restatement of Gravo MG's hierarchy-construction algorithms over raw arrays. Every
 * function cites the reference file:line it follows (paths relative to /root/reference).
 *
 * PARITY PINNING: the reference ships no tests, golden vectors or fixtures,
 * so this oracle is pinned by executing the reference itself: tests/test_oracle.py compares
 * every orc_* function bit-for-bit with ref_* (the unmodified reference sources compiled against
 * oracle/eigen_shim, oracle/_ref/libgravomg_ref.so), and tests/golden/ holds outputs of ref_*
 * generated in the build container by tests/golden/make_golden.py. The arithmetic order of 3-vector
 * reductions is the shim's ((x+y)+z), which real Eigen 3.4 is believed, not verified, to share.
 *
 * Build: gcc -std=c11 -O2 -ffp-contract=off (no FMA contraction, no -march).
 */
#define ORACLE_PFX(name) orc_##name
#include "oracle_api.h"

#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

/* ---- 3-vector helpers; association order (x+y)+z as in oracle/eigen_shim/Eigen/Core ---------- */
typedef struct { double x, y, z; } v3;

static v3 pt(const double* p, int64_t n, int64_t i) { v3 r = {p[i], p[n + i], p[2 * n + i]}; return r; }
static v3 sub(v3 a, v3 b) { v3 r = {a.x - b.x, a.y - b.y, a.z - b.z}; return r; }
static v3 scale(v3 a, double s) { v3 r = {a.x * s, a.y * s, a.z * s}; return r; }
static double dot(v3 a, v3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
static double norm(v3 a) { return sqrt(dot(a, a)); }
static v3 cross(v3 a, v3 b) {
    v3 r = {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
    return r;
}
static v3 normalized(v3 a) {
    double z = dot(a, a);
    if (z > 0.0) { double s = sqrt(z); v3 r = {a.x / s, a.y / s, a.z / s}; return r; }
    return a;
}
static double edge_len(const double* p, int64_t n, int64_t a, int64_t b) { return norm(sub(pt(p, n, a), pt(p, n, b))); }

/* value at (row, col) of a compressed column-major matrix, 0 when absent (Eigen coeff()) */
static double csc_coeff(const int32_t* colptr, const int32_t* rowidx, const double* val, int64_t row, int64_t col) {
    int32_t lo = colptr[col], hi = colptr[col + 1];
    while (lo < hi) {
        int32_t mid = lo + (hi - lo) / 2;
        if (rowidx[mid] < row) lo = mid + 1; else hi = mid;
    }
    if (lo < colptr[col + 1] && rowidx[lo] == row) return val[lo];
    return 0.0;
}

void orc_free(void* p) { free(p); }

/* src/utility.cpp:50-56 -- every stored entry (row r, column c) becomes |p_r - p_c| */
void orc_to_edge_distance_matrix(int64_t n, const int32_t* colptr, const int32_t* rowidx, const double* points,
                                 double* out_val) {
    for (int64_t c = 0; c < n; ++c)
        for (int32_t k = colptr[c]; k < colptr[c + 1]; ++k) out_val[k] = edge_len(points, n, rowidx[k], c);
}

/* src/utility.cpp:58-71 -- row e of the edge list is (col, row) of stored entry e */
void orc_extract_edges(int64_t n, const int32_t* colptr, const int32_t* rowidx, const double* val,
                       int32_t* out_edges, double* out_values) {
    int64_t nnz = colptr[n], e = 0;
    for (int64_t c = 0; c < n; ++c)
        for (int32_t k = colptr[c]; k < colptr[c + 1]; ++k, ++e) {
            out_values[e] = val[k];
            out_edges[e] = (int32_t)c;
            out_edges[nnz + e] = rowidx[k];
        }
}

/* src/multigrid.cpp:127-133 -- std::transform_reduce over a random-access range. libstdc++ 13
 * (<numeric>:442-457) consumes four elements at a time as init += ((e0+e1)+(e2+e3)) and then adds
 * the tail one by one. Row e of the edge list is (i, j) = (edges[e], edges[n_edges+e]). */
double orc_average_edge_length(int64_t n, const double* points, int64_t n_edges, const int32_t* edges) {
    double init = 0.0;
    int64_t i = 0;
#define LEN(e) edge_len(points, n, edges[n_edges + (e)], edges[(e)])
    for (; n_edges - i >= 4; i += 4) {
        double v1 = LEN(i) + LEN(i + 1);
        double v2 = LEN(i + 2) + LEN(i + 3);
        init = init + (v1 + v2);
    }
    for (; i < n_edges; ++i) init = init + LEN(i);
#undef LEN
    return init / (double)(n_edges - n);
}

/* src/sampling.cpp:7-53 -- greedy in index order; one hop within radius or two hops whose summed
 * lengths stay within radius mark a vertex visited. Stored values are never read. */
int64_t orc_fast_disc_sample(int64_t n, const double* points, const int32_t* colptr, const int32_t* rowidx,
                             double radius, int64_t* selection_out) {
    unsigned char* visited = (unsigned char*)calloc((size_t)(n ? n : 1), 1);
    int64_t m = 0;
    for (int64_t i = 0; i < n; ++i) {                                   /* :22 */
        if (visited[i]) continue;                                       /* :25 */
        selection_out[m++] = i;                                         /* :28 */
        for (int32_t a = colptr[i]; a < colptr[i + 1]; ++a) {           /* :31 */
            int64_t j = rowidx[a];
            double d_ij = edge_len(points, n, j, i);                    /* :34 */
            if (d_ij < radius) {                                        /* :35 */
                visited[j] = 1;                                         /* :38 */
                for (int32_t b = colptr[j]; b < colptr[j + 1]; ++b) {   /* :43 */
                    int64_t k = rowidx[b];
                    if (d_ij + edge_len(points, n, k, j) < radius) visited[k] = 1; /* :44-45 */
                }
            }
        }
    }
    free(visited);
    return m;
}

/* The option the reference marks as a todo (src/sampling.cpp:11-12: "std::size_t min_selection = 0, std::size_t
 * max_selection = max"), with the semantics the CUDA path implements [not in reference]:
 *   max: the greedy loop stops selecting once max_selection vertices are selected = the first max_selection entries
 *        of the ascending selection;
 *   min: a radius that yields fewer than min_selection samples is halved and the sampling repeated, until there are
 *        enough (or every vertex is selected). *radius_used receives the radius of the run that was kept. */
int64_t orc_fast_disc_sample_bounded(int64_t n, const double* points, const int32_t* colptr, const int32_t* rowidx,
                                     double radius, int64_t min_selection, int64_t max_selection, int64_t* selection_out,
                                     double* radius_used) {
    int64_t m = 0;
    for (int pass = 0; pass < 64; ++pass) {
        m = orc_fast_disc_sample(n, points, colptr, rowidx, radius, selection_out);
        if (m >= min_selection || m >= n) break;
        radius = radius / 2;
    }
    if (radius_used) *radius_used = radius;
    return m > max_selection ? max_selection : m;
}

/* ---- binary min-heap of (distance, vertex), ordered like std::greater<pair<double,Index>> ---- */
typedef struct { double d; int64_t v; } heap_item;
typedef struct { heap_item* a; size_t n, cap; } heap;
static int item_less(heap_item x, heap_item y) { return x.d < y.d || (!(y.d < x.d) && x.v < y.v); }
static void heap_push(heap* h, heap_item it) {
    if (h->n == h->cap) { h->cap = h->cap ? 2 * h->cap : 1024; h->a = (heap_item*)realloc(h->a, h->cap * sizeof(heap_item)); }
    size_t i = h->n++;
    while (i > 0) {
        size_t p = (i - 1) / 2;
        if (!item_less(it, h->a[p])) break;
        h->a[i] = h->a[p]; i = p;
    }
    h->a[i] = it;
}
static heap_item heap_pop(heap* h) {
    heap_item top = h->a[0], last = h->a[--h->n];
    size_t i = 0;
    for (;;) {
        size_t c = 2 * i + 1;
        if (c >= h->n) break;
        if (c + 1 < h->n && item_less(h->a[c + 1], h->a[c])) ++c;
        if (!item_less(h->a[c], last)) break;
        h->a[i] = h->a[c]; i = c;
    }
    if (h->n) h->a[i] = last;
    return top;
}

/* src/multigrid.cpp:77-125 -- multi-source Dijkstra; pops are the total order (distance, index),
 * which is all std::priority_queue guarantees, so any correct heap gives the same pop sequence.
 * Edge lengths are recomputed from positions (:107); candidate = edge + path (:108); strict < (:111). */
void orc_assign_parents(int64_t n, const double* points, const int32_t* colptr, const int32_t* rowidx, int64_t m,
                        const int64_t* samples, int64_t* parents_out) {
    double* dist = (double*)malloc(sizeof(double) * (size_t)(n ? n : 1));
    heap h = {0, 0, 0};
    for (int64_t i = 0; i < n; ++i) { parents_out[i] = 0; dist[i] = DBL_MAX; }   /* :82-84 */
    for (int64_t c = 0; c < m; ++c) {                                             /* :89-93 */
        parents_out[samples[c]] = c;
        heap_item it = {0.0, samples[c]};
        heap_push(&h, it);
        dist[samples[c]] = 0.0;
    }
    while (h.n) {                                                                 /* :96 */
        heap_item top = heap_pop(&h);
        int64_t fine = top.v;
        for (int32_t k = colptr[fine]; k < colptr[fine + 1]; ++k) {               /* :104 */
            int64_t nb = rowidx[k];
            double cand = edge_len(points, n, fine, nb) + top.d;                  /* :107-108 */
            if (cand < dist[nb]) {                                                /* :111 */
                parents_out[nb] = parents_out[fine];
                dist[nb] = cand;
                heap_item it = {cand, nb};
                heap_push(&h, it);
            }
        }
    }
    free(h.a);
    free(dist);
}

/* ---- growable per-column lists, standing in for Eigen's uncompressed insertion mode ---------- */
typedef struct { int32_t* idx; double* val; int32_t n, cap; } col_list;
static int32_t col_find(const col_list* c, int32_t key, int* found) {
    int32_t lo = 0, hi = c->n;
    while (lo < hi) { int32_t mid = lo + (hi - lo) / 2; if (c->idx[mid] < key) lo = mid + 1; else hi = mid; }
    *found = (lo < c->n && c->idx[lo] == key);
    return lo;
}
static void col_insert(col_list* c, int32_t pos, int32_t key, double v) {
    if (c->n == c->cap) {
        c->cap = c->cap ? 2 * c->cap : 8;
        c->idx = (int32_t*)realloc(c->idx, sizeof(int32_t) * (size_t)c->cap);
        c->val = (double*)realloc(c->val, sizeof(double) * (size_t)c->cap);
    }
    memmove(c->idx + pos + 1, c->idx + pos, sizeof(int32_t) * (size_t)(c->n - pos));
    memmove(c->val + pos + 1, c->val + pos, sizeof(double) * (size_t)(c->n - pos));
    c->idx[pos] = key; c->val[pos] = v; c->n++;
}

/* src/multigrid.cpp:135-169. For each fine column `fine` and stored neighbour nb with parents
 * P != Q: t = coeff(fine, P) + value(nb, fine)  (:150-151; P, a COARSE index, is used as a FINE
 * column number -- reproduced as written). coarse(P,Q): if the current coefficient reads 0 (absent
 * or stored zero) it is overwritten by t, else min-reduced (:156-164). */
void orc_extract_coarse_edges(int64_t n, const int32_t* colptr, const int32_t* rowidx, const double* val, int64_t m,
                              const int64_t* parents, int32_t* out_colptr, int32_t** out_rowidx, double** out_val) {
    col_list* cols = (col_list*)calloc((size_t)(m ? m : 1), sizeof(col_list));
    for (int64_t fine = 0; fine < n; ++fine) {
        int64_t P = parents[fine];
        for (int32_t k = colptr[fine]; k < colptr[fine + 1]; ++k) {
            int64_t Q = parents[rowidx[k]];
            if (P == Q) continue;                                                   /* :148 */
            double t = csc_coeff(colptr, rowidx, val, fine, P) + val[k];            /* :150-151 */
            col_list* c = &cols[Q];                                                 /* entry (row P, col Q) */
            int found;
            int32_t pos = col_find(c, (int32_t)P, &found);
            double cur = found ? c->val[pos] : 0.0;
            if (cur == 0) {                                                         /* :156-159 */
                if (found) c->val[pos] = t; else col_insert(c, pos, (int32_t)P, t);
            } else {                                                                /* :160-164 */
                c->val[pos] = (t < cur) ? t : cur;                                  /* std::min(cur, t) */
            }
        }
    }
    int64_t nnz = 0;
    for (int64_t q = 0; q < m; ++q) { out_colptr[q] = (int32_t)nnz; nnz += cols[q].n; }
    out_colptr[m] = (int32_t)nnz;
    *out_rowidx = (int32_t*)malloc(sizeof(int32_t) * (size_t)(nnz ? nnz : 1));
    *out_val = (double*)malloc(sizeof(double) * (size_t)(nnz ? nnz : 1));
    for (int64_t q = 0; q < m; ++q) {
        memcpy(*out_rowidx + out_colptr[q], cols[q].idx, sizeof(int32_t) * (size_t)cols[q].n);
        memcpy(*out_val + out_colptr[q], cols[q].val, sizeof(double) * (size_t)cols[q].n);
        free(cols[q].idx); free(cols[q].val);
    }
    free(cols);
}

/* src/multigrid.cpp:171-207. Children of c = ordered set of fine vertices with parent c (:179-181);
 * a set of exactly one element also takes every stored neighbour of that element (:185-191); the
 * coarse point is the component-wise sum in ascending fine index from (0,0,0), divided by the set
 * size (:198-203; std::set iterators are bidirectional, so transform_reduce runs its plain loop). */
void orc_coarse_from_mean_of_fine_children(int64_t n, const double* points, const int32_t* colptr,
                                           const int32_t* rowidx, const int64_t* parents, int64_t m,
                                           double* out_points) {
    int64_t* start = (int64_t*)calloc((size_t)(m + 2), sizeof(int64_t));
    int64_t* child = (int64_t*)malloc(sizeof(int64_t) * (size_t)(n ? n : 1));
    for (int64_t f = 0; f < n; ++f) start[parents[f] + 2]++;
    for (int64_t c = 0; c < m; ++c) start[c + 2] += start[c + 1];
    for (int64_t f = 0; f < n; ++f) child[start[parents[f] + 1]++] = f;      /* ascending within a cluster */
    /* now start[c] .. start[c+1] is cluster c */
    for (int64_t c = 0; c < m; ++c) {
        int64_t b = start[c], e = start[c + 1];
        double sx = 0.0, sy = 0.0, sz = 0.0;
        size_t count = 0;
        if (e - b == 1) {
            int64_t only = child[b];
            int self_done = 0;
            for (int32_t k = colptr[only]; k <= colptr[only + 1]; ++k) {     /* merge {only} into its sorted column */
                int64_t v = (k < colptr[only + 1]) ? rowidx[k] : INT64_MAX;
                if (!self_done && only <= v) {
                    sx = sx + points[only]; sy = sy + points[n + only]; sz = sz + points[2 * n + only];
                    ++count; self_done = 1;
                    if (v == only) continue;
                }
                if (k < colptr[only + 1]) {
                    sx = sx + points[v]; sy = sy + points[n + v]; sz = sz + points[2 * n + v];
                    ++count;
                }
            }
        } else {
            for (int64_t i = b; i < e; ++i) {
                int64_t f = child[i];
                sx = sx + points[f]; sy = sy + points[n + f]; sz = sz + points[2 * n + f];
            }
            count = (size_t)(e - b);
        }
        out_points[c] = sx / (double)count;             /* Point / size_t (:203) */
        out_points[m + c] = sy / (double)count;
        out_points[2 * m + c] = sz / (double)count;
    }
    free(start); free(child);
}

/* src/multigrid.cpp:209-263. v0 ascending; v1 over column v0 with v1 >= v0 (:225, strict skip);
 * v2 over the REST of column v0 with v2 >= v0 (:228-232); triangle iff coeff(v1, v2) != 0 (:237);
 * normal = normalize((p1-p0) x (p2-p0)) (:240-242). Ids are emission order; every id is appended
 * to the lists of its three vertices (:254-256). */
int64_t orc_construct_voronoi_triangles(int64_t m, const double* points, const int32_t* colptr, const int32_t* rowidx,
                                        const double* val, int64_t** out_tri, double** out_normal,
                                        int64_t* out_fan_ptr, int64_t** out_fan) {
    size_t cap = 1024, T = 0;
    int64_t* tri = (int64_t*)malloc(sizeof(int64_t) * 3 * cap);
    double* nrm = (double*)malloc(sizeof(double) * 3 * cap);
    for (int64_t v0 = 0; v0 < m; ++v0)
        for (int32_t a = colptr[v0]; a < colptr[v0 + 1]; ++a) {
            int64_t v1 = rowidx[a];
            if (v1 < v0) continue;
            for (int32_t b = a + 1; b < colptr[v0 + 1]; ++b) {
                int64_t v2 = rowidx[b];
                if (v2 < v0) continue;
                if (csc_coeff(colptr, rowidx, val, v1, v2) != 0) {
                    if (T == cap) {
                        cap *= 2;
                        tri = (int64_t*)realloc(tri, sizeof(int64_t) * 3 * cap);
                        nrm = (double*)realloc(nrm, sizeof(double) * 3 * cap);
                    }
                    v3 p0 = pt(points, m, v0);
                    v3 nn = normalized(cross(sub(pt(points, m, v1), p0), sub(pt(points, m, v2), p0)));
                    tri[3 * T] = v0; tri[3 * T + 1] = v1; tri[3 * T + 2] = v2;
                    nrm[3 * T] = nn.x; nrm[3 * T + 1] = nn.y; nrm[3 * T + 2] = nn.z;
                    ++T;
                }
            }
        }
    int64_t* fan = (int64_t*)malloc(sizeof(int64_t) * (3 * T + 1));
    int64_t* fill = (int64_t*)calloc((size_t)(m + 1), sizeof(int64_t));
    for (int64_t i = 0; i <= m; ++i) out_fan_ptr[i] = 0;
    for (size_t i = 0; i < 3 * T; ++i) out_fan_ptr[tri[i] + 1]++;
    for (int64_t v = 0; v < m; ++v) out_fan_ptr[v + 1] += out_fan_ptr[v];
    for (size_t t = 0; t < T; ++t)                      /* ids ascending => each list ascending */
        for (int k = 0; k < 3; ++k) { int64_t v = tri[3 * t + k]; fan[out_fan_ptr[v] + fill[v]++] = (int64_t)t; }
    free(fill);
    *out_tri = tri; *out_normal = nrm; *out_fan = fan;
    return (int64_t)T;
}

/* src/multigrid.cpp:63-75 */
static void inverse_distance_weights(const double* cp, int64_t m, v3 p, const int64_t* ids, int count, double* w) {
    double sum = 0.0;
    for (int j = 0; j < count; ++j) {
        double d = norm(sub(p, pt(cp, m, ids[j])));
        w[j] = 1.0 / ((1e-8 < d) ? d : 1e-8);           /* std::max(1e-8, d) */
        sum += w[j];
    }
    for (int j = 0; j < count; ++j) w[j] = w[j] / sum;
}

/* weight of the far end of edge (c, q) for point p: src/multigrid.cpp:309-315 and :424-429 */
static double edge_weight(v3 p, v3 pc, v3 pq) {
    v3 e = sub(pq, pc);
    double len = norm(e);
    len = (len < 1e-8) ? 1e-8 : len;                      /* std::max(len, 1e-8) */
    double w = dot(sub(p, pc), normalized(e)) / len;
    /* std::clamp(w,0,1) and std::min(std::max(w,0.),1.) agree on every input incl. NaN (-> NaN) */
    if (w < 0.0) w = 0.0;
    if (1.0 < w) w = 1.0;
    return w;
}

typedef struct { int32_t* col; double* val; size_t n, cap; int32_t* rowptr; } p_builder;
/* emit one row the way setFromTriplets leaves it: duplicates summed in order, columns ascending */
static void emit_row(p_builder* b, int64_t row, const int64_t* ids, const double* w, int count) {
    int64_t c[3]; double v[3]; int k = 0;
    for (int i = 0; i < count; ++i) {
        int j; for (j = 0; j < k; ++j) if (c[j] == ids[i]) break;
        if (j < k) v[j] = v[j] + w[i]; else { c[k] = ids[i]; v[k] = w[i]; ++k; }
    }
    for (int i = 1; i < k; ++i)
        for (int j = i; j > 0 && c[j] < c[j - 1]; --j) {
            int64_t tc = c[j]; c[j] = c[j - 1]; c[j - 1] = tc;
            double tv = v[j]; v[j] = v[j - 1]; v[j - 1] = tv;
        }
    if (b->n + 3 > b->cap) {
        b->cap = b->cap ? 2 * b->cap : 4096;
        b->col = (int32_t*)realloc(b->col, sizeof(int32_t) * b->cap);
        b->val = (double*)realloc(b->val, sizeof(double) * b->cap);
    }
    for (int i = 0; i < k; ++i) { b->col[b->n] = (int32_t)c[i]; b->val[b->n] = v[i]; b->n++; }
    b->rowptr[row + 1] = (int32_t)b->n;
}

/* std::map<Index,float> insideEdge of src/multigrid.cpp:353, as a small sorted array */
typedef struct { int64_t* key; float* val; int n, cap; } edge_map;
static int emap_find(const edge_map* mp, int64_t key, int* found) {
    int lo = 0, hi = mp->n;
    while (lo < hi) { int mid = (lo + hi) / 2; if (mp->key[mid] < key) lo = mid + 1; else hi = mid; }
    *found = (lo < mp->n && mp->key[lo] == key);
    return lo;
}
static void emap_set(edge_map* mp, int64_t key, float v, int only_if_absent) {
    int found, pos = emap_find(mp, key, &found);
    if (found) { if (!only_if_absent) mp->val[pos] = v; return; }
    if (mp->n == mp->cap) {
        mp->cap = mp->cap ? 2 * mp->cap : 16;
        mp->key = (int64_t*)realloc(mp->key, sizeof(int64_t) * (size_t)mp->cap);
        mp->val = (float*)realloc(mp->val, sizeof(float) * (size_t)mp->cap);
    }
    memmove(mp->key + pos + 1, mp->key + pos, sizeof(int64_t) * (size_t)(mp->n - pos));
    memmove(mp->val + pos + 1, mp->val + pos, sizeof(float) * (size_t)(mp->n - pos));
    mp->key[pos] = key; mp->val[pos] = v; mp->n++;
}

/* src/multigrid.cpp:18-55. tri is already rotated so tri[0] is the parent. Returns >= 0 iff the
 * projection of p lies inside (all three barycentrics >= 0), else -1. */
static double in_triangle(v3 p, const int64_t* tri, v3 nrm, const double* cp, int64_t m, double* bary, edge_map* mp) {
    v3 v1 = pt(cp, m, tri[0]), v2 = pt(cp, m, tri[1]), v3_ = pt(cp, m, tri[2]);
    v3 v1ToP = sub(p, v1), e12 = sub(v2, v1), e13 = sub(v3_, v1);
    double dist = dot(sub(p, v1), nrm);                                         /* :29 */
    v3 pProj = sub(p, scale(nrm, dist));                                        /* :30 (dist * n) */
    double dblArea = dot(cross(sub(v2, v1), sub(v3_, v1)), nrm);                /* :32 */
    bary[0] = dot(cross(sub(v3_, v2), sub(pProj, v2)), nrm) / dblArea;          /* :33 */
    bary[1] = dot(cross(sub(v1, v3_), sub(pProj, v3_)), nrm) / dblArea;         /* :34 */
    bary[2] = 1. - bary[0] - bary[1];                                           /* :35 */
    emap_set(mp, tri[1], (float)norm(sub(v1ToP, scale(e12, dot(v1ToP, e12)))), 1);  /* :37-39 */
    emap_set(mp, tri[2], (float)norm(sub(v1ToP, scale(e13, dot(v1ToP, e13)))), 1);  /* :40-42 */
    if (bary[0] < 0. || bary[1] < 0.) emap_set(mp, tri[1], -1.f, 0);            /* :43-45 */
    if (bary[0] < 0. || bary[2] < 0.) emap_set(mp, tri[2], -1.f, 0);            /* :46-48 */
    if (bary[0] >= 0. && bary[1] >= 0. && bary[2] >= 0.) return fabs(dist);     /* :50-52 */
    return -1.;
}

/* src/multigrid.cpp:265-498 */
int orc_construct_prolongation(int64_t n, const double* fp, int64_t m, const double* cp, const int32_t* c_colptr,
                               const int32_t* c_rowidx, const double* c_val, const int64_t* parents, int weighting,
                               int32_t* out_rowptr, int32_t** out_colidx, double** out_val) {
    int64_t *tris, *fan, *fan_ptr = (int64_t*)malloc(sizeof(int64_t) * (size_t)(m + 1));
    double* nrms;
    orc_construct_voronoi_triangles(m, cp, c_colptr, c_rowidx, c_val, &tris, &nrms, fan_ptr, &fan);   /* :272-275 */
    p_builder pb = {0, 0, 0, 0, out_rowptr};
    out_rowptr[0] = 0;
    edge_map mp = {0, 0, 0, 0};
    int64_t fallback = 0;
    size_t cdcap = 0; double* cd_d = 0; int64_t* cd_i = 0;
    /* nested mode [the reference's todo at :291-292, not in reference]: weighting | 0x100 -- a fine point that coincides
     * with its parent's coarse point gets the row (parent, 1.0) */
    const int nested = (weighting & 0x100) != 0;
    weighting &= ~0x100;
    for (int64_t f = 0; f < n; ++f) {                                           /* :286 */
        int64_t c = parents[f];
        v3 p = pt(fp, n, f), pc = pt(cp, m, c);
        int32_t cb = c_colptr[c], ce = c_colptr[c + 1];
        int64_t ids[3]; double w[3];
        if (nested && p.x == pc.x && p.y == pc.y && p.z == pc.z) {              /* :291-292 */
            ids[0] = c; w[0] = 1.0; emit_row(&pb, f, ids, w, 1);
        } else if (cb == ce) {                                                  /* :295-299 */
            ids[0] = c; w[0] = 1.0; emit_row(&pb, f, ids, w, 1);
        } else if (ce - cb == 1) {                                              /* :301-334 */
            int64_t q = c_rowidx[cb];
            double wq = edge_weight(p, pc, pt(cp, m, q));
            ids[0] = c; ids[1] = q;
            if (weighting == 0) { w[0] = 1 - wq; w[1] = wq; }
            else if (weighting == 1) { w[0] = w[1] = 1. / 2; }
            else inverse_distance_weights(cp, m, p, ids, 2, w);
            emit_row(&pb, f, ids, w, 2);
        } else {
            int found = 0; int64_t chosen[3]; double cb_[3];
            mp.n = 0;
            for (int64_t i = fan_ptr[c]; i < fan_ptr[c + 1]; ++i) {             /* :356 */
                int64_t t = fan[i], tri[3] = {tris[3 * t], tris[3 * t + 1], tris[3 * t + 2]};
                while (tri[0] != c) { int64_t t0 = tri[0]; tri[0] = tri[1]; tri[1] = tri[2]; tri[2] = t0; }  /* :360 */
                v3 nn = {nrms[3 * t], nrms[3 * t + 1], nrms[3 * t + 2]};
                double bary[3];
                double d = in_triangle(p, tri, nn, cp, m, bary, &mp);           /* :364 */
                if (d >= 0. && d < DBL_MAX) {                                   /* :374-380: first hit, break */
                    found = 1; memcpy(chosen, tri, sizeof tri); memcpy(cb_, bary, sizeof bary);
                    break;
                }
            }
            if (found) {                                                        /* :385-405 */
                if (weighting == 0) memcpy(w, cb_, sizeof w);
                else if (weighting == 1) { w[0] = w[1] = w[2] = 1. / 3; }
                else inverse_distance_weights(cp, m, p, chosen, 3, w);
                emit_row(&pb, f, chosen, w, 3);
            } else {
                int found_edge = 0; int64_t q = 0;
                for (int i = 0; i < mp.n; ++i)                                  /* :414-421: first key with value >= 0 */
                    if ((double)mp.val[i] >= 0. && (double)mp.val[i] < DBL_MAX) { found_edge = 1; q = mp.key[i]; break; }
                if (found_edge) {                                               /* :422-448 */
                    double w2 = edge_weight(p, pc, pt(cp, m, q));
                    ids[0] = c; ids[1] = q;
                    if (weighting == 0) { w[0] = 1. - w2; w[1] = w2; }
                    else if (weighting == 1) { w[0] = w[1] = 1. / 2; }
                    else inverse_distance_weights(cp, m, p, ids, 2, w);
                    emit_row(&pb, f, ids, w, 2);
                } else {                                                        /* :449-483 */
                    size_t cnt = 0;
                    if ((size_t)(ce - cb) > cdcap) {
                        cdcap = (size_t)(ce - cb) * 2;
                        cd_d = (double*)realloc(cd_d, sizeof(double) * cdcap);
                        cd_i = (int64_t*)realloc(cd_i, sizeof(int64_t) * cdcap);
                    }
                    for (int32_t k = cb; k < ce; ++k) {                         /* :458-467 */
                        int64_t nb = c_rowidx[k];
                        if (nb == c) continue;
                        cd_d[cnt] = norm(sub(p, pt(cp, m, nb))); cd_i[cnt] = nb; ++cnt;
                    }
                    /* two smallest by (distance, index) == first two of the std::sort at :469 */
                    ids[0] = c;
                    for (int j = 0; j < 2; ++j) {
                        size_t best = j;
                        for (size_t i = j + 1; i < cnt; ++i)
                            if (cd_d[i] < cd_d[best] || (!(cd_d[best] < cd_d[i]) && cd_i[i] < cd_i[best])) best = i;
                        double td = cd_d[j]; cd_d[j] = cd_d[best]; cd_d[best] = td;
                        int64_t ti = cd_i[j]; cd_i[j] = cd_i[best]; cd_i[best] = ti;
                        ids[j + 1] = cd_i[j];
                    }
                    inverse_distance_weights(cp, m, p, ids, 3, w);              /* :478 */
                    emit_row(&pb, f, ids, w, 3);
                    ++fallback;
                }
            }
        }
    }
    free(tris); free(nrms); free(fan); free(fan_ptr); free(mp.key); free(mp.val); free(cd_d); free(cd_i);
    if (!pb.col) { pb.col = (int32_t*)malloc(sizeof(int32_t)); pb.val = (double*)malloc(sizeof(double)); }
    *out_colidx = pb.col; *out_val = pb.val;
    return ((float)fallback / (float)n < 0.5f) ? 0 : 1;                          /* assert at :488 */
}

/* src/multigrid.cpp:500-510 -- row f = sum over stored entries, ascending column, of coarse * w */
void orc_projected_points(int64_t n, int64_t m, const int32_t* rowptr, const int32_t* colidx, const double* val,
                          const double* coarse_points, double* out_points) {
    for (int64_t f = 0; f < n; ++f) {
        double x = 0.0, y = 0.0, z = 0.0;
        for (int32_t k = rowptr[f]; k < rowptr[f + 1]; ++k) {
            int64_t c = colidx[k];
            x = x + coarse_points[c] * val[k];
            y = y + coarse_points[m + c] * val[k];
            z = z + coarse_points[2 * m + c] * val[k];
        }
        out_points[f] = x; out_points[n + f] = y; out_points[2 * n + f] = z;
    }
}
