/* TEST INFRASTRUCTURE ONLY -- never linked into, imported by or called from the product path.
 *
 * Single-thread CPU checker and timed CPU baseline for the parts of the path the reference does NOT
 * contain (README.md:6 of the reference: only the prolongation/restriction builders were kept):
 * the Galerkin product A_c = P^T A P, SpMV and multicolour Gauss-Seidel, which BASELINE.json's north
 * star adds. "parity unpinned by the reference": these are pinned against scipy (tests/test_oracle.py),
 * with Eigen's product semantics -- structural (un-pruned) pattern, sorted inner indices.
 *
 * Matrices are compressed by OUTER index like Eigen::SparseMatrix<double> (include/gravomg/utility.h:15):
 * A and A_c column-major (outer = column), P row-major (include/gravomg/utility.h:18).
 *
 * orc_galerkin computes the two products the CUDA path computes (gravo-mg_b200/csrc/k_galerkin.cu), each sum in
 * the plain sequential order:
 *     T(i,k)   = sum_l A(l,k) * P(l,i)     entries of A's outer vector k in stored (ascending l) order, entries of
 *                                          P's row l in stored order; the first product of a key is stored as it is
 *     A_c(i,j) = sum_k T(i,k) * P(k,j)     k ascending over the rows of P that store column j
 * The kernel follows the same order (one thread per outer vector, no atomics), so the comparison is bit-exact
 * instead of 1e-12.
 *
 * Build: gcc -std=c11 -O2 -ffp-contract=off (no FMA contraction, no -march).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

void orc_solver_free(void* p) { free(p); }

static int cmp_i32(const void* a, const void* b) {
    const int32_t x = *(const int32_t*)a, y = *(const int32_t*)b;
    return (x > y) - (x < y);
}

/* A: n x n by outer vector (a_ptr[n+1], a_idx, a_val). P: n x m row-major. Result: m outer vectors,
 * out_ptr[m+1] caller-allocated, *out_idx / *out_val malloc'ed (release with orc_solver_free). */
void orc_galerkin(int64_t n, const int32_t* a_ptr, const int32_t* a_idx, const double* a_val, int64_t m,
                  const int32_t* p_ptr, const int32_t* p_col, const double* p_val, int32_t* out_ptr,
                  int32_t** out_idx, double** out_val) {
    /* R = P^T, fine rows ascending within every coarse row (stable counting sort) */
    int32_t* r_ptr = (int32_t*)calloc((size_t)m + 2, sizeof(int32_t));
    const int64_t nnzp = p_ptr[n];
    int32_t* r_col = (int32_t*)malloc(sizeof(int32_t) * (size_t)(nnzp ? nnzp : 1));
    double* r_val = (double*)malloc(sizeof(double) * (size_t)(nnzp ? nnzp : 1));
    for (int64_t e = 0; e < nnzp; ++e) r_ptr[p_col[e] + 2]++;
    for (int64_t c = 0; c < m; ++c) r_ptr[c + 2] += r_ptr[c + 1];
    for (int64_t r = 0; r < n; ++r)
        for (int32_t e = p_ptr[r]; e < p_ptr[r + 1]; ++e) {
            const int32_t pos = r_ptr[p_col[e] + 1]++;
            r_col[pos] = (int32_t)r; r_val[pos] = p_val[e];
        }
    /* sparse accumulator over the coarse index space, shared by both products */
    int32_t* slot = (int32_t*)malloc(sizeof(int32_t) * (size_t)(m ? m : 1));
    for (int64_t i = 0; i < m; ++i) slot[i] = -1;
    size_t cap = 1024;
    int32_t* list = (int32_t*)malloc(sizeof(int32_t) * cap);
    double* acc = (double*)malloc(sizeof(double) * cap);
    /* ---- T = P^T A by fine column */
    int64_t* t_ptr = (int64_t*)malloc(sizeof(int64_t) * ((size_t)n + 1));
    size_t tcap = (size_t)(8 * n + 16), tnnz = 0;
    int32_t* t_idx = (int32_t*)malloc(sizeof(int32_t) * tcap);
    double* t_val = (double*)malloc(sizeof(double) * tcap);
    t_ptr[0] = 0;
    for (int64_t k = 0; k < n; ++k) {
        size_t cnt = 0;
        for (int32_t q = a_ptr[k]; q < a_ptr[k + 1]; ++q) {
            const int32_t l = a_idx[q];
            const double w = a_val[q];
            for (int32_t t = p_ptr[l]; t < p_ptr[l + 1]; ++t) {
                const int32_t i = p_col[t];
                const double v = w * p_val[t];
                if (slot[i] >= 0) { acc[slot[i]] = acc[slot[i]] + v; continue; }
                if (cnt == cap) { cap *= 2; list = (int32_t*)realloc(list, sizeof(int32_t) * cap); acc = (double*)realloc(acc, sizeof(double) * cap); }
                slot[i] = (int32_t)cnt; list[cnt] = i; acc[cnt] = v; ++cnt;
            }
        }
        if (tnnz + cnt > tcap) { while (tnnz + cnt > tcap) tcap *= 2; t_idx = (int32_t*)realloc(t_idx, sizeof(int32_t) * tcap); t_val = (double*)realloc(t_val, sizeof(double) * tcap); }
        for (size_t s = 0; s < cnt; ++s) { t_idx[tnnz] = list[s]; t_val[tnnz] = acc[s]; ++tnnz; slot[list[s]] = -1; }
        t_ptr[k + 1] = (int64_t)tnnz;
    }
    /* ---- A_c = T P by coarse column */
    size_t ocap = (size_t)(8 * m + 16), onnz = 0;
    int32_t* oi = (int32_t*)malloc(sizeof(int32_t) * ocap);
    double* ov = (double*)malloc(sizeof(double) * ocap);
    int32_t* order = (int32_t*)malloc(sizeof(int32_t) * cap);
    size_t order_cap = cap;
    out_ptr[0] = 0;
    for (int64_t j = 0; j < m; ++j) {
        size_t cnt = 0;
        for (int32_t e = r_ptr[j]; e < r_ptr[j + 1]; ++e) {
            const int32_t k = r_col[e];
            const double w = r_val[e];
            for (int64_t s = t_ptr[k]; s < t_ptr[k + 1]; ++s) {
                const int32_t i = t_idx[s];
                const double v = t_val[s] * w;
                if (slot[i] >= 0) { acc[slot[i]] = acc[slot[i]] + v; continue; }
                if (cnt == cap) { cap *= 2; list = (int32_t*)realloc(list, sizeof(int32_t) * cap); acc = (double*)realloc(acc, sizeof(double) * cap); }
                slot[i] = (int32_t)cnt; list[cnt] = i; acc[cnt] = v; ++cnt;
            }
        }
        if (order_cap < cap) { order_cap = cap; order = (int32_t*)realloc(order, sizeof(int32_t) * order_cap); }
        memcpy(order, list, sizeof(int32_t) * cnt);
        qsort(order, cnt, sizeof(int32_t), cmp_i32);
        if (onnz + cnt > ocap) { while (onnz + cnt > ocap) ocap *= 2; oi = (int32_t*)realloc(oi, sizeof(int32_t) * ocap); ov = (double*)realloc(ov, sizeof(double) * ocap); }
        for (size_t s = 0; s < cnt; ++s) { oi[onnz] = order[s]; ov[onnz] = acc[slot[order[s]]]; ++onnz; }
        for (size_t s = 0; s < cnt; ++s) slot[list[s]] = -1;
        out_ptr[j + 1] = (int32_t)onnz;
    }
    free(r_ptr); free(r_col); free(r_val); free(slot); free(list); free(acc); free(order); free(t_ptr); free(t_idx); free(t_val);
    *out_idx = oi; *out_val = ov;
}

/* y[i] = sum_k val[k] * x[inner[k]] over outer vector i, summed in stored order (the kernel uses 8 strided partial
 * sums and a shuffle tree: compare with a tolerance) */
void orc_spmv(int64_t n, const int32_t* outer, const int32_t* inner, const double* val, const double* x, double* y) {
    for (int64_t i = 0; i < n; ++i) {
        double s = 0.0;
        for (int32_t k = outer[i]; k < outer[i + 1]; ++k) s += val[k] * x[inner[k]];
        y[i] = s;
    }
}

/* `sweeps` Gauss-Seidel sweeps in the given row order (rows[0..n): the rows bucketed by colour, ascending inside a
 * bucket = the (colour, row) order of the multicolour kernel): x[i] = (b[i] - sum_{j != i} a_ij x[j]) / a_ii */
void orc_gauss_seidel(int64_t n, const int32_t* outer, const int32_t* inner, const double* val, const double* b,
                      const int32_t* rows, int sweeps, double* x) {
    for (int s = 0; s < sweeps; ++s)
        for (int64_t r = 0; r < n; ++r) {
            const int32_t i = rows[r];
            double diag = 0.0, acc = 0.0;
            for (int32_t k = outer[i]; k < outer[i + 1]; ++k) {
                if (inner[k] == i) diag += val[k];
                else acc += val[k] * x[inner[k]];
            }
            x[i] = (b[i] - acc) / diag;
        }
}
