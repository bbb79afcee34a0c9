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
 * orc_galerkin computes, per outer vector i of A_c,
 *     c(l)   = sum_k P(k,i) * A(l,k)      k ascending over the rows of P that store column i,
 *                                         entries of A's outer vector k in stored (ascending) order
 *     out(j) = sum_l c(l) * P(l,j)        l ascending over the distinct l, entries of P's row l in stored order
 * with the SAME association as the CUDA kernel (gravo-mg_b200/csrc/k_galerkin.cu), which processes
 * the products one warp-wide window at a time (step 1: 4 rows k x 8 entries; step 2: 32 interleaved l, one
 * entry of their rows of P) and adds the partial sum of equal keys inside a window before it touches the
 * accumulator. Following that order makes the comparison bit-exact instead of 1e-12.
 *
 * Build: gcc -std=c11 -O2 -ffp-contract=off (no FMA contraction, no -march).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define WINDOW 32

void orc_solver_free(void* p) { free(p); }

/* Add the window's values into acc[slot[key]] with the kernel's association: for every distinct key of the
 * window, in order of first appearance, partial = v_a + v_b + ... (appearance order), then acc += partial. */
static void window_accumulate(int count, const int32_t* keys, const double* vals, const int32_t* slot, double* acc) {
    unsigned char done[WINDOW];
    memset(done, 0, sizeof done);
    for (int a = 0; a < count; ++a) {
        if (done[a]) continue;
        double partial = vals[a];
        for (int b = a + 1; b < count; ++b)
            if (!done[b] && keys[b] == keys[a]) { partial = partial + vals[b]; done[b] = 1; }
        acc[slot[keys[a]]] = acc[slot[keys[a]]] + partial;
    }
}

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
    /* sparse accumulators: slot maps of the fine and the coarse index space */
    int32_t* fslot = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n ? n : 1));
    int32_t* cslot = (int32_t*)malloc(sizeof(int32_t) * (size_t)(m ? m : 1));
    for (int64_t i = 0; i < n; ++i) fslot[i] = -1;
    for (int64_t i = 0; i < m; ++i) cslot[i] = -1;
    size_t fcap = 1024, ccap = 1024, ocap = (size_t)(8 * m + 16), onnz = 0;
    int32_t* flist = (int32_t*)malloc(sizeof(int32_t) * fcap);
    double* fval = (double*)malloc(sizeof(double) * fcap);
    int32_t* clist = (int32_t*)malloc(sizeof(int32_t) * ccap);
    double* cval = (double*)malloc(sizeof(double) * ccap);
    int32_t* oi = (int32_t*)malloc(sizeof(int32_t) * ocap);
    double* ov = (double*)malloc(sizeof(double) * ocap);
    int32_t wkey[WINDOW];
    double wval[WINDOW];
    out_ptr[0] = 0;
    for (int64_t i = 0; i < m; ++i) {
        /* step 1, keys: distinct fine indices l, ascending */
        size_t kf = 0;
        for (int32_t e = r_ptr[i]; e < r_ptr[i + 1]; ++e) {
            const int32_t k = r_col[e];
            for (int32_t q = a_ptr[k]; q < a_ptr[k + 1]; ++q) {
                const int32_t l = a_idx[q];
                if (fslot[l] >= 0) continue;
                if (kf == fcap) { fcap *= 2; flist = (int32_t*)realloc(flist, sizeof(int32_t) * fcap); fval = (double*)realloc(fval, sizeof(double) * fcap); }
                fslot[l] = 0; flist[kf++] = l;
            }
        }
        qsort(flist, kf, sizeof(int32_t), cmp_i32);
        for (size_t f = 0; f < kf; ++f) { fslot[flist[f]] = (int32_t)f; fval[f] = 0.0; }
        /* step 1, values: a window holds 4 rows k of R (ascending) times entries [q0, q0 + 8) of A's outer vectors
         * k, ordered (k, q) */
        int w = 0;
        for (int32_t e0 = r_ptr[i]; e0 < r_ptr[i + 1]; e0 += 4) {
            const int32_t e1 = e0 + 4 < r_ptr[i + 1] ? e0 + 4 : r_ptr[i + 1];
            int32_t longest = 0;
            for (int32_t e = e0; e < e1; ++e) { const int32_t len = a_ptr[r_col[e] + 1] - a_ptr[r_col[e]]; if (len > longest) longest = len; }
            for (int32_t q0 = 0; q0 < longest; q0 += 8) {
                w = 0;
                for (int32_t e = e0; e < e1; ++e) {
                    const int32_t k = r_col[e];
                    for (int32_t q = a_ptr[k] + q0; q < a_ptr[k + 1] && q < a_ptr[k] + q0 + 8; ++q) {
                        wkey[w] = a_idx[q]; wval[w] = r_val[e] * a_val[q]; ++w;
                    }
                }
                window_accumulate(w, wkey, wval, fslot, fval);
            }
        }
        /* step 2, keys: distinct coarse indices j, ascending */
        size_t kc = 0;
        for (size_t f = 0; f < kf; ++f) {
            const int32_t l = flist[f];
            for (int32_t t = p_ptr[l]; t < p_ptr[l + 1]; ++t) {
                const int32_t j = p_col[t];
                if (cslot[j] >= 0) continue;
                if (kc == ccap) { ccap *= 2; clist = (int32_t*)realloc(clist, sizeof(int32_t) * ccap); cval = (double*)realloc(cval, sizeof(double) * ccap); }
                cslot[j] = 0; clist[kc++] = j;
            }
        }
        qsort(clist, kc, sizeof(int32_t), cmp_i32);
        for (size_t s = 0; s < kc; ++s) { cslot[clist[s]] = (int32_t)s; cval[s] = 0.0; }
        /* step 2, values: the sorted distinct l are dealt out interleaved over 32 lanes (round r takes
         * f = lane * rounds + r); within a round entry t = 0, 1, 2 of every row of P in turn, each turn one window
         * in lane order */
        {
            const size_t rounds = (kf + WINDOW - 1) / WINDOW;
            for (size_t r = 0; r < rounds; ++r) {
                int32_t longest = 0;
                for (size_t lane = 0; lane < WINDOW; ++lane) {
                    const size_t f = lane * rounds + r;
                    if (f >= kf) continue;
                    const int32_t len = p_ptr[flist[f] + 1] - p_ptr[flist[f]];
                    if (len > longest) longest = len;
                }
                for (int32_t t = 0; t < longest; ++t) {
                    w = 0;
                    for (size_t lane = 0; lane < WINDOW; ++lane) {
                        const size_t f = lane * rounds + r;
                        if (f >= kf) continue;
                        const int32_t l = flist[f];
                        if (p_ptr[l] + t < p_ptr[l + 1]) { wkey[w] = p_col[p_ptr[l] + t]; wval[w] = fval[f] * p_val[p_ptr[l] + t]; ++w; }
                    }
                    window_accumulate(w, wkey, wval, cslot, cval);
                }
            }
        }
        if (onnz + kc > ocap) { while (onnz + kc > ocap) ocap *= 2; oi = (int32_t*)realloc(oi, sizeof(int32_t) * ocap); ov = (double*)realloc(ov, sizeof(double) * ocap); }
        for (size_t s = 0; s < kc; ++s) { oi[onnz] = clist[s]; ov[onnz] = cval[s]; ++onnz; cslot[clist[s]] = -1; }
        for (size_t f = 0; f < kf; ++f) fslot[flist[f]] = -1;
        out_ptr[i + 1] = (int32_t)onnz;
    }
    free(r_ptr); free(r_col); free(r_val); free(fslot); free(cslot); free(flist); free(fval); free(clist); free(cval);
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
