// C ABI of SpMV, multicolour Gauss-Seidel and the V-cycle (include/gravomg_b200.h, last section).
#include "hierarchy_types.h"

#include <cmath>
#include <stdexcept>

struct SolverLevel {
    int32_t n = 0;
    const int32_t* outer = nullptr;   // operator of this level (views into input / hierarchy)
    const int32_t* inner = nullptr;
    const double* val = nullptr;
    Colouring col;
    DeviceCsr R;                       // P^T of the prolongation from this level to the next
    const DeviceCsr* P = nullptr;
    DevBuf<double> x, b, r;
};

struct gmg_solver {
    gmg_ctx* ctx = nullptr;
    std::vector<std::unique_ptr<SolverLevel>> levels;   // levels.size() = prolongations + 1
};

namespace {

template <typename F> int guarded_s(gmg_ctx* c, F&& body) {
    if (!c) return GMG_ERR_STATE;
    try {
        GMG_CUDA(cudaSetDevice(c->device));
        body();
        c->last_error.clear();
        return GMG_OK;
    } catch (const Error& e) {
        c->last_error = e.what;
        abort_sync(c);
        return e.status;
    } catch (const std::exception& e) {
        c->last_error = e.what();
        abort_sync(c);
        return GMG_ERR_INVALID;
    }
}

struct Timer {
    gmg_ctx* c;
    explicit Timer(gmg_ctx* ctx) : c(ctx) { cudaEventRecord(c->ev0, c->stream); }
    void stop() {
        cudaEventRecord(c->ev1, c->stream);
        cudaEventSynchronize(c->ev1);
        float ms = 0.f;
        cudaEventElapsedTime(&ms, c->ev0, c->ev1);
        c->last_kernel_ms = ms;
    }
};

// symmetric: the coarsest level sweeps forward then backward, so that with pre == post the cycle is a
// symmetric operator (what conjugate gradients needs of its preconditioner)
void vcycle(gmg_ctx* c, gmg_solver* s, size_t l, int pre, int post, int coarse_sweeps, bool symmetric = false) {
    SolverLevel& L = *s->levels[l];
    if (l + 1 == s->levels.size()) {
        gauss_seidel(c, L.col, L.outer, L.inner, L.val, L.b.p, L.x.p, coarse_sweeps, false);
        if (symmetric) gauss_seidel(c, L.col, L.outer, L.inner, L.val, L.b.p, L.x.p, coarse_sweeps, true);
        return;
    }
    SolverLevel& C = *s->levels[l + 1];
    gauss_seidel(c, L.col, L.outer, L.inner, L.val, L.b.p, L.x.p, pre, false);
    spmv(c, L.n, L.outer, L.inner, L.val, L.x.p, L.r.p, -1.0, L.b.p);                     // r = b - A x
    spmv(c, C.n, L.R.rowptr.p, L.R.colidx.p, L.R.val.p, L.r.p, C.b.p, 1.0, nullptr);      // b_c = P^T r
    C.x.zero();
    vcycle(c, s, l + 1, pre, post, coarse_sweeps, symmetric);
    spmv(c, L.n, L.P->rowptr.p, L.P->colidx.p, L.P->val.p, C.x.p, L.x.p, 1.0, L.x.p);     // x += P e
    gauss_seidel(c, L.col, L.outer, L.inner, L.val, L.b.p, L.x.p, post, true);
}

}  // namespace

extern "C" {

int gmg_spmv(gmg_ctx* c, int64_t n, int64_t n_inner, const int32_t* outer, const int32_t* inner, const double* val,
             const double* x, double* y) {
    return guarded_s(c, [&] {
        GMG_REQUIRE(n >= 0 && outer && outer[0] == 0, GMG_ERR_INVALID, "bad matrix");
        const int64_t nnz = outer[n];
        auto d_o = to_device(c, outer, n + 1); auto d_i = to_device(c, inner, nnz); auto d_v = to_device(c, val, nnz);
        auto d_x = to_device(c, x, n_inner);
        check_outer_pointers(c, n, d_o.p, "spmv");
        check_index_range(c, nnz, d_i.p, n_inner, "spmv");
        DevBuf<double> d_y(c, n);
        Timer t(c);
        spmv(c, int32_t(n), d_o.p, d_i.p, d_v.p, d_x.p, d_y.p, 1.0, nullptr);
        t.stop();
        d_y.download(y, n);
        sync(c);
    });
}

int gmg_gauss_seidel(gmg_ctx* c, int64_t n, const int32_t* outer, const int32_t* inner, const double* val,
                     const double* b, double* x, int sweeps, int32_t* colours_out, int32_t* num_colours_out) {
    return guarded_s(c, [&] {
        GMG_REQUIRE(n >= 0 && outer && outer[0] == 0, GMG_ERR_INVALID, "bad matrix");
        const int64_t nnz = outer[n];
        auto d_o = to_device(c, outer, n + 1); auto d_i = to_device(c, inner, nnz); auto d_v = to_device(c, val, nnz);
        auto d_b = to_device(c, b, n); auto d_x = to_device(c, x, n);
        check_outer_pointers(c, n, d_o.p, "gaussSeidel");
        check_index_range(c, nnz, d_i.p, n, "gaussSeidel");
        Colouring col;
        colour_graph(c, int32_t(n), d_o.p, d_i.p, col);
        Timer t(c);
        gauss_seidel(c, col, d_o.p, d_i.p, d_v.p, d_b.p, d_x.p, sweeps, false);
        t.stop();
        d_x.download(x, n);
        if (colours_out) col.colour.download(colours_out, n);
        if (num_colours_out) *num_colours_out = col.num_colours;
        sync(c);
    });
}

int gmg_solver_create(gmg_ctx* c, const gmg_input* in, const gmg_hierarchy* h, gmg_solver** out) {
    if (!out) return GMG_ERR_INVALID;
    *out = nullptr;
    return guarded_s(c, [&] {
        GMG_REQUIRE(in && h && in->has_a, GMG_ERR_INVALID, "solver needs an input with A values");
        await_values(c, in);
        std::unique_ptr<gmg_solver> s(new gmg_solver());
        s->ctx = c;
        const size_t nl = h->levels.size();
        for (size_t l = 0; l <= nl; ++l) {
            std::unique_ptr<SolverLevel> L(new SolverLevel());
            if (l == 0) { L->n = in->n; L->outer = in->colptr.p; L->inner = in->rowidx.p; L->val = in->a_val.p; }
            else {
                const Level& prev = *h->levels[l - 1];
                GMG_REQUIRE(prev.Ac.colptr.p != nullptr, GMG_ERR_STATE, "hierarchy was built without the Galerkin operator");
                L->n = prev.Ac.n; L->outer = prev.Ac.colptr.p; L->inner = prev.Ac.rowidx.p; L->val = prev.Ac.val.p;
            }
            colour_graph(c, L->n, L->outer, L->inner, L->col);
            L->x.alloc(c, L->n); L->b.alloc(c, L->n); L->r.alloc(c, L->n);
            if (l < nl) { L->P = &h->levels[l]->P; transpose_csr(c, *L->P, L->R); }
            s->levels.push_back(std::move(L));
        }
        sync(c);
        *out = s.release();
    });
}

void gmg_solver_destroy(gmg_solver* s) {
    if (!s) return;
    cudaSetDevice(s->ctx->device);
    delete s;
}

int gmg_solver_vcycle(gmg_ctx* c, gmg_solver* s, const double* b, double* x, int cycles, int pre, int post,
                      int coarse_sweeps, double* residual_out) {
    return guarded_s(c, [&] {
        GMG_REQUIRE(s && b && x && cycles >= 0, GMG_ERR_INVALID, "bad arguments");
        SolverLevel& F = *s->levels[0];
        F.b.upload(b, F.n);
        F.x.upload(x, F.n);
        auto resnorm = [&] {
            spmv(c, F.n, F.outer, F.inner, F.val, F.x.p, F.r.p, -1.0, F.b.p);
            return norm2(c, F.n, F.r.p);
        };
        if (residual_out) residual_out[0] = resnorm();
        double total_ms = 0.0;
        for (int k = 0; k < cycles; ++k) {
            Timer t(c);
            vcycle(c, s, 0, pre, post, coarse_sweeps);
            t.stop();
            total_ms += c->last_kernel_ms;
            if (residual_out) residual_out[k + 1] = resnorm();
        }
        c->last_kernel_ms = total_ms;
        F.x.download(x, F.n);
        sync(c);
    });
}

/* Conjugate gradients on A x = b preconditioned by one symmetric V(sweeps, sweeps) cycle from a zero guess.
 * Stops when ||b - A x||_2 <= rel_tol * ||b||_2 or after max_iter iterations. residual_out (optional,
 * max_iter + 1 entries) receives the residual norm before the first and after every iteration. */
int gmg_solver_pcg(gmg_ctx* c, gmg_solver* s, const double* b, double* x, double rel_tol, int max_iter, int sweeps,
                   int coarse_sweeps, int* iterations_out, double* residual_out) {
    return guarded_s(c, [&] {
        GMG_REQUIRE(s && b && x && max_iter >= 0 && sweeps >= 1, GMG_ERR_INVALID, "bad arguments");
        SolverLevel& F = *s->levels[0];
        const int32_t n = F.n;
        DevBuf<double> sol(c, n), res(c, n), p(c, n), Ap(c, n), rhs(c, n);
        rhs.upload(b, n);
        sol.upload(x, n);
        spmv(c, n, F.outer, F.inner, F.val, sol.p, res.p, -1.0, rhs.p);            // r = b - A x
        const double bnorm = norm2(c, n, rhs.p);
        double rnorm = norm2(c, n, res.p);
        if (residual_out) residual_out[0] = rnorm;
        auto precondition = [&] {                                                   // F.x = M^-1 res
            GMG_CUDA(cudaMemcpyAsync(F.b.p, res.p, sizeof(double) * size_t(n), cudaMemcpyDeviceToDevice, c->stream));
            F.x.zero();
            vcycle(c, s, 0, sweeps, sweeps, coarse_sweeps, true);
        };
        Timer t(c);
        int it = 0;
        double rz = 0.0;
        if (rnorm > rel_tol * bnorm) {
            precondition();
            GMG_CUDA(cudaMemcpyAsync(p.p, F.x.p, sizeof(double) * size_t(n), cudaMemcpyDeviceToDevice, c->stream));
            rz = dot(c, n, res.p, F.x.p);
        }
        while (it < max_iter && rnorm > rel_tol * bnorm) {
            spmv(c, n, F.outer, F.inner, F.val, p.p, Ap.p, 1.0, nullptr);
            const double pAp = dot(c, n, p.p, Ap.p);
            GMG_REQUIRE(pAp > 0.0, GMG_ERR_INVALID, "pcg: the operator is not positive definite along the search direction");
            const double alpha = rz / pAp;
            axpy(c, n, alpha, p.p, sol.p);
            axpy(c, n, -alpha, Ap.p, res.p);
            rnorm = norm2(c, n, res.p);
            ++it;
            if (residual_out) residual_out[it] = rnorm;
            if (rnorm <= rel_tol * bnorm) break;
            precondition();
            const double rz_new = dot(c, n, res.p, F.x.p);
            xpby(c, n, F.x.p, rz_new / rz, p.p);                                    // p = z + beta p
            rz = rz_new;
        }
        t.stop();
        if (iterations_out) *iterations_out = it;
        sol.download(x, n);
        sync(c);
    });
}

}  // extern "C"
