// Procedural level-0 inputs generated on the device [not in reference; SURVEY.md 8(f) rank 2: the step
// before the hot path]. At 10^8 vertices building the mesh with numpy on every rank would dwarf the
// build being measured, so the torus of configs C3/C5 has a closed-form generator:
// vertex (i,j) = i*nv + j; every quad (i,j),(i+1,j),(i+1,j+1),(i,j+1) is split on the (i,j)-(i+1,j+1)
// diagonal, so a vertex has the six neighbours (i+-1,j), (i,j+-1), (i+1,j+1), (i-1,j-1) plus the
// diagonal entry; each position is displaced by amp*h*(u0,u1,u2) with the counter-based splitmix64
// noise of gravo-mg_b200/meshgen.py (same keys, same bits; only sin/cos may differ in the last ulp
// from numpy's, which is why parity tests feed both sides the same host arrays instead).
#include "hierarchy_types.h"

#include <cmath>

using namespace gmg;

namespace {

__device__ __forceinline__ unsigned long long splitmix64(unsigned long long x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
__device__ __forceinline__ double uniform_pm1(unsigned long long seed, unsigned long long index, unsigned component) {
    const unsigned long long key = index * 4ull + component + seed * 0x100000001B3ull;
    const unsigned long long bits = splitmix64(splitmix64(key));
    return double(bits >> 11) * (2.0 / 9007199254740992.0) - 1.0;
}

__global__ void __launch_bounds__(kBlock) torus_kernel(int32_t nu, int32_t nv, double R, double r, double amp_h,
                                                       unsigned long long seed, double* __restrict__ pts,
                                                       int32_t* __restrict__ colptr, int32_t* __restrict__ rowidx) {
    const int64_t n = int64_t(nu) * nv;
    const int64_t idx = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (idx > n) return;
    if (idx <= n) colptr[idx] = int32_t(7 * idx);
    if (idx == n) return;
    const int32_t i = int32_t(idx / nv), j = int32_t(idx % nv);
    const double kTwoPi = 6.283185307179586476925286766559;
    const double u = kTwoPi * double(i) / double(nu), v = kTwoPi * double(j) / double(nv);
    double su, cu, sv, cv;
    sincos(u, &su, &cu);
    sincos(v, &sv, &cv);
    pts[idx] = (R + r * cv) * cu + amp_h * uniform_pm1(seed, (unsigned long long)idx, 0);
    pts[n + idx] = (R + r * cv) * su + amp_h * uniform_pm1(seed, (unsigned long long)idx, 1);
    pts[2 * n + idx] = r * sv + amp_h * uniform_pm1(seed, (unsigned long long)idx, 2);
    const int32_t ip = (i + 1) % nu, im = (i + nu - 1) % nu, jp = (j + 1) % nv, jm = (j + nv - 1) % nv;
    int32_t e[7] = {int32_t(idx), ip * nv + j, im * nv + j, i * nv + jp, i * nv + jm, ip * nv + jp, im * nv + jm};
#pragma unroll
    for (int a = 1; a < 7; ++a) {                    // insertion sort of seven registers
#pragma unroll
        for (int b = a; b > 0; --b) {
            if (e[b] < e[b - 1]) { const int32_t t = e[b]; e[b] = e[b - 1]; e[b - 1] = t; }
        }
    }
#pragma unroll
    for (int a = 0; a < 7; ++a) rowidx[7 * idx + a] = e[a];
}

// Cotangent Laplacian + mass_shift * lumped mass of the torus grid, on the pattern torus_kernel wrote
// (the system matrix of configs C3/C5; same formula as gravo-mg_b200/meshgen.py cotangent_laplacian).
// The neighbours of v in cyclic order are n0=(i+1,j), n1=(i+1,j+1), n2=(i,j+1), n3=(i-1,j), n4=(i-1,j-1),
// n5=(i,j-1); the six triangles around v are (v, n_k, n_k+1), so edge (v, n_k) is opposite n_k-1 and n_k+1.
__global__ void __launch_bounds__(kBlock) torus_operator_kernel(int32_t nu, int32_t nv, double mass_shift,
                                                                const double* __restrict__ pts, double* __restrict__ a_val) {
    const int64_t n = int64_t(nu) * nv;
    const int64_t idx = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (idx >= n) return;
    const int32_t i = int32_t(idx / nv), j = int32_t(idx % nv);
    const int32_t ip = (i + 1) % nu, im = (i + nu - 1) % nu, jp = (j + 1) % nv, jm = (j + nv - 1) % nv;
    const int32_t nb[6] = {ip * nv + j, ip * nv + jp, i * nv + jp, im * nv + j, im * nv + jm, i * nv + jm};
    const V3 pv = ld3(pts, n, int32_t(idx));
    V3 q[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) q[k] = ld3(pts, n, nb[k]);
    double w[6] = {0, 0, 0, 0, 0, 0};
    double area = 0.0;
#pragma unroll
    for (int k = 0; k < 6; ++k) {
        const int k1 = (k + 1) % 6;
        // triangle (v, n_k, n_k1): the angle at n_k1 is opposite edge (v, n_k), the angle at n_k opposite (v, n_k1)
        const V3 a1 = pv - q[k1], b1 = q[k] - q[k1];
        const V3 a0 = pv - q[k], b0 = q[k1] - q[k];
        const double c1 = fmax(norm(cross(a1, b1)), 1e-300), c0 = fmax(norm(cross(a0, b0)), 1e-300);
        w[k] += 0.5 * (dot(a1, b1) / c1);
        w[k1] += 0.5 * (dot(a0, b0) / c0);
        area += norm(cross(q[k] - pv, q[k1] - pv)) / 6.0;
    }
    double diag = mass_shift * area;
#pragma unroll
    for (int k = 0; k < 6; ++k) diag += w[k];
    int32_t e[7] = {int32_t(idx), nb[0], nb[1], nb[2], nb[3], nb[4], nb[5]};
    double v[7] = {diag, -w[0], -w[1], -w[2], -w[3], -w[4], -w[5]};
#pragma unroll
    for (int a = 1; a < 7; ++a) {
#pragma unroll
        for (int b = a; b > 0; --b) {
            if (e[b] < e[b - 1]) {
                const int32_t t = e[b]; e[b] = e[b - 1]; e[b - 1] = t;
                const double tv = v[b]; v[b] = v[b - 1]; v[b - 1] = tv;
            }
        }
    }
#pragma unroll
    for (int a = 0; a < 7; ++a) a_val[7 * idx + a] = v[a];
}

template <typename F> int guarded_g(gmg_ctx* c, F&& body) {
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

}  // namespace

extern "C" {

int gmg_generate_torus(gmg_ctx* c, int32_t nu, int32_t nv, double R, double r, double amplitude, uint64_t seed,
                       gmg_input** out) {
    if (!out) return GMG_ERR_INVALID;
    *out = nullptr;
    return guarded_g(c, [&] {
        GMG_REQUIRE(nu >= 3 && nv >= 3, GMG_ERR_INVALID, "torus grid needs at least 3 x 3 vertices (distinct neighbours)");
        const int64_t n = int64_t(nu) * nv;
        GMG_REQUIRE(7 * n < (int64_t(1) << 31), GMG_ERR_INVALID, "torus too large for 32-bit entry offsets");
        std::unique_ptr<gmg_input> in(new gmg_input());
        in->ctx = c; in->n = int32_t(n); in->nnz = 7 * n;
        in->pts.alloc(c, 3 * n); in->colptr.alloc(c, n + 1); in->rowidx.alloc(c, 7 * n);
        const double kTwoPi = 6.283185307179586476925286766559;
        const double h = std::fmin(kTwoPi * (R + r) / nu, kTwoPi * r / nv);   // the shorter grid step
        GMG_LAUNCH(c, torus_kernel, div_up(n + 1, kBlock), kBlock, 0, nu, nv, R, r, amplitude * h, (unsigned long long)seed,
                   in->pts.p, in->colptr.p, in->rowidx.p);
        sync(c);
        *out = in.release();
    });
}

/* System matrix of a torus made by gmg_generate_torus (same nu, nv): cotangent Laplacian + mass_shift * lumped mass */
int gmg_input_torus_operator(gmg_ctx* c, gmg_input* in, int32_t nu, int32_t nv, double mass_shift) {
    return guarded_g(c, [&] {
        GMG_REQUIRE(in && int64_t(nu) * nv == in->n && in->nnz == 7 * int64_t(in->n), GMG_ERR_INVALID, "not a torus input of these dimensions");
        in->a_val.alloc(c, in->nnz);
        GMG_LAUNCH(c, torus_operator_kernel, div_up(in->n, kBlock), kBlock, 0, nu, nv, mass_shift, in->pts.p, in->a_val.p);
        in->has_a = true;
        sync(c);
    });
}

/* which: 0 points (f64[3n], column-major n x 3), 1 colptr (int32[n+1]), 2 rowidx (int32[nnz]), 3 A values (f64[nnz]) */
int gmg_input_fetch(gmg_ctx* c, const gmg_input* in, int which, void* host_out) {
    return guarded_g(c, [&] {
        GMG_REQUIRE(in && host_out, GMG_ERR_STATE, "null input");
        GMG_REQUIRE(which >= 0 && which <= 3, GMG_ERR_INVALID, "unknown input array");
        GMG_REQUIRE(which != 3 || in->has_a, GMG_ERR_STATE, "input has no A values");
        if (which == 3) await_values(c, in);
        const void* src = which == 0 ? (const void*)in->pts.p : which == 1 ? (const void*)in->colptr.p
                        : which == 2 ? (const void*)in->rowidx.p : (const void*)in->a_val.p;
        const size_t bytes = which == 0 ? sizeof(double) * 3 * size_t(in->n)
                           : which == 1 ? sizeof(int32_t) * (size_t(in->n) + 1)
                           : which == 2 ? sizeof(int32_t) * size_t(in->nnz) : sizeof(double) * size_t(in->nnz);
        GMG_CUDA(cudaMemcpyAsync(host_out, src, bytes, cudaMemcpyDeviceToHost, c->stream));
        sync(c);
    });
}

int gmg_input_size(const gmg_input* in, int64_t* n, int64_t* nnz) {
    if (!in || !n || !nnz) return GMG_ERR_STATE;
    *n = in->n; *nnz = in->nnz;
    return GMG_OK;
}

}  // extern "C"
