/* Study (test infrastructure, CPU only): how the residual correlation c may be formed without changing the
 * reference's behaviour on ill-conditioned dictionaries.
 *
 * The reference re-forms  c = A^T (y - A x)  every iteration (src/solvers/homotopy-cpu.cpp:87-98, called at :255).
 * The restatement under oracle/ does the same; here its loop is repeated with the alternatives a GPU kernel could use,
 * on the reference's own benchmark workload (src/solvers/homotopy_bench.cpp:14-60), to see which of them keep its
 * iteration counts and which stall ("|c|_inf freezes above the tolerance while atoms keep joining").
 *
 *   mode 0  reference form:  r = y - A x (all non-zero x),   c = A^T r                        (= the oracle)
 *   mode 1  incremental c:   c -= gamma q                                                     (round 1 product)
 *   mode 2  incremental r:   r -= gamma p,                   c = A^T r
 *   mode 3  active-set r:    r = y - A_G x_G (from scratch), c = A^T r
 *   mode 4  stacked:         c' = A^T r (r as mode 3) is what the step search and lambda use; the direction signs, the
 *                            stop test and the report use  c' - gamma q  (one step from a fresh base, never accumulated)
 *   mode 5  Gram form:       c = A^T y - G_G x_G,  q = G_G d  with G = fl(A^T A)
 *   mode 6  mixed:           mode 5 while iter < switch_iter, mode 4 from then on      (what the lockstep path does)
 *   mode 8  mode 0 with the x update FUSED, x = fma(gamma, d, x): what a GPU compiler makes of `x += gamma * d` by default.  The
 *           reference's build rounds the product and the sum separately, which leaves a LEAVING atom's coefficient at exactly
 *           zero most of the time; fused, the division's rounding error stays behind, and when its sign is "wrong" the atom
 *           is thrown out again the moment it rejoins, stays outside the active set with the largest |c| -- and the solve stalls
 *   mode 7  mode 4 with the step candidates ORDERED by an approximate quotient num * (1/den) (two roundings, like the GPU's
 *           reciprocal-based __fdividef) and the winning step then recomputed with IEEE division
 *
 * Build: gcc -O2 -fopenmp -o tests/studies/_build/residual_modes tests/studies/residual_modes.c -lm
 * Driver: tests/studies/residual_modes.py
 */
#include <math.h>
#include <float.h>
#include <stddef.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define T float
#define FN(name) name##_f32
#define T_EPS FLT_EPSILON
#define T_MAX FLT_MAX
#include "../../oracle/ss_oracle_impl.h"

#define A_(i, j) A[(size_t)(i) * n + (j)]

static void correlate(const float* A, size_t m, size_t n, const float* v, float* out)
{
    for (size_t j = 0; j < n; ++j) out[j] = 0;
    /* column sums in row order: same per-element order as the oracle's loop */
    for (size_t j = 0; j < n; ++j) {
        float acc = 0;
        for (size_t i = 0; i < m; ++i) acc += A_(i, j) * v[i];
        out[j] = acc;
    }
}

static void active_residual(const float* A, size_t m, size_t n, const float* y, const float* x, const rank_index_f32* lam, float* r)
{
    for (size_t i = 0; i < m; ++i) {
        float acc = 0;
        for (size_t k = 0; k < lam->n; ++k) { size_t j = lam->v[k]; acc += A_(i, j) * x[j]; }
        r[i] = y[i] - acc;
    }
}

/* returns iterations; *err = final |c|_inf */
uint32_t solve_mode(int mode, int switch_iter, const float* A, const float* G, size_t m, size_t n, const float* y, float tol,
                    uint32_t max_iter, float* x, double* err)
{
    float* direction = calloc(n, sizeof(float));
    float* c = calloc(n, sizeof(float));
    float* c0 = calloc(n, sizeof(float));
    float* cs = calloc(n, sizeof(float));     /* the vector the step search reads */
    float* c_gamma = calloc(n, sizeof(float));
    float* p = calloc(m, sizeof(float));
    float* q = calloc(n, sizeof(float));
    float* r = calloc(m, sizeof(float));
    float* colbuf = calloc(m, sizeof(float));
    rank_index_f32 lam = { 0, 0, 0 };
    oinv_f32 inv = { 0, 0, m, 0, 0 };
    float c_inf = 0;
    const ptrdiff_t rs = (ptrdiff_t)n, cst = 1;

    for (size_t j = 0; j < n; ++j) x[j] = 0;
    memcpy(r, y, m * sizeof(float));
    correlate(A, m, n, r, c);
    memcpy(c0, c, n * sizeof(float));
    {
        size_t idx;
        c_inf = inf_norm_f32(c, n, &idx);
        inverse_add_or_remove_f32(A, m, rs, cst, idx, &lam, &inv, colbuf);
        direction[0] = sign_tol_f32(c_inf, tol) * inv.inv[0];
        expand_f32(direction, n, &lam);
    }
    uint32_t iter = 0;
    do {
        iter++;
        const int md = mode == 6 ? ((int)iter < switch_iter ? 5 : 4) : mode == 7 ? 4 : mode == 8 ? 0 : mode;
        float min; size_t idx;
        const float* csearch = c;
        float lam_search = c_inf;
        if (md == 4) {
            /* fresh base for this iteration: r from the active set (x and G as they stand before the step) */
            active_residual(A, m, n, y, x, &lam, r);
            correlate(A, m, n, r, cs);
            csearch = cs;
            lam_search = inf_norm_f32(cs, n, 0);
        }
        if (md == 5) {
            /* q = G_G d ; p is not needed */
            for (size_t j = 0; j < n; ++j) {
                float acc = 0;
                for (size_t k = 0; k < lam.n; ++k) { size_t g = lam.v[k]; acc += G[g * n + j] * direction[g]; }
                q[j] = acc;
            }
            /* the search itself, as find_max_gamma part 2 */
            float mn = FLT_MAX; size_t ix = 0, l = 0;
            for (size_t i = 0; i < n; ++i) {
                const float prev = mn;
                if (l < lam.n && lam.v[l] == i) { float t = -x[i] / direction[i]; if (t > 0.0 && t < mn) mn = t; l++; }
                else {
                    float dl = 1.0f - q[i], dr = 1.0f + q[i];
                    if (dl != 0.0) { float t = (c_inf - c[i]) / dl; if (t > 0.0 && t < mn) mn = t; }
                    if (dr != 0.0) { float t = (c_inf + c[i]) / dr; if (t > 0.0 && t < mn) mn = t; }
                }
                if (prev > mn) ix = i;
            }
            min = mn; idx = ix;
        } else if (mode == 7) {
            /* p, q as find_max_gamma; candidates ordered by num * (1 / den), the winner's step in IEEE arithmetic */
            for (size_t i = 0; i < m; ++i) {
                float acc = 0;
                for (size_t k = 0; k < lam.n; ++k) { size_t j = lam.v[k]; acc += A_(i, j) * direction[j]; }
                p[i] = acc;
            }
            correlate(A, m, n, p, q);
            float mn = FLT_MAX; size_t ix = 0, l = 0; int join = 0;
            for (size_t i = 0; i < n; ++i) {
                const float prev = mn;
                if (l < lam.n && lam.v[l] == i) { float t = -x[i] / direction[i]; if (t > 0.0 && t < mn) { mn = t; join = 0; } l++; }
                else {
                    float dl = 1.0f - q[i], dr = 1.0f + q[i];
                    volatile float rl = 1.0f / dl, rr = 1.0f / dr;
                    if (dl != 0.0) { float t = (lam_search - csearch[i]) * rl; if (t > 0.0 && t < mn) { mn = t; join = 1; } }
                    if (dr != 0.0) { float t = (lam_search + csearch[i]) * rr; if (t > 0.0 && t < mn) { mn = t; join = 1; } }
                }
                if (prev > mn) ix = i;
            }
            if (join && mn < FLT_MAX) {
                float best = FLT_MAX, dl = 1.0f - q[ix], dr = 1.0f + q[ix];
                if (dl != 0.0) { float t = (lam_search - csearch[ix]) / dl; if (t > 0.0 && t < best) best = t; }
                if (dr != 0.0) { float t = (lam_search + csearch[ix]) / dr; if (t > 0.0 && t < best) best = t; }
                if (best < FLT_MAX) mn = best;
            }
            min = mn; idx = ix;
        } else {
            /* find_max_gamma with the chosen search vector (it forms p and q itself) */
            find_max_gamma_f32(A, m, n, rs, cst, csearch, x, direction, lam_search, &lam, p, q, &min, &idx);
        }
        inverse_add_or_remove_f32(A, m, rs, cst, idx, &lam, &inv, colbuf);
        size_t K = lam.n;
        if (K == 0) break;
        if (mode == 8) for (size_t j = 0; j < n; ++j) x[j] = fmaf(min, direction[j], x[j]);
        else for (size_t j = 0; j < n; ++j) x[j] += min * direction[j];
        switch (md) {
        case 0: residual_vector_f32(A, m, n, rs, cst, y, x, c, r); break;
        case 1: for (size_t j = 0; j < n; ++j) c[j] = c[j] - min * q[j]; break;
        case 2: for (size_t i = 0; i < m; ++i) r[i] = r[i] - min * p[i]; correlate(A, m, n, r, c); break;
        case 3: active_residual(A, m, n, y, x, &lam, r); correlate(A, m, n, r, c); break;
        case 4: for (size_t j = 0; j < n; ++j) c[j] = cs[j] - min * q[j]; break;
        case 5:
            for (size_t j = 0; j < n; ++j) {
                float acc = 0;
                for (size_t k = 0; k < lam.n; ++k) { size_t g = lam.v[k]; acc += G[g * n + j] * x[g]; }
                c[j] = c0[j] - acc;
            }
            break;
        }
        for (size_t k = 0; k < K; ++k) c_gamma[k] = sign_tol_f32(c[lam.v[k]], tol);
        for (size_t a = 0; a < K; ++a) {
            float acc = 0;
            for (size_t b = 0; b < K; ++b) acc += inv.inv[a * K + b] * c_gamma[b];
            direction[a] = acc;
        }
        expand_f32(direction, n, &lam);
        c_inf = inf_norm_f32(c, n, 0);
    } while (iter < max_iter && c_inf > tol);
    *err = (double)c_inf;
    free(direction); free(c); free(c0); free(cs); free(c_gamma); free(p); free(q); free(r); free(colbuf);
    free(lam.v); free(inv.At); free(inv.inv);
    return iter;
}

/* batch entry for the Python driver: Y [batch][m], X [batch][n] */
void solve_modes(int mode, int switch_iter, const float* A, const float* G, size_t m, size_t n, const float* Y, size_t batch,
                 float tol, uint32_t max_iter, float* X, uint32_t* iters, double* errs)
{
#pragma omp parallel for schedule(dynamic)
    for (long b = 0; b < (long)batch; ++b)
        iters[b] = solve_mode(mode, switch_iter, A, G, m, n, Y + (size_t)b * m, tol, max_iter, X + (size_t)b * n, errs + b);
}

/* G = fl(A^T A) in float32, row order */
void gram(const float* A, size_t m, size_t n, float* G)
{
#pragma omp parallel for
    for (long a = 0; a < (long)n; ++a)
        for (size_t b = 0; b < n; ++b) {
            float acc = 0;
            for (size_t i = 0; i < m; ++i) acc += A_((size_t)i, (size_t)a) * A_((size_t)i, b);
            G[(size_t)a * n + b] = acc;
        }
}

/* One cell of the reference bench (homotopy_bench.cpp:38-48): before solve i the pattern (+1 on every second row) is added
 * to column i % n of the haystack and taken away afterwards.  Every solve works on its own copy. */
void grid_cell(int mode, int switch_iter, const float* hay, const float* sig, size_t m, size_t n, size_t solves, float tol,
               uint32_t max_iter, uint32_t* iters, double* errs)
{
    float* G0 = NULL;
    if (mode >= 5) { G0 = malloc(n * n * sizeof(float)); gram(hay, m, n, G0); }
#pragma omp parallel for schedule(dynamic)
    for (long s = 0; s < (long)solves; ++s) {
        float* A = malloc(m * n * sizeof(float));
        float* x = malloc(n * sizeof(float));
        float* G = NULL;
        memcpy(A, hay, m * n * sizeof(float));
        const size_t col = (size_t)s % n;
        for (size_t i = 0; i < m; i += 2) A[i * n + col] += 1.0f;
        if (mode >= 5) {
            G = malloc(n * n * sizeof(float));
            memcpy(G, G0, n * n * sizeof(float));
            for (size_t b = 0; b < n; ++b) {
                float acc = 0;
                for (size_t i = 0; i < m; ++i) acc += A[i * n + col] * A[i * n + b];
                G[col * n + b] = acc; G[b * n + col] = acc;
            }
        }
        iters[s] = solve_mode(mode, switch_iter, A, G, m, n, sig, tol, max_iter, x, errs + s);
        free(A); free(x); free(G);
    }
    free(G0);
}
