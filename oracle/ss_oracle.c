/* CPU restatement of the reference's l1 hot path (ss::homotopy / ss::irls solve
 * loops and the linear-algebra helpers they call), in plain C.
 *
 * THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may load it, and only as the checker or the timed CPU baseline.  Nothing
 * under sparse-solvers_b200/ links, loads or calls it; the product path fails
 * loudly when the CUDA library is missing.
 *
 * Parity is PINNED: tests/test_oracle.py checks this restatement against
 *   (i)  outputs of the unmodified reference itself (oracle/_ref, built from
 *        /root/reference by oracle/ref_build/Makefile) on seeded inputs, and
 *   (ii) the committed fixtures tests/golden/ (.npz files) generated from that same
 *        reference build by tests/golden/make_golden.py, and
 *   (iii) the reference's own known-answer tests for this path
 *        (online_inverse_test.cpp, rank_index_test.cpp,
 *         cholesky_decomposition_test.cpp, qr_decomposition_test.cpp,
 *         test_util.h smoke tests, bindings/python/reference/main.py).
 *
 * Arithmetic differs from the reference only where the reference calls
 * OpenBLAS (summation order inside dot/gemv/ger); see DESIGN.md "noise floor".
 *
 * Build: gcc -O2 -fopenmp -fPIC -shared -o oracle/_build/libss_oracle.so oracle/ss_oracle.c -lm
 */
#include <math.h>
#include <float.h>
#include <stddef.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define T float
#define FN(name) name##_f32
#define T_EPS FLT_EPSILON
#define T_MAX FLT_MAX
#include "ss_oracle_impl.h"
#undef T
#undef FN
#undef T_EPS
#undef T_MAX

#define T double
#define FN(name) name##_f64
#define T_EPS DBL_EPSILON
#define T_MAX DBL_MAX
#include "ss_oracle_impl.h"
#undef T
#undef FN
#undef T_EPS
#undef T_MAX

/* --- thin test hooks over the helpers (known-answer tests of the reference) --- */
void ssor_square_permute_f32(float* A, size_t N, size_t src, size_t dest) { square_permute_f32(A, N, src, dest); }

/* inserts/removes a sequence of columns and returns the maintained inverse;
   ops[i] >= 0: insert column ops[i] at its sorted rank; ops[i] < 0: remove column -ops[i]-1 */
#define ONLINE_INVERSE_HOOK(SFX, T_) \
size_t ssor_online_inverse_##SFX(const T_* A, size_t m, size_t n, const int* ops, size_t nops, T_* inv_out, uint32_t* idx_out) \
{ \
    rank_index_##SFX lam = { 0, 0, 0 }; oinv_##SFX inv = { 0, 0, m, 0, 0 }; \
    T_* col = (T_*)malloc(m * sizeof(T_)); \
    for (size_t i = 0; i < nops; ++i) { \
        size_t c = ops[i] >= 0 ? (size_t)ops[i] : (size_t)(-ops[i] - 1); \
        inverse_add_or_remove_##SFX(A, m, (ptrdiff_t)n, 1, c, &lam, &inv, col); \
    } \
    for (size_t i = 0; i < inv.n * inv.n; ++i) inv_out[i] = inv.inv[i]; \
    for (size_t i = 0; i < lam.n; ++i) idx_out[i] = lam.v[i]; \
    size_t k = inv.n; free(col); free(lam.v); free(inv.At); free(inv.inv); return k; \
}
ONLINE_INVERSE_HOOK(f32, float)
ONLINE_INVERSE_HOOK(f64, double)

int ssor_cholesky_f32(const float* G, size_t N, float* L)   { return cholesky_f32(L, G, N); }
int ssor_cholesky_f64(const double* G, size_t N, double* L) { return cholesky_f64(L, G, N); }
void ssor_chol_solve_f32(const float* L, size_t N, float* x)   { chol_solve_f32(L, N, x); }
void ssor_chol_solve_f64(const double* L, size_t N, double* x) { chol_solve_f64(L, N, x); }
