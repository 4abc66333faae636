/* Type-generic body of the CPU restatement; included twice by ss_oracle.c with
 *   #define T float / double,  #define FN(name) name##_f32 / _f64,  #define T_EPS, T_MAX
 *
 * TEST INFRASTRUCTURE ONLY (see ss_oracle.c header).  Every function cites the
 * reference file:line it restates.  Plain loops; no BLAS.
 */

/* ------------------------------------------------------------------------ */
/* sorted active-set index      reference: src/linalg/rank_index.h:53-98     */

typedef struct { uint32_t* v; size_t n, cap; } FN(rank_index);

static int FN(ri_lower_bound)(const FN(rank_index)* r, uint32_t item)
{
    size_t lo = 0, hi = r->n;
    while (lo < hi) { size_t mid = (lo + hi) / 2; if (r->v[mid] < item) lo = mid + 1; else hi = mid; }
    return (int)lo;
}
/* rank_index.h:76-82 */
static int FN(ri_rank_of)(const FN(rank_index)* r, uint32_t item)
{
    int b = FN(ri_lower_bound)(r, item);
    return ((size_t)b == r->n || r->v[b] != item) ? -1 : b;
}
/* rank_index.h:64-74 */
static int FN(ri_insert)(FN(rank_index)* r, uint32_t item)
{
    int b = FN(ri_lower_bound)(r, item);
    if ((size_t)b == r->n || r->v[b] != item) {
        if (r->n == r->cap) { r->cap = r->cap ? 2 * r->cap : 16; r->v = (uint32_t*)realloc(r->v, r->cap * sizeof(uint32_t)); }
        memmove(r->v + b + 1, r->v + b, (r->n - b) * sizeof(uint32_t));
        r->v[b] = item; r->n++;
    }
    return b;
}
/* rank_index.h:89-98: erases the lower-bound element (no equality check) */
static void FN(ri_erase)(FN(rank_index)* r, uint32_t item)
{
    int b = FN(ri_lower_bound)(r, item);
    if ((size_t)b != r->n) { memmove(r->v + b, r->v + b + 1, (r->n - b - 1) * sizeof(uint32_t)); r->n--; }
}

/* ------------------------------------------------------------------------ */
/* online (A_G^T A_G)^-1        reference: src/linalg/online_inverse.h        */

typedef struct {
    T* At;    /* n x m, row-major: the active columns, in rank order   (:56) */
    T* inv;   /* n x n, row-major, leading dimension = n                (:58) */
    size_t m, n, cap;
} FN(oinv);

static void FN(oinv_reserve)(FN(oinv)* o, size_t n)
{
    if (n <= o->cap) return;
    size_t cap = o->cap ? o->cap : 8;
    while (cap < n) cap *= 2;
    o->At  = (T*)realloc(o->At,  cap * o->m * sizeof(T));
    o->inv = (T*)realloc(o->inv, cap * cap * sizeof(T));
    o->cap = cap;
}

/* online_inverse.h:76-117 square_permute: row+column `src` moves to `dest`,
   the ones in between shift by one.  Restated through an index map. */
static void FN(square_permute)(T* A, size_t N, size_t src, size_t dest)
{
    if (N == 1 || src == dest) return;
    size_t* map = (size_t*)malloc(N * sizeof(size_t));
    T* tmp = (T*)malloc(N * N * sizeof(T));
    /* map[new position] = old position */
    for (size_t i = 0; i < N; ++i) map[i] = i;
    if (dest > src) { for (size_t i = src; i < dest; ++i) map[i] = i + 1; map[dest] = src; }
    else            { for (size_t i = src; i > dest; --i) map[i] = i - 1; map[dest] = src; }
    for (size_t a = 0; a < N; ++a)
        for (size_t b = 0; b < N; ++b)
            tmp[a * N + b] = A[map[a] * N + map[b]];
    memcpy(A, tmp, N * N * sizeof(T));
    free(tmp); free(map);
}

/* online_inverse.h:183-251 insert(rank, column) */
static void FN(oinv_insert)(FN(oinv)* o, size_t idx, const T* col)
{
    const size_t m = o->m, n = o->n;
    FN(oinv_reserve)(o, n + 1);

    if (n == 0) {
        /* :193-201  inv = 1 / nrm2(v)^2 */
        memcpy(o->At, col, m * sizeof(T));
        T ss = 0;
        for (size_t i = 0; i < m; ++i) ss += col[i] * col[i];
        T nrm = (T)sqrt((double)ss);
        o->inv[0] = (T)1 / (nrm * nrm);
    }
    else {
        T* u1 = (T*)malloc(n * sizeof(T));
        T* u2 = (T*)malloc(n * sizeof(T));
        /* :217-218  dot = v.v ; u1 = A_G^T v */
        T dot = 0;
        for (size_t i = 0; i < m; ++i) dot += col[i] * col[i];
        for (size_t r = 0; r < n; ++r) {
            T acc = 0; const T* row = o->At + r * m;
            for (size_t i = 0; i < m; ++i) acc += row[i] * col[i];
            u1[r] = acc;
        }
        /* :209,:221  the new row lands at rank idx */
        memmove(o->At + (idx + 1) * m, o->At + idx * m, (n - idx) * m * sizeof(T));
        memcpy(o->At + idx * m, col, m * sizeof(T));

        /* :225  u2 = inv u1 */
        for (size_t r = 0; r < n; ++r) {
            T acc = 0;
            for (size_t k = 0; k < n; ++k) acc += o->inv[r * n + k] * u1[k];
            u2[r] = acc;
        }
        /* :228-229  d = 1/(dot - u1.u2) ; inv += d u2 u2^T */
        T u1u2 = 0;
        for (size_t k = 0; k < n; ++k) u1u2 += u1[k] * u2[k];
        T d = (T)1 / (dot - u1u2);
        for (size_t r = 0; r < n; ++r) {
            T t = d * u2[r];
            for (size_t k = 0; k < n; ++k) o->inv[r * n + k] += t * u2[k];
        }
        /* :232-245  grow to (n+1)^2 with border -d u2 and corner d */
        T* ni = (T*)malloc((n + 1) * (n + 1) * sizeof(T));
        for (size_t r = 0; r < n; ++r) {
            for (size_t k = 0; k < n; ++k) ni[r * (n + 1) + k] = o->inv[r * n + k];
            T u3 = -d * u2[r];
            ni[r * (n + 1) + n] = u3;
            ni[n * (n + 1) + r] = u3;
        }
        ni[n * (n + 1) + n] = d;
        /* :248  last row/col -> rank idx */
        FN(square_permute)(ni, n + 1, n, idx);
        memcpy(o->inv, ni, (n + 1) * (n + 1) * sizeof(T));
        free(ni); free(u1); free(u2);
    }
    o->n++;
}

/* online_inverse.h:253-293 remove(rank) */
static void FN(oinv_remove)(FN(oinv)* o, size_t idx)
{
    const size_t m = o->m, n = o->n;
    if (n == 1) { o->n = 0; return; }

    /* :269-272 drop row idx of A_G^T */
    memmove(o->At + idx * m, o->At + (idx + 1) * m, (n - idx - 1) * m * sizeof(T));
    /* :275 rank idx -> last */
    FN(square_permute)(o->inv, n, idx, n - 1);
    /* :278-279  d = inv[n-1,n-1] ; last column *= -(1/d) */
    T d = o->inv[(n - 1) * n + (n - 1)];
    T sc = -((T)1 / d);
    for (size_t r = 0; r + 1 < n; ++r) o->inv[r * n + (n - 1)] *= sc;
    /* :284-287  leading block += (-d) u u^T, u = scaled last column */
    for (size_t r = 0; r + 1 < n; ++r) {
        T t = -d * o->inv[r * n + (n - 1)];
        for (size_t k = 0; k + 1 < n; ++k) o->inv[r * n + k] += t * o->inv[k * n + (n - 1)];
    }
    /* :290 shrink to (n-1)^2 */
    for (size_t r = 0; r + 1 < n; ++r)
        memmove(o->inv + r * (n - 1), o->inv + r * n, (n - 1) * sizeof(T));
    o->n--;
}

/* ------------------------------------------------------------------------ */
/* homotopy                      reference: src/solvers/homotopy-cpu.cpp      */

#define A_(i, j) A[(ptrdiff_t)(i) * rs + (ptrdiff_t)(j) * cs]

/* :87-98  c = A^T (y - A x) */
static void FN(residual_vector)(const T* A, size_t m, size_t n, ptrdiff_t rs, ptrdiff_t cs,
                                const T* y, const T* x, T* c, T* r)
{
    for (size_t i = 0; i < m; ++i) {
        T acc = 0;
        for (size_t j = 0; j < n; ++j) if (x[j] != 0) acc += A_(i, j) * x[j];
        r[i] = y[i] - acc;
    }
    for (size_t j = 0; j < n; ++j) {
        T acc = 0;
        for (size_t i = 0; i < m; ++i) acc += A_(i, j) * r[i];
        c[j] = acc;
    }
}

/* :32-37  cblas i?amax: FIRST index of max |v| */
static T FN(inf_norm)(const T* v, size_t n, size_t* idx)
{
    size_t best = 0; T bv = (T)fabs((double)v[0]);
    for (size_t i = 1; i < n; ++i) { T a = (T)fabs((double)v[i]); if (a > bv) { bv = a; best = i; } }
    if (idx) *idx = best;
    return bv;
}

/* :59-67 */
static T FN(sign_tol)(T v, T tol) { return v > tol ? (T)1 : (v < -tol ? (T)-1 : (T)0); }

/* :100-164 */
static void FN(find_max_gamma)(const T* A, size_t m, size_t n, ptrdiff_t rs, ptrdiff_t cs,
                               const T* c, const T* x, const T* direction, T c_inf,
                               const FN(rank_index)* lam, T* p, T* q, T* out_min, size_t* out_idx)
{
    /* :114-120  p = A d ; q = A^T p */
    for (size_t i = 0; i < m; ++i) {
        T acc = 0;
        for (size_t k = 0; k < lam->n; ++k) { size_t j = lam->v[k]; acc += A_(i, j) * direction[j]; }
        p[i] = acc;
    }
    for (size_t j = 0; j < n; ++j) {
        T acc = 0;
        for (size_t i = 0; i < m; ++i) acc += A_(i, j) * p[i];
        q[j] = acc;
    }
    /* :122-161 */
    T min = T_MAX; size_t idx = 0; size_t l = 0;
    for (size_t i = 0; i < n; ++i) {
        const T prev = min;
        if (l < lam->n && lam->v[l] == i) {
            T minT = -x[i] / direction[i];
            if (minT > 0.0 && minT < min) min = minT;
            l++;
        } else {
            T di_left = (T)1 - q[i], di_right = (T)1 + q[i];
            if (di_left != 0.0)  { T t = (c_inf - c[i]) / di_left;  if (t > 0.0 && t < min) min = t; }
            if (di_right != 0.0) { T t = (c_inf + c[i]) / di_right; if (t > 0.0 && t < min) min = t; }
        }
        if (prev > min) idx = i;
    }
    *out_min = min; *out_idx = idx;
}

/* :166-183 */
static void FN(inverse_add_or_remove)(const T* A, size_t m, ptrdiff_t rs, ptrdiff_t cs, size_t col,
                                      FN(rank_index)* lam, FN(oinv)* inv, T* colbuf)
{
    int rank = FN(ri_rank_of)(lam, (uint32_t)col);
    if (rank >= 0) {
        FN(ri_erase)(lam, (uint32_t)col);
        FN(oinv_remove)(inv, (size_t)rank);
    } else {
        rank = FN(ri_insert)(lam, (uint32_t)col);
        for (size_t i = 0; i < m; ++i) colbuf[i] = A_(i, col);
        FN(oinv_insert)(inv, (size_t)rank, colbuf);
    }
}

/* :69-85  scatter the K compact direction values to their N-vector positions */
static void FN(expand)(T* direction, size_t n, const FN(rank_index)* lam)
{
    ptrdiff_t i = (ptrdiff_t)n - 1, j = (ptrdiff_t)lam->n - 1;
    for (ptrdiff_t it = (ptrdiff_t)lam->n - 1; it >= 0; --it) {
        while (i > (ptrdiff_t)lam->v[it]) direction[i--] = 0;
        direction[i--] = direction[j--];
    }
    while (i >= 0) direction[i--] = 0;
}

/* :185-275 run_solver */
void FN(ssor_homotopy)(const T* A, size_t m, size_t n, ptrdiff_t rs, ptrdiff_t cs,
                       const T* y, T tol, uint32_t max_iter, T* x, uint32_t* out_iter, double* out_err)
{
    T* direction = (T*)calloc(n, sizeof(T));
    T* c = (T*)calloc(n, sizeof(T));
    T* c_gamma = (T*)calloc(n, sizeof(T));
    T* p = (T*)calloc(m, sizeof(T));
    T* q = (T*)calloc(n, sizeof(T));
    T* r = (T*)calloc(m, sizeof(T));
    T* colbuf = (T*)calloc(m, sizeof(T));
    FN(rank_index) lam = { 0, 0, 0 };
    FN(oinv) inv = { 0, 0, m, 0, 0 };
    T c_inf = 0;

    for (size_t j = 0; j < n; ++j) x[j] = 0;                          /* :204 */
    FN(residual_vector)(A, m, n, rs, cs, y, x, c, r);                  /* :215 */
    {
        size_t idx;
        c_inf = FN(inf_norm)(c, n, &idx);                              /* :219 */
        FN(inverse_add_or_remove)(A, m, rs, cs, idx, &lam, &inv, colbuf); /* :221 */
        T cg = FN(sign_tol)(c_inf, tol);                               /* :223-224: sign of |c|, a quirk */
        direction[0] = cg * inv.inv[0];                                /* :227 */
        FN(expand)(direction, n, &lam);                                /* :228 */
    }

    uint32_t iter = 0;
    do {
        iter++;
        T min; size_t idx;
        FN(find_max_gamma)(A, m, n, rs, cs, c, x, direction, c_inf, &lam, p, q, &min, &idx); /* :241 */
        FN(inverse_add_or_remove)(A, m, rs, cs, idx, &lam, &inv, colbuf);                    /* :246 */
        size_t K = lam.n;
        if (K == 0) break;                                             /* :248-249 */
        for (size_t j = 0; j < n; ++j) x[j] += min * direction[j];     /* :252 (old direction) */
        FN(residual_vector)(A, m, n, rs, cs, y, x, c, r);              /* :255 */
        for (size_t k = 0; k < K; ++k) c_gamma[k] = FN(sign_tol)(c[lam.v[k]], tol); /* :259-260 */
        for (size_t a = 0; a < K; ++a) {                               /* :263 */
            T acc = 0;
            for (size_t b = 0; b < K; ++b) acc += inv.inv[a * K + b] * c_gamma[b];
            direction[a] = acc;
        }
        FN(expand)(direction, n, &lam);                                /* :266 */
        c_inf = FN(inf_norm)(c, n, 0);                                 /* :270 */
    } while (iter < max_iter && c_inf > tol);                          /* :272 */

    *out_iter = iter; *out_err = (double)c_inf;                        /* :274 */
    free(direction); free(c); free(c_gamma); free(p); free(q); free(r); free(colbuf);
    free(lam.v); free(inv.At); free(inv.inv);
}
#undef A_

/* ------------------------------------------------------------------------ */
/* Householder QR               reference: src/linalg/qr_decomposition.h      */

typedef struct { T* qr; T* rdiag; size_t M, N; } FN(qr_t);

/* :65-88 next_householder */
static void FN(next_householder)(T* A, size_t M, size_t N, size_t n, T* hvec, T* nrm2)
{
    if (n < N) {
        *nrm2 = 0;
        if (n > 0) A[(n - 1) * N + (n - 1)] = hvec[n - 1];
        for (size_t m = n; m < M; ++m) {
            if (n > 0) A[m * N + (n - 1)] = hvec[m];
            hvec[m] = A[m * N + n];
            *nrm2 = (T)hypot((double)*nrm2, (double)hvec[m]);
        }
    } else {
        for (size_t m = n - 1; m < M; ++m) A[m * N + (n - 1)] = hvec[m];
    }
}

/* :93-137 constructor (A copied row-major M x N, requires M >= N) */
static void FN(qr_factor)(FN(qr_t)* f, const T* A, size_t M, size_t N, ptrdiff_t rs, ptrdiff_t cs)
{
    f->M = M; f->N = N;
    f->qr = (T*)malloc(M * N * sizeof(T));
    f->rdiag = (T*)calloc(N, sizeof(T));
    for (size_t i = 0; i < M; ++i) for (size_t j = 0; j < N; ++j) f->qr[i * N + j] = A[(ptrdiff_t)i * rs + (ptrdiff_t)j * cs];
    T* s = (T*)calloc(N, sizeof(T));
    T* hvec = (T*)calloc(M, sizeof(T));
    T nrm2;
    for (size_t k = 0; k < N; ++k) {
        FN(next_householder)(f->qr, M, N, k, hvec, &nrm2);            /* :111 */
        if (nrm2 != (T)0) {
            if (hvec[k] < 0) nrm2 = -nrm2;                             /* :113 */
            for (size_t m = k; m < M; ++m) hvec[m] /= nrm2;            /* :116 */
            hvec[k] += (T)1;                                           /* :117 */
            for (size_t n = 0; n < N; ++n) s[n] = 0;                   /* :120 */
            for (size_t m = k; m < M; ++m)                             /* :121-125 */
                for (size_t n = k + 1; n < N; ++n) s[n] += hvec[m] * f->qr[m * N + n];
            for (size_t n = 0; n < N; ++n) s[n] /= -hvec[k];           /* :127 */
            for (size_t m = k; m < M; ++m)                             /* :128-132 */
                for (size_t n = k + 1; n < N; ++n) f->qr[m * N + n] += hvec[m] * s[n];
        }
        f->rdiag[k] = -nrm2;                                           /* :134 */
    }
    FN(next_householder)(f->qr, M, N, N, hvec, &nrm2);                 /* :136 */
    free(s); free(hvec);
}

/* :139-173 q(): thin Q, M x N row-major */
static void FN(qr_q)(const FN(qr_t)* f, T* q)
{
    const size_t M = f->M, N = f->N;
    for (ptrdiff_t k = (ptrdiff_t)N - 1; k >= 0; --k) {
        for (size_t m = 0; m < M; ++m) q[m * N + k] = 0;               /* :153 */
        q[k * N + k] = (T)1;                                           /* :154 */
        for (size_t n = (size_t)k; n < N; ++n) {
            if (f->qr[k * N + k] != (T)0) {
                T s = 0;
                for (size_t m = (size_t)k; m < M; ++m) s += f->qr[m * N + k] * q[m * N + n];
                s = -s / f->qr[k * N + k];
                for (size_t m = (size_t)k; m < M; ++m) q[m * N + n] += s * f->qr[m * N + k];
            }
        }
    }
}

/* :175-190 r(): N x N upper */
static void FN(qr_r)(const FN(qr_t)* f, T* r)
{
    const size_t N = f->N;
    for (size_t i = 0; i < N * N; ++i) r[i] = 0;
    for (size_t m = 0; m < N; ++m) {
        r[m * N + m] = f->rdiag[m];
        for (size_t n = m + 1; n < N; ++n) r[m * N + n] = f->qr[m * N + n];
    }
}

/* ------------------------------------------------------------------------ */
/* Cholesky                     reference: src/linalg/cholesky_decomposition.h */

/* :57-86: left-looking on the LOWER triangle of G; returns isspd */
static int FN(cholesky)(T* l, const T* G, size_t N)
{
    int isspd = 1;
    for (size_t i = 0; i < N; ++i) for (size_t j = 0; j < N; ++j) l[i * N + j] = (j <= i) ? G[i * N + j] : (T)0; /* :69 */
    for (size_t j = 0; j < N; ++j) {
        if (j > 0)                                                     /* :74-78  v -= L[j:,0:j] L[j,0:j] */
            for (size_t i = j; i < N; ++i) {
                T acc = 0;
                for (size_t k = 0; k < j; ++k) acc += l[i * N + k] * l[j * N + k];
                l[i * N + j] = l[i * N + j] - acc;
            }
        T ajj = (T)sqrt((double)l[j * N + j]);                         /* :80 */
        if (ajj <= T_EPS) isspd = 0;                                   /* :81-83 (NaN pivots pass!) */
        T inv = (T)1 / ajj;                                            /* :84 */
        for (size_t i = j; i < N; ++i) l[i * N + j] *= inv;
    }
    return isspd;
}

/* :93-103 solve: L z = b, L^T x = z */
static void FN(chol_solve)(const T* l, size_t N, T* x)
{
    for (size_t i = 0; i < N; ++i) {
        T acc = x[i];
        for (size_t k = 0; k < i; ++k) acc -= l[i * N + k] * x[k];
        x[i] = acc / l[i * N + i];
    }
    for (ptrdiff_t i = (ptrdiff_t)N - 1; i >= 0; --i) {
        T acc = x[i];
        for (size_t k = (size_t)i + 1; k < N; ++k) acc -= l[k * N + i] * x[k];
        x[i] = acc / l[i * N + i];
    }
}

/* ------------------------------------------------------------------------ */
/* IRLS                          reference: src/solvers/irls-cpu.cpp          */

/* :39-60 irls_newton */
static int FN(irls_newton)(const T* R, const T* Q, size_t M, size_t N, const T* y, const T* w,
                           T* x, T* G, T* L, T* s, T* t)
{
    /* :47-48  G = Q^T (Q o w) : column j of Q scaled by w_j */
    for (size_t i = 0; i < N; ++i)
        for (size_t j = 0; j < N; ++j) {
            T acc = 0;
            for (size_t m = 0; m < M; ++m) acc += Q[m * N + i] * (Q[m * N + j] * w[j]);
            G[i * N + j] = acc;
        }
    if (!FN(cholesky)(L, G, N)) return 0;                              /* :50-51 */
    for (size_t j = 0; j < N; ++j) {                                   /* :53 s = Q^T y */
        T acc = 0;
        for (size_t m = 0; m < M; ++m) acc += Q[m * N + j] * y[m];
        s[j] = acc;
    }
    FN(chol_solve)(L, N, s);                                           /* :54 */
    for (size_t m = 0; m < M; ++m) {                                   /* :55 t = Q s */
        T acc = 0;
        for (size_t j = 0; j < N; ++j) acc += Q[m * N + j] * s[j];
        t[m] = acc;
    }
    for (size_t j = 0; j < N; ++j) {                                   /* :57 x = Q^T t */
        T acc = 0;
        for (size_t m = 0; m < M; ++m) acc += Q[m * N + j] * t[m];
        x[j] = acc;
    }
    for (ptrdiff_t i = (ptrdiff_t)N - 1; i >= 0; --i) {                /* :58 R x = x (upper) */
        T acc = x[i];
        for (size_t k = (size_t)i + 1; k < N; ++k) acc -= R[i * N + k] * x[k];
        x[i] = acc / R[i * N + i];
    }
    return 1;
}

/* :62-124 run_solver, given Q (M x N) and R (N x N) */
void FN(ssor_irls_qr)(const T* Q, const T* R, size_t M, size_t N, const T* y, T tol, uint32_t max_iter,
                      T* x, uint32_t* out_iter, double* out_err, uint8_t* out_spd)
{
    const T p = (T)0.9;                                                /* :70 */
    T* w = (T*)malloc(N * sizeof(T));
    T* xnext = (T*)malloc(N * sizeof(T));
    T* G = (T*)malloc(N * N * sizeof(T));
    T* L = (T*)malloc(N * N * sizeof(T));
    T* s = (T*)malloc(N * sizeof(T));
    T* t = (T*)malloc(M * sizeof(T));
    for (size_t j = 0; j < N; ++j) { x[j] = 0; w[j] = 1; xnext[j] = 1; } /* :82-85 */
    uint32_t iter = 0; int spd_error = 0; T abstol = 1, eps = 1, second = 1;

    do {
        if (!FN(irls_newton)(R, Q, M, N, y, w, xnext, G, L, s, t)) { spd_error = 1; break; } /* :94-97 */
        T mx = xnext[0];                                               /* :100 signed max */
        for (size_t j = 1; j < N; ++j) if (xnext[j] > mx) mx = xnext[j];
        abstol = mx * tol;
        for (size_t j = 0; j < N; ++j) { if (xnext[j] < abstol) xnext[j] = 0; x[j] = xnext[j]; } /* :103-104 */
        /* :107  nth_element(.., begin+1, .., greater) -> second largest value */
        {
            T a = xnext[0], b = (N > 1) ? xnext[1] : xnext[0];
            if (b > a) { T tmp = a; a = b; b = tmp; }
            for (size_t j = 2; j < N; ++j) { T v = xnext[j]; if (v > a) { b = a; a = v; } else if (v > b) b = v; }
            second = b;
        }
        if (second / (T)N < eps) eps = second / (T)N;                  /* :110 */
        /* :113  w = (x*x + eps)^(p/2 - 1): base in T, exponent and pow in double */
        const double ex = ((double)p / 2.0) - 1.0;
        for (size_t j = 0; j < N; ++j) w[j] = (T)pow((double)(T)(x[j] * x[j] + eps), ex);
        T sw = 0; for (size_t j = 0; j < N; ++j) sw += w[j];           /* :114 */
        for (size_t j = 0; j < N; ++j) w[j] /= sw;
        iter++;
    } while (iter < max_iter && second > abstol);                      /* :118 */

    T sx = 0; for (size_t j = 0; j < N; ++j) sx += x[j];               /* :121 */
    for (size_t j = 0; j < N; ++j) x[j] /= sx;
    *out_iter = iter; *out_err = (double)eps; *out_spd = (uint8_t)spd_error; /* :123 */
    free(w); free(xnext); free(G); free(L); free(s); free(t);
}

/* irls_state ctor + solve (src/lib.cpp:51-73): factor, form Q and R, run */
void FN(ssor_irls)(const T* A, size_t m, size_t n, ptrdiff_t rs, ptrdiff_t cs,
                   const T* Y, size_t batch, size_t ldy, T tol, uint32_t max_iter,
                   T* X, size_t ldx, uint32_t* iters, double* errs, uint8_t* spd)
{
    FN(qr_t) f;
    FN(qr_factor)(&f, A, m, n, rs, cs);
    T* Q = (T*)malloc(m * n * sizeof(T));
    T* R = (T*)malloc(n * n * sizeof(T));
    FN(qr_q)(&f, Q); FN(qr_r)(&f, R);
    for (size_t b = 0; b < batch; ++b)
        FN(ssor_irls_qr)(Q, R, m, n, Y + b * ldy, tol, max_iter, X + b * ldx, &iters[b], &errs[b], &spd[b]);
    free(Q); free(R); free(f.qr); free(f.rdiag);
}

/* exposes Q and R so tests can pin them (qr_decomposition_test.cpp:36-85) */
void FN(ssor_qr)(const T* A, size_t m, size_t n, ptrdiff_t rs, ptrdiff_t cs, T* Q, T* R)
{
    FN(qr_t) f;
    FN(qr_factor)(&f, A, m, n, rs, cs);
    FN(qr_q)(&f, Q); FN(qr_r)(&f, R);
    free(f.qr); free(f.rdiag);
}

void FN(ssor_homotopy_batch)(const T* A, size_t m, size_t n, ptrdiff_t rs, ptrdiff_t cs,
                             const T* Y, size_t batch, size_t ldy, T tol, uint32_t max_iter,
                             T* X, size_t ldx, uint32_t* iters, double* errs)
{
    /* pack A column-major once so both products stream contiguously (layout only) */
    T* Ac = (T*)malloc(m * n * sizeof(T));
    for (size_t j = 0; j < n; ++j) for (size_t i = 0; i < m; ++i) Ac[j * m + i] = A[(ptrdiff_t)i * rs + (ptrdiff_t)j * cs];
    #pragma omp parallel for schedule(dynamic, 1)
    for (ptrdiff_t b = 0; b < (ptrdiff_t)batch; ++b)
        FN(ssor_homotopy)(Ac, m, n, 1, (ptrdiff_t)m, Y + b * ldy, tol, max_iter, X + b * ldx, &iters[b], &errs[b]);
    free(Ac);
}
