/* movement_oracle_c.c -- C/OpenMP restatement of the Monge-Ampere relaxation hot path.
 *
 * TEST / MEASUREMENT INFRASTRUCTURE ONLY (same status as movement_oracle.py): it is the timed
 * CPU baseline of bench.py ("cpu_baseline", kind "port") and of `bench.py --impl reference`, and
 * is itself checked against the numpy oracle in tests/test_oracle_c.py.  Nothing under
 * movement_b200/ links or loads it.  Parity vs a live Firedrake run: unpinned in this container
 * (Firedrake/PETSc are not installable here); it follows, like the numpy oracle,
 *   movement/monge_ampere.py:311-342 (L2 projection), :226-244 (x = xi + grad phi),
 *   :405-410,514-517 (theta, residual), :419-449 (pseudo-timestep), :451-485 (Hessian recovery),
 *   :500-538 (loop), movement/mover.py:199-217 (diagnostics), movement/tangling.py:26-46,
 * using the P1 closed forms of SURVEY 8a-R and Jacobi-preconditioned CG for every inner solve
 * (the reference: PETSc CG + ILU / GAMG; the converged answers agree to the solver tolerance).
 *
 * Build: make -C oracle   (gcc -O3 -fopenmp -shared)
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef struct {
  int dim;
  int64_t V, C, F;
  double* xi;      /* [V,d] */
  int32_t* cells;  /* [C,d+1] */
  double* vol;     /* [C] */
  double* grad;    /* [C,d+1,d] */
  int64_t* v2c_ptr;
  int32_t* v2c;    /* entries cell*(d+1)+loc, ascending */
  /* CSR of M and K (same pattern) */
  int64_t* rp;
  int32_t* ci;
  double *M, *K, *ml, *dinvM, *dinvK;
  /* boundary facets */
  int32_t* bf;      /* [F,d] */
  int32_t* bfcell;  /* [F] */
  double *bfn, *bfa;
  int64_t* v2f_ptr;
  int32_t* v2f;     /* entries f*d+loc */
  uint16_t* bcmask; /* [V] */
  double total_volume, ml_sumsq;
  /* state */
  double *x, *phi, *phi_old, *H, *H_old, *g, *mon, *a, *res, *rhs, *volume, *vol0;
  /* cg work */
  double *r, *p, *q, *z;
  int nthreads;
} orc_ma;

static void* xmalloc(size_t n) {
  void* p = calloc(n ? n : 1, 1);
  if (!p) { fprintf(stderr, "oracle_c: out of memory (%zu bytes)\n", n); abort(); }
  return p;
}

/* ---------------------------------------------------------------- P1 geometry */
static void geom2(const double* x0, const double* x1, const double* x2, double* vol, double* g) {
  const double a = x1[0] - x0[0], b = x2[0] - x0[0], c = x1[1] - x0[1], d = x2[1] - x0[1];
  const double det = a * d - b * c, inv = 1.0 / det;
  *vol = 0.5 * fabs(det);
  g[2] = d * inv; g[3] = -b * inv; g[4] = -c * inv; g[5] = a * inv;
  g[0] = -(g[2] + g[4]); g[1] = -(g[3] + g[5]);
}
static void geom3(const double* x0, const double* x1, const double* x2, const double* x3, double* vol, double* g) {
  double e1[3], e2[3], e3[3];
  for (int a = 0; a < 3; ++a) { e1[a] = x1[a] - x0[a]; e2[a] = x2[a] - x0[a]; e3[a] = x3[a] - x0[a]; }
  const double c1[3] = {e2[1] * e3[2] - e2[2] * e3[1], e2[2] * e3[0] - e2[0] * e3[2], e2[0] * e3[1] - e2[1] * e3[0]};
  const double c2[3] = {e3[1] * e1[2] - e3[2] * e1[1], e3[2] * e1[0] - e3[0] * e1[2], e3[0] * e1[1] - e3[1] * e1[0]};
  const double c3[3] = {e1[1] * e2[2] - e1[2] * e2[1], e1[2] * e2[0] - e1[0] * e2[2], e1[0] * e2[1] - e1[1] * e2[0]};
  const double det = e1[0] * c1[0] + e1[1] * c1[1] + e1[2] * c1[2], inv = 1.0 / det;
  *vol = fabs(det) / 6.0;
  for (int a = 0; a < 3; ++a) {
    g[3 + a] = c1[a] * inv; g[6 + a] = c2[a] * inv; g[9 + a] = c3[a] * inv;
    g[a] = -(g[3 + a] + g[6 + a] + g[9 + a]);
  }
}
static void cell_geom(int d, const double* X, const int32_t* cv, double* vol, double* g) {
  if (d == 2) geom2(X + 2 * (int64_t)cv[0], X + 2 * (int64_t)cv[1], X + 2 * (int64_t)cv[2], vol, g);
  else geom3(X + 3 * (int64_t)cv[0], X + 3 * (int64_t)cv[1], X + 3 * (int64_t)cv[2], X + 3 * (int64_t)cv[3], vol, g);
}

/* signed detJ, cofactor expansion in the order of movement_oracle.signed_detJ (no FMA: -ffp-contract=off) */
static double signed_detJ(int d, const double* X, const int32_t* cv) {
  if (d == 2) {
    const double* x0 = X + 2 * (int64_t)cv[0]; const double* x1 = X + 2 * (int64_t)cv[1]; const double* x2 = X + 2 * (int64_t)cv[2];
    const double e00 = x1[0] - x0[0], e01 = x1[1] - x0[1], e10 = x2[0] - x0[0], e11 = x2[1] - x0[1];
    return e00 * e11 - e01 * e10;
  }
  const double* x0 = X + 3 * (int64_t)cv[0];
  double a[3], b[3], c[3];
  for (int k = 0; k < 3; ++k) {
    a[k] = X[3 * (int64_t)cv[1] + k] - x0[k]; b[k] = X[3 * (int64_t)cv[2] + k] - x0[k]; c[k] = X[3 * (int64_t)cv[3] + k] - x0[k];
  }
  const double m0 = b[1] * c[2] - b[2] * c[1], m1 = b[0] * c[2] - b[2] * c[0], m2 = b[0] * c[1] - b[1] * c[0];
  return (a[0] * m0 - a[1] * m1) + a[2] * m2;
}

/* ---------------------------------------------------------------- inverse maps */
static void inverse_map(const int32_t* keys, int64_t n, int64_t nkeys, int64_t** ptr_out, int32_t** list_out) {
  int64_t* ptr = (int64_t*)xmalloc((nkeys + 1) * sizeof(int64_t));
  for (int64_t i = 0; i < n; ++i) ptr[keys[i] + 1]++;
  for (int64_t k = 0; k < nkeys; ++k) ptr[k + 1] += ptr[k];
  int32_t* list = (int32_t*)xmalloc((size_t)n * sizeof(int32_t));
  int64_t* cur = (int64_t*)xmalloc(nkeys * sizeof(int64_t));
  memcpy(cur, ptr, nkeys * sizeof(int64_t));
  for (int64_t i = 0; i < n; ++i) list[cur[keys[i]]++] = (int32_t)i; /* ascending i within a key */
  free(cur);
  *ptr_out = ptr; *list_out = list;
}

static int cmp_i32(const void* a, const void* b) { return (*(const int32_t*)a > *(const int32_t*)b) - (*(const int32_t*)a < *(const int32_t*)b); }

/* coords [V,d], cells [C,d+1], boundary facets with adjacent cell / outward normal / measure and the
 * Dirichlet component mask of the L2 projector (bit a = component a constrained) are inputs: they are
 * produced by the numpy oracle's MAOperators (monge_ampere.py:246-272 axis-aligned branch). */
orc_ma* orc_ma_create(int dim, int64_t V, int64_t C, const double* coords, const int32_t* cells, int64_t F,
                      const int32_t* bf, const int32_t* bfcell, const double* bfn, const double* bfa,
                      const uint16_t* bcmask, int nthreads) {
  orc_ma* o = (orc_ma*)xmalloc(sizeof(orc_ma));
  const int d = dim, d1 = dim + 1, dd = dim * dim;
  o->dim = d; o->V = V; o->C = C; o->F = F;
#ifdef _OPENMP
  o->nthreads = nthreads > 0 ? nthreads : omp_get_max_threads();
  omp_set_num_threads(o->nthreads);
#else
  o->nthreads = 1;
#endif
  o->xi = (double*)xmalloc(V * d * sizeof(double)); memcpy(o->xi, coords, V * d * sizeof(double));
  o->cells = (int32_t*)xmalloc(C * d1 * sizeof(int32_t)); memcpy(o->cells, cells, C * d1 * sizeof(int32_t));
  o->vol = (double*)xmalloc(C * sizeof(double));
  o->grad = (double*)xmalloc(C * d1 * d * sizeof(double));
#pragma omp parallel for schedule(static)
  for (int64_t c = 0; c < C; ++c) cell_geom(d, o->xi, o->cells + c * d1, &o->vol[c], o->grad + c * d1 * d);
  double tv = 0.0;
  for (int64_t c = 0; c < C; ++c) tv += o->vol[c];
  o->total_volume = tv;
  inverse_map(o->cells, C * d1, V, &o->v2c_ptr, &o->v2c);
  /* pattern */
  o->rp = (int64_t*)xmalloc((V + 1) * sizeof(int64_t));
#pragma omp parallel
  {
    int32_t nb[512];
#pragma omp for schedule(static)
    for (int64_t v = 0; v < V; ++v) {
      int n = 0; nb[n++] = (int32_t)v;
      for (int64_t t = o->v2c_ptr[v]; t < o->v2c_ptr[v + 1]; ++t) {
        const int32_t* cv = o->cells + (int64_t)(o->v2c[t] / d1) * d1;
        for (int j = 0; j < d1; ++j) if (n < 512) nb[n++] = cv[j];
      }
      qsort(nb, n, sizeof(int32_t), cmp_i32);
      int u = 0;
      for (int i = 0; i < n; ++i) if (i == 0 || nb[i] != nb[i - 1]) ++u;
      o->rp[v + 1] = u;
    }
  }
  o->rp[0] = 0;
  for (int64_t v = 0; v < V; ++v) o->rp[v + 1] += o->rp[v];
  const int64_t nnz = o->rp[V];
  o->ci = (int32_t*)xmalloc(nnz * sizeof(int32_t));
  o->M = (double*)xmalloc(nnz * sizeof(double));
  o->K = (double*)xmalloc(nnz * sizeof(double));
  o->ml = (double*)xmalloc(V * sizeof(double));
  o->dinvM = (double*)xmalloc(V * sizeof(double));
  o->dinvK = (double*)xmalloc(V * sizeof(double));
  const double mfac = 1.0 / ((d + 1) * (d + 2));
#pragma omp parallel
  {
    int32_t nb[512];
#pragma omp for schedule(static)
    for (int64_t v = 0; v < V; ++v) {
      int n = 0; nb[n++] = (int32_t)v;
      for (int64_t t = o->v2c_ptr[v]; t < o->v2c_ptr[v + 1]; ++t) {
        const int32_t* cv = o->cells + (int64_t)(o->v2c[t] / d1) * d1;
        for (int j = 0; j < d1; ++j) if (n < 512) nb[n++] = cv[j];
      }
      qsort(nb, n, sizeof(int32_t), cmp_i32);
      int64_t b = o->rp[v]; int u = 0;
      for (int i = 0; i < n; ++i) if (i == 0 || nb[i] != nb[i - 1]) o->ci[b + u++] = nb[i];
      double mlv = 0, dM = 0, dK = 0;
      for (int64_t t = o->v2c_ptr[v]; t < o->v2c_ptr[v + 1]; ++t) {
        const int64_t c = o->v2c[t] / d1; const int loc = o->v2c[t] % d1;
        const int32_t* cv = o->cells + c * d1; const double* g = o->grad + c * d1 * d; const double vol = o->vol[c];
        mlv += vol / d1;
        for (int j = 0; j < d1; ++j) {
          double kij = 0; for (int a = 0; a < d; ++a) kij += g[loc * d + a] * g[j * d + a];
          kij *= vol;
          const double mij = vol * mfac * (j == loc ? 2.0 : 1.0);
          int pos = 0; while (o->ci[b + pos] != cv[j]) ++pos;
          o->M[b + pos] += mij; o->K[b + pos] += kij;
          if (j == loc) { dM += mij; dK += kij; }
        }
      }
      o->ml[v] = mlv; o->dinvM[v] = 1.0 / dM; o->dinvK[v] = 1.0 / dK;
    }
  }
  double s2 = 0; for (int64_t v = 0; v < V; ++v) s2 += o->ml[v] * o->ml[v];
  o->ml_sumsq = s2;
  /* boundary */
  o->bf = (int32_t*)xmalloc(F * d * sizeof(int32_t)); memcpy(o->bf, bf, F * d * sizeof(int32_t));
  o->bfcell = (int32_t*)xmalloc(F * sizeof(int32_t)); memcpy(o->bfcell, bfcell, F * sizeof(int32_t));
  o->bfn = (double*)xmalloc(F * d * sizeof(double)); memcpy(o->bfn, bfn, F * d * sizeof(double));
  o->bfa = (double*)xmalloc(F * sizeof(double)); memcpy(o->bfa, bfa, F * sizeof(double));
  inverse_map(o->bf, F * d, V, &o->v2f_ptr, &o->v2f);
  o->bcmask = (uint16_t*)xmalloc(V * sizeof(uint16_t)); memcpy(o->bcmask, bcmask, V * sizeof(uint16_t));
  /* state */
  o->x = (double*)xmalloc(V * d * sizeof(double)); memcpy(o->x, coords, V * d * sizeof(double));
  o->phi = (double*)xmalloc(V * sizeof(double)); o->phi_old = (double*)xmalloc(V * sizeof(double));
  o->H = (double*)xmalloc(V * dd * sizeof(double)); o->H_old = (double*)xmalloc(V * dd * sizeof(double));
  o->g = (double*)xmalloc(V * d * sizeof(double)); o->mon = (double*)xmalloc(V * sizeof(double));
  o->a = (double*)xmalloc(V * sizeof(double)); o->res = (double*)xmalloc(V * sizeof(double));
  o->rhs = (double*)xmalloc(V * dd * sizeof(double));
  o->volume = (double*)xmalloc(C * sizeof(double)); o->vol0 = (double*)xmalloc(C * sizeof(double));
  memcpy(o->vol0, o->vol, C * sizeof(double));
  o->r = (double*)xmalloc(V * dd * sizeof(double)); o->p = (double*)xmalloc(V * dd * sizeof(double));
  o->q = (double*)xmalloc(V * dd * sizeof(double)); o->z = (double*)xmalloc(V * dd * sizeof(double));
  return o;
}

void orc_ma_destroy(orc_ma* o) {
  if (!o) return;
  void* ps[] = {o->xi, o->cells, o->vol, o->grad, o->v2c_ptr, o->v2c, o->rp, o->ci, o->M, o->K, o->ml, o->dinvM, o->dinvK,
                o->bf, o->bfcell, o->bfn, o->bfa, o->v2f_ptr, o->v2f, o->bcmask, o->x, o->phi, o->phi_old, o->H, o->H_old,
                o->g, o->mon, o->a, o->res, o->rhs, o->volume, o->vol0, o->r, o->p, o->q, o->z};
  for (size_t i = 0; i < sizeof(ps) / sizeof(ps[0]); ++i) free(ps[i]);
  free(o);
}

/* ---------------------------------------------------------------- Jacobi-PCG, nc interleaved right-hand sides
 * Same algorithm and stopping rule as the GPU path: stop when r.D^-1.r <= rtol^2 * max(r0.D^-1.r0, b.D^-1.b)
 * for every component.  mask bit c => dof fixed (kept).  project: constant nullspace (rhs, x -> zero mean). */
static void spmv(const orc_ma* o, const double* A, int nc, const double* x, double* y) {
  const int64_t V = o->V;
#pragma omp parallel for schedule(static)
  for (int64_t v = 0; v < V; ++v) {
    double acc[9] = {0};
    for (int64_t k = o->rp[v]; k < o->rp[v + 1]; ++k) {
      const double a = A[k]; const double* xp = x + (int64_t)o->ci[k] * nc;
      for (int c = 0; c < nc; ++c) acc[c] += a * xp[c];
    }
    for (int c = 0; c < nc; ++c) y[v * nc + c] = acc[c];
  }
}

static int pcg(orc_ma* o, const double* A, const double* dinv, int nc, const double* b, double* x, const uint16_t* mask,
               double rtol, int maxit, int project) {
  const int64_t V = o->V;
  double *r = o->r, *p = o->p, *q = o->q;
  double mean[9] = {0};
  if (project && b) {
    for (int c = 0; c < nc; ++c) { double s = 0; for (int64_t v = 0; v < V; ++v) s += b[v * nc + c]; mean[c] = s / (double)V; }
  }
  spmv(o, A, nc, x, q);
  double rz[9] = {0}, bb[9] = {0}, thr[9];
  for (int c = 0; c < nc; ++c) {
    double s1 = 0, s2 = 0;
#pragma omp parallel for reduction(+ : s1, s2) schedule(static)
    for (int64_t v = 0; v < V; ++v) {
      const int64_t i = v * nc + c;
      const int fix = mask && ((mask[v] >> c) & 1);
      const double bv = b ? b[i] - mean[c] : 0.0;
      const double rr = fix ? 0.0 : bv - q[i];
      const double zz = dinv[v] * rr;
      r[i] = rr; p[i] = zz;
      s1 += rr * zz; if (!fix) s2 += bv * dinv[v] * bv;
    }
    rz[c] = s1; bb[c] = s2;
    const double ref = s1 > s2 ? s1 : s2;
    thr[c] = rtol * rtol * ref;
  }
  int it = 0;
  for (;; ++it) {
    int done = 1;
    for (int c = 0; c < nc; ++c) if (!(rz[c] <= thr[c])) done = 0;
    if (done || it >= maxit) break;
    spmv(o, A, nc, p, q);
    double alpha[9], beta[9];
    for (int c = 0; c < nc; ++c) {
      double pq = 0;
#pragma omp parallel for reduction(+ : pq) schedule(static)
      for (int64_t v = 0; v < V; ++v) pq += p[v * nc + c] * q[v * nc + c];
      alpha[c] = pq > 0 ? rz[c] / pq : 0.0;
    }
    for (int c = 0; c < nc; ++c) {
      double s = 0; const double al = alpha[c];
#pragma omp parallel for reduction(+ : s) schedule(static)
      for (int64_t v = 0; v < V; ++v) {
        const int64_t i = v * nc + c;
        x[i] += al * p[i];
        const double rr = r[i] - al * q[i];
        r[i] = rr;
        const int fix = mask && ((mask[v] >> c) & 1);
        if (!fix) s += rr * dinv[v] * rr;
      }
      beta[c] = rz[c] > 0 ? s / rz[c] : 0.0; rz[c] = s;
    }
#pragma omp parallel for schedule(static)
    for (int64_t v = 0; v < V; ++v)
      for (int c = 0; c < nc; ++c) {
        const int64_t i = v * nc + c;
        const int fix = mask && ((mask[v] >> c) & 1);
        p[i] = (fix ? 0.0 : dinv[v] * r[i]) + beta[c] * p[i];
      }
  }
  if (project)
    for (int c = 0; c < nc; ++c) {
      double s = 0; for (int64_t v = 0; v < V; ++v) s += x[v * nc + c];
      s /= (double)V; for (int64_t v = 0; v < V; ++v) x[v * nc + c] -= s;
    }
  return it;
}

/* ---------------------------------------------------------------- quadrature (degree d+2): Dunavant-6 / 14-point tet */
static const double Q2A = 0.44594849091596488631832925388305, Q2B = 0.09157621350977074345957146340220;
static const double Q2WA = 0.22338158967801146569500700843312, Q2WB = 0.10995174365532186763832632490021;
static int quad_rule(int d, double lam[][4], double* w) {
  if (d == 2) {
    const double a = Q2A, b = Q2B;
    const double L[6][3] = {{1 - 2 * a, a, a}, {a, 1 - 2 * a, a}, {a, a, 1 - 2 * a}, {1 - 2 * b, b, b}, {b, 1 - 2 * b, b}, {b, b, 1 - 2 * b}};
    for (int q = 0; q < 6; ++q) { for (int j = 0; j < 3; ++j) lam[q][j] = L[q][j]; w[q] = q < 3 ? Q2WA : Q2WB; }
    return 6;
  }
  const double a1 = 0.31088591926330060980, a2 = 0.09273525031089122640, b = 0.04550370412564964949;
  const double c1 = 1 - 3 * a1, c2 = 1 - 3 * a2, dd = 0.5 - b;
  const double L[14][4] = {{c1, a1, a1, a1}, {a1, c1, a1, a1}, {a1, a1, c1, a1}, {a1, a1, a1, c1}, {c2, a2, a2, a2}, {a2, c2, a2, a2},
                           {a2, a2, c2, a2}, {a2, a2, a2, c2}, {b, b, dd, dd},   {b, dd, b, dd},   {b, dd, dd, b},   {dd, b, b, dd},
                           {dd, b, dd, b},   {dd, dd, b, b}};
  for (int q = 0; q < 14; ++q) {
    for (int j = 0; j < 4; ++j) lam[q][j] = L[q][j];
    w[q] = q < 4 ? 0.11268792571801585080 : (q < 8 ? 0.07349304311636194954 : 0.04254602077708146644);
  }
  return 14;
}
static double det_I_plus(int d, const double* h) {
  if (d == 2) return (1 + h[0]) * (1 + h[3]) - h[1] * h[2];
  const double a = 1 + h[0], e = 1 + h[4], i = 1 + h[8];
  return a * (e * i - h[5] * h[7]) - h[1] * (h[3] * i - h[5] * h[6]) + h[2] * (h[3] * h[7] - e * h[6]);
}

/* ---------------------------------------------------------------- monitors (kinds as MV_MON_*) */
static double monitor_raw(int kind, const double* P, int d, const double* x) {
  if (kind == 1 || kind == 2) {
    double dist = 0; for (int a = 0; a < d; ++a) dist += (x[a] - P[a]) * (x[a] - P[a]);
    const double rr = kind == 1 ? P[3] * P[3] : P[3];
    const double ch = cosh(P[5] * (dist - rr));
    return 1.0 + P[4] / (ch * ch);
  }
  if (kind == 3) {
    const double w2 = P[2] * P[2], dx = x[0] - P[0], dy = x[1] - P[1];
    const double u = exp(-(dx * dx + dy * dy) / w2);
    const double uxx = u * (4 * dx * dx / (w2 * w2) - 2 / w2), uyy = u * (4 * dy * dy / (w2 * w2) - 2 / w2), uxy = u * (4 * dx * dy / (w2 * w2));
    return sqrt(uxx * uxx + uyy * uyy + 2 * uxy * uxy);
  }
  if (kind == 4) {
    const double X = x[0], Y = x[1], Z = x[2];
    const double sn = sin(50 * X * Z), cs = cos(50 * X * Z), s5 = sin(5 * Y), c5 = cos(5 * Y);
    double h[3][3] = {{-250 * Z * Z * sn, 0, 5 * cs - 250 * X * Z * sn}, {0, 0, 0}, {5 * cs - 250 * X * Z * sn, 0, -250 * X * X * sn}};
    const double q = s5 - 2 * X * Z, D = 0.01 + q * q, g = -0.1 / D, gp = 0.2 * q / (D * D);
    const double gq[3] = {-2 * Z, 5 * c5, -2 * X};
    double hh = 0;
    for (int a = 0; a < 3; ++a) for (int b = 0; b < 3; ++b) {
      double v = h[a][b] + gp * gq[a] * gq[b];
      if (a == 1 && b == 1) v += g * (-25 * s5);
      if ((a == 0 && b == 2) || (a == 2 && b == 0)) v += g * (-2.0);
      hh += v * v;
    }
    return hh;
  }
  return 1.0;
}

/* ---------------------------------------------------------------- one relaxation iteration */
static void grad_on_cell(const orc_ma* o, int64_t c, const double* phi, double* gk) {
  const int d = o->dim, d1 = d + 1;
  const int32_t* cv = o->cells + c * d1; const double* g = o->grad + c * d1 * d;
  for (int a = 0; a < d; ++a) { double s = 0; for (int j = 0; j < d1; ++j) s += phi[cv[j]] * g[j * d + a]; gk[a] = s; }
}

/* trace4 = {volume ratio, coefficient of variation, residual, theta}; iters3 = {l2, poisson, hessian} CG iterations.
 * do_update = 0 evaluates only the first half (diagnostics of the current state). */
int orc_ma_step(orc_ma* o, int mon_kind, const double* mon_params, double dt, double inner_rtol, int do_update,
                double* trace4, int* iters3, int64_t* tangled) {
  const int d = o->dim, d1 = d + 1, dd = d * d;
  const int64_t V = o->V, C = o->C;
  /* A1: b_a,i = sum_K vol/(d+1) (grad phi_old)_a ; M g = b with Dirichlet mask */
#pragma omp parallel for schedule(static)
  for (int64_t v = 0; v < V; ++v) {
    double acc[3] = {0, 0, 0}, gk[3];
    for (int64_t t = o->v2c_ptr[v]; t < o->v2c_ptr[v + 1]; ++t) {
      const int64_t c = o->v2c[t] / d1;
      grad_on_cell(o, c, o->phi_old, gk);
      for (int a = 0; a < d; ++a) acc[a] += o->vol[c] / d1 * gk[a];
    }
    for (int a = 0; a < d; ++a) o->rhs[v * d + a] = acc[a];
  }
  iters3[0] = pcg(o, o->M, o->dinvM, d, o->rhs, o->g, o->bcmask, inner_rtol, 100000, 0);
  /* A2 + A3: x = xi + g ; monitor (normalised kinds need the global max) */
  double mmax = 0;
#pragma omp parallel for reduction(max : mmax) schedule(static)
  for (int64_t v = 0; v < V; ++v) {
    for (int a = 0; a < d; ++a) o->x[v * d + a] = o->xi[v * d + a] + o->g[v * d + a];
    const double r = monitor_raw(mon_kind, mon_params, d, o->x + v * d);
    o->mon[v] = r; if (r > mmax) mmax = r;
  }
  if (mon_kind == 3 || mon_kind == 4) {
    const double s = (mon_kind == 3 ? mon_params[3] : mon_params[0]) / mmax;
#pragma omp parallel for schedule(static)
    for (int64_t v = 0; v < V; ++v) o->mon[v] = 1.0 + s * o->mon[v];
  }
  /* A9 on the computational mesh */
  int64_t nt = 0;
#pragma omp parallel for reduction(+ : nt) schedule(static)
  for (int64_t c = 0; c < C; ++c) {
    const double r = signed_detJ(d, o->xi, o->cells + c * d1) / signed_detJ(d, o->xi, o->cells + c * d1);
    if (r <= 0) ++nt;
  }
  if (tangled) *tangled = nt;
  /* A4 volume diagnostic */
  double vs = 0, vmax = 0, vmin = 1e300;
#pragma omp parallel for reduction(+ : vs) reduction(max : vmax) reduction(min : vmin) schedule(static)
  for (int64_t c = 0; c < C; ++c) {
    double vol, g[12]; const int32_t* cv = o->cells + c * d1;
    cell_geom(d, o->x, cv, &vol, g);
    double mm = 0; for (int j = 0; j < d1; ++j) mm += o->mon[cv[j]];
    const double v = vol * (mm / d1) / o->vol0[c];
    o->volume[c] = v; vs += v; if (v > vmax) vmax = v; if (v < vmin) vmin = v;
  }
  const double mean = vs / (double)C; double ssd = 0;
#pragma omp parallel for reduction(+ : ssd) schedule(static)
  for (int64_t c = 0; c < C; ++c) ssd += (o->volume[c] - mean) * (o->volume[c] - mean);
  /* A5: a_i = int psi_i m det(I+H_old) */
  double lam[14][4], w[14]; const int Q = quad_rule(d, lam, w);
  double suma = 0;
#pragma omp parallel for reduction(+ : suma) schedule(static)
  for (int64_t v = 0; v < V; ++v) {
    double acc = 0;
    for (int64_t t = o->v2c_ptr[v]; t < o->v2c_ptr[v + 1]; ++t) {
      const int64_t c = o->v2c[t] / d1; const int loc = o->v2c[t] % d1; const int32_t* cv = o->cells + c * d1;
      double cell = 0;
      for (int q = 0; q < Q; ++q) {
        double mq = 0, hq[9] = {0};
        for (int j = 0; j < d1; ++j) {
          mq += lam[q][j] * o->mon[cv[j]];
          const double* Hj = o->H_old + (int64_t)cv[j] * dd;
          for (int k = 0; k < dd; ++k) hq[k] += lam[q][j] * Hj[k];
        }
        cell += w[q] * lam[q][loc] * mq * det_I_plus(d, hq);
      }
      acc += o->vol[c] * cell;
    }
    o->a[v] = acc; suma += acc;
  }
  const double theta = suma / o->total_volume; double r2 = 0;
#pragma omp parallel for reduction(+ : r2) schedule(static)
  for (int64_t v = 0; v < V; ++v) { const double r = o->a[v] - theta * o->ml[v]; o->res[v] = r; r2 += r * r; }
  trace4[0] = vmax / vmin; trace4[1] = sqrt(ssd / (double)C) / mean;
  trace4[2] = sqrt(r2) / (fabs(theta) * sqrt(o->ml_sumsq)); trace4[3] = theta;
  iters3[1] = iters3[2] = 0;
  if (!do_update) return 0;
  /* A7: K dphi = dt * r (constant nullspace) ; phi = phi_old + dphi */
  double* dphi = o->z;
#pragma omp parallel for schedule(static)
  for (int64_t v = 0; v < V; ++v) { o->rhs[v] = dt * o->res[v]; dphi[v] = 0.0; }
  iters3[1] = pcg(o, o->K, o->dinvK, 1, o->rhs, dphi, NULL, inner_rtol, 10000000, 1);
#pragma omp parallel for schedule(static)
  for (int64_t v = 0; v < V; ++v) o->phi[v] = o->phi_old[v] + dphi[v];
  /* A8: M H_ab = -sum_K vol d_b psi_i (grad phi)_a + sum_F area/d t_a n_b */
#pragma omp parallel for schedule(static)
  for (int64_t v = 0; v < V; ++v) {
    double acc[9] = {0}, gk[3];
    for (int64_t t = o->v2c_ptr[v]; t < o->v2c_ptr[v + 1]; ++t) {
      const int64_t c = o->v2c[t] / d1; const int loc = o->v2c[t] % d1;
      grad_on_cell(o, c, o->phi, gk);
      const double* gl = o->grad + c * d1 * d + loc * d;
      for (int a = 0; a < d; ++a) for (int b = 0; b < d; ++b) acc[a * d + b] -= o->vol[c] * gl[b] * gk[a];
    }
    for (int64_t t = o->v2f_ptr[v]; t < o->v2f_ptr[v + 1]; ++t) {
      const int64_t f = o->v2f[t] / d;
      grad_on_cell(o, o->bfcell[f], o->phi, gk);
      const double* n = o->bfn + f * d; double gn = 0;
      for (int a = 0; a < d; ++a) gn += gk[a] * n[a];
      const double wgt = o->bfa[f] / d;
      for (int a = 0; a < d; ++a) { const double ta = gk[a] - gn * n[a]; for (int b = 0; b < d; ++b) acc[a * d + b] += wgt * ta * n[b]; }
    }
    for (int k = 0; k < dd; ++k) o->rhs[v * dd + k] = acc[k];
  }
  iters3[2] = pcg(o, o->M, o->dinvM, dd, o->rhs, o->H, NULL, inner_rtol, 100000, 0);
  memcpy(o->phi_old, o->phi, V * sizeof(double));
  memcpy(o->H_old, o->H, V * dd * sizeof(double));
  return 0;
}

/* field access for the tests: what = 0 x, 1 phi_old, 2 H_old, 3 grad_phi, 4 monitor */
int orc_ma_get(const orc_ma* o, int what, double* dst) {
  const int d = o->dim; const int64_t V = o->V;
  switch (what) {
    case 0: memcpy(dst, o->x, V * d * sizeof(double)); return 0;
    case 1: memcpy(dst, o->phi_old, V * sizeof(double)); return 0;
    case 2: memcpy(dst, o->H_old, V * d * d * sizeof(double)); return 0;
    case 3: memcpy(dst, o->g, V * d * sizeof(double)); return 0;
    case 4: memcpy(dst, o->mon, V * sizeof(double)); return 0;
  }
  return 1;
}
int orc_ma_threads(const orc_ma* o) { return o->nthreads; }
