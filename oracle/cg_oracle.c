/* TEST INFRASTRUCTURE ONLY — not part of the product path.
 *
 * Plain-C restatement of the reference's CG hot path, used as the parity oracle by tests/,
 * by __graft_entry__.smoke() and by bench.py's cpu_baseline leg (kind "port") when the
 * compiled reference (oracle/_ref) is unavailable. Nothing under conjugategradient_b200/ or
 * include/ may call into this file.
 *
 * PINNING: this restatement is checked against the reference's own sources compiled here
 * (oracle/_ref/libref_serial.so, see oracle/ref_bridge.cpp) in tests/test_oracle_pinning.py and
 * against the committed fixtures in tests/golden/ that were generated from that build
 * (tests/golden/make_golden.py). The reference ships no golden vectors of its own.
 *
 * Every function cites the reference lines it follows. Arithmetic is IEEE fp64, sequential,
 * left-to-right, compiled with -ffp-contract=off so no FMA contraction changes the rounding.
 */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* Result[i] = sum_{j=rows[i]}^{rows[i+1]-1} data[j]*vec[columns[j]], sequential per row.
 * reference: src/VectorOperations.hpp:455-462 */
void oracle_spmv(long n, const double *data, const int *columns, const int *rows, const double *vec,
                 double *result) {
  for (long i = 0; i < n; i++) {
    double single_result = 0;
    for (int j = rows[i]; j < rows[i + 1]; j++) single_result += data[j] * vec[columns[j]];
    result[i] = single_result;
  }
}

/* *result += sum x[i]*y[i]   (accumulates: no initialize_to_identity).
 * reference: src/VectorOperations.hpp:300-305 */
void oracle_dot_acc(long n, const double *x, const double *y, double *result) {
  double s = 0;
  for (long i = 0; i < n; i++) s += x[i] * y[i];
  *result = *result + s;
}

/* *result += sum x[i]^2 (no sqrt).  reference: src/VectorOperations.hpp:323-327 */
void oracle_norm_acc(long n, const double *x, double *result) {
  double s = 0;
  for (long i = 0; i < n; i++) s += x[i] * x[i];
  *result = *result + s;
}

/* result = a*x + b*y.  reference: src/VectorOperations.hpp:361-363 */
void oracle_saxpby(long n, const double *x, const double *y, double a, double b, double *result) {
  for (long i = 0; i < n; i++) result[i] = a * x[i] + b * y[i];
}
/* result = x + b*y.  reference: src/VectorOperations.hpp:422-424 (alias-safe: elementwise) */
void oracle_sapbx(long n, const double *x, const double *y, double b, double *result) {
  for (long i = 0; i < n; i++) result[i] = x[i] + b * y[i];
}
/* result = x - b*y.  reference: src/VectorOperations.hpp:391-393 */
void oracle_sambx(long n, const double *x, const double *y, double b, double *result) {
  for (long i = 0; i < n; i++) result[i] = x[i] - b * y[i];
}

/* |sum (b-Ax)^2 / sum x^2|.  reference: src/CG.hpp:463-515 */
double oracle_accuracy(long n, const double *data, const int *columns, const int *rows, const double *x,
                       const double *b, double *res2_out, double *x2_out) {
  double normres = 0, normx = 0;
  for (long i = 0; i < n; i++) {
    double single_result = 0;
    for (int j = rows[i]; j < rows[i + 1]; j++) single_result += data[j] * x[columns[j]];
    double a = b[i] - single_result;
    normres += a * a;
    normx += x[i] * x[i];
  }
  if (res2_out) *res2_out = normres;
  if (x2_out) *x2_out = normx;
  return fabs(normres / normx);
}

/* memoryFootprint.  reference: src/CG.hpp:555-558 */
long oracle_memory_footprint(long n, long nnz) {
  return (2 * nnz + 4 * n) * (long)sizeof(double) + 2 * n * (long)sizeof(int);
}

/* Hestenes–Stiefel CG with the reference's exact stop rule.
 * reference: src/CG.hpp:255-454.  Set-up :313-341 (r = b - A x0, p = r, rxr = r.r);
 * loop body :359-436. Stop rule (:396-404, :436): the check reads rxr BEFORE it is replaced
 * by the new r.r (lag of one iteration), tolerance is absolute, NaN stops, at most max_iter
 * trips (reference: N+1, from `while (counter++ < N && !done)`).
 *
 * x        in: initial guess (zeros when the caller set none, :291-297); out: solution
 * hist     optional, length >= max_iter+1: hist[k] = rxr after k iterations (hist[0] = r0.r0)
 * returns  number of loop trips executed */
long oracle_cg_solve(long n, const double *data, const int *columns, const int *rows, const double *b,
                     double *x, double improvement, long max_iter, double *final_rxr, double *hist) {
  double *helper = (double *)calloc((size_t)n, sizeof(double));
  double *rnext = (double *)calloc((size_t)n, sizeof(double));
  double *p = (double *)calloc((size_t)n, sizeof(double));
  double rxr = 0, value2, value3, alpha, beta;
  int is_done = 0;
  long counter = 0, trips = 0;
  if (max_iter < 0) max_iter = n + 1;

  for (long i = 0; i < n; i++) { /* :324-332 */
    double single_result = 0;
    for (int j = rows[i]; j < rows[i + 1]; j++) single_result += data[j] * x[columns[j]];
    rnext[i] = b[i] - single_result;
    p[i] = rnext[i];
  }
  oracle_dot_acc(n, rnext, rnext, &rxr); /* :341 */
  if (hist) hist[0] = rxr;

  do {
    value2 = 0; value3 = 0; alpha = 0; beta = 0;             /* :365-370 */
    oracle_spmv(n, data, columns, rows, p, helper);           /* :374-375 */
    oracle_dot_acc(n, helper, p, &value2);                    /* :378-379 */
    alpha = rxr / value2;                                     /* :385-386 */
    oracle_sapbx(n, x, p, alpha, x);                          /* :390 */
    oracle_sambx(n, rnext, helper, alpha, rnext);             /* :392-393 */
    if (isnan(rxr) || sqrt(rxr) <= improvement) is_done = 1;  /* :400-403 (old rxr) */
    oracle_dot_acc(n, rnext, rnext, &value3);                 /* :406-407 */
    beta = value3 / rxr;                                      /* :414 */
    rxr = value3;                                             /* :415 */
    oracle_sapbx(n, rnext, p, beta, p);                       /* :418 */
    trips++;
    if (hist) hist[trips] = rxr;
  } while (counter++ < max_iter - 1 && !is_done);             /* :436 with N -> max_iter-1 */

  if (final_rxr) *final_rxr = rxr;
  free(helper); free(rnext); free(p);
  return trips;
}

/* ---- Matrix Market reader, dialect-faithful ------------------------------------------------
 * reference: test/mm_reader.cpp:154-171 (read_file), :45-107 (read_real_coordinate_matrix).
 * Quirks kept on purpose: banner result ignored; the line after the banner is discarded
 * unconditionally (:163-164); '%' lines skipped (:146-152); the size line is consumed but unused
 * (:49-52); triples are read until extraction fails (:62-66); EVERY off-diagonal is mirrored
 * regardless of the banner (:68-74); sort by (row, col) (:76-86); a new row offset is pushed only
 * when the row id increases, so empty rows vanish (:94-104). */
typedef struct { int r, c; double v; } coo_t;
static int coo_cmp(const void *a, const void *b) {
  const coo_t *x = (const coo_t *)a, *y = (const coo_t *)b;
  if (x->r != y->r) return x->r < y->r ? -1 : 1;
  if (x->c != y->c) return x->c < y->c ? -1 : 1;
  return 0;
}
typedef struct { long n, nnz; double *vals; int *cols; int *rows; } oracle_csr_t;

static int skip_line(FILE *f) {
  int ch;
  while ((ch = fgetc(f)) != EOF && ch != '\n') {}
  return ch != EOF;
}

void *oracle_mm_open(const char *path, long *n_out, long *nnz_out) {
  FILE *f = fopen(path, "r");
  if (!f) return NULL;
  skip_line(f);                    /* banner (parse_header result is never used) */
  skip_line(f);                    /* :163-164 unconditional discard */
  for (;;) {                       /* skip_comments */
    int ch = fgetc(f);
    if (ch == '%') skip_line(f);
    else { if (ch != EOF) ungetc(ch, f); break; }
  }
  skip_line(f);                    /* size line: read, values unused */
  size_t cap = 1024, cnt = 0;
  coo_t *co = (coo_t *)malloc(cap * sizeof(coo_t));
  int a, b; double v;
  while (fscanf(f, "%d %d %lf", &a, &b, &v) == 3) {
    if (cnt == cap) { cap *= 2; co = (coo_t *)realloc(co, cap * sizeof(coo_t)); }
    co[cnt].r = a - 1; co[cnt].c = b - 1; co[cnt].v = v; cnt++;
  }
  fclose(f);
  size_t h = cnt;
  for (size_t i = 0; i < h; i++) {
    if (co[i].r != co[i].c) {
      if (cnt == cap) { cap *= 2; co = (coo_t *)realloc(co, cap * sizeof(coo_t)); }
      co[cnt].r = co[i].c; co[cnt].c = co[i].r; co[cnt].v = co[i].v; cnt++;
    }
  }
  qsort(co, cnt, sizeof(coo_t), coo_cmp);
  oracle_csr_t *m = (oracle_csr_t *)calloc(1, sizeof(oracle_csr_t));
  m->vals = (double *)malloc((cnt ? cnt : 1) * sizeof(double));
  m->cols = (int *)malloc((cnt ? cnt : 1) * sizeof(int));
  m->rows = (int *)malloc((cnt + 2) * sizeof(int));
  long nr = 0;
  m->rows[nr++] = 0;
  for (size_t i = 0; i < cnt; i++) {
    m->vals[i] = co[i].v; m->cols[i] = co[i].c;
    if (i > 0 && co[i].r > co[i - 1].r) m->rows[nr++] = (int)i;
  }
  m->rows[nr++] = (int)cnt;
  m->n = nr - 1; m->nnz = (long)cnt;
  free(co);
  *n_out = m->n; *nnz_out = m->nnz;
  return m;
}
void oracle_mm_copy(void *h, double *vals, int *cols, int *rows) {
  oracle_csr_t *m = (oracle_csr_t *)h;
  memcpy(vals, m->vals, sizeof(double) * (size_t)m->nnz);
  memcpy(cols, m->cols, sizeof(int) * (size_t)m->nnz);
  memcpy(rows, m->rows, sizeof(int) * (size_t)(m->n + 1));
}
void oracle_mm_close(void *h) {
  oracle_csr_t *m = (oracle_csr_t *)h;
  if (!m) return;
  free(m->vals); free(m->cols); free(m->rows); free(m);
}
