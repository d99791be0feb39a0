/* CPU restatement ("port") of the hot path's arithmetic in plain C + OpenMP.  TEST INFRASTRUCTURE ONLY:
 * used by tests/ as a checker for sizes numpy is too slow for, and by bench.py's cpu_baseline /
 * --impl reference legs as the timed CPU baseline.  Never linked into or called by the product.
 *
 * What it restates: the reference hands QuTiP a CSR matrix H (dipolar_ensemble_with_rare.py:570) and zvode
 * calls rhs(t, y) = CSR(-iH) @ y once per evaluation.  At N >~ 20 that CSR cannot be stored
 * (1.2 TB at N = 28), so the same row-by-row gather product is formed matrix-free from the coupling table:
 *   out[x] = d(x) psi[x] + sum_b (hx_b - i s_b hy_b)/2 psi[x^m_b] + sum_{b<c} (Jxx - s_b s_c Jyy)/4 psi[x^m_b^m_c]
 * (appendix A.2).  One call = one CSR matvec's worth of work, restricted to rows [row0, row1).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef struct { double re, im; } cplx;

int oc_max_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

/* The team size of every later parallel region (bench.py's reference arm runs under torchrun, which exports
 * OMP_NUM_THREADS=1 to its workers; the CPU baseline is meant to use all host cores). */
void oc_set_threads(int n) {
#ifdef _OPENMP
  if (n > 0) omp_set_num_threads(n);
#else
  (void)n;
#endif
}

/* diag[x - row0] for x in [row0,row1): sum_b hz_b s_b/2 + sum_{b<c} Jzz_bc s_b s_c/4 */
void oc_diagonal(int n, const double* h, const double* J, uint64_t row0, uint64_t row1, double* diag) {
  const double* hz = h + 2 * n;
  const double* jz = J + 2 * (size_t)n * n;
#pragma omp parallel for schedule(static)
  for (int64_t xi = (int64_t)row0; xi < (int64_t)row1; ++xi) {
    uint64_t x = (uint64_t)xi;
    double acc = 0.0;
    for (int b = 0; b < n; ++b) {
      double sb = (x >> b & 1) ? -1.0 : 1.0;
      acc += 0.5 * hz[b] * sb;
      for (int c = b + 1; c < n; ++c) acc += 0.25 * jz[(size_t)b * n + c] * ((x >> c & 1) ? -sb : sb);
    }
    diag[x - row0] = acc;
  }
}

/* out[x - row0] = (H psi)[x] for x in [row0,row1); psi is the full 2^n vector. */
void oc_apply_h_rows(int n, const double* h, const double* J, const double* diag, const cplx* psi, cplx* out,
                     uint64_t row0, uint64_t row1) {
  const double *hx = h, *hy = h + n;
  const double *jx = J, *jy = J + (size_t)n * n;
  /* compact term lists, as a CSR builder would keep only structural non-zeros */
  int nf = 0, np = 0;
  uint64_t* fm = malloc(sizeof(uint64_t) * n);
  double *fx = malloc(sizeof(double) * n), *fy = malloc(sizeof(double) * n);
  uint64_t* pm = malloc(sizeof(uint64_t) * n * n);
  double *ps = malloc(sizeof(double) * n * n), *pd = malloc(sizeof(double) * n * n);
  for (int b = 0; b < n; ++b) {
    if (hx[b] != 0.0 || hy[b] != 0.0) { fm[nf] = 1ull << b; fx[nf] = 0.5 * hx[b]; fy[nf] = 0.5 * hy[b]; ++nf; }
    for (int c = b + 1; c < n; ++c) {
      double xx = jx[(size_t)b * n + c], yy = jy[(size_t)b * n + c];
      if (xx != 0.0 || yy != 0.0) {
        pm[np] = (1ull << b) | (1ull << c); ps[np] = 0.25 * (xx - yy); pd[np] = 0.25 * (xx + yy); ++np;
      }
    }
  }
#pragma omp parallel for schedule(static)
  for (int64_t xi = (int64_t)row0; xi < (int64_t)row1; ++xi) {
    uint64_t x = (uint64_t)xi;
    double d = diag[x - row0];
    double ar = d * psi[x].re, ai = d * psi[x].im;
    for (int f = 0; f < nf; ++f) {
      cplx u = psi[x ^ fm[f]];
      double cr = fx[f], ci = (x & fm[f]) ? fy[f] : -fy[f];
      ar += cr * u.re - ci * u.im;
      ai += cr * u.im + ci * u.re;
    }
    for (int p = 0; p < np; ++p) {
      double c = (__builtin_popcountll(x & pm[p]) & 1) ? pd[p] : ps[p];
      if (c != 0.0) {
        cplx u = psi[x ^ pm[p]];
        ar += c * u.re;
        ai += c * u.im;
      }
    }
    out[x - row0].re = ar;
    out[x - row0].im = ai;
  }
  free(fm); free(fx); free(fy); free(pm); free(ps); free(pd);
}
