/*
 * CPU ORACLE for any number of states (C restatement) — test infrastructure and timed CPU
 * baseline only; never linked into or called from the product path.
 *
 * PARITY UNPINNED, like oracle/phylo_oracle.c (whose header names the sources restated): the
 * same analytic two-pass algorithm with the state count S a run-time argument, so the codon
 * configuration (61 states, BASELINE.json configs[4]) can be checked at full size and timed
 * on all host cores.  Tip data are state CODES (0..S-1; anything else = missing), as
 * b2p_set_tip_states takes them for n_states != 4.
 *
 *   post_v[k,p,i] = (sum_j P_l,k[i,j] post_l[k,p,j]) * (sum_j P_r,k[i,j] post_r[k,p,j])
 *   pre_root = pi ;  pre_c = P_c^T (pre_v o (P_s post_s))
 *   g_c = sum_p w_p [sum_k w_k r_k pre_c.(Q P_c post_c)] / [sum_k w_k pre_c.(P_c post_c)]
 *
 * Validated against oracle/oracle.py by tests/test_oracle.py.  Patterns are processed in
 * blocks (OpenMP across blocks); per-pattern rescaling by 2^200 keeps deep trees in range.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define SCALE_THRESHOLD 6.223015277861142e-61 /* 2^-200 */
#define SCALE_FACTOR 1.6069380442589903e+60   /* 2^+200 */
#define SCALE_LOG 138.62943611198907          /* 200 ln 2 */
#define PB 8                                  /* patterns per block */

int oracle_dense_version(void) { return 1; }

/* y = M x for one pattern; M row-major [S][S] */
static inline void matvec(int S, const double *restrict m, const double *restrict x, double *restrict y) {
  for (int i = 0; i < S; ++i) {
    const double *row = m + (size_t)i * S;
    double acc = 0.0;
    for (int j = 0; j < S; ++j) acc += row[j] * x[j];
    y[i] = acc;
  }
}
/* y = M^T x */
static inline void matvec_t(int S, const double *restrict m, const double *restrict x, double *restrict y) {
  for (int j = 0; j < S; ++j) y[j] = 0.0;
  for (int i = 0; i < S; ++i) {
    const double *row = m + (size_t)i * S;
    const double xi = x[i];
    for (int j = 0; j < S; ++j) y[j] += row[j] * xi;
  }
}

/*
 * triples: int32 [N-1][3] (node,left,right) post-order; tips 0..N-1, root 2N-2.
 * codes:   uint8 [N][P] state codes.   bl: double [2N-2] by child node id.
 * evec/ivec: [S][S] row-major, evals/freqs: [S].  grad: double [2N-2] or NULL.
 */
int oracle_dense_loglik_grad(int n_states, int n_tips, int n_pat, int n_cat, const int32_t *triples,
                             const uint8_t *codes, const double *weights, const double *bl,
                             const double *evec, const double *ivec, const double *evals,
                             const double *freqs, const double *rates, const double *cat_w,
                             double *logl_out, double *grad, int n_threads) {
  const int S = n_states, N = n_tips, P = n_pat, K = n_cat;
  const int n_nodes = 2 * N - 1, n_br = 2 * N - 2, root = 2 * N - 2;
  const size_t SS = (size_t)S * S;
  if (S < 2 || N < 2 || P < 1 || K < 1) return 1;
#ifdef _OPENMP
  if (n_threads > 0) omp_set_num_threads(n_threads);
#else
  (void)n_threads;
#endif
  /* transition matrices P[b][k] and Q */
  double *pm = (double *)malloc(sizeof(double) * (size_t)n_br * K * SS);
  double *q = (double *)malloc(sizeof(double) * SS);
  int *left = (int *)malloc(sizeof(int) * n_nodes), *right = (int *)malloc(sizeof(int) * n_nodes);
  int *sib = (int *)malloc(sizeof(int) * n_nodes), *par = (int *)malloc(sizeof(int) * n_nodes);
  if (!pm || !q || !left || !right || !sib || !par) return 2;
  for (int i = 0; i < S; ++i)
    for (int l = 0; l < S; ++l) {
      double acc = 0.0;
      for (int j = 0; j < S; ++j) acc += evec[(size_t)i * S + j] * evals[j] * ivec[(size_t)j * S + l];
      q[(size_t)i * S + l] = acc;
    }
#pragma omp parallel
  {
    double *ex = (double *)malloc(sizeof(double) * S), *tmp = (double *)malloc(sizeof(double) * SS);
#pragma omp for schedule(static)
    for (int bk = 0; bk < n_br * K; ++bk) {
      const int b = bk / K, k = bk % K;
      const double t = bl[b] * rates[k];
      for (int j = 0; j < S; ++j) ex[j] = exp(evals[j] * t);
      for (int i = 0; i < S; ++i)
        for (int j = 0; j < S; ++j) tmp[(size_t)i * S + j] = evec[(size_t)i * S + j] * ex[j];
      double *p = pm + (size_t)bk * SS;
      for (int i = 0; i < S; ++i)
        for (int l = 0; l < S; ++l) p[(size_t)i * S + l] = 0.0;
      for (int i = 0; i < S; ++i)
        for (int j = 0; j < S; ++j) {
          const double a = tmp[(size_t)i * S + j];
          const double *iv = ivec + (size_t)j * S;
          double *row = p + (size_t)i * S;
          for (int l = 0; l < S; ++l) row[l] += a * iv[l];
        }
    }
    free(ex); free(tmp);
  }
  for (int v = 0; v < n_nodes; ++v) left[v] = right[v] = sib[v] = par[v] = -1;
  for (int e = 0; e < N - 1; ++e) {
    const int v = triples[3 * e], l = triples[3 * e + 1], r = triples[3 * e + 2];
    left[v] = l; right[v] = r; sib[l] = r; sib[r] = l; par[l] = par[r] = v;
  }
  const int n_blocks = (P + PB - 1) / PB;
  double logl = 0.0;
  int n_thr = 1;
#ifdef _OPENMP
  n_thr = omp_get_max_threads();
#endif
  double *gacc = grad ? (double *)calloc((size_t)n_thr * n_br, sizeof(double)) : NULL;
  int fail = 0;
#pragma omp parallel reduction(+ : logl)
  {
    int tid = 0;
#ifdef _OPENMP
    tid = omp_get_thread_num();
#endif
    /* per block: m[node] = P_node post_node (message to the parent), post of internal nodes, pre */
    const size_t per_node = (size_t)K * PB * S;
    double *msg = (double *)malloc(sizeof(double) * per_node * n_nodes);
    double *pre = grad ? (double *)malloc(sizeof(double) * per_node * n_nodes) : NULL;
    double *x = (double *)malloc(sizeof(double) * S), *y = (double *)malloc(sizeof(double) * S);
    int *scount = (int *)malloc(sizeof(int) * PB);
    double *site = (double *)malloc(sizeof(double) * PB);
    if (!msg || (grad && !pre) || !x || !y || !scount || !site) fail = 1;
    double *g = gacc ? gacc + (size_t)tid * n_br : NULL;
#pragma omp for schedule(dynamic, 1)
    for (int blk = 0; blk < n_blocks; ++blk) {
      if (fail) continue;
      const int p0 = blk * PB, np = (P - p0 < PB) ? P - p0 : PB;
      for (int i = 0; i < PB; ++i) scount[i] = 0;
      /* tip messages: column `code` of P (ones if missing) */
      for (int t = 0; t < N; ++t)
        for (int k = 0; k < K; ++k) {
          const double *p = pm + ((size_t)t * K + k) * SS;
          for (int i = 0; i < np; ++i) {
            double *m = msg + (size_t)t * per_node + ((size_t)k * PB + i) * S;
            const int c = codes[(size_t)t * P + p0 + i];
            if (c < S) for (int s = 0; s < S; ++s) m[s] = p[(size_t)s * S + c];
            else for (int s = 0; s < S; ++s) m[s] = 1.0;
          }
        }
      /* post-order */
      for (int e = 0; e < N - 1; ++e) {
        const int v = triples[3 * e], l = left[v], r = right[v];
        for (int i = 0; i < np; ++i) {
          double mx = 0.0;
          for (int k = 0; k < K; ++k) {
            const double *ml = msg + (size_t)l * per_node + ((size_t)k * PB + i) * S;
            const double *mr = msg + (size_t)r * per_node + ((size_t)k * PB + i) * S;
            double *xv = msg + (size_t)v * per_node + ((size_t)k * PB + i) * S; /* x_v for now */
            for (int s = 0; s < S; ++s) { xv[s] = ml[s] * mr[s]; if (xv[s] > mx) mx = xv[s]; }
          }
          if (mx < SCALE_THRESHOLD && mx > 0.0) {
            for (int k = 0; k < K; ++k) {
              double *xv = msg + (size_t)v * per_node + ((size_t)k * PB + i) * S;
              for (int s = 0; s < S; ++s) xv[s] *= SCALE_FACTOR;
            }
            scount[i] += 1;
          }
          if (v == root) continue;
          for (int k = 0; k < K; ++k) {
            double *xv = msg + (size_t)v * per_node + ((size_t)k * PB + i) * S;
            memcpy(x, xv, sizeof(double) * S);
            matvec(S, pm + ((size_t)v * K + k) * SS, x, xv);
          }
        }
      }
      /* root: site likelihoods (msg[root] holds x_root) */
      for (int i = 0; i < np; ++i) {
        double lk = 0.0;
        for (int k = 0; k < K; ++k) {
          const double *xr = msg + (size_t)root * per_node + ((size_t)k * PB + i) * S;
          double acc = 0.0;
          for (int s = 0; s < S; ++s) acc += freqs[s] * xr[s];
          lk += cat_w[k] * acc;
        }
        site[i] = lk;
        logl += weights[p0 + i] * (log(lk) - scount[i] * SCALE_LOG);
      }
      if (!grad) continue;
      /* pre-order: pre[root] = pi; pre_c = P_c^T (pre_v o msg_sibling); gradient per branch */
      for (int k = 0; k < K; ++k)
        for (int i = 0; i < np; ++i)
          memcpy(pre + (size_t)root * per_node + ((size_t)k * PB + i) * S, freqs, sizeof(double) * S);
      for (int e = N - 2; e >= 0; --e) {
        const int v = triples[3 * e];
        for (int side = 0; side < 2; ++side) {
          const int c = side ? right[v] : left[v], s2 = sib[c];
          for (int i = 0; i < np; ++i) {
            double num = 0.0, den = 0.0, mx = 0.0;
            for (int k = 0; k < K; ++k) {
              const double *u = pre + (size_t)v * per_node + ((size_t)k * PB + i) * S;
              const double *ms = msg + (size_t)s2 * per_node + ((size_t)k * PB + i) * S;
              const double *mc = msg + (size_t)c * per_node + ((size_t)k * PB + i) * S;
              for (int s = 0; s < S; ++s) x[s] = u[s] * ms[s]; /* q_c */
              double d = 0.0;
              for (int s = 0; s < S; ++s) d += x[s] * mc[s];
              matvec(S, q, mc, y); /* Q (P_c post_c) */
              double nn = 0.0;
              for (int s = 0; s < S; ++s) nn += x[s] * y[s];
              den += cat_w[k] * d;
              num += cat_w[k] * rates[k] * nn;
              if (c >= N) {
                double *uc = pre + (size_t)c * per_node + ((size_t)k * PB + i) * S;
                matvec_t(S, pm + ((size_t)c * K + k) * SS, x, uc);
                for (int s = 0; s < S; ++s) if (uc[s] > mx) mx = uc[s];
              }
            }
            if (den > 0.0) g[c] += weights[p0 + i] * num / den;
            if (c >= N && mx < SCALE_THRESHOLD && mx > 0.0)
              for (int k = 0; k < K; ++k) {
                double *uc = pre + (size_t)c * per_node + ((size_t)k * PB + i) * S;
                for (int s = 0; s < S; ++s) uc[s] *= SCALE_FACTOR;
              }
          }
        }
      }
    }
    free(msg); free(pre); free(x); free(y); free(scount); free(site);
  }
  if (grad) {
    for (int b = 0; b < n_br; ++b) {
      double acc = 0.0;
      for (int t = 0; t < n_thr; ++t) acc += gacc[(size_t)t * n_br + b];
      grad[b] = acc;
    }
    free(gacc);
  }
  *logl_out = logl;
  free(pm); free(q); free(left); free(right); free(sib); free(par);
  return fail ? 2 : 0;
}
