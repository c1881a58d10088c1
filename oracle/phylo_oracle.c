/*
 * CPU ORACLE (C restatement) — test infrastructure and timed CPU baseline only;
 * never linked into or called from the product path.
 *
 * PARITY UNPINNED: /root/reference ships no likelihood code and no golden values
 * (SURVEY.md §0, §8(c)).  This file restates the analytic two-pass algorithm that
 * the reference's CPU contenders run — physher@b19ff2f "treelikelihood" lower/upper
 * partials with SSE 4-state kernels (configured by
 * /root/reference/flu_H3N2/physher-JC69.template:6-57, invoked at
 * /root/reference/modules/micro.nf:20-24 and /root/reference/modules/macro_flu.nf:104)
 * and BEAGLE's beagleUpdatePartials / beagleUpdatePrePartials /
 * beagleCalculateEdgeDerivatives as driven by bito@cc0806a through torchtree-bito
 * (/root/reference/modules/micro.nf:47-53, /root/reference/modules/macro_flu.nf:119-122).
 * None of those sources is in /root/reference, so the formulas are the published
 * ones (Felsenstein 1981; Yang 1994; Ji et al. 2020) as written in SURVEY.md §8(a):
 *
 *   post_v[k,p,i] = (sum_j P_l,k[i,j] post_l[k,p,j]) * (sum_j P_r,k[i,j] post_r[k,p,j])
 *   pre_root = pi ;  pre_c = P_c^T (pre_v o (P_s post_s))
 *   g_c = sum_p w_p [sum_k w_k r_k pre_c.(Q post_c)] / [sum_k w_k pre_c.post_c]
 *
 * It is validated against oracle/oracle.py (three independent formulations) by
 * tests/test_oracle.py.  4-state models only; patterns are processed in blocks so a
 * block's working set stays in cache, with OpenMP across blocks.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define S 4
#define SCALE_THRESHOLD 6.223015277861142e-61 /* 2^-200 */
#define SCALE_FACTOR 1.6069380442589903e+60   /* 2^+200 */
#define SCALE_LOG 138.62943611198907          /* 200 ln 2 */

static void transition(const double *evec, const double *ivec, const double *evals,
                       double t, double *p /* [S*S] row-major */) {
  double ex[S];
  for (int j = 0; j < S; ++j) ex[j] = exp(evals[j] * t);
  for (int i = 0; i < S; ++i)
    for (int l = 0; l < S; ++l) {
      double acc = 0.0;
      for (int j = 0; j < S; ++j) acc += evec[i * S + j] * ex[j] * ivec[j * S + l];
      p[i * S + l] = acc;
    }
}

static inline void matvec(const double *restrict p, const double *restrict x,
                          double *restrict y) {
  for (int i = 0; i < S; ++i)
    y[i] = p[i * S] * x[0] + p[i * S + 1] * x[1] + p[i * S + 2] * x[2] + p[i * S + 3] * x[3];
}

static inline void matvec_t(const double *restrict p, const double *restrict x,
                            double *restrict y) {
  for (int j = 0; j < S; ++j)
    y[j] = p[j] * x[0] + p[S + j] * x[1] + p[2 * S + j] * x[2] + p[3 * S + j] * x[3];
}

int oracle_version(void) { return 1; }

int oracle_max_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

/*
 * triples: int32 [N-1][3] (node,left,right) post-order; tips 0..N-1, root 2N-2.
 * masks:   uint8 [N][P] IUPAC bit masks (bit i = state i compatible).
 * bl:      double [2N-2] by child node id.   grad: double [2N-2] or NULL (logL only).
 * Returns 0 on success.
 */
int oracle_loglik_grad(int n_tips, int n_pat, int n_cat, const int32_t *triples,
                       const uint8_t *masks, const double *weights, const double *bl,
                       const double *evec, const double *ivec, const double *evals,
                       const double *freqs, const double *rates, const double *cat_w,
                       double *logl_out, double *grad, int n_threads) {
  const int n_nodes = 2 * n_tips - 1;
  const int n_br = n_nodes - 1;
  const int root = n_nodes - 1;
  const int K = n_cat;
  const int BLK = 256;
  const int n_blk = (n_pat + BLK - 1) / BLK;
  if (n_tips < 2 || n_pat < 1 || K < 1) return -1;

  double *pm = (double *)malloc(sizeof(double) * (size_t)n_br * K * S * S);
  double q[S * S];
  if (!pm) return -2;
  for (int i = 0; i < S; ++i)
    for (int l = 0; l < S; ++l) {
      double acc = 0.0;
      for (int j = 0; j < S; ++j) acc += evec[i * S + j] * evals[j] * ivec[j * S + l];
      q[i * S + l] = acc;
    }
  for (int b = 0; b < n_br; ++b)
    for (int k = 0; k < K; ++k)
      transition(evec, ivec, evals, bl[b] * rates[k], pm + ((size_t)b * K + k) * S * S);

#ifdef _OPENMP
  if (n_threads > 0) omp_set_num_threads(n_threads);
#endif
  int n_thr = 1;
#ifdef _OPENMP
  n_thr = n_threads > 0 ? n_threads : omp_get_max_threads();
#endif
  double *grad_thr = NULL;
  if (grad) {
    grad_thr = (double *)calloc((size_t)n_thr * n_br, sizeof(double));
    if (!grad_thr) { free(pm); return -2; }
  }
  double logl = 0.0;
  int err = 0;

#pragma omp parallel reduction(+ : logl)
  {
#ifdef _OPENMP
    const int tid = omp_get_thread_num();
#else
    const int tid = 0;
#endif
    /* per-thread block buffers: [node][k][BLK][S] */
    const size_t node_stride = (size_t)K * BLK * S;
    double *post = (double *)malloc(sizeof(double) * node_stride * n_nodes);
    double *pre = grad ? (double *)malloc(sizeof(double) * node_stride * n_nodes) : NULL;
    int *nscale = (int *)malloc(sizeof(int) * BLK);
    double *gl = grad ? grad_thr + (size_t)tid * n_br : NULL;
    if (!post || (grad && !pre) || !nscale) {
#pragma omp atomic write
      err = -2;
    } else {
#pragma omp for schedule(dynamic, 1)
      for (int blk = 0; blk < n_blk; ++blk) {
        const int p0 = blk * BLK;
        const int np = (n_pat - p0 < BLK) ? n_pat - p0 : BLK;
        memset(nscale, 0, sizeof(int) * BLK);
        /* tips */
        for (int t = 0; t < n_tips; ++t)
          for (int k = 0; k < K; ++k) {
            double *dst = post + t * node_stride + (size_t)k * BLK * S;
            const uint8_t *m = masks + (size_t)t * n_pat + p0;
            for (int p = 0; p < np; ++p)
              for (int i = 0; i < S; ++i) dst[p * S + i] = (double)((m[p] >> i) & 1);
          }
        /* post-order */
        for (int e = 0; e < n_tips - 1; ++e) {
          const int v = triples[3 * e], l = triples[3 * e + 1], r = triples[3 * e + 2];
          for (int k = 0; k < K; ++k) {
            const double *pl = pm + ((size_t)l * K + k) * S * S;
            const double *pr = pm + ((size_t)r * K + k) * S * S;
            const double *xl = post + l * node_stride + (size_t)k * BLK * S;
            const double *xr = post + r * node_stride + (size_t)k * BLK * S;
            double *xv = post + v * node_stride + (size_t)k * BLK * S;
            for (int p = 0; p < np; ++p) {
              double a[S], b[S];
              matvec(pl, xl + p * S, a);
              matvec(pr, xr + p * S, b);
              for (int i = 0; i < S; ++i) xv[p * S + i] = a[i] * b[i];
            }
          }
          /* lazy power-of-two rescale, factor shared across states and categories */
          for (int p = 0; p < np; ++p) {
            double mx = 0.0;
            for (int k = 0; k < K; ++k) {
              const double *xv = post + v * node_stride + (size_t)k * BLK * S + p * S;
              for (int i = 0; i < S; ++i) mx = xv[i] > mx ? xv[i] : mx;
            }
            if (mx < SCALE_THRESHOLD && mx > 0.0) {
              for (int k = 0; k < K; ++k) {
                double *xv = post + v * node_stride + (size_t)k * BLK * S + p * S;
                for (int i = 0; i < S; ++i) xv[i] *= SCALE_FACTOR;
              }
              nscale[p] += 1;
            }
          }
        }
        /* root */
        for (int p = 0; p < np; ++p) {
          double lp = 0.0;
          for (int k = 0; k < K; ++k) {
            const double *xv = post + root * node_stride + (size_t)k * BLK * S + p * S;
            double s = 0.0;
            for (int i = 0; i < S; ++i) s += freqs[i] * xv[i];
            lp += cat_w[k] * s;
          }
          logl += weights[p0 + p] * (log(lp) - nscale[p] * SCALE_LOG);
        }
        if (!grad) continue;
        /* pre-order + edge derivatives */
        for (int k = 0; k < K; ++k) {
          double *xr = pre + root * node_stride + (size_t)k * BLK * S;
          for (int p = 0; p < np; ++p)
            for (int i = 0; i < S; ++i) xr[p * S + i] = freqs[i];
        }
        for (int e = n_tips - 2; e >= 0; --e) {
          const int v = triples[3 * e];
          for (int side = 0; side < 2; ++side) {
            const int c = triples[3 * e + 1 + side], sb = triples[3 * e + 2 - side];
            for (int k = 0; k < K; ++k) {
              const double *pc = pm + ((size_t)c * K + k) * S * S;
              const double *ps = pm + ((size_t)sb * K + k) * S * S;
              const double *xs = post + sb * node_stride + (size_t)k * BLK * S;
              const double *uv = pre + v * node_stride + (size_t)k * BLK * S;
              double *uc = pre + c * node_stride + (size_t)k * BLK * S;
              for (int p = 0; p < np; ++p) {
                double ms[S], qq[S];
                matvec(ps, xs + p * S, ms);
                for (int i = 0; i < S; ++i) qq[i] = uv[p * S + i] * ms[i];
                matvec_t(pc, qq, uc + p * S);
              }
            }
            double g = 0.0;
            for (int p = 0; p < np; ++p) {
              double num = 0.0, den = 0.0, mx = 0.0;
              for (int k = 0; k < K; ++k) {
                const double *uc = pre + c * node_stride + (size_t)k * BLK * S + p * S;
                const double *xc = post + c * node_stride + (size_t)k * BLK * S + p * S;
                double qx[S], a = 0.0, b = 0.0;
                matvec(q, xc, qx);
                for (int i = 0; i < S; ++i) {
                  a += uc[i] * qx[i];
                  b += uc[i] * xc[i];
                  mx = uc[i] > mx ? uc[i] : mx;
                }
                num += cat_w[k] * rates[k] * a;
                den += cat_w[k] * b;
              }
              g += weights[p0 + p] * num / den;
              if (mx < SCALE_THRESHOLD && mx > 0.0)
                for (int k = 0; k < K; ++k) {
                  double *uc = pre + c * node_stride + (size_t)k * BLK * S + p * S;
                  for (int i = 0; i < S; ++i) uc[i] *= SCALE_FACTOR;
                }
            }
            gl[c] += g;
          }
        }
      }
    }
    free(post);
    free(pre);
    free(nscale);
  }
  if (grad && !err) {
    for (int b = 0; b < n_br; ++b) {
      double s = 0.0;
      for (int t = 0; t < n_thr; ++t) s += grad_thr[(size_t)t * n_br + b];
      grad[b] = s;
    }
  }
  free(grad_thr);
  free(pm);
  *logl_out = logl;
  return err;
}
