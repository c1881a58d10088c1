"""Among-site rate heterogeneity (row a7).

``ConstantSiteModel`` is what the benchmark uses
(/root/reference/flu_H3N2/phylotorch-JC69.template:45-48); ``GammaSiteModel`` is
Yang's (1994) discretised Gamma(alpha, alpha) with K equal-weight categories and
mean-of-bin rates, normalised to mean 1 (alpha=0.5, K=4 →
0.0334, 0.2519, 0.8203, 2.8944).
"""
from __future__ import annotations

import numpy as np

__all__ = ["ConstantSiteModel", "GammaSiteModel", "WeibullSiteModel"]


class ConstantSiteModel:
    def __init__(self):
        self.rates = np.ones(1)
        self.weights = np.ones(1)


class GammaSiteModel:
    def __init__(self, alpha: float = 0.5, categories: int = 4, median: bool = False):
        from scipy.special import gammainc
        from scipy.stats import gamma

        k = int(categories)
        a = float(alpha)
        if median:
            q = (2.0 * np.arange(k) + 1.0) / (2.0 * k)
            rates = gamma.ppf(q, a, scale=1.0 / a)
            rates = rates / rates.mean()
        else:
            cuts = gamma.ppf(np.arange(1, k) / k, a, scale=1.0 / a)
            upper = np.concatenate([gammainc(a + 1.0, cuts * a), [1.0]])
            lower = np.concatenate([[0.0], upper[:-1]])
            rates = (upper - lower) * k
        self.alpha = a
        self.rates = np.asarray(rates, dtype=np.float64)
        self.weights = np.full(k, 1.0 / k)


class WeibullSiteModel:
    """Discretised Weibull(shape, scale 1) rate heterogeneity as torchtree parameterises it
    (upstream recall): K equal-weight categories at the quantiles of the bin
    midpoints, ``r_k = (-log(1 - (2k+1)/(2K)))^(1/shape)``, normalised to mean 1.  The Weibull
    quantile is closed-form, which is why autodiff frameworks prefer it to the Gamma."""

    def __init__(self, shape: float = 1.0, categories: int = 4):
        k = int(categories)
        a = float(shape)
        q = (2.0 * np.arange(k) + 1.0) / (2.0 * k)
        rates = np.power(-np.log1p(-q), 1.0 / a)
        self.shape = a
        self.rates = rates / rates.mean()
        self.weights = np.full(k, 1.0 / k)
