"""Reversible nucleotide substitution models → rate matrix + eigen system (row a2).

Selected in the reference by ``"substitution_model": {"type": "JC69"}``
(/root/reference/flu_H3N2/phylotorch-JC69.template:70-73), ``"model": "jc69"``
(/root/reference/flu_H3N2/physher-JC69.template:28) and ``--gtr``
(/root/reference/modules/micro.nf:52).  Q is normalised to one expected
substitution per unit time (``-sum_i pi_i Q_ii = 1``).  The engine consumes the
eigen system ``Q = V diag(lambda) V^-1`` exactly like BEAGLE's
``beagleSetEigenDecomposition``.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

__all__ = ["SubstitutionModel", "JC69", "HKY", "GTR", "GeneralReversible", "CodonGY94"]


@dataclass
class EigenSystem:
    evec: np.ndarray  # [S, S]  V
    ivec: np.ndarray  # [S, S]  V^-1
    evals: np.ndarray  # [S]
    freqs: np.ndarray  # [S]
    q: np.ndarray  # [S, S]


class SubstitutionModel:
    """Base: subclasses provide exchangeabilities (upper triangle) and frequencies."""

    n_states = 4

    def __init__(self, rates, freqs):
        self.freqs = np.asarray(freqs, dtype=np.float64)
        self.freqs = self.freqs / self.freqs.sum()
        self.rates = np.asarray(rates, dtype=np.float64)

    def q_matrix(self) -> np.ndarray:
        s = self.freqs.shape[0]
        r = np.zeros((s, s))
        iu = np.triu_indices(s, 1)
        r[iu] = self.rates
        r = r + r.T
        q = r * self.freqs[None, :]
        np.fill_diagonal(q, 0.0)
        np.fill_diagonal(q, -q.sum(axis=1))
        norm = -(self.freqs * np.diag(q)).sum()
        return q / norm

    def eigen(self) -> EigenSystem:
        """Symmetrised eigendecomposition (real for any reversible model)."""
        q = self.q_matrix()
        sq = np.sqrt(self.freqs)
        sym = (sq[:, None] * q) / sq[None, :]
        sym = 0.5 * (sym + sym.T)
        lam, u = np.linalg.eigh(sym)
        evec = u / sq[:, None]
        ivec = u.T * sq[None, :]
        return EigenSystem(evec, ivec, lam, self.freqs.copy(), q)

    def p_t(self, t: np.ndarray) -> np.ndarray:
        """P(t) for an array of times → [..., S, S] (host-side helper/simulator)."""
        e = self.eigen()
        t = np.asarray(t, dtype=np.float64)
        ex = np.exp(t[..., None] * e.evals)
        return np.einsum("ij,...j,jk->...ik", e.evec, ex, e.ivec)


class JC69(SubstitutionModel):
    def __init__(self):
        super().__init__(np.ones(6), np.full(4, 0.25))


class HKY(SubstitutionModel):
    def __init__(self, kappa: float = 3.0, freqs=(0.25, 0.25, 0.25, 0.25)):
        # order A C G T; transitions A<->G and C<->T
        k = float(kappa)
        super().__init__([1.0, k, 1.0, 1.0, k, 1.0], freqs)
        self.kappa = k


class GTR(SubstitutionModel):
    def __init__(self, rates=(1.0, 1.0, 1.0, 1.0, 1.0, 1.0), freqs=(0.25, 0.25, 0.25, 0.25)):
        super().__init__(rates, freqs)


class GeneralReversible(SubstitutionModel):
    """Any state count (e.g. 61-state codon) from exchangeabilities + frequencies."""

    def __init__(self, rates, freqs):
        super().__init__(rates, freqs)
        self.n_states = len(self.freqs)


class CodonGY94(GeneralReversible):
    """61-state Goldman-Yang-style codon model (BASELINE.json configs[4]): sense codons of the
    standard code; exchangeability 0 for multi-nucleotide changes, ``kappa`` for transitions,
    ``omega`` for non-synonymous changes; codon frequencies from nucleotide frequencies (F1x4)."""

    _AA = "FFLLSSSSYY**CC*WLLLLPPPPHHQQRRRRIIIMTTTTNNKKSSRRVVVVAAAADDEEGGGG"  # TCAG order

    def __init__(self, kappa: float = 2.0, omega: float = 0.5, nuc_freqs=(0.25, 0.25, 0.25, 0.25)):
        nuc = "TCAG"
        pi_n = dict(zip("ACGT", np.asarray(nuc_freqs, dtype=np.float64)))
        codons = [a + b + c for a in nuc for b in nuc for c in nuc]
        sense = [(cd, aa) for cd, aa in zip(codons, self._AA) if aa != "*"]
        self.codons = [cd for cd, _ in sense]
        purines = {"A", "G"}
        n = len(sense)
        rates = []
        for i in range(n):
            for j in range(i + 1, n):
                (ci, ai), (cj, aj) = sense[i], sense[j]
                diff = [(x, y) for x, y in zip(ci, cj) if x != y]
                if len(diff) != 1:
                    rates.append(0.0)
                    continue
                x, y = diff[0]
                r = kappa if (x in purines) == (y in purines) else 1.0
                rates.append(r * (1.0 if ai == aj else omega))
        freqs = [pi_n[c[0]] * pi_n[c[1]] * pi_n[c[2]] for c in self.codons]
        super().__init__(rates, freqs)
        self.kappa, self.omega = float(kappa), float(omega)
