"""Alignments, IUPAC tip masks and site-pattern compression.

The reference selects this step with ``"site_pattern": {"type": "SitePattern",
"alignment": ...}`` (/root/reference/flu_H3N2/phylotorch-JC69.template:59-69) and
``-i <fasta>`` on the micro CLI (/root/reference/modules/micro.nf:20,47).  physher
is configured with ``"tipstates": false`` (tip *partials*,
/root/reference/flu_H3N2/physher-JC69.template:11), so ambiguity codes are IUPAC
state subsets and gaps/N are all-ones; a tip is stored as a 4-bit mask per pattern
(bit i set = state i compatible; A=1 C=2 G=4 T=8).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, List, Sequence, Tuple

import numpy as np

__all__ = ["Alignment", "SitePattern", "read_fasta", "write_fasta", "compress_alignment", "read_beast_xml",
           "simulate_alignment", "simulate_patterns", "IUPAC_MASK"]

_A, _C, _G, _T = 1, 2, 4, 8
IUPAC_MASK: Dict[str, int] = {
    "A": _A, "C": _C, "G": _G, "T": _T, "U": _T,
    "R": _A | _G, "Y": _C | _T, "S": _C | _G, "W": _A | _T, "K": _G | _T, "M": _A | _C,
    "B": _C | _G | _T, "D": _A | _G | _T, "H": _A | _C | _T, "V": _A | _C | _G,
    "N": 15, "-": 15, "?": 15, ".": 15, "X": 15,
}
_LUT = np.full(256, 15, dtype=np.uint8)
for _ch, _m in IUPAC_MASK.items():
    _LUT[ord(_ch)] = _m
    _LUT[ord(_ch.lower())] = _m


@dataclass
class Alignment:
    names: List[str]
    sequences: List[str]

    @property
    def n_taxa(self) -> int:
        return len(self.names)

    @property
    def n_sites(self) -> int:
        return len(self.sequences[0]) if self.sequences else 0

    def masks(self) -> np.ndarray:
        """uint8 [n_taxa, n_sites] IUPAC masks."""
        length = self.n_sites
        out = np.empty((self.n_taxa, length), dtype=np.uint8)
        for i, s in enumerate(self.sequences):
            if len(s) != length:
                raise ValueError(f"sequence {self.names[i]} has length {len(s)} != {length}")
            out[i] = _LUT[np.frombuffer(s.encode("ascii"), dtype=np.uint8)]
        return out


@dataclass
class SitePattern:
    """Unique alignment columns: ``masks`` uint8 [N, P], ``weights`` float64 [P].

    ``n_states`` 4: ``masks`` are IUPAC bit masks.  More states (codon models): ``masks``
    holds the state CODE of each tip (0..n_states-1; anything else = missing)."""

    names: List[str]
    masks: np.ndarray
    weights: np.ndarray
    n_states: int = 4

    @property
    def n_patterns(self) -> int:
        return self.masks.shape[1]

    def partials(self) -> np.ndarray:
        """float64 [N, P, S] tip partials (0/1) — what the oracle consumes."""
        if self.n_states != 4:
            out = (self.masks[..., None] == np.arange(self.n_states, dtype=np.uint8)).astype(np.float64)
            out[self.masks >= self.n_states] = 1.0
            return out
        bits = (self.masks[..., None] >> np.arange(4, dtype=np.uint8)) & 1
        return bits.astype(np.float64)

    def reorder(self, taxa: Sequence[str]) -> "SitePattern":
        index = {nm: i for i, nm in enumerate(self.names)}
        rows = [index[t] for t in taxa]
        return SitePattern(list(taxa), self.masks[rows], self.weights, self.n_states)

    def slice(self, start: int, stop: int) -> "SitePattern":
        """Contiguous pattern shard (multi-GPU sharding)."""
        return SitePattern(self.names, np.ascontiguousarray(self.masks[:, start:stop]),
                           self.weights[start:stop].copy(), self.n_states)


def read_fasta(path: str) -> Alignment:
    names: List[str] = []
    seqs: List[List[str]] = []
    with open(path) as fh:
        for line in fh:
            line = line.strip()
            if not line:
                continue
            if line.startswith(">"):
                names.append(line[1:].split()[0])
                seqs.append([])
            else:
                seqs[-1].append(line)
    return Alignment(names, ["".join(s) for s in seqs])


def read_beast_xml(path: str) -> Alignment:
    """The ``<alignment>`` block of a BEAST 1 XML: ``<sequence><taxon idref="NAME"/> ACGT... </sequence>``
    (the form of /root/reference/treetime_validation/resources/beast/bedford_et_al_2015__h3_small_tree.xml:4184+,
    which SURVEY.md §8(c) lists as reference-held input data for this path)."""
    import re
    with open(path) as fh:
        text = fh.read()
    m = re.search(r"<alignment\b.*?</alignment>", text, flags=re.S)
    if not m:
        raise ValueError(f"{path}: no <alignment> block")
    names, seqs = [], []
    for sm in re.finditer(r"<sequence>\s*<taxon\s+idref=\"([^\"]+)\"\s*/>(.*?)</sequence>", m.group(0), flags=re.S):
        names.append(sm.group(1))
        seqs.append("".join(sm.group(2).split()))
    if not names:
        raise ValueError(f"{path}: <alignment> holds no sequences")
    return Alignment(names, seqs)


def write_fasta(path: str, aln: Alignment) -> None:
    with open(path, "w") as fh:
        for nm, s in zip(aln.names, aln.sequences):
            fh.write(f">{nm}\n{s}\n")


def compress_alignment(aln: Alignment) -> SitePattern:
    """Unique columns + counts, first-occurrence order (deterministic)."""
    m = aln.masks()
    cols, first, counts = np.unique(m.T, axis=0, return_index=True, return_counts=True)
    order = np.argsort(first, kind="stable")
    return SitePattern(list(aln.names), np.ascontiguousarray(cols[order].T),
                       counts[order].astype(np.float64))


def _evolve(tree, p_mats: np.ndarray, freqs: np.ndarray, n_sites: int, rng,
            site_cat: np.ndarray) -> np.ndarray:
    """States uint8 [n_nodes, n_sites] evolved root→tips. p_mats [B, K, S, S]."""
    n_st = p_mats.shape[-1]
    states = np.empty((tree.n_nodes, n_sites), dtype=np.uint8)
    states[tree.root] = rng.choice(n_st, size=n_sites, p=freqs)
    for v in range(tree.n_nodes - 1, tree.n_tips - 1, -1):
        for c in tree.children[v]:
            pm = p_mats[c][site_cat, states[v]]  # [n_sites, 4] row of P for parent state
            cum = np.cumsum(pm, axis=1)
            u = rng.random(n_sites)[:, None]
            states[c] = np.minimum((u > cum).sum(axis=1), n_st - 1).astype(np.uint8)
    return states


def simulate_alignment(tree, p_mats: np.ndarray, freqs: np.ndarray, n_sites: int,
                       seed: int = 20261105, cat_weights=None) -> Alignment:
    """Evolve ACGT-only sequences down ``tree`` (cf. the reference's simulator,
    /root/reference/treetime_validation/utility_functions_simulated_data.py:88-123)."""
    rng = np.random.default_rng(seed)
    k = p_mats.shape[1]
    w = np.full(k, 1.0 / k) if cat_weights is None else np.asarray(cat_weights)
    site_cat = rng.choice(k, size=n_sites, p=w)
    states = _evolve(tree, p_mats, freqs, n_sites, rng, site_cat)[: tree.n_tips]
    alphabet = np.frombuffer(b"ACGT", dtype=np.uint8)
    seqs = [alphabet[row].tobytes().decode("ascii") for row in states]
    names = tree.names or [f"t{i}" for i in range(tree.n_tips)]
    return Alignment(list(names), seqs)


def simulate_patterns(tree, p_mats: np.ndarray, freqs: np.ndarray, n_patterns: int,
                      seed: int = 20261105, cat_weights=None, batch: int = 0) -> SitePattern:
    """Simulate sites until ``n_patterns`` unique columns exist, then truncate
    (configs that name a *pattern* count)."""
    rng = np.random.default_rng(seed)
    k = p_mats.shape[1]
    w = np.full(k, 1.0 / k) if cat_weights is None else np.asarray(cat_weights)
    batch = batch or max(256, n_patterns // 2)
    seen: Dict[bytes, int] = {}
    cols: List[np.ndarray] = []
    counts: List[int] = []
    guard = 0
    while len(cols) < n_patterns:
        site_cat = rng.choice(k, size=batch, p=w)
        st = _evolve(tree, p_mats, freqs, batch, rng, site_cat)[: tree.n_tips]
        for j in range(batch):
            key = st[:, j].tobytes()
            idx = seen.get(key)
            if idx is None:
                if len(cols) >= n_patterns:
                    continue
                seen[key] = len(cols)
                cols.append(st[:, j].copy())
                counts.append(1)
            else:
                counts[idx] += 1
        guard += 1
        if guard > 10000:
            raise RuntimeError("could not reach the requested number of unique patterns")
    n_st = p_mats.shape[-1]
    codes = np.stack(cols, axis=1)
    masks = (1 << codes).astype(np.uint8) if n_st == 4 else codes.astype(np.uint8)
    names = tree.names or [f"t{i}" for i in range(tree.n_tips)]
    return SitePattern(list(names), masks, np.asarray(counts, dtype=np.float64), n_st)
