"""Fixtures derived from the reference's own data files (run in the authoring container, where
/root/reference exists; the GPU box only sees the committed outputs).

    python tests/golden/make_reference_fixtures.py

ref_h3_patterns.npz    site patterns of the 1391-taxon x 1698-site H3 alignment in
                       /root/reference/treetime_validation/resources/beast/bedford_et_al_2015__h3_small_tree.xml
                       (masks uint8 [1391, 1113], weights; IUPAC R/Y/W/K/M/S/D/?/-), SURVEY.md §8(c)
ref_h3n2_tree.npz      /root/reference/treetime_validation/resources/flu_H3N2/H3N2_HA_2011_2013.nwk read by
                       b200phylo.tree.Tree.from_newick: post-order triples + branch lengths (3561 tips,
                       polytomies resolved with 1e-5 branches, zero-length branches kept)
ref_phylotorch_jc69.json   /root/reference/flu_H3N2/phylotorch-JC69.template with its *_TEMPLATE placeholders
                       substituted the way /root/reference/SConstruct:75-122 (prepare_phylotorch, non-libsbn
                       branch) does, for a small simulated data set (24 dated taxa)
"""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "gradient-benchmark_b200"))

from b200phylo.alignment import compress_alignment, read_beast_xml  # noqa: E402
from b200phylo.tree import Tree  # noqa: E402
from b200phylo.workloads import make_workload  # noqa: E402

REF = "/root/reference"
H3_XML = f"{REF}/treetime_validation/resources/beast/bedford_et_al_2015__h3_small_tree.xml"
H3N2_NWK = f"{REF}/treetime_validation/resources/flu_H3N2/H3N2_HA_2011_2013.nwk"
TEMPLATE = f"{REF}/flu_H3N2/phylotorch-JC69.template"


def substituted_template(template_text: str, names, dates, sequences, newick: str, iterations: int = 10) -> str:
    """prepare_phylotorch of /root/reference/SConstruct:75-122 (the branch that inlines the data)."""
    taxa = [{"id": nm, "type": "phylotorch.evolution.taxa.Taxon", "attributes": {"date": float(d)}}
            for nm, d in zip(names, dates)]
    seqs = [{"taxon": nm, "sequence": s} for nm, s in zip(names, sequences)]
    root_shift = max(dates) - min(dates)
    return (template_text.replace("SEQUENCES_TEMPLATE", json.dumps(seqs))
            .replace("TAXA_TEMPLATE", json.dumps(taxa))
            .replace("TREE_TEMPLATE", '"' + newick + '"')
            .replace("ITERATION_TEMPLATE", str(iterations))
            .replace("ROOT_SHIFT_TEMPLATE", str(root_shift))
            .replace("DIM_TEMPLATE", str(len(dates) - 2)))


def main():
    sp = compress_alignment(read_beast_xml(H3_XML))
    np.savez_compressed(os.path.join(HERE, "ref_h3_patterns.npz"), masks=sp.masks, weights=sp.weights,
                        names_sha256=np.frombuffer(hashlib.sha256("\n".join(sp.names).encode()).digest(), dtype=np.uint8))
    print("H3 alignment:", sp.masks.shape, "patterns, weight sum", sp.weights.sum())

    with open(H3N2_NWK) as fh:
        t = Tree.from_newick(fh.read())
    np.savez_compressed(os.path.join(HERE, "ref_h3n2_tree.npz"), post_triples=t.post_triples.astype(np.int32),
                        branch_lengths=t.branch_lengths)
    bl = t.branch_lengths
    print("H3N2 tree:", t.n_tips, "tips,", int((bl == 0).sum()), "zero-length branches,", int(np.isclose(bl, 1e-5).sum()),
          "polytomy-resolution branches")

    w = make_workload("jc69_20x1000", n_tips=24, n_patterns=120)
    n = w.n_tips
    dates = [round(1990.0 + 21.0 * i / (n - 1), 3) for i in range(n)]
    names = [f"A/x{i}_{dates[i]}" for i in range(n)]
    w.tree.names = names
    code = {1: "A", 2: "C", 4: "G", 8: "T"}
    seqs = ["".join(code.get(int(m), "N") for m in np.repeat(w.patterns.masks[i], w.patterns.weights.astype(int))) for i in range(n)]
    with open(TEMPLATE) as fh:
        text = substituted_template(fh.read(), names, dates, seqs, w.tree.to_newick())
    json.loads(text)  # must be valid JSON after substitution
    with open(os.path.join(HERE, "ref_phylotorch_jc69.json"), "w") as fh:
        fh.write(text)
    print("template:", len(text), "bytes")


if __name__ == "__main__":
    main()
