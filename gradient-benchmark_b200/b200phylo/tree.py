"""Rooted binary trees: Newick I/O, node numbering, post-order triples, simulators.

Node numbering (the contract of ``include/b200phylo.h``): tips are ``0..N-1`` in
taxon order, internal nodes are ``N..2N-2`` numbered in post-order, so the root is
always ``2N-2`` and a branch is identified by its child node id (``0..2N-3``).
This mirrors how the reference's contenders index branch lengths (one entry per
non-root node; torchtree ``tree_model.branch_lengths()`` of length ``2N-2``,
selected through ``"tree_model"`` in
/root/reference/flu_H3N2/phylotorch-JC69.template:14-44).

Input trees come as Newick with ``name_date`` tip labels
(/root/reference/bin/helper.py:44,52) and may contain polytomies, which the
reference's data preparation breaks into binary splits joined by 1e-5 branches
(/root/reference/treetime_validation/utility_functions_general.py:18-46); the
parser does the same.
"""
from __future__ import annotations

import re
from dataclasses import dataclass, field
from typing import List, Optional, Sequence

import numpy as np

__all__ = ["Tree", "simulate_tree", "parse_newick", "read_lsd_nexus"]


@dataclass
class Tree:
    """Rooted strictly-binary tree in array form."""

    n_tips: int
    children: np.ndarray  # int32 [2N-1, 2]; -1 for tips
    branch_lengths: np.ndarray  # float64 [2N-2], indexed by child node id
    names: List[str] = field(default_factory=list)

    # ------------------------------------------------------------------ basic
    @property
    def n_nodes(self) -> int:
        return 2 * self.n_tips - 1

    @property
    def root(self) -> int:
        return 2 * self.n_tips - 2

    @property
    def parent(self) -> np.ndarray:
        par = np.full(self.n_nodes, -1, dtype=np.int32)
        for v in range(self.n_tips, self.n_nodes):
            par[self.children[v, 0]] = v
            par[self.children[v, 1]] = v
        return par

    @property
    def post_triples(self) -> np.ndarray:
        """int32 [N-1, 3] rows ``(node, left, right)`` in post-order."""
        n = self.n_tips
        out = np.empty((n - 1, 3), dtype=np.int32)
        out[:, 0] = np.arange(n, 2 * n - 1, dtype=np.int32)
        out[:, 1:] = self.children[n:]
        return out

    def levels(self) -> np.ndarray:
        """Height of each node above its deepest tip (tips = 0)."""
        h = np.zeros(self.n_nodes, dtype=np.int32)
        for v in range(self.n_tips, self.n_nodes):
            l, r = self.children[v]
            h[v] = max(h[l], h[r]) + 1
        return h

    def validate(self) -> None:
        n = self.n_tips
        seen = np.zeros(self.n_nodes, dtype=bool)
        for v in range(n, self.n_nodes):
            for c in self.children[v]:
                if not (0 <= c < v or c < n):
                    raise ValueError(f"node {v}: child {c} not in post-order")
                if seen[c]:
                    raise ValueError(f"node {c} has two parents")
                seen[c] = True
        if seen[self.root] or not seen[: self.root].all():
            raise ValueError("tree is not a single rooted binary tree")
        if self.branch_lengths.shape != (2 * n - 2,):
            raise ValueError("branch_lengths must have length 2N-2")

    # ------------------------------------------------------------------ newick
    def to_newick(self, precision: int = 10) -> str:
        names = self.names or [f"t{i}" for i in range(self.n_tips)]
        strs: List[Optional[str]] = [None] * self.n_nodes
        for v in range(self.n_nodes):
            if v < self.n_tips:
                s = names[v]
            else:
                l, r = self.children[v]
                s = f"({strs[l]},{strs[r]})"
                strs[l] = strs[r] = None
            if v != self.root:
                s += ":" + format(float(self.branch_lengths[v]), f".{precision}g")
            strs[v] = s
        return strs[self.root] + ";"

    @classmethod
    def from_newick(cls, text: str, taxa: Optional[Sequence[str]] = None,
                    polytomy_branch: float = 1e-5) -> "Tree":
        return parse_newick(text, taxa=taxa, polytomy_branch=polytomy_branch)

    # ------------------------------------------------------------------ misc
    def scaled(self, factor: float) -> "Tree":
        return Tree(self.n_tips, self.children.copy(), self.branch_lengths * factor, list(self.names))

    def node_heights(self) -> np.ndarray:
        """Heights (time before the most recent tip) implied by branch lengths."""
        depth = np.zeros(self.n_nodes)
        # parents have larger ids than children, so walk internal nodes downwards
        for v in range(self.n_nodes - 1, self.n_tips - 1, -1):
            for c in self.children[v]:
                depth[c] = depth[v] + self.branch_lengths[c]
        return depth.max() - depth


_TOKEN = re.compile(r"\s*(\(|\)|,|;|:|\[[^\]]*\]|'[^']*'|[^\s\(\),:;\[\]]+)")


def parse_newick(text: str, taxa: Optional[Sequence[str]] = None,
                 polytomy_branch: float = 1e-5) -> Tree:
    """Parse one Newick tree; break polytomies; drop unary nodes; renumber."""
    # raw parse into nested lists: node = [children(list) | None, name, length]
    toks = [t for t in _TOKEN.findall(text) if not t.startswith("[")]
    pos = 0

    def new_node():
        return {"ch": [], "name": None, "len": 0.0}

    root = new_node()
    stack = [root]
    cur = root
    expect_len = False
    for t in toks:
        if t == "(":
            child = new_node()
            stack[-1]["ch"].append(child)
            stack.append(child)
            cur = child
        elif t == ",":
            stack.pop()
            child = new_node()
            stack[-1]["ch"].append(child)
            stack.append(child)
            cur = child
        elif t == ")":
            stack.pop()
            cur = stack[-1]
        elif t == ":":
            expect_len = True
        elif t == ";":
            break
        else:
            if expect_len:
                cur["len"] = float(t)
                expect_len = False
            else:
                cur["name"] = t.strip("'")
        pos += 1
    # the outermost "(" made `root` the parent of the real root clade
    if len(root["ch"]) == 1 and root["name"] is None:
        root = root["ch"][0]

    # iterative clean-up: collapse unary nodes, binarise polytomies
    def fix(node):
        work = [node]
        order = []
        while work:
            x = work.pop()
            order.append(x)
            work.extend(x["ch"])
        for x in reversed(order):
            new_ch = []
            for c in x["ch"]:
                while len(c["ch"]) == 1:  # unary: merge into child
                    g = c["ch"][0]
                    g["len"] += c["len"]
                    c = g
                new_ch.append(c)
            while len(new_ch) > 2:  # polytomy: peel off pairs
                a = new_ch.pop()
                b = new_ch.pop()
                j = new_node()
                j["ch"] = [b, a]
                j["len"] = polytomy_branch
                new_ch.append(j)
            x["ch"] = new_ch

    fix(root)
    while len(root["ch"]) == 1:
        root = root["ch"][0]

    # collect tips
    tips = []
    work = [root]
    while work:
        x = work.pop()
        if x["ch"]:
            work.extend(reversed(x["ch"]))
        else:
            tips.append(x)
    names = [t["name"] for t in tips]
    if taxa is not None:
        index = {nm: i for i, nm in enumerate(taxa)}
        missing = [nm for nm in names if nm not in index]
        if missing:
            raise ValueError(f"tips not in taxa: {missing[:5]}")
        names_out = list(taxa)
    else:
        index = {nm: i for i, nm in enumerate(names)}
        names_out = names
    n = len(tips)
    if n < 2:
        raise ValueError("need at least two tips")
    children = np.full((2 * n - 1, 2), -1, dtype=np.int32)
    bl = np.zeros(2 * n - 2)
    next_id = n
    # iterative post-order
    work = [(root, False)]
    while work:
        x, done = work.pop()
        if not x["ch"]:
            x["id"] = index[x["name"]]
        elif not done:
            work.append((x, True))
            work.append((x["ch"][1], False))
            work.append((x["ch"][0], False))
            continue
        else:
            x["id"] = next_id
            children[next_id] = (x["ch"][0]["id"], x["ch"][1]["id"])
            next_id += 1
        if x is not root:
            bl[x["id"]] = max(x["len"], 0.0)
    tree = Tree(n, children, bl, names_out)
    tree.validate()
    return tree


def read_lsd_nexus(path: str, taxa: Optional[Sequence[str]] = None) -> Tree:
    """Dated tree written by lsd2 (``*.date.nexus``): the ``tree 1 = ...`` line with its
    ``[&date="..."]`` node annotations dropped — what ``helper.py convert`` does before the
    contenders see the tree (/root/reference/bin/helper.py:169-176, /root/reference/main.nf:18-40)."""
    newick = None
    with open(path) as fh:
        for line in fh:
            if line.lstrip().lower().startswith("tree 1 = "):
                newick = line.strip()[len("tree 1 = "):].strip()
    if newick is None:
        raise ValueError(f"{path}: no 'tree 1 = ' line")
    newick = re.sub(r'\[&date="\d+\.?\d*"\]', "", newick)
    return parse_newick(newick, taxa=taxa)


def _renumber(n: int, raw_children: dict, raw_len: dict, root) -> Tree:
    """Renumber an arbitrary-id binary tree (tips 0..n-1) into post-order ids."""
    children = np.full((2 * n - 1, 2), -1, dtype=np.int32)
    bl = np.zeros(2 * n - 2)
    new_id = {}
    next_id = n
    work = [(root, False)]
    while work:
        x, done = work.pop()
        if x < n and x not in raw_children:
            new_id[x] = x
        elif not done:
            work.append((x, True))
            l, r = raw_children[x]
            work.append((r, False))
            work.append((l, False))
            continue
        else:
            l, r = raw_children[x]
            new_id[x] = next_id
            children[next_id] = (new_id[l], new_id[r])
            next_id += 1
        if x != root:
            bl[new_id[x]] = raw_len[x]
    return Tree(n, children, bl, [f"t{i}_{2000 + i % 14}" for i in range(n)])


def simulate_tree(n_tips: int, shape: str = "coalescent", seed: int = 20261104,
                  mean_branch: float = 0.02) -> Tree:
    """Seeded synthetic trees.

    shapes: ``coalescent`` (Kingman topology, the flu-like default), ``yule``,
    ``balanced``, ``caterpillar`` (worst-case depth).  Branch lengths are
    Exp(mean_branch) substitutions/site, independent per branch.
    """
    rng = np.random.default_rng(seed)
    n = n_tips
    raw_children = {}
    next_raw = n
    if shape in ("coalescent", "yule"):
        # random-joining topology (Kingman); "yule" = random splitting, same
        # topology distribution on labelled histories but built top-down
        active = list(range(n))
        if shape == "coalescent":
            while len(active) > 1:
                i, j = rng.choice(len(active), size=2, replace=False)
                a, b = active[i], active[j]
                for idx in sorted((i, j), reverse=True):
                    active.pop(idx)
                raw_children[next_raw] = (a, b)
                active.append(next_raw)
                next_raw += 1
            root = active[0]
        else:
            # top-down: split a uniformly chosen leaf until n leaves
            leaves = [("L", 0)]
            kids = {}
            cnt = 1
            while len(leaves) < n:
                k = int(rng.integers(len(leaves)))
                node = leaves.pop(k)
                a, b = ("L", cnt), ("L", cnt + 1)
                cnt += 2
                kids[node] = (a, b)
                leaves += [a, b]
            tip_id = {leaf: i for i, leaf in enumerate(leaves)}
            ids = dict(tip_id)
            # assign raw ids bottom-up
            order = []
            work = [("L", 0)]
            while work:
                x = work.pop()
                order.append(x)
                if x in kids:
                    work.extend(kids[x])
            for x in reversed(order):
                if x in kids:
                    ids[x] = next_raw
                    raw_children[next_raw] = (ids[kids[x][0]], ids[kids[x][1]])
                    next_raw += 1
            root = ids[("L", 0)]
    elif shape == "balanced":
        level = list(range(n))
        while len(level) > 1:
            nxt = []
            for i in range(0, len(level) - 1, 2):
                raw_children[next_raw] = (level[i], level[i + 1])
                nxt.append(next_raw)
                next_raw += 1
            if len(level) % 2:
                nxt.append(level[-1])
            level = nxt
        root = level[0]
    elif shape == "caterpillar":
        cur = 0
        for i in range(1, n):
            raw_children[next_raw] = (cur, i)
            cur = next_raw
            next_raw += 1
        root = cur
    else:
        raise ValueError(f"unknown tree shape {shape!r}")
    raw_len = {x: float(rng.exponential(mean_branch)) for x in range(next_raw)}
    tree = _renumber(n, raw_children, raw_len, root)
    tree.validate()
    return tree
