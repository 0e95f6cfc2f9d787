"""TEST INFRASTRUCTURE - not part of the product.

Plain-Python restatement (small cases only) of GLIA's greedy tree resolution, the checker for
glia_b200_segment_greedy.  Follows, line for line in behaviour:
  genTree / genTreeWithNodePotentials   /root/reference/code/hmt/tree_build.hxx:12-62
  pickTreeNode / resolveTreeGreedy      hmt/tree_greedy.hxx:80-152
  genLabelTransform                     hmt/tree_segment.hxx:24-35
  transformImage(lmap, mask, fill)      util/image.hxx:211-242
  operation()                           hmt/main_segment_greedy.cxx:30-95
Parity: pinned against the compiled reference (oracle/ref_driver.cxx glia_ref_segment_greedy) by
tests/test_oracle_vs_ref.py::test_segment_greedy_matches_reference and the fixtures tests/golden/seg/*.npz.
"""
import numpy as np

FEPS = 2.22e-16  # glia_base.hxx:59


class Tree:
    def __init__(self):
        self.label, self.parent, self.children, self.potential = [], [], [], []

    def add(self, label, children=()):
        self.label.append(int(label)); self.parent.append(-1); self.children.append(list(children)); self.potential.append(1.0)
        return len(self.label) - 1

    def is_leaf(self, i):
        return not self.children[i]

    def descendants(self, i):  # type/tree.hxx:120-124: breadth first below each child in turn
        out = []
        for c in self.children[i]:
            q = [c]
            k = 0
            while k < len(q):
                v = q[k]; k += 1
                out.append(v)
                q.extend(self.children[v])
        return out

    def leaves(self, i):  # type/tree.hxx:154-162
        return [i] if self.is_leaf(i) else [v for v in self.descendants(i) if self.is_leaf(v)]


def build_tree(order, merge_probs=None):
    t, nmap = Tree(), {}
    for k, (x0, x1, x2) in enumerate(np.asarray(order).tolist()):
        for x in (x0, x1):
            if x not in nmap:
                nmap[x] = t.add(x)
        n0, n1 = nmap[x0], nmap[x1]
        z = t.add(x2, (n0, n1))
        t.parent[n0] = t.parent[n1] = z
        if merge_probs is not None:
            t.potential[z] = float(merge_probs[k])
            ps = 1.0 - float(merge_probs[k])
            for c in (n0, n1):
                t.potential[c] = ps * ps if t.is_leaf(c) else t.potential[c] * ps
        nmap.setdefault(x2, z)
    if merge_probs is not None and t.label:
        t.potential[-1] *= t.potential[-1]
    return t


def segment_greedy(labels, orders, merge_probs=None, region_probs=None, mask=None, ignore=True):
    """Returns (final label volume, picks [(tree, node)], potentials per tree)."""
    trees = []
    for i, o in enumerate(orders):
        t = build_tree(o, None if merge_probs is None else merge_probs[i])
        if region_probs is not None:
            t.potential = [p * max(float(r), FEPS) for p, r in zip(t.potential, region_probs[i])]
        trees.append(t)
    valid = [[True] * len(t.label) for t in trees]
    lnmap = [{t.label[v]: v for v in t.leaves(len(t.label) - 1)} if t.label else {} for t in trees]
    picks = []
    while True:
        best = None
        for i, t in enumerate(trees):  # pickTreeNode: replaced only by a strictly larger potential
            for v in range(len(t.label)):
                if valid[i][v] and (best is None or trees[best[0]].potential[best[1]] < t.potential[v]):
                    best = (i, v)
        if best is None:
            break
        picks.append(best)
        i, v = best
        t = trees[i]
        valid[i][v] = False
        a = t.parent[v]
        while a >= 0:
            valid[i][a] = False; a = t.parent[a]
        ll = []
        for d in t.descendants(v):
            valid[i][d] = False
            if t.is_leaf(d):
                ll.append(t.label[d])
        for lab in ll:
            for j, u in enumerate(trees):
                if j != i and lab in lnmap[j]:
                    w = lnmap[j][lab]
                    valid[j][w] = False
                    a = u.parent[w]
                    while a >= 0:
                        valid[j][a] = False; a = u.parent[a]
    lmap, key = {}, 1
    for i, v in picks:
        for leaf in trees[i].leaves(v):
            lmap[trees[i].label[leaf]] = key
        key += 1
    lab = np.asarray(labels)
    out = lab.copy().ravel()
    flat = lab.ravel()
    m = None if mask is None else np.asarray(mask).ravel()
    for p in range(flat.size):
        if m is not None and m[p] == 0:
            continue
        if int(flat[p]) in lmap:
            out[p] = lmap[int(flat[p])]
        elif ignore:
            out[p] = 0
    return out.reshape(lab.shape), picks, [t.potential for t in trees]
