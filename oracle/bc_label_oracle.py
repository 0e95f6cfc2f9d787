"""TEST INFRASTRUCTURE - not part of the product.

Plain Python / numpy restatement (small cases only) of GLIA's boundary-classifier label generation from a truth image,
the checker for glia_b200_bc_label.  It follows the reference's definitions voxel set by voxel set - no shortcut
through per-node summaries, so that it checks the summaries the product uses:
  pairStats (contingency map -> TP, TN, FP, FN)      /root/reference/code/util/stats.hxx:186-230
  pairStats over regions, truth BG_VAL excluded       util/image_stats.hxx:173-193
  randIndex / precision / recall / f1                 util/stats.hxx:232-265
  pairF1                                              util/image_stats.hxx:222-245
  genBoundaryClassificationLabelF1 / ...RI            hmt/bc_label.hxx:44-122
  operation(): optSplit, globalOpt 0 / 1 / 2          hmt/main_bc_label_ri.cxx:31-155
  genTree (node numbering)                            hmt/tree_build.hxx:12-38
Python integers are unbounded, as the reference's int512_t is for these purposes.
Parity: pinned against the compiled reference (oracle/ref_driver.cxx glia_ref_bc_label, the reference's templates
instantiated with __int128) by tests/test_oracle_vs_ref.py::test_bc_label_matches_reference and tests/golden/bclabel/.
"""
from collections import Counter

import numpy as np

FEPS = 2.22e-16
BC_LABEL_SPLIT, BC_LABEL_MERGE = 1, -1


def _isfeq(a, b):  # glia_base.hxx:71-72
    return abs(a - b) < FEPS


def pair_stats(regions, truth_flat):
    """regions: list of voxel-index arrays (pRegions).  Returns TP, TN, FP, FN (util/stats.hxx:186-230)."""
    cmap = Counter()
    for i, idx in enumerate(regions):
        t = truth_flat[idx]
        t = t[t != 0]  # excluded = {BG_VAL}
        for key, c in zip(*np.unique(t, return_counts=True)):
            cmap[(i, int(key))] += int(c)
    tp, n = 0, 0
    keys0, keys1 = set(), set()
    for (k0, k1), c in cmap.items():
        n += c
        keys0.add(k0); keys1.add(k1)
        tp += c * (c - 1) // 2
    npair = n * (n - 1) // 2
    same1 = 0
    for k1 in keys1:
        m = sum(cmap.get((k0, k1), 0) for k0 in keys0)
        same1 += m * (m - 1) // 2
    same0 = 0
    for k0 in keys0:
        m = sum(cmap.get((k0, k1), 0) for k1 in keys1)
        same0 += m * (m - 1) // 2
    return tp, npair - same1 + tp - same0, same0 - tp, same1 - tp


def _ratio(num, den):
    den = float(den)
    with np.errstate(all="ignore"):
        return float(np.float64(float(num)) / np.float64(den if not _isfeq(den, 0.0) else den + FEPS))


def pair_f1(regions, truth_flat):
    tp, tn, fp, fn = pair_stats(regions, truth_flat)
    prec, rec = _ratio(tp, tp + fp), _ratio(tp, tp + fn)
    with np.errstate(all="ignore"):
        f = float(np.float64(2.0) * prec * rec / (np.float64(prec) + np.float64(rec)))  # 0 / 0 -> NaN, as in C++
    return f, prec, rec


def rand_index(regions, truth_flat):
    tp, tn, fp, fn = pair_stats(regions, truth_flat)
    num = float(tp + tn)
    den = float(fp + fn) + num
    return num / (den if not _isfeq(den, 0.0) else den + FEPS)


def label_f1(split_regions, merge_regions, truth_flat, tweak, max_prec_drop):
    sf, sp, sr = pair_f1(split_regions, truth_flat)
    mf, mp, mr = pair_f1(merge_regions, truth_flat)
    if max_prec_drop < 1.0 and sp - mp > max_prec_drop:
        return BC_LABEL_SPLIT, mf, sf
    if tweak:
        merge = (mf > sf or (sp < FEPS and sr < FEPS and mp < FEPS and mr < FEPS) or (sf == mf and sp > 0.9 and mp > 0.9))
        return (BC_LABEL_MERGE if merge else BC_LABEL_SPLIT), mf, sf
    return (BC_LABEL_MERGE if mf > sf else BC_LABEL_SPLIT), mf, sf


def bc_labels(labels, truth, order, mask=None, use_f1=True, opt_split=False, global_opt=0, tweak=False, max_prec_drop=1.0):
    """Returns (bc labels [n], merge scores [n], split scores [n]) - scores are F1 (or RI) for global_opt == 0."""
    lab = np.asarray(labels).ravel()
    tr = np.asarray(truth).ravel()
    ok = np.ones(lab.shape, bool) if mask is None else (np.asarray(mask).ravel() != 0)
    vox = {}  # region key -> voxel indices (TRegionMap with the order applied, type/region_map.hxx:67-68)
    for l in np.unique(lab[ok]):
        vox[int(l)] = np.flatnonzero((lab == l) & ok)
    order = np.asarray(order).tolist()
    for x0, x1, x2 in order:
        vox[x2] = np.concatenate([vox[x0], vox[x1]])
    n = len(order)
    mscore, sscore = [0.0] * n, [0.0] * n
    if global_opt == 0:
        out = [0] * n
        if use_f1:
            for i, (x0, x1, x2) in enumerate(order):
                out[i], mscore[i], sscore[i] = label_f1([vox[x0], vox[x1]], [vox[x2]], tr, tweak, max_prec_drop)
            if opt_split:  # main_bc_label_ri.cxx:58-82
                smap = {}
                for i, (x0, x1, x2) in enumerate(order):
                    s0 = smap.setdefault(x0, [vox[x0]])
                    s1 = smap.setdefault(x1, [vox[x1]])
                    split = s0 + s1
                    if out[i] == BC_LABEL_SPLIT:
                        smap[x2] = smap.get(x2, []) + split
                    else:
                        f, _, _ = pair_f1(split, tr)
                        if mscore[i] > f:
                            smap[x2] = smap.get(x2, []) + [vox[x2]]
                        else:
                            out[i] = BC_LABEL_SPLIT
                            smap[x2] = smap.get(x2, []) + split
        else:
            for i, (x0, x1, x2) in enumerate(order):
                sscore[i] = rand_index([vox[x0], vox[x1]], tr)
                mscore[i] = rand_index([vox[x2]], tr)
                out[i] = BC_LABEL_MERGE if mscore[i] > sscore[i] else BC_LABEL_SPLIT
        return np.array(out, np.int32), np.array(mscore), np.array(sscore)
    # global optimum: genTree numbering, best splits bottom-up, then path consistency (:84-146)
    nmap, nodes = {}, []  # node = [label, parent, children, bcLabel, bestSplits]
    for x0, x1, x2 in order:
        for x in (x0, x1):
            if x not in nmap:
                nmap[x] = len(nodes); nodes.append([x, -1, [], BC_LABEL_MERGE, []])
        z = len(nodes)
        nodes.append([x2, -1, [nmap[x0], nmap[x1]], BC_LABEL_MERGE, []])
        nodes[nmap[x0]][1] = nodes[nmap[x1]][1] = z
        nmap.setdefault(x2, z)
    for i, nd in enumerate(nodes):
        if not nd[2]:
            nd[3] = BC_LABEL_MERGE; nd[4] = [i]
        else:
            split = [vox[nodes[j][0]] for c in nd[2] for j in nodes[c][4]]
            mf, _, _ = pair_f1([vox[nd[0]]], tr)
            sf, _, _ = pair_f1(split, tr)
            if mf > sf:
                nd[3] = BC_LABEL_MERGE; nd[4] = [i]
            else:
                nd[3] = BC_LABEL_SPLIT; nd[4] = [j for c in nd[2] for j in nodes[c][4]]
    if global_opt == 1:  # a merge node makes everything below it a merge
        q = [len(nodes) - 1]
        while q:
            ni = q.pop(0)
            if nodes[ni][3] == BC_LABEL_MERGE:
                stack = list(nodes[ni][2])
                while stack:
                    v = stack.pop(); nodes[v][3] = BC_LABEL_MERGE; stack.extend(nodes[v][2])
            else:
                q.extend(nodes[ni][2])
    else:  # a split node makes every ancestor a split
        for i, nd in enumerate(nodes):
            if nd[3] == BC_LABEL_SPLIT:
                a = nd[1]
                while a >= 0:
                    nodes[a][3] = BC_LABEL_SPLIT; a = nodes[a][1]
    return np.array([nd[3] for nd in nodes if nd[2]], np.int32), np.array(mscore), np.array(sscore)
