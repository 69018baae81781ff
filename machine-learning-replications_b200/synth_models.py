"""Synthetic model variants for the BASELINE.json configurations that have no published parameters.

* ``stress_stack`` -- "scaled stress ensemble: 2000 trees, depth 16": boosted trees with general node tables far larger
  than shared memory (L2-resident path), stacked with the published logistic member.
* ``without_member`` -- the same stack minus one member kind (e.g. the SVC-free stack, the part of the ensemble that is
  genuinely HBM-bound: SURVEY.md section 8d).

Trees are built as plain ``TreeTable`` arrays (what ``Tree.__getstate__`` holds); the tests hand the same arrays to
scikit-learn through ``Tree.__setstate__``, the GPU gets them through the same blob as any other model.
"""
from __future__ import annotations

import copy

import numpy as np

from .export import GbtMember, ModelArrays, TreeTable

# plausible thresholds per schema column (predict_hf.py:6-26 domains)
_THRESHOLDS = {6: [1.5], 13: [x + 0.5 for x in range(9, 36)], 15: [0.5, 1.5, 2.5, 3.5],
               16: [x + 0.25 for x in range(40, 80)]}


def random_tree(rng, max_depth: int, n_leaves: int, n_features: int = 17, scale: float = 0.05) -> TreeTable:
    """Random unbalanced binary tree: expand random frontier leaves (depth < max_depth) until n_leaves exist."""
    cap = 2 * n_leaves - 1
    left = np.full(cap, -1, np.int32)
    right = np.full(cap, -1, np.int32)
    feat = np.full(cap, -2, np.int32)
    thr = np.full(cap, -2.0)
    depth = np.zeros(cap, np.int32)
    frontier = [0]
    n_nodes, leaves = 1, 1
    while leaves < n_leaves and frontier:
        k = int(rng.integers(len(frontier)))
        frontier[k], frontier[-1] = frontier[-1], frontier[k]
        i = frontier.pop()
        f = int(rng.integers(n_features))
        choices = _THRESHOLDS.get(f, [0.5])
        left[i], right[i], feat[i] = n_nodes, n_nodes + 1, f
        thr[i] = choices[int(rng.integers(len(choices)))]
        depth[n_nodes] = depth[n_nodes + 1] = depth[i] + 1
        if depth[i] + 1 < max_depth:
            frontier += [n_nodes, n_nodes + 1]
        n_nodes += 2
        leaves += 1
    value = rng.normal(0.0, scale, n_nodes)
    return TreeTable(left=left[:n_nodes].copy(), right=right[:n_nodes].copy(), feature=feat[:n_nodes].copy(),
                     threshold=thr[:n_nodes].copy(), value=value)


def stress_forest(n_trees=2000, max_depth=16, n_leaves=512, seed=2020, learning_rate=0.1) -> GbtMember:
    rng = np.random.default_rng(seed)
    trees = [random_tree(rng, max_depth, n_leaves) for _ in range(n_trees)]
    return GbtMember(name="gbc_stress", init_raw=-1.4, learning_rate=learning_rate, trees=trees)


def without_member(arrays: ModelArrays, kind: str) -> ModelArrays:
    keep = [i for i, m in enumerate(arrays.members) if m.kind != kind]
    out = copy.deepcopy(arrays)
    out.members = [out.members[i] for i in keep]
    out.final_coef = np.asarray(arrays.final_coef)[keep].copy()
    out.source = f"{arrays.source} without {kind}"
    return out


def stress_stack(base: ModelArrays, with_svc=False, **kw) -> ModelArrays:
    """Published linear member (+ optionally the SVC) stacked with a 2000 x depth-16 forest."""
    out = copy.deepcopy(base)
    forest = stress_forest(**kw)
    out.members = [forest if m.kind == "gbt" else m for m in out.members]
    if not with_svc:
        out = without_member(out, "svc")
    out.source = "stress forest"
    return out
