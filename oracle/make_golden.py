"""Generate the committed fixtures under tests/golden/.  Run HERE (needs /root/reference and scikit-learn):

    python -m oracle.make_golden

TEST INFRASTRUCTURE ONLY.  Produces

* ``published_model.blob``  - static decode of the reference's hf_predict_model.pkl (never unpickled) -> flat blob;
* ``retrained_model.blob``  - the stack refitted by ``oracle/retrain.py`` (replay of train_ensemble_public.py:36-63 on
  a synthetic cohort) walked into the same blob format, plus ``retrained_columns.npy`` (which of the 64 candidate
  columns LassoCV/SelectFromModel kept, in order);
* ``golden_published.npz`` / ``golden_retrained.npz`` - inputs (float32-representable) and what **scikit-learn 1.9.0
  itself** returns for them: ``predict_proba``, ``predict``, ``transform`` and the boosting member's ``apply``.  For
  the published parameters scikit-learn runs on estimators rebuilt from the decoded arrays (``oracle/sk_rebuild.py``);
  for the retrained model it is the live fitted object.

The rows include the three known-answer patients of SURVEY.md Appendix A, a block of synthetic cohort rows, and rows
picked because they sit next to a discontinuity of the path (the libsvm coupling solver's early-exit plateau and
iteration-count edges; ``z_final`` near 0 where the hard label flips).
"""
from __future__ import annotations

import importlib
import os
import tempfile
import warnings

import numpy as np

from . import retrain, sk_rebuild, stack_oracle as so

export = importlib.import_module("machine-learning-replications_b200.export")
synth = importlib.import_module("machine-learning-replications_b200.synth")

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(os.path.dirname(HERE), "tests", "golden")
PUBLISHED_PKL = os.path.join(retrain.REFERENCE_DIR, "hf_predict_model.pkl")

APPENDIX_ROWS = np.array([
    [1, 1, 0, 0, 1, 0, 1, 1, 0, 0, 0, 0, 0, 13, 0, 0, 55],      # predict_hf.py:6-26 defaults
    [0, 0, 0, 0, 0, 0, 1, 0, 0, 0, 0, 0, 0, 18, 0, 0, 63],
    [1, 0, 1, 1, 1, 1, 2, 1, 1, 1, 1, 1, 1, 30, 1, 4, 35],
], dtype=np.float64)


def _edge_rows(model, X, per_edge=24):
    """Rows of X closest to the path's discontinuities, judged with the float64 oracle."""
    picks = []
    svc = next((m for m in model.members if m.kind == "svc"), None)
    if svc is not None:
        r = np.clip(so.sigmoid_predict(so.svc_decision(svc, X), svc.prob_a, svc.prob_b), 1e-7, 1 - 1e-7)
        _, iters = so.coupling_2class(r)
        picks.append(np.argsort(np.abs(r - 0.495))[:per_edge])
        picks.append(np.argsort(np.abs(r - 0.505))[:per_edge])
        picks.append(np.argsort(np.abs(r - 0.5))[:per_edge])
        order = np.argsort(r)
        change = np.nonzero(np.diff(iters[order]) != 0)[0]
        for c in change[:16]:
            picks.append(order[max(c - 2, 0):c + 3])
    z = so.decision_function(model, X)
    picks.append(np.argsort(np.abs(z))[:per_edge])
    return np.unique(np.concatenate(picks))


def _golden_for(model, sk_model, X_rows, seed, specs, path):
    pool = synth.host_rows(specs, seed, 10_000_000, 60_000).astype(np.float64)
    edge = pool[_edge_rows(model, pool)]
    block = synth.host_rows(specs, seed, 0, 1024).astype(np.float64)
    X = np.concatenate([X_rows, block, edge]).astype(np.float32).astype(np.float64)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        proba = sk_model.predict_proba(X)
        label = sk_model.predict(X)
        meta = sk_model.transform(X)
        gbt_idx = next(i for i, m in enumerate(model.members) if m.kind == "gbt")
        leaves = sk_model.estimators_[gbt_idx].apply(X)[:, :, 0]
    assert np.abs(proba - so.predict_proba(model, X)).max() < 1e-12
    assert np.array_equal(label, so.predict(model, X))
    assert np.array_equal(leaves, so.gbt_apply(model.members[gbt_idx], X))
    np.savez_compressed(path, X=X.astype(np.float32), proba=proba, label=label, meta=meta,
                        leaves=leaves.astype(np.uint16), n_appendix=np.int64(X_rows.shape[0]))
    print(f"{os.path.basename(path)}: {X.shape[0]} rows, {edge.shape[0]} edge rows")


def main():
    os.makedirs(GOLDEN, exist_ok=True)
    published = export.arrays_from_published_pickle(PUBLISHED_PKL)
    with open(os.path.join(GOLDEN, "published_model.blob"), "wb") as fh:
        fh.write(export.write_blob(published))
    published = export.read_blob(open(os.path.join(GOLDEN, "published_model.blob"), "rb").read())
    _golden_for(published, sk_rebuild.rebuild_stack(published), APPENDIX_ROWS, 2020, synth.schema_specs(),
                os.path.join(GOLDEN, "golden_published.npz"))

    def risk(X17):
        # outcome model of the synthetic cohort: the published stack's own logit, with the two echocardiographic
        # measurements (wall thickness, ejection fraction) given the strong per-unit effect they need to survive
        # LassoCV's raw-coefficient ranking at train:51-52, as they did on the real cohort
        z = so.decision_function(published, X17) + 0.25 * (X17[:, 13] - 18.6) - 0.20 * (X17[:, 16] - 63.2) - 0.4
        return 1.0 / (1.0 + np.exp(-z))

    with tempfile.TemporaryDirectory() as tmp:
        clf, mask, X_sel, y_sel, names = retrain.replay_training(tmp, risk, verbose=True)
    cols = np.nonzero(mask)[0].astype(np.int32)
    np.save(os.path.join(GOLDEN, "retrained_columns.npy"), cols)
    retrained = export.arrays_from_stack(clf, source="retrain replay")
    with open(os.path.join(GOLDEN, "retrained_model.blob"), "wb") as fh:
        fh.write(export.write_blob(retrained))
    print("retrained: n_sv =", retrained.members[0].sv.shape[0], "columns", cols.tolist())
    specs = [synth.candidate_specs()[c] for c in cols]
    _golden_for(retrained, clf, X_sel[:64], 2020, specs, os.path.join(GOLDEN, "golden_retrained.npz"))


if __name__ == "__main__":
    main()
