"""Feature schema of the path: cohort ``.mat`` files -> (X, Y, names), and the 17-variable patient record.

Equivalent of the reference's load_data_public.py:4-14 (MATLAB file with ``data_tb`` whose last column is the outcome
and ``clin_var_names``; everything returned as float64) and of the ``Patient_Params`` dictionary order in
predict_hf.py:5-31.
"""
from __future__ import annotations

import numpy as np

from .synth import SCHEMA_NAMES

N_FEATURES = len(SCHEMA_NAMES)


def load_cohort_mat(dataset_path):
    """Same contract as ``load_data`` (load_data_public.py:4-14): returns ``(X float64 [n,F], Y float64 [n], names)``."""
    import scipy.io as sio

    dataset = sio.loadmat(dataset_path)
    table = np.asarray(dataset["data_tb"])
    names = dataset["clin_var_names"]
    return table[:, :-1].astype(np.float64), table[:, -1].astype(np.float64), names


def patient_vector(params: dict) -> np.ndarray:
    """``Patient_Params`` mapping -> ``[1, 17]`` array in predict_hf.py:6-26 order (what predict_hf.py:29-36 builds)."""
    missing = [k for k in SCHEMA_NAMES if k not in params]
    if missing:
        raise ValueError(f"patient record lacks {missing}")
    return np.reshape([params[k] for k in SCHEMA_NAMES], (1, -1))
