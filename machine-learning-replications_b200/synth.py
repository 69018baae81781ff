"""Synthetic patient rows in the load_data_public.py schema -- host mirror of the ``hfe_synth_rows`` device kernel.

The reference's cohort files (develop_data.mat / model_select_data.mat, train_ensemble_public.py:36,39) are not
published, so every row this repository scores is synthetic.  A value is a pure function of ``(seed, row, column)``:
a splitmix64-style integer hash gives 32 uniform bits ``u``; the column's value is ``base + step * #{i : u >= thr[i]}``
with a short ascending table of u32 thresholds (an inverse-CDF lookup done with integer compares only).  No
floating-point transcendental is involved, so the CUDA kernel and this numpy mirror agree bit for bit and any row of a
10^9-row device-resident matrix can be regenerated on the host for a parity check.

Column distributions follow SURVEY.md section 8(d): the 17 model inputs (predict_hf.py:6-26) use the published scaler
means as Bernoulli rates, NYHA in {1,2}, mitral regurgitation in {0..4}, wall thickness and ejection fraction as
clipped integer-valued normals (cross-checked against the reference's "Table 1.DOCX": wall 19+-5 mm, EF 64+-5 %).
The remaining candidate columns (the cohort has 64 candidates before LassoCV picks 17) are generic binary / integer
covariates.
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import numpy as np

N_CANDIDATES = 64
SCHEMA_NAMES = [
    "Obstructive HCM", "Gender", "Syncope", "Dyspnea", "Fatigue", "Presyncope", "NYHA_Class", "Atrial_Fibrillation",
    "Hypertension", "Beta_blocker", "Ca_Channel_Blockers", "ACEI_ARB", "Coumadin", "Max_Wall_Thick",
    "Septal_Anterior_Motion", "Mitral_Regurgitation", "Ejection_Fraction",
]

_M64 = (1 << 64) - 1
_GOLD = 0x9E3779B97F4A7C15


@dataclass
class ColumnSpec:
    col_id: int          # identity of the column inside the 64-candidate cohort (enters the hash)
    base: float
    step: float
    thresholds: np.ndarray   # ascending u32; value = base + step * count(u >= thr)
    name: str = ""


def _bernoulli(p, base=0.0):
    # P(u >= thr) = p  ->  thr = round((1-p) * 2^32)
    return base, 1.0, np.array([min(max(int(round((1.0 - p) * 2 ** 32)), 1), 2 ** 32 - 1)], dtype=np.uint64)


def _categorical(probs, base=0.0):
    cdf = np.cumsum(np.asarray(probs, dtype=np.float64))
    thr = [min(max(int(round(c * 2 ** 32)), 1), 2 ** 32 - 1) for c in cdf[:-1]]
    return base, 1.0, np.array(thr, dtype=np.uint64)


def _clipped_normal(mean, sd, lo, hi):
    # integer-valued: value v in [lo, hi]; P(value <= v) = Phi((v + 0.5 - mean) / sd)
    thr = []
    for v in range(lo, hi):
        c = 0.5 * (1.0 + math.erf((v + 0.5 - mean) / (sd * math.sqrt(2.0))))
        thr.append(min(max(int(round(c * 2 ** 32)), 1), 2 ** 32 - 1))
    return float(lo), 1.0, np.array(thr, dtype=np.uint64)


def candidate_specs():
    """The 64 candidate columns of the synthetic cohort; the first 17 are the published model's inputs."""
    rates = [0.533, 0.708, 0.097, 0.442, 0.137, 0.056, None, 0.125, 0.331, 0.561, 0.217, 0.229, 0.055, None, 0.682,
             None, None]
    specs = []
    for j, name in enumerate(SCHEMA_NAMES):
        if name == "NYHA_Class":
            b, s, t = _bernoulli(0.442, base=1.0)
        elif name == "Max_Wall_Thick":
            b, s, t = _clipped_normal(18.6, 4.36, 8, 45)
        elif name == "Mitral_Regurgitation":
            b, s, t = _categorical([0.62, 0.27, 0.08, 0.02, 0.01])
        elif name == "Ejection_Fraction":
            b, s, t = _clipped_normal(63.2, 5.23, 20, 80)
        else:
            b, s, t = _bernoulli(rates[j])
        specs.append(ColumnSpec(j, b, s, t, name))
    for j in range(len(SCHEMA_NAMES), N_CANDIDATES):
        k = j - len(SCHEMA_NAMES)
        if k % 6 == 5:
            b, s, t = _clipped_normal(40.0 + 3.0 * (k % 7), 6.0 + (k % 5), 10, 90)
        elif k % 6 == 4:
            b, s, t = _categorical([0.5, 0.25, 0.15, 0.1])
        else:
            b, s, t = _bernoulli(0.08 + 0.017 * ((7 * k) % 47))
        specs.append(ColumnSpec(j, b, s, t, f"cand_{j}"))
    return specs


def schema_specs():
    """Specs of the 17 model inputs in predict_hf.py:6-26 order."""
    return candidate_specs()[:len(SCHEMA_NAMES)]


def hash_u32(seed: int, rows: np.ndarray, col_id: int) -> np.ndarray:
    """32 uniform bits per (seed, row, column); identical to ``hfe_synth_hash`` in csrc/hfe_kernels.cuh."""
    with np.errstate(over="ignore"):
        x = (rows.astype(np.uint64) << np.uint64(8)) | np.uint64(col_id & 0xFF)
        x = x ^ np.uint64((seed * _GOLD) & _M64)
        x = x + np.uint64(_GOLD)
        x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        x = x ^ (x >> np.uint64(31))
    return (x >> np.uint64(32)).astype(np.uint32)


def column_values(spec: ColumnSpec, seed: int, rows: np.ndarray) -> np.ndarray:
    u = hash_u32(seed, rows, spec.col_id).astype(np.uint64)
    count = (u[:, None] >= spec.thresholds[None, :]).sum(axis=1).astype(np.float32)
    return (np.float32(spec.base) + np.float32(spec.step) * count).astype(np.float32)


def host_rows(specs, seed: int, row0: int, n: int, rows: np.ndarray | None = None) -> np.ndarray:
    """Row-major float32 ``[n, len(specs)]`` for rows ``row0 .. row0+n-1`` (or the explicit ``rows`` indices)."""
    idx = np.arange(row0, row0 + n, dtype=np.uint64) if rows is None else np.asarray(rows, dtype=np.uint64)
    return np.stack([column_values(s, seed, idx) for s in specs], axis=1)


def pack_specs(specs):
    """Flatten specs for the C-ABI: ``(col_id i32[F], base f32[F], step f32[F], thr_off i32[F+1], thr u32[])``."""
    col_id = np.array([s.col_id for s in specs], dtype=np.int32)
    base = np.array([s.base for s in specs], dtype=np.float32)
    step = np.array([s.step for s in specs], dtype=np.float32)
    off = np.zeros(len(specs) + 1, dtype=np.int32)
    off[1:] = np.cumsum([len(s.thresholds) for s in specs])
    thr = np.concatenate([s.thresholds for s in specs]).astype(np.uint32)
    return col_id, base, step, off, thr
