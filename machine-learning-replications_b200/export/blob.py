"""Flat model blob: the on-disk / over-the-C-ABI replacement for ``hf_predict_model.pkl``.

``ModelArrays`` -> one little-endian byte string that ``hfe_model_create`` (include/hfe.h) validates and uploads.  The
blob holds only the *canonical float64 / int32 parameters* exactly as scikit-learn stores them; every derived table the
kernels read (f32 support-vector rows, thresholds rounded toward -inf, merged stump table, level-ordered node arrays)
is built by the C++ side at create time so that there is one place where rounding decisions are made.

Layout (all integers little-endian, every section 16-byte aligned)::

    0    char[8]  magic "HFEBLOB1"
    8    u32      version (=1)
    12   u32      header_bytes (=64)
    16   u64      total_bytes
    24   u32      crc32 (zlib polynomial) of bytes[header_bytes + 32*n_sections : total_bytes]
    28   u32      n_features
    32   u32      n_members
    36   u32      n_sections
    40   u8[24]   reserved (zero)
    64   section directory: n_sections x { u32 kind, u32 member, u64 offset, u64 nbytes, u64 count }
    ...  payload sections

Section kinds (``member`` = index into the stack's ``estimators_``; 0xFFFFFFFF for stack-level sections):

    1  FINAL          f64[n_members+1]   combiner coef_ ..., intercept_         (linear_model/_base.py:366-396)
    2  CLASSES        f64[2]             classes_
    10 SVC_PARAMS     f64[8]             gamma, intercept(-rho), probA, probB, n_sv, n_support[0], n_support[1], 0
    11 SVC_MEAN       f64[F]             StandardScaler.mean_
    12 SVC_SCALE      f64[F]             StandardScaler.scale_
    13 SVC_SV         f64[n_sv*F]        support_vectors_, row-major
    14 SVC_DUAL       f64[n_sv]          _dual_coef_[0]
    20 GBT_PARAMS     f64[4]             init_raw, learning_rate, n_trees, n_nodes_total
    21 GBT_OFFSETS    u32[n_trees+1]     first node of each tree in the concatenated node arrays
    22 GBT_FEATURE    i32[n_nodes]       split feature, -2 at leaves
    23 GBT_LEFT       i32[n_nodes]       left child (tree-local), -1 at leaves
    24 GBT_RIGHT      i32[n_nodes]       right child (tree-local), -1 at leaves
    25 GBT_THRESHOLD  f64[n_nodes]
    26 GBT_VALUE      f64[n_nodes]       values[:,0,0]
    30 LIN_PARAMS     f64[F+1]           coef_ ..., intercept_
"""
from __future__ import annotations

import struct
import zlib

import numpy as np

from .walker import GbtMember, LinearMember, ModelArrays, ModelShapeError, SvcMember, TreeTable

MAGIC = b"HFEBLOB1"
VERSION = 1
HEADER_BYTES = 64
DIRENT_BYTES = 32
STACK_LEVEL = 0xFFFFFFFF

K_FINAL, K_CLASSES = 1, 2
K_SVC_PARAMS, K_SVC_MEAN, K_SVC_SCALE, K_SVC_SV, K_SVC_DUAL = 10, 11, 12, 13, 14
K_GBT_PARAMS, K_GBT_OFFSETS, K_GBT_FEATURE, K_GBT_LEFT, K_GBT_RIGHT, K_GBT_THRESHOLD, K_GBT_VALUE = \
    20, 21, 22, 23, 24, 25, 26
K_LIN_PARAMS = 30

_DTYPES = {
    K_FINAL: "<f8", K_CLASSES: "<f8", K_SVC_PARAMS: "<f8", K_SVC_MEAN: "<f8", K_SVC_SCALE: "<f8", K_SVC_SV: "<f8",
    K_SVC_DUAL: "<f8", K_GBT_PARAMS: "<f8", K_GBT_OFFSETS: "<u4", K_GBT_FEATURE: "<i4", K_GBT_LEFT: "<i4",
    K_GBT_RIGHT: "<i4", K_GBT_THRESHOLD: "<f8", K_GBT_VALUE: "<f8", K_LIN_PARAMS: "<f8",
}


class BlobError(ValueError):
    """The byte string is not a well-formed model blob."""


def _sections_of(model: ModelArrays):
    yield K_FINAL, STACK_LEVEL, np.concatenate([model.final_coef, [model.final_intercept]])
    yield K_CLASSES, STACK_LEVEL, np.asarray(model.classes, dtype=np.float64)
    for idx, m in enumerate(model.members):
        if isinstance(m, SvcMember):
            yield K_SVC_PARAMS, idx, np.array([m.gamma, m.intercept, m.prob_a, m.prob_b, m.sv.shape[0],
                                               m.n_support[0], m.n_support[1], 0.0])
            yield K_SVC_MEAN, idx, m.mean
            yield K_SVC_SCALE, idx, m.scale
            yield K_SVC_SV, idx, np.ascontiguousarray(m.sv).ravel()
            yield K_SVC_DUAL, idx, m.dual_coef
        elif isinstance(m, GbtMember):
            sizes = np.array([t.left.shape[0] for t in m.trees], dtype=np.int64)
            offs = np.zeros(len(m.trees) + 1, dtype=np.int64)
            np.cumsum(sizes, out=offs[1:])
            if offs[-1] >= 2 ** 31:
                raise ModelShapeError("too many tree nodes for 32-bit offsets")
            yield K_GBT_PARAMS, idx, np.array([m.init_raw, m.learning_rate, len(m.trees), offs[-1]])
            yield K_GBT_OFFSETS, idx, offs.astype("<u4")
            yield K_GBT_FEATURE, idx, np.concatenate([t.feature for t in m.trees]).astype("<i4")
            yield K_GBT_LEFT, idx, np.concatenate([t.left for t in m.trees]).astype("<i4")
            yield K_GBT_RIGHT, idx, np.concatenate([t.right for t in m.trees]).astype("<i4")
            yield K_GBT_THRESHOLD, idx, np.concatenate([t.threshold for t in m.trees])
            yield K_GBT_VALUE, idx, np.concatenate([t.value for t in m.trees])
        elif isinstance(m, LinearMember):
            yield K_LIN_PARAMS, idx, np.concatenate([m.coef, [m.intercept]])
        else:
            raise ModelShapeError(f"unknown member type {type(m).__name__}")


def write_blob(model: ModelArrays) -> bytes:
    """Serialise ``ModelArrays`` into the flat blob ``hfe_model_create`` accepts."""
    sections = []
    for kind, member, arr in _sections_of(model):
        raw = np.ascontiguousarray(arr, dtype=_DTYPES[kind]).tobytes()
        sections.append((kind, member, raw, len(raw) // np.dtype(_DTYPES[kind]).itemsize))
    offset = HEADER_BYTES + DIRENT_BYTES * len(sections)
    offset = (offset + 15) & ~15
    directory, payload = [], bytearray()
    base = offset
    for kind, member, raw, count in sections:
        directory.append(struct.pack("<IIQQQ", kind, member, offset, len(raw), count))
        payload += raw
        pad = (-len(raw)) % 16
        payload += b"\0" * pad
        offset += len(raw) + pad
    dir_bytes = b"".join(directory)
    dir_bytes += b"\0" * (base - HEADER_BYTES - len(dir_bytes))
    crc = zlib.crc32(bytes(payload)) & 0xFFFFFFFF
    header = struct.pack("<8sIIQIIII24x", MAGIC, VERSION, HEADER_BYTES, offset, crc, model.n_features,
                         len(model.members), len(sections))
    assert len(header) == HEADER_BYTES
    return header + dir_bytes + bytes(payload)


def read_blob(data: bytes) -> ModelArrays:
    """Parse and validate a blob back into ``ModelArrays`` (the Python mirror of the C++ loader's checks)."""
    if len(data) < HEADER_BYTES:
        raise BlobError("blob shorter than its header")
    magic, version, hbytes, total, crc, n_features, n_members, n_sections = struct.unpack_from("<8sIIQIIII", data, 0)
    if magic != MAGIC or version != VERSION or hbytes != HEADER_BYTES:
        raise BlobError("bad magic, version or header size")
    if total != len(data):
        raise BlobError("total_bytes does not match the buffer length")
    if not (1 <= n_features <= 4096 and 1 <= n_members <= 16 and 2 <= n_sections <= 256):
        raise BlobError("implausible header counts")
    pay0 = (HEADER_BYTES + DIRENT_BYTES * n_sections + 15) & ~15
    if pay0 > total:
        raise BlobError("section directory runs past the end")
    if (zlib.crc32(data[pay0:]) & 0xFFFFFFFF) != crc:
        raise BlobError("checksum mismatch")
    secs = {}
    for i in range(n_sections):
        kind, member, off, nbytes, count = struct.unpack_from("<IIQQQ", data, HEADER_BYTES + DIRENT_BYTES * i)
        if kind not in _DTYPES:
            raise BlobError(f"unknown section kind {kind}")
        dt = np.dtype(_DTYPES[kind])
        if off % 16 or off < pay0 or off + nbytes > total or nbytes != count * dt.itemsize:
            raise BlobError("section out of bounds or mis-sized")
        if (kind, member) in secs:
            raise BlobError("duplicate section")
        secs[(kind, member)] = np.frombuffer(data, dtype=dt, count=count, offset=off).copy()

    def need(kind, member, count=None):
        arr = secs.get((kind, member))
        if arr is None or (count is not None and arr.shape[0] != count):
            raise BlobError(f"section {kind} of member {member} missing or mis-sized")
        return arr

    final = need(K_FINAL, STACK_LEVEL, n_members + 1)
    classes = need(K_CLASSES, STACK_LEVEL, 2)
    members = []
    for idx in range(n_members):
        if (K_SVC_PARAMS, idx) in secs:
            p = need(K_SVC_PARAMS, idx, 8)
            n_sv = int(p[4])
            if n_sv < 1 or n_sv != p[4] or int(p[5]) + int(p[6]) != n_sv:
                raise BlobError("SVC support counts are inconsistent")
            members.append(SvcMember(
                name=f"m{idx}", mean=need(K_SVC_MEAN, idx, n_features), scale=need(K_SVC_SCALE, idx, n_features),
                gamma=float(p[0]), sv=need(K_SVC_SV, idx, n_sv * n_features).reshape(n_sv, n_features),
                dual_coef=need(K_SVC_DUAL, idx, n_sv), intercept=float(p[1]), prob_a=float(p[2]), prob_b=float(p[3]),
                n_support=np.array([int(p[5]), int(p[6])], dtype=np.int64)))
        elif (K_GBT_PARAMS, idx) in secs:
            p = need(K_GBT_PARAMS, idx, 4)
            n_trees, n_nodes = int(p[2]), int(p[3])
            offs = need(K_GBT_OFFSETS, idx, n_trees + 1).astype(np.int64)
            if offs[0] != 0 or offs[-1] != n_nodes or np.any(np.diff(offs) < 1):
                raise BlobError("tree offsets are inconsistent")
            feat, left, right = (need(k, idx, n_nodes) for k in (K_GBT_FEATURE, K_GBT_LEFT, K_GBT_RIGHT))
            thr, val = need(K_GBT_THRESHOLD, idx, n_nodes), need(K_GBT_VALUE, idx, n_nodes)
            trees = []
            for t in range(n_trees):
                s = slice(offs[t], offs[t + 1])
                trees.append(TreeTable(left=left[s].copy(), right=right[s].copy(), feature=feat[s].copy(),
                                       threshold=thr[s].copy(), value=val[s].copy()))
            members.append(GbtMember(name=f"m{idx}", init_raw=float(p[0]), learning_rate=float(p[1]), trees=trees))
        elif (K_LIN_PARAMS, idx) in secs:
            p = need(K_LIN_PARAMS, idx, n_features + 1)
            members.append(LinearMember(name=f"m{idx}", coef=p[:-1].copy(), intercept=float(p[-1])))
        else:
            raise BlobError(f"member {idx} has no parameter section")
    return ModelArrays(n_features=n_features, members=members, final_coef=final[:-1].copy(),
                       final_intercept=float(final[-1]), classes=classes, source="blob")
