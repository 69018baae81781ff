"""Static decoder for the published ``hf_predict_model.pkl`` -- no reference class is ever imported or run.

The reference loads its model with ``pickle.load`` (predict_hf.py:33-34).  That file is untrusted input, and it was
written by scikit-learn 0.23.2 whose module layout no longer exists, so loading it the reference's way is neither safe
nor possible.  This module walks the pickle opcode stream with a ``pickle.Unpickler`` whose ``find_class`` hands out
*inert stand-ins* for every GLOBAL: a stand-in records the constructor arguments and the state it is given and does
nothing else.  numpy arrays are rebuilt from the raw ``(version, shape, dtype, fortran, bytes)`` state tuple with
``np.frombuffer`` after the dtype string has been checked against a small whitelist.

The result is a tree of :class:`Inert` nodes whose ``.state`` dictionaries mirror the fitted estimators' ``__dict__``;
``walker.arrays_from_inert`` turns that into the plain ``ModelArrays`` the blob writer consumes.
"""
from __future__ import annotations

import io
import pickle
import re

import numpy as np

# every GLOBAL the published file may name; anything else aborts the decode
ALLOWED_GLOBALS = frozenset({
    ("numpy", "dtype"),
    ("numpy", "ndarray"),
    ("numpy.core.multiarray", "_reconstruct"),
    ("numpy.core.multiarray", "scalar"),
    ("numpy._core.multiarray", "_reconstruct"),
    ("numpy._core.multiarray", "scalar"),
    ("numpy.random._pickle", "__randomstate_ctor"),
    ("sklearn.dummy", "DummyClassifier"),
    ("sklearn.ensemble._gb", "GradientBoostingClassifier"),
    ("sklearn.ensemble._gb_losses", "BinomialDeviance"),
    ("sklearn.ensemble._stacking", "StackingClassifier"),
    ("sklearn.linear_model._logistic", "LogisticRegression"),
    ("sklearn.pipeline", "Pipeline"),
    ("sklearn.preprocessing._data", "StandardScaler"),
    ("sklearn.preprocessing._label", "LabelEncoder"),
    ("sklearn.svm._classes", "SVC"),
    ("sklearn.tree._classes", "DecisionTreeRegressor"),
    ("sklearn.tree._tree", "Tree"),
    ("sklearn.utils", "Bunch"),
})

_SIMPLE_DTYPE = re.compile(r"^[<>|=]?(?:[iuf][1248]|b1|O8?|O)$")
_VOID_DTYPE = re.compile(r"^[<>|=]?V\d+$")
_MAX_ARRAY_BYTES = 64 << 20


class StaticDecodeError(ValueError):
    """The byte stream is not a pickle this decoder is willing to interpret."""


class Inert:
    """Stand-in for one pickled object: remembers what it was called and what state it was handed."""

    def __init__(self, qualname: str, args: tuple = ()):  # noqa: D401
        self.qualname = qualname
        self.args = args
        self.state = None

    def __setstate__(self, state):
        self.state = state

    def get(self, key, default=None):
        if isinstance(self.state, dict):
            return self.state.get(key, default)
        return default

    def __getitem__(self, key):
        if not isinstance(self.state, dict) or key not in self.state:
            raise KeyError(f"{self.qualname} carries no attribute {key!r}")
        return self.state[key]

    def __repr__(self):
        return f"<Inert {self.qualname}>"


class _InertDict(Inert, dict):
    """``sklearn.utils.Bunch`` is a dict subclass; its items arrive through SETITEMS."""

    def __init__(self, qualname: str, args: tuple = ()):
        dict.__init__(self)
        Inert.__init__(self, qualname, args)


class _DType(Inert):
    """``numpy.dtype`` stand-in; finalised into a real (whitelisted) dtype by ``resolve``."""

    def resolve(self) -> np.dtype:
        code = self.args[0] if self.args else None
        if not isinstance(code, str):
            raise StaticDecodeError("dtype constructor without a type string")
        st = self.state
        if _SIMPLE_DTYPE.match(code):
            order = "|"
            if isinstance(st, tuple) and len(st) >= 2 and st[1] in ("<", ">", "|", "="):
                order = st[1]
            base = code.lstrip("<>|=")
            if base in ("O", "O8"):
                return np.dtype(object)
            if order == ">":
                raise StaticDecodeError("big-endian arrays are not expected in this model")
            return np.dtype(base)
        if _VOID_DTYPE.match(code):
            # structured record (tree nodes): state = (ver, order, subdtype, names, fields, itemsize, align, flags)
            if not (isinstance(st, tuple) and len(st) >= 6 and st[3] is not None):
                raise StaticDecodeError("void dtype without field table")
            names, fields, itemsize = st[3], st[4], st[5]
            spec = {"names": [], "formats": [], "offsets": [], "itemsize": int(itemsize)}
            for nm in names:
                sub, off = fields[nm][0], fields[nm][1]
                if not isinstance(nm, str) or not isinstance(sub, _DType):
                    raise StaticDecodeError("malformed record field")
                spec["names"].append(nm)
                spec["formats"].append(sub.resolve())
                spec["offsets"].append(int(off))
            return np.dtype(spec)
        raise StaticDecodeError(f"dtype string {code!r} is outside the whitelist")


class _NDArray(Inert):
    """``numpy.ndarray`` stand-in; ``.array`` is filled in when the BUILD opcode delivers the state tuple."""

    array = None

    def __setstate__(self, state):
        if not (isinstance(state, tuple) and len(state) == 5):
            raise StaticDecodeError("unexpected ndarray state layout")
        _ver, shape, dt, fortran, raw = state
        if not isinstance(dt, _DType):
            raise StaticDecodeError("ndarray state without dtype stand-in")
        dtype = dt.resolve()
        shape = tuple(int(s) for s in shape)
        count = int(np.prod(shape, dtype=np.int64)) if shape else 1
        order = "F" if fortran else "C"
        if dtype == np.dtype(object):
            if not isinstance(raw, list) or len(raw) != count:
                raise StaticDecodeError("object array payload mismatch")
            arr = np.empty(count, dtype=object)
            for i, item in enumerate(raw):
                arr[i] = item
            self.array = arr.reshape(shape, order=order)
            return
        if not isinstance(raw, (bytes, bytearray)):
            raise StaticDecodeError("ndarray payload is not bytes")
        if len(raw) > _MAX_ARRAY_BYTES or len(raw) != count * dtype.itemsize:
            raise StaticDecodeError("ndarray payload size does not match shape*itemsize")
        self.array = np.frombuffer(bytes(raw), dtype=dtype, count=count).reshape(shape, order=order).copy()


def _reconstruct(cls, shape, code):  # numpy.core.multiarray._reconstruct(ndarray, (0,), b'b')
    if cls is not _NDArrayFactory:
        raise StaticDecodeError("_reconstruct asked for something other than ndarray")
    return _NDArray("numpy.ndarray")


def _scalar(dt, raw):  # numpy.core.multiarray.scalar(dtype, bytes)
    if not isinstance(dt, _DType) or not isinstance(raw, (bytes, bytearray)):
        raise StaticDecodeError("malformed numpy scalar")
    dtype = dt.resolve()
    if dtype == np.dtype(object) or len(raw) != dtype.itemsize:
        raise StaticDecodeError("malformed numpy scalar payload")
    return np.frombuffer(bytes(raw), dtype=dtype, count=1)[0]


class _Factory:
    """What ``find_class`` returns: calling it (REDUCE) or ``__new__``-ing it (NEWOBJ) yields an inert node."""

    def __init__(self, qualname, node_cls=Inert):
        self.qualname = qualname
        self.node_cls = node_cls

    def __call__(self, *args):
        return self.node_cls(self.qualname, args)


class _NewobjMeta(type):
    """NEWOBJ runs ``cls.__new__(cls, *args)``, so stand-ins reached that way must be real classes."""


def _make_newobj_class(qualname, dict_like=False):
    base = _InertDict if dict_like else Inert

    class _Node(base):  # type: ignore[misc, valid-type]
        def __new__(cls, *args):
            obj = (dict.__new__(cls) if dict_like else object.__new__(cls))
            base.__init__(obj, qualname, args)
            return obj

        def __init__(self, *args):  # state arrives later through BUILD
            pass

    _Node.__name__ = qualname.rsplit(".", 1)[-1]
    return _Node


_NDArrayFactory = _Factory("numpy.ndarray")


class _StaticUnpickler(pickle.Unpickler):
    def find_class(self, module, name):
        key = (module, name)
        if key not in ALLOWED_GLOBALS:
            raise StaticDecodeError(f"GLOBAL {module}.{name} is not on the allow-list")
        qual = f"{module}.{name}"
        if key == ("numpy", "dtype"):
            return _Factory(qual, _DType)
        if key == ("numpy", "ndarray"):
            return _NDArrayFactory
        if name == "_reconstruct":
            return _reconstruct
        if name == "scalar":
            return _scalar
        if key == ("numpy.random._pickle", "__randomstate_ctor"):
            return _Factory(qual)
        return _make_newobj_class(qual, dict_like=(key == ("sklearn.utils", "Bunch")))

    def persistent_load(self, pid):
        raise StaticDecodeError("persistent ids are not expected in this model")


def decode_pickle_bytes(data: bytes) -> Inert:
    """Decode the pickle into a tree of inert stand-ins.  Executes no code from the modules the pickle names."""
    if len(data) > (32 << 20):
        raise StaticDecodeError("refusing to decode a model file this large")
    root = _StaticUnpickler(io.BytesIO(data)).load()
    if not isinstance(root, Inert):
        raise StaticDecodeError("top-level object is not an estimator stand-in")
    return root


def decode_pickle_file(path: str) -> Inert:
    with open(path, "rb") as fh:
        return decode_pickle_bytes(fh.read())


def unwrap(value):
    """Inert ndarray -> real ndarray (recursing into object arrays is left to the caller)."""
    if isinstance(value, _NDArray):
        if value.array is None:
            raise StaticDecodeError("ndarray stand-in never received its state")
        return value.array
    return value
