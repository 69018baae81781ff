"""The C-ABI library builds, loads, and exports every symbol include/hfe.h declares.  No compute; runs without a GPU."""
import ctypes as C
import os
import re

from conftest import ROOT


def declared_functions():
    text = open(os.path.join(ROOT, "include", "hfe.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(hfe_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(cuda_lib, pkg):
    names = declared_functions()
    assert len(names) >= 13 and "hfe_predict_proba" in names
    for name in names:
        assert hasattr(cuda_lib, name), f"{name} is declared in hfe.h but not exported by libhfe.so"
    assert sorted(pkg.capi.EXPORTS) == names
    assert cuda_lib.hfe_abi_version() == 2


def test_bad_blob_is_refused_before_any_device_work(cuda_lib, pkg, published):
    handle = C.c_void_p()
    junk = b"\0" * 128
    assert cuda_lib.hfe_model_create(junk, len(junk), 0, C.byref(handle)) == 2          # HFE_ERR_BAD_BLOB
    assert b"magic" in cuda_lib.hfe_last_error_string()
    blob = bytearray(published["blob"])
    blob[-5] ^= 1
    assert cuda_lib.hfe_model_create(bytes(blob), len(blob), 0, C.byref(handle)) == 2
    assert b"checksum" in cuda_lib.hfe_last_error_string()
    assert cuda_lib.hfe_model_create(None, 0, 0, C.byref(handle)) == 1                  # HFE_ERR_INVALID_ARGUMENT
    assert not handle.value


def test_product_does_not_import_the_oracle():
    pkg_dir = os.path.join(ROOT, "machine-learning-replications_b200")
    for base, _, files in os.walk(pkg_dir):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(base, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), f"{f} imports the oracle"
                assert "stack_oracle" not in src and "sk_rebuild" not in src, f"{f} reaches into the oracle package"
