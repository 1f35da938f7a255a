"""CPU: the C-ABI library loads, exports every symbol include/fcm.h declares, and fails loudly
(no CPU fallback) when no CUDA device is present."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from flag_complex_mcmc_b200 import _native as N

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, "include", "fcm.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(fcm_[a-z_0-9]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    L = N.lib()
    names = header_symbols()
    assert len(names) >= 30
    for name in names:
        assert hasattr(L, name), name
    assert sorted(N.SYMBOLS) == names
    assert L.fcm_abi_version() == 2


def test_rust_shim_declares_every_entry_point():
    """rust/fcm.rs (the binding a maintainer of the reference drops into src/) cannot be compiled here (no rustc): at least
    keep it in step with the header - every function of include/fcm.h is declared exactly once, and nothing else is."""
    rs = open(os.path.join(ROOT, "rust", "fcm.rs")).read()
    declared = re.findall(r"pub fn (fcm_[a-z_0-9]+)\s*\(", rs)
    assert sorted(declared) == header_symbols()
    # argument counts agree with the C declarations
    hdr = re.sub(r"/\*.*?\*/", "", open(os.path.join(ROOT, "include", "fcm.h")).read(), flags=re.S)
    for name in header_symbols():
        c_args = re.search(r"\b%s\s*\(([^;]*?)\)\s*;" % name, hdr, flags=re.S).group(1)
        r_args = re.search(r"pub fn %s\s*\((.*?)\)\s*(->|;)" % name, rs, flags=re.S).group(1)
        nc = 0 if c_args.strip() in ("", "void") else len([a for a in re.sub(r"\([^()]*\)", "", c_args).split(",")])
        nr = 0 if not r_args.strip() else len([a for a in r_args.split(",") if ":" in a])
        assert nc == nr, (name, c_args, r_args)


def test_ctypes_signatures_have_the_headers_argument_counts():
    """the Python host side binds through ctypes: a drifted argtypes list would corrupt a call silently"""
    L = N.lib()
    hdr = re.sub(r"/\*.*?\*/", "", open(os.path.join(ROOT, "include", "fcm.h")).read(), flags=re.S)
    checked = 0
    for name in header_symbols():
        f = getattr(L, name)
        if f.argtypes is None:
            continue
        c_args = re.search(r"\b%s\s*\(([^;]*?)\)\s*;" % name, hdr, flags=re.S).group(1)
        nc = 0 if c_args.strip() in ("", "void") else len(re.sub(r"\([^()]*\)", "", c_args).split(","))
        assert len(f.argtypes) == nc, (name, c_args, f.argtypes)
        checked += 1
    assert checked >= 40


def test_no_cpu_fallback_without_device():
    L = N.lib()
    if L.fcm_device_count() > 0:
        pytest.skip("a CUDA device is present")
    h = C.c_void_p()
    assert L.fcm_ctx_create(0, C.byref(h)) == N.FCM_ENODEVICE
    assert b"no CPU fallback" in L.fcm_last_error()
    e = np.array([[0, 1]], dtype=np.uint32)
    sz = C.c_size_t()
    assert not L.fcm_count_unweighted(2, 1, N.ptr(e), C.byref(sz))
    import flag_complex_mcmc_b200 as fcm
    with pytest.raises(fcm.FcmError):
        fcm.Context(0)
    with pytest.raises(fcm.FcmError):
        fcm.flagser_count(2, e)


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "flag_complex_mcmc_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", ".hpp")) or f == "Makefile":
                txt = open(os.path.join(dirpath, f)).read()
                assert "oracle_binding" not in txt and "libfcm_oracle" not in txt, f
                assert not re.search(r"#include\s*[<\"][^>\"]*oracle", txt), f
