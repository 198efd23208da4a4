"""CPU: the drop-in boundary -- the C-ABI library loads without a GPU, exports exactly what
include/adacom_gpu.h declares, fails loudly (no CPU fallback) and never touches oracle/."""
import ctypes as C
import os
import re
import subprocess

import pytest

import duckdb_adaptive_compression_b200 as adacom
from duckdb_adaptive_compression_b200 import _capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "duckdb_adaptive_compression_b200")


def test_library_exports_every_declared_symbol():
    lib = _capi.load()
    declared = _capi.declared_symbols()
    assert len(declared) >= 38
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/adacom_gpu.h but not exported"
    assert set(declared) == set(_capi._PROTOTYPES), "ctypes prototypes out of sync with the header"
    out = subprocess.check_output(["nm", "-D", "--defined-only", _capi.LIB_PATH], text=True)
    exported = set(re.findall(r" T (adacom_[a-z0-9_]+)", out))
    assert exported == set(declared)


def test_abi_version_and_struct_sizes():
    lib = _capi.load()
    assert lib.adacom_abi_version() == 1
    assert C.sizeof(_capi.SegmentMeta) == 64 and C.sizeof(_capi.Agg) == 32


def test_no_cpu_fallback_without_a_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(adacom.AdacomError) as e:
        adacom.Context(0)
    assert e.value.status == _capi.ERR_NO_DEVICE
    assert "no CPU fallback" in str(e.value)


def test_null_handles_are_refused_before_any_cuda_call():
    """argument validation of the entry points added this round needs no device: a NULL context / column is
    ADACOM_ERR_INVALID, never a crash and never a silent success"""
    lib = _capi.load()
    out = C.c_void_p()
    assert lib.adacom_device_alloc(None, 64, C.byref(out)) == _capi.ERR_INVALID
    lib.adacom_device_free(None, None)
    agg = _capi.Agg()
    assert lib.adacom_scan_select(None, 0, 0, 0, None, 0, 0, None, 0, None, None, C.byref(agg)) == _capi.ERR_INVALID
    assert lib.adacom_scan_sum_valid(None, 0, 0, 0, None, 0, C.byref(agg)) == _capi.ERR_INVALID
    assert lib.adacom_fetch_rows(None, 0, None, None, 0, None, 0) == _capi.ERR_INVALID


def test_product_never_references_the_oracle():
    for dirpath, _, files in os.walk(PKG):
        if "build" in dirpath:
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h", "Makefile")):
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "oracle" not in text.lower(), f"{f} mentions the oracle"
    out = subprocess.check_output(["ldd", _capi.LIB_PATH], text=True)
    assert "oracle" not in out and "adacom_ref" not in out


def test_kernels_are_sm100a_with_tma():
    """The shipped cubin is sm_100a and the decode/encode kernels use the TMA bulk-copy engine (UBLKCP)."""
    try:
        sass = subprocess.check_output(["cuobjdump", "-sass", _capi.LIB_PATH], text=True, stderr=subprocess.DEVNULL)
    except (OSError, subprocess.CalledProcessError):
        pytest.skip("cuobjdump not available")
    assert "sm_100a" in sass
    assert "UBLKCP" in sass, "no TMA bulk copy in the SASS"
    assert "SYNCS" in sass, "no mbarrier in the SASS"


def test_segment_layout_helpers():
    import numpy as np
    c = adacom.row_group_segment_counts(1 << 27, np.int64)
    assert len(c) == 4370 and int(c.sum()) == 1 << 27
    assert list(c[:4]) == [32767, 32767, 32767, 24579]
    c = adacom.row_group_segment_counts(200000, np.uint32)
    assert list(c) == [65534, 57346, 65534, 11586]
    assert adacom.segment_capacity(np.int8) == 262136


def test_a_segment_never_has_more_than_16_scan_chunks():
    """kScanParts (csrc/adacom_device.cuh): the table-free scan kernels enumerate <= 16 chunks per segment."""
    worst = 0
    for size in (1, 2, 4, 8):
        cap = 262136 // size
        for w in range(1, 8 * size + 1):
            gran = 2048 if w >= 43 else 128 * (1024 // w)
            worst = max(worst, -(-cap // gran))
    assert worst == 16
