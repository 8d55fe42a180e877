"""The C-ABI library loads and exports every symbol include/mbgfe.h declares (no compute without a GPU)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "mbgfe.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(mbgfe_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_the_path():
    syms = declared_symbols()
    for needed in ("mbgfe_create", "mbgfe_add_reads", "mbgfe_add_file", "mbgfe_upload", "mbgfe_process", "mbgfe_run", "mbgfe_data",
                   "mbgfe_count", "mbgfe_exchange", "mbgfe_exchange_group", "mbgfe_comm_init_nccl", "mbgfe_destroy"):
        assert needed in syms


def test_cuda_library_builds_and_exports_all_symbols():
    from verkko_b200 import build
    path = build.build_mbgfe()
    lib = ctypes.CDLL(path)
    for s in declared_symbols():
        assert hasattr(lib, s), f"{s} declared in include/mbgfe.h but not exported by {path}"


def test_no_cpu_fallback_without_gpu():
    """on a machine without a GPU creating a handle must fail loudly, not fall back"""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from verkko_b200 import MbgFrontEnd, MbgfeError
    with pytest.raises(MbgfeError) as e:
        MbgFrontEnd(1001, 100)
    assert e.value.code == -2


def test_parameter_validation_matches_reference_rules():
    # MBG/main.cpp:158-198; validation happens before any device is touched
    from verkko_b200 import MbgFrontEnd, MbgfeError
    for k, w in ((1000, 100), (29, 1), (1001, 0), (1001, 972)):
        with pytest.raises(MbgfeError) as e:
            MbgFrontEnd(k, w)
        assert e.value.code == -1, (k, w)


def test_product_does_not_touch_the_oracle():
    """oracle/ is test infrastructure: nothing under verkko_b200/ or include/ may reference it"""
    for base in ("verkko_b200", "include"):
        for dirpath, _, files in os.walk(os.path.join(ROOT, base)):
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                    text = open(os.path.join(dirpath, f)).read()
                    assert "libmbg_oracle" not in text and "oracle_lib" not in text and "libmbgref" not in text, os.path.join(dirpath, f)


def test_product_does_not_import_torch():
    """the product path is C++ / CUDA behind a ctypes mirror: torch (and torch.distributed) is test and bench plumbing only"""
    for dirpath, _, files in os.walk(os.path.join(ROOT, "verkko_b200")):
        for f in files:
            if f.endswith(".py"):
                text = open(os.path.join(dirpath, f)).read()
                assert "import torch" not in text and "from torch" not in text, os.path.join(dirpath, f)
