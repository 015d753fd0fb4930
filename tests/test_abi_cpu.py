"""CPU-side checks of the product boundary: the C-ABI library builds, loads without a GPU, exports
every symbol include/gomoku_b200.h declares, fails loudly without a device, and the product package
never touches the oracle."""
import ctypes as C
import os
import re

import pytest

from conftest import PKG, ROOT


@pytest.fixture(scope="module")
def lib():
    import importlib.util
    spec = importlib.util.spec_from_file_location("gz_build", os.path.join(PKG, "build.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    m.build()
    import engine
    return engine.load_library()


def test_exports_match_header(lib):
    import engine
    hdr = open(os.path.join(ROOT, "include", "gomoku_b200.h")).read()
    declared = set(re.findall(r"\b(gz_[a-z_0-9]+)\s*\(", hdr))
    assert declared == set(engine.EXPORTS)
    for name in declared:
        assert hasattr(lib, name), name


def test_struct_sizes_and_counts(lib):
    import engine
    import oracle as orc
    assert C.sizeof(engine.GzSearchParams) == 72 and C.sizeof(engine.GzConfig) == 24
    assert lib.gz_pv_param_count() == orc.lib().gzo_pv_param_count() == 711591 + 5 * 256
    assert lib.gz_graphnet_param_count() == 267749 and lib.gz_dqn_param_count() == 296673


def test_no_device_fails_loudly(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    import engine
    cfg = engine.GzConfig(0, 4, 8, 0, 0, 0)
    h = C.c_void_p()
    assert lib.gz_create(C.byref(cfg), C.byref(h)) == -2
    assert b"no CPU path" in lib.gz_last_error()
    with pytest.raises(engine.EngineError):
        engine.Engine(4, 8)


def test_product_never_references_oracle():
    for dirpath, _, files in os.walk(PKG):
        if "build" in os.path.relpath(dirpath, PKG).split(os.sep):
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "gomoku_oracle" not in src and "import oracle" not in src, f
                assert "/root/reference" not in src, f
