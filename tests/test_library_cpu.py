"""CPU-side checks: the C-ABI library loads and exports every symbol include/ksched_b200.h declares,
refuses to run without a GPU, and the wire/JSON harness parses clusters. No compute calls here."""
import ctypes
import os
import re

import pytest

from kubernetes_b200 import lib, workloads

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _built():
    if not (os.path.exists(lib.ENGINE_SO) and os.path.exists(lib.HOST_SO)):
        lib.build()


def test_header_symbols_are_exported():
    _built()
    header = open(os.path.join(ROOT, "include", "ksched_b200.h")).read()
    declared = set(re.findall(r"\b(ksb_[a-z_]+)\s*\(", header))
    assert declared == set(lib.CABI_SYMBOLS), declared ^ set(lib.CABI_SYMBOLS)
    so = lib.engine()
    for name in declared:
        assert hasattr(so, name), name


def test_create_fails_loudly_without_gpu():
    _built()
    try:
        import torch
        if torch.cuda.is_available():
            pytest.skip("a GPU is present")
    except ImportError:
        pass
    so = lib.engine()
    cfg = lib.KsbConfig()
    so.ksb_config_default(ctypes.byref(cfg))
    assert cfg.w_prefer_avoid == 10000 and cfg.label_words == 4
    err = ctypes.create_string_buffer(512)
    h = so.ksb_create(ctypes.byref(cfg), err, 512)
    assert not h
    assert b"CUDA" in err.value or b"cuda" in err.value


def test_session_parses_cluster_and_reports_missing_gpu():
    _built()
    cluster = workloads.generate(3, 20, 30)
    s = lib.Session(cluster=cluster)
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except ImportError:
        has_gpu = False
    if not has_gpu:
        with pytest.raises(RuntimeError):
            s.run()
    s.close()


def test_generator_is_deterministic_and_shaped():
    a = workloads.generate(3, 50, 80)
    b = workloads.generate(3, 50, 80)
    assert a == b
    assert len(a["nodes"]) == 50 and len(a["pods"]) == 80 and len(a["services"]) == 200
    c = workloads.generate(2, 10, 20)
    assert all(len(n["extendedResources"]["nvidia.com/gpu"]) == 16 for n in c["nodes"])
    d = workloads.generate(1, 12, 5, padded=False)
    assert d["nodes"][10]["name"] == "node-10"
