"""CPU-side checks: the C-ABI library loads and exports every symbol include/ksched_b200.h declares,
refuses to run without a GPU, and the wire/JSON harness parses clusters. No compute calls here."""
import ctypes
import json
import os
import re

import pytest

from kubernetes_b200 import lib, wire, workloads

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


def test_go_shim_binds_only_declared_symbols():
    """integration/gpusched/*.go cannot be compiled here (no Go toolchain): at least every C symbol, constant and type the
    shim names must exist in include/ksched_b200.h, and the Go encoders must write the records field for field as
    kubernetes_b200/wire.py does (same sequence of writer calls per record kind)."""
    header = open(os.path.join(ROOT, "include", "ksched_b200.h")).read()
    declared = set(re.findall(r"\b(ksb_[a-z_0-9]+)\b", header)) | set(re.findall(r"\b(KSB_[A-Z_0-9]+)\b", header))
    gdir = os.path.join(ROOT, "integration", "gpusched")
    used = set()
    for fn in sorted(os.listdir(gdir)):
        if fn.endswith(".go"):
            used |= set(re.findall(r"\bC\.((?:ksb|KSB)_[A-Za-z_0-9]+)\b", open(os.path.join(gdir, fn)).read()))
    assert used, "the shim binds nothing?"
    assert used <= declared, sorted(used - declared)
    # the shim covers the path's entry points
    for name in ("ksb_create", "ksb_schedule", "ksb_schedule_extended", "ksb_preempt", "ksb_node_upsert", "ksb_node_remove",
                 "ksb_pod_add", "ksb_pod_remove", "ksb_selector_owner_upsert", "ksb_failed_predicates", "ksb_resync_begin",
                 "ksb_policy_set"):
        assert name in used, name


def test_go_and_python_pod_encoders_agree_on_field_order():
    """Both writers are sequences of u8 / u32 / i32 / i64 / str / map calls: the pod record's tail (extended resources,
    volume count, priority, container images) and the node record's head are compared by the order of the field names they mention."""
    go = open(os.path.join(ROOT, "integration", "gpusched", "encode.go")).read()
    py = open(os.path.join(ROOT, "kubernetes_b200", "wire.py")).read()
    go_pod = go[go.index("func encodePod"):go.index("// selector owners for SelectorSpread")]
    py_pod = py[py.index("def pod_record"):]
    order_go = [go_pod.index(k) for k in ("p.Namespace", "p.Name", "p.Labels", "DeletionTimestamp", "NodeName", "Containers",
                                          "InitContainers", "NodeSelector", "Tolerations", "ExtendedResources", "Volumes", "prio",
                                          ".Image")]
    assert order_go == sorted(order_go), order_go
    order_py = [py_pod.index(k) for k in ('"namespace"', '"name"', '"labels"', '"deleting"', '"nodeName"', '"controller"',
                                          '"containers"', '"nodeSelector"', '"tolerations"', '"extendedResources"', '"volumes"',
                                          '"priority"', '"image"')]
    assert order_py == sorted(order_py), order_py


def _records_agree(cluster, label):
    """Every record of `cluster` as the C++ encoder (JSON loader of csrc/host/model.hpp + include/ksb_wire.hpp) and as the
    Python encoder (kubernetes_b200/wire.py, written against the layout comment alone) produce it."""
    s = lib.Session(cluster=cluster)
    try:
        n = 0
        for i, node in enumerate(cluster.get("nodes") or []):
            assert s.record(0, i) == wire.node_record(node), (label, "node", i)
            n += 1
        for i, pod in enumerate(cluster.get("pods") or []):
            assert s.record(1, i) == wire.pod_record(pod), (label, "pod", i)
            n += 1
        for i, pod in enumerate(cluster.get("existingPods") or []):
            assert s.record(2, i) == wire.pod_record(pod), (label, "existing pod", i)
            n += 1
        k = 0
        for kind, key in enumerate(("services", "rcs", "rss", "sss")):
            for i, o in enumerate(cluster.get(key) or []):
                o = dict(o, name=o.get("name") or "owner-%d-%d" % (kind, i))
                assert s.record(3, k) == wire.owner_record(o, kind >= 2), (label, key, i)
                k += 1
                n += 1
        assert s.record(0, len(cluster.get("nodes") or [])) is None
        return n
    finally:
        s.close()


def test_cpp_and_python_encoders_write_the_same_bytes():
    """The oracle reads its clusters through the same JSON loader as the product's host side (model.hpp): a field dropped
    there would be invisible to the parity tests. The Python encoder shares no code with it, so byte-equal records for
    the generated workloads and for every transcribed reference vector pin the loader and both encoders to the layout
    in include/ksb_wire.hpp (a pod or node field that one side forgets shifts every byte after it)."""
    _built()
    total = 0
    for cfg in (1, 2, 3, 4, 5):
        total += _records_agree(workloads.generate(cfg, 50, 150), "cfg%d" % cfg)
    with open(os.path.join(ROOT, "tests", "golden", "reference_vectors.json")) as f:
        cases = json.load(f)["cases"]
    for c in cases:
        cl = c.get("cluster")
        if not isinstance(cl, dict):
            continue
        cl = dict(cl)
        if isinstance(c.get("pod"), dict):
            cl["pods"] = list(cl.get("pods") or []) + [c["pod"]]
        total += _records_agree(cl, "%s / %s" % (c.get("source"), c.get("test")))
    assert total > 2000, total
