"""ksb_resync_begin / ksb_resync_end through the raw C-ABI: after the caller has replayed its cache (nodes, selector
owners, the pods placed so far with their node names) the engine must go on exactly as an engine that was never
interrupted: same hosts, filtered counts and scores for the rest of the stream."""
import ctypes

import pytest

from kubernetes_b200 import lib, wire, workloads

pytestmark = pytest.mark.gpu


def _create(so):
    cfg = lib.KsbConfig()
    so.ksb_config_default(ctypes.byref(cfg))
    err = ctypes.create_string_buffer(512)
    e = so.ksb_create(ctypes.byref(cfg), err, 512)
    assert e, err.value
    return e


def _load(so, e, cluster):
    for n in cluster["nodes"]:
        b = wire.node_record(n)
        assert so.ksb_node_upsert(e, b, len(b)) == 0, so.ksb_last_error(e)
    for i, s in enumerate(cluster.get("services", [])):
        b = wire.owner_record(dict(s, name=s.get("name") or "svc-%d" % i), False)
        assert so.ksb_selector_owner_upsert(e, 0, b, len(b)) == 0, so.ksb_last_error(e)


def _batch(so, e, pods):
    recs = [wire.pod_record(p) for p in pods]
    offs = [0]
    for r in recs:
        offs.append(offs[-1] + len(r))
    out = (lib.KsbResult * len(recs))()
    assert so.ksb_schedule_batch(e, b"".join(recs), (ctypes.c_uint64 * (len(recs) + 1))(*offs), len(recs), out) == 0, so.ksb_last_error(e)
    return [(r.status, r.node_index, r.num_filtered, r.score) for r in out]


@pytest.mark.parametrize("cfg,nodes,pods", [(3, 500, 1600), (5, 200, 800)])
def test_resync_then_continue_matches_uninterrupted_run(cfg, nodes, pods):
    so = lib.engine()
    so.ksb_selector_owner_upsert.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.c_char_p, ctypes.c_size_t]
    so.ksb_resync_begin.argtypes = [ctypes.c_void_p]
    so.ksb_resync_end.argtypes = [ctypes.c_void_p]
    so.ksb_set_last_node_index.argtypes = [ctypes.c_void_p, ctypes.c_uint64]
    cluster = workloads.generate(cfg, nodes, pods)
    half = pods // 2
    ref = _create(so)
    eng = _create(so)
    try:
        _load(so, ref, cluster)
        want_a = _batch(so, ref, cluster["pods"][:half])
        want_b = _batch(so, ref, cluster["pods"][half:])
        _load(so, eng, cluster)
        got_a = _batch(so, eng, cluster["pods"][:half])
        assert got_a == want_a
        # the caller's cache: every node, owner, and the pods placed so far (NodeName set, as AssumePod records them)
        names = sorted(n["name"] for n in cluster["nodes"])
        assert so.ksb_resync_begin(eng) == 0, so.ksb_last_error(eng)
        _load(so, eng, cluster)
        for p, r in zip(cluster["pods"][:half], got_a):
            if r[0] != 0:
                continue
            placed = dict(p, nodeName=names[r[1]])
            b = wire.pod_record(placed)
            assert so.ksb_pod_add(eng, b, len(b)) == 0, so.ksb_last_error(eng)
        assert so.ksb_resync_end(eng) == 0, so.ksb_last_error(eng)
        got_b = _batch(so, eng, cluster["pods"][half:])
        assert got_b == want_b
        assert so.ksb_get_last_node_index(eng) == so.ksb_get_last_node_index(ref)
    finally:
        so.ksb_destroy(ref)
        so.ksb_destroy(eng)
