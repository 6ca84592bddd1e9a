"""The exported cache state machine (ksb_cache_*, schedulercache/cache.go:118-462 and factory.go:306-353) through the raw
C-ABI: pods placed by ksb_schedule_batch are assumed, skipPodUpdate ignores a NodeName-only update, an assumed pod whose
binding finished expires after the ttl and can be added back by the informer's AddPod."""
import ctypes
import json

import pytest

from kubernetes_b200 import lib, wire, workloads

pytestmark = pytest.mark.gpu


def _node_pods(so, e, name):
    buf = ctypes.create_string_buffer(1 << 16)
    ln = ctypes.c_size_t()
    assert so.ksb_node_info(e, name.encode(), len(name), buf, 1 << 16, ctypes.byref(ln)) == 0
    return json.loads(buf.value.decode())["pods"]


def test_batch_placements_enter_the_state_machine():
    so = lib.engine()
    so.ksb_selector_owner_upsert.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.c_char_p, ctypes.c_size_t]
    so.ksb_node_info.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_size_t, ctypes.c_char_p, ctypes.c_size_t,
                                 ctypes.POINTER(ctypes.c_size_t)]
    so.ksb_node_name.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.POINTER(ctypes.c_char_p), ctypes.POINTER(ctypes.c_size_t)]
    so.ksb_cache_set_ttl.argtypes = [ctypes.c_void_p, ctypes.c_double]
    so.ksb_cache_set_ttl.restype = None
    for name in ("ksb_cache_assume_pod", "ksb_cache_forget_pod", "ksb_cache_add_pod", "ksb_cache_remove_pod"):
        getattr(so, name).argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_size_t]
    so.ksb_cache_finish_binding.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_size_t, ctypes.c_double]
    so.ksb_cache_cleanup_assumed.argtypes = [ctypes.c_void_p, ctypes.c_double]
    for name in ("ksb_cache_is_assumed_pod", "ksb_cache_skip_pod_update"):
        getattr(so, name).argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_size_t, ctypes.POINTER(ctypes.c_int32)]
    cluster = workloads.generate(3, 60, 40)
    cfg = lib.KsbConfig()
    so.ksb_config_default(ctypes.byref(cfg))
    err = ctypes.create_string_buffer(512)
    e = so.ksb_create(ctypes.byref(cfg), err, 512)
    assert e, err.value
    try:
        for n in cluster["nodes"]:
            b = wire.node_record(n)
            assert so.ksb_node_upsert(e, b, len(b)) == 0
        for i, s in enumerate(cluster.get("services", [])):
            b = wire.owner_record(dict(s, name=s.get("name") or "svc-%d" % i), False)
            assert so.ksb_selector_owner_upsert(e, 0, b, len(b)) == 0
        so.ksb_cache_set_ttl(e, 30.0)
        recs = [wire.pod_record(p) for p in cluster["pods"]]
        offs = [0]
        for r in recs:
            offs.append(offs[-1] + len(r))
        out = (lib.KsbResult * len(recs))()
        assert so.ksb_schedule_batch(e, b"".join(recs), (ctypes.c_uint64 * (len(recs) + 1))(*offs), len(recs), out) == 0
        i = next(k for k in range(len(recs)) if out[k].status == 0)
        nm, nl = ctypes.c_char_p(), ctypes.c_size_t()
        assert so.ksb_node_name(e, out[i].node_index, ctypes.byref(nm), ctypes.byref(nl)) == 0
        host = nm.value[:nl.value].decode()
        placed = dict(cluster["pods"][i], nodeName=host)
        flag = ctypes.c_int32(-1)
        b = wire.pod_record(placed)
        assert so.ksb_cache_is_assumed_pod(e, b, len(b), ctypes.byref(flag)) == 0 and flag.value == 1
        # the informer's echo of the bind: NodeName now set, nothing else changed -> skipped (factory.go:306-353)
        assert so.ksb_cache_skip_pod_update(e, b, len(b), ctypes.byref(flag)) == 0 and flag.value == 1
        relabeled = wire.pod_record(dict(placed, labels=dict(placed.get("labels") or {}, extra="1")))
        assert so.ksb_cache_skip_pod_update(e, relabeled, len(relabeled), ctypes.byref(flag)) == 0 and flag.value == 0
        # assuming it again is the reference's error
        assert so.ksb_cache_assume_pod(e, b, len(b)) != 0
        assert b"state wasn't initial but get assumed" in so.ksb_last_error(e)
        before = _node_pods(so, e, host)
        # binding finished at t=0, ttl 30 s: still there at t=29, expired at t=31 (cache.go:431-462)
        assert so.ksb_cache_finish_binding(e, b, len(b), 0.0) == 0
        assert so.ksb_cache_cleanup_assumed(e, 29.0) == 0
        assert _node_pods(so, e, host) == before
        assert so.ksb_cache_cleanup_assumed(e, 31.0) == 0
        assert _node_pods(so, e, host) == before - 1
        assert so.ksb_cache_is_assumed_pod(e, b, len(b), ctypes.byref(flag)) == 0 and flag.value == 0
        # the informer reports it after all: added (back), cache.go:248-252
        assert so.ksb_cache_add_pod(e, b, len(b)) == 0, so.ksb_last_error(e)
        assert _node_pods(so, e, host) == before
        assert so.ksb_cache_add_pod(e, b, len(b)) != 0 and b"already in added state" in so.ksb_last_error(e)
        assert so.ksb_cache_remove_pod(e, b, len(b)) == 0
        assert _node_pods(so, e, host) == before - 1
        # a pod that was never assumed cannot be forgotten
        assert so.ksb_cache_forget_pod(e, b, len(b)) != 0 and b"wasn't assumed but get forgotten" in so.ksb_last_error(e)
    finally:
        so.ksb_destroy(e)
