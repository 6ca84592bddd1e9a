"""Mutators never wait for a scheduling call (the delta queue of SURVEY.md §8b): while one thread runs a large
ksb_schedule_batch, cache updates from another thread return at once, and they are in force, in order, when the next
call starts."""
import ctypes
import json
import threading
import time

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


def _encode(pods):
    recs = [wire.pod_record(p) for p in pods]
    offs = [0]
    for r in recs:
        offs.append(offs[-1] + len(r))
    return b"".join(recs), (ctypes.c_uint64 * (len(recs) + 1))(*offs), len(recs)


def _batch(so, e, enc):
    blob, offs, n = enc
    out = (lib.KsbResult * n)()
    rc = so.ksb_schedule_batch(e, blob, offs, n, out)
    assert rc == 0, so.ksb_last_error(e)
    return [(r.status, r.node_index, r.num_filtered, r.score) for r in out]


def test_mutators_do_not_wait_for_a_running_batch():
    so = lib.engine()
    so.ksb_selector_owner_upsert.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.c_char_p, ctypes.c_size_t]
    so.ksb_pod_remove.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_size_t, ctypes.c_char_p, ctypes.c_size_t]
    so.ksb_node_info.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_size_t, ctypes.c_char_p, ctypes.c_size_t,
                                 ctypes.POINTER(ctypes.c_size_t)]
    so.ksb_deferred_errors.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_size_t]
    so.ksb_flush.argtypes = [ctypes.c_void_p]
    cluster = workloads.generate(3, 5000, 30000)
    first, rest = _encode(cluster["pods"][:24000]), _encode(cluster["pods"][24000:26000])  # encoded up front: the thread below starts in the call

    def load(e):
        for n in cluster["nodes"]:
            b = wire.node_record(n)
            assert so.ksb_node_upsert(e, b, len(b)) == 0
        for i, s in enumerate(cluster.get("services", [])):
            b = wire.owner_record(dict(s, name=s.get("name") or "svc-%d" % i), False)
            assert so.ksb_selector_owner_upsert(e, 0, b, len(b)) == 0
        assert so.ksb_flush(e) == 0

    target = sorted(n["name"] for n in cluster["nodes"])[7]
    extra = [dict(workloads.generate(1, 4, 40)["pods"][i], name="late-%d" % i, nodeName=target) for i in range(40)]

    # reference engine: the batch, then the updates, then more pods
    ref = _create(so)
    eng = _create(so)
    try:
        load(ref)
        want_a = _batch(so, ref, first)
        for p in extra:
            b = wire.pod_record(p)
            assert so.ksb_pod_add(ref, b, len(b)) == 0, so.ksb_last_error(ref)
        want_b = _batch(so, ref, rest)

        load(eng)
        got_a = []
        t = threading.Thread(target=lambda: got_a.extend(_batch(so, eng, first)))
        t0 = time.perf_counter()
        t.start()
        time.sleep(0.005)  # let the batch take the engine (it runs for some 100 ms)
        slow, queued_while_busy = 0.0, 0
        for p in extra:
            b = wire.pod_record(p)
            c0 = time.perf_counter()
            assert so.ksb_pod_add(eng, b, len(b)) == 0
            c1 = time.perf_counter()
            slow = max(slow, c1 - c0)
            queued_while_busy += 1 if t.is_alive() else 0
        # a duplicate arrives queued: its failure is deferred, not lost
        b = wire.pod_record(extra[0])
        dup_rc = so.ksb_pod_add(eng, b, len(b))
        dup_was_queued = t.is_alive()
        t.join()
        batch_s = time.perf_counter() - t0
        assert got_a == want_a  # the updates did not leak into the batch they arrived next to
        assert queued_while_busy >= 30, (queued_while_busy, batch_s)
        assert slow < batch_s / 4, (slow, batch_s)
        got_b = _batch(so, eng, rest)
        assert got_b == want_b  # ... and are all in force for the next call
        buf = ctypes.create_string_buffer(4096)
        nerr = so.ksb_deferred_errors(eng, buf, 4096)
        if dup_was_queued:
            assert dup_rc == 0 and nerr == 1 and b"already in the cache" in buf.value, (dup_rc, nerr, buf.value)
        else:
            assert dup_rc != 0
        info = ctypes.create_string_buffer(1 << 16)
        ln = ctypes.c_size_t()
        assert so.ksb_node_info(eng, target.encode(), len(target), info, 1 << 16, ctypes.byref(ln)) == 0
        assert json.loads(info.value.decode())["pods"] >= 40
    finally:
        so.ksb_destroy(ref)
        so.ksb_destroy(eng)
