"""ksb_schedule_extended: Schedule with the verdicts of scheduler extenders (core/generic_scheduler.go:333-351, :608-632)
against the oracle: Filter verdicts cut the filtered list before the fork's device selection and before the Reduce
halves normalise over it, Prioritize verdicts are added to the weighted sum."""
import ctypes
import json
import os

import pytest

from kubernetes_b200 import lib, wire, workloads
from tests import oracle_lib

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("cfg,nodes,pods", [(3, 300, 60), (5, 150, 40)])
def test_extender_verdicts_match_oracle(cfg, nodes, pods):
    so = lib.engine()
    so.ksb_selector_owner_upsert.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.c_char_p, ctypes.c_size_t]
    so.ksb_schedule_extended.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_size_t, ctypes.POINTER(ctypes.c_uint32),
                                         ctypes.POINTER(ctypes.c_int32), ctypes.POINTER(lib.KsbResult)]
    so.ksb_node_name.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.POINTER(ctypes.c_char_p), ctypes.POINTER(ctypes.c_size_t)]
    so.ksb_fit_error_message.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.c_char_p, ctypes.c_size_t]
    cluster = workloads.generate(cfg, nodes, pods + 400)
    warm, probes = cluster["pods"][:400], cluster["pods"][400:]
    names = sorted(n["name"] for n in cluster["nodes"])
    rng = workloads.SplitMix64(4242 + cfg)
    cfgs = lib.KsbConfig()
    so.ksb_config_default(ctypes.byref(cfgs))
    err = ctypes.create_string_buffer(512)
    e = so.ksb_create(ctypes.byref(cfgs), err, 512)
    assert e, err.value
    try:
        for n in cluster["nodes"]:
            b = wire.node_record(n)
            assert so.ksb_node_upsert(e, b, len(b)) == 0
        for i, s in enumerate(cluster.get("services", [])):
            b = wire.owner_record(dict(s, name=s.get("name") or "svc-%d" % i), False)
            assert so.ksb_selector_owner_upsert(e, 0, b, len(b)) == 0
        # some load first, placed by the engine itself
        recs = [wire.pod_record(p) for p in warm]
        offs = [0]
        for r in recs:
            offs.append(offs[-1] + len(r))
        out = (lib.KsbResult * len(recs))()
        assert so.ksb_schedule_batch(e, b"".join(recs), (ctypes.c_uint64 * (len(recs) + 1))(*offs), len(recs), out) == 0
        placed = []
        for i, p in enumerate(warm):
            if out[i].status != 0:
                continue
            q = dict(p, nodeName=names[out[i].node_index])
            placed.append(q)
        checked = fit_errors = 0
        for k, pod in enumerate(probes):
            keep = None
            if k % 3 != 2:  # a Filter verdict: keep about a third of the nodes (sometimes none of the fitting ones)
                keep = [nm for nm in names if rng.below(100) < (35 if k % 5 else 2)]
            scores = {nm: rng.below(40) for nm in names} if k % 2 == 0 else None
            ext = {}
            if keep is not None:
                ext["keep"] = keep
            if scores is not None:
                ext["scores"] = scores
            want = oracle_lib.call({"op": "stream", "cluster": dict(cluster, pods=[pod], existingPods=placed), "extender": ext,
                                    "lastNodeIndex": int(so.ksb_get_last_node_index(e))})["results"][0]
            words = (len(names) + 31) // 32
            karr = None
            if keep is not None:
                kset = set(keep)
                bits = [0] * words
                for r, nm in enumerate(names):
                    if nm in kset:
                        bits[r >> 5] |= 1 << (r & 31)
                karr = (ctypes.c_uint32 * words)(*bits)
            sarr = (ctypes.c_int32 * len(names))(*[scores[nm] for nm in names]) if scores is not None else None
            b = wire.pod_record(pod)
            res = lib.KsbResult()
            assert so.ksb_schedule_extended(e, b, len(b), karr, sarr, ctypes.byref(res)) == 0, so.ksb_last_error(e)
            assert res.status == want["status"], (k, res.status, want)
            assert res.num_filtered == want["filtered"], (k, res.num_filtered, want)
            if want["status"] == 0:
                assert names[res.node_index] == want["host"] and res.score == want["score"], (k, names[res.node_index], res.score, want)
            elif want["status"] == 2:
                msg = ctypes.create_string_buffer(4096)
                assert so.ksb_fit_error_message(e, 0, msg, 4096) == 0
                assert msg.value.decode() == want["message"], (k, msg.value, want["message"])
                fit_errors += 1
            checked += 1
        assert checked == len(probes)
    finally:
        so.ksb_destroy(e)


def test_extender_reference_vectors_on_gpu():
    """The reference's own vectors that reach the engine as extender verdicts, through the C-ABI: TestSelectHost
    (core/generic_scheduler_test.go:117-181: the HostPriorityList arrives as the extenders' combined score with every
    priority weighted 0, ten Schedule + AssumePod rounds per list; the host must be one of possibleHosts every time and
    follow the round-robin order selectHost implies) and TestGenericSchedulerWithExtenders (core/extender_test.go:188-337:
    Filter verdicts as the rows kept, Prioritize verdicts as scores, EqualPriority weight 1 on the built-in side)."""
    so = lib.engine()
    so.ksb_schedule_extended.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_size_t, ctypes.POINTER(ctypes.c_uint32),
                                         ctypes.POINTER(ctypes.c_int32), ctypes.POINTER(lib.KsbResult)]
    so.ksb_policy_set.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.c_char_p, ctypes.c_size_t, ctypes.c_int32, ctypes.c_int32]
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_vectors.json")) as f:
        cases = [c for c in json.load(f)["cases"] if c["op"] == "stream" and "extender" in c]
    assert len(cases) == 9
    for c in cases:
        cfg = lib.KsbConfig()
        so.ksb_config_default(ctypes.byref(cfg))
        for f_ in ("w_least_requested", "w_balanced", "w_selector_spread", "w_node_affinity", "w_taint_toleration",
                   "w_inter_pod_affinity", "w_prefer_avoid"):
            setattr(cfg, f_, 0)
        err = ctypes.create_string_buffer(512)
        e = so.ksb_create(ctypes.byref(cfg), err, 512)
        assert e, err.value
        try:
            equal = c["weights"].get("EqualPriority", 0)
            if equal:
                assert so.ksb_policy_set(e, 7, None, 0, 0, equal) == 0, so.ksb_last_error(e)
            names = sorted((n["name"] for n in c["cluster"]["nodes"]), key=lambda s: s.encode())
            for n in c["cluster"]["nodes"]:
                b = wire.node_record(n)
                assert so.ksb_node_upsert(e, b, len(b)) == 0
            ext = c["extender"]
            sarr = karr = None
            if "scores" in ext:
                sarr = (ctypes.c_int32 * len(names))(*[ext["scores"][nm] for nm in names])
            if "keep" in ext:
                bits = [0] * ((len(names) + 31) // 32)
                for r, nm in enumerate(names):
                    if nm in ext["keep"]:
                        bits[r >> 5] |= 1 << (r & 31)
                karr = (ctypes.c_uint32 * len(bits))(*bits)
            hosts, status, scores = [], [], []
            for pod in c["cluster"]["pods"]:
                b = wire.pod_record(pod)
                res = lib.KsbResult()
                assert so.ksb_schedule_extended(e, b, len(b), karr, sarr, ctypes.byref(res)) == 0, so.ksb_last_error(e)
                status.append(res.status)
                scores.append(res.score)
                hosts.append(names[res.node_index] if res.status == 0 else "")
                if res.status == 0:
                    placed = wire.pod_record(dict(pod, nodeName=names[res.node_index]))  # AssumePod (scheduler.go:478)
                    assert so.ksb_pod_add(e, placed, len(placed)) == 0, so.ksb_last_error(e)
            exp = c["expect"]
            if "status" in exp:
                assert status == exp["status"], (c["test"], status)
            for g, allowed in zip(hosts, exp.get("host_in", [])):
                assert g in allowed, (c["test"], hosts)
            if "hosts" in exp:
                assert hosts == exp["hosts"], (c["test"], hosts)
            if "scores" in exp:
                assert scores == exp["scores"], (c["test"], scores)
        finally:
            so.ksb_destroy(e)
