"""ImageLocalityPriority (priorities/image_locality.go:32-82), enabled through a Policy: whole streams against the oracle,
and the score column of an image list following the Node objects (created on first use, refreshed by a node update)."""
import copy
import ctypes

import pytest

from kubernetes_b200 import lib, wire, workloads
from tests import oracle_lib
from tests.test_gpu_parity import compare

pytestmark = pytest.mark.gpu

MB = 1024 * 1024


def _with_images(cluster, n_images=12):
    """Every node holds four of `n_images` images (sizes 60 MB ... 720 MB, two names each), every pod asks for one or two."""
    for i, node in enumerate(cluster["nodes"]):
        ids = sorted({(i + j * (2 + i % 5)) % n_images for j in range(4)})
        node["images"] = [{"names": ["reg/img-%d" % k, "reg/img-%d:v1" % k], "sizeBytes": (k + 1) * 60 * MB} for k in ids]
    for i, pod in enumerate(cluster["pods"]):
        pod["containers"][0]["image"] = "reg/img-%d" % (i % n_images)
        if i % 4 == 0:
            pod["containers"].append({"requests": {}, "image": "reg/img-%d:v1" % ((5 * i + 1) % n_images)})
        if i % 9 == 0:  # the same image twice counts twice (totalImageSize walks the containers)
            pod["containers"].append({"requests": {}, "image": pod["containers"][0]["image"]})
    return cluster


OPTS = {"weights": {"LeastRequestedPriority": 1, "BalancedResourceAllocation": 1, "SelectorSpreadPriority": 1,
                    "NodeAffinityPriority": 1, "TaintTolerationPriority": 1, "InterPodAffinityPriority": 1,
                    "NodePreferAvoidPodsPriority": 10000, "ImageLocalityPriority": 3}}


def test_image_locality_streams_match_oracle():
    cluster = _with_images(workloads.generate(3, 300, 1500))
    compare(cluster, "image locality", gpu_opts=OPTS, oracle_opts=OPTS)
    compare(_with_images(workloads.generate(3, 120, 300)), "image locality single", gpu_opts=dict(OPTS, batch=0), oracle_opts=OPTS)
    both = copy.deepcopy(OPTS)  # next to the other Policy priorities
    both["weights"].update({"MostRequestedPriority": 2, "ResourceLimitsPriority": 1})
    compare(_with_images(workloads.generate(5, 100, 400)), "image locality + most requested", gpu_opts=both, oracle_opts=both)


def _scores(so, e, pod, n):
    rec = wire.pod_record(pod)
    r = lib.KsbResult()
    assert so.ksb_schedule(e, rec, len(rec), ctypes.byref(r)) == 0, so.ksb_last_error(e)
    assert r.status == 0 and r.num_filtered == n, (r.status, r.num_filtered)
    sc = (ctypes.c_int64 * n)()
    so.ksb_scores.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_int64), ctypes.c_int32]
    assert so.ksb_scores(e, sc, n) == 0, so.ksb_last_error(e)
    return list(sc)


def test_image_scores_follow_node_updates():
    """Only ImageLocality weighted: the whole HostPriorityList before and after a node's image list changes, for an
    image list seen before the update and one first seen after it."""
    so = lib.engine()
    so.ksb_policy_set.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.c_char_p, ctypes.c_size_t, ctypes.c_int32, ctypes.c_int32]
    cfg = lib.KsbConfig()
    so.ksb_config_default(ctypes.byref(cfg))
    for f in ("w_least_requested", "w_balanced", "w_selector_spread", "w_node_affinity", "w_taint_toleration",
              "w_inter_pod_affinity", "w_prefer_avoid"):
        setattr(cfg, f, 0)
    err = ctypes.create_string_buffer(512)
    e = so.ksb_create(ctypes.byref(cfg), err, 512)
    assert e, err.value
    try:
        assert so.ksb_policy_set(e, 5, None, 0, 0, 1) == 0
        cluster = _with_images(workloads.generate(1, 40, 4))
        nodes = sorted(cluster["nodes"], key=lambda n: n["name"].encode())
        for n in nodes:
            b = wire.node_record(n)
            assert so.ksb_node_upsert(e, b, len(b)) == 0
        pod_a = dict(cluster["pods"][0], containers=[{"requests": {}, "image": "reg/img-3"}, {"requests": {}, "image": "reg/img-7:v1"}])
        pod_b = dict(cluster["pods"][1], containers=[{"requests": {}, "image": "reg/img-11"}, {"requests": {}, "image": "reg/new"}])

        def want(pod):
            out = oracle_lib.call({"op": "priority", "name": "ImageLocalityPriority", "pod": pod,
                                   "cluster": {"nodes": nodes, "pods": []}})
            by = {s["host"]: s["score"] for s in out["scores"]}
            return [by[n["name"]] for n in nodes]

        got = _scores(so, e, pod_a, len(nodes))
        assert got == want(pod_a) and max(got) > 0, got
        # node 5 loses its images, node 6 gains a large one under a new name and the old one at another size
        nodes[5] = dict(nodes[5], images=[])
        nodes[6] = dict(nodes[6], images=nodes[6]["images"] + [{"names": ["reg/new"], "sizeBytes": 400 * MB},
                                                               {"names": ["reg/img-3"], "sizeBytes": 990 * MB}])
        for n in (nodes[5], nodes[6]):
            b = wire.node_record(n)
            assert so.ksb_node_upsert(e, b, len(b)) == 0
        got2 = _scores(so, e, pod_a, len(nodes))
        assert got2 == want(pod_a), (got2, want(pod_a))
        assert got2[5] == 0 and got2[6] == 10 and got2 != got
        got_b = _scores(so, e, pod_b, len(nodes))  # an image list first seen after the update
        assert got_b == want(pod_b) and got_b[6] > 0, got_b
    finally:
        so.ksb_destroy(e)
