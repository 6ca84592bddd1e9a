"""The headline configuration under pytest: cfg4 at its full 50,000 nodes through ksb_schedule_batch (the
automatic pipelined schedule and the chunked host pipeline, exactly what bench.py times), compared pod by pod with
the literal oracle's outcome for the first 50,000 pods of the stream. The oracle needed 95 minutes for that prefix
(scripts/oracle_prefix.py), so its outcome is a committed fixture (tests/golden/cfg4_prefix_oracle.json:
status, host, filtered count, score, sha1 prefix of the FitError text per pod, and lastNodeIndex); the GPU side
takes about a second. Two live-oracle cases cover what the fixture cannot: a table that crosses the 16,384-row
threshold at which pipelining switches itself on, and a batch with more no-fit pods than KSB_MAX_FITERR."""
import hashlib
import json
import os

import pytest

from kubernetes_b200 import lib, workloads
from tests import oracle_lib
from tests.test_gpu_parity import compare

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))


def test_cfg4_full_node_count_prefix_matches_committed_oracle():
    ref = json.load(open(os.path.join(HERE, "golden", "cfg4_prefix_oracle.json")))
    assert (ref["config"], ref["nodes"], ref["stream_pods"], ref["prefix"]) == (4, 50000, 100000, 50000)
    cluster = workloads.generate(4, ref["nodes"], ref["stream_pods"])
    cluster["pods"] = cluster["pods"][:ref["prefix"]]
    sess = lib.Session(cluster=cluster)
    got = sess.run()
    sess.close()
    assert len(got["results"]) == len(ref["results"]) == 50000
    bad = []
    for i, (g, w) in enumerate(zip(got["results"], ref["results"])):
        msg = hashlib.sha1(g["message"].encode()).hexdigest()[:12] if g["status"] == 2 else ""
        mine = [g["status"], g["host"], g["filtered"], g["score"] if g["status"] == 0 else w[3], msg]
        if mine != w:
            bad.append((i, mine, w))
    assert not bad, bad[:3]
    assert got["lastNodeIndex"] == ref["lastNodeIndex"]
    assert got["groups"] > 0 and got["kernel_launches"] >= 12  # 50,000 pods: the chunked pipeline (4,096-pod launches)


def test_cfg4_crossing_the_automatic_pipeline_threshold_live_oracle():
    """17,000 rows: above the 16,384 rows from which the engine pipelines commit and evaluation by itself."""
    cluster = workloads.generate(4, 17000, 400)
    got, _ = compare(cluster, "cfg4 17000x400")
    assert got["groups"] > 0


def test_more_no_fit_pods_than_the_histogram_floor():
    """cfg2 on a tiny cluster: most of 5,000 pods end in a FitError, more than KSB_MAX_FITERR (4,096) in one launch;
    every one of them must still carry the reference's error text."""
    cluster = workloads.generate(2, 12, 5000)
    got, want = compare(cluster, "cfg2 12x5000")
    assert got["fit_errors"] > 4096
