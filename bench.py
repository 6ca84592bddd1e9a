#!/usr/bin/env python3
"""bench.py — pods scheduled per second through the GPU engine (BASELINE.json's metric).

One "step" = one pass of the hot path (genericScheduler.Schedule + assume, for every pod of the
stream in FIFO order) over one synthetic cluster. Default workload: BASELINE.json configs[3]
(50,000 nodes x 100,000 pods, kubemark distributions); `--config 3` selects the
5,000-node case the metric is also quoted on.

  value     pods/s with the node table, the pod requirement blocks and all scratch already resident
            in HBM (ksb_replay_last_batch: CUDA-event time of the persistent kernel only)
  e2e       pods/s through the C-ABI call a Go shim makes (ksb_schedule_batch) with HOST buffers:
            decode + encode on the host, H2D of the requirement blocks, kernel, D2H of the results
  roofline  algorithmic bytes (B_cfg x nodes x pods) / kernel time vs measured HBM peak
  cpu_baseline / --impl reference: the CPU restatement of the reference's Go scheduler (oracle/,
            16 worker threads like workqueue.Parallelize(16)) on a bounded prefix of the same stream.

Multi-GPU (torchrun): ONE stream, the node table sharded by row range over the ranks' GPUs; phases 1-2 of
the kernel run on every GPU, the Reduce inputs are combined with NVLink atomics, rank 0 commits (DESIGN.md
"Multi-GPU"). Total work is fixed, so scaling is "strong"; the stream is sequential and latency-bound, so
more GPUs do not make it faster at these sizes -- the numbers say so.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "pods_scheduled_per_sec"
UNIT = "pods/s"


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.path = tempfile.mktemp(prefix="clocks_", suffix=".csv")
        self.proc = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(gpu_index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except OSError:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        try:
            for line in open(self.path):
                f = [x.strip() for x in line.split(",")]
                if len(f) < 6:
                    continue
                try:
                    sm.append(float(f[0]))
                    mx.append(float(f[1]))
                except ValueError:
                    continue
                for n, v in zip(names, f[2:6]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            os.unlink(self.path)
        except OSError:
            pass
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def oracle_baseline(cluster, sample_pods, workers):
    from tests import oracle_lib
    t0 = time.time()
    out = oracle_lib.call({"op": "stream", "cluster": cluster, "limit": sample_pods, "workers": workers})
    wall = time.time() - t0
    if "error" in out:
        raise RuntimeError(out["error"])
    return out["pods"] / out["seconds"], out["seconds"], wall


def ncu_traffic(cfg, n_nodes, n_pods):
    """DRAM bytes per launch of the batch kernel from the committed `ncu --set full` summary of this workload, else None."""
    path = os.path.join(ROOT, "profiles", "r02_ncu_summary.json")
    if not os.path.exists(path):
        return None
    try:
        with open(path) as f:
            for rec in json.load(f)["captures"]:
                if (rec["config"], rec["nodes"], rec["pods"]) == (cfg, n_nodes, n_pods):
                    return rec["dram_bytes_read"] + rec["dram_bytes_write"]
    except (OSError, KeyError, ValueError):
        pass
    return None


def one_pod_latency(sess, device, n_pods):
    """p50 / p99 of the one-pod-per-call path (Schedule + AssumePod per pod, the reference's own call pattern): over
    EVERY call after ksb_flush, first uses of labels and selector classes included, and over the warm tail alone."""
    n_cold = min(n_pods // 2, 500)
    n_lat = min(n_pods - n_cold, 2000)
    every = sess.run(device=device, results=False, batch=0, limit=n_cold + n_lat, lat_skip=0)
    warm = sess.run(device=device, results=False, batch=0, limit=n_cold + n_lat, lat_skip=n_cold)
    return {"p50_us": every["p50_us"], "p99_us": every["p99_us"], "max_us": every["max_us"],
            "warm_p50_us": warm["p50_us"], "warm_p99_us": warm["p99_us"], "warm_max_us": warm["max_us"],
            "unit": "us per ksb_schedule call (host to host)",
            "sample": "pods 0..%d of the stream, one C-ABI call per pod, every call counted; warm_* leaves the first %d "
                      "calls out" % (n_cold + n_lat, n_cold)}


def batch_latency(sess, device, n_pods, bsz=64):
    """The queue drained `bsz` pods per call: a call's duration is the latency of each of its pods."""
    n_b = min(n_pods, bsz * 160)
    bl = sess.run(device=device, results=False, batch=bsz, limit=n_b)
    if "batch_call_p50_us" not in bl:
        return None
    return {"call_p50_us": bl["batch_call_p50_us"], "call_p99_us": bl["batch_call_p99_us"], "calls": bl["batch_calls"],
            "pods_per_s": n_b / (bl["e2e_ms"] / 1000.0)}


def side_record(cfg, device, steps):
    """The other half of BASELINE.json's metric ("at 5k and 50k nodes"): the 5,000-node config measured in the same
    run, same method (end-to-end pass on a fresh engine, then a resident replay with L2 flushed)."""
    from kubernetes_b200 import lib, workloads
    n_nodes, n_pods = workloads.CONFIG_SIZES[cfg]
    cluster = workloads.generate(cfg, n_nodes, n_pods)
    sess = lib.Session(cluster=cluster)
    sess.run(device=device, results=False)
    e2e, rep, hashes = [], [], set()
    for _ in range(max(2, steps)):
        out = sess.run(device=device, results=False, replay_iters=1)
        e2e.append(out["e2e_ms"])
        rep.append(out["replay_ms"])
        hashes.add(out["placement_hash"])
    assert len(hashes) == 1, "placements changed between steps"
    step_ms, e2e_ms = sum(rep) / len(rep), sum(e2e) / len(e2e)
    peak, _ = measured_peak()
    achieved = workloads.ALGO_BYTES[cfg] * n_nodes * n_pods / (step_ms / 1000.0) / 1e9
    rec = {"workload": "cfg%d: %d nodes x %d pods (BASELINE.json configs[%d])" % (cfg, n_nodes, n_pods, cfg - 1),
           "value": n_pods / (step_ms / 1000.0), "unit": UNIT, "ms_per_step": step_ms,
           "e2e": {"value": n_pods / (e2e_ms / 1000.0), "unit": UNIT, "ms_per_step": e2e_ms},
           "roofline": {"achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak},
           "placement_hash": sorted(hashes)[0], "latency": one_pod_latency(sess, device, n_pods)}
    sess.close()
    return rec


def pick_sample(n_nodes):
    # ~20 s of CPU work: the literal oracle on 16 threads does about 0.6 M (pod, node) evaluations per second
    return max(20, min(4000, int(1.25e7 / max(1, n_nodes))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", type=int, default=4, choices=[1, 2, 3, 4, 5])
    ap.add_argument("--nodes", type=int, default=None)
    ap.add_argument("--pods", type=int, default=None)
    ap.add_argument("--cpu-sample", type=int, default=None)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-side-config", action="store_true", help="skip the 5,000-node sub-record of the default run")
    args = ap.parse_args()
    # stdout carries exactly one JSON line: anything a library prints there (NCCL's version banner) goes to stderr
    json_out = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    from kubernetes_b200 import workloads
    n_nodes, n_pods = workloads.CONFIG_SIZES[args.config]
    n_nodes = args.nodes or n_nodes
    n_pods = args.pods or n_pods
    algo_bytes = workloads.ALGO_BYTES[args.config]
    workload = "cfg%d: %d nodes x %d pods (BASELINE.json configs[%d])" % (args.config, n_nodes, n_pods, args.config - 1)
    cores = os.cpu_count() or 1
    workers = min(16, cores)

    if args.impl == "reference":
        if rank != 0:
            return
        cluster = workloads.generate(args.config, n_nodes, n_pods)
        sample = args.cpu_sample or pick_sample(n_nodes)
        times = []
        for i in range(args.warmup + args.steps):
            pps, secs, _ = oracle_baseline(cluster, sample, workers)
            if i >= args.warmup:
                times.append((pps, secs))
        v = sum(p for p, _ in times) / len(times)
        ms = 1000.0 * sum(s for _, s in times) / len(times)
        sample_txt = "first %d pods of the stream on all %d nodes, %d worker threads (Parallelize(16))" % (sample, n_nodes, workers)
        json_out.write(json.dumps({
            "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": 0, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "int64", "data": "synthetic",
            "config": {"workload": workload, "note": "reference is Go and no Go toolchain exists here: this is the C++ "
                       "restatement of its algorithm (oracle/), literal mode"},
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": workers, "kind": "port", "sample": sample_txt},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        }) + "\n")
        json_out.flush()
        return

    import torch
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("cpu:gloo,cuda:nccl", device_id=torch.device("cuda", local_rank))
    torch.cuda.set_device(local_rank)

    from kubernetes_b200 import lib, sharding
    cluster = workloads.generate(args.config, n_nodes, n_pods)
    sess = lib.Session(cluster=cluster)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def one_pass(replay_iters):
        """One end-to-end pass on a fresh engine (cluster load untimed); N>1: node table sharded over the ranks."""
        if world == 1:
            return sess.run(device=local_rank, results=False, replay_iters=replay_iters)
        eng = lib.OpenEngine(sess, device=local_rank, shard_rank=rank, shard_count=world)
        try:
            sharding.connect(eng, rank, world, dist.all_gather_object, n_pods)
            dist.barrier()
            return eng.run(results=False, replay_iters=replay_iters)
        finally:
            dist.barrier()
            eng.close()

    # warm-up: full end-to-end passes
    for _ in range(max(3, args.warmup)):
        one_pass(0)
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    e2e_ms, h2d, d2h, kernel_ms, launches = [], 0, 0, [], 0
    replay_ms = []
    hashes = set()
    groups = restarts = 0
    for _ in range(args.steps):
        # one step: e2e pass on a fresh engine, then one resident-HBM replay of the same batch
        out = one_pass(1)
        e2e_ms.append(out["e2e_ms"])
        kernel_ms.append(out["kernel_ms"])
        replay_ms.append(out["replay_ms"])
        h2d, d2h = out["h2d_bytes"], out["d2h_bytes"]
        launches += int(out.get("kernel_launches", 1))  # the engine's own count: chunk launches of the e2e pass + the replay
        hashes.add(out["placement_hash"])
        result_hash = out.get("result_hash")
        scheduled, fit_errors = out["scheduled"], out["fit_errors"]
        groups, restarts = out["groups"], out["group_restarts"]
    barrier()
    clocks = sampler.stop() if sampler else None
    assert len(hashes) == 1, "placements changed between steps"

    step_ms = sum(replay_ms) / len(replay_ms)
    e2e_step_ms = sum(e2e_ms) / len(e2e_ms)
    if dist is not None:
        t = torch.tensor([step_ms, e2e_step_ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        step_ms, e2e_step_ms = float(t[0]), float(t[1])
    matches_unsharded = None
    if dist is not None:
        # the unsharded pass once on rank 0 (the sharded engines are closed, the other ranks wait): same placements,
        # filtered counts and scores (result_hash covers status, host, device ids, filtered count and score per pod)
        if rank == 0:
            solo = sess.run(device=local_rank, results=False)
            matches_unsharded = (solo["placement_hash"] == sorted(hashes)[0] and solo.get("result_hash") == result_hash)
        dist.barrier()
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    total_pods = n_pods  # N>1: one stream, node table sharded over the ranks (strong scaling)
    value = total_pods / (step_ms / 1000.0)
    e2e_value = total_pods / (e2e_step_ms / 1000.0)
    peak, peak_src = measured_peak()
    achieved = algo_bytes * n_nodes * n_pods / (step_ms / 1000.0) / 1e9
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup),
        "ms_per_step": step_ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "int64",
        "data": "synthetic",
        "config": {"workload": workload, "nodes": n_nodes, "pods": n_pods, "scheduled": scheduled, "fit_errors": fit_errors,
                   "parallelism": ("node table sharded over %d GPUs (row ranges), NVLink P2P exchange in-kernel" % world)
                   if world > 1 else "1 GPU",
                   "groups": groups, "group_restarts": restarts,
                   "l2": "L2 flushed (256 MiB memset) before every timed replay; the node table itself is L2-resident by design",
                   "placement_hash": sorted(hashes)[0], "result_hash": result_hash},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": e2e_step_ms},
        # ksb_schedule_batch_kernel launches inside the timed steps, as counted by the engine: one per 4,096-pod chunk
        # of each end-to-end pass plus one per resident replay
        "gpu_launches": launches,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": ncu_traffic(args.config, n_nodes, n_pods) if world == 1 else None, "peak_source": peak_src, "kernel": "ksb_schedule_batch_kernel",
                     "algorithmic_bytes_per_eval": algo_bytes, "evals_per_launch": n_nodes * n_pods,
                     "kernel_ms_in_e2e": sum(kernel_ms) / len(kernel_ms)},
        "clocks": clocks,
    }
    if world > 1:
        line["matches_unsharded"] = matches_unsharded
    if world == 1:
        # p99 per-pod latency (BASELINE.json's metric)
        line["latency"] = one_pod_latency(sess, local_rank, n_pods)
        b64 = batch_latency(sess, local_rank, n_pods)
        if b64:
            line["latency"]["batch64"] = b64
        if args.config == 4 and not args.no_side_config:
            line["cfg3"] = side_record(3, local_rank, args.steps)
    if not args.no_cpu_baseline:
        sample = args.cpu_sample or pick_sample(n_nodes)
        pps, secs, _ = oracle_baseline(cluster, sample, workers)
        line["cpu_baseline"] = {"value": pps, "unit": UNIT, "cores": workers, "kind": "port",
                                "sample": "first %d pods of the same stream on all %d nodes, C++ restatement of the Go "
                                          "scheduler (literal mode), %d worker threads, %.1f s" % (sample, n_nodes, workers, secs)}
    json_out.write(json.dumps(line) + "\n")
    json_out.flush()
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
