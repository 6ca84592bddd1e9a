"""Seeded synthetic clusters and pod streams for the five BASELINE.json configs
(SURVEY.md §8d). The generator is splitmix64 with seed 0xB2000000 + config number and
draws in a fixed, documented order, so the oracle and the GPU engine always see the
same objects. Output is the JSON interchange format of csrc/host/model.hpp.

Shapes follow the reference's own perf fixtures where they exist
(test/integration/scheduler_perf/scheduler_test.go:49-68 fake nodes,
test/utils/runners.go:979-997 pods) and the kubemark scale named in BASELINE.json.
"""
import json

MASK = (1 << 64) - 1
GIB = 1 << 30
MIB = 1 << 20
ZONE = "failure-domain.beta.kubernetes.io/zone"
REGION = "failure-domain.beta.kubernetes.io/region"
HOSTNAME = "kubernetes.io/hostname"
GPU = "nvidia.com/gpu"

CONFIG_SIZES = {1: (100, 1000), 2: (1000, 10000), 3: (5000, 30000), 4: (50000, 100000), 5: (5000, 20000)}
# algorithmic bytes per (pod,node) evaluation, SURVEY.md §8d / BASELINE.md §4
ALGO_BYTES = {1: 68, 2: 120, 3: 136, 4: 136, 5: 108}


class SplitMix64:
    def __init__(self, seed):
        self.s = seed & MASK

    def next(self):
        self.s = (self.s + 0x9E3779B97F4A7C15) & MASK
        z = self.s
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & MASK
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & MASK
        return z ^ (z >> 31)

    def below(self, n):
        return self.next() % n

    def chance(self, pct):  # pct in percent, integer or float with 0.1 resolution
        return self.below(1000) < int(pct * 10)

    def pick(self, seq):
        return seq[self.below(len(seq))]


def _ready(status="True"):
    return [{"type": "Ready", "status": status}]


def _node_name(i, padded=True):
    return ("node-%06d" % i) if padded else ("node-%d" % i)


def _basic_pod(rng, i, ns="default"):
    cpu = 100 * (1 + rng.below(10))
    mem = 128 * MIB * (1 + rng.below(32))
    p = {"name": "pod-%06d" % i, "namespace": ns}
    if rng.chance(10):
        p["containers"] = [{"requests": {}}]
    else:
        p["containers"] = [{"requests": {"cpu": cpu, "memory": mem}}]
    return p


def gen_cfg1(n_nodes, n_pods, rng, padded=True):
    nodes = []
    for i in range(n_nodes):
        nodes.append({"name": _node_name(i, padded), "labels": {HOSTNAME: _node_name(i, padded)},
                      "allocatable": {"cpu": rng.pick([4000, 8000, 16000]), "memory": rng.pick([16, 32, 64]) * GIB,
                                      "pods": 110},
                      "conditions": _ready()})
    pods = [_basic_pod(rng, i) for i in range(n_pods)]
    return {"nodes": nodes, "pods": pods}


def gen_cfg2(n_nodes, n_pods, rng, hetero_pct=0):
    mems = ["16384", "32768", "81920"]
    ccs = ["7.0", "8.0", "9.0"]
    eccs = ["true", "false"]
    nodes = []
    for i in range(n_nodes):
        name = _node_name(i)
        cls = {"nvidia.com/memory": rng.pick(mems), "nvidia.com/compute-capabilities": rng.pick(ccs),
               "nvidia.com/ecc": rng.pick(eccs)}
        hetero = hetero_pct and rng.chance(hetero_pct)
        devs = []
        for k in range(16):
            attrs = dict(cls)
            if hetero and k >= 8:
                attrs["nvidia.com/memory"] = rng.pick(mems)
                attrs["nvidia.com/ecc"] = rng.pick(eccs)
            devs.append({"id": "GPU-%06d-%02d" % (i, k), "health": "Healthy", "attributes": attrs})
        nodes.append({"name": name, "labels": {HOSTNAME: name},
                      "allocatable": {"cpu": 128000, "memory": 1024 * GIB, "pods": 110, GPU: 16},
                      "conditions": _ready(), "extendedResources": {GPU: devs}})
    pods = []
    for i in range(n_pods):
        r = rng.below(10)
        count = 1 if r < 7 else (2 if r < 9 else 4)
        required = []
        if not rng.chance(20):
            if rng.chance(50):
                required.append({"key": "nvidia.com/memory", "op": "Gt", "values": [rng.pick(["16000", "32000", "80000"])]})
            if rng.chance(30):
                required.append({"key": "nvidia.com/ecc", "op": "In", "values": ["true"]})
            if rng.chance(40) or not required:
                k = 1 + rng.below(3)
                start = rng.below(3)
                required.append({"key": "nvidia.com/compute-capabilities", "op": "In",
                                 "values": sorted({ccs[(start + j) % 3] for j in range(k)})})
        pods.append({"name": "pod-%06d" % i, "namespace": "default",
                     "containers": [{"requests": {"cpu": 4000 * count, "memory": 32 * GIB * count, GPU: count}}],
                     "extendedResources": [{"name": "gpu", "limits": {GPU: count}, "required": required}]})
    return {"nodes": nodes, "pods": pods}


LABEL_KEYS = ["lk%02d" % i for i in range(20)]
TAINT_KEYS = ["tk%d" % i for i in range(4)]


def _kubemark_nodes(n_nodes, rng):
    nodes = []
    for i in range(n_nodes):
        name = _node_name(i)
        labels = {HOSTNAME: name}
        if not rng.chance(10):
            r = rng.below(3)
            labels[REGION] = "region-%d" % r
            labels[ZONE] = "zone-%d-%d" % (r, rng.below(3))
        for _ in range(6):
            labels[rng.pick(LABEL_KEYS)] = "v%d" % rng.below(8)
        labels["disk-gb"] = str(100 * (1 + rng.below(20)))
        n = {"name": name, "labels": labels, "allocatable": {"cpu": 32000, "memory": 128 * GIB, "pods": 110}}
        taints = []
        if rng.chance(10):
            taints.append({"key": rng.pick(TAINT_KEYS), "value": "tv%d" % rng.below(2), "effect": "NoSchedule"})
        if rng.chance(10):
            taints.append({"key": rng.pick(TAINT_KEYS), "value": "tv%d" % rng.below(2), "effect": "PreferNoSchedule"})
        if taints:
            n["taints"] = taints
        if rng.chance(1):
            n["unschedulable"] = True
        conds = _ready("False" if rng.chance(1) else "True")
        if rng.chance(2):
            conds.append({"type": "MemoryPressure", "status": "True"})
        n["conditions"] = conds
        nodes.append(n)
    return nodes


def _node_expr(rng):
    op = rng.pick(["In", "In", "NotIn", "Exists", "DoesNotExist", "Gt", "Lt"])
    if op in ("Gt", "Lt"):
        return {"key": "disk-gb", "op": op, "values": [str(100 * (1 + rng.below(20)))]}
    key = rng.pick(LABEL_KEYS)
    if op in ("Exists", "DoesNotExist"):
        return {"key": key, "op": op, "values": []}
    k = 1 + rng.below(4)
    start = rng.below(8)
    return {"key": key, "op": op, "values": ["v%d" % ((start + j) % 8) for j in range(k)]}


def gen_cfg3(n_nodes, n_pods, rng, n_services=200):
    nodes = _kubemark_nodes(n_nodes, rng)
    services = [{"namespace": "default", "name": "svc-%d" % k, "selector": {"app": "svc-%d" % k}} for k in range(n_services)]
    pods = []
    for i in range(n_pods):
        p = _basic_pod(rng, i)
        p["labels"] = {"app": "svc-%d" % (i % n_services)}
        if rng.chance(5):
            p["containers"] = [{"requests": {}}]  # BestEffort: no requests, no limits
        if rng.chance(30):
            sel = {}
            for _ in range(1 + rng.below(2)):
                sel[rng.pick(LABEL_KEYS)] = "v%d" % rng.below(8)
            p["nodeSelector"] = sel
        node_aff = {}
        if rng.chance(20):
            node_aff["required"] = [{"matchExpressions": [_node_expr(rng) for _ in range(1 + rng.below(3))]}
                                    for _ in range(1 + rng.below(2))]
        if rng.chance(20):
            node_aff["preferred"] = [{"weight": 1 + rng.below(100),
                                      "matchExpressions": [_node_expr(rng) for _ in range(1 + rng.below(2))]}
                                     for _ in range(1 + rng.below(3))]
        if node_aff:
            p["affinity"] = {"node": node_aff}
        if rng.chance(15):
            tols = []
            if rng.chance(5):
                tols.append({"key": "", "operator": "Exists", "value": "", "effect": ""})
            else:
                for _ in range(1 + rng.below(2)):
                    if rng.chance(50):
                        tols.append({"key": rng.pick(TAINT_KEYS), "operator": "Exists", "value": "",
                                     "effect": rng.pick(["", "NoSchedule", "PreferNoSchedule"])})
                    else:
                        tols.append({"key": rng.pick(TAINT_KEYS), "operator": "Equal", "value": "tv%d" % rng.below(2),
                                     "effect": rng.pick(["", "NoSchedule", "PreferNoSchedule"])})
            p["tolerations"] = tols
        if rng.chance(2):
            port = {"hostPort": 30000 + rng.below(50), "protocol": rng.pick(["TCP", "TCP", "UDP"])}
            if rng.chance(50):
                port["hostIP"] = "127.0.0.1"
            p["containers"][0]["ports"] = [port]
        pods.append(p)
    return {"nodes": nodes, "services": services, "pods": pods}


def gen_cfg5(n_nodes, n_pods, rng, n_apps=100):
    nodes = []
    for i in range(n_nodes):
        name = _node_name(i)
        r = rng.below(3)
        labels = {HOSTNAME: name, REGION: "region-%d" % r, ZONE: "zone-%d-%d" % (r, rng.below(3))}
        nodes.append({"name": name, "labels": labels, "allocatable": {"cpu": 32000, "memory": 128 * GIB, "pods": 110},
                      "conditions": _ready()})
    pods = []
    for i in range(n_pods):
        p = _basic_pod(rng, i)
        app = rng.below(n_apps)
        p["labels"] = {"app": "app-%d" % app}
        own = {"matchLabels": {"app": "app-%d" % app}}
        partner = {"matchLabels": {"app": "app-%d" % ((app + 1) % n_apps)}}
        r = rng.below(100)
        aff = {}
        if r < 40:
            aff["podAntiAffinity"] = {"required": [{"selector": own, "topologyKey": HOSTNAME}]}
        elif r < 60:
            aff["podAffinity"] = {"required": [{"selector": partner, "topologyKey": ZONE}]}
        elif r < 90:
            kind = "podAffinity" if rng.chance(50) else "podAntiAffinity"
            terms = []
            for _ in range(1 + rng.below(2)):
                terms.append({"weight": 1 + rng.below(100),
                              "term": {"selector": partner if rng.chance(50) else own,
                                       "topologyKey": ZONE if rng.chance(50) else HOSTNAME}})
            aff[kind] = {"preferred": terms}
        if aff:
            p["affinity"] = aff
        pods.append(p)
    return {"nodes": nodes, "pods": pods}


def generate(cfg, n_nodes=None, n_pods=None, seed=None, **kw):
    """Return the cluster dict for BASELINE.json configs[cfg-1]; sizes may be reduced for tests."""
    dn, dp = CONFIG_SIZES[cfg]
    n_nodes = dn if n_nodes is None else n_nodes
    n_pods = dp if n_pods is None else n_pods
    rng = SplitMix64((0xB2000000 + cfg) if seed is None else seed)
    if cfg == 1:
        return gen_cfg1(n_nodes, n_pods, rng, **kw)
    if cfg == 2:
        return gen_cfg2(n_nodes, n_pods, rng, **kw)
    if cfg in (3, 4):
        return gen_cfg3(n_nodes, n_pods, rng, **kw)
    if cfg == 5:
        return gen_cfg5(n_nodes, n_pods, rng, **kw)
    raise ValueError("unknown config %r" % cfg)


def write(cluster, path):
    with open(path, "w") as f:
        json.dump(cluster, f, separators=(",", ":"))
    return path
