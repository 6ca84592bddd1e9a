"""More transcriptions of the reference's tests (see build_fixtures.py). Each builder cites its source."""


def builders(g):
    node, pod, ctr, expr = g["node"], g["pod"], g["ctr"], g["expr"]
    priority, predicate, CASES = g["priority"], g["predicate"], g["CASES"]
    ZONE, REGION, HOSTNAME = g["ZONE"], g["REGION"], g["HOSTNAME"]

    def sel(*exprs, **match_labels):
        s = {}
        if match_labels:
            s["matchLabels"] = match_labels
        if exprs:
            s["matchExpressions"] = list(exprs)
        return s

    def term(selector, key, namespaces=None):
        t = {"selector": selector, "topologyKey": key}
        if namespaces is not None:
            t["namespaces"] = namespaces
        return t

    def wterm(weight, selector, key):
        return {"weight": weight, "term": term(selector, key)}

    # ----------------------------------------------------------------- InterPodAffinityPriority
    # algorithm/priorities/interpod_affinity_test.go:42-527
    def build_interpod_priority():
        src = "algorithm/priorities/interpod_affinity_test.go:270-507"
        rgChina, rgIndia = {"region": "China"}, {"region": "India"}
        az1, az2 = {"az": "az1"}, {"az": "az2"}
        rgChinaAz1 = {"region": "China", "az": "az1"}
        S1, S2 = {"security": "S1"}, {"security": "S2"}
        stayS1Region = {"podAffinity": {"preferred": [wterm(5, sel(expr("security", "In", "S1")), "region")]}}
        stayS2Region = {"podAffinity": {"preferred": [wterm(6, sel(expr("security", "In", "S2")), "region")]}}
        affinity3 = {"podAffinity": {"preferred": [
            wterm(8, sel(expr("security", "NotIn", "S1"), expr("security", "In", "S2")), "region"),
            wterm(2, sel(expr("security", "Exists"), expr("wrongkey", "DoesNotExist")), "region")]}}
        hardAffinity = {"podAffinity": {"required": [
            term(sel(expr("security", "In", "S1", "value2")), "region"),
            term(sel(expr("security", "Exists"), expr("wrongkey", "DoesNotExist")), "region")]}}
        awayS1Az = {"podAntiAffinity": {"preferred": [wterm(5, sel(expr("security", "In", "S1")), "az")]}}
        awayS2Az = {"podAntiAffinity": {"preferred": [wterm(5, sel(expr("security", "In", "S2")), "az")]}}
        stayS1AwayS2 = {"podAffinity": {"preferred": [wterm(8, sel(expr("security", "In", "S1")), "region")]},
                        "podAntiAffinity": {"preferred": [wterm(5, sel(expr("security", "In", "S2")), "az")]}}
        N = lambda n, l: node(n, labels=l)
        P = lambda n, l=None, a=None: pod(n, labels=l, **({"affinity": a} if a else {}))
        I = "InterPodAffinityPriority"
        m = ["machine%d" % i for i in range(6)]
        priority(I, src, "all machines are same priority as Affinity is nil", pod(labels=S1),
                 [N(m[1], rgChina), N(m[2], rgIndia), N(m[3], az1)], [], [(m[1], 0), (m[2], 0), (m[3], 0)])
        priority(I, src, "Affinity: pod that matches topology key & pods in nodes gets high score",
                 pod(labels=S1, affinity=stayS1Region), [N(m[1], rgChina), N(m[2], rgIndia), N(m[3], az1)],
                 [P(m[1], S1), P(m[2], S2), P(m[3], S1)], [(m[1], 10), (m[2], 0), (m[3], 0)])
        priority(I, src, "All nodes with same topology key & value as a node with a matching pod have the same score",
                 pod(affinity=stayS1Region), [N(m[1], rgChina), N(m[2], rgChinaAz1), N(m[3], rgIndia)], [P(m[1], S1)],
                 [(m[1], 10), (m[2], 10), (m[3], 0)])
        priority(I, src, "Affinity: region with more matching pods gets high score", pod(labels=S1, affinity=stayS2Region),
                 [N(m[1], rgChina), N(m[2], rgIndia), N(m[3], rgChina), N(m[4], rgChina), N(m[5], rgIndia)],
                 [P(m[1], S2), P(m[1], S2), P(m[2], S2), P(m[3], S2), P(m[4], S2), P(m[5], S2)],
                 [(m[1], 10), (m[2], 5), (m[3], 10), (m[4], 10), (m[5], 5)])
        priority(I, src, "Affinity: different Label operators and values, including some match failures",
                 pod(labels=S1, affinity=affinity3), [N(m[1], rgChina), N(m[2], rgIndia), N(m[3], az1)],
                 [P(m[1], S1), P(m[2], S2), P(m[3], S1)], [(m[1], 2), (m[2], 10), (m[3], 0)])
        priority(I, src, "Affinity symmetry: preferredDuringScheduling in pod affinity symmetry", pod(labels=S2),
                 [N(m[1], rgChina), N(m[2], rgIndia), N(m[3], az1)],
                 [P(m[1], S1, stayS1Region), P(m[2], S2, stayS2Region)], [(m[1], 0), (m[2], 10), (m[3], 0)])
        priority(I, src, "Affinity symmetry: RequiredDuringScheduling in pod affinity symmetry", pod(labels=S1),
                 [N(m[1], rgChina), N(m[2], rgIndia), N(m[3], az1)],
                 [P(m[1], S1, hardAffinity), P(m[2], S2, hardAffinity)], [(m[1], 10), (m[2], 10), (m[3], 0)])
        priority(I, src, "Anti Affinity: pod that does not match existing pods in node gets high score",
                 pod(labels=S1, affinity=awayS1Az), [N(m[1], az1), N(m[2], rgChina)], [P(m[1], S1), P(m[2], S2)],
                 [(m[1], 0), (m[2], 10)])
        priority(I, src, "Anti Affinity: node without topology key scores higher", pod(labels=S1, affinity=awayS1Az),
                 [N(m[1], az1), N(m[2], rgChina)], [P(m[1], S1), P(m[2], S1)], [(m[1], 0), (m[2], 10)])
        priority(I, src, "Anti Affinity: node with more unmatches gets high score", pod(labels=S1, affinity=awayS1Az),
                 [N(m[1], az1), N(m[2], rgIndia)], [P(m[1], S1), P(m[1], S1), P(m[2], S2)], [(m[1], 0), (m[2], 10)])
        priority(I, src, "Anti Affinity symmetry", pod(labels=S2), [N(m[1], az1), N(m[2], az2)],
                 [P(m[1], S1, awayS2Az), P(m[2], S2, awayS1Az)], [(m[1], 0), (m[2], 10)])
        priority(I, src, "Affinity and Anti Affinity: preferred only", pod(labels=S1, affinity=stayS1AwayS2),
                 [N(m[1], rgChina), N(m[2], az1)], [P(m[1], S1), P(m[2], S1)], [(m[1], 10), (m[2], 0)])
        priority(I, src, "Affinity and Anti Affinity: same labels", pod(labels=S1, affinity=stayS1AwayS2),
                 [N(m[1], rgChinaAz1), N(m[2], rgIndia), N(m[3], rgChina), N(m[4], rgChina), N(m[5], rgIndia)],
                 [P(m[1], S1), P(m[1], S1), P(m[2], S1), P(m[3], S1), P(m[3], S1), P(m[4], S1), P(m[5], S1)],
                 [(m[1], 10), (m[2], 4), (m[3], 10), (m[4], 10), (m[5], 4)])
        priority(I, src, "Affinity and Anti Affinity and symmetry", pod(labels=S1, affinity=stayS1AwayS2),
                 [N(m[1], rgChina), N(m[2], az1), N(m[3], rgIndia), N(m[4], az2)],
                 [P(m[1], S1), P(m[2], S2), P(m[3], None, stayS1AwayS2), P(m[4], None, awayS1Az)],
                 [(m[1], 10), (m[2], 0), (m[3], 10), (m[4], 0)])

    # ----------------------------------------------------------------- PodFitsResources
    # algorithm/predicates/predicates_test.go:94-456
    EXT_A, EXT_B, HUGE_A = "example.com/aaa", "example.com/bbb", "hugepages-2Mi"

    def R(cpu=0, mem=0, eph=0, **scalars):
        # schedulercache.Resource.ResourceList(): always emits the five standard keys
        r = {"cpu": cpu, "memory": mem, "alpha.kubernetes.io/nvidia-gpu": 0, "pods": 0, "ephemeral-storage": eph}
        for k, v in scalars.items():
            r[{"a": EXT_A, "b": EXT_B, "h": HUGE_A}[k]] = v
        return {"requests": r}

    def rpod(*usage, init=None):
        kw = {}
        if init is not None:
            kw["initContainers"] = list(init)
        return pod(containers=list(usage), **kw)

    def build_fits_resources():
        src = "algorithm/predicates/predicates_test.go:94-456"
        alloc = lambda cpu, mem, gpu, pods, a, st, h: {
            "cpu": cpu, "memory": mem, "alpha.kubernetes.io/nvidia-gpu": gpu, "pods": pods, EXT_A: a,
            "ephemeral-storage": st, HUGE_A: h}
        big = lambda: node("n", alloc=alloc(10, 20, 0, 32, 5, 20, 5))
        I = lambda r: "Insufficient " + r
        T = [
            (pod(), R(10, 20), True, [], "no resources requested always fits"),
            (rpod(R(1, 1)), R(10, 20), False, [I("cpu"), I("memory")], "too many resources fails"),
            (rpod(R(1, 1), init=[R(3, 1)]), R(8, 19), False, [I("cpu")], "too many resources fails due to init container cpu"),
            (rpod(R(1, 1), init=[R(3, 1), R(2, 1)]), R(8, 19), False, [I("cpu")], "fails due to highest init container cpu"),
            (rpod(R(1, 1), init=[R(1, 3)]), R(9, 19), False, [I("memory")], "fails due to init container memory"),
            (rpod(R(1, 1), init=[R(1, 3), R(1, 2)]), R(9, 19), False, [I("memory")], "fails due to highest init container memory"),
            (rpod(R(1, 1), init=[R(1, 1)]), R(9, 19), True, [], "init container fits because it's the max, not sum"),
            (rpod(R(1, 1), init=[R(1, 1), R(1, 1)]), R(9, 19), True, [], "multiple init containers fit"),
            (rpod(R(1, 1)), R(5, 5), True, [], "both resources fit"),
            (rpod(R(2, 1)), R(9, 5), False, [I("cpu")], "one resource memory fits"),
            (rpod(R(1, 2)), R(5, 19), False, [I("memory")], "one resource cpu fits"),
            (rpod(R(5, 1)), R(5, 19), True, [], "equal edge case"),
            (rpod(R(4, 1), init=[R(5, 1)]), R(5, 19), True, [], "equal edge case for init container"),
            (rpod(R(a=1)), R(), True, [], "extended resource fits"),
            (rpod(R(), init=[R(a=1)]), R(), True, [], "extended resource fits for init container"),
            (rpod(R(1, 1, a=10)), R(0, 0, a=0), False, [I(EXT_A)], "extended resource capacity enforced"),
            (rpod(R(), init=[R(1, 1, a=10)]), R(0, 0, a=0), False, [I(EXT_A)], "extended resource capacity enforced for init container"),
            (rpod(R(1, 1, a=1)), R(0, 0, a=5), False, [I(EXT_A)], "extended resource allocatable enforced"),
            (rpod(R(), init=[R(1, 1, a=1)]), R(0, 0, a=5), False, [I(EXT_A)], "extended resource allocatable enforced for init container"),
            (rpod(R(1, 1, a=3), R(1, 1, a=3)), R(0, 0, a=2), False, [I(EXT_A)], "extended resource allocatable enforced for multiple containers"),
            (rpod(R(), init=[R(1, 1, a=3), R(1, 1, a=3)]), R(0, 0, a=2), True, [], "extended resource allocatable admits multiple init containers"),
            (rpod(R(), init=[R(1, 1, a=6), R(1, 1, a=3)]), R(0, 0, a=2), False, [I(EXT_A)], "extended resource allocatable enforced for multiple init containers"),
            (rpod(R(1, 1, b=1)), R(0, 0), False, [I(EXT_B)], "extended resource allocatable enforced for unknown resource"),
            (rpod(R(), init=[R(1, 1, b=1)]), R(0, 0), False, [I(EXT_B)], "unknown resource for init container"),
            (rpod(R(1, 1, h=10)), R(0, 0, h=0), False, [I(HUGE_A)], "hugepages resource capacity enforced"),
            (rpod(R(), init=[R(1, 1, h=10)]), R(0, 0, h=0), False, [I(HUGE_A)], "hugepages resource capacity enforced for init container"),
            (rpod(R(1, 1, h=3), R(1, 1, h=3)), R(0, 0, h=2), False, [I(HUGE_A)], "hugepages resource allocatable enforced for multiple containers"),
        ]
        for p, existing, fits, reasons, test in T:
            predicate("PodFitsResources", src, test, p, big(), [pod(containers=[existing])], fits, reasons)
        small = lambda: node("n", alloc=alloc(10, 20, 0, 1, 0, 0, 0))
        for p, existing, test in [
            (pod(), R(10, 20), "fails when there's no space for additional pod"),
            (rpod(R(1, 1)), R(5, 5), "both resources fit but no space for additional pod"),
            (rpod(R(5, 1)), R(5, 19), "equal edge case but no space for additional pod"),
            (rpod(R(5, 1), init=[R(5, 1)]), R(5, 19), "equal edge case, init container, no space for additional pod"),
        ]:
            predicate("PodFitsResources", src, test, p, small(), [pod(containers=[existing])], False, [I("pods")])
        for p, existing, fits, reasons, test in [
            (rpod(R(1, 1)), R(10, 10), False, [I("cpu")], "due to container scratch disk"),
            (rpod(R(1, 1)), R(2, 10), True, [], "pod fit"),
            (rpod(R(eph=25)), R(2, 2), False, [I("ephemeral-storage")], "ephemeral storage request exceeds allocatable"),
            (rpod(R(eph=10)), R(2, 2), True, [], "pod fits (storage)"),
        ]:
            predicate("PodFitsResources", src, test, p, big(), [pod(containers=[existing])], fits, reasons)

    # ----------------------------------------------------------------- PodFitsHostPorts
    # algorithm/predicates/predicates_test.go:542-655
    def ports_pod(*infos):
        ports = []
        for i in infos:
            proto, ip, port = i.split("/")
            ports.append({"protocol": proto, "hostIP": ip, "hostPort": int(port)})
        return pod(containers=[{"ports": ports}])

    def build_host_ports():
        src = "algorithm/predicates/predicates_test.go:542-655"
        T = [
            (pod(), [], True, "nothing running"),
            (ports_pod("UDP/127.0.0.1/8080"), ["UDP/127.0.0.1/9090"], True, "other port"),
            (ports_pod("UDP/127.0.0.1/8080"), ["UDP/127.0.0.1/8080"], False, "same udp port"),
            (ports_pod("TCP/127.0.0.1/8080"), ["TCP/127.0.0.1/8080"], False, "same tcp port"),
            (ports_pod("TCP/127.0.0.1/8080"), ["TCP/127.0.0.2/8080"], True, "different host ip"),
            (ports_pod("UDP/127.0.0.1/8080"), ["TCP/127.0.0.1/8080"], True, "different protocol"),
            (ports_pod("UDP/127.0.0.1/8000", "UDP/127.0.0.1/8080"), ["UDP/127.0.0.1/8080"], False, "second udp port conflict"),
            (ports_pod("TCP/127.0.0.1/8001", "UDP/127.0.0.1/8080"), ["TCP/127.0.0.1/8001", "UDP/127.0.0.1/8081"], False,
             "first tcp port conflict"),
            (ports_pod("TCP/0.0.0.0/8001"), ["TCP/127.0.0.1/8001"], False, "first tcp port conflict due to 0.0.0.0 hostIP"),
            (ports_pod("TCP/10.0.10.10/8001", "TCP/0.0.0.0/8001"), ["TCP/127.0.0.1/8001"], False,
             "TCP hostPort conflict due to 0.0.0.0 hostIP"),
            (ports_pod("TCP/127.0.0.1/8001"), ["TCP/0.0.0.0/8001"], False, "second tcp port conflict to 0.0.0.0 hostIP"),
            (ports_pod("UDP/127.0.0.1/8001"), ["TCP/0.0.0.0/8001"], True, "second different protocol"),
            (ports_pod("UDP/127.0.0.1/8001"), ["TCP/0.0.0.0/8001", "UDP/0.0.0.0/8001"], False,
             "UDP hostPort conflict due to 0.0.0.0 hostIP"),
        ]
        for p, existing, fits, test in T:
            predicate("PodFitsHostPorts", src, test, p, node("m1"), [ports_pod(*existing)] if existing else [], fits,
                      [] if fits else ["PodFitsHostPorts"])

    # ----------------------------------------------------------------- PodMatchNodeSelector
    # algorithm/predicates/predicates_test.go:916-1385
    def build_fits_selector():
        src = "algorithm/predicates/predicates_test.go:916-1385"
        req = lambda *terms: {"node": {"required": [{"matchExpressions": list(t)} for t in terms]}}
        foo = {"foo": "bar"}
        T = [
            (pod(), None, True, "no selector"),
            (pod(nodeSelector=foo), None, False, "missing labels"),
            (pod(nodeSelector=foo), foo, True, "same labels"),
            (pod(nodeSelector=foo), {"foo": "bar", "baz": "blah"}, True, "node labels are superset"),
            (pod(nodeSelector={"foo": "bar", "baz": "blah"}), foo, False, "node labels are subset"),
            (pod(affinity=req([expr("foo", "In", "bar", "value2")])), foo, True, "In operator that matches"),
            (pod(affinity=req([expr("kernel-version", "Gt", "0204")])), {"kernel-version": "0206"}, True, "Gt operator that matches"),
            (pod(affinity=req([expr("mem-type", "NotIn", "DDR", "DDR2")])), {"mem-type": "DDR3"}, True, "NotIn operator that matches"),
            (pod(affinity=req([expr("GPU", "Exists")])), {"GPU": "NVIDIA-GRID-K1"}, True, "Exists operator that matches"),
            (pod(affinity=req([expr("foo", "In", "value1", "value2")])), foo, False, "affinity that doesn't match node's labels"),
            (pod(affinity={"node": {"required": []}}), foo, False, "nil []NodeSelectorTerm matches nothing"),
            # (a nil and an empty slice are the same record on this wire: a term count of 0; the reference has one case each)
            (pod(affinity={"node": {"required": []}}), foo, False, "empty []NodeSelectorTerm matches nothing"),
            (pod(affinity=req([])), foo, False, "empty MatchExpressions matches no objects"),
            (pod(), foo, True, "Pod with no Affinity will schedule onto a node"),
            (pod(affinity={"node": {"required": None}}), foo, True, "Affinity but nil NodeSelector will schedule onto a node"),
            (pod(affinity=req([expr("GPU", "Exists"), expr("GPU", "NotIn", "AMD", "INTER")])), {"GPU": "NVIDIA-GRID-K1"}, True,
             "multiple matchExpressions ANDed that matches"),
            (pod(affinity=req([expr("GPU", "Exists"), expr("GPU", "In", "AMD", "INTER")])), {"GPU": "NVIDIA-GRID-K1"}, False,
             "multiple matchExpressions ANDed that doesn't match"),
            (pod(affinity=req([expr("foo", "In", "bar", "value2")], [expr("diffkey", "In", "wrong", "value2")])), foo, True,
             "multiple NodeSelectorTerms ORed"),
            (pod(nodeSelector=foo, affinity=req([expr("foo", "Exists")])), foo, True, "Affinity and PodSpec.NodeSelector both satisfied"),
            (pod(nodeSelector=foo, affinity=req([expr("foo", "Exists")])), {"foo": "barrrrrr"}, False,
             "Affinity matches but PodSpec.NodeSelector is not satisfied"),
        ]
        for p, labels, fits, test in T:
            predicate("PodMatchNodeSelector", src, test, p, node("n", labels=labels), [], fits,
                      [] if fits else ["MatchNodeSelector"])

    # ----------------------------------------------------------------- PodToleratesNodeTaints
    # algorithm/predicates/predicates_test.go:3215-3410
    def build_tolerates_taints():
        src = "algorithm/predicates/predicates_test.go:3215-3410"
        Tn = lambda k, v, e: {"key": k, "value": v, "effect": e}
        Tl = lambda k, op, v, e: {"key": k, "operator": op, "value": v, "effect": e}
        NS, PNS = "NoSchedule", "PreferNoSchedule"
        T = [
            ([], [Tn("dedicated", "user1", NS)], False, "no tolerations, node with nonempty taints"),
            ([Tl("dedicated", "", "user1", NS)], [Tn("dedicated", "user1", NS)], True, "dedicated node assigned to user1"),
            ([Tl("dedicated", "Equal", "user2", NS)], [Tn("dedicated", "user1", NS)], False, "dedicated node assigned to user2"),
            ([Tl("foo", "Exists", "", NS)], [Tn("foo", "bar", NS)], True, "operator Exists tolerates the taints"),
            ([Tl("dedicated", "Equal", "user2", NS), Tl("foo", "Exists", "", NS)],
             [Tn("dedicated", "user2", NS), Tn("foo", "bar", NS)], True, "multiple tolerations, multiple taints, all tolerated"),
            ([Tl("foo", "Equal", "bar", PNS)], [Tn("foo", "bar", NS)], False, "non-empty effect doesn't match"),
            ([Tl("foo", "Equal", "bar", "")], [Tn("foo", "bar", NS)], True, "empty toleration effect matches NoSchedule"),
            ([Tl("dedicated", "Equal", "user2", NS)], [Tn("dedicated", "user1", PNS)], True, "taint effect is PreferNoSchedule"),
            ([], [Tn("dedicated", "user1", PNS)], True, "no toleration, taint effect PreferNoSchedule"),
        ]
        for tols, taints, fits, test in T:
            predicate("PodToleratesNodeTaints", src, test, pod(tolerations=tols), node("n", taints=taints), [], fits,
                      [] if fits else ["PodToleratesNodeTaints"])

    # ----------------------------------------------------------------- MatchInterPodAffinity
    # algorithm/predicates/predicates_test.go:2162-2775 and :2777-3213
    def build_interpod_predicate():
        src = "algorithm/predicates/predicates_test.go:2162-2775"
        podLabel, podLabel2 = {"service": "securityscan"}, {"security": "S1"}
        n1 = lambda: node("machine1", labels={"region": "r1", "zone": "z11"})
        aff = lambda *terms: {"required": list(terms)}
        svc_in = lambda *v: sel(expr("service", "In", *v))
        A, B, C = "MatchInterPodAffinity", "PodAffinityRulesNotMatch", "PodAntiAffinityRulesNotMatch"
        D = "ExistingPodsAntiAffinityRulesNotMatch"
        M = "MatchInterPodAffinity"
        ex = lambda **kw: [pod("machine1", labels=podLabel, **kw)]
        predicate(M, src, "no required pod affinity rules, no existing pods", pod(), n1(), [], True, [])
        predicate(M, src, "PodAffinity In operator matches the existing pod",
                  pod(labels=podLabel2, affinity={"podAffinity": aff(term(svc_in("securityscan", "value2"), "region"))}),
                  n1(), ex(), True, [])
        predicate(M, src, "PodAffinity NotIn operator matches the existing pod",
                  pod(labels=podLabel2, affinity={"podAffinity": aff(term(sel(expr("service", "NotIn", "securityscan3", "value3")), "region"))}),
                  n1(), ex(), True, [])
        # (the next two terms carry no topologyKey in the reference's table, predicates_test.go:2246-2308: the predicate
        #  returns its reasons AND an error, :1293-1297; the Go test drops the error, a Schedule call would fail on it)
        predicate(M, src, "Does not satisfy the PodAffinity with labelSelector because of diff Namespace",
                  pod(labels=podLabel2, affinity={"podAffinity": aff(term(svc_in("securityscan", "value2"), "", ["DiffNameSpace"]))}),
                  n1(), ex(ns="ns"), False, [A, B])
        CASES[-1]["expect"]["error"] = "Cannot schedule pod onto node, because of PodAffinityTerm"
        predicate(M, src, "Doesn't satisfy the PodAffinity because of unmatching labelSelector with the existing pod",
                  pod(labels=podLabel, affinity={"podAffinity": aff(term(svc_in("antivirusscan", "value2"), ""))}),
                  n1(), ex(), False, [A, B])
        CASES[-1]["expect"]["error"] = "Cannot schedule pod onto node, because of PodAffinityTerm"
        multi_ok = aff(term(sel(expr("service", "Exists"), expr("wrongkey", "DoesNotExist")), "region"),
                       term(sel(expr("service", "In", "securityscan"), expr("service", "NotIn", "WrongValue")), "region"))
        predicate(M, src, "different label Operators in multiple required terms",
                  pod(labels=podLabel2, affinity={"podAffinity": multi_ok}), n1(), ex(), True, [])
        multi_bad = aff(term(sel(expr("service", "Exists"), expr("wrongkey", "DoesNotExist")), "region"),
                        term(sel(expr("service", "In", "securityscan2"), expr("service", "NotIn", "WrongValue")), "region"))
        predicate(M, src, "matchExpressions are ANDed; one item doesn't match",
                  pod(labels=podLabel2, affinity={"podAffinity": multi_bad}), n1(), ex(), False, [A, B])
        both = {"podAffinity": aff(term(svc_in("securityscan", "value2"), "region")),
                "podAntiAffinity": aff(term(svc_in("antivirusscan", "value2"), "node"))}
        predicate(M, src, "satisfies the PodAffinity and PodAntiAffinity with the existing pod",
                  pod(labels=podLabel2, affinity=both), n1(), ex(), True, [])
        predicate(M, src, "satisfies PodAffinity, PodAntiAffinity and PodAntiAffinity symmetry",
                  pod(labels=podLabel2, affinity=both), n1(),
                  ex(affinity={"podAntiAffinity": aff(term(svc_in("antivirusscan", "value2"), "node"))}), True, [])
        predicate(M, src, "satisfies the PodAffinity but not the PodAntiAffinity with the existing pod",
                  pod(labels=podLabel2, affinity={"podAffinity": aff(term(svc_in("securityscan", "value2"), "region")),
                                                  "podAntiAffinity": aff(term(svc_in("securityscan", "value2"), "zone"))}),
                  n1(), ex(), False, [A, C])
        predicate(M, src, "doesn't satisfy PodAntiAffinity symmetry with the existing pod",
                  pod(labels=podLabel, affinity=both), n1(),
                  ex(affinity={"podAntiAffinity": aff(term(svc_in("securityscan", "value2"), "zone"))}), False, [A, D])
        predicate(M, src, "pod matches its own Label in PodAffinity and that matches the existing pod Labels",
                  pod(labels=podLabel, affinity={"podAffinity": aff(term(sel(expr("service", "NotIn", "securityscan", "value2")), "region"))}),
                  n1(), [], False, [A, B],
                  other_nodes=[node("machine2", labels={"region": "r1", "zone": "z11"})],
                  other_pods=[pod("machine2", labels=podLabel)])
        predicate(M, src, "existing pod's PodAntiAffinity respected when pod has no constraints (violated)",
                  pod(labels=podLabel), n1(),
                  ex(affinity={"podAntiAffinity": aff(term(svc_in("securityscan", "value2"), "zone"))}), False, [A, D])
        predicate(M, src, "existing pod's PodAntiAffinity respected when pod has no constraints (satisfied)",
                  pod(labels=podLabel), n1(),
                  ex(affinity={"podAntiAffinity": aff(term(sel(expr("service", "NotIn", "securityscan", "value2")), "zone"))}),
                  True, [])

        src = "algorithm/predicates/predicates_test.go:2777-3213"
        rgChina, rgChinaAz1, rgIndia, rgUS = {"region": "China"}, {"region": "China", "az": "az1"}, {"region": "India"}, {"region": "US"}
        foo_in = lambda *v: sel(expr("foo", "In", *v))

        def multi(test, p, pods, nodes, expect):
            for i, (name, fits, reasons) in enumerate(expect):
                others = [n for n in nodes if n["name"] != name]
                me = [n for n in nodes if n["name"] == name][0]
                predicate(M, src, test + " [" + name + "]", p, me, [], fits, reasons, other_nodes=others, other_pods=pods)

        multi("same topology key & value as a node with a matching pod",
              pod(affinity={"podAffinity": aff(term(foo_in("bar"), "region"))}), [pod("machine1", labels={"foo": "bar"})],
              [node("machine1", labels=rgChina), node("machine2", labels=rgChinaAz1), node("machine3", labels=rgIndia)],
              [("machine1", True, []), ("machine2", True, []), ("machine3", False, [A, B])])
        multi("first pod of the collection is not blocked",
              pod(labels={"foo": "bar"}, affinity={"podAffinity": aff(term(foo_in("bar"), "zone"))}), [],
              [node("nodeA", labels={"zone": "az1", "hostname": "h1"}), node("nodeB", labels={"zone": "az2", "hostname": "h2"})],
              [("nodeA", True, []), ("nodeB", True, [])])
        multi("anti-affinity: nodeA and nodeB share the region of a matching pod",
              pod(affinity={"podAntiAffinity": aff(term(foo_in("abc"), "region"))}), [pod("nodeA", labels={"foo": "abc"})],
              [node("nodeA", labels={"region": "r1", "hostname": "nodeA"}), node("nodeB", labels={"region": "r1", "hostname": "nodeB"})],
              [("nodeA", False, [A, C]), ("nodeB", False, [A, C])])
        multi("anti-affinity: nodeC in another region is fine",
              pod(affinity={"podAntiAffinity": aff(term(foo_in("abc"), "region"))}), [pod("nodeA", labels={"foo": "abc"})],
              [node("nodeA", labels=rgChina), node("nodeB", labels=rgChinaAz1), node("nodeC", labels=rgIndia)],
              [("nodeA", False, [A, C]), ("nodeB", False, [A, C]), ("nodeC", True, [])])
        multi("anti-affinity both directions",
              pod(labels={"foo": "123"}, affinity={"podAntiAffinity": aff(term(foo_in("bar"), "region"))}),
              [pod("nodeA", labels={"foo": "bar"}),
               pod("nodeC", affinity={"podAntiAffinity": aff(term(foo_in("123"), "region"))})],
              [node("nodeA", labels=rgChina), node("nodeB", labels=rgChinaAz1), node("nodeC", labels=rgIndia), node("nodeD", labels=rgUS)],
              [("nodeA", False, [A, C]), ("nodeB", False, [A, C]), ("nodeC", False, [A, D]), ("nodeD", True, [])])
        multi("anti-affinity with namespaces",
              pod(labels={"foo": "123"}, ns="NS1", affinity={"podAntiAffinity": aff(term(foo_in("bar"), "region"))}),
              [pod("nodeA", labels={"foo": "bar"}, ns="NS1"),
               pod("nodeC", ns="NS2", affinity={"podAntiAffinity": aff(term(foo_in("123"), "region"))})],
              [node("nodeA", labels=rgChina), node("nodeB", labels=rgChinaAz1), node("nodeC", labels=rgIndia)],
              [("nodeA", False, [A, C]), ("nodeB", False, [A, C]), ("nodeC", True, [])])

    # ----------------------------------------------------------------- fork: device matching
    # core/extended_resources_test.go:84-294
    def build_devices():
        src = "core/extended_resources_test.go:84-137"
        MEM, CC, ECC = "nvidia.com/memory", "nvidia.com/compute-capabilities", "nvidia.com/ecc"
        dev = lambda i, mem="4096", cc="4.2", ecc="true": {"id": i, "health": "Healthy", "attributes": {MEM: mem, CC: cc, ECC: ecc}}
        for reqs, match, err, test in [
            ([expr(MEM, "Gt", "4095")], True, False, "memory Gt 4095"),
            ([expr(MEM, "Lt", "4095")], False, False, "memory Lt 4095"),
            ([expr(MEM, "Gt", "4095"), expr(ECC, "In", "true")], True, False, "memory Gt and ecc In true"),
            ([expr(MEM, "Gt", "4095"), expr(ECC, "In", "false")], False, False, "memory Gt and ecc In false"),
            ([expr(MEM, "Gt", "4095"), expr(ECC, "Gt", "false", "true")], False, True, "Gt with two values is an error"),
            ([], True, False, "empty affinity matches"),
        ]:
            CASES.append({"op": "device_match", "source": src, "test": test, "device": dev("someId1"), "required": reqs,
                          "expect": {"match": match, "error": err}})
        src = "core/extended_resources_test.go:193-294"
        R1, R2 = "nvidia.com/foo", "nvidia.com/bar"
        gt = [expr(MEM, "Gt", "4095")]
        devs = [dev("someId1"), dev("someId2"), dev("someId3", "2048"), dev("someId4", "2048")]
        er = lambda name, rname, n, required=None: {"name": name, "limits": {rname: n}, "required": required or []}

        def hx(test, ers, nres, fits, devices=None, reasons=None):
            c = {"op": "predicate", "name": "HasExtendedResources", "source": src, "test": test,
                 "cluster": {"nodes": [node("n", extendedResources=nres)], "existingPods": []},
                 "pod": pod(extendedResources=ers), "node": "n", "expect": {"fits": fits, "reasons": reasons}}
            if devices is not None:
                c["expect"]["devices"] = devices
            CASES.append(c)

        avail = {R1: [devs[2], devs[0], devs[1]], R2: [devs[2], devs[3], devs[0]]}
        NOT = "nvidia.com/baz"
        hx("an extended resource entry with no limits", [{"name": "", "limits": {}, "required": []}], avail, False, None,
           ["unexpected limits length: 0 != 1"])
        hx("an extended resource not available", [er("", NOT, 2)], avail, False, None, ["Insufficient " + NOT])
        hx("not enough resources", [er("", R1, 4)], avail, False, None, ["Insufficient " + R1])
        hx("successful request with no affinity", [er("foo", R1, 3)], avail, True, {"foo": ["someId1", "someId2", "someId3"]})
        hx("successful request with affinity", [er("foo", R1, 2, gt)], avail, True, {"foo": ["someId1", "someId2"]})
        hx("successful complex request with affinity", [er("foo", R1, 2, gt), er("bar", R2, 1, gt)], avail, True,
           {"foo": ["someId1", "someId2"], "bar": ["someId1"]})
        # derived from the same code path (core/extended_resources.go:83-111): entries drain one pool in order
        src = "core/extended_resources.go:83-150 (derived)"
        hx("two entries drain the same pool", [er("a", R1, 1, gt), er("b", R1, 1, gt)], {R1: devs}, True,
           {"a": ["someId1"], "b": ["someId2"]})
        hx("second entry finds the pool drained", [er("a", R1, 2, gt), er("b", R1, 1, gt)], {R1: devs}, False, None,
           ["Insufficient " + R1])

    # ----------------------------------------------------------------- TestZeroRequest (whole PrioritizeNodes)
    # core/generic_scheduler_test.go:433-561: LeastRequested + Balanced + SelectorSpread, all weight 1, total 25.
    # Run through the whole Schedule path: the other four priorities are weighted 0 and nodes get a pod
    # capacity (the Go test calls PrioritizeNodes directly, so its nodes need none).
    def build_zero_request():
        src = "core/generic_scheduler_test.go:433-561"
        DC, DM = 100, 200 * 1024 * 1024
        no_res = [{}]
        small = [ctr(DC, DM)]
        large = [ctr(DC * 3, DM * 3)]
        nodes = lambda: [node("machine1", 1000, DM * 10, pods=110), node("machine2", 1000, DM * 10, pods=110)]
        existing = lambda: [pod("machine1", containers=large), pod("machine1", containers=no_res),
                            pod("machine2", containers=large), pod("machine2", containers=small)]
        w = {"NodeAffinityPriority": 0, "TaintTolerationPriority": 0, "InterPodAffinityPriority": 0,
             "NodePreferAvoidPodsPriority": 0}
        for test, c, score in [("zero-request pod", no_res, 25), ("nonzero-request pod", small, 25)]:
            CASES.append({"op": "stream", "source": src, "test": "test priority of " + test, "weights": w,
                          "cluster": {"nodes": nodes(), "existingPods": existing(), "pods": [pod(containers=c)]},
                          "expect": {"scores": [score], "host_in": [["machine1", "machine2"]]}})
        # "we're not just getting the same score no matter what we schedule" (:505-514, :548-551)
        CASES.append({"op": "stream", "source": src, "test": "test priority of larger pod", "weights": w,
                      "cluster": {"nodes": nodes(), "existingPods": existing(), "pods": [pod(containers=large)]},
                      "expect": {"scores_not": [25], "host_in": [["machine1", "machine2"]]}})

    return [build_zero_request, build_interpod_priority, build_fits_resources, build_host_ports, build_fits_selector, build_tolerates_taints,
            build_interpod_predicate, build_devices] + g.get("_EXTRA_BUILDERS", [])
