"""Fifth batch of transcriptions of the reference's own tests (see build_fixtures.py): preemption
(core/generic_scheduler_test.go). Only the cases that plug in the real PodFitsResources predicate are transcribed; the
ones built on the test's stub predicates (falsePredicate, truePredicate, matchesPredicate), the anti-affinity case and
TestPreempt's two cases with fake extenders have no counterpart here."""

CPU, MEM = 100, 200 * 1024 * 1024  # priorityutil.DefaultMilliCpuRequest, DefaultMemoryRequest (non_zero.go:29-34)
NEG, LOW, MID, HIGH, VERY_HIGH = -100, 0, 100, 1000, 10000


def builders(g):
    node, pod, ctr, CASES = g["node"], g["pod"], g["ctr"], g["CASES"]
    small, medium, large, very_large = ([ctr(CPU * k, MEM * k)] for k in (1, 2, 3, 5))

    def mk_nodes(names):  # makeNode(n, DefaultMilliCpuRequest*5, DefaultMemoryRequest*5), 100 pods
        return [node(n, CPU * 5, MEM * 5, pods=100) for n in names]

    def case(test, src, nodes, p, pods, expect, **kw):
        CASES.append(dict({"op": "preempt", "name": "Preempt", "source": src, "test": test, "resources_only": True,
                           "cluster": {"nodes": nodes, "existingPods": pods}, "pod": p, "expect": expect}, **kw))

    def P(name, containers, prio, on=None):
        return pod(on, name=name, containers=containers, priority=prio)

    # ----------------------------------------------------------------- TestSelectNodesForPreemption
    # core/generic_scheduler_test.go:666-802 (cases with algorithmpredicates.PodFitsResources)
    def build_select_nodes():
        src = "core/generic_scheduler_test.go:706-747"
        n2 = lambda: mk_nodes(["machine1", "machine2"])
        case("a pod that fits on both machines when lower priority pods are preempted", src, n2(), P("machine1", large, HIGH),
             [P("a", large, MID, "machine1"), P("b", large, MID, "machine2")],
             {"candidates": {"machine1": ["a"], "machine2": ["b"]}})
        case("a pod that would fit on the machines, but other pods running are higher priority", src, n2(),
             P("machine1", large, LOW), [P("a", large, MID, "machine1"), P("b", large, MID, "machine2")], {"candidates": {}})
        case("medium priority pod is preempted, but lower priority one stays as it is small", src, n2(), P("machine1", large, HIGH),
             [P("a", small, LOW, "machine1"), P("b", large, MID, "machine1"), P("c", large, MID, "machine2")],
             {"candidates": {"machine1": ["b"], "machine2": ["c"]}})
        case("mixed priority pods are preempted", src, n2(), P("machine1", large, HIGH),
             [P("a", small, MID, "machine1"), P("b", small, LOW, "machine1"), P("c", medium, MID, "machine1"),
              P("d", small, HIGH, "machine1"), P("e", large, HIGH, "machine2")],
             {"candidates": {"machine1": ["b", "c"]}})

    # ----------------------------------------------------------------- TestPickOneNodeForPreemption
    # core/generic_scheduler_test.go:804-963: `expected` lists the acceptable nodes
    def build_pick_one():
        src = "core/generic_scheduler_test.go:813-940"
        n = lambda k: mk_nodes(["machine%d" % i for i in range(1, k + 1)])
        case("No node needs preemption", src, n(1), P("machine1", large, HIGH), [P("m1.1", small, MID, "machine1")],
             {"node_in": ["machine1"]})
        case("a pod that fits on both machines when lower priority pods are preempted", src, n(2), P("machine1", large, HIGH),
             [P("m1.1", large, MID, "machine1"), P("m2.1", large, MID, "machine2")], {"node_in": ["machine1", "machine2"]})
        case("a pod that fits on a machine with no preemption", src, n(3), P("machine1", large, HIGH),
             [P("m1.1", large, MID, "machine1"), P("m2.1", large, MID, "machine2")], {"node_in": ["machine3"]})
        case("machine with min highest priority pod is picked", src, n(3), P("machine1", very_large, HIGH),
             [P("m1.1", medium, MID, "machine1"), P("m1.2", large, MID, "machine1"), P("m2.1", medium, MID, "machine2"),
              P("m2.2", medium, LOW, "machine2"), P("m3.1", medium, LOW, "machine3"), P("m3.2", medium, LOW, "machine3")],
             {"node_in": ["machine3"]})
        case("when highest priorities are the same, minimum sum of priorities is picked", src, n(3), P("machine1", very_large, HIGH),
             [P("m1.1", medium, MID, "machine1"), P("m1.2", large, MID, "machine1"), P("m2.1", large, MID, "machine2"),
              P("m2.2", medium, LOW, "machine2"), P("m3.1", medium, MID, "machine3"), P("m3.2", medium, MID, "machine3")],
             {"node_in": ["machine2"]})
        case("when highest priority and sum are the same, minimum number of pods is picked", src, n(3),
             P("machine1", very_large, HIGH),
             [P("m1.1", small, MID, "machine1"), P("m1.2", small, NEG, "machine1"), P("m1.3", small, MID, "machine1"),
              P("m1.4", small, NEG, "machine1"), P("m2.1", large, MID, "machine2"), P("m2.2", medium, NEG, "machine2"),
              P("m3.1", medium, MID, "machine3"), P("m3.2", small, NEG, "machine3"), P("m3.3", small, LOW, "machine3")],
             {"node_in": ["machine2"]})
        case("sum of adjusted priorities is considered", src, n(3), P("machine1", very_large, HIGH),
             [P("m1.1", small, MID, "machine1"), P("m1.2", small, NEG, "machine1"), P("m1.3", small, NEG, "machine1"),
              P("m2.1", large, MID, "machine2"), P("m2.2", medium, NEG, "machine2"), P("m3.1", medium, MID, "machine3"),
              P("m3.2", small, NEG, "machine3"), P("m3.3", small, LOW, "machine3")],
             {"node_in": ["machine2"]})
        case("non-overlapping lowest high priority, sum priorities, and number of pods", src, n(4), P("pod1", very_large, VERY_HIGH),
             [P("m1.1", small, MID, "machine1"), P("m1.2", small, LOW, "machine1"), P("m1.3", small, LOW, "machine1"),
              P("m2.1", large, HIGH, "machine2"), P("m3.1", medium, MID, "machine3"), P("m3.2", small, LOW, "machine3"),
              P("m3.3", small, LOW, "machine3"), P("m3.4", medium, LOW, "machine3"), P("m4.1", medium, MID, "machine4"),
              P("m4.2", small, MID, "machine4"), P("m4.3", small, MID, "machine4"), P("m4.4", small, NEG, "machine4")],
             {"node_in": ["machine1"]})

    # ----------------------------------------------------------------- TestNodesWherePreemptionMightHelp
    # core/generic_scheduler_test.go:965-1081 (reasons by the texts this repository uses for them)
    def build_might_help():
        src = "core/generic_scheduler_test.go:977-1066"
        n4 = lambda: mk_nodes(["machine1", "machine2", "machine3", "machine4"])
        case("No node should be attempted", src, n4(), P("pod1", [], 0), [],
             {"potential": []},
             failed={"machine1": ["MatchNodeSelector"], "machine2": ["HostName"], "machine3": ["PodToleratesNodeTaints"],
                     "machine4": ["CheckNodeLabelPresence"]})
        case("pod affinity should be tried", src, n4(), P("pod1", [], 0), [], {"potential": ["machine1", "machine4"]},
             failed={"machine1": ["MatchInterPodAffinity", "PodAffinityRulesNotMatch"], "machine2": ["HostName"],
                     "machine3": ["NodeUnschedulable"]})
        case("pod with both pod affinity and anti-affinity should be tried", src, n4(), P("pod1", [], 0), [],
             {"potential": ["machine1", "machine3", "machine4"]},
             failed={"machine1": ["MatchInterPodAffinity", "PodAffinityRulesNotMatch"], "machine2": ["HostName"]})
        case("Mix of failed predicates works fine", src, n4(), P("pod1", [], 0), [], {"potential": ["machine3", "machine4"]},
             failed={"machine1": ["MatchNodeSelector", "NodeOutOfDisk", "Insufficient memory"], "machine2": ["HostName", "NoDiskConflict"],
                     "machine3": ["Insufficient memory"], "machine4": []})

    # ----------------------------------------------------------------- TestPreempt
    # core/generic_scheduler_test.go:1083-1217, the two cases without extenders; `failed` is the test's own failedPredMap
    # (:1084-1088: machine2 carries a reason preemption cannot resolve)
    def build_preempt():
        src = "core/generic_scheduler_test.go:1102-1133"
        n3 = lambda: mk_nodes(["machine1", "machine2", "machine3"])
        failed = {"machine1": ["Insufficient memory"], "machine2": ["NoDiskConflict"], "machine3": ["Insufficient memory"]}
        case("basic preemption logic", src, n3(), P("pod1", very_large, HIGH),
             [P("m1.1", small, LOW, "machine1"), P("m1.2", small, LOW, "machine1"), P("m2.1", large, HIGH, "machine2"),
              P("m3.1", medium, MID, "machine3")],
             {"node_in": ["machine1"], "victims": ["m1.1", "m1.2"]}, failed=failed)
        case("One node doesn't need any preemption", src, n3(), P("pod1", very_large, HIGH),
             [P("m1.1", small, LOW, "machine1"), P("m1.2", small, LOW, "machine1"), P("m2.1", large, HIGH, "machine2")],
             {"node_in": ["machine3"], "victims": []}, failed=failed)

    return [build_select_nodes, build_pick_one, build_might_help, build_preempt]
