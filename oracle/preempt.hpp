// ORACLE — TEST INFRASTRUCTURE ONLY. CPU restatement of the reference's preemption path. Nothing under
// kubernetes_b200/ may include or link this.
//
// Follows (paths relative to /root/reference/plugin/pkg/scheduler):
//   core/generic_scheduler.go:199-256 (Preempt), :654-758 (pickOneNodeForPreemption),
//     :760-793 (selectNodesForPreemption), :883-953 (selectVictimsOnNode), :957-991 (nodesWherePreemptionMightHelp)
//   util/utils.go:76-131 (GetPodPriority, SortableList, HigherPriorityPod)
// and, for the order of equal-priority victims, the Go standard library the reference is built with:
//   sort.Sort of go1.9 (src/sort/sort.go: insertionSort, siftDown, heapSort, medianOfThree, doPivot, quickSort,
//   maxDepth). The Go toolchain is absent here, so this restatement of a dependency OUTSIDE /root/reference is
//   "parity unpinned" beyond what the reference's own preemption tests exercise (at most five victims per node, i.e.
//   the insertion-sort path).
//
// Scope of this restatement (the engine's ksb_preempt has the same): podFitsOnNode is evaluated without the
// inter-pod-affinity metadata updates of meta.RemovePod / meta.AddPod, so a preemptor with pod (anti-)affinity terms
// (or host ports: the engine keeps port bitmaps per node, not per pod), or a cluster in which an existing pod carries
// required anti-affinity terms, is refused with an error. No
// PodDisruptionBudgets (numPDBViolations is 0), no nominated pods (the scheduling queue is nil), no extenders.
//
// Where the reference iterates a Go map (pickOneNodeForPreemption over nodesToVictims) its result is one of several
// equally good nodes at random; here, and in the engine, candidates are visited in node-list order and "the first"
// means the first in that order.
#pragma once
#include <climits>
#include <functional>
#include <map>
#include <string>
#include <vector>

#include "scheduler.hpp"

namespace oracle {

// ---- go1.9 sort.Sort over indices [0, n) with data.Less(i, j) / data.Swap(i, j)
struct GoSort {
  std::function<bool(int, int)> less;
  std::function<void(int, int)> swap;

  void insertionSort(int a, int b) {
    for (int i = a + 1; i < b; i++)
      for (int j = i; j > a && less(j, j - 1); j--) swap(j, j - 1);
  }
  void siftDown(int lo, int hi, int first) {
    int root = lo;
    for (;;) {
      int child = 2 * root + 1;
      if (child >= hi) return;
      if (child + 1 < hi && less(first + child, first + child + 1)) child++;
      if (!less(first + root, first + child)) return;
      swap(first + root, first + child);
      root = child;
    }
  }
  void heapSort(int a, int b) {
    int first = a, lo = 0, hi = b - a;
    for (int i = (hi - 1) / 2; i >= 0; i--) siftDown(i, hi, first);
    for (int i = hi - 1; i >= 0; i--) {
      swap(first, first + i);
      siftDown(lo, i, first);
    }
  }
  void medianOfThree(int m1, int m0, int m2) {
    if (less(m1, m0)) swap(m1, m0);
    if (less(m2, m1)) {
      swap(m2, m1);
      if (less(m1, m0)) swap(m1, m0);
    }
  }
  void doPivot(int lo, int hi, int* midlo, int* midhi) {
    int m = (int)((unsigned)(lo + hi) >> 1);
    if (hi - lo > 40) {
      int s = (hi - lo) / 8;
      medianOfThree(lo, lo + s, lo + 2 * s);
      medianOfThree(m, m - s, m + s);
      medianOfThree(hi - 1, hi - 1 - s, hi - 1 - 2 * s);
    }
    medianOfThree(lo, m, hi - 1);
    int pivot = lo;
    int a = lo + 1, c = hi - 1;
    for (; a < c && less(a, pivot); a++) {
    }
    int b = a;
    for (;;) {
      for (; b < c && !less(pivot, b); b++) {
      }
      for (; b < c && less(pivot, c - 1); c--) {
      }
      if (b >= c) break;
      swap(b, c - 1);
      b++;
      c--;
    }
    bool protect = hi - c < 5;
    if (!protect && hi - c < (hi - lo) / 4) {
      int dups = 0;
      if (!less(pivot, hi - 1)) {
        swap(c, hi - 1);
        c++;
        dups++;
      }
      if (!less(b - 1, pivot)) {
        b--;
        dups++;
      }
      if (!less(m, pivot)) {
        swap(m, b - 1);
        b--;
        dups++;
      }
      protect = dups > 1;
    }
    if (protect) {
      for (;;) {
        for (; a < b && !less(b - 1, pivot); b--) {
        }
        for (; a < b && less(a, pivot); a++) {
        }
        if (a >= b) break;
        swap(a, b - 1);
        a++;
        b--;
      }
    }
    swap(pivot, b - 1);
    *midlo = b - 1;
    *midhi = c;
  }
  void quickSort(int a, int b, int maxDepth) {
    while (b - a > 12) {
      if (maxDepth == 0) {
        heapSort(a, b);
        return;
      }
      maxDepth--;
      int mlo, mhi;
      doPivot(a, b, &mlo, &mhi);
      if (mlo - a < b - mhi) {
        quickSort(a, mlo, maxDepth);
        a = mhi;
      } else {
        quickSort(mhi, b, maxDepth);
        b = mlo;
      }
    }
    if (b - a > 1) {
      for (int i = a + 6; i < b; i++)
        if (less(i, i - 6)) swap(i, i - 6);
      insertionSort(a, b);
    }
  }
  void sort(int n) {
    int depth = 0;
    for (int i = n; i > 0; i >>= 1) depth++;
    quickSort(0, n, depth * 2);
  }
};

struct Victims {
  std::vector<PodPtr> pods;  // decreasing priority (the order they were reprieve-tested in)
  int numPDBViolations = 0;
};

struct PreemptResult {
  std::string error;                            // non-empty: Preempt returned an error (or the case is out of scope)
  std::vector<std::string> potential;           // nodesWherePreemptionMightHelp, node-list order
  std::vector<std::pair<std::string, Victims>> candidates;  // selectNodesForPreemption, node-list order
  std::string node;                             // picked node ("" = none)
  std::vector<std::string> victims;             // ns/name of the picked node's victims, decreasing priority
  bool clear_nomination = false;                // no potential node: the pod's own nomination is to be cleared
};

inline bool unresolvable_reason(const std::string& r) {  // generic_scheduler.go:969-981
  return r == "MatchNodeSelector" || r == "HostName" || r == "PodToleratesNodeTaints" || r == "CheckNodeLabelPresence" ||
         r == "NodeNotReady" || r == "NodeNetworkUnavailable" || r == "NodeUnschedulable" || r == "NodeUnknownCondition" ||
         r == "NoVolumeZoneConflict" || r == "NoVolumeNodeConflict" || r == "VolumeBindingNoMatch";
}

// generic_scheduler.go:957-991
inline std::vector<NodePtr> nodes_where_preemption_might_help(const std::vector<NodePtr>& nodes,
                                                              const std::map<std::string, std::vector<std::string>>& failed) {
  std::vector<NodePtr> out;
  for (auto& n : nodes) {
    bool unresolvable = false;
    auto it = failed.find(n->name);
    if (it != failed.end())
      for (auto& r : it->second)
        if (unresolvable_reason(r)) unresolvable = true;
    if (it == failed.end() || !unresolvable) out.push_back(n);
  }
  return out;
}

// fit check used inside selectVictimsOnNode: "resources" = PodFitsResources alone (what the reference's preemption
// tests plug in), otherwise the configured predicate set of podFitsOnNode
using FitFn = std::function<bool(const NodeInfo&)>;

// generic_scheduler.go:883-953
inline bool select_victims_on_node(const Pod& pod, const NodeInfo& nodeInfo, const FitFn& fits, Victims* out) {
  NodeInfo copy = nodeInfo;  // nodeInfo.Clone()
  const int32_t podPriority = pod.priority;
  std::vector<PodPtr> potential;
  // `for _, p := range nodeInfoCopy.Pods()` while RemovePod swaps the last element into the freed slot: the range
  // walks the ORIGINAL length over the same backing array, so every pod is visited once, in its original position
  // (a moved pod is met again at its old slot, which RemovePod never overwrites), and the victims keep NodeInfo order.
  const std::vector<PodPtr> original = copy.pods;
  for (auto& p : original)
    if (p->priority < podPriority) {
      potential.push_back(p);
      copy.RemovePod(*p);
    }
  {  // potentialVictims.Sort(): sort.Sort with HigherPriorityPod as Less
    GoSort gs;
    gs.less = [&](int i, int j) { return potential[i]->priority > potential[j]->priority; };
    gs.swap = [&](int i, int j) { std::swap(potential[i], potential[j]); };
    gs.sort((int)potential.size());
  }
  if (!fits(copy)) return false;
  out->pods.clear();
  out->numPDBViolations = 0;
  // no PDBs: every potential victim is "non-violating"; reprieve from the highest priority down
  for (auto& p : potential) {
    copy.AddPod(p);
    if (!fits(copy)) {
      copy.RemovePod(*p);
      out->pods.push_back(p);
    }
  }
  return true;
}

// generic_scheduler.go:654-758, candidates visited in node-list order
inline int pick_one_node_for_preemption(const std::vector<std::pair<std::string, Victims>>& cands) {
  if (cands.empty()) return -1;
  for (size_t i = 0; i < cands.size(); ++i)
    if (cands[i].second.pods.empty()) return (int)i;  // a node that needs no preemption
  std::vector<int> cur;
  int minViol = INT_MAX;
  for (size_t i = 0; i < cands.size(); ++i) {
    int v = cands[i].second.numPDBViolations;
    if (v < minViol) {
      minViol = v;
      cur.clear();
    }
    if (v == minViol) cur.push_back((int)i);
  }
  if (cur.size() == 1) return cur[0];
  std::vector<int> next;
  int32_t minHighest = INT32_MAX;
  for (int i : cur) {
    int32_t h = cands[i].second.pods[0]->priority;
    if (h < minHighest) {
      minHighest = h;
      next.clear();
    }
    if (h == minHighest) next.push_back(i);
  }
  if (next.size() == 1) return next[0];
  cur.swap(next);
  next.clear();
  int64_t minSum = INT64_MAX;
  for (int i : cur) {
    int64_t sum = 0;
    for (auto& p : cands[i].second.pods) sum += (int64_t)p->priority + ((int64_t)INT32_MAX + 1);
    if (sum < minSum) {
      minSum = sum;
      next.clear();
    }
    if (sum == minSum) next.push_back(i);
  }
  if (next.size() == 1) return next[0];
  cur.swap(next);
  next.clear();
  int minPods = INT_MAX;
  for (int i : cur) {
    int np = (int)cands[i].second.pods.size();
    if (np < minPods) {
      minPods = np;
      next.clear();
    }
    if (np == minPods) next.push_back(i);
  }
  return next.empty() ? -1 : next[0];
}

// genericScheduler.Preempt, generic_scheduler.go:199-256, after a Schedule of `pod` that ended in a FitError with
// `failed` as its FailedPredicateMap. `resources_only`: plug in PodFitsResources alone, as the reference's tests do.
inline PreemptResult preempt(Oracle& o, const Pod& pod, const std::vector<NodePtr>& nodes,
                             const std::map<std::string, std::vector<std::string>>& failed, bool resources_only) {
  PreemptResult res;
  if (nodes.empty()) {
    res.error = "no nodes available to schedule pods";
    return res;
  }
  if (!resources_only) {
    if (pod.affinity.present && (pod.affinity.pod.present || pod.affinity.anti.present)) {
      res.error = "preemption for a pod with inter-pod (anti-)affinity terms is not implemented";
      return res;
    }
    for (auto& c : pod.containers)
      for (auto& port : c.ports)
        if (port.host_port != 0) {
          res.error = "preemption for a pod with host ports is not implemented";
          return res;
        }
    for (auto& kv : o.cachedNodeInfoMap)
      for (auto& ep : kv.second.podsWithAffinity)
        if (ep->affinity.present && !required_terms(ep->affinity.anti).empty()) {
          res.error = "preemption next to pods with required anti-affinity terms is not implemented";
          return res;
        }
  }
  std::vector<NodePtr> potential = nodes_where_preemption_might_help(nodes, failed);
  for (auto& n : potential) res.potential.push_back(n->name);
  if (potential.empty()) {
    res.clear_nomination = true;
    return res;
  }
  PredicateMeta meta = o.make_predicate_meta(pod);
  std::vector<PodPtr> allPods;  // only read by the inter-pod-affinity predicate, which is out of scope here
  std::vector<Oracle::TermProbe> none;
  FitFn fits = [&](const NodeInfo& ni) {
    std::vector<std::string> fails;
    std::string err;
    if (resources_only) Oracle::pod_fits_resources(meta, ni, &fails);
    else if (!o.pod_fits_on_node(pod, meta, ni, allPods, none, none, &fails, &err)) return false;
    return fails.empty();
  };
  for (auto& n : potential) {
    Victims v;
    if (select_victims_on_node(pod, o.cachedNodeInfoMap.at(n->name), fits, &v)) res.candidates.push_back({n->name, v});
  }
  int pick = pick_one_node_for_preemption(res.candidates);
  if (pick < 0) return res;
  res.node = res.candidates[pick].first;
  for (auto& p : res.candidates[pick].second.pods) res.victims.push_back(p->key());
  return res;
}

}  // namespace oracle
