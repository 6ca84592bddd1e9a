package gpusched

// scheduler.go: algorithm.ScheduleAlgorithm (plugin/pkg/scheduler/algorithm/scheduler_interface.go:48-61) over the
// C-ABI. Constructed at plugin/pkg/scheduler/factory/factory.go:936 in place of core.NewGenericScheduler; the
// generic scheduler is still built and kept for Preempt / Predicates / Prioritizers, never for Schedule.
// Source only: see the package comment in encode.go.

/*
#cgo LDFLAGS: -lksched_b200
#include <stdlib.h>
#include "ksched_b200.h"
*/
import "C"

import (
	"fmt"
	"sync"
	"unsafe"

	"k8s.io/api/core/v1"
	"k8s.io/kubernetes/plugin/pkg/scheduler/algorithm"
	"k8s.io/kubernetes/plugin/pkg/scheduler/algorithm/predicates"
	"k8s.io/kubernetes/plugin/pkg/scheduler/core"
	"k8s.io/kubernetes/plugin/pkg/scheduler/schedulercache"
)

// the predicate / priority keys the engine implements (algorithmprovider/defaults/defaults.go:119-184, 217-257)
var supportedPredicates = map[string]bool{
	"GeneralPredicates": true, "PodToleratesNodeTaints": true, "CheckNodeCondition": true,
	"CheckNodeMemoryPressure": true, "CheckNodeDiskPressure": true, "MatchInterPodAffinity": true,
	// volume predicates are registered by the DefaultProvider and are trivially true for pods without volumes; the
	// engine refuses pods that have volumes (no CPU fallback), so accepting the keys is safe
	"CheckNodeLabelPresence": true, // with its labels passed through AddLabelPresence
	"NoVolumeZoneConflict": true, "MaxEBSVolumeCount": true, "MaxGCEPDVolumeCount": true,
	"MaxAzureDiskVolumeCount": true, "NoDiskConflict": true, "CheckVolumeBinding": true,
}

var weightField = map[string]func(*C.ksb_config, C.int32_t){
	"LeastRequestedPriority":      func(c *C.ksb_config, w C.int32_t) { c.w_least_requested = w },
	"BalancedResourceAllocation":  func(c *C.ksb_config, w C.int32_t) { c.w_balanced = w },
	"SelectorSpreadPriority":      func(c *C.ksb_config, w C.int32_t) { c.w_selector_spread = w },
	"NodeAffinityPriority":        func(c *C.ksb_config, w C.int32_t) { c.w_node_affinity = w },
	"TaintTolerationPriority":     func(c *C.ksb_config, w C.int32_t) { c.w_taint_toleration = w },
	"InterPodAffinityPriority":    func(c *C.ksb_config, w C.int32_t) { c.w_inter_pod_affinity = w },
	"NodePreferAvoidPodsPriority": func(c *C.ksb_config, w C.int32_t) { c.w_prefer_avoid = w },
}

// priorities a Policy can enable by name that take no arguments (ksb_policy_set)
var policyWeightKind = map[string]C.int32_t{
	"MostRequestedPriority":    C.KSB_POLICY_MOST_REQUESTED,
	"ResourceLimitsPriority":   C.KSB_POLICY_RESOURCE_LIMITS,
	"ImageLocalityPriority":    C.KSB_POLICY_IMAGE_LOCALITY,
	"ServiceSpreadingPriority": C.KSB_POLICY_SERVICE_SPREADING,
	"EqualPriority":            C.KSB_POLICY_EQUAL,
}

// GPUScheduler implements algorithm.ScheduleAlgorithm.
type GPUScheduler struct {
	mu      sync.Mutex // Schedule is called from one goroutine (scheduler.go:179); the lock documents it
	e       *C.ksb_engine
	generic algorithm.ScheduleAlgorithm
}

// New refuses any predicate or priority key outside the DefaultProvider set: there is no CPU fallback for Schedule.
func New(predicateKeys []string, priorityWeights map[string]int, hardPodAffinityWeight int32,
	generic algorithm.ScheduleAlgorithm) (*GPUScheduler, error) {
	for _, k := range predicateKeys {
		if !supportedPredicates[k] {
			return nil, fmt.Errorf("gpusched: predicate %q is not implemented by the GPU engine", k)
		}
	}
	var cfg C.ksb_config
	C.ksb_config_default(&cfg)
	for name, set := range weightField {
		set(&cfg, 0) // a priority that is not configured does not count
		_ = name
	}
	policyWeights := map[string]int{}
	for name, w := range priorityWeights {
		if _, ok := policyWeightKind[name]; ok {
			policyWeights[name] = w
			continue
		}
		set, ok := weightField[name]
		if !ok {
			return nil, fmt.Errorf("gpusched: priority %q is not implemented by the GPU engine", name)
		}
		set(&cfg, C.int32_t(w))
	}
	if len(priorityWeights) == 0 {
		// PrioritizeNodes without any priority config (and without extenders) applies EqualPriorityMap: every node
		// scores 1 (core/generic_scheduler.go:517-529)
		policyWeights["EqualPriority"] = 1
	}
	cfg.hard_pod_affinity_weight = C.int32_t(hardPodAffinityWeight)
	var errbuf [512]C.char
	e := C.ksb_create(&cfg, &errbuf[0], 512)
	if e == nil {
		return nil, fmt.Errorf("ksb_create: %s", C.GoString(&errbuf[0]))
	}
	for name, w := range policyWeights {
		if C.ksb_policy_set(e, policyWeightKind[name], nil, 0, 0, C.int32_t(w)) != 0 {
			return nil, fmt.Errorf("ksb_policy_set: %s", C.GoString(C.ksb_last_error(e)))
		}
	}
	return &GPUScheduler{e: e, generic: generic}, nil
}

// AddNodeLabelPriority / AddLabelPresence: the Policy plugins that carry arguments (api/types.go:64-125:
// LabelPreference, LabelsPresence). Called by the factory for each such entry of the Policy, before the first Schedule.
func (g *GPUScheduler) AddNodeLabelPriority(label string, presence bool, weight int) error {
	p := C.int32_t(0)
	if presence {
		p = 1
	}
	lp, ll := strArg(label)
	if C.ksb_policy_set(g.e, C.KSB_POLICY_NODE_LABEL_PRIORITY, lp, ll, p, C.int32_t(weight)) != 0 {
		return g.lastError()
	}
	return nil
}

func (g *GPUScheduler) AddLabelPresence(labels []string, presence bool) error {
	p := C.int32_t(0)
	if presence {
		p = 1
	}
	for _, l := range labels {
		lp, ll := strArg(l)
		if C.ksb_policy_set(g.e, C.KSB_POLICY_LABEL_PRESENCE, lp, ll, p, 0) != 0 {
			return g.lastError()
		}
	}
	return nil
}

func (g *GPUScheduler) lastError() error { return fmt.Errorf("%s", C.GoString(C.ksb_last_error(g.e))) }

// strArg: a string's bytes for the duration of one C call (the engine copies what it keeps). A copy, because the
// tree this is written for builds with go1.9, which has no unsafe.StringData; never nil, so that the C side may
// build a std::string from it even when the length is 0.
var emptyArg [1]byte

func strArg(s string) (*C.char, C.size_t) {
	if len(s) == 0 {
		return (*C.char)(unsafe.Pointer(&emptyArg[0])), 0
	}
	b := []byte(s)
	return (*C.char)(unsafe.Pointer(&b[0])), C.size_t(len(b))
}

func bytesPtr(b []byte) (*C.uint8_t, C.size_t) {
	if len(b) == 0 {
		return nil, 0
	}
	return (*C.uint8_t)(unsafe.Pointer(&b[0])), C.size_t(len(b))
}

// Schedule: genericScheduler.Schedule (core/generic_scheduler.go:109-161) on the device.
func (g *GPUScheduler) Schedule(pod *v1.Pod, nodeLister algorithm.NodeLister) (string, v1.ExtendedResourceBinding, error) {
	g.mu.Lock()
	defer g.mu.Unlock()
	buf := encodePod(pod)
	p, n := bytesPtr(buf)
	var r C.ksb_result
	if C.ksb_schedule(g.e, p, n, &r) != 0 {
		return "", nil, g.lastError()
	}
	switch r.status {
	case C.KSB_NO_NODES:
		return "", nil, core.ErrNoNodesAvailable
	case C.KSB_FIT_ERROR:
		return "", nil, g.fitError(pod, int(r.num_nodes))
	case C.KSB_SCHEDULE_ERROR:
		return "", nil, g.lastError()
	}
	var name *C.char
	var nameLen C.size_t
	if C.ksb_node_name(g.e, r.node_index, &name, &nameLen) != 0 {
		return "", nil, fmt.Errorf("gpusched: node row %d has no name", int(r.node_index))
	}
	host := C.GoStringN(name, C.int(nameLen))
	return host, g.binding(pod, &r), nil
}

// ScheduleWithExtenders: Schedule when algorithm.SchedulerExtender instances are configured
// (core/generic_scheduler.go:333-351, :608-632). The extenders are called from Go exactly as the generic scheduler
// calls them (HTTPExtender.Filter / Prioritize, core/extender.go; wire form api/types.go:127-202); the engine gets their
// verdict: `keep` = the node names their Filter calls left (nil: no Filter extender), `scores` = combinedScores, the sum
// over extenders of weight * score by node name (nil: no Prioritize extender). The caller obtains the list to show the
// extenders from a first Schedule: the nodes without a failed predicate (FitError.FailedPredicates otherwise).
func (g *GPUScheduler) ScheduleWithExtenders(pod *v1.Pod, keep map[string]bool, scores map[string]int) (string, v1.ExtendedResourceBinding, error) {
	g.mu.Lock()
	defer g.mu.Unlock()
	p, n := bytesPtr(encodePod(pod))
	numNodes := 0
	for ; ; numNodes++ { // rows are the nodes in name order; ksb_node_name fails past the last one
		var name *C.char
		var nameLen C.size_t
		if C.ksb_node_name(g.e, C.int32_t(numNodes), &name, &nameLen) != 0 {
			break
		}
	}
	var keepRows []C.uint32_t
	var rowScores []C.int32_t
	if keep != nil {
		keepRows = make([]C.uint32_t, (numNodes+31)/32)
	}
	if scores != nil {
		rowScores = make([]C.int32_t, numNodes)
	}
	for row := 0; row < numNodes; row++ {
		var name *C.char
		var nameLen C.size_t
		C.ksb_node_name(g.e, C.int32_t(row), &name, &nameLen)
		host := C.GoStringN(name, C.int(nameLen))
		if keep != nil && keep[host] {
			keepRows[row/32] |= 1 << uint(row%32)
		}
		if scores != nil {
			rowScores[row] = C.int32_t(scores[host])
		}
	}
	var kp *C.uint32_t
	var sp *C.int32_t
	if len(keepRows) > 0 {
		kp = &keepRows[0]
	}
	if len(rowScores) > 0 {
		sp = &rowScores[0]
	}
	var r C.ksb_result
	if C.ksb_schedule_extended(g.e, p, n, kp, sp, &r) != 0 {
		return "", nil, g.lastError()
	}
	switch r.status {
	case C.KSB_NO_NODES:
		return "", nil, core.ErrNoNodesAvailable
	case C.KSB_FIT_ERROR:
		return "", nil, g.fitError(pod, int(r.num_nodes)) // "extender" stands for the extenders' own messages
	case C.KSB_SCHEDULE_ERROR:
		return "", nil, g.lastError()
	}
	var name *C.char
	var nameLen C.size_t
	if C.ksb_node_name(g.e, r.node_index, &name, &nameLen) != 0 {
		return "", nil, fmt.Errorf("gpusched: node row %d has no name", int(r.node_index))
	}
	return C.GoStringN(name, C.int(nameLen)), g.binding(pod, &r), nil
}

// binding: device bit masks -> v1.ExtendedResourceBinding (resource entry name -> device ids), what
// core.GetExtendedResources returns (core/extended_resources.go:42-81).
func (g *GPUScheduler) binding(pod *v1.Pod, r *C.ksb_result) v1.ExtendedResourceBinding {
	if len(pod.Spec.ExtendedResources) == 0 {
		return nil
	}
	out := v1.ExtendedResourceBinding{}
	for i, e := range pod.Spec.ExtendedResources {
		if i >= C.KSB_MAX_EXT_ENTRIES || len(e.Resources.Limits) != 1 {
			continue
		}
		var rname string
		for k := range e.Resources.Limits {
			rname = string(k)
		}
		crname := C.CString(rname)
		var ids []string
		mask := uint32(r.device_mask[i])
		for bit := 0; bit < C.KSB_MAX_DEVICES; bit++ {
			if mask&(1<<uint(bit)) == 0 {
				continue
			}
			var id *C.char
			var idLen C.size_t
			if C.ksb_device_id(g.e, r.node_index, crname, C.size_t(len(rname)), C.int32_t(bit), &id, &idLen) == 0 {
				ids = append(ids, C.GoStringN(id, C.int(idLen)))
			}
		}
		C.free(unsafe.Pointer(crname))
		out[e.Name] = v1.ExtendedResourceList{Resources: ids}
	}
	return out
}

// fitError rebuilds *core.FitError (core/generic_scheduler.go:47-51) from the per-node reason masks: the text of
// Error() (:70-88) and nodesWherePreemptionMightHelp (:957-991) both read FailedPredicates.
func (g *GPUScheduler) fitError(pod *v1.Pod, numNodes int) error {
	masks := make([]C.uint32_t, numNodes)
	failed := core.FailedPredicateMap{}
	if numNodes > 0 && C.ksb_failed_predicates(g.e, &masks[0], C.int32_t(numNodes)) == 0 {
		for row := 0; row < numNodes; row++ {
			if masks[row] == 0 {
				continue
			}
			var name *C.char
			var nameLen C.size_t
			if C.ksb_node_name(g.e, C.int32_t(row), &name, &nameLen) != 0 {
				continue
			}
			var reasons []algorithm.PredicateFailureReason
			for bit := 0; bit < 32; bit++ {
				if uint32(masks[row])&(1<<uint(bit)) == 0 {
					continue
				}
				var text [160]C.char
				if C.ksb_reason_name(g.e, C.int32_t(bit), &text[0], 160) == 0 {
					reasons = append(reasons, predicates.NewFailureReason(C.GoString(&text[0])))
				}
			}
			failed[C.GoStringN(name, C.int(nameLen))] = reasons
		}
	}
	return &core.FitError{Pod: pod, NumAllNodes: numNodes, FailedPredicates: failed}
}

// PreemptOnGPU: genericScheduler.Preempt's decision (core/generic_scheduler.go:199-256) from the engine: the node to
// preempt on, the victims as namespace/name keys in decreasing priority, and whether the pod's own nomination is to be
// cleared. An error means the case is outside the engine's scope (inter-pod affinity in play, host ports on the
// preemptor, PDBs, nominated pods): Preempt below then asks the generic scheduler, as before.
func (g *GPUScheduler) PreemptOnGPU(pod *v1.Pod) (string, []string, bool, error) {
	g.mu.Lock()
	defer g.mu.Unlock()
	p, n := bytesPtr(encodePod(pod))
	var r C.ksb_preempt_result
	if C.ksb_preempt(g.e, p, n, &r) != 0 {
		return "", nil, false, g.lastError()
	}
	if r.status == 2 {
		return "", nil, true, nil
	}
	if r.status != 0 {
		return "", nil, false, nil
	}
	var name *C.char
	var nameLen C.size_t
	if C.ksb_node_name(g.e, r.node_index, &name, &nameLen) != 0 {
		return "", nil, false, fmt.Errorf("gpusched: node row %d has no name", int(r.node_index))
	}
	victims := make([]string, 0, int(r.num_victims))
	for i := 0; i < int(r.num_victims); i++ {
		var key *C.char
		var keyLen C.size_t
		if C.ksb_preempt_victim(g.e, C.int32_t(i), &key, &keyLen) == 0 {
			victims = append(victims, C.GoStringN(key, C.int(keyLen)))
		}
	}
	return C.GoStringN(name, C.int(nameLen)), victims, false, nil
}

// Preempt asks the generic scheduler (which resolves victims to *v1.Pod objects from its own cache, honours PDBs and
// nominated pods); a factory that wants the engine's decision calls PreemptOnGPU first and maps the keys through its
// pod lister. Predicates and Prioritizers stay with the generic scheduler as well.
func (g *GPUScheduler) Preempt(pod *v1.Pod, nodeLister algorithm.NodeLister, scheduleErr error) (*v1.Node, []*v1.Pod, []*v1.Pod, error) {
	return g.generic.Preempt(pod, nodeLister, scheduleErr)
}
func (g *GPUScheduler) Predicates() map[string]algorithm.FitPredicate { return g.generic.Predicates() }
func (g *GPUScheduler) Prioritizers() []algorithm.PriorityConfig     { return g.generic.Prioritizers() }

// Resync replays a schedulercache snapshot into the engine after a device error (ksb_resync_begin/end).
func (g *GPUScheduler) Resync(nodes []*v1.Node, infos map[string]*schedulercache.NodeInfo) error {
	g.mu.Lock()
	defer g.mu.Unlock()
	if C.ksb_resync_begin(g.e) != 0 {
		return g.lastError()
	}
	for _, n := range nodes {
		p, l := bytesPtr(encodeNode(n))
		if C.ksb_node_upsert(g.e, p, l) != 0 {
			return g.lastError()
		}
	}
	for _, info := range infos {
		for _, pod := range info.Pods() {
			p, l := bytesPtr(encodePod(pod))
			if C.ksb_pod_add(g.e, p, l) != 0 {
				return g.lastError()
			}
		}
	}
	if C.ksb_resync_end(g.e) != 0 {
		return g.lastError()
	}
	return nil
}

// Close releases the engine.
func (g *GPUScheduler) Close() { C.ksb_destroy(g.e) }
