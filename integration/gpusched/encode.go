// Package gpusched binds the B200 scheduling engine (libksched_b200.so, include/ksched_b200.h) into
// rotorliu/kubernetes' kube-scheduler behind algorithm.ScheduleAlgorithm. This image has no Go toolchain, so
// these files are shipped as source and are NOT built or tested here; tests/ drives the same C-ABI through
// ctypes with kubernetes_b200/wire.py, which this file mirrors field for field.
//
// encode.go: v1.Node / v1.Pod / selector owners -> the flat little-endian records of include/ksb_wire.hpp.
// A record holds no Go pointers, so it may be handed to C (cgo pointer-passing rules).
package gpusched

import (
	"encoding/json"
	"sort"

	"k8s.io/api/core/v1"
	metav1 "k8s.io/apimachinery/pkg/apis/meta/v1"
)

type writer struct{ b []byte }

func (w *writer) u8(v bool) {
	if v {
		w.b = append(w.b, 1)
	} else {
		w.b = append(w.b, 0)
	}
}
func (w *writer) u32(v uint32) { w.b = append(w.b, byte(v), byte(v>>8), byte(v>>16), byte(v>>24)) } // little-endian
func (w *writer) i32(v int32)  { w.u32(uint32(v)) }
func (w *writer) i64(v int64)  { w.u32(uint32(uint64(v))); w.u32(uint32(uint64(v) >> 32)) }
func (w *writer) str(s string) { w.u32(uint32(len(s))); w.b = append(w.b, s...) }

// maps are written with sorted keys, as wire.py does (the engine does not depend on the order)
func (w *writer) strMap(m map[string]string) {
	keys := make([]string, 0, len(m))
	for k := range m {
		keys = append(keys, k)
	}
	sort.Strings(keys)
	w.u32(uint32(len(keys)))
	for _, k := range keys {
		w.str(k)
		w.str(m[k])
	}
}

// resList: cpu in millicores, everything else as a plain integer (schedulercache.Resource, node_info.go:68-78,
// reads MilliValue() for cpu and Value() for the rest: node_info.go:81-111).
func (w *writer) resList(l v1.ResourceList) {
	keys := make([]string, 0, len(l))
	for k := range l {
		keys = append(keys, string(k))
	}
	sort.Strings(keys)
	w.u32(uint32(len(keys)))
	for _, k := range keys {
		q := l[v1.ResourceName(k)]
		w.str(k)
		if k == string(v1.ResourceCPU) {
			w.i64(q.MilliValue())
		} else {
			w.i64(q.Value())
		}
	}
}

func (w *writer) strs(v []string) {
	w.u32(uint32(len(v)))
	for _, s := range v {
		w.str(s)
	}
}

func (w *writer) nodeReqs(reqs []v1.NodeSelectorRequirement) {
	w.u32(uint32(len(reqs)))
	for _, r := range reqs {
		w.str(r.Key)
		w.str(string(r.Operator))
		w.strs(r.Values)
	}
}

func (w *writer) labelSelector(sel *metav1.LabelSelector) {
	w.u8(sel != nil)
	if sel == nil {
		w.strMap(nil)
		w.u32(0)
		return
	}
	w.strMap(sel.MatchLabels)
	w.u32(uint32(len(sel.MatchExpressions)))
	for _, r := range sel.MatchExpressions {
		w.str(r.Key)
		w.str(string(r.Operator))
		w.strs(r.Values)
	}
}

// encodeNode: what NodeInfo.SetNode reads (schedulercache/node_info.go:469-491) plus the fork's
// Status.ExtendedResources (schedulercache/extended_resources.go:154-207).
func encodeNode(n *v1.Node) []byte {
	w := &writer{}
	w.str(n.Name)
	w.u8(n.Labels != nil)
	w.strMap(n.Labels)
	w.resList(n.Status.Allocatable)
	w.u32(uint32(len(n.Spec.Taints)))
	for _, t := range n.Spec.Taints {
		w.str(t.Key)
		w.str(t.Value)
		w.str(string(t.Effect))
	}
	w.u32(uint32(len(n.Status.Conditions)))
	for _, c := range n.Status.Conditions {
		w.str(string(c.Type))
		w.str(string(c.Status))
	}
	w.u8(n.Spec.Unschedulable)
	// preferAvoidPods annotation (node_prefer_avoid_pods.go:47-57 via helpers.go:335-346): parsed once per
	// node here instead of once per (pod, node) in the reference
	var avoid v1.AvoidPods
	invalid := false
	if s := n.Annotations[v1.PreferAvoidPodsAnnotationKey]; s != "" {
		if err := json.Unmarshal([]byte(s), &avoid); err != nil {
			invalid = true
		}
	}
	w.u8(invalid)
	type ref struct{ kind, uid string }
	var refs []ref
	for _, e := range avoid.PreferAvoidPods {
		if c := e.PodSignature.PodController; c != nil {
			refs = append(refs, ref{c.Kind, string(c.UID)})
		}
	}
	w.u32(uint32(len(refs)))
	for _, r := range refs {
		w.str(r.kind)
		w.str(r.uid)
	}
	names := make([]string, 0, len(n.Status.ExtendedResources))
	for k := range n.Status.ExtendedResources {
		names = append(names, string(k))
	}
	sort.Strings(names)
	w.u32(uint32(len(names)))
	for _, name := range names {
		dom := n.Status.ExtendedResources[v1.ResourceName(name)]
		ids := make([]string, 0, len(dom.Resources))
		for id := range dom.Resources {
			ids = append(ids, id)
		}
		sort.Strings(ids) // ascending device id = the engine's bit order
		w.str(name)
		w.u32(uint32(len(ids)))
		for _, id := range ids {
			d := dom.Resources[id]
			w.str(d.ID)
			w.str(d.Health)
			w.strMap(d.Attributes)
		}
	}
	w.u32(uint32(len(n.Status.Images))) // priorities/image_locality.go:66-73
	for _, im := range n.Status.Images {
		w.strs(im.Names)
		w.i64(im.SizeBytes)
	}
	return w.b
}

func (w *writer) container(c *v1.Container) {
	w.resList(c.Resources.Requests)
	w.resList(c.Resources.Limits)
	w.u32(uint32(len(c.Ports)))
	for _, p := range c.Ports {
		w.i32(p.HostPort)
		w.str(string(p.Protocol))
		w.str(p.HostIP)
	}
}

func (w *writer) podTerm(t *v1.PodAffinityTerm) {
	w.labelSelector(t.LabelSelector)
	w.strs(t.Namespaces)
	w.str(t.TopologyKey)
}

func (w *writer) podAffinity(present bool, req []v1.PodAffinityTerm, pref []v1.WeightedPodAffinityTerm) {
	w.u8(present)
	w.u32(uint32(len(req)))
	for i := range req {
		w.podTerm(&req[i])
	}
	w.u32(uint32(len(pref)))
	for i := range pref {
		w.i32(pref[i].Weight)
		w.podTerm(&pref[i].PodAffinityTerm)
	}
}

// encodePod: what predicateMetadata / priorityMetadata (predicates/metadata.go:45-101, priorities/metadata.go:45-66)
// and NodeInfo.AddPod (node_info.go:329-356) read. Volumes are counted only: the engine refuses pods that have any.
func encodePod(p *v1.Pod) []byte {
	w := &writer{}
	w.str(p.Namespace)
	w.str(p.Name)
	w.strMap(p.Labels)
	w.u8(p.DeletionTimestamp != nil)
	w.str(p.Spec.NodeName)
	ctl := metav1.GetControllerOf(p)
	w.u8(ctl != nil)
	if ctl != nil {
		w.str(ctl.Kind)
		w.str(string(ctl.UID))
	} else {
		w.str("")
		w.str("")
	}
	w.u32(uint32(len(p.Spec.Containers)))
	for i := range p.Spec.Containers {
		w.container(&p.Spec.Containers[i])
	}
	w.u32(uint32(len(p.Spec.InitContainers)))
	for i := range p.Spec.InitContainers {
		w.container(&p.Spec.InitContainers[i])
	}
	w.strMap(p.Spec.NodeSelector)
	aff := p.Spec.Affinity
	w.u8(aff != nil)
	var na *v1.NodeAffinity
	if aff != nil {
		na = aff.NodeAffinity
	}
	w.u8(na != nil)
	var req *v1.NodeSelector
	if na != nil {
		req = na.RequiredDuringSchedulingIgnoredDuringExecution
	}
	w.u8(req != nil)
	if req != nil {
		w.u32(uint32(len(req.NodeSelectorTerms)))
		for _, t := range req.NodeSelectorTerms {
			w.nodeReqs(t.MatchExpressions)
		}
	} else {
		w.u32(0)
	}
	if na != nil {
		w.u32(uint32(len(na.PreferredDuringSchedulingIgnoredDuringExecution)))
		for _, t := range na.PreferredDuringSchedulingIgnoredDuringExecution {
			w.i32(t.Weight)
			w.nodeReqs(t.Preference.MatchExpressions)
		}
	} else {
		w.u32(0)
	}
	if aff != nil && aff.PodAffinity != nil {
		w.podAffinity(true, aff.PodAffinity.RequiredDuringSchedulingIgnoredDuringExecution,
			aff.PodAffinity.PreferredDuringSchedulingIgnoredDuringExecution)
	} else {
		w.podAffinity(false, nil, nil)
	}
	if aff != nil && aff.PodAntiAffinity != nil {
		w.podAffinity(true, aff.PodAntiAffinity.RequiredDuringSchedulingIgnoredDuringExecution,
			aff.PodAntiAffinity.PreferredDuringSchedulingIgnoredDuringExecution)
	} else {
		w.podAffinity(false, nil, nil)
	}
	w.u32(uint32(len(p.Spec.Tolerations)))
	for _, t := range p.Spec.Tolerations {
		w.str(t.Key)
		w.str(string(t.Operator))
		w.str(t.Value)
		w.str(string(t.Effect))
	}
	w.u32(uint32(len(p.Spec.ExtendedResources)))
	for _, e := range p.Spec.ExtendedResources {
		w.str(e.Name)
		w.resList(e.Resources.Limits)
		reqs := make([]v1.NodeSelectorRequirement, len(e.Affinity.Required))
		for i, r := range e.Affinity.Required {
			reqs[i] = v1.NodeSelectorRequirement(r) // ResourceSelector is a NodeSelectorRequirement (types.go)
		}
		w.nodeReqs(reqs)
		w.strs(e.Assigned)
	}
	w.i32(int32(len(p.Spec.Volumes)))
	prio := int32(0) // util.GetPodPriority
	if p.Spec.Priority != nil {
		prio = *p.Spec.Priority
	}
	w.i32(prio)
	w.u32(uint32(len(p.Spec.Containers))) // priorities/image_locality.go:75-80
	for i := range p.Spec.Containers {
		w.str(p.Spec.Containers[i].Image)
	}
	return w.b
}

// selector owners for SelectorSpread (priorities/metadata.go:69-96): services and RCs select by map,
// ReplicaSets and StatefulSets by LabelSelector.
func encodeMapOwner(ns, name string, selector map[string]string) []byte {
	w := &writer{}
	w.str(ns)
	w.str(name)
	w.strMap(selector)
	w.u8(false)
	w.strMap(nil)
	w.u32(0)
	return w.b
}

func encodeLabelSelectorOwner(ns, name string, sel *metav1.LabelSelector) []byte {
	w := &writer{}
	w.str(ns)
	w.str(name)
	w.strMap(nil)
	w.labelSelector(sel)
	return w.b
}
