package gpusched

// cache.go: a decorator of schedulercache.Cache (plugin/pkg/scheduler/schedulercache/interface.go:61-121). Every
// mutator goes to the real Go cache first (Preempt, the listers and skipPodUpdate, factory/factory.go:312, still need
// it) and, when that succeeded, to the engine. The pod state machine (assumed / added / expired, cache.go:118-312,
// :426-462) stays in the Go cache: the engine only needs the net effect, which is "this pod is on that node" or
// "it is not". Source only: see the package comment in encode.go.

/*
#include "ksched_b200.h"
*/
import "C"

import (
	"k8s.io/api/core/v1"
	"k8s.io/apimachinery/pkg/labels"
	"k8s.io/kubernetes/plugin/pkg/scheduler/schedulercache"
)

// Cache wraps the scheduler cache handed to factory.NewConfigFactory.
type Cache struct {
	schedulercache.Cache
	g *GPUScheduler
}

// NewCache decorates `inner`; pass the result wherever the factory uses its schedulerCache.
func NewCache(inner schedulercache.Cache, g *GPUScheduler) *Cache { return &Cache{Cache: inner, g: g} }

func (c *Cache) podAdd(pod *v1.Pod) error {
	p, l := bytesPtr(encodePod(pod))
	if C.ksb_pod_add(c.g.e, p, l) != 0 {
		return c.g.lastError()
	}
	return nil
}

func (c *Cache) podRemove(pod *v1.Pod) error {
	ns, nsLen := strArg(pod.Namespace)
	name, nameLen := strArg(pod.Name)
	if C.ksb_pod_remove(c.g.e, ns, nsLen, name, nameLen) != 0 {
		return c.g.lastError()
	}
	return nil
}

// AssumePod (cache.go:118-137): scheduler.go:478 calls it with NodeName set; the bind echo of device ids
// (pkg/registry/core/pod/storage/storage.go:180-187) is anticipated by copying the binding into Assigned first,
// which is what makes devices deplete before the informer echo arrives.
func (c *Cache) AssumePod(pod *v1.Pod) error {
	if err := c.Cache.AssumePod(pod); err != nil {
		return err
	}
	return c.podAdd(pod)
}

// ForgetPod (cache.go:170-197): a failed bind.
func (c *Cache) ForgetPod(pod *v1.Pod) error {
	if err := c.Cache.ForgetPod(pod); err != nil {
		return err
	}
	return c.podRemove(pod)
}

// AddPod (cache.go:223-260): the informer confirms an assumed pod (no change for the engine unless the node
// differs) or reports a pod scheduled by someone else.
func (c *Cache) AddPod(pod *v1.Pod) error {
	wasAssumed, _ := c.Cache.IsAssumedPod(pod)
	if err := c.Cache.AddPod(pod); err != nil {
		return err
	}
	if wasAssumed {
		_ = c.podRemove(pod) // re-add below with the informer's copy (node name and Assigned ids as stored)
	}
	return c.podAdd(pod)
}

// UpdatePod (cache.go:262-281) = remove the old copy, add the new one.
func (c *Cache) UpdatePod(oldPod, newPod *v1.Pod) error {
	if err := c.Cache.UpdatePod(oldPod, newPod); err != nil {
		return err
	}
	if err := c.podRemove(oldPod); err != nil {
		return err
	}
	return c.podAdd(newPod)
}

// RemovePod (cache.go:283-312).
func (c *Cache) RemovePod(pod *v1.Pod) error {
	if err := c.Cache.RemovePod(pod); err != nil {
		return err
	}
	return c.podRemove(pod)
}

func (c *Cache) nodeUpsert(node *v1.Node) error {
	p, l := bytesPtr(encodeNode(node))
	if C.ksb_node_upsert(c.g.e, p, l) != 0 {
		return c.g.lastError()
	}
	return nil
}

// AddNode / UpdateNode (cache.go:347-371) -> NodeInfo.SetNode.
func (c *Cache) AddNode(node *v1.Node) error {
	if err := c.Cache.AddNode(node); err != nil {
		return err
	}
	return c.nodeUpsert(node)
}

func (c *Cache) UpdateNode(oldNode, newNode *v1.Node) error {
	if err := c.Cache.UpdateNode(oldNode, newNode); err != nil {
		return err
	}
	return c.nodeUpsert(newNode)
}

// RemoveNode (cache.go:373-387).
func (c *Cache) RemoveNode(node *v1.Node) error {
	if err := c.Cache.RemoveNode(node); err != nil {
		return err
	}
	name, nameLen := strArg(node.Name)
	if C.ksb_node_remove(c.g.e, name, nameLen) != 0 {
		return c.g.lastError()
	}
	return nil
}

// Expired assumed pods (cache.go:426-462) leave the Go cache on a timer, not through this interface: the factory's
// cleanup goroutine must call Expired for each pod it drops (one added line next to cache.go:449).
func (c *Cache) Expired(pod *v1.Pod) { _ = c.podRemove(pod) }

// The listers SelectorSpread reads (priorities/metadata.go:69-96) reach the engine from the informers' event
// handlers: kind 0 service, 1 replication controller (map selectors), 2 ReplicaSet, 3 StatefulSet (label selectors).
func (g *GPUScheduler) UpsertMapSelectorOwner(kind int, ns, name string, selector map[string]string) error {
	p, l := bytesPtr(encodeMapOwner(ns, name, selector))
	if C.ksb_selector_owner_upsert(g.e, C.int32_t(kind), p, l) != 0 {
		return g.lastError()
	}
	return nil
}

func (g *GPUScheduler) RemoveSelectorOwner(kind int, ns, name string) error {
	nsp, nsLen := strArg(ns)
	namep, nameLen := strArg(name)
	if C.ksb_selector_owner_remove(g.e, C.int32_t(kind), nsp, nsLen, namep, nameLen) != 0 {
		return g.lastError()
	}
	return nil
}

// the remaining methods (List, FilteredList, UpdateNodeNameToInfoMap, IsAssumedPod, FinishBinding) are the embedded
// cache's own
var _ = labels.Everything
