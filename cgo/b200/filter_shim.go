package b200

import (
	"errors"
	"strconv"
	"time"

	corev1 "k8s.io/api/core/v1"
	extenderv1 "k8s.io/kube-scheduler/extender/v1"

	"github.com/Project-HAMi/HAMi/pkg/device"
	"github.com/Project-HAMi/HAMi/pkg/util"
)

// GoPath is what the shim needs from *scheduler.Scheduler for the cases the engine hands back: the reference's own
// Filter for a pod or cluster outside the engine's scope, and the pieces of Filter's tail that stay in Go.
type GoPath interface {
	// FilterGo is the reference's unmodified Scheduler.Filter (pkg/scheduler/scheduler.go:1067-1154).
	FilterGo(args extenderv1.ExtenderArgs) (*extenderv1.ExtenderFilterResult, error)
	// FilterSimulationGo is the cluster-autoscaler path with node OBJECTS in the request (scheduler.go:1156-1199).
	FilterSimulationGo(args extenderv1.ExtenderArgs, reqs device.PodDeviceRequests) (*extenderv1.ExtenderFilterResult, error)
	// RecordFilterEvent is recordScheduleFilterResultEvent (event.go:45-78).
	RecordFilterEvent(pod *corev1.Pod, reason, msg string, err error)
	// QuotaAdd / QuotaRm keep quotaManager's own books for the webhook's pre-check (webhook.go:145-192).
	QuotaAdd(pod *corev1.Pod, devs device.PodDevices)
	QuotaRm(pod *corev1.Pod, devs device.PodDevices)
	// PodManagerAdd / PodManagerDel keep podManager for Bind, metrics and the informer handlers.
	PodManagerAdd(pod *corev1.Pod, node string, devs device.PodDevices) bool
	PodManagerTakeAndDelete(pod *corev1.Pod) (device.PodDevices, bool)
	PodManagerDel(pod *corev1.Pod)
}

// Filter is the drop-in for Scheduler.Filter: everything above line 1092 and below line 1118 of scheduler.go is the
// reference's own code, restated here only as far as the control flow needs it; the block in between — getNodesUsage,
// calcScore, sort.Sort and take-last — is ONE call into the engine.
//
//	func (s *Scheduler) Filter(args extenderv1.ExtenderArgs) (*extenderv1.ExtenderFilterResult, error) {
//	    return b200.Filter(s.engine, s, args)
//	}
func Filter(e *Engine, g GoPath, args extenderv1.ExtenderArgs) (*extenderv1.ExtenderFilterResult, error) {
	reqs := device.Resourcereqs(args.Pod) // scheduler.go:1069
	has := false
	for _, m := range reqs {
		if len(m) > 0 {
			has = true
			break
		}
	}
	if !has { // scheduler.go:1080-1087
		return &extenderv1.ExtenderFilterResult{NodeNames: args.NodeNames}, nil
	}
	if args.Nodes != nil { // scheduler.go:1088-1090: node objects arrive with the request; hs_filter_simulate is the
		// engine's entry for it once the caller maps node annotations to registrations — until then the Go path answers
		return g.FilterSimulationGo(args, reqs)
	}
	var names []string
	if args.NodeNames != nil {
		names = *args.NodeNames
	}
	// podManager.TakeAndDeletePod + quotaManager.RmUsage (scheduler.go:1092-1094): the engine forgets the pod's earlier
	// decision itself (hs_filter releases it first); Go's own books follow
	if old, ok := g.PodManagerTakeAndDelete(args.Pod); ok {
		g.QuotaRm(args.Pod, old)
	}
	res, err := e.Filter(args.Pod, reqs, names)
	if errors.Is(err, ErrUnsupported) { // > 8 devices on a node, a vendor the engine does not hold, a topology-aware pod beside MIG GPUs
		return g.FilterGo(args)
	}
	if err != nil { // scheduler.go:1096-1107
		g.RecordFilterEvent(args.Pod, "FilteringFailed", "", err)
		return nil, err
	}
	if res.PassThrough {
		return &extenderv1.ExtenderFilterResult{NodeNames: args.NodeNames}, nil
	}
	if !res.Fit { // scheduler.go:1109-1114
		g.RecordFilterEvent(args.Pod, "FilteringFailed", "", errors.New("no available node, "+strconv.Itoa(len(names))+" nodes do not meet"))
		return &extenderv1.ExtenderFilterResult{FailedNodes: res.FailedNodes}, nil
	}
	m := res.Node
	// scheduler.go:1124-1130: the engine returns every vendor's PatchAnnotations output; the two timestamps are Go's
	annotations := res.Annotations
	annotations[util.AssignedNodeAnnotations] = m.NodeID
	now := strconv.FormatInt(time.Now().Unix(), 10)
	annotations[util.AssignedTimeAnnotations] = now
	for _, k := range []string{"predicate-time", "mthreads.com/predicate-time"} {
		if _, ok := annotations[k]; ok {
			annotations[k] = now
		}
	}
	effective := device.CollapseInitContainerUsage(args.Pod, m.Devices) // scheduler.go:1132-1138
	added := g.PodManagerAdd(args.Pod, m.NodeID, effective)
	if added {
		g.QuotaAdd(args.Pod, effective)
	}
	if err := util.PatchPodAnnotations(args.Pod, annotations); err != nil { // scheduler.go:1139-1147
		g.RecordFilterEvent(args.Pod, "FilteringFailed", "", err)
		if added {
			g.QuotaRm(args.Pod, effective)
		}
		g.PodManagerDel(args.Pod)
		e.DeletePod(string(args.Pod.UID)) // hs_rollback
		return nil, err
	}
	g.RecordFilterEvent(args.Pod, "FilteringSucceed", "find fit node("+m.NodeID+")", nil)
	return &extenderv1.ExtenderFilterResult{NodeNames: &[]string{m.NodeID}}, nil
}
