// Package b200 binds the B200 Filter/Score engine (libhami_sched.so, include/hami_sched.h) for
// pkg/scheduler.  It is a thin cgo layer: every exported method marshals Go values into the C structs of
// the header, makes ONE call, and copies the result back into Go values before it returns — the C side
// keeps no Go pointer and never calls back into Go (cgo rules).  All C memory built for a call is freed
// by the arena that built it.
//
// The engine has no CPU fallback: New fails when no CUDA device is usable, and the scheduler then keeps
// its own Go path (see filter_shim.go).
//
// This file cannot be compiled in the builder's container (no Go toolchain there); it is written
// against include/hami_sched.h as committed and against the reference's types at
// pkg/device/devices.go:106-236 and pkg/scheduler/policy/node_policy.go:27-33.
package b200

/*
#cgo CFLAGS: -I${SRCDIR}/../../include
#cgo LDFLAGS: -L${SRCDIR}/../../k8s-vgpu-scheduler_b200 -lhami_sched -Wl,-rpath,${SRCDIR}/../../k8s-vgpu-scheduler_b200
#include <stdlib.h>
#include <string.h>
#include "hami_sched.h"
*/
import "C"

import (
	"errors"
	"fmt"
	"sort"
	"sync"
	"unsafe"

	corev1 "k8s.io/api/core/v1"

	"github.com/Project-HAMi/HAMi/pkg/device"
	"github.com/Project-HAMi/HAMi/pkg/device/nvidia"
	"github.com/Project-HAMi/HAMi/pkg/scheduler/policy"
)

// ErrNoFit is not an error of the call: the result carries FailedNodes instead of a node.
var ErrUnsupported = errors.New("b200: request or node outside the engine's scope, use the Go path")

// Engine is one hs_handle.  The handle is thread-compatible, not thread-safe: every method takes mu, which
// also gives the deterministic order the reference's sequential Filter calls have.
type Engine struct {
	mu sync.Mutex
	h  *C.hs_handle
}

// arena owns the C allocations of one call.
type arena struct{ ptrs []unsafe.Pointer }

func (a *arena) str(s string) *C.char {
	p := C.CString(s)
	a.ptrs = append(a.ptrs, unsafe.Pointer(p))
	return p
}

func (a *arena) alloc(n, size uintptr) unsafe.Pointer {
	if n == 0 {
		n = 1
	}
	p := C.calloc(C.size_t(n), C.size_t(size))
	a.ptrs = append(a.ptrs, p)
	return p
}

func (a *arena) free() {
	for _, p := range a.ptrs {
		C.free(p)
	}
	a.ptrs = nil
}

// AscendConfig mirrors one VNPUConfig of the device config file (pkg/device/ascend, config.go:340-460).
type AscendConfig struct {
	CommonWord        string
	MemoryAllocatable int64
	MemoryCapacity    int64
	Templates         []AscendTemplate
}
type AscendTemplate struct {
	Name   string
	Memory int64
}

// VendorConfig registers a further vendor with a plain Fit (hs_vendor_config; FitKind is one of C.HS_FIT_*).
type VendorConfig struct {
	CommonWord                                                     string
	FitKind                                                        int32
	UseUUIDKey, NoUseUUIDKey, ToAllocateKey, AllocatedKey          string
	UseTypeKey, NoUseTypeKey                                       string
	Option                                                         int32 // HS_FIT_METAX_SGPU: 1 = MetaxConfig.TopologyAware
}

// New replaces newScheduler + config.InitDevicesWithConfig for this path (cmd/scheduler/main.go:70-71 flags).
func New(nodePolicy, gpuPolicy string, ascend []AscendConfig, vendors []VendorConfig, ascendHamiVnpuCore bool, reserveNodes uint32) (*Engine, error) {
	var a arena
	defer a.free()
	var cfg C.hs_config
	cfg.node_scheduler_policy = a.str(nodePolicy)
	cfg.gpu_scheduler_policy = a.str(gpuPolicy)
	cfg.cuda_device = -1
	cfg.reserve_nodes = C.uint32_t(reserveNodes)
	if ascendHamiVnpuCore {
		cfg.ascend_hami_vnpu_core = 1
	}
	if len(ascend) > 0 {
		arr := (*[1 << 16]C.hs_ascend_config)(a.alloc(uintptr(len(ascend)), unsafe.Sizeof(C.hs_ascend_config{})))
		for i, ac := range ascend {
			arr[i].common_word = a.str(ac.CommonWord)
			arr[i].memory_allocatable = C.int64_t(ac.MemoryAllocatable)
			arr[i].memory_capacity = C.int64_t(ac.MemoryCapacity)
			if len(ac.Templates) > 0 {
				t := (*[1 << 16]C.hs_ascend_template)(a.alloc(uintptr(len(ac.Templates)), unsafe.Sizeof(C.hs_ascend_template{})))
				for j, tp := range ac.Templates {
					t[j].name = a.str(tp.Name)
					t[j].memory = C.int64_t(tp.Memory)
				}
				arr[i].templates = &t[0]
				arr[i].n_templates = C.uint32_t(len(ac.Templates))
			}
		}
		cfg.ascend = &arr[0]
		cfg.n_ascend = C.uint32_t(len(ascend))
	}
	if len(vendors) > 0 {
		arr := (*[1 << 16]C.hs_vendor_config)(a.alloc(uintptr(len(vendors)), unsafe.Sizeof(C.hs_vendor_config{})))
		opt := func(s string) *C.char {
			if s == "" {
				return nil
			}
			return a.str(s)
		}
		for i, v := range vendors {
			arr[i].common_word = a.str(v.CommonWord)
			arr[i].fit_kind = C.int32_t(v.FitKind)
			arr[i].use_uuid_key = opt(v.UseUUIDKey)
			arr[i].nouse_uuid_key = opt(v.NoUseUUIDKey)
			arr[i].to_allocate_key = opt(v.ToAllocateKey)
			arr[i].allocated_key = opt(v.AllocatedKey)
			arr[i].use_type_key = opt(v.UseTypeKey)
			arr[i].nouse_type_key = opt(v.NoUseTypeKey)
			arr[i].option = C.int32_t(v.Option)
		}
		cfg.vendors = &arr[0]
		cfg.n_vendors = C.uint32_t(len(vendors))
	}
	h := C.hs_create(&cfg)
	if h == nil {
		return nil, errors.New(C.GoString(C.hs_last_error(nil))) // no CPU fallback
	}
	return &Engine{h: h}, nil
}

func (e *Engine) Close() {
	e.mu.Lock()
	defer e.mu.Unlock()
	if e.h != nil {
		C.hs_destroy(e.h)
		e.h = nil
	}
}

func (e *Engine) lastError(rc C.int32_t) error {
	if rc == C.HS_ERR_UNSUPPORTED {
		return fmt.Errorf("%w: %s", ErrUnsupported, C.GoString(C.hs_last_error(e.h)))
	}
	return fmt.Errorf("b200: rc %d: %s", int(rc), C.GoString(C.hs_last_error(e.h)))
}

// UpsertNode mirrors nodeManager.addNode (pkg/scheduler/nodes.go:76-121) into HBM.  Devices are passed vendor by
// vendor in sorted vendor-key order, registration order inside a vendor (the engine's canonical device order).
// cordon is the value of hami.io/device-cordon ("" with hasCordon=false when absent); vnpuCore is the node's
// hami-vnpu-core annotation: -1 absent, 1 "true", 0 anything else.
func (e *Engine) UpsertNode(info *device.NodeInfo, cordon string, hasCordon bool, vnpuCore int8, metaxLosses, metaxScores *string) error {
	var a arena
	defer a.free()
	vendors := make([]string, 0, len(info.Devices))
	for v := range info.Devices {
		vendors = append(vendors, v)
	}
	sort.Strings(vendors)
	var devs []device.DeviceInfo
	for _, v := range vendors {
		devs = append(devs, info.Devices[v]...)
	}
	n := len(devs)
	arr := (*[1 << 12]C.hs_device_info)(a.alloc(uintptr(n), unsafe.Sizeof(C.hs_device_info{})))
	index := make(map[string]int, n)
	anyPair := false
	for i, d := range devs {
		index[d.ID] = i
		arr[i].id = a.str(d.ID)
		arr[i].index = C.uint32_t(d.Index)
		arr[i].count = C.int32_t(d.Count)
		arr[i].devmem = C.int32_t(d.Devmem)
		arr[i].devcore = C.int32_t(d.Devcore)
		arr[i]._type = a.str(d.Type)
		arr[i].numa = C.int32_t(d.Numa)
		arr[i].mode = a.str(d.Mode)
		if d.Health {
			arr[i].health = 1
		}
		arr[i].vendor = a.str(d.DeviceVendor)
		arr[i].network_id = -1
		if v, ok := d.CustomInfo["NetworkID"]; ok { // ascend/device.go: CustomInfo["NetworkID"]
			switch t := v.(type) {
			case int:
				arr[i].network_id = C.int32_t(t)
			case float64:
				arr[i].network_id = C.int32_t(t)
			}
		}
		if len(d.DevicePairScore.Scores) > 0 {
			anyPair = true
		}
		if qos, ok := d.CustomInfo["QosPolicy"]; ok { // metax sGPU: CustomInfo of convertMetaxSDeviceToHAMIDevice (metax/protocol.go:108-112)
			sg := (*C.hs_metax_sgpu_info)(a.alloc(1, unsafe.Sizeof(C.hs_metax_sgpu_info{})))
			sg.qos_policy = a.str(fmt.Sprint(qos))
			switch z := d.CustomInfo["LinkZone"].(type) {
			case int32:
				sg.link_zone = C.int32_t(z)
			case int:
				sg.link_zone = C.int32_t(z)
			case float64:
				sg.link_zone = C.int32_t(z)
			}
			arr[i].sgpu = sg
		}
		if profiles, ok := d.CustomInfo["profiles"]; ok { // enflame shared vGCU: CustomInfo of GetNodeDevices (enflame/device.go:235-252)
			ef := (*C.hs_enflame_info)(a.alloc(1, unsafe.Sizeof(C.hs_enflame_info{})))
			ef.minor = a.str(fmt.Sprint(d.CustomInfo["minor"]))
			ef.index = a.str(fmt.Sprint(d.CustomInfo["index"]))
			if ms, ok := d.CustomInfo["maxSlice"].(int); ok {
				ef.max_slice = C.int32_t(ms)
			}
			var names, ids []string
			switch m := profiles.(type) {
			case map[string]string:
				for k, v := range m {
					names, ids = append(names, k), append(ids, v)
				}
			case map[string]any:
				for k, v := range m {
					if sv, ok := v.(string); ok {
						names, ids = append(names, k), append(ids, sv)
					}
				}
			}
			if len(names) > 0 {
				kv := (*[1 << 8]C.hs_kv)(a.alloc(uintptr(len(names)), unsafe.Sizeof(C.hs_kv{})))
				for k := range names {
					kv[k].key = a.str(names[k])
					kv[k].value = a.str(ids[k])
				}
				ef.profiles = &kv[0]
				ef.n_profiles = C.uint32_t(len(names))
			}
			arr[i].enflame = ef
		}
		if d.Mode == "mig" && len(d.MIGProfiles) > 0 { // DeviceInfo.MIGProfiles, devices.go:141: what dynamic MIG selects from
			profs := (*[1 << 8]C.hs_mig_profile)(a.alloc(uintptr(len(d.MIGProfiles)), unsafe.Sizeof(C.hs_mig_profile{})))
			for k, mp := range d.MIGProfiles {
				profs[k].name = a.str(mp.Name)
				profs[k].memory_mb = C.int32_t(mp.MemoryMB)
				profs[k].core = C.int32_t(mp.Core)
				profs[k].slice_count = C.uint32_t(mp.SliceCount)
				if len(mp.Placements) > 0 {
					pl := (*[1 << 8]C.hs_mig_placement)(a.alloc(uintptr(len(mp.Placements)), unsafe.Sizeof(C.hs_mig_placement{})))
					for q, p := range mp.Placements {
						pl[q].start = C.uint32_t(p.Start)
						pl[q].size = C.uint32_t(p.Size)
					}
					profs[k].placements = &pl[0]
					profs[k].n_placements = C.uint32_t(len(mp.Placements))
				}
			}
			arr[i].mig_profiles = &profs[0]
			arr[i].n_mig_profiles = C.uint32_t(len(d.MIGProfiles))
		}
	}
	var ni C.hs_node_info
	ni.name = a.str(info.ID)
	if n > 0 {
		ni.devices = &arr[0]
	}
	ni.n_devices = C.uint32_t(n)
	if anyPair { // hami.io/node-nvidia-score, nvidia/device.go:319-340: row-major n x n
		ps := (*[1 << 24]C.int32_t)(a.alloc(uintptr(n*n), unsafe.Sizeof(C.int32_t(0))))
		for i, d := range devs {
			for other, sc := range d.DevicePairScore.Scores {
				if j, ok := index[other]; ok {
					ps[i*n+j] = C.int32_t(sc)
				}
			}
		}
		ni.pair_scores = &ps[0]
	}
	if hasCordon {
		ni.device_cordon = a.str(cordon)
	}
	ni.hami_vnpu_core = C.int8_t(vnpuCore)
	if metaxLosses != nil {
		ni.metax_losses = a.str(*metaxLosses)
	}
	if metaxScores != nil {
		ni.metax_scores = a.str(*metaxScores)
	}
	e.mu.Lock()
	defer e.mu.Unlock()
	if rc := C.hs_node_upsert(e.h, &ni); rc != C.HS_OK {
		return e.lastError(rc)
	}
	return nil
}

// RemoveNode mirrors nodeManager.rmNodeDevices (nodes.go:123-148).
func (e *Engine) RemoveNode(name string) {
	cs := C.CString(name)
	defer C.free(unsafe.Pointer(cs))
	e.mu.Lock()
	defer e.mu.Unlock()
	C.hs_node_remove(e.h, cs)
}

// AddPod mirrors podManager.AddPod fed by the informer (scheduler.go:137-171): the collapsed devices of a pod that
// is already bound (restart replay, pods another replica placed).
func (e *Engine) AddPod(uid, node string, devices device.PodDevices) error {
	cu, cn := C.CString(uid), C.CString(node)
	defer C.free(unsafe.Pointer(cu))
	defer C.free(unsafe.Pointer(cn))
	e.mu.Lock()
	defer e.mu.Unlock()
	for _, single := range devices {
		for _, ctr := range single {
			for _, d := range ctr {
				cd := C.CString(d.UUID)
				rc := C.hs_usage_apply(e.h, cu, cn, cd, C.int32_t(d.Slots), C.int32_t(d.Usedmem), C.int32_t(d.Usedcores))
				C.free(unsafe.Pointer(cd))
				if rc != C.HS_OK && rc != C.HS_ERR_NOTFOUND { // a UUID the node no longer advertises contributes nothing (join by UUID)
					return e.lastError(rc)
				}
			}
		}
	}
	return nil
}

// SetPodMigAllocations hands the engine the decoded hami.io/vgpu-mig-allocations annotation of a cached pod
// (nvidia.DecodeMigAllocations; read by getNodesUsage, scheduler.go:757-842).  Call it from onAddPod / onUpdatePod
// after AddPod: a GPU in MIG mode that Filter just picked stays unhealthy until its pod's reservation arrives here,
// as in the reference, where Filter caches the unpatched pod (scheduler.go:1132-1139).  A decode error in the caller
// counts as no reservation: pass nil.
func (e *Engine) SetPodMigAllocations(uid string, allocations []nvidia.MigAllocation) error {
	var a arena
	defer a.free()
	cu := a.str(uid)
	var first *C.hs_mig_reservation
	if len(allocations) > 0 {
		arr := (*[1 << 10]C.hs_mig_reservation)(a.alloc(uintptr(len(allocations)), unsafe.Sizeof(C.hs_mig_reservation{})))
		for i, al := range allocations {
			arr[i].gpu_uuid = a.str(al.GPUUUID)
			arr[i].profile = a.str(al.Profile)
			arr[i].start = C.uint32_t(al.Placement.Start)
			arr[i].size = C.uint32_t(al.Placement.Size)
		}
		first = &arr[0]
	}
	e.mu.Lock()
	defer e.mu.Unlock()
	if rc := C.hs_pod_set_mig(e.h, cu, first, C.uint32_t(len(allocations))); rc != C.HS_OK && rc != C.HS_ERR_NOTFOUND {
		return e.lastError(rc)
	}
	return nil
}

// DeletePod mirrors podManager.DelPod + the usage release (scheduler.go:226-253); also the rollback of a Filter whose
// annotation PATCH failed (scheduler.go:1140-1147).
func (e *Engine) DeletePod(uid string) {
	cu := C.CString(uid)
	defer C.free(unsafe.Pointer(cu))
	e.mu.Lock()
	defer e.mu.Unlock()
	C.hs_pod_delete(e.h, cu)
}

// Result is what the replaced block of Scheduler.Filter produces: the chosen NodeScore and the annotation map to
// PATCH (without the two timestamps, which the shim fills in), or the FailedNodes map.
type Result struct {
	Fit         bool
	PassThrough bool // the pod asks for no HAMi resource: every candidate passes (scheduler.go:1080-1087)
	Node        *policy.NodeScore
	Annotations map[string]string
	FailedNodes map[string]string
}

func buildPod(a *arena, pod *corev1.Pod, reqs device.PodDeviceRequests) *C.hs_pod {
	p := (*C.hs_pod)(a.alloc(1, unsafe.Sizeof(C.hs_pod{})))
	p.uid = a.str(string(pod.UID))
	p.ns = a.str(pod.Namespace)
	p.name = a.str(pod.Name)
	if n := len(pod.Annotations); n > 0 {
		kv := (*[1 << 16]C.hs_kv)(a.alloc(uintptr(n), unsafe.Sizeof(C.hs_kv{})))
		i := 0
		for k, v := range pod.Annotations {
			kv[i].key = a.str(k)
			kv[i].value = a.str(v)
			i++
		}
		p.annotations = &kv[0]
		p.n_annotations = C.uint32_t(n)
	}
	// device.Resourcereqs(pod) (devices.go:639-691): one map per container, init containers first
	if n := len(reqs); n > 0 {
		cs := (*[1 << 10]C.hs_container)(a.alloc(uintptr(n), unsafe.Sizeof(C.hs_container{})))
		nInit := len(pod.Spec.InitContainers)
		for ci, m := range reqs {
			// Container.Name by position: enflame's assigned-containers annotation is keyed by it (enflame/device.go:692-705)
			if ci < nInit {
				cs[ci].name = a.str(pod.Spec.InitContainers[ci].Name)
			} else if ci-nInit < len(pod.Spec.Containers) {
				cs[ci].name = a.str(pod.Spec.Containers[ci-nInit].Name)
			}
			if len(m) == 0 {
				continue
			}
			rq := (*[1 << 10]C.hs_container_request)(a.alloc(uintptr(len(m)), unsafe.Sizeof(C.hs_container_request{})))
			j := 0
			for key, r := range m {
				rq[j].key = a.str(key)
				rq[j]._type = a.str(r.Type)
				rq[j].nums = C.int32_t(r.Nums)
				rq[j].memreq = C.int32_t(r.Memreq)
				rq[j].mempct = C.int32_t(r.MemPercentagereq)
				rq[j].coresreq = C.int32_t(r.Coresreq)
				j++
			}
			cs[ci].reqs = &rq[0]
			cs[ci].n_reqs = C.uint32_t(len(m))
		}
		p.containers = &cs[0]
		p.n_containers = C.uint32_t(n)
	}
	p.n_init = C.uint32_t(len(pod.Spec.InitContainers))
	return p
}

// readResult copies an hs_filter_result (owned by the handle, valid until its next call) into Go values.
func readResult(res *C.hs_filter_result, nContainers int) *Result {
	out := &Result{}
	switch res.status {
	case C.HS_PASSTHROUGH:
		out.PassThrough = true
		return out
	case C.HS_NOFIT:
		out.FailedNodes = make(map[string]string, int(res.n_failed))
		failed := unsafe.Slice(res.failed, int(res.n_failed))
		for _, f := range failed {
			out.FailedNodes[C.GoString(f.node)] = C.GoString(f.reason)
		}
		return out
	}
	out.Fit = true
	devs := device.PodDevices{}
	entries := unsafe.Slice(res.entries, int(res.n_entries))
	for _, en := range entries {
		key := C.GoString(en.type_key)
		single := devs[key]
		for len(single) <= int(en.container) {
			single = append(single, device.ContainerDevices{})
		}
		if en.uuid != nil { // uuid == NULL marks an empty ContainerDevices at that position (score.go:266-282 padding)
			cd := device.ContainerDevice{
				Idx: int(en.idx), UUID: C.GoString(en.uuid), Type: key, Usedmem: int32(en.usedmem), Usedcores: int32(en.usedcores)}
			if en.mig_profile != nil { // the reservation AddResourceUsage made on a GPU in MIG mode (nvidia/device.go:653-657)
				cd.CustomInfo = map[string]any{
					nvidia.MigProfileCustomInfo:   C.GoString(en.mig_profile),
					nvidia.MigPlacementCustomInfo: device.MigPlacement{Start: uint32(en.mig_start), Size: uint32(en.mig_size)},
				}
			}
			single[en.container] = append(single[en.container], cd)
		}
		devs[key] = single
	}
	out.Node = &policy.NodeScore{NodeID: C.GoString(res.node_name), Devices: devs, Score: float32(res.score)}
	out.Annotations = make(map[string]string, int(res.n_annotations))
	annos := unsafe.Slice(res.annotations, int(res.n_annotations))
	for _, kv := range annos {
		out.Annotations[C.GoString(kv.key)] = C.GoString(kv.value)
	}
	return out
}

// Filter is the replaced block of Scheduler.Filter (scheduler.go:1092-1138): forget the pod's previous decision, sweep
// every candidate node, pick the arg-best by the canonical rule, commit the collapsed usage.  nodeNames == nil means
// "every registered node" (nodeCacheCapable extender configuration passes the names it wants considered).
func (e *Engine) Filter(pod *corev1.Pod, reqs device.PodDeviceRequests, nodeNames []string) (*Result, error) {
	return e.filter(pod, reqs, nodeNames, false)
}

// FilterDry decides like Filter but releases nothing, commits nothing and remembers nothing
// (filterSimulation on the live inventory, scheduler.go:1088-1090,1156-1199).
func (e *Engine) FilterDry(pod *corev1.Pod, reqs device.PodDeviceRequests, nodeNames []string) (*Result, error) {
	return e.filter(pod, reqs, nodeNames, true)
}

func (e *Engine) filter(pod *corev1.Pod, reqs device.PodDeviceRequests, nodeNames []string, dry bool) (*Result, error) {
	var a arena
	defer a.free()
	p := buildPod(&a, pod, reqs)
	var names **C.char
	if nodeNames != nil {
		arr := (*[1 << 24]*C.char)(a.alloc(uintptr(len(nodeNames)), unsafe.Sizeof((*C.char)(nil))))
		for i, s := range nodeNames {
			arr[i] = a.str(s)
		}
		names = &arr[0]
	}
	var res C.hs_filter_result
	e.mu.Lock()
	defer e.mu.Unlock()
	var rc C.int32_t
	if dry {
		rc = C.hs_filter_dry(e.h, p, names, C.uint32_t(len(nodeNames)), &res)
	} else {
		rc = C.hs_filter(e.h, p, names, C.uint32_t(len(nodeNames)), &res)
	}
	if rc < 0 {
		return nil, e.lastError(rc) // HS_ERR_WEIGHTS: invalid hami.io/device-scoring-weights, the whole Filter errors (score.go:361-364)
	}
	return readResult(&res, len(reqs)), nil // copied before the lock is released: the result belongs to the handle
}

// SetQuota / DelQuota mirror QuotaManager.AddQuota / DelQuota for one managed key (quota.go:243-289): which = 0 memory, 1 cores.
func (e *Engine) SetQuota(namespace, vendor string, which int32, limit int64) error {
	cn, cv := C.CString(namespace), C.CString(vendor)
	defer C.free(unsafe.Pointer(cn))
	defer C.free(unsafe.Pointer(cv))
	e.mu.Lock()
	defer e.mu.Unlock()
	if rc := C.hs_quota_set(e.h, cn, cv, C.int32_t(which), C.int64_t(limit)); rc != C.HS_OK {
		return e.lastError(rc)
	}
	return nil
}

func (e *Engine) DelQuota(namespace, vendor string, which int32) {
	cn, cv := C.CString(namespace), C.CString(vendor)
	defer C.free(unsafe.Pointer(cn))
	defer C.free(unsafe.Pointer(cv))
	e.mu.Lock()
	defer e.mu.Unlock()
	C.hs_quota_del(e.h, cn, cv, C.int32_t(which))
}

// NodeUsage reads one node's usage rows back from HBM (metrics, scheduler.go:630-639 overview).
func (e *Engine) NodeUsage(name string) ([]device.DeviceUsage, error) {
	cn := C.CString(name)
	defer C.free(unsafe.Pointer(cn))
	var rows [16]C.hs_device_usage_row
	e.mu.Lock()
	defer e.mu.Unlock()
	n := C.hs_snapshot_node(e.h, cn, &rows[0], 16)
	if n < 0 {
		return nil, e.lastError(n)
	}
	out := make([]device.DeviceUsage, int(n))
	for i := range out {
		out[i].Used, out[i].Usedmem, out[i].Usedcores = int32(rows[i].used), int32(rows[i].usedmem), int32(rows[i].usedcores)
	}
	return out, nil
}
