"""ctypes mirror of include/hami_sched.h (the C ABI a cgo shim binds).

Plain data only: builders turn Python dicts into the caller-owned buffers the
ABI takes, readers turn hs_filter_result back into Python values.  Nothing here
computes anything; the engine lives in csrc/ (CUDA) behind libhami_sched.so.
"""
import ctypes as C

HS_FIT, HS_NOFIT, HS_PASSTHROUGH, HS_FIT_REMOTE = 0, 1, 2, 3
HS_ERR_INVALID, HS_ERR_CUDA, HS_ERR_UNSUPPORTED, HS_ERR_WEIGHTS, HS_ERR_NOTFOUND = -1, -2, -3, -4, -5
HS_OP_FILTER, HS_OP_DELETE = 0, 1

REASONS = [
    "CardTypeMismatch", "CardUuidMismatch", "CardTimeSlicingExhausted", "CardComputeUnitsExhausted",
    "CardInsufficientMemory", "CardInsufficientCore", "CardNotHealth", "CardCordoned", "NumaNotFit",
    "ExclusiveDeviceAllocateConflict", "CardNotFoundCustomFilterRule", "CardMigTopologyInfeasible",
    "AllocatedCardsInsufficientRequest", "ResourceQuotaNotFit", "ModeNotFit",
]


class hs_kv(C.Structure):
    _fields_ = [("key", C.c_char_p), ("value", C.c_char_p)]


class hs_enflame_info(C.Structure):
    _fields_ = [("minor", C.c_char_p), ("index", C.c_char_p), ("max_slice", C.c_int32),
                ("profiles", C.POINTER(hs_kv)), ("n_profiles", C.c_uint32)]


class hs_metax_sgpu_info(C.Structure):
    _fields_ = [("qos_policy", C.c_char_p), ("link_zone", C.c_int32)]


class hs_mig_placement(C.Structure):
    _fields_ = [("start", C.c_uint32), ("size", C.c_uint32)]


class hs_mig_profile(C.Structure):
    _fields_ = [("name", C.c_char_p), ("memory_mb", C.c_int32), ("core", C.c_int32), ("slice_count", C.c_uint32),
                ("placements", C.POINTER(hs_mig_placement)), ("n_placements", C.c_uint32)]


class hs_mig_reservation(C.Structure):
    _fields_ = [("gpu_uuid", C.c_char_p), ("profile", C.c_char_p), ("start", C.c_uint32), ("size", C.c_uint32)]


class hs_device_info(C.Structure):
    _fields_ = [("id", C.c_char_p), ("index", C.c_uint32), ("count", C.c_int32), ("devmem", C.c_int32),
                ("devcore", C.c_int32), ("type", C.c_char_p), ("numa", C.c_int32), ("mode", C.c_char_p),
                ("health", C.c_uint8), ("vendor", C.c_char_p), ("network_id", C.c_int32),
                ("mig_profiles", C.POINTER(hs_mig_profile)), ("n_mig_profiles", C.c_uint32),
                ("enflame", C.POINTER(hs_enflame_info)), ("sgpu", C.POINTER(hs_metax_sgpu_info))]


def build_sgpu_info(info, keep):
    """{"qos_policy", "link_zone"} -> POINTER(hs_metax_sgpu_info) or None"""
    if info is None:
        return None
    e = hs_metax_sgpu_info(_b(info.get("qos_policy", "")), int(info.get("link_zone", 0)))
    keep.append(e)
    return C.pointer(e)


def build_enflame_info(info, keep):
    """{"minor", "index", "max_slice", "profiles": {name: id}} -> POINTER(hs_enflame_info) or None"""
    if not info:
        return None
    profs = list((info.get("profiles") or {}).items())
    kv = (hs_kv * max(1, len(profs)))()
    for i, (k, v) in enumerate(profs):
        kv[i] = hs_kv(_b(k), _b(v))
    e = hs_enflame_info(_b(info.get("minor", "")), _b(info.get("index", "")), int(info.get("max_slice", 0)), kv, len(profs))
    keep += [kv, e]
    return C.pointer(e)


def build_mig_profiles(profiles, keep):
    """[{name, memory_mb, core, slice_count, placements: [(start, size)]}] -> (hs_mig_profile array, n)"""
    if not profiles:
        return None, 0
    arr = (hs_mig_profile * len(profiles))()
    for i, p in enumerate(profiles):
        pl = p.get("placements") or []
        parr = (hs_mig_placement * max(1, len(pl)))()
        for j, (st, sz) in enumerate(pl):
            parr[j] = hs_mig_placement(st, sz)
        keep.append(parr)
        arr[i] = hs_mig_profile(_b(p["name"]), p.get("memory_mb", 0), p.get("core", 0), p.get("slice_count", 0),
                                parr, len(pl))
    keep.append(arr)
    return arr, len(profiles)


def build_mig_reservations(res, keep):
    """[(gpu_uuid, profile, start, size)] -> (hs_mig_reservation array, n)"""
    arr = (hs_mig_reservation * max(1, len(res)))()
    for i, (u, p, st, sz) in enumerate(res):
        arr[i] = hs_mig_reservation(_b(u), _b(p), st, sz)
    keep.append(arr)
    return arr, len(res)


class hs_node_info(C.Structure):
    _fields_ = [("name", C.c_char_p), ("devices", C.POINTER(hs_device_info)), ("n_devices", C.c_uint32),
                ("pair_scores", C.POINTER(C.c_int32)), ("device_cordon", C.c_char_p), ("hami_vnpu_core", C.c_int8),
                ("metax_losses", C.c_char_p), ("metax_scores", C.c_char_p)]


class hs_vendor_registration(C.Structure):
    _fields_ = [("vendor", C.c_char_p), ("format", C.c_int32), ("annotation", C.c_char_p), ("pair_scores", C.c_char_p),
                ("node_name", C.c_char_p), ("capacity_cores", C.c_int64), ("capacity_memory", C.c_int64),
                ("capacity_aux", C.c_int64)]


class hs_pod_device(C.Structure):
    _fields_ = [("key", C.c_char_p), ("container", C.c_uint32), ("uuid", C.c_char_p), ("type", C.c_char_p),
                ("usedmem", C.c_int32), ("usedcores", C.c_int32), ("slots", C.c_int32)]


class hs_sim_node(C.Structure):
    _fields_ = [("name", C.c_char_p), ("regs", C.POINTER(hs_vendor_registration)), ("n_regs", C.c_uint32),
                ("device_cordon", C.c_char_p), ("hami_vnpu_core", C.c_int8)]


HS_REG_CSV, HS_REG_JSON, HS_REG_CAPACITY_MLU, HS_REG_CAPACITY_MTHREADS, HS_REG_CAPACITY_METAX, HS_REG_CAPACITY_NEURON = 0, 1, 2, 3, 4, 5
HS_REG_CAPACITY_KUNLUN, HS_REG_ENFLAME_DRS, HS_REG_METAX_SGPU, HS_REG_CAPACITY_GCU = 6, 7, 8, 9


class hs_container_request(C.Structure):
    _fields_ = [("key", C.c_char_p), ("type", C.c_char_p), ("nums", C.c_int32), ("memreq", C.c_int32),
                ("mempct", C.c_int32), ("coresreq", C.c_int32)]


class hs_container(C.Structure):
    _fields_ = [("reqs", C.POINTER(hs_container_request)), ("n_reqs", C.c_uint32), ("name", C.c_char_p)]


class hs_pod(C.Structure):
    _fields_ = [("uid", C.c_char_p), ("ns", C.c_char_p), ("name", C.c_char_p),
                ("annotations", C.POINTER(hs_kv)), ("n_annotations", C.c_uint32),
                ("containers", C.POINTER(hs_container)), ("n_containers", C.c_uint32), ("n_init", C.c_uint32)]


class hs_alloc_entry(C.Structure):
    _fields_ = [("container", C.c_uint32), ("type_key", C.c_char_p), ("uuid", C.c_char_p),
                ("idx", C.c_int32), ("usedmem", C.c_int32), ("usedcores", C.c_int32),
                ("mig_profile", C.c_char_p), ("mig_start", C.c_uint32), ("mig_size", C.c_uint32)]


class hs_failed_node(C.Structure):
    _fields_ = [("node", C.c_char_p), ("reason", C.c_char_p)]


class hs_filter_result(C.Structure):
    _fields_ = [("status", C.c_int32), ("node_index", C.c_int32), ("node_name", C.c_char_p), ("score", C.c_float),
                ("entries", C.POINTER(hs_alloc_entry)), ("n_entries", C.c_uint32),
                ("annotations", C.POINTER(hs_kv)), ("n_annotations", C.c_uint32),
                ("failed", C.POINTER(hs_failed_node)), ("n_failed", C.c_uint32)]


class hs_device_usage_row(C.Structure):
    _fields_ = [("used", C.c_int32), ("usedmem", C.c_int32), ("usedcores", C.c_int32)]


class hs_pod_usage_row(C.Structure):
    _fields_ = [("uuid", C.c_char_p), ("slots", C.c_int32), ("usedmem", C.c_int32), ("usedcores", C.c_int32)]


class hs_ascend_template(C.Structure):
    _fields_ = [("name", C.c_char_p), ("memory", C.c_int64)]


class hs_ascend_config(C.Structure):
    _fields_ = [("common_word", C.c_char_p), ("memory_allocatable", C.c_int64), ("memory_capacity", C.c_int64),
                ("templates", C.POINTER(hs_ascend_template)), ("n_templates", C.c_uint32)]


class hs_quantity(C.Structure):
    _fields_ = [("name", C.c_char_p), ("as_int64", C.c_int64), ("as_int64_ok", C.c_uint8), ("value", C.c_int64), ("sign", C.c_int8),
                ("has_suffix", C.c_uint8)]


class hs_container_resources(C.Structure):
    _fields_ = [("limits", C.POINTER(hs_quantity)), ("n_limits", C.c_uint32), ("requests", C.POINTER(hs_quantity)), ("n_requests", C.c_uint32)]


class hs_vendor_resources(C.Structure):
    _fields_ = [("common_word", C.c_char_p), ("kind", C.c_int32), ("count_name", C.c_char_p), ("memory_name", C.c_char_p),
                ("memory_percentage_name", C.c_char_p), ("cores_name", C.c_char_p), ("default_memory", C.c_int32),
                ("default_cores", C.c_int32), ("memory_factor", C.c_int32)]


HS_REQ_NVIDIA, HS_REQ_DCU, HS_REQ_MLU, HS_REQ_ASCEND = 0, 1, 2, 3
HS_REQ_ILUVATAR, HS_REQ_COUNT, HS_REQ_AMD, HS_REQ_MTHREADS, HS_REQ_WHOLE_CARD, HS_REQ_METAX, HS_REQ_NEURON = 4, 5, 6, 7, 8, 9, 10
HS_REQ_KUNLUN_VXPU, HS_REQ_ENFLAME, HS_REQ_METAX_SGPU = 11, 12, 13


class hs_vendor_config(C.Structure):
    _fields_ = [("common_word", C.c_char_p), ("fit_kind", C.c_int32), ("use_uuid_key", C.c_char_p),
                ("nouse_uuid_key", C.c_char_p), ("to_allocate_key", C.c_char_p), ("allocated_key", C.c_char_p),
                ("use_type_key", C.c_char_p), ("nouse_type_key", C.c_char_p), ("option", C.c_int32)]


HS_FIT_TEMPLATE, HS_FIT_COUNT_ONLY, HS_FIT_AMD, HS_FIT_MTHREADS, HS_FIT_WHOLE_CARD, HS_FIT_METAX, HS_FIT_NEURON, HS_FIT_VASTAI = 0, 1, 2, 3, 4, 5, 6, 7
HS_FIT_KUNLUN, HS_FIT_KUNLUN_VXPU, HS_FIT_ENFLAME_VGCU, HS_FIT_METAX_SGPU = 8, 9, 10, 11

# The plain-Fit vendors of the reference's default device config: iluvatar's CommonWords
# (pkg/scheduler/config/config.go iluvatars: BI-V100 / BI-V150 / MR-V100 / MR-V50) and biren (biren/device.go:40-47).
DEFAULT_VENDORS = [
    ("BI-V100", HS_FIT_TEMPLATE, None, None, None, None),
    ("BI-V150", HS_FIT_TEMPLATE, None, None, None, None),
    ("MR-V100", HS_FIT_TEMPLATE, None, None, None, None),
    ("MR-V50", HS_FIT_TEMPLATE, None, None, None, None),
    ("Biren", HS_FIT_COUNT_ONLY, "birentech.com/use-gpuuuid", "birentech.com/nouse-gpuuuid", None, None),
    # amd/device.go:40-52,62-66
    ("AMD", HS_FIT_AMD, "amd.com/use-gpu-uuid", "amd.com/nouse-gpu-uuid", "hami.io/amd-devices-to-allocate",
     "hami.io/amd-devices-allocated", "amd.com/use-gputype", "amd.com/nouse-gputype"),
    # vastai in card mode (vastai/device.go:228-291 without DeviceStrategy "die") has biren's gate list
    ("Vastai", HS_FIT_VASTAI, "vastaitech.com/use-gpuuuid", "vastaitech.com/nouse-gpuuuid", None, None),
    # enflame whole-card GCU (enflame/gcu.go:31-46)
    ("GCU", HS_FIT_WHOLE_CARD, None, None, "hami.io/enflame-gcu-devices-to-allocate", "hami.io/enflame-gcu-devices-allocated"),
    # mthreads/device.go:44-58,76-80
    ("Mthreads", HS_FIT_MTHREADS, "mthreads.ai/use-gpuuuid", "mthreads.ai/nouse-gpuuuid",
     "hami.io/mthreads-vgpu-devices-to-allocate", "hami.io/mthreads-vgpu-devices-allocated"),
    # metax whole GPUs (metax/device.go:35-56): no uuid lists
    ("Metax-GPU", HS_FIT_METAX, None, None, "hami.io/metax-gpu-devices-to-allocate", "hami.io/metax-gpu-devices-allocated"),
    # awsneuron/device.go:42-58,68-79 (SupportDevices only: no to-allocate key is written)
    ("AWSNeuron", HS_FIT_NEURON, "aws.amazon.com/use-neuron-uuid", "aws.amazon.com/nouse-neuron-uuid", None,
     "hami.io/aws-neuron-devices-allocated"),
    # kunlun whole XPUs (kunlun/device.go:32-56): SupportDevices only; the second uuid pair (baidu.com/...) is fixed in the Fit
    ("kunlun", HS_FIT_KUNLUN, "hami.io/use-xpu-uuid", "hami.io/no-use-xpu-uuid", None, "hami.io/kunlun-allocated"),
    # kunlun shared XPUs (kunlun/vdevice.go:34-63)
    # metax sGPU (metax/sdevice.go:39-46,63-84); option = MetaxConfig.TopologyAware
    ("Metax-SGPU", HS_FIT_METAX_SGPU, "metax-tech.com/use-gpuuuid", "metax-tech.com/nouse-gpuuuid",
     "hami.io/metax-sgpu-devices-to-allocate", "hami.io/metax-sgpu-devices-allocated", None, None, 0),
    # enflame shared vGCU (enflame/device.go:38-46,86-104): SupportDevices only
    ("Enflame", HS_FIT_ENFLAME_VGCU, "enflame.com/use-gpuuuid", "enflame.com/nouse-gpuuuid", None, "hami.io/enflame-vgpu-devices-allocated"),
    ("XPU", HS_FIT_KUNLUN_VXPU, "hami.io/use-xpu-uuid", "hami.io/no-use-xpu-uuid", "hami.io/xpu-devices-to-allocate",
     "hami.io/xpu-devices-allocated"),
]


class hs_config(C.Structure):
    _fields_ = [("node_scheduler_policy", C.c_char_p), ("gpu_scheduler_policy", C.c_char_p),
                ("ascend", C.POINTER(hs_ascend_config)), ("n_ascend", C.c_uint32),
                ("ascend_hami_vnpu_core", C.c_uint8), ("cuda_device", C.c_int32), ("reserve_nodes", C.c_uint32),
                ("vendors", C.POINTER(hs_vendor_config)), ("n_vendors", C.c_uint32)]


class hs_op(C.Structure):
    _fields_ = [("kind", C.c_int32), ("pod", C.POINTER(hs_pod))]


def _b(s):
    if s is None:
        return None
    return s if isinstance(s, bytes) else str(s).encode()


# Ascend VNPU configs of the reference's built-in default device config
# (pkg/scheduler/config/config.go:340-460): commonWord, allocatable, capacity, templates.
DEFAULT_ASCEND = [
    ("Ascend910A", 32768, 32768, [("vir02", 2184), ("vir04", 4369), ("vir08", 8738), ("vir16", 17476)]),
    ("Ascend910B2", 65536, 65536, [("vir03_1c_8g", 8192), ("vir06_1c_16g", 16384), ("vir12_3c_32g", 32768)]),
    ("Ascend910B3", 65536, 65536, [("vir05_1c_16g", 16384), ("vir10_3c_32g", 32768)]),
    ("Ascend910B4", 32768, 32768, [("vir05_1c_8g", 8192), ("vir10_3c_16g", 16384)]),
    ("Ascend910C", 65536, 65536, []),
    ("Ascend310P", 21527, 24576, [("vir01", 3072), ("vir02", 6144), ("vir04", 12288)]),
]


def build_config(node_policy=None, gpu_policy=None, ascend=None, ascend_hami_vnpu_core=False,
                 cuda_device=-1, reserve_nodes=0, vendors=None):
    keep = []
    vendors = DEFAULT_VENDORS if vendors is None else vendors
    varr = (hs_vendor_config * max(1, len(vendors)))()
    for i, v in enumerate(vendors):
        word, kind, use, nouse, toalloc, alloc = v[:6]
        use_t, nouse_t = (v[6], v[7]) if len(v) > 6 else (None, None)
        varr[i] = hs_vendor_config(_b(word), kind, _b(use), _b(nouse), _b(toalloc), _b(alloc), _b(use_t), _b(nouse_t),
                                   int(v[8]) if len(v) > 8 else 0)
    keep.append(varr)
    ascend = DEFAULT_ASCEND if ascend is None else ascend
    arr = (hs_ascend_config * max(1, len(ascend)))()
    for i, (word, alloc, cap, templates) in enumerate(ascend):
        t = (hs_ascend_template * max(1, len(templates)))()
        for j, (name, mem) in enumerate(templates):
            t[j].name = _b(name)
            t[j].memory = mem
        keep.append(t)
        arr[i].common_word = _b(word)
        arr[i].memory_allocatable = alloc
        arr[i].memory_capacity = cap
        arr[i].templates = t
        arr[i].n_templates = len(templates)
    cfg = hs_config(_b(node_policy), _b(gpu_policy), arr, len(ascend), 1 if ascend_hami_vnpu_core else 0,
                    cuda_device, reserve_nodes, varr, len(vendors))
    keep.append(arr)
    return cfg, keep


def build_node(node):
    """node: {"name", "devices": [{id,index,count,devmem,devcore,type,numa,mode,health,vendor,network_id}],
    "pair_scores": [[int]] | None, "device_cordon": str | None, "hami_vnpu_core": -1|0|1}"""
    devs = node["devices"]
    arr = (hs_device_info * max(1, len(devs)))()
    keep = [arr]
    for i, d in enumerate(devs):
        mp, nmp = build_mig_profiles(d.get("mig_profiles"), keep)
        arr[i] = hs_device_info(_b(d["id"]), d.get("index", i), d.get("count", 0), d.get("devmem", 0),
                                d.get("devcore", 0), _b(d.get("type", "")), d.get("numa", 0), _b(d.get("mode", "")),
                                1 if d.get("health", True) else 0, _b(d.get("vendor", "NVIDIA")),
                                d.get("network_id", -1), mp, nmp, build_enflame_info(d.get("enflame"), keep),
                                build_sgpu_info(d.get("sgpu"), keep))
    ps = None
    if node.get("pair_scores") is not None:
        n = len(devs)
        flat = (C.c_int32 * max(1, n * n))()
        for i in range(n):
            for j in range(n):
                flat[i * n + j] = int(node["pair_scores"][i][j])
        ps = C.cast(flat, C.POINTER(C.c_int32))
        keep.append(flat)
    ni = hs_node_info(_b(node["name"]), arr, len(devs), ps, _b(node.get("device_cordon")),
                      node.get("hami_vnpu_core", -1), _b(node.get("metax_losses")), _b(node.get("metax_scores")))
    return ni, keep


def build_pod(pod):
    """pod: {"uid","ns","name","annotations": {k: v}, "containers": [[{key,type,nums,memreq,mempct,coresreq}]], "n_init"}"""
    keep = []
    annos = pod.get("annotations") or {}
    kv = (hs_kv * max(1, len(annos)))()
    for i, (k, v) in enumerate(annos.items()):
        kv[i] = hs_kv(_b(k), _b(v))
    ctrs = pod.get("containers") or []
    carr = (hs_container * max(1, len(ctrs)))()
    for i, reqs in enumerate(ctrs):
        rarr = (hs_container_request * max(1, len(reqs)))()
        for j, r in enumerate(reqs):
            rarr[j] = hs_container_request(_b(r.get("key", r["type"])), _b(r["type"]), r.get("nums", 0),
                                           r.get("memreq", 0), r.get("mempct", 0), r.get("coresreq", 0))
        keep.append(rarr)
        names = pod.get("container_names") or []
        carr[i] = hs_container(rarr, len(reqs), _b(names[i]) if i < len(names) and names[i] is not None else None)
    keep += [kv, carr]
    p = hs_pod(_b(pod.get("uid", pod.get("name", ""))), _b(pod.get("ns", "default")), _b(pod.get("name", "")),
               kv, len(annos), carr, len(ctrs), pod.get("n_init", 0))
    return p, keep


def _s(p):
    return p.decode() if p is not None else None


def read_result(res):
    """hs_filter_result -> dict. devices: {type_key: [[(uuid, idx, usedmem, usedcores), ...] per container]}; a pick
    on a GPU in MIG mode carries its reservation too: (uuid, idx, usedmem, usedcores, profile, start, size)"""
    out = {"status": res.status, "node_index": res.node_index, "node": _s(res.node_name), "score": res.score,
           "devices": {}, "annotations": {}, "failed": {}}
    for i in range(res.n_entries):
        e = res.entries[i]
        lst = out["devices"].setdefault(_s(e.type_key), [])
        while len(lst) <= e.container:
            lst.append([])
        if e.uuid is not None:
            row = (_s(e.uuid), e.idx, e.usedmem, e.usedcores)
            if e.mig_profile is not None:
                row += (_s(e.mig_profile), e.mig_start, e.mig_size)
            lst[e.container].append(row)
    for i in range(res.n_annotations):
        out["annotations"][_s(res.annotations[i].key)] = _s(res.annotations[i].value)
    for i in range(res.n_failed):
        out["failed"][_s(res.failed[i].node)] = _s(res.failed[i].reason)
    return out
