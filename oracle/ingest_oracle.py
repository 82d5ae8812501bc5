"""TEST INFRASTRUCTURE ONLY — CPU restatement of the reference's ingest codecs.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this; the product decoders
are k8s-vgpu-scheduler_b200/csrc/hs_ingest.cc, written independently in C++.  Pinned against the reference's
own test vectors (tests/golden/decode_node_devices.json, unmarshal_node_devices.json, decode_pod_devices.json,
encode_node_devices.json, marshal_node_devices.json, ingest_funcs.json) by tests/test_ingest.py.

Small inputs, so plain Python; JSON goes through the stdlib parser with Go's encoding/json typing rules
re-imposed on top (case-insensitive keys, integer literals only for integer fields, range checks).
"""
import json
import re

_INT_RE = re.compile(r"^[+-]?[0-9]+$")


class DecodeError(Exception):
    pass


def _quote(s):
    return '"' + s.replace("\\", "\\\\").replace('"', '\\"') + '"'


def _parse_int(s, bits, fn="ParseInt"):
    """strconv.ParseInt(s, 10, bits) / strconv.Atoi (devices.go:331-364 call sites)."""
    if not _INT_RE.match(s):
        raise DecodeError("strconv.%s: parsing %s: invalid syntax" % (fn, _quote(s)))
    v = int(s)
    if not -(1 << (bits - 1)) <= v <= (1 << (bits - 1)) - 1:
        raise DecodeError("strconv.%s: parsing %s: value out of range" % (fn, _quote(s)))
    return v


def _parse_bool(s):
    if s in ("1", "t", "T", "TRUE", "true", "True"):
        return True
    if s in ("0", "f", "F", "FALSE", "false", "False"):
        return False
    raise DecodeError("strconv.ParseBool: parsing %s: invalid syntax" % _quote(s))


def _wrap(prefix, fn, *a):
    try:
        return fn(*a)
    except DecodeError as e:
        raise DecodeError("%s: %s" % (prefix, e))


def decode_node_devices(s):
    """device.DecodeNodeDevices, pkg/device/devices.go:317-381.  Returns a list of dicts (DeviceInfo fields)."""
    if ":" not in s:
        raise DecodeError("node annotation missing device separator")
    out = []
    for val in s.split(":"):
        if val == "":
            continue
        if "," not in val:
            raise DecodeError("malformed node annotation segment: %s" % _quote(val))
        it = val.split(",")
        if len(it) not in (7, 9):
            raise DecodeError("unexpected field count %d in node annotation" % len(it))
        count = _wrap("invalid count field", _parse_int, it[1], 32)
        devmem = _wrap("invalid memory field", _parse_int, it[2], 32)
        devcore = _wrap("invalid core field", _parse_int, it[3], 32)
        health = _wrap("invalid health field", _parse_bool, it[6])
        numa = _wrap("invalid numa field", _parse_int, it[5], 64, "Atoi")
        mode, index = "hami-core", 0
        if len(it) == 9:
            index = _wrap("invalid index field", _parse_int, it[7], 64, "Atoi")
            if index < 0:
                raise DecodeError("index field must not be negative: %d" % index)
            mode = it[8]
        out.append({"ID": it[0], "Count": count, "Devmem": devmem, "Devcore": devcore, "Type": it[4], "Numa": numa,
                    "Health": health, "Mode": mode, "Index": index})
    return out


class _Num(str):
    """a JSON number kept as its literal, so integer-ness is decided the way encoding/json decides it"""


class _Obj(list):
    """a JSON object as its (key, value) pairs in document order — duplicates kept, because encoding/json applies
    every member in turn (a later null does NOT undo an earlier value)"""


def _loads(text):
    try:
        return json.loads(text, parse_int=_Num, parse_float=_Num, object_pairs_hook=_Obj,
                          parse_constant=lambda c: (_ for _ in ()).throw(ValueError(c)))
    except ValueError as e:
        raise DecodeError("invalid JSON: %s" % e)


def _field(obj, name):
    """encoding/json matches keys case-insensitively and applies members in document order; null is a no-op for
    non-pointer fields, so the value that sticks is the last non-null one."""
    hit = [v for k, v in obj if k.lower() == name.lower() and v is not None]
    return hit[-1] if hit else None


def _as_int(v, bits, unsigned=False):
    if not isinstance(v, _Num) or not _INT_RE.match(v) or (unsigned and v.startswith("-")):
        raise DecodeError("json: cannot unmarshal into integer field")
    n = int(v)
    lo, hi = (0, (1 << 63) - 1) if unsigned else (-(1 << (bits - 1)), (1 << (bits - 1)) - 1)
    if not lo <= n <= hi:
        raise DecodeError("json: number out of range")
    return n


def _as(v, ty):
    if isinstance(v, _Num) or not isinstance(v, ty) or (ty is not bool and isinstance(v, bool)) or (ty is list and isinstance(v, _Obj)):
        raise DecodeError("json: cannot unmarshal into %s field" % ty.__name__)
    return v


def unmarshal_node_devices(text):
    """device.UnMarshalNodeDevices, pkg/device/devices.go:445-449 (struct tags devices.go:127-141)."""
    doc = _loads(text)
    if doc is None:
        return []
    if not isinstance(doc, list) or isinstance(doc, _Obj):
        raise DecodeError("json: cannot unmarshal into []*DeviceInfo")
    out = []
    for el in doc:
        if not isinstance(el, _Obj):
            raise DecodeError("json: bad device entry")
        d = {"ID": "", "Index": 0, "Count": 0, "Devmem": 0, "Devcore": 0, "Type": "", "Numa": 0, "Mode": "", "Health": False,
             "NetworkID": -1, "HasMigProfiles": False}
        for name, key, conv in (("id", "ID", lambda v: _as(v, str)), ("type", "Type", lambda v: _as(v, str)),
                                ("mode", "Mode", lambda v: _as(v, str)), ("health", "Health", lambda v: _as(v, bool)),
                                ("index", "Index", lambda v: _as_int(v, 64, True)), ("count", "Count", lambda v: _as_int(v, 32)),
                                ("devmem", "Devmem", lambda v: _as_int(v, 32)), ("devcore", "Devcore", lambda v: _as_int(v, 32)),
                                ("numa", "Numa", lambda v: _as_int(v, 64))):
            v = _field(el, name)
            if v is not None:
                d[key] = conv(v)
        for name, ty in (("devicevendor", str), ("migtemplate", list)):
            v = _field(el, name)
            if v is not None:
                _as(v, ty)
        v = _field(el, "migProfiles")
        if v is not None:
            d["HasMigProfiles"] = len(_as(v, list)) > 0
        v = _field(el, "custominfo")
        if v is not None:
            ci = dict(_as(v, _Obj))
            if isinstance(ci.get("NetworkID"), _Num):      # CustomInfo["NetworkID"].(float64) → int (ascend/device.go:402-404)
                d["NetworkID"] = int(float(ci["NetworkID"]))
        out.append(d)
    return out


def pair_score_matrix(text, devices):
    """device.DecodePairScores (devices.go:383-389) + the attach loop of nvidia/device.go:319-340, as the n x n
    row-major matrix hs_node_info.pair_scores wants."""
    n = len(devices)
    m = [0] * (n * n)
    doc = _loads(text)
    if doc is None:
        return m
    if not isinstance(doc, list) or isinstance(doc, _Obj):
        raise DecodeError("json: cannot unmarshal into DevicePairScores")
    recs = []
    for el in doc:
        if el is None:
            recs.append(("", {}))
            continue
        if not isinstance(el, _Obj):
            raise DecodeError("json: bad pair score entry")
        u = _field(el, "uuid")
        sc = _field(el, "score")
        scores = {}
        if sc is not None:
            for k, v in _as(sc, _Obj):
                scores[k] = 0 if v is None else _as_int(v, 64)
        recs.append((_as(u, str) if u is not None else "", scores))
    for i, d in enumerate(devices):
        for uuid, scores in recs:
            if uuid == d["ID"]:
                for j, p in enumerate(devices):
                    m[i * n + j] = scores.get(p["ID"], 0)
                break
    return m


def decode_container_devices(s):
    """device.DecodeContainerDevices, pkg/device/devices.go:496-524."""
    if len(s) == 0:
        return []
    out = []
    for val in s.split(":"):
        if "," in val:
            f = val.split(",")
            if len(f) < 4:
                raise DecodeError("pod annotation format error, missing fields, do not use nodeName in task spec")
            mem = _wrap("invalid memory field", _parse_int, f[2], 32)
            cores = _wrap("invalid core field", _parse_int, f[3], 32)
            out.append({"UUID": f[0], "Type": f[1], "Usedmem": mem, "Usedcores": cores})
    return out


def decode_pod_devices(checklist, annos):
    """device.DecodePodDevices, pkg/device/devices.go:526-556: {devID: [container devices, …]} with empty containers kept."""
    if len(annos) == 0:
        return {}
    pd = {}
    for dev_id, anno in sorted(checklist.items()):   # Go ranges the map in random order: which of two bad annotations is reported is not defined; canonical = key order
        if anno not in annos:
            continue
        pd[dev_id] = [decode_container_devices(seg) for seg in annos[anno].split(";")]
    return pd


def collapse_init_container_usage(n_init, raw):
    """device.CollapseInitContainerUsage, pkg/device/initContainer.go:36-114."""
    if raw is None:
        return None
    init_peak, app_sum = {}, {}
    for dev_type, single in raw.items():
        for cidx, ctr in enumerate(single):
            for dev in ctr or []:
                key = (dev_type, dev["UUID"])
                if cidx < n_init:
                    cur = init_peak.get(key, [0, 0, 0])
                    init_peak[key] = [max(cur[0], dev["Usedmem"]), max(cur[1], dev["Usedcores"]), 1]
                else:
                    cur = app_sum.get(key, [0, 0, 0])
                    app_sum[key] = [cur[0] + dev["Usedmem"], cur[1] + dev["Usedcores"], cur[2] + 1]
    out = {}
    for dev_type in raw:
        uuids = sorted({k[1] for k in list(init_peak) + list(app_sum) if k[0] == dev_type})
        devs = []
        for u in uuids:
            i = init_peak.get((dev_type, u), [0, 0, 0])
            a = app_sum.get((dev_type, u), [0, 0, 0])
            devs.append({"UUID": u, "Type": dev_type, "Usedmem": max(i[0], a[0]), "Usedcores": max(i[1], a[1]),
                         "Slots": max(i[2], a[2], 1)})
        out[dev_type] = [devs]
    return out


def app_containers_only_device_usage(n_init, raw):
    """device.AppContainersOnlyDeviceUsage, pkg/device/initContainer.go:118-157."""
    if raw is None:
        return None
    out = {}
    for dev_type, single in raw.items():
        sums = {}
        for cidx, ctr in enumerate(single):
            if cidx < n_init:
                continue
            for dev in ctr or []:
                s = sums.get(dev["UUID"], [0, 0, 0])
                sums[dev["UUID"]] = [s[0] + dev["Usedmem"], s[1] + dev["Usedcores"], s[2] + 1]
        out[dev_type] = [[{"UUID": u, "Type": dev_type, "Usedmem": s[0], "Usedcores": s[1], "Slots": s[2]}
                          for u, s in sorted(sums.items())]]
    return out


def _i64(v):
    v &= (1 << 64) - 1
    return v - (1 << 64) if v >> 63 else v


def _i32(v):
    v &= (1 << 32) - 1
    return v - (1 << 32) if v >> 31 else v


def _go_div(a, b):
    q = abs(a) // abs(b)
    return q if (a < 0) == (b < 0) else -q


def neuron_node_devices(node, counts, cores_total, cores_per, instance_type):
    """AWSNeuronDevices.GetNodeDevices, pkg/device/awsneuron/device.go:156-198.  cores_per = the vendor object's
    coresPerAWSNeuron so far (0: derived from this node); instance_type = the node.kubernetes.io/instance-type label ("" when
    absent).  NetworkID carries CustomInfo["AWSNodeType"] as graphSelect (:414) reads it: 1 = contains "inf"/"Inf", else 0."""
    if counts == 0:
        raise ValueError("device not found aws.amazon.com/neuron")
    u64 = lambda v: v & ((1 << 64) - 1)   # noqa: E731
    per = u64(cores_per) or u64(cores_total) // u64(counts)
    bits = _i32(per) if _i32(per) > 0 else 0   # for i < int(coresPerAWSNeuron); a mask is all ones after 64 rounds anyway
    mask = 0
    for _ in range(min(bits, 64)):
        mask = u64(mask * 2 + 1)
    ring = "inf" in (instance_type or "") or "Inf" in (instance_type or "")
    return [{"Index": i, "ID": "%s-AWSNeuron-%d" % (node, i), "Count": _i32(per), "Devmem": 0, "Devcore": _i32(mask), "Type": "AWSNeuron",
             "Numa": 0, "Health": True, "DeviceVendor": "AWSNeuron", "NetworkID": 1 if ring else 0} for i in range(max(counts, 0))]


def capacity_node_devices(vendor, node, cores, memory):
    """GetNodeDevices of the two vendors that register through node CAPACITY instead of an annotation:
    cambricon (pkg/device/cambricon/device.go:209-232) and mthreads (pkg/device/mthreads/device.go:130-153).
    cores / memory = AsInt64 of Status.Capacity[<cores resource>] / [<memory resource>] (0 when absent)."""
    if vendor == "Metax-GPU":          # metax/device.go:67-89: `cores` is the GPU count
        if cores == 0:
            raise ValueError("device not found metax-tech.com/gpu")
        return [{"Index": i, "ID": "%s-metax-%d" % (node, i), "Count": 1, "Devmem": 65536, "Devcore": 100, "Type": vendor, "Numa": 0,
                 "Health": True, "DeviceVendor": vendor} for i in range(max(cores, 0))]
    per_card, factor, infix, res = {"MLU": (100, 256, "-cambricon-mlu-", "cambricon.com/mlu.smlu.vcore"),
                                    "Mthreads": (16, 512, "-mthreads-", "mthreads.com/sgpu-core")}[vendor]
    if cores == 0:
        raise ValueError("device not found " + res)
    out, i = [], 0
    while i * per_card < cores:
        out.append({"Index": i, "ID": "%s%s%d" % (node, infix, i), "Count": 100,
                    "Devmem": _i32(_go_div(_i64(memory * factor * per_card), cores)), "Devcore": per_card, "Type": vendor,
                    "Numa": 0, "Health": True, "DeviceVendor": vendor})
        i += 1
    return out


def encode_node_devices(devs):
    """device.EncodeNodeDevices, pkg/device/devices.go:391-417 (used for round-trip checks)."""
    return "".join("%s,%d,%d,%d,%s,%d,%s,%d,%s:" % (d["ID"], d["Count"], d["Devmem"], d["Devcore"], d["Type"], d["Numa"],
                                                    "true" if d["Health"] else "false", d.get("Index", 0), d.get("Mode", ""))
                   for d in devs)
