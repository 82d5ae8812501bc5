#!/usr/bin/env python3
"""Extract the reference's table-driven known-answer tests into JSON fixtures.

Run HERE (the container that has /root/reference); the GPU box never reads the
reference.  It parses the Go composite literals of the `tests := []struct{...}{...}`
tables in the reference's *_test.go files for the hot path into
plain JSON under tests/golden/.  Only DATA is extracted (inputs and expected
outputs of the reference's own tests); no reference code is copied.

    python tools/extract_golden.py [/root/reference] [tests/golden]
"""
import json
import os
import re
import sys

# --------------------------------------------------------------------------- tokenizer
TOKEN_RE = re.compile(r"""
    (?P<ws>\s+)
  | (?P<lcomment>//[^\n]*)
  | (?P<bcomment>/\*.*?\*/)
  | (?P<raw>`[^`]*`)
  | (?P<str>"(?:\\.|[^"\\])*")
  | (?P<char>'(?:\\.|[^'\\])')
  | (?P<num>0[xX][0-9a-fA-F]+|\d+\.\d*(?:[eE][-+]?\d+)?|\.\d+|\d+(?:[eE][-+]?\d+)?)
  | (?P<id>[A-Za-z_][A-Za-z_0-9]*)
  | (?P<op>:=|\.\.\.|&&|\|\||==|!=|<=|>=|<-|[{}()\[\],:&*.\-+/=<>!;|%^])
""", re.X | re.S)


def tokenize(src):
    toks = []
    pos = 0
    while pos < len(src):
        m = TOKEN_RE.match(src, pos)
        if not m:
            raise SyntaxError("cannot tokenize at %d: %r" % (pos, src[pos:pos + 40]))
        pos = m.end()
        k = m.lastgroup
        if k in ("ws", "lcomment", "bcomment"):
            continue
        toks.append((k, m.group(k)))
    return toks


def unquote(s):
    if s[0] == "`":
        return s[1:-1]
    return json.loads(s) if "\\x" not in s else bytes(s[1:-1], "utf-8").decode("unicode_escape")


# --------------------------------------------------------------------------- parser
class P:
    def __init__(self, toks):
        self.t = toks
        self.i = 0

    def peek(self, o=0):
        return self.t[self.i + o] if self.i + o < len(self.t) else ("eof", "")

    def next(self):
        tok = self.peek()
        self.i += 1
        return tok

    def accept(self, v):
        if self.peek()[1] == v and self.peek()[0] in ("op", "id"):
            self.i += 1
            return True
        return False

    def expect(self, v):
        if not self.accept(v):
            raise SyntaxError("expected %r got %r at tok %d" % (v, self.peek(), self.i))

    def skip_balanced(self, open_, close):
        depth = 0
        while True:
            k, v = self.next()
            if k == "eof":
                raise SyntaxError("unbalanced")
            if k == "op" and v == open_:
                depth += 1
            elif k == "op" and v == close:
                depth -= 1
                if depth == 0:
                    return

    # ---- types
    def parse_type(self):
        k, v = self.peek()
        if v == "[" and k == "op":
            self.next()
            if self.accept("]"):
                return "[]" + self.parse_type()
            n = self.next()[1]
            self.expect("]")
            return "[%s]" % n + self.parse_type()
        if v == "*" and k == "op":
            self.next()
            return "*" + self.parse_type()
        if v == "map" and k == "id":
            self.next()
            self.expect("[")
            kt = self.parse_type()
            self.expect("]")
            return "map[%s]%s" % (kt, self.parse_type())
        if v == "struct" and k == "id":
            self.next()
            self.skip_balanced("{", "}")
            return "struct"
        if v == "interface" and k == "id":
            self.next()
            self.skip_balanced("{", "}")
            return "interface"
        if v == "func" and k == "id":
            self.next()
            self.skip_balanced("(", ")")
            # optional result
            if self.peek()[1] == "(":
                self.skip_balanced("(", ")")
            elif self.peek()[0] == "id" or self.peek()[1] in ("*", "["):
                self.parse_type()
            return "func"
        if k == "id":
            name = self.next()[1]
            while self.peek()[1] == "." and self.peek(1)[0] == "id":
                self.next()
                name += "." + self.next()[1]
            return name
        raise SyntaxError("type expected at %r" % (self.peek(),))

    # ---- expressions
    def parse_expr(self):
        return self.parse_binary()

    def parse_binary(self):
        left = self.parse_unary()
        while self.peek()[0] == "op" and self.peek()[1] in ("+", "-", "*", "/", "==", "!=", "<", ">", "<=", ">=", "&&", "||", "%", "|"):
            op = self.next()[1]
            right = self.parse_unary()
            if isinstance(left, (int, float)) and isinstance(right, (int, float)) and op in "+-*/":
                left = {"+": left + right, "-": left - right, "*": left * right,
                        "/": (left // right if isinstance(left, int) and isinstance(right, int) else left / right)}[op]
            elif isinstance(left, str) and isinstance(right, str) and op == "+":
                left = left + right
            else:
                left = {"_bin": op, "l": left, "r": right}
        return left

    def parse_unary(self):
        k, v = self.peek()
        if k == "op" and v in ("&", "*"):
            self.next()
            return self.parse_unary()
        if k == "op" and v == "-":
            self.next()
            x = self.parse_unary()
            return -x if isinstance(x, (int, float)) else {"_neg": x}
        if k == "op" and v == "!":
            self.next()
            return {"_not": self.parse_unary()}
        return self.parse_postfix(self.parse_primary())

    def parse_primary(self):
        k, v = self.peek()
        if k in ("str", "raw"):
            self.next()
            return unquote(v)
        if k == "char":
            self.next()
            return v[1:-1]
        if k == "num":
            self.next()
            if v.lower().startswith("0x"):
                return int(v, 16)
            return float(v) if any(c in v for c in ".eE") else int(v)
        if k == "op" and v == "(":
            self.next()
            e = self.parse_expr()
            self.expect(")")
            return e
        if k == "op" and v == "{":
            return self.parse_body(None)
        if k == "id" and v == "func":
            self.parse_type()
            if self.peek()[1] == "{":
                self.skip_balanced("{", "}")
            return {"_func": True}
        if (k == "op" and v == "[") or (k == "id" and v in ("map", "struct")):
            ty = self.parse_type()
            if self.peek()[1] == "{":
                return self.parse_body(ty)
            if self.peek()[1] == "(":  # conversion like []byte("x")
                self.next()
                e = self.parse_expr()
                self.expect(")")
                return e
            return {"_type": ty}
        if k == "id":
            if v == "true":
                self.next(); return True
            if v == "false":
                self.next(); return False
            if v == "nil":
                self.next(); return None
            name = self.next()[1]
            return {"_id": name}
        raise SyntaxError("unexpected token %r at %d" % (self.peek(), self.i))

    def parse_postfix(self, e):
        while True:
            k, v = self.peek()
            if k == "op" and v == "." and self.peek(1)[0] == "id":
                self.next()
                name = self.next()[1]
                if isinstance(e, dict) and "_id" in e and len(e) == 1:
                    e = {"_id": e["_id"] + "." + name}
                else:
                    e = {"_sel": name, "x": e}
            elif k == "op" and v == "." and self.peek(1)[1] == "(":  # type assertion
                self.next()
                self.skip_balanced("(", ")")
            elif k == "op" and v == "(":
                self.next()
                args = []
                while not self.accept(")"):
                    if self.peek()[0] == "op" and self.peek()[1] == "[" or self.peek()[1] in ("map",):
                        # make([]T, n) / make(map[K]V)
                        save = self.i
                        try:
                            ty = self.parse_type()
                            if self.peek()[1] == "{":
                                args.append(self.parse_body(ty))
                            else:
                                args.append({"_type": ty})
                        except SyntaxError:
                            self.i = save
                            args.append(self.parse_expr())
                    else:
                        args.append(self.parse_expr())
                    self.accept(",")
                fn = e["_id"] if isinstance(e, dict) and "_id" in e and len(e) == 1 else e
                e = {"_call": fn, "args": args}
            elif k == "op" and v == "[":
                self.next()
                idx = self.parse_expr() if self.peek()[1] != ":" else None
                if self.accept(":"):
                    if self.peek()[1] != "]":
                        self.parse_expr()
                self.expect("]")
                e = {"_index": idx, "x": e}
            elif k == "op" and v == "{" and isinstance(e, dict) and "_id" in e and len(e) == 1:
                e = self.parse_body(e["_id"])
            else:
                return e

    def parse_body(self, ty):
        self.expect("{")
        elems = []
        while not self.accept("}"):
            first = self.parse_expr()
            if self.accept(":"):
                val = self.parse_expr()
                elems.append((first, val))
            else:
                elems.append((None, first))
            if not self.accept(","):
                self.expect("}")
                break
        if elems and all(k is not None for k, _ in elems):
            d = {}
            for k, v in elems:
                if isinstance(k, dict) and "_id" in k and len(k) == 1:
                    k = k["_id"]
                    # map keys written as Go constants resolve to their string value;
                    # struct field names (unqualified, non-map literal) stay as they are
                    if k in IDENTS and ("." in k or (ty or "").startswith("map")):
                        k = IDENTS[k]
                elif isinstance(k, dict) and "_call" in k:
                    # e.g. corev1.ResourceName("x") used as a key
                    k = k["args"][0] if k["args"] else str(k)
                    if isinstance(k, dict) and "_id" in k:
                        k = k["_id"]
                elif not isinstance(k, (str, int, float)):
                    k = json.dumps(k, sort_keys=True)
                d[str(k) if not isinstance(k, str) else k] = v
            out = {"_t": ty, "f": d} if ty else {"f": d}
        else:
            out = {"_t": ty, "e": [v for _, v in elems]} if ty else {"e": [v for _, v in elems]}
        return out


# --------------------------------------------------------------------------- simplify / resolve
IDENTS = {
    "nvidia.NvidiaGPUDevice": "NVIDIA",
    "NvidiaGPUDevice": "NVIDIA",
    "nvidia.GPUInUse": "nvidia.com/use-gputype",
    "nvidia.GPUNoUse": "nvidia.com/nouse-gputype",
    "GPUInUse": "nvidia.com/use-gputype",
    "GPUNoUse": "nvidia.com/nouse-gputype",
    "nvidia.NumaBind": "nvidia.com/numa-bind",
    "NumaBind": "nvidia.com/numa-bind",
    "nvidia.GPUUseUUID": "nvidia.com/use-gpuuuid",
    "nvidia.GPUNoUseUUID": "nvidia.com/nouse-gpuuuid",
    "GPUUseUUID": "nvidia.com/use-gpuuuid",
    "GPUNoUseUUID": "nvidia.com/nouse-gpuuuid",
    "nvidia.AllocateMode": "nvidia.com/vgpu-mode",
    "AllocateMode": "nvidia.com/vgpu-mode",
    "nvidia.MigMode": "mig", "MigMode": "mig",
    "nvidia.HamiCoreMode": "hami-core", "HamiCoreMode": "hami-core",
    "nvidia.MpsMode": "mps", "MpsMode": "mps",
    "nvidia.DeviceCordonAnnotation": "hami.io/device-cordon",
    "DeviceCordonAnnotation": "hami.io/device-cordon",
    "nvidia.RegisterGPUPairScore": "hami.io/node-nvidia-score",
    "hygon.HygonDCUDevice": "DCU", "HygonDCUDevice": "DCU",
    "DCUInUse": "hygon.com/use-dcutype", "DCUNoUse": "hygon.com/nouse-dcutype",
    "DCUUseUUID": "hygon.com/use-gpuuuid", "DCUNoUseUUID": "hygon.com/nouse-gpuuuid",
    "metax.MetaxGPUDevice": "Metax-GPU",
    "cambricon.CambriconMLUDevice": "MLU", "CambriconMLUDevice": "MLU",
    "MLUUseUUID": "cambricon.com/use-gpuuuid", "MLUNoUseUUID": "cambricon.com/nouse-gpuuuid",
    "util.GPUSchedulerPolicyAnnotationKey": "hami.io/gpu-scheduler-policy",
    "util.NodeSchedulerPolicyAnnotationKey": "hami.io/node-scheduler-policy",
    "util.DeviceScoringWeightsAnnotationKey": "hami.io/device-scoring-weights",
    "policy.GPUSchedulerPolicyAnnotationKey": "hami.io/gpu-scheduler-policy",
    "policy.NodeSchedulerPolicyAnnotationKey": "hami.io/node-scheduler-policy",
    "util.AssignedNodeAnnotations": "hami.io/vgpu-node",
    "util.GPUSchedulerPolicySpread": "spread", "util.GPUSchedulerPolicyBinpack": "binpack",
    "util.GPUSchedulerPolicyTopology": "topology-aware", "util.GPUSchedulerPolicyMutex": "mutex",
    "util.GPUSchedulerPolicyNuma": "numa",
    "util.NodeSchedulerPolicySpread": "spread", "util.NodeSchedulerPolicyBinpack": "binpack",
    "Ascend910CType": "Ascend910C", "Ascend910Prefix": "Ascend910",
    "VNPUModeAnnotation": "huawei.com/vnpu-mode", "VNPUModeHamiCore": "hami-core",
    "VNPUNodeSelectorAnnotation": "hami-vnpu-core",
    "AWSNeuronDevice": "AWSNeuron", "MetaxGPUDevice": "Metax-GPU", "MetaxGPUCommonWord": "Metax-GPU",
    "MthreadsGPUDevice": "Mthreads", "MthreadsUseUUID": "mthreads.ai/use-gpuuuid", "MthreadsNoUseUUID": "mthreads.ai/nouse-gpuuuid",
    "VastaiDevice": "Vastai", "VastaiUseUUID": "vastaitech.com/use-gpuuuid", "VastaiNoUseUUID": "vastaitech.com/nouse-gpuuuid",
    "EnflameGCUDevice": "GCU", "KunlunGPUDevice": "kunlun", "KunlunGPUCommonWord": "kunlun", "kunlun.KunlunGPUDevice": "kunlun",
    "XPUDevice": "XPU", "KunlunResourceVCount": "kunlunxin.com/vxpu", "KunlunResourceVMemory": "kunlunxin.com/vxpu-memory",
    "KunlunMaxMemory": 98304,
    "MetaxSGPUDevice": "Metax-SGPU", "MetaxSGPUCommonWord": "Metax-SGPU", "MetaxSGPUAppClass": "metax-tech.com/sgpu-app-class",
    "MetaxSGPUQosPolicy": "metax-tech.com/sgpu-qos-policy", "MetaxSGPUTopologyAware": "metax-tech.com/sgpu-topology-aware",
    "Online": "online", "Offline": "offline", "BestEffort": "best-effort", "FixedShare": "fixed-share", "BurstShare": "burst-share",
    "MetaxUseUUID": "metax-tech.com/use-gpuuuid", "MetaxNoUseUUID": "metax-tech.com/nouse-gpuuuid",
    "BirenDevice": "Biren", "BirenCommonWord": "Biren",
    "BirenUseUUID": "birentech.com/use-gpuuuid", "BirenNoUseUUID": "birentech.com/nouse-gpuuuid",
}
for k in list(IDENTS):
    if k.startswith("util."):
        IDENTS[k[5:]] = IDENTS[k]
for c in ["CardTypeMismatch", "CardUUIDMismatch", "CardTimeSlicingExhausted", "CardComputeUnitsExhausted",
          "CardInsufficientMemory", "CardInsufficientCore", "CardNotHealth", "CardCordoned", "NumaNotFit",
          "ExclusiveDeviceAllocateConflict", "CardNotFoundCustomFilterRule", "CardMigTopologyInfeasible",
          "NodeInsufficientDevice", "AllocatedCardsInsufficientRequest", "NodeUnfitPod", "NodeFitPod",
          "ResourceQuotaNotFit", "ModeNotFit"]:
    IDENTS["common." + c] = "CardUuidMismatch" if c == "CardUUIDMismatch" else c

PASSTHROUGH_CALLS = {"int32", "int64", "int", "uint", "uint32", "float32", "float64", "string", "k8stypes.UID",
                     "types.UID", "corev1.ResourceName", "ptr.To", "new", "float32Ptr", "SchedulerPolicyName",
                     "util.SchedulerPolicyName"}


def simplify(x):
    if isinstance(x, list):
        return [simplify(v) for v in x]
    if not isinstance(x, dict):
        return x
    if "_id" in x and len(x) == 1:
        return IDENTS.get(x["_id"], x)
    if "_call" in x:
        fn = x["_call"]
        args = [simplify(a) for a in x["args"]]
        if isinstance(fn, str):
            if fn.endswith(".String") and fn[:-7] in IDENTS:
                return IDENTS[fn[:-7]]
            if fn in PASSTHROUGH_CALLS and len(args) == 1:
                return args[0]
            if fn in ("resource.NewQuantity", "resource.NewMilliQuantity") and args:
                return args[0]
            if fn == "resource.MustParse" and args:
                return {"_quantity": args[0]}
            if fn == "make":
                return {} if (args and isinstance(args[0], dict) and str(args[0].get("_type", "")).startswith("map")) else []
            if fn == "fmt.Sprintf" and args and isinstance(args[0], str) and all(isinstance(a, (str, int, float)) for a in args[1:]):
                try:
                    return args[0] % tuple(args[1:])
                except Exception:
                    pass
        return {"_call": simplify(fn) if isinstance(fn, dict) else fn, "args": args}
    if "f" in x and set(x) <= {"_t", "f"}:
        d = {k: simplify(v) for k, v in x["f"].items()}
        return d
    if "e" in x and set(x) <= {"_t", "e"}:
        return [simplify(v) for v in x["e"]]
    if "_type" in x and len(x) == 1:
        return x
    return {k: simplify(v) for k, v in x.items()}


# --------------------------------------------------------------------------- table extraction
def find_func(src, name):
    m = re.search(r"^func %s\(" % re.escape(name), src, re.M)
    if not m:
        raise KeyError(name)
    # body ends at the next top-level "\n}\n"
    end = src.index("\n}\n", m.start())
    return src[m.start():end + 3]


def extract_table(body, var="tests"):
    if var is None:      # `for _, ts := range []struct{…}{…}`: the table literal sits in the range clause
        m = re.search(r"\brange\s+(\[\]struct\s*\{)", body)
    else:
        m = re.search(r"\b%s\s*:?=\s*(\[\]struct\s*\{)" % var, body)
    if not m:
        raise KeyError("no table %s" % var)
    toks = tokenize(body[m.start(1):])
    p = P(toks)
    ty = p.parse_type()
    lit = p.parse_body(ty)
    return simplify(lit)


TABLES = [
    # (fixture name, file, func, table var)
    ("calc_score", "pkg/scheduler/score_test.go", "Test_calcScore", "tests"),
    ("fit_in_certain_device", "pkg/scheduler/score_test.go", "Test_fitInCertainDevice", "tests"),
    ("fit_in_devices", "pkg/scheduler/score_test.go", "Test_fitInDevices", "tests"),
    ("nvidia_gpu_topology", "pkg/scheduler/score_test.go", "Test_Nvidia_GPU_Topology", "tests"),
    ("device_less", "pkg/scheduler/policy/gpu_policy_test.go", "TestDeviceUsageList_Less", "tests"),
    ("compute_score", "pkg/scheduler/policy/gpu_policy_test.go", "TestComputeScore", "tests"),
    ("node_less", "pkg/scheduler/policy/node_policy_test.go", "TestLess", "tests"),
    ("node_default_score", "pkg/scheduler/policy/node_policy_test.go", "TestComputeDefaultScore", "tests"),
    ("nvidia_fit", "pkg/device/nvidia/device_test.go", "TestDevices_Fit", "tests"),
    ("hygon_fit", "pkg/device/hygon/device_test.go", "TestDevices_Fit", "tests"),
    ("cambricon_fit", "pkg/device/cambricon/device_test.go", "TestDevices_Fit", "tests"),
    ("ascend_fit", "pkg/device/ascend/device_test.go", "TestDevices_Fit", "tests"),
    ("ascend_fit_910c", "pkg/device/ascend/device_test.go", "TestDevices_Fit_910C", "tests"),
    ("check_uuid", "pkg/device/devices_test.go", "TestCheckUUID", "tests"),
    ("check_type", "pkg/device/devices_test.go", "TestCheckType", "tests"),
    ("filter", "pkg/scheduler/scheduler_test.go", "Test_Filter", "tests"),
    ("weights_parse", "pkg/util/device_scoring_weights_test.go", "TestParseDeviceScoringWeights", "tests"),
    # ingest codecs
    ("decode_node_devices", "pkg/device/devices_test.go", "Test_DecodeNodeDevices", "tests"),
    ("unmarshal_node_devices", "pkg/device/devices_test.go", "TestUnMarshalNodeDevices", "tests"),
    ("decode_pod_devices", "pkg/device/devices_test.go", "Test_DecodePodDevices", "tests"),
    ("encode_node_devices", "pkg/device/devices_test.go", "Test_EncodeNodeDevices", "tests"),
    ("marshal_node_devices", "pkg/device/devices_test.go", "TestMarshalNodeDevices", "tests"),
    # namespace quota: quota = limits.nvidia.com/gpumem 2048 in namespace "default" (device_test.go:1324-1339)
    ("nvidia_fit_quota", "pkg/device/nvidia/device_test.go", "TestFitQuota", "tests"),
    # vendors with a plain Fit
    # request extraction: <vendor>.GenerateResourceRequests
    ("nvidia_generate_requests", "pkg/device/nvidia/device_test.go", "TestGenerateResourceRequests", "tests"),
    ("hygon_generate_requests", "pkg/device/hygon/device_test.go", "Test_GenerateResourceRequests", "tests"),
    ("hygon_generate_requests_range", "pkg/device/hygon/device_test.go", "Test_GenerateResourceRequests_OutOfRangeValues", "tests"),
    ("cambricon_generate_requests", "pkg/device/cambricon/device_test.go", "Test_GenerateResourceRequests", "tests"),
    ("ascend_generate_requests", "pkg/device/ascend/device_test.go", "Test_GenerateResourceRequests", "tests"),
    ("ascend_generate_requests_range", "pkg/device/ascend/device_test.go", "Test_GenerateResourceRequests_OutOfRangeValues", "tests"),
    ("iluvatar_generate_requests", "pkg/device/iluvatar/device_test.go", "Test_GenerateResourceRequests", "tests"),
    ("biren_generate_requests", "pkg/device/biren/device_test.go", "Test_GenerateResourceRequests", "tests"),
    ("amd_generate_requests", "pkg/device/amd/device_test.go", "Test_GenerateResourceRequests", "tests"),
    ("mthreads_generate_requests", "pkg/device/mthreads/device_test.go", "Test_GenerateResourceRequests", "tests"),
    ("vastai_generate_requests", "pkg/device/vastai/device_test.go", "Test_GenerateResourceRequests", "tests"),
    ("gcu_generate_requests", "pkg/device/enflame/gcu_test.go", "TestGCUDevices_GenerateResourceRequests", "tests"),
    ("iluvatar_fit", "pkg/device/iluvatar/device_test.go", "Test_Fit", "tests"),
    ("biren_fit", "pkg/device/biren/device_test.go", "TestDevices_Fit", "tests"),
    ("mthreads_fit", "pkg/device/mthreads/device_test.go", "TestDevices_Fit", "tests"),
    ("vastai_fit", "pkg/device/vastai/device_test.go", "TestDevices_Fit", "tests"),
    ("gcu_fit", "pkg/device/enflame/gcu_test.go", "TestGCUDevices_Fit", "tests"),
    ("awsneuron_generate_requests", "pkg/device/awsneuron/device_test.go", "Test_GenerateResourceRequests", "tests"),
    ("metax_fit", "pkg/device/metax/device_test.go", "TestMetaxDevices_Fit", "tests"),
    ("metax_score_node", "pkg/device/metax/device_test.go", "Test_ScoreNode", "tests"),
    ("metax_parse_annos", "pkg/device/metax/device_test.go", "TestParseMetaxAnnos", "tests"),
    ("metax_generate_requests", "pkg/device/metax/device_test.go", "Test_GenerateResourceRequests", "tests"),
    ("metax_get_node_devices", "pkg/device/metax/device_test.go", "TestGetNodeDevices", "tests"),
    # kunlun whole XPUs: graghSelect over FitXPU, ScoreNode by wings
    ("kunlun_graph_select", "pkg/device/kunlun/device_test.go", "Test_graphSelect", "tests"),
    ("kunlun_score_node", "pkg/device/kunlun/device_test.go", "Test_ScoreNode", "tests"),
    # kunlun shared XPUs ("XPU"): the same graghSelect with FitVXPU, and the request extraction with trimMemory
    ("kunlun_graph_select_vxpu", "pkg/device/kunlun/device_test.go", "Test_graphSelectVXPU", "tests"),
    ("kunlun_vxpu_generate_requests", "pkg/device/kunlun/vdevice_test.go", "Test_KunlunVDevices_GenerateResourceRequests", "tests"),
    ("kunlun_fit_vxpu_direct", "pkg/device/kunlun/vdevice_test.go", "Test_FitVXPU_direct", "tests"),
    # metax sGPU ("Metax-SGPU"): filter-then-prioritise, application classes, QoS, MetaLink scores
    ("metax_sgpu_fit", "pkg/device/metax/sdevice_test.go", "TestMetaxSDevices_Fit", "tests"),
    ("metax_sgpu_generate_requests", "pkg/device/metax/sdevice_test.go", "TestGenerateResourceRequests", None),
    ("metax_sgpu_check_qos", "pkg/device/metax/sdevice_test.go", "TestCheckDeviceQos", None),
    ("metax_sgpu_prioritize_exclusive", "pkg/device/metax/sdevice_test.go", "TestPrioritizeExclusiveDevices", None),
    ("metax_sgpu_score_exclusive", "pkg/device/metax/sdevice_test.go", "TestScoreExclusiveDevices", None),
    ("metax_sgpu_check_app_class", "pkg/device/metax/sdevice_test.go", "TestCheckAppClass", None),
    ("metax_sgpu_prioritize_online", "pkg/device/metax/sdevice_test.go", "TestPrioritizeOnlineDevices", None),
    ("metax_sgpu_score_online", "pkg/device/metax/sdevice_test.go", "TestScoreOnlineDevices", None),
]


SOURCE_PATCHES = {
    ("pkg/device/nvidia/device_test.go", "TestGenerateResourceRequests"): [("int64(1)<<32+50", "4294967346")],
    ("pkg/device/awsneuron/device_test.go", "Test_GenerateResourceRequests"): [("int64(math.MaxInt32)+1", "2147483648"),
                                                                               ("int64(math.MaxInt32)*2+2", "4294967296")],
}


def main():
    ref = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
    out = sys.argv[2] if len(sys.argv) > 2 else os.path.join(os.path.dirname(__file__), "..", "tests", "golden")
    os.makedirs(out, exist_ok=True)
    for name, rel, func, var in TABLES:
        path = os.path.join(ref, rel)
        try:
            src = open(path).read()
            body = find_func(src, func)
            line = src[:src.index(body)].count("\n") + 1
            for old, new in SOURCE_PATCHES.get((rel, func), []):   # Go expressions the literal parser does not evaluate
                body = body.replace(old, new)
            table = extract_table(body, var)
        except (KeyError, SyntaxError, FileNotFoundError) as e:
            print("SKIP %-28s %s:%s  (%s: %s)" % (name, rel, func, type(e).__name__, e))
            continue
        doc = {"source": "%s:%d (%s)" % (rel, line, func),
               "note": "data extracted from the reference's own test table by tools/extract_golden.py",
               "cases": table}
        with open(os.path.join(out, name + ".json"), "w") as f:
            json.dump(doc, f, indent=1, sort_keys=True)
        print("OK   %-28s %3d cases  %s:%d" % (name, len(table), rel, line))


if __name__ == "__main__":
    main()
