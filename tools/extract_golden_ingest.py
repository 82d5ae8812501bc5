#!/usr/bin/env python3
"""Golden vectors for the ingest codecs that the reference tests as stand-alone functions (not tables):
pkg/device/initContainer_test.go (CollapseInitContainerUsage / AppContainersOnlyDeviceUsage) and the
GetNodeDevices pair-score tests of pkg/device/nvidia/device_test.go.  Writes tests/golden/ingest_funcs.json.

    python tools/extract_golden_ingest.py [/root/reference]
"""
import json
import os
import re
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import extract_golden as X  # noqa: E402


def init_container_cases(ref):
    rel = "pkg/device/initContainer_test.go"
    src = open(os.path.join(ref, rel)).read()
    out = []
    for m in re.finditer(r"^func (Test(CollapseInitContainerUsage|AppContainersOnlyDeviceUsage)_\w+)\(", src, re.M):
        name = m.group(1)
        body = X.find_func(src, name)
        mp = re.search(r'makePod\("\w+",\s*(\d+),\s*(\d+)\)', body)
        case = {"name": name, "func": m.group(2), "source": "%s:%d" % (rel, src[:m.start()].count("\n") + 1)}
        if mp:
            case["n_init"], case["n_app"] = int(mp.group(1)), int(mp.group(2))
        for var in ("raw", "expected"):
            mv = re.search(r"\b%s\s*:=\s*" % var, body)
            if mv:
                case[var] = X.simplify(X.P(X.tokenize(body[mv.end():])).parse_expr())
        # the two *_MultiAppSameGPUSlots tests assert one field instead of a whole value
        ms = re.search(r"assert\.Equal\(t, result\[\"(\w+)\"\]\[0\]\[0\]\.Slots, int32\((\d+)\)\)", body)
        if ms:
            case["expect_slots"] = {"key": ms.group(1), "slots": int(ms.group(2))}
        if "raw" in case:
            out.append(case)
    return out


def pair_score_cases(ref):
    rel = "pkg/device/nvidia/device_test.go"
    src = open(os.path.join(ref, rel)).read()
    out = []
    for name in ("TestGetNodeDevices_InvalidJSON", "TestGetNodeDevices_MigProfilesFromNode", "TestGetNodeDevices_PairScores",
                 "TestGetNodeDevices_InvalidPairScores"):
        body = X.find_func(src, name)
        case = {"name": name, "source": "%s:%d" % (rel, src[:src.index(body)].count("\n") + 1)}
        for key, field in (("RegisterAnnos", "register"), ("RegisterGPUPairScore", "pair_scores")):
            mv = re.search(r"%s:\s*(`[^`]*`|\"(?:[^\"\\]|\\.)*\")" % key, body)
            if mv:
                lit = mv.group(1)
                case[field] = lit[1:-1] if lit[0] == "`" else X.unquote(lit)
        case["want_err"] = "err != nil" in body
        me = re.search(r"assert\.Equal\(t, len\(result\), (\d+)\)", body)
        if me:
            case["want_len"] = int(me.group(1))
        ms = re.search(r"result\[(\d+)\]\.DevicePairScore\.Scores\[\"([^\"]+)\"\], (\d+)\)", body)
        if ms:
            case["want_score"] = {"device": int(ms.group(1)), "peer": ms.group(2), "score": int(ms.group(3))}
        out.append(case)
    return out


def amd_fit_cases(ref):
    """pkg/device/amd/device_test.go TestDevices_Fit is a list of t.Run blocks, not a table."""
    rel = "pkg/device/amd/device_test.go"
    src = open(os.path.join(ref, rel)).read()
    body = X.find_func(src, "TestDevices_Fit")
    base = src[:src.index(body)].count("\n") + 1
    X.IDENTS.update({"AMDDevice": "AMD", "AMDUseUUID": "amd.com/use-gpu-uuid", "AMDNoUseUUID": "amd.com/nouse-gpu-uuid",
                     "AMDInUse": "amd.com/use-gputype", "AMDNoUse": "amd.com/nouse-gputype"})
    out = []
    starts = [m.start() for m in re.finditer(r'\bt\.Run\("', body)] + [len(body)]
    for a, b in zip(starts, starts[1:]):
        blk = body[a:b]
        name = re.match(r't\.Run\("((?:[^"\\]|\\.)*)"', blk).group(1)
        mp = re.search(r'productType :?= "([^"]*)"', blk)
        if mp:
            X.IDENTS["productType"] = mp.group(1)

        def lit(pattern):
            m = re.search(pattern, blk)
            return X.simplify(X.P(X.tokenize(blk[m.end():])).parse_expr()) if m else None
        case = {"name": name, "source": "%s:%d" % (rel, base + body[:a].count("\n")),
                "devices": lit(r"devices := "), "request": lit(r"req := "), "annos": lit(r"Annotations:\s*") or {}}
        case["wantFit"] = re.search(r"assert\.Equal\(t, (true|false), ok\)", blk).group(1) == "true"
        case["wantUUIDs"] = [u for _, u in sorted((int(i), u) for u, i in re.findall(r'assert\.Equal\(t, "([^"]+)", got\[AMDDevice\]\[(\d+)\]\.UUID\)', blk))]
        for field in ("Usedmem", "Usedcores"):
            m = re.search(r"assert\.Equal\(t, int32\((\d+)\), got\[AMDDevice\]\[0\]\.%s\)" % field, blk)
            if m:
                case["want" + field] = int(m.group(1))
        if re.search(r"assert\.Equal\(t, productType, got\[AMDDevice\]\[0\]\.Type\)", blk):
            case["wantType"] = X.IDENTS["productType"]
        m = re.search(r"strings\.Contains\(reason, common\.(\w+)\)", blk)
        if m:
            case["wantReasonContains"] = X.IDENTS["common." + m.group(1)]
        elif re.search(r'assert\.Equal\(t, "", reason\)', blk):
            case["wantReason"] = ""
        m = re.search(r"assert\.Equal\(t, (\d+), len\(got\[AMDDevice\]\)\)", blk)
        if m:
            case["wantLen"] = int(m.group(1))
        out.append(case)
    return out


def amd_request_cases(ref):
    """pkg/device/amd/device_test.go Test_GenerateResourceRequests: three t.Run blocks and one loop table."""
    rel = "pkg/device/amd/device_test.go"
    src = open(os.path.join(ref, rel)).read()
    body = X.find_func(src, "Test_GenerateResourceRequests")
    base = src[:src.index(body)].count("\n") + 1
    num = lambda s: int(eval(s.replace("math.MaxInt32", "2147483647"), {"__builtins__": {}}))   # noqa: E731
    qty = r'"([^"]+)":\s*\*resource\.NewQuantity\(([^,]+), resource\.DecimalSI\)'
    out = []
    loop_at = body.index("for _, tc := range")
    starts = [m.start() for m in re.finditer(r'\bt\.Run\("', body[:loop_at])] + [loop_at]
    for a, b in zip(starts, starts[1:]):
        blk = body[a:b]
        name = re.match(r't\.Run\("((?:[^"\\]|\\.)*)"', blk).group(1)
        lim_at, req_at = blk.find("Limits:"), blk.find("Requests:")
        lim_txt = blk[lim_at:req_at] if req_at > lim_at >= 0 else blk[lim_at:]
        req_txt = blk[req_at:] if req_at >= 0 else ""
        case = {"name": name, "source": "%s:%d" % (rel, base + body[:a].count("\n")),
                "limits": {k: num(v) for k, v in re.findall(qty, lim_txt)}, "requests": {k: num(v) for k, v in re.findall(qty, req_txt)}}
        full = re.search(r"Nums:\s*(\d+),\s*Type:\s*AMDDevice,\s*Memreq:\s*(\d+),\s*MemPercentagereq:\s*(\d+),\s*Coresreq:\s*(\d+)", blk)
        part = re.search(r"assert\.Equal\(t, int32\((\d+)\), got\.Coresreq\)", blk)
        if full:
            case["want"] = {"Nums": int(full.group(1)), "Type": "AMD", "Memreq": int(full.group(2)),
                            "MemPercentagereq": int(full.group(3)), "Coresreq": int(full.group(4))}
        elif part:
            case["want_partial"] = {"Coresreq": int(part.group(1))}
        else:
            assert "device.ContainerDeviceRequest{}, got" in blk, name
            case["want"] = None
        out.append(case)
    loop = body[loop_at:]
    base_limits = {k: num(v) for k, v in re.findall(qty, loop)}
    for m in re.finditer(r'\{name:\s*"([^"]+)",\s*resource:\s*"([^"]+)",\s*value:\s*([^}]+)\}', loop):
        out.append({"name": m.group(1), "source": "%s:%d" % (rel, base + body[:loop_at + m.start()].count("\n")),
                    "limits": dict(base_limits, **{m.group(2): num(m.group(3))}), "requests": {}, "want": None})
    return out


def neuron_cases(ref):
    """pkg/device/awsneuron/device_test.go: TestDevices_Fit builds its devices with makeAWSDeviceUsage(id, index, used, count,
    totalcore, usedcores, nodeType, health) and Test_graphSelect with makeDeviceUsages(nodeType, {index: used}, health) (16
    devices, the node type on device 0 only) — both helper calls are expanded here."""
    rel = "pkg/device/awsneuron/device_test.go"
    src = open(os.path.join(ref, rel)).read()
    out = {"fit": [], "graph_select": []}
    body = X.find_func(src, "TestDevices_Fit")
    base = src[:src.index(body)].count("\n") + 1
    starts = [m.start() for m in re.finditer(r'\n\t\t\{\n\t\t\tname:', body)] + [body.index("for _, test := range tests")]
    for a, b in zip(starts, starts[1:]):
        blk = body[a:b]
        devs = [{"ID": m.group(1), "Index": int(m.group(2)), "Used": int(m.group(3)), "Count": int(m.group(4)), "Totalcore": int(m.group(5)),
                 "Usedcores": int(m.group(6)), "Type": "AWSNeuron", "Health": m.group(8) == "true", "CustomInfo": {"AWSNodeType": m.group(7)}}
                for m in re.finditer(r'makeAWSDeviceUsage\("([^"]+)", (\d+), (\d+), (\d+), (\d+), (\d+), "([^"]*)", (true|false)\)', blk)]
        req = {k: int(v) for k, v in re.findall(r'\b(Nums|Memreq|MemPercentagereq|Coresreq):\s*(\d+)', blk)}
        t = re.search(r'Type:\s*(AWSNeuronDevice|"[^"]*")', blk).group(1)
        req["Type"] = "AWSNeuron" if t == "AWSNeuronDevice" else t.strip('"')
        annos = dict(re.findall(r'"([^"]+)":\s*"([^"]*)"', re.search(r'annos:\s*map\[string\]string\{([^}]*)\}', blk).group(1)))
        out["fit"].append({"name": re.search(r'name:\s*"([^"]+)"', blk).group(1), "source": "%s:%d" % (rel, base + body[:a].count("\n") + 1),
                           "devices": devs, "request": req, "annos": annos,
                           "wantFit": re.search(r'wantFit:\s*(true|false)', blk).group(1) == "true",
                           "wantLen": int(re.search(r'wantLen:\s*(\d+)', blk).group(1)),
                           "wantDevIDs": re.findall(r'"([^"]+)"', re.search(r'wantDevIDs:\s*\[\]string\{([^}]*)\}', blk).group(1)),
                           "wantReason": re.search(r'wantReason:\s*"([^"]*)"', blk).group(1)})
    body = X.find_func(src, "Test_graphSelect")
    base = src[:src.index(body)].count("\n") + 1
    for m in re.finditer(r'name:\s*"([^"]+)".*?makeDeviceUsages\("([^"]*)", (nil|map\[int\]int32\{[^}]*\}), (true|false)\),\s*c:\s*(\d+),.*?want1:\s*\[\]int\{([^}]*)\}', body, re.S):
        used = {} if m.group(3) == "nil" else {int(k): int(v) for k, v in re.findall(r'(\d+):\s*(\d+)', m.group(3))}
        out["graph_select"].append({"name": m.group(1), "source": "%s:%d" % (rel, base + body[:m.start()].count("\n")), "node_type": m.group(2),
                                    "used": {str(k): v for k, v in used.items()}, "health": m.group(4) == "true", "count": int(m.group(5)),
                                    "want": [int(x) for x in re.findall(r'\d+', m.group(6))]})
    return out


def capacity_node_cases(ref):
    """cambricon / mthreads Test_GetNodeDevices: the node's capacity in, the derived DeviceInfo list out."""
    out = []
    for vendor, rel, fn, cores_res, mem_res in (
            ("MLU", "pkg/device/cambricon/device_test.go", "Test_GetNodeDevices", "cambricon.com/mlu.smlu.vcore", "cambricon.com/mlu.smlu.vmemory"),
            ("Mthreads", "pkg/device/mthreads/device_test.go", "Test_GetNodeDevices", "mthreads.com/sgpu-core", "mthreads.com/sgpu-memory")):
        src = open(os.path.join(ref, rel)).read()
        body = X.find_func(src, fn)
        base = src[:src.index(body)].count("\n") + 1
        for m in re.finditer(r'name:\s*"((?:[^"\\]|\\.)*)"', body):
            blk = body[m.start():]
            nxt = re.search(r'\n\t\t\},\n', blk)
            blk = blk[:nxt.end()] if nxt else blk

            def qty(res):
                q = re.search(r'"%s":\s*\*resource\.NewQuantity\((-?\d+), resource\.DecimalSI\)' % re.escape(res), blk)
                return int(q.group(1)) if q else 0
            want = []
            for w in re.finditer(r'Index:\s*(?:uint\()?(\d+)\)?,\s*ID:\s*"([^"]+)",\s*Count:\s*int32\((\d+)\),\s*Devmem:\s*int32\((\d+)\),\s*Devcore:\s*int32\((\d+)\)', blk):
                want.append({"Index": int(w.group(1)), "ID": w.group(2), "Count": int(w.group(3)), "Devmem": int(w.group(4)),
                             "Devcore": int(w.group(5)), "Type": vendor, "Numa": 0, "Health": True, "DeviceVendor": vendor})
            out.append({"name": m.group(1), "source": "%s:%d" % (rel, base + body[:m.start()].count("\n")), "vendor": vendor,
                        "node": re.search(r'Name:\s*"([^"]*)"', blk).group(1), "cores": qty(cores_res), "memory": qty(mem_res), "want": want})
    return out


def main():
    ref = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
    doc = {"note": "extracted from the reference's own tests by tools/extract_golden_ingest.py",
           "init_container": init_container_cases(ref), "nvidia_get_node_devices": pair_score_cases(ref),
           "amd_fit": amd_fit_cases(ref), "capacity_node_devices": capacity_node_cases(ref),
           "amd_generate_requests": amd_request_cases(ref), "awsneuron": neuron_cases(ref)}
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden", "ingest_funcs.json")
    with open(path, "w") as f:
        json.dump(doc, f, indent=1, sort_keys=True)
    print("OK ingest_funcs: %d init-container cases, %d GetNodeDevices cases" % (len(doc["init_container"]), len(doc["nvidia_get_node_devices"])))


if __name__ == "__main__":
    main()
