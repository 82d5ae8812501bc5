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


def main():
    ref = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
    doc = {"note": "extracted from the reference's own tests by tools/extract_golden_ingest.py",
           "init_container": init_container_cases(ref), "nvidia_get_node_devices": pair_score_cases(ref)}
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden", "ingest_funcs.json")
    with open(path, "w") as f:
        json.dump(doc, f, indent=1, sort_keys=True)
    print("OK ingest_funcs: %d init-container cases, %d GetNodeDevices cases" % (len(doc["init_container"]), len(doc["nvidia_get_node_devices"])))


if __name__ == "__main__":
    main()
