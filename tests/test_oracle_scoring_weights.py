"""hami.io/device-scoring-weights at the calcScore level (pkg/scheduler/score_test.go:4048-4209), replayed through the oracle's
Filter: the weights reach ComputeScore for app AND init containers, an invalid annotation fails the whole Filter, and under the
topology-aware policy the utilisation score only breaks ties between equal pair scores."""
import pytest

import oracle_abi as oa
from oracle_abi import abi

NV = "NVIDIA"
GPU_POLICY, WEIGHTS = "hami.io/gpu-scheduler-policy", "hami.io/device-scoring-weights"


def cluster(pair_scores=None):
    # newDeviceScoringWeightTestNodes (score_test.go:4057-4080): gpu-a {used 1, cores 90, mem 10}, gpu-b {used 7, cores 10, mem 20}
    o = oa.Oracle("binpack", "binpack")
    node = {"name": "node1", "devices": [
        {"id": "gpu-a", "index": 0, "count": 10, "devmem": 100, "devcore": 100, "type": NV, "numa": 0, "mode": "hami-core", "health": True, "vendor": NV},
        {"id": "gpu-b", "index": 1, "count": 10, "devmem": 100, "devcore": 100, "type": NV, "numa": 0, "mode": "hami-core", "health": True, "vendor": NV}]}
    if pair_scores is not None:
        node["pair_scores"] = pair_scores
    o.node_upsert(node)
    o.usage_apply("held-a", "node1", "gpu-a", 1, 10, 90)
    o.usage_apply("held-b", "node1", "gpu-b", 7, 20, 10)
    return o


def pod(annos, init=False):
    req = [{"key": NV, "type": NV, "nums": 1, "memreq": 0, "mempct": 40, "coresreq": 0}]
    return {"uid": "w", "ns": "default", "name": "w", "annotations": annos, "containers": [req, []] if init else [req],
            "n_init": 1 if init else 0}


def picked(r):
    assert r["status"] == abi.HS_FIT, r
    return [[d[0] for d in c] for c in r["devices"][NV]]


def test_invalid_weights_fail_the_filter():
    # TestCalcScoreRejectsInvalidDeviceScoringWeights (:4048-4056)
    o = cluster()
    r = o.filter(pod({WEIGHTS: "slot=1,core=-1,memory=3"}))
    assert r["status"] < 0 and '"core" weight must not be negative' in o.L.oracle_last_error(o.h).decode()
    o.close()


@pytest.mark.parametrize("init", [False, True], ids=["app", "init"])
def test_pod_weights_reach_compute_score(init):
    # TestCalcScoreUsesPodDeviceScoringWeights (:4083-4110) and …ForInitContainers (:4112-4149): default weights rank gpu-a
    # (16) above gpu-b (15), memory-heavy weights flip it (26 against 27); the init container's pick follows them too
    o = cluster()
    got = picked(o.filter(pod({GPU_POLICY: "binpack"}, init)))
    assert got == ([["gpu-a"], []] if init else [["gpu-a"]])
    o.close()
    o = cluster()
    got = picked(o.filter(pod({GPU_POLICY: "binpack", WEIGHTS: "slot=1,core=1,memory=3"}, init)))
    assert got == ([["gpu-b"], []] if init else [["gpu-b"]])
    o.close()


@pytest.mark.parametrize("weights,want", [("", "gpu-b"), ("slot=1,core=1,memory=3", "gpu-a")])
def test_weights_break_topology_ties_only(weights, want):
    # TestCalcScoreUsesDeviceScoringWeightsAsTopologyTieBreaker (:4151-4209): equal pair scores (1 both ways), one card asked
    o = cluster(pair_scores=[[0, 1], [1, 0]])
    annos = {GPU_POLICY: "topology-aware"}
    if weights:
        annos[WEIGHTS] = weights
    assert picked(o.filter(pod(annos))) == [[want]]
    o.close()


BAD = [("slot=1,core=-1,memory=3", '"core" weight must not be negative'), ("slot=1,core=1", "expected slot, core, and memory weights"),
       ("slot=1,core,memory=3", "expected key=value entries"), ("slot=1,slot=2,memory=3", 'duplicate "slot" weight'),
       ("slot=1,core=x,memory=3", '"core" weight must be an integer'), ("slot=1,core=1,mem=3", 'unknown weight "mem"'),
       ("slot=0,core=0,memory=0", "at least one weight must be positive"), ('slot=1,co"re=1,memory=3', 'unknown weight "co\\"re"')]


@pytest.mark.parametrize("value,detail", BAD, ids=[b[1][:24] for b in BAD])
def test_weight_errors_carry_the_reference_text(value, detail):
    # util/device_scoring_weights.go:58-102: fmt.Errorf("invalid %s annotation %q: …", key, value[, key])
    o = cluster()
    r = o.filter(pod({WEIGHTS: value}))
    quoted = '"' + value.replace("\\", "\\\\").replace('"', '\\"') + '"'
    assert r["status"] == abi.HS_ERR_WEIGHTS
    assert o.L.oracle_last_error(o.h).decode() == "invalid hami.io/device-scoring-weights annotation %s: %s" % (quoted, detail)
    o.close()


@pytest.mark.gpu
def test_engine_reports_the_same_weight_errors():
    import k8s_vgpu_scheduler_b200.engine as eng
    o = cluster()
    e = eng.Engine("binpack", "binpack")
    e.node_upsert({"name": "node1", "devices": [
        {"id": "gpu-a", "index": 0, "count": 10, "devmem": 100, "devcore": 100, "type": NV, "numa": 0, "mode": "hami-core", "health": True, "vendor": NV}]})
    for value, _ in BAD:
        p = pod({WEIGHTS: value})
        assert o.filter(p)["status"] == abi.HS_ERR_WEIGHTS
        assert e.filter(p)["status"] == abi.HS_ERR_WEIGHTS
        assert e.error() == o.L.oracle_last_error(o.h).decode(), value
    e.close()
    o.close()
