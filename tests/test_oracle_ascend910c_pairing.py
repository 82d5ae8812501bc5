"""pkg/device/ascend/device_910c_pairing_test.go through the oracle's Fit hook: an Ascend910C request is served in whole
modules (the two NPUs whose indices are 2k and 2k+1), with or without NetworkIDs."""
import pytest

import oracle_abi as oa
from test_oracle_golden import run_fit

W = "Ascend910C"


@pytest.fixture(scope="module")
def orc():
    o = oa.Oracle()
    yield o
    o.close()


def fit(orc, indices, nums, net=0):
    devices = [{"ID": "dev-%d" % i, "Index": i, "Count": 1, "Used": 0, "Totalmem": 32000, "Health": True, "Type": W} for i in indices]
    if net is not None:
        for d in devices:
            d["CustomInfo"] = {"NetworkID": net}
    node = {"name": "node", "devices": [{"id": "dev-%d" % i, "index": i, "vendor": W, "type": W, "network_id": -1 if net is None else net}
                                        for i in indices]}
    ok, got, reason = run_fit(orc, W, devices, {"Nums": nums, "Type": W}, None, node)
    return ok, sorted(g[0] for g in got), reason


def test_partial_modules_cannot_fill_the_request(orc):
    # :31-65 TestAscend910C_FitPartialAllocationBug: dev-0/1 form a module, dev-2, dev-4, dev-6 are halves; 4 asked
    ok, got, reason = fit(orc, [0, 1, 2, 4, 6], 4)
    assert not ok and "AllocatedCardsInsufficientRequest" in reason


def test_exact_count_of_two_halves_is_refused(orc):
    # :67-95 TestAscend910C_FitExactCountBypassBug
    ok, got, reason = fit(orc, [0, 2], 2)
    assert not ok and "AllocatedCardsInsufficientRequest" in reason


def test_one_full_module_fits(orc):
    # :97-122 TestAscend910C_FitFullPairSucceeds
    ok, got, reason = fit(orc, [0, 1], 2)
    assert ok and got == ["dev-0", "dev-1"], reason


def test_pairing_is_validated_without_network_ids(orc):
    # :124-163 TestAscend910C_FitWithoutNetworkID_ValidatesPairing
    ok, got, reason = fit(orc, [0, 2], 2, net=None)
    assert not ok
    ok, got, reason = fit(orc, [0, 1], 2, net=None)
    assert ok and got == ["dev-0", "dev-1"], reason


def test_four_halves_give_no_combination(orc):
    # :165-208 TestComputeBestCombination910C_NoFullPairsReturnsEmpty
    ok, got, reason = fit(orc, [0, 2, 4, 6], 4, net=None)
    assert not ok
