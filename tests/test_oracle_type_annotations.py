"""The use / nouse TYPE annotations of amd and hygon (amd/device_test.go:479-511 TestCheckAMDType, hygon/device_test.go:85-200
Test_checkDCUtype) through the oracle's Fit hook: one idle card of the given type, one device asked."""
import pytest

import oracle_abi as oa
from test_oracle_golden import run_fit


@pytest.fixture(scope="module")
def orc():
    o = oa.Oracle()
    yield o
    o.close()


USE, NOUSE = "amd.com/use-gputype", "amd.com/nouse-gputype"
AMD_CASES = [
    ({}, "MI300X", True), ({USE: "MI300"}, "MI300X", True), ({USE: "MI250"}, "MI300X", False), ({NOUSE: "MI300"}, "MI300X", False),
    ({NOUSE: "MI250"}, "MI300X", True), ({NOUSE: ""}, "MI300X", True), ({USE: ""}, "MI300X", True), ({NOUSE: "   "}, "MI300X", True),
    ({USE: "   "}, "MI300X", True), ({NOUSE: "MI250,"}, "MI300X", True), ({NOUSE: ",MI250"}, "MI300X", True),
    ({NOUSE: " MI250 , "}, "MI250X", False), ({NOUSE: ", "}, "MI300X", True), ({USE: "MI300,"}, "MI250X", False), ({USE: ", "}, "MI300X", False),
]


@pytest.mark.parametrize("annos,card,want", AMD_CASES, ids=["%s=%r on %s" % ((list(a) or ["-"])[0].split("/")[-1], (list(a.values()) or [""])[0], c)
                                                              for a, c, _ in AMD_CASES])
def test_check_amd_type(orc, annos, card, want):
    devs = [{"ID": "gpu-0", "Count": 1, "Totalmem": 65536, "Totalcore": 100, "Type": "AMD-" + card, "Health": True}]
    fit, _, reason = run_fit(orc, "AMD", devs, {"Nums": 1, "Memreq": 1024, "Coresreq": 10, "Type": "AMD"}, annos)
    assert fit == want, (annos, card, reason)


@pytest.mark.parametrize("annos,card,want", [
    # the table's shape with model names in place of its "dcu" / "test" (checkDCUtype is called on the bare card type there; through
    # Fit the device type must also contain the request's "DCU")
    ({"hygon.com/use-dcutype": "Z100"}, "Z100", True), ({"hygon.com/use-dcutype": "Z100"}, "K100", False), ({}, "Z100", True),
    ({"hygon.com/use-dcutype": "Z100,K100"}, "Z100", True), ({"hygon.com/nouse-dcutype": "Z100"}, "Z100", False),
    ({"hygon.com/nouse-dcutype": "K100,Z100"}, "Z100", False), ({"hygon.com/nouse-dcutype": "Z100"}, "K100", True)])
def test_check_dcu_type(orc, annos, card, want):
    devs = [{"ID": "dcu-0", "Count": 4, "Totalmem": 65536, "Totalcore": 100, "Type": "DCU-" + card, "Health": True}]
    fit, _, reason = run_fit(orc, "DCU", devs, {"Nums": 1, "Memreq": 1024, "Coresreq": 10, "Type": "DCU"}, annos)
    assert fit == want, (annos, card, reason)
