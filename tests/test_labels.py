"""Label parsing (filter.go:60-86): host Python, the C ABI helper and the oracle agree on the KATs."""
from oracle import oracle as ora
from oracle import pyref
from yoda_scheduler_b200 import labels
from yoda_scheduler_b200.api import parse_label
from yoda_scheduler_b200 import _lib


def test_parser_kat(kat):
    for s, want in kat["parser"]:
        assert labels.str_to_uint64(s) == want, s
        assert parse_label(s) == want, s
        assert ora.str_to_uint64(s) == want, s
        assert pyref.str_to_u64(s) == want, s


def test_uint64_to_int64_kat(kat):
    L = _lib.load()
    for u, want in kat["uint64_to_int64"]:
        assert labels.uint64_to_int64(u) == want
        assert L.yoda_uint64_to_int64(u) == want
        assert ora.uint64_to_int64(u) == want
        assert pyref.u64_to_i64(u) == want


def test_atoi_error_kinds():
    assert labels.go_atoi("12") == (12, None)
    assert labels.go_atoi("x")[1] == "syntax"
    assert labels.go_atoi("9223372036854775808") == (2**63 - 1, "range")
    assert labels.go_atoi("-9223372036854775809") == (-2**63, "range")


def test_pod_priority():
    # sort.go:12-18 keeps Atoi's value and drops the error
    assert labels.pod_priority({}) == 0
    assert labels.pod_priority({"scv/priority": "7"}) == 7
    assert labels.pod_priority({"scv/priority": "-3"}) == -3
    assert labels.pod_priority({"scv/priority": "zz"}) == 0
    assert labels.pod_priority({"scv/priority": "99999999999999999999"}) == 2**63 - 1


def test_requests_from_labels():
    r = labels.pods_to_requests([{"scv/memory": "1000"}, {"scv/number": "-1", "scv/clock": "x"}, {}])
    assert r["flags"].tolist() == [2, 5, 0]
    assert r["memory"].tolist() == [1000, 0, 0]
    assert r["number"].tolist() == [0, 2**64 - 1, 0]
    assert r["clock"].tolist() == [0, 0, 0]
