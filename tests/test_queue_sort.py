"""QueueSort (pkg/yoda/sort/sort.go): scv/priority, higher first; SURVEY.md 8(f) row 3."""
import random

import numpy as np
import pytest

from oracle import oracle as ora
from yoda_scheduler_b200 import labels


CASES = [None, "0", "7", "-3", "zz", "", "+5", "99999999999999999999", "-99999999999999999999", "9223372036854775807", "007", "1_0"]


def test_priority_parsing_matches_oracle():
    for lab in CASES:
        want = ora.pod_priority(lab)
        got = labels.pod_priority({} if lab is None else {"scv/priority": lab})
        assert got == want, lab
    assert ora.pod_priority("99999999999999999999") == 2**63 - 1        # Atoi's clamped value is kept (sort.go:14)
    assert ora.pod_priority("zz") == 0 and ora.pod_priority(None) == 0


def test_oracle_order_is_stable_and_descending():
    labs = ["1", None, "5", "5", "-2", "zz", "5", "9"]
    assert ora.queue_order(labs).tolist() == [7, 2, 3, 6, 0, 1, 5, 4]


@pytest.mark.gpu
def test_device_queue_order_matches_oracle():
    from yoda_scheduler_b200 import api
    rng = random.Random(11)
    with api.Context(4) as ctx:
        for n in (1, 2, 33, 1000, 20000):
            labs = [rng.choice(CASES + [str(rng.randrange(-50, 50)) for _ in range(4)]) for _ in range(n)]
            pri = [labels.pod_priority({} if l is None else {"scv/priority": l}) for l in labs]
            got = ctx.queue_order(pri)
            if n <= 1000:
                assert np.array_equal(got, ora.queue_order(labs))
            p = np.array(pri, dtype=np.int64)
            assert sorted(got.tolist()) == list(range(n))
            assert (np.diff(p[got]) <= 0).all()                             # descending priority
            same = np.diff(p[got]) == 0
            assert (np.diff(got.astype(np.int64))[same] > 0).all()          # arrival order among equals
