"""CPU tier: the random-agent protocol's statistics (deciles.py:4-5, KS test) and env sharding."""
import numpy as np
import pytest

from cartpoleplusplus_b200 import protocol
from cartpoleplusplus_b200.parallel import shard


def test_deciles_from_histogram_equals_deciles_py():
    rng = np.random.default_rng(0)
    for n in (1, 2, 7, 100, 1001):
        lengths = rng.integers(1, 201, size=n)
        hist = np.bincount(lengths, minlength=201)
        want = np.percentile(lengths, np.linspace(0, 100, 11))  # deciles.py:4-5
        assert np.allclose(protocol.deciles_from_histogram(hist), want, rtol=0, atol=1e-12)
        assert np.array_equal(protocol.lengths_from_histogram(hist), np.sort(lengths))
    assert np.array_equal(protocol.deciles_from_histogram(np.bincount([200] * 100, minlength=201)), [200] * 11)
    with pytest.raises(ValueError):
        protocol.deciles_from_histogram(np.zeros(201, dtype=np.int64))


def test_ks_matches_scipy():
    from scipy import stats
    rng = np.random.default_rng(1)
    a = rng.geometric(0.08, size=5000).clip(1, 200)
    for shift, same in ((0.08, True), (0.10, False)):
        b = rng.geometric(shift, size=7000).clip(1, 200)
        d, p = protocol.ks_two_sample(np.bincount(a, minlength=201), np.bincount(b, minlength=201))
        ref = stats.ks_2samp(a, b, method="asymp")
        assert abs(d - ref.statistic) < 1e-12
        assert (p > 0.05) == (ref.pvalue > 0.05) == same


def test_readme_windows_present():
    assert protocol.README_DECILES[(0.0, (0,))] == [200] * 11
    assert len(protocol.README_DECILES) == 4 and all(len(v) == 11 for v in protocol.README_DECILES.values())


def test_shard_partitions_the_envs():
    for total, world in ((1 << 20, 8), (65536, 4), (10, 4), (7, 8)):
        spans = [shard(total, r, world) for r in range(world)]
        assert spans[0][0] == 0 and sum(n for _, n in spans) == total
        for (o0, n0), (o1, _) in zip(spans, spans[1:]):
            assert o0 + n0 == o1
    with pytest.raises(ValueError):
        shard(8, 2, 2)
