"""When does the split of the replicated writes of the previous take stand (partitioner.previous_assignment_stands)?
Decided from gathered throughputs alone; the rule is: unjudged while nothing moved by more than 3 %, otherwise
given up only if a fresh split would end more than 4 % earlier."""
import pytest

import tsb200_loader

tsb200_loader.load()
from torchsnapshot_b200 import partitioner as P  # noqa: E402


def _negotiation(speeds, can_repeat=True, digest="d", nonrep=0):
    return P.Negotiation([(nonrep, digest, s, "paths", can_repeat) for s in speeds], rank=0)


def _split(loads, speeds):
    """What partition_write_reqs does after the exchange: assign, remember."""
    neg = _negotiation(speeds)
    out = P.assign_write_loads(loads, [0] * len(speeds), speeds)
    P._remember_assignment(neg, out, {(p, i): sz for p, i, sz, _sub in loads}, loads)
    return out


@pytest.fixture(autouse=True)
def _clean():
    saved = P._last_assignment
    P._last_assignment = None
    yield
    P._last_assignment = saved


def _fine_loads(n=400):
    return [(f"p{i}", 0, 1_000_000 + 997 * (i % 13), False) for i in range(n)]


def test_nothing_remembered_or_not_everybody_can_repeat():
    assert not P.previous_assignment_stands(_negotiation([40.0, 40.0]))
    _split(_fine_loads(), [40.0, 40.0])
    assert P.previous_assignment_stands(_negotiation([40.0, 40.0]))
    assert not P.previous_assignment_stands(_negotiation([40.0, 40.0], can_repeat=False))
    assert not P.previous_assignment_stands(_negotiation([40.0, 40.0], digest="other loads"))
    assert not P.previous_assignment_stands(_negotiation([40.0, 40.0, 40.0]))  # another world


def test_small_moves_are_not_even_judged():
    _split(_fine_loads(), [25.0, 23.0, 43.0, 44.0])
    assert P.previous_assignment_stands(_negotiation([25.4, 22.8, 43.3, 43.7]))
    assert "judged_speeds" not in P._last_assignment
    # every rank faster by the same factor: shares unchanged
    assert P.previous_assignment_stands(_negotiation([50.0, 46.0, 86.0, 88.0]))


def test_a_split_out_of_balance_is_given_up_when_a_fresh_one_pays():
    loads = _fine_loads()
    _split(loads, [28.0, 23.0, 39.0, 44.0])  # what a probe said
    now = [25.1, 22.8, 43.0, 43.8]  # what the takes measured: shares moved by 9-12 %
    assert not P.previous_assignment_stands(_negotiation(now))
    _split(loads, now)
    assert P.previous_assignment_stands(_negotiation(now))
    assert P.previous_assignment_stands(_negotiation([25.6, 22.5, 43.2, 43.6]))


def test_a_split_that_cannot_be_improved_stands_and_is_not_judged_twice(monkeypatch):
    loads = [(f"big{i}", 0, 1 << 30, False) for i in range(4)]  # one unit per rank whatever the speeds
    _split(loads, [40.0, 40.0, 40.0, 40.0])
    calls = {"n": 0}
    real = P.assign_write_loads

    def counting(*a, **k):
        calls["n"] += 1
        return real(*a, **k)

    monkeypatch.setattr(P, "assign_write_loads", counting)
    now = [36.0, 40.0, 42.0, 40.0]
    assert P.previous_assignment_stands(_negotiation(now))
    assert calls["n"] == 1 and P._last_assignment["judged_speeds"] == now
    assert P.previous_assignment_stands(_negotiation([36.2, 39.9, 42.1, 40.0]))
    assert calls["n"] == 1  # within 3 % of what was judged: not judged again


def test_ranks_without_a_measured_speed():
    loads = _fine_loads(50)
    neg = _negotiation([None, None])
    out = P.assign_write_loads(loads, [0, 0], None)
    P._remember_assignment(neg, out, {(p, i): sz for p, i, sz, _sub in loads}, loads)
    assert P.previous_assignment_stands(_negotiation([None, None]))
    assert not P.previous_assignment_stands(_negotiation([40.0, 41.0]))  # speeds have appeared: split by them


def test_speed_estimate_is_the_median_of_the_last_takes():
    saved, ctx = dict(P._speed_estimates), P.speed_context()
    try:
        P._speed_estimates.clear()
        P.set_speed_context("SinkA")
        assert P.observed_speed() is None
        P._speed_estimates[""] = [30.0]  # the probed seed
        assert P.observed_speed() == 30.0
        P.record_take_speed(4 << 30, 100.0)  # 42.9 GB/s
        assert abs(P.observed_speed() - (4 << 30) / 100.0 / 1e6) < 1e-9  # a measurement replaces the seed
        for ms in (100.0, 101.0, 99.0):
            P.record_take_speed(4 << 30, ms)
        before = P.observed_speed()
        P.record_take_speed(4 << 30, 160.0)  # one slow take
        assert abs(P.observed_speed() - before) / before < 0.011
        for ms in (160.0, 161.0, 159.0):  # the rank really is slower now: the estimate follows within three takes
            P.record_take_speed(4 << 30, ms)
        assert abs(P.observed_speed() - (4 << 30) / 160.0 / 1e6) / P.observed_speed() < 0.01
        assert len(P._speed_estimates["SinkA"]) == P._SPEED_SAMPLES
        P.record_take_speed(1 << 20, 1.0)  # too small to say anything
        assert len(P._speed_estimates["SinkA"]) == P._SPEED_SAMPLES
        P.set_speed_context("SinkB")  # another kind of sink starts from the seed again
        assert P.observed_speed() == 30.0
    finally:
        P._speed_estimates.clear()
        P._speed_estimates.update(saved)
        P.set_speed_context(ctx)
