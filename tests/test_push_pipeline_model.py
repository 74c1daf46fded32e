"""The queueing model quoted in DESIGN.md 6.a (tools/sim_push_pipeline.py) says what it is quoted for: a lock-step push
transport loses the spread of pages per owner every round, two multi-gets in flight recover it."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))


def test_two_in_flight_recover_the_lock_step_loss():
    from sim_push_pipeline import efficiency
    one = efficiency(n=8, rounds=200, depth=1)
    two = efficiency(n=8, rounds=200, depth=2)
    assert 0.88 < one < 0.94 and two > 0.985 and two - one > 0.05
    assert efficiency(n=8, rounds=200, depth=1, per_requester=True) > 0.985  # serving requesters as they arrive needs no depth
    assert efficiency(n=2, rounds=200, depth=2) > 0.99
