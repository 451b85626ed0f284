"""The hand-off protocol of the CTA-pair GEMM (csrc/gemm_tc.cu, tc_pair_gemm_kernel) checked WITHOUT a GPU: tests/tools/pair_gemm_model.py
models its mbarriers, rings and roles and runs them under random schedules with random, occasionally very long, TMA latencies.

Background: on B200 the inference bench faulted (`unspecified launch failure`) about once per 10^5 GEMM launches.  The two converter
groups take alternate k-blocks; with three slots the k-blocks of a slot alternate between the groups, so on a "tile landed" barrier
shared by both a group tested a phase it had not seen the predecessor of, and an mbarrier parity test also succeeds one whole phase
early.  The model reproduces all three outcomes of that layout (stale operands, a second expect_tx on a barrier with no pending
arrival = the fault, the odd deadlock) and none of them with a barrier per (group, slot), which is what the kernel has now."""
import os
import random
import re
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), 'tools'))
import pair_gemm_model as model                                  # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_model_constants_are_the_kernels():
    src = open(os.path.join(ROOT, 'deltaconv_b200', 'csrc', 'gemm_tc.cu')).read()
    for name in ('PA_SLOTS', 'PB_SLOTS', 'PT_SLOTS', 'PD_SLOTS'):
        value = int(re.search(rf'constexpr int {name} = (\d+);', src).group(1))
        assert value == getattr(model, name), name
    # the kernel has a "tile landed" barrier per converter group and slot, and the phase arithmetic that goes with it
    assert 'auto a_full = [&](int grp, int q)' in src
    assert 'auto b_full = [&](int grp, int sb)' in src
    assert '(it / (2 * PA_SLOTS)) & 1' in src and '(it / (2 * PB_SLOTS)) & 1' in src


def test_barrier_per_group_and_slot_holds_under_random_schedules():
    total = 0
    for units in (1, 2, 3, 5):
        for handoffs in (1, 2, 3, 4, 7):
            for slow in (0.0, 0.05, 0.3):
                bad, first = model.run_schedules(False, 25, seed=units * 100 + handoffs * 10 + int(slow * 10), units=units,
                                                 handoffs=handoffs, slow_load_prob=slow)
                total += 25
                assert bad == 0, (units, handoffs, slow, first)
    assert total == 1500


def test_shared_barriers_reproduce_the_fault_seen_on_b200():
    bad, first = model.run_schedules(True, 300, seed=1, units=4, handoffs=4)
    assert bad > 0 and 'holds k-block' in first, (bad, first)          # a group consumes a slot whose load has not landed
    # left running past the stale read (as the GPU does), some schedules arrive on a barrier that has no pending arrival left
    outcomes = set()
    for s in range(400):
        m = model.PairGemmModel(4, 5, True, random.Random(99000 + s), 0.08, data_checks=False)
        try:
            m.run()
            outcomes.add('stale' if m.stale_reads else 'clean')
        except model.ProtocolViolation as exc:
            outcomes.add('fault' if 'no pending arrival' in str(exc) else 'stuck')
    assert {'clean', 'stale', 'fault'} <= outcomes, outcomes


def test_slow_roles_do_not_break_it_either():
    """Roles that are scheduled fifty times less often than the others (a converter group behind the rest, a lagging producer, a
    slow epilogue): a parity wait is also wrong when the waiter is LATE by two phases — it would then block for ever."""
    names = ['mma'] + [f'{r}{c}' for c in (0, 1) for r in ('a_prod', 'b_prod', 'epi')] + [f'conv{c}{g}' for c in (0, 1) for g in (0, 1)]
    for s in range(800):
        rng = random.Random(5000 + s)
        weights = {n: rng.choice([1.0, 1.0, 1.0, 0.02]) for n in names}
        m = model.PairGemmModel(rng.randint(1, 5), rng.randint(1, 7), False, rng, rng.choice([0.0, 0.05, 0.3]))
        m.run(role_weights=weights)                              # raises ProtocolViolation on a stale read, a barrier fault or a deadlock
