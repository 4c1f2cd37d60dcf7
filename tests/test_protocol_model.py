"""The cross-GPU protocol of the training step, checked as a happens-before graph (CPU only).

csrc/allreduce.cu ends its kernels without exit barriers; pytorch_mnist_ddp_b200/parallel/protocol_model.py restates
the step (three communication kernels per rank, the branch streams of csrc/engine.cu:enqueue_train, rank-0-only
evaluation between steps) as events and proves that no load can see a stale or overwritten value and that no wait can
deadlock.  Each ordering the argument rests on is then removed in turn and must be reported."""
import pytest

from pytorch_mnist_ddp_b200.parallel import protocol_model as pm


@pytest.mark.parametrize("world", [2, 3, 4, 8])
@pytest.mark.parametrize("fc_mode", [pm.ONESHOT, pm.TWOSHOT])
@pytest.mark.parametrize("conv2_mode", [pm.ONESHOT, pm.TWOSHOT])
@pytest.mark.parametrize("ordered", [True, False])
def test_step_protocol_has_no_race_and_no_deadlock(world, fc_mode, conv2_mode, ordered):
    st = pm.verify(world, steps=3, fc_mode=fc_mode, conv2_mode=conv2_mode, ordered_conv2=ordered)
    assert st["waits"] > 0 and st["reads"] > 0
    # rank-0-only work after a later step / never, more chunks per owner, more CTAs in the fc kernel
    pm.verify(world, steps=3, fc_mode=fc_mode, conv2_mode=conv2_mode, ordered_conv2=ordered, eval_after=1,
              chunks=2 * world + 1, fc_ctas=3)
    pm.verify(world, steps=2, fc_mode=fc_mode, conv2_mode=conv2_mode, ordered_conv2=ordered, eval_after=None)


# mutation -> (fc mode, conv2 mode) in which that ordering carries weight, and what its removal must look like
CASES = [
    ("no_finalize_barrier", pm.TWOSHOT, pm.TWOSHOT, "early overwrite"),
    ("no_finalize_barrier", pm.ONESHOT, pm.ONESHOT, "early overwrite"),
    ("no_r_complete_wait", pm.TWOSHOT, pm.TWOSHOT, "stale read"),
    ("no_r_complete_signal_order", pm.TWOSHOT, pm.TWOSHOT, "stale read"),
    ("fc_before_dgrad", pm.TWOSHOT, pm.TWOSHOT, "early overwrite"),
    ("fc_before_dgrad", pm.ONESHOT, pm.TWOSHOT, "early overwrite"),
    ("no_join_fc", pm.TWOSHOT, pm.TWOSHOT, "early overwrite"),
    ("no_join_conv2", pm.TWOSHOT, pm.ONESHOT, "early overwrite"),
    ("no_gather_wait", pm.TWOSHOT, pm.TWOSHOT, "stale read"),
    ("no_fc_entry_barrier", pm.TWOSHOT, pm.TWOSHOT, "early overwrite"),
    ("no_fc_entry_barrier", pm.ONESHOT, pm.TWOSHOT, "stale read"),
    ("stale_token", pm.TWOSHOT, pm.TWOSHOT, "token reuse"),
]


@pytest.mark.parametrize("world", [2, 4])
@pytest.mark.parametrize("mutation,fc_mode,conv2_mode,what", CASES)
def test_every_removed_ordering_is_reported(world, mutation, fc_mode, conv2_mode, what):
    with pytest.raises(pm.Violation, match=what):
        pm.verify(world, steps=3, fc_mode=fc_mode, conv2_mode=conv2_mode, mutate=mutation)


def test_every_mutation_has_a_case():
    assert {c[0] for c in CASES} == set(pm.MUTATIONS)


def test_a_wait_nobody_signals_is_a_deadlock():
    m = pm.StepModel(2, 1)
    m.g.add(0, "WAIT", ("fc", 0, 1, 0), 99, "extra wait")
    with pytest.raises(pm.Violation, match="deadlock"):
        m.check()


def test_a_cyclic_wait_is_a_deadlock():
    """Two ranks that each wait for the other's signal BEFORE sending their own."""
    g = pm.Graph()
    w0 = g.add(0, "WAIT", ("x", 1, 0), 1, "rank 0 waits first")
    g.add(0, "SIG", ("x", 0, 1), 1, "then signals", (w0,))
    w1 = g.add(1, "WAIT", ("x", 0, 1), 1, "rank 1 waits first")
    g.add(1, "SIG", ("x", 1, 0), 1, "then signals", (w1,))
    m = pm.StepModel.__new__(pm.StepModel)
    m.g = g
    with pytest.raises(pm.Violation, match="deadlock"):
        m.check()


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("mode", [pm.ONESHOT, pm.TWOSHOT])
def test_standalone_calls_need_the_exit_barrier(world, mode):
    """Back-to-back `b200_comm_run` calls (tests, microbenchmark) have no finalize kernel between them: with the exit
    barrier they are clean, without it a fast rank's next gradient fill races a slow rank's loads -- which is why the
    stand-alone entry point keeps the barrier that the training step drops."""
    assert pm.verify_standalone(world, calls=3, mode=mode)["waits"] > 0
    with pytest.raises(pm.Violation, match="early overwrite"):
        pm.verify_standalone(world, calls=3, mode=mode, exit_barrier=False)
