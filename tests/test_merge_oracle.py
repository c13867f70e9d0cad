"""K-AVG semantics on the host: the CPU oracle (ref-cpu) and the merge barrier with a shrinking
participant set."""
import threading

import numpy as np
import pytest

from kubeml_b200.exceptions import MergeError
from kubeml_b200.model import Model, ParallelSGD, TensorStore, weight_key
from kubeml_b200.model.model import kavg_oracle
from kubeml_b200.train.barrier import MergeBarrier


def test_oracle_average_float_and_truncating_int64():
    reps = [{"w": np.array([1.0, 2.0, 3.5], np.float32), "n": np.array([3], np.int64)},
            {"w": np.array([3.0, 0.0, 0.5], np.float32), "n": np.array([4], np.int64)},
            {"w": np.array([2.0, 1.0, 2.0], np.float32), "n": np.array([4], np.int64)}]
    out = kavg_oracle(reps)
    np.testing.assert_array_equal(out["w"], (np.float32(1) * (reps[0]["w"] + reps[1]["w"] + reps[2]["w"])) / np.float32(3))
    assert out["n"].dtype == np.int64 and out["n"][0] == 11 // 3
    neg = kavg_oracle([{"n": np.array([-7], np.int64)}, {"n": np.array([0], np.int64)}])
    assert neg["n"][0] == -3                         # Go's integer division truncates toward zero


def test_model_build_update_save_clear_key_layout():
    store = TensorStore()
    names = ["conv.weight", "bn.num_batches_tracked"]
    store.multi_set({weight_key("j", "conv.weight"): np.ones(4, np.float32),
                     weight_key("j", "bn.num_batches_tracked"): np.array(0, np.int64)})
    for f in range(2):
        store.multi_set({weight_key("j", "conv.weight", f): np.full(4, f + 1, np.float32),
                         weight_key("j", "bn.num_batches_tracked", f): np.array(10 + f, np.int64)})
    assert weight_key("j", "conv.weight", 1) == "j:conv.weight/1" and weight_key("j", "conv.weight") == "j:conv.weight"
    m = Model("j", names, store)
    m.build()
    assert m.state_dict["bn.num_batches_tracked"].dtype == "INT64" and m.num_parameters() == 5
    m.clear()
    m.update(0); m.update(1)
    ParallelSGD().average(m, 2)
    m.save()
    np.testing.assert_array_equal(store.tensorget("j:conv.weight"), np.full(4, 1.5, np.float32))
    assert int(store.tensorget("j:bn.num_batches_tracked")) == 10
    assert "FLOAT" in m.summary()


def _publish(store, job, fid, value):
    store.multi_set({weight_key(job, "w", fid): np.full(8, value, np.float32)})


def test_barrier_shrinking_participants_and_divisor():
    """3 workers: worker 2 has one round, workers 0/1 have two.  Round 1 averages {0,1,2}, round 2
    averages {0,1} (job.go:414-438)."""
    store = TensorStore()
    store.multi_set({weight_key("j", "w"): np.zeros(8, np.float32)})
    m = Model("j", ["w"], store)
    merges = []
    b = MergeBarrier(m, ParallelSGD(), 3, on_merge=lambda parts, dt: merges.append(sorted(parts)))

    def worker(fid, rounds):
        for r in range(rounds):
            _publish(store, "j", fid, float(10 * (r + 1) + fid))
            if r != rounds - 1:
                b.arrive(fid)
        b.finished(fid, True)

    ts = [threading.Thread(target=worker, args=(f, 2 if f < 2 else 1)) for f in range(3)]
    [t.start() for t in ts]; [t.join(10) for t in ts]
    assert b.done.is_set() and b.failed is None
    assert merges == [[0, 1, 2], [0, 1]]
    np.testing.assert_allclose(store.tensorget("j:w"), np.full(8, (20 + 21) / 2, np.float32))


def test_failed_worker_does_not_count_in_divisor():
    store = TensorStore()
    store.multi_set({weight_key("j", "w"): np.zeros(8, np.float32)})
    m = Model("j", ["w"], store)
    merges = []
    b = MergeBarrier(m, ParallelSGD(), 3, on_merge=lambda parts, dt: merges.append(sorted(parts)))
    _publish(store, "j", 0, 3.0); _publish(store, "j", 1, 5.0)
    b.finished(0, True)
    b.finished(2, False)          # crashed before publishing: released from the barrier, not summed
    b.finished(1, True)
    assert merges == [[0, 1]]
    np.testing.assert_allclose(store.tensorget("j:w"), np.full(8, 4.0, np.float32))


def test_barrier_with_no_contribution_fails_the_merge():
    store = TensorStore()
    m = Model("j", ["w"], store)
    b = MergeBarrier(m, ParallelSGD(), 1)
    b.finished(0, False)
    assert b.failed is not None and b.done.is_set()


def test_barrier_stress_random_arrival_order():
    """10^2 rounds x 4 workers with random sleeps: every round merges everyone exactly once."""
    import random
    import time
    store = TensorStore()
    store.multi_set({weight_key("j", "w"): np.zeros(4, np.float32)})
    m = Model("j", ["w"], store)
    merges = []
    b = MergeBarrier(m, ParallelSGD(), 4, on_merge=lambda parts, dt: merges.append(tuple(sorted(parts))))
    R = 100

    def worker(fid):
        rng = random.Random(fid)
        for r in range(R):
            time.sleep(rng.random() * 1e-4)
            _publish(store, "j", fid, float(r))
            if r != R - 1:
                b.arrive(fid)
        b.finished(fid, True)

    ts = [threading.Thread(target=worker, args=(f,)) for f in range(4)]
    [t.start() for t in ts]; [t.join(60) for t in ts]
    assert b.failed is None and len(merges) == R and set(merges) == {(0, 1, 2, 3)}
    np.testing.assert_allclose(store.tensorget("j:w"), np.full(8, R - 1, np.float32))
