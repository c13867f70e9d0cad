"""Pure-Python semantics: sharding helpers, wire types, errors, scheduler policy."""
import json
import math

import pytest
from hypothesis import given, settings, strategies as st

from kubeml_b200.api.types import (History, JobHistory, JobInfo, JobState, MetricUpdate, TrainOptions,
                                   TrainRequest, TrainTask, latest_metrics)
from kubeml_b200.args import _KubeArgs
from kubeml_b200.exceptions import (DataError, DatasetNotFoundError, InvalidArgsError, KubeMLException, MergeError,
                                    error_from_response, to_error_payload)
from kubeml_b200.scheduler.policy import TaskOperation, ThroughputBasedPolicy
from kubeml_b200.util import get_subset_period, round_schedule, rounds_for, split_minibatches


@given(st.integers(0, 500), st.integers(1, 16))
@settings(max_examples=200, deadline=None)
def test_split_minibatches_is_a_balanced_contiguous_partition(n, parts):
    rs = split_minibatches(range(n), parts)
    assert len(rs) == parts
    assert [i for r in rs for i in r] == list(range(n))            # contiguous, in order, complete
    sizes = [len(r) for r in rs]
    assert max(sizes) - min(sizes) <= 1
    assert sizes == sorted(sizes, reverse=True)                    # the first n % parts get the extra one


def test_subset_period_formula():
    assert get_subset_period(-1, 64, range(10, 37)) == 27          # K=-1: the whole shard, one sync per epoch
    assert get_subset_period(16, 64, range(100)) == 16
    assert get_subset_period(8, 128, range(100)) == 16
    assert get_subset_period(3, 50, range(100)) == math.ceil(150 / 64)


@given(st.integers(1, 300), st.integers(1, 8), st.sampled_from([-1, 1, 2, 8, 16]), st.sampled_from([16, 64, 128]))
@settings(max_examples=150, deadline=None)
def test_round_schedule_shrinks_monotonically(docs, n, k, batch):
    sched = round_schedule(docs, n, k, batch)
    counts = [len(rounds_for(docs, n, r, k, batch)) for r in range(n)]
    assert len(sched) == max(counts)
    for i, parts in enumerate(sched):
        assert parts == [r for r in range(n) if counts[r] > i]
    for a, b in zip(sched, sched[1:]):
        assert set(b) <= set(a)                                    # workers only ever drop out
    if k == -1:
        assert len(sched) <= 1


def test_train_request_golden_json():
    req = TrainRequest(model_type="resnet34", batch_size=128, epochs=30, dataset="cifar10", lr=0.1,
                       function_name="resnet34",
                       options=TrainOptions(default_parallelism=8, static_parallelism=True, validate_every=1, k=8,
                                            goal_accuracy=70.0))
    want = {"model_type": "resnet34", "batch_size": 128, "epochs": 30, "dataset": "cifar10", "lr": 0.1,
            "function_name": "resnet34",
            "options": {"default_parallelism": 8, "static_parallelism": True, "validate_every": 1, "k": 8,
                        "goal_accuracy": 70.0}}
    assert req.to_dict() == want
    assert TrainRequest.from_json(json.dumps(want)) == req
    task = TrainTask(request=req, job=JobInfo(id="abcd1234", state=JobState(parallelism=4, elapsed_time=1.5)))
    d = task.to_dict()
    assert d["job"] == {"id": "abcd1234", "state": {"parallelism": 4, "elapsed_time": 1.5}}
    assert TrainTask.from_dict(d).job.state.parallelism == 4


def test_metric_update_keeps_reference_spelling():
    h = JobHistory(validation_loss=[1.0, 0.5], accuracy=[10.0, 55.0], train_loss=[2.0], parallelism=[2.0],
                   epoch_duration=[3.0])
    m = latest_metrics(h)
    d = m.to_dict()
    assert d["validations_loss"] == 0.5 and d["accuracy"] == 55.0 and d["train_loss"] == 2.0
    assert MetricUpdate.from_dict(d) == m
    hist = History(id="x", task=TrainRequest(), data=h)
    assert History.from_dict(hist.to_dict()).data.accuracy == [10.0, 55.0]


def test_kubeargs_round_trip_and_errors():
    q = {"task": "train", "jobId": "j1", "N": "4", "K": "-1", "funcId": "3", "batchSize": "64", "lr": "0.01",
         "epoch": "2"}
    a = _KubeArgs.parse(q)
    assert (a._job_id, a._N, a._K, a._task, a._func_id, a.batch_size, a.lr, a.epoch) == ("j1", 4, -1, "train", 3, 64, 0.01, 2)
    assert _KubeArgs.parse(a.to_query()).to_query() == a.to_query()
    with pytest.raises(InvalidArgsError):
        _KubeArgs.parse(dict(q, N="four"))


def test_error_contract():
    assert MergeError(ValueError("x")).status_code == 500
    assert DataError().status_code == 400 and DatasetNotFoundError().status_code == 404
    assert to_error_payload(DatasetNotFoundError()) == {"error": "Dataset not found in storage service", "code": 404}
    assert to_error_payload(ValueError("boom")) == {"error": "ValueError('boom')", "code": 500}
    assert error_from_response(200, b"") is None
    e = error_from_response(500, b'{"error": "Error merging model: x", "code": 500}')
    assert isinstance(e, KubeMLException) and e.message == "Error merging model: x" and e.status_code == 500
    assert error_from_response(502, b"  bad gateway \n").message == "bad gateway"


def _task(jid, parallelism, elapsed, default=5):
    return TrainTask(request=TrainRequest(options=TrainOptions(default_parallelism=default)),
                     job=JobInfo(id=jid, state=JobState(parallelism=parallelism, elapsed_time=elapsed)))


def test_throughput_policy_transition_table():
    p = ThroughputBasedPolicy(max_parallelism=8)
    assert p.calculate_parallelism(_task("a", 0, 0.0, default=3)) == (3, TaskOperation.CREATE)
    assert p.calculate_parallelism(_task("a", 3, 10.0)) == (4, TaskOperation.UPDATE)      # no reference yet: +1
    assert p.reference_time("a") == 10.0
    assert p.calculate_parallelism(_task("a", 4, 10.4)) == (5, TaskOperation.UPDATE)      # <= 1.05x: +1, re-reference
    assert p.reference_time("a") == 10.4
    assert p.calculate_parallelism(_task("a", 5, 11.5)) == (5, TaskOperation.UPDATE)      # in between: keep
    assert p.reference_time("a") == 10.4
    assert p.calculate_parallelism(_task("a", 5, 12.5)) == (4, TaskOperation.UPDATE)      # >= 1.2x: -1, re-reference
    assert p.reference_time("a") == 12.5
    p.task_finished("a")
    assert p.reference_time("a") is None


def test_throughput_policy_is_clamped():
    p = ThroughputBasedPolicy(max_parallelism=2)
    p.calculate_parallelism(_task("b", 0, 0.0, default=9))
    assert p.calculate_parallelism(_task("b", 2, 5.0))[0] == 2            # would be 3
    assert p.calculate_parallelism(_task("b", 1, 50.0))[0] == 1           # would be 0 in the reference
