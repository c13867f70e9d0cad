import math
import os

import numpy as np

from kubeml_b200.experiments import analysis
from kubeml_b200.experiments.harness import (KubemlExperiment, grid_points, point_title, retry, run_grid,
                                             to_dataframe)
from kubeml_b200.experiments.sysmetrics import MetricsCollector


def test_tta_definitions():
    acc = [50, 69, 71, 68, 72, 73, 74]
    t = [1, 2, 3, 4, 5, 6, 7]
    assert analysis.tta(acc, t, 70) == 3
    assert analysis.tta_crossbow(acc, t, 70) == 6            # median of the last five first reaches 70 at epoch 6
    assert math.isnan(analysis.tta(acc, t, 99))
    assert analysis.total_time(t) == 7 and analysis.mean_parallelism([2, 4]) == 3


def test_retry_backoff():
    calls = {"n": 0}

    @retry(ValueError, total_tries=3, initial_wait=0.001)
    def flaky():
        calls["n"] += 1
        if calls["n"] < 3:
            raise ValueError
        return "ok"
    assert flaky() == "ok" and calls["n"] == 3


def test_grid_resume_and_dataframe(kube_home, monkeypatch, tmp_path):
    from kubeml_b200.controller import Cluster, make_kubeml_client
    monkeypatch.setenv("KUBEML_CUDA_GRAPH", "0")
    c = Cluster(home=kube_home, max_parallelism=2, options={"force_host_sync": True})
    rng = np.random.default_rng(0)
    c.shard_store.create("mnist", rng.integers(0, 255, (256, 28, 28), dtype=np.uint8), rng.integers(0, 10, (256,)),
                         rng.integers(0, 255, (64, 28, 28), dtype=np.uint8), rng.integers(0, 10, (64,)))
    grid = {"batch": [64], "k": [-1, 2], "parallelism": [1, 2]}
    assert len(grid_points(grid)) == 4 and point_title("lenet", grid_points(grid)[0]) == "lenet-b64-k-1-p1"
    mc = MetricsCollector(interval=0.05)
    recs = run_grid(make_kubeml_client(c), "lenet", "mnist", 0.01, grid, str(tmp_path / "out"), epochs=1,
                    metrics=mc, timeout=300)
    assert len(recs) == 4 and all(r.error is None for r in recs)
    assert "cpu" in recs[0].system_metrics and len(recs[0].system_metrics["cpu"]) >= 1
    n_files = len(os.listdir(tmp_path / "out"))
    again = run_grid(make_kubeml_client(c), "lenet", "mnist", 0.01, grid, str(tmp_path / "out"), epochs=1)
    assert len(again) == 4 and len(os.listdir(tmp_path / "out")) == n_files
    assert [r.task_id for r in again] == [r.task_id for r in recs]          # nothing was re-run
    df = to_dataframe(recs)
    assert list(df["k"]) == [-1, -1, 2, 2] and len(df["train_loss"][0]) == 1
    c.shutdown()
    assert KubemlExperiment._fake_history(3).data.accuracy == [10.0, 20.0, 30.0]


def test_k_optimizer_learns_a_trend():
    ko = analysis.KOptimizer()
    for _ in range(30):
        for k, t in ((8, 1480.0), (16, 925.0), (32, 665.0), (-1, 450.0)):      # BASELINE.md, batch 128
            ko.update(128, 0.1, 4, k, t, 70.0)
    assert ko.best_k(128, 0.1, 4) == -1
    t8, _ = ko.predict(128, 0.1, 4, 8)
    tinf, _ = ko.predict(128, 0.1, 4, -1)
    assert t8 > tinf


def test_stock_torch_experiment_records_like_a_kubeml_experiment(kube_home):
    """The comparison system in the harness (the reference's TensorflowExperiment + TimeHistory): a plain-PyTorch
    LeNet run on a stored dataset, recorded with the same history keys, cumulative epoch times and system metrics,
    and usable by the same analysis code."""
    from kubeml_b200.experiments import analysis, harness
    from kubeml_b200.experiments.sysmetrics import MetricsCollector
    from kubeml_b200.functions.function_lenet import LeNet
    from kubeml_b200.storage import default_store
    rng = np.random.default_rng(0)
    store = default_store()
    x = rng.integers(0, 255, (256, 28, 28), dtype=np.uint8)
    y = (x.reshape(256, -1).mean(1) > 127).astype(np.int64)          # learnable: bright vs dark images
    store.create("mnist2", x[:192], y[:192], x[192:], y[192:])
    exp = harness.StockTorchExperiment("lenet-batch32-epochs3", LeNet, "mnist2", batch=32, epochs=3, lr=0.05,
                                       store=store, device="cpu", metrics=MetricsCollector(interval=0.2),
                                       transform=lambda t: t.float().div(255.0).sub(0.5).unsqueeze(1))
    rec = exp.run()
    assert rec.error is None, rec.error
    h = rec.history["data"]
    assert set(h) == {"train_loss", "validation_loss", "accuracy", "parallelism", "epoch_duration"}
    assert all(len(v) == 3 for v in h.values()) and h["parallelism"] == [1.0] * 3
    assert rec.history["times"] == sorted(rec.history["times"]) and len(exp.times) == 3
    assert abs(exp.times[-1] - sum(h["epoch_duration"])) < 0.5
    assert np.isfinite(h["train_loss"]).all() and h["train_loss"][-1] <= h["train_loss"][0] + 0.05
    assert "cpu" in rec.system_metrics and rec.wall_seconds > 0
    df = harness.to_dataframe([rec])
    assert list(df["epochs"]) == [3] and len(df["accuracy"][0]) == 3
    assert analysis.total_time(h["epoch_duration"]) > 0


def test_experiment_driver_runs_a_grid_resumes_and_summarises(kube_home, tmp_path, capsys):
    """`python -m kubeml_b200.experiments` (the reference's train.py / time_to_accuracy.py drivers): grid from the
    command line, one result file per point, a second invocation recomputes nothing, summary with both TTA values."""
    import json
    from kubeml_b200.experiments.__main__ import main
    from kubeml_b200.storage import default_store
    rng = np.random.default_rng(0)
    x = rng.integers(0, 255, (64 * 6, 28, 28), dtype=np.uint8)
    y = (x.reshape(len(x), -1).mean(1) > 127).astype(np.int64)
    default_store().create("mnist", x[:64 * 4], y[:64 * 4], x[64 * 4:], y[64 * 4:])
    out = str(tmp_path / "grid")
    argv = ["--function", "lenet", "--dataset", "mnist", "--lr", "0.01", "--epochs", "2", "--batch", "64",
            "--k", "-1", "2", "--parallelism", "1", "2", "--goal", "30", "--out", out, "--local", "--timeout", "300"]
    assert main(argv) == 0
    files = sorted(f for f in os.listdir(out) if f.startswith("lenet-"))
    assert files == ["lenet-b64-k-1-p1.json", "lenet-b64-k-1-p2.json", "lenet-b64-k2-p1.json", "lenet-b64-k2-p2.json"]
    summ = json.load(open(os.path.join(out, "summary.json")))
    assert len(summ["rows"]) == 4 and all(r["error"] is None and r["total_time_s"] > 0 for r in summ["rows"])
    assert {(r["k"], r["parallelism"]) for r in summ["rows"]} == {(-1, 1), (-1, 2), (2, 1), (2, 2)}
    assert all("tta_s" in r and "tta_crossbow_s" in r for r in summ["rows"])
    stamp = {f: os.path.getmtime(os.path.join(out, f)) for f in files}
    capsys.readouterr()
    assert main(argv) == 0                               # resume: every point already has its file
    assert {f: os.path.getmtime(os.path.join(out, f)) for f in files} == stamp
    assert "4 of 4 grid points" in capsys.readouterr().out
