"""Storage service, controller REST surface over real HTTP, typed client and CLI."""
import io
import json
import os

import numpy as np
import pytest

from kubeml_b200.exceptions import KubeMLException
from kubeml_b200.httpd import decode_multipart, encode_multipart, http_call
from kubeml_b200.storage import ShardStore, StorageService


def _npy(arr):
    b = io.BytesIO()
    np.save(b, arr)
    return b.getvalue()


def test_shard_store_subsets_and_contiguous_ranges(tmp_path):
    st = ShardStore(str(tmp_path))
    x = np.arange(200 * 3, dtype=np.float32).reshape(200, 3)
    y = np.arange(200).reshape(200, 1)
    s = st.create("d", x, y, x[:70], y[:70])
    assert (s.train_set_size, s.test_set_size) == (200, 70)
    assert st.count_docs("d", "train") == 4 and st.count_docs("d", "test") == 2      # ceil(n/64)
    xs, ys = st.load_docs("d", "train", 1, 3)
    np.testing.assert_array_equal(xs, x[64:192])
    np.testing.assert_array_equal(ys, np.arange(64, 192))                             # labels flattened
    xs, _ = st.load_docs("d", "train", 3, 4)
    assert len(xs) == 200 - 192                                                       # short last subset
    with pytest.raises(KubeMLException) as e:
        st.create("d", x, y, x, y)
    assert e.value.status_code == 400
    st.delete("d")
    assert st.list() == []
    with pytest.raises(KubeMLException) as e:
        st.delete("d")
    assert e.value.status_code == 404


def test_storage_service_upload_contract(tmp_path):
    svc = StorageService(ShardStore(str(tmp_path)))
    x, y = np.zeros((10, 2), np.float32), np.zeros(10, np.int64)
    files = {k: (k + ".npy", _npy(v)) for k, v in (("x-train", x), ("y-train", y), ("x-test", x), ("y-test", y))}
    assert svc.upload("m", files) == {"result": "Dataset created"}
    with pytest.raises(KubeMLException) as e:
        svc.upload("m", files)
    assert e.value.status_code == 400 and "already exists" in e.value.message
    with pytest.raises(KubeMLException) as e:
        svc.upload("n", {})
    assert e.value.status_code == 400
    bad = dict(files, **{"x-train": ("x-train.csv", b"1,2")})
    with pytest.raises(KubeMLException) as e:
        svc.upload("n", bad)
    assert "extension" in e.value.message
    body, ctype = encode_multipart(files)
    assert set(decode_multipart(body, ctype)) == set(files)


@pytest.fixture()
def served_cluster(kube_home):
    from kubeml_b200.controller import Cluster
    c = Cluster(home=kube_home, max_parallelism=2, options={"force_host_sync": True})
    c.serve(host="127.0.0.1", controller_port=0)
    yield c, f"http://127.0.0.1:{c.controller_port}"
    c.shutdown()


def test_controller_rest_and_cli_end_to_end(served_cluster, tmp_path, capsys, monkeypatch):
    from kubeml_b200.cli.main import main as cli
    cluster, url = served_cluster
    assert http_call("GET", url + "/health") is None
    rng = np.random.default_rng(0)
    paths = {}
    for name, arr in (("xtr", rng.integers(0, 255, (256, 28, 28), dtype=np.uint8)), ("ytr", rng.integers(0, 10, (256,))),
                      ("xte", rng.integers(0, 255, (64, 28, 28), dtype=np.uint8)), ("yte", rng.integers(0, 10, (64,)))):
        p = tmp_path / f"{name}.npy"
        np.save(p, arr)
        paths[name] = str(p)
    monkeypatch.setenv("KUBEML_CONTROLLER_URL", url)
    assert cli(["dataset", "create", "--name", "mnist", "--traindata", paths["xtr"], "--trainlabels", paths["ytr"],
                "--testdata", paths["xte"], "--testlabels", paths["yte"]]) == 0
    assert cli(["dataset", "list"]) == 0
    assert "mnist" in capsys.readouterr().out
    # validation done by the CLI before it submits (cmd/train.go:89-119)
    assert cli(["train", "-d", "mnist", "-f", "lenet", "-e", "1", "-b", "4096", "--lr", "0.01"]) == 1
    assert cli(["train", "-d", "nope", "-f", "lenet", "-e", "1", "-b", "64", "--lr", "0.01"]) == 1
    assert cli(["train", "-d", "mnist", "-f", "nofn", "-e", "1", "-b", "64", "--lr", "0.01"]) == 1
    capsys.readouterr()
    assert cli(["train", "-d", "mnist", "-f", "lenet", "-e", "1", "-b", "64", "--lr", "0.01", "--parallelism", "2",
                "--static", "--K", "2", "--wait"]) == 0
    out = capsys.readouterr().out.strip().splitlines()
    job_id = out[0]
    hist = json.loads(out[-1])
    assert hist["id"] == job_id and len(hist["data"]["train_loss"]) == 1 and hist["data"]["parallelism"] == [2.0]
    assert cli(["history", "list"]) == 0 and job_id in capsys.readouterr().out
    assert cli(["history", "get", "--network", job_id]) == 0          # the harness' drifted flag name
    assert json.loads(capsys.readouterr().out)["id"] == job_id
    # per-job log file written by the parameter server, read by `kubeml logs` (cmd/log.go:28-65)
    assert cli(["logs", "--id", job_id]) == 0
    logged = capsys.readouterr().out
    assert "Starting train job" in logged and "Training finished" in logged
    assert cli(["logs", "--id", "deadbeef"]) == 1
    capsys.readouterr()
    assert cli(["task", "list", "--short"]) == 0
    assert cli(["task", "prune"]) == 0 and "Deleted 1 finished task logs" in capsys.readouterr().out
    assert cli(["function", "list"]) == 0 and "function_lenet" in capsys.readouterr().out
    assert cli(["history", "delete", "--id", job_id]) == 0
    assert http_call("GET", url + "/history") == []
    with pytest.raises(KubeMLException) as e:
        http_call("GET", url + "/history/" + job_id)
    assert e.value.status_code == 404
    metrics = http_call("GET", url + "/metrics").decode()
    assert "kubeml_job_running_total" in metrics
    assert cli(["dataset", "delete", "--name", "mnist"]) == 0


def test_function_registry_user_file(kube_home, tmp_path):
    from kubeml_b200.runtime.functions import FunctionRegistry
    reg = FunctionRegistry()
    src = tmp_path / "f.py"
    src.write_text("def main():\n    return {'ok': 1}, 200\n")
    reg.create("myfn", str(src))
    assert reg.exists("myfn") and "myfn" in reg.list() and reg.load("myfn")() == ({"ok": 1}, 200)
    with pytest.raises(KubeMLException):
        reg.create("myfn", str(src))
    reg.delete("myfn")
    assert not reg.exists("myfn") and reg.exists("lenet") and reg.exists("builtin:function_lenet")


def test_kubeml_command_is_packaged_and_launchable():
    """`kubeml` as a command: the launcher script of the source tree and the console-script entry of pyproject.toml
    (the reference ships a Go binary named kubeml, ml/pkg/kubeml-cli/)."""
    import importlib
    import os
    import subprocess
    import tomllib
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    meta = tomllib.load(open(os.path.join(root, "pyproject.toml"), "rb"))
    mod, fn = meta["project"]["scripts"]["kubeml"].split(":")
    assert callable(getattr(importlib.import_module(mod), fn))
    out = subprocess.run([os.path.join(root, "scripts", "kubeml"), "train", "--help"], capture_output=True, text=True,
                         timeout=120, cwd="/")
    assert out.returncode == 0 and "--sparse-avg" in out.stdout and "--resume" in out.stdout
