"""deploy/ is what replaces the reference's Helm chart (ml/charts/kubeml: Deployments + Services, a Prometheus
ServiceMonitor scraped every 10 s, the Grafana dashboard).  These tests start what the unit files start, with the
environment file's values, and check that Prometheus' scrape target and every dashboard query are really served."""
import json
import os
import re
import shlex
import socket
import subprocess
import sys
import time

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DEPLOY = os.path.join(ROOT, "deploy")


def _env_file():
    env = {}
    for line in open(os.path.join(DEPLOY, "kubeml.env")):
        line = line.strip()
        if line and not line.startswith("#"):
            k, _, v = line.partition("=")
            env[k] = v
    return env


def _exec_start(unit):
    text = open(os.path.join(DEPLOY, "systemd", unit)).read()
    assert "[Unit]" in text and "[Service]" in text and "[Install]" in text
    assert "EnvironmentFile=/etc/kubeml_b200/kubeml.env" in text
    return re.search(r"^ExecStart=(.*)$", text, re.M).group(1)


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _wait_http(url, timeout=240):
    from kubeml_b200.httpd import http_call
    deadline = time.time() + timeout
    while time.time() < deadline:
        try:
            return http_call("GET", url)
        except Exception:  # noqa: BLE001
            time.sleep(0.3)
    raise AssertionError(f"{url} never answered")


def _start(cmd, env, tmp_path):
    e = dict(os.environ)
    e.update(env)
    e.update(KUBEML_HOME=str(tmp_path), CUDA_VISIBLE_DEVICES="", PYTHONPATH=ROOT)
    argv = shlex.split(cmd)
    assert argv[:2] == ["/usr/bin/env", "python"]
    return subprocess.Popen([sys.executable] + argv[2:], env=e, cwd=str(tmp_path), stdout=subprocess.PIPE,
                            stderr=subprocess.STDOUT, text=True)


@pytest.mark.parametrize("unit", ["kubeml.service", "kubeml-workers.service"])
def test_unit_file_command_brings_up_the_control_plane(unit, tmp_path):
    from kubeml_b200.httpd import http_call
    env = _env_file()
    port = _free_port()
    env["KUBEML_CONTROLLER_PORT"] = str(port)
    cmd = _exec_start(unit).replace("${KUBEML_CONTROLLER_PORT}", str(port))
    if "torch.distributed.run" in cmd:          # one worker on this GPU-less box instead of eight
        cmd = cmd.replace("--nproc-per-node 8", "--nproc-per-node 1").replace("--master-port 29400",
                                                                             f"--master-port {_free_port()}")
    proc = _start(cmd, env, tmp_path)
    try:
        url = f"http://127.0.0.1:{port}"
        assert _wait_http(url + "/health") is None
        # what deploy/prometheus.yml scrapes: the target and the path must be served by this process
        prom = open(os.path.join(DEPLOY, "prometheus.yml")).read()
        assert f'targets: ["127.0.0.1:{_env_file()["KUBEML_CONTROLLER_PORT"]}"]' in prom
        path = re.search(r"metrics_path:\s*(\S+)", prom).group(1)
        metrics = http_call("GET", url + path).decode()
        assert "kubeml_job_running_total" in metrics
        assert http_call("GET", url + "/tasks") == [] and http_call("GET", url + "/dataset") == []
        if "torch.distributed.run" in cmd:
            # the process-per-GPU deployment must also *train*: dataset upload, a LeNet job through the scheduler
            # and the worker rank, history back over HTTP (on this GPU-less box the merge is the collective path)
            import numpy as np
            from kubeml_b200.api.types import TrainOptions, TrainRequest
            from kubeml_b200.controller.client import make_kubeml_client
            rng = np.random.default_rng(0)
            files = {}
            for name, arr in (("xtr", rng.integers(0, 255, (256, 28, 28), dtype=np.uint8)),
                              ("ytr", rng.integers(0, 10, (256,))),
                              ("xte", rng.integers(0, 255, (64, 28, 28), dtype=np.uint8)),
                              ("yte", rng.integers(0, 10, (64,)))):
                files[name] = str(tmp_path / f"{name}.npy")
                np.save(files[name], arr)
            v1 = make_kubeml_client(url=url).v1()
            v1.datasets().create("mnist", files["xtr"], files["ytr"], files["xte"], files["yte"])
            job_id = v1.networks().train(TrainRequest(
                model_type="lenet", batch_size=64, epochs=2, dataset="mnist", lr=0.01, function_name="lenet",
                options=TrainOptions(default_parallelism=1, static_parallelism=True, validate_every=1, k=2)))
            deadline, hist = time.time() + 240, None
            while time.time() < deadline and hist is None:
                try:
                    hist = v1.histories().get(job_id)
                except Exception:  # noqa: BLE001
                    time.sleep(0.5)
            assert hist is not None, "the job never produced a history"
            assert len(hist.data.train_loss) == 2 and len(hist.data.accuracy) == 2
    finally:
        proc.terminate()
        try:
            proc.wait(timeout=20)
        except subprocess.TimeoutExpired:
            proc.kill()


def test_every_dashboard_query_names_a_metric_the_parameter_server_exports():
    from kubeml_b200.ps.metrics import MetricsRegistry
    from kubeml_b200.api.types import MetricUpdate
    dash = json.load(open(os.path.join(DEPLOY, "grafana", "KubeML.json")))
    exprs = re.findall(r'"expr":\s*"([a-zA-Z_:][a-zA-Z0-9_:]*)', json.dumps(dash))
    assert len(exprs) >= 6
    reg = MetricsRegistry()
    reg.task_started("train")
    reg.update("abcd1234", MetricUpdate(validation_loss=1.0, accuracy=50.0, train_loss=2.0, parallelism=2,
                                        epoch_duration=3.0))
    text = reg.render()
    for name in set(exprs):
        assert re.search(rf"^{name}\{{", text, re.M), f"dashboard queries {name}, /metrics does not export it"


def test_kubernetes_manifests_and_image_recipe_name_real_commands_ports_and_paths():
    """deploy/k8s/kubeml.yaml + deploy/Dockerfile (the chart and the image builds of the reference):
    every command must be one the entry points accept, every probe / scrape path one the control plane serves."""
    import argparse
    import yaml
    docs = [d for d in yaml.safe_load_all(open(os.path.join(DEPLOY, "k8s", "kubeml.yaml"))) if d]
    kinds = [d["kind"] for d in docs]
    assert kinds == ["Namespace", "PersistentVolumeClaim", "Deployment", "Service", "ServiceMonitor"]
    dep = docs[2]
    c = dep["spec"]["template"]["spec"]["containers"][0]
    assert c["resources"]["limits"]["nvidia.com/gpu"] == 8
    cmd = c["command"]
    assert cmd[:3] == ["python", "-m", "torch.distributed.run"] and "kubeml_b200.runtime.spmd_main" in cmd
    assert int(cmd[cmd.index("--nproc-per-node") + 1]) == c["resources"]["limits"]["nvidia.com/gpu"]
    # the worker entry accepts exactly these flags (same parser as runtime/spmd_main.py)
    tail = cmd[cmd.index("kubeml_b200.runtime.spmd_main") + 1:]
    ap = argparse.ArgumentParser(); ap.add_argument("--host"); ap.add_argument("--port", type=int)
    args = ap.parse_args(tail)
    port = {p["name"]: p["containerPort"] for p in c["ports"]}["http"]
    assert args.port == port == int(_env_file()["KUBEML_CONTROLLER_PORT"])
    # probes, Service and ServiceMonitor point at routes the controller serves
    from kubeml_b200.controller import Cluster
    import tempfile
    cl = Cluster(home=tempfile.mkdtemp(), start=False)
    routes = cl.controller.router()._routes
    paths = {c["readinessProbe"]["httpGet"]["path"], c["livenessProbe"]["httpGet"]["path"],
             docs[4]["spec"]["endpoints"][0]["path"]}
    assert paths == {"/health", "/metrics"}
    for p in paths:
        assert any(m == "GET" and rx.fullmatch(p) for m, rx, _ in routes), f"GET {p} is not served"
    assert docs[3]["metadata"]["name"] == "controller" and docs[3]["spec"]["ports"][0]["targetPort"] == "http"
    assert docs[4]["spec"]["endpoints"][0]["interval"] == "10s"
    assert docs[4]["spec"]["selector"]["matchLabels"].items() <= docs[3]["metadata"]["labels"].items()
    cl.shutdown()
    # the image: builds the native library for sm_100a without a GPU and starts the same entry binary
    df = open(os.path.join(DEPLOY, "Dockerfile")).read()
    assert "__graft_entry__" in df and 'ENTRYPOINT ["python", "-m", "kubeml_b200"]' in df
    from kubeml_b200.cli.main import build_parser
    m = re.search(r"^CMD (\[.*\])$", df, re.M)
    ns = build_parser().parse_args(json.loads(m.group(1)))
    assert ns.port == port
