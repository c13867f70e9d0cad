"""The reference arm's stand-in services must not be the thing the arm measures.

Round 1's arm spent 98 % of its timed region in a Nagle / delayed-ACK stall of the stand-in tensor
store (two small writes per reply on a socket without TCP_NODELAY): 8.2 s to read one ResNet-34
state_dict.  These tests pin the fix: a 218-tensor / 83 MB state_dict round-trips through the store in
well under a second, and the arm itself runs end to end on CPU with the reference's LeNet function and
reports where its time went.
"""
import json
import os
import socket
import socketserver
import subprocess
import sys
import threading
import time

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SHIMS = os.path.join(ROOT, "baseline", "shims")


def _free_port() -> int:
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.fixture()
def store_server():
    sys.path.insert(0, ROOT)
    from baseline.reference_arm import TensorStore, _StoreHandler
    port = _free_port()
    srv = socketserver.ThreadingTCPServer(("127.0.0.1", port), _StoreHandler, bind_and_activate=False)
    srv.allow_reuse_address = True
    srv.daemon_threads = True
    srv.server_bind()
    srv.server_activate()
    srv.store = TensorStore()
    threading.Thread(target=srv.serve_forever, daemon=True).start()
    yield port
    srv.shutdown()
    srv.server_close()


def _load_shim():
    import importlib.util
    spec = importlib.util.spec_from_file_location("_kml_redisai_shim", os.path.join(SHIMS, "redisai", "__init__.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_state_dict_round_trip_is_not_a_tcp_stall(store_server):
    shim = _load_shim()
    # ResNet-34-like: 218 entries, ~83 MB fp32, 36 of them int64 scalars
    rng = np.random.default_rng(0)
    tensors = {}
    big = (83 * 2 ** 20 // 4) // 182
    for i in range(182):
        tensors[f"job:layer{i}.weight"] = rng.standard_normal(big, dtype=np.float32)
    for i in range(36):
        tensors[f"job:bn{i}.num_batches_tracked"] = np.array(i, dtype=np.int64)
    assert len(tensors) == 218
    c = shim.Client("127.0.0.1", store_server)
    t0 = time.perf_counter()
    for k, v in tensors.items():
        c.tensorset(k, v, dtype="float32")
    t_set = time.perf_counter() - t0
    t0 = time.perf_counter()
    got = {k: c.tensorget(k) for k in tensors}
    t_get = time.perf_counter() - t0
    c.close()
    for k, v in tensors.items():
        assert got[k].dtype == v.dtype and got[k].shape == v.shape
        assert np.array_equal(got[k], v)
    assert got["job:layer0.weight"].flags.writeable            # torch.from_numpy wants a writable array
    # 40 ms x 218 tensors = 8.7 s is the stall; loopback moves 83 MB in ~0.1-0.3 s
    assert t_get < 1.0, f"tensorget of a 218-tensor state_dict took {t_get:.2f} s"
    assert t_set < 1.5, f"tensorset of a 218-tensor state_dict took {t_set:.2f} s"
    assert shim.STATS["get_n"] == 218 and shim.STATS["set_n"] == 218


@pytest.mark.skipif(not os.path.isdir(os.path.join(ROOT, "baseline", "_ref", "kubeml")),
                    reason="baseline/_ref not installed")
def test_reference_arm_runs_lenet_on_cpu_and_reports_breakdown():
    env = dict(os.environ, KUBEML_REF_FUNCTION="function_lenet", MASTER_PORT=str(_free_port()),
               CUDA_VISIBLE_DEVICES="")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "1",
                          "--steps", "16", "--warmup", "3"], env=env, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    line = [l for l in out.stdout.splitlines() if l.startswith("{")][-1]
    rec = json.loads(line)
    assert rec["impl"] == "reference" and "unavailable" not in rec, rec
    b = rec["breakdown"]
    assert b["job_merges"] == b["rounds"] == 2
    assert b["load_model_tensors"] == b["save_model_tensors"] == 2 * 10          # LeNet: 10 state_dict entries
    # the store is not the bottleneck any more: < 5 ms per tensor even on a loaded CI box
    assert b["load_model_s"] / b["load_model_tensors"] < 0.005, b
    parts = b["load_model_s"] + b["save_model_s"] + b["job_merge_s"] + b["train_and_data_s"]
    assert abs(parts - b["timed_region_s"]) < 1e-2
