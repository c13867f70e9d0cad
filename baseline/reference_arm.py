"""`bench.py --impl reference`: the UNMODIFIED reference on this box.

What runs unmodified (from baseline/_ref, installed by baseline/install_reference.sh):
  * the reference Python library `kubeml` (KubeModel.start -> __train: Mongo fetch, DataLoader,
    load_model from the tensor store, user train(), save_model, POST /next, ...), and
  * the reference's own example function ml/experiments/kubeml/function_resnet34.py (torchvision
    resnet34(), per-sample PIL transforms, F.cross_entropy, torch.optim.SGD, loss.item()).
What cannot run here and is stood in for (none of kubeml_b200 is on this path):
  * flask / redisai / redis / pymongo are not installable offline -> baseline/shims (request
    context, a TCP tensor store with one blob per tensor like RedisAI, pickled 64-sample documents
    like the storage service writes);
  * the Go train job (no `go`, no module cache) -> `JobServer` below restates the part of it the
    function talks to: POST /next/{funcId} barrier, Model.Update (sum replicas), ParallelSGD.Average
    (divide; integer division for int64), Model.Save (ml/pkg/train/api.go:100-126, job.go:368-442,
    ml/pkg/model/model.go:249-302, parallelSGD.go:26-54).  The function reaches it through the HTTP
    proxy environment variables, exactly as it would reach http://job-<id>.kubeml.
One process per GPU (GPU_ID pins the device, python/kubeml/kubeml/util.py:21-33), as one Fission pod
per function.  Same metric and config as the own arm: ResNet-34 (1000 logits), synthetic CIFAR-10-shape
data, batch 128 per worker, K=8, SGD lr 0.1.  The reference computes in fp32 on its stock path.
"""
from __future__ import annotations

import json
import os
import pickle
import socketserver
import struct
import sys
import threading
import time
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "baseline", "_ref")
SHIMS = os.path.join(ROOT, "baseline", "shims")
BATCH, K_AVG = 128, 8
BASELINE_IMG_S = 3200.0


def unavailable(why: str) -> int:
    print(json.dumps({"impl": "reference", "unavailable": why}), flush=True)
    return 0


# ---------------------------------------------------------------------------------------------
# infrastructure stand-ins hosted by rank 0
# ---------------------------------------------------------------------------------------------
class TensorStore:
    def __init__(self):
        self.d = {}
        self.lock = threading.Lock()


class _StoreHandler(socketserver.BaseRequestHandler):
    def setup(self):
        # Without this the accepted socket keeps Nagle on: a reply whose header segment is still
        # un-ACKed holds the blob back for the client's delayed-ACK timer (~40 ms per tensor).
        import socket as _socket
        self.request.setsockopt(_socket.IPPROTO_TCP, _socket.TCP_NODELAY, 1)

    def handle(self):
        sys.path.insert(0, SHIMS)
        from redisai import _recv, _send
        store: TensorStore = self.server.store
        try:
            while True:
                header, blob = _recv(self.request)
                if header[0] == "set":
                    _, key, dt, shape = header
                    with store.lock:
                        store.d[key] = (dt, tuple(shape), bytes(blob))
                    _send(self.request, ("ok",))
                elif header[0] == "get":
                    with store.lock:
                        ent = store.d.get(header[1])
                    if ent is None:
                        _send(self.request, ("missing", None, None))
                    else:
                        _send(self.request, ("ok", ent[0], ent[1]), ent[2])
        except (ConnectionError, struct.error, OSError):
            return


class JobServer:
    """The slice of the Go train job a function interacts with during one epoch."""

    def __init__(self, store: TensorStore, job_id: str):
        self.store, self.job_id = store, job_id
        self.cv = threading.Condition()
        self.layers = []
        self.stats = {"merge_s": 0.0, "merges": 0, "barrier_wait_s": 0.0}
        self.reset(0)

    def reset(self, parallelism: int):
        with self.cv:
            self.running = parallelism
            self.pending = parallelism
            self.arrived = []
            self.generation = 0

    def _merge_locked(self):
        import numpy as np
        funcs = self.arrived
        t0 = time.perf_counter()
        # One thread, under the store lock: the Go job serialises Model.Update behind one mutex too
        # (ml/pkg/model/model.go:273-274); numpy releases the GIL inside the adds.
        with self.store.lock:
            for name in self.layers:
                acc = None
                for f in funcs:                                   # Model.Update: sum in arrival order
                    dt, shape, blob = self.store.d[f"{self.job_id}:{name}/{f}"]
                    a = np.frombuffer(blob, dtype=np.dtype(dt)).reshape(shape)
                    acc = a.copy() if acc is None else acc + a
                if np.issubdtype(acc.dtype, np.integer):          # ParallelSGD.Average
                    acc = (acc // len(funcs)).astype(acc.dtype)
                else:
                    acc = (acc / np.float32(len(funcs))).astype(acc.dtype)
                self.store.d[f"{self.job_id}:{name}"] = (acc.dtype.str, acc.shape, acc.tobytes())   # Model.Save
        self.stats["merge_s"] += time.perf_counter() - t0
        self.stats["merges"] += 1
        self.arrived = []
        self.pending = self.running
        self.generation += 1
        self.cv.notify_all()

    def arrive(self, func_id: int, returning: bool):
        with self.cv:
            self.arrived.append(func_id)
            if returning:
                self.running -= 1
            self.pending -= 1
            gen = self.generation
            if self.pending == 0:
                self._merge_locked()
            else:
                t0 = time.perf_counter()
                while self.generation == gen:
                    self.cv.wait()
                self.stats["barrier_wait_s"] += time.perf_counter() - t0


def _make_job_http(job: JobServer):
    class H(BaseHTTPRequestHandler):
        def do_POST(self):
            path = self.path.split("://", 1)[-1]
            path = path[path.index("/"):] if "/" in path else "/"
            parts = [p for p in path.split("/") if p]
            if len(parts) == 2 and parts[0] in ("next", "finished"):
                job.arrive(int(parts[1]), returning=parts[0] == "finished")
                self.send_response(200)
            elif len(parts) == 2 and parts[0] == "reset":            # benchmark control: new epoch with N functions
                job.reset(int(parts[1]))
                self.send_response(200)
            elif len(parts) == 1 and parts[0] == "stats":            # benchmark control: merge timers (and reset)
                body = json.dumps(job.stats).encode()
                job.stats = {"merge_s": 0.0, "merges": 0, "barrier_wait_s": 0.0}
                self.send_response(200)
                self.send_header("Content-Length", str(len(body)))
                self.end_headers()
                self.wfile.write(body)
                return
            elif len(parts) == 1 and parts[0] == "layers":           # benchmark control: layer list from init
                n = int(self.headers.get("Content-Length") or 0)
                job.layers = json.loads(self.rfile.read(n).decode())
                self.send_response(200)
            else:
                self.send_response(404)
            self.send_header("Content-Length", "0")
            self.end_headers()

        def log_message(self, *a):
            pass
    return H


def serve_infrastructure(store_port: int, job_port: int, job_id: str) -> None:
    """Runs in its own process (like the RedisAI and job pods): tensor store + job endpoints."""
    store = TensorStore()
    srv = socketserver.ThreadingTCPServer(("127.0.0.1", store_port), _StoreHandler, bind_and_activate=False)
    srv.allow_reuse_address = True
    srv.daemon_threads = True
    srv.server_bind(); srv.server_activate()
    srv.store = store
    threading.Thread(target=srv.serve_forever, daemon=True).start()
    job = JobServer(store, job_id)
    http = ThreadingHTTPServer(("127.0.0.1", job_port), _make_job_http(job))
    http.daemon_threads = True
    http.serve_forever()


# ---------------------------------------------------------------------------------------------
def run_reference_arm(args) -> int:
    if not os.path.isdir(os.path.join(REF, "kubeml")) or not os.path.isfile(
            os.path.join(REF, "experiments", "function_resnet34.py")):
        return unavailable("baseline/_ref is missing: run baseline/install_reference.sh (offline pip install "
                           "--no-deps of /root/reference/python/kubeml)")
    try:
        import numpy as np
        import requests  # noqa: F401
        import torch
        import torch.distributed as dist
        import torchvision  # noqa: F401
        import PIL  # noqa: F401
    except ImportError as e:
        return unavailable(f"a dependency of the reference's stock path is missing: {e}")
    fn_name = os.environ.get("KUBEML_REF_FUNCTION", "function_resnet34")     # function_lenet: CPU plumbing test
    lenet = fn_name == "function_lenet"
    cuda = torch.cuda.is_available()
    if not cuda and not lenet:
        return unavailable("no CUDA device")

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", str(rank)))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            import subprocess
            port = os.environ.get("MASTER_PORT", "29519")
            return subprocess.call([sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
                                    f"--nproc-per-node={args.gpus}", "--master-addr", "127.0.0.1",
                                    "--master-port", port, os.path.join(ROOT, "bench.py")] + sys.argv[1:])
        return unavailable(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    N = world
    if cuda:
        torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("gloo")
    base_port = int(os.environ.get("MASTER_PORT", "29519")) + 101
    store_port, job_port = base_port, base_port + 1
    job_id = "refbench"

    infra = None
    if rank == 0:
        import multiprocessing as mp
        import socket as _socket
        infra = mp.get_context("spawn").Process(target=serve_infrastructure, args=(store_port, job_port, job_id),
                                                daemon=True)
        infra.start()
        for port in (store_port, job_port):
            for _ in range(600):
                try:
                    _socket.create_connection(("127.0.0.1", port), timeout=0.5).close()
                    break
                except OSError:
                    time.sleep(0.05)
    if world > 1:
        dist.barrier()

    # the environment contract of the reference (network.py:19-26, dataset.py:14-21, util.py:21-33)
    os.environ.update({"REDIS_URL": "127.0.0.1", "REDIS_PORT": str(store_port), "MONGO_IP": "127.0.0.1",
                       "MONGO_PORT": "27017", "GPU_ID": str(local),
                       "http_proxy": f"http://127.0.0.1:{job_port}", "HTTP_PROXY": f"http://127.0.0.1:{job_port}",
                       "no_proxy": "", "NO_PROXY": ""})
    sys.path.insert(0, SHIMS)
    sys.path.insert(0, REF)
    sys.path.insert(0, os.path.join(REF, "experiments"))
    import flask
    import pymongo
    import importlib
    fn = importlib.import_module(fn_name)    # the reference's own function, unmodified
    import kubeml as ref_kubeml
    assert os.path.realpath(ref_kubeml.__file__).startswith(os.path.realpath(REF)), ref_kubeml.__file__

    def make_docs(steps: int):
        n = N * steps * BATCH
        rng = np.random.default_rng(1234)
        x = rng.integers(0, 256, size=(n, 28, 28) if lenet else (n, 32, 32, 3), dtype=np.uint8)
        y = rng.integers(0, 10, size=(n,), dtype=np.int64)
        docs = [{"_id": i, "data": pickle.dumps(x[i * 64:(i + 1) * 64], pickle.HIGHEST_PROTOCOL),
                 "labels": pickle.dumps(y[i * 64:(i + 1) * 64], pickle.HIGHEST_PROTOCOL)}
                for i in range((n + 63) // 64)]
        return {"train": docs, "test": docs[:2]}

    def invoke(task: str, func_id: int, n: int, epoch: int):
        flask.set_request({"task": task, "jobId": job_id, "N": n, "K": K_AVG, "funcId": func_id,
                           "batchSize": BATCH, "lr": 0.1, "epoch": epoch})
        return fn.main()

    db_name = "mnist" if lenet else "cifar10"
    sync = torch.cuda.synchronize if cuda else (lambda: None)

    def run_epoch(steps: int, epoch: int) -> float:
        pymongo.DATABASES[db_name] = make_docs(steps)
        import requests as rq
        if rank == 0:
            rq.post(f"http://job-{job_id}.kubeml/reset/{N}")
        if world > 1:
            dist.barrier()
        sync()
        t0 = time.perf_counter()
        invoke("train", rank, N, epoch)
        # the deferred finish path of launchFunction (function.go:180-190)
        rq.post(f"http://job-{job_id}.kubeml/finished/{rank}")
        sync()
        if world > 1:
            dist.barrier()
        dt = time.perf_counter() - t0
        if world > 1:
            t = torch.tensor([dt], dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t.item())
        return dt

    # init (rank 0 only, like the job's invokeInitFunction), then the layer list for the merger
    pymongo.DATABASES[db_name] = make_docs(1)
    if rank == 0:
        import requests as rq
        layers, _ = invoke("init", 0, 1, 1)
        rq.post(f"http://job-{job_id}.kubeml/layers", data=json.dumps(list(layers)))
    if world > 1:
        dist.barrier()

    warm_steps = max(args.warmup, 3)
    run_epoch(warm_steps, 1)
    if os.environ.get("KUBEML_REF_PROFILE") and rank == 0:
        import cProfile
        import pstats
        pr = cProfile.Profile()
        pr.enable()
        run_epoch(8, 1)
        pr.disable()
        pstats.Stats(pr, stream=sys.stderr).sort_stats("cumulative").print_stats(35)
    from bench import ClockSampler
    import redisai as _store_shim
    import requests as rq
    sampler = ClockSampler(local).start() if rank == 0 else None
    for k in _store_shim.STATS:
        _store_shim.STATS[k] = 0
    if rank == 0:
        rq.post(f"http://job-{job_id}.kubeml/stats")
    dt = run_epoch(args.steps, 2)
    clocks = sampler.stop() if sampler else None
    if rank == 0:
        # where rank 0's wall time went (seconds over the whole timed region): tensor-store traffic of
        # __load_model / __save_model as seen by the client, the job's merge, and the rest (Mongo-document
        # unpickling, PIL transforms, H2D, forward/backward/step, loss.item(), waiting at /next)
        job_stats = rq.post(f"http://job-{job_id}.kubeml/stats").json()
        st = dict(_store_shim.STATS)
        rounds = max(1, (args.steps + K_AVG - 1) // K_AVG)
        breakdown = {
            "rounds": rounds, "timed_region_s": round(dt, 4),
            "load_model_s": round(st["get_s"], 4), "load_model_tensors": st["get_n"], "load_model_bytes": st["get_bytes"],
            "save_model_s": round(st["set_s"], 4), "save_model_tensors": st["set_n"], "save_model_bytes": st["set_bytes"],
            "job_merge_s": round(job_stats["merge_s"], 4), "job_merges": job_stats["merges"],
            "job_barrier_wait_s_all_ranks": round(job_stats["barrier_wait_s"], 4),
            "train_and_data_s": round(dt - st["get_s"] - st["set_s"] - job_stats["merge_s"], 4),
            "merge_note": "single Python thread under the store lock (restates the Go job's model mutex, "
                          "ml/pkg/model/model.go:273-274)"}
        imgs = N * args.steps * BATCH
        value = imgs / dt
        print(json.dumps({
            "metric": "images/sec ResNet-34 K-AVG (K=8, batch 128/worker)" if not lenet else "images/sec LeNet (plumbing test)",
            "impl": "reference",
            "value": round(value, 1), "unit": "images/s", "n_gpus": N, "steps": args.steps, "warmup": warm_steps,
            "ms_per_step": round(1e3 * dt / args.steps, 3), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": round(value / BASELINE_IMG_S, 2), "dtype": "fp32 (the reference's stock path)",
            "data": "synthetic",
            "config": {"model": "torchvision resnet34() 1000 logits on 32x32x3 (reference function_resnet34.py)",
                       "global_batch": BATCH * N, "seq_len": None, "parallelism": f"kavg-dp{N}", "k": K_AVG,
                       "timing": "wall clock around the train fan-out incl. the final merge, max over ranks "
                                 "(the reference has no device timers); includes its per-batch loss.item()",
                       "stand_ins": "flask/redisai/pymongo shims + Python restatement of the Go job's merge"},
            "breakdown": breakdown,
            "e2e": {"value": round(value, 1), "unit": "images/s",
                    "note": "the reference only has an end-to-end path (CPU transforms, pageable H2D, .item())"},
            "clocks": clocks}), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if infra is not None:
        infra.terminate()
    return 0
