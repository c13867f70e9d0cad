"""Run under torchrun with 4 GPUs: two independent LeNet K-AVG jobs at the same time, each on its own pair of
GPUs (ranks [0,1] and [2,3]), driven by two threads of rank 0.  Used by tests/test_multigpu.py."""
import json, os, sys, threading, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
import torch


def main():
    home = sys.argv[1]
    os.environ["KUBEML_HOME"] = home
    from kubeml_b200.api.types import JobInfo, JobState, TrainOptions, TrainRequest, TrainTask
    from kubeml_b200.runtime.spmd import SpmdRuntime
    from kubeml_b200.storage.store import ShardStore
    from kubeml_b200.train.job import TrainJob
    rank = int(os.environ.get("RANK", "0"))
    store = ShardStore(os.path.join(home, "datasets"))
    rt = SpmdRuntime(shard_store=store)
    if rank == 0 and not store.exists("mnist"):
        rng = np.random.default_rng(0)
        store.create("mnist", rng.integers(0, 255, (2048, 28, 28), dtype=np.uint8), rng.integers(0, 10, (2048,)),
                     rng.integers(0, 255, (256, 28, 28), dtype=np.uint8), rng.integers(0, 10, (256,)))
    torch.distributed.barrier()
    if rank != 0:
        rt.serve(); rt.close(); return
    out, spans = {}, {}

    def run(name, lr):
        ranks = rt.allocate(2)
        req = TrainRequest(model_type="lenet", batch_size=64, epochs=3, dataset="mnist", lr=lr, function_name="lenet",
                           options=TrainOptions(default_parallelism=2, static_parallelism=True, validate_every=1, k=2))
        task = TrainTask(request=req, job=JobInfo(id=name, state=JobState(parallelism=2)))
        job = TrainJob(task, rt.invoker(ranks), rt.tensor_store, device_mode=True)
        t0 = time.time()
        job.train()
        spans[name] = (t0, time.time())
        rt.free(ranks)
        out[name] = {"ranks": ranks, "error": None if job.exit_err is None else str(job.exit_err),
                     "train_loss": job.history.train_loss, "accuracy": job.history.accuracy}

    threads = [threading.Thread(target=run, args=("jobA", 0.01)), threading.Thread(target=run, args=("jobB", 0.02))]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    a, b = spans["jobA"], spans["jobB"]
    out["overlap_s"] = max(0.0, min(a[1], b[1]) - max(a[0], b[0]))
    rt.stop()
    print("RESULT " + json.dumps(out), flush=True)
    rt.close()


if __name__ == "__main__":
    main()
