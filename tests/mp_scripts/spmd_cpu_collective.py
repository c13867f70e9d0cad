"""Run under torchrun on the CPU (gloo): the SPMD invoker + CollectiveSync K-AVG path."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
import torch


def main():
    home = sys.argv[1]
    os.environ["KUBEML_HOME"] = home
    os.environ["KUBEML_CUDA_GRAPH"] = "0"
    torch.set_num_threads(1)
    from kubeml_b200.api.types import JobInfo, JobState, TrainOptions, TrainRequest, TrainTask
    from kubeml_b200.runtime.spmd import SpmdRuntime
    from kubeml_b200.storage.store import ShardStore
    from kubeml_b200.train.job import TrainJob
    from kubeml_b200.train.sync import CollectiveSync
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    store = ShardStore(os.path.join(home, "datasets"))
    rt = SpmdRuntime(shard_store=store, use_device_sync=False)
    rt.services.device_sync = None
    rt.services.options["sync_factory"] = lambda: CollectiveSync(rt.tensor_store)
    if rank == 0 and not store.exists("mnist"):
        rng = np.random.default_rng(0)
        store.create("mnist", rng.integers(0, 255, (64 * 8, 28, 28), dtype=np.uint8), rng.integers(0, 10, (64 * 8,)),
                     rng.integers(0, 255, (128, 28, 28), dtype=np.uint8), rng.integers(0, 10, (128,)))
    torch.distributed.barrier()
    if rank != 0:
        rt.serve(); return
    req = TrainRequest(model_type="lenet", batch_size=64, epochs=2, dataset="mnist", lr=0.01, function_name="lenet",
                       options=TrainOptions(default_parallelism=world, static_parallelism=True, validate_every=1, k=2))
    job = TrainJob(TrainTask(request=req, job=JobInfo(id="gloojob", state=JobState(parallelism=world))),
                   rt.invoker(), rt.tensor_store, device_mode=True)
    try:
        job.train()
    finally:
        rt.stop()
    print("RESULT " + json.dumps({"error": None if job.exit_err is None else str(job.exit_err),
                                  "history": job.history.to_dict()}), flush=True)


if __name__ == "__main__":
    main()
