"""Run under torchrun: LeNet K-AVG job on N GPUs through the SPMD runtime with device-resident
weights; rank 0 prints the history as JSON.  Used by tests/test_multigpu.py."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
import torch


def main():
    home = sys.argv[1]
    epochs = int(sys.argv[2]) if len(sys.argv) > 2 else 2
    os.environ["KUBEML_HOME"] = home
    from kubeml_b200.api.types import JobInfo, JobState, TrainOptions, TrainRequest, TrainTask
    from kubeml_b200.runtime.spmd import SpmdRuntime
    from kubeml_b200.storage.store import ShardStore
    from kubeml_b200.train.job import TrainJob
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    store = ShardStore(os.path.join(home, "datasets"))
    rt = SpmdRuntime(shard_store=store)
    if rank == 0 and not store.exists("mnist"):
        rng = np.random.default_rng(0)
        store.create("mnist", rng.integers(0, 255, (1024 + 64, 28, 28), dtype=np.uint8), rng.integers(0, 10, (1024 + 64,)),
                     rng.integers(0, 255, (256, 28, 28), dtype=np.uint8), rng.integers(0, 10, (256,)))
    if world > 1:
        torch.distributed.barrier()
    if rank != 0:
        rt.serve(); rt.close(); return
    req = TrainRequest(model_type="lenet", batch_size=64, epochs=epochs, dataset="mnist", lr=0.01, function_name="lenet",
                       options=TrainOptions(default_parallelism=world, static_parallelism=True, validate_every=1, k=2))
    task = TrainTask(request=req, job=JobInfo(id="spmdtest", state=JobState(parallelism=world)))
    job = TrainJob(task, rt.invoker(), rt.tensor_store, device_mode=rt.group is not None)
    job.train()
    rt.stop()
    print("RESULT " + json.dumps({"error": None if job.exit_err is None else str(job.exit_err),
                                  "history": job.history.to_dict(), "stats": job.epoch_stats,
                                  "nvls": bool(rt.group and rt.group.rs.peers.mcast)}), flush=True)
    rt.close()


if __name__ == "__main__":
    main()
