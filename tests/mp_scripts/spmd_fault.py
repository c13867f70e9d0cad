"""Run under torchrun: LeNet K-AVG job on N GPUs where one worker fails (or stalls) in the middle of an epoch.
Rank 0 prints the history and what every rank's merge engine saw.  Used by tests/test_multigpu.py."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
import torch


def main():
    home, spec = sys.argv[1], sys.argv[2]
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
        n = world * 8 * 64
        store.create("mnist", rng.integers(0, 255, (n, 28, 28), dtype=np.uint8), rng.integers(0, 10, (n,)),
                     rng.integers(0, 255, (256, 28, 28), dtype=np.uint8), rng.integers(0, 10, (256,)))
    if world > 1:
        torch.distributed.barrier()
    if rank != 0:
        rt.serve(); rt.close(); return
    req = TrainRequest(model_type="lenet", batch_size=64, epochs=2, dataset="mnist", lr=0.01, function_name="lenet",
                       options=TrainOptions(default_parallelism=world, static_parallelism=True, validate_every=0, k=2))
    task = TrainTask(request=req, job=JobInfo(id="faulttest", state=JobState(parallelism=world)))
    inv = rt.invoker()
    inv.options = {"inject_fault": spec}
    job = TrainJob(task, inv, rt.tensor_store, device_mode=rt.group is not None, skip_final_validation=True)
    job.release_on_finish = False
    job.train()
    rs = rt.group.engine_for("faulttest", 0) if rt.group is not None else None
    rt.stop()
    print("RESULT " + json.dumps({"error": None if job.exit_err is None else str(job.exit_err),
                                  "history": job.history.to_dict(), "stats": job.epoch_stats,
                                  "recovered": getattr(rs, "recovered", None), "merges": getattr(rs, "merges", None)}),
          flush=True)
    rt.close()


if __name__ == "__main__":
    main()
