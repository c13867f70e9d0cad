"""Run under torchrun with >= 2 GPUs: dynamic parallelism through the scheduler policy over
process-per-GPU workers (BASELINE.md config 4 in miniature: N grows between epochs, the new rank
joins the peer group live and starts from the merged model)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
import torch


def main():
    home = sys.argv[1]
    os.environ["KUBEML_HOME"] = home
    from kubeml_b200.api.types import JobInfo, JobState, TrainOptions, TrainRequest, TrainTask
    from kubeml_b200.runtime.spmd import SpmdRuntime
    from kubeml_b200.scheduler.scheduler import Scheduler
    from kubeml_b200.storage.store import ShardStore
    from kubeml_b200.train.job import TrainJob
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    store = ShardStore(os.path.join(home, "datasets"))
    rt = SpmdRuntime(shard_store=store)
    if rank == 0 and not store.exists("mnist"):
        rng = np.random.default_rng(0)
        store.create("mnist", rng.integers(0, 255, (64 * 16, 28, 28), dtype=np.uint8), rng.integers(0, 10, (64 * 16,)),
                     rng.integers(0, 255, (256, 28, 28), dtype=np.uint8), rng.integers(0, 10, (256,)))
    torch.distributed.barrier()
    if rank != 0:
        rt.serve(); rt.close(); return
    req = TrainRequest(model_type="lenet", batch_size=64, epochs=3, dataset="mnist", lr=0.01, function_name="lenet",
                       options=TrainOptions(default_parallelism=1, static_parallelism=False, validate_every=1, k=2))
    task = TrainTask(request=req, job=JobInfo(id="elastic", state=JobState(parallelism=1)))

    class PS:        # the scheduler talks to the job through the parameter server's update_task
        def __init__(self): self.job = None
        def start_task(self, t): pass
        def update_task(self, t): self.job.update_task(JobState(parallelism=t.job.state.parallelism, elapsed_time=t.job.state.elapsed_time))
        def update_metrics(self, *a): pass
        def job_finished(self, *a): pass
    ps = PS()
    sched = Scheduler(ps=ps, max_parallelism=world).start()
    sched.policy.calculate_parallelism(task)          # first sight: CREATE with the default parallelism
    job = TrainJob(task, rt.invoker(), rt.tensor_store, scheduler=sched, ps=ps, device_mode=True)
    ps.job = job
    try:
        job.train()
    finally:
        sched.stop()
        rt.stop()
    print("RESULT " + json.dumps({"error": None if job.exit_err is None else str(job.exit_err),
                                  "history": job.history.to_dict()}), flush=True)
    rt.close()


if __name__ == "__main__":
    main()
