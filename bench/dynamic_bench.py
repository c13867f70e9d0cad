"""ResNet-34 with dynamic parallelism.  Run under torchrun with 8 GPUs:

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 \
      bench/dynamic_bench.py [epochs] [images]

The job starts on 2 workers; after every epoch it asks the scheduler, whose throughput policy
(ml/pkg/scheduler/policy.go:50-94, clamped to the GPU count here) grows the group while epochs keep
getting faster.  Joining ranks map the peers' arenas live and start from the merged model.  Prints one
JSON line with per-epoch parallelism, device-timed train seconds and images/s."""
import json, os, sys, tempfile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch


def main():
    epochs = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    images = int(sys.argv[2]) if len(sys.argv) > 2 else 840 * 64        # divisible by 2..8 workers x 64-sample subsets
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    home = os.path.join(tempfile.gettempdir(), f"kubeml_dyn_{os.getuid()}_{os.environ.get('MASTER_PORT', '0')}")
    os.environ["KUBEML_HOME"] = home
    os.environ["KUBEML_RESNET_DATASET"] = "cifar10_dyn"
    from kubeml_b200.api.types import JobInfo, JobState, TrainOptions, TrainRequest, TrainTask
    from kubeml_b200.runtime.spmd import SpmdRuntime
    from kubeml_b200.scheduler.scheduler import Scheduler
    from kubeml_b200.storage.store import ShardStore
    from kubeml_b200.train.job import TrainJob
    store = ShardStore(os.path.join(home, "datasets"))
    rt = SpmdRuntime(shard_store=store)
    if rank == 0 and not store.exists("cifar10_dyn"):
        rng = np.random.default_rng(0)
        x = rng.integers(0, 256, (images, 32, 32, 3), dtype=np.uint8)
        y = rng.integers(0, 10, (images,), dtype=np.int64)
        store.create("cifar10_dyn", x, y, x[:1024], y[:1024])
    torch.distributed.barrier()
    store.pinned("cifar10_dyn", "train")
    if rank != 0:
        rt.serve(); rt.close(); return
    req = TrainRequest(model_type="resnet34", batch_size=128, epochs=epochs, dataset="cifar10_dyn", lr=0.05,
                       function_name="resnet34",
                       options=TrainOptions(default_parallelism=2, static_parallelism=False, validate_every=0, k=8,
                                            goal_accuracy=101.0))
    task = TrainTask(request=req, job=JobInfo(id="dynamic", state=JobState(parallelism=2)))

    class PS:
        def __init__(self): self.job = None
        def start_task(self, t): pass
        def update_task(self, t): self.job.update_task(JobState(parallelism=t.job.state.parallelism, elapsed_time=t.job.state.elapsed_time))
        def update_metrics(self, *a): pass
        def job_finished(self, *a): pass
    ps = PS()
    sched = Scheduler(ps=ps, max_parallelism=world).start()
    sched.policy.calculate_parallelism(task)
    inv = rt.invoker()
    inv.env = {"KUBEML_RESNET_DATASET": "cifar10_dyn"}
    job = TrainJob(task, inv, rt.tensor_store, scheduler=sched, ps=ps, device_mode=True, skip_final_validation=True)
    ps.job = job
    try:
        job.train()
    finally:
        sched.stop()
        rt.stop()
    rows = []
    for st, n in zip(job.epoch_stats, job.history.parallelism):
        imgs = images
        rows.append({"parallelism": int(n), "steps_per_worker": int(st["steps"]),
                     "train_seconds": round(st["train_seconds"], 4),
                     "images_per_s": round(imgs / max(st["train_seconds"], 1e-9), 1),
                     "merge_ms_per_sync": round(1e3 * st.get("merge_seconds", 0.0) / max(int(st["steps"]) // 8, 1), 3)})
    print(json.dumps({"bench": "dynamic parallelism 2->%d, ResNet-34 K=8 batch 128" % world, "images_per_epoch": images,
                      "error": None if job.exit_err is None else str(job.exit_err),
                      "epoch_wall_seconds": [round(v, 3) for v in job.history.epoch_duration],
                      "train_loss": [round(v, 4) for v in job.history.train_loss], "epochs": rows}), flush=True)
    rt.close()


if __name__ == "__main__":
    main()
