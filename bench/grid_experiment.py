"""The reference's experiment grid (ml/experiments/train.py:17-61,76-96: batch x K x N, static parallelism,
validation every epoch, time-to-accuracy read from the History, ml/experiments/insights.ipynb cells 3-4) run for
real on the GPUs of this box through the experiment harness: controller client -> scheduler -> parameter server ->
train job -> process-per-GPU workers.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29537 \
        bench/grid_experiment.py [epochs]

Data: CIFAR-shaped images with teacher labels (bench/loss_parity.make_data) so that accuracy is meaningful; the grid
is batch {128} x K {8, -1 (one merge per epoch)} x N {1, 2, 4, ... <= GPUs}.  Prints one JSON line with, per grid
point, total time, final accuracy and time-to-accuracy for two definitions the reference's notebooks use
(`tta`: first epoch at or above the goal; `tta_crossbow`: median of the last 5 epochs at or above it)."""
import importlib.util, json, os, sys, tempfile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch


def main():
    epochs = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    goal = float(sys.argv[2]) if len(sys.argv) > 2 else 30.0
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    home = os.path.join(tempfile.gettempdir(), f"kubeml_grid_{os.getuid()}_{os.environ.get('MASTER_PORT', '0')}")
    os.environ["KUBEML_HOME"] = home
    os.environ["KUBEML_RESNET_DATASET"] = "teacher"
    os.environ["KUBEML_RESNET_NOAUG"] = "1"
    # bench.py at the repository root shadows the bench/ directory as a module name: load the file itself
    spec = importlib.util.spec_from_file_location(
        "kubeml_bench_loss_parity", os.path.join(os.path.dirname(os.path.abspath(__file__)), "loss_parity.py"))
    lp = importlib.util.module_from_spec(spec); spec.loader.exec_module(lp)
    make_data = lp.make_data
    from kubeml_b200.controller.client import make_kubeml_client
    from kubeml_b200.controller.cluster import Cluster
    from kubeml_b200.experiments import analysis, harness
    from kubeml_b200.runtime.spmd import SpmdRuntime
    from kubeml_b200.storage.store import default_store
    rt = SpmdRuntime(shard_store=default_store())
    if rank != 0:
        rt.serve(); rt.close(); return
    cluster = Cluster(home=home, max_parallelism=world, invoker_factory=lambda c, task: _invoker(rt),
                      tensor_store=rt.tensor_store)
    rt.services.shard_store = cluster.shard_store
    if not cluster.shard_store.exists("teacher"):
        xtr, ytr, xte, yte = make_data(64 * 128, 2048)
        cluster.shard_store.create("teacher", xtr, ytr, xte, yte)
    client = make_kubeml_client(cluster)
    ns = [n for n in (1, 2, 4, 8) if n <= world]
    grid = {"batch": [128], "k": [8, -1], "parallelism": ns}
    recs = harness.run_grid(client, "resnet34", "teacher", 0.1, grid, os.path.join(home, "grid"), epochs=epochs,
                            validate_every=1, timeout=900)
    rows = []
    for r in recs:
        h = r.history.get("data", {}) if r.history else {}
        acc, dur = h.get("accuracy", []), h.get("epoch_duration", [])
        # the final validation has no epoch of its own: pair accuracies with the epoch durations they follow
        n = min(len(acc), len(dur))
        rows.append({"title": r.title, "error": r.error, "k": r.request["options"]["k"],
                     "parallelism": r.request["options"]["default_parallelism"],
                     "total_time_s": round(analysis.total_time(dur), 3) if dur else None,
                     "final_accuracy": round(acc[-1], 2) if acc else None,
                     "tta_s": _num(analysis.tta(acc[:n], dur[:n], goal)),
                     "tta_crossbow_s": _num(analysis.tta_crossbow(acc[:n], dur[:n], goal)),
                     "train_loss": [round(v, 3) for v in h.get("train_loss", [])]})
    cluster.shutdown()
    rt.stop()
    print(json.dumps({"bench": f"experiment grid, ResNet-34, teacher-labelled CIFAR-shape data, {epochs} epochs, "
                               f"time to {goal:.0f} % accuracy", "gpus": world, "rows": rows}), flush=True)
    rt.close()


def _num(v):
    return None if v != v else round(float(v), 3)       # NaN (goal never reached) -> null


def _invoker(rt):
    inv = rt.invoker()
    inv.env = {"KUBEML_RESNET_DATASET": "teacher", "KUBEML_RESNET_NOAUG": "1", "KUBEML_RESNET_IMPL": "native"}
    return inv


if __name__ == "__main__":
    main()
