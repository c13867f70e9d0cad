"""Do the three ways of taking a round's last SGD step move the weights the same way?  Five short ResNet-34 jobs from
one initial state_dict: partial deferral (default) twice, full deferral twice, no deferral once; prints, per layer,
the relative difference of the weight movement between every pair of runs (run-to-run noise vs a real difference)."""
import json, os, sys, tempfile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch


def main():
    from kubeml_b200.api import TrainOptions, TrainRequest
    from kubeml_b200.controller import Cluster, make_kubeml_client
    from kubeml_b200.models import resnet34
    from kubeml_b200.utils.checkpoint import load_checkpoint
    lr = float(sys.argv[1]) if len(sys.argv) > 1 else 0.005
    tmp = tempfile.mkdtemp(prefix="kubeml_pd_")
    os.environ["KUBEML_RESNET_CLASSES"] = "10"
    os.environ["KUBEML_RESNET_NOAUG"] = "1"
    torch.manual_seed(31)
    init = {k: v.clone() for k, v in resnet34(num_classes=10).state_dict().items()}
    torch.save(init, os.path.join(tmp, "init.pt"))
    os.environ["KUBEML_RESNET_INIT"] = os.path.join(tmp, "init.pt")
    req = TrainRequest(model_type="resnet34", batch_size=32, epochs=1, dataset="cifar10", lr=lr, function_name="resnet34",
                       options=TrainOptions(default_parallelism=1, static_parallelism=True, validate_every=0, k=4))
    runs = [("partial_a", {"KUBEML_PARTIAL_DEFER": "1"}), ("partial_b", {"KUBEML_PARTIAL_DEFER": "1"}),
            ("full_a", {"KUBEML_PARTIAL_DEFER": "0"}), ("full_b", {"KUBEML_PARTIAL_DEFER": "0"}),
            ("unfused", {"KUBEML_FUSE_LAST_STEP": "0"}), ("eager_unfused", {"KUBEML_FUSE_LAST_STEP": "0", "KUBEML_CUDA_GRAPH": "0"})]
    moved, info = {}, {}
    for name, env in runs:
        for k in ("KUBEML_PARTIAL_DEFER", "KUBEML_FUSE_LAST_STEP", "KUBEML_CUDA_GRAPH"):
            os.environ.pop(k, None)
        os.environ.update(env)
        home = os.path.join(tmp, name)
        os.environ["KUBEML_HOME"] = home
        c = Cluster(home=home, max_parallelism=1)
        rng = np.random.default_rng(0)
        c.shard_store.create("cifar10", rng.integers(0, 255, (512, 32, 32, 3), dtype=np.uint8), rng.integers(0, 10, (512,)),
                             rng.integers(0, 255, (64, 32, 32, 3), dtype=np.uint8), rng.integers(0, 10, (64,)))
        jid = make_kubeml_client(c).v1().networks().train(req)
        err = c.wait(jid, 300)
        job = c.jobs[jid]
        sd, _ = load_checkpoint(c.checkpoint_dir, jid)
        moved[name] = {k: torch.as_tensor(sd[k]).float().cpu() - init[k].float().cpu() for k in init
                       if k.endswith("weight") and init[k].dim() >= 2}
        info[name] = {"error": err, "train_loss": job.history.train_loss, "fused": [e["fused_last_steps"] for e in job.epoch_stats]}
        c.shutdown()
    layers = ["conv1.weight", "layer1.0.conv1.weight", "layer2.1.conv2.weight", "layer3.2.conv2.weight",
              "layer4.0.conv1.weight", "layer4.2.conv2.weight", "fc.weight"]
    names = [n for n, _ in runs]
    out = {"lr": lr, "runs": info, "relative_difference_of_movement": {},
           "movement_norm": {k: {n: round(float(moved[n][k].norm()), 6) for n in names} for k in layers}}
    for k in layers:
        row = {}
        for i, a in enumerate(names):
            for b in names[i + 1:]:
                d = float((moved[a][k] - moved[b][k]).norm() / moved[b][k].norm().clamp_min(1e-12))
                row[f"{a}|{b}"] = round(d, 4)
        out["relative_difference_of_movement"][k] = row
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
