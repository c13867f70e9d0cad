"""Control plane + one worker process per GPU, launched by torchrun:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
        -m kubeml_b200.runtime.spmd_main [--port 10100]

Rank 0 hosts the Cluster (controller, scheduler, parameter server, storage; HTTP on --port) whose
jobs fan out over the ranks through `SpmdInvoker`; the other ranks are warm workers that execute what
rank 0 broadcasts (the Fission `torch` environment pool of the reference, ml/charts/kubeml/values.yaml:24)."""
from __future__ import annotations

import argparse
import logging
import os
import time


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--host", default="127.0.0.1")
    ap.add_argument("--port", type=int, default=int(os.environ.get("KUBEML_CONTROLLER_PORT", "10100")))
    args = ap.parse_args()
    logging.basicConfig(level=logging.INFO, format="[%(asctime)s] %(levelname)-7s %(name)s: %(message)s")
    from ..controller.cluster import Cluster
    from ..storage.store import default_store
    from .spmd import SpmdRuntime

    rt = SpmdRuntime(shard_store=default_store())
    if rt.rank != 0:
        rt.serve()
        rt.close()
        return 0
    cluster = Cluster(max_parallelism=rt.world, invoker_factory=lambda c, task: rt.invoker(),
                      tensor_store=rt.tensor_store)
    rt.services.shard_store = cluster.shard_store
    cluster.serve(host=args.host, controller_port=args.port)
    print(f"kubeml control plane on http://{args.host}:{cluster.controller_port} with {rt.world} GPU workers", flush=True)
    try:
        while True:
            time.sleep(3600)
    except KeyboardInterrupt:
        pass
    finally:
        cluster.shutdown()
        rt.stop()
        rt.close()
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
