"""`kubeml` command line (reference: ml/pkg/kubeml-cli/cmd/*.go, cobra).

    kubeml serve [--port P]                      start the control plane on this host
    kubeml train -d DATASET -f FUNCTION -e EPOCHS -b BATCH --lr LR
                 [--validate-every N] [--parallelism N] [--static] [--K K] [--sparse-avg]
                 [--goal-accuracy A] [--local] [--wait]
    kubeml infer --id JOB --datafile FILE.json
    kubeml dataset create --name N --traindata X --trainlabels Y --testdata X --testlabels Y
    kubeml dataset delete --name N | dataset list
    kubeml function create --name N --code FILE.py | function delete --name N | function list
    kubeml history get --id ID | history delete --id ID | history list | history prune
    kubeml task list [--short] | task stop --id ID | task prune
    kubeml logs --id ID [-f]

Flag names, defaults and validation follow the reference (`train`: epochs 1, batch 64, lr 0.01,
validate-every 0, parallelism 2, static false, K -1, goal-accuracy 100; 0 < batch <= 1024;
`--sparse-avg` forces K = -1; cmd/train.go:44-119,150-172).  The drifted spellings used by the
reference's own experiment harness are accepted as aliases (`--default-parallelism`,
`history get --network`).  `--local` runs the whole control plane inside this process
instead of talking to `kubeml serve`.
"""
from __future__ import annotations

import argparse
import json
import math
import os
import sys
import time
from typing import List, Optional

from ..api import const
from ..api.types import InferRequest, TrainOptions, TrainRequest
from ..exceptions import KubeMLException


def _client(args):
    from ..controller.client import make_kubeml_client
    if getattr(args, "local", False):
        from ..controller.cluster import Cluster
        cluster = Cluster()
        args._cluster = cluster
        return make_kubeml_client(cluster)
    return make_kubeml_client(url=getattr(args, "url", None))


def _table(header: List[str], rows: List[List]) -> str:
    cols = [header] + [[str(c) for c in r] for r in rows]
    widths = [max(len(r[i]) for r in cols) for i in range(len(header))]
    return "\n".join("  ".join(c.ljust(w) for c, w in zip(r, widths)).rstrip() for r in cols)


def _last(xs) -> float:
    return xs[-1] if xs else math.nan


# ---------------------------------------------------------------------------------------- commands
def cmd_serve(args) -> int:
    from ..controller.cluster import Cluster
    import logging
    logging.basicConfig(level=logging.INFO, format="[%(asctime)s] %(levelname)-7s %(name)s: %(message)s")
    log_dir = os.path.join(const.home(), "logs")
    os.makedirs(log_dir, exist_ok=True)
    cluster = Cluster().serve(host=args.host, controller_port=args.port,
                              scheduler_port=args.scheduler_port, ps_port=args.ps_port,
                              storage_port=args.storage_port)
    print(f"kubeml control plane listening on http://{args.host}:{cluster.controller_port} "
          f"(workers: up to {cluster.max_parallelism})", flush=True)
    try:
        while True:
            time.sleep(3600)
    except KeyboardInterrupt:
        cluster.shutdown()
    return 0


def cmd_train(args) -> int:
    if not (0 < args.batch <= 1024):
        raise KubeMLException("batch size must be in (0, 1024]", 400)
    if args.epochs <= 0:
        raise KubeMLException("epochs must be positive", 400)
    if args.lr <= 0:
        raise KubeMLException("learning rate must be positive", 400)
    K = -1 if args.sparse_avg else args.K
    c = _client(args)
    c.v1().datasets().get(args.dataset)                   # 404 if the dataset is missing
    if not c.v1().functions().exists(args.function):
        raise KubeMLException(f"function {args.function} does not exist", 404)
    req = TrainRequest(model_type=args.function, batch_size=args.batch, epochs=args.epochs,
                       dataset=args.dataset, lr=args.lr, function_name=args.function,
                       options=TrainOptions(default_parallelism=args.parallelism,
                                            static_parallelism=args.static,
                                            validate_every=args.validate_every, k=K,
                                            goal_accuracy=args.goal_accuracy,
                                            resume_from=getattr(args, "resume", "") or ""))
    job_id = c.v1().networks().train(req)
    print(job_id)
    if getattr(args, "local", False) or args.wait:
        cluster = getattr(args, "_cluster", None)
        if cluster is not None:
            err = cluster.wait(job_id)
            if err:
                print(f"job failed: {err}", file=sys.stderr)
                return 1
            print(c.v1().histories().get(job_id).to_json())
        else:
            # The task appears in `task list` only once the scheduler has handed it to the parameter
            # server (its queue is polled every 10 ms), so "not listed" right after submission does not
            # mean "finished": the history record is what marks the end.
            seen, deadline, hist = False, time.monotonic() + 120.0, None
            while hist is None:
                if any(t.job.id == job_id for t in c.v1().tasks().list()):
                    seen = True
                else:
                    try:
                        hist = c.v1().histories().get(job_id)
                        break
                    except KubeMLException as e:
                        if e.status_code != 404:
                            raise
                        if seen:
                            time.sleep(0.2)         # the job may be between de-registration and its history write
                            try:
                                hist = c.v1().histories().get(job_id)
                                break
                            except KubeMLException:
                                print(f"job {job_id} ended without a history record (see `kubeml logs --id {job_id}`)",
                                      file=sys.stderr)
                                return 1
                        if time.monotonic() > deadline:
                            print(f"job {job_id} never started", file=sys.stderr)
                            return 1
                time.sleep(0.1 if not seen else 0.5)
            print(hist.to_json())
    return 0


def cmd_infer(args) -> int:
    with open(args.datafile) as f:
        data = json.load(f)
    print(json.dumps(_client(args).v1().networks().infer(InferRequest(model_id=args.id, data=data))))
    return 0


def cmd_dataset(args) -> int:
    ds = _client(args).v1().datasets()
    if args.action == "create":
        ds.create(args.name, args.traindata, args.trainlabels, args.testdata, args.testlabels)
        print(f"Dataset {args.name} created")
    elif args.action == "delete":
        ds.delete(args.name)
        print(f"Dataset {args.name} deleted")
    else:
        print(_table(["NAME", "TRAINSET", "TESTSET"],
                     [[d.name, d.train_set_size, d.test_set_size] for d in ds.list()]))
    return 0


def cmd_function(args) -> int:
    fns = _client(args).v1().functions()
    if args.action == "create":
        fns.create(args.name, args.code)
        print(f"Function {args.name} created")
    elif args.action == "delete":
        fns.delete(args.name)
        print(f"Function {args.name} deleted")
    else:
        got = fns.list()
        print(_table(["NAME", "KIND"], [[n, "user"] for n in got["functions"]] +
                     [[n, "builtin"] for n in got["builtin"]]))
    return 0


def cmd_history(args) -> int:
    hs = _client(args).v1().histories()
    if args.action == "get":
        print(json.dumps(hs.get(args.id).to_dict(), indent=2))
    elif args.action == "delete":
        hs.delete(args.id)
        print(f"History {args.id} deleted")
    elif args.action == "prune":
        hs.prune()
        print("Deleted all histories")
    else:
        rows = []
        for h in hs.list():
            par = h.data.parallelism
            rows.append([h.id, h.task.model_type, h.task.dataset, h.task.epochs, h.task.batch_size, h.task.lr,
                         round(sum(par) / len(par), 2) if par else math.nan, h.task.options.k,
                         h.task.options.static_parallelism, _last(h.data.accuracy),
                         _last(h.data.validation_loss), _last(h.data.epoch_duration)])
        print(_table(["NAME", "MODEL", "DATASET", "EPOCHS", "BATCH", "LR", "PARALLELISM", "K", "STATIC",
                      "ACCURACY", "LOSS", "TIME (s)"], rows))
    return 0


def cmd_task(args) -> int:
    ts = _client(args).v1().tasks()
    if args.action == "stop":
        ts.stop(args.id)
        print(f"Task {args.id} stopping")
    elif args.action == "prune":
        # job pods/services do not exist here; finished jobs leave only their log file behind
        log_dir = os.path.join(const.home(), "logs")
        running = {t.job.id for t in ts.list()}
        n = 0
        if os.path.isdir(log_dir):
            for fn in os.listdir(log_dir):
                if fn.startswith("job-") and fn[4:-4] not in running:
                    os.remove(os.path.join(log_dir, fn))
                    n += 1
        print(f"Deleted {n} finished task logs")
    else:
        tasks = ts.list()
        if args.short:
            for t in tasks:
                print(t.job.id)
        else:
            print(_table(["NAME", "MODEL", "DATASET", "EPOCHS", "BATCH", "LR", "PARALLELISM", "K", "STATIC"],
                         [[t.job.id, t.request.model_type, t.request.dataset, t.request.epochs,
                           t.request.batch_size, t.request.lr, t.job.state.parallelism, t.request.options.k,
                           t.request.options.static_parallelism] for t in tasks]))
    return 0


def cmd_logs(args) -> int:
    """`kubectl logs job-<id> [-f]` (cmd/log.go:28-65) -> the job's log file."""
    path = os.path.join(const.home(), "logs", f"job-{args.id}.log")
    if not os.path.isfile(path):
        raise KubeMLException(f"no logs for task {args.id}", 404)
    with open(path) as f:
        sys.stdout.write(f.read())
        while args.follow:
            line = f.readline()
            if line:
                sys.stdout.write(line)
                sys.stdout.flush()
            else:
                time.sleep(0.5)
    return 0


# ---------------------------------------------------------------------------------------- parser
def build_parser() -> argparse.ArgumentParser:
    ap = argparse.ArgumentParser(prog="kubeml", description="kubeml_b200 command line")
    ap.add_argument("--url", default=None, help="controller URL (default $KUBEML_CONTROLLER_URL or 127.0.0.1:10100)")
    ap.add_argument("--local", action="store_true", help="run the control plane inside this process")
    sub = ap.add_subparsers(dest="cmd", required=True)

    p = sub.add_parser("serve", help="start controller + scheduler + parameter server + storage")
    p.add_argument("--host", default=const.HOST)
    p.add_argument("--port", type=int, default=const.CONTROLLER_PORT)
    p.add_argument("--scheduler-port", type=int, default=None)
    p.add_argument("--ps-port", type=int, default=None)
    p.add_argument("--storage-port", type=int, default=None)
    p.set_defaults(fn=cmd_serve)

    p = sub.add_parser("train", help="create a train task")
    p.add_argument("-d", "--dataset", required=True)
    p.add_argument("-f", "--function", required=True)
    p.add_argument("-e", "--epochs", type=int, default=1)
    p.add_argument("-b", "--batch", type=int, default=64)
    p.add_argument("--lr", type=float, default=0.01)
    p.add_argument("--validate-every", type=int, default=0)
    p.add_argument("--parallelism", "--default-parallelism", dest="parallelism", type=int,
                   default=const.DEBUG_PARALLELISM)
    p.add_argument("--static", action="store_true")
    p.add_argument("--K", "-K", dest="K", type=int, default=-1)
    p.add_argument("--sparse-avg", action="store_true")
    p.add_argument("--goal-accuracy", type=float, default=100.0)
    p.add_argument("--resume", default="", metavar="TASK_ID",
                   help="start from the checkpoint of an earlier task ($KUBEML_HOME/checkpoints/<id>)")
    p.add_argument("--wait", action="store_true", help="block until the task finishes, print its history")
    p.set_defaults(fn=cmd_train)

    p = sub.add_parser("infer", help="run inference with a running task's model")
    p.add_argument("--id", "--network", dest="id", required=True)
    p.add_argument("--datafile", required=True)
    p.set_defaults(fn=cmd_infer)

    p = sub.add_parser("dataset")
    s2 = p.add_subparsers(dest="action", required=True)
    q = s2.add_parser("create")
    q.add_argument("--name", "-n", required=True)
    q.add_argument("--traindata", required=True)
    q.add_argument("--trainlabels", required=True)
    q.add_argument("--testdata", required=True)
    q.add_argument("--testlabels", required=True)
    q = s2.add_parser("delete")
    q.add_argument("--name", "-n", required=True)
    s2.add_parser("list")
    p.set_defaults(fn=cmd_dataset)

    p = sub.add_parser("function", aliases=["fn"])
    s2 = p.add_subparsers(dest="action", required=True)
    q = s2.add_parser("create")
    q.add_argument("--name", "-n", required=True)
    q.add_argument("--code", required=True)
    q = s2.add_parser("delete")
    q.add_argument("--name", "-n", required=True)
    s2.add_parser("list")
    p.set_defaults(fn=cmd_function)

    p = sub.add_parser("history")
    s2 = p.add_subparsers(dest="action", required=True)
    for name in ("get", "delete"):
        q = s2.add_parser(name)
        q.add_argument("--id", "--network", dest="id", required=True)
    s2.add_parser("list")
    s2.add_parser("prune")
    p.set_defaults(fn=cmd_history)

    p = sub.add_parser("task")
    s2 = p.add_subparsers(dest="action", required=True)
    q = s2.add_parser("list")
    q.add_argument("--short", action="store_true")
    q = s2.add_parser("stop")
    q.add_argument("--id", required=True)
    s2.add_parser("prune")
    p.set_defaults(fn=cmd_task, short=False)

    p = sub.add_parser("logs")
    p.add_argument("--id", required=True)
    p.add_argument("-f", "--follow", action="store_true")
    p.set_defaults(fn=cmd_logs)
    return ap


def main(argv: Optional[List[str]] = None) -> int:
    args = build_parser().parse_args(argv)
    try:
        return int(args.fn(args) or 0)
    except KubeMLException as e:
        print(f"Error: {e.message} (code {e.status_code})", file=sys.stderr)
        return 1
    finally:
        cluster = getattr(args, "_cluster", None)
        if cluster is not None:
            cluster.shutdown()


if __name__ == "__main__":
    sys.exit(main())
