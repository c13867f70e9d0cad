"""`python -m kubeml_b200 ...` — the single entry binary (reference: ml/cmd/ml/main.go:60-156, one
executable whose role is chosen by flag).  `--controllerPort P`, `--schedulerPort P`, `--psPort P`
start the whole control plane with that role's routes on P (the roles are one process here);
`--jobPort P --jobId ID` serves one train job (what the parameter server spawns with STANDALONE_JOBS=true);
anything else is the `kubeml` CLI."""
import sys

from .cli.main import main as cli_main


def main(argv=None) -> int:
    argv = list(sys.argv[1:] if argv is None else argv)
    role_flags = {"--controllerPort": "--port", "--schedulerPort": "--scheduler-port", "--psPort": "--ps-port",
                  "--storagePort": "--storage-port"}
    if "--jobPort" in argv:          # a train job as a process of its own (main.go:60-156, train/server.py)
        import argparse
        ap = argparse.ArgumentParser(prog="kubeml_b200 --jobPort")
        ap.add_argument("--jobPort", type=int, required=True)
        ap.add_argument("--jobId", required=True)
        ap.add_argument("--host", default="127.0.0.1")
        a = ap.parse_args(argv)
        from .train.server import main as job_main
        return job_main(a.jobId, a.jobPort, a.host)
    if argv and argv[0] in role_flags:
        out = ["serve"]
        it = iter(argv)
        for a in it:
            if a in role_flags:
                out += [role_flags[a], next(it)]
        return cli_main(out)
    return cli_main(argv)


if __name__ == "__main__":
    sys.exit(main())
