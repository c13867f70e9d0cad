"""Two plain processes (no torchrun, so that one may die without the launcher tearing the other down):
rank 1 exits hard in the middle of a fan-out; rank 0 must get an error entry for it within the heartbeat
timeout instead of blocking, and later fan-outs must not wait for it at all."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))


def main():
    from kubeml_b200.runtime.spmd import SpmdRuntime
    rank = int(os.environ["RANK"])
    rt = SpmdRuntime(use_device_sync=False)
    inner = rt._execute

    def execute(msg):
        if msg.get("die") and rank == 1:
            os._exit(17)                       # no goodbye, no result
        if msg.get("echo") is not None:
            return ({"echo": msg["echo"], "rank": rank}, 200)
        return inner(msg)
    rt._execute = execute
    if rank != 0:
        rt.serve()
        return
    out = {}
    out["first"] = rt.run({"echo": 1})
    t0 = time.time()
    out["after_death"] = rt.run({"die": True, "echo": 2})
    out["detect_s"] = time.time() - t0
    t0 = time.time()
    out["later"] = rt.run({"echo": 3})
    out["later_s"] = time.time() - t0
    print("RESULT " + json.dumps(out), flush=True)
    os._exit(0)                                # the process group cannot be torn down cleanly with a dead peer


if __name__ == "__main__":
    main()
