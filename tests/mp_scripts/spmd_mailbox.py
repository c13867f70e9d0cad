"""Three plain processes: rank 0 sends two fan-outs at the same time, one to rank 1 and one to rank 2; each
takes 1 s to execute.  With a mailbox per rank they overlap."""
import json, os, sys, threading, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))


def main():
    from kubeml_b200.runtime.spmd import SpmdRuntime
    rank = int(os.environ["RANK"])
    rt = SpmdRuntime(use_device_sync=False)

    def execute(msg):
        if msg.get("nap"):
            time.sleep(msg["nap"])
        return ({"echo": msg.get("echo"), "rank": rank, "members": msg.get("members")}, 200)
    rt._execute = execute
    if rank != 0:
        rt.serve()
        return
    res = {}

    def fan(i):
        res[i] = rt.run({"nap": 1.0, "echo": i}, ranks=[i])
    t0 = time.time()
    ts = [threading.Thread(target=fan, args=(i,)) for i in (1, 2)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    both = time.time() - t0
    held = rt.allocate(2)
    try:
        rt.allocate(2)
        over = False
    except Exception:
        over = True
    rt.free(held)
    everyone = rt.run({"echo": 9})
    rt.stop()
    print("RESULT " + json.dumps({"both_s": both, "res": {str(k): v for k, v in res.items()}, "held": held,
                                  "refused_when_short": over, "everyone": everyone}), flush=True)


if __name__ == "__main__":
    main()
