#!/usr/bin/env python
"""Create synthetic MNIST / CIFAR-10 / CIFAR-100-shaped datasets in the local store (the reference's
ml/hack/*.sh upload the real ones through `kubeml dataset create`; there is no network here)."""
import argparse, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from kubeml_b200.storage import default_store

SHAPES = {"mnist": ((28, 28), 10, 60000, 10000), "cifar10": ((32, 32, 3), 10, 50000, 10000),
          "cifar100": ((32, 32, 3), 100, 50000, 10000)}

if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("names", nargs="*", default=list(SHAPES))
    ap.add_argument("--scale", type=float, default=1.0, help="fraction of the real dataset size")
    a = ap.parse_args()
    st = default_store()
    rng = np.random.default_rng(0)
    for n in a.names:
        shape, classes, ntr, nte = SHAPES[n]
        ntr, nte = max(64, int(ntr * a.scale)), max(64, int(nte * a.scale))
        if st.exists(n):
            print(n, "exists"); continue
        st.create(n, rng.integers(0, 256, (ntr, *shape), dtype=np.uint8), rng.integers(0, classes, (ntr,)),
                  rng.integers(0, 256, (nte, *shape), dtype=np.uint8), rng.integers(0, classes, (nte,)))
        print(n, st.summary(n).to_json())
