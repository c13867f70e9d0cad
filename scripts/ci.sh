#!/bin/sh
# What CI runs (the reference's workflows only build images, .github/workflows/*.yml).
set -e
cd "$(dirname "$0")/.."
python -c "import __graft_entry__ as g; g.build()"
python scripts/check_names.py kubeml_b200 bench bench.py baseline/ref_torch.py baseline/reference_arm.py __graft_entry__.py tests
python -m pytest tests -x -q -m "not gpu"
if python -c "import torch,sys; sys.exit(0 if torch.cuda.is_available() else 1)"; then
  python -m pytest tests -x -q -m gpu
  python bench.py --steps 16 --warmup 4
fi
