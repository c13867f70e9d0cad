#!/bin/sh
# The one offline install of the reference.  /root/reference has no
# setup.py at its root and is read-only: the Python package it contains is installed from a copy,
# without dependencies (torch>=1.7 is not a wheel in the wheelhouse; flask/redisai/pymongo are absent).
# The reference's own example function is placed next to it, unmodified, so that it travels to the
# GPU box with baseline/_ref (git-ignored).
set -e
cd "$(dirname "$0")/.."
rm -rf /tmp/kubeml_refcopy && cp -r /root/reference/python/kubeml /tmp/kubeml_refcopy
python -m pip install --no-index --no-build-isolation --no-deps --find-links /opt/wheelhouse \
    --target baseline/_ref /tmp/kubeml_refcopy
mkdir -p baseline/_ref/experiments
cp /root/reference/ml/experiments/kubeml/function_resnet34.py baseline/_ref/experiments/function_resnet34.py
cp /root/reference/ml/experiments/kubeml/function_lenet.py baseline/_ref/experiments/function_lenet.py
echo "reference installed under baseline/_ref"
