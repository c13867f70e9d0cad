import os
import sys

import pytest

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on a B200 via gpurun)")
    config.addinivalue_line("markers", "multigpu: needs >= 2 CUDA devices")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        have = torch.cuda.is_available()
        ngpu = torch.cuda.device_count() if have else 0
    except Exception:
        have, ngpu = False, 0
    skip_gpu = pytest.mark.skip(reason="no CUDA device")
    skip_multi = pytest.mark.skip(reason="needs >= 2 CUDA devices")
    for item in items:
        if "gpu" in item.keywords and not have:
            item.add_marker(skip_gpu)
        if "multigpu" in item.keywords and ngpu < 2:
            item.add_marker(skip_multi)


@pytest.fixture()
def kube_home(tmp_path, monkeypatch):
    monkeypatch.setenv("KUBEML_HOME", str(tmp_path))
    monkeypatch.delenv("GPU_ID", raising=False)
    from kubeml_b200.storage import set_default_store
    set_default_store(None)
    yield str(tmp_path)
    set_default_store(None)
