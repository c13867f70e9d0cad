import sys, os
sys.path.insert(0, os.getcwd())
import torch
from kubeml_b200 import ops
from kubeml_b200.models.native import fused_plan, pick_block_n
dev = torch.device("cuda", 0); torch.cuda.set_device(dev)
for bn in (64, 128):
    for bmn in (False, True):
        print(bn, bmn, {S: ops.max_coresident(bn, False, bmn, S) for S in (1, 2, 3, 4, 5, 6, 7, 8, 12, 16)})
for name, m_tiles, n, chunks in (("layer1", 64, 64, 9), ("layer2", 16, 128, 18), ("layer3", 4, 256, 36), ("layer4", 1, 512, 8), ("layer4.0.c1 s2", 1, 512, 16), ("stem", 256, 64, 3)):
    print(name, "fwd", fused_plan(dev, m_tiles, n, chunks, pick_block_n(m_tiles, n), False), "bwd", fused_plan(dev, m_tiles, n, chunks, pick_block_n(m_tiles, n), True))

from kubeml_b200 import _native
lib = _native.lib()
import ctypes
lib.kml_gemm_blocks_per_sm.restype = ctypes.c_int
lib.kml_gemm_blocks_per_sm.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int]
for bmn, mode in ((0, 0), (1, 0), (0, 1), (1, 2)):
    v = lib.kml_gemm_blocks_per_sm(64, bmn, mode)
    print("blocks/SM", bmn, mode, v, lib.kml_last_error().decode())
