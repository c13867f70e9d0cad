"""Hand-written sm_100a ops.  See basic.py (flat optimizers, fused softmax-CE) and gemm.py
(tcgen05/TMEM/TMA GEMM + implicit-GEMM convolution)."""
from .basic import (sgd_step_, adam_step_, fill_, cast_bf16_, softmax_ce_raw, cross_entropy,
                    accuracy_count)
from .gemm import (gemm_bf16, conv2d_fprop, conv2d_dgrad, conv2d_wgrad, conv2d_fprop_supported, BnFwd, BnBwd,
                   max_coresident)
from . import nn

__all__ = ["sgd_step_", "adam_step_", "fill_", "cast_bf16_", "softmax_ce_raw", "cross_entropy",
           "accuracy_count", "gemm_bf16", "conv2d_fprop", "conv2d_dgrad", "conv2d_wgrad", "nn", "conv2d_fprop_supported", "BnFwd", "BnBwd", "max_coresident"]
