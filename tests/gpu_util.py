"""Helpers of the GPU parity tests: build the product's decoders / grids from golden fixtures."""
import numpy as np
import torch

from conftest import sub
from nice_slam_annotation_b200.conv_onet.models.decoder import NICE

DEV = "cuda:0"


def nice_from_state(state, bound, coarse=True, enlarge=2.0):
    """Product NICE module loaded with a reference state_dict, bounds set as
    src/NICE_SLAM.py:202-208 does."""
    dec = NICE(coarse=coarse)
    dec.load_state_dict({k: torch.as_tensor(v) for k, v in state.items()})
    b = torch.as_tensor(np.asarray(bound), dtype=torch.float64)
    dec.middle_decoder.bound = b
    dec.fine_decoder.bound = b
    dec.color_decoder.bound = b
    if coarse:
        dec.coarse_decoder.bound = b * enlarge
    return dec.to(DEV)


def grids_from(d, prefix="c.", requires_grad=False):
    out = {}
    for k, v in sub(d, prefix).items():
        g = v.to(DEV).contiguous(memory_format=torch.channels_last_3d)
        out[k] = g.requires_grad_(requires_grad)
    return out


def oracle_params(state):
    return {k: torch.as_tensor(v) for k, v in state.items()}
