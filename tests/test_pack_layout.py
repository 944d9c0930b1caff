"""CPU checks of the host logic of the CUDA path: the packed decoder layout and the warp-level
register choreography of csrc/nsb_decoder.cu, emulated lane by lane (tests/frag_emul.py) against
the oracle MLP and torch autograd; plus the C-ABI surface of the built library."""
import re

import numpy as np
import pytest
import torch

from nice_slam_annotation_b200 import _lib
from oracle import nice_oracle as O
import frag_emul as FE

KINDS = {0: "coarse_decoder.", 1: "middle_decoder.", 2: "fine_decoder.", 3: "color_decoder."}


def canonical(prm, kind):
    names = [f"pts_linears.{i}.weight" for i in range(5)] + [f"pts_linears.{i}.bias" for i in range(5)]
    if kind != 0:
        names += [f"fc_c.{i}.weight" for i in range(5)] + [f"fc_c.{i}.bias" for i in range(5)]
    names += ["output_linear.weight", "output_linear.bias"]
    if kind != 0:
        names.append("embedder._B")
    return [prm[n] for n in names]


def torch_mlp(prm, kind, e, c):
    h = c if kind == 0 else e
    first = h
    for i in range(5):
        h = torch.relu(h @ prm[f"pts_linears.{i}.weight"].t() + prm[f"pts_linears.{i}.bias"])
        if kind != 0:
            h = h + c @ prm[f"fc_c.{i}.weight"].t() + prm[f"fc_c.{i}.bias"]
        if i == 2:
            h = torch.cat([first, h], -1)
    return h @ prm["output_linear.weight"].t() + prm["output_linear.bias"]


def test_library_exports_every_declared_symbol():
    hdr = open(_lib.HERE + "/../include/nice_slam_b200.h").read()
    declared = set(re.findall(r"\b(nsb_[a-z0-9_]+)\s*\(", hdr))
    L = _lib.lib()
    for name in declared:
        assert hasattr(L, name), f"{name} declared in the header but not exported"
    assert declared == set(_lib.EXPORTS), declared ^ set(_lib.EXPORTS)
    assert L.nsb_version() == 1
    # parameter counts of the reference decoders
    assert [L.nsb_flat_numel(k) for k in range(4)] == [6337, 15800, 20920, 15899]


@pytest.mark.parametrize("kind", [0, 1, 2, 3])
def test_fragment_emulation_matches_mlp(kind):
    torch.manual_seed(kind)
    full = O.init_decoder_params(seed=3 + kind)
    pre = KINDS[kind]
    prm = {k[len(pre):]: v.double() + 0.05 * torch.randn(v.shape).double() for k, v in full.items()
           if k.startswith(pre)}
    cdim = 64 if kind == 2 else 32
    c = torch.randn(16, cdim, dtype=torch.float64, requires_grad=True)
    e = None
    if kind != 0:
        p = torch.rand(16, 3, dtype=torch.float64)
        e = torch.sin(p @ prm["embedder._B"] * 0.05).detach().requires_grad_(True)
    out = torch_mlp(prm, kind, e, c)
    nout = out.shape[1]
    gout = torch.randn(16, nout, dtype=torch.float64)
    (out * gout).sum().backward()

    blob = FE.pack(kind, [t.numpy() for t in canonical(prm, kind)])
    e_pad = None
    if kind != 0:
        e_pad = np.zeros((16, 96))
        e_pad[:, :93] = e.detach().numpy()
    o, masks = FE.forward(kind, blob, e_pad, c.detach().numpy())
    assert np.allclose(o[:, :nout], out.detach().numpy(), rtol=1e-10, atol=1e-10)
    assert np.allclose(o[:, nout:], 0.0)
    dout = np.zeros((16, 8))
    dout[:, :nout] = gout.numpy()
    dE, dc = FE.backward(kind, blob, masks, dout)
    assert np.allclose(dc, c.grad.numpy()[:, :32], rtol=1e-9, atol=1e-10)
    if kind != 0:
        assert np.allclose(dE[:, :93], e.grad.numpy(), rtol=1e-9, atol=1e-10)
        assert np.allclose(dE[:, 93:], 0.0)


def test_imap_workspace_sizes_hold_the_weight_images():
    """Host-only size queries of the iMAP* path (no GPU work): the forward workspace and the backward scratch hold
    the activations AND the per-launch operand images of the weight matrices (csrc/nsb_imap.cu: pack_floats), at
    offsets that keep the 16-byte alignment the bulk copies need."""
    L = _lib.lib()

    def pack(N, K):                     # csrc/nsb_gemm.cuh: gemm3x_pack_floats (hi + lo image of every 16-k chunk)
        return (K + 15) // 16 * N * 32

    for hidden in (16, 64, 256):
        need = pack(hidden, 93) + 3 * pack(hidden, hidden) + pack(16, hidden) + pack(hidden, 4) + pack(96, hidden)
        for P in (1, 777, 220000, 300000):
            acts = P * (96 + 4 * hidden)
            assert L.nsb_imap_workspace_floats(P, hidden, 1) == acts + need
            assert L.nsb_imap_backward_scratch_floats(P, hidden) == acts + need
            chunk = min(P, 262144)
            assert L.nsb_imap_workspace_floats(P, hidden, 0) == chunk * (96 + 2 * hidden) + need
            assert acts % 4 == 0 and (chunk * (96 + 2 * hidden)) % 4 == 0      # the images start 16-byte aligned
    assert L.nsb_gemm3x_pack_floats(93, 777) == pack(96, 777)                  # N padded to the MMA width
    assert L.nsb_gemm3x_pack_floats(256, 16) == 256 * 32
