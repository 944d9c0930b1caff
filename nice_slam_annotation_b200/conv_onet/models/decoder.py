"""Decoders of the render hot path with the reference's module tree and parameter names
(src/conv_onet/models/decoder.py), evaluated by the fused sm_100a kernels.

The classes hold parameters only (same ``state_dict`` keys, same initialisation, same
constructor signatures: ``load_state_dict`` of a reference checkpoint, ``copy.deepcopy``,
``.to(device)``, ``.share_memory()`` and optimiser parameter groups all work unchanged).
``forward`` does not run torch layers: it hands the points to the CUDA extension
(``fused.eval_points_raw``); there is no CPU path.
"""
import torch
import torch.nn as nn

from ... import fused


class GaussianFourierFeatureTransform(nn.Module):
    """Parameter holder of the embedding ``sin(x @ _B)`` (decoder.py:7-30); evaluated inside
    the decoder kernel in fp32."""

    def __init__(self, num_input_channels, mapping_size=93, scale=25, learnable=True):
        super().__init__()
        if learnable:
            self._B = nn.Parameter(torch.randn((num_input_channels, mapping_size)) * scale)
        else:
            self._B = torch.randn((num_input_channels, mapping_size)) * scale


class DenseLayer(nn.Linear):
    """nn.Linear with xavier-uniform weights and zero bias (decoder.py:70-85)."""

    def __init__(self, in_dim, out_dim, activation="relu", *args, **kwargs):
        self.activation = activation
        super().__init__(in_dim, out_dim, *args, **kwargs)

    def reset_parameters(self):
        torch.nn.init.xavier_uniform_(self.weight, gain=torch.nn.init.calculate_gain(self.activation))
        if self.bias is not None:
            torch.nn.init.zeros_(self.bias)


def _check_supported(name, hidden_size, n_blocks, skips, c_dim, dims, pos_embedding_method=None):
    ok = hidden_size == 32 and n_blocks == 5 and list(skips) == [2] and c_dim in dims
    if pos_embedding_method is not None:
        ok = ok and pos_embedding_method == 'fourier'
    if not ok:
        raise NotImplementedError(
            f"decoder {name!r}: the sm_100a kernels cover the NICE-SLAM configuration (hidden 32, 5 "
            f"blocks, skip at 2, c_dim 32, fourier embedding); got hidden={hidden_size}, n_blocks={n_blocks}, "
            f"skips={skips}, c_dim={c_dim}, embedding={pos_embedding_method}")


class MLP(nn.Module):
    """decoder.py:93-219.  ``name`` selects the feature grid (``c_grid['grid_' + name]``)."""

    def __init__(self, name='', dim=3, c_dim=128, hidden_size=256, n_blocks=5, leaky=False,
                 sample_mode='bilinear', color=False, skips=[2], grid_len=0.16,
                 pos_embedding_method='fourier', concat_feature=False):
        super().__init__()
        self.name = name
        self.color = color
        self.no_grad_feature = False
        self.c_dim = c_dim
        self.grid_len = grid_len
        self.concat_feature = concat_feature
        self.n_blocks = n_blocks
        self.skips = skips
        self.hidden_size = hidden_size
        self.pos_embedding_method = pos_embedding_method
        if leaky or sample_mode != 'bilinear':
            raise NotImplementedError("leaky ReLU / nearest sampling are not part of the render hot path")
        if c_dim != 0:
            self.fc_c = nn.ModuleList([nn.Linear(c_dim, hidden_size) for i in range(n_blocks)])
        if pos_embedding_method != 'fourier':
            raise NotImplementedError("only the 'fourier' embedding is implemented (configs use no other)")
        embedding_size = 93
        self.embedder = GaussianFourierFeatureTransform(dim, mapping_size=embedding_size, scale=25)
        self.pts_linears = nn.ModuleList(
            [DenseLayer(embedding_size, hidden_size, activation="relu")] +
            [DenseLayer(hidden_size, hidden_size, activation="relu") if i not in self.skips
             else DenseLayer(hidden_size + embedding_size, hidden_size, activation="relu")
             for i in range(n_blocks - 1)])
        self.output_linear = DenseLayer(hidden_size, 4 if self.color else 1, activation="linear")
        self.sample_mode = sample_mode

    def kind(self):
        if self.c_dim == 0:
            raise RuntimeError("the single-MLP (iMAP*) decoder is not a NICE sub-decoder")
        _check_supported(self.name, self.hidden_size, self.n_blocks, self.skips, self.c_dim, (32, 64),
                         self.pos_embedding_method)
        if self.color:
            return fused.DEC_COLOR
        return fused.DEC_FINE if self.concat_feature else fused.DEC_MIDDLE

    def forward(self, p, c_grid=None):
        """Raw decoder output for points p ([1,N,3] or [N,3]): [N] or [N,4] (colour)."""
        if self.c_dim == 0:                 # iMAP*: no feature grid (config.py:33-37)
            from ... import imap
            return imap.eval_points(p, self)
        return fused.decoder_forward_single(self, self.kind(), p, c_grid)


class MLP_no_xyz(nn.Module):
    """decoder.py:222-297 (coarse level: the point only addresses the grid)."""

    def __init__(self, name='', dim=3, c_dim=128, hidden_size=256, n_blocks=5, leaky=False,
                 sample_mode='bilinear', color=False, skips=[2], grid_len=0.16):
        super().__init__()
        self.name = name
        self.no_grad_feature = False
        self.color = color
        self.grid_len = grid_len
        self.c_dim = c_dim
        self.n_blocks = n_blocks
        self.skips = skips
        self.hidden_size = hidden_size
        if leaky or sample_mode != 'bilinear' or color:
            raise NotImplementedError("leaky ReLU / nearest sampling / colour output are not used by NICE")
        self.pts_linears = nn.ModuleList(
            [DenseLayer(hidden_size, hidden_size, activation="relu")] +
            [DenseLayer(hidden_size, hidden_size, activation="relu") if i not in self.skips
             else DenseLayer(hidden_size + c_dim, hidden_size, activation="relu")
             for i in range(n_blocks - 1)])
        self.output_linear = DenseLayer(hidden_size, 1, activation="linear")
        self.sample_mode = sample_mode

    def kind(self):
        _check_supported(self.name, self.hidden_size, self.n_blocks, self.skips, self.c_dim, (32,))
        return fused.DEC_COARSE

    def forward(self, p, c_grid, **kwargs):
        return fused.decoder_forward_single(self, self.kind(), p, c_grid)


class NICE(nn.Module):
    """Neural Implicit Scalable Encoding (decoder.py:300-370): coarse / middle / fine / colour
    decoders over the hierarchical feature grids."""

    def __init__(self, dim=3, c_dim=32, coarse_grid_len=2.0, middle_grid_len=0.16, fine_grid_len=0.16,
                 color_grid_len=0.16, hidden_size=32, coarse=False, pos_embedding_method='fourier'):
        super().__init__()
        if coarse:
            self.coarse_decoder = MLP_no_xyz(name='coarse', dim=dim, c_dim=c_dim, color=False,
                                             hidden_size=hidden_size, grid_len=coarse_grid_len)
        self.middle_decoder = MLP(name='middle', dim=dim, c_dim=c_dim, color=False, skips=[2], n_blocks=5,
                                  hidden_size=hidden_size, grid_len=middle_grid_len,
                                  pos_embedding_method=pos_embedding_method)
        self.fine_decoder = MLP(name='fine', dim=dim, c_dim=c_dim * 2, color=False, skips=[2], n_blocks=5,
                                hidden_size=hidden_size, grid_len=fine_grid_len, concat_feature=True,
                                pos_embedding_method=pos_embedding_method)
        self.color_decoder = MLP(name='color', dim=dim, c_dim=c_dim, color=True, skips=[2], n_blocks=5,
                                 hidden_size=hidden_size, grid_len=color_grid_len,
                                 pos_embedding_method=pos_embedding_method)

    def scene_bound(self):
        """The Renderer's bound: every non-coarse decoder carries it (NICE_SLAM.py:202-206)."""
        return self.middle_decoder.bound

    def forward(self, p, c_grid, stage='middle', **kwargs):
        """raw [N,4] = [rgb | occupancy] of the stage (decoder.py:336-370), WITHOUT the
        out-of-bound rule (that lives in Renderer.eval_points)."""
        return fused.nice_forward(self, p, c_grid, stage)
