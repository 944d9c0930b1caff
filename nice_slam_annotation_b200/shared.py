"""Preparing the objects the reference shares between its processes (src/NICE_SLAM.py:117-148,357-376).

The reference builds ``shared_c`` / ``shared_decoders`` in ``NICE_SLAM.__init__``, moves them to the mapping
device, calls ``share_memory()`` and hands the whole object to three spawned processes; CUDA tensors then reach
the tracker and the coarse mapper through CUDA IPC, i.e. every process maps THE STORAGES THAT EXISTED AT SPAWN.
The fused mapping loop wants the grids channels-last (one 128-byte line per voxel) and every trained decoder as
one flat buffer; converting lazily inside the mapper process would re-allocate them and the tracker's
``update_para_from_mapping`` (src/Tracker.py:137-149) would keep copying the stale storages.

``prepare_shared(slam)`` does both conversions up front, in place of the objects, BEFORE they are shared:

    self.shared_c[key] = val.to(device)            # NICE_SLAM.py:117-121
    self.shared_decoders = self.shared_decoders.to(device)
    nice_slam_annotation_b200.shared.prepare_shared(self)      # <-- the one added line
    self.shared_decoders.share_memory()            # NICE_SLAM.py:134

After it, ``Mapper`` (which reads ``slam.shared_layout_prepared``) raises instead of re-allocating.  Without it
the mapping loop still works across processes: it keeps the original storages and copies the trained values
back into them at the end of every ``optimize_map`` call (``MappingLoop.finish``), at the price of one layout
copy per grid per call.
"""
import torch

from .fused import DEC_NAMES
from .map_loop import flatten_decoder


def prepare_shared(slam, train=("fine_decoder", "color_decoder")):
    """Grids of ``slam.shared_c`` -> channels_last_3d (same logical shape and values); the decoders named in
    ``train`` (the ones ``Mapper.optimize_map`` may optimise, Mapper.py:419-428) -> one flat buffer each."""
    if getattr(slam, "nice", True):
        for key, g in list(slam.shared_c.items()):
            if g.dim() == 5 and not g.is_contiguous(memory_format=torch.channels_last_3d):
                slam.shared_c[key] = g.detach().contiguous(memory_format=torch.channels_last_3d)
        for kind, name in enumerate(DEC_NAMES):
            if name in train and hasattr(slam.shared_decoders, name):
                flatten_decoder(getattr(slam.shared_decoders, name), kind)
    slam.shared_layout_prepared = True
    return slam
