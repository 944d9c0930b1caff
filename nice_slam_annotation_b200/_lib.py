"""ctypes binding of the C-ABI library ``csrc/libnsb200.so`` (include/nice_slam_b200.h).

The library is loaded lazily, once per process, the first time a kernel is needed: the
reference pickles Renderer / decoders into three spawned worker processes
(src/NICE_SLAM.py:357-376), so no handle may be created at construction time.  There is no
CPU fallback: a missing library is a hard error.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("NSB_LIB") or os.path.join(HERE, "csrc", "libnsb200.so")   # NSB_LIB: instrumented builds

NSB_EINVAL, NSB_ENOSYS = -22, -38
DEC_COARSE, DEC_MIDDLE, DEC_FINE, DEC_COLOR = 0, 1, 2, 3
STAGES = {"coarse": 0, "middle": 1, "fine": 2, "color": 3}
GRAD_GRID, GRAD_POINT, GRAD_WEIGHT = 1, 2, 4

c_f32p = C.c_void_p      # device pointers travel as integers
c_ptr = C.c_void_p


class NsbGrid(C.Structure):
    _fields_ = [("data", c_ptr), ("grad", c_ptr), ("X", C.c_int32), ("Y", C.c_int32), ("Z", C.c_int32)]


class NsbPoints(C.Structure):
    _fields_ = [("rays_o", c_ptr), ("rays_d", c_ptr), ("rays_o_stride", C.c_int64),
                ("rays_d_stride", C.c_int64), ("z", c_ptr), ("N", C.c_int64), ("S", C.c_int32),
                ("p64", c_ptr), ("p32", c_ptr), ("P", C.c_int64), ("relu_masks", c_ptr),
                ("x_scratch", c_ptr), ("x_scratch_bytes", C.c_int64), ("x_mask", C.c_uint32), ("point_cache", c_ptr),
                ("line_marks", c_ptr), ("line_marks_base", c_ptr)]


class NsbScene(C.Structure):
    _fields_ = [("bound", (C.c_double * 2) * 3), ("coarse_bound", (C.c_double * 2) * 3),
                ("grid", NsbGrid * 4), ("packed", c_ptr * 4)]


class NsbRenderCall(C.Structure):
    _fields_ = [("rays_o", c_ptr), ("rays_d", c_ptr), ("rays_o_stride", C.c_int64), ("rays_d_stride", C.c_int64),
                ("N", C.c_int64), ("gt_depth", c_ptr), ("keep", c_ptr), ("depth_max", c_ptr),
                ("N_samples", C.c_int32), ("N_surface", C.c_int32), ("lindisp", C.c_int32), ("stage", C.c_int32),
                ("occupancy", C.c_int32),
                ("t_uniform", c_ptr), ("t_surface", c_ptr), ("u", c_ptr), ("scene", NsbScene),
                ("workspace", c_ptr), ("workspace_bytes", C.c_int64),
                ("depth", c_ptr), ("var", c_ptr), ("color", c_ptr),
                ("g_depth", c_ptr), ("g_var", c_ptr), ("g_color", c_ptr),
                ("flags", C.c_uint32), ("dweight_mask", C.c_uint32), ("d_weight", c_ptr * 4),
                ("scratch", c_ptr), ("scratch_bytes", C.c_int64), ("g_rays_o", c_ptr), ("g_rays_d", c_ptr),
                ("gt_color", c_ptr), ("w_color", C.c_float), ("use_color", C.c_int32), ("loss", c_ptr),
                ("deposit_dweight_inputs", C.c_int32), ("line_marks", c_ptr), ("line_marks_base", c_ptr),
                ("dweight_phase", C.c_int32), ("dweight_max_ctas", C.c_int32)]


class NsbImapDecoder(C.Structure):
    _fields_ = [("B", c_ptr), ("W", c_ptr * 4), ("b", c_ptr * 4), ("Wo", c_ptr), ("bo", c_ptr),
                ("hidden", C.c_int32)]


class NsbImapGrads(C.Structure):
    _fields_ = [("B", c_ptr), ("W", c_ptr * 4), ("b", c_ptr * 4), ("Wo", c_ptr), ("bo", c_ptr)]


NSB_MAX_PEERS = 8


class NsbPeers(C.Structure):
    _fields_ = [("rank", C.c_int32), ("world", C.c_int32), ("bucket", c_ptr * NSB_MAX_PEERS),
                ("bucket_multicast", c_ptr), ("param", c_ptr * NSB_MAX_PEERS), ("param_multicast", c_ptr),
                ("signal", c_ptr * NSB_MAX_PEERS), ("error", c_ptr), ("scalar", c_ptr * NSB_MAX_PEERS),
                ("touched", c_ptr * NSB_MAX_PEERS), ("line_marks", c_ptr), ("line_marks_floats", C.c_int64)]


_SIGS = {
    "nsb_version": (C.c_int, []),
    "nsb_last_error": (C.c_char_p, []),
    "nsb_launch_count": (C.c_int64, [C.c_int]),
    "nsb_timing": (C.c_int, [C.c_int]),
    "nsb_timing_read": (C.c_int, [c_ptr, c_ptr, c_ptr, c_ptr]),
    "nsb_timing_read_kernel": (C.c_int, [C.c_int, c_ptr, c_ptr]),
    "nsb_param_count": (C.c_int, [C.c_int]),
    "nsb_param_numel": (C.c_int64, [C.c_int, C.c_int]),
    "nsb_flat_numel": (C.c_int64, [C.c_int]),
    "nsb_packed_numel": (C.c_int64, [C.c_int]),
    "nsb_pack_table_host": (C.c_int, [C.c_int, c_ptr]),
    "nsb_layout_offsets": (C.c_int, [C.c_int, c_ptr]),
    "nsb_pack_weights": (C.c_int, [C.c_int, c_ptr, c_ptr, c_ptr, c_ptr]),
    "nsb_sample_rays": (C.c_int, [c_ptr, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                  C.c_float, C.c_float, C.c_float, C.c_float, c_ptr, c_ptr, c_ptr, c_ptr,
                                  C.c_int, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr]),
    "nsb_all_rays": (C.c_int, [C.c_int, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float, c_ptr,
                               c_ptr, c_ptr, c_ptr]),
    "nsb_depth_max": (C.c_int, [c_ptr, c_ptr, C.c_int64, c_ptr, c_ptr]),
    "nsb_sample_z": (C.c_int, [C.POINTER(NsbPoints), c_ptr, c_ptr, C.c_int, C.c_int, C.c_int, c_ptr,
                               c_ptr, c_ptr, c_ptr, c_ptr, c_ptr]),
    "nsb_importance_z": (C.c_int, [c_ptr, c_ptr, C.c_int64, C.c_int, C.c_int, c_ptr, C.c_int64, c_ptr,
                                   c_ptr, c_ptr]),
    "nsb_box_prefilter": (C.c_int, [c_ptr, c_ptr, c_ptr, C.c_int64, c_ptr, c_ptr, c_ptr]),
    "nsb_regulation_points": (C.c_int, [c_ptr, c_ptr, c_ptr, C.c_int64, C.c_int, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr]),
    "nsb_regulation_points_backward": (C.c_int, [c_ptr, c_ptr, C.c_int64, C.c_int, c_ptr, c_ptr, c_ptr]),
    "nsb_sample_pdf": (C.c_int, [c_ptr, c_ptr, C.c_int64, C.c_int, C.c_int, c_ptr, C.c_int64, c_ptr, c_ptr]),
    "nsb_decoders_forward": (C.c_int, [C.POINTER(NsbPoints), C.POINTER(NsbScene), C.c_int, c_ptr, c_ptr,
                                       c_ptr, c_ptr]),
    "nsb_assemble_raw": (C.c_int, [C.POINTER(NsbPoints), c_ptr, C.c_int, c_ptr, c_ptr, c_ptr, c_ptr,
                                   c_ptr, c_ptr]),
    "nsb_dweight_scratch_bytes": (C.c_int64, [C.c_int, C.c_int64]),
    "nsb_decoders_backward": (C.c_int, [C.POINTER(NsbPoints), C.POINTER(NsbScene), C.c_int, c_ptr,
                                        C.c_uint, C.c_uint, c_ptr, c_ptr, c_ptr, C.c_int64, c_ptr]),
    "nsb_composite_forward": (C.c_int, [C.POINTER(NsbPoints), c_ptr, C.c_int, C.c_int, c_ptr, c_ptr,
                                        c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr]),
    "nsb_composite_backward": (C.c_int, [C.POINTER(NsbPoints), c_ptr, C.c_int, C.c_int, c_ptr, c_ptr,
                                         c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr]),
    "nsb_ray_grad_reduce": (C.c_int, [C.POINTER(NsbPoints), c_ptr, C.c_int, c_ptr, c_ptr, c_ptr, c_ptr]),
    "nsb_pose_grad": (C.c_int, [c_ptr, c_ptr, c_ptr, c_ptr, C.c_int64, C.c_int, C.c_int, C.c_int,
                                C.c_float, C.c_float, C.c_float, C.c_float, c_ptr, c_ptr, c_ptr]),
    "nsb_tc_selftest": (C.c_int, [c_ptr, c_ptr, c_ptr, C.c_int, C.c_int, C.c_int, c_ptr]),
    "nsb_tc_mma_latency": (C.c_int, [c_ptr, C.c_int, C.c_int, C.c_int, C.c_int, c_ptr]),
    "nsb_load_batch": (C.c_int, [c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, C.c_int64, C.c_int64, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr,
                                 c_ptr]),
    "nsb_load_batch_packed": (C.c_int, [c_ptr, c_ptr, C.c_int64, C.c_int64, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr]),
    "nsb_render_workspace_bytes": (C.c_int64, [C.c_int64, C.c_int]),
    "nsb_render_forward": (C.c_int, [C.POINTER(NsbRenderCall), c_ptr]),
    "nsb_render_backward": (C.c_int, [C.POINTER(NsbRenderCall), c_ptr]),
    "nsb_render_live_tiles": (C.c_int, [C.POINTER(NsbRenderCall), C.c_int, c_ptr, c_ptr]),
    "nsb_mapping_loss": (C.c_int, [c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, C.c_int64, C.c_float, C.c_int, c_ptr,
                                   c_ptr, c_ptr, c_ptr]),
    "nsb_composite_mapping_partials": (C.c_int64, [C.c_int64]),
    "nsb_composite_mapping": (C.c_int, [C.POINTER(NsbPoints), c_ptr, C.c_int, C.c_int, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr,
                                        c_ptr, C.c_float, C.c_int, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr,
                                        c_ptr]),
    "nsb_composite_mapping_static": (C.c_int, [C.POINTER(NsbPoints), c_ptr, C.c_int, C.c_int, c_ptr, c_ptr, c_ptr, c_ptr,
                                               c_ptr, c_ptr, C.c_float, C.c_int, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr,
                                               c_ptr]),
    "nsb_tracking_loss": (C.c_int, [c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, C.c_int64, C.c_float, C.c_int,
                                    C.c_int, c_ptr, c_ptr, c_ptr, c_ptr]),
    "nsb_adam": (C.c_int, [c_ptr, c_ptr, c_ptr, c_ptr, C.c_int64, c_ptr, c_ptr, c_ptr, C.c_float, C.c_float, C.c_float,
                           c_ptr]),
    "nsb_frustum_mask": (C.c_int, [c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, C.c_int, C.c_int, C.c_int, c_ptr, C.c_int, C.c_int,
                                   C.c_double, C.c_double, C.c_double, C.c_double, c_ptr, c_ptr, c_ptr, c_ptr]),
    "nsb_keyframe_overlap": (C.c_int, [c_ptr, c_ptr, c_ptr, C.c_int64, C.c_int, c_ptr, c_ptr, C.c_int, C.c_int, C.c_int,
                                       C.c_double, C.c_double, C.c_double, C.c_double, C.c_int, c_ptr, c_ptr]),
    "nsb_mesh_depth_workspace_bytes": (C.c_int64, [C.c_int64, C.c_int]),
    "nsb_mesh_depth": (C.c_int, [c_ptr, C.c_int64, c_ptr, C.c_int64, c_ptr, C.c_int, C.c_int, C.c_int, C.c_float,
                                 C.c_float, C.c_float, C.c_float, C.c_float, C.c_float, c_ptr, C.c_int64, c_ptr, c_ptr]),
    "nsb_exchange_signal_words": (C.c_int64, []),
    "nsb_bucket_allreduce": (C.c_int, [C.POINTER(NsbPeers), C.c_int64, C.c_int64, c_ptr]),
    "nsb_exchange_slice": (C.c_int64, [C.c_int, C.c_int, C.c_int64, c_ptr]),
    "nsb_exchange_touched_words": (C.c_int64, [C.c_int64]),
    "nsb_depth_max_peers": (C.c_int, [C.POINTER(NsbPeers), c_ptr, c_ptr, C.c_int64, c_ptr, c_ptr]),
    "nsb_bucket_allreduce_sparse": (C.c_int, [C.POINTER(NsbPeers), C.c_int64, C.c_int64, c_ptr]),
    "nsb_bucket_allreduce_sparse_ctas": (C.c_int, [C.POINTER(NsbPeers), C.c_int64, C.c_int64, C.c_int, c_ptr]),
    "nsb_bucket_reduce_adam": (C.c_int, [C.POINTER(NsbPeers), C.c_int64, C.c_int64, C.c_int64, c_ptr, c_ptr, c_ptr,
                                         c_ptr, c_ptr, C.c_float, C.c_float, C.c_float, c_ptr]),
    "nsb_imap_workspace_floats": (C.c_int64, [C.c_int64, C.c_int, C.c_int]),
    "nsb_imap_forward": (C.c_int, [C.POINTER(NsbPoints), C.POINTER(NsbImapDecoder), c_ptr, c_ptr, C.c_int, c_ptr]),
    "nsb_imap_backward_scratch_floats": (C.c_int64, [C.c_int64, C.c_int]),
    "nsb_imap_backward": (C.c_int, [C.POINTER(NsbPoints), C.POINTER(NsbImapDecoder), c_ptr, c_ptr, C.c_uint,
                                    C.POINTER(NsbImapGrads), c_ptr, c_ptr, c_ptr]),
    "nsb_gemm3x": (C.c_int, [c_ptr, C.c_int64, C.c_int, c_ptr, C.c_int64, C.c_int, c_ptr, C.c_int64, C.c_int64,
                             C.c_int, C.c_int64, C.c_int, c_ptr, c_ptr, C.c_int64, c_ptr, C.c_int, c_ptr, c_ptr]),
    "nsb_gemm3x_pack_floats": (C.c_int64, [C.c_int, C.c_int64]),
    "nsb_pose_grad_adam": (C.c_int, [c_ptr, c_ptr, c_ptr, c_ptr, C.c_int64, C.c_int, C.c_int, C.c_int,
                                     C.c_float, C.c_float, C.c_float, C.c_float, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr,
                                     C.c_float, C.c_float, C.c_float, C.c_float, C.c_float, c_ptr, c_ptr, c_ptr, c_ptr]),
    "nsb_cam_adam": (C.c_int, [c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, C.c_float, C.c_float, C.c_float, C.c_float,
                               C.c_float, c_ptr, c_ptr, c_ptr, c_ptr]),
}

EXPORTS = tuple(_SIGS)
_lib = None


def lib():
    """The loaded library (raises if it has not been built: there is no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build the CUDA extension first "
                "(python -m nice_slam_annotation_b200.build); there is no CPU fallback")
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGS.items():
            fn = getattr(handle, name)
            fn.restype, fn.argtypes = res, args
        _lib = handle
    return _lib


def check(rc, what=""):
    if rc != 0:
        msg = lib().nsb_last_error().decode(errors="replace")
        raise RuntimeError(f"nice_slam_b200 {what} failed (code {rc}): {msg}")


def __getstate__():      # pragma: no cover - module objects are not pickled
    return {}
