"""ctypes binding of include/absolutetrack_b200.h (the C-ABI of libabsolutetrack_b200.so).

There is NO fallback: if the shared library is missing or a call fails, an exception is raised.
The reference has no FFI layer (it is pure Python); the functions bound here replace the library
calls behind lib/tracker/tracker.py:61-89, lib/models/umetrack_model.py:188-242 and
lib/common/hand_skinning.py:189-209.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libabsolutetrack_b200.so")

AT_U8, AT_F32, AT_BF16 = 0, 1, 2
AT_MAP_FP32, AT_MAP_FP64 = 0, 1
AT_PREC_FP32, AT_PREC_BF16, AT_PREC_FP16 = 0, 1, 2
AT_HEAD_KNOWN_SKELETON, AT_HEAD_PREDICT_SCALE = 0, 1
AT_STATUS_BAD_SAMPLE_RANGE, AT_STATUS_FP16_OVERFLOW = 1, 2


class AtCropParams(C.Structure):
    _fields_ = [
        ("frame", C.c_int32),
        ("view", C.c_int32),
        ("src_w", C.c_int32),
        ("src_h", C.c_int32),
        ("M", C.c_double * 9),
        ("d", C.c_double * 3),
        ("f_dst", C.c_double * 2),
        ("c_dst", C.c_double * 2),
        ("f_src", C.c_double * 2),
        ("c_src", C.c_double * 2),
        ("k", C.c_double * 8),
        ("reserved", C.c_double * 2),
    ]


CROP_PARAMS_BYTES = 256
assert C.sizeof(AtCropParams) == CROP_PARAMS_BYTES


class AtHandModel(C.Structure):
    _fields_ = [
        ("joint_rotation_axes", C.c_void_p),
        ("joint_rest_positions", C.c_void_p),
        ("landmark_rest_positions", C.c_void_p),
        ("landmark_rest_bone_weights", C.c_void_p),
        ("landmark_rest_bone_indices", C.c_void_p),
        ("n_weights", C.c_int32),
        ("per_sample", C.c_int32),
    ]


class AtLayerInfo(C.Structure):
    _fields_ = [
        ("name", C.c_char * 48),
        ("cin", C.c_int32),
        ("cout", C.c_int32),
        ("ksize", C.c_int32),
        ("stride", C.c_int32),
        ("weight_offset", C.c_int64),
        ("bias_offset", C.c_int64),
    ]


class AtTemporalState(C.Structure):
    _fields_ = [("mem_features", C.c_void_p), ("prev_extrinsics", C.c_void_p), ("capacity", C.c_int64)]


class AtTailInputs(C.Structure):
    _fields_ = [
        ("features", C.c_void_p),
        ("intrinsics", C.c_void_p),
        ("extrinsics", C.c_void_p),
        ("sample_range", C.c_void_p),
        ("memory_idx", C.c_void_p),
        ("use_memory", C.c_void_p),
        ("hand_idx", C.c_void_p),
        ("skel_axes", C.c_void_p),
        ("skel_rest", C.c_void_p),
        ("n", C.c_int32),
        ("c", C.c_int32),
        ("n_skel", C.c_int32),
        ("head_mode", C.c_int32),
    ]


class AtTailOutputs(C.Structure):
    _fields_ = [
        ("joint_angles", C.c_void_p),
        ("wrist_xfs", C.c_void_p),
        ("skel_scales", C.c_void_p),
        ("sigmas", C.c_void_p),
        ("status", C.c_void_p),
    ]


class AtRigCamera(C.Structure):
    _fields_ = [
        ("f", C.c_double * 2),
        ("c", C.c_double * 2),
        ("k", C.c_double * 8),
        ("angle_deg", C.c_double),
        ("width", C.c_int32),
        ("height", C.c_int32),
    ]


CROP_CAMERAS_STATIC_MAX_POSES = 2048  # AT_CROP_CAMERAS_STATIC_MAX_POSES
RIG_CAMERA_BYTES = 112
assert C.sizeof(AtRigCamera) == RIG_CAMERA_BYTES


class AtCropCameraOpts(C.Structure):
    _fields_ = [
        ("num_crop_points", C.c_int32),
        ("min_required_vis_landmarks", C.c_int32),
        ("max_view_num", C.c_int32),
        ("min_num_crops", C.c_int32),
        ("sort_camera_index", C.c_int32),
        ("mirror_right_hand", C.c_int32),
        ("crop_w", C.c_int32),
        ("crop_h", C.c_int32),
        ("focal_multiplier", C.c_double),
        ("min_focal", C.c_double),
    ]


class AtCropCameraInputs(C.Structure):
    _fields_ = [
        ("joint_angles", C.c_void_p),
        ("wrist_xforms", C.c_void_p),
        ("pose_frame", C.c_void_p),
        ("pose_hand", C.c_void_p),
        ("pose_valid", C.c_void_p),
        ("neutral_joint_angles", C.c_void_p),
        ("camera_to_world", C.c_void_p),
        ("cameras", C.c_void_p),
        ("n_poses", C.c_int32),
        ("n_frames", C.c_int32),
        ("n_cams", C.c_int32),
        ("reserved", C.c_int32),
    ]


class AtCropCameraOutputs(C.Structure):
    _fields_ = [
        ("crop_params", C.c_void_p),
        ("intrinsics", C.c_void_p),
        ("extrinsics", C.c_void_p),
        ("crop_camera_to_world", C.c_void_p),
        ("crop_camera_index", C.c_void_p),
        ("sample_range", C.c_void_p),
        ("hand_idx", C.c_void_p),
        ("memory_idx", C.c_void_p),
        ("use_memory", C.c_void_p),
        ("pose_index", C.c_void_p),
        ("tracking_history", C.c_void_p),
        ("counts", C.c_void_p),
    ]


# every symbol include/absolutetrack_b200.h declares: name -> (restype, argtypes)
_VP, _I, _I64, _F = C.c_void_p, C.c_int, C.c_int64, C.c_float
SIGNATURES = {
    "at_version": (_I, []),
    "at_last_error_string": (C.c_char_p, []),
    "at_has_tcgen05": (_I, []),
    "at_launch_count": (_I64, []),
    "at_crop_resample_fisheye62_u8": (_I, [_VP, _I64, _I64, _I64, _VP, _I, _I, _I, _VP, _I, _I, _VP]),
    "at_crop_resample_fisheye62_u8_ex": (_I, [_VP, _I64, _I64, _I64, _VP, _I, _I, _I, _VP, _I, _I, _VP, _VP]),
    "at_crop_pull_source_boxes_u8": (_I, [_VP, _VP, _I64, _I64, _I64, _I64, _VP, _I, _I, _I, _I, _VP, _VP]),
    "at_resample_homography_f32": (_I, [_VP, _I, _I, _I, _VP, _I, _I, _VP, _F, _VP]),
    "at_gen_crop_matrices_f32": (_I, [_VP, _VP, _VP, _VP, _I64, _I, _I, _I, _I, _F, _VP, _VP, _VP, _VP, _VP]),
    "at_fk_skin_f32": (_I, [_VP, _VP, _VP, C.POINTER(AtHandModel), _I64, _VP, _VP]),
    "at_fk_skin_scaled_f32": (_I, [_VP, _VP, _VP, C.POINTER(AtHandModel), _I64, _F, _VP, _VP, _VP]),
    "at_eval_accumulate_f32": (_I, [_VP, _VP, _VP, _VP, _VP, _I64, _VP, _VP]),
    "at_eval_recording_f32": (_I, [_VP, _VP, _VP, _I64, _I64, _VP, _I, _VP, _VP, _VP]),
    "at_crop_cameras_workspace_bytes": (_I64, [_I, _I]),
    "at_gen_crop_cameras": (
        _I,
        [C.POINTER(AtCropCameraInputs), C.POINTER(AtHandModel), C.POINTER(AtCropCameraOpts),
         C.POINTER(AtCropCameraOutputs), _VP, _I64, _VP],
    ),
    "at_gen_crop_cameras_static": (
        _I,
        [C.POINTER(AtCropCameraInputs), C.POINTER(AtHandModel), C.POINTER(AtCropCameraOpts),
         C.POINTER(AtCropCameraOutputs), _VP],
    ),
    "at_model_layer_table": (_I, [C.POINTER(AtLayerInfo), _I]),
    "at_model_blob_floats": (_I64, []),
    "at_model_bf16_blob_bytes": (_I64, []),
    "at_model_prepack_bf16": (_I, [_VP, _VP, _VP]),
    "at_model_prepack_tc": (_I, [_VP, _VP, _I, _VP]),
    "at_debug_set_prof": (None, [_VP]),
    "at_debug_set_prof_seq": (None, [_VP]),
    "at_debug_conv_tc_workspace_bytes": (_I64, [_I, _I, _I]),
    "at_debug_conv_tc": (_I, [_VP, _VP, _I, _VP, _I, _I, _I, _VP, _I, _I, _I, _VP, _VP, _I64, _VP]),
    "at_encoder_workspace_bytes": (_I64, [_I, _I]),
    "at_encoder_forward": (_I, [_VP, _VP, _VP, _I, _I, _VP, _VP, _I64, _I, _VP]),
    "at_encoder_chunk_crops": (_I, []),
    "at_encoder_forward_prepared": (_I, [_VP, _VP, _VP, _I, _I, _VP, _VP, _I64, _I, _I, _VP]),
    "at_encoder_forward_status": (_I, [_VP, _VP, _VP, _I, _I, _VP, _VP, _I64, _I, _I, _VP, _VP]),
    "at_encoder_forward_release": (_I, [_VP, _VP, _VP, _I, _I, _VP, _VP, _I64, _I, _I, _VP, _VP, _VP]),
    "at_encoder_forward_front": (_I, [_VP, _VP, _VP, _I, _I, _VP, _VP, _I64, _I, _I, _VP, _VP, _VP, _VP]),
    "at_tail_workspace_bytes": (_I64, [_I]),
    "at_tail_forward": (
        _I,
        [_VP, C.POINTER(AtTailInputs), C.POINTER(AtTemporalState), C.POINTER(AtTailOutputs), _VP, _I64, _VP],
    ),
    "at_tail_f16_blob_bytes": (_I64, []),
    "at_tail_prepack_f16": (_I, [_VP, _VP, _VP]),
    "at_tail_forward_f16": (
        _I,
        [_VP, _VP, C.POINTER(AtTailInputs), C.POINTER(AtTemporalState), C.POINTER(AtTailOutputs), _VP, _I64, _VP],
    ),
}

_lib = None


class AbsoluteTrackError(RuntimeError):
    pass


def lib():
    """Load (once) and return the C-ABI library; raises if it is not built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise AbsoluteTrackError(
                f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                "(or ./build.sh).  absolutetrack_b200 has no CPU or PyTorch fallback."
            )
        import torch  # noqa: F401  (loads libcudart first so the SONAME resolves to torch's copy)

        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(rc, what=""):
    if rc != 0:
        msg = lib().at_last_error_string().decode(errors="replace")
        raise AbsoluteTrackError(f"{what or 'absolutetrack_b200 call'} failed ({rc}): {msg}")


def layer_table():
    L = lib()
    n = L.at_model_layer_table(None, 0)
    arr = (AtLayerInfo * n)()
    L.at_model_layer_table(arr, n)
    return [
        dict(
            name=a.name.decode(),
            cin=a.cin,
            cout=a.cout,
            ksize=a.ksize,
            stride=a.stride,
            weight_offset=a.weight_offset,
            bias_offset=a.bias_offset,
        )
        for a in arr
    ]


def ptr(t):
    """Device (or host) address of a contiguous torch tensor, None -> NULL."""
    if t is None:
        return None
    assert t.is_contiguous(), "absolutetrack_b200: tensor must be contiguous"
    return t.data_ptr()


def stream_ptr(device=None):
    import torch

    return torch.cuda.current_stream(device).cuda_stream
