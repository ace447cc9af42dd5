"""ctypes binding of libmgb200.so (the C ABI in include/mgb200.h).  No fallback: if the library is
missing and cannot be built, importing this module raises."""
from __future__ import annotations

import ctypes as C
import os

from . import build as _build

MG_OK, MG_EINVAL, MG_ECUDA, MG_ENOMEM, MG_ESTATE, MG_EACTION = 0, -1, -2, -3, -4, -5
OBS_BYTES = 147


class ClassDesc(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("max_steps", C.c_int32),
                ("see_through_walls", C.c_int32), ("rule", C.c_int32), ("reserved", C.c_int32 * 3)]


class Config(C.Structure):
    _fields_ = [("device", C.c_int32), ("num_envs", C.c_int32), ("num_classes", C.c_int32),
                ("autoreset", C.c_int32), ("classes", C.POINTER(ClassDesc))]


def _load():
    path = os.environ.get("MGB200_LIB") or _build.LIB        # MGB200_LIB: an experimental build variant (build.py -D...)
    if path != _build.LIB:
        if not os.path.exists(path):
            raise ImportError(f"MGB200_LIB={path} does not exist")
    elif not os.path.exists(path) or _build.needs_build():
        try:
            _build.build()
        except Exception as e:  # no nvcc on the box and no prebuilt library: fail loudly
            if not os.path.exists(path):
                raise ImportError(f"libmgb200.so is missing and could not be built: {e}") from e
    lib = C.CDLL(path)
    vp, i32, i64 = C.c_void_p, C.c_int32, C.c_int64
    lib.mg_last_error.restype = C.c_char_p
    lib.mg_create.argtypes = [C.POINTER(vp), C.POINTER(Config)]
    lib.mg_destroy.argtypes = [vp]
    lib.mg_destroy.restype = None
    lib.mg_num_envs.argtypes = [vp]
    lib.mg_plane_stride.argtypes = [vp]
    lib.mg_launch_count.argtypes = [vp]
    lib.mg_launch_count.restype = i64
    lib.mg_reset_stride.argtypes = [vp]
    lib.mg_upload_layouts.argtypes = [vp, vp, vp, i32]
    lib.mg_step_occupancy.argtypes = [vp, C.POINTER(i32), C.POINTER(i32), C.POINTER(i64)]
    lib.mg_reset.argtypes = [vp, vp, vp, vp, vp, vp]
    lib.mg_step_many.argtypes = [vp, vp, i32, vp, vp, vp, vp, vp, vp]
    lib.mg_step_many_final.argtypes = [vp, vp, i32, vp, vp, vp, vp, vp, vp, vp, vp]
    lib.mg_step.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp]
    lib.mg_step_many_host.argtypes = [vp, vp, i32, vp, vp, vp, vp, vp]
    lib.mg_poll_error.argtypes = [vp, C.POINTER(i64), C.POINTER(i32)]
    lib.mg_get_state.argtypes = [vp, vp, vp]
    lib.mg_set_state.argtypes = [vp, vp, vp]
    lib.mg_render_pov.argtypes = [vp, i32, vp, vp]
    lib.mg_render_full.argtypes = [vp, i32, i32, vp, vp]
    lib.mg_render_full_stride.argtypes = [vp, i32]
    lib.mg_render_full_stride.restype = i64
    lib.mg_render_atlas.argtypes = [vp, i32, vp]
    lib.mg_obs_fully.argtypes = [vp, vp, vp]
    lib.mg_obs_symbolic.argtypes = [vp, vp, vp]
    lib.mg_obs_onehot.argtypes = [vp, vp, i64, vp, vp]
    lib.mg_obs_flat.argtypes = [vp, vp, vp, vp, i32, vp, vp]
    lib.mg_upload_dyn.argtypes = [vp, vp, i32]
    lib.mg_get_dyn.argtypes = [vp, vp]
    lib.mg_set_dyn.argtypes = [vp, vp]
    lib.mg_set_no_death.argtypes = [vp, C.c_uint32, C.c_double]
    lib.mg_obs_view.argtypes = [vp, i32, vp, vp]
    lib.mg_layout_ids.argtypes = [vp, vp, vp]
    lib.mg_obs_goal_direction.argtypes = [vp, vp, i32, vp, vp]
    lib.mg_stochastic_actions.argtypes = [vp, vp, vp, vp, C.c_double, i32, vp, vp]
    lib.mg_bonus.argtypes = [vp, i32, vp, vp, vp, vp]
    f32 = C.c_float
    lib.mgdp_launch_count.restype = i64
    lib.mgdp_sweep.argtypes = [vp, i64, i32, i64, i64, vp, vp, vp, vp, f32, vp, vp]
    lib.mgdp_solve.argtypes = [vp, i64, i32, i64, vp, vp, vp, vp, f32, f32, i32, vp, vp]
    lib.mgdp_rowmask.argtypes = [vp, i64, i32, i64, vp, vp]
    lib.mgdp_sweep_masked.argtypes = [vp, vp, i64, i32, i64, i64, vp, vp, vp, f32, vp, vp]
    lib.mgdp_solve_masked.argtypes = [vp, vp, i64, i32, i64, vp, vp, vp, f32, f32, i32, vp, vp]
    lib.mgdp_solve_stream.argtypes = [vp, vp, i64, i32, i64, vp, vp, vp, f32, f32, i32, vp, vp]
    lib.mgdp_solve_static.argtypes = [vp, i64, i32, i64, vp, vp, vp, f32, f32, i32, vp, vp]
    lib.mgdp_static_max_states.restype = i64
    lib.mgdp_sweep_push.argtypes = [vp, i64, i32, i64, i64, vp, vp, vp, i32, vp, f32, vp, vp]
    lib.mgdp_solve_push.argtypes = [vp, i64, i32, i64, i64, vp, vp, vp, vp, vp, vp, i32, i32, vp, f32, f32, i32, vp, vp, vp]
    lib.mgdp_solve_push_mc.argtypes = [vp, i64, i32, i64, i64, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, vp, f32, f32, i32, vp, vp, vp]
    lib.mgdp_solve_push_masked.argtypes = [vp, vp, i64, i32, i64, i64, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, vp, f32, f32, i32, vp, vp, vp]
    lib.mgdp_finite_horizon.argtypes = [vp, i64, i32, i64, i32, C.c_double, vp, vp, vp, vp]
    lib.mgdp_enumerate.argtypes = [C.POINTER(vp), i32, vp, i32, i32, i32, i32, i32, i32, vp, i32, i64]
    lib.mgdp_enumerate_batch.argtypes = [C.POINTER(vp), i32, i32, vp, i32, i32, vp, i32, vp, i32, i64]
    lib.mgdp_num_layouts.argtypes = [vp]
    lib.mgdp_key_words.argtypes = [vp]
    lib.mgdp_layout_offsets.argtypes = [vp, vp]
    lib.mgdp_free.argtypes = [vp]
    lib.mgdp_free.restype = None
    lib.mgdp_num_states.argtypes = [vp]
    lib.mgdp_num_states.restype = i64
    lib.mgdp_padded_states.argtypes = [vp]
    lib.mgdp_padded_states.restype = i64
    lib.mgdp_num_levels.argtypes = [vp]
    lib.mgdp_succ_dev.argtypes = [vp]
    lib.mgdp_succ_dev.restype = vp
    lib.mgdp_info_dev.argtypes = [vp]
    lib.mgdp_info_dev.restype = vp
    lib.mgdp_copy_out.argtypes = [vp, vp, vp, vp]
    return lib


lib = _load()


class MgError(RuntimeError):
    pass


def check(rc, what=""):
    if rc != MG_OK:
        msg = lib.mg_last_error().decode()
        if rc == MG_EACTION:
            raise ValueError(msg)
        raise MgError(f"{what} failed ({rc}): {msg}")
