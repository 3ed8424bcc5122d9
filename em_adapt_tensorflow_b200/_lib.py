"""ctypes binding of libweaklabel_cuda.so (include/weaklabel_cuda.h).

The library is the product; this module only declares prototypes.  It fails loudly when the
shared object is missing and cannot be built -- there is no CPU fallback anywhere."""
import ctypes as C
import os

from . import build as _build

_f32p = C.POINTER(C.c_float)
_i32p = C.POINTER(C.c_int)
_u8p = C.POINTER(C.c_ubyte)
_lib = None


class WeakLabelError(RuntimeError):
    pass


def lib_path():
    return _build.LIB


def load():
    """Load (building in-tree first if stale and nvcc is present) and return the CDLL."""
    global _lib
    if _lib is not None:
        return _lib
    path = _build.LIB
    if not os.path.exists(path) or os.environ.get("WEAKLABEL_REBUILD") == "1":
        path = _build.build()
    else:
        try:
            path = _build.build()          # no-op when up to date; rebuilds stale objects
        except RuntimeError:
            pass                           # no nvcc on this box: use the prebuilt object
    L = C.CDLL(path)
    L.e_step.argtypes = [_f32p, _i32p, _i32p, _i32p, C.c_bool, C.c_int, C.c_float, C.c_float, C.c_float]
    L.e_step.restype = C.c_void_p
    L.weaklabel_last_error.restype = C.c_char_p
    L.weaklabel_version.restype = C.c_int
    L.weaklabel_device_count.restype = C.c_int
    L.weaklabel_rng_create.argtypes = [C.c_uint]
    L.weaklabel_rng_create.restype = C.c_void_p
    L.weaklabel_rng_destroy.argtypes = [C.c_void_p]
    L.weaklabel_rng_next.argtypes = [C.c_void_p]
    L.weaklabel_rng_next.restype = C.c_int
    L.weaklabel_rng_skip.argtypes = [C.c_void_p, C.c_longlong]
    L.weaklabel_rng_draws.argtypes = [C.c_void_p]
    L.weaklabel_rng_draws.restype = C.c_longlong
    L.weaklabel_draws_for.argtypes = [C.c_int, C.c_int]
    L.weaklabel_draws_for.restype = C.c_longlong
    L.weaklabel_label_set_host.argtypes = [_i32p, C.c_int, C.c_int, C.c_int, C.c_int, _u8p]
    L.weaklabel_draw_order_host.argtypes = [_u8p, C.c_int, C.c_int, C.c_int, C.c_void_p, _i32p]
    L.weaklabel_plan_create.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]
    L.weaklabel_plan_destroy.argtypes = [C.c_void_p]
    L.weaklabel_plan_destroy.restype = None
    L.weaklabel_plan_prepare.argtypes = [C.c_void_p, _u8p, C.c_bool, C.c_int, C.c_float, C.c_float,
                                         C.c_float, C.c_void_p]
    L.weaklabel_plan_prepare_async.argtypes = [C.c_void_p, _u8p, C.c_bool, C.c_int, C.c_float, C.c_float,
                                               C.c_float, C.c_void_p, C.c_void_p]
    L.weaklabel_label_set_device.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                             C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    L.weaklabel_plan_prepare_device_gt.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_bool,
                                                   C.c_int, C.c_float, C.c_float, C.c_float, C.c_void_p, C.c_void_p]
    L.weaklabel_plan_status.argtypes = [C.c_void_p, _i32p]
    L.weaklabel_plan_set_outputs.argtypes = [C.c_void_p, C.c_int]
    L.weaklabel_fixed_bias_labels_device.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_float,
                                                     C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]
    L.weaklabel_estep_ce_device.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                            C.c_void_p, C.c_void_p]
    L.weaklabel_estep_labels_dlpack.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.weaklabel_estep_labels_device.argtypes = [C.c_void_p, C.c_void_p, C.c_longlong, C.c_longlong,
                                                C.c_void_p, C.c_void_p]
    L.weaklabel_estep_finalize_map_device.argtypes = [C.c_void_p, C.c_void_p, C.c_longlong,
                                                      C.c_longlong, C.c_void_p]
    L.weaklabel_plan_probes.argtypes = [C.c_void_p, _i32p, _f32p, _f32p, C.POINTER(C.c_double)]
    L.weaklabel_plan_tune.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int]
    L.weaklabel_plan_set_path.argtypes = [C.c_void_p, C.c_int]
    L.weaklabel_plan_last_path.argtypes = [C.c_void_p]
    L.weaklabel_plan_launches.argtypes = [C.c_void_p]
    L.weaklabel_plan_profile.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
    L.weaklabel_plan_launches.restype = C.c_longlong
    L.e_step_labels_device.argtypes = [C.c_void_p, _i32p, _i32p, _i32p, C.c_bool, C.c_int, C.c_float,
                                       C.c_float, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.weaklabel_softmax_ce_device.argtypes = [C.c_void_p, C.c_void_p, C.c_longlong, C.c_int, C.c_int,
                                              C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.weaklabel_softmax_ce_static.argtypes = [C.c_void_p, C.c_void_p, C.c_longlong, C.c_int, C.c_int,
                                              C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.weaklabel_upsample_bilinear_device.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                                     C.c_void_p, C.c_void_p]
    L.weaklabel_estep_labels_upsampled_device.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    _lib = L
    return L


def check(rc):
    if rc != 0:
        raise WeakLabelError("libweaklabel_cuda error %d: %s" % (rc, load().weaklabel_last_error().decode()))


EXPORTS = ["e_step", "e_step_labels_device", "weaklabel_last_error", "weaklabel_version",
           "weaklabel_device_count", "weaklabel_rng_create", "weaklabel_rng_destroy",
           "weaklabel_rng_next", "weaklabel_rng_skip", "weaklabel_rng_draws", "weaklabel_draws_for",
           "weaklabel_label_set_host", "weaklabel_draw_order_host", "weaklabel_plan_create", "weaklabel_plan_destroy",
           "weaklabel_plan_prepare", "weaklabel_estep_labels_device",
           "weaklabel_estep_finalize_map_device", "weaklabel_plan_probes", "weaklabel_plan_tune",
           "weaklabel_plan_launches", "weaklabel_plan_set_path", "weaklabel_plan_last_path",
           "weaklabel_plan_profile",
           "weaklabel_softmax_ce_device", "weaklabel_plan_prepare_async", "weaklabel_label_set_device",
           "weaklabel_plan_prepare_device_gt", "weaklabel_plan_status", "weaklabel_estep_ce_device",
           "weaklabel_estep_labels_dlpack", "weaklabel_plan_set_outputs", "weaklabel_fixed_bias_labels_device",
           "weaklabel_softmax_ce_static", "weaklabel_upsample_bilinear_device",
           "weaklabel_estep_labels_upsampled_device"]
