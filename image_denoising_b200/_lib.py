"""ctypes binding of libb200denoise.so (the C ABI in include/b200denoise.h).

The library is built in-tree by ``build()`` (``make -C csrc``: nvcc, sm_100a
only).  There is no CPU fallback anywhere in this package: if the shared
library is missing or no CUDA device is present, calls raise.
"""
import ctypes
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B2D_LIB") or os.path.join(_HERE, "libb200denoise.so")   # B2D_LIB: kernel-tuning variants
CSRC = os.path.join(_HERE, "csrc")

OK, ERR_INVALID, ERR_UNSUPPORTED, ERR_CUDA, ERR_PARSE, ERR_NO_DEVICE = 0, -1, -2, -3, -4, -5

# enums (mirror include/b200denoise.h)
NOISE = {"none": 0, "gaussian": 1, "poisson": 2, "sap": 3, "speckle": 4, "periodic": 5, "brownian": 6,
         "quant": 7, "uniform": 8, "gamma": 9, "rayleigh": 10, "bloom": 11, "shader": 12}
DENOISE = {"none": 0, "gaus_blur": 1, "mean": 2, "median": 3, "bilateral": 4, "wavelet": 5}
WAVELET = {"db1": 0, "haar": 0, "bior1.5": 1}
COMPAT_SHADER_BGR = 1
COMPAT_SAP_TRAIN06 = 2
COMPAT_GAUSSIAN_RANDOM_LEVEL = 16     # host front-end flag (spec.resolve)
LEVEL_SIZE = -1.0


class B2DSpec(ctypes.Structure):
    _fields_ = [("noise", ctypes.c_int32), ("denoise", ctypes.c_int32), ("level", ctypes.c_double),
                ("ksize", ctypes.c_int32), ("d", ctypes.c_int32), ("sigma_color", ctypes.c_float),
                ("sigma_space", ctypes.c_float), ("wavelet", ctypes.c_int32), ("levels", ctypes.c_int32),
                ("compat_flags", ctypes.c_int32), ("reserved", ctypes.c_int32)]


class B2DFields(ctypes.Structure):
    _fields_ = [("f0", ctypes.c_void_p), ("f1", ctypes.c_void_p)]


class B2DError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("libb200denoise error %d: %s" % (code, msg))
        self.code = code


EXPORTS = ["b2d_parse_spec", "b2d_plan_create", "b2d_plan_workspace_bytes", "b2d_plan_destroy", "b2d_run",
           "b2d_run_host", "b2d_host_register", "b2d_host_unregister", "b2d_noise_field", "b2d_wavelet_stats", "b2d_stats", "b2d_lab_convert",
           "b2d_poisson_table", "b2d_gamma_table", "b2d_blob_shape", "b2d_blob_prep", "b2d_jpeg_info", "b2d_jpeg_create", "b2d_jpeg_decode", "b2d_jpeg_destroy", "b2d_jpeg_entropy", "b2d_nms", "b2d_nms_workspace_bytes", "b2d_nms_device", "b2d_nms_device_static",
           "b2d_launch_count", "b2d_reset_launch_count", "b2d_last_error", "b2d_version"]

_lib = None


def build(verbose=False):
    """Compile libb200denoise.so in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    cmd = ["make", "-C", CSRC, "-j", str(min(16, os.cpu_count() or 4))]
    res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if verbose or res.returncode != 0:
        print(res.stdout)
    if res.returncode != 0:
        raise RuntimeError("building libb200denoise.so failed")
    return LIB_PATH


def lib():
    """Load the shared library (never builds implicitly, never falls back)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError("libb200denoise.so is not built: run `python -c 'import __graft_entry__ as g; "
                          "g.build()'` (there is no CPU fallback)")
    L = ctypes.CDLL(LIB_PATH)
    vp, u64, sz = ctypes.c_void_p, ctypes.c_uint64, ctypes.c_size_t
    L.b2d_parse_spec.argtypes = [ctypes.c_char_p, ctypes.POINTER(B2DSpec)]
    L.b2d_plan_create.argtypes = [ctypes.POINTER(vp), ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                  ctypes.POINTER(B2DSpec), ctypes.c_int]
    L.b2d_plan_workspace_bytes.argtypes = [vp, ctypes.POINTER(sz)]
    L.b2d_plan_destroy.argtypes = [vp]
    L.b2d_plan_destroy.restype = None
    L.b2d_run.argtypes = [vp, vp, vp, vp, ctypes.POINTER(B2DFields), u64, u64, vp]
    L.b2d_run_host.argtypes = [vp, vp, vp, u64, u64]
    L.b2d_host_register.argtypes = [vp, sz]
    L.b2d_host_unregister.argtypes = [vp]
    L.b2d_noise_field.argtypes = [vp, vp, vp, u64, u64, vp]
    L.b2d_stats.argtypes = [vp, vp, vp, vp, vp, vp]
    L.b2d_wavelet_stats.argtypes = [vp, ctypes.c_int, vp, vp]
    L.b2d_lab_convert.argtypes = [vp, vp, sz, ctypes.c_int, vp]
    L.b2d_poisson_table.argtypes = [ctypes.c_int, vp]
    L.b2d_gamma_table.argtypes = [ctypes.c_double, vp, vp, vp]
    L.b2d_blob_shape.argtypes = [ctypes.c_int] * 4 + [ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_int),
                                                      ctypes.POINTER(ctypes.c_int)]
    L.b2d_blob_prep.argtypes = [vp, ctypes.c_int, ctypes.c_int, ctypes.c_int, vp, ctypes.c_int, ctypes.c_int,
                                ctypes.c_double, ctypes.POINTER(ctypes.c_double), vp, ctypes.c_int, vp]
    ip = ctypes.POINTER(ctypes.c_int)
    L.b2d_jpeg_info.argtypes = [vp, sz, ip, ip, ip, ip, ip]
    L.b2d_jpeg_create.argtypes = [ctypes.POINTER(vp), ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int]
    L.b2d_jpeg_decode.argtypes = [vp, ctypes.c_int, ctypes.POINTER(vp), ctypes.POINTER(sz), vp, vp]
    L.b2d_jpeg_destroy.argtypes = [vp]
    L.b2d_jpeg_entropy.argtypes = [vp, sz, vp, sz, ip, ip]
    L.b2d_jpeg_entropy.restype = ctypes.c_int
    L.b2d_jpeg_destroy.restype = None
    L.b2d_nms.argtypes = [vp, ip, vp, ctypes.c_int, ctypes.c_int, ctypes.c_float, ctypes.c_int]
    L.b2d_nms_workspace_bytes.argtypes = [ctypes.c_int, ctypes.POINTER(sz)]
    L.b2d_nms_device.argtypes = [vp, ctypes.c_int, ctypes.c_int, ctypes.c_float, vp, vp, vp, vp]
    L.b2d_nms_device_static.argtypes = [vp, ctypes.c_int, ctypes.c_int, ctypes.c_float, vp, vp, vp]
    for name in ("b2d_nms", "b2d_nms_workspace_bytes", "b2d_nms_device", "b2d_nms_device_static"):
        getattr(L, name).restype = ctypes.c_int
    L.b2d_launch_count.restype = u64
    L.b2d_reset_launch_count.restype = None
    L.b2d_last_error.restype = ctypes.c_char_p
    L.b2d_version.restype = ctypes.c_char_p
    for name in ("b2d_parse_spec", "b2d_plan_create", "b2d_plan_workspace_bytes", "b2d_run", "b2d_run_host",
                 "b2d_host_register", "b2d_host_unregister", "b2d_noise_field", "b2d_wavelet_stats", "b2d_stats",
                 "b2d_lab_convert", "b2d_poisson_table", "b2d_gamma_table", "b2d_blob_shape", "b2d_blob_prep", "b2d_jpeg_info", "b2d_jpeg_create",
                 "b2d_jpeg_decode"):
        getattr(L, name).restype = ctypes.c_int
    _lib = L
    return L


def check(rc):
    if rc != 0:
        raise B2DError(rc, lib().b2d_last_error().decode("utf-8", "replace"))
