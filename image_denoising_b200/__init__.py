"""image_denoising_b200 - B200 (sm_100a) implementation of the image-corruption +
denoising preprocessing path of mahesh-kashyap-ml/image-denoising, behind the
reference's spec-string grammar and Python entry points.

Layout: ``csrc/`` hand-written CUDA kernels + the C ABI (``include/b200denoise.h``),
``_lib`` ctypes binding, ``spec`` the grammar, ``api`` batch/host calls,
``minibatch`` / ``test_hook`` the reference-named shims.
"""
from .spec import Spec, SpecError, parse, resolve, resolve_mix  # noqa: F401


def __getattr__(name):          # torch / CUDA are needed only by the compute entry points
    if name in ("apply", "corrupt", "corrupt_host", "Plan", "get_plan"):
        from . import api
        return getattr(api, name)
    raise AttributeError(name)
