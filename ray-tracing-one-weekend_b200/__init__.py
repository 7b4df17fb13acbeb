"""rtow_b200 -- Python test/bench driver over the C ABI of the B200-native path tracer.

The product is the CUDA library `csrc/librtow_b200.so` (include/rtow_b200.h); this package is a
thin ctypes binding used by tests/, bench.py and __graft_entry__.py.  It contains no rendering
or intersection code and NO fallback: if the shared library is missing or no CUDA device is
visible, compute calls raise.

The directory name carries a hyphen (it is the repository's package directory), so it is
loaded under the module name ``rtow_b200`` by ``__graft_entry__.load_package()``.
"""
from . import _abi as abi  # noqa: F401
from . import raysets  # noqa: F401
from . import sharding  # noqa: F401
from .progressive import ProgressiveRender  # noqa: F401
from .binding import (  # noqa: F401
    RtowError,
    lib,
    lib_path,
    HostScene,
    decode_image,
    image_dir,
    DeviceScene,
    MultiDeviceScene,
    make_camera,
    main_camera,
    write_ppm,
    write_pfm,
    write_exr,
    oriented_camera,
)

__all__ = ["abi", "RtowError", "lib", "lib_path", "HostScene", "decode_image", "image_dir", "DeviceScene", "MultiDeviceScene", "make_camera", "main_camera",
           "write_ppm", "write_pfm", "write_exr", "oriented_camera", "ProgressiveRender"]
