"""B200-native vGPU enforcement / oversubscription data path.

Host-side mirror of the reference plugin's interface for the hot path
(SURVEY.md §8): `plugin` mirrors server.go / vdevice.go, `capi` binds the C-ABI
in include/vgpu.h (libvgpu_dp.so, built in-tree by `make` / __graft_entry__.build()).

The directory name is not a Python identifier; import it through the repo-root
helper `vgpu_b200.py`, which registers this package as ``vgpu_b200``.
"""
from . import capi  # noqa: F401  (raises loudly if libvgpu_dp.so is missing)
from . import plugin  # noqa: F401

try:  # tensor helpers need torch; the C-ABI and plugin mirror do not
    from . import datapath  # noqa: F401
except ImportError:  # pragma: no cover
    datapath = None

__all__ = ["capi", "plugin"]
