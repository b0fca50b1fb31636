"""B200-native `volpath` (mono) hot path of eradiate-kernel: host-side mirror of the
reference's load_dict / render entry points over the C-ABI of include/b2vp.h."""
from . import abi
from .abi import B2vpError
from .dist import reduce_film, shard_interleaved, shard_range
from .render import DeviceScene, RenderResult, read_exr, render, render_defaults, shard_streams
from .scene import Scene, load_dict, read_vol, write_vol
from .transform import Transform

__all__ = ["abi", "B2vpError", "DeviceScene", "RenderResult", "render", "render_defaults", "shard_streams", "read_exr", "Scene", "load_dict", "read_vol", "write_vol",
           "Transform", "shard_range", "shard_interleaved", "reduce_film"]


def ext():
    """The pybind11 front end (`_b2vp`: load_dict / DeviceScene.render in C++ over include/b2vp.hpp), built by
    __graft_entry__.build() next to libb2vp.so.  Imported on demand: the ctypes host needs no compiled module."""
    import importlib
    return importlib.import_module("._b2vp", __name__)
