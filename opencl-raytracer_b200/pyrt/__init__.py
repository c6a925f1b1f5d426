"""pyrt -- Python host-side mirror of the render path's C ABI (include/rt_b200.h).

`types`, `scenes`, `glm32` are pure host logic and import without a GPU.
`pyrt.lib` (the ctypes binding of librt_b200.so) is imported lazily; it fails loudly
when the CUDA library is missing -- there is no CPU fallback.
"""
from . import glm32, scenes, shard, types  # noqa: F401
