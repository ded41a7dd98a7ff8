"""mandheling_dsp_training_b200 — B200-native NITI int8 training hot path.

`lib`     ctypes binding of the C ABI (include/niti_b200.h), device pointers in / out
`step`    the int8 training step (LeNet / VGG-11 / ResNet-18 layer lists) and its data-parallel form
`build`   nvcc recipe for libniti_b200.so (sm_100a, in-tree)
The MNN Backend/Execution creator set that mirrors the reference's operator interface lives in mnn_backend/ (C++).
"""
from .lib import NitiB200, ConvDesc, NitiError  # noqa: F401
