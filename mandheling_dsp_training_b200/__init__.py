"""mandheling_dsp_training_b200 — B200-native NITI int8 training hot path.

`lib`     ctypes binding of the C ABI (include/niti_b200.h), device pointers in / out
`ops`     host-side mirror of the reference's NITI executions (same names, inputs, outputs)
`step`    the int8 training step (LeNet / VGG-11 / ResNet-18 layer lists) and its data-parallel form
"""
from .lib import NitiB200, ConvDesc, NitiError  # noqa: F401
