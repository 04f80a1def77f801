"""amber_b200 -- B200 (sm_100a) implementation of amber's path-tracing hot path.

Only what that path needs: ``csrc/`` (CUDA kernels + the C ABI of
``include/amber_cuda.h``), ``scene`` (flat scene description with the reference's
factory verbs), ``render`` (Sensor / Context / CudaPathTracing: the host-side mirror
of ``rendering::Algorithm<RGB>::Render``), ``dist`` (sample partition + film
all-reduce), ``image_io`` (PNG / PFM output of the resolved film).  The product never imports ``oracle/``.
"""
from . import capi, image_io, importer, scene  # noqa: F401
from .render import (Context, CudaBidirectionalPathTracing, CudaLightTracing, CudaPathTracing, Device,  # noqa: F401
                     Sensor)
from .scene import FlatScene, FlatSceneBuilder, cornel_box, cornel_mesh  # noqa: F401
