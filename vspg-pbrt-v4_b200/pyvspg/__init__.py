"""pyvspg: Python harness over libvspg.so, the B200-native render pass of GuidedVolPathVSPGIntegrator.

`api` binds the C ABI (include/vspg.h); `scenes`, `synthetic`, `configs` build the POD inputs; `dist` shards sample
passes over ranks (torch.distributed).  The CUDA library is mandatory: nothing here computes radiance on the CPU.
"""
from . import _abi  # noqa: F401
