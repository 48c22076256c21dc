"""foremast_b200 -- host-side mirror of the foremast-brain scoring pass on top of libforemast_gpu.

The product path is the C-ABI library built from foremast_b200/csrc (include/foremast_gpu.h);
this package only binds it (ctypes), mirrors the brain's job-level interface and shards batches
across GPUs.  Nothing here computes scores on the CPU.
"""
from . import _abi as abi  # noqa: F401

__version__ = "0.1.0"
