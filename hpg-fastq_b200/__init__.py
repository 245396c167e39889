"""hpg-fastq_b200: B200-native per-read QC hot path of hpg-fastq.

The directory name carries the reference's hyphen, so it is imported through
`__graft_entry__.load_package()` (registers it as module `hpg_fastq_b200`).
Contents: csrc/ (sm_100a kernels + the C ABI of include/hpg_fastq_cuda.h), host/ (the C
`hpg-fastq stats|filter|edit` front end) and hpgq.py (ctypes plumbing for tests and bench).
"""
from . import hpgq  # noqa: F401
from .hpgq import *  # noqa: F401,F403
