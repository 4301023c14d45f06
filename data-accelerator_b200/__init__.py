"""data-accelerator_b200 — B200 (sm_100a) flow processor for Data Accelerator.

Only the per-micro-batch hot path lives here: csrc/ (CUDA kernels + C ABI, built into
libdxgpu.so) and the host-side mirror of the reference's processor interface.  The directory
name carries a hyphen; import it as `import dxgpu` (root shim) or via importlib.
"""
from ._native import DxError, GenParams, lib, LIB_PATH, declared_symbols  # noqa: F401
from .device import Context, Decoder, Table  # noqa: F401
from .processor import (CommonProcessorFactory, CommonProcessor, BatchResult, rules_of, rule_hits, sql_canonical,  # noqa: F401
                        transform_split, string_to_timestamp, duration_seconds)
from . import flowgen  # noqa: F401
from . import parallel  # noqa: F401
