"""Enum values of the C ABI (include/flextree.h) and dtype tables."""
import numpy as np

UINT8, INT8, UINT16, INT16, INT32, INT64, FLOAT, DOUBLE, BOOL, LONG_LONG, UINT32, UINT64, HALF, BFLOAT16, FP8_BLOCK = range(15)
SUM, BAND, MAX, MIN, BOR = range(5)
ALGO_AUTO, ALGO_ONESHOT, ALGO_STAGES, ALGO_NVLS, ALGO_RING = range(5)
MODE_STAGING, MODE_REGISTERED = 0, 1

OPS = {"sum": SUM, "band": BAND, "max": MAX, "min": MIN, "bor": BOR}
ALGO_TWOSHOT = 6
ALGOS = {"auto": ALGO_AUTO, "oneshot": ALGO_ONESHOT, "stages": ALGO_STAGES, "nvls": ALGO_NVLS, "ring": ALGO_RING,
         "twoshot": ALGO_TWOSHOT}
ALGO_NAMES = {v: k for k, v in ALGOS.items()}
# CPU communicators only: the staged algorithm with the bulk (vectorised chunk) tiles of csrc/host/cpu_backend.h
HOST_ALGOS = dict(ALGOS, stages_bulk=5, twoshot=6)  # 6: two-shot LL (reduce-scatter + all-gather over LL lines)

DTYPE_SIZE = {UINT8: 1, INT8: 1, BOOL: 1, UINT16: 2, INT16: 2, HALF: 2, BFLOAT16: 2, INT32: 4, UINT32: 4,
              FLOAT: 4, INT64: 8, UINT64: 8, DOUBLE: 8, LONG_LONG: 8, FP8_BLOCK: 1}

NUMPY_DTYPE = {
    np.dtype(np.uint8): UINT8, np.dtype(np.int8): INT8, np.dtype(np.uint16): UINT16,
    np.dtype(np.int16): INT16, np.dtype(np.int32): INT32, np.dtype(np.int64): INT64,
    np.dtype(np.float32): FLOAT, np.dtype(np.float64): DOUBLE, np.dtype(np.bool_): BOOL,
    np.dtype(np.uint32): UINT32, np.dtype(np.uint64): UINT64, np.dtype(np.float16): HALF,
}


_TORCH_TABLE = None


def torch_dtype_code(dt):
    global _TORCH_TABLE
    if _TORCH_TABLE is not None:
        try:
            return _TORCH_TABLE[dt]
        except KeyError:
            raise TypeError(f"flextree: unsupported dtype {dt}")
    import torch
    table = {torch.uint8: UINT8, torch.int8: INT8, torch.int16: INT16, torch.int32: INT32,
             torch.int64: INT64, torch.float32: FLOAT, torch.float64: DOUBLE, torch.bool: BOOL,
             torch.float16: HALF, torch.bfloat16: BFLOAT16}
    for name, code in (("uint16", UINT16), ("uint32", UINT32), ("uint64", UINT64)):
        if hasattr(torch, name):
            table[getattr(torch, name)] = code
    _TORCH_TABLE = table
    if dt not in table:
        raise TypeError(f"flextree: unsupported dtype {dt}")
    return table[dt]


__all__ = [n for n in dir() if n.isupper()] + ["torch_dtype_code"]


def default_wire(dt, op, env=None):
    """Wire dtype code when the caller passes none: the environment variable FT_WIRE =
    ``native`` (default: the tensor's own dtype), ``bf16`` (fp32 tensors travel as bf16) or ``fp8`` (fp32 and bf16
    tensors travel as block-scaled e4m3). Narrow wires are lossy, so they only ever apply to floating-point SUM."""
    import os
    v = (os.environ if env is None else env).get("FT_WIRE", "native").strip().lower()
    opc = OPS[op] if isinstance(op, str) else int(op)
    if opc != OPS["sum"] or v in ("", "native"):
        return dt
    if v == "bf16":
        return BFLOAT16 if dt == FLOAT else dt
    if v in ("fp8", "fp8-block16", "fp8_block"):
        return FP8_BLOCK if dt in (FLOAT, BFLOAT16) else dt
    raise ValueError(f"FT_WIRE={v!r}: expected native, bf16 or fp8")
