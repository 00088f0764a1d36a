"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py): the REST JSON ndarray decode of the reference, and a CPU
restatement of the decimal-to-binary64 conversion the device kernel uses.

Reference path: python/seldon_core/utils.py:499-505 (`extract_request_parts_json`): `json.loads` and
`np.array(datadef["ndarray"])`; CPython's json module converts every number with `float()`.  `decode()` below is that
path (it IS the reference's code path: stdlib json + numpy), narrowed to fp32 like every other decode.

`eisel_lemire()` restates csrc/jsoncodec.cu::decimal_to_double_bits (128-bit product with the tabulated power of
five from tools/gen_pow5_table.py); the CPU tests check it against `float()` so that a wrong table or a wrong
rounding rule is caught without a GPU.
"""
import json
import os
import struct
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
from gen_pow5_table import SMALLEST, table  # noqa: E402

_T = None
_M64 = (1 << 64) - 1


def decode(payload):
    with np.errstate(over="ignore"):
        return np.array(json.loads(payload)["data"]["ndarray"]).astype(np.float32)


def eisel_lemire(q, w):
    """binary64 nearest to w * 10**q (0 < w < 2**64), as a Python float."""
    global _T
    if _T is None:
        _T = table()
    if q < -342:
        return 0.0
    if q > 308:
        return float("inf")
    lz = 64 - w.bit_length()
    w <<= lz
    hi5, lo5 = _T[q - SMALLEST]
    p = w * hi5
    phi, plo = p >> 64, p & _M64
    mask = _M64 >> 55
    if (phi & mask) == mask:
        shi = (w * lo5) >> 64
        plo = (plo + shi) & _M64
        if shi > plo:
            phi += 1
    upper = phi >> 63
    mant = phi >> (upper + 64 - 52 - 3)
    power2 = ((((152170 + 65536) * q) >> 16) + 63) + upper - lz + 1023
    if power2 <= 0:
        if -power2 + 1 >= 64:
            return 0.0
        mant >>= (-power2 + 1)
        mant += mant & 1
        mant >>= 1
        bits = mant
    else:
        if plo <= 1 and -4 <= q <= 23 and (mant & 3) == 1 and (mant << (upper + 64 - 52 - 3)) == phi:
            mant &= ~1
        mant += mant & 1
        mant >>= 1
        if mant >= (2 << 52):
            mant = 1 << 52
            power2 += 1
        mant &= ~(1 << 52)
        bits = (0x7ff << 52) if power2 >= 0x7ff else ((power2 << 52) | mant)
    return struct.unpack("<d", struct.pack("<Q", bits))[0]
