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


# ------------------------------------------------------------------------------------------------ encode direction
# Reference path: python/seldon_core/utils.py:306-407 (`construct_response_json`): the result array goes through
# `.tolist()` and the wrapper's `json.dumps`, i.e. every float is written by CPython's float.__repr__ (the shortest
# decimal string that reads back as the same binary64, the closest one when there are several) and every int as an int.
# `encode()` IS that path.  `shortest_decimal()` / `format_repr()` restate csrc/jsonencode.cu (Schubfach with the
# tabulated powers of ten of tools/gen_pow10_table.py, then CPython's repr layout rules, Python/pystrtod.c
# format_float_short with mode 'r': exponent form iff decpt <= -4 or decpt > 16); the CPU tests check them against
# repr() so that a wrong table or rounding rule is caught without a GPU.

def encode(arr):
    """The text json.dumps writes for the "ndarray" member of a response."""
    return json.dumps(np.asarray(arr).tolist()).encode()


_G = None
_M63 = (1 << 63) - 1


def _g(k):
    global _G
    if _G is None:
        from gen_pow10_table import K_MIN, table as t10

        _G = (K_MIN, t10())
    return _G[1][k - _G[0]]


def _rop(g1, g0, cp):
    x1 = (g0 * cp) >> 64
    y0 = (g1 * cp) & _M64
    y1 = (g1 * cp) >> 64
    z = (y0 >> 1) + x1
    vbp = y1 + (z >> 63)
    return vbp | (((z & _M63) + _M63) >> 63)


def shortest_decimal(bits):
    """(digits, exponent) with value = digits * 10**exponent, digits without trailing zeros, for a finite positive
    binary64 given by its bit pattern."""
    t = bits & ((1 << 52) - 1)
    bq = (bits >> 52) & 0x7FF
    if bq:
        c, q = t | (1 << 52), bq - 1075
    else:
        c, q = t, -1074
    out = c & 1
    cb = c << 2
    cbr = cb + 2
    if c != (1 << 52) or q == -1074:
        cbl = cb - 2
        k = (q * 661971961083) >> 41
    else:
        cbl = cb - 1
        k = (q * 661971961083 - 274743187321) >> 41
    h = q + ((-k * 913124641741) >> 38) + 2
    g1, g0 = _g(-k)
    vb = _rop(g1, g0, cb << h)
    vbl = _rop(g1, g0, cbl << h)
    vbr = _rop(g1, g0, cbr << h)
    lower, upper = vbl + out, vbr - out
    s = vb >> 2
    done = False
    if s >= 10:
        sp = s // 10
        up_in = lower <= 40 * sp
        wp_in = 40 * sp + 40 <= upper
        if up_in != wp_in:
            dig, e = sp + int(wp_in), k + 1
            done = True
    if not done:
        u_in = lower <= 4 * s
        w_in = 4 * s + 4 <= upper
        if u_in != w_in:
            dig, e = s + int(w_in), k
        else:
            mid = 4 * s + 2
            dig, e = s + int(vb > mid or (vb == mid and (s & 1))), k
    while dig % 10 == 0:
        dig //= 10
        e += 1
    return dig, e


def format_repr(x):
    """float.__repr__ from shortest_decimal (json.dumps spellings for the non-finite values)."""
    if x != x:
        return "NaN"
    if x in (float("inf"), float("-inf")):
        return "Infinity" if x > 0 else "-Infinity"
    bits = struct.unpack("<Q", struct.pack("<d", x))[0]
    sign = "-" if bits >> 63 else ""
    bits &= (1 << 63) - 1
    if bits == 0:
        return sign + "0.0"
    dig, e = shortest_decimal(bits)
    ds = str(dig)
    nd = len(ds)
    decpt = nd + e
    if decpt <= -4 or decpt > 16:
        ex = decpt - 1
        mant = ds[0] + ("." + ds[1:] if nd > 1 else "")
        return "%s%se%s%02d" % (sign, mant, "-" if ex < 0 else "+", abs(ex))
    if decpt <= 0:
        return sign + "0." + "0" * (-decpt) + ds
    if decpt >= nd:
        return sign + ds + "0" * (decpt - nd) + ".0"
    return sign + ds[:decpt] + "." + ds[decpt:]
