"""ctypes front-end of the CPU checkers.  TEST INFRASTRUCTURE ONLY.

Two checkers live here:

* ``port``  — oracle/flrl_oracle.c, our C restatement of the reference's
  ``cpuCompress`` / ``cpuDecompress`` (/root/reference/src/fl/fl_cpu.cu:9-147)
  and of the run-length spec (IMPLEMENTATION-PLAN.md:81-179).
* ``ref``   — oracle/_ref/libflref.so, the reference's own fl_cpu.cu compiled in
  place by oracle/Makefile (absent when /root/reference was never available).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / ``--impl
reference`` legs may import this module.  The product path (the CUDA library
behind include/flrl.h) never does.
"""
from __future__ import annotations

import ctypes as C
import os
import struct
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
PORT_SO = os.path.join(HERE, "libflrl_oracle.so")
REF_SO = os.path.join(HERE, "_ref", "libflref.so")
REF_CLI = os.path.join(HERE, "_ref", "compress_ref")

FRAME = 128
_u8p = C.POINTER(C.c_uint8)


def build(ref: bool = True) -> None:
    """Compile the checkers (gcc for the port; nvcc on the reference sources for _ref)."""
    targets = ["port"] + (["ref"] if ref else [])
    subprocess.run(["make", "-C", HERE, "-s"] + targets, check=True)


def _ptr(a: np.ndarray):
    return a.ctypes.data_as(_u8p)


def _as_u8(data) -> np.ndarray:
    a = np.frombuffer(data, dtype=np.uint8) if isinstance(data, (bytes, bytearray, memoryview)) else data
    a = np.ascontiguousarray(a, dtype=np.uint8)
    return a


_port = None


def port():
    global _port
    if _port is None:
        if not os.path.exists(PORT_SO):
            build(ref=False)
        lib = C.CDLL(PORT_SO)
        lib.flrl_oracle_fl_frames.restype = C.c_size_t
        lib.flrl_oracle_fl_frames.argtypes = [C.c_size_t]
        lib.flrl_oracle_fl_compress.restype = C.c_uint64
        lib.flrl_oracle_fl_compress.argtypes = [_u8p, C.c_size_t, _u8p, _u8p]
        lib.flrl_oracle_fl_decompress.restype = C.c_uint64
        lib.flrl_oracle_fl_decompress.argtypes = [_u8p, C.c_size_t, _u8p, C.c_size_t, _u8p, C.c_size_t]
        lib.flrl_oracle_rl_compress.restype = C.c_uint64
        lib.flrl_oracle_rl_compress.argtypes = [_u8p, C.c_size_t, _u8p, _u8p]
        lib.flrl_oracle_rl_count_pairs.restype = C.c_uint64
        lib.flrl_oracle_rl_count_pairs.argtypes = [_u8p, C.c_size_t]
        lib.flrl_oracle_rl_decompress.restype = C.c_uint64
        lib.flrl_oracle_rl_decompress.argtypes = [_u8p, _u8p, C.c_size_t, _u8p, C.c_size_t]
        lib.flrl_oracle_rl_decoded_size.restype = C.c_uint64
        lib.flrl_oracle_rl_decoded_size.argtypes = [_u8p, C.c_size_t]
        _port = lib
    return _port


def have_ref() -> bool:
    return os.path.exists(REF_SO)


_ref = None


def ref():
    global _ref
    if _ref is None:
        lib = C.CDLL(REF_SO)
        lib.flref_fl_cpu_compress.restype = C.c_int
        lib.flref_fl_cpu_compress.argtypes = [_u8p, C.c_size_t, C.POINTER(_u8p), C.POINTER(C.c_size_t),
                                              C.POINTER(_u8p), C.POINTER(C.c_size_t)]
        lib.flref_fl_cpu_decompress.restype = C.c_int
        lib.flref_fl_cpu_decompress.argtypes = [C.c_size_t, _u8p, C.c_size_t, _u8p, C.c_size_t,
                                                C.POINTER(_u8p), C.POINTER(C.c_size_t)]
        lib.flref_free.restype = None
        lib.flref_free.argtypes = [C.c_void_p]
        _ref = lib
    return _ref


# ----------------------------------------------------------------------------- port
def fl_compress(data):
    """-> (bits u8[F], values u8[valuesSize]); fl_cpu.cu:9-90."""
    a = _as_u8(data)
    n = a.size
    frames = (n + FRAME - 1) // FRAME
    bits = np.empty(frames, dtype=np.uint8)
    values = np.empty(max(n, 1), dtype=np.uint8)
    nv = port().flrl_oracle_fl_compress(_ptr(a), n, _ptr(bits), _ptr(values))
    return bits, values[:nv].copy()


def fl_decompress(bits, values, n):
    """-> u8[n] (empty when bits or values is empty); fl_cpu.cu:92-147."""
    b, v = _as_u8(bits), _as_u8(values)
    out = np.empty(max(n, 1), dtype=np.uint8)
    got = port().flrl_oracle_fl_decompress(_ptr(b), b.size, _ptr(v), v.size, _ptr(out), n)
    return out[:got].copy()


def rl_compress(data):
    """-> (counts u8[P], values u8[P]); IMPLEMENTATION-PLAN.md:83-152."""
    a = _as_u8(data)
    n = a.size
    p = port().flrl_oracle_rl_count_pairs(_ptr(a), n)
    counts = np.empty(max(p, 1), dtype=np.uint8)
    values = np.empty(max(p, 1), dtype=np.uint8)
    got = port().flrl_oracle_rl_compress(_ptr(a), n, _ptr(counts), _ptr(values))
    assert got == p
    return counts[:p].copy(), values[:p].copy()


def rl_decompress(counts, values):
    """-> u8[sum(counts)]; IMPLEMENTATION-PLAN.md:154-179."""
    c, v = _as_u8(counts), _as_u8(values)
    assert c.size == v.size
    n = port().flrl_oracle_rl_decoded_size(_ptr(c), c.size)
    out = np.empty(max(n, 1), dtype=np.uint8)
    got = port().flrl_oracle_rl_decompress(_ptr(c), _ptr(v), c.size, _ptr(out), n)
    assert got == n
    return out[:n].copy()


# ------------------------------------------------------------------------------ ref
def ref_fl_compress(data):
    """The reference's own cpuCompress (compiled in place).  Prints its [TIMER] lines to fd 1."""
    a = _as_u8(data)
    bits, values = _u8p(), _u8p()
    nb, nv = C.c_size_t(), C.c_size_t()
    rc = ref().flref_fl_cpu_compress(_ptr(a), a.size, C.byref(bits), C.byref(nb), C.byref(values), C.byref(nv))
    if rc:
        raise RuntimeError("reference cpuCompress threw")
    try:
        b = np.ctypeslib.as_array(bits, (nb.value,)).copy() if nb.value else np.empty(0, np.uint8)
        v = np.ctypeslib.as_array(values, (nv.value,)).copy() if nv.value else np.empty(0, np.uint8)
    finally:
        ref().flref_free(bits)
        ref().flref_free(values)
    return b, v


def ref_fl_decompress(bits, values, n):
    b, v = _as_u8(bits), _as_u8(values)
    out, no = _u8p(), C.c_size_t()
    rc = ref().flref_fl_cpu_decompress(n, _ptr(b), b.size, _ptr(v), v.size, C.byref(out), C.byref(no))
    if rc:
        raise RuntimeError("reference cpuDecompress threw")
    try:
        o = np.ctypeslib.as_array(out, (no.value,)).copy() if no.value else np.empty(0, np.uint8)
    finally:
        ref().flref_free(out)
    return o


# ----------------------------------------------------------------- container (tests)
def fl_container(n, bits, values) -> bytes:
    """FL file image: u64 inputSize, u64 bitsSize, u64 valuesSize, bits, values
    (/root/reference/src/file_io.cu:222-280)."""
    b, v = _as_u8(bits), _as_u8(values)
    return struct.pack("<QQQ", n, b.size, v.size) + b.tobytes() + v.tobytes()


def fl_container_parse(blob: bytes):
    n, nb, nv = struct.unpack_from("<QQQ", blob, 0)
    b = np.frombuffer(blob, np.uint8, nb, 24)
    v = np.frombuffer(blob, np.uint8, nv, 24 + nb)
    return n, b, v


def rl_container(n, counts, values) -> bytes:
    """RL file image (BUILD DECISION, the reference has none — SURVEY.md §8c):
    u64 inputSize, u64 pairCount, counts[pairCount], values[pairCount]."""
    c, v = _as_u8(counts), _as_u8(values)
    return struct.pack("<QQ", n, c.size) + c.tobytes() + v.tobytes()


def rl_container_parse(blob: bytes):
    n, p = struct.unpack_from("<QQ", blob, 0)
    c = np.frombuffer(blob, np.uint8, p, 16)
    v = np.frombuffer(blob, np.uint8, p, 16 + p)
    return n, c, v
