"""Python binding of the B200-native FL/RL codec (ctypes over the C ABI in include/flrl.h).

The directory name mirrors the reference repo and is not a Python identifier; import it with
``importlib.import_module("fl-rl-compression-mpi_b200")`` (see ``flrl_import()`` in
tests/conftest.py).  PyTorch is used only for device memory and streams.

There is no CPU fallback: if ``libflrl.so`` is missing or fails to load, importing this
module raises.  Build it with ``python -c "import __graft_entry__ as g; g.build()"`` or
``make -C fl-rl-compression-mpi_b200``.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("FLRL_LIB_PATH") or os.path.join(HERE, "libflrl.so")  # override: kernel A/B experiments
FRAME = 128

if not os.path.exists(LIB_PATH):
    raise ImportError(f"{LIB_PATH} not built — run `make -C {HERE}` (no CPU fallback exists)")
lib = C.CDLL(LIB_PATH)

_vp, _sz, _u8p = C.c_void_p, C.c_size_t, C.POINTER(C.c_uint8)

lib.flrl_last_error.restype = C.c_char_p
lib.flrl_version.restype = C.c_int
lib.flrl_fl_frames.restype = _sz
lib.flrl_fl_frames.argtypes = [_sz]
lib.flrl_fl_workspace_bytes.restype = _sz
lib.flrl_fl_workspace_bytes.argtypes = [_sz]
lib.flrl_fl_encode.restype = C.c_int
lib.flrl_fl_encode.argtypes = [_vp, _sz, _vp, _vp, _sz, _vp, _vp, _sz, _vp]
lib.flrl_fl_decode.restype = C.c_int
lib.flrl_fl_decode.argtypes = [_vp, _sz, _vp, _sz, _vp, _sz, _vp, _vp, _sz, _vp]
lib.flrl_fl_decode_static.restype = C.c_int  # no workspace, no allocation (static device scratch inside the library)
lib.flrl_fl_decode_static.argtypes = [_vp, _sz, _vp, _sz, _vp, _sz, _vp, _vp]
STATIC_MAX_BYTES = 1 << 30


class FlrlError(RuntimeError):
    pass


def _check(rc: int, what: str) -> None:
    if rc != 0:
        raise FlrlError(f"{what}: rc={rc}: {lib.flrl_last_error().decode(errors='replace')}")


def _torch():
    import torch

    if not torch.cuda.is_available():
        raise FlrlError("CUDA device required: the fl / rl methods have no CPU fallback")
    return torch


def _stream_ptr(torch) -> int:
    return torch.cuda.current_stream().cuda_stream


def _ptr(t) -> int:
    return t.data_ptr() if t.numel() else 0


# ------------------------------------------------------------------ device layer: FL
class FLEncoder:
    """Reusable buffers for flrl_fl_encode over inputs of up to `n` bytes on the current device."""

    def __init__(self, n: int):
        torch = _torch()
        self.n = n
        dev = torch.device("cuda", torch.cuda.current_device())
        self.bits = torch.empty(max(lib.flrl_fl_frames(n), 1), dtype=torch.uint8, device=dev)
        self.cap = (n + 15) // 16 * 16
        self.values = torch.empty(max(self.cap, 16), dtype=torch.uint8, device=dev)
        self.size = torch.zeros(1, dtype=torch.int64, device=dev)
        self.ws_bytes = lib.flrl_fl_workspace_bytes(n)
        self.ws = torch.empty(max(self.ws_bytes, 16), dtype=torch.uint8, device=dev)

    def launch(self, d_in, n: int | None = None) -> None:
        """Asynchronous on the current torch stream."""
        torch = _torch()
        n = d_in.numel() if n is None else n
        assert n <= self.n and d_in.dtype == torch.uint8 and d_in.is_cuda
        rc = lib.flrl_fl_encode(_ptr(d_in), n, _ptr(self.bits), _ptr(self.values), self.cap, _ptr(self.size), _ptr(self.ws),
                                self.ws_bytes, _stream_ptr(torch))
        _check(rc, "flrl_fl_encode")

    def result(self, n: int):
        """(bits, values) as device tensors (synchronises to read the size)."""
        nv = int(self.size.item())
        return self.bits[: lib.flrl_fl_frames(n)], self.values[:nv]


class FLDecoder:
    def __init__(self, n: int):
        torch = _torch()
        self.n = n
        dev = torch.device("cuda", torch.cuda.current_device())
        self.out = torch.empty(max((n + 15) // 16 * 16, 16), dtype=torch.uint8, device=dev)
        self.status = torch.zeros(1, dtype=torch.int32, device=dev)
        self.ws_bytes = lib.flrl_fl_workspace_bytes(n)
        self.ws = torch.empty(max(self.ws_bytes, 16), dtype=torch.uint8, device=dev)

    def launch(self, d_bits, d_values, n: int) -> None:
        torch = _torch()
        assert n <= self.n
        rc = lib.flrl_fl_decode(_ptr(d_bits), d_bits.numel(), _ptr(d_values), d_values.numel(), _ptr(self.out), n,
                                _ptr(self.status), _ptr(self.ws), self.ws_bytes, _stream_ptr(torch))
        _check(rc, "flrl_fl_decode")

    def result(self, n: int):
        st = int(self.status.item())
        if st:
            raise FlrlError(f"flrl_fl_decode: malformed container (status {st})")
        return self.out[:n]


def fl_encode_device(d_in):
    """One-shot: uint8 CUDA tensor -> (bits, values) CUDA tensors."""
    enc = FLEncoder(d_in.numel())
    enc.launch(d_in)
    return enc.result(d_in.numel())


def fl_decode_device(d_bits, d_values, n: int):
    dec = FLDecoder(n)
    dec.launch(d_bits, d_values, n)
    return dec.result(n)


def _aligned_values(torch, v):
    """Device copies of compressed arrays must be 16-byte aligned for the TMA bulk loads."""
    t = torch.empty((v.numel() + 15) // 16 * 16 + 16, dtype=torch.uint8, device=v.device)
    t[: v.numel()] = v
    return t[: v.numel()]


def to_device(a):
    torch = _torch()
    a = np.ascontiguousarray(np.frombuffer(a, dtype=np.uint8) if isinstance(a, (bytes, bytearray, memoryview)) else a)
    t = torch.empty(max((a.size + 15) // 16 * 16, 16), dtype=torch.uint8, device="cuda")
    if a.size:
        t[: a.size] = torch.from_numpy(a.copy())
    return t[: a.size]


# ------------------------------------------------------------------ device layer: RL
lib.flrl_rl_workspace_bytes.restype = _sz
lib.flrl_rl_workspace_bytes.argtypes = [_sz]
lib.flrl_rl_decode_workspace_bytes.restype = _sz
lib.flrl_rl_decode_workspace_bytes.argtypes = [_sz]
lib.flrl_rl_encode.restype = C.c_int
lib.flrl_rl_encode.argtypes = [_vp, _sz, _sz, C.c_uint64, _vp, _vp, _sz, _vp, _vp, _sz, _vp]
lib.flrl_rl_decode.restype = C.c_int
lib.flrl_rl_decode.argtypes = [_vp, _vp, _sz, _vp, _sz, _vp, _vp, _vp, _sz, _vp]


class RLEncoder:
    """Reusable buffers for flrl_rl_encode over shards of up to `n` bytes on the current device."""

    def __init__(self, n: int):
        torch = _torch()
        self.n = n
        dev = torch.device("cuda", torch.cuda.current_device())
        self.cap = max((n + 15) // 16 * 16, 16)
        self.counts = torch.empty(self.cap, dtype=torch.uint8, device=dev)
        self.values = torch.empty(self.cap, dtype=torch.uint8, device=dev)
        self.size = torch.zeros(1, dtype=torch.int64, device=dev)
        self.ws_bytes = lib.flrl_rl_workspace_bytes(n)
        self.ws = torch.empty(max(self.ws_bytes, 16), dtype=torch.uint8, device=dev)

    def launch(self, d_in, n: int | None = None, prev_run: int = 0) -> None:
        """Encode d_in[:n]; bytes d_in[n:] are look-ahead only (shard semantics of include/flrl.h)."""
        torch = _torch()
        n_avail = d_in.numel()
        n = n_avail if n is None else n
        assert n <= self.n and d_in.dtype == torch.uint8 and d_in.is_cuda
        rc = lib.flrl_rl_encode(_ptr(d_in), n, n_avail, prev_run, _ptr(self.counts), _ptr(self.values), self.cap,
                                _ptr(self.size), _ptr(self.ws), self.ws_bytes, _stream_ptr(torch))
        _check(rc, "flrl_rl_encode")

    def result(self):
        p = int(self.size.item())
        return self.counts[:p], self.values[:p]


class RLDecoder:
    def __init__(self, out_cap: int, max_pairs: int | None = None):
        torch = _torch()
        dev = torch.device("cuda", torch.cuda.current_device())
        self.out_cap = out_cap
        self.out = torch.empty(max((out_cap + 15) // 16 * 16, 16), dtype=torch.uint8, device=dev)
        self.size = torch.zeros(1, dtype=torch.int64, device=dev)
        self.status = torch.zeros(1, dtype=torch.int32, device=dev)
        self.ws_bytes = lib.flrl_rl_decode_workspace_bytes(out_cap if max_pairs is None else max_pairs)
        self.ws = torch.empty(max(self.ws_bytes, 16), dtype=torch.uint8, device=dev)

    def launch(self, d_counts, d_values) -> None:
        torch = _torch()
        assert d_counts.numel() == d_values.numel()
        rc = lib.flrl_rl_decode(_ptr(d_counts), _ptr(d_values), d_counts.numel(), _ptr(self.out), self.out_cap, _ptr(self.size),
                                _ptr(self.status), _ptr(self.ws), self.ws_bytes, _stream_ptr(torch))
        _check(rc, "flrl_rl_decode")

    def result(self):
        st = int(self.status.item())
        if st:
            raise FlrlError(f"flrl_rl_decode: malformed pair stream or output too small (status {st})")
        return self.out[: int(self.size.item())]


def rl_encode_device(d_in, n: int | None = None, prev_run: int = 0):
    enc = RLEncoder(d_in.numel() if n is None else n)
    enc.launch(d_in, n, prev_run)
    return enc.result()


def rl_decode_device(d_counts, d_values, out_cap: int):
    dec = RLDecoder(out_cap, max(d_counts.numel(), 1))
    dec.launch(d_counts, d_values)
    return dec.result()


# ------------------------------------------------------------------------- host layer
_szp = C.POINTER(C.c_size_t)
lib.flrl_host_alloc.restype = _vp
lib.flrl_host_alloc.argtypes = [_sz]
lib.flrl_host_free.argtypes = [_vp]
lib.flrl_free.argtypes = [_vp]
lib.flrl_last_timing.argtypes = [C.POINTER(C.c_float), C.POINTER(C.c_float)]
lib.flrl_set_concat_mode.argtypes = [C.c_int]
lib.flrl_set_pipe_chunk.argtypes = [_sz]
lib.flrl_prepare.restype = C.c_int
lib.flrl_prepare.argtypes = [C.c_int]
lib.flrl_last_gpu_count.restype = C.c_int
lib.flrl_last_timing_gpu.restype = C.c_int
lib.flrl_last_timing_gpu.argtypes = [C.c_int, C.POINTER(C.c_float), C.POINTER(C.c_float), C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
lib.flrl_fl_gpu_compress_into.restype = C.c_int
lib.flrl_fl_gpu_compress_into.argtypes = [_vp, _sz, C.c_int, _vp, _sz, _vp, _sz, _szp, _szp]
lib.flrl_fl_gpu_decompress_into.restype = C.c_int
lib.flrl_fl_gpu_decompress_into.argtypes = [_sz, _vp, _sz, _vp, _sz, C.c_int, _vp, _sz, _szp]
lib.flrl_rl_gpu_compress_into.restype = C.c_int
lib.flrl_rl_gpu_compress_into.argtypes = [_vp, _sz, C.c_int, _vp, _vp, _sz, _szp]
lib.flrl_rl_gpu_decompress_into.restype = C.c_int
lib.flrl_rl_gpu_decompress_into.argtypes = [_vp, _vp, _sz, C.c_int, _vp, _sz, _szp]
_u8pp = C.POINTER(_u8p)
lib.flrl_fl_gpu_compress.restype = C.c_int
lib.flrl_fl_gpu_compress.argtypes = [_vp, _sz, C.c_int, _u8pp, _szp, _u8pp, _szp]
lib.flrl_fl_gpu_decompress.restype = C.c_int
lib.flrl_fl_gpu_decompress.argtypes = [_sz, _vp, _sz, _vp, _sz, C.c_int, _u8pp, _szp]
lib.flrl_rl_gpu_compress.restype = C.c_int
lib.flrl_rl_gpu_compress.argtypes = [_vp, _sz, C.c_int, _u8pp, _u8pp, _szp]
lib.flrl_rl_gpu_decompress.restype = C.c_int
lib.flrl_rl_gpu_decompress.argtypes = [_vp, _vp, _sz, _sz, C.c_int, _u8pp, _szp]
for _n in ("fl", "rl"):
    getattr(lib, f"flrl_{_n}_cpu_compress").restype = C.c_int
    getattr(lib, f"flrl_{_n}_cpu_decompress").restype = C.c_int
lib.flrl_fl_cpu_compress.argtypes = [_vp, _sz, _u8pp, _szp, _u8pp, _szp]
lib.flrl_fl_cpu_decompress.argtypes = [_sz, _vp, _sz, _vp, _sz, _u8pp, _szp]
lib.flrl_rl_cpu_compress.argtypes = [_vp, _sz, _u8pp, _u8pp, _szp]
lib.flrl_rl_cpu_decompress.argtypes = [_vp, _vp, _sz, _u8pp, _szp]


def _take(ptr, n) -> np.ndarray:
    """Copy a malloc-owned result out and release it with flrl_free (free(), as the reference's callers do)."""
    try:
        return np.ctypeslib.as_array(ptr, (n,)).copy() if n else np.empty(0, np.uint8)
    finally:
        lib.flrl_free(ptr)


def _np_u8(a) -> np.ndarray:
    return np.ascontiguousarray(np.frombuffer(a, dtype=np.uint8) if isinstance(a, (bytes, bytearray, memoryview)) else a, dtype=np.uint8)


def fl_gpu_compress(data, ngpus: int = 1):
    """flrl_fl_gpu_compress (the reference's gpuCompress shape): host bytes -> (bits, values) numpy arrays."""
    a = _np_u8(data)
    bits, values, nb, nv = _u8p(), _u8p(), C.c_size_t(), C.c_size_t()
    _check(lib.flrl_fl_gpu_compress(a.ctypes.data, a.size, ngpus, C.byref(bits), C.byref(nb), C.byref(values), C.byref(nv)),
           "flrl_fl_gpu_compress")
    return _take(bits, nb.value), _take(values, nv.value)


def fl_gpu_decompress(n: int, bits, values, ngpus: int = 1):
    b, v = _np_u8(bits), _np_u8(values)
    out, no = _u8p(), C.c_size_t()
    _check(lib.flrl_fl_gpu_decompress(n, b.ctypes.data, b.size, v.ctypes.data, v.size, ngpus, C.byref(out), C.byref(no)),
           "flrl_fl_gpu_decompress")
    return _take(out, no.value)


def rl_gpu_compress(data, ngpus: int = 1):
    a = _np_u8(data)
    counts, values, p = _u8p(), _u8p(), C.c_size_t()
    _check(lib.flrl_rl_gpu_compress(a.ctypes.data, a.size, ngpus, C.byref(counts), C.byref(values), C.byref(p)), "flrl_rl_gpu_compress")
    return _take(counts, p.value), _take(values, p.value)


def rl_gpu_decompress(counts, values, expected_size: int = 0, ngpus: int = 1):
    c, v = _np_u8(counts), _np_u8(values)
    out, no = _u8p(), C.c_size_t()
    _check(lib.flrl_rl_gpu_decompress(c.ctypes.data, v.ctypes.data, c.size, expected_size, ngpus, C.byref(out), C.byref(no)),
           "flrl_rl_gpu_decompress")
    return _take(out, no.value)


def fl_cpu_compress(data):
    a = _np_u8(data)
    bits, values, nb, nv = _u8p(), _u8p(), C.c_size_t(), C.c_size_t()
    _check(lib.flrl_fl_cpu_compress(a.ctypes.data, a.size, C.byref(bits), C.byref(nb), C.byref(values), C.byref(nv)), "flrl_fl_cpu_compress")
    return _take(bits, nb.value), _take(values, nv.value)


def fl_cpu_decompress(n: int, bits, values):
    b, v = _np_u8(bits), _np_u8(values)
    out, no = _u8p(), C.c_size_t()
    _check(lib.flrl_fl_cpu_decompress(n, b.ctypes.data, b.size, v.ctypes.data, v.size, C.byref(out), C.byref(no)), "flrl_fl_cpu_decompress")
    return _take(out, no.value)


def rl_cpu_compress(data):
    a = _np_u8(data)
    counts, values, p = _u8p(), _u8p(), C.c_size_t()
    _check(lib.flrl_rl_cpu_compress(a.ctypes.data, a.size, C.byref(counts), C.byref(values), C.byref(p)), "flrl_rl_cpu_compress")
    return _take(counts, p.value), _take(values, p.value)


def rl_cpu_decompress(counts, values):
    c, v = _np_u8(counts), _np_u8(values)
    out, no = _u8p(), C.c_size_t()
    _check(lib.flrl_rl_cpu_decompress(c.ctypes.data, v.ctypes.data, c.size, C.byref(out), C.byref(no)), "flrl_rl_cpu_decompress")
    return _take(out, no.value)


class PinnedBuffer:
    """flrl_host_alloc'ed host memory viewed as a numpy array (the host layer copies straight to / from it)."""

    def __init__(self, nbytes: int):
        self.ptr = lib.flrl_host_alloc(max(nbytes, 1))
        if not self.ptr:
            raise FlrlError("flrl_host_alloc failed: " + lib.flrl_last_error().decode(errors="replace"))
        self.nbytes = nbytes
        self.array = np.ctypeslib.as_array(C.cast(self.ptr, _u8p), (max(nbytes, 1),))[:nbytes]

    def close(self):
        if self.ptr:
            lib.flrl_host_free(self.ptr)
            self.ptr = None

    def __del__(self):
        self.close()


class HostCodec:
    """End-to-end round trip through the host layer with pinned HOST buffers (bench.py's e2e leg, the CLI's shape):
    compress() = flrl_{fl,rl}_gpu_compress_into, decompress() = flrl_{fl,rl}_gpu_decompress_into."""

    def __init__(self, kind: str, n: int, ngpus: int = 1):
        assert kind in ("fl", "rl")
        self.kind, self.n, self.ngpus = kind, n, ngpus
        self.src = PinnedBuffer(n)
        self.back = PinnedBuffer(n)
        self.meta = PinnedBuffer((n + 127) // 128 if kind == "fl" else n)  # bits | counts
        self.payload = PinnedBuffer(n + 16)                                # values
        self.s0, self.s1 = C.c_size_t(), C.c_size_t()
        self.out_size = C.c_size_t()

    def load_input(self, src):
        """src: numpy array or CUDA tensor of n bytes."""
        if hasattr(src, "cpu"):
            src = src.cpu().numpy()
        self.src.array[:] = _np_u8(src)

    def compress(self):
        if self.kind == "fl":
            _check(lib.flrl_fl_gpu_compress_into(self.src.ptr, self.n, self.ngpus, self.meta.ptr, self.meta.nbytes, self.payload.ptr,
                                                 self.payload.nbytes, C.byref(self.s0), C.byref(self.s1)), "flrl_fl_gpu_compress_into")
        else:
            _check(lib.flrl_rl_gpu_compress_into(self.src.ptr, self.n, self.ngpus, self.meta.ptr, self.payload.ptr, self.n,
                                                 C.byref(self.s0)), "flrl_rl_gpu_compress_into")
            self.s1.value = self.s0.value

    def decompress(self):
        if self.kind == "fl":
            _check(lib.flrl_fl_gpu_decompress_into(self.n, self.meta.ptr, self.s0.value, self.payload.ptr, self.s1.value, self.ngpus,
                                                   self.back.ptr, self.n, C.byref(self.out_size)), "flrl_fl_gpu_decompress_into")
        else:
            _check(lib.flrl_rl_gpu_decompress_into(self.meta.ptr, self.payload.ptr, self.s0.value, self.ngpus, self.back.ptr, self.n,
                                                   C.byref(self.out_size)), "flrl_rl_gpu_decompress_into")

    def compressed(self):
        return self.meta.array[: self.s0.value], self.payload.array[: self.s1.value]

    def compressed_bytes(self) -> int:
        return self.s0.value + self.s1.value

    def check_round_trip(self) -> bool:
        return self.out_size.value == self.n and bool(np.array_equal(self.back.array, self.src.array))

    def timing(self):
        k, t = C.c_float(), C.c_float()
        lib.flrl_last_timing(C.byref(k), C.byref(t))
        return k.value, t.value

    def timing_per_gpu(self):
        """[(kernel_ms, total_ms, h2d_bytes, d2h_bytes)] for every GPU of the last call."""
        res = []
        for g in range(lib.flrl_last_gpu_count()):
            k, t, bi, bo = C.c_float(), C.c_float(), C.c_uint64(), C.c_uint64()
            _check(lib.flrl_last_timing_gpu(g, C.byref(k), C.byref(t), C.byref(bi), C.byref(bo)), "flrl_last_timing_gpu")
            res.append((k.value, t.value, bi.value, bo.value))
        return res
