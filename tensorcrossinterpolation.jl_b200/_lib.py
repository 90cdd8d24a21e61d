"""ctypes binding of libtci_b200.so (the C ABI declared in include/tci_b200.h).

There is no CPU fallback: if the shared library is missing or no CUDA device is present, creating a
Handle raises.  Device buffers are torch CUDA tensors (torch is plumbing here: memory and streams).
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(_HERE, "libtci_b200.so")

TCI_OK = 0
TCI_ERR_INVALID, TCI_ERR_CUDA, TCI_ERR_NAN_L, TCI_ERR_NAN_U = -1, -2, -3, -4
TCI_ERR_NOT_SQUARE, TCI_ERR_NAN_TENSOR, TCI_ERR_NVRTC, TCI_ERR_NOMEM = -5, -6, -7, -8


class TCIError(RuntimeError):
    """Raised where the reference throws ErrorException / ArgumentError; `.code` is the tci_status."""

    def __init__(self, code, msg):
        super().__init__(msg)
        self.code = code


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(SO_PATH):
            raise ImportError(
                f"{SO_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(nvcc, sm_100a). There is no CPU fallback.")
        L = C.CDLL(SO_PATH)
        L.tci_last_error.restype = C.c_char_p
        L.tci_last_error.argtypes = [C.c_void_p]
        L.tci_version.restype = C.c_char_p
        L.tci_launch_count.restype = C.c_int64
        L.tci_launch_count.argtypes = [C.c_void_p]
        _lib = L
    return _lib


def p_f64(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def p_i64(a):
    return a.ctypes.data_as(C.POINTER(C.c_int64))


def p_i32(a):
    return a.ctypes.data_as(C.POINTER(C.c_int32))


def _raw_stream(device):
    """torch's current stream on `device` as an integer.  torch._C._cuda_getCurrentRawStream is the sub-microsecond accessor
    (torch.cuda.current_stream builds a Stream object: ~10 us, and Handle.h is read before EVERY library call — 1.5 M times in
    one contract_TCI run at BASELINE C5 size)."""
    import torch
    get = getattr(torch._C, "_cuda_getCurrentRawStream", None)
    if get is not None:
        return int(get(int(device)))
    return int(torch.cuda.current_stream(device).cuda_stream)


class Handle:
    """One tci_handle = one CUDA device + stream + workspaces (not thread-safe)."""

    def __init__(self, device=0):
        self._h = C.c_void_p()
        rc = lib().tci_create(C.byref(self._h), int(device))
        if rc != TCI_OK:
            raise TCIError(rc, lib().tci_last_error(None).decode())
        self.device = int(device)
        self._stream = None
        # torch owns the device buffers, so run on torch's current stream: its copies and our kernels stay ordered
        self.use_torch_stream()

    @property
    def h(self):
        """The raw tci_handle for a library call.  Every call site reads this property right before the call, so it is also
        where the handle follows torch's CURRENT stream: under `with torch.cuda.stream(s)` torch's copies go to `s`, and
        library kernels left on the stream bound at construction would race with them (read-before-write on Pi and the
        factors).  One pointer compare per call when nothing changed."""
        try:
            s = _raw_stream(self.device)
        except Exception:                                   # no CUDA device: calls fail loudly in the library itself
            return self._h
        if s != self._stream:
            self._stream = s
            lib().tci_set_stream(self._h, C.c_void_p(s))
        return self._h

    def check(self, rc):
        if rc != TCI_OK:
            raise TCIError(rc, lib().tci_last_error(self._h).decode())

    def use_torch_stream(self):
        """Launch on torch's current stream so torch.cuda.Event timing brackets our kernels."""
        import torch
        s = torch.cuda.current_stream(self.device).cuda_stream
        self._stream = s
        self.check(lib().tci_set_stream(self._h, C.c_void_p(s)))

    def synchronize(self):
        self.check(lib().tci_synchronize(self._h))

    def launch_count(self):
        return int(lib().tci_launch_count(self._h))

    def close(self):
        if self._h:
            lib().tci_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


_default = {}


def default_handle(device=0):
    if device not in _default:
        _default[device] = Handle(device)
    return _default[device]
