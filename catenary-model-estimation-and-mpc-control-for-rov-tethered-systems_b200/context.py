"""Device context: one CUDA device + one stream + library-owned scratch (cat_init / cat_destroy)."""
from __future__ import annotations

import ctypes as C
import threading

import numpy as np

from . import _lib


class Context:
    def __init__(self, device=0):
        self.lib = _lib.load()
        h = C.c_void_p()
        rc = self.lib.cat_init(int(device), C.byref(h))
        if rc != _lib.CAT_OK:
            raise _lib.CatenaryLibraryError(
                {_lib.CAT_ENODEV: "no CUDA device (libcatenary_b200 has no CPU fallback)",
                 _lib.CAT_ECUDA: "CUDA runtime error in cat_init",
                 _lib.CAT_EINVAL: f"bad device index {device}"}.get(rc, f"cat_init failed ({rc})"))
        self.handle = h
        self.device = int(device)
        self._pinned = []

    # -- plumbing --------------------------------------------------------------------------
    def check(self, rc):
        if rc != _lib.CAT_OK:
            msg = self.lib.cat_last_error(self.handle)
            msg = msg.decode() if msg else ""
            if rc == _lib.CAT_EINVAL:
                raise ValueError(msg or "invalid argument")
            if rc == _lib.CAT_ENOMEM:
                raise MemoryError(msg)
            raise _lib.CatenaryLibraryError(f"libcatenary_b200 error {rc}: {msg}")

    def sync(self):
        self.check(self.lib.cat_sync(self.handle))

    @property
    def launch_count(self):
        return int(self.lib.cat_launch_count(self.handle))

    @property
    def stream(self):
        return self.lib.cat_stream(self.handle)

    def set_stream(self, cuda_stream):
        """Run subsequent calls on a caller-owned CUDA stream (an integer handle; 0 = the legacy default
        stream).  ``None`` restores the context's own stream."""
        if cuda_stream is None:
            self.check(self.lib.cat_set_stream(self.handle, None, 1))
        else:
            self.check(self.lib.cat_set_stream(self.handle, C.c_void_p(int(cuda_stream)), 0))

    def timer_start(self):
        self.check(self.lib.cat_timer_start(self.handle))

    def timer_stop_ms(self):
        ms = C.c_float()
        self.check(self.lib.cat_timer_stop_ms(self.handle, C.byref(ms)))
        return float(ms.value)

    def pinned_empty(self, shape, dtype=np.float64):
        """ndarray backed by page-locked host memory (cat_host_alloc); freed with the context."""
        dtype = np.dtype(dtype)
        n = int(np.prod(shape)) * dtype.itemsize
        p = C.c_void_p()
        self.check(self.lib.cat_host_alloc(self.handle, max(n, 16), C.byref(p)))
        self._pinned.append(p.value)
        buf = (C.c_char * max(n, 1)).from_address(p.value)
        return np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)

    def measure_peak(self, kind):
        """kind: 'dfma' | 'ffma' | 'mufu' -> operations per second (FMA = 1 op)."""
        v = C.c_double()
        self.check(self.lib.cat_measure_peak(self.handle, {"dfma": 0, "ffma": 1, "mufu": 2}[kind], C.byref(v)))
        return float(v.value)

    def close(self):
        if getattr(self, "handle", None):
            for p in self._pinned:
                self.lib.cat_host_free(self.handle, C.c_void_p(p))
            self._pinned = []
            self.lib.cat_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


_default = {}
_dlock = threading.Lock()


def default_context(device=0):
    """Process-wide context of a device, created on first use."""
    with _dlock:
        ctx = _default.get(device)
        if ctx is None or ctx.handle is None:
            ctx = _default[device] = Context(device)
        return ctx
