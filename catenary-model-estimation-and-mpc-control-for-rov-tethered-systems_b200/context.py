"""Device context: one CUDA device + one stream + library-owned scratch (cat_init / cat_destroy)."""
from __future__ import annotations

import ctypes as C
import threading
import weakref

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
        self._registered = {}     # address -> (nbytes, weakref.finalize) of page-locked caller arrays
        self._unlockable = set()  # addresses cat_host_register refused (already page-locked, ...): do not retry

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

    def set_overlap(self, on):
        """Programmatic dependent launch of the rollout kernels (cat_set_overlap; default on)."""
        self.check(self.lib.cat_set_overlap(self.handle, 1 if on else 0))

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

    # -- caller arrays ---------------------------------------------------------------------
    REGISTER_MIN_BYTES = 4 << 20

    def register(self, array):
        """Page-lock a caller-owned, C-contiguous ndarray in place (cat_host_register) so the host entry points copy
        it at PCIe speed; the registration ends when the array is garbage-collected (or at ``close``).  Returns True
        if the array is (now) registered.  Arrays that do not own their memory are registered through their base."""
        base = array
        while isinstance(getattr(base, "base", None), np.ndarray):
            base = base.base
        if not isinstance(base, np.ndarray) or not base.flags["C_CONTIGUOUS"] or base.nbytes == 0:
            return False
        addr, nbytes = base.ctypes.data, base.nbytes
        have = self._registered.get(addr)
        if have is not None and have[0] >= nbytes:
            return True
        if have is not None:
            have[1]()                                    # grown in place: re-register
        if addr in self._unlockable:
            return False
        rc = self.lib.cat_host_register(self.handle, C.c_void_p(addr), nbytes)
        if rc != _lib.CAT_OK:                            # pinned_empty memory, 64 ranges in use, unlockable pages
            self._unlockable.add(addr)
            return False
        lib, handle, table = self.lib, self.handle, self._registered

        def _release(addr=addr):
            if table.pop(addr, None) is not None and handle.value:
                lib.cat_host_unregister(handle, C.c_void_p(addr))
        try:
            fin = weakref.finalize(base, _release)
        except TypeError:                                # not weak-referenceable: do not keep a dangling lock
            lib.cat_host_unregister(handle, C.c_void_p(addr))
            return False
        self._registered[addr] = (nbytes, fin)
        return True

    def maybe_register(self, array):
        """``register`` for arrays large enough to be worth the ~0.1 ms/MB page-locking cost, seen for the first
        time; a no-op for small or already page-locked ones."""
        if array is None or array.nbytes < self.REGISTER_MIN_BYTES:
            return False
        return self.register(array)

    def stats(self, enable=True, reset=True):
        """Cold-path counters of the rollout kernel (cat_mpc_stats): dict(solve_cold, sine_cold) warp-level entries
        since the last reset; switches counting on/off for the launches that follow."""
        out = (C.c_uint64 * 4)()
        self.check(self.lib.cat_mpc_stats(self.handle, 1 if enable else 0, 1 if reset else 0, out))
        return {"solve_cold": int(out[0]), "sine_cold": int(out[1])}

    def measure_peak(self, kind):
        """kind: 'dfma' | 'ffma' | 'mufu' -> operations per second (FMA = 1 op)."""
        v = C.c_double()
        self.check(self.lib.cat_measure_peak(self.handle, {"dfma": 0, "ffma": 1, "mufu": 2}[kind], C.byref(v)))
        return float(v.value)

    def close(self):
        if getattr(self, "handle", None):
            for _, fin in list(self._registered.values()):
                fin()
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
