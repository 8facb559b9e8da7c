"""ctypes binding of libfdb200.so (the C ABI declared in include/fdb200.h).

There is no fallback: if the shared object is missing the import fails loudly, and if no
CUDA device is present every compute call raises ``RuntimeError``.
"""

from __future__ import annotations

import ctypes as C
import os
import threading

import numpy as np

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("FDB_LIB_PATH") or os.path.join(PKG_DIR, "libfdb200.so")   # override: A/B builds (tools/variant_build.py)

F32, F64 = 0, 1
N_FEATURES = 30

OK, ERR_ARG, ERR_ALIGN, ERR_CUDA, ERR_NOMEM, ERR_STATE = 0, -1, -2, -3, -4, -5


class FdbError(RuntimeError):
    """CUDA / state error raised by the native library."""


def _load() -> C.CDLL:
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build the CUDA extension first "
            "(python fraud_detection_b200/build.py). There is no CPU fallback.")
    return C.CDLL(LIB_PATH, mode=C.RTLD_LOCAL)


lib = _load()

_vp, _i, _i64, _dp = C.c_void_p, C.c_int, C.c_int64, C.POINTER(C.c_double)

_SIGS = {
    "fdb_version": (C.c_int, []),
    "fdb_last_error": (C.c_char_p, []),
    "fdb_ctx_create": (_i, [_i, C.POINTER(_vp)]),
    "fdb_ctx_destroy": (_i, [_vp]),
    "fdb_ctx_device": (_i, [_vp]),
    "fdb_ctx_sm_count": (_i, [_vp]),
    "fdb_pinned_alloc": (_i, [C.c_size_t, C.POINTER(_vp)]),
    "fdb_pinned_free": (_i, [_vp]),
    "fdb_transform_dev": (_i, [_vp, _vp, _i, _i64, _vp, _dp, _dp, _vp]),
    "fdb_transform_host": (_i, [_vp, _i, _i64, _vp, _dp, _dp, _vp]),
    "fdb_linear_pass_dev": (_i, [_vp, _vp, _i, _i64, _vp, _dp, _dp, C.c_double, _vp, _i, _vp, _vp, _vp, _vp]),
    "fdb_linear_pass_host": (_i, [_vp, _i, _i64, _vp, _dp, _dp, C.c_double, _vp, _i, _vp, _vp, _vp, _vp]),
    "fdb_linear_pass_noalloc": (_i, [_vp, _i, _i64, _vp, _dp, _dp, C.c_double, _vp, _i, _vp, _vp, _vp, _vp]),
    "fdb_kshap_plan_create": (_i, [_vp, _dp, C.c_double, _i, _vp, _i, _i, _vp, _vp, _i, C.POINTER(_vp)]),
    "fdb_kshap_plan_create_varying": (_i, [_vp, _dp, C.c_double, _i, _vp, _i, _i, _vp, _vp, _i, C.c_uint32, C.POINTER(_vp)]),
    "fdb_kshap_plan_create_sampled": (_i, [_vp, _dp, C.c_double, _i, _vp, _i, _i, C.c_uint64, _i, C.POINTER(_vp)]),
    "fdb_kshap_plan_get_masks": (_i, [_vp, _i, _vp, _vp]),
    "fdb_kshap_plan_resample": (_i, [_vp, C.c_uint64]),
    "fdb_kshap_plan_resample_dev": (_i, [_vp, _vp, C.c_uint64]),
    "fdb_kshap_plan_status": (_i, [_vp, C.POINTER(C.c_int)]),
    "fdb_kshap_plan_destroy": (_i, [_vp]),
    "fdb_kshap_expected_value": (C.c_double, [_vp]),
    "fdb_kshap_plan_info": (_i, [_vp, C.POINTER(C.c_int)]),
    "fdb_kshap_plan_set_num_features": (_i, [_vp, _i]),
    "fdb_kshap_plan_set_l1": (_i, [_vp, _i, C.c_double]),
    "fdb_format_json_rows": (_i, [_vp, _i64, _i, _vp, _i64, _vp]),
    "fdb_kshap_explain_dev": (_i, [_vp, _vp, _i64, _vp, _vp, _i, _vp, _vp, _vp]),
    "fdb_kshap_explain_host": (_i, [_vp, _i64, _vp, _vp, _i, _vp, _vp]),
    "fdb_probe_pipes": (_i, [_vp, _dp]),
}


def bind(sigs: dict) -> None:
    for name, (res, args) in sigs.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args


bind(_SIGS)


def check(rc: int) -> None:
    if rc == OK:
        return
    msg = (lib.fdb_last_error() or b"").decode("utf-8", "replace")
    if rc in (ERR_ARG, ERR_ALIGN):
        raise ValueError(msg)
    if rc == ERR_NOMEM:
        raise MemoryError(msg)
    raise FdbError(msg)


# ------------------------------------------------------------------ contexts
# One native context per (process, device).  Keyed by pid so that a forked gunicorn/Celery
# worker creates its own CUDA context lazily on first use instead of inheriting the parent's.
_ctx_lock = threading.Lock()
_ctxs: dict[tuple[int, int], int] = {}


def get_ctx(device: int | None = None) -> int:
    if device is None:
        device = int(os.environ.get("FDB_DEVICE", os.environ.get("LOCAL_RANK", "0")))
    key = (os.getpid(), int(device))
    with _ctx_lock:
        h = _ctxs.get(key)
        if h is None:
            out = _vp()
            check(lib.fdb_ctx_create(int(device), C.byref(out)))
            h = out.value
            _ctxs[key] = h
        return h


def sm_count(device: int | None = None) -> int:
    return int(lib.fdb_ctx_sm_count(get_ctx(device)))


def probe_pipes(device: int | None = None) -> dict:
    """Measured instruction-pipe rates of the device (fdb_probe_pipes): the Kernel SHAP roofline denominators."""
    out = (C.c_double * 8)()
    check(lib.fdb_probe_pipes(get_ctx(device), out))
    v = list(out)
    return {"ex2_per_s": v[0], "rcp_per_s": v[1], "ffma_per_s": v[2], "k2f_quad_16warps_sigmoids_per_s": v[3],
            "k2f_quad_lds_sigmoids_per_s": v[4], "sm_count": v[6]}


# ------------------------------------------------------------------ small helpers
def dvec(a) -> "C.Array":
    """30 doubles as a ctypes array (host parameter vectors of the ABI)."""
    a = np.ascontiguousarray(a, dtype=np.float64).reshape(-1)
    if a.shape[0] != N_FEATURES:
        raise ValueError(f"expected {N_FEATURES} parameters, got {a.shape[0]}")
    return (C.c_double * N_FEATURES)(*a.tolist())


def np_ptr(a: np.ndarray | None):
    return None if a is None else a.ctypes.data_as(_vp)


def dtype_code(dt) -> int:
    dt = np.dtype(dt)
    if dt == np.float32:
        return F32
    if dt == np.float64:
        return F64
    raise ValueError(f"unsupported dtype {dt}; use float32 or float64")


class PinnedBuffer:
    """Page-locked host memory exposed as a numpy array (freed with the object)."""

    def __init__(self, shape, dtype):
        self.shape = tuple(int(s) for s in np.atleast_1d(shape))
        self.dtype = np.dtype(dtype)
        nbytes = int(np.prod(self.shape)) * self.dtype.itemsize
        p = _vp()
        check(lib.fdb_pinned_alloc(max(nbytes, 1), C.byref(p)))
        self._ptr = p.value
        buf = (C.c_char * max(nbytes, 1)).from_address(self._ptr)
        self.array = np.frombuffer(buf, dtype=self.dtype, count=int(np.prod(self.shape))).reshape(self.shape)

    def __del__(self):
        ptr, self._ptr = getattr(self, "_ptr", None), None
        if ptr:
            try:
                lib.fdb_pinned_free(ptr)
            except Exception:
                pass


class host_near_gpu:
    """``with host_near_gpu(device): buf = PinnedBuffer(...)`` - run the enclosed host allocations on the CPUs next to
    the GPU (NVML's ideal CPU affinity), so page-locked buffers land on the socket the GPU's PCIe link hangs off
    instead of wherever the scheduler put the process; the previous affinity is restored on exit.  A no-op when
    NVML or the affinity call is not available (``.cpus`` is then None)."""

    def __init__(self, device=None):
        self.device = device
        self.cpus = None
        self._saved = None

    def __enter__(self):
        import os
        try:
            import pynvml
            import torch
            idx = torch.cuda.current_device() if self.device is None else int(self.device)
            pynvml.nvmlInit()
            try:
                h = pynvml.nvmlDeviceGetHandleByUUID("GPU-" + str(torch.cuda.get_device_properties(idx).uuid))
            except Exception:
                h = pynvml.nvmlDeviceGetHandleByIndex(idx)
            n_words = (os.cpu_count() + 63) // 64
            words = pynvml.nvmlDeviceGetCpuAffinity(h, n_words)
            ideal = {w * 64 + b for w, word in enumerate(words) for b in range(64) if (int(word) >> b) & 1}
            allowed = os.sched_getaffinity(0)
            cpus = ideal & allowed
            if cpus:
                self._saved = allowed
                os.sched_setaffinity(0, cpus)
                self.cpus = sorted(cpus)
        except Exception:
            self.cpus = None
        return self

    def __exit__(self, *exc):
        import os
        if self._saved is not None:
            try:
                os.sched_setaffinity(0, self._saved)
            except Exception:
                pass
        return False
