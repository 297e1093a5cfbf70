"""ctypes binding of include/mtldas.h (the drop-in boundary)."""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libmtldas_b200.so")


class MtldasError(RuntimeError):
    pass


def library_path():
    return _SO


def _load():
    if not os.path.exists(_SO):
        raise MtldasError(
            f"{_SO} is missing: build it with `python mtl-das.pytorch_b200/build.py` "
            "(or __graft_entry__.build()). There is no CPU or PyTorch fallback for this path.")
    L = ctypes.CDLL(_SO)
    c = ctypes
    vp, ci, cf, sz = c.c_void_p, c.c_int, c.c_float, c.c_size_t
    sig = {
        "mtldas_version": (c.c_char_p, []),
        "mtldas_last_error": (c.c_char_p, []),
        "mtldas_device_ok": (ci, []),
        "mtldas_plan_create": (vp, [ci, ci, ci]),
        "mtldas_plan_destroy": (None, [vp]),
        "mtldas_plan_num_state_entries": (ci, [vp]),
        "mtldas_plan_state_key": (c.c_char_p, [vp, ci]),
        "mtldas_plan_workspace_bytes": (sz, [vp]),
        "mtldas_plan_bind_workspace": (ci, [vp, vp, sz, vp]),
        "mtldas_plan_describe": (sz, [vp, c.c_char_p, sz]),
        "mtldas_plan_pack_eval": (ci, [vp, c.POINTER(vp), ci, vp]),
        "mtldas_plan_forward_eval": (ci, [vp, vp, ci, vp, vp, vp]),
        "mtldas_plan_forward_eval_host": (ci, [vp, vp, ci, vp, vp, vp]),
        "mtldas_plan_forward_eval_host_bf16": (ci, [vp, vp, ci, vp, vp, vp]),
        "mtldas_plan_profile_eval": (ci, [vp, vp, ci, vp, vp, vp, c.POINTER(cf), ci]),
        "mtldas_plan_last_launch_count": (ci, [vp]),
        "mtldas_plan_read_tensor": (ci, [vp, c.c_char_p, ci, vp, vp]),
        "mtldas_conv2d_nchw": (ci, [vp, vp, vp, vp, vp, vp, ci, ci, ci, ci, ci, ci, ci, ci, ci, ci, ci]),
        "mtldas_conv2d_nchw_static": (ci, [vp, vp, vp, vp, vp, vp, ci, ci, ci, ci, ci, ci, ci, ci, ci, ci, ci, vp]),
        "mtldas_train_create": (vp, [ci, ci, ci]),
        "mtldas_train_destroy": (None, [vp]),
        "mtldas_train_num_state_entries": (ci, [vp]),
        "mtldas_train_num_params": (sz, [vp]),
        "mtldas_train_workspace_bytes": (sz, [vp]),
        "mtldas_train_bind_workspace": (ci, [vp, vp, sz, vp]),
        "mtldas_train_forward": (ci, [vp, c.POINTER(vp), ci, vp, ci, vp, vp, vp]),
        "mtldas_train_backward": (ci, [vp, c.POINTER(vp), ci, vp, vp, vp, vp]),
        "mtldas_train_last_launch_count": (ci, [vp]),
        "mtldas_train_read_tensor": (ci, [vp, c.c_char_p, ci, vp, vp]),
        "mtldas_train_describe": (sz, [vp, c.c_char_p, sz]),
        "mtldas_plan_forward_eval_windows": (ci, [vp, vp, c.c_longlong, c.c_longlong, ci, ci, ci, vp, vp, vp]),
        "mtldas_window_gather": (ci, [vp, c.c_longlong, c.c_longlong, ci, ci, ci, vp, vp]),
        "mtldas_nll_loss": (ci, [vp, vp, ci, vp, vp, vp, ci, vp, ci, vp, vp, vp]),
        "mtldas_adam_step": (ci, [vp, vp, vp, vp, sz, ci, cf, cf, cf, cf, cf, vp]),
        "mtldas_reduce_adam_multimem": (ci, [vp, vp, vp, vp, vp, sz, ci, ci, ci, cf, cf, cf, cf, cf, vp]),
        "mtldas_eval_metrics": (ci, [vp, vp, ci, vp, vp, vp, ci, vp, ci, vp, vp, vp, vp, vp]),
        "mtldas_c_pack_conv": (ci, [vp, vp, vp, vp, vp, cf, ci, ci, ci, ci, vp, vp, vp]),
        "mtldas_c_conv": (ci, [vp, ci, ci, ci, ci, vp, vp, vp, ci, ci, ci, ci, ci, ci, ci, ci, ci, ci, vp]),
        "mtldas_c_pool": (ci, [vp, ci, ci, ci, ci, vp, ci, ci, ci, ci, ci, vp]),
        "mtldas_c_head": (ci, [vp, ci, ci, ci, vp, vp, ci, vp, vp]),
        "mtldas_c_read": (ci, [vp, ci, ci, ci, ci, ci, vp, vp]),
        "mtldas_umma_selftest": (cf, [ci, ci, ci]),
        "mtldas_umma_rate": (ci, [ci, ci, ci, c.POINTER(cf)]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype = res
        fn.argtypes = args
    L._declared = sorted(sig)
    return L


class _LazyLib:
    """The shared library is mapped on first use, not at package import: importing `mtldas_b200` (or its synthetic-data
    helpers) from a CPU-only process -- the oracle, the reference arm of bench.py -- must not load any native code."""

    _real = None

    def _get(self):
        if _LazyLib._real is None:
            _LazyLib._real = _load()
        return _LazyLib._real

    def __getattr__(self, name):
        return getattr(self._get(), name)


lib = _LazyLib()


def loaded():
    """True once libmtldas_b200.so has been mapped into this process."""
    return _LazyLib._real is not None


def check(rc, what):
    if rc != 0:
        raise MtldasError(f"{what} failed ({rc}): {lib.mtldas_last_error().decode()}")
