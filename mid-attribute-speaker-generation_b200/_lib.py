"""ctypes binding of libmasg.so (include/masg.h).  Fails loudly when the CUDA library is missing or no
B200 is present: there is no CPU fallback and no alternate backend."""
import ctypes
import os
import threading

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libmasg.so")

PREC_BF16, PREC_FP32, PREC_BF16X3 = 0, 1, 2
_PREC_NAMES = {"bf16": PREC_BF16, "fp32": PREC_FP32, "bf16x3": PREC_BF16X3, PREC_BF16: PREC_BF16, PREC_FP32: PREC_FP32, PREC_BF16X3: PREC_BF16X3}

c_void_p, c_int, c_float, c_i64, c_u64, c_char_p = (ctypes.c_void_p, ctypes.c_int, ctypes.c_float, ctypes.c_int64,
                                                     ctypes.c_uint64, ctypes.c_char_p)
P = c_void_p

# name → (restype, argtypes); must list every symbol include/masg.h declares (tests check this)
SIGNATURES = {
    "masg_version": (c_char_p, []),
    "masg_ctx_create": (c_int, [c_int, c_int, ctypes.POINTER(P)]),
    "masg_ctx_destroy": (None, [P]),
    "masg_last_error": (c_char_p, [P]),
    "masg_load_tensor": (c_int, [P, c_char_p, c_char_p, P, c_int, ctypes.POINTER(c_i64)]),
    "masg_fs2_commit": (c_int, [P, P]),
    "masg_hifigan_commit": (c_int, [P, P]),
    "masg_fs2_phase_a": (c_int, [P, P, P, P, c_int, P, c_int, c_int, P, c_float, c_float, c_float, P, P, P,
                                 P, P, P, P, P, P, ctypes.POINTER(c_i64)]),
    "masg_fs2_phase_b": (c_int, [P, P, c_int, P, P, P]),
    "masg_duration_round": (c_int, [P, P, P, c_float, c_i64, P]),
    "masg_mel_lengths": (c_int, [P, P, P, c_int, c_int, P]),
    "masg_length_regulate": (c_int, [P, P, P, P, c_int, c_int, c_int, c_int, P, P, P]),
    "masg_bucketize_embed_add": (c_int, [P, P, P, P, c_int, P, c_i64, c_int, P, P]),
    "masg_hifigan_forward": (c_int, [P, P, P, c_int, c_int, c_int, P, P]),
    "masg_hifigan_forward_packed": (c_int, [P, P, P, c_int, c_int, c_int, P, c_int, P, P]),
    "masg_gmm_params": (c_int, [P, P, P, c_int, P, P, P]),
    "masg_gmm_sample": (c_int, [P, P, P, P, P, c_int, c_int, c_int, c_int, c_u64, c_u64, P, P, P, P]),
    "masg_gmm_interpolate": (c_int, [P, P, P, P, P, P, P, P, c_int, c_int, c_float, c_int, P, P, P, P]),
    "masg_gmm_barycenter": (c_int, [P, P, P, P, P, P, c_int, c_int, c_int, P, P, P, P]),
    "masg_gmm_barycenter_noalloc": (c_int, [P, P, P, P, P, P, c_int, c_int, c_int, P, P, P, P]),
    "masg_synthesize_host": (c_int, [P, P, P, P, P, c_int, c_int, P, c_float, c_float, c_float, c_int, P, P,
                                     ctypes.POINTER(c_i64)]),
    "masg_synthesize_host_packed": (c_int, [P, P, P, P, P, c_int, c_int, P, c_float, c_float, c_float, c_i64, P, P, P,
                                            ctypes.POINTER(c_i64)]),
    "masg_synthesize_host_packed_async": (c_int, [P, P, P, P, P, c_int, c_int, P, c_float, c_float, c_float, c_i64, P, P, P,
                                                  ctypes.POINTER(c_i64), ctypes.POINTER(c_int)]),
    "masg_synthesize_host_wait": (c_int, [P, c_int]),
    "masg_conv1d": (c_int, [P, P, P, c_int, c_int, c_int, P, P, c_int, c_int, c_int, c_int, c_float, c_int, c_int, c_int, P,
                            ctypes.POINTER(c_float)]),
    "masg_attention": (c_int, [P, P, P, P, c_int, c_int, P]),
    "masg_attention_planar": (c_int, [P, P, P, P, c_int, c_int, P]),
    "masg_respair": (c_int, [P, P, P, P, c_int, c_int, c_int, P, P, P, P, c_int, c_int, c_float, c_float, c_int, P, P]),
    "masg_profile": (c_int, [P, c_int]),
    "masg_profile_read": (c_int, [P, P, ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_double), ctypes.POINTER(c_i64)]),
    "masg_launch_count": (c_i64, [P, c_int]),
    "masg_debug_enable_taps": (c_int, [P, c_int]),
    "masg_debug_tap": (c_int, [P, P, c_char_p, P, c_i64]),
    "masg_frontend_create": (c_int, [ctypes.POINTER(c_char_p), c_int, ctypes.POINTER(P)]),
    "masg_frontend_destroy": (None, [P]),
    "masg_frontend_last_error": (c_char_p, [P]),
    "masg_frontend_tokenize": (c_int, [P, c_char_p, c_i64, P, c_int, c_int, c_int, c_int, P, P, P]),
}

_lib = None
_lock = threading.Lock()


def load_library():
    """dlopen libmasg.so and bind every entry point; raises if the extension is not built."""
    global _lib
    with _lock:
        if _lib is None:
            if not os.path.exists(LIB_PATH):
                raise RuntimeError(
                    "libmasg.so is not built (%s): run `python __graft_entry__.py` / build.py. The CUDA extension is "
                    "required; there is no CPU fallback." % LIB_PATH)
            lib = ctypes.CDLL(LIB_PATH)
            for name, (res, args) in SIGNATURES.items():
                fn = getattr(lib, name)
                fn.restype = res
                fn.argtypes = args
            _lib = lib
    return _lib


def ptr(t):
    """device/host pointer of a contiguous tensor (None → NULL)."""
    if t is None:
        return None
    assert t.is_contiguous(), "non-contiguous tensor passed to libmasg"
    return ctypes.c_void_p(t.data_ptr())


def stream_ptr(device):
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


class Context:
    """One masg_ctx per (device, precision) — owns repacked weights and the workspace arena."""

    def __init__(self, device, precision="bf16"):
        self.lib = load_library()
        if not torch.cuda.is_available():
            raise RuntimeError("libmasg needs a CUDA device (B200, sm_100a); no CPU fallback exists")
        dev = torch.device(device)
        if dev.type != "cuda":
            raise RuntimeError("libmasg contexts live on CUDA devices, got %s" % (device,))
        self.index = dev.index if dev.index is not None else torch.cuda.current_device()
        self.device = torch.device("cuda", self.index)
        self.precision = _PREC_NAMES[precision]
        h = P()
        rc = self.lib.masg_ctx_create(self.index, self.precision, ctypes.byref(h))
        if rc != 0:
            raise RuntimeError("masg_ctx_create: " + self.lib.masg_last_error(None).decode())
        self.h = h
        self._keep = {}          # keeps registered tensors alive while the context references them

    def check(self, rc, what):
        if rc != 0:
            raise RuntimeError("%s: %s" % (what, self.lib.masg_last_error(self.h).decode()))

    def stream(self):
        return stream_ptr(self.device)

    def load_state(self, model, state_dict):
        keep = {}
        for k, v in state_dict.items():
            if not torch.is_floating_point(v):
                continue                                   # num_batches_tracked
            t = v.detach()
            if t.dtype != torch.float32 or not t.is_contiguous():
                t = t.float().contiguous()
            if t.device != self.device:
                raise RuntimeError("parameter %s is on %s, context is on %s (no CPU fallback)" % (k, t.device, self.device))
            shape = (c_i64 * max(t.dim(), 1))(*t.shape)
            self.check(self.lib.masg_load_tensor(self.h, model.encode(), k.encode(), ptr(t), t.dim(), shape), "masg_load_tensor(%s)" % k)
            keep[k] = t
        self._keep[model] = keep
        if model == "fs2":
            self.check(self.lib.masg_fs2_commit(self.h, self.stream()), "masg_fs2_commit")
        else:
            self.check(self.lib.masg_hifigan_commit(self.h, self.stream()), "masg_hifigan_commit")

    def launch_count(self, reset=False):
        return int(self.lib.masg_launch_count(self.h, 1 if reset else 0))

    def profile(self, enable):
        self.check(self.lib.masg_profile(self.h, 1 if enable else 0), "masg_profile")

    def profile_read(self):
        """→ (device ms, algorithmic FLOPs, launches) of the dense-contraction kernel since the last read."""
        ms, fl, n = ctypes.c_double(0), ctypes.c_double(0), c_i64(0)
        self.check(self.lib.masg_profile_read(self.h, self.stream(), ctypes.byref(ms), ctypes.byref(fl), ctypes.byref(n)),
                   "masg_profile_read")
        return ms.value, fl.value, int(n.value)

    def enable_taps(self, on=True):
        """Keep copies of the intermediate tensors of the next forwards for ``tap`` (off by default)."""
        self.check(self.lib.masg_debug_enable_taps(self.h, 1 if on else 0), "masg_debug_enable_taps")

    def tap(self, name, shape):
        out = torch.empty(shape, dtype=torch.float32, device=self.device)
        self.check(self.lib.masg_debug_tap(self.h, self.stream(), name.encode(), ptr(out), out.numel()), "masg_debug_tap")
        return out

    def close(self):
        if getattr(self, "h", None):
            self.lib.masg_ctx_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


_default = {}


def default_context(device, precision=None):
    """Shared context for stateless helpers (GMM sampling/interpolation, standalone kernels)."""
    precision = precision or os.environ.get("MASG_PRECISION", "bf16")
    dev = torch.device(device)
    idx = dev.index if dev.index is not None else torch.cuda.current_device()
    key = (idx, _PREC_NAMES[precision])
    if key not in _default:
        _default[key] = Context(torch.device("cuda", idx), precision)
    return _default[key]
