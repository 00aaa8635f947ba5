"""Drop-ins for ``model.distributions.InterpolateGMM`` / ``BarycenterGMM``
(model/distributions.py:12-77, 79-192): same constructors, ``interpolate_rate`` / ``barycenter_rate``
and ``.sample()``; both stay ``torch.distributions.MixtureSameFamily`` subclasses and picklable with
``torch.save`` (examples_gen_distri.py:211 loads them back).

The reference builds these on the CPU with scipy ``sqrtm``/``inv`` on 256x256 diagonal matrices and
Python triple loops (seconds per construction).  Here the diagonal closed forms run as CUDA kernels
(masg_gmm_interpolate / masg_gmm_barycenter); only the K×K exact optimal-transport plan (``ot.emd``,
:22) is solved on the host inside the C library.  The reference's quirks are reproduced on purpose
(SURVEY.md Appendix B-8/9): the cost is the as-written ``_w2sq`` variant, mixture weights are
``gamma.flatten()`` (index i·K+j) while components are ordered (j, i), and the component "scale" is
the McCann-interpolated *variance*.
"""
import ctypes
import itertools

import numpy as np
import torch
import torch.distributions as D

from . import _lib
from .fastspeech2 import gmm_sample


def _unwrap(distri):
    comp = distri.component_distribution
    assert type(comp) == D.Normal or type(comp.base_dist) == D.Normal
    normal = comp if type(comp) == D.Normal else comp.base_dist
    return distri.mixture_distribution.probs, normal


class _DeviceSampleMixin:
    def sample(self, sample_shape=torch.Size()):
        """MixtureSameFamily.sample() through the fused Philox categorical+normal kernel; shape
        sample_shape + [1, D] like the reference's batch-1 mixtures."""
        n = int(np.prod(tuple(sample_shape))) if len(tuple(sample_shape)) else 1
        comp = self.component_distribution.base_dist
        dev = comp.loc.device
        ctx = _lib.default_context(dev)
        out = gmm_sample(ctx, self.mixture_distribution.probs.reshape(1, -1), comp.loc.reshape(1, comp.loc.shape[-2], -1),
                         comp.scale.reshape(1, comp.scale.shape[-2], -1), n)
        return out.reshape(tuple(sample_shape) + (1, out.shape[-1]))


class InterpolateGMM(_DeviceSampleMixin, D.MixtureSameFamily):
    def __init__(self, distri_a: D.MixtureSameFamily, distri_b: D.MixtureSameFamily, _plan=None, _t=0.5, _cost=None):
        pa, na = _unwrap(distri_a)
        pb, nb = _unwrap(distri_b)
        dev = na.loc.device
        if dev.type != "cuda":
            raise RuntimeError("InterpolateGMM (B200-native) needs CUDA distributions; no CPU fallback")
        # only batch element [0] of each input GMM is used (distributions.py:22,41)
        f32 = dict(device=dev, dtype=torch.float32)
        self._a = (pa[0].to(**f32).contiguous(), na.loc[0].to(**f32).contiguous(), na.scale[0].to(**f32).contiguous())
        self._b = (pb[0].to(**f32).contiguous(), nb.loc[0].to(**f32).contiguous(), nb.scale[0].to(**f32).contiguous())
        self.t = _t
        if _plan is None:
            w, mu, sc, cost = self._run(recompute_plan=True)
            self.ot_Cost = cost.tolist()
            self.ot_Matrix = w.reshape(self._a[0].shape[0], -1).double().cpu().numpy()
        else:
            # a transport plan that was solved elsewhere (a reference-made pickle carries POT's own ot.emd result): keep it
            _, mu, sc, _ = self._run(recompute_plan=False)
            self.ot_Matrix = np.asarray(_plan, dtype=np.float64)
            self.ot_Cost = _cost
            w = torch.from_numpy(self.ot_Matrix).flatten().to(**f32)
        self.mix = D.Categorical(w.unsqueeze(0))
        comp = D.Independent(D.Normal(mu.unsqueeze(0), sc.unsqueeze(0)), 1)
        super(InterpolateGMM, self).__init__(self.mix, comp)

    def _run(self, recompute_plan):
        (pa, ma, sa), (pb, mb, sb) = self._a, self._b
        K, Dm = ma.shape
        dev = ma.device
        ctx = _lib.default_context(dev)
        w = torch.empty(K * K, device=dev)
        mu = torch.empty(K * K, Dm, device=dev)
        sc = torch.empty(K * K, Dm, device=dev)
        cost = (ctypes.c_double * (K * K))()
        ctx.check(ctx.lib.masg_gmm_interpolate(ctx.h, ctx.stream(), _lib.ptr(pa), _lib.ptr(ma), _lib.ptr(sa), _lib.ptr(pb),
                                               _lib.ptr(mb), _lib.ptr(sb), K, Dm, float(self.t), 1 if recompute_plan else 0,
                                               _lib.ptr(w), _lib.ptr(mu), _lib.ptr(sc), ctypes.cast(cost, ctypes.c_void_p)),
                  "masg_gmm_interpolate")
        return w, mu, sc, np.array(cost).reshape(K, K)

    def interpolate_rate(self, t):
        """model/distributions.py:31-38: re-derive the components at rate t (the OT plan is kept)."""
        self.t = t
        _, mu, sc, _ = self._run(recompute_plan=False)
        self._component_distribution = D.Independent(D.Normal(mu.unsqueeze(0), sc.unsqueeze(0)), 1)


class BarycenterGMM(_DeviceSampleMixin, D.MixtureSameFamily):
    """Multi-attribute barycentre GMM (model/distributions.py:79-192).  The reference constructor raises
    TypeError as written (:92 vs :136); this class implements the maths it states:
    K^A barycentric Gaussians (mean Σ r_i μ_i, std by the 60-step fixed-point loop), each original
    component sending r_i·π_ij to its W2-nearest barycentre, unused barycentres dropped, components in
    first-use order."""

    def __init__(self, model, device="cpu"):
        self.metadata_list = model.speaker_enc.metadata_list
        dev = next(model.parameters()).device
        if dev.type != "cuda":
            raise RuntimeError("BarycenterGMM (B200-native) needs the model on CUDA; no CPU fallback")
        self.device = dev
        pools = [np.eye(len(mt))[list(mt.values())] for mt in self.metadata_list.values()]
        metas = [np.concatenate(c, 0) for c in itertools.product(*pools)]
        meta = torch.tensor(np.stack(metas), dtype=torch.float32, device=dev)
        self._meta = meta
        pi, mu, sigma = model.speaker_enc.params(meta)                      # [A,K], [A,K,D], [A,K,D]
        self._params = (pi.contiguous(), mu.contiguous(), sigma.contiguous())
        self.rate = [1 / meta.shape[0]] * meta.shape[0]
        mix, comp = self._build()
        super(BarycenterGMM, self).__init__(mix, comp)

    def _build(self):
        pi, mu, sigma = self._params
        A, K, Dm = mu.shape
        nb = K ** A
        dev = mu.device
        ctx = _lib.default_context(dev)
        bmean = torch.empty(nb, Dm, device=dev); bstd = torch.empty(nb, Dm, device=dev)
        bw = torch.empty(nb, device=dev); assign = torch.empty(A * K, dtype=torch.int32, device=dev)
        rate = (ctypes.c_float * A)(*[float(r) for r in self.rate])
        ctx.check(ctx.lib.masg_gmm_barycenter(ctx.h, ctx.stream(), _lib.ptr(pi), _lib.ptr(mu), _lib.ptr(sigma),
                                              ctypes.cast(rate, ctypes.c_void_p), A, K, Dm, _lib.ptr(bmean), _lib.ptr(bstd),
                                              _lib.ptr(bw), _lib.ptr(assign)), "masg_gmm_barycenter")
        order = []
        for a in assign.cpu().tolist():                                      # dict insertion order (:94,98)
            if a not in order:
                order.append(a)
        self.used = order
        idx = torch.tensor(order, device=dev, dtype=torch.long)
        mix = D.Categorical(bw[idx].unsqueeze(0))
        comp = D.Independent(D.Normal(bmean[idx].unsqueeze(0), bstd[idx].unsqueeze(0)), 1)
        return mix, comp

    def barycenter_rate(self, rate, _print=True):
        assert hasattr(rate, "__len__") and len(rate) == self._meta.shape[0]
        assert abs(sum(rate) - 1) < 1e-6
        self.rate = list(rate)
        if _print:
            print("rate: ", rate)
        mix, comp = self._build()
        self._component_distribution = comp
        self._mixture_distribution = mix


# ------------------------------------------------------------------------------------------------
# reference-pickled distribution artefacts: result/<...>/distributions/distri_*.pth are
# ``torch.save(InterpolateGMM | BarycenterGMM)`` of the REFERENCE classes (class path model.distributions.*), loaded back by
# examples_gen_distri.py:211 with torch.load.  Without the reference package on sys.path that unpickling fails; the loader
# below maps the class path onto plain holders, reads the parts and rebinds them to the drop-in classes above.
# ------------------------------------------------------------------------------------------------
class _RefHolder:
    """Receives the pickled __dict__ of a reference InterpolateGMM / BarycenterGMM (no constructor runs)."""


class _RefInterpolateGMM(_RefHolder):
    pass


class _RefBarycenterGMM(_RefHolder):
    pass


_REF_CLASSES = {"InterpolateGMM": _RefInterpolateGMM, "BarycenterGMM": _RefBarycenterGMM}


def _ref_pickle_module():
    import pickle
    import types

    class _Unpickler(pickle.Unpickler):
        def find_class(self, module, name):
            if module == "model.distributions" and name in _REF_CLASSES:      # the reference's class path
                return _REF_CLASSES[name]
            return super().find_class(module, name)

    m = types.ModuleType("pickle")
    m.Unpickler = _Unpickler
    m.load = lambda f, **kw: _Unpickler(f, **kw).load()
    m.loads = lambda b, **kw: _Unpickler(__import__("io").BytesIO(b), **kw).load()
    return m


def read_reference_distribution(path):
    """Parts of a reference-made ``distri_*.pth`` as CPU tensors (no GPU needed):
    {"kind": "interpolate", "a": (pi [K], mu [K,D], sigma [K,D]), "b": (...), "t", "plan" [K,K] float64, "cost"} or
    {"kind": "barycenter", "pi" [A,K], "mu" [A,K,D], "sigma" [A,K,D], "rate", "metadata_list"}; an object saved by the drop-in
    classes of this module is returned as {"kind": "dropin", "object": obj}."""
    obj = torch.load(path, map_location="cpu", weights_only=False, pickle_module=_ref_pickle_module())
    cpu = lambda t: t.detach().to("cpu", torch.float32).contiguous()
    if isinstance(obj, _RefInterpolateGMM):
        (pa, na), (pb, nb) = obj.distri_a, obj.distri_b
        return {"kind": "interpolate", "a": (cpu(pa[0]), cpu(na.loc[0]), cpu(na.scale[0])), "b": (cpu(pb[0]), cpu(nb.loc[0]), cpu(nb.scale[0])),
                "t": float(obj.t), "plan": np.asarray(obj.ot_Matrix, dtype=np.float64), "cost": obj.ot_Cost}
    if isinstance(obj, _RefBarycenterGMM):
        ds = list(obj.original_distri.values())
        return {"kind": "barycenter", "pi": torch.stack([cpu(d["probs"].reshape(-1)) for d in ds]),
                "mu": torch.stack([cpu(d["distributes"].loc[0]) for d in ds]), "sigma": torch.stack([cpu(d["distributes"].scale[0]) for d in ds]),
                "rate": [float(r) for r in obj.rate], "metadata_list": obj.metadata_list}
    if isinstance(obj, (InterpolateGMM, BarycenterGMM)):
        return {"kind": "dropin", "object": obj}
    raise TypeError("%s holds a %s, not a pickled InterpolateGMM / BarycenterGMM" % (path, type(obj).__name__))


def _mixture(pi, mu, sigma, dev):
    f = dict(device=dev, dtype=torch.float32)
    return D.MixtureSameFamily(D.Categorical(pi.to(**f).unsqueeze(0)), D.Independent(D.Normal(mu.to(**f).unsqueeze(0), sigma.to(**f).unsqueeze(0)), 1))


def load_reference_distribution(path, device="cuda"):
    """examples_gen_distri.py:211 ``dist = torch.load(path)`` for a file written by the REFERENCE: returns the drop-in
    InterpolateGMM / BarycenterGMM on ``device`` with the same parameters, interpolation rate and — for InterpolateGMM — the very
    transport plan the reference's ``ot.emd`` solved (it is read from the pickle, not recomputed), so ``.sample()``,
    ``interpolate_rate`` and ``barycenter_rate`` continue from where the reference left off."""
    parts = read_reference_distribution(path)
    dev = torch.device(device)
    if parts["kind"] == "dropin":
        return parts["object"]
    if parts["kind"] == "interpolate":
        return InterpolateGMM(_mixture(*parts["a"], dev), _mixture(*parts["b"], dev), _plan=parts["plan"], _t=parts["t"], _cost=parts["cost"])
    obj = BarycenterGMM.__new__(BarycenterGMM)
    obj.metadata_list = parts["metadata_list"]
    obj.device = dev
    obj._meta = torch.zeros(parts["pi"].shape[0], 1, device=dev)                # only its row count is used (barycenter_rate)
    f = dict(device=dev, dtype=torch.float32)
    obj._params = (parts["pi"].to(**f).contiguous(), parts["mu"].to(**f).contiguous(), parts["sigma"].to(**f).contiguous())
    obj.rate = list(parts["rate"])
    mix, comp = obj._build()
    D.MixtureSameFamily.__init__(obj, mix, comp)
    return obj
