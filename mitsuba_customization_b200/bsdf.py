"""Host-side mirror of the Mitsuba 3 plugin interface for the `merl` and
`customized_measurment` BSDFs.

The reference's C++ plugins (named by /root/reference/README.md:1; their sources are
un-vendored submodules, /root/reference/.SUBMODULES.json:8-23) derive from Mitsuba 3's
`BSDF` and override eval / pdf / eval_pdf / sample with the signatures mirrored below.  Here
torch CUDA tensors stand in for Dr.Jit CUDA arrays: a Vector3f / Color3f is a float32
tensor of shape (3, n) whose rows are the SoA component arrays, a Float is shape (n,), a
Mask is a bool/uint8 tensor of shape (n,).  Every method hands raw device pointers to
libmerl_b200 through ctypes on torch's current stream; nothing here computes.

The C++ host sources a Mitsuba maintainer would compile are under plugins/ (see
INTEGRATION.md); this module is the harness that can run in this repository.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import Optional

import numpy as np
import torch

from . import _cabi
from ._cabi import MerlError, check


# --- the handful of Mitsuba 3 types the four entry points touch -------------------------------
class BSDFFlags:
    Empty = 0x0
    DiffuseReflection = 0x2
    GlossyReflection = 0x8
    FrontSide = 0x10000
    BackSide = 0x20000
    Reflection = 0x2 | 0x8 | 0x20 | 0x80
    All = 0x1 | 0x2 | 0x4 | 0x8 | 0x10 | 0x20 | 0x40 | 0x80 | 0x100


class TransportMode:
    Radiance = 0
    Importance = 1


@dataclass
class BSDFContext:
    mode: int = TransportMode.Radiance
    type_mask: int = BSDFFlags.All
    component: int = -1  # uint32(-1) upstream: all components

    def is_enabled(self, type_: int, component: int = 0) -> bool:
        return (self.component == -1 or self.component == component) and (self.type_mask & type_) != 0


@dataclass
class SurfaceInteraction3f:
    wi: torch.Tensor  # (3, n) incident direction in the local shading frame


@dataclass
class BSDFSample3f:
    wo: torch.Tensor            # (3, n)
    pdf: torch.Tensor           # (n,)
    eta: float = 1.0
    sampled_type: int = BSDFFlags.GlossyReflection
    sampled_component: int = 0
    lobe: Optional[torch.Tensor] = None  # (n,) uint32-as-int32: 0 cosine, 1 tabulated (build-specific)


class Properties(dict):
    """Plugin parameters as the XML/dict loader would deliver them."""

    def string(self, k, default=None):
        return str(self.get(k, default))


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else C.c_void_p(t.data_ptr())


def _vec(t: torch.Tensor, what: str) -> torch.Tensor:
    if t.dtype != torch.float32 or t.dim() != 2 or t.shape[0] != 3 or not t.is_cuda:
        raise MerlError(f"{what} must be a float32 CUDA tensor of shape (3, n)")
    return t if t.is_contiguous() else t.contiguous()


def _mask(active, n, device):
    if active is None or active is True:
        return None
    if active is False:
        return torch.zeros(n, dtype=torch.uint8, device=device)
    if active.dtype == torch.bool:
        return active.view(torch.uint8) if active.is_contiguous() else active.contiguous().view(torch.uint8)
    return active.to(torch.uint8).contiguous()


class _EvalOp(torch.autograd.Function):
    """The table lookup as one autograd node (the torch twin of plugins/mitsuba3/merl_eval_op.h): forward =
    merl_eval, backward = merl_backward into a zeroed (ncell, 4) gradient whose hit-count lane is cleared
    (that lane of the table holds validity bits).  `cells` is the plugin's own table view; the kernel reads
    it through the table handle, the argument only ties the node to the leaf."""

    @staticmethod
    def forward(fctx, cells, plug, ctx, si, wo, active):
        fctx.plug, fctx.si, fctx.wo, fctx.active = plug, si, wo, active
        fctx.enabled = plug._enabled(ctx)
        return plug.eval(ctx, si, wo, active)

    @staticmethod
    def backward(fctx, grad_out):
        plug = fctx.plug
        grad = torch.zeros(plug.info.ncell, 4, dtype=torch.float32, device=grad_out.device)
        if fctx.enabled:
            plug.backward(fctx.si, fctx.wo, grad_out.contiguous(), active=fctx.active, grad_cells=grad)
            check(_cabi.lib().merl_grad_clear_counts(plug._h, C.c_void_p(grad.data_ptr()), _stream(grad.device)))
        return grad, None, None, None, None, None


def _stream(device):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


class _DevArray:
    """Zero-copy view of library-owned device memory via __cuda_array_interface__."""

    def __init__(self, ptr, shape, typestr="<f4"):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False),
                                         "version": 2}


class MERL:
    """`<bsdf type="merl">` -- parameters: filename (MERL .binary), optional spec_weight.

    Construct from a file (`filename`) as the reference plugin does, or from an in-memory
    planar table (`data`, shape (3, ncell) float64 in .binary units) for tests.
    """

    plugin_name = "merl"

    def __init__(self, props=None, **kw):
        props = Properties(props or {})
        props.update(kw)
        self.props = props
        self.device = torch.device(props.get("device", "cuda:0"))
        if self.device.type != "cuda" or not torch.cuda.is_available():
            raise MerlError("the merl plugin runs on a CUDA device only (no CPU fallback)")
        dev_index = self.device.index or 0
        p = _cabi.default_params()
        self._configure(p, props)
        h = C.c_void_p()
        if "data" in props:
            raw = np.ascontiguousarray(props["data"], dtype=np.float64)
            if raw.size != 3 * p.nh * p.nd * p.np:
                raise MerlError(f"data has {raw.size} values, expected 3*{p.nh}*{p.nd}*{p.np}")
            check(_cabi.lib().merl_table_create_from_host(raw.ctypes.data, C.byref(p), dev_index, C.byref(h)))
        elif "packed" in props:
            # sidecar written by save_packed(): device-format table + CDF, parameters included
            check(_cabi.lib().merl_table_create_from_packed(str(props["packed"]).encode(), dev_index, C.byref(h)))
        elif "filename" in props:
            check(_cabi.lib().merl_table_create_from_file(str(props["filename"]).encode(), C.byref(p), dev_index,
                                                          C.byref(h)))
        else:
            raise MerlError(f"{self.plugin_name}: parameter 'filename' is required")
        self._h = h
        self.info = _cabi.TableInfo()
        check(_cabi.lib().merl_table_info(self._h, C.byref(self.info)))
        self.m_flags = BSDFFlags.GlossyReflection | BSDFFlags.FrontSide
        self.m_components = [self.m_flags]

    def _configure(self, p, props):
        """MERL fixes resolution and scales; only sampler knobs are parameters."""
        if "spec_weight" in props:
            p.spec_weight = float(props["spec_weight"])
        if "ni" in props:
            p.ni = int(props["ni"])
        if "l2_persist" in props:
            p.l2_persist = int(bool(props["l2_persist"]))

    # -- Mitsuba Object plumbing ---------------------------------------------------------------
    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                _cabi.lib().merl_table_destroy(h)
            except Exception:
                pass

    def flags(self):
        return self.m_flags

    def component_count(self):
        return len(self.m_components)

    def to_string(self):
        i = self.info
        return (f"{type(self).__name__}[\n  resolution = {i.nh}x{i.nd}x{i.np},\n  table_bytes = {i.table_bytes},\n"
                f"  spec_weight = {i.spec_weight:g},\n  device = {i.device}\n]")

    __repr__ = to_string

    def traverse(self):
        """Differentiable parameters, as `traverse(TraversalCallback*)` would expose them."""
        return {"data": self.cells()}

    def parameters_changed(self, keys=(), data=None):
        """After `data` (cells()) was updated in place, e.g. by an optimiser step: rebuild the sampler.
        `data`: the new values in another (ncell, 4) float32 CUDA tensor (an optimiser that replaces
        the parameter instead of writing into it); copied into the table first."""
        if not keys or "data" in keys:
            if data is not None:
                if data.dtype != torch.float32 or tuple(data.shape) != (self.info.ncell, 4) or not data.is_cuda:
                    raise ValueError("data must be a float32 CUDA tensor of shape (ncell, 4)")
                data = data.contiguous()
                check(_cabi.lib().merl_table_update_cells(self._h, C.c_void_p(data.data_ptr()), _stream(self.device)))
            else:
                check(_cabi.lib().merl_table_refresh_sampler(self._h, _stream(self.device)))

    def cells(self) -> torch.Tensor:
        """The prescaled float4 table on the device, (ncell, 4), zero-copy."""
        return torch.as_tensor(_DevArray(self.info.d_cells, (self.info.ncell, 4)), device=self.device)

    def cdf(self) -> torch.Tensor:
        return torch.as_tensor(_DevArray(self.info.d_cdf, (self.info.ni, self.info.nh + 1)), device=self.device)

    def dens(self) -> torch.Tensor:
        return torch.as_tensor(_DevArray(self.info.d_dens, (self.info.ni, self.info.nh)), device=self.device)

    def save_packed(self, path):
        """Write the device-format sidecar (float4 cells + CDF rows + parameters); load with packed=path."""
        check(_cabi.lib().merl_table_save_packed(self._h, str(path).encode()))

    def set_cdf(self, cdf):
        cdf = np.ascontiguousarray(cdf, dtype=np.float32)
        assert cdf.shape == (self.info.ni, self.info.nh + 1)
        check(_cabi.lib().merl_table_set_cdf(self._h, cdf.ctypes.data))

    def _enabled(self, ctx):
        return ctx.is_enabled(BSDFFlags.GlossyReflection, 0)

    # -- the four BSDF entry points ------------------------------------------------------------
    def bins(self, si, wo, active=True):
        wi, wo = _vec(si.wi, "si.wi"), _vec(wo, "wo")
        n = wi.shape[1]
        idx = torch.empty(n, dtype=torch.int32, device=wi.device)
        m = _mask(active, n, wi.device)
        check(_cabi.lib().merl_bins(self._h, _ptr(wi[0]), _ptr(wi[1]), _ptr(wi[2]), _ptr(wo[0]), _ptr(wo[1]),
                                    _ptr(wo[2]), _ptr(m), n, _ptr(idx), _stream(wi.device)))
        return idx

    def eval(self, ctx, si, wo, active=True):
        wi, wo = _vec(si.wi, "si.wi"), _vec(wo, "wo")
        n = wi.shape[1]
        if not self._enabled(ctx):
            return torch.zeros(3, n, dtype=torch.float32, device=wi.device)
        out = torch.empty(3, n, dtype=torch.float32, device=wi.device)
        m = _mask(active, n, wi.device)
        check(_cabi.lib().merl_eval(self._h, _ptr(wi[0]), _ptr(wi[1]), _ptr(wi[2]), _ptr(wo[0]), _ptr(wo[1]),
                                    _ptr(wo[2]), _ptr(m), n, _ptr(out[0]), _ptr(out[1]), _ptr(out[2]),
                                    _stream(wi.device)))
        return out

    def eval_ad(self, ctx, si, wo, cells, active=True):
        """eval() differentiable with respect to the table: `cells` is `self.cells()` with requires_grad
        set (the parameter `data` of traverse()); `loss.backward()` leaves d loss / d cells in `cells.grad`
        ((ncell, 4), validity lane zero).  After an in-place optimiser step call parameters_changed()."""
        if cells.data_ptr() != self.info.d_cells or tuple(cells.shape) != (self.info.ncell, 4):
            raise MerlError("cells must be this plugin's own table view (self.cells())")
        return _EvalOp.apply(cells, self, ctx, si, wo, active)

    def pdf(self, ctx, si, wo, active=True):
        wi, wo = _vec(si.wi, "si.wi"), _vec(wo, "wo")
        n = wi.shape[1]
        if not self._enabled(ctx):
            return torch.zeros(n, dtype=torch.float32, device=wi.device)
        out = torch.empty(n, dtype=torch.float32, device=wi.device)
        m = _mask(active, n, wi.device)
        check(_cabi.lib().merl_pdf(self._h, _ptr(wi[0]), _ptr(wi[1]), _ptr(wi[2]), _ptr(wo[0]), _ptr(wo[1]),
                                   _ptr(wo[2]), _ptr(m), n, _ptr(out), _stream(wi.device)))
        return out

    def eval_pdf(self, ctx, si, wo, active=True, out=None):
        wi, wo = _vec(si.wi, "si.wi"), _vec(wo, "wo")
        n = wi.shape[1]
        if not self._enabled(ctx):
            return (torch.zeros(3, n, dtype=torch.float32, device=wi.device),
                    torch.zeros(n, dtype=torch.float32, device=wi.device))
        if out is None:
            out = torch.empty(4, n, dtype=torch.float32, device=wi.device)
        m = _mask(active, n, wi.device)
        check(_cabi.lib().merl_eval_pdf(self._h, _ptr(wi[0]), _ptr(wi[1]), _ptr(wi[2]), _ptr(wo[0]), _ptr(wo[1]),
                                        _ptr(wo[2]), _ptr(m), n, _ptr(out[0]), _ptr(out[1]), _ptr(out[2]),
                                        _ptr(out[3]), _stream(wi.device)))
        return out[:3], out[3]

    def sample(self, ctx, si, sample1, sample2, active=True):
        wi = _vec(si.wi, "si.wi")
        n = wi.shape[1]
        dev = wi.device
        if not self._enabled(ctx):
            z = torch.zeros(3, n, dtype=torch.float32, device=dev)
            return BSDFSample3f(wo=z, pdf=torch.zeros(n, dtype=torch.float32, device=dev)), z.clone()
        u1 = sample1.contiguous()
        u2 = sample2.contiguous()
        if u2.shape != (2, n) or u1.shape != (n,):
            raise MerlError("sample1 must have shape (n,), sample2 shape (2, n)")
        out = torch.empty(7, n, dtype=torch.float32, device=dev)
        lobe = torch.empty(n, dtype=torch.int32, device=dev)
        m = _mask(active, n, dev)
        check(_cabi.lib().merl_sample(self._h, _ptr(wi[0]), _ptr(wi[1]), _ptr(wi[2]), _ptr(u1), _ptr(u2[0]),
                                      _ptr(u2[1]), _ptr(m), n, _ptr(out[0]), _ptr(out[1]), _ptr(out[2]),
                                      _ptr(out[3]), _ptr(out[4]), _ptr(out[5]), _ptr(out[6]), _ptr(lobe),
                                      _stream(dev)))
        bs = BSDFSample3f(wo=out[0:3], pdf=out[3], lobe=lobe)
        return bs, out[4:7]

    # -- differentiable backward (row A10) -----------------------------------------------------
    def backward(self, si, wo, grad_rgb, active=True, grad_cells=None, warp_aggregate=False, defer_mask=False,
                 cta_aggregate=False):
        """grad_cells (ncell, 4) += d eval / d cells . grad_rgb; returns grad_cells.
        warp_aggregate / cta_aggregate: for coherent batches (gradients of a rendered image); off by
        default, as in every host of this repository (include/merl_b200.h).  defer_mask: accumulate
        unmasked, the caller applies the validity mask once (apply_grad_mask / PeerBuffer.allreduce_)."""
        wi, wo, g = _vec(si.wi, "si.wi"), _vec(wo, "wo"), _vec(grad_rgb, "grad_rgb")
        n = wi.shape[1]
        if grad_cells is None:
            grad_cells = torch.zeros(self.info.ncell, 4, dtype=torch.float32, device=wi.device)
        m = _mask(active, n, wi.device)
        check(_cabi.lib().merl_backward(self._h, _ptr(wi[0]), _ptr(wi[1]), _ptr(wi[2]), _ptr(wo[0]), _ptr(wo[1]),
                                        _ptr(wo[2]), _ptr(g[0]), _ptr(g[1]), _ptr(g[2]), _ptr(m), n,
                                        _ptr(grad_cells),
                                        (_cabi.BWD_WARP_AGGREGATE if warp_aggregate else 0)
                                        | (_cabi.BWD_DEFER_MASK if defer_mask else 0)
                                        | (_cabi.BWD_CTA_AGGREGATE if cta_aggregate else 0),
                                        _stream(wi.device)))
        return grad_cells

    def apply_grad_mask(self, grad_cells):
        """Zero the channels of invalid cells in a gradient accumulated with defer_mask=True."""
        check(_cabi.lib().merl_grad_apply_mask(self._h, _ptr(grad_cells), _stream(grad_cells.device)))
        return grad_cells

    # -- host-buffer path (scalar variants / end-to-end) ---------------------------------------
    def eval_pdf_host(self, wi_host, wo_host, out_host):
        """wi_host, wo_host: (3, n) float32 CPU tensors (pinned for full PCIe rate);
        out_host: (4, n) float32 CPU tensor receiving r, g, b, pdf."""
        n = wi_host.shape[1]
        a, b, o = wi_host, wo_host, out_host
        check(_cabi.lib().merl_eval_pdf_host(self._h, _ptr(a[0]), _ptr(a[1]), _ptr(a[2]), _ptr(b[0]), _ptr(b[1]),
                                             _ptr(b[2]), n, _ptr(o[0]), _ptr(o[1]), _ptr(o[2]), _ptr(o[3])))
        return out_host


class CustomizedMeasurement(MERL):
    """`<bsdf type="customized_measurment">` (spelling per /root/reference/README.md:1).

    Assumption B7: a MERL-layout table whose resolution comes from the file
    header (or `resolution`), with per-channel `scale` and a `theta_h_mapping` of "sqrt"
    (MERL's nonlinear bins) or "linear".
    """

    plugin_name = "customized_measurment"

    def _configure(self, p, props):
        super()._configure(p, props)
        if "resolution" in props:
            p.nh, p.nd, p.np = (int(x) for x in props["resolution"])
        elif "filename" in props:
            p.nh = p.nd = p.np = 0  # take the header's
        if "scale" in props:
            s = props["scale"]
            s = (s, s, s) if np.isscalar(s) else s
            p.scale = (C.c_double * 3)(*[float(x) for x in s])
        mapping = props.get("theta_h_mapping", "sqrt")
        if mapping not in ("sqrt", "linear"):
            raise MerlError("theta_h_mapping must be 'sqrt' or 'linear'")
        p.theta_h_sqrt = 1 if mapping == "sqrt" else 0
        if "clamp_negative" in props:
            p.clamp_negative = int(bool(props["clamp_negative"]))


PLUGINS = {"merl": MERL, "customized_measurment": CustomizedMeasurement}


def load_dict(d: dict):
    """Tiny stand-in for mi.load_dict({'type': 'merl', 'filename': ...})."""
    d = dict(d)
    kind = d.pop("type")
    if kind not in PLUGINS:
        raise MerlError(f"unknown BSDF plugin '{kind}'")
    return PLUGINS[kind](d)
