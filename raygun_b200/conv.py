"""Convolution dispatch: tcgen05 implicit GEMM (bf16 NDHWC) when eligible, direct kernels otherwise.

Weights are always the checkpoint-layout fp32 tensors (``[Cout, Cin/g, kD, kH, kW]``); the bf16
K-major packs the tensor-core kernels read are derived caches, re-packed when the parameter's
version counter changes (i.e. after every optimizer step), never serialised.
"""
from __future__ import annotations

import ctypes as C
import weakref

import torch

from . import _C
from . import kernels as K
from ._C import ACT_NONE, ConvGeom, IgemmDesc, Upconv2Desc, check, count_launch, i3, lib, ptr, require_device, stream_ptr, t5

stats = {"igemm": 0, "direct": 0}     # which path convolutions took (tests / bench inspect this)

# Optional per-launch device timing of the tensor-core kernels (bench.py's roofline): when `profile`
# is a list, every igemm launch appends (kind, flops, start_event, end_event) recorded on the
# launching stream.
profile = None


class _Timed:
    def __init__(self, kind, flops, device, kernel="igemm_conv_kernel"):
        self.on = profile is not None
        if self.on:
            self.kind, self.flops, self.kernel = kind, flops, kernel
            self.e0 = torch.cuda.Event(enable_timing=True)
            self.e1 = torch.cuda.Event(enable_timing=True)
            self.stream = torch.cuda.current_stream(device)

    def __enter__(self):
        if self.on:
            self.e0.record(self.stream)
        return self

    def __exit__(self, *exc):
        if self.on:
            self.e1.record(self.stream)
            profile.append((self.kind, self.flops, self.e0, self.e1, self.kernel))
        return False


def _geom(kernel, stride, pad, groups=1, act=ACT_NONE, slope=0.0):
    g = ConvGeom()
    g.kernel[:] = [int(k) for k in kernel]
    g.stride[:] = [int(s) for s in stride]
    g.pad[:] = [int(p) for p in pad]
    g.groups, g.act, g.slope = int(groups), int(act), float(slope)
    return g


def out_size(i, k, s, p):
    return (i + 2 * p - k) // s + 1


# --------------------------------------------------------------------------------------------
# derived weight tensors (bf16 packs) and plan caches
# --------------------------------------------------------------------------------------------
# Every tensor the tensor-core kernels read instead of the fp32 checkpoint weight -- K-major fprop / dgrad packs, slice-ring
# packs, the stem's [Cout][64] rows, the up-convolution's [8][Cout][Cin] classes, zero-padded variants for narrow layers --
# is a DERIVED tensor registered here per source parameter.  The buffers are persistent: a refresh re-packs INTO them, so
#   * a captured CUDA graph keeps reading valid addresses, and the re-pack kernels recorded after the optimiser step
#     (optim.FlatAdam.step -> repack) refresh them on every replay -- an eager forward between replays never sees stale
#     packs;
#   * torch-side in-place updates (load_state_dict, torch.optim, init_weights) are caught through the parameter's
#     version counter / data pointer, and a replayed graph that updated parameters behind Python's back through
#     `invalidate_packs()` (a global epoch that is part of every entry's validity).
_pack_epoch = 0


def invalidate_packs():
    """Mark every derived weight tensor stale (they are re-packed in place on next use)."""
    global _pack_epoch
    _pack_epoch += 1


class _Derived:
    __slots__ = ("ref", "items")

    def __init__(self, w):
        self.ref = weakref.ref(w)
        self.items = {}          # tag -> [tensor(s), build, version, data_ptr, epoch]


_derived: "dict[int, _Derived]" = {}


def derived(w: torch.Tensor, tag, build):
    """The tensor(s) `build(w, out)` derives from parameter `w` (out = None: allocate; else refresh in place), cached until
    `w` changes."""
    key = id(w)
    ent = _derived.get(key)
    if ent is None or ent.ref() is not w:
        for k in [k for k, e in _derived.items() if e.ref() is None]:      # purge dead entries on every insert
            del _derived[k]
        ent = _derived[key] = _Derived(w)
    it = ent.items.get(tag)
    ver, dptr = w._version, w.data_ptr()
    if it is not None and it[2] == ver and it[3] == dptr and it[4] == _pack_epoch:
        return it[0]
    out = build(w, it[0] if it is not None else None)
    ent.items[tag] = [out, build, ver, dptr, _pack_epoch]
    return out


def repack(params):
    """Refresh, in place, every derived tensor of `params` (called by the optimiser right after it updated the fp32
    weights through the C ABI, which does not touch Python's version counters)."""
    n = 0
    for w in params:
        ent = _derived.get(id(w))
        if ent is None or ent.ref() is not w:
            continue
        for it in ent.items.values():
            it[0] = it[1](w, it[0])
            it[2], it[3], it[4] = w._version, w.data_ptr(), _pack_epoch
            n += 1
    return n


def _contig(w):
    w = w.detach()
    return w if w.is_contiguous() else w.contiguous()


def _pad_rows(w, rows):
    """w [Cout, ...] zero-padded to `rows` output channels (narrow layers run on the tensor cores with padded channels)."""
    wp = torch.zeros(rows, *w.shape[1:], dtype=torch.float32, device=w.device)
    wp[: w.shape[0]] = w.detach()
    return wp


def _pad_cin(w, cin):
    wp = torch.zeros(w.shape[0], cin, *w.shape[2:], dtype=torch.float32, device=w.device)
    wp[:, : w.shape[1]] = w.detach()
    return wp


class WeightSrc:
    """A convolution weight as the kernels see it: `param` (the fp32 checkpoint tensor the packs are cached under), an
    optional transform `fn` (zero-padding of channels, the stem's H/W-tap fold) with the resulting `shape`, and `tag`,
    which names the transform in the cache.  Plain parameters are wrapped as WeightSrc(param)."""
    __slots__ = ("param", "fn", "shape", "tag")

    def __init__(self, param, fn=None, shape=None, tag=""):
        self.param, self.fn, self.tag = param, fn, tag
        self.shape = tuple(shape) if shape is not None else tuple(param.shape)

    def fp32(self):
        return _fp32_of(self.param, self.fn)

    @property
    def device(self):
        return self.param.device


def _fp32_of(param, fn):
    return _contig(param) if fn is None else fn(param).contiguous()


def _src(w):
    return w if isinstance(w, WeightSrc) else WeightSrc(w)


def packed_weights(w):
    """(w_fprop [Cout][tap][Cin], w_dgrad [Cin][tap][Cout]) bf16 packs of a conv-layout weight."""
    ws = _src(w)
    cout, cin = ws.shape[0], ws.shape[1]
    taps = ws.shape[2] * ws.shape[3] * ws.shape[4]
    fn = ws.fn           # the builders must not hold the parameter itself (the registry keeps them; entries die with it)

    def build(p_, out):
        wf, wd = out if out is not None else (torch.empty(cout, taps, cin, dtype=torch.bfloat16, device=p_.device),
                                              torch.empty(cin, taps, cout, dtype=torch.bfloat16, device=p_.device))
        check(lib().rg_pack_weights(ptr(_fp32_of(p_, fn)), C.c_int(cout), C.c_int(cin), C.c_int(taps), ptr(wf), ptr(wd),
                                    stream_ptr(p_.device)), "rg_pack_weights")
        count_launch()
        return wf, wd

    return derived(ws.param, ("igemm", ws.tag), build)


class _Plan:
    def __init__(self, handle):
        self.handle = handle

    def __del__(self):
        try:
            if self.handle:
                lib().rg_plan_destroy(self.handle)
        except Exception:
            pass


_plans: "dict[tuple, _Plan | None]" = {}


def _get_plan(mode, n, cin, cout, in_dhw, out_dhw, kernel, stride, pad, xs, ys, act, slope):
    key = (mode, n, cin, cout, tuple(in_dhw), tuple(out_dhw), tuple(kernel), tuple(stride), tuple(pad), tuple(xs),
           tuple(ys), act, slope)
    if key in _plans:
        return _plans[key]
    d = IgemmDesc()
    d.mode, d.n, d.cin, d.cout = mode, n, cin, cout
    d.in_dhw[:] = list(in_dhw)
    d.out_dhw[:] = list(out_dhw)
    d.kernel[:] = list(kernel)
    d.stride[:] = list(stride)
    d.pad[:] = list(pad)
    d.x_strides[:] = list(xs)
    d.y_strides[:] = list(ys)
    d.act, d.slope = act, slope
    handle = C.c_void_p()
    rc = lib().rg_plan_create(C.byref(d), C.byref(handle))
    if rc == _C.E_UNSUPPORTED:
        _plans[key] = None
        return None
    check(rc, "rg_plan_create")
    _plans[key] = _Plan(handle)
    return _plans[key]


# --------------------------------------------------------------------------------------------
# slice-ring kernel (stride-1 layers producing 64-channel blocks): descriptor, support cache, weight packs
# --------------------------------------------------------------------------------------------
import os

_ring_descs: "dict[tuple, IgemmDesc | None]" = {}
RING_MAX_PRODUCED = int(os.environ.get("RAYGUN_B200_RING_MAX_PRODUCED", "192"))   # 0 disables the ring kernel


def _ring_desc(mode, n, cin, cout, in_dhw, out_dhw, kernel, stride, pad, xs, ys, act, slope):
    """IgemmDesc when the slice-ring kernel should run this layer, else None."""
    produced = cout if mode == 0 else cin
    if tuple(stride) != (1, 1, 1) or produced > RING_MAX_PRODUCED or produced % 64 or (cin if mode == 0 else cout) % 64:
        return None
    key = (mode, n, cin, cout, tuple(in_dhw), tuple(out_dhw), tuple(kernel), tuple(pad), tuple(xs), tuple(ys), act, slope)
    if key in _ring_descs:
        return _ring_descs[key]
    d = IgemmDesc()
    d.mode, d.n, d.cin, d.cout = mode, n, cin, cout
    d.in_dhw[:] = list(in_dhw)
    d.out_dhw[:] = list(out_dhw)
    d.kernel[:] = list(kernel)
    d.stride[:] = [1, 1, 1]
    d.pad[:] = list(pad)
    d.x_strides[:] = list(xs)
    d.y_strides[:] = list(ys)
    d.act, d.slope = act, slope
    ok = lib().rg_ring_supported(C.byref(d)) == 1
    _ring_descs[key] = d if ok else None
    return _ring_descs[key]


STEM_DIRECT = os.environ.get("RAYGUN_B200_STEM_DIRECT", "1") != "0"
FUSE_HEAD = os.environ.get("RAYGUN_B200_FUSE_HEAD", "1") != "0"


class HeadRequest:
    """The generator's head -- Conv3d(ngf -> output_nc, 1x1x1) [+ activation] (BaseSystem.py:209-224) -- offered to the
    convolution that produces its input (`weight` identifies that layer).  When that layer runs on the slice-ring kernel
    under no_grad, its epilogue reduces every voxel's 64 activated channels against the head weights and only the head's
    output leaves the SM (rg_ring_run_head): the 64-channel tensor is neither written nor read back.  `done` tells the
    caller whether the head was consumed; if not it runs the head layers itself."""

    def __init__(self, weight, head_w, head_b, act=ACT_NONE, slope=0.0):
        self.weight, self.head_w, self.head_b, self.act, self.slope = weight, head_w, head_b, act, float(slope)
        self.done = False


_head_request: "HeadRequest | None" = None


class offer_head:
    """Context manager: `with offer_head(HeadRequest(...)) as h: y = net(x)`; afterwards h.done."""

    def __init__(self, req):
        self.req = req

    def __enter__(self):
        global _head_request
        self.prev, _head_request = _head_request, self.req
        return self.req

    def __exit__(self, *exc):
        global _head_request
        _head_request = self.prev
        return False


def _take_head(w, cout, want_stats, out):
    h = _head_request
    if (h is None or h.done or not FUSE_HEAD or torch.is_grad_enabled() or want_stats or out is not None or cout != 64
            or not isinstance(w, torch.Tensor) or w.data_ptr() != h.weight.data_ptr() or w.shape != h.weight.shape
            or not 1 <= h.head_w.shape[0] <= 4 or h.head_w.shape[1] != 64):
        return None
    return h


def stem_direct_ok(w, want_stats):
    """The small-K stem kernel (stem_conv_sm100.cu) takes a stride-1 conv whose whole reduction fits one 64-wide K block."""
    cout, cin, kd, kh, kw = w.shape
    return (STEM_DIRECT and not want_stats and cin * kd * kh * kw <= 64 and cout % 32 == 0 and 32 <= cout <= 256
            and K.compute_dtype() == torch.bfloat16)


def stem_conv_fprop(x, w, bias, pad, act=ACT_NONE, slope=0.0, out=None, pad_mode=0):
    """Forward of the first U-Net convolution (UNet.py:52-87, input_nc -> ngf) without the folded intermediate.
    `pad_mode` (functional.PAD_*): zeros, or reflect / replicate read through mirrored indices (inference only)."""
    require_device(x)
    n, cin, d, h, wd_ = x.shape
    cout = w.shape[0]
    kernel = tuple(w.shape[2:])
    od = [out_size(i, k, 1, p) for i, k, p in zip((d, h, wd_), kernel, pad)]
    if min(od) < 1:
        raise RuntimeError(f"conv3d: kernel {kernel} does not fit input {tuple(x.shape)}")
    xb = x if x.dtype == torch.bfloat16 else K.convert(x, torch.bfloat16)
    if out is not None and tuple(out.shape) == (n, cout, *od) and out.dtype == torch.bfloat16 and K.channels_inner_bf16(out):
        y = out
    elif K.arena_enabled():
        y = K.empty5_arena((n, cout, *od), torch.bfloat16, x.device)
    else:
        y = K.empty5((n, cout, *od), torch.bfloat16, x.device)
    def build(p_, out):
        w64 = out if out is not None else torch.zeros(cout, 64, dtype=torch.bfloat16, device=p_.device)
        w64[:, : cin * kernel[0] * kernel[1] * kernel[2]] = p_.detach().reshape(cout, -1)
        return w64

    w64 = derived(w, "stem64", build)
    d_ = IgemmDesc()
    d_.mode, d_.n, d_.cin, d_.cout = 0, n, cin, cout
    d_.in_dhw[:] = [d, h, wd_]
    d_.out_dhw[:] = od
    d_.kernel[:] = list(kernel)
    d_.stride[:] = [1, 1, 1]
    d_.pad[:] = list(pad)
    d_.x_strides[:] = list(_sp_strides(xb))
    d_.y_strides[:] = list(_sp_strides(y))
    d_.act, d_.slope = act, float(slope)
    d_.has_bias = 1 if bias is not None else 0
    bias_f = bias.detach().float().contiguous() if bias is not None else None
    with _Timed("fprop", 2.0 * y.numel() * cin * kernel[0] * kernel[1] * kernel[2], x.device, "stem_conv_kernel"):
        check(lib().rg_stem_conv_run(C.byref(d_), ptr(xb), C.c_longlong(xb.stride(1)), ptr(w64), ptr(bias_f), ptr(y),
                                     C.c_int(int(pad_mode)), stream_ptr(x.device)), "rg_stem_conv_run")
    count_launch()
    stats["stem_direct"] = stats.get("stem_direct", 0) + 1
    return y


UPCONV2 = os.environ.get("RAYGUN_B200_UPCONV2", "1") != "0"


def upconv2_ok(x, w, stride, pad, out_pad, want_stats):
    """ConvTranspose3d with kernel = stride = 2, no padding (U-Net `Upsample.up`, UNet.py:265-269): one fused launch."""
    cin, cout = w.shape[0], w.shape[1]
    return (UPCONV2 and not want_stats and tuple(w.shape[2:]) == (2, 2, 2) and tuple(stride) == (2, 2, 2)
            and tuple(pad) == (0, 0, 0) and tuple(out_pad) == (0, 0, 0) and cin % 64 == 0 and 64 <= cin <= 512
            and cout % 32 == 0 and 32 <= cout <= 512 and K.compute_dtype() == torch.bfloat16 and K.use_tensor_cores())


def upconv2_fprop(x, w, bias, act=ACT_NONE, slope=0.0, out=None):
    """y = ConvTranspose3d(x; w [Cin][Cout][2][2][2], stride 2): every input voxel writes its 2x2x2 output voxels."""
    require_device(x)
    n, cin, d, h, wd_ = x.shape
    cout = w.shape[1]
    xb = K.conv_operand(x)
    shape = (n, cout, 2 * d, 2 * h, 2 * wd_)
    if out is not None and tuple(out.shape) == shape and K.channels_inner_bf16(out):
        y = out
    elif K.arena_enabled():
        y = K.empty5_arena(shape, torch.bfloat16, x.device)
    else:
        y = K.empty5(shape, torch.bfloat16, x.device)
    def build(p_, out):
        wp = out if out is not None else torch.empty(8 * cout, cin, dtype=torch.bfloat16, device=p_.device)
        wp.view(2, 2, 2, cout, cin).copy_(p_.detach().permute(2, 3, 4, 1, 0))
        return wp

    wpk = derived(w, "upconv2", build)
    d_ = Upconv2Desc()
    d_.n, d_.cin, d_.cout = n, cin, cout
    d_.in_dhw[:] = [d, h, wd_]
    d_.x_strides[:] = list(_sp_strides(xb))
    d_.y_strides[:] = list(_sp_strides(y))
    d_.act, d_.slope = act, float(slope)
    d_.has_bias = 1 if bias is not None else 0
    bias_f = bias.detach().float().contiguous() if bias is not None else None
    with _Timed("fprop", 2.0 * xb.numel() * cout * 8, x.device, "upconv2_kernel"):
        check(lib().rg_upconv2_run(C.byref(d_), ptr(xb), ptr(wpk), ptr(bias_f), ptr(y), stream_ptr(x.device)),
              "rg_upconv2_run")
    count_launch()
    stats["upconv2"] = stats.get("upconv2", 0) + 1
    return y


_halo_descs: "dict[tuple, IgemmDesc | None]" = {}
HALO_WGRAD = os.environ.get("RAYGUN_B200_HALO_WGRAD", "1") != "0"


def _halo_wgrad_desc(n, cin, cout, in_dhw, out_dhw, kernel, stride, pad, xs, ys):
    """IgemmDesc (mode 2) when the halo-resident weight-gradient kernel should run this layer, else None."""
    if not HALO_WGRAD or tuple(stride) != (1, 1, 1) or cin % 64 or cout % 64:
        return None
    key = (n, cin, cout, tuple(in_dhw), tuple(out_dhw), tuple(kernel), tuple(pad), tuple(xs), tuple(ys))
    if key in _halo_descs:
        return _halo_descs[key]
    d = IgemmDesc()
    d.mode, d.n, d.cin, d.cout = 2, n, cin, cout
    d.in_dhw[:] = list(in_dhw)
    d.out_dhw[:] = list(out_dhw)
    d.kernel[:] = list(kernel)
    d.stride[:] = [1, 1, 1]
    d.pad[:] = list(pad)
    d.x_strides[:] = list(xs)
    d.y_strides[:] = list(ys)
    ok = lib().rg_halo_wgrad_supported(C.byref(d)) == 1
    _halo_descs[key] = d if ok else None
    return _halo_descs[key]


def ring_packed_weights(w, mode: int):
    """bf16 slice-ring pack of a conv-layout weight for fprop (mode 0) or dgrad (mode 1); persistent, refreshed in place."""
    ws = _src(w)
    shape, fn = ws.shape, ws.fn

    def build(p_, out):
        out = out if out is not None else torch.empty(shape[0] * shape[1] * shape[2] * shape[3] * shape[4],
                                                      dtype=torch.bfloat16, device=p_.device)
        check(lib().rg_ring_pack_weights(ptr(_fp32_of(p_, fn)), C.c_int(shape[0]), C.c_int(shape[1]), i3(shape[2:]), C.c_int(mode),
                                         ptr(out), stream_ptr(p_.device)), "rg_ring_pack_weights")
        count_launch()
        return out

    return derived(ws.param, ("ring", mode, ws.tag), build)


def _stats_buffers(n_groups, nbytes, device):
    """(stats [2][n*c] outputs, partial-row workspace) of a tensor-core convolution that also produces norm statistics."""
    return torch.empty(2, n_groups, dtype=torch.float32, device=device), K.ws_bytes(nbytes, device)


_ws_bytes_cache: "dict[tuple, int]" = {}


def _sp_strides(t):
    s = t.stride()
    return (s[0], s[2], s[3], s[4])


def _pw_ok(cin, cout, kernel, stride, pad, groups):
    cg = cin // 8
    return (K.get_precision() == "bf16" and K.use_tensor_cores() and groups == 1 and tuple(kernel) == (1, 1, 1)
            and tuple(stride) == (1, 1, 1) and tuple(pad) == (0, 0, 0) and 1 <= cout <= 4 and cin % 8 == 0
            and 1 <= cg <= 32 and (cg & (cg - 1)) == 0)


def _igemm_ok(cin, cout, kernel, stride, groups):
    return (K.get_precision() == "bf16" and K.use_tensor_cores() and groups == 1 and cin % 64 == 0
            and cout % 64 == 0 and all(s in (1, 2) for s in stride) and all(1 <= k <= 8 for k in kernel))


# --------------------------------------------------------------------------------------------
# the three convolution kernels, as raw tensor functions
# --------------------------------------------------------------------------------------------
def conv_fprop(x, w, bias, stride, pad, groups=1, act=ACT_NONE, slope=0.0, want_stats=False, out_dtype=None, out=None):
    """y = act(conv3d(x, w) + bias) with zero padding `pad`.  Returns (y, sum, sqsum) -- the two
    per-(n,cout) statistics are None unless want_stats and the tensor-core epilogue produced them."""
    require_device(x)
    n, cin_t, d, h, wd_ = x.shape
    cout = w.shape[0]
    kernel = tuple(w.shape[2:])
    od = [out_size(i, k, s, p) for i, k, s, p in zip((d, h, wd_), kernel, stride, pad)]
    if min(od) < 1:
        raise RuntimeError(f"conv3d: kernel {kernel} does not fit input {tuple(x.shape)}")
    cdt = out_dtype or K.compute_dtype()
    bias_f = bias.detach() if bias is not None else None
    if out is not None and (tuple(out.shape) != (n, cout, *od) or out.dtype != cdt or out.stride(1) != 1):
        out = None              # caller falls back to a copy when the result is not `out`

    def _alloc(dtype):
        if out is not None and out.dtype == dtype:
            return out
        if dtype == torch.bfloat16 and tuple(stride) == (1, 1, 1) and K.arena_enabled():
            return K.empty5_arena((n, cout, *od), dtype, x.device)
        return K.empty5((n, cout, *od), dtype, x.device)

    if _igemm_ok(cin_t, cout, kernel, stride, groups) and cdt == torch.bfloat16:
        xb = K.conv_operand(x)
        head = _take_head(w, cout, want_stats, out)
        if head is not None:
            # strides of the (never materialised) 64-channel output only select the kernel; use the dense layout
            rd = _ring_desc(0, n, cin_t, cout, (d, h, wd_), od, kernel, stride, pad, _sp_strides(xb),
                            (od[0] * od[1] * od[2] * cout, od[1] * od[2] * cout, od[2] * cout, cout), act, float(slope))
            if rd is not None:
                hn = head.head_w.shape[0]
                yh = K.empty5((n, hn, *od), torch.bfloat16, x.device)
                hw = _src(head.head_w).fp32().reshape(hn, 64)
                hb = head.head_b.detach() if head.head_b is not None else None
                hs = (C.c_int64 * 4)(*_sp_strides(yh))
                with _Timed("fprop", 2.0 * n * od[0] * od[1] * od[2] * cout * (cin_t * kernel[0] * kernel[1] * kernel[2] + hn),
                            x.device, "ring_conv_kernel"):
                    check(lib().rg_ring_run_head(C.byref(rd), ptr(xb), ptr(ring_packed_weights(w, 0)), ptr(bias_f), ptr(hw),
                                                 ptr(hb), C.c_int(hn), C.c_int(head.act), C.c_float(head.slope), ptr(yh), hs,
                                                 stream_ptr(x.device)), "rg_ring_run_head")
                count_launch()
                stats["igemm"] += 1
                stats["ring"] = stats.get("ring", 0) + 1
                stats["ring_head"] = stats.get("ring_head", 0) + 1
                head.done = True
                return yh, None, None
        y = _alloc(torch.bfloat16)
        rd = _ring_desc(0, n, cin_t, cout, (d, h, wd_), od, kernel, stride, pad, _sp_strides(xb), _sp_strides(y),
                        act, float(slope))
        if rd is not None:
            wr = ring_packed_weights(w, 0)
            st, sws = _stats_buffers(n * cout, lib().rg_ring_workspace_bytes(C.byref(rd)), x.device) if want_stats else (None, None)
            with _Timed("fprop", 2.0 * y.numel() * cin_t * kernel[0] * kernel[1] * kernel[2], x.device, "ring_conv_kernel"):
                check(lib().rg_ring_run(C.byref(rd), ptr(xb), ptr(wr), ptr(bias_f), ptr(y),
                                        ptr(st[0]) if want_stats else None, ptr(st[1]) if want_stats else None, ptr(sws),
                                        stream_ptr(x.device)), "rg_ring_run(fprop)")
            count_launch(3 if want_stats else 1)
            stats["igemm"] += 1
            stats["ring"] = stats.get("ring", 0) + 1
            return (y, st[0], st[1]) if want_stats else (y, None, None)
        plan = _get_plan(0, n, cin_t, cout, (d, h, wd_), od, kernel, stride, pad, _sp_strides(xb), _sp_strides(y),
                         act, float(slope))
        if plan is not None:
            wf, _ = packed_weights(w)
            st, sws = _stats_buffers(n * cout, lib().rg_plan_workspace_bytes(plan.handle), x.device) if want_stats else (None, None)
            with _Timed("fprop", 2.0 * y.numel() * cin_t * kernel[0] * kernel[1] * kernel[2], x.device):
                check(lib().rg_igemm_run(plan.handle, ptr(xb), ptr(wf), ptr(bias_f), ptr(y),
                                         ptr(st[0]) if want_stats else None, ptr(st[1]) if want_stats else None, ptr(sws),
                                         stream_ptr(x.device)), "rg_igemm_run(fprop)")
            count_launch(3 if want_stats else 1)
            stats["igemm"] += 1
            return (y, st[0], st[1]) if want_stats else (y, None, None)
    if _pw_ok(cin_t, cout, kernel, stride, pad, groups):
        xb = K.conv_operand(x)
        y = _alloc(cdt)
        check(lib().rg_pw_conv_fprop(C.byref(t5(xb)), ptr(xb), ptr(_src(w).fp32()), ptr(bias_f), C.c_int(act),
                                     C.c_float(slope), C.byref(t5(y)), ptr(y), stream_ptr(x.device)), "rg_pw_conv_fprop")
        count_launch()
        stats["pointwise"] = stats.get("pointwise", 0) + 1
        return y, None, None
    y = _alloc(cdt)
    g = _geom(kernel, stride, pad, groups, act, slope)
    wc = _src(w).fp32()
    check(lib().rg_conv3d_direct_fprop(C.byref(t5(x)), ptr(x), ptr(wc), ptr(bias_f), C.byref(t5(y)), ptr(y),
                                       C.byref(g), stream_ptr(x.device)), "rg_conv3d_direct_fprop")
    count_launch(groups)
    stats["direct"] += 1
    return y, None, None


def conv_dgrad(dy, w, x_shape, stride, pad, groups=1, bias=None, act=ACT_NONE, slope=0.0, want_stats=False,
               out_dtype=None, out=None):
    """dx of the conv whose weight is `w` ([Cout,Cin/g,k..]); also ConvTranspose3d forward
    (then `bias`/`act` are the transposed conv's epilogue and x_shape its output shape)."""
    require_device(dy)
    n, cout_t, od, oh, ow = dy.shape
    cin_t = x_shape[1]
    kernel = tuple(w.shape[2:])
    cdt = out_dtype or K.compute_dtype()
    bias_f = bias.detach() if bias is not None else None
    if _igemm_ok(cin_t, cout_t, kernel, stride, groups) and cdt == torch.bfloat16:
        dyb = K.conv_operand(dy)
        if out is not None and tuple(out.shape) == tuple(x_shape) and out.dtype == torch.bfloat16 and out.stride(1) == 1:
            dx = out
        else:
            dx = K.empty5(tuple(x_shape), torch.bfloat16, dy.device)
        rd = _ring_desc(1, n, cin_t, cout_t, tuple(x_shape[2:]), (od, oh, ow), kernel, stride, pad, _sp_strides(dx),
                        _sp_strides(dyb), act, float(slope))
        if rd is not None:
            wr = ring_packed_weights(w, 1)
            st, sws = _stats_buffers(n * cin_t, lib().rg_ring_workspace_bytes(C.byref(rd)), dy.device) if want_stats else (None, None)
            with _Timed("dgrad", 2.0 * dyb.numel() * cin_t * kernel[0] * kernel[1] * kernel[2], dy.device, "ring_conv_kernel"):
                check(lib().rg_ring_run(C.byref(rd), ptr(dyb), ptr(wr), ptr(bias_f), ptr(dx),
                                        ptr(st[0]) if want_stats else None, ptr(st[1]) if want_stats else None, ptr(sws),
                                        stream_ptr(dy.device)), "rg_ring_run(dgrad)")
            count_launch(3 if want_stats else 1)
            stats["igemm"] += 1
            stats["ring"] = stats.get("ring", 0) + 1
            return (dx, st[0], st[1]) if want_stats else (dx, None, None)
        plan = _get_plan(1, n, cin_t, cout_t, tuple(x_shape[2:]), (od, oh, ow), kernel, stride, pad, _sp_strides(dx),
                         _sp_strides(dyb), act, float(slope))
        if plan is not None:
            _, wdp = packed_weights(w)
            st, sws = _stats_buffers(n * cin_t, lib().rg_plan_workspace_bytes(plan.handle), dy.device) if want_stats else (None, None)
            with _Timed("dgrad", 2.0 * dyb.numel() * cin_t * kernel[0] * kernel[1] * kernel[2], dy.device):
                check(lib().rg_igemm_run(plan.handle, ptr(dyb), ptr(wdp), ptr(bias_f), ptr(dx),
                                         ptr(st[0]) if want_stats else None, ptr(st[1]) if want_stats else None, ptr(sws),
                                         stream_ptr(dy.device)), "rg_igemm_run(dgrad)")
            count_launch(stride[0] * stride[1] * stride[2] + (2 if want_stats else 0))
            stats["igemm"] += 1
            return (dx, st[0], st[1]) if want_stats else (dx, None, None)
    if (_pw_ok(cin_t, cout_t, kernel, stride, pad, groups) and bias is None and act == ACT_NONE
            and cdt == torch.bfloat16):
        dx = K.empty5(tuple(x_shape), torch.bfloat16, dy.device)
        check(lib().rg_pw_conv_dgrad(C.byref(t5(dy)), ptr(dy), ptr(_src(w).fp32()), C.byref(t5(dx)), ptr(dx),
                                     stream_ptr(dy.device)), "rg_pw_conv_dgrad")
        count_launch()
        stats["pointwise"] = stats.get("pointwise", 0) + 1
        return dx, None, None
    if (cin_t < 64 and cout_t % 64 == 0 and groups == 1 and tuple(stride) == (1, 1, 1) and bias is None and act == ACT_NONE
            and kernel[0] * kernel[1] * kernel[2] > 8 and cdt == torch.bfloat16 and K.use_tensor_cores()
            and _igemm_ok(64, cout_t, kernel, stride, groups)):
        # few input channels, many taps (the U-Net stem under a cycle loss, UNet.py:52-87): tensor cores with the
        # produced channels zero-padded to 64 instead of streaming Cout * taps values per voxel
        base = _src(w)
        inner = base.fn
        wp = WeightSrc(base.param, (lambda p_: _pad_cin(inner(p_) if inner else p_, 64)), (cout_t, 64, *kernel),
                       base.tag + "+cin64")
        dxp, _, _ = conv_dgrad(dy, wp, (n, 64, *x_shape[2:]), stride, pad, groups, out_dtype=torch.bfloat16)
        dx = K.copy_into(K.empty5(tuple(x_shape), cdt, dy.device), dxp[:, :cin_t])
        return dx, None, None
    if (cin_t <= 4 and cout_t % 8 == 0 and groups == 1 and bias is None and act == ACT_NONE and K.channels_inner_bf16(dy)
            and cin_t * cout_t * kernel[0] * kernel[1] * kernel[2] * 4 <= 48 * 1024):
        # few input channels (discriminator stem under the generator loss): streaming kernel instead of the tiled GEMM
        dx = K.empty5(tuple(x_shape), cdt, dy.device)
        wc = _src(w).fp32()
        check(lib().rg_smallcin_dgrad(C.byref(t5(dy)), ptr(dy), ptr(wc), C.byref(t5(dx)), ptr(dx),
                                      C.byref(_geom(kernel, stride, pad, groups)), stream_ptr(dy.device)), "rg_smallcin_dgrad")
        count_launch()
        stats["smallcin_dgrad"] = stats.get("smallcin_dgrad", 0) + 1
        return dx, None, None
    dx = K.empty5(tuple(x_shape), cdt, dy.device)
    g = _geom(kernel, stride, pad, groups, act, slope)
    wc = _src(w).fp32()
    check(lib().rg_conv3d_direct_dgrad(C.byref(t5(dy)), ptr(dy), ptr(wc), ptr(bias_f), C.byref(t5(dx)), ptr(dx),
                                       C.byref(g), stream_ptr(dy.device)), "rg_conv3d_direct_dgrad")
    count_launch(groups)
    stats["direct"] += 1
    return dx, None, None


SMALLK_WGRAD = os.environ.get("RAYGUN_B200_SMALLK_WGRAD", "1") != "0"


def smallk_wgrad_ok(x, dy, w_shape, groups=1):
    """The CUDA-core small-reduction weight gradient (smallk_wgrad.cu): Cin * taps <= 64 onto a multiple of 64 channels."""
    cout, cin_g = w_shape[0], w_shape[1]
    return (SMALLK_WGRAD and K.get_precision() == "bf16" and groups == 1 and cout % 64 == 0
            and 1 < cin_g * w_shape[2] * w_shape[3] * w_shape[4] <= 64 and x.dim() == 5 and x.is_cuda
            and x.dtype in (torch.bfloat16, torch.float32) and K.channels_inner_bf16(dy))


def conv_wgrad(x, dy, w_shape, stride, pad, groups=1, want_bias=False, db=None, dw_out=None, accumulate=False):
    """dw (fp32, conv layout) and optionally dbias = sum(dy) (`db`: already computed by the caller's activation pass).
    `dw_out`: write (or, with `accumulate`, add) the weight gradient into this fp32 tensor of shape `w_shape` -- a view of
    a flat gradient bucket (optim.FlatAdam) -- instead of a fresh one.  Every reduction is two-stage and ordered: the result
    is bit-reproducible."""
    require_device(x)
    cout, cin_g = w_shape[0], w_shape[1]
    kernel = tuple(w_shape[2:])
    acc = C.c_int(int(bool(accumulate and dw_out is not None)))

    def _dw():
        if dw_out is not None:
            if tuple(dw_out.shape) != tuple(w_shape) or dw_out.dtype != torch.float32 or not dw_out.is_contiguous():
                raise RuntimeError("conv_wgrad: dw_out must be a contiguous fp32 tensor of the weight's shape")
            return dw_out
        return torch.empty(tuple(w_shape), dtype=torch.float32, device=x.device)

    if _igemm_ok(cin_g, cout, kernel, stride, groups) and groups == 1:
        xb, dyb = K.conv_operand(x), K.conv_operand(dy)      # strided channel-inner views (arena interiors) in place
        n = x.shape[0]
        hd = _halo_wgrad_desc(n, cin_g, cout, tuple(x.shape[2:]), tuple(dy.shape[2:]), kernel, stride, pad,
                              _sp_strides(xb), _sp_strides(dyb))
        if hd is not None:
            dw = _dw()
            key = ("halo", n, cin_g, cout, tuple(x.shape[2:]), kernel, tuple(pad))
            nb = _ws_bytes_cache.get(key)
            if nb is None:
                nb = _ws_bytes_cache[key] = int(lib().rg_halo_wgrad_workspace_bytes(C.byref(hd)))
            ws = K.ws_bytes(nb, x.device)
            with _Timed("wgrad", 2.0 * dyb.numel() * cin_g * kernel[0] * kernel[1] * kernel[2], x.device, "halo_wgrad_kernel"):
                check(lib().rg_halo_wgrad(C.byref(hd), ptr(xb), ptr(dyb), ptr(dw), ptr(ws), acc, stream_ptr(x.device)),
                      "rg_halo_wgrad")
            count_launch(2)
            stats["igemm"] += 1
            stats["halo_wgrad"] = stats.get("halo_wgrad", 0) + 1
            db = (db if db is not None else K.norm_stats(dyb, _C.NORM_BATCH)[0]) if want_bias else None
            return dw, db
        plan = _get_plan(2, n, cin_g, cout, tuple(x.shape[2:]), tuple(dy.shape[2:]), kernel, stride, pad,
                         _sp_strides(xb), _sp_strides(dyb), ACT_NONE, 0.0)
        if plan is not None:
            dw = _dw()
            ws = K.ws_bytes(lib().rg_plan_workspace_bytes(plan.handle), x.device)
            with _Timed("wgrad", 2.0 * dyb.numel() * cin_g * kernel[0] * kernel[1] * kernel[2], x.device, "igemm_wgrad_kernel"):
                check(lib().rg_igemm_wgrad(plan.handle, ptr(xb), ptr(dyb), ptr(dw), None, ptr(ws), acc,
                                           stream_ptr(x.device)), "rg_igemm_wgrad")
            count_launch(2)
            stats["igemm"] += 1
            db = (db if db is not None else K.norm_stats(dyb, _C.NORM_BATCH)[0]) if want_bias else None
            return dw, db
    if smallk_wgrad_ok(x, dy, w_shape, groups):
        g = _geom(kernel, stride, pad, 1)
        xd, dyd = t5(x), t5(dy)
        if lib().rg_smallk_wgrad_supported(C.byref(xd), C.byref(dyd), C.byref(g)) == 1:
            kk = cin_g * kernel[0] * kernel[1] * kernel[2]
            dw = _dw()
            own_db = want_bias and db is None
            if own_db:
                db = torch.empty(cout, dtype=torch.float32, device=x.device)
            ws = K.ws_bytes(lib().rg_smallk_wgrad_workspace_bytes(C.c_int(cout), C.c_int(kk)), x.device)
            check(lib().rg_smallk_wgrad(C.byref(xd), ptr(x), C.byref(dyd), ptr(dy), C.byref(g), ptr(dw),
                                        ptr(db) if own_db else None, ptr(ws), acc, stream_ptr(x.device)), "rg_smallk_wgrad")
            count_launch(3 if own_db else 2)
            stats["smallk_wgrad"] = stats.get("smallk_wgrad", 0) + 1
            return dw, (db if want_bias else None)
    if _pw_ok(cin_g, cout, kernel, stride, pad, groups):
        xb = K.conv_operand(x)
        dw = _dw()
        db = torch.empty(cout, dtype=torch.float32, device=x.device) if want_bias else None
        ws = K.ws_bytes(lib().rg_pw_conv_wgrad_workspace_bytes(C.c_int(cin_g), C.c_int(cout)), x.device)
        check(lib().rg_pw_conv_wgrad(C.byref(t5(xb)), ptr(xb), C.byref(t5(dy)), ptr(dy), ptr(dw), ptr(db), ptr(ws), acc,
                                     stream_ptr(x.device)), "rg_pw_conv_wgrad")
        count_launch(3 if want_bias else 2)
        stats["pointwise"] = stats.get("pointwise", 0) + 1
        return dw, db
    g = _geom(tuple(w_shape[2:]), stride, pad, groups)
    xd, dyd = t5(x), t5(dy)
    ws = K.ws_bytes(lib().rg_conv3d_direct_wgrad_workspace_bytes(C.byref(xd), C.byref(dyd), C.byref(g)), x.device)
    dw = _dw()
    db = torch.empty(w_shape[0], dtype=torch.float32, device=x.device) if want_bias else None
    check(lib().rg_conv3d_direct_wgrad(C.byref(xd), ptr(x), C.byref(dyd), ptr(dy), ptr(dw), ptr(db), C.byref(g), acc,
                                       ptr(ws), stream_ptr(x.device)), "rg_conv3d_direct_wgrad")
    count_launch((3 if want_bias else 2) * groups)
    stats["direct"] += 1
    return dw, db
