"""Encoder / Decoder of the VAE: parameter containers that are state_dict-compatible with the
reference (src/models/encoder_decoder.py) plus the hand-scheduled forward/backward that drives
the sm_100a kernels.

The nn.Conv2d / nn.GroupNorm children are used ONLY as parameter holders (same names, shapes,
construction order and default init as the reference, so a seeded construction reproduces the
reference's initial weights and `load_state_dict` works both ways).  Their `.forward` is never
called: `Encoder.forward` / `Decoder.forward` run the NHWC kernel schedule below inside one
`torch.autograd.Function`, whose backward is the reverse schedule (dgrad + wgrad per conv, fused
GroupNorm+SiLU backward with the residual-gradient add, attention backward).
"""
from __future__ import annotations

import os
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.nn as nn

from . import _lib as L
from . import ops

GN_EPS = 1e-6
# GroupNorm reductions ride in the epilogues of the neighbouring tcgen05 convs (bf16 mode); set
# MVAE_GN_FUSE=0 to run the stand-alone two-pass GroupNorm kernels instead (A/B and cross-checks).
FUSE_GN = os.environ.get("MVAE_GN_FUSE", "1") != "0"


def _norm(c: int) -> nn.GroupNorm:
    """encoder_decoder.py:28-33 — GroupNorm(min(32, C), C, eps=1e-6, affine)."""
    return nn.GroupNorm(num_groups=min(32, c), num_channels=c, eps=GN_EPS, affine=True)


# ----------------------------------------------------------------------------------------------
# parameter containers
# ----------------------------------------------------------------------------------------------
class ResnetBlock(nn.Module):
    """encoder_decoder.py:110-170 (temb_channels == 0, conv_shortcut == False always)."""

    def __init__(self, in_channels: int, out_channels: Optional[int] = None, dropout: float = 0.0):
        super().__init__()
        out_channels = in_channels if out_channels is None else out_channels
        self.in_channels, self.out_channels = in_channels, out_channels
        self.norm1 = _norm(in_channels)
        self.conv1 = nn.Conv2d(in_channels, out_channels, 3, 1, 1)
        self.norm2 = _norm(out_channels)
        self.dropout = nn.Dropout(dropout)
        self.conv2 = nn.Conv2d(out_channels, out_channels, 3, 1, 1)
        if in_channels != out_channels:
            self.nin_shortcut = nn.Conv2d(in_channels, out_channels, 1, 1, 0)


class AttnBlock(nn.Module):
    """encoder_decoder.py:68-107."""

    def __init__(self, in_channels: int):
        super().__init__()
        self.in_channels = in_channels
        self.norm = _norm(in_channels)
        self.q = nn.Conv2d(in_channels, in_channels, 1, 1, 0)
        self.k = nn.Conv2d(in_channels, in_channels, 1, 1, 0)
        self.v = nn.Conv2d(in_channels, in_channels, 1, 1, 0)
        self.proj_out = nn.Conv2d(in_channels, in_channels, 1, 1, 0)


class Downsample(nn.Module):
    """encoder_decoder.py:173-191 (with_conv=True: pad (0,1,0,1) + conv3x3 s2 p0)."""

    def __init__(self, in_channels: int, with_conv: bool = True):
        super().__init__()
        if not with_conv:
            raise NotImplementedError("Downsample(with_conv=False) is never built by the reference models")
        self.with_conv = with_conv
        self.conv = nn.Conv2d(in_channels, in_channels, 3, 2, 0)


class Upsample(nn.Module):
    """encoder_decoder.py:194-209 (nearest 2x + conv3x3)."""

    def __init__(self, in_channels: int, with_conv: bool = True):
        super().__init__()
        if not with_conv:
            raise NotImplementedError("Upsample(with_conv=False) is never built by the reference models")
        self.with_conv = with_conv
        self.conv = nn.Conv2d(in_channels, in_channels, 3, 1, 1)


class LinearAttention(nn.Module):
    """encoder_decoder.py:36-59 (parameter container; heads * dim_head hidden channels)."""

    def __init__(self, dim: int, heads: int = 4, dim_head: int = 32):
        super().__init__()
        self.heads = heads
        hidden_dim = dim_head * heads
        self.to_qkv = nn.Conv2d(dim, hidden_dim * 3, 1, bias=False)
        self.to_out = nn.Conv2d(hidden_dim, dim, 1)


class LinAttnBlock(LinearAttention):
    """encoder_decoder.py:62-65: heads=1, dim_head=in_channels; no GroupNorm, no residual."""

    def __init__(self, in_channels: int):
        super().__init__(dim=in_channels, heads=1, dim_head=in_channels)
        self.in_channels = in_channels


def make_attn(in_channels: int, attn_type: str = "vanilla") -> nn.Module:
    """encoder_decoder.py:18-25."""
    if attn_type == "vanilla":
        return AttnBlock(in_channels)
    if attn_type == "linear":
        return LinAttnBlock(in_channels)
    raise NotImplementedError(f"Attention type {attn_type} not implemented")


# ----------------------------------------------------------------------------------------------
# kernel schedule
# ----------------------------------------------------------------------------------------------
class Run:
    """State of one forward (+ its backward) through a stack: dtype, saved activations and the
    parameter gradients produced so far."""

    def __init__(self, dtype: torch.dtype, cache: dict, training: bool, sink=None, views=None):
        self.sink = sink              # callable(param, grad): takes the gradient instead of autograd
        self.views = views            # train.GradBuckets: kernels accumulate straight into its flat views
        self.dtype = dtype
        self.code = L.code_of(dtype)
        self.cache = cache            # packed-operand cache owned by the nn.Module
        self.training = training
        self.fuse_gn = FUSE_GN and dtype == torch.bfloat16
        self.grads: Dict[nn.Parameter, torch.Tensor] = {}
        self.tape: list = []
        self.dropout_seed = 0          # set by _StackFn from the owning stack
        self.dropout_offset = 0
        self.sample_base = 0           # first sample of this forward in the global batch (Philox element counter)
        self.dropout_masks = None      # verification hook: {id(ResnetBlock): uint8 NHWC keep mask}
        self.step_state = None         # device-side stream position (TrainStep graph mode)
        self.taps = None               # verification hook: {reference module path: NHWC activation} (stack.tap_sink)
        self.tap_names = None          # {id(module): path}

    def tap(self, module, suffix: str, t: torch.Tensor):
        """record a per-layer activation under the reference's module path (parity tests compare these with the
        oracle's / the golden fixtures' taps); no-op unless the owning stack has a tap_sink"""
        if self.taps is not None:
            name = self.tap_names.get(id(module))
            if name is not None:
                self.taps[name + suffix] = t

    def next_dropout(self, blk):
        """-> (seed, offset, mask override or None) for the next dropout site of this forward"""
        self.dropout_offset += 1
        m = self.dropout_masks.get(id(blk)) if self.dropout_masks else None
        return self.dropout_seed, self.dropout_offset, m

    def add_grad(self, p: nn.Parameter, g: torch.Tensor):
        if self.sink is not None:
            self.sink(p, g)
            return
        g = g.view(p.shape)
        if p in self.grads:
            self.grads[p] = self.grads[p] + g
        else:
            self.grads[p] = g


# include/medvae_b200.h: mvae_pack_job (80 bytes)
_PACK_JOB = np.dtype([("w", "<u8"), ("wf", "<u8"), ("wd", "<u8"), ("bsrc", "<u8"), ("bdst", "<u8"), ("out_c", "<i4"),
                      ("in_c", "<i4"), ("ksize", "<i4"), ("mode", "<i4"), ("n_slabs", "<i4"), ("ctot", "<i4"),
                      ("coff", "<i4"), ("cinp", "<i4"), ("tile_begin", "<i4"), ("reserved", "<i4")])
_pack_tables: dict = {}


def pack_entries(entries, device):
    """(re)pack the operands of these cache entries with ONE launch and mark them current"""
    if not entries:
        return
    code = entries[0]["code"]
    key = (tuple(id(e) for e in entries), tuple(c.weight.data_ptr() for e in entries for c, _ in e["jobs"]), str(device))
    tab = _pack_tables.get(key)
    if tab is None:
        rows, tile = [], 0
        for e in entries:
            assert e["code"] == code
            op = e["op"]
            for conv, off in e["jobs"]:
                cout, cin = conv.out_channels, conv.in_channels
                rows.append((conv.weight.data_ptr(), e["wf"].data_ptr(), e["wd"].data_ptr(),
                             conv.bias.data_ptr() if conv.bias is not None else 0,
                             e["bias"].data_ptr() if e["own_bias"] else 0, cout, cin, op.k, op.mode, e["nslabs"], e["ctot"],
                             off, e["cinp"], tile, 0))
                tile += -(-cout // 32) * -(-cin // 32)
        table = torch.from_numpy(np.array(rows, dtype=_PACK_JOB).view(np.uint8).copy()).to(device)
        tab = (table, len(rows), tile)
        if len(entries) > 1:          # the whole-model table of the per-step refresh is reused every step
            _pack_tables.clear()
            _pack_tables[key] = tab
    ops.conv_pack_weights_batched(tab[0], tab[1], tab[2], code)
    for e in entries:
        e["ver"] = e["op"]._versions()


def mark_packed_current(stacks):
    """the packed operands were refreshed behind Python's back (a CUDA-graph replay ends with the re-pack launch): record
    the parameters' current versions so eager consumers do not pack again"""
    for st in stacks:
        for e in st._pack_cache.values():
            e["ver"] = e["op"]._versions()


def refresh_packed(stacks):
    """one launch re-packing every conv operand of these stacks whose master weights changed (after the optimizer step)"""
    stale, dev = [], None
    for st in stacks:
        for e in st._pack_cache.values():
            if e["ver"] != e["op"]._versions():
                stale.append(e)
                dev = e["dev"]
    by_code = {}
    for e in stale:
        by_code.setdefault(e["code"], []).append(e)
    for es in by_code.values():
        pack_entries(es, dev)


PAD_C = 64   # skinny convs are zero-padded to one tcgen05 K / N tile in bf16 mode


class ConvOp:
    """One (or several channel-concatenated) nn.Conv2d as packed fprop/dgrad operands.

    pad_in / pad_out: in bf16 mode a conv whose Cin (encoder conv_in: 1..6 channels) or Cout
    (decoder conv_out: 1..3 channels) is below one tensor-core tile is run with that dimension
    zero-padded to 64, so it takes the tcgen05 path instead of the SIMT one (these two
    layers are bandwidth-bound; the padded MMA work is negligible).  The caller supplies / receives
    activations with the padded channel count (`cin_of`, `cout_of`)."""

    def __init__(self, convs: Sequence[nn.Conv2d], mode: int, pad_in: bool = False, pad_out: bool = False):
        self.convs = list(convs)
        c0 = self.convs[0]
        self.cin = c0.in_channels
        self.couts = [c.out_channels for c in self.convs]
        self.cout = sum(self.couts)
        self.k = c0.kernel_size[0]
        self.mode = mode
        self.pad_in, self.pad_out = pad_in, pad_out

    def cin_of(self, dtype) -> int:
        if self.pad_in and dtype == torch.bfloat16 and self.cin < PAD_C and self.cout % 64 == 0:
            return PAD_C
        return self.cin

    def cout_of(self, dtype) -> int:
        if self.pad_out and dtype == torch.bfloat16 and self.cout < PAD_C and self.cin % 64 == 0:
            return PAD_C
        return self.cout

    def desc(self, run: Run, b: int, h: int, w: int):
        d = ops.conv_desc(b, h, w, self.cin_of(run.dtype), self.cout_of(run.dtype), self.k, self.mode, run.code)
        d.real_channels = (self.cin, self.cout)     # for the FLOP accounting of the profiler (ops.conv_flops)
        return d

    def params(self):
        out = []
        for c in self.convs:
            out += [c.weight] + ([c.bias] if c.bias is not None else [])
        return out

    def _versions(self):
        return tuple((c.weight._version, c.bias._version if c.bias is not None else 0) for c in self.convs)

    def packed(self, run: Run, device):
        """-> cache entry {"wf", "wd", "bias"}: the packed bf16 / fp32 operands of this conv (group), refreshed when the
        master weights changed.  train.TrainStep refreshes ALL entries of a model with one launch right after the
        optimizer step (refresh_packed); this per-conv path serves the first use and externally modified weights."""
        key = (tuple(id(c.weight) for c in self.convs), run.code, self.mode)
        ent = run.cache.get(key)
        if ent is None or ent["dev"] != device:
            cin_p, cout_p = self.cin_of(run.dtype), self.cout_of(run.dtype)
            d1 = ops.conv_desc(1, 8, 8, cin_p, cout_p, self.k, self.mode, run.code)
            nslabs = ops.conv_num_slabs(d1)
            padded = cin_p != self.cin or cout_p != self.cout
            alloc = torch.zeros if padded else torch.empty      # the tile padding of a skinny conv stays zero
            wf = alloc((nslabs, cout_p, cin_p), dtype=run.dtype, device=device)
            wd = alloc((nslabs, cin_p, cout_p), dtype=run.dtype, device=device)
            single = len(self.convs) == 1 and not padded and self.convs[0].bias is not None
            if single:
                bias, own_bias = self.convs[0].bias.detach(), False       # the parameter itself: always current
            else:
                bias, own_bias = torch.zeros((cout_p,), dtype=torch.float32, device=device), True
            jobs, off = [], 0
            for c in self.convs:
                jobs.append((c, off))
                off += c.out_channels
            ent = {"ver": None, "dev": device, "wf": wf, "wd": wd, "bias": bias, "own_bias": own_bias, "jobs": jobs,
                   "op": self, "nslabs": nslabs, "ctot": cout_p, "cinp": cin_p, "code": run.code}
            run.cache[key] = ent
        ver = self._versions()
        if ent["ver"] != ver:
            pack_entries([ent], device)
        return ent

    def fwd(self, run: Run, x: torch.Tensor, residual: Optional[torch.Tensor] = None, gn_groups: int = 0):
        """-> (y, sums): sums = per-(sample, group) (sum, sum of squares) of y for the GroupNorm with
        `gn_groups` groups that consumes y, when the conv kernel can produce them; else None."""
        b, h, w, _ = x.shape
        ent = self.packed(run, x.device)
        d = self.desc(run, b, h, w)
        if gn_groups and run.fuse_gn and ops.conv_fuses_gn(d, gn_groups, 0):
            return ops.conv_fprop_gn(d, x, ent["wf"], ent["bias"], residual, gn_groups)
        return ops.conv_fprop(d, x, ent["wf"], ent["bias"], residual), None

    def bwd(self, run: Run, x: torch.Tensor, dy: torch.Tensor, need_dx: bool = True, dy_colsum=None, gn=None):
        """parameter gradients + (dx, None), or, with gn = (gn_x, stats, norm, silu) — x is the output
        of [SiLU](GroupNorm(gn_x)) — (du, s12) from the fused dgrad epilogue when available.
        dy_colsum: the column sums of dy if the producer of dy already has them (bias gradient)."""
        b, h, w, _ = x.shape
        d = self.desc(run, b, h, w)
        ent = self.packed(run, x.device)
        vw = vb = None
        if len(self.convs) == 1 and self.convs[0].bias is None:
            dw, _ = ops.conv_wgrad(d, x, dy, False)
            run.add_grad(self.convs[0].weight, dw)
        elif run.views is not None and len(self.convs) == 1 and d.in_c == self.cin and d.out_c == self.cout:
            vw, vb = run.views.view_of(self.convs[0].weight), run.views.view_of(self.convs[0].bias)
        if len(self.convs) == 1 and self.convs[0].bias is None:
            pass
        elif vw is not None and vb is not None:
            # the wgrad finalisation adds into the flat gradient buffer (no temporary, no add pass)
            c = self.convs[0]
            ops.conv_wgrad_acc(d, x, dy, vw, vb if dy_colsum is None else None, 1.0)
            run.views.mark_ready(c.weight)
            if dy_colsum is None:
                run.views.mark_ready(c.bias)
            else:
                run.add_grad(c.bias, dy_colsum)
        else:
            dw, db = ops.conv_wgrad(d, x, dy, dy_colsum is None)
            if db is None:
                db = dy_colsum
            if d.in_c != self.cin or d.out_c != self.cout:
                dw = dw[:self.cout, :self.cin].contiguous()
                db = db[:self.cout]
            off = 0
            for c in self.convs:
                run.add_grad(c.weight, dw[off:off + c.out_channels])
                run.add_grad(c.bias, db[off:off + c.out_channels])
                off += c.out_channels
        if not need_dx:
            return None, None
        if gn is not None and run.fuse_gn:
            gn_x, stats, norm, silu = gn
            if ops.conv_fuses_gn(d, norm.num_groups, 1):
                return ops.conv_dgrad_gn(d, dy, ent["wd"], gn_x, stats, norm.weight.detach(), norm.bias.detach(),
                                         norm.num_groups, silu)
        return ops.conv_dgrad(d, dy, ent["wd"]), None


def gn_fwd(run: Run, norm: nn.GroupNorm, x, silu: bool, sums=None):
    """-> (y, stats).  sums: the group sums of x from the conv that produced it (one-pass kernel)."""
    if sums is not None:
        return ops.gn_apply_fwd(x, norm.weight.detach(), norm.bias.detach(), sums, norm.num_groups, norm.eps, silu)
    return ops.gn_fwd(x, norm.weight.detach(), norm.bias.detach(), norm.num_groups, norm.eps, silu)


def gn_bwd(run: Run, norm: nn.GroupNorm, dy, x, stats, dx_add, silu: bool, s12=None):
    """-> (dx, colsum(dx) or None).  With s12, `dy` is already du (fused dgrad epilogue) and the
    one-pass kernel also returns the column sums of dx."""
    if s12 is not None:
        dx, dg, db, cs = ops.gn_apply_bwd(dy, x, norm.weight.detach(), stats, s12, dx_add, norm.num_groups)
    else:
        dx, dg, db = ops.gn_bwd(dy, x, norm.weight.detach(), norm.bias.detach(), stats, dx_add, norm.num_groups, silu)
        cs = None
    run.add_grad(norm.weight, dg)
    run.add_grad(norm.bias, db)
    return dx, cs


class ResOp:
    """x + conv2(silu(gn2(conv1(silu(gn1(x)))))) with an optional 1x1 shortcut
    (encoder_decoder.py:148-170).  The residual add is fused into conv2's epilogue; in the
    backward the shortcut gradient is fused into the GroupNorm-1 backward."""

    def __init__(self, blk: ResnetBlock):
        self.blk = blk
        self.in_groups, self.out_groups = blk.norm1.num_groups, 0
        self.c1 = ConvOp([blk.conv1], L.CONV_S1)
        self.c2 = ConvOp([blk.conv2], L.CONV_S1)
        self.nin = ConvOp([blk.nin_shortcut], L.CONV_S1) if hasattr(blk, "nin_shortcut") else None
        if blk.dropout.p > 0.0:
            self.p_drop = blk.dropout.p
        else:
            self.p_drop = 0.0

    def params_backward_order(self):
        b = self.blk
        return (self.c2.params() + [b.norm2.weight, b.norm2.bias] + self.c1.params()
                + (self.nin.params() if self.nin is not None else []) + [b.norm1.weight, b.norm1.bias])

    def fwd(self, run: Run, x, x_sums=None):
        blk = self.blk
        a1, st1 = gn_fwd(run, blk.norm1, x, True, x_sums)
        h1, h1_sums = self.c1.fwd(run, a1, gn_groups=blk.norm2.num_groups)
        run.tap(blk, ".conv1", h1)
        a2, st2 = gn_fwd(run, blk.norm2, h1, True, h1_sums)
        drop = None
        if self.p_drop > 0.0 and run.training:
            # nn.Dropout between swish(norm2(h)) and conv2 (encoder_decoder.py:164): Philox keep mask, regenerated
            # in the backward from (seed, offset); a2 below is the dropped tensor conv2 consumes
            drop = run.next_dropout(blk) + (run.sample_base * (a2.numel() // a2.shape[0]),)
            a2, _ = ops.dropout_apply(a2, self.p_drop, drop[0], drop[1], mask_in=drop[2], step_state=run.step_state,
                                      elem_base=drop[3])
        xs = self.nin.fwd(run, x)[0] if self.nin is not None else x
        out, out_sums = self.c2.fwd(run, a2, residual=xs, gn_groups=self.out_groups)
        run.tape.append((self, (x, st1, a1, h1, st2, a2, drop)))
        run.tap(blk, "", out)
        return out, out_sums

    def bwd(self, run: Run, saved, dout, dout_cs=None):
        x, st1, a1, h1, st2, a2, drop = saved
        blk = self.blk
        if drop is not None:
            # the dropout sits between conv2's dgrad and the GroupNorm backward: no fused epilogue here
            d_a2, _ = self.c2.bwd(run, a2, dout, dy_colsum=dout_cs)
            d_a2, _ = ops.dropout_apply(d_a2, self.p_drop, drop[0], drop[1], mask_in=drop[2], step_state=run.step_state,
                                        elem_base=drop[3])
            s12 = None
        else:
            d_a2, s12 = self.c2.bwd(run, a2, dout, dy_colsum=dout_cs, gn=(h1, st2, blk.norm2, True))
        d_h1, h1_cs = gn_bwd(run, blk.norm2, d_a2, h1, st2, None, True, s12)
        del d_a2
        d_a1, s12 = self.c1.bwd(run, a1, d_h1, dy_colsum=h1_cs, gn=(x, st1, blk.norm1, True))
        del d_h1
        addend = self.nin.bwd(run, x, dout, dy_colsum=dout_cs)[0] if self.nin is not None else dout
        return gn_bwd(run, blk.norm1, d_a1, x, st1, addend, True, s12)


class AttnOp:
    """x + proj_out(attn(q,k,v of gn(x)))  (encoder_decoder.py:83-107); q,k,v are one N=3C GEMM."""

    def __init__(self, blk: AttnBlock):
        self.blk = blk
        self.in_groups, self.out_groups = blk.norm.num_groups, 0
        self.qkv = ConvOp([blk.q, blk.k, blk.v], L.CONV_S1)
        self.proj = ConvOp([blk.proj_out], L.CONV_S1)

    def params_backward_order(self):
        return self.proj.params() + self.qkv.params() + [self.blk.norm.weight, self.blk.norm.bias]

    def fwd(self, run: Run, x, x_sums=None):
        hn, st = gn_fwd(run, self.blk.norm, x, False, x_sums)
        qkv, _ = self.qkv.fwd(run, hn)
        o, probs = ops.attn_fwd(qkv)
        out, out_sums = self.proj.fwd(run, o, residual=x, gn_groups=self.out_groups)
        run.tape.append((self, (x, st, hn, qkv, o, probs)))
        run.tap(self.blk, "", out)
        return out, out_sums

    def bwd(self, run: Run, saved, dout, dout_cs=None):
        x, st, hn, qkv, o, probs = saved
        d_o, _ = self.proj.bwd(run, o, dout, dy_colsum=dout_cs)
        d_qkv = ops.attn_bwd(qkv, probs, d_o)
        del d_o
        d_hn, s12 = self.qkv.bwd(run, hn, d_qkv, gn=(x, st, self.blk.norm, False))
        del d_qkv
        return gn_bwd(run, self.blk.norm, d_hn, x, st, dout, False, s12)


class LinAttnOp:
    """to_out(linear_attention(to_qkv(x)))  (encoder_decoder.py:46-59 with heads=1, dim_head=C): softmax of k over
    the tokens, context = k v^T, out = context^T q.  No GroupNorm and no residual, as in the reference."""

    def __init__(self, blk: LinAttnBlock):
        self.blk = blk
        self.in_groups, self.out_groups = 0, 0
        self.qkv = ConvOp([blk.to_qkv], L.CONV_S1)
        self.proj = ConvOp([blk.to_out], L.CONV_S1)

    def params_backward_order(self):
        return self.proj.params() + self.qkv.params()

    def fwd(self, run: Run, x, x_sums=None):
        qkv, _ = self.qkv.fwd(run, x)
        o, ks, ctx = ops.linattn_fwd(qkv)
        out, out_sums = self.proj.fwd(run, o, gn_groups=self.out_groups)
        run.tape.append((self, (x, qkv, o, ks, ctx)))
        run.tap(self.blk, "", out)
        return out, out_sums

    def bwd(self, run: Run, saved, dout, dout_cs=None):
        x, qkv, o, ks, ctx = saved
        d_o, _ = self.proj.bwd(run, o, dout, dy_colsum=dout_cs)
        d_qkv = ops.linattn_bwd(qkv, ks, ctx, d_o)
        del d_o
        return self.qkv.bwd(run, x, d_qkv)[0], None


def attn_stage(blk: nn.Module):
    return LinAttnOp(blk) if isinstance(blk, LinAttnBlock) else AttnOp(blk)


class PlainConvOp:
    """a conv that is its own stage (conv_in, Downsample, Upsample)."""

    def __init__(self, conv: nn.Conv2d, mode: int, need_dx: bool = True, pad_in: bool = False):
        self.op = ConvOp([conv], mode, pad_in=pad_in)
        self.need_dx = need_dx
        self.in_groups, self.out_groups = 0, 0

    def params_backward_order(self):
        return self.op.params()

    def fwd(self, run: Run, x, x_sums=None):
        y, y_sums = self.op.fwd(run, x, gn_groups=self.out_groups)
        run.tape.append((self, (x,)))
        run.tap(self.op.convs[0], "", y)
        return y, y_sums

    def bwd(self, run: Run, saved, dout, dout_cs=None):
        return self.op.bwd(run, saved[0], dout, self.need_dx, dy_colsum=dout_cs)[0], None


class HeadOp:
    """norm_out -> swish -> conv_out (encoder_decoder.py:325-327, :446-450)."""

    def __init__(self, norm: nn.GroupNorm, conv: nn.Conv2d, pad_out: bool = False):
        self.norm = norm
        self.op = ConvOp([conv], L.CONV_S1, pad_out=pad_out)
        self.in_groups, self.out_groups = norm.num_groups, 0

    def params_backward_order(self):
        return self.op.params() + [self.norm.weight, self.norm.bias]

    def fwd(self, run: Run, x, x_sums=None):
        a, st = gn_fwd(run, self.norm, x, True, x_sums)
        y, _ = self.op.fwd(run, a)
        run.tape.append((self, (x, st, a)))
        run.tap(self.op.convs[0], "", y)
        return y, None

    def bwd(self, run: Run, saved, dout, dout_cs=None):
        x, st, a = saved
        d_a, s12 = self.op.bwd(run, a, dout, dy_colsum=dout_cs, gn=(x, st, self.norm, True))
        return gn_bwd(run, self.norm, d_a, x, st, None, True, s12)


def link_stages(stages):
    """tell every stage how many GroupNorm groups the NEXT stage normalises its input with, so
    its last conv can produce that GroupNorm's statistics in its epilogue."""
    for cur, nxt in zip(stages[:-1], stages[1:]):
        cur.out_groups = nxt.in_groups
    return stages


def run_forward(stages, run: Run, x):
    sums = None
    for s in stages:
        x, sums = s.fwd(run, x, sums)
    return x


def run_backward(run: Run, dout):
    cs = None
    while run.tape:
        stage, saved = run.tape.pop()
        dout, cs = stage.bwd(run, saved, dout, cs)
        del saved
    return dout


# ----------------------------------------------------------------------------------------------
# Encoder / Decoder modules
# ----------------------------------------------------------------------------------------------
class _Stack(nn.Module):
    """common machinery: dtype selection, packed-operand cache, autograd bridge."""

    def _init_runtime(self):
        self._pack_cache: dict = {}
        self._grad_sink = None        # set by train.TrainStep: gradients go to its flat buckets
        self._grad_views = None       # set by train.TrainStep: the GradBuckets object itself
        self.compute_dtype = torch.bfloat16
        self.dropout_seed: Optional[int] = None    # None: torch.initial_seed()
        self._dropout_step = 0
        self._rng_sample_base = 0                  # first sample of this forward in the global batch (train.TrainStep)
        self.dropout_mask_override = None          # verification hook: {ResnetBlock module: uint8 NHWC keep mask}
        self._step_state = None                    # device-side Philox position (train.TrainStep graph mode)
        # verification hook: a dict that receives every stage's output (NHWC, compute dtype, padded channels) under the
        # reference's module path, e.g. "down.0.block.1", "down.0.block.1.conv1", "mid.attn_1", "down.1.downsample",
        # "conv_in", "conv_out" (the names oracle.vae_oracle's `taps` and the golden fixtures use, minus the
        # "encoder." / "decoder." prefix)
        self.tap_sink: Optional[dict] = None

    def stages(self, need_input_grad: bool):
        raise NotImplementedError

    def param_list(self) -> List[nn.Parameter]:
        return list(self.parameters())

    def forward_nhwc(self, x_nhwc: torch.Tensor, need_input_grad: bool):
        """x NHWC (compute dtype) -> y NHWC, differentiable w.r.t. x and the parameters."""
        L.require_device()
        params = self.param_list()
        return _StackFn.apply(self, need_input_grad, x_nhwc, *params)


class _StackFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, stack: "_Stack", need_input_grad: bool, x, *params):
        run = Run(x.dtype, stack._pack_cache, stack.training, stack._grad_sink, stack._grad_views)
        if stack.training:
            stack._dropout_step += 1
            seed = stack.dropout_seed if stack.dropout_seed is not None else torch.initial_seed()
            # one Philox stream per (stack, optimizer step); sites within the call take consecutive offsets, and the
            # element counter is the element's index in the GLOBAL batch (sample_base): data-parallel ranks and
            # micro-batches draw different masks, and a sample's mask does not depend on the split (train.TrainStep
            # holds _dropout_step fixed over the micro-batches of a step)
            run.dropout_seed = (seed + 0x9E3779B97F4A7C15 * (2 if type(stack).__name__ == "Decoder" else 1)) & 0xFFFFFFFFFFFFFFFF
            run.dropout_offset = stack._dropout_step << 16
            run.sample_base = int(stack._rng_sample_base)
            run.step_state = stack._step_state
            if stack.dropout_mask_override:
                run.dropout_masks = {id(k): v for k, v in stack.dropout_mask_override.items()}
        if stack.tap_sink is not None:
            run.taps = stack.tap_sink
            # a Downsample / Upsample is tapped under its own path (its only child is `.conv`)
            run.tap_names = {id(m): (n[:-5] if n.endswith("sample.conv") else n) for n, m in stack.named_modules()}
        with torch.no_grad():
            y = run_forward(stack.stages(need_input_grad), run, x)
        ctx.run = run
        ctx.params = params
        ctx.need_input_grad = need_input_grad
        if not any(ctx.needs_input_grad):
            run.tape.clear()   # inference: nothing is kept
        return y

    @staticmethod
    def backward(ctx, dy):
        run: Run = ctx.run
        if not run.tape:
            raise RuntimeError("backward through a stack whose tape was released (double backward?)")
        with torch.no_grad():
            dx = run_backward(run, dy.contiguous())
        grads = []
        for p in ctx.params:
            g = run.grads.get(p)
            if g is None and p.requires_grad and run.sink is None:
                g = torch.zeros_like(p)
            grads.append(g)
        run.grads = {}
        ctx.run = None
        return (None, None, dx if ctx.need_input_grad else None, *grads)


class Encoder(_Stack):
    """encoder_decoder.py:212-328."""

    def __init__(self, *, ch: int, out_ch: int, ch_mult: Tuple[int, ...] = (1, 2, 4, 8), num_res_blocks: int,
                 attn_resolutions: List[int], dropout: float = 0.0, resamp_with_conv: bool = True,
                 in_channels: int, resolution: int, z_channels: int, double_z: bool = True,
                 use_linear_attn: bool = False, attn_type: str = "vanilla", **ignore_kwargs):
        super().__init__()
        if use_linear_attn:
            attn_type = "linear"
        self.ch, self.temb_ch = ch, 0
        self.num_resolutions = len(ch_mult)
        self.num_res_blocks = num_res_blocks
        self.resolution = resolution
        self.in_channels = in_channels
        self.conv_in = nn.Conv2d(in_channels, ch, 3, 1, 1)
        self.in_ch_mult = (1,) + tuple(ch_mult)
        curr_res = resolution
        self.down = nn.ModuleList()
        block_in = ch
        for lv in range(self.num_resolutions):
            level = nn.Module()
            level.block, level.attn = nn.ModuleList(), nn.ModuleList()
            block_in, block_out = ch * self.in_ch_mult[lv], ch * ch_mult[lv]
            for _ in range(num_res_blocks):
                level.block.append(ResnetBlock(block_in, block_out, dropout))
                block_in = block_out
                if curr_res in attn_resolutions:
                    level.attn.append(make_attn(block_in, attn_type))
            if lv != self.num_resolutions - 1:
                level.downsample = Downsample(block_in, resamp_with_conv)
                curr_res //= 2
            self.down.append(level)
        self.mid = nn.Module()
        self.mid.block_1 = ResnetBlock(block_in, block_in, dropout)
        self.mid.attn_1 = make_attn(block_in, attn_type)
        self.mid.block_2 = ResnetBlock(block_in, block_in, dropout)
        self.norm_out = _norm(block_in)
        self.conv_out = nn.Conv2d(block_in, 2 * z_channels if double_z else z_channels, 3, 1, 1)
        self._init_runtime()

    def stages(self, need_input_grad: bool):
        st = [PlainConvOp(self.conv_in, L.CONV_S1, need_dx=need_input_grad, pad_in=True)]
        for lv, level in enumerate(self.down):
            for ib, blk in enumerate(level.block):
                st.append(ResOp(blk))
                if len(level.attn) > 0:
                    st.append(attn_stage(level.attn[ib]))
            if hasattr(level, "downsample"):
                st.append(PlainConvOp(level.downsample.conv, L.CONV_DOWN))
        st += [ResOp(self.mid.block_1), attn_stage(self.mid.attn_1), ResOp(self.mid.block_2),
               HeadOp(self.norm_out, self.conv_out)]
        return link_stages(st)

    def in_channels_nhwc(self, dtype) -> int:
        """channel count forward_nhwc() expects (conv_in zero-padded to 64 in bf16 mode)"""
        return ConvOp([self.conv_in], L.CONV_S1, pad_in=True).cin_of(dtype)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        """x NCHW fp32 -> (B, 2*z_channels, R, R) NCHW fp32 (same contract as the reference)."""
        L.require_device()
        xin = _ToNHWC.apply(x.float().contiguous(), self.compute_dtype, self.in_channels_nhwc(self.compute_dtype))
        y = self.forward_nhwc(xin, x.requires_grad)
        return _ToNCHW.apply(y, y.shape[-1])


class Decoder(_Stack):
    """encoder_decoder.py:331-451."""

    def __init__(self, *, ch: int, out_ch: int, ch_mult: Tuple[int, ...] = (1, 2, 4, 8), num_res_blocks: int,
                 attn_resolutions: List[int], dropout: float = 0.0, resamp_with_conv: bool = True,
                 in_channels: int, resolution: int, z_channels: int, give_pre_end: bool = False,
                 tanh_out: bool = False, use_linear_attn: bool = False, attn_type: str = "vanilla", **ignorekwargs):
        super().__init__()
        if use_linear_attn:
            attn_type = "linear"
        if give_pre_end or tanh_out:
            raise NotImplementedError("give_pre_end / tanh_out are never enabled by the reference models")
        self.ch, self.temb_ch = ch, 0
        self.num_resolutions = len(ch_mult)
        self.num_res_blocks = num_res_blocks
        self.resolution = resolution
        self.in_channels = in_channels
        self.give_pre_end, self.tanh_out = give_pre_end, tanh_out
        block_in = ch * ch_mult[self.num_resolutions - 1]
        curr_res = resolution // 2 ** (self.num_resolutions - 1)
        self.z_shape = (1, z_channels, curr_res, curr_res)
        self.conv_in = nn.Conv2d(z_channels, block_in, 3, 1, 1)
        self.mid = nn.Module()
        self.mid.block_1 = ResnetBlock(block_in, block_in, dropout)
        self.mid.attn_1 = make_attn(block_in, attn_type)
        self.mid.block_2 = ResnetBlock(block_in, block_in, dropout)
        self.up = nn.ModuleList()
        for lv in reversed(range(self.num_resolutions)):
            level = nn.Module()
            level.block, level.attn = nn.ModuleList(), nn.ModuleList()
            block_out = ch * ch_mult[lv]
            for _ in range(num_res_blocks + 1):
                level.block.append(ResnetBlock(block_in, block_out, dropout))
                block_in = block_out
                if curr_res in attn_resolutions:
                    level.attn.append(make_attn(block_in, attn_type))
            if lv != 0:
                level.upsample = Upsample(block_in, resamp_with_conv)
                curr_res *= 2
            self.up.insert(0, level)
        self.norm_out = _norm(block_in)
        self.conv_out = nn.Conv2d(block_in, out_ch, 3, 1, 1)
        self._init_runtime()

    def stages(self, need_input_grad: bool):
        st = [PlainConvOp(self.conv_in, L.CONV_S1, need_dx=need_input_grad),
              ResOp(self.mid.block_1), attn_stage(self.mid.attn_1), ResOp(self.mid.block_2)]
        for lv in reversed(range(self.num_resolutions)):
            level = self.up[lv]
            for ib, blk in enumerate(level.block):
                st.append(ResOp(blk))
                if len(level.attn) > 0:
                    st.append(attn_stage(level.attn[ib]))
            if hasattr(level, "upsample"):
                st.append(PlainConvOp(level.upsample.conv, L.CONV_UP))
        st.append(HeadOp(self.norm_out, self.conv_out, pad_out=True))
        return link_stages(st)

    def out_channels_nhwc(self, dtype) -> int:
        """channel count of forward_nhwc()'s output (conv_out zero-padded to 64 in bf16 mode)"""
        return ConvOp([self.conv_out], L.CONV_S1, pad_out=True).cout_of(dtype)

    def forward(self, z: torch.Tensor) -> torch.Tensor:
        """z NCHW fp32 -> reconstruction NCHW fp32."""
        L.require_device()
        zin = _ToNHWC.apply(z.float().contiguous(), self.compute_dtype, z.shape[1])
        y = self.forward_nhwc(zin, z.requires_grad)
        return _ToNCHW.apply(y, self.conv_out.out_channels)


class _ToNHWC(torch.autograd.Function):
    """NCHW fp32 -> NHWC compute dtype, channels zero-padded to c_out"""

    @staticmethod
    def forward(ctx, x_nchw, dtype, c_out):
        ctx.c = x_nchw.shape[1]
        return ops.input_pack_fwd(x_nchw, None, None, None, c_out, dtype)

    @staticmethod
    def backward(ctx, dy):
        return ops.nhwc_to_nchw_f32(dy.contiguous(), 0, ctx.c), None, None


class _ToNCHW(torch.autograd.Function):
    """NHWC -> NCHW fp32, first c_real channels (the rest is tcgen05 tile padding)"""

    @staticmethod
    def forward(ctx, x_nhwc, c_real):
        ctx.dtype = x_nhwc.dtype
        ctx.c_pad = x_nhwc.shape[-1]
        return ops.nhwc_to_nchw_f32(x_nhwc, 0, c_real)

    @staticmethod
    def backward(ctx, dy):
        return ops.input_pack_fwd(dy.float().contiguous(), None, None, None, ctx.c_pad, ctx.dtype), None
