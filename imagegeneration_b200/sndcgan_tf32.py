"""The fp32 / tf32 path of the SNDCGAN models (north_star: "the fp32/tf32 path must match per-layer outputs and gradients
within 1e-3 relative L2"): `SNDCGANEngine(dtype="tf32")`, `bench.py --dtype tf32`.

Same classes, same step schedule (sndcgan.py), different storage and kernels:
  * activations, activation gradients and packed conv weights are fp32 (weights rounded to tf32 while packing);
  * the nine wide convolutions run the per-tap tcgen05 kernel with `kind::tf32` (forward, input gradient, weight gradient);
  * the 3-channel edge layers (D.conv1, G.convT4, all roles) and the discriminator head run EXACT fp32 on CUDA cores
    (b2g_conv3x3_*_f32, b2g_wgrad3x3_c3_c64_f32, b2g_dhead_*_f32) — they are HBM-bound, tensor cores buy nothing there;
  * the generator stem: forward kind::tf32 with fp32 output, backward with the fp32 gradient split into bf16 hi + lo
    (products exact to 2^-17);
  * bias / beta gradients are column sums of the stored fp32 gradients (b2g_colsum_f32) instead of fused epilogue sums.
It is the parity configuration (BASELINE.md config C2'), not the throughput one: steps stay eager (no CUDA graph).
"""
from __future__ import annotations

import ctypes as C

import torch

from . import lib as _lib
from . import ops
from .lib import ACT_LEAKY, ACT_NONE, ACT_RELU, ACT_TANH, CONV_K3S1, CONV_K4S2, CONVT_K4S2, B200GanError, check
from .sndcgan import D_CONVS, INV_STD, LEAKY_ALPHA, Z_SIZE, Discriminator, Generator, _p, _stream

F32 = torch.float32


def _colsum_f32(x: torch.Tensor, out: torch.Tensor) -> None:
    c = out.numel()
    check(_lib.load().b2g_colsum_f32(_p(x), _p(out), x.numel() // c, c, _stream()), "b2g_colsum_f32")


class GeneratorTF32(Generator):
    dtype = "tf32"

    def repack(self):
        if self._pack_plan is None:
            v = self.arena.variables
            plan = ops.PackPlan()
            for i, k in enumerate((3, 6, 9)):
                gamma = v[k + 1]
                self._wp_f[i] = plan.add(CONVT_K4S2, v[k], n_is_d3=False, scale_scalar=self._inv_std, scale_d2=gamma,
                                         tf32=True)
                self._wp_d[i] = plan.add(CONV_K4S2, v[k], n_is_d3=True, scale_scalar=self._inv_std, scale_d2=gamma,
                                         tf32=True)
            self._pack_plan = plan
        self._pack_plan.run()
        self._packed_dirty = False

    def forward(self, z):
        lib = _lib.load()
        if z.dtype != F32 or not z.is_cuda or z.shape[1] != Z_SIZE:
            raise B200GanError("generator input must be a CUDA fp32 tensor [B, 128]")
        z = z.contiguous()
        if self._packed_dirty:
            self.repack()
        B = z.shape[0]
        v = self.arena.variables
        a0 = self._buf("a0", (B, self.h0, self.w0, 512), F32)
        check(lib.b2g_dense_g_fwd_f32(_p(z), _p(v[0]), _p(v[1]), _p(v[2]), INV_STD, _p(a0), B, self.F, _stream()),
              "b2g_dense_g_fwd_f32")
        acts = [a0]
        x = a0
        for i, k in enumerate((3, 6, 9)):
            cout = v[k].shape[2]
            out = self._buf(f"a{i + 1}", (B, x.shape[1] * 2, x.shape[2] * 2, cout), F32)
            ops.conv_igemm(CONVT_K4S2, x, self._wp_f[i], out, bias=v[k + 2], act=ACT_RELU, tf32=True)
            acts.append(out)
            x = out
        H, W, _ = self.output_size
        g_out = self._buf("g_out", (B, H, W, 3), F32)
        check(lib.b2g_conv3x3_c64_to_c3_f32(_p(x), _p(v[12]), None, _p(g_out), B, H, W, ACT_TANH, None, ACT_NONE,
                                            _stream()), "b2g_conv3x3_c64_to_c3_f32")
        self._acts = acts
        self._z = z
        self.g_out = g_out
        return g_out

    def backward(self, d_pre_tanh):
        lib = _lib.load()
        v, g = self.arena.variables, self.arena.grads
        a0, a1, a2, a3 = self._acts
        B, H, W, _ = d_pre_tanh.shape
        self.zero_grads()
        d_pre_tanh = d_pre_tanh.contiguous()
        check(lib.b2g_wgrad3x3_c3_c64_f32(_p(d_pre_tanh), _p(a3), _p(g[12]), None, B, H, W, _stream()),
              "b2g_wgrad3x3_c3_c64_f32")
        dq = self._buf("dq3", tuple(a3.shape), F32)
        check(lib.b2g_conv3x3_c3_to_c64_f32(_p(d_pre_tanh), _p(v[12]), None, None, _p(dq), B, H, W, ACT_NONE, 0.0, 0, 0,
                                            _p(a3), ACT_RELU, _stream()), "b2g_conv3x3_c3_to_c64_f32")
        _colsum_f32(dq, g[11])
        acts_in = [a0, a1, a2]
        for i in (2, 1, 0):
            k = (3, 6, 9)[i]
            x_in = acts_in[i]
            cout, cin = v[k].shape[2], v[k].shape[3]
            ops.conv_wgrad(CONV_K4S2, dq, x_in, g[k], tf32=True)
            check(lib.b2g_bn_fold_finalize(_p(v[k]), _p(g[k]), _p(v[k + 1]), INV_STD, _p(g[k + 1]), 16, cout, cin,
                                           _stream()), "b2g_bn_fold_finalize")
            dprev = self._buf(f"dq{i}", tuple(x_in.shape), F32)
            ops.conv_igemm(CONV_K4S2, dq, self._wp_d[i], dprev, yprev=x_in, dact=ACT_RELU, tf32=True)
            if i > 0:
                _colsum_f32(dprev, g[(3, 6, 9)[i - 1] + 2])
            dq = dprev
        scratch = self._buf("stem_hilo", (2, B, self.F), torch.bfloat16)
        check(lib.b2g_dense_g_bwd_f32(_p(self._z), _p(v[0]), _p(v[1]), INV_STD, _p(dq), _p(scratch), _p(g[0]), _p(g[1]),
                                      _p(g[2]), B, self.F, _stream()), "b2g_dense_g_bwd_f32", 1 + 2 * ((B + 63) // 64))


class DiscriminatorTF32(Discriminator):
    dtype = "tf32"

    def repack(self):
        if self._pack_plan is None:
            v = self.arena.variables
            plan = ops.PackPlan()
            for i in range(1, 7):
                _name, k, s, _ci, _co = D_CONVS[i]
                sc = self._inv_sigma(i)
                if s == 1:
                    self._wp_f[i] = plan.add(CONV_K3S1, v[2 * i], n_is_d3=True, scale_scalar=sc, tf32=True)
                    self._wp_d[i] = plan.add(CONV_K3S1, v[2 * i], n_is_d3=False, flip=True, scale_scalar=sc, tf32=True)
                else:
                    self._wp_f[i] = plan.add(CONV_K4S2, v[2 * i], n_is_d3=True, scale_scalar=sc, tf32=True)
                    self._wp_d[i] = plan.add(CONVT_K4S2, v[2 * i], n_is_d3=False, scale_scalar=sc, tf32=True)
            self._pack_plan = plan
        self._pack_plan.run()
        self._packed_dirty = False

    def forward(self, x, training=False, dropout_seeds=None, x27=None):
        lib = _lib.load()
        if x.dtype != F32 or not x.is_cuda or tuple(x.shape[1:]) != self.input_size:
            raise B200GanError(f"discriminator input must be a CUDA fp32 tensor [B,{self.input_size}]")
        x = x.contiguous()
        if self.spectral_norm and self._sn_stale:
            self.sn_power_iteration()
        if self._packed_dirty:
            self.repack()
        B, H, W, _ = x.shape
        v = self.arena.variables
        t8 = self.thresh8 if training else 0
        if t8 and dropout_seeds is None:
            raise B200GanError("training-mode forward needs dropout_seeds")
        seeds = dropout_seeds if t8 else [0] * 7
        if any(isinstance(s, torch.Tensor) for s in seeds):
            raise B200GanError("the fp32/tf32 path takes host dropout seeds (it does not replay CUDA graphs)")
        c1 = self._buf("c1", (B, H, W, 64), F32)
        check(lib.b2g_conv3x3_c3_to_c64_f32(_p(x), _p(v[0]), _p(self._inv_sigma(0)), _p(v[1]), _p(c1), B, H, W, ACT_LEAKY,
                                            LEAKY_ALPHA, t8, C.c_ulonglong(seeds[0] & 0xFFFFFFFFFFFFFFFF), None, ACT_NONE,
                                            _stream()), "b2g_conv3x3_c3_to_c64_f32")
        self._x_in, self._x27 = x, None
        acts = [c1]
        cur = c1
        for i in range(1, 7):
            _name, k, s, _ci, cout = D_CONVS[i]
            out = self._buf(f"c{i + 1}", (B, cur.shape[1] // s, cur.shape[2] // s, cout), F32)
            ops.conv_igemm(CONV_K3S1 if s == 1 else CONV_K4S2, cur, self._wp_f[i], out, bias=v[2 * i + 1],
                           act=ACT_LEAKY, alpha=LEAKY_ALPHA, dropout_thresh8=t8, dropout_seed=seeds[i], tf32=True)
            acts.append(out)
            cur = out
        self._x = x
        self._acts = acts
        self._train_fwd = bool(t8)
        self._B = B
        self.logits = self._buf("logits", (B, 1), F32)
        self._dlogit = self._buf("dlogit", (B,), F32)
        self._loss = torch.empty(2, dtype=F32, device=self.device)
        self._ws = self._buf("head_ws", (64 * B,), F32)
        return self

    def head_loss(self, loss_kind: int, label: float, inv_batch: float):
        v = self.arena.variables
        check(_lib.load().b2g_dhead_fwd_loss_f32(_p(self._acts[6]), _p(v[14]), _p(self._inv_sigma(7)), _p(v[15]),
                                                 self._B, self.F, loss_kind, label, inv_batch, _p(self._ws),
                                                 self._ws.numel(), _p(self.logits), _p(self._dlogit), _p(self._loss),
                                                 _stream()), "b2g_dhead_fwd_loss_f32", 2)
        return self._loss[0], self.logits

    def backward(self, need_wgrad: bool, need_input_grad: bool, g_out=None, on_tail_grads_ready=None):
        lib = _lib.load()
        v, g = self.arena.variables, self.arena.grads
        acts = self._acts
        B = self._B
        dscale = self.keep_scale if self._train_fwd else 1.0
        dzero = 0.0 if self._train_fwd else LEAKY_ALPHA
        if need_wgrad:
            self.arena.g.zero_()
            self._sn_grads_final = False
        c7 = acts[6]
        dpre = self._buf("dpre7", tuple(c7.shape), F32)
        check(lib.b2g_dhead_bwd_f32(_p(c7), _p(v[14]), _p(self._inv_sigma(7)), _p(self._dlogit), B, self.F, LEAKY_ALPHA,
                                    dscale, dzero, _p(dpre), _p(g[14]) if need_wgrad else None,
                                    _p(g[15]) if need_wgrad else None, _stream()), "b2g_dhead_bwd_f32")
        for i in range(6, 0, -1):
            _name, k, s, _ci, cout = D_CONVS[i]
            x_in = acts[i - 1]
            if need_wgrad:
                _colsum_f32(dpre, g[2 * i + 1])
                ops.conv_wgrad(CONV_K3S1 if s == 1 else CONV_K4S2, x_in, dpre, g[2 * i], tf32=True)
                if i == 5 and on_tail_grads_ready is not None and not self.spectral_norm:
                    on_tail_grads_ready()
            dprev = self._buf(f"dpre{i}", tuple(x_in.shape), F32)
            ops.conv_igemm(CONV_K3S1 if s == 1 else CONVT_K4S2, dpre, self._wp_d[i], dprev, yprev=x_in,
                           dact=ACT_LEAKY, dact_alpha=LEAKY_ALPHA, dact_scale=dscale, dact_zero=dzero, tf32=True)
            dpre = dprev
        H, W = self.input_size[0], self.input_size[1]
        if need_wgrad:
            check(lib.b2g_wgrad3x3_c3_c64_f32(_p(self._x_in), _p(dpre), _p(g[0]), _p(g[1]), B, H, W, _stream()),
                  "b2g_wgrad3x3_c3_c64_f32")
        d_in = None
        if need_input_grad:
            d_in = self._buf("d_in", (B, H, W, 3), F32)
            check(lib.b2g_conv3x3_c64_to_c3_f32(_p(dpre), _p(v[0]), _p(self._inv_sigma(0)), _p(d_in), B, H, W, ACT_NONE,
                                                _p(g_out), ACT_TANH if g_out is not None else ACT_NONE, _stream()),
                  "b2g_conv3x3_c64_to_c3_f32")
        return d_in
