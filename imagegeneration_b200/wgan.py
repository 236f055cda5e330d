"""Host-side mirror of the reference's WGAN API (wasserstein_gan/WGAN.py) driving the sm_100a kernels —
SURVEY.md section 8(f) row N1.

Same names and call shapes as the reference:
  define_critic(in_shape)                      WGAN.py:53-101
  define_generator(latent_dim, output_size)    WGAN.py:105-134
  define_gan(generator, critic)                WGAN.py:138-152
  WGAN(...).train loop                         WGAN.py:279-326  ->  WGANEngine.train_iteration / WGAN.train
The critic / generator convolutions run on the same tcgen05 implicit-GEMM kernels as the SNDCGAN path
(b2g_conv_igemm / b2g_conv_wgrad, raw outputs with the bias fused); what is new here is training-mode
BatchNormalization (b2g_bn_*: batch statistics, moving statistics, fused LeakyReLU, backward), RMSprop with the
ClipConstraint fused in (b2g_rmsprop) and the Wasserstein loss in the critic head (b2g_dhead_fwd_loss kind 3).
There is no CPU fallback.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import lib as _lib
from . import ops
from .lib import (ACT_LEAKY, ACT_NONE, ACT_TANH, CONV_K1S1, CONV_K3S1, CONV_K3S1_C3, CONV_K4S2, CONVT_K4S2,
                  B200GanError, check)
from .sndcgan import ParamArena, _glorot_uniform, _p, _stream

BN_EPS = 1e-3          # Keras BatchNormalization defaults (WGAN.py:62 ...)
BN_MOMENTUM = 0.99
CLIP_VALUE = 0.01      # ClipConstraint(0.01), WGAN.py:57
LATENT_DIM = 128       # WGAN.py:174
LOSS_WASSERSTEIN = 3
GEN_ALPHA = 0.2        # LeakyReLU(alpha=0.2) in the generator, WGAN.py:114-128
# k, stride, cin, cout, LeakyReLU alpha  (WGAN.py:60-93)
CRITIC_CONVS = [(3, 1, 3, 64, 0.2), (4, 2, 64, 128, 0.2), (3, 1, 128, 128, 0.2), (4, 2, 128, 256, 0.2),
                (3, 1, 256, 256, 0.2), (4, 2, 256, 512, 0.1), (3, 1, 512, 512, 0.1)]


def _off_ptr(t: torch.Tensor, off: int):
    return C.c_void_p(t.data_ptr() + 4 * off)


class RMSprop:
    """tf.keras.optimizers.RMSprop(learning_rate) (WGAN.py:99,150): rho 0.9, momentum 0, epsilon 1e-7."""

    def __init__(self, learning_rate=0.001, rho=0.9, epsilon=1e-7):
        self.learning_rate, self.rho, self.epsilon = learning_rate, rho, epsilon
        self.iterations = 0

    def apply(self, w, g, rms, off, n, gscale=1.0, clip_n=0, clip=0.0, count=True):
        if count:
            self.iterations += 1
        check(_lib.load().b2g_rmsprop(_off_ptr(w, off), _off_ptr(g, off), _off_ptr(rms, 0), n, self.learning_rate,
                                      self.rho, self.epsilon, gscale, clip_n, clip, _stream()), "b2g_rmsprop")


class _BatchNormSet:
    """Statistics buffers of the BatchNormalization layers of one model: per layer stat = [mean | invstd] of the
    last forward, moving mean / variance (non-trainable), and the shared reduction scratch."""

    def __init__(self, channels, device):
        self.channels = list(channels)
        self.stat = [torch.zeros(2 * c, dtype=torch.float32, device=device) for c in channels]
        self.moving_mean = [torch.zeros(c, dtype=torch.float32, device=device) for c in channels]
        self.moving_var = [torch.ones(c, dtype=torch.float32, device=device) for c in channels]
        self.scratch = torch.zeros(16 * max(channels), dtype=torch.float32, device=device)   # [8][2C] partial sums

    def stat_targets(self, i):
        """(colsum, colsumsq) views for the conv in front of BatchNorm i: copy 0 of the [8][2C] partial sums, zeroed
        here — the conv epilogue accumulates the batch statistics while it stores its output."""
        c = self.channels[i]
        self.scratch[:16 * c].zero_()
        return self.scratch[:c], self.scratch[c:2 * c]

    def forward(self, i, raw, gamma, beta, alpha, out, training, bias=None, fused_stats=False):
        """bias: the conv ran without its bias (cancelled by the mean subtraction); the moving mean still tracks it.
        fused_stats: the sums are already in scratch (stat_targets + conv epilogue) — no reduction pass over raw."""
        lib = _lib.load()
        c = self.channels[i]
        p = raw.numel() // c
        if training:
            if not fused_stats:
                check(lib.b2g_bn_stats(_p(raw), _p(self.scratch), p, c, _stream()), "b2g_bn_stats")
            check(lib.b2g_bn_finalize(_p(self.scratch), p, c, BN_EPS, BN_MOMENTUM, _p(self.stat[i]),
                                      _p(self.moving_mean[i]), _p(self.moving_var[i]), _p(bias), _stream()),
                  "b2g_bn_finalize")
        else:
            check(lib.b2g_bn_inference_stat(_p(self.moving_mean[i]), _p(self.moving_var[i]), c, BN_EPS,
                                            _p(self.stat[i]), _p(bias), _stream()), "b2g_bn_inference_stat")
        check(lib.b2g_bn_apply_act(_p(raw), _p(self.stat[i]), _p(gamma), _p(beta), alpha, _p(out), p, c, _stream()),
              "b2g_bn_apply_act")
        return out

    def backward(self, i, g, raw, gamma, dx, dgamma, dbeta):
        c = self.channels[i]
        p = raw.numel() // c
        check(_lib.load().b2g_bn_bwd(_p(g), _p(raw), _p(self.stat[i]), _p(gamma), _p(self.scratch), _p(dx),
                                     _p(dgamma), _p(dbeta), p, c, _stream()), "b2g_bn_bwd", 2)
        return dx


# ----------------------------------------------------------------------------------------------------------
# Critic  (WGAN.py:53-101)
# ----------------------------------------------------------------------------------------------------------
class Critic:
    name = "critic"

    def __init__(self, in_shape, device="cuda", seed=63):
        self.input_size = tuple(in_shape)
        h, w, c = self.input_size
        if c != 3 or h % 8 or w % 8:
            raise B200GanError("critic needs an (H, W, 3) input with H, W divisible by 8")
        self.device = torch.device(device)
        self.F = 512 * (h // 8) * (w // 8)
        # storage order: the 7 clipped conv kernels first (one clamp range for b2g_rmsprop), then biases + Dense,
        # then the BatchNorm gamma/beta (the range gan_model's optimizer also updates, WGAN.py:140-142)
        shapes = [(f"conv{i + 1}/kernel", (k, k, ci, co)) for i, (k, _s, ci, co, _a) in enumerate(CRITIC_CONVS)]
        shapes += [(f"conv{i + 1}/bias", (co,)) for i, (_k, _s, _ci, co, _a) in enumerate(CRITIC_CONVS)]
        shapes += [("dense/kernel", (self.F, 1)), ("dense/bias", (1,))]
        for i, (_k, _s, _ci, co, _a) in enumerate(CRITIC_CONVS):
            shapes += [(f"bn{i + 1}/gamma", (co,)), (f"bn{i + 1}/beta", (co,))]
        self.arena = ParamArena(shapes, self.device)
        self.clip_n = self.arena.offsets[7]
        self.bn_off = self.arena.offsets[16]
        self.bn_n = self.arena.numel - self.bn_off
        # Keras order of critic.trainable_variables: conv/kernel, conv/bias, bn/gamma, bn/beta per block, then Dense
        self._keras_idx = []
        for i in range(7):
            self._keras_idx += [i, 7 + i, 16 + 2 * i, 17 + 2 * i]
        self._keras_idx += [14, 15]
        gen = torch.Generator().manual_seed(seed)
        v = self.arena.variables
        for i in range(7):   # RandomNormal(stddev=0.02), WGAN.py:55
            v[i].copy_((torch.randn(tuple(v[i].shape), generator=gen, dtype=torch.float64) * 0.02).to(torch.float32))
            v[16 + 2 * i].fill_(1.0)
        v[14].copy_(_glorot_uniform(tuple(v[14].shape), gen))   # Dense keeps glorot_uniform
        self.bn = _BatchNormSet([c[3] for c in CRITIC_CONVS], self.device)
        self._pack_plan = None
        self._wp_f = [None] * 7
        self._wp_d = [None] * 7
        self._packed_dirty = True
        self._bufs = {}

    @property
    def trainable_variables(self):
        v = self.arena.variables
        return [v[i] for i in self._keras_idx]

    @property
    def trainable_gradients(self):
        g = self.arena.grads
        return [g[i] for i in self._keras_idx]

    def set_weights(self, tensors):
        for v, t in zip(self.trainable_variables, tensors):
            v.copy_(t.to(self.device, torch.float32))
        self._packed_dirty = True

    def _buf(self, key, shape, dtype):
        t = self._bufs.get(key)
        if t is None or tuple(t.shape) != tuple(shape) or t.dtype != dtype:
            t = torch.empty(shape, dtype=dtype, device=self.device)
            self._bufs[key] = t
        return t

    def repack(self):
        if self._pack_plan is None:
            v = self.arena.variables
            plan = ops.PackPlan()
            self._wp_f[0] = plan.add(CONV_K1S1, v[0].view(1, 1, 27, 64), n_is_d3=True, ck_pad=64)
            self._wp_d[0] = plan.add(CONV_K3S1, v[0], n_is_d3=False, flip=True)
            for i in range(1, 7):
                s = CRITIC_CONVS[i][1]
                if s == 1:
                    self._wp_f[i] = plan.add(CONV_K3S1, v[i], n_is_d3=True)
                    self._wp_d[i] = plan.add(CONV_K3S1, v[i], n_is_d3=False, flip=True)
                else:
                    self._wp_f[i] = plan.add(CONV_K4S2, v[i], n_is_d3=True)
                    self._wp_d[i] = plan.add(CONVT_K4S2, v[i], n_is_d3=False)
            self._pack_plan = plan
        self._pack_plan.run()
        self._packed_dirty = False

    def forward(self, x, training: bool):
        """x: fp32 CUDA [B,H,W,3].  training=True: batch statistics (and the moving statistics move)."""
        if x.dtype != torch.float32 or not x.is_cuda or tuple(x.shape[1:]) != self.input_size:
            raise B200GanError(f"critic input must be a CUDA fp32 tensor [B,{self.input_size}]")
        x = x.contiguous()
        if self._packed_dirty:
            self.repack()
        B, H, W, _ = x.shape
        v = self.arena.variables
        raws, acts = [], []
        cur = x
        for i, (k, s, _ci, co, alpha) in enumerate(CRITIC_CONVS):
            oh, ow = cur.shape[1] // s, cur.shape[2] // s
            raw = self._buf(f"raw{i + 1}", (B, oh, ow, co), torch.bfloat16)
            kind = CONV_K3S1_C3 if i == 0 else (CONV_K3S1 if s == 1 else CONV_K4S2)
            # the conv bias is NOT added: BatchNorm subtracts the channel mean, so normalised outputs and every
            # gradient are identical without it (only the moving mean sees it: bn_finalize adds it back).  That
            # leaves a bias-free epilogue, which in training mode also accumulates sum x / sum x^2 per channel.
            if training:
                cs, cq = self.bn.stat_targets(i)
                ops.conv_igemm(kind, cur, self._wp_f[i], raw, act=ACT_NONE, colsum=cs, colsumsq=cq)
            else:
                ops.conv_igemm(kind, cur, self._wp_f[i], raw, act=ACT_NONE)
            y = self._buf(f"c{i + 1}", (B, oh, ow, co), torch.bfloat16)
            self.bn.forward(i, raw, v[16 + 2 * i], v[17 + 2 * i], alpha, y, training, bias=v[7 + i],
                            fused_stats=training)
            raws.append(raw)
            acts.append(y)
            cur = y
        self._x_in, self._raws, self._acts, self._B = x, raws, acts, B
        self.logits = self._buf("logits", (B, 1), torch.float32)
        self._dlogit = self._buf("dlogit", (B,), torch.float32)
        self._loss = torch.empty(2, dtype=torch.float32, device=self.device)
        self._ws = self._buf("head_ws", (64 * B,), torch.float32)
        return self

    def head_loss(self, label: float):
        """Dense(1) + wasserstein_loss(label, logits) = mean(label * logits) (WGAN.py:48-49,96)."""
        v = self.arena.variables
        check(_lib.load().b2g_dhead_fwd_loss(_p(self._acts[6]), _p(v[14]), None, _p(v[15]), self._B, self.F,
                                             LOSS_WASSERSTEIN, label, 1.0 / self._B, _p(self._ws), self._ws.numel(),
                                             _p(self.logits), _p(self._dlogit), _p(self._loss), _stream()),
              "b2g_dhead_fwd_loss", 2)
        return self._loss[0], self.logits

    def __call__(self, x, training=False):
        self.forward(x, training)
        _l, logits = self.head_loss(1.0)
        return logits.clone()

    def backward(self, need_wgrad: bool, need_input_grad: bool, g_out=None):
        """Backward from the dlogit of the last head_loss.  BatchNorm gamma/beta gradients are always produced;
        need_wgrad adds the conv / Dense kernel and bias gradients (critic_model.train_on_batch); need_input_grad
        returns dL/d(pre-tanh generator output) (gan_model.train_on_batch, conv/Dense frozen: WGAN.py:140-142)."""
        lib = _lib.load()
        v, g = self.arena.variables, self.arena.grads
        acts, raws, B = self._acts, self._raws, self._B
        self.arena.g.zero_()
        a7 = CRITIC_CONVS[6][4]
        dcur = self._buf("d7", tuple(acts[6].shape), torch.bfloat16)
        check(lib.b2g_dhead_bwd(_p(acts[6]), _p(v[14]), None, _p(self._dlogit), B, self.F, a7, 1.0, a7, _p(dcur),
                                _p(g[14]) if need_wgrad else None, _p(g[15]) if need_wgrad else None, _stream()),
              "b2g_dhead_bwd")
        for i in range(6, -1, -1):
            k, s, _ci, co, _alpha = CRITIC_CONVS[i]
            # dcur = dL/d(BN output of layer i) -> dL/d(raw conv output), in place; gamma / beta gradients
            self.bn.backward(i, dcur, raws[i], v[16 + 2 * i], dcur, g[16 + 2 * i], g[17 + 2 * i])
            if need_wgrad:
                # conv bias gradient = sum over pixels of dL/d(raw) = 0 identically under BatchNorm (the mean is
                # subtracted).  The reference's autodiff value is fp32 rounding noise (~1e-9, far below RMSprop's
                # epsilon 1e-7, so its biases stay put); g[7 + i] keeps the exact zero written by arena.g.zero_().
                if i == 0:
                    ops.conv_wgrad(CONV_K3S1_C3, self._x_in, dcur, g[0])
                else:
                    ops.conv_wgrad(CONV_K3S1 if s == 1 else CONV_K4S2, acts[i - 1], dcur, g[i])
            if i == 0:
                break
            a_prev = CRITIC_CONVS[i - 1][4]
            dprev = self._buf(f"d{i}", tuple(acts[i - 1].shape), torch.bfloat16)
            ops.conv_igemm(CONV_K3S1 if s == 1 else CONVT_K4S2, dcur, self._wp_d[i], dprev, yprev=acts[i - 1],
                           dact=ACT_LEAKY, dact_alpha=a_prev, dact_scale=1.0, dact_zero=a_prev)
            dcur = dprev
        d_in = None
        if need_input_grad:
            H, W = self.input_size[0], self.input_size[1]
            d_in = self._buf("d_in", (B, H, W, 3), torch.float32)
            ops.conv_igemm(CONV_K3S1, dcur, self._wp_d[0], d_in, yprev=g_out,
                           dact=ACT_TANH if g_out is not None else ACT_NONE, out_f32=True)
        return d_in


# ----------------------------------------------------------------------------------------------------------
# Generator  (WGAN.py:105-134)
# ----------------------------------------------------------------------------------------------------------
class WGenerator:
    name = "generator"

    def __init__(self, latent_dim, output_size, device="cuda", seed=62):
        if latent_dim != LATENT_DIM:
            raise B200GanError("latent_dim must be 128 (WGAN.py:174)")
        self.output_size = tuple(output_size)
        h, w, c = self.output_size
        if c != 3 or h % 8 or w % 8:
            raise B200GanError("generator needs an (H, W, 3) output with H, W divisible by 8")
        self.h0, self.w0 = h // 8, w // 8
        self.F = 512 * self.h0 * self.w0
        self.device = torch.device(device)
        shapes = [("dense/kernel", (LATENT_DIM, self.F)),
                  ("convt1/kernel", (4, 4, 256, 512)), ("bn1/gamma", (256,)), ("bn1/beta", (256,)),
                  ("convt2/kernel", (4, 4, 128, 256)), ("bn2/gamma", (128,)), ("bn2/beta", (128,)),
                  ("convt3/kernel", (4, 4, 64, 128)), ("bn3/gamma", (64,)), ("bn3/beta", (64,)),
                  ("conv4/kernel", (3, 3, 64, 3))]
        self.arena = ParamArena(shapes, self.device)
        gen = torch.Generator().manual_seed(seed)
        for name, v in zip(self.arena.names, self.arena.variables):
            if name == "dense/kernel":
                v.copy_(_glorot_uniform(tuple(v.shape), gen))
            elif name.endswith("kernel"):
                v.copy_((torch.randn(tuple(v.shape), generator=gen, dtype=torch.float64) * 0.02).to(torch.float32))
            elif name.endswith("gamma"):
                v.fill_(1.0)
        self.bn = _BatchNormSet([256, 128, 64], self.device)
        self._ones = torch.ones(self.F, dtype=torch.float32, device=self.device)
        self._zeros = torch.zeros(self.F, dtype=torch.float32, device=self.device)
        self._stem_scratch = torch.empty(2 * self.F, dtype=torch.float32, device=self.device)
        self._w4t = torch.empty((3, 3, 3, 64), dtype=torch.float32, device=self.device)   # conv4 as [ky',kx',co,ci]
        self._dw4t = torch.empty((3, 3, 3, 64), dtype=torch.float32, device=self.device)
        self._pack_plan = None
        self._wp_f = [None] * 3
        self._wp_d = [None] * 3
        self._packed_dirty = True
        self._bufs = {}

    @property
    def trainable_variables(self):
        return self.arena.variables

    def set_weights(self, tensors):
        for v, t in zip(self.arena.variables, tensors):
            v.copy_(t.to(self.device, torch.float32))
        self._packed_dirty = True

    def _buf(self, key, shape, dtype):
        t = self._bufs.get(key)
        if t is None or tuple(t.shape) != tuple(shape) or t.dtype != dtype:
            t = torch.empty(shape, dtype=dtype, device=self.device)
            self._bufs[key] = t
        return t

    def repack(self):
        v = self.arena.variables
        # the 64 -> 3 Conv2D's input gradient runs as a 3 -> 64 edge conv over dY with the taps flipped and the
        # channel axes swapped (1 728 weights: a device-side index shuffle, no arithmetic)
        self._w4t.copy_(v[10].flip(0, 1).permute(0, 1, 3, 2))
        if self._pack_plan is None:
            plan = ops.PackPlan()
            for i, k in enumerate((1, 4, 7)):
                self._wp_f[i] = plan.add(CONVT_K4S2, v[k], n_is_d3=False)
                self._wp_d[i] = plan.add(CONV_K4S2, v[k], n_is_d3=True)
            self._wp4_f = plan.add(CONV_K3S1, v[10], n_is_d3=True)
            self._wp4_d = plan.add(CONV_K1S1, self._w4t.view(1, 1, 27, 64), n_is_d3=True, ck_pad=64)
            self._pack_plan = plan
        self._pack_plan.run()
        self._packed_dirty = False

    def forward(self, z, training: bool):
        """training=True: batch statistics (gan_model.train_on_batch, WGAN.py:317); False: moving statistics
        (generator_model.predict, WGAN.py:224)."""
        lib = _lib.load()
        if z.dtype != torch.float32 or not z.is_cuda or z.shape[1] != LATENT_DIM:
            raise B200GanError("generator input must be a CUDA fp32 tensor [B, 128]")
        z = z.contiguous()
        if self._packed_dirty:
            self.repack()
        B = z.shape[0]
        v = self.arena.variables
        a0 = self._buf("a0", (B, self.h0, self.w0, 512), torch.bfloat16)
        check(lib.b2g_dense_g_fwd_leaky(_p(z), _p(v[0]), _p(self._ones), _p(self._zeros), 1.0, GEN_ALPHA, _p(a0), B,
                                        self.F, _stream()), "b2g_dense_g_fwd_leaky")
        acts, raws = [a0], []
        x = a0
        for i, k in enumerate((1, 4, 7)):
            cout = v[k].shape[2]
            shape = (B, x.shape[1] * 2, x.shape[2] * 2, cout)
            raw = self._buf(f"graw{i + 1}", shape, torch.bfloat16)
            if training:
                cs, cq = self.bn.stat_targets(i)
                ops.conv_igemm(CONVT_K4S2, x, self._wp_f[i], raw, act=ACT_NONE, colsum=cs, colsumsq=cq)
            else:
                ops.conv_igemm(CONVT_K4S2, x, self._wp_f[i], raw, act=ACT_NONE)
            y = self._buf(f"a{i + 1}", shape, torch.bfloat16)
            self.bn.forward(i, raw, v[k + 1], v[k + 2], GEN_ALPHA, y, training, fused_stats=training)
            raws.append(raw)
            acts.append(y)
            x = y
        H, W, _ = self.output_size
        g_out = self._buf("g_out", (B, H, W, 3), torch.float32)
        ops.conv_igemm(CONV_K3S1, x, self._wp4_f, g_out, act=ACT_TANH, out_f32=True)
        self._acts, self._raws, self._z, self.g_out = acts, raws, z, g_out
        return g_out

    def __call__(self, z, training=False):
        return self.forward(z, training)

    def predict(self, z):
        return self.forward(z, False)

    def backward(self, d_pre_tanh):
        """d_pre_tanh: fp32 [B,H,W,3] = dL/d(conv4 output before tanh).  Fills arena.g."""
        lib = _lib.load()
        v, g = self.arena.variables, self.arena.grads
        a0, a1, a2, a3 = self._acts
        B = d_pre_tanh.shape[0]
        self.arena.g[self.arena.offsets[1]:].zero_()
        self._dw4t.zero_()
        ops.conv_wgrad(CONV_K3S1_C3, d_pre_tanh, a3, self._dw4t)
        g[10].copy_(self._dw4t.flip(0, 1).permute(0, 1, 3, 2))
        dcur = self._buf("dq3", tuple(a3.shape), torch.bfloat16)
        ops.conv_igemm(CONV_K3S1_C3, d_pre_tanh, self._wp4_d, dcur, yprev=a3, dact=ACT_LEAKY, dact_alpha=GEN_ALPHA,
                       dact_scale=1.0, dact_zero=GEN_ALPHA)
        acts_in = [a0, a1, a2]
        for i in (2, 1, 0):
            k = (1, 4, 7)[i]
            self.bn.backward(i, dcur, self._raws[i], v[k + 1], dcur, g[k + 1], g[k + 2])
            x_in = acts_in[i]
            ops.conv_wgrad(CONV_K4S2, dcur, x_in, g[k])
            dprev = self._buf(f"dq{i}", tuple(x_in.shape), torch.bfloat16)
            ops.conv_igemm(CONV_K4S2, dcur, self._wp_d[i], dprev, yprev=x_in, dact=ACT_LEAKY, dact_alpha=GEN_ALPHA,
                           dact_scale=1.0, dact_zero=GEN_ALPHA)
            dcur = dprev
        sc = self._stem_scratch
        check(lib.b2g_dense_g_bwd(_p(self._z), _p(v[0]), _p(self._ones), 1.0, _p(dcur), _p(g[0]), _off_ptr(sc, 0),
                                  _off_ptr(sc, self.F), B, self.F, _stream()), "b2g_dense_g_bwd", (B + 31) // 32)


def define_critic(in_shape=(144, 256, 3), device="cuda", seed=63) -> Critic:
    """wasserstein_gan/WGAN.py:53."""
    return Critic(in_shape, device=device, seed=seed)


def define_generator(latent_dim=LATENT_DIM, output_size=(144, 256, 3), device="cuda", seed=62) -> WGenerator:
    """wasserstein_gan/WGAN.py:105."""
    return WGenerator(latent_dim, output_size, device=device, seed=seed)


class GanModel:
    """define_gan (WGAN.py:138-152): generator followed by the critic whose conv / Dense layers are frozen; its
    own RMSprop holds slots for the generator's variables and for the critic's BatchNorm gamma / beta."""

    def __init__(self, generator: WGenerator, critic: Critic, learning_rate=5e-5):
        self.generator, self.critic = generator, critic
        self.optimizer = RMSprop(learning_rate)
        self.rms_cbn = torch.zeros(critic.bn_n, dtype=torch.float32, device=critic.device)


def define_gan(generator: WGenerator, critic: Critic) -> GanModel:
    return GanModel(generator, critic)


# ----------------------------------------------------------------------------------------------------------
# WGAN: state + the three train_on_batch flavours of WGAN.train (WGAN.py:279-326)
# ----------------------------------------------------------------------------------------------------------
class WGANEngine:
    def __init__(self, image_size=(144, 256, 3), bach_size=64, critic_learn_iterations=5, device="cuda", seed=62,
                 learning_rate=5e-5, process_group=None):
        _lib.load()
        self.image_size = tuple(image_size)
        self.bach_size = bach_size                      # the reference's spelling, WGAN.py:172
        self.critic_learn_iterations = critic_learn_iterations
        self.latent_dim = LATENT_DIM
        self.device = torch.device(device)
        self.generator_model = define_generator(self.latent_dim, image_size, device=device, seed=seed)
        self.critic_model = define_critic(image_size, device=device, seed=seed + 1)
        self.critic_optimizer = RMSprop(learning_rate)                      # WGAN.py:99
        self.gan_model = define_gan(self.generator_model, self.critic_model)
        self.gan_model.optimizer.learning_rate = learning_rate              # WGAN.py:150
        self.pg = process_group
        self.world, self.rank = 1, 0
        if process_group is not None:
            import torch.distributed as dist
            self.world = dist.get_world_size(process_group)
            self.rank = dist.get_rank(process_group)
            dist.broadcast(self.generator_model.arena.w, src=0, group=process_group)
            dist.broadcast(self.critic_model.arena.w, src=0, group=process_group)

    def _allreduce(self, flat):
        if self.world > 1:
            from .dp import allreduce_sum_
            allreduce_sum_(flat, self.pg, False)

    def average_bn_statistics(self):
        """Data parallel keeps BatchNorm's batch statistics per GPU (north_star item 4), so the MOVING mean / variance of
        the replicas drift apart while the weights stay identical.  Before anything that reads the moving statistics is
        shared between ranks — generator.predict for saved samples, a checkpoint, a model export — this averages them
        across the ranks (moving statistics are running averages of per-batch values: the mean over ranks is the
        running average over the global batches' per-shard statistics).  No-op on one GPU."""
        if self.world == 1:
            return
        import torch.distributed as dist
        for model in (self.generator_model, self.critic_model):
            for t in model.bn.moving_mean + model.bn.moving_var:
                dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.pg)
                t.mul_(1.0 / self.world)

    def train_critic_on_batch(self, x, label: float):
        """critic_model.train_on_batch(x, label*ones) (WGAN.py:298,303): returns the loss (CUDA scalar)."""
        Cm = self.critic_model
        Cm.forward(x, training=True)
        loss, _ = Cm.head_loss(label)
        Cm.backward(need_wgrad=True, need_input_grad=False)
        self._allreduce(Cm.arena.g)
        a = Cm.arena
        self.critic_optimizer.apply(a.w, a.g, a.m, 0, a.numel, 1.0 / self.world, Cm.clip_n, CLIP_VALUE)
        Cm._packed_dirty = True
        return loss

    def generate_fake_samples(self, z):
        """generator_model.predict(z) (WGAN.py:219-227): BatchNorm on the moving statistics."""
        return self.generator_model.predict(z)

    def train_gan_on_batch(self, z):
        """gan_model.train_on_batch(z, -ones) (WGAN.py:317)."""
        G, Cm, gan = self.generator_model, self.critic_model, self.gan_model
        fake = G.forward(z, training=True)
        Cm.forward(fake, training=True)
        loss, _ = Cm.head_loss(-1.0)
        d_fake = Cm.backward(need_wgrad=False, need_input_grad=True, g_out=fake)
        G.backward(d_fake)
        self._allreduce(G.arena.g)
        self._allreduce(Cm.arena.g[Cm.bn_off:])
        gs = 1.0 / self.world
        gan.optimizer.apply(G.arena.w, G.arena.g, G.arena.m, 0, G.arena.numel, gs)
        gan.optimizer.apply(Cm.arena.w, Cm.arena.g, gan.rms_cbn, Cm.bn_off, Cm.bn_n, gs, count=False)
        G._packed_dirty = True
        return loss

    def train_iteration(self, reals, zs_fake, z_gan):
        """critic_learn_iterations x (critic on a real batch, critic on predicted fakes) + one generator update:
        the body of WGAN.train between two g_hist entries (WGAN.py:296-318).
        Returns (mean real loss, mean fake loss, generator loss)."""
        c1, c2 = [], []
        for x, z in zip(reals, zs_fake):
            c1.append(self.train_critic_on_batch(x, -1.0).clone())
            fake = self.generate_fake_samples(z)
            c2.append(self.train_critic_on_batch(fake, 1.0).clone())
        g_loss = self.train_gan_on_batch(z_gan).clone()
        return torch.stack(c1).mean(), torch.stack(c2).mean(), g_loss


class WGAN(WGANEngine):
    """Drop-in for wasserstein_gan/WGAN.py's WGAN class without its filesystem side effects: `dataset` is an
    iterable of fp32 NHWC batches in [-1, 1] (wasserstein_gan/Trainer.py:15-31)."""

    def __init__(self, dataset, image_size, bach_size, critic_learn_iterations, device="cuda", seed=62,
                 process_group=None):
        super().__init__(image_size, bach_size, critic_learn_iterations, device=device, seed=seed,
                         process_group=process_group)
        self.dataset = dataset
        self.epoch = 0
        self.loss_hist = {"c1_hist": [], "c2_hist": [], "g_hist": []}
        self._gen = torch.Generator(device=self.device).manual_seed(seed + 1000 * (self.rank + 1))

    def generate_latent_points(self, n_samples):
        """randn(latent_dim * n).reshape(n, latent_dim) (WGAN.py:211-217), drawn on the device."""
        return torch.randn((n_samples, self.latent_dim), generator=self._gen, device=self.device, dtype=torch.float32)

    def train(self, epochs):
        count = 0
        c1_tmp, c2_tmp = [], []
        h = self.loss_hist
        for _ in range(epochs - self.epoch):
            self.epoch += 1
            for batch in self.dataset:
                batch = batch.to(self.device, torch.float32)
                c1_tmp.append(self.train_critic_on_batch(batch, -1.0).clone())
                fake = self.generate_fake_samples(self.generate_latent_points(self.bach_size))
                c2_tmp.append(self.train_critic_on_batch(fake, 1.0).clone())
                count += 1
                if count == self.critic_learn_iterations:
                    count = 0
                    h["c1_hist"].append(float(torch.stack(c1_tmp).mean()))
                    h["c2_hist"].append(float(torch.stack(c2_tmp).mean()))
                    c1_tmp, c2_tmp = [], []
                    g_loss = self.train_gan_on_batch(self.generate_latent_points(self.bach_size))
                    h["g_hist"].append(float(g_loss))
        return h
