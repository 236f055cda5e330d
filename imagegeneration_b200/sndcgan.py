"""Host-side mirror of the reference's SNDCGAN API (sndcgan/SNDCGAN.py) driving the sm_100a kernels.

Same names and call shapes as the reference so it is a drop-in for the training-step path:
  make_dcgan_generator(output_size)                      SNDCGAN.py:25-66
  make_dcgan_discriminator(dropout_rate, input_size)     SNDCGAN.py:69-128
  SNDCGAN(...).train_step(input_real, input_z) -> 6-tuple SNDCGAN.py:241-269
Models are callables `model(x, training=bool)` exposing `.trainable_variables` in Keras order and Keras
layouts (Conv2D HWIO, Conv2DTranspose HWOI, Dense [in,out]); tensors are torch CUDA tensors, NHWC.

Everything numerical happens in libb200gan.so (imagegeneration_b200/csrc); torch only owns device memory,
streams and (for data parallel runs) the NCCL process group.  There is no CPU fallback.
"""
from __future__ import annotations

import ctypes as C
import math
import os

import torch

from . import lib as _lib
from . import ops
from .lib import (ACT_LEAKY, ACT_NONE, ACT_RELU, ACT_TANH, CONV_K1S1, CONV_K3S1, CONV_K3S1_C3, CONV_K4S2, CONVT_K4S2,
                  B200GanError, check)

BN_EPS = 1e-3                    # Keras BatchNormalization default epsilon
INV_STD = 1.0 / math.sqrt(1.0 + BN_EPS)   # inference BN with moving_var = 1
LEAKY_ALPHA = 0.1                # SNDCGAN.py:77
Z_SIZE = 128                     # Trainer.py:8

_MASK64 = 0xFFFFFFFFFFFFFFFF


def _splitmix64(x: int) -> int:
    x = (x + 0x9E3779B97F4A7C15) & _MASK64
    z = x
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & _MASK64
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & _MASK64
    return z ^ (z >> 31)


def step_dropout_seeds(seed: int, rank: int, step: int, pass_idx: int) -> list[int]:
    """The 7 per-layer dropout seeds the engine uses for discriminator pass `pass_idx` (0: generator step, 1: real,
    2: stale fake) of training step `step` — a pure function of (engine seed, rank, step), restated on the device by
    b2g_step_prologue, so tests and bench.py can regenerate the masks (oracle/dropout_hash.py) without a GPU."""
    base = (seed * 0x1000003 + rank * 0x10001 + 1) & _MASK64
    return [_splitmix64(base ^ ((step * 3 + pass_idx) * 8 + layer)) for layer in range(7)]


def rate_to_thresh8(rate: float) -> int:
    return max(0, min(255, int(round(rate * 256.0))))


def _p(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _glorot_uniform(shape, gen):
    if len(shape) == 2:
        fan_in, fan_out = shape
    else:
        rf = 1
        for d in shape[:-2]:
            rf *= d
        fan_in, fan_out = shape[-2] * rf, shape[-1] * rf
    limit = math.sqrt(6.0 / (fan_in + fan_out))
    return ((torch.rand(shape, generator=gen, dtype=torch.float64) * 2 - 1) * limit).to(torch.float32)


class ParamArena:
    """All variables of one network in flat fp32 device buffers (weights, grads, Adam m and v) so that one
    fused Adam launch and one NCCL all-reduce cover the whole network.  Variables are views into the arena."""

    def __init__(self, shapes: list[tuple[str, tuple[int, ...]]], device):
        self.names = [n for n, _ in shapes]
        self.shapes = [s for _, s in shapes]
        self.offsets = []
        off = 0
        for s in self.shapes:
            self.offsets.append(off)
            n = 1
            for d in s:
                n *= d
            off += (n + 3) // 4 * 4
        self.numel = off
        self.w = torch.zeros(off, dtype=torch.float32, device=device)
        self.g = torch.zeros(off, dtype=torch.float32, device=device)
        self.m = torch.zeros(off, dtype=torch.float32, device=device)
        self.v = torch.zeros(off, dtype=torch.float32, device=device)

    def _views(self, flat):
        out = []
        for s, o in zip(self.shapes, self.offsets):
            n = 1
            for d in s:
                n *= d
            out.append(flat[o:o + n].view(s))
        return out

    @property
    def variables(self):
        return self._views(self.w)

    @property
    def grads(self):
        return self._views(self.g)


class Adam:
    """tf.keras.optimizers.Adam(learning_rate) (SNDCGAN.py:165,168): beta 0.9/0.999, epsilon 1e-7."""

    def __init__(self, learning_rate=0.001, beta_1=0.9, beta_2=0.999, epsilon=1e-7):
        self.learning_rate = learning_rate
        self.beta_1, self.beta_2, self.epsilon = beta_1, beta_2, epsilon
        self.iterations = 0

    def apply_arena(self, arena: ParamArena, gscale: float = 1.0, lr_t_dev: torch.Tensor | None = None) -> None:
        """lr_t_dev (1-element fp32 CUDA tensor written by b2g_step_prologue): the bias-corrected rate is read on
        the device, so the launch can sit in a replayed CUDA graph; `iterations` is then advanced by the caller."""
        if lr_t_dev is not None:
            check(_lib.load().b2g_adam_dev(_p(arena.w), _p(arena.g), _p(arena.m), _p(arena.v), arena.numel,
                                           _p(lr_t_dev), self.beta_1, self.beta_2, self.epsilon, gscale, _stream()),
                  "b2g_adam_dev")
            return
        self.iterations += 1
        check(_lib.load().b2g_adam(_p(arena.w), _p(arena.g), _p(arena.m), _p(arena.v), arena.numel, self._lr_t(),
                                   self.beta_1, self.beta_2, self.epsilon, gscale, _stream()), "b2g_adam")

    def apply_arena_nvls(self, arena: ParamArena, nv: dict, gscale: float, lr_t_dev: torch.Tensor | None = None) -> None:
        """Data-parallel apply_gradients over NVSwitch multicast memory (b2g_adam_nvls): this rank reduces, updates and
        broadcasts its slice of the arena; the barriers make all gradients final before and all weights visible after."""
        lr_t = 0.0
        if lr_t_dev is None:
            self.iterations += 1
            lr_t = self._lr_t()
        nv["hg"].barrier(channel=0)
        check(_lib.load().b2g_adam_nvls(C.c_void_p(nv["w_mc"]), C.c_void_p(nv["g_mc"]), _p(nv["w_full"]), _p(nv["m_full"]),
                                        _p(nv["v_full"]), nv["begin"], nv["end"], lr_t, _p(lr_t_dev), self.beta_1,
                                        self.beta_2, self.epsilon, gscale, _stream()), "b2g_adam_nvls")
        nv["hw"].barrier(channel=1)

    def _lr_t(self) -> float:
        t = self.iterations
        return self.learning_rate * math.sqrt(1.0 - self.beta_2 ** t) / (1.0 - self.beta_1 ** t)

    def apply_model(self, model, gscale: float = 1.0, lr_t_dev: torch.Tensor | None = None) -> None:
        """apply_gradients for a whole model.  A spectrally-normalised discriminator takes the fused path: one dot
        launch (<G,W> per matrix) + ONE pass over the arena that maps the raw gradients through d(W/sigma)/dW, applies
        Adam and runs the next power iteration on the updated rows (b2g_adam_sn) — no stand-alone spectral-norm pass."""
        if not getattr(model, "spectral_norm", False) or model._sn_grads_final:
            self.apply_arena(model.arena, gscale, lr_t_dev)
            if getattr(model, "spectral_norm", False):
                model._sn_stale = True          # weights moved: the next forward runs a stand-alone power iteration
            model._packed_dirty = True
            return
        lib, a, sn = _lib.load(), model.arena, model._sn_args
        check(lib.b2g_sn_dot_multi(_p(a.w), _p(a.g), 8, model._sn_off, sn[5], sn[6], _p(model._sn_dots), _stream()),
              "b2g_sn_dot_multi")
        lr_t = 0.0
        if lr_t_dev is None:
            self.iterations += 1
            lr_t = self._lr_t()
        check(lib.b2g_adam_sn(_p(a.w), _p(a.g), _p(a.m), _p(a.v), a.numel, 8, model._sn_off, sn[5], sn[6], sn[1], sn[2],
                              sn[3], sn[4], _p(model._sn_dots), lr_t, _p(lr_t_dev), self.beta_1, self.beta_2,
                              self.epsilon, gscale, _stream()), "b2g_adam_sn")
        model._sn_stale = False                 # sigma / u now belong to the updated weights
        model._packed_dirty = True


# ----------------------------------------------------------------------------------------------------------
# Generator  (SNDCGAN.py:25-66)
# ----------------------------------------------------------------------------------------------------------
class Generator:
    name = "SNDC_Generator"

    def __init__(self, output_size, device="cuda", seed=62):
        self.output_size = tuple(output_size)
        h, w, c = self.output_size
        if c != 3 or h % 8 or w % 8:
            raise B200GanError("generator needs an (H, W, 3) output with H, W divisible by 8")
        self.h0, self.w0 = h // 8, w // 8
        self.F = 512 * self.h0 * self.w0
        self.device = torch.device(device)
        shapes = [("dense/kernel", (Z_SIZE, self.F)), ("bn0/gamma", (self.F,)), ("bn0/beta", (self.F,)),
                  ("convt1/kernel", (4, 4, 256, 512)), ("bn1/gamma", (256,)), ("bn1/beta", (256,)),
                  ("convt2/kernel", (4, 4, 128, 256)), ("bn2/gamma", (128,)), ("bn2/beta", (128,)),
                  ("convt3/kernel", (4, 4, 64, 128)), ("bn3/gamma", (64,)), ("bn3/beta", (64,)),
                  ("convt4/kernel", (3, 3, 3, 64))]
        self.arena = ParamArena(shapes, self.device)
        gen = torch.Generator().manual_seed(seed)
        for name, v in zip(self.arena.names, self.arena.variables):
            if name.endswith("kernel"):
                v.copy_(_glorot_uniform(tuple(v.shape), gen))
            elif name.endswith("gamma"):
                v.fill_(1.0)
        self._inv_std = torch.full((1,), INV_STD, dtype=torch.float32, device=self.device)
        self._pack_plan = None            # ops.PackPlan: all packings in one launch
        self._wp_f = [None] * 3           # packed fprop weights
        self._wp_d = [None] * 3           # packed dgrad weights
        self._wp4_f = None                # convT4 forward  (64 -> 3, flipped 3x3 taps), [1][3][576]
        self._wp4_d = None                # convT4 input-gradient (im2col27 -> 64), [1][64][64]
        self._packed_dirty = True
        self._bufs = {}
        self._z = None
        self._params = None

    # -- Keras-like surface
    @property
    def trainable_variables(self):
        return self.arena.variables

    def set_weights(self, tensors):
        for v, t in zip(self.arena.variables, tensors):
            v.copy_(t.to(self.device, torch.float32))
        self._packed_dirty = True

    def parameters(self):
        """The trainable variables as autograd LEAVES (views of the weight arena, requires_grad=True).  Once a caller
        has asked for them, `model(z)` under torch.enable_grad() builds an autograd graph out of the b200gan custom ops
        (torch_ops.py) and `.backward()` fills their `.grad` — any loss can be put on top."""
        if self._params is None:
            self._params = [v.requires_grad_(True) for v in self.arena.variables]
        return self._params

    def __call__(self, z, training=False):
        # BatchNorm runs in inference mode whatever `training` says: the moving statistics stay at (0, 1) in the
        # reference because its train_step never passes training=True to the generator (SNDCGAN.py:243).
        if self._params is not None and torch.is_grad_enabled():
            from .torch_ops import generator_graph
            return generator_graph(self, z)
        return self.forward(z)

    # -- engine
    def _buf(self, key, shape, dtype):
        # one buffer per (name, shape, dtype), never freed: a captured CUDA graph keeps the addresses of the
        # full-batch buffers, so a ragged batch in between must not make the allocator recycle them
        k = (key, tuple(shape), dtype)
        t = self._bufs.get(k)
        if t is None:
            t = torch.empty(shape, dtype=dtype, device=self.device)
            self._bufs[k] = t
        return t

    def repack(self):
        """One launch re-packs all eight packed weights (BatchNorm scale gamma * inv_std folded in)."""
        if self._pack_plan is None:
            v = self.arena.variables
            plan = ops.PackPlan()
            for i, k in enumerate((3, 6, 9)):
                gamma = v[k + 1]
                self._wp_f[i] = plan.add(CONVT_K4S2, v[k], n_is_d3=False, scale_scalar=self._inv_std, scale_d2=gamma)
                self._wp_d[i] = plan.add(CONV_K4S2, v[k], n_is_d3=True, scale_scalar=self._inv_std, scale_d2=gamma)
            self._wp4_f = plan.add(CONV_K3S1, v[12], n_is_d3=False, flip=True)
            self._wp4_d = plan.add(CONV_K1S1, v[12].view(1, 1, 27, 64), n_is_d3=True, ck_pad=64)
            self._pack_plan = plan
        self._pack_plan.run()
        self._packed_dirty = False

    def forward(self, z):
        lib = _lib.load()
        if z.dtype != torch.float32 or not z.is_cuda or z.shape[1] != Z_SIZE:
            raise B200GanError("generator input must be a CUDA fp32 tensor [B, 128]")
        z = z.contiguous()
        if self._packed_dirty:
            self.repack()
        B = z.shape[0]
        v = self.arena.variables
        h0, w0 = self.h0, self.w0
        a0 = self._buf("a0", (B, h0, w0, 512), torch.bfloat16)
        check(lib.b2g_dense_g_fwd(_p(z), _p(v[0]), _p(v[1]), _p(v[2]), INV_STD, _p(a0), B, self.F, _stream()),
              "b2g_dense_g_fwd")
        acts = [a0]
        x = a0
        for i, k in enumerate((3, 6, 9)):
            cout = v[k].shape[2]
            out = self._buf(f"a{i + 1}", (B, x.shape[1] * 2, x.shape[2] * 2, cout), torch.bfloat16)
            ops.conv_igemm(CONVT_K4S2, x, self._wp_f[i], out, bias=v[k + 2], act=ACT_RELU)
            acts.append(out)
            x = out
        H, W, _ = self.output_size
        g_out = self._buf("g_out", (B, H, W, 3), torch.float32)
        ops.conv_igemm(CONV_K3S1, x, self._wp4_f, g_out, act=ACT_TANH, out_f32=True)
        self._acts = acts
        self._z = z
        self.g_out = g_out
        return g_out

    def zero_grads(self):
        # the dense kernel / bn0 gradients are overwritten by the stem backward; conv gradients are accumulated
        tail = self.arena.offsets[3]
        self.arena.g[tail:].zero_()

    def backward(self, d_pre_tanh):
        """d_pre_tanh: fp32 [B,H,W,3] = dL/d(convT4 output before tanh).  Fills arena.g."""
        lib = _lib.load()
        v, g = self.arena.variables, self.arena.grads
        a0, a1, a2, a3 = self._acts
        B, H, W, _ = d_pre_tanh.shape
        self.zero_grads()
        # convT4 (64 -> 3, k3 s1) on tensor cores through the 27-channel im2col of dY: wgrad + input gradient
        # (ReLU' of a3 in the epilogue)
        ops.conv_wgrad(CONV_K3S1_C3, d_pre_tanh, a3, g[12])
        # every input-gradient call also leaves the per-channel sums of what it stored = the beta gradient of the
        # BatchNorm in front of it (fused into the epilogue: no separate pass over dq)
        dq = self._buf("dq3", tuple(a3.shape), torch.bfloat16)
        ops.conv_igemm(CONV_K3S1_C3, d_pre_tanh, self._wp4_d, dq, yprev=a3, dact=ACT_RELU, colsum=g[9 + 2])
        acts_in = [a0, a1, a2]
        for i in (2, 1, 0):
            k = (3, 6, 9)[i]
            x_in = acts_in[i]
            cout, cin = v[k].shape[2], v[k].shape[3]
            # raw weight gradient (w.r.t. the folded weight), then gamma gradient + rescale
            ops.conv_wgrad(CONV_K4S2, dq, x_in, g[k])
            check(lib.b2g_bn_fold_finalize(_p(v[k]), _p(g[k]), _p(v[k + 1]), INV_STD, _p(g[k + 1]), 16, cout, cin,
                                           _stream()), "b2g_bn_fold_finalize")
            dprev = self._buf(f"dq{i}", tuple(x_in.shape), torch.bfloat16)
            ops.conv_igemm(CONV_K4S2, dq, self._wp_d[i], dprev, yprev=x_in, dact=ACT_RELU,
                           colsum=g[(3, 6, 9)[i - 1] + 2] if i > 0 else None)
            dq = dprev
        check(lib.b2g_dense_g_bwd(_p(self._z), _p(v[0]), _p(v[1]), INV_STD, _p(dq), _p(g[0]), _p(g[1]), _p(g[2]),
                                  B, self.F, _stream()), "b2g_dense_g_bwd", (B + 31) // 32)


# ----------------------------------------------------------------------------------------------------------
# Discriminator  (SNDCGAN.py:69-128)
# ----------------------------------------------------------------------------------------------------------
D_CONVS = [("conv1", 3, 1, 3, 64), ("conv2", 4, 2, 64, 128), ("conv3", 3, 1, 128, 128), ("conv4", 4, 2, 128, 256),
           ("conv5", 3, 1, 256, 256), ("conv6", 4, 2, 256, 512), ("conv7", 3, 1, 512, 512)]


class Discriminator:
    name = "SNDC_Discriminator"

    def __init__(self, dropout_rate, input_size, device="cuda", seed=63, spectral_norm=False):
        self.input_size = tuple(input_size)
        h, w, c = self.input_size
        if c != 3 or h % 8 or w % 8:
            raise B200GanError("discriminator needs an (H, W, 3) input with H, W divisible by 8")
        self.dropout_rate = float(dropout_rate)
        self.thresh8 = rate_to_thresh8(self.dropout_rate)
        self.keep_scale = 256.0 / (256.0 - self.thresh8)
        self.device = torch.device(device)
        self.F = 512 * (h // 8) * (w // 8)
        shapes = []
        for name, k, _s, cin, cout in D_CONVS:
            shapes += [(f"{name}/kernel", (k, k, cin, cout)), (f"{name}/bias", (cout,))]
        shapes += [("dense/kernel", (self.F, 1)), ("dense/bias", (1,))]
        self.arena = ParamArena(shapes, self.device)
        gen = torch.Generator().manual_seed(seed)
        for name, v in zip(self.arena.names, self.arena.variables):
            if name.endswith("kernel"):
                v.copy_(_glorot_uniform(tuple(v.shape), gen))
        self._pack_plan = None
        self._wp_f = [None] * 7
        self._wp_d = [None] * 7
        self._packed_dirty = True
        self._bufs = {}
        self._params = None
        self.spectral_norm = bool(spectral_norm)
        self._sn_stale = False
        self._sn_grads_final = False
        if self.spectral_norm:
            self._init_sn(seed + 1)

    # -- spectral norm state (north-star feature; absent from the reference)
    def _init_sn(self, seed):
        gen = torch.Generator().manual_seed(seed)
        v = self.arena.variables
        self.sn_u, self.sn_vraw, self.sn_traw, self.sn_scal = [], [], [], []
        Ks, Cs = [], []
        for i in range(8):
            wt = v[2 * i]
            cout = wt.shape[-1]
            k = wt.numel() // cout
            u = torch.randn(cout, generator=gen, dtype=torch.float64)
            self.sn_u.append((u / u.norm()).to(torch.float32).to(self.device))
            self.sn_vraw.append(torch.zeros(k, dtype=torch.float32, device=self.device))
            self.sn_traw.append(torch.zeros(cout, dtype=torch.float32, device=self.device))
            self.sn_scal.append(torch.zeros(8, dtype=torch.float32, device=self.device))
            Ks.append(k)
            Cs.append(cout)
        arr = lambda ts: (C.c_void_p * 8)(*[t.data_ptr() for t in ts])
        self._sn_args = (arr([v[2 * i] for i in range(8)]), arr(self.sn_u), arr(self.sn_vraw), arr(self.sn_traw),
                         arr(self.sn_scal), (C.c_int * 8)(*Ks), (C.c_int * 8)(*Cs))
        self._sn_K, self._sn_C = Ks, Cs
        self._sn_dot = torch.zeros(1, dtype=torch.float32, device=self.device)
        self._sn_dots = torch.zeros(8, dtype=torch.float32, device=self.device)      # <G,W> per matrix (fused path)
        self._sn_off = (C.c_long * 8)(*[self.arena.offsets[2 * i] for i in range(8)])
        # sigma is undefined until a power iteration has seen the current weights: the first forward (training or
        # not) runs a stand-alone one; afterwards every optimizer update carries the next iteration (b2g_adam_sn)
        self._sn_stale = True

    def sn_power_iteration(self):
        """Stand-alone power iteration (one launch, all 8 matrices): used when the weights were set from outside
        (initialisation, set_weights, checkpoint restore).  Training updates get theirs inside b2g_adam_sn."""
        a = self._sn_args
        check(_lib.load().b2g_spectral_norm_multi(8, a[0], a[1], a[2], a[3], a[4], a[5], a[6], _stream()),
              "b2g_spectral_norm_multi", 2)
        self._sn_stale = False
        self._packed_dirty = True

    def finalize_sn_grads(self):
        """Maps the raw kernel gradients of the last backward (dL/d(W/sigma)) to dL/dW in place — only for callers
        that want to READ the gradients; train_step leaves them raw because b2g_adam_sn applies the same map."""
        if not self.spectral_norm or self._sn_grads_final:
            return
        lib, v, g = _lib.load(), self.arena.variables, self.arena.grads
        for i in range(8):
            check(lib.b2g_sn_grad_finalize(_p(v[2 * i]), _p(g[2 * i]), _p(self.sn_u[i]), _p(self.sn_vraw[i]),
                                           _p(self.sn_scal[i]), _p(self._sn_dot), self._sn_K[i], self._sn_C[i],
                                           _stream()), "b2g_sn_grad_finalize", 3)
        self._sn_grads_final = True

    def _inv_sigma(self, i):
        return self.sn_scal[i][2:3] if self.spectral_norm else None

    # -- Keras-like surface
    @property
    def trainable_variables(self):
        return self.arena.variables

    def set_weights(self, tensors):
        for v, t in zip(self.arena.variables, tensors):
            v.copy_(t.to(self.device, torch.float32))
        self._packed_dirty = True
        self._sn_stale = self.spectral_norm

    def parameters(self):
        """See Generator.parameters: autograd leaves over the weight arena; enables the torch.library graph path."""
        if self._params is None:
            self._params = [v.requires_grad_(True) for v in self.arena.variables]
        return self._params

    def __call__(self, x, training=False, dropout_seeds=None):
        """Keras-style call: returns the logits [B,1] (SNDCGAN.py:244,252,260).  When autograd is recording (the
        input requires grad, or parameters() has been handed out) the call builds a differentiable graph of b200gan
        custom ops instead of using the engine's buffers."""
        if training and dropout_seeds is None and self.thresh8:
            self._call_count = getattr(self, "_call_count", 0) + 1
            dropout_seeds = [_splitmix64(0xD15C0000 + self._call_count * 8 + i) for i in range(7)]
        if torch.is_grad_enabled() and (self._params is not None or x.requires_grad):
            from .torch_ops import discriminator_graph
            self.parameters()
            return discriminator_graph(self, x, training, dropout_seeds)
        self.forward(x, training=training, dropout_seeds=dropout_seeds)
        _loss, logits = self.head_loss(0, 1.0, 1.0 / x.shape[0])
        return logits.clone()

    def _buf(self, key, shape, dtype):
        # one buffer per (name, shape, dtype), never freed: a captured CUDA graph keeps the addresses of the
        # full-batch buffers, so a ragged batch in between must not make the allocator recycle them
        k = (key, tuple(shape), dtype)
        t = self._bufs.get(k)
        if t is None:
            t = torch.empty(shape, dtype=dtype, device=self.device)
            self._bufs[k] = t
        return t

    def repack(self):
        """One launch re-packs the forward and input-gradient packings of all seven conv layers (1/sigma folded in)."""
        if self._pack_plan is None:
            v = self.arena.variables
            plan = ops.PackPlan()
            sc0 = self._inv_sigma(0)
            self._wp_f[0] = plan.add(CONV_K1S1, v[0].view(1, 1, 27, 64), n_is_d3=True, ck_pad=64,
                                     scale_scalar=sc0)                                    # im2col27 -> 64
            self._wp_d[0] = plan.add(CONV_K3S1, v[0], n_is_d3=False, flip=True, scale_scalar=sc0)   # 64 -> 3
            for i in range(1, 7):
                _name, k, s, _ci, _co = D_CONVS[i]
                sc = self._inv_sigma(i)
                if s == 1:
                    self._wp_f[i] = plan.add(CONV_K3S1, v[2 * i], n_is_d3=True, scale_scalar=sc)
                    self._wp_d[i] = plan.add(CONV_K3S1, v[2 * i], n_is_d3=False, flip=True, scale_scalar=sc)
                else:
                    self._wp_f[i] = plan.add(CONV_K4S2, v[2 * i], n_is_d3=True, scale_scalar=sc)
                    self._wp_d[i] = plan.add(CONVT_K4S2, v[2 * i], n_is_d3=False, scale_scalar=sc)
            self._pack_plan = plan
        self._pack_plan.run()
        self._packed_dirty = False

    def forward(self, x, training=False, dropout_seeds=None, x27=None):
        """x: fp32 CUDA [B,H,W,3].  With training=True dropout is active; `dropout_seeds` (7 ints) makes the
        masks reproducible (tests regenerate them with oracle/dropout_hash.py).  x27: optional cached
        ops.im2col27(x) (train_step feeds the same generated batch to two passes)."""
        lib = _lib.load()
        if x.dtype != torch.float32 or not x.is_cuda or tuple(x.shape[1:]) != self.input_size:
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
        # conv1 (3 -> 64): the 27-wide im2col patch is built inside the kernel (B2G_CONV_K3S1_C3); the im2col27 tensor
        # is only materialised in backward(), for the weight gradient
        c1 = self._buf("c1", (B, H, W, 64), torch.bfloat16)
        ops.conv_igemm(CONV_K3S1_C3, x, self._wp_f[0], c1, bias=v[1], act=ACT_LEAKY, alpha=LEAKY_ALPHA,
                       dropout_thresh8=t8, dropout_seed=seeds[0])
        self._x_in, self._x27 = x, x27
        acts = [c1]
        cur = c1
        for i in range(1, 7):
            _name, k, s, _ci, cout = D_CONVS[i]
            out = self._buf(f"c{i + 1}", (B, cur.shape[1] // s, cur.shape[2] // s, cout), torch.bfloat16)
            ops.conv_igemm(CONV_K3S1 if s == 1 else CONV_K4S2, cur, self._wp_f[i], out, bias=v[2 * i + 1],
                           act=ACT_LEAKY, alpha=LEAKY_ALPHA, dropout_thresh8=t8, dropout_seed=seeds[i])
            acts.append(out)
            cur = out
        self._x = x
        self._acts = acts
        self._train_fwd = bool(t8)
        self._B = B
        self.logits = self._buf("logits", (B, 1), torch.float32)
        self._dlogit = self._buf("dlogit", (B,), torch.float32)
        self._loss = torch.empty(2, dtype=torch.float32, device=self.device)  # fresh: returned to the caller
        self._ws = self._buf("head_ws", (64 * B,), torch.float32)
        self._head_pending = True
        return self

    def head_loss(self, loss_kind: int, label: float, inv_batch: float):
        """Dense(1) + loss.  Returns (loss scalar tensor, logits [B,1])."""
        v = self.arena.variables
        check(_lib.load().b2g_dhead_fwd_loss(_p(self._acts[6]), _p(v[14]), _p(self._inv_sigma(7)), _p(v[15]),
                                             self._B, self.F, loss_kind, label, inv_batch, _p(self._ws),
                                             self._ws.numel(), _p(self.logits), _p(self._dlogit), _p(self._loss),
                                             _stream()), "b2g_dhead_fwd_loss", 2)
        return self._loss[0], self.logits

    # first variable of the gradient range that is complete once conv6's weight gradient has been issued
    # (dense, conv7, conv6: 19 of the 24.6 MB) — data-parallel runs start reducing it while conv5..conv1 still run
    EARLY_GRAD_FIRST_VAR = 10

    def backward(self, need_wgrad: bool, need_input_grad: bool, g_out=None, on_tail_grads_ready=None):
        """Backward from the dlogit of the last head_loss.  need_wgrad fills arena.g (zeroed first);
        need_input_grad returns dL/d(pre-tanh generator output) fp32 [B,H,W,3] (uses tanh'(g_out)).
        on_tail_grads_ready: called once the gradients of variables EARLY_GRAD_FIRST_VAR.. are final (never with
        spectral norm, whose finalisation rewrites every kernel gradient at the end)."""
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
        dpre = self._buf("dpre7", tuple(c7.shape), torch.bfloat16)
        check(lib.b2g_dhead_bwd(_p(c7), _p(v[14]), _p(self._inv_sigma(7)), _p(self._dlogit), B, self.F, LEAKY_ALPHA,
                                dscale, dzero, _p(dpre), _p(g[14]) if need_wgrad else None,
                                _p(g[15]) if need_wgrad else None, _stream()), "b2g_dhead_bwd")
        for i in range(6, 0, -1):
            _name, k, s, _ci, cout = D_CONVS[i]
            x_in = acts[i - 1]
            if need_wgrad:
                if i == 6:  # conv7's output gradient comes from the head kernel; the others carry fused sums
                    check(lib.b2g_colsum_bf16(_p(dpre), _p(g[2 * i + 1]), dpre.numel() // cout, cout, _stream()),
                          "b2g_colsum_bf16")
                ops.conv_wgrad(CONV_K3S1 if s == 1 else CONV_K4S2, x_in, dpre, g[2 * i])
                if i == 5 and on_tail_grads_ready is not None and not self.spectral_norm:
                    on_tail_grads_ready()
            dprev = self._buf(f"dpre{i}", tuple(x_in.shape), torch.bfloat16)
            # the input-gradient call also sums what it stores per channel = bias gradient of the layer below
            # (conv1's comes out of its weight-gradient GEMM through the ones channel of the im2col operand)
            ops.conv_igemm(CONV_K3S1 if s == 1 else CONVT_K4S2, dpre, self._wp_d[i], dprev, yprev=x_in,
                           dact=ACT_LEAKY, dact_alpha=LEAKY_ALPHA, dact_scale=dscale, dact_zero=dzero,
                           colsum=g[2 * (i - 1) + 1] if (need_wgrad and i >= 2) else None)
            dpre = dprev
        H, W = self.input_size[0], self.input_size[1]
        if need_wgrad:
            if self._x27 is not None:   # caller supplied the im2col operand (b2g_im2col27): K1S1 GEMM route
                ops.conv_wgrad(CONV_K1S1, self._x27, dpre, g[0], ca_valid=27, dbias=g[1])
            else:                       # patch built inside the kernel from the fp32 input
                ops.conv_wgrad(CONV_K3S1_C3, self._x_in, dpre, g[0], dbias=g[1])
            # with spectral norm the kernel gradients stay RAW (dL/d(W/sigma)): the optimizer's fused pass maps them
            # through d(W/sigma)/dW (Adam.apply_model); finalize_sn_grads() does it in place for callers that read them
        d_in = None
        if need_input_grad:
            d_in = self._buf("d_in", (B, H, W, 3), torch.float32)
            ops.conv_igemm(CONV_K3S1, dpre, self._wp_d[0], d_in, yprev=g_out,
                           dact=ACT_TANH if g_out is not None else ACT_NONE, out_f32=True)
        return d_in


def _check_dtype(dtype):
    if dtype not in ("bf16", "tf32"):
        raise B200GanError("dtype must be 'bf16' (training path) or 'tf32' (fp32 storage / kind::tf32 parity path)")


def make_dcgan_generator(output_size, device="cuda", seed=62, dtype="bf16") -> Generator:
    """sndcgan/SNDCGAN.py:25.  dtype="tf32": the fp32-storage / kind::tf32 parity path (sndcgan_tf32.py)."""
    _check_dtype(dtype)
    if dtype == "tf32":
        from .sndcgan_tf32 import GeneratorTF32
        return GeneratorTF32(output_size, device=device, seed=seed)
    return Generator(output_size, device=device, seed=seed)


def make_dcgan_discriminator(dropout_rate, input_size, device="cuda", seed=63, spectral_norm=False,
                             dtype="bf16") -> Discriminator:
    """sndcgan/SNDCGAN.py:69."""
    _check_dtype(dtype)
    if dtype == "tf32":
        from .sndcgan_tf32 import DiscriminatorTF32
        return DiscriminatorTF32(dropout_rate, input_size, device=device, seed=seed, spectral_norm=spectral_norm)
    return Discriminator(dropout_rate, input_size, device=device, seed=seed, spectral_norm=spectral_norm)


# ----------------------------------------------------------------------------------------------------------
# SNDCGAN: state + train_step  (SNDCGAN.py:148-269)
# ----------------------------------------------------------------------------------------------------------
LOSS_BCE, LOSS_HINGE_D, LOSS_HINGE_G = 0, 1, 2


class SNDCGANEngine:
    """Models, optimizers and the exact step ordering of SNDCGAN.train_step (SNDCGAN.py:241-269), without the
    reference's filesystem / dataset side effects (those live in trainer.SNDCGAN, which subclasses this)."""

    def __init__(self, batch_size, dropout=0.5, learning_rate_disc=2e-4, learning_rate_gen=2e-4,
                 image_size=(144, 256, 3), z_size=Z_SIZE, device="cuda", seed=62, spectral_norm=False, loss="bce",
                 process_group=None, cuda_graph=None, dtype="bf16"):
        if z_size != Z_SIZE:
            raise B200GanError("z_size must be 128 (sndcgan/Trainer.py:8)")
        if loss not in ("bce", "hinge"):
            raise B200GanError("loss must be 'bce' (reference) or 'hinge'")
        _lib.load()
        self.batch_size = batch_size
        self.z_size = z_size
        self.image_size = tuple(image_size)
        self.device = torch.device(device)
        self.loss = loss
        _check_dtype(dtype)
        self.dtype = dtype
        if dtype == "tf32":
            cuda_graph = False            # parity configuration: eager steps, host dropout seeds
        self.gen_model = make_dcgan_generator(output_size=image_size, device=device, seed=seed, dtype=dtype)   # :164
        self.g_optimizer = Adam(learning_rate=learning_rate_gen)                                         # :165
        self.disc_model = make_dcgan_discriminator(dropout, input_size=image_size, device=device, seed=seed + 1,
                                                   spectral_norm=spectral_norm, dtype=dtype)             # :167
        self.d_optimizer = Adam(learning_rate=learning_rate_disc)                                        # :168
        from .torch_ops import BinaryCrossentropy
        self.loss_fn = BinaryCrossentropy(from_logits=True)                                              # :170
        self.seed = seed
        self.step_count = 0
        self.pg = process_group
        self.world = 1
        self.rank = 0
        if process_group is not None:
            import torch.distributed as dist
            self.world = dist.get_world_size(process_group)
            self.rank = dist.get_rank(process_group)
            # identical initial weights on every rank
            dist.broadcast(self.gen_model.arena.w, src=0, group=process_group)
            dist.broadcast(self.disc_model.arena.w, src=0, group=process_group)
            self.gen_model._packed_dirty = True
            self.disc_model._packed_dirty = True
        # The generator's 164 MB gradient all-reduce hides behind both discriminator steps, but its NCCL kernel holds SMs
        # meanwhile, and a persistent grid of one CTA per SM then runs in TWO waves: the CTAs that found their SM taken
        # start when the others finish, with a full static share of the tiles (per-rank timeline, profiles/r2_dp_timeline.md:
        # every kernel that overlaps an NCCL kernel takes 1.4-1.9x its time, +0.32 ms per step at 2 GPUs).  So the hidden
        # all-reduce gets its OWN communicator capped at B2G_DP_G_CTAS CTAs (slow, but it has ~5 ms to finish), and the
        # kernels launched while it may be in flight leave exactly that many SMs free (b2g_set_sm_limit).
        self.pg_g = self.pg
        self._g_ctas = 0
        if self.world > 1:
            import torch.distributed as dist
            n = int(os.environ.get("B2G_DP_G_CTAS", "4"))
            if n > 0 and dist.get_backend(process_group) == "nccl":
                opts = dist.ProcessGroupNCCL.Options()
                opts.config.max_ctas = n
                opts.config.min_ctas = n
                self.pg_g = dist.new_group(ranks=list(range(self.world)), backend="nccl", pg_options=opts)
                self._g_ctas = n
                dist.all_reduce(torch.zeros(1, device=self.device), group=self.pg_g)   # communicator set up before any capture
        self._nvls = None
        if self.world > 1:
            self._setup_nvls()
        self.gpu_launches = 0
        # Single-GPU steps with engine-chosen dropout seeds are replayed as ONE CUDA graph after two eager steps
        # (the per-step scalars — dropout seeds, Adam's bias-corrected rates — then come from device memory, written
        # by b2g_step_prologue).  cuda_graph=False or B2G_CUDA_GRAPH=0 keeps every step eager.
        if cuda_graph is None:
            cuda_graph = os.environ.get("B2G_CUDA_GRAPH", "1") != "0"
        self.cuda_graph = bool(cuda_graph)
        self._graph = None
        self._graph_failed = None
        self._eager_steps = 0

    def _setup_nvls(self) -> None:
        """Data parallel on an NVSwitch box: the discriminator's two EXPOSED gradient all-reduces + Adam passes of a step
        become b2g_adam_nvls launches (reduce in the switch, sharded update, multicast weight store).  Its weight and
        gradient arenas move into symmetric memory bound to a multicast object (torch.distributed._symmetric_memory does
        the cuMem / cuMulticast plumbing); m and v stay local, each rank keeping its own slice current
        (gather_optimizer_state() before a checkpoint).  Falls back to NCCL all-reduce + b2g_adam when multicast memory is
        unavailable, for a spectrally-normalised discriminator (its fused b2g_adam_sn pass needs whole matrices), on the
        fp32/tf32 parity path, or with B2G_DP_NVLS=0."""
        import torch.distributed as dist
        D = self.disc_model
        if (os.environ.get("B2G_DP_NVLS", "1") == "0" or D.spectral_norm or self.dtype != "bf16"
                or dist.get_backend(self.pg) != "nccl"):
            return
        try:
            import torch.distributed._symmetric_memory as symm
            a = D.arena
            per = (a.numel + 4 * self.world - 1) // (4 * self.world) * 4
            padded = per * self.world
            w_s = symm.empty(padded, dtype=torch.float32, device=self.device)
            g_s = symm.empty(padded, dtype=torch.float32, device=self.device)
            hw = symm.rendezvous(w_s, self.pg)
            hg = symm.rendezvous(g_s, self.pg)
            if not hw.multicast_ptr or not hg.multicast_ptr:
                return
        except Exception as e:      # no multicast support in this driver / container: NCCL path
            if os.environ.get("B2G_DP_NVLS") == "1":
                raise
            self._nvls_error = repr(e)
            return
        w_s.zero_()
        g_s.zero_()
        w_s[:a.numel].copy_(a.w)
        m_full = torch.zeros(padded, dtype=torch.float32, device=self.device)
        v_full = torch.zeros(padded, dtype=torch.float32, device=self.device)
        a.w, a.g, a.m, a.v = w_s[:a.numel], g_s[:a.numel], m_full[:a.numel], v_full[:a.numel]
        D._pack_plan = None
        D._packed_dirty = True
        self._nvls = {"hw": hw, "hg": hg, "w_mc": hw.multicast_ptr, "g_mc": hg.multicast_ptr, "w_full": w_s, "g_full": g_s,
                      "m_full": m_full, "v_full": v_full, "begin": self.rank * per, "end": (self.rank + 1) * per, "per": per}
        torch.cuda.synchronize()
        dist.barrier(group=self.pg)

    def gather_optimizer_state(self) -> None:
        """NVLS data parallel keeps Adam's m / v of the discriminator sharded by rank; this all-gathers the slices so that
        every rank holds the full state (call before saving a checkpoint).  No-op otherwise."""
        if self._nvls is None:
            return
        import torch.distributed as dist
        nv = self._nvls
        for full in (nv["m_full"], nv["v_full"]):
            dist.all_gather_into_tensor(full, full[nv["begin"]:nv["end"]].clone(), group=self.pg)

    def close(self) -> None:
        """Drops the captured CUDA graph (it references the NCCL kernels of a data-parallel step) and waits for the
        device: call before torch.distributed.destroy_process_group()."""
        self._graph = None
        self._graph_outs = None
        import gc
        gc.collect()
        torch.cuda.synchronize()

    def _seed_base(self) -> int:
        return (self.seed * 0x1000003 + self.rank * 0x10001 + 1) & _MASK64

    def _dropout_seeds(self, pass_idx: int):
        return step_dropout_seeds(self.seed, self.rank, self.step_count, pass_idx)

    def _allreduce(self, flat, async_op=False, group=None):
        if self.world == 1:
            return None
        from .dp import allreduce_sum_
        return allreduce_sum_(flat, group if group is not None else self.pg, async_op)

    def _d_backward_allreduce(self, D):
        """Discriminator backward with its gradient all-reduce (data parallel): the tail of the arena (conv6, conv7,
        Dense: 19 of 24.6 MB) is reduced asynchronously as soon as it is final, behind the conv5..conv1 backward
        kernels; only the 5.5 MB head is reduced after the last kernel."""
        if self.world == 1:
            D.backward(need_wgrad=True, need_input_grad=False)
            return
        split = D.arena.offsets[D.EARLY_GRAD_FIRST_VAR]
        pending = []
        # off by default since round 2: the timeline showed the overlapped weight-gradient kernels losing more (two waves:
        # +71 / +44 us) than the 19 MB reduction saves when it is hidden
        early = ((not D.spectral_norm) and not torch.cuda.is_current_stream_capturing()
                 and os.environ.get("B2G_DP_EARLY_ALLREDUCE", "0") != "0")
        D.backward(need_wgrad=True, need_input_grad=False,
                   on_tail_grads_ready=(lambda: pending.append(self._allreduce(D.arena.g[split:], async_op=True)))
                   if early else None)
        if pending:
            self._allreduce(D.arena.g[:split])
            for w in pending:
                if w is not None:
                    w.wait()
        else:
            self._allreduce(D.arena.g)

    def _d_update(self, D, gs, lr_dev):
        """tape.gradient + apply_gradients of the discriminator (SNDCGAN.py:256-257 / :266-267), data parallel included."""
        if self._nvls is not None:
            D.backward(need_wgrad=True, need_input_grad=False)
            self.d_optimizer.apply_arena_nvls(D.arena, self._nvls, gs, lr_dev)
            D._packed_dirty = True
            return
        self._d_backward_allreduce(D)
        self.d_optimizer.apply_model(D, gs, lr_dev)

    def train_step(self, input_real, input_z, dropout_seeds=None):
        """One G update + two D updates, in the reference's order (SNDCGAN.py:241-269).

        Returns (g_loss, d_loss, d_loss_real, d_loss_fake, d_logits_real, d_logits_fake) as CUDA tensors.
        dropout_seeds: optional (seeds_a, seeds_b, seeds_c), 7 ints each, for reproducible masks."""
        if self._graph_usable(input_real, input_z, dropout_seeds):
            if self._graph is None:
                self._capture_graph()
            if self._graph is not None:
                return self._train_step_graph(input_real, input_z)
        sa, sb, sc = dropout_seeds if dropout_seeds is not None else [self._dropout_seeds(i) for i in range(3)]
        out = self._step_body(input_real, input_z, sa, sb, sc, None)
        self.step_count += 1
        self._eager_steps += 1
        return out

    # -- CUDA-graph replay of the whole step ------------------------------------------------------------------
    def _graph_usable(self, input_real, input_z, dropout_seeds) -> bool:
        return (self.cuda_graph and self._graph_failed is None and dropout_seeds is None
                and (self.world == 1 or os.environ.get("B2G_CUDA_GRAPH_DP", "0") == "1")
                and _lib.profile_sink is None and self._eager_steps >= 2
                and input_real.shape[0] == self.batch_size and input_z.shape[0] == self.batch_size
                and input_real.dtype == torch.float32 and input_z.dtype == torch.float32
                and self.d_optimizer.iterations == 2 * self.step_count
                and self.g_optimizer.iterations == self.step_count
                and not torch.cuda.is_current_stream_capturing())

    def _capture_graph(self) -> None:
        lib, dev, B = _lib.load(), self.device, self.batch_size
        try:
            self._gs_ctr = torch.full((1,), self.step_count, dtype=torch.int64, device=dev)
            self._gs_ctr_host = self.step_count
            self._gs_seeds = torch.zeros(21, dtype=torch.int64, device=dev)
            self._gs_lr = torch.zeros(3, dtype=torch.float32, device=dev)
            self._g_real = torch.empty((B, *self.image_size), dtype=torch.float32, device=dev)
            self._g_z = torch.empty((B, self.z_size), dtype=torch.float32, device=dev)
            seeds = [[self._gs_seeds[p * 7 + i:p * 7 + i + 1] for i in range(7)] for p in range(3)]
            lrs = [self._gs_lr[k:k + 1] for k in range(3)]
            g_opt, d_opt = self.g_optimizer, self.d_optimizer
            side = torch.cuda.Stream(device=dev)
            side.wait_stream(torch.cuda.current_stream())
            l0 = _lib.launch_count
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.stream(side):
                with torch.cuda.graph(graph, stream=side):
                    check(lib.b2g_step_prologue(_p(self._gs_ctr), C.c_ulonglong(self._seed_base()),
                                                _p(self._gs_seeds), _p(self._gs_lr), g_opt.learning_rate,
                                                d_opt.learning_rate, d_opt.beta_1, d_opt.beta_2, _stream()),
                          "b2g_step_prologue")
                    outs = self._step_body(self._g_real, self._g_z, seeds[0], seeds[1], seeds[2], lrs)
            torch.cuda.current_stream().wait_stream(side)
            self._graph, self._graph_outs = graph, outs
            self._graph_launches = _lib.launch_count - l0
            _lib.launch_count = l0          # nothing ran during capture; replays add the count
        except Exception as exc:            # capture unsupported here: stay eager (same results, more launches)
            self._graph = None
            self._graph_failed = repr(exc)

    def _train_step_graph(self, input_real, input_z):
        if self._gs_ctr_host != self.step_count:     # eager steps ran in between: re-align the device step index
            self._gs_ctr.fill_(self.step_count)
        if input_real.data_ptr() != self._g_real.data_ptr():
            self._g_real.copy_(input_real)
        if input_z.data_ptr() != self._g_z.data_ptr():
            self._g_z.copy_(input_z)
        self._graph.replay()
        _lib.launch_count += self._graph_launches
        self.step_count += 1
        self._gs_ctr_host = self.step_count
        self.g_optimizer.iterations += 1
        self.d_optimizer.iterations += 2
        self.gen_model._packed_dirty = True
        self.disc_model._packed_dirty = True
        return tuple(o.clone() for o in self._graph_outs)   # the graph's own buffers are rewritten by the next replay

    @staticmethod
    def _nvtx(name=None):
        """NVTX ranges around the three phases of the step when B2G_NVTX=1 (nsys / ncu --nvtx timelines): _nvtx(name) opens
        a range, _nvtx() closes it.  No-op otherwise."""
        if os.environ.get("B2G_NVTX", "0") != "1":
            return
        if name is None:
            torch.cuda.nvtx.range_pop()
        else:
            torch.cuda.nvtx.range_push(name)

    def _step_body(self, input_real, input_z, sa, sb, sc, lr_dev):
        """The launches of one step.  sa/sb/sc: 7 dropout seeds per discriminator pass (ints, or 1-element device
        tensors under graph capture); lr_dev: None, or the three device-resident Adam rates (G, D first, D second)."""
        G, D = self.gen_model, self.disc_model
        # the reference draws batch_size latent vectors even for a ragged last real batch (SNDCGAN.py:283): the two
        # batches may differ in size, each loss is a mean over its own
        inv_b = 1.0 / input_z.shape[0]
        inv_b_real = 1.0 / input_real.shape[0]
        gs = 1.0 / self.world
        lr_g, lr_d1, lr_d2 = lr_dev if lr_dev is not None else (None, None, None)
        g_kind = LOSS_BCE if self.loss == "bce" else LOSS_HINGE_G
        d_kind = LOSS_BCE if self.loss == "bce" else LOSS_HINGE_D

        # (1) generator step (:242-249)
        self._nvtx("b2g:generator_step")
        g_output = G.forward(input_z)                                               # :243
        D.forward(g_output, training=True, dropout_seeds=sa)                        # :244
        g_loss, _ = D.head_loss(g_kind, 1.0, inv_b)                                 # :245-246
        d_gout = D.backward(need_wgrad=False, need_input_grad=True, g_out=g_output)
        G.backward(d_gout)                                                          # :248
        self._nvtx()
        g_work = self._allreduce(G.arena.g, async_op=True, group=self.pg_g)
        # G's Adam (:249) is issued at the end of the step: nothing below reads the generator's weights, so the
        # result is identical and its gradient all-reduce overlaps both discriminator steps.

        # (2) discriminator step on the real batch (:251-257)
        # While the generator's 164 MB all-reduce is in flight its NCCL kernel holds B2G_DP_G_CTAS SMs; a persistent grid of
        # one CTA per SM would wait for them (two waves), so every kernel launched in that window runs on that many fewer
        # SMs.  (B2G_DP_SM_RESERVE overrides the count; B2G_DP_G_WINDOW = 1: step (2) only, 2 (default): steps (2) and (3).)
        reserve = int(os.environ.get("B2G_DP_SM_RESERVE", str(self._g_ctas))) if g_work is not None else 0
        window = int(os.environ.get("B2G_DP_G_WINDOW", "2"))
        lib = _lib.load()
        if reserve:
            check(lib.b2g_set_sm_limit(max(8, lib.b2g_num_sms() - reserve)), "b2g_set_sm_limit", 0)
        try:
            self._nvtx("b2g:discriminator_step_real")
            D.forward(input_real, training=True, dropout_seeds=sb)                  # :252
            d_loss_real, d_logits_real = D.head_loss(d_kind, 1.0, inv_b_real)       # :253-254
            d_logits_real = d_logits_real.clone()
            self._d_update(D, gs, lr_d1)                                            # :256-257
            self._nvtx()
            if reserve and window < 2:
                check(lib.b2g_set_sm_limit(0), "b2g_set_sm_limit", 0)

            # (3) discriminator step on the stale fake batch (:259-267)
            self._nvtx("b2g:discriminator_step_fake")
            D.forward(g_output, training=True, dropout_seeds=sc)                    # :260
            d_loss_fake, d_logits_fake = D.head_loss(d_kind, 0.0, inv_b)            # :261-262
            d_logits_fake = d_logits_fake.clone()   # D.logits is a reused buffer: the caller gets its own tensor
            self._d_update(D, gs, lr_d2)                                            # :266-267
            self._nvtx()
        finally:
            if reserve:
                check(lib.b2g_set_sm_limit(0), "b2g_set_sm_limit", 0)

        if g_work is not None:
            g_work.wait()
        self.g_optimizer.apply_arena(G.arena, gs, lr_g)                             # :249 (deferred)
        G._packed_dirty = True

        d_loss = d_loss_fake + d_loss_real                                          # :264
        return (g_loss, d_loss, d_loss_real, d_loss_fake, d_logits_real, d_logits_fake)


def create_samples(g_model: Generator, input_z, batch_size):
    """sndcgan/generator_output.py:25-28: G(z, training=False) -> reshape (B,H,W,3) -> (x+1)/2."""
    g_output = g_model(input_z, training=False)
    images = g_output.reshape(batch_size, *g_model.output_size)
    return (images + 1) / 2.0


class DeviceFeeder:
    """Double-buffered host -> device staging of (input_real, input_z) batches on a copy stream — the analogue of
    the reference's `train_ds.prefetch(AUTOTUNE)` (SNDCGAN.py:143): batch i+1 is copied while step i computes."""

    def __init__(self, batch_size, image_size=(144, 256, 3), z_size=Z_SIZE, device="cuda"):
        self.device = torch.device(device)
        self.copy_stream = torch.cuda.Stream(device=self.device)
        self.real = [torch.empty((batch_size, *image_size), dtype=torch.float32, device=self.device) for _ in range(2)]
        self.z = [torch.empty((batch_size, z_size), dtype=torch.float32, device=self.device) for _ in range(2)]
        self.ready = [torch.cuda.Event(), torch.cuda.Event()]
        self.free = [torch.cuda.Event(), torch.cuda.Event()]
        self.put_idx = 0
        self.get_idx = 0
        self.bytes_per_batch = self.real[0].numel() * 4 + self.z[0].numel() * 4

    def put(self, real_host: torch.Tensor, z_host: torch.Tensor) -> None:
        """Start the asynchronous copy of one (pinned) host batch into the next staging slot."""
        i = self.put_idx
        self.put_idx ^= 1
        with torch.cuda.stream(self.copy_stream):
            self.copy_stream.wait_event(self.free[i])      # the step that last used this slot has been issued
            self.real[i].copy_(real_host, non_blocking=True)
            self.z[i].copy_(z_host, non_blocking=True)
            self.ready[i].record(self.copy_stream)

    def get(self):
        """Device tensors of the oldest pending batch; the compute stream waits for its copy."""
        i = self.get_idx
        self.get_idx ^= 1
        torch.cuda.current_stream().wait_event(self.ready[i])
        return self.real[i], self.z[i], i

    def release(self, i: int) -> None:
        self.free[i].record(torch.cuda.current_stream())
