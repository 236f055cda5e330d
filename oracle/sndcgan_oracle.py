"""ORACLE — TEST INFRASTRUCTURE ONLY (tests/, __graft_entry__.smoke(), bench.py cpu_baseline / --impl reference).

Line-by-line CPU restatement (PyTorch, fp32 or fp64) of the reference's SNDCGAN hot path:
  make_dcgan_generator      sndcgan/SNDCGAN.py:25-66
  make_dcgan_discriminator  sndcgan/SNDCGAN.py:69-128
  optimizers + loss         sndcgan/SNDCGAN.py:165-170
  SNDCGAN.train_step        sndcgan/SNDCGAN.py:241-269
  create_samples            sndcgan/generator_output.py:25-28
The arithmetic of those lines lives in TensorFlow/Keras 2.x (not vendored, not installable here, no version
pinned by the reference) and the reference holds no tests / golden vectors for it ==> PARITY UNPINNED at the
TensorFlow boundary; Keras semantics are taken from its documentation (see keras_ops.py).  Dropout masks are
explicit inputs because TF's RNG stream cannot be reproduced.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import torch

from . import keras_ops as K

Z_SIZE = 128  # sndcgan/Trainer.py:8


# ----------------------------------------------------------------------------------------------------------
# parameters, in Keras `trainable_variables` order
# ----------------------------------------------------------------------------------------------------------
def generator_var_shapes(image_size=(144, 256, 3)) -> list[tuple[str, tuple[int, ...]]]:
    h, w = image_size[0] // 8, image_size[1] // 8                      # SNDCGAN.py:26
    feat = 512 * h * w
    return [
        ("dense/kernel", (Z_SIZE, feat)),                               # :31 (use_bias=False)
        ("bn0/gamma", (feat,)), ("bn0/beta", (feat,)),                  # :32
        ("convt1/kernel", (4, 4, 256, 512)),                            # :36-39  [kh,kw,Cout,Cin]
        ("bn1/gamma", (256,)), ("bn1/beta", (256,)),                    # :40
        ("convt2/kernel", (4, 4, 128, 256)),                            # :43-46
        ("bn2/gamma", (128,)), ("bn2/beta", (128,)),                    # :47
        ("convt3/kernel", (4, 4, 64, 128)),                             # :50-53
        ("bn3/gamma", (64,)), ("bn3/beta", (64,)),                      # :54
        ("convt4/kernel", (3, 3, image_size[2], 64)),                   # :57-61
    ]


D_CONVS = [  # (name, k, stride, Cin, Cout)  SNDCGAN.py:73-120
    ("conv1", 3, 1, 3, 64), ("conv2", 4, 2, 64, 128), ("conv3", 3, 1, 128, 128), ("conv4", 4, 2, 128, 256),
    ("conv5", 3, 1, 256, 256), ("conv6", 4, 2, 256, 512), ("conv7", 3, 1, 512, 512),
]


def discriminator_var_shapes(image_size=(144, 256, 3)) -> list[tuple[str, tuple[int, ...]]]:
    out = []
    for name, k, _s, cin, cout in D_CONVS:
        cin = image_size[2] if name == "conv1" else cin
        out += [(f"{name}/kernel", (k, k, cin, cout)), (f"{name}/bias", (cout,))]
    feat = 512 * (image_size[0] // 8) * (image_size[1] // 8)
    out += [("dense/kernel", (feat, 1)), ("dense/bias", (1,))]       # :122-123
    return out


def init_vars(shapes, seed: int, dtype=torch.float32) -> list[torch.Tensor]:
    """Keras default initialisers: glorot_uniform kernels, zero biases, BN gamma=1 / beta=0."""
    gen = torch.Generator().manual_seed(seed)
    out = []
    for name, shape in shapes:
        if name.endswith("kernel"):
            out.append(K.glorot_uniform(shape, gen, dtype))
        elif name.endswith("gamma"):
            out.append(torch.ones(shape, dtype=dtype))
        else:
            out.append(torch.zeros(shape, dtype=dtype))
    return out


# ----------------------------------------------------------------------------------------------------------
# optional emulation of the CUDA path's storage precision (bf16 activations / activation gradients, bf16
# packed conv weights with the BatchNorm / spectral-norm scale folded in, fp32 accumulation and master weights).
# With emu=False (default) the oracle is the plain fp32/fp64 restatement of the reference.
# ----------------------------------------------------------------------------------------------------------
class _QAct(torch.autograd.Function):
    """round to bf16 in forward (stored activation) and in backward (stored activation gradient)"""

    @staticmethod
    def forward(ctx, x):
        return x.to(torch.bfloat16).to(x.dtype)

    @staticmethod
    def backward(ctx, g):
        return g.to(torch.bfloat16).to(g.dtype)


class _QW(torch.autograd.Function):
    """round a (folded) weight to bf16 in forward; straight-through in backward (fp32 weight gradients)"""

    @staticmethod
    def forward(ctx, w):
        return w.to(torch.bfloat16).to(w.dtype)

    @staticmethod
    def backward(ctx, g):
        return g


class _QTf32(torch.autograd.Function):
    """what tcgen05 kind::tf32 reads of an fp32 operand: the low 13 mantissa bits are dropped (truncation);
    straight-through in backward.  The CUDA path's generator stem (b2g_dense_tc.cu) multiplies z and the fp32
    dense weight this way in forward; its backward is computed from the unrounded values."""

    @staticmethod
    def forward(ctx, x):
        x32 = x.to(torch.float32).contiguous()
        return (x32.view(torch.int32) & ~0x1FFF).view(torch.float32).to(x.dtype)

    @staticmethod
    def backward(ctx, g):
        return g


def _qa(x, emu):
    return _QAct.apply(x) if emu else x


def _qw(w, emu):
    return _QW.apply(w) if emu else w


# ----------------------------------------------------------------------------------------------------------
# spectral normalisation state (north-star option, NOT in the reference: SURVEY F1)
# ----------------------------------------------------------------------------------------------------------
class SNState:
    """Persistent u (and the matching v) of the 8 discriminator matrices.  Miyato et al. (2018) Alg. 1: ONE power
    iteration per weight update — it runs when the weights changed since the last one (`fresh` False: initial
    weights, after every optimizer step), not once per forward pass; passes that see the same weights share sigma."""

    def __init__(self, u):
        self.u = list(u)
        self.v = [None] * len(self.u)
        self.fresh = False

    def refresh(self, dv):
        if self.fresh:
            return
        with torch.no_grad():
            for i in range(len(self.u)):
                wt = dv[2 * i]
                _s, u, v = K.spectral_norm_power_iteration(wt.reshape(-1, wt.shape[-1]), self.u[i].to(wt.dtype), 1)
                self.u[i], self.v[i] = u, v
        self.fresh = True


# ----------------------------------------------------------------------------------------------------------
# forward passes
# ----------------------------------------------------------------------------------------------------------
def generator_forward(gv: list[torch.Tensor], z: torch.Tensor, image_size=(144, 256, 3), collect: dict | None = None,
                      emu: bool = False):
    """SNDCGAN.py:28-61 with BatchNorm in INFERENCE mode (moving mean 0 / var 1): train_step calls
    self.gen_model(input_z) without training=True (:243) and create_samples passes training=False."""
    h, w = image_size[0] // 8, image_size[1] // 8
    x = K.dense(_QTf32.apply(z), _QTf32.apply(gv[0])) if emu else K.dense(z, gv[0])
    x = _qa(torch.relu(K.batchnorm_inference(x, gv[1], gv[2])), emu)
    x = x.reshape(-1, h, w, 512)                                        # :34 row-major NHWC
    if collect is not None:
        collect["g_a0"] = x
    for i, (kidx, stride) in enumerate([(3, 2), (6, 2), (9, 2)]):
        if emu:  # same math, BatchNorm scale folded into the bf16 weight like the CUDA path
            gs = gv[kidx + 1] / math.sqrt(1.0 + K.BN_EPS)
            x = K.conv2d_transpose_same(x, _qw(gv[kidx] * gs[None, None, :, None], True), stride) + gv[kidx + 2]
            x = _qa(torch.relu(x), True)
        else:
            x = K.conv2d_transpose_same(x, gv[kidx], stride)
            x = torch.relu(K.batchnorm_inference(x, gv[kidx + 1], gv[kidx + 2]))
        if collect is not None:
            collect[f"g_a{i + 1}"] = x
    x = torch.tanh(K.conv2d_transpose_same(x, gv[12], 1))              # :57-61
    if collect is not None:
        collect["g_out"] = x
    return x


def discriminator_forward(dv: list[torch.Tensor], x: torch.Tensor, rate: float, masks: list | None,
                          collect: dict | None = None, sn_state: list | None = None, tag: str = "d",
                          emu: bool = False):
    """SNDCGAN.py:70-124 with training=True dropout given as explicit keep masks (one per conv, or None).

    sn_state (optional, NOT in the reference): an SNState (or a list of 8 `u` vectors, wrapped into a stale one);
    every kernel is divided by sigma = v^T W u with u, v from the state's power iteration (constants for autograd)."""
    if isinstance(sn_state, list):
        sn_state = SNState(sn_state)
    if sn_state is not None:
        sn_state.refresh(dv)

    def eff(i, wt):
        if sn_state is None:
            return wt
        w2d = wt.reshape(-1, wt.shape[-1])
        u, v = sn_state.u[i], sn_state.v[i]
        sigma = v @ (w2d @ u)
        if collect is not None:
            collect[f"{tag}_sigma{i}"] = sigma.detach()
        return wt / sigma

    for i, (_name, _k, stride, _ci, _co) in enumerate(D_CONVS):
        x = K.conv2d_same(x, _qw(eff(i, dv[2 * i]), emu and i > 0), dv[2 * i + 1], stride)
        x = K.leaky_relu(x)
        x = _qa(K.dropout(x, None if masks is None else masks[i], rate), emu)
        if collect is not None:
            collect[f"{tag}_c{i + 1}"] = x
    x = x.reshape(x.shape[0], -1)                                       # Flatten :122
    return K.dense(x, eff(7, dv[14]), dv[15])                          # Dense(1) :123


def create_samples(gv, z, batch_size, image_size=(144, 256, 3)):
    """sndcgan/generator_output.py:25-28."""
    with torch.no_grad():
        g = generator_forward(gv, z, image_size)
    return (g.reshape(batch_size, *image_size) + 1) / 2.0


# ----------------------------------------------------------------------------------------------------------
# losses
# ----------------------------------------------------------------------------------------------------------
def g_loss_fn(kind: str, d_fake_logits):
    if kind == "bce":
        return K.bce_with_logits(d_fake_logits, 1.0)                    # SNDCGAN.py:245-246
    return -d_fake_logits.mean()                                        # hinge GAN generator loss (option)


def d_loss_fn(kind: str, logits, label: float):
    if kind == "bce":
        return K.bce_with_logits(logits, label)                         # :254, :262
    return K.hinge_d_loss(logits, label)


# ----------------------------------------------------------------------------------------------------------
# the training step
# ----------------------------------------------------------------------------------------------------------
@dataclass
class AdamState:
    lr: float
    m: list = field(default_factory=list)
    v: list = field(default_factory=list)
    iterations: int = 0

    def apply(self, variables: list[torch.Tensor], grads: list[torch.Tensor]) -> None:
        if not self.m:
            self.m = [torch.zeros_like(p) for p in variables]
            self.v = [torch.zeros_like(p) for p in variables]
        self.iterations += 1
        with torch.no_grad():
            for p, g, m, v in zip(variables, grads, self.m, self.v):
                K.adam_step(p, g, m, v, self.iterations, self.lr)


class OracleSNDCGAN:
    """State + train_step of the reference's SNDCGAN class (model/optimizer part of SNDCGAN.py:149-170)."""

    def __init__(self, image_size=(144, 256, 3), dropout=0.5, lr_d=2e-4, lr_g=2e-4, seed=62,
                 dtype=torch.float32, loss="bce", spectral_norm=False, emu=False):
        self.emu = emu
        self.image_size = tuple(image_size)
        self.dropout = dropout
        self.loss = loss
        self.dtype = dtype
        self.gv = init_vars(generator_var_shapes(image_size), seed, dtype)
        self.dv = init_vars(discriminator_var_shapes(image_size), seed + 1, dtype)
        self.g_opt = AdamState(lr_g)
        self.d_opt = AdamState(lr_d)
        self.sn_state = None
        if spectral_norm:
            gen = torch.Generator().manual_seed(seed + 2)
            self.sn_state = []
            for i in range(8):
                cout = self.dv[2 * i].shape[-1]
                u = torch.randn(cout, generator=gen, dtype=torch.float64).to(dtype)
                self.sn_state.append(u / u.norm())

    def train_step(self, input_real, input_z, masks_a=None, masks_b=None, masks_c=None, collect: dict | None = None):
        """SNDCGAN.py:241-269.  masks_a/b/c: the 7 dropout keep-masks of the three D passes."""
        gv, dv = self.gv, self.dv
        if isinstance(self.sn_state, list):
            self.sn_state = SNState(self.sn_state)
        for p in gv + dv:
            p.requires_grad_(True)
        # (1) generator step, :242-249
        g_output = generator_forward(gv, input_z, self.image_size, collect, self.emu)
        d_logits_fake = discriminator_forward(dv, g_output, self.dropout, masks_a, collect, self.sn_state, "da",
                                              self.emu)
        g_loss = g_loss_fn(self.loss, d_logits_fake)
        g_grads = torch.autograd.grad(g_loss, gv)
        g_output = g_output.detach()
        if collect is not None:
            collect["g_grads"] = [g.clone() for g in g_grads]
        self.g_opt.apply(gv, g_grads)
        # (2) discriminator step on real, :251-257
        d_logits_real = discriminator_forward(dv, input_real, self.dropout, masks_b, collect, self.sn_state, "db",
                                              self.emu)
        d_loss_real = d_loss_fn(self.loss, d_logits_real, 1.0)
        d_grads1 = torch.autograd.grad(d_loss_real, dv)
        if collect is not None:
            collect["d_grads1"] = [g.clone() for g in d_grads1]
        self.d_opt.apply(dv, d_grads1)
        if self.sn_state is not None:
            self.sn_state.fresh = False
        # (3) discriminator step on the stale fake batch, :259-267
        d_logits_fake2 = discriminator_forward(dv, g_output, self.dropout, masks_c, collect, self.sn_state, "dc",
                                               self.emu)
        d_loss_fake = d_loss_fn(self.loss, d_logits_fake2, 0.0)
        d_loss = d_loss_fake + d_loss_real                               # :264
        d_grads2 = torch.autograd.grad(d_loss_fake, dv)
        if collect is not None:
            collect["d_grads2"] = [g.clone() for g in d_grads2]
        self.d_opt.apply(dv, d_grads2)
        if self.sn_state is not None:
            self.sn_state.fresh = False
        for p in gv + dv:
            p.requires_grad_(False)
        return (g_loss.detach(), d_loss.detach(), d_loss_real.detach(), d_loss_fake.detach(),
                d_logits_real.detach(), d_logits_fake2.detach())                       # :269


def flops_per_image(image_size=(144, 256, 3)) -> dict:
    """Algorithmic FLOPs (2*MAC) of G and D forward per image."""
    h, w = image_size[0], image_size[1]
    g = Z_SIZE * 512 * (h // 8) * (w // 8)
    hh, ww, cin = h // 8, w // 8, 512
    for cout in (256, 128, 64):
        g += hh * ww * cin * cout * 16
        hh, ww, cin = hh * 2, ww * 2, cout
    g += hh * ww * 64 * image_size[2] * 9
    d = 0
    hh, ww = h, w
    for name, k, s, ci, co in D_CONVS:
        ci = image_size[2] if name == "conv1" else ci
        hh, ww = hh // s, ww // s
        d += hh * ww * ci * co * k * k
    d += 512 * (h // 8) * (w // 8)
    return {"g_fwd": 2 * g, "d_fwd": 2 * d, "train_step": 2 * (3 * g + 8 * d)}
