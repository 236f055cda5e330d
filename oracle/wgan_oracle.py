"""ORACLE — TEST INFRASTRUCTURE ONLY (imported by tests/ and bench.py's cpu_baseline / reference arm).

CPU restatement (plain PyTorch autograd, fp32/fp64) of the reference's WGAN path, SURVEY.md section 8(f) row N1:
  define_critic      wasserstein_gan/WGAN.py:53-101   (7 x [Conv2D + bias, BatchNormalization, LeakyReLU], Dense(1))
  define_generator   wasserstein_gan/WGAN.py:105-134  (Dense, LeakyReLU, 3 x [Conv2DTranspose, BN, LeakyReLU], Conv2D tanh)
  define_gan         wasserstein_gan/WGAN.py:138-152  (critic conv/dense frozen, critic BatchNorm gamma/beta trainable)
  train loop         wasserstein_gan/WGAN.py:296-318  (critic on real y=-1, critic on predicted fakes y=+1, every
                                                       critic_learn_iterations batches one generator update y=-1)
Keras-2 semantics assumed (TensorFlow is not installable here ==> PARITY UNPINNED):
  * train_on_batch runs the model with training=True: BatchNormalization normalises with the batch mean and the
    BIASED batch variance, eps 1e-3, and moves moving_mean / moving_variance with momentum 0.99; on the fused
    4-D path the moving variance is fed the Bessel-corrected variance (n/(n-1)).
  * model.predict runs BatchNormalization on the moving statistics (WGAN.py:224).
  * RespectCompiledTrainableState: critic.train_on_batch updates every critic variable although define_gan later
    set the conv/dense layers untrainable; gan_model.train_on_batch updates the generator and the critic's gamma/beta.
  * RMSprop(learning_rate=5e-5): rho 0.9, momentum 0, epsilon 1e-7, not centred:
        rms <- rho rms + (1-rho) g^2 ;  w <- w - lr g / (sqrt(rms) + eps)        (WGAN.py:99,150; two optimizers)
  * ClipConstraint(0.01) clips the 7 critic conv KERNELS (not biases, not Dense) after each critic update (WGAN.py:33-44,57).
  * wasserstein_loss = mean(y_true * y_pred) (WGAN.py:48-49).
  * kernel_initializer RandomNormal(stddev=0.02) on the conv / conv-transpose kernels; Dense layers keep glorot_uniform.
"""
from __future__ import annotations

import torch

from . import keras_ops as K
from .sndcgan_oracle import _qa, _qw

BN_EPS = 1e-3
BN_MOMENTUM = 0.99
CLIP = 0.01
LATENT = 128
CRITIC_CONVS = [(3, 1, 3, 64, 0.2), (4, 2, 64, 128, 0.2), (3, 1, 128, 128, 0.2), (4, 2, 128, 256, 0.2),
                (3, 1, 256, 256, 0.2), (4, 2, 256, 512, 0.1), (3, 1, 512, 512, 0.1)]   # k, stride, cin, cout, alpha
GEN_ALPHA = 0.2


def critic_var_shapes(image_size=(144, 256, 3)):
    """critic.trainable_variables in Keras order (WGAN.py:60-96)."""
    h, w, _ = image_size
    out = []
    for i, (k, _s, cin, cout, _a) in enumerate(CRITIC_CONVS):
        out += [(f"conv{i + 1}/kernel", (k, k, cin, cout)), (f"conv{i + 1}/bias", (cout,)),
                (f"bn{i + 1}/gamma", (cout,)), (f"bn{i + 1}/beta", (cout,))]
    out += [("dense/kernel", (512 * (h // 8) * (w // 8), 1)), ("dense/bias", (1,))]
    return out


def generator_var_shapes(image_size=(144, 256, 3)):
    """generator.trainable_variables in Keras order (WGAN.py:112-132)."""
    h, w, _ = image_size
    f = 512 * (h // 8) * (w // 8)
    return [("dense/kernel", (LATENT, f)),
            ("convt1/kernel", (4, 4, 256, 512)), ("bn1/gamma", (256,)), ("bn1/beta", (256,)),
            ("convt2/kernel", (4, 4, 128, 256)), ("bn2/gamma", (128,)), ("bn2/beta", (128,)),
            ("convt3/kernel", (4, 4, 64, 128)), ("bn3/gamma", (64,)), ("bn3/beta", (64,)),
            ("conv4/kernel", (3, 3, 64, 3))]


def init_vars(shapes, seed, dtype=torch.float32):
    gen = torch.Generator().manual_seed(seed)
    out = []
    for name, s in shapes:
        if name.startswith("dense") and name.endswith("kernel"):
            out.append(K.glorot_uniform(s, gen, dtype))
        elif name.endswith("kernel"):
            out.append((torch.randn(s, generator=gen, dtype=torch.float64) * 0.02).to(dtype))
        elif name.endswith("gamma"):
            out.append(torch.ones(s, dtype=dtype))
        else:
            out.append(torch.zeros(s, dtype=dtype))
    return out


def batchnorm(x, gamma, beta, moving, training: bool, eps=BN_EPS, momentum=BN_MOMENTUM):
    """BatchNormalization(axis=-1) on NHWC.  moving = [mean, var] (updated in place when training)."""
    if not training:
        return gamma * (x - moving[0]) / torch.sqrt(moving[1] + eps) + beta
    c = x.shape[-1]
    flat = x.reshape(-1, c)
    n = flat.shape[0]
    mean = flat.mean(0)
    var = ((flat - mean) ** 2).mean(0)
    with torch.no_grad():
        moving[0].mul_(momentum).add_(mean.detach() * (1 - momentum))
        moving[1].mul_(momentum).add_(var.detach() * (n / max(n - 1, 1)) * (1 - momentum))
    return gamma * (x - mean) / torch.sqrt(var + eps) + beta


def critic_forward(cv, moving, x, training: bool, emu=False, collect=None):
    """cv: variables in critic_var_shapes order; moving: 7 x [mean, var]; x NHWC.  Returns logits [B,1]."""
    a = x
    for i, (_k, s, _ci, _co, alpha) in enumerate(CRITIC_CONVS):
        w, b, g, be = cv[4 * i:4 * i + 4]
        if emu:   # the CUDA path stores the conv output WITHOUT the bias in bf16 (BatchNorm cancels the bias exactly;
            raw = _qa(K.conv2d_same(a if i else _qa(a, emu), _qw(w, emu), None, s), emu) + b   # only the moving mean sees it)
        else:
            raw = K.conv2d_same(a, w, b, s)
        a = _qa(K.leaky_relu(batchnorm(raw, g, be, moving[i], training), alpha), emu)
        if collect is not None:
            collect[f"raw{i + 1}"], collect[f"c{i + 1}"] = raw, a
    return K.dense(a.reshape(a.shape[0], -1), cv[28], cv[29])


def generator_forward(gv, moving, z, training: bool, image_size=(144, 256, 3), emu=False, collect=None):
    h, w, _ = image_size
    a = _qa(K.leaky_relu(K.dense(z, gv[0]), GEN_ALPHA), emu).reshape(z.shape[0], h // 8, w // 8, 512)
    for i in range(3):
        wt, g, be = gv[1 + 3 * i:4 + 3 * i]
        raw = _qa(K.conv2d_transpose_same(a, _qw(wt, emu), 2), emu)
        a = _qa(K.leaky_relu(batchnorm(raw, g, be, moving[i], training), GEN_ALPHA), emu)
        if collect is not None:
            collect[f"graw{i + 1}"], collect[f"a{i + 1}"] = raw, a
    return torch.tanh(K.conv2d_same(a, _qw(gv[10], emu), None, 1))


def wasserstein_loss(y_true: float, y_pred):
    return (y_true * y_pred).mean()


def rmsprop_step(w, g, rms, lr=5e-5, rho=0.9, eps=1e-7):
    rms.mul_(rho).add_(g * g * (1 - rho))
    w.sub_(lr * g / (torch.sqrt(rms) + eps))


class OracleWGAN:
    """State + the three train_on_batch flavours of WGAN.train (WGAN.py:296-318)."""

    def __init__(self, image_size=(144, 256, 3), seed=62, lr=5e-5, dtype=torch.float32, emu=False):
        self.image_size, self.lr, self.emu, self.dtype = tuple(image_size), lr, emu, dtype
        self.cv = init_vars(critic_var_shapes(image_size), seed + 1, dtype)
        self.gv = init_vars(generator_var_shapes(image_size), seed, dtype)
        self.c_moving = [[torch.zeros(c[3], dtype=dtype), torch.ones(c[3], dtype=dtype)] for c in CRITIC_CONVS]
        self.g_moving = [[torch.zeros(c, dtype=dtype), torch.ones(c, dtype=dtype)] for c in (256, 128, 64)]
        self.c_rms = [torch.zeros_like(v) for v in self.cv]             # critic_model's optimizer (WGAN.py:99)
        self.gan_rms_g = [torch.zeros_like(v) for v in self.gv]         # gan_model's optimizer (WGAN.py:150)
        self.gan_rms_cbn = {i: torch.zeros_like(self.cv[i]) for i in self.critic_bn_indices()}

    @staticmethod
    def critic_bn_indices():
        return [4 * i + j for i in range(7) for j in (2, 3)]

    def train_critic_on_batch(self, x, label: float, collect=None):
        """critic_model.train_on_batch(x, label * ones) (WGAN.py:298,303).  Returns the loss."""
        vs = [v.clone().requires_grad_(True) for v in self.cv]
        logits = critic_forward(vs, self.c_moving, x.to(self.dtype), True, self.emu, collect)
        loss = wasserstein_loss(label, logits)
        grads = torch.autograd.grad(loss, vs)
        if collect is not None:
            collect["grads"], collect["logits"] = grads, logits.detach()
        with torch.no_grad():
            for i, (v, g) in enumerate(zip(self.cv, grads)):
                rmsprop_step(v, g, self.c_rms[i], self.lr)
                if i < 28 and i % 4 == 0:
                    v.clamp_(-CLIP, CLIP)
        return loss.detach()

    def predict(self, z):
        """generator_model.predict(z) (WGAN.py:224): BatchNorm on the moving statistics."""
        with torch.no_grad():
            return generator_forward(self.gv, self.g_moving, z.to(self.dtype), False, self.image_size, self.emu)

    def train_gan_on_batch(self, z, collect=None):
        """gan_model.train_on_batch(z, -ones) (WGAN.py:317)."""
        gvs = [v.clone().requires_grad_(True) for v in self.gv]
        cvs = [v.clone().requires_grad_(i in self.gan_rms_cbn) for i, v in enumerate(self.cv)]
        fake = generator_forward(gvs, self.g_moving, z.to(self.dtype), True, self.image_size, self.emu, collect)
        logits = critic_forward(cvs, self.c_moving, fake, True, self.emu, collect)
        loss = wasserstein_loss(-1.0, logits)
        idx = self.critic_bn_indices()
        grads = torch.autograd.grad(loss, gvs + [cvs[i] for i in idx])
        if collect is not None:
            collect["g_grads"], collect["cbn_grads"], collect["fake"] = grads[:len(gvs)], grads[len(gvs):], fake.detach()
        with torch.no_grad():
            for v, g, r in zip(self.gv, grads[:len(gvs)], self.gan_rms_g):
                rmsprop_step(v, g, r, self.lr)
            for i, g in zip(idx, grads[len(gvs):]):
                rmsprop_step(self.cv[i], g, self.gan_rms_cbn[i], self.lr)
        return loss.detach()

    def train_iteration(self, reals, zs_fake, z_gan):
        """critic_learn_iterations x (critic on real, critic on predicted fakes) then one generator update.
        Returns (mean real loss, mean fake loss, generator loss) like the c1/c2/g history entries."""
        c1, c2 = [], []
        for x, z in zip(reals, zs_fake):
            c1.append(self.train_critic_on_batch(x, -1.0))
            c2.append(self.train_critic_on_batch(self.predict(z), 1.0))
        g = self.train_gan_on_batch(z_gan)
        return torch.stack(c1).mean(), torch.stack(c2).mean(), g
