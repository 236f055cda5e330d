"""ORACLE — TEST INFRASTRUCTURE ONLY (tests/, tools/cyclegan_bench.py CPU leg).

CPU restatement (PyTorch, fp32 / fp64) of the reference's CycleGAN path, SURVEY.md section 8(f) row N2:
  ReflectionPadding2D     cyclegan/CycleGAN.py:37-57
  ResBlock                cyclegan/CycleGAN.py:60-92   (conv, norm, relu, conv, ADD, relu, norm — in that order)
  k_conv / discriminator  cyclegan/CycleGAN.py:95-126  (4x4 s2 VALID convs, PatchGAN output [N,3,3,1] at 128x128)
  conv_c7_s1/d_conv/u_conv cyclegan/CycleGAN.py:129-157
  define_generator        cyclegan/CycleGAN.py:161-183
  losses                  cyclegan/CycleGAN.py:186-208 (BCE from logits, L1 cycle x10, L1 identity x5)
  train_step              cyclegan/CycleGAN.py:325-382 (6 generator + 4 discriminator forwards, 4 Adam(2e-4, beta_1 0.5))

PARITY UNPINNED: the arithmetic lives in TensorFlow/Keras 2.x and tensorflow_addons (InstanceNormalization), neither
vendored nor installable here, no versions pinned by the reference, no tests or vectors in the reference.  Semantics
restated from their documentation:
  * `InstanceNormalization(axis=1)` on an NHWC tensor is tfa GroupNormalization(groups=-1, axis=1): the "groups" are the H
    image ROWS — mean / variance over (W, C) per (n, h), epsilon 1e-3, gamma / beta of length H indexed by the row, both
    initialised U(-0.05, 0.05) ("random_uniform").  (per image row, not per channel.)
  * Conv2D 'valid' / 'same' (7x7 s1 'same' = symmetric pad 3), ReflectionPadding2D((1,1)) = tf.pad REFLECT by 1 pixel;
  * Conv2DTranspose 3x3 s2 'same': output 2H x 2W, y[i] = sum_{o,k: i = 2 o + k} x[o] w[k] (the gradient of a SAME conv whose
    single padding row/column sits at the END) = torch conv_transpose2d(padding=0) cropped to the first 2H x 2W
    (checked against autograd of that conv in tests/test_cyclegan_oracle.py);
  * kernels glorot_uniform, biases zero; Adam as in keras_ops.adam_step with beta_1 = 0.5.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

from . import keras_ops as K

IN_EPS = 1e-3        # tfa GroupNormalization default epsilon
LAMBDA = 10.0        # cyclegan/CycleGAN.py:186
LR, BETA1 = 2e-4, 0.5  # :229-233


# ----------------------------------------------------------------------------------------------------------
# layers
# ----------------------------------------------------------------------------------------------------------
def row_norm(x: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor, eps: float = IN_EPS) -> torch.Tensor:
    """InstanceNormalization(axis=1) on NHWC: statistics over (W, C) per (n, h); gamma, beta [H]."""
    mean = x.mean(dim=(2, 3), keepdim=True)
    var = x.var(dim=(2, 3), unbiased=False, keepdim=True)
    return (x - mean) / torch.sqrt(var + eps) * gamma.view(1, -1, 1, 1) + beta.view(1, -1, 1, 1)


def conv2d(x, w, b, stride=1, pad=0):
    """x NHWC, w [kh,kw,Cin,Cout] (Keras Conv2D), explicit symmetric zero padding."""
    y = F.conv2d(x.permute(0, 3, 1, 2), w.permute(3, 2, 0, 1), b, stride=stride, padding=pad)
    return y.permute(0, 2, 3, 1)


def reflect_pad(x, p=1):
    return F.pad(x.permute(0, 3, 1, 2), (p, p, p, p), mode="reflect").permute(0, 2, 3, 1)


def conv2d_transpose_k3s2_same(x, w, b):
    """Keras Conv2DTranspose(3x3, strides 2, 'same'): w [kh,kw,Cout,Cin]."""
    n, h, wd, _ = x.shape
    y = F.conv_transpose2d(x.permute(0, 3, 1, 2), w.permute(3, 2, 0, 1), b, stride=2, padding=0)
    return y[:, :, :2 * h, :2 * wd].permute(0, 2, 3, 1)


# ----------------------------------------------------------------------------------------------------------
# variables, in Keras `trainable_variables` order
# ----------------------------------------------------------------------------------------------------------
def generator_var_shapes(image_size=(128, 128, 3)):
    h = image_size[0]
    out = [("c7s1_in/kernel", (7, 7, image_size[2], 64)), ("c7s1_in/bias", (64,)),
           ("c7s1_in/gamma", (h,)), ("c7s1_in/beta", (h,)),
           ("d1/kernel", (3, 3, 64, 128)), ("d1/bias", (128,)), ("d1/gamma", (h // 2,)), ("d1/beta", (h // 2,)),
           ("d2/kernel", (3, 3, 128, 256)), ("d2/bias", (256,)), ("d2/gamma", (h // 4,)), ("d2/beta", (h // 4,))]
    for i in range(9):                                                     # :166-174
        for j in (1, 2):
            out += [(f"res{i}/conv{j}/kernel", (3, 3, 256, 256)), (f"res{i}/conv{j}/bias", (256,)),
                    (f"res{i}/in{j}/gamma", (h // 4,)), (f"res{i}/in{j}/beta", (h // 4,))]
    out += [("u1/kernel", (3, 3, 128, 256)), ("u1/bias", (128,)), ("u1/gamma", (h // 2,)), ("u1/beta", (h // 2,)),
            ("u2/kernel", (3, 3, 64, 128)), ("u2/bias", (64,)), ("u2/gamma", (h,)), ("u2/beta", (h,)),
            ("c7s1_out/kernel", (7, 7, 64, image_size[2])), ("c7s1_out/bias", (image_size[2],)),
            ("c7s1_out/gamma", (h,)), ("c7s1_out/beta", (h,))]
    return out


def disc_sizes(h):
    """Row counts after the four 4x4 s2 VALID convs and the final 4x4 s1 VALID conv."""
    s = [h]
    for _ in range(4):
        s.append((s[-1] - 4) // 2 + 1)
    s.append(s[-1] - 3)
    return s


def discriminator_var_shapes(image_size=(128, 128, 3)):
    hs = disc_sizes(image_size[0])
    if hs[-1] < 1:
        raise ValueError("image too small for the PatchGAN discriminator (needs >= 94 pixels per side)")
    out = [("k1/kernel", (4, 4, image_size[2], 64)), ("k1/bias", (64,))]
    cin = 64
    for i, cout in enumerate((128, 256, 512)):
        out += [(f"k{i + 2}/kernel", (4, 4, cin, cout)), (f"k{i + 2}/bias", (cout,)),
                (f"k{i + 2}/gamma", (hs[i + 2],)), (f"k{i + 2}/beta", (hs[i + 2],))]
        cin = cout
    out += [("out/kernel", (4, 4, 512, 1)), ("out/bias", (1,))]
    return out


def init_vars(shapes, seed, dtype=torch.float32):
    gen = torch.Generator().manual_seed(seed)
    out = []
    for name, shape in shapes:
        if name.endswith("kernel"):
            out.append(K.glorot_uniform(shape, gen, dtype))
        elif name.endswith("bias"):
            out.append(torch.zeros(shape, dtype=dtype))
        else:  # gamma / beta: keras "random_uniform" = U(-0.05, 0.05)
            out.append(((torch.rand(shape, generator=gen, dtype=torch.float64) - 0.5) * 0.1).to(dtype))
    return out


# ----------------------------------------------------------------------------------------------------------
# models
# ----------------------------------------------------------------------------------------------------------
def generator_forward(v, x, collect=None, tag="g"):
    """define_generator (cyclegan/CycleGAN.py:161-183); v in generator_var_shapes order."""
    def keep(name, t):
        if collect is not None:
            collect[f"{tag}_{name}"] = t
        return t
    a = torch.relu(row_norm(conv2d(x, v[0], v[1], 1, 3), v[2], v[3]))                       # conv_c7_s1(64)
    keep("c7", a)
    i = 4
    for name in ("d1", "d2"):                                                               # d_conv
        a = torch.relu(row_norm(conv2d(reflect_pad(a), v[i], v[i + 1], 2, 0), v[i + 2], v[i + 3]))
        keep(name, a)
        i += 4
    for r in range(9):                                                                      # ResBlock :84-92
        fx = conv2d(a, v[i], v[i + 1], 1, 1)
        fx = torch.relu(row_norm(fx, v[i + 2], v[i + 3]))
        fx = conv2d(fx, v[i + 4], v[i + 5], 1, 1)
        a = row_norm(torch.relu(a + fx), v[i + 6], v[i + 7])
        keep(f"res{r}", a)
        i += 8
    for name in ("u1", "u2"):                                                               # u_conv
        a = torch.relu(row_norm(conv2d_transpose_k3s2_same(a, v[i], v[i + 1]), v[i + 2], v[i + 3]))
        keep(name, a)
        i += 4
    a = torch.tanh(row_norm(conv2d(a, v[i], v[i + 1], 1, 3), v[i + 2], v[i + 3]))           # conv_c7_s1(3, tanh)
    return keep("out", a)


def discriminator_forward(v, x, collect=None, tag="d"):
    """define_discriminator (:112-126): logits [N, h', w', 1]."""
    a = F.leaky_relu(conv2d(x, v[0], v[1], 2, 0), 0.2)
    if collect is not None:
        collect[f"{tag}_k1"] = a
    i = 2
    for j in range(3):
        a = F.leaky_relu(row_norm(conv2d(a, v[i], v[i + 1], 2, 0), v[i + 2], v[i + 3]), 0.2)
        if collect is not None:
            collect[f"{tag}_k{j + 2}"] = a
        i += 4
    return conv2d(a, v[i], v[i + 1], 1, 0)


# ----------------------------------------------------------------------------------------------------------
# losses (:186-208) and the training step (:325-382)
# ----------------------------------------------------------------------------------------------------------
def bce(logits, label):
    return K.bce_with_logits(logits, label)


def discriminator_loss(real, generated):
    return (bce(real, 1.0) + bce(generated, 0.0)) * 0.5


def generator_loss(generated):
    return bce(generated, 1.0)


def calc_cycle_loss(real_image, cycled_image):
    return LAMBDA * (real_image - cycled_image).abs().mean()


def identity_loss(real_image, same_image):
    return LAMBDA * 0.5 * (real_image - same_image).abs().mean()


class _Adam:
    def __init__(self):
        self.m, self.v, self.t = None, None, 0

    def apply(self, variables, grads):
        if self.m is None:
            self.m = [torch.zeros_like(p) for p in variables]
            self.v = [torch.zeros_like(p) for p in variables]
        self.t += 1
        with torch.no_grad():
            for p, g, m, v in zip(variables, grads, self.m, self.v):
                K.adam_step(p, g, m, v, self.t, LR, beta1=BETA1)


class OracleCycleGAN:
    def __init__(self, image_size=(128, 128, 3), seed=62, dtype=torch.float32):
        self.image_size = tuple(image_size)
        self.g = init_vars(generator_var_shapes(image_size), seed, dtype)            # generator_g: X -> Y
        self.f = init_vars(generator_var_shapes(image_size), seed + 1, dtype)        # generator_f: Y -> X
        self.dx = init_vars(discriminator_var_shapes(image_size), seed + 2, dtype)
        self.dy = init_vars(discriminator_var_shapes(image_size), seed + 3, dtype)
        self.opt = {k: _Adam() for k in ("g", "f", "dx", "dy")}

    def train_step(self, real_x, real_y, collect=None, apply=True):
        allv = self.g + self.f + self.dx + self.dy
        for p in allv:
            p.requires_grad_(True)
        fake_y = generator_forward(self.g, real_x, collect, "g_x")
        cycled_x = generator_forward(self.f, fake_y, collect, "f_fy")
        fake_x = generator_forward(self.f, real_y, collect, "f_y")
        cycled_y = generator_forward(self.g, fake_x, collect, "g_fx")
        same_x = generator_forward(self.f, real_x, collect, "f_x")
        same_y = generator_forward(self.g, real_y, collect, "g_y")
        disc_real_x = discriminator_forward(self.dx, real_x)
        disc_real_y = discriminator_forward(self.dy, real_y)
        disc_fake_x = discriminator_forward(self.dx, fake_x)
        disc_fake_y = discriminator_forward(self.dy, fake_y)
        gen_g_loss = generator_loss(disc_fake_y)
        gen_f_loss = generator_loss(disc_fake_x)
        total_cycle_loss = calc_cycle_loss(real_x, cycled_x) + calc_cycle_loss(real_y, cycled_y)
        identity_loss_g = identity_loss(real_y, same_y)
        identity_loss_f = identity_loss(real_x, same_x)
        total_gen_g_loss = gen_g_loss + total_cycle_loss + identity_loss_g
        total_gen_f_loss = gen_f_loss + total_cycle_loss + identity_loss_f
        disc_x_loss = discriminator_loss(disc_real_x, disc_fake_x)
        disc_y_loss = discriminator_loss(disc_real_y, disc_fake_y)
        gg = torch.autograd.grad(total_gen_g_loss, self.g, retain_graph=True)
        gf = torch.autograd.grad(total_gen_f_loss, self.f, retain_graph=True)
        gdx = torch.autograd.grad(disc_x_loss, self.dx, retain_graph=True)
        gdy = torch.autograd.grad(disc_y_loss, self.dy)
        if collect is not None:
            collect.update({"grads_g": gg, "grads_f": gf, "grads_dx": gdx, "grads_dy": gdy,
                            "fake_y": fake_y.detach(), "fake_x": fake_x.detach(),
                            "disc_x_loss": disc_x_loss.detach(), "disc_y_loss": disc_y_loss.detach()})
        for p in allv:
            p.requires_grad_(False)
        if apply:
            self.opt["g"].apply(self.g, gg)
            self.opt["f"].apply(self.f, gf)
            self.opt["dx"].apply(self.dx, gdx)
            self.opt["dy"].apply(self.dy, gdy)
        names = ("gen_g_loss", "gen_f_loss", "identity_loss_g", "identity_loss_f", "total_gen_g_loss", "total_gen_f_loss",
                 "total_cycle_loss")
        vals = (gen_g_loss, gen_f_loss, identity_loss_g, identity_loss_f, total_gen_g_loss, total_gen_f_loss,
                total_cycle_loss)
        return {k: float(v.detach()) for k, v in zip(names, vals)}
