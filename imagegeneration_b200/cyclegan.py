"""CycleGAN on the B200 kernels — SURVEY.md section 8(f) row N2, BASELINE config 5 (cyclegan/CycleGAN.py).

Same surface as the reference: `define_generator()`, `define_discriminator()`, the five loss functions and
`CycleGAN.train_step(real_x, real_y)` returning the reference's dict of seven losses (cyclegan/CycleGAN.py:325-382).

How it runs.  The model's convolutions (7x7 s1 'same', reflection-padded 3x3 s2 'valid', 3x3 s1 residual convs, 3x3 s2
'same' transposed, 4x4 s2 / s1 'valid' PatchGAN convs, feature maps of 63, 30, 14, 6 pixels) do not fit the 128-pixel TMA
tilings of the SNDCGAN kernels, so every one of them is   b2g_im2col (gather, bf16)  ->  tcgen05 GEMM (b2g_conv_igemm kind
K1S1, bias / LeakyReLU in its epilogue)  and, backward,  b2g_conv_wgrad kind K1S1  +  a second gather + GEMM for the input
gradient.  The row-wise InstanceNormalization of the reference, the activation behind it and the ResBlock's add + ReLU
are one kernel each way (b2g_rownorm_fwd / _bwd).  This is the first correct path of this row, not a tuned one.

The step has 6 generator and 6 discriminator forward passes whose graphs share tensors (fake_y feeds the cycle, the
identity... ), so the backward ORDER is left to torch.autograd: every fused layer is a torch.autograd.Function over the C
ABI, the step builds one graph and calls backward once.  torch does plumbing only: tensor allocation, the graph, the
accumulation of gradients that meet at a shared tensor, scalar loss sums.  No arithmetic of a layer happens in torch.
There is no CPU fallback.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import lib as _lib
from . import ops
from .lib import ACT_LEAKY, ACT_NONE, ACT_RELU, ACT_TANH, CONV_K1S1, B200GanError, check
from .sndcgan import Adam, ParamArena, _glorot_uniform

IN_EPS = 1e-3          # tensorflow_addons InstanceNormalization / GroupNormalization default epsilon
LAMBDA = 10.0          # cyclegan/CycleGAN.py:186
BF16, F32 = torch.bfloat16, torch.float32


def _p(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _up(n, m):
    return (n + m - 1) // m * m


class ConvSpec:
    """Geometry of one convolution layer.  transposed=False: Conv2D (k, stride, zero pad `pad`, or reflect=1: the
    ReflectionPadding2D((1,1)) in front of a 'valid' conv); transposed=True: Conv2DTranspose(3x3, strides 2, 'same')."""

    def __init__(self, k, stride, pad, cin, cout, transposed=False, reflect=0):
        self.k, self.stride, self.pad, self.cin, self.cout = k, stride, pad, cin, cout
        self.transposed, self.reflect = transposed, reflect
        self.npad = _up(cout, 64)                      # GEMM N (output channels, padded)
        self.cin_pad = _up(cin, 64)                    # GEMM N of the input-gradient GEMM
        self.kf = _up(k * k * cin, 64)                 # GEMM K forward
        self.kb = _up(k * k * cout, 64)                # GEMM K of the input gradient

    def out_hw(self, h, w):
        if self.transposed:
            return 2 * h, 2 * w
        p = 1 if self.reflect else self.pad
        return (h + 2 * p - self.k) // self.stride + 1, (w + 2 * p - self.k) // self.stride + 1


def _im2col(src, src_f32, n, hin, win, c, ldc, k, mode, stride, pad, reflect, hout, wout, kpad):
    """Gathered GEMM operand [Mpad, Kpad] bf16 (Mpad: multiple of 256, padding zero)."""
    m = n * hout * wout
    out = torch.empty((_up(m, 256), kpad), dtype=BF16, device=src.device)
    check(_lib.load().b2g_im2col(_p(src), int(src_f32), n, hin, win, c, ldc, k, k, mode, stride, pad, pad, reflect, hout,
                                 wout, _p(out), out.shape[0], kpad, _stream()), "b2g_im2col")
    return out


def _pack(w, taps, d2, d3, n_is_d3, npad, kpad):
    out = torch.empty((1, npad, kpad), dtype=BF16, device=w.device)
    check(_lib.load().b2g_pack_gemm_weight(_p(w), taps, d2, d3, int(n_is_d3), _p(out), npad, kpad, _stream()),
          "b2g_pack_gemm_weight")
    return out


def _gemm(a, wp, bias=None, act=ACT_NONE, alpha=0.0):
    """[Mpad, K] x [1, N, K]^T -> bf16 [Mpad, N] on the tcgen05 conv kernels (a 1x1 'convolution' over Mpad pixels)."""
    mpad, k = a.shape
    out = torch.empty((mpad, wp.shape[1]), dtype=BF16, device=a.device)
    ops.conv_igemm(CONV_K1S1, a.view(1, mpad // 128, 128, k), wp, out.view(1, mpad // 128, 128, wp.shape[1]), bias=bias,
                   act=act, alpha=alpha)
    return out


def _wgrad(shifted, unshifted, k_real, shape):
    """dw[a][b] = sum_m shifted[m][a] * unshifted[m][b] -> fp32 tensor of `shape` (= [k_real, Cb] flattened)."""
    mpad, ka = shifted.shape
    cb = unshifted.shape[1]
    dw = torch.zeros((k_real, cb), dtype=F32, device=shifted.device)
    ops.conv_wgrad(CONV_K1S1, shifted.view(1, mpad // 64, 64, ka), unshifted.view(1, mpad // 64, 64, cb), dw,
                   ca_valid=k_real if k_real < ka else 0)
    return dw if shape is None else dw.view(shape)


def _padded_rows(t, m, ld):
    """bf16 [m, ld] matrix whose storage continues, zero-filled, up to a multiple of 256 rows (GEMM / wgrad operand)."""
    buf = torch.zeros((_up(m, 256), ld), dtype=BF16, device=t.device)
    buf[:m].copy_(t.reshape(m, ld))
    return buf


def _colsum(x2d, m, ld, cout):
    out = torch.zeros(ld, dtype=F32, device=x2d.device)
    check(_lib.load().b2g_colsum_bf16(_p(x2d), _p(out), m, ld, _stream()), "b2g_colsum_bf16")
    return out[:cout]


def _rows_ok(t, m, ld):
    """True when `t` is a contiguous bf16 [m * ld] tensor at the start of a storage that extends to the padded row count."""
    return (t.dtype == BF16 and t.is_contiguous() and t.numel() == m * ld and t.storage_offset() == 0
            and t.untyped_storage().nbytes() >= _up(m, 256) * ld * 2)


def _new_act(n, h, w, c, dtype, device):
    """Activation tensor [n,h,w,c] whose storage continues (zeros) up to a multiple of 256 pixel rows, so that it can be
    a GEMM / weight-gradient operand as it is."""
    m = n * h * w
    buf = torch.zeros(_up(m, 256) * c, dtype=dtype, device=device)
    return buf[:m * c].view(n, h, w, c)


class ConvBlock(torch.autograd.Function):
    """One fused layer: [ReflectionPadding2D] -> Conv2D / Conv2DTranspose (+bias) -> [row-wise InstanceNormalization,
    optionally behind the ResBlock's add + ReLU] -> activation.   x: [N,H,W,Cin] bf16 (fp32 for 3-channel images)."""

    @staticmethod
    def forward(ctx, x, w, b, gamma, beta, res, spec, act, alpha, logits_out):
        lib = _lib.load()
        s = spec
        x = x.contiguous()
        n, h, wd, cin = x.shape
        ho, wo = s.out_hw(h, wd)
        m = n * ho * wo
        x_f32 = x.dtype == F32
        if s.transposed:
            a_mat = _im2col(x, x_f32, n, h, wd, cin, cin, s.k, 1, 2, 0, 0, ho, wo, s.kf)
            wp = _pack(w, s.k * s.k, s.cout, s.cin, False, s.npad, s.kf)
        else:
            a_mat = _im2col(x, x_f32, n, h, wd, cin, cin, s.k, 0, s.stride, 1 if s.reflect else s.pad, s.reflect, ho, wo, s.kf)
            wp = _pack(w, s.k * s.k, s.cin, s.cout, True, s.npad, s.kf)
        bp = b
        if s.npad != s.cout:
            bp = torch.zeros(s.npad, dtype=F32, device=x.device)
            bp[:s.cout].copy_(b)
        norm = gamma is not None
        pre = _gemm(a_mat, wp, bias=bp, act=ACT_NONE if norm else act, alpha=alpha)      # [Mpad, npad]
        del a_mat
        stat = None
        if norm:
            y = _new_act(n, ho, wo, s.cout, F32 if s.cout == 3 else BF16, x.device)
            stat = torch.empty(2 * n * ho, dtype=F32, device=x.device)
            check(lib.b2g_rownorm_fwd(_p(pre), _p(res), _p(gamma), _p(beta), _p(y), _p(stat), n, ho, wo, s.cout, s.npad,
                                      act, alpha, IN_EPS, int(y.dtype == F32), _stream()), "b2g_rownorm_fwd")
        elif logits_out:                       # PatchGAN head: one real column of the padded GEMM output
            y = pre[:m, :s.cout].to(F32).view(n, ho, wo, s.cout)
        else:
            if s.npad != s.cout:
                raise B200GanError("a layer without normalisation needs Cout % 64 == 0")
            y = pre[:m].view(n, ho, wo, s.cout)
        ctx.save_for_backward(x, w, gamma, pre, res, y, stat)
        ctx.spec, ctx.act, ctx.alpha, ctx.logits_out, ctx.geom = s, act, alpha, logits_out, (n, h, wd, cin, ho, wo, m)
        return y

    @staticmethod
    def backward(ctx, dy):
        lib = _lib.load()
        x, w, gamma, pre, res, y, stat = ctx.saved_tensors
        s, act, alpha = ctx.spec, ctx.act, ctx.alpha
        n, h, wd, cin, ho, wo, m = ctx.geom
        dev = x.device
        need_x, need_w, need_b, need_g = ctx.needs_input_grad[0], ctx.needs_input_grad[1], ctx.needs_input_grad[2], \
            ctx.needs_input_grad[3]
        dy = dy.contiguous()
        da = torch.zeros((_up(m, 256), s.npad), dtype=BF16, device=dev)          # dL/d(conv output), GEMM operand
        dgamma = dbeta = dres = None
        if gamma is not None:
            dgamma = torch.zeros(ho, dtype=F32, device=dev)
            dbeta = torch.zeros(ho, dtype=F32, device=dev)
            check(lib.b2g_rownorm_bwd(_p(dy.to(y.dtype)), _p(y), _p(pre), _p(res), _p(gamma), _p(stat), _p(da), _p(dgamma),
                                      _p(dbeta), n, ho, wo, s.cout, s.npad, act, alpha, int(y.dtype == F32), _stream()),
                  "b2g_rownorm_bwd")
            if res is not None:
                dres = da[:m].view(n, ho, wo, s.cout)
        elif ctx.logits_out:
            da[:m, :s.cout].copy_(dy.reshape(m, s.cout))
        else:   # LeakyReLU derivative from the saved output; Keras: 1 for x > 0, alpha otherwise
            check(lib.b2g_dact(_p(dy.to(BF16)), _p(y), _p(da), m * s.npad, 0, act, alpha, 1.0, alpha if act == ACT_LEAKY else 1.0,
                               _stream()), "b2g_dact")
        db = _colsum(da, m, s.npad, s.cout) if need_b else None
        dw = dx = None
        x_f32 = x.dtype == F32
        if s.transposed:
            # gather of the output gradient at stride 2: operand of BOTH the input gradient and the weight gradient
            g0 = _im2col(da, False, n, ho, wo, s.cout, s.npad, s.k, 0, 2, 0, 0, h, wd, s.kb)
            if need_w:
                mx = n * h * wd
                xm = x if _rows_ok(x, mx, cin) else _padded_rows(x, mx, cin)
                xmat = torch.as_strided(xm, (_up(mx, 256), cin), (cin, 1))      # the zero-padded rows behind the tensor
                dw = _wgrad(g0, xmat, s.k * s.k * s.cout, tuple(w.shape))       # [ky,kx,co][ci] = Keras [k,k,Cout,Cin]
            if need_x:
                wpd = _pack(w, s.k * s.k, s.cout, s.cin, True, s.cin_pad, s.kb)
                dx = _gemm(g0, wpd)[:n * h * wd].view(n, h, wd, cin)
        else:
            if need_w:
                a_mat = _im2col(x, x_f32, n, h, wd, cin, cin, s.k, 0, s.stride, 1 if s.reflect else s.pad, s.reflect, ho, wo,
                                s.kf)
                dwm = _wgrad(a_mat, da, s.k * s.k * cin, None)                       # [K, npad]
                del a_mat
                dw = dwm.view(s.k, s.k, cin, s.npad)
                if s.npad != s.cout:
                    dw = dw[..., :s.cout].contiguous()
            if need_x:
                hp, wp_ = (h + 2, wd + 2) if s.reflect else (h, wd)
                g1 = _im2col(da, False, n, ho, wo, s.cout, s.npad, s.k, 1, s.stride, 0 if s.reflect else s.pad, 0, hp, wp_, s.kb)
                wpd = _pack(w, s.k * s.k, s.cin, s.cout, False, s.cin_pad, s.kb)
                dxm = _gemm(g1, wpd)                                                  # [., cin_pad]
                del g1
                if s.reflect:
                    dx = _new_act(n, h, wd, cin, BF16, dev)
                    check(lib.b2g_reflect_fold(_p(dxm), _p(dx), n, h, wd, cin, s.cin_pad, 1, _stream()), "b2g_reflect_fold")
                elif s.cin_pad != cin:
                    dx = dxm[:n * h * wd, :cin].to(x.dtype).view(n, h, wd, cin)
                else:
                    dx = dxm[:n * h * wd].view(n, h, wd, cin)
                if dx.dtype != x.dtype:
                    dx = dx.to(x.dtype)
        return dx, dw, db, dgamma if need_g else None, dbeta if need_g else None, dres, None, None, None, None


class L1Loss(torch.autograd.Function):
    """scale * mean|a - b| on fp32 images (calc_cycle_loss / identity_loss, cyclegan/CycleGAN.py:199-207); gradient w.r.t. a."""

    @staticmethod
    def forward(ctx, a, b, scale):
        a, b = a.contiguous(), b.contiguous()
        loss = torch.zeros(1, dtype=F32, device=a.device)
        da = torch.empty_like(a)
        check(_lib.load().b2g_l1_loss(_p(a), _p(b), a.numel(), scale, _p(loss), _p(da), _stream()), "b2g_l1_loss")
        ctx.save_for_backward(da)
        return loss[0]

    @staticmethod
    def backward(ctx, g):
        (da,) = ctx.saved_tensors
        return da * g, None, None


# ----------------------------------------------------------------------------------------------------------
# models (cyclegan/CycleGAN.py:95-183)
# ----------------------------------------------------------------------------------------------------------
def disc_sizes(h):
    s = [h]
    for _ in range(4):
        s.append((s[-1] - 4) // 2 + 1)
    s.append(s[-1] - 3)
    return s


class _Model:
    """Variables in Keras `trainable_variables` order inside one flat arena (one fused Adam launch per model)."""

    def __init__(self, shapes, device, seed):
        self.device = torch.device(device)
        self.arena = ParamArena(shapes, self.device)
        gen = torch.Generator().manual_seed(seed)
        for name, v in zip(self.arena.names, self.arena.variables):
            if name.endswith("kernel"):
                v.copy_(_glorot_uniform(tuple(v.shape), gen))
            elif not name.endswith("bias"):      # gamma / beta: keras "random_uniform" = U(-0.05, 0.05)
                v.copy_(((torch.rand(tuple(v.shape), generator=gen, dtype=torch.float64) - 0.5) * 0.1).to(F32))
        self._params = None

    @property
    def trainable_variables(self):
        return self.arena.variables

    def set_weights(self, tensors):
        for v, t in zip(self.arena.variables, tensors):
            v.copy_(t.to(self.device, F32))

    def parameters(self):
        """Autograd leaves over the weight arena; their .grad are views of the gradient arena."""
        if self._params is None:
            self._params = [v.requires_grad_(True) for v in self.arena.variables]
            for p, g in zip(self._params, self.arena.grads):
                p.grad = g
        return self._params


class CycleGenerator(_Model):
    """define_generator (cyclegan/CycleGAN.py:161-183): c7s1-64, d128, d256, 9 x ResBlock(256), u128, u64, c7s1-3 (tanh)."""

    def __init__(self, image_size=(128, 128, 3), device="cuda", seed=62):
        h, w, c = image_size
        if h % 4 or w % 4 or c != 3:
            raise B200GanError("CycleGAN generator needs an (H, W, 3) image with H, W divisible by 4")
        self.image_size = tuple(image_size)
        shapes = [("c7s1_in/kernel", (7, 7, 3, 64)), ("c7s1_in/bias", (64,)), ("c7s1_in/gamma", (h,)), ("c7s1_in/beta", (h,)),
                  ("d1/kernel", (3, 3, 64, 128)), ("d1/bias", (128,)), ("d1/gamma", (h // 2,)), ("d1/beta", (h // 2,)),
                  ("d2/kernel", (3, 3, 128, 256)), ("d2/bias", (256,)), ("d2/gamma", (h // 4,)), ("d2/beta", (h // 4,))]
        for i in range(9):
            for j in (1, 2):
                shapes += [(f"res{i}/conv{j}/kernel", (3, 3, 256, 256)), (f"res{i}/conv{j}/bias", (256,)),
                           (f"res{i}/in{j}/gamma", (h // 4,)), (f"res{i}/in{j}/beta", (h // 4,))]
        shapes += [("u1/kernel", (3, 3, 128, 256)), ("u1/bias", (128,)), ("u1/gamma", (h // 2,)), ("u1/beta", (h // 2,)),
                   ("u2/kernel", (3, 3, 64, 128)), ("u2/bias", (64,)), ("u2/gamma", (h,)), ("u2/beta", (h,)),
                   ("c7s1_out/kernel", (7, 7, 64, 3)), ("c7s1_out/bias", (3,)), ("c7s1_out/gamma", (h,)),
                   ("c7s1_out/beta", (h,))]
        super().__init__(shapes, device, seed)
        self.specs = {"c7in": ConvSpec(7, 1, 3, 3, 64), "d1": ConvSpec(3, 2, 0, 64, 128, reflect=1),
                      "d2": ConvSpec(3, 2, 0, 128, 256, reflect=1), "res": ConvSpec(3, 1, 1, 256, 256),
                      "u1": ConvSpec(3, 2, 0, 256, 128, transposed=True), "u2": ConvSpec(3, 2, 0, 128, 64, transposed=True),
                      "c7out": ConvSpec(7, 1, 3, 64, 3)}

    def __call__(self, x, training=True, params=None):
        """x: fp32 CUDA [N,H,W,3] in [-1,1] -> fp32 [N,H,W,3].  `training` changes nothing (no dropout, the row norm has
        no moving statistics), it is there for the reference's call shape."""
        v = params if params is not None else (self._params if self._params is not None else self.arena.variables)
        sp = self.specs
        a = ConvBlock.apply(x, v[0], v[1], v[2], v[3], None, sp["c7in"], ACT_RELU, 0.0, False)
        a = ConvBlock.apply(a, v[4], v[5], v[6], v[7], None, sp["d1"], ACT_RELU, 0.0, False)
        a = ConvBlock.apply(a, v[8], v[9], v[10], v[11], None, sp["d2"], ACT_RELU, 0.0, False)
        i = 12
        for _ in range(9):                                  # ResBlock.call, :84-92
            fx = ConvBlock.apply(a, v[i], v[i + 1], v[i + 2], v[i + 3], None, sp["res"], ACT_RELU, 0.0, False)
            a = ConvBlock.apply(fx, v[i + 4], v[i + 5], v[i + 6], v[i + 7], a, sp["res"], ACT_NONE, 0.0, False)
            i += 8
        a = ConvBlock.apply(a, v[i], v[i + 1], v[i + 2], v[i + 3], None, sp["u1"], ACT_RELU, 0.0, False)
        a = ConvBlock.apply(a, v[i + 4], v[i + 5], v[i + 6], v[i + 7], None, sp["u2"], ACT_RELU, 0.0, False)
        return ConvBlock.apply(a, v[i + 8], v[i + 9], v[i + 10], v[i + 11], None, sp["c7out"], ACT_TANH, 0.0, False)


class CycleDiscriminator(_Model):
    """define_discriminator (:112-126): k64 (no norm), k128, k256, k512 (4x4 s2 'valid' + row norm + LeakyReLU 0.2), 4x4 s1
    conv to one logit per patch."""

    def __init__(self, image_size=(128, 128, 3), device="cuda", seed=64):
        hs = disc_sizes(image_size[0])
        if hs[-1] < 1 or disc_sizes(image_size[1])[-1] < 1:
            raise B200GanError("image too small for the PatchGAN discriminator (needs >= 94 pixels per side)")
        shapes = [("k1/kernel", (4, 4, 3, 64)), ("k1/bias", (64,))]
        cin = 64
        for i, cout in enumerate((128, 256, 512)):
            shapes += [(f"k{i + 2}/kernel", (4, 4, cin, cout)), (f"k{i + 2}/bias", (cout,)),
                       (f"k{i + 2}/gamma", (hs[i + 2],)), (f"k{i + 2}/beta", (hs[i + 2],))]
            cin = cout
        shapes += [("out/kernel", (4, 4, 512, 1)), ("out/bias", (1,))]
        super().__init__(shapes, device, seed)
        self.specs = [ConvSpec(4, 2, 0, 3, 64), ConvSpec(4, 2, 0, 64, 128), ConvSpec(4, 2, 0, 128, 256),
                      ConvSpec(4, 2, 0, 256, 512), ConvSpec(4, 1, 0, 512, 1)]

    def __call__(self, x, training=True, params=None):
        """x: fp32 [N,H,W,3] -> logits fp32 [N,h',w',1]."""
        v = params if params is not None else (self._params if self._params is not None else self.arena.variables)
        sp = self.specs
        a = ConvBlock.apply(x, v[0], v[1], None, None, None, sp[0], ACT_LEAKY, 0.2, False)
        i = 2
        for j in range(1, 4):
            a = ConvBlock.apply(a, v[i], v[i + 1], v[i + 2], v[i + 3], None, sp[j], ACT_LEAKY, 0.2, False)
            i += 4
        return ConvBlock.apply(a, v[i], v[i + 1], None, None, None, sp[4], ACT_NONE, 0.0, True)


def define_generator(image_size=(128, 128, 3), device="cuda", seed=62):
    return CycleGenerator(image_size, device, seed)


def define_discriminator(image_size=(128, 128, 3), device="cuda", seed=64):
    return CycleDiscriminator(image_size, device, seed)


# ----------------------------------------------------------------------------------------------------------
# losses (cyclegan/CycleGAN.py:186-208)
# ----------------------------------------------------------------------------------------------------------
def _bce(logits, label):
    from .torch_ops import bce_with_logits
    return bce_with_logits(torch.full_like(logits, label), logits)


def discriminator_loss(real, generated):
    return (_bce(real, 1.0) + _bce(generated, 0.0)) * 0.5


def generator_loss(generated):
    return _bce(generated, 1.0)


def calc_cycle_loss(real_image, cycled_image):
    return L1Loss.apply(cycled_image, real_image, LAMBDA)


def identity_loss(real_image, same_image):
    return L1Loss.apply(same_image, real_image, LAMBDA * 0.5)


# ----------------------------------------------------------------------------------------------------------
# CycleGAN (cyclegan/CycleGAN.py:211-382, without the reference's data loader / plots / checkpoint files)
# ----------------------------------------------------------------------------------------------------------
class CycleGAN:
    def __init__(self, batch_size, image_size=(128, 128, 3), device="cuda", seed=62, cuda_graph=None):
        _lib.load()
        self.batch_size, self.image_size = batch_size, tuple(image_size)
        # The step is ~1 900 small launches (gather + GEMM per conv, row norms, autograd glue): launched eagerly the HOST
        # is the bottleneck (55 ms per step at batch 8 for ~20 ms of kernels).  After two eager steps the whole step —
        # forward passes, torch.autograd's backward, the four Adam updates (bias-corrected rates read from device memory)
        # — is captured into ONE CUDA graph and replayed.  cuda_graph=False / B2G_CUDA_GRAPH=0 keeps every step eager.
        import os
        if cuda_graph is None:
            cuda_graph = os.environ.get("B2G_CUDA_GRAPH", "1") != "0"
        self.cuda_graph = bool(cuda_graph)
        self._graph = None
        self._graph_failed = None
        self._eager_steps = 0
        self.generator_g = define_generator(image_size, device, seed)            # X -> Y   (:235)
        self.generator_f = define_generator(image_size, device, seed + 1)        # Y -> X
        self.discriminator_x = define_discriminator(image_size, device, seed + 2)
        self.discriminator_y = define_discriminator(image_size, device, seed + 3)
        mk = lambda: Adam(2e-4, beta_1=0.5)                                       # :229-233
        self.generator_g_optimizer, self.generator_f_optimizer = mk(), mk()
        self.discriminator_x_optimizer, self.discriminator_y_optimizer = mk(), mk()
        self.losses = {k: [] for k in ("gen_g_loss", "gen_f_loss", "identity_loss_g", "identity_loss_f", "total_gen_g_loss",
                                       "total_gen_f_loss", "total_cycle_loss")}

    def _models(self):
        return ((self.generator_g, self.generator_g_optimizer), (self.generator_f, self.generator_f_optimizer),
                (self.discriminator_x, self.discriminator_x_optimizer), (self.discriminator_y, self.discriminator_y_optimizer))

    def compute_gradients(self, real_x, real_y):
        """Forward passes and the four gradient sets of train_step (:328-369); fills every model's gradient arena and
        returns the loss dict (CUDA scalars).  One autograd graph, one backward:
          * d(total_gen_g_loss)/d(G) and d(total_gen_f_loss)/d(F) come from the single sum gen_g + gen_f + cycle + id_g + id_f
            (gen_f / id_f do not depend on G, gen_g / id_g not on F, the cycle term is in both totals);
          * the generator losses see the discriminators through DETACHED weights (no discriminator weight gradient from
            them, as tape.gradient(..., generator variables) gives), the discriminator losses see DETACHED fakes."""
        G, F, DX, DY = self.generator_g, self.generator_f, self.discriminator_x, self.discriminator_y
        for m, _ in self._models():
            m.parameters()
            m.arena.g.zero_()
        frozen = lambda m: [p.detach() for p in m.parameters()]
        with torch.enable_grad():
            fake_y = G(real_x)
            cycled_x = F(fake_y)
            fake_x = F(real_y)
            cycled_y = G(fake_x)
            same_x = F(real_x)
            same_y = G(real_y)
            gen_g_loss = generator_loss(DY(fake_y, params=frozen(DY)))
            gen_f_loss = generator_loss(DX(fake_x, params=frozen(DX)))
            total_cycle_loss = calc_cycle_loss(real_x, cycled_x) + calc_cycle_loss(real_y, cycled_y)
            identity_loss_g = identity_loss(real_y, same_y)
            identity_loss_f = identity_loss(real_x, same_x)
            disc_x_loss = discriminator_loss(DX(real_x), DX(fake_x.detach()))
            disc_y_loss = discriminator_loss(DY(real_y), DY(fake_y.detach()))
            total = (gen_g_loss + gen_f_loss + total_cycle_loss + identity_loss_g + identity_loss_f
                     + disc_x_loss + disc_y_loss)
            total.backward()
        for m, _ in self._models():          # autograd accumulates in place into the arena views; make sure it did
            for p, g in zip(m.parameters(), m.arena.grads):
                if p.grad is None:
                    raise B200GanError("a CycleGAN variable received no gradient")
                if p.grad.data_ptr() != g.data_ptr():
                    g.copy_(p.grad)
                    p.grad = g
        self.fake_x, self.fake_y = fake_x.detach(), fake_y.detach()
        return {"gen_g_loss": gen_g_loss.detach(), "gen_f_loss": gen_f_loss.detach(),
                "identity_loss_g": identity_loss_g.detach(), "identity_loss_f": identity_loss_f.detach(),
                "total_gen_g_loss": (gen_g_loss + total_cycle_loss + identity_loss_g).detach(),
                "total_gen_f_loss": (gen_f_loss + total_cycle_loss + identity_loss_f).detach(),
                "total_cycle_loss": total_cycle_loss.detach(),
                "disc_x_loss": disc_x_loss.detach(), "disc_y_loss": disc_y_loss.detach()}

    def _step_body(self, real_x, real_y, lr_dev=None):
        out = self.compute_gradients(real_x, real_y)
        with torch.no_grad():
            for i, (m, opt) in enumerate(self._models()):                          # :372-377
                opt.apply_arena(m.arena, 1.0, None if lr_dev is None else lr_dev[i:i + 1])
        return {k: v for k, v in out.items() if not k.startswith("disc_")}

    def _capture_graph(self, real_x, real_y):
        """Captures one whole step (autograd backward included) into a CUDA graph over static input buffers.  The Adam
        rates are read from a 4-float device tensor refreshed before every replay."""
        try:
            self._g_x, self._g_y = real_x.clone(), real_y.clone()
            self._g_lr = torch.zeros(4, dtype=F32, device=real_x.device)
            snap = [(m.arena.w.clone(), m.arena.m.clone(), m.arena.v.clone()) for m, _ in self._models()]
            # capture must not change the training state: the weights / Adam slots it touches are restored afterwards
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):                  # warm-up on a side stream (allocator + autograd state)
                self._step_body(self._g_x, self._g_y, self._g_lr)
            torch.cuda.current_stream().wait_stream(side)
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                self._g_out = self._step_body(self._g_x, self._g_y, self._g_lr)
            for (m, _), (w, mm, vv) in zip(self._models(), snap):
                m.arena.w.copy_(w)
                m.arena.m.copy_(mm)
                m.arena.v.copy_(vv)
            self._graph = graph
        except Exception as e:          # capture is an optimisation: fall back to eager launches, keep the reason
            self._graph = None
            self._graph_failed = repr(e)
            self.cuda_graph = False
            torch.cuda.synchronize()

    def train_step(self, real_x, real_y):
        """cyclegan/CycleGAN.py:325-382: returns the reference's seven losses (CUDA scalars)."""
        if not real_x.is_cuda or not real_y.is_cuda:
            raise B200GanError("b200gan ops need CUDA tensors: there is no CPU fallback")
        real_x, real_y = real_x.to(F32), real_y.to(F32)
        usable = (self.cuda_graph and self._eager_steps >= 2 and tuple(real_x.shape) == (self.batch_size, *self.image_size)
                  and real_y.shape == real_x.shape)
        if usable and self._graph is None:
            self._capture_graph(real_x, real_y)
        if usable and self._graph is not None:
            lrs = []
            for _, opt in self._models():
                opt.iterations += 1
                lrs.append(opt._lr_t())
            self._g_lr.copy_(torch.tensor(lrs, dtype=F32), non_blocking=False)
            self._g_x.copy_(real_x)
            self._g_y.copy_(real_y)
            self._graph.replay()
            return {k: v.clone() for k, v in self._g_out.items()}
        self._eager_steps += 1
        return self._step_body(real_x, real_y)

    def train(self, loader, epochs):
        """The inner loop of :384-402 over any iterable of (batch_x, batch_y) fp32 NHWC CUDA tensors in [-1, 1]."""
        for _ in range(epochs):
            local = {k: [] for k in self.losses}
            for batch1, batch2 in loader:
                losses = self.train_step(batch1, batch2)
                for k, v in losses.items():
                    local[k].append(v)
            for k, v in local.items():
                if v:
                    self.losses[k].append(float(torch.stack(v).mean()))
        return self.losses
