"""GPU parity of the evaluation path (reference sndcgan/generator_evaluation.py:36-55,134-138) vs
oracle/evaluation_oracle.py, whose calculate_fid is the reference's numpy/scipy arithmetic call for call."""
import ctypes as C

import numpy as np
import pytest
import torch

from oracle import evaluation_oracle as E

pytestmark = pytest.mark.gpu


def _rel_l2(a, b):
    a, b = torch.as_tensor(a).double().flatten().cpu(), torch.as_tensor(b).double().flatten().cpu()
    return float((a - b).norm() / (b.norm() + 1e-30))


@pytest.mark.parametrize("shape", [(3, 18, 32, 512), (2, 8, 8, 64), (1, 17, 25, 128)])
def test_avgpool8_matches_keras_valid_pooling(shape):
    from imagegeneration_b200 import lib as L
    lib = L.load()
    g = torch.Generator().manual_seed(sum(shape))
    x = torch.randn(shape, generator=g).to(torch.bfloat16)
    n, h, w, c = shape
    out = torch.empty((n, h // 8, w // 8, c), device="cuda")
    xc = x.cuda()
    s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    L.check(lib.b2g_avgpool8_bf16(C.c_void_p(xc.data_ptr()), C.c_void_p(out.data_ptr()), n, h, w, c, s), "avgpool")
    torch.cuda.synchronize()
    assert _rel_l2(out, E.avg_pool_8x8(x.double())) < 1e-6
    assert lib.b2g_avgpool8_bf16(C.c_void_p(xc.data_ptr()), C.c_void_p(out.data_ptr()), n, 4, w, c, s) != 0


def test_fid_arithmetic_matches_reference_numpy_scipy():
    from imagegeneration_b200.evaluation import fid_from_features
    rng = np.random.default_rng(3)
    for n, m, d in ((40, 50, 24), (12, 12, 96), (6, 9, 1024)):
        a = rng.normal(size=(n, d)) * 0.7 + 0.1
        b = rng.normal(size=(m, d)) * 1.1 - 0.2
        ref = E.calculate_fid_from_features(a, b)
        got = fid_from_features(torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda())
        # scipy's sqrtm of a singular product (n < d) is itself only accurate to ~1e-4 relative
        assert abs(got - ref) < 2e-4 * abs(ref) + 1e-6, (n, m, d, got, ref)


def test_calculate_fid_end_to_end():
    from imagegeneration_b200.evaluation import calculate_fid, discriminator_features
    from imagegeneration_b200.sndcgan import make_dcgan_discriminator
    image = (64, 128, 3)
    D = make_dcgan_discriminator(0.5, image, seed=7)
    dv = [v.detach().cpu().clone() for v in D.trainable_variables]
    g = torch.Generator().manual_seed(1)
    real = torch.rand((6, *image), generator=g) * 2 - 1
    fake = torch.rand((6, *image), generator=g) * 0.5 + 0.25           # create_samples range [0, 1] (the reference's quirk)
    fe = discriminator_features(D, fake.cuda()).clone()
    assert tuple(fe.shape) == (6, 1 * 2 * 512)
    assert _rel_l2(fe, E.features(dv, fake)) < 2e-2
    got = calculate_fid(D, fake.cuda(), real.cuda())
    ref = E.calculate_fid(dv, fake, real)
    assert abs(got - ref) < 5e-2 * abs(ref), (got, ref)


def test_evaluate_fid_loop_over_saved_generators():
    """generator_evaluation.py:156-175: one FID per (saved generator, real batch); identical generators give identical
    scores, a perturbed generator a different one; every score matches the oracle on the same fakes."""
    from imagegeneration_b200.evaluation import evaluate_fid_models as evaluate_fid
    from imagegeneration_b200.sndcgan import create_samples, make_dcgan_discriminator, make_dcgan_generator
    from oracle import sndcgan_oracle as O
    image = (64, 128, 3)
    D = make_dcgan_discriminator(0.5, image, seed=7)
    G = make_dcgan_generator(image, seed=8)
    w0 = [v.detach().clone() for v in G.trainable_variables]
    w1 = [v * 1.5 if v.dim() > 1 else v.clone() for v in w0]
    g = torch.Generator().manual_seed(2)
    reals = [(torch.rand((4, *image), generator=g) * 2 - 1).cuda() for _ in range(2)]
    zs = [(torch.rand((4, 128), generator=g) * 2 - 1).cuda() for _ in range(2)]
    epochs, fids = evaluate_fid(D, G, {5: w0, 10: w1, 15: w0}, reals, zs)
    assert epochs == [5, 10, 15] and [len(f) for f in fids] == [2, 2, 2]
    assert all(abs(a - b) < 1e-6 * abs(a) + 1e-9 for a, b in zip(fids[0], fids[2]))
    assert any(abs(a - b) > 1e-3 * abs(a) for a, b in zip(fids[0], fids[1]))
    dv = [v.detach().cpu() for v in D.trainable_variables]
    fake0 = O.create_samples([v.cpu() for v in w0], zs[0].cpu(), 4, image)
    ref = E.calculate_fid(dv, fake0, reals[0].cpu())
    assert abs(fids[0][0] - ref) < 5e-2 * abs(ref), (fids[0][0], ref)
