"""GPU: the Trainer-level drop-in (sndcgan/SNDCGAN.py:148-212, :272-335): directory handling, epoch loop, losses
pickle, checkpoint cadence, --continue resume; plus hinge / spectral-norm options against the oracle."""
import os
import pickle

import numpy as np
import pytest
import torch

from oracle import sndcgan_oracle as O

pytestmark = pytest.mark.gpu


def test_train_loop_checkpoint_and_resume(tmp_path):
    from imagegeneration_b200.trainer import SNDCGAN
    d = str(tmp_path / "training")
    live = str(tmp_path / "live")
    size = (16, 32, 3)
    m = SNDCGAN(d, "synthetic:12", 4, 0.5, 2e-4, 2e-4, live, False, size, 128)
    m.train(2 + 1, 2)                                  # epochs 0,1,2 -> checkpoints at 0 and 2 (:317)
    ck = sorted(os.listdir(os.path.join(d, "checkpoints")))
    assert ck == ["ckpt-0.pt", "ckpt-2.pt"]
    losses = pickle.load(open(os.path.join(d, "losses.pickle"), "rb"))
    assert losses["epoch"] == [0, 1, 2] and len(losses["avg_g_loss"]) == 3
    assert all(np.isfinite(v) for v in losses["avg_d_loss"])
    assert m.d_optimizer.iterations == 2 * m.g_optimizer.iterations == 2 * 9
    assert os.path.exists(live + ".png")
    w_before = m.gen_model.arena.w.clone()
    m2 = SNDCGAN(d, "synthetic:12", 4, 0.5, 2e-4, 2e-4, live, True, size, 128)      # --continue
    assert m2.start_epoch == 3
    assert torch.equal(m2.gen_model.arena.w, w_before)
    assert m2.g_optimizer.iterations == 9
    m3 = SNDCGAN(d, "synthetic:12", 4, 0.5, 2e-4, 2e-4, live, False, size, 128)     # fresh start wipes the dir (:152)
    assert m3.start_epoch == 0 and os.listdir(os.path.join(d, "checkpoints")) == []


def test_ragged_last_batch_model_exports_and_fid_from_saved_files(tmp_path):
    """VERDICT r1 boundary items: the ragged last batch is TRAINED ON like the reference does (SNDCGAN.py:282-284, with
    batch_size latent vectors), every checkpointed epoch exports models/generator/gen_model-<epoch>.pt and
    models/discriminator/disc_model-<epoch>.pt (:327-331), `ckpt_manager` / `loss_fn` exist, and the reference's
    evaluation + sampling entry points run from those files alone."""
    import imagegeneration_b200 as P
    from imagegeneration_b200 import evaluation, generator_output
    d = str(tmp_path / "training")
    live = str(tmp_path / "live")
    size = (64, 64, 3)                                   # AveragePooling2D(8, 8) of the FID features needs >= 64 x 64
    m = P.SNDCGAN(d, "synthetic:10", 4, 0.5, 2e-4, 2e-4, live, False, size, 128)
    assert len(m.train_ds) == 3                          # 4 + 4 + 2: the last batch is ragged
    assert m.ckpt_manager.latest_checkpoint is None and callable(m.loss_fn)
    m.train(2 + 1, 1)                                    # epochs 0..2, checkpoint every epoch, max_to_keep = 2
    assert m.g_optimizer.iterations == 9                 # 3 batches per epoch INCLUDING the ragged one
    assert [os.path.basename(c) for c in m.ckpt_manager.checkpoints] == ["ckpt-1.pt", "ckpt-2.pt"]
    gens = sorted(os.listdir(os.path.join(d, "models", "generator")))
    discs = sorted(os.listdir(os.path.join(d, "models", "discriminator")))
    assert gens == [f"gen_model-{e}.pt" for e in range(3)] and discs == [f"disc_model-{e}.pt" for e in range(3)]
    assert os.path.exists(os.path.join(d, "plot_line_plot_loss.png"))
    # a saved generator reproduces the live one
    g2 = P.load_model(os.path.join(d, "models", "generator", "gen_model-2.pt"))
    z = torch.rand((3, 128), device="cuda") * 2 - 1
    assert torch.equal(P.create_samples(g2, z, 3), P.create_samples(m.gen_model, z, 3))
    # generator_evaluation.evaluate_fid's signature, from the saved files only
    epochs, fids = evaluation.evaluate_fid(d, "synthetic:10", 4, str(tmp_path), 1, 0, 2, False, image_size=size)
    assert epochs == [0, 1, 2] and len(fids) == 3 and all(len(f) == 3 and np.isfinite(f).all() for f in fids)
    res = pickle.load(open(os.path.join(str(tmp_path), "evaluation", "fid_results.pickle"), "rb"))
    assert res["epochs"] == [0, 1, 2]
    epochs2, fids2 = evaluation.evaluate_fid(d, "synthetic:10", 4, str(tmp_path), 1, 0, 2, True, image_size=size)
    assert epochs2 == epochs and np.allclose(fids2, fids)                     # --continue: nothing left to do
    png = generator_output.output_results_models(3, d, 1, "grid", 0)
    from PIL import Image
    assert Image.open(png).size == (3 * size[1], 3 * size[0])
    png2 = generator_output.output_results_ckpts(3, d, 1, "grid_ck", 0)
    assert Image.open(png2).size == (3 * size[1], 2 * size[0])


def test_device_image_dataset_matches_rescaling(tmp_path):
    """get_dataset (SNDCGAN.py:131-145): decode -> centre crop to the aspect ratio -> bilinear resize -> x/127.5 - 1,
    served from the device-resident uint8 cache by b2g_gather_u8_to_f32; every image appears once per epoch."""
    from PIL import Image
    from imagegeneration_b200.trainer import DeviceImageDataset, get_dataset
    imgs = torch.randint(0, 256, (10, 16, 32, 3), dtype=torch.uint8, generator=torch.Generator().manual_seed(0))
    ds = DeviceImageDataset(imgs, 4)
    seen = []
    for x, y in ds:
        assert x.is_cuda and x.dtype == torch.float32 and y.shape[0] == x.shape[0]
        seen.append(x.cpu())
    got = torch.cat(seen)
    assert got.shape[0] == 10 and [s.shape[0] for s in seen] == [4, 4, 2]
    ref = imgs.float() / 127.5 - 1.0
    # same multiset of images, exact arithmetic up to one fp32 rounding (x * (1/127.5) - 1 vs x / 127.5 - 1)
    order = [int(((ref - g[None]).abs().flatten(1).max(1).values).argmin()) for g in got]
    assert sorted(order) == list(range(10))
    assert float((got - ref[order]).abs().max()) < 1e-6
    # folder path: a 40x40 image, target 16x32 -> centre crop 40x20, resize without antialiasing
    root = tmp_path / "data" / "class0"
    root.mkdir(parents=True)
    arr = np.random.RandomState(1).randint(0, 256, (40, 40, 3), dtype=np.uint8)
    Image.fromarray(arr).save(str(root / "a.png"))
    x, _ = next(iter(get_dataset(str(tmp_path / "data"), 1, (16, 32, 3))))
    crop = torch.from_numpy(arr[10:30]).permute(2, 0, 1)[None].float()
    want = torch.nn.functional.interpolate(crop, size=(16, 32), mode="bilinear", align_corners=False, antialias=False)
    want = want[0].permute(1, 2, 0).round() / 127.5 - 1.0
    assert float((x[0].cpu() - want).abs().max()) < 1e-6


def test_spectral_norm_inference_before_training_and_after_restore(tmp_path):
    """a fresh spectrally-normalised discriminator must give meaningful inference logits (sigma used
    to be 0 until the first TRAINING forward), also right after a checkpoint restore."""
    from imagegeneration_b200.sndcgan import make_dcgan_discriminator
    from imagegeneration_b200.trainer import SNDCGAN
    size, B = (16, 32, 3), 4
    D = make_dcgan_discriminator(0.5, size, seed=63, spectral_norm=True)
    x = torch.rand((B, *size), generator=torch.Generator().manual_seed(4)) * 2 - 1
    logits = D(x.cuda(), training=False)
    torch.cuda.synchronize()
    dv = [v.detach().cpu().clone() for v in D.trainable_variables]
    u0 = []
    gen = torch.Generator().manual_seed(64)               # the engine's u0 (seed + 1), see Discriminator._init_sn
    for i in range(8):
        u = torch.randn(dv[2 * i].shape[-1], generator=gen, dtype=torch.float64)
        u0.append((u / u.norm()).to(torch.float32))
    ref = O.discriminator_forward(dv, x, 0.0, None, sn_state=O.SNState(u0))
    assert float((logits.cpu() - ref).abs().max()) < 2e-2 * (1 + float(ref.abs().max()))
    assert float(logits.std()) > 0 and all(float(s[1]) > 0 for s in D.sn_scal)
    # trainer round trip: sigma / u / v survive a checkpoint, the restored model needs no extra power iteration
    d = str(tmp_path / "training")
    live = str(tmp_path / "live")
    m = SNDCGAN(d, "synthetic:8", 4, 0.5, 2e-4, 2e-4, live, False, size, 128, spectral_norm=True)
    m.train(1, 1)
    want = m.disc_model(x.cuda(), training=False).clone()
    m2 = SNDCGAN(d, "synthetic:8", 4, 0.5, 2e-4, 2e-4, live, True, size, 128, spectral_norm=True)
    assert not m2.disc_model._sn_stale
    got = m2.disc_model(x.cuda(), training=False)
    assert torch.equal(got, want)


@pytest.mark.parametrize("loss,sn", [("hinge", False), ("bce", True), ("hinge", True)])
def test_hinge_and_spectral_norm_options(loss, sn):
    """north_star options that the reference lacks: losses vs the oracle <= 2e-2, sigma <= 1e-5."""
    from imagegeneration_b200.sndcgan import SNDCGANEngine
    size, B = (16, 32, 3), 4
    eng = SNDCGANEngine(B, dropout=0.0, image_size=size, seed=62, loss=loss, spectral_norm=sn)
    orc = O.OracleSNDCGAN(size, dropout=0.0, seed=62, loss=loss, spectral_norm=sn)
    orc.gv = [v.detach().cpu().clone() for v in eng.gen_model.trainable_variables]
    orc.dv = [v.detach().cpu().clone() for v in eng.disc_model.trainable_variables]
    if sn:
        orc.sn_state = [u.detach().cpu().clone() for u in eng.disc_model.sn_u]
    g = torch.Generator().manual_seed(3)
    real = torch.rand((B, *size), generator=g) * 2 - 1
    z = torch.rand((B, 128), generator=g) * 2 - 1
    col = {}
    ref = orc.train_step(real, z, collect=col)
    out = eng.train_step(real.cuda(), z.cuda())
    torch.cuda.synchronize()
    for name, a, b in zip(("g_loss", "d_loss", "d_loss_real", "d_loss_fake"), out[:4], ref[:4]):
        assert abs(float(a) - float(b)) <= 2e-2 * max(1.0, abs(float(b))), (name, float(a), float(b))
    if sn:
        # Spectral norm runs ONE power iteration per weight update (Miyato Alg. 1), inside the optimizer pass: after the
        # step the engine's sigma belongs to the twice-updated weights, i.e. to the oracle's NEXT lazy refresh.
        # Tolerance 1e-3, not 1e-5: the weights took Adam steps of lr * sign(m) per element, and elements whose gradient
        # is ~0 may step the other way than in the oracle (fp32 atomics in the weight-gradient kernels make that run
        # dependent; observed 0.5e-4 .. 2.1e-4).  The 1e-5 bound of north_star is checked below on identical weights.
        orc.sn_state.refresh(orc.dv)
        for i in range(8):
            w2d = orc.dv[2 * i].reshape(-1, orc.dv[2 * i].shape[-1])
            s_ref = float(orc.sn_state.v[i] @ (w2d @ orc.sn_state.u[i]))
            assert abs(float(eng.disc_model.sn_scal[i][1]) - s_ref) / s_ref < 1e-3, (
                i, float(eng.disc_model.sn_scal[i][1]), s_ref)
        from oracle import keras_ops as K
        # identical u0: re-create the engine's initial vectors
        eng3 = SNDCGANEngine(B, dropout=0.0, image_size=size, seed=62, loss=loss, spectral_norm=True)
        u0s = [u.detach().cpu().double() for u in eng3.disc_model.sn_u]
        eng3.disc_model.sn_power_iteration()
        torch.cuda.synchronize()
        for i in range(8):
            w2d = eng3.disc_model.trainable_variables[2 * i].detach().cpu().double()
            w2d = w2d.reshape(-1, w2d.shape[-1])
            sigma, _u, _v = K.spectral_norm_power_iteration(w2d, u0s[i], 1)
            assert abs(float(eng3.disc_model.sn_scal[i][1]) - float(sigma)) / float(sigma) < 1e-5, i
