"""FID-like score of the reference (sndcgan/generator_evaluation.py), SURVEY.md section 8(f) row N3.

  calculate_fid(disc_model, images_fake, images_real)     generator_evaluation.py:36-55
  feature extractor = discriminator minus Flatten/Dense + AveragePooling2D(8, 8) + Flatten   :134-138

The discriminator forward runs on the tcgen05 conv kernels (inference: no dropout), the pooling on b2g_avgpool8_bf16.
The Fréchet distance itself is small dense linear algebra on [B, 4096] features: means / traces by torch reductions,
and tr sqrtm(cov_fake cov_real) as the nuclear norm of the B x B matrix Xc Yc^T / sqrt((n-1)(m-1)) (exactly the trace
of the principal square root, without forming or square-rooting the singular 4096 x 4096 product the reference hands to
scipy.linalg.sqrtm) — torch.linalg on the device, library code off the training hot path."""
from __future__ import annotations

import torch

from . import lib as _lib
from .lib import B200GanError, check
from .sndcgan import Discriminator, _p, _stream


def discriminator_features(disc_model: Discriminator, images: torch.Tensor) -> torch.Tensor:
    """disc_model.predict(images) of the popped + pooled model: fp32 [B, (H/64)*(W/64)*512] (4096 at 144x256)."""
    if not isinstance(disc_model, Discriminator):
        raise B200GanError("discriminator_features needs an imagegeneration_b200 Discriminator")
    disc_model.forward(images.to(torch.float32), training=False)
    c7 = disc_model._acts[6]
    n, h, w, c = c7.shape
    if h < 8 or w < 8:
        raise B200GanError("AveragePooling2D(8, 8) needs a conv7 activation of at least 8 x 8 (images >= 64 x 64)")
    out = torch.empty((n, h // 8, w // 8, c), dtype=torch.float32, device=c7.device)
    check(_lib.load().b2g_avgpool8_bf16(_p(c7), _p(out), n, h, w, c, _stream()), "b2g_avgpool8_bf16")
    return out.reshape(n, -1)


def fid_from_features(predict_fake: torch.Tensor, predict_real: torch.Tensor) -> float:
    """generator_evaluation.py:41-55 on device tensors (fp64 accumulation)."""
    f, r = predict_fake.double(), predict_real.double()
    n, m = f.shape[0], r.shape[0]
    if n < 2 or m < 2:
        raise B200GanError("calculate_fid needs at least two images per set (np.cov)")
    mu_f, mu_r = f.mean(0), r.mean(0)
    xc, yc = f - mu_f, r - mu_r
    ssdiff = ((mu_f - mu_r) ** 2).sum()
    tr_f = (xc * xc).sum() / (n - 1)          # trace of np.cov(predict_fake, rowvar=False)
    tr_r = (yc * yc).sum() / (m - 1)
    cross = xc @ yc.t() / ((n - 1) * (m - 1)) ** 0.5
    tr_covmean = torch.linalg.svdvals(cross).sum()
    return float(ssdiff + tr_f + tr_r - 2.0 * tr_covmean)


def calculate_fid(disc_model: Discriminator, images_fake: torch.Tensor, images_real: torch.Tensor) -> float:
    """sndcgan/generator_evaluation.py:36."""
    return fid_from_features(discriminator_features(disc_model, images_fake).clone(),
                             discriminator_features(disc_model, images_real))


def evaluate_fid_models(disc_model: Discriminator, gen_model, generator_weights: dict, img_real_used, random_z_used):
    """The inner loop of sndcgan/generator_evaluation.py:156-175 on in-memory models: for every generator (epoch -> list of
    weights in trainable_variables order) and every real batch i, images_fake = create_samples(gen_model,
    random_z_used[i], B) (range [0, 1], as the reference feeds them) and fid = calculate_fid(disc_model, images_fake,
    images_real).  Returns (epochs, [fids per epoch])."""
    from .sndcgan import create_samples
    epochs, epoch_fids = [], []
    for epoch in sorted(generator_weights):
        gen_model.set_weights(generator_weights[epoch])
        fids = []
        for images_real, z in zip(img_real_used, random_z_used):
            images_fake = create_samples(gen_model, z, images_real.shape[0])
            fids.append(calculate_fid(disc_model, images_fake, images_real))
        epochs.append(epoch)
        epoch_fids.append(fids)
    return epochs, epoch_fids


MAX_BATCHES = 50          # generator_evaluation.py:27 (max_batches)


def _saved_epochs(model_dir, prefix):
    import glob
    import os
    files = glob.glob(os.path.join(model_dir, prefix + "*.pt"))
    return sorted(int(os.path.basename(f)[len(prefix):-3]) for f in files)


def init_fid_evaluation(dir_path, dataset, batch_size, output, step_size, start_epoch, disc_epoch,
                        image_size=(144, 256, 3), z_size=128):
    """generator_evaluation.py:58-102: which saved generators to score, and the fixed real batches / latent batches."""
    import os
    import pickle
    from .trainer import get_dataset
    epochs_used = [e for e in _saved_epochs(os.path.join(dir_path, "models", "generator"), "gen_model-")
                   if e >= start_epoch][::step_size]
    train_ds = get_dataset(dataset, batch_size, image_size)
    batches_used = min(MAX_BATCHES, len(train_ds))
    img_real_used, random_z_used = [], []
    for i, (images_real, _) in enumerate(train_ds):
        img_real_used.append(images_real.cpu())
        random_z_used.append(torch.rand((images_real.shape[0], z_size)) * 2.0 - 1.0)
        if i == batches_used - 1:
            break
    init = {"epochs_used": epochs_used, "img_real_used": img_real_used, "random_z_used": random_z_used,
            "batches_used": batches_used, "disc_epoch": disc_epoch}
    with open(os.path.join(output, "fid_tmp_init.pickle"), "wb") as f:
        pickle.dump(init, f)
    return init


def evaluate_fid(dir_path, dataset, batch_size, output, step_size, start_epoch, disc_epoch, continue_,
                 image_size=(144, 256, 3), z_size=128):
    """sndcgan/generator_evaluation.py:107-199, same signature: scores every saved generator
    `<dir_path>/models/generator/gen_model-<epoch>.pt` (written by SNDCGAN.train at each checkpoint) with the FID-like
    distance on the features of `<dir_path>/models/discriminator/disc_model-<disc_epoch>.pt`, resumable through the same
    pickles (fid_tmp_init / fid_tmp_results / fid_results under <output>/evaluation).  The matplotlib box / line plots
    (:202-245) are not drawn (no matplotlib, no display).  Returns (epochs_used, epoch_fids)."""
    import os
    import pickle
    import shutil
    from .sndcgan import create_samples
    from .trainer import load_model
    output = os.path.join(output, "evaluation")
    if not continue_ and os.path.exists(output):
        shutil.rmtree(output)
    os.makedirs(output, exist_ok=True)
    if not continue_:
        init = init_fid_evaluation(dir_path, dataset, batch_size, output, step_size, start_epoch, disc_epoch,
                                   image_size, z_size)
    else:
        with open(os.path.join(output, "fid_tmp_init.pickle"), "rb") as f:
            init = pickle.load(f)
    epochs_used, disc_epoch = init["epochs_used"], init["disc_epoch"]
    disc_model = load_model(os.path.join(dir_path, "models", "discriminator", f"disc_model-{disc_epoch}.pt"))
    dev = disc_model.device
    tmp_results = os.path.join(output, "fid_tmp_results.pickle")
    epochs_finished, epoch_fids = [], []
    if continue_ and os.path.exists(tmp_results):
        with open(tmp_results, "rb") as f:
            d = pickle.load(f)
        epochs_finished, epoch_fids = d["epochs_finished"], d["epoch_fids"]
    for epoch in [e for e in epochs_used if e not in epochs_finished]:
        print("\n## Start FID calculation of epoch", epoch)
        gen_model = load_model(os.path.join(dir_path, "models", "generator", f"gen_model-{epoch}.pt"))
        fids = []
        for i, images_real in enumerate(init["img_real_used"]):
            print(f"\r### Calculate FID of Batch {i + 1}/{init['batches_used']}", end="", flush=True)
            images_fake = create_samples(gen_model, init["random_z_used"][i].to(dev), images_real.shape[0])
            fids.append(calculate_fid(disc_model, images_fake, images_real.to(dev)))
        epochs_finished.append(epoch)
        epoch_fids.append(fids)
        with open(tmp_results, "wb") as f:
            pickle.dump({"epochs_finished": epochs_finished, "epoch_fids": epoch_fids}, f)
    print("\n\n## Calculation finished.")
    with open(os.path.join(output, "fid_results.pickle"), "wb") as f:
        pickle.dump({"epochs": epochs_used, "fids": epoch_fids}, f)
    return epochs_used, epoch_fids
