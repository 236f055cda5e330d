"""Drop-in for the sampling side of sndcgan/generator_output.py: `create_samples` (the hot-path function, :25-28) and the
sample-grid writers over the files SNDCGAN.train leaves behind (`output_results_models` :51-70, `output_results_ckpts`
:73-100).  The reference draws matplotlib PDFs; there is no matplotlib / display here, so a grid is one PNG (Pillow):
row = saved epoch, column = sample.

    python -m imagegeneration_b200.generator_output <every> [-b N] [-d DIR] [-o NAME] [-s EPOCH]
"""
from __future__ import annotations

import argparse
import glob
import os

import numpy as np
import torch

from .sndcgan import Z_SIZE, create_samples  # noqa: F401  (create_samples is part of this module's surface)

z_size = Z_SIZE


def create_plot(dir_path, epoch_samples, batch_size, output_image, epochs_used):
    """:37-48 — one row per epoch, `batch_size` images per row; returns the PNG path."""
    from PIL import Image
    rows = [np.concatenate(list((e[:batch_size].clamp(0, 1) * 255).to(torch.uint8).cpu().numpy()), axis=1)
            for e in epoch_samples]
    path = os.path.join(dir_path, output_image + ".png")
    Image.fromarray(np.concatenate(rows, axis=0)).save(path)
    return path


def _epochs(pattern, prefix, start_epoch, every):
    found = sorted(int(os.path.basename(f)[len(prefix):].split(".")[0]) for f in glob.glob(pattern))
    return [e for e in found if e >= start_epoch][::every]


def output_results_models(batch_size, dir_path, every, output_image, start_epoch):
    """:51-70 — samples of every `every`-th exported generator (models/generator/gen_model-<epoch>.pt), same latent
    batch for all."""
    from .trainer import load_model
    model_path = os.path.join(dir_path, "models", "generator")
    epochs_used = _epochs(os.path.join(model_path, "gen_model-*.pt"), "gen_model-", start_epoch, every)
    fixed_z = torch.rand((batch_size, z_size), device="cuda") * 2.0 - 1.0
    epoch_samples = []
    for i, epoch in enumerate(epochs_used):
        print(f"\r Load Model {i}", end="", flush=True)
        gen_model = load_model(os.path.join(model_path, f"gen_model-{epoch}.pt"))
        epoch_samples.append(create_samples(gen_model, fixed_z, batch_size).clone())
    return create_plot(dir_path, epoch_samples, batch_size, output_image, epochs_used)


def output_results_ckpts(batch_size, dir_path, every, output_image, start_epoch):
    """:73-100 — the same from the rolling training checkpoints (checkpoints/ckpt-<epoch>.pt: generator weights only)."""
    from .trainer import load_model
    gen_model = load_model(os.path.join(dir_path, "models", "generator", "gen_model-0.pt"))
    ckpt_path = os.path.join(dir_path, "checkpoints")
    epochs_used = _epochs(os.path.join(ckpt_path, "ckpt-*.pt"), "ckpt-", start_epoch, every)
    fixed_z = torch.rand((batch_size, z_size), device="cuda") * 2.0 - 1.0
    epoch_samples = []
    for i, epoch in enumerate(epochs_used):
        print(f"\r Load Checkpoint {i}", end="", flush=True)
        st = torch.load(os.path.join(ckpt_path, f"ckpt-{epoch}.pt"), map_location="cpu")
        gen_model.arena.w.copy_(st["g_w"])
        gen_model._packed_dirty = True
        epoch_samples.append(create_samples(gen_model, fixed_z, batch_size).clone())
    return create_plot(dir_path, epoch_samples, batch_size, output_image, epochs_used)


def main(argv=None):
    p = argparse.ArgumentParser(description="Train GAN to generate landscapes")
    p.add_argument("every", type=int, help="Produce example for every xth checkpoint")
    p.add_argument("-b", "--bSize", type=int, dest="bSize", default=3)
    p.add_argument("-d", "--directory", type=str, dest="dirPath", default="training")
    p.add_argument("-o", "--output", type=str, dest="output", default="training")
    p.add_argument("-s", "--start", type=int, dest="start", default=0)
    a = p.parse_args(argv)
    print(output_results_models(a.bSize, a.dirPath, a.every, a.output, a.start))


if __name__ == "__main__":
    main()
