"""Drop-in for the reference's `sndcgan/Trainer.py` CLI and the non-kernel parts of its `SNDCGAN` class
(sndcgan/SNDCGAN.py:131-212, :272-335): dataset iterator, output directory handling, losses pickle, checkpoints,
epoch loop.  The training step itself is SNDCGANEngine.train_step (imagegeneration_b200/sndcgan.py).

    python -m imagegeneration_b200.trainer <bSize> <epochs> [-cf N] [-d DIR] [-x DATA] [-r RATE] [-ld LR] [-lg LR]
                                                           [-lo NAME] [-ct]   (+ opt-in extras, defaults = reference)
"""
from __future__ import annotations

import argparse
import glob
import math
import os
import pickle
import shutil
from time import gmtime, strftime, time

import numpy as np
import torch

from .sndcgan import SNDCGANEngine, create_samples

img_height = 144          # sndcgan/Trainer.py:5-8
img_width = 256
image_size = (img_height, img_width, 3)
z_size = 128


# ----------------------------------------------------------------------------------------------------------
# dataset  (SNDCGAN.py:131-145)
# ----------------------------------------------------------------------------------------------------------
def _decode_folder(root, size):
    """image_dataset_from_directory(image_size=..., crop_to_aspect_ratio=True): every file is decoded ONCE (PIL), the
    largest centred window with the target aspect ratio is cropped and resized with half-pixel BILINEAR interpolation
    WITHOUT antialiasing (what Keras' smart_resize / tf.image.resize do; PIL's BILINEAR would antialias when shrinking).
    Returns uint8 [N, H, W, 3] — the resized floats are rounded to 8 bits for the cache (<= 0.5/255 per pixel)."""
    from PIL import Image
    exts = ("*.jpg", "*.jpeg", "*.png", "*.bmp")
    files = sorted(f for e in exts for f in glob.glob(os.path.join(root, "**", e), recursive=True))
    if not files:
        raise FileNotFoundError(f"no images under {root!r} (use --data synthetic for generated data)")
    h, w = size[:2]
    imgs = torch.empty((len(files), h, w, 3), dtype=torch.uint8)
    for i, f in enumerate(files):
        im = Image.open(f).convert("RGB")
        iw, ih = im.size
        if iw * h > ih * w:      # too wide: crop width
            cw = int(ih * w / h)
            x0 = (iw - cw) // 2
            im = im.crop((x0, 0, x0 + cw, ih))
        else:
            ch = int(iw * h / w)
            y0 = (ih - ch) // 2
            im = im.crop((0, y0, iw, y0 + ch))
        t = torch.from_numpy(np.asarray(im).copy()).permute(2, 0, 1)[None].float()
        t = torch.nn.functional.interpolate(t, size=(h, w), mode="bilinear", align_corners=False, antialias=False)
        imgs[i] = t[0].permute(1, 2, 0).round().clamp(0, 255).to(torch.uint8)
    return imgs


class DeviceImageDataset:
    """get_dataset's `Rescaling(1/127.5, offset=-1)` + `.cache().shuffle(10000).prefetch()` (SNDCGAN.py:141-143) with the
    cache RESIDENT IN HBM as uint8 (the reference's 7 244 usable images at 144x256x3 are 0.8 GB): an epoch is a device-side
    permutation of indices, a batch is one b2g_gather_u8_to_f32 launch (gather + /127.5 - 1 into a fresh fp32 NHWC
    tensor).  No host work, no H2D copy per step.  Like tf.data (drop_remainder=False) the last batch may be ragged.
    Yields (images fp32 CUDA [b,H,W,3] in [-1,1], labels int64 zeros [b])."""

    def __init__(self, images_u8: torch.Tensor, batch_size: int, device="cuda", seed=123):
        if images_u8.dtype != torch.uint8 or images_u8.dim() != 4 or images_u8.shape[3] != 3:
            raise ValueError("DeviceImageDataset needs uint8 [N,H,W,3] images")
        self.device = torch.device(device)
        self.images = images_u8.contiguous().to(self.device)
        self.batch_size = batch_size
        self.image_size = tuple(self.images.shape[1:])
        self.per_image = int(self.images[0].numel())
        if self.per_image % 16:
            raise ValueError("image bytes (H*W*3) must be a multiple of 16")
        self.gen = torch.Generator(device=self.device).manual_seed(seed)

    def __len__(self):
        return (self.images.shape[0] + self.batch_size - 1) // self.batch_size

    def __iter__(self):
        from . import lib as _lib
        n = self.images.shape[0]
        perm = torch.randperm(n, generator=self.gen, device=self.device)
        for i in range(0, n, self.batch_size):
            idx = perm[i:i + self.batch_size].contiguous()
            b = idx.numel()
            out = torch.empty((b, *self.image_size), dtype=torch.float32, device=self.device)
            _lib.check(_lib.load().b2g_gather_u8_to_f32(
                self.images.data_ptr(), idx.data_ptr(), out.data_ptr(), b, self.per_image,
                torch.cuda.current_stream().cuda_stream), "b2g_gather_u8_to_f32")
            yield out, torch.zeros(b, dtype=torch.int64, device=self.device)


def get_dataset(dataset, batch_size, image_size, device="cuda"):
    """SNDCGAN.py:131: directory of images -> iterable of (fp32 NHWC in [-1,1], labels).  "synthetic[:N]" -> N (default
    7 244 = the reference's usable set, dataset_creator/results_open_images/stats.json) generated uint8 images, because
    Open Images is not available offline.  Either way the cache lives on the device (DeviceImageDataset)."""
    if dataset == "synthetic" or str(dataset).startswith("synthetic:"):
        n = int(str(dataset).split(":")[1]) if ":" in str(dataset) else 7244
        g = torch.Generator().manual_seed(123)
        imgs = torch.randint(0, 256, (n, *tuple(image_size)), generator=g, dtype=torch.uint8)
    else:
        imgs = _decode_folder(dataset, tuple(image_size))
    return DeviceImageDataset(imgs, batch_size, device=device)


# ----------------------------------------------------------------------------------------------------------
# persistence  (SNDCGAN.py:181-199, :317-331)
# ----------------------------------------------------------------------------------------------------------
def save_model(model, path):
    """`model.save(path, include_optimizer=False)` (SNDCGAN.py:327-331): architecture arguments + the trainable variables
    in Keras order and Keras layouts, as one torch file (h5py / TensorFlow are not available: no .h5 byte compatibility)."""
    os.makedirs(os.path.dirname(path), exist_ok=True)
    from .sndcgan import Generator
    state = {"kind": "generator" if isinstance(model, Generator) else "discriminator",
             "image_size": tuple(model.output_size if isinstance(model, Generator) else model.input_size),
             "names": list(model.arena.names), "weights": [v.detach().cpu().clone() for v in model.trainable_variables]}
    if state["kind"] == "discriminator":
        state["dropout_rate"] = model.dropout_rate
        state["spectral_norm"] = model.spectral_norm
        if model.spectral_norm:
            if model._sn_stale:
                model.sn_power_iteration()
            state["sn_u"] = [u.cpu() for u in model.sn_u]
    torch.save(state, path)


def load_model(path, device="cuda"):
    """tf.keras.models.load_model(path) for files written by save_model (generator_output.py:67, generator_evaluation.py:
    134,160): returns a ready Generator / Discriminator."""
    from .sndcgan import make_dcgan_discriminator, make_dcgan_generator
    st = torch.load(path, map_location="cpu")
    if st["kind"] == "generator":
        model = make_dcgan_generator(st["image_size"], device=device)
    else:
        model = make_dcgan_discriminator(st["dropout_rate"], st["image_size"], device=device,
                                         spectral_norm=st.get("spectral_norm", False))
        if st.get("sn_u"):
            for u, s in zip(model.sn_u, st["sn_u"]):
                u.copy_(s)
    model.set_weights(st["weights"])
    return model


class CheckpointManager:
    """tf.train.CheckpointManager(ckpt, directory, max_to_keep=2) as the reference uses it (SNDCGAN.py:181-199,318):
    `latest_checkpoint`, `checkpoints`, `save(checkpoint_number=epoch)`; plus `restore(path)` (Checkpoint.restore).
    Files are torch.save archives `ckpt-<n>.pt` holding both models, both optimizers' slots and counters."""

    def __init__(self, owner, directory, max_to_keep=2):
        self.owner, self.directory, self.max_to_keep = owner, directory, max_to_keep
        os.makedirs(directory, exist_ok=True)

    @property
    def checkpoints(self):
        files = glob.glob(os.path.join(self.directory, "ckpt-*.pt"))
        return sorted(files, key=lambda f: int(os.path.basename(f).split("-")[-1].split(".")[0]))

    @property
    def latest_checkpoint(self):
        c = self.checkpoints
        return c[-1] if c else None

    def save(self, checkpoint_number):
        path = os.path.join(self.directory, f"ckpt-{checkpoint_number}.pt")
        self.owner._save_state(path, checkpoint_number)
        if self.max_to_keep:
            for old in self.checkpoints[:-self.max_to_keep]:
                os.remove(old)
        return path

    def restore(self, path):
        self.owner._restore(path)


# ----------------------------------------------------------------------------------------------------------
# SNDCGAN  (SNDCGAN.py:148-335)
# ----------------------------------------------------------------------------------------------------------
class SNDCGAN(SNDCGANEngine):
    def __init__(self, dir_path, dataset, batch_size, dropout, learning_rate_disc, learning_rate_gen, live_output,
                 continue_, image_size, z_size, **engine_kwargs):
        if not continue_ and os.path.exists(dir_path):          # :152-153 (yes, it deletes the directory)
            shutil.rmtree(dir_path)
        if not os.path.exists(dir_path):
            os.mkdir(dir_path)
        self.dir_path = dir_path
        self.train_ds = get_dataset(dataset, batch_size, image_size)
        super().__init__(batch_size, dropout=dropout, learning_rate_disc=learning_rate_disc,
                         learning_rate_gen=learning_rate_gen, image_size=image_size, z_size=z_size, **engine_kwargs)
        self.live_preview_file = os.path.join(live_output + ".png")
        self.losses_file = os.path.join(dir_path, "losses.pickle")
        if os.path.exists(self.losses_file):
            with open(self.losses_file, mode="rb") as f:
                self.losses = pickle.load(f)
        else:
            self.losses = {"epoch": [], "avg_g_loss": [], "avg_d_loss": [], "d_real": [], "d_fake": []}
        self.checkpoint_path = os.path.join(dir_path, "checkpoints")
        self.ckpt_manager = CheckpointManager(self, self.checkpoint_path, max_to_keep=2)      # :187-188
        latest = self.ckpt_manager.latest_checkpoint
        if continue_ and latest:                                 # :192-196
            self.start_epoch = int(os.path.basename(latest).split("-")[-1].split(".")[0]) + 1
            self.ckpt_manager.restore(latest)
            print("Latest checkpoint restored!!")
        else:
            self.start_epoch = 0
            print("No checkpoints were restored!!")
        print("\nInitialized SNDCGAN successfully!\n")

    # -- checkpoint contents (written / read through self.ckpt_manager)
    def _save_state(self, path, epoch):
        self.gather_optimizer_state()       # NVLS data parallel shards the discriminator's Adam state (no-op otherwise)
        G, D = self.gen_model.arena, self.disc_model.arena
        state = {"epoch": epoch, "g_w": G.w.cpu(), "g_m": G.m.cpu(), "g_v": G.v.cpu(), "d_w": D.w.cpu(),
                 "d_m": D.m.cpu(), "d_v": D.v.cpu(), "g_iterations": self.g_optimizer.iterations,
                 "d_iterations": self.d_optimizer.iterations, "step_count": self.step_count}
        if self.disc_model.spectral_norm:
            D = self.disc_model
            if D._sn_stale:                 # never trained / weights set by hand: make sigma valid before saving it
                D.sn_power_iteration()
            state["sn_u"] = [u.cpu() for u in D.sn_u]
            state["sn_vraw"] = [x.cpu() for x in D.sn_vraw]
            state["sn_scal"] = [x.cpu() for x in D.sn_scal]
        torch.save(state, path)

    def _restore(self, path):
        st = torch.load(path, map_location="cpu")
        G, D = self.gen_model.arena, self.disc_model.arena
        for dst, key in ((G.w, "g_w"), (G.m, "g_m"), (G.v, "g_v"), (D.w, "d_w"), (D.m, "d_m"), (D.v, "d_v")):
            dst.copy_(st[key])
        self.g_optimizer.iterations, self.d_optimizer.iterations = st["g_iterations"], st["d_iterations"]
        self.step_count = st["step_count"]
        D = self.disc_model
        if D.spectral_norm:
            D._sn_stale = True              # without a saved sigma the first forward runs a power iteration
            if "sn_u" in st:
                for u, s in zip(D.sn_u, st["sn_u"]):
                    u.copy_(s)
            if "sn_scal" in st and "sn_vraw" in st:   # sigma / v of the saved weights: resume is exact
                for dst, src in zip(D.sn_vraw + D.sn_scal, st["sn_vraw"] + st["sn_scal"]):
                    dst.copy_(src)
                D._sn_stale = False
        self.gen_model._packed_dirty = True
        self.disc_model._packed_dirty = True

    def plot_history(self):
        """:216-225 draws the loss curves with matplotlib (absent here, and there is no display): the same four curves
        go to <dir>/plot_line_plot_loss.png as a plain Pillow line plot."""
        try:
            from PIL import Image, ImageDraw
            W, H, pad = 800, 400, 30
            img = Image.new("RGB", (W, H), "white")
            dr = ImageDraw.Draw(img)
            series = [(k, v) for k, v in self.losses.items() if k != "epoch" and len(v)]
            vals = [x for _k, v in series for x in v if math.isfinite(x)]
            if vals:
                lo, hi = min(vals), max(vals)
                hi = hi if hi > lo else lo + 1.0
                colors = ["#1f77b4", "#ff7f0e", "#2ca02c", "#d62728"]
                for ci, (k, v) in enumerate(series):
                    n = max(len(v) - 1, 1)
                    pts = [(pad + (W - 2 * pad) * i / n, H - pad - (H - 2 * pad) * (x - lo) / (hi - lo))
                           for i, x in enumerate(v) if math.isfinite(x)]
                    if len(pts) > 1:
                        dr.line(pts, fill=colors[ci % 4], width=2)
                    dr.text((pad + 5, 5 + 12 * ci), f"{k} [{lo:.3g}, {hi:.3g}]", fill=colors[ci % 4])
            img.save(os.path.join(self.dir_path, "plot_line_plot_loss.png"))
        except Exception as e:  # a plot is a convenience, never fatal
            print(f"(loss plot not written: {e})")

    def display_examples(self, samples, number_of_images, info_text):
        """:228-238 writes a matplotlib PDF; headless here: the preview strip goes to a PNG (Pillow)."""
        try:
            from PIL import Image
            strip = (samples.clamp(0, 1) * 255).to(torch.uint8).cpu().numpy()
            Image.fromarray(np.concatenate(list(strip[:number_of_images]), axis=1)).save(self.live_preview_file)
        except Exception as e:  # previews are a convenience, never fatal
            print(f"(preview not written: {e})")

    def train(self, num_epochs, checkpoint_frequency):
        """SNDCGAN.py:272-335 — same loop, logging keys and checkpoint cadence."""
        start_time = time()
        local_losses = {"epoch": [], "avg_g_loss": [], "avg_d_loss": [], "d_real": [], "d_fake": []}
        dev = self.device
        for epoch in range(self.start_epoch, num_epochs):
            epoch_losses, epoch_d_vals = [], []
            batch_start_time = time()
            for i, (input_real, _) in enumerate(self.train_ds):
                # a ragged last batch is trained on like any other, with batch_size latent vectors (:282-284)
                input_real = input_real.to(dev, non_blocking=True)
                input_z = torch.rand((self.batch_size, self.z_size), device=dev) * 2.0 - 1.0        # :283
                out = self.train_step(input_real, input_z)                                          # :284
                epoch_losses.append(torch.stack(out[:4]))
                epoch_d_vals.append(torch.stack((torch.sigmoid(out[4]).mean(), torch.sigmoid(out[5]).mean())))
                print(f"\r>Batch {i:03d}, passed time: {strftime('%M:%S', gmtime(time() - batch_start_time))}",
                      end="", flush=True)
            if not epoch_losses:
                raise RuntimeError("dataset yielded no batch")
            avg_losses = torch.stack(epoch_losses).mean(dim=0).cpu().tolist()     # one host sync per epoch
            for key, val in zip(("epoch", "avg_g_loss", "avg_d_loss", "d_real", "d_fake"), [epoch] + avg_losses):
                local_losses[key].append(val)
            epoch_duration = strftime("%H:%M:%S", gmtime(time() - start_time))
            info_text = "Epoch {:04d} | ET {} min | Avg Losses G/D {:.4f}/{:.4f} [D-Real: {:.4f} D-Fake {:.4f}]".format(
                epoch, epoch_duration, *avg_losses)
            print("\r" + info_text)
            n_prev = 3
            fixed_z = torch.rand((n_prev, self.z_size), device=dev) * 2.0 - 1.0
            self.display_examples(create_samples(self.gen_model, fixed_z, n_prev), n_prev, info_text)
            if epoch % checkpoint_frequency == 0:                                                   # :317
                self.ckpt_manager.save(checkpoint_number=epoch)                                     # :318
                for key, val in local_losses.items():
                    self.losses[key] = self.losses[key] + val
                with open(self.losses_file, mode="wb") as f:
                    pickle.dump(self.losses, f)
                # :327-331 — per-checkpoint model exports, consumed by generator_output / generator_evaluation
                save_model(self.gen_model, os.path.join(self.dir_path, "models", "generator", f"gen_model-{epoch}.pt"))
                save_model(self.disc_model,
                           os.path.join(self.dir_path, "models", "discriminator", f"disc_model-{epoch}.pt"))
                self.plot_history()                                                                 # :333
                local_losses = {"epoch": [], "avg_g_loss": [], "avg_d_loss": [], "d_real": [], "d_fake": []}


def build_parser():
    """sndcgan/Trainer.py:12-31 flag for flag; extras after the blank line default to the reference's behaviour."""
    p = argparse.ArgumentParser(description="GAN Trainer to generate landscape images.")
    p.add_argument("bSize", type=int, help="Batch Size to use.")
    p.add_argument("epochs", type=int, help="Number of epochs to train.")
    p.add_argument("-cf", "--checkpointFrequency", type=int, dest="ckptFreq", default=5)
    p.add_argument("-d", "--directory", type=str, dest="dirPath", default="training")
    p.add_argument("-x", "--data", type=str, dest="data", default="dataset")
    p.add_argument("-r", "--dropout", type=float, dest="dropout", default=0.5)
    p.add_argument("-ld", "--learnRateDisc", type=float, dest="learnRateDisc", default=0.0002)
    p.add_argument("-lg", "--learnRateGen", type=float, dest="learnRateGen", default=0.0002)
    p.add_argument("-lo", "--liveOutput", type=str, dest="liveOutput", default="live")
    p.add_argument("-ct", "--continue", dest="continue_", action="store_true", default=False)

    p.add_argument("--spectral-norm", action="store_true", help="spectral normalisation of D (not in the reference)")
    p.add_argument("--loss", choices=["bce", "hinge"], default="bce")
    p.add_argument("--seed", type=int, default=62)
    p.add_argument("--dtype", choices=["bf16", "tf32"], default="bf16",
                   help="bf16 activations (default) or the fp32-storage / kind::tf32 parity path")
    return p


def main(argv=None):
    args = build_parser().parse_args(argv)
    sndcgan = SNDCGAN(args.dirPath, args.data, args.bSize, args.dropout, args.learnRateDisc, args.learnRateGen,
                      args.liveOutput, args.continue_, image_size, z_size, spectral_norm=args.spectral_norm,
                      loss=args.loss, seed=args.seed, dtype=args.dtype)
    sndcgan.train(args.epochs + 1, args.ckptFreq)


if __name__ == "__main__":
    main()
