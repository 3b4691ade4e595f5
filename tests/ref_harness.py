"""Harness that runs the reference's OWN ``esrgan.train(opt)`` unmodified.  TEST INFRASTRUCTURE.

    python tests/ref_harness.py --ref /root/reference --out trace.json                 # reference models, CPU (fixture)
    python tests/ref_harness.py --ref baseline/_ref --models super-resolution_b200 \
           --out trace.json                                                             # drop-in models on the GPU

Nothing in the reference tree is edited; everything below is harness-side:
  * CWD = the reference root (options/default.py:6 opens a CWD-relative path);
  * stub modules for the plotting / physics imports the train path never calls (matplotlib, energyflow, h5py, tables);
  * torch.optim.lr_scheduler.ReduceLROnPlateau without the ``verbose=`` keyword torch 2.11 dropped (esrgan.py:308,310);
  * a ``txt`` dataset (datasets.py:148-167) of synthetic sparse jet images written to a temp directory from a seed;
  * ``opt`` = options/default.json + the keys esrgan.py reads that the json lacks + the overrides given here;
  * ``--models DIR`` puts DIR before the reference root on sys.path, so ``from models import *`` (esrgan.py:17,
    datasets.py:12, evaluation/eval.py:11) resolves to the drop-in.
The harness records what the run consumed (batch order, the interpolation draws of esrgan.py:598, checksums of the
initial weights) so the fused train step can be replayed on exactly the same inputs.
"""
import argparse
import contextlib
import io
import json
import os
import sys
import tempfile
from collections import namedtuple
from unittest import mock

import numpy as np

EXTRA_KEYS = dict(use_transposed_conv=False, fully_transposed_conv=False, num_final_res_blocks=0, wasserstein=-1,
                  set_zero_def=[], set_zero_pow=[], split_eval=False, nth_jet_eval_mode="hr")
STUBS = ["matplotlib", "matplotlib.pyplot", "matplotlib.colors", "matplotlib.ticker", "matplotlib.patches",
         "matplotlib.lines", "matplotlib.cm", "matplotlib.gridspec", "mpl_toolkits", "mpl_toolkits.axes_grid1",
         "energyflow", "h5py", "tables", "pyjet"]
REF_MODULES = ("models", "esrgan", "datasets", "utils", "options", "options.default", "evaluation", "evaluation.eval",
               "evaluation.PSNR_SSIM_metric", "srb200")


def synthetic_hits(n_images, hr, seed):
    """Per image: 20..60 distinct flat indices and Exp(5) energies (legacy RandomState: stable across numpy versions)."""
    rs = np.random.RandomState(seed)
    out = []
    for _ in range(n_images):
        n = int(rs.randint(20, 61))
        idx = rs.choice(hr * hr, size=n, replace=False)
        e = rs.exponential(5.0, size=n).astype(np.float32)
        out.append((idx.astype(np.int64), e))
    return out


def hits_to_batch(hits, hr, factor):
    """The images EventDatasetText.__getitem__ builds from the files (datasets.py:136-145,160-166), as a torch batch."""
    import torch
    imgs = torch.zeros(len(hits), 1, hr, hr)
    for b, (idx, e) in enumerate(hits):
        for i, v in zip(idx.tolist(), e.tolist()):
            imgs[b, 0, i % hr, i // hr] += float(np.float32(v))
    lr = (factor * factor) * torch.nn.functional.avg_pool2d(imgs, factor)
    return imgs, lr


def write_txt_dataset(path, hits):
    os.makedirs(path, exist_ok=True)
    for i, (idx, e) in enumerate(hits):
        with open(os.path.join(path, "img_%05d.txt" % i), "w") as f:
            for a, b in zip(idx.tolist(), e.tolist()):
                f.write("%d 0 %.9g\n" % (a, b))


def run(ref_root, models_dir=None, *, n_batches=20, batch_size=16, seed=4321, factor=4, residual_blocks=10,
        overrides=None, quiet=True, preload=False):
    """``preload``: build every dataset item once before train() starts and serve it from memory afterwards, so the
    per-step times in the result (``step_t``: perf_counter at the first gradient-penalty draw of each step) measure
    the training iteration, not np.loadtxt (bench.py's reference arm uses this)."""
    import torch
    ref_root = os.path.abspath(ref_root)
    if not os.path.isdir(ref_root):
        raise FileNotFoundError(ref_root)
    tmp = tempfile.mkdtemp(prefix="srb_refrun_")
    hits = synthetic_hits(n_batches * batch_size, 80, seed)
    write_txt_dataset(os.path.join(tmp, "data"), hits)

    saved_path, saved_cwd = list(sys.path), os.getcwd()
    saved_mods = {k: sys.modules.pop(k) for k in list(sys.modules)
                  if k in REF_MODULES or k.startswith(("evaluation.", "options.", "srb200."))}
    stubs = {}
    for name in STUBS:
        if name not in sys.modules:
            stubs[name] = sys.modules[name] = mock.MagicMock(name=name)
    real_sched = torch.optim.lr_scheduler.ReduceLROnPlateau

    class Sched(real_sched):
        def __init__(self, optimizer, *a, verbose=None, **kw):
            super().__init__(optimizer, *a, **kw)

    import time as _time
    rec = {"epsilons": [], "order": [], "t": []}
    real_rand = torch.rand

    def rec_rand(*a, **kw):
        t = real_rand(*a, **kw)
        if t.dim() == 4 and t.shape[1:] == (1, 1, 1):       # esrgan.py:598
            rec["epsilons"].append(t.reshape(-1).tolist())
            rec["t"].append(_time.perf_counter())
        return t

    try:
        os.chdir(ref_root)
        sys.path.insert(0, ref_root)
        if models_dir:
            sys.path.insert(0, os.path.abspath(models_dir))
        torch.optim.lr_scheduler.ReduceLROnPlateau = Sched
        torch.rand = rec_rand
        out = io.StringIO()
        with (contextlib.redirect_stdout(out) if quiet else contextlib.nullcontext()):
            import esrgan
            models_file = sys.modules["models"].__file__
            real_get = esrgan.get_dataset

            def get_dataset(*a, **kw):
                ds = real_get(*a, **kw)
                cls = type(ds)
                real_item = cls.__getitem__

                cache = {}

                class Recording(cls):
                    def __getitem__(self, i):
                        rec["order"].append(int(i))
                        if i in cache:
                            return {k: v.clone() for k, v in cache[i].items()}
                        return real_item(self, i)
                if preload:
                    for i in range(len(ds)):
                        cache[i] = real_item(ds, i)
                ds.__class__ = Recording
                return ds

            esrgan.get_dataset = get_dataset
            # initial weights are a function of the seed and of the constructors' draw order: checksum them
            real_adam = torch.optim.Adam
            sums = []

            def adam(params, *a, **kw):
                params = list(params)
                sums.append([float(sum(p.detach().double().sum().item() for p in params)),
                             float(sum(p.detach().double().abs().sum().item() for p in params))])
                return real_adam(params, *a, **kw)

            torch.optim.Adam = adam
            d = dict(esrgan.default_dict)
            d.update(EXTRA_KEYS)
            d.update(dataset_type="txt", dataset_path=os.path.join(tmp, "data"), factor=factor, batch_size=batch_size,
                     residual_blocks=residual_blocks, warmup_batches=0, sample_interval=-1, save=False, save_info=True,
                     n_batches=n_batches, set_seed=seed, report_freq=1, root=tmp, n_epochs=1,
                     model_path=os.path.join(tmp, "saved_models"))      # info.json (the loss trace) goes there
            d.update(overrides or {})
            opt = namedtuple("opt", d.keys())(*d.values())
            try:
                esrgan.train(opt)
            finally:
                torch.optim.Adam = real_adam
        # train() returns info only when batches_done reaches total_batches (esrgan.py:810), one batch more than the
        # data holds; the trace is what save_info() wrote at the end of the epoch (esrgan.py:366-371,813)
        with open(os.path.join(tmp, "saved_models", "_info.json")) as f:      # model_name = opt.name + '_' (esrgan.py:153)
            info = json.load(f)
        loss = {k: [float(x) for x in v] for k, v in (info.get("loss") or {}).items()}
        extra = {k: v for k, v in info.items() if k.startswith("binedges")}
        if not loss:
            raise RuntimeError("reference train() returned no loss trace")
        return {"loss": loss, "order": rec["order"], "epsilons": rec["epsilons"], "init_checksums": sums,
                "step_t": rec["t"][0::2], "info_extra": extra,
                "cfg": dict(n_batches=n_batches, batch_size=batch_size, seed=seed, factor=factor,
                            residual_blocks=residual_blocks, overrides=overrides or {}),
                "models_file": models_file, "device": "cuda" if torch.cuda.is_available() else "cpu",
                "stdout_tail": out.getvalue()[-2000:]}
    finally:
        torch.rand = real_rand
        torch.optim.lr_scheduler.ReduceLROnPlateau = real_sched
        os.chdir(saved_cwd)
        sys.path[:] = saved_path
        for k in list(sys.modules):
            if k in REF_MODULES or k.startswith(("evaluation.", "options.")):
                sys.modules.pop(k, None)
        for k in stubs:
            sys.modules.pop(k, None)
        sys.modules.update(saved_mods)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ref", required=True)
    ap.add_argument("--models", default=None)
    ap.add_argument("--out", required=True)
    ap.add_argument("--n-batches", type=int, default=20)
    ap.add_argument("--batch-size", type=int, default=16)
    ap.add_argument("--seed", type=int, default=4321)
    ap.add_argument("--factor", type=int, default=4)
    ap.add_argument("--residual-blocks", type=int, default=10)
    ap.add_argument("--overrides", default="{}")
    ap.add_argument("--verbose", action="store_true")
    ap.add_argument("--preload", action="store_true")
    ap.add_argument("--threads", type=int, default=0)
    a = ap.parse_args()
    if a.threads:
        import torch
        torch.set_num_threads(a.threads)
    res = run(a.ref, a.models, n_batches=a.n_batches, batch_size=a.batch_size, seed=a.seed, factor=a.factor,
              residual_blocks=a.residual_blocks, overrides=json.loads(a.overrides), quiet=not a.verbose,
              preload=a.preload)
    with open(a.out, "w") as f:
        json.dump(res, f)
    print("wrote", a.out, {k: v[-1] for k, v in res["loss"].items() if v and k in ("g_loss", "d_loss_def", "d_loss_pow")})


if __name__ == "__main__":
    main()
