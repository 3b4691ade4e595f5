"""Loss trace of the reference's OWN esrgan.train (unmodified, imported from /root/reference) -> golden_trace_v1.json.

Run in the build container only:  python tests/golden/make_golden_trace.py
B = 16, 20 GAN steps, options/default.json with factor 4 and no warm-up (SURVEY.md section 8c harness, see
tests/ref_harness.py).  The fixture holds info['loss'] (23 series), the batch order, the gradient-penalty draws and
checksums of the initial weights; images and weights are re-derived from the seed.  It pins oracle.generator_loss /
discriminator_loss (row a8/a9 of SURVEY.md section 8) and is what the GPU loss-trace tests compare against.
"""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import ref_harness  # noqa: E402

VARIANTS = {      # small runs of the non-default training recipes the fused step also covers
    "scaling_power_0.5": dict(scaling_power=0.5),                   # esrgan.py:470-473: the "pow" branch on x ** 0.5
    "second_discriminators": dict(second_discr_reset_interval=4),   # esrgan.py:221-237,495-505,789-795
    "optional_losses": dict(lambda_nnz=1e-4, lambda_mask=0.5, lambda_hit=2.0),          # esrgan.py:530-537,546-550
    "hist": dict(lambda_hist=0.1, warmup_batches=2, bins=6),        # esrgan.py:433-452,538-545 (2 warm-up steps first)
}

if __name__ == "__main__":
    if "--variants" not in sys.argv:
        res = ref_harness.run("/root/reference", None, n_batches=20, batch_size=16, seed=4321, factor=4, residual_blocks=10)
        res.pop("stdout_tail")
        res.pop("step_t", None)
        res["models_file"] = "reference models.py"
        with open(os.path.join(HERE, "golden_trace_v1.json"), "w") as f:
            json.dump(res, f)
        print({k: (v[0], v[-1]) for k, v in res["loss"].items() if any(v)})
    out = {}
    for name, ov in VARIANTS.items():
        r = ref_harness.run("/root/reference", None, n_batches=6, batch_size=4, seed=977, factor=4, residual_blocks=1,
                            overrides=ov)
        r.pop("stdout_tail")
        r.pop("step_t", None)
        r["models_file"] = "reference models.py"
        out[name] = r
        print(name, {k: (v[0], v[-1]) for k, v in r["loss"].items() if any(v)})
    with open(os.path.join(HERE, "golden_trace_variants_v1.json"), "w") as f:
        json.dump(out, f)
