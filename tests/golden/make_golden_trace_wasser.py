"""Loss trace of the reference's OWN esrgan.train (unmodified, imported from /root/reference) with the sorted-Wasserstein
/ Sinkhorn term switched on (esrgan.py:539-542, utils.py:147-255) -> golden_trace_wasser_v1.json.

Run in the build container only:  python tests/golden/make_golden_trace_wasser.py
Same harness and layout as make_golden_trace.py --variants (B = 4, one RRDB, 6 GAN steps).
"""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import ref_harness  # noqa: E402

if __name__ == "__main__":
    r = ref_harness.run("/root/reference", None, n_batches=6, batch_size=4, seed=977, factor=4, residual_blocks=1,
                        overrides=dict(lambda_wasser=0.002))
    r.pop("stdout_tail")
    r.pop("step_t", None)
    r["models_file"] = "reference models.py"
    print({k: v for k, v in r["loss"].items() if any(v)})
    with open(os.path.join(HERE, "golden_trace_wasser_v1.json"), "w") as f:
        json.dump({"wasser": r}, f)
