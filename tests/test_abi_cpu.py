"""CPU-side checks: the C-ABI library builds/loads and exports every symbol the binding declares; the drop-in
modules keep the reference's state_dict layout; the hot path refuses to run without CUDA (no fallback)."""
import ctypes
import os
import re

import pytest
import torch

from oracle import oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_declared_symbols():
    from srb200 import _lib as L
    L.build()
    lib = L.lib()
    assert lib.srb_abi_version() == 3
    hdr = open(os.path.join(ROOT, "include", "srb200.h")).read()
    declared = set(re.findall(r"\b(srb_[a-z0-9_]+)\s*\(", hdr))
    assert declared, "header declares no entry points?"
    for name in declared | set(L.exported_symbols()):
        assert hasattr(lib, name), f"libsrb200.so lacks {name}"
    assert declared >= set(L.exported_symbols()), \
        f"binding uses symbols the header does not declare: {set(L.exported_symbols()) - declared}"


def test_state_dict_layout_matches_reference_shapes():
    import models
    g = models.GeneratorRRDB(1, filters=64, num_res_blocks=2, num_upsample=2, res_scale=0.1)
    shapes = O.generator_shapes(1, 64, 2, 2)
    sd = g.state_dict()
    assert list(sd) == list(shapes)
    assert all(tuple(sd[k].shape) == tuple(shapes[k]) for k in shapes)
    d = models.Markovian_Discriminator((1, 80, 80), [16, 32, 32, 64])
    ds = O.discriminator_shapes()
    assert list(d.state_dict()) == list(ds) and d.output_shape == (1, 5, 5)
    assert [n for n, p in g.named_parameters() if not p.requires_grad] == ["power", "multiplier"]


def test_default_generator_has_314_entries():
    import models
    g = models.GeneratorRRDB(1, filters=64, num_res_blocks=10, num_upsample=2)
    assert len(g.state_dict()) == 314       # SURVEY.md section 5, checkpoint layout


def test_no_cpu_fallback():
    import models
    g = models.GeneratorRRDB(1, filters=64, num_res_blocks=1, num_upsample=1)
    with pytest.raises(RuntimeError, match="CUDA"):
        g(torch.zeros(1, 1, 8, 8))
    d = models.Markovian_Discriminator((1, 16, 16), [16])
    with pytest.raises(RuntimeError, match="CUDA"):
        d(torch.zeros(1, 1, 16, 16))
    with pytest.raises(RuntimeError, match="CUDA"):
        models.SumPool2d(2)(torch.zeros(1, 1, 4, 4))


def test_weight_reset_and_uniform_reset_apply():
    import models
    g = models.GeneratorRRDB(1, filters=64, num_res_blocks=1, num_upsample=1, uniform_init=True)
    assert float(g.conv1.bias.abs().sum()) == 0.0
    before = g.conv2.weight.clone()
    g.apply(models.weight_reset)
    assert not torch.equal(before, g.conv2.weight)
    g.apply(models.uniform_reset)
    assert float(g.conv3[0].bias.abs().sum()) == 0.0


def test_chain_flag_arena_and_default():
    """host logic of the chained launches: the flag arena hands out disjoint zeroed slices and refuses to overrun;
    SRB_CONV_CHAIN parses to a clamped integer (default 5 = one launch per dense block)."""
    import os
    import torch
    from srb200 import generator as G
    a = G._FlagArena(10, "cpu")
    x, y = a.take(4), a.take(6)
    assert x.numel() == 4 and y.numel() == 6 and x.data_ptr() != y.data_ptr()
    x.fill_(7)
    with pytest.raises(RuntimeError, match="exhausted"):
        a.take(1)
    a.reset()
    assert int(a.t.abs().sum()) == 0 and a.take(10).numel() == 10
    old = os.environ.get("SRB_CONV_CHAIN")
    try:
        os.environ.pop("SRB_CONV_CHAIN", None)
        assert G._chain_default() == 5
        os.environ["SRB_CONV_CHAIN"] = "99"
        assert G._chain_default() == 16
        os.environ["SRB_CONV_CHAIN"] = "0"
        assert G._chain_default() == 0
    finally:
        if old is None:
            os.environ.pop("SRB_CONV_CHAIN", None)
        else:
            os.environ["SRB_CONV_CHAIN"] = old


def test_data_path_and_eval_stats_have_no_cpu_fallback():
    """the widened rows follow the same rule as the hot path: CPU tensors are refused loudly"""
    from srb200 import data
    pairs = O.synthetic_constituents(2, 16, 80, 80, 1)
    with pytest.raises(RuntimeError, match="CUDA"):
        data.extract_batch(pairs, 80, 80, 4)
    with pytest.raises(RuntimeError, match="CUDA"):
        data.image_stats(torch.zeros(2, 1, 80, 80), 0.0, 2)
