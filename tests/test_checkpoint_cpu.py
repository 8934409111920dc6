"""On-disk format: the reference's checkpoints (utils.py:84-128) map onto the engine's parameter keys.
No GPU: only the host-side key / order logic is exercised here; tests/test_score_gpu.py loads a checkpoint into the
engine."""
import hashlib
import json
import os

import pytest
import torch

from b_pinn_kalman_filter_b200 import configs
from b_pinn_kalman_filter_b200.models import utils as mutils
import b_pinn_kalman_filter_b200.models  # noqa: F401  (registers ncsnpp / ddpm)

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "param_order.json")


@pytest.mark.parametrize("name", ["ve/cifar10_ncsnpp_continuous", "vp/nc_ddpmpp", "ve/celebahq_256_ncsnpp_continuous",
                                  "vp/cifar10_ncsnpp_continuous"])
def test_trainable_parameter_order_matches_reference_registration_order(name):
    """Golden = names of the reference model's `parameters()` (requires_grad) in order, made by
    oracle/make_golden.py gen_param_order from the unmodified reference."""
    with open(GOLDEN) as f:
        gold = json.load(f)[name]
    order = mutils.trainable_parameter_order(configs.get_config(name))
    assert len(order) == gold["count"]
    assert order[:12] == gold["first"] and order[-4:] == gold["last"]
    assert hashlib.sha1("\n".join(order).encode()).hexdigest() == gold["sha1"]


def _fake_checkpoint(cfg, seed=0):
    """A checkpoint dict laid out as utils.save_checkpoint writes it, with distinct model / EMA values."""
    from b_pinn_kalman_filter_b200.models.engine import parameter_spec
    spec, _ = parameter_spec(cfg, mutils.get_model(cfg.model.name))
    g = torch.Generator().manual_seed(seed)
    model = {"module." + k: torch.randn(shape, generator=g) for k, shape, _ in spec}
    model["module.sigmas"] = torch.tensor(mutils.get_sigmas(cfg), dtype=torch.float32)
    order = mutils.trainable_parameter_order(cfg)
    shapes = {k: s for k, s, _ in spec}
    shadow = [torch.randn(shapes[k], generator=g) for k in order]
    return {"info": 1, "optimizer": {}, "model": model, "step": 7,
            "ema": {"decay": 0.999, "num_updates": 7, "shadow_params": shadow}}, order, shadow


def test_checkpoint_state_dict_model_and_ema(tmp_path):
    cfg = configs.get_config("test/tiny_ncsnpp")
    ck, order, shadow = _fake_checkpoint(cfg)
    path = tmp_path / "checkpoint_3.pth"
    torch.save(ck, path)
    sd = mutils.checkpoint_state_dict(str(path), cfg)
    assert set(sd) == {k[len("module."):] for k in ck["model"]}
    for k, v in sd.items():
        assert torch.equal(v, ck["model"]["module." + k])
    se = mutils.checkpoint_state_dict(str(path), cfg, use_ema=True)
    assert set(se) == set(sd)
    for k, v in zip(order, shadow):
        assert torch.equal(se[k], v)
    # what EMA does not track comes from the model entry: sigmas and (Fourier nets) the frozen projection
    assert torch.equal(se["sigmas"], sd["sigmas"])
    for k in set(sd) - set(order):
        assert torch.equal(se[k], sd[k])
    bad = dict(ck)
    bad["ema"] = dict(ck["ema"], shadow_params=shadow[:-1])
    with pytest.raises(RuntimeError):
        mutils.ema_state_dict(cfg, bad["ema"], base=ck["model"])


def test_load_checkpoint_missing_file_returns_model_unchanged(tmp_path, caplog):
    class Dummy:
        config = None
    m = Dummy()
    assert mutils.load_checkpoint(str(tmp_path / "nope.pth"), m) is m


def test_ema_mapping_equals_reference_copy_to(reference):
    """With the real reference: ExponentialMovingAverage.copy_to(model.parameters()) then model.state_dict() ==
    checkpoint_state_dict(..., use_ema=True) of the checkpoint utils.save_checkpoint would write."""
    import importlib
    from oracle import ref_import
    ema_mod = importlib.import_module("models.ema")
    rcfg = ref_import.get_config("configs.vp.nc_ddpmpp")
    rcfg.device = torch.device("cpu")
    rcfg.model.update(dict(nf=32, ch_mult=(1, 2), num_res_blocks=1, attn_resolutions=(8,)))
    cfg = configs.get_config("vp/nc_ddpmpp")
    cfg.model.update(dict(nf=32, ch_mult=(1, 2), num_res_blocks=1, attn_resolutions=(8,)))
    torch.manual_seed(3)
    model = torch.nn.DataParallel(reference.mutils.get_model(rcfg.model.name)(rcfg))
    ema = ema_mod.ExponentialMovingAverage(model.parameters(), decay=0.5)
    with torch.no_grad():
        for p in model.parameters():
            p.add_(torch.randn_like(p))
    ema.update(model.parameters())
    ck = {"info": 1, "optimizer": {}, "model": {k: v.clone() for k, v in model.state_dict().items()},
          "ema": ema.state_dict(), "step": 1}
    mine = mutils.checkpoint_state_dict(ck, cfg, use_ema=True)
    ema.copy_to(model.parameters())
    ref_sd = model.state_dict()
    assert set(mine) == {k[len("module."):] for k in ref_sd}
    for k, v in mine.items():
        assert torch.equal(v, ref_sd["module." + k]), k
