"""Import the UNMODIFIED reference (/root/reference) in the authoring container.

TEST INFRASTRUCTURE ONLY -- never imported by the product package.  Used by
`oracle/make_golden.py` (fixture generation) and by the `not gpu` tests that pin
the oracle restatements against the real reference when /root/reference is
present.  /root/reference does not exist on the GPU box, so nothing marked
`gpu`, `smoke()` or `bench.py` may import this module.

Shims, none of which edits the reference:
  1. `ml_collections` / `imageio` attribute-dict stubs on sys.path (oracle/shims).
  2. `torch.utils.cpp_extension.load` is stubbed while `op` is imported so the
     reference does not JIT-build its CUDA extensions (op/upfirdn2d.py:10-16,
     op/fused_act.py:11-17); on CPU tensors the reference only ever runs its own
     pure-PyTorch branches (`upfirdn2d_native`, the CPU branch of
     `fused_leaky_relu`), which is what we want as the oracle anchor.
  3. `VESDE.marginal_coef` subclass (the shipped VESDE is abstract, sde_lib.py:34-37
     vs :253-307) -- see `VESDEShim`.
"""
import importlib
import os
import sys
import types

REF = os.environ.get("B200_REFERENCE_DIR", "/root/reference")
_SHIMS = os.path.join(os.path.dirname(os.path.abspath(__file__)), "shims")


def available():
    return os.path.isdir(os.path.join(REF, "op")) and os.path.isfile(os.path.join(REF, "sampling.py"))


_loaded = {}


def load():
    """Returns a namespace with the reference modules: op, sde_lib, sampling, mutils, ncsnpp, ddpm, ..."""
    if _loaded:
        return _loaded["ns"]
    if not available():
        raise RuntimeError("reference tree not present at %s" % REF)
    import torch
    import torch.utils.cpp_extension as cpp_ext

    for p in (REF, _SHIMS):
        if p not in sys.path:
            sys.path.insert(0, p)
    # Our own package also has modules called `sampling`, `sde_lib`, `op` but only
    # INSIDE the package namespace, so the top-level names below are the reference's.
    real_load = cpp_ext.load

    def fake_load(name, sources, **kw):
        m = types.ModuleType(name)
        m.__dict__["_stub_for"] = sources
        return m

    cpp_ext.load = fake_load
    try:
        op = importlib.import_module("op")
    finally:
        cpp_ext.load = real_load
    sde_lib = importlib.import_module("sde_lib")
    sampling = importlib.import_module("sampling")
    mutils = importlib.import_module("models.utils")
    ncsnpp = importlib.import_module("models.ncsnpp")
    ddpm = importlib.import_module("models.ddpm")
    layers = importlib.import_module("models.layers")
    layerspp = importlib.import_module("models.layerspp")
    updown = importlib.import_module("models.up_or_down_sampling")
    upfirdn2d_mod = importlib.import_module("op.upfirdn2d")
    fused_act_mod = importlib.import_module("op.fused_act")

    class VESDEShim(sde_lib.VESDE):
        """Adds the missing abstract method, consistent with marginal_prob (sde_lib.py:285-288)."""

        def marginal_coef(self, t):
            return torch.ones_like(t), self.sigma_min * (self.sigma_max / self.sigma_min) ** t

    ns = types.SimpleNamespace(
        op=op, sde_lib=sde_lib, sampling=sampling, mutils=mutils, ncsnpp=ncsnpp, ddpm=ddpm,
        layers=layers, layerspp=layerspp, updown=updown, upfirdn2d_mod=upfirdn2d_mod,
        fused_act_mod=fused_act_mod, VESDEShim=VESDEShim, torch=torch)
    _loaded["ns"] = ns
    return ns


def get_config(module_name):
    """e.g. get_config('configs.ve.cifar10_ncsnpp_continuous') -> ConfigDict on CPU."""
    ns = load()
    cfg = importlib.import_module(module_name).get_config()
    cfg.device = ns.torch.device("cpu")
    return cfg


def get_sde(ns, config):
    """Mirror of run_lib._get_sde (run_lib.py:45-58); run_lib itself needs absent deps."""
    name = config.training.sde.lower()
    m = config.model
    if name == "vpsde":
        return ns.sde_lib.VPSDE(beta_min=m.beta_min, beta_max=m.beta_max, N=m.num_scales), 1e-3
    if name == "subvpsde":
        return ns.sde_lib.subVPSDE(beta_min=m.beta_min, beta_max=m.beta_max, N=m.num_scales), 1e-3
    if name == "vesde":
        return ns.VESDEShim(sigma_min=m.sigma_min, sigma_max=m.sigma_max, N=m.num_scales), 1e-5
    raise NotImplementedError(name)
