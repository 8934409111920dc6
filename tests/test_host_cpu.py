"""Host-side contract of the reference-facing Python layer that needs no GPU: plugin registries (reference
sampling.py:30-77), and that the product path refuses to run without CUDA instead of falling back to anything."""
import pytest
import torch


def test_registry_contract():
    from b_pinn_kalman_filter_b200 import sampling
    assert sampling.get_predictor("reverse_diffusion").__name__ == "ReverseDiffusionPredictor"
    assert sampling.get_predictor("euler_maruyama").__name__ == "EulerMaruyamaPredictor"
    assert sampling.get_predictor("ancestral_sampling").__name__ == "AncestralSamplingPredictor"
    assert sampling.get_corrector("langevin").__name__ == "LangevinCorrector"
    assert sampling.get_corrector("ald").__name__ == "AnnealedLangevinDynamics"
    assert sampling.get_predictor("none") is sampling.NonePredictor
    assert sampling.get_corrector("none") is sampling.NoneCorrector
    with pytest.raises(ValueError):                      # duplicate registration, sampling.py:42-43
        sampling.register_predictor(name="euler_maruyama")(sampling.EulerMaruyamaPredictor)
    with pytest.raises(ValueError):
        sampling.register_corrector(name="langevin")(sampling.LangevinCorrector)

    @sampling.register_predictor(name="user_defined_for_test")
    class UserPredictor(sampling.Predictor):
        def update_fn(self, x, t):
            return x, x

    assert sampling.get_predictor("user_defined_for_test") is UserPredictor


def test_ops_refuse_cpu_tensors():
    """No CPU fallback: CPU tensors are an error, not a slow path."""
    from b_pinn_kalman_filter_b200 import op
    from b_pinn_kalman_filter_b200.op import ns_step
    x = torch.zeros(1, 1, 8, 8)
    k = torch.ones(4, 4)
    with pytest.raises(RuntimeError):
        op.upfirdn2d(x, k, up=1, down=2, pad=(1, 1))
    with pytest.raises(RuntimeError):
        ns_step.update_density(x, torch.ones(1, 2, 8, 8), 0.0025, 0.005)


def test_sde_tables_match_oracle():
    """sde_lib mirrors reference sde_lib.py: discrete tables used by the fused update kernels."""
    from b_pinn_kalman_filter_b200 import sde_lib
    from oracle import sampling_ref
    ve = sde_lib.VESDE(0.01, 50, 10)
    rve = sampling_ref.RefSDE("vesde", 10)
    t = torch.linspace(1.0, 1e-5, 10)
    torch.testing.assert_close(ve.marginal_prob(torch.zeros(10, 1), t)[1], rve.marginal_std(t))
    vp = sde_lib.VPSDE(0.1, 20, 10)
    rvp = sampling_ref.RefSDE("vpsde", 10)
    torch.testing.assert_close(vp.marginal_prob(torch.zeros(10, 1), t)[1], rvp.marginal_std(t))
