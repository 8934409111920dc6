"""The Navier-Stokes dynamics of the unscented Kalman filter (reference pinn_kalman/ukf_utils.py:8-22, 85-119) without
the `torchfilter` base classes (absent dependency): `patch` / `unpatch` and the body of `NSDynamics.forward` as a
function.  State layout: channels (density, u, v, pressure) cut into p x p patches, one patch per row."""
import torch

from ..op import ns_step


def patch(x, p_size):
    """[B, C, H, W] -> [C*B*N, p*p]  (ukf_utils.py:8-15)."""
    x = x.transpose(0, 1)
    x = x.unfold(2, p_size, p_size).unfold(3, p_size, p_size)
    return x.reshape(-1, p_size ** 2)


def unpatch(x, p_size, f_size, channel_num=6):
    """Inverse of `patch` (ukf_utils.py:17-22; torchvision's make_grid with padding 0 is a plain tiling)."""
    num = f_size // p_size
    x = x.reshape(channel_num, -1, num, num, p_size, p_size)          # [C, B, ny, nx, p, p]
    x = x.permute(0, 1, 2, 4, 3, 5).reshape(channel_num, -1, f_size, f_size)
    return x.transpose(0, 1)


def ns_dynamics_forward(initial_states, p_size, f_size):
    """NSDynamics.forward (ukf_utils.py:96-119): unpatch -> one Navier-Stokes step -> patch, plus the constant
    1e-8 * I process covariance."""
    unpatched = unpatch(initial_states, p_size, f_size, 4)
    f = unpatched[:, 0:1].contiguous()
    v = unpatched[:, 1:3].contiguous()
    p = unpatched[:, 3:4].contiguous()
    dt = 0.0005 * 5
    dx = 1 / 200
    v = ns_step.update_velocity(v, p, dt, dx)
    p = ns_step.update_pressure(p, v, dt, dx)
    f = ns_step.update_density(f, v, dt, dx)
    state = patch(torch.cat([f, v, p], dim=1), p_size)
    uncer = torch.eye(p_size ** 2, device=state.device).unsqueeze(0).repeat(state.shape[0], 1, 1) * 1e-8
    return state, uncer
