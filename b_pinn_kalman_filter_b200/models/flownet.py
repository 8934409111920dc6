"""`project` of the reference's flow network (models/flownet.py:7-25): semi-Lagrangian warp of a field by a velocity
field through op.grid_sample (border padding, align_corners).  The rest of FlowNet is out of scope."""
import torch

from ..op import grid_sample

temp_grid = {}


def project(f, u, dt):
    key = (tuple(u.shape), str(u.device))
    if key not in temp_grid:
        B, C, H, W = u.shape
        grid_h = torch.linspace(-1.0, 1.0, W, device=u.device).view(1, 1, 1, W).expand(B, -1, H, -1)
        grid_v = torch.linspace(-1.0, 1.0, H, device=u.device).view(1, 1, H, 1).expand(B, -1, -1, W)
        temp_grid[key] = torch.cat([grid_h, grid_v], 1)
    grid = temp_grid[key]
    u = torch.cat([u[:, 1:2, :, :] / ((f.size(2) - 1.0) / 2.0), u[:, 0:1, :, :] / ((f.size(3) - 1.0) / 2.0)], 1)
    return grid_sample.grid_sample_2d(input=f, grid=(grid - u * dt).permute(0, 2, 3, 1).contiguous(),
                                      padding_mode='border', align_corners=True)
