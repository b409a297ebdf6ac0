"""NeRFNetwork with the fully-fused MLPs -- the autograd-level module of /root/reference/nerf/network_ff.py (:11-124)
built from this package's plugin modules (GridEncoder, SHEncoder, FFMLP).  NGPTrainer runs the same network on raw
buffers; this module is the drop-in for code written against the reference's class, and the twin its tests compare to."""
import torch
import torch.nn as nn
import torch.nn.functional as F
from torch.autograd import Function
from torch.amp import custom_bwd, custom_fwd

from ffmlp import FFMLP
from gridencoder import GridEncoder
from shencoder import SHEncoder


class _trunc_exp(Function):  # activation.py:5-17
    @staticmethod
    @custom_fwd(device_type="cuda", cast_inputs=torch.float32)
    def forward(ctx, x):
        ctx.save_for_backward(x)
        return torch.exp(x)

    @staticmethod
    @custom_bwd(device_type="cuda")
    def backward(ctx, g):
        x = ctx.saved_tensors[0]
        return g * torch.exp(x.clamp(-15, 15))


trunc_exp = _trunc_exp.apply


class NeRFNetwork(nn.Module):
    def __init__(self, bound=1, num_layers=2, hidden_dim=64, geo_feat_dim=15, num_layers_color=3, hidden_dim_color=64,
                 density_scale=1, num_levels=16, log2_hashmap_size=19, base_resolution=16, cuda_ray=True, grid_size=128,
                 min_near=0.2, density_thresh=0.01):
        """forward / density return the RAW trunc_exp density, as the reference's module does (network_ff.py:51-74, :76-91);
        `density_scale` is applied by whoever renders (renderer.py:295, :412, :536, :572) -- here NGPTrainer, which takes its own
        density_scale argument.  The buffers NeRFRenderer registers (renderer.py:83-100: aabb_train / aabb_infer and, with
        cuda_ray, density_grid / density_bitfield / step_counter) are registered too, so a checkpoint written by the reference
        loads with strict=True and NGPTrainer.load_state_dict / state_dict exchange the same keys."""
        super().__init__()
        import math
        self.bound, self.density_scale, self.geo_feat_dim = bound, density_scale, geo_feat_dim
        self.cascade = 1 + math.ceil(math.log2(bound))
        self.grid_size, self.min_near, self.density_thresh, self.cuda_ray = grid_size, min_near, density_thresh, cuda_ray
        aabb = torch.tensor([-bound, -bound, -bound, bound, bound, bound], dtype=torch.float32)
        self.register_buffer("aabb_train", aabb)
        self.register_buffer("aabb_infer", aabb.clone())
        if cuda_ray:
            self.register_buffer("density_grid", torch.zeros(self.cascade, grid_size ** 3))
            self.register_buffer("density_bitfield", torch.zeros(self.cascade * grid_size ** 3 // 8, dtype=torch.uint8))
            self.register_buffer("step_counter", torch.zeros(16, 2, dtype=torch.int32))
            self.mean_density = self.iter_density = self.mean_count = self.local_step = 0
        self.encoder = GridEncoder(input_dim=3, num_levels=num_levels, level_dim=2, base_resolution=base_resolution,
                                   log2_hashmap_size=log2_hashmap_size, desired_resolution=2048 * bound)
        self.in_dim = self.encoder.output_dim
        self.sigma_net = FFMLP(input_dim=self.in_dim, output_dim=1 + geo_feat_dim, hidden_dim=hidden_dim, num_layers=num_layers)
        self.encoder_dir = SHEncoder(degree=4)
        self.in_dim_color = self.encoder_dir.output_dim + geo_feat_dim + 1  # a manual fixing to make it 32
        self.color_net = FFMLP(input_dim=self.in_dim_color, output_dim=3, hidden_dim=hidden_dim_color, num_layers=num_layers_color)

    def forward(self, x, d):
        x = self.encoder(x, bound=self.bound)
        h = self.sigma_net(x)
        sigma = trunc_exp(h[..., 0])
        geo_feat = h[..., 1:]
        d = self.encoder_dir(d)
        p = torch.zeros_like(geo_feat[..., :1])
        h = torch.cat([d, geo_feat, p], dim=-1)
        h = self.color_net(h)
        return sigma, torch.sigmoid(h)

    def density(self, x):
        x = self.encoder(x, bound=self.bound)
        h = self.sigma_net(x)
        return {"sigma": trunc_exp(h[..., 0]), "geo_feat": h[..., 1:]}

    def color(self, x, d, mask=None, geo_feat=None, **kwargs):
        if mask is not None:
            rgbs = torch.zeros(mask.shape[0], 3, dtype=x.dtype, device=x.device)
            if not mask.any():
                return rgbs
            x, d, geo_feat = x[mask], d[mask], geo_feat[mask]
        d = self.encoder_dir(d)
        p = torch.zeros_like(geo_feat[..., :1])
        h = self.color_net(torch.cat([d, geo_feat, p], dim=-1))
        h = torch.sigmoid(h)
        if mask is not None:
            rgbs[mask] = h.to(rgbs.dtype)
            return rgbs
        return h

    def get_params(self, lr):
        return [{"params": self.encoder.parameters(), "lr": lr}, {"params": self.sigma_net.parameters(), "lr": lr},
                {"params": self.color_net.parameters(), "lr": lr}]
