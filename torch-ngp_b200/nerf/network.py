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
                 density_scale=1, num_levels=16, log2_hashmap_size=19, base_resolution=16):
        super().__init__()
        self.bound, self.density_scale, self.geo_feat_dim = bound, density_scale, geo_feat_dim
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
        sigma = trunc_exp(h[..., 0]) * self.density_scale
        geo_feat = h[..., 1:]
        d = self.encoder_dir(d)
        p = torch.zeros_like(geo_feat[..., :1])
        h = torch.cat([d, geo_feat, p], dim=-1)
        h = self.color_net(h)
        return sigma, torch.sigmoid(h)

    def density(self, x):
        x = self.encoder(x, bound=self.bound)
        h = self.sigma_net(x)
        return {"sigma": trunc_exp(h[..., 0]) * self.density_scale, "geo_feat": h[..., 1:]}

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
