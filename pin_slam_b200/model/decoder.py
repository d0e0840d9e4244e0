"""Decoder -- the 11 -> 64 -> 1 SDF MLP (reference: model/decoder.py:15-88).

Keeps the reference's constructor signature and state_dict layout (`layers.0.weight [64,11]`,
`layers.0.bias [64]`, `lout.weight [1,64]`, `lout.bias [1]`) so maps saved by either side load in
the other.  `sdf()` is the un-fused API form (plain nn.Linear on the [N,K,11] tensor produced by
`NeuralPoints.query_feature`); Tracker / Mapper / Mesher use `NeuralPoints.query_sdf`, which runs
gather + MLP + weighted combine in one CUDA kernel.  Semantic / colour / time-conditioned heads are
out of scope.
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from .. import _pinq as P


class Decoder(nn.Module):
    def __init__(self, config, hidden_dim, hidden_level, out_dim, is_time_conditioned=False):
        super().__init__()
        if is_time_conditioned:
            raise NotImplementedError("time-conditioned decoder is out of scope")
        self.out_dim = out_dim
        self.use_leaky_relu = False
        self.num_bands = config.pos_encoding_band
        self.dimensionality = config.pos_input_dim
        if config.use_gaussian_pe:
            position_dim = config.pos_input_dim + 2 * config.pos_encoding_band
        else:
            position_dim = config.pos_input_dim * (2 * config.pos_encoding_band + 1)
        in_dim = config.feature_dim + position_dim
        layers = []
        for i in range(hidden_level):
            layers.append(nn.Linear(in_dim if i == 0 else hidden_dim, hidden_dim, config.mlp_bias_on))
        self.layers = nn.ModuleList(layers)
        self.lout = nn.Linear(hidden_dim, out_dim, config.mlp_bias_on)
        if config.main_loss_type == "bce":
            self.sdf_scale = config.logistic_gaussian_ratio * config.sigma_sigmoid_m
        else:
            self.sdf_scale = 1.0
        self.to(config.device)

    def forward(self, feature):
        return self.sdf(feature)

    def sdf(self, features):
        """reference: model/decoder.py:66-88."""
        h = features
        for layer in self.layers:
            h = F.relu(layer(h))
        out = self.lout(h).squeeze(1)
        out = out * self.sdf_scale
        return out

    def occupancy(self, features):
        return torch.sigmoid(self.sdf(features) / -self.sdf_scale)

    # ---- fused-kernel view ----------------------------------------------------------------
    def fused_supported(self) -> bool:
        return (len(self.layers) == 1 and self.out_dim == 1 and self.layers[0].bias is not None
                and self.lout.bias is not None and self.layers[0].out_features == 64
                and self.layers[0].in_features == 11)

    def fused_params(self):
        """(W1, b1, W2, b2) parameter tensors in the order of pinq_decoder_view."""
        if not self.fused_supported():
            raise P.PinqError("fused query kernels are specialised for a 11->64->1 decoder with biases")
        return self.layers[0].weight, self.layers[0].bias, self.lout.weight, self.lout.bias

    def view(self) -> P.DecoderView:
        W1, b1, W2, b2 = self.fused_params()
        return P.DecoderView(P.ptr(W1.detach()), P.ptr(b1.detach()), P.ptr(W2.detach()), P.ptr(b2.detach()),
                             W1.shape[0], W1.shape[1], float(self.sdf_scale))
