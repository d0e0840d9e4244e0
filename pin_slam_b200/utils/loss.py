"""Losses on the mapping path (reference: utils/loss.py:40-47).  Element-wise on [N] tensors: kept in
PyTorch; they only provide the upstream gradient dL/dsdf of the fused backward kernel."""
import torch
import torch.nn.functional as F


def sdf_bce_loss(pred, label, sigma, weight, weighted=False, bce_reduction="mean"):
    label_op = torch.sigmoid(label / sigma)
    return F.binary_cross_entropy_with_logits(pred / sigma, label_op, weight=weight if weighted else None,
                                              reduction=bce_reduction)
