"""The `kine` model's two token backbones and their positional encoding, mirroring the
reference's src/trackformer/models/backbone.py:110-187 (``LayerBackboneFC``, ``Kinet_Backbone``, ``Joiner``) and
src/trackformer/models/position_encoding.py:152-185 (``PositionEmbeddingSineDetections``): same class names,
constructor arguments, parameter names and outputs, so ``KinematicDetectorTransformer`` runs from raw detections --
``samples.detections`` = NestedTensor((BS, T, 4) boxes, (BS, T) padding mask), ``samples.metadata`` = NestedTensor((BS, T, 1
or 2) confidence [+ class], mask) -- exactly as the reference's ``build_backbone`` (:188-212) wires it.

What runs differently: the three Linear layers go through functions/layer_ops.linear (ReLU in the GEMM epilogue for
``activation='relu'``).  Everything here is a few thousand FLOP per token; it exists so that a frame can be recorded from
the raw detections on (one CUDA graph: backbones, transformer, heads).
"""
import math

import torch
import torch.nn.functional as F
from torch import nn

from ..functions.layer_ops import linear
from .deformable_transformer import _get_activation_fn


class NestedTensor:
    """The reference's util.misc.NestedTensor (:407-427): a padded batch and its (BS, T) bool padding mask."""

    def __init__(self, tensors, mask=None):
        self.tensors, self.mask = tensors, mask

    def to(self, device):
        return NestedTensor(self.tensors.to(device), None if self.mask is None else self.mask.to(device))

    def decompose(self):
        return self.tensors, self.mask

    def __repr__(self):
        return str(self.tensors)


class NestedTensorKinet:
    """util.misc.NestedTensorKinet (:445-459): the pair of NestedTensors a `kine` sample consists of."""

    def __init__(self, detections, metadata, img=None):
        self.detections, self.metadata = detections, metadata

    def to(self, device):
        return NestedTensorKinet(self.detections.to(device), self.metadata.to(device))


class LayerBackboneFC(nn.Module):
    """backbone.py:110-124: norm(linear3(dropout(act(linear2(dropout(linear1(x))))))) -- no activation behind linear1."""

    def __init__(self, input_dim, hidden_dim, activation='relu', dropout=0.1):
        super().__init__()
        self._relu = activation == 'relu'
        self.activation = _get_activation_fn(activation)
        self.linear1 = nn.Linear(input_dim, hidden_dim)
        self.linear2 = nn.Linear(hidden_dim, hidden_dim)
        self.linear3 = nn.Linear(hidden_dim, hidden_dim)
        self.dropout = nn.Dropout(dropout)
        self.norm = nn.LayerNorm(hidden_dim)

    def forward(self, x):
        h = self.dropout(linear(x, self.linear1))
        h = linear(h, self.linear2, relu=True) if self._relu else self.activation(linear(h, self.linear2))
        h = linear(self.dropout(h), self.linear3)
        return F.layer_norm(h.float(), self.norm.normalized_shape, self.norm.weight, self.norm.bias, self.norm.eps)


class Kinet_Backbone(nn.Module):
    """backbone.py:127-165.  ``hidden_dims`` only contributes its LAST entry (the one layer maps input_dim -> hidden_dims[-1];
    the reference keeps the list as ``num_channels``).  Returns {"0": NestedTensor(tokens, mask)}."""

    def __init__(self, input_dim, hidden_dims, activation='relu', return_interm_layers=False):
        super().__init__()
        self.return_interm_layers = return_interm_layers
        self.num_channels = hidden_dims
        self.layers = nn.ModuleList([LayerBackboneFC(input_dim, hidden_dims[-1], activation)])

    def forward(self, tensor_list):
        x, mask = tensor_list.tensors, tensor_list.mask
        out, count = {}, 0
        for layer in self.layers:
            x = layer(x)
            if self.return_interm_layers:
                out["{:d}".format(count)] = NestedTensor(x, mask)
                count += 1
        if self.return_interm_layers:
            return out
        out["{:d}".format(count)] = NestedTensor(x, mask)
        return out


class PositionEmbeddingSineDetections(nn.Module):
    """position_encoding.py:152-185: sine / cosine embedding of a detection's RANK among the unpadded detections of its
    sample (modulo ``max_detections``) -> (BS, T, 2 * num_pos_feats).  Note the reference's frequencies:
    temperature ** (i / num_pos_feats) for i = 0 .. num_pos_feats-1 (not the paired 2 * (i // 2) of the image variants),
    and sin / cos interleaved per frequency."""

    def __init__(self, num_pos_feats=64, temperature=10000, scale=None, max_detections=60):
        super().__init__()
        self.num_pos_feats = num_pos_feats
        self.temperature = temperature
        self.scale = 2 * math.pi if scale is None else scale
        self.max_detections = max_detections

    def forward(self, tensor_list):
        mask, x = tensor_list.mask, tensor_list.tensors
        y_embed = (~mask).cumsum(1, dtype=torch.float32) % self.max_detections
        y_embed = (y_embed - 0.5) / self.max_detections * self.scale
        dim_t = torch.arange(self.num_pos_feats, dtype=torch.float32, device=x.device)
        dim_t = self.temperature ** (dim_t / self.num_pos_feats)
        pos_y = y_embed[:, :, None] / dim_t
        return torch.stack((pos_y.sin(), pos_y.cos()), dim=3).flatten(2)


class Joiner(nn.Sequential):
    """backbone.py:177-190: backbone + position encoding -> ([NestedTensor per level], [pos per level])."""

    def __init__(self, backbone, position_embedding):
        super().__init__(backbone, position_embedding)
        self.num_channels = backbone.num_channels

    def forward(self, tensor_list):
        xs = self[0](tensor_list)
        out, pos = [], []
        for x in xs.values():
            out.append(x)
            pos.append(self[1](x).to(x.tensors.dtype))
        return out, pos


def build_kine_backbone(hidden_dim, activation='relu', use_encoding_dets=False, encoding_dim_detections=0, use_class=False,
                        max_number_detection=60, return_interm_layers=False):
    """The `kine` branch of the reference's build_backbone (backbone.py:188-212) with its build_position_encoding
    (position_encoding.py:187-207, ``position_embedding: sine_detection``) on explicit arguments:
    -> [Joiner(detections), Joiner(metadata)], what KinematicDetectorTransformer takes as ``backbone``."""
    input_dim_det = encoding_dim_detections * 4 if use_encoding_dets else 4
    input_dim_meta = 2 if use_class else 1
    det = Kinet_Backbone(input_dim_det, [256, 512, hidden_dim], activation, return_interm_layers)
    meta = Kinet_Backbone(input_dim_meta, [16, 64, hidden_dim], activation, return_interm_layers)
    # (the reference builds ONE embedding module and shares it between the two Joiners; it has no parameters)
    emb = PositionEmbeddingSineDetections(hidden_dim // 2, max_detections=max_number_detection)
    return [Joiner(det, emb), Joiner(meta, emb)]
