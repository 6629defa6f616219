"""Backbone registry with the reference's interface (modules/backbones/__init__.py:6-18)."""
import inspect

import torch.nn

from .wavenet import WaveNet
from .lynxnet import LYNXNet

BACKBONES = {
    'wavenet': WaveNet,
    'lynxnet': LYNXNet,
}


def filter_kwargs(dict_to_filter, kwarg_obj):
    """Same contract as utils/__init__.py:149-163: keep only keys the callable's signature names."""
    sig = inspect.signature(kwarg_obj)
    if any(p.kind == p.VAR_KEYWORD for p in sig.parameters.values()):
        return dict_to_filter.copy()
    keys = [p.name for p in sig.parameters.values() if p.kind in (p.POSITIONAL_OR_KEYWORD, p.KEYWORD_ONLY)]
    return {k: dict_to_filter[k] for k in keys if k in dict_to_filter}


def build_backbone(out_dims: int, num_feats: int, backbone_type: str, backbone_args: dict) -> torch.nn.Module:
    backbone = BACKBONES[backbone_type]
    kwargs = filter_kwargs(backbone_args, backbone)
    return backbone(out_dims, num_feats, **kwargs)


def install_into_reference():
    """Drop-in: replace ``modules.backbones.BACKBONES['wavenet']`` of an importable DiffSinger tree."""
    import modules.backbones as ref_backbones  # the reference package, must be on sys.path
    ref_backbones.BACKBONES['wavenet'] = WaveNet
    ref_backbones.BACKBONES['lynxnet'] = LYNXNet
    return ref_backbones.BACKBONES
