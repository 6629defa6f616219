"""One-call installation of this package at the reference's plugin boundary.

``install_dropin()`` must run inside a DiffSinger checkout (the reference's ``modules`` package importable), BEFORE the
top-level model is built.  It swaps
  * ``modules.backbones.BACKBONES['wavenet' | 'lynxnet']``                       (modules/backbones/__init__.py:6-9)
  * ``modules.aux_decoder.AUX_DECODERS['convnext']``                              (modules/aux_decoder/__init__.py:7-9)
  * the six sampler classes where the reference binds them BY NAME at import time: ``modules.core``,
    ``modules.toplevel`` (toplevel.py:15-18) and ``modules.fastspeech.param_adaptor`` (param_adaptor.py:6; the default
    ``cls=`` of ``build_adaptor`` is re-pointed too)
  * ``modules.toplevel.FastSpeech2Acoustic`` (the condition producer, toplevel.py:19,40-42) for rotary-embedding configurations
and returns a callable that undoes all of it.
"""
from __future__ import annotations

SAMPLER_CLASSES = ('GaussianDiffusion', 'PitchDiffusion', 'MultiVarianceDiffusion',
                   'RectifiedFlow', 'PitchRectifiedFlow', 'MultiVarianceRectifiedFlow')


def install_dropin(samplers: bool = True, aux_decoder: bool = True, encoder: bool = True):
    import diffsinger_b200 as dsb
    import modules.backbones as mb  # the reference package, must be on sys.path
    undo = []

    def swap_item(d, k, v):
        old = d.get(k)
        undo.append(lambda: d.__setitem__(k, old))
        d[k] = v

    def swap_attr(mod, name, v):
        old = getattr(mod, name)
        undo.append(lambda: setattr(mod, name, old))
        setattr(mod, name, v)

    swap_item(mb.BACKBONES, 'wavenet', dsb.WaveNet)
    swap_item(mb.BACKBONES, 'lynxnet', dsb.LYNXNet)
    if aux_decoder:
        import modules.aux_decoder as ma
        swap_item(ma.AUX_DECODERS, 'convnext', dsb.ConvNeXtDecoder)
    if samplers:
        import modules.core as mc
        import modules.toplevel as mt
        import modules.fastspeech.param_adaptor as pa
        for name in SAMPLER_CLASSES:
            for mod in (mc, mt, pa):
                if hasattr(mod, name):
                    swap_attr(mod, name, getattr(dsb, name))
        if encoder:
            # the condition producer: toplevel.py:19 binds FastSpeech2Acoustic by name and calls it as FastSpeech2Acoustic(vocab_size=...).
            # Configurations this package does not build (rel_pos / sinusoidal position embeddings) keep the reference's class.
            ref_fs2 = mt.FastSpeech2Acoustic

            def fs2_factory(vocab_size):
                try:
                    return dsb.FastSpeech2Acoustic(vocab_size)
                except NotImplementedError:
                    return ref_fs2(vocab_size)

            swap_attr(mt, 'FastSpeech2Acoustic', fs2_factory)
        # build_adaptor(self, cls=MultiVarianceDiffusion) captured the class object as a default argument
        fn = pa.ParameterAdaptorModule.build_adaptor
        if fn.__defaults__:
            old_defaults = fn.__defaults__
            undo.append(lambda: setattr(fn, '__defaults__', old_defaults))
            fn.__defaults__ = tuple(dsb.MultiVarianceDiffusion if getattr(d, '__name__', '') == 'MultiVarianceDiffusion' else d
                                    for d in old_defaults)

    def uninstall():
        while undo:
            undo.pop()()

    return uninstall
