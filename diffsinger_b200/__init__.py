"""diffsinger_b200 -- B200-native (sm_100a) drop-in for DiffSinger's diffusion-denoiser sampling path:
the WaveNet backbone (BACKBONES['wavenet']) and the GaussianDiffusion / RectifiedFlow samplers.
Host code is Python/PyTorch plumbing over the C ABI in include/dsb200.h (libdsb200.so)."""
from .hparams import hparams, get_hparams
from .backbones import BACKBONES, build_backbone, WaveNet, LYNXNet
from .segments import encode_segments, infer_segments, sample_segments
from .dropin import install_dropin
from .sharding import sample_sharded
from .fastspeech import FastSpeech2Acoustic
from .aux_decoder import AUX_DECODERS, AuxDecoderAdaptor, ConvNeXtDecoder, build_aux_decoder
from .vocoder import VOCODERS, NsfHifiGAN
from .core import (GaussianDiffusion, PitchDiffusion, MultiVarianceDiffusion,
                   RectifiedFlow, PitchRectifiedFlow, MultiVarianceRectifiedFlow)

__all__ = ["hparams", "get_hparams", "BACKBONES", "build_backbone", "WaveNet", "LYNXNet", "GaussianDiffusion",
           "PitchDiffusion", "MultiVarianceDiffusion", "RectifiedFlow", "PitchRectifiedFlow",
           "MultiVarianceRectifiedFlow", "sample_segments", "encode_segments", "infer_segments", "install_dropin", "sample_sharded", "FastSpeech2Acoustic", "AUX_DECODERS", "AuxDecoderAdaptor", "ConvNeXtDecoder", "build_aux_decoder",
           "VOCODERS", "NsfHifiGAN"]
