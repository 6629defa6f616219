from .ddpm import GaussianDiffusion, PitchDiffusion, MultiVarianceDiffusion
from .reflow import RectifiedFlow, PitchRectifiedFlow, MultiVarianceRectifiedFlow
