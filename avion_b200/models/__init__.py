from . import model_clip, model_videomae, transformer  # noqa: F401
from .model_clip import CLIP, CLIP_VITB16, CLIP_VITL14, VideoClassifier  # noqa: F401
from .transformer import TextTransformer, VisionTransformer  # noqa: F401
from .model_videomae import VIDEOMAE_VITB16, PretrainVisionTransformer  # noqa: F401
