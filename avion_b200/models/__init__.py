from . import model_clip, transformer  # noqa: F401
from .model_clip import CLIP, CLIP_VITB16, CLIP_VITL14  # noqa: F401
from .transformer import TextTransformer, VisionTransformer  # noqa: F401
