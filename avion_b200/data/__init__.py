from .transforms import GPUClipTransform  # noqa: F401
