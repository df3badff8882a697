"""src/modules/losses/__init__.py of the reference re-exports PoseLoss."""
from src.modules.losses.contperceptual import PoseLoss  # noqa: F401
