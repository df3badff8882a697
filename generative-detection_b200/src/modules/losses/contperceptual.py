"""src/modules/losses/contperceptual.py of the reference -> B200-native PoseLoss."""
from odvae_b200.losses import PoseLoss  # noqa: F401
