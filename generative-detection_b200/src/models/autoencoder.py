"""src/models/autoencoder.py of the reference -> B200-native classes (same names)."""
from odvae_b200.autoencoder import Autoencoder, PoseAutoencoder  # noqa: F401
