from odvae_b200.modules import Encoder


class FeatEncoder(Encoder):
    """src/modules/autoencodermodules/feat_encoder.py:4-6."""
