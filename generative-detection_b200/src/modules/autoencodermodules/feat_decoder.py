from odvae_b200.modules import Decoder


class FeatDecoder(Decoder):
    """src/modules/autoencodermodules/feat_decoder.py:4-6."""
