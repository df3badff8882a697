from odvae_b200.pose import PoseDecoderSpatialVAE  # noqa: F401
