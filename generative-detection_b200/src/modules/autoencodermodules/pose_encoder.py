from odvae_b200.pose import PoseEncoderSpatialVAE  # noqa: F401
