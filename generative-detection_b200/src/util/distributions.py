from odvae_b200.distributions import DiagonalGaussianDistribution  # noqa: F401
