"""Drop-in `src` package: the dotted paths the reference's configs name (`target: src.models.autoencoder.
PoseAutoencoder`, `src.modules.losses.PoseLoss`, ... -- configs/autoencoder/pose/autoencoder_kl_16x16x16.yaml:3,16,34,46)
resolve here to the B200-native classes of odvae_b200 when this directory precedes the reference on sys.path."""
# let the reference checkout (later on sys.path) supply the sub-modules this package does not shadow
# (src.data.*, src.util.{misc,callbacks,cameras}, ...): the hot-path modules defined here win, the rest resolve there.
from pkgutil import extend_path
__path__ = extend_path(__path__, __name__)
