# let the reference checkout (later on sys.path) supply the sub-modules this package does not shadow
# (src.data.*, src.util.{misc,callbacks,cameras}, ...): the hot-path modules defined here win, the rest resolve there.
from pkgutil import extend_path
__path__ = extend_path(__path__, __name__)
