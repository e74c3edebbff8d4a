from . import utils, spaces  # noqa: F401
