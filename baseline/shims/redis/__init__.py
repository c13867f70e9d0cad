from . import exceptions  # noqa: F401
