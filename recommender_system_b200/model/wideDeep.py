"""Drop-in for the reference's ``model/wideDeep.py`` (class WideDeep); implementation in criteo.py."""
from .criteo import WideDeep  # noqa: F401
