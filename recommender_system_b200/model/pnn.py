"""Drop-in for the reference's ``model/pnn.py`` (class PNN); implementation in criteo.py."""
from .criteo import PNN  # noqa: F401
