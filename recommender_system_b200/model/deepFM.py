"""Drop-in for the reference's ``model/deepFM.py`` (class DeepFM); implementation in criteo.py."""
from .criteo import DeepFM  # noqa: F401
