"""Drop-in for the reference's ``model/fm.py`` (class FM); implementation in criteo.py."""
from .criteo import FM  # noqa: F401
