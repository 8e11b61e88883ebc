"""Drop-in for the reference's ``model/dcn.py`` (class DCN); implementation in criteo.py."""
from .criteo import DCN  # noqa: F401
