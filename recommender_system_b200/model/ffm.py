"""Drop-in for the reference's ``model/ffm.py`` (class FFM); implementation in criteo.py."""
from .criteo import FFM  # noqa: F401
