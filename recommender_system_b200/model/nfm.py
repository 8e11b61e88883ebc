"""Drop-in for the reference's ``model/nfm.py`` (class NFM); implementation in criteo.py."""
from .criteo import NFM  # noqa: F401
