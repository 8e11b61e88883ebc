"""Drop-in for the reference's ``model/afm.py`` (class AFM) and ``layer/interaction.py``'s AFMLayer; implementation in criteo.py."""
from .criteo import AFM, AFMLayer  # noqa: F401
