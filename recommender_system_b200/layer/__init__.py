"""Hot-path layers of the reference's ``layer`` package (interaction.py, core.py) on libctrhot."""
from .base import Layer, Model, set_seed  # noqa: F401
from .core import EmbedLayer  # noqa: F401
from .interaction import (Attention, CrossLayer, Dice, DNNLayer, FFMLayer, FMLayer, InnerProductLayer,  # noqa: F401
                          OuterProductLayer, WideLayer)
