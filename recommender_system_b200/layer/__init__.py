"""Hot-path layers of the reference's ``layer`` package (interaction.py, core.py) on libctrhot."""
from .base import Layer, Model, set_seed  # noqa: F401
from .core import DNN, EmbedLayer, LocalActivationUnit  # noqa: F401
from .interaction import (Attention, CrossLayer, Dice, DNNLayer, FFMLayer, FMLayer, InnerProductLayer,  # noqa: F401
                          OuterProductLayer, ResLayer, WideLayer, mmoe_layer, tower_layer)
from .utils import Hash  # noqa: F401
