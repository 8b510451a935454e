from .mlp import MLP
from .gin import GIN
from .spline import SplineCNN
from .rel import RelCNN
from .dgmc import DGMC

__all__ = ['MLP', 'GIN', 'SplineCNN', 'RelCNN', 'DGMC']
