from .mlp import MLP
from .gin import GIN
from .spline import SplineCNN
from .rel import RelCNN
from .dgmc import DGMC
from .sharded import RowShardedDGMC

__all__ = ['MLP', 'GIN', 'SplineCNN', 'RelCNN', 'DGMC', 'RowShardedDGMC']
