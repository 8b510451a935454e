from .data import PairDataset, ValidPairDataset, PairData
from .graphed import GraphedTrainStep

__all__ = ['PairDataset', 'ValidPairDataset', 'GraphedTrainStep']
