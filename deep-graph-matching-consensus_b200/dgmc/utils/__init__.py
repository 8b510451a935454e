from .data import PairDataset, ValidPairDataset, PairData

__all__ = ['PairDataset', 'ValidPairDataset']
