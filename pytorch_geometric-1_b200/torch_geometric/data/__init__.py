from .data import Data, Batch
from .dataset import Dataset, InMemoryDataset
from .loader import DataLoader, DataListLoader, DenseDataLoader

__all__ = ['Data', 'Batch', 'Dataset', 'InMemoryDataset', 'DataLoader', 'DataListLoader',
           'DenseDataLoader']
