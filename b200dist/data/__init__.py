"""Synthetic datasets of the shapes the reference recipes train on."""
from .datasets import ERA5Dataset, RandomTokenDataset, SimpleDataset, SyntheticImageDataset, ToyRegressionDataset

__all__ = ["ERA5Dataset", "RandomTokenDataset", "SimpleDataset", "SyntheticImageDataset", "ToyRegressionDataset"]
