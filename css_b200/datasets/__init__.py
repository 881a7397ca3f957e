"""Dataset registry (reference: css/datasets/__init__.py)."""
from .separation import ContinuousSpeechSeparationDataset, UtterancePool

DATASETS = {"ContinuousSpeechSeparationDataset": ContinuousSpeechSeparationDataset}
