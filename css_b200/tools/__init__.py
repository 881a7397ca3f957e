"""Command-line tools that sit beside the training path (reference: scripts/python/)."""
