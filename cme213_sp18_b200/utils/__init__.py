"""Framework utilities: parameter files, synthetic data, timing, checkpoints, comparisons."""
