"""Baselines: the reference's own code driven for comparison. Never imported by the product."""
