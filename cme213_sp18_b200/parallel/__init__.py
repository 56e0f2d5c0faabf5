"""Parallel layer: process bootstrap, symmetric memory, data-parallel trainer, multi-GPU workloads."""
