"""Model / workload layer: the MLP and the homework workloads as Python objects over the ops."""
