"""cme213b: the CME213 (spring 2018) workloads rebuilt for B200 (sm_100a).

Layers:
  csrc/, include/      hand-written CUDA behind a C ABI (libcme213b.so) + the C++/OpenMP host library
  ops/                 ctypes bindings: `ops.device` (kernels on torch CUDA tensors), `ops.host`
  models/              MLP oracle and engines, heat, PageRank, cipher
  parallel/            process-group bootstrap, symmetric memory, data-parallel trainers, slabs, shards
  utils/, cli/         data, grading, checkpoints, timing/tracing; one front-end per reference binary
"""
__version__ = "0.1.0"
