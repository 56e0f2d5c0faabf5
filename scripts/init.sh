#!/bin/bash
# Set-up script (ref fpcode/init.sh:8-39 downloads MNIST and Armadillo and creates obj/ and
# Outputs/CPUmats/).  There is no network on the GPU box, so this creates the directories and,
# if the real IDX files are not already under data/, writes synthetic MNIST-shape IDX files there.
cd "$(dirname "$0")/.."
mkdir -p Outputs/CPUmats data
python - <<'PY'
from pathlib import Path
from cme213_sp18_b200.utils import mnist
d = Path("data")
if not (d / "train-images-idx3-ubyte").exists():
    x, lab = mnist.synthetic_mnist(mnist.NUM_TRAIN, seed=0)
    xt, _ = mnist.synthetic_mnist(mnist.NUM_TEST, seed=1)
    mnist.write_idx_images(d / "train-images-idx3-ubyte", x)
    mnist.write_idx_labels(d / "train-labels-idx1-ubyte", lab)
    mnist.write_idx_images(d / "t10k-images-idx3-ubyte", xt)
    print("wrote synthetic MNIST-shape IDX files to data/")
else:
    print("data/ already holds IDX files")
PY
python -m cme213_sp18_b200.build
