"""Two-layer MLP: model object, fp64 CPU math and the serial trainer.

This is the oracle layer: the same math the reference runs on Armadillo (ref
fpcode/neural_network.cpp:91-279), written on torch CPU fp64 tensors (BLAS-backed like
Armadillo).  The GPU engines in `mlp_engine.py` are checked against it.

Conventions
-----------
* `X` is [N, D] (one sample per row).  The reference's `arma::mat X` is D x N column-major,
  which is the same bytes.  Labels are integer class ids [N]; `label_to_y` builds the one-hot
  matrix the reference passes around (ref fpcode/utils/common.cpp:61-70).
* `nn.W[i]` is [H[i+1], H[i]] and `nn.b[i]` is [H[i+1]], the reference's W/b (ref
  fpcode/inc/neural_network.h:8-31).  `checkpoint.save_raw_ascii` writes them in Armadillo's
  raw_ascii layout.
"""
from __future__ import annotations

import math
import os
from dataclasses import dataclass, field
from pathlib import Path

import torch

F64 = torch.float64


class NeuralNetwork:
    """Weights of a 2-layer network. Init as the reference: seed = layer index, W = 0.01*randn,
    b = 0, identical on every rank so no initial broadcast is needed
    (ref fpcode/inc/neural_network.h:20-30)."""

    num_layers = 2

    def __init__(self, H=(784, 1000, 10)):
        H = list(H)
        if len(H) != 3:
            raise ValueError("H must be [inputs, hidden, classes]")
        self.H = H
        self.W, self.b = [], []
        for i in range(self.num_layers):
            g = torch.Generator().manual_seed(i)
            self.W.append(0.01 * torch.randn((H[i + 1], H[i]), dtype=F64, generator=g))
            self.b.append(torch.zeros(H[i + 1], dtype=F64))

    def clone(self) -> "NeuralNetwork":
        other = NeuralNetwork.__new__(NeuralNetwork)
        other.H = list(self.H)
        other.W = [w.clone() for w in self.W]
        other.b = [b.clone() for b in self.b]
        return other

    def state_dict(self) -> dict:
        return {"H": list(self.H), "W0": self.W[0], "b0": self.b[0], "W1": self.W[1], "b1": self.b[1]}

    def load_state_dict(self, sd: dict) -> None:
        self.H = list(sd["H"])
        self.W = [sd["W0"].to(F64).clone(), sd["W1"].to(F64).clone()]
        self.b = [sd["b0"].to(F64).clone(), sd["b1"].to(F64).clone()]


@dataclass
class Cache:
    """Forward intermediates (ref fpcode/utils/common.h:15-20 struct cache)."""
    X: torch.Tensor
    z: list = field(default_factory=list)
    a: list = field(default_factory=list)
    yc: torch.Tensor | None = None


@dataclass
class Grads:
    """ref fpcode/utils/common.h:10-13 struct grads."""
    dW: list
    db: list


def label_to_y(labels: torch.Tensor, num_classes: int) -> torch.Tensor:
    """One-hot [N, C] (ref fpcode/utils/common.cpp:61-70)."""
    return torch.nn.functional.one_hot(labels.long(), num_classes).to(F64)


def sigmoid(z: torch.Tensor) -> torch.Tensor:
    return 1.0 / (1.0 + torch.exp(-z))


def relu(z: torch.Tensor) -> torch.Tensor:
    """ref fpcode/utils/common.cpp:73-76 (defined there, never used by the trainer)."""
    return torch.clamp_min(z, 0.0)


def softmax(z: torch.Tensor) -> torch.Tensor:
    """Row-wise softmax with the max subtracted (the reference does not subtract it, ref
    fpcode/utils/common.cpp:13-18; identical in exact arithmetic)."""
    e = torch.exp(z - z.max(dim=1, keepdim=True).values)
    return e / e.sum(dim=1, keepdim=True)


def feedforward(nn: NeuralNetwork, X: torch.Tensor) -> Cache:
    """ref fpcode/neural_network.cpp:91-114."""
    X = X.to(F64)
    z1 = X @ nn.W[0].T + nn.b[0]
    a1 = sigmoid(z1)
    z2 = a1 @ nn.W[1].T + nn.b[1]
    a2 = softmax(z2)
    return Cache(X=X, z=[z1, z2], a=[a1, a2], yc=a2)


def backprop(nn: NeuralNetwork, y: torch.Tensor, reg: float, cache: Cache, norm: float | None = None,
             reg_scale: float = 1.0) -> Grads:
    """ref fpcode/neural_network.cpp:123-139.  `norm` overrides the 1/N factor and `reg_scale`
    the weight of the L2 term: the data-parallel path uses norm = in_proc*P and reg/P per rank
    (ref fpcode/neural_network.cpp:330-334)."""
    N = y.shape[0] if norm is None else norm
    diff = (cache.yc - y) / N
    a1 = cache.a[0]
    dW2 = diff.T @ a1 + (reg * reg_scale) * nn.W[1]
    db2 = diff.sum(dim=0)
    da1 = diff @ nn.W[1]
    dz1 = da1 * a1 * (1.0 - a1)
    dW1 = dz1.T @ cache.X + (reg * reg_scale) * nn.W[0]
    db1 = dz1.sum(dim=0)
    return Grads(dW=[dW1, dW2], db=[db1, db2])


def loss(nn: NeuralNetwork, yc: torch.Tensor, y: torch.Tensor, reg: float) -> float:
    """Mean cross-entropy + 0.5*reg*sum ||W||^2 (ref fpcode/neural_network.cpp:144-154)."""
    N = y.shape[0]
    ce = -(y * torch.log(yc)).sum().item()
    l2 = sum((w * w).sum().item() for w in nn.W)
    return ce / N + 0.5 * reg * l2


def predict(nn: NeuralNetwork, X: torch.Tensor) -> torch.Tensor:
    """Arg-max class per sample (ref fpcode/neural_network.cpp:159-169)."""
    return feedforward(nn, X).yc.argmax(dim=1)


def precision(pred: torch.Tensor, labels: torch.Tensor) -> float:
    """Accuracy (ref fpcode/utils/common.cpp:72-80)."""
    return float((pred.long() == labels.long()).double().mean().item())


def numgrad(nn: NeuralNetwork, X: torch.Tensor, y: torch.Tensor, reg: float, h: float = 1e-5) -> Grads:
    """Central differences over every parameter (ref fpcode/neural_network.cpp:174-214)."""
    out = Grads(dW=[torch.zeros_like(w) for w in nn.W], db=[torch.zeros_like(b) for b in nn.b])

    def f():
        return loss(nn, feedforward(nn, X).yc, y, reg)

    for params, grads in ((nn.W, out.dW), (nn.b, out.db)):
        for p, g in zip(params, grads):
            flat_p, flat_g = p.view(-1), g.view(-1)
            for j in range(flat_p.numel()):
                old = flat_p[j].item()
                flat_p[j] = old + h
                fp = f()
                flat_p[j] = old - h
                fm = f()
                flat_p[j] = old
                flat_g[j] = (fp - fm) / (2 * h)
    return out


def rel_error(a: torch.Tensor, b: torch.Tensor) -> float:
    """max |a-b| / max(|a|,|b|) elementwise max (ref fpcode/utils/common.cpp:23-28)."""
    denom = torch.maximum(a.abs(), b.abs()).clamp_min(1e-300)
    return float(((a - b).abs() / denom).max().item())


def gradcheck(analytic: Grads, numeric: Grads, threshold: float = 1e-6) -> bool:
    """A real check (the reference's threshold of 1000 never fires)."""
    worst = 0.0
    for a, n in zip(analytic.dW + analytic.db, numeric.dW + numeric.db):
        worst = max(worst, rel_error(a, n))
    return worst <= threshold


def write_cpudata_tofile(nn: NeuralNetwork, it: int, outdir: str | os.PathLike = "Outputs/CPUmats") -> None:
    """Debug dump of the serial weights (ref fpcode/neural_network.cpp:27-40)."""
    from ..utils.checkpoint import save_raw_ascii

    d = Path(outdir)
    d.mkdir(parents=True, exist_ok=True)
    save_raw_ascii(nn.W[0], d / f"SequentialW0-{it}.mat")
    save_raw_ascii(nn.W[1], d / f"SequentialW1-{it}.mat")
    save_raw_ascii(nn.b[0].view(-1, 1), d / f"Sequentialb0-{it}.mat")
    save_raw_ascii(nn.b[1].view(-1, 1), d / f"Sequentialb1-{it}.mat")


def print_flag_for(batch: int, it: int, print_every: int) -> bool:
    """ref fpcode/neural_network.cpp:267-275 and :543-547."""
    return batch == 0 if print_every <= 0 else it % print_every == 0


def train(nn: NeuralNetwork, X: torch.Tensor, labels: torch.Tensor, learning_rate: float = 0.001,
          reg: float = 0.0, epochs: int = 1, batch_size: int = 800, grad_check: bool = False,
          print_every: int = -1, debug: bool = False, outdir: str = "Outputs/CPUmats",
          log=print, tracer=None) -> dict:
    """Serial fp64 trainer (ref fpcode/neural_network.cpp:219-279): epochs x batches, the last
    batch may be short, plain SGD.  Returns {'iterations', 'losses'}."""
    N = X.shape[0]
    C = nn.H[2]
    Xd = X.to(F64)
    y = label_to_y(labels, C)
    it = 0
    losses = []
    num_batches = (N + batch_size - 1) // batch_size

    def one_epoch(epoch: int) -> None:
        nonlocal it
        for batch in range(num_batches):
            lo = batch * batch_size
            hi = min(N, lo + batch_size)
            cache = feedforward(nn, Xd[lo:hi])
            grads = backprop(nn, y[lo:hi], reg, cache)
            if print_every > 0 and it % print_every == 0:
                if grad_check:
                    ok = gradcheck(grads, numgrad(nn, Xd[lo:hi], y[lo:hi], reg))
                    log(f"Gradient check at iteration {it}: {'passed' if ok else 'FAILED'}")
                val = loss(nn, cache.yc, y[lo:hi], reg)
                losses.append((it, val))
                log(f"Loss at iteration {it} of epoch {epoch}/{epochs} = {val}")
            for i in range(nn.num_layers):
                nn.W[i] -= learning_rate * grads.dW[i]
                nn.b[i] -= learning_rate * grads.db[i]
            if debug and print_flag_for(batch, it, print_every):
                write_cpudata_tofile(nn, it, outdir)
            it += 1

    for epoch in range(epochs):
        if tracer is None:
            one_epoch(epoch)
        else:
            with tracer.span("serial epoch", epoch=epoch):
                one_epoch(epoch)
    return {"iterations": it, "losses": losses}


def train_native(nn: NeuralNetwork, X: torch.Tensor, labels: torch.Tensor, learning_rate: float = 0.001,
                 reg: float = 0.0, epochs: int = 1, batch_size: int = 800, print_every: int = -1,
                 log=print) -> dict:
    """The same serial schedule as train() run by the C++/OpenMP trainer of the host library
    (csrc/host/mlp_cpu.cpp, ref fpcode/neural_network.cpp:219-279): no Python and no BLAS on the
    per-step path.  `nn` is updated in place; returns {'iterations', 'losses'}."""
    from ..ops import host

    w = [nn.W[0].contiguous().numpy().copy(), nn.b[0].contiguous().numpy().copy(),
         nn.W[1].contiguous().numpy().copy(), nn.b[1].contiguous().numpy().copy()]
    it, losses = host.mlp_train(X.to(F64).contiguous().numpy(), labels.to(torch.int32).contiguous().numpy(), *w,
                                learning_rate, reg, epochs, batch_size, loss_every=max(print_every, 0))
    nn.W = [torch.from_numpy(w[0]), torch.from_numpy(w[2])]
    nn.b = [torch.from_numpy(w[1]), torch.from_numpy(w[3])]
    out = [(i * print_every, float(v)) for i, v in enumerate(losses)] if print_every > 0 else []
    num_batches = (X.shape[0] + batch_size - 1) // batch_size
    for i, v in out:
        log(f"Loss at iteration {i} of epoch {i // max(num_batches, 1)}/{epochs} = {v}")
    return {"iterations": it, "losses": out}


def split_train_dev(X: torch.Tensor, labels: torch.Tensor, dev_fraction: float = 0.1):
    """Last 10 % of the samples are the dev set (ref fpcode/main.cpp:197-204)."""
    n = X.shape[0]
    dev = int(dev_fraction * n)
    return (X[: n - dev], labels[: n - dev]), (X[n - dev:], labels[n - dev:])


def flops_per_sample(H) -> float:
    """Multiply-adds x2 of one training sample: forward + the two weight-gradient products and
    the da1 product (3 GEMMs per layer pair as in ref fpcode/neural_network.cpp:281-394)."""
    d, h, c = H
    return 2.0 * (d * h + h * c) + 2.0 * (2 * h * c + d * h)


def sgd_reference_step(nn: NeuralNetwork, grads: Grads, lr: float) -> None:
    for i in range(nn.num_layers):
        nn.W[i] -= lr * grads.dW[i]
        nn.b[i] -= lr * grads.db[i]


def dp_in_proc(batch_size: int, n_remaining: int, num_procs: int) -> int:
    """Samples per rank in one minibatch: min(batch, N-start)/P with integer division; the
    remainder is dropped (ref fpcode/neural_network.cpp:457-458)."""
    return min(batch_size, n_remaining) // num_procs


def is_finite_model(nn: NeuralNetwork) -> bool:
    return all(bool(torch.isfinite(t).all()) for t in nn.W + nn.b) and not math.isnan(0.0)
