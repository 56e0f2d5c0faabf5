"""Correctness reports of the MLP grade modes.

checkNNErrors (ref fpcode/utils/tests.cpp:17-75): element mismatches above 1e-4 listed to
Outputs/NNErrors.txt, relative inf-norm and L2 per tensor, "Correctness test failed" above 1e-7.
write_diff_gpu_cpu (ref fpcode/neural_network.cpp:42-88): per-iteration drift table in
Outputs/CpuGpuDiff.txt against the serial dumps.
"""
from __future__ import annotations

from pathlib import Path

import torch

from .checkpoint import load_raw_ascii


def inf_norm(m: torch.Tensor) -> float:
    """Matrix inf-norm (max absolute row sum), as arma::norm(X, "inf")."""
    m = m if m.dim() == 2 else m.view(-1, 1)
    return float(m.abs().sum(dim=1).max().item())


def l2_norm(m: torch.Tensor) -> float:
    """Spectral norm for matrices, Euclidean for vectors, as arma::norm(X, 2)."""
    if m.dim() == 2 and min(m.shape) > 1:
        return float(torch.linalg.matrix_norm(m, ord=2).item())
    return float(m.norm().item())


def rel_errors(seq: torch.Tensor, par: torch.Tensor):
    d = par - seq
    return inf_norm(d) / max(inf_norm(seq), 1e-300), l2_norm(d) / max(l2_norm(seq), 1e-300)


def check_nn_errors(seq_nn, par_nn, filename="Outputs/NNErrors.txt", abs_tol: float = 1e-4,
                    rel_tol: float = 1e-7, log=print) -> dict:
    Path(filename).parent.mkdir(parents=True, exist_ok=True)
    report = {"passed": True, "tensors": {}}
    with open(filename, "w") as f:
        for name, s, p in (("W[0]", seq_nn.W[0], par_nn.W[0]), ("b[0]", seq_nn.b[0], par_nn.b[0]),
                           ("W[1]", seq_nn.W[1], par_nn.W[1]), ("b[1]", seq_nn.b[1], par_nn.b[1])):
            bad = ((s - p).abs() > abs_tol).nonzero()
            for idx in bad[:10000].tolist():
                i = tuple(idx)
                f.write(f"Mismatch in {name} at {i}: seq={s[i].item():.12e} par={p[i].item():.12e}\n")
            rinf, rl2 = rel_errors(s, p)
            ok = rinf <= rel_tol
            report["tensors"][name] = {"rel_inf": rinf, "rel_l2": rl2, "mismatches": int(bad.shape[0]), "ok": ok}
            report["passed"] &= ok
            log(f"Max norm of diff b/w seq and par: {name}: {rinf:.6e}, L2 norm: {rl2:.6e}"
                + ("" if ok else "  ** Correctness test failed **"))
    return report


class CpuGpuDiffLog:
    """Per-iteration relative error table; only rank 0 opens the file (the reference let every
    rank truncate it)."""

    COLS = ["Iteration", "Max Err W0", "Max Err W1", "Max Err b0", "Max Err b1",
            "L2 Err W0", "L2 Err W1", "L2 Err b0", "L2 Err b1"]

    def __init__(self, filename="Outputs/CpuGpuDiff.txt", cpumats="Outputs/CPUmats"):
        self.cpumats = Path(cpumats)
        Path(filename).parent.mkdir(parents=True, exist_ok=True)
        self.f = open(filename, "w")
        self.rows = []

    def write(self, nn, it: int) -> dict | None:
        names = [("W0", nn.W[0]), ("W1", nn.W[1]), ("b0", nn.b[0].view(-1, 1)), ("b1", nn.b[1].view(-1, 1))]
        errs = {}
        for key, cur in names:
            path = self.cpumats / f"Sequential{key}-{it}.mat"
            if not path.exists():
                return None
            ref = load_raw_ascii(path)
            errs[key] = rel_errors(ref, cur.cpu().double())
        if not self.rows:
            self.f.write("".join(f"{c:<15}" for c in self.COLS) + "\n")
        vals = [errs[k][0] for k in ("W0", "W1", "b0", "b1")] + [errs[k][1] for k in ("W0", "W1", "b0", "b1")]
        self.f.write(f"{it:<15}" + "".join(f"{v:<15.6g}" for v in vals) + "\n")
        self.f.flush()
        self.rows.append((it, vals))
        return errs

    def close(self):
        self.f.close()


def save_label(path, labels: torch.Tensor, separator: str = "") -> None:
    """Predicted labels; the reference concatenates them with no separator (ref
    fpcode/utils/common.cpp:82-94).  A one-per-line copy goes to <path>.lines."""
    Path(path).parent.mkdir(parents=True, exist_ok=True)
    vals = [str(int(v)) for v in labels.tolist()]
    Path(path).write_text(separator.join(vals))
    Path(str(path) + ".lines").write_text("\n".join(vals) + "\n")
