"""hw4 driver: `create <textfile> <period>` then `solve <cipherfile>`
(ref hw4code/hw4.sh:39-45: ./create_cipher mobydick.txt 8 ; ./solve_cipher cipher_text.txt)."""
from __future__ import annotations

import argparse
import sys
from pathlib import Path

import torch

from ..models import cipher as C
from ..utils.text import load_text
from ..utils.timing import EventTimer


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="cipher", description=__doc__)
    sub = ap.add_subparsers(dest="cmd", required=True)
    c = sub.add_parser("create")
    c.add_argument("text", nargs="?", default=None, help="plain text file (synthetic English if absent)")
    c.add_argument("period", type=int)
    c.add_argument("--out", default="cipher_text.txt")
    c.add_argument("--seed", type=int, default=123)
    c.add_argument("--wrap26", action="store_true")
    s = sub.add_parser("solve")
    s.add_argument("cipher", nargs="?", default="cipher_text.txt")
    s.add_argument("--out", default="plain_text.txt")
    s.add_argument("--wrap26", action="store_true")
    args = ap.parse_args(argv)
    if args.cmd == "create":
        if args.period < 4:
            print("Period must be at least 4!", file=sys.stderr)
            return 1
        raw = load_text(args.text)
        t = EventTimer().start()
        res = C.create_cipher(raw, args.period, args.seed, args.wrap26)
        ms = t.stop().elapsed_ms()
        ok = res.freq_cpu[0] == res.freq_gpu[0] and all(
            abs(a - b) <= 1e-14 for a, b in zip(res.freq_cpu[1], res.freq_gpu[1]))
        print("Before ciphering!\n")
        for ch, f in zip(*res.freq_gpu):
            print(f"{ch} {f}")
        print("Letter frequency GPU vs CPU:", "PASSED" if ok else "FAILED")
        print("\nEncryption key:", C.key_string(res.shifts))
        Path(args.out).write_bytes(res.cipher.cpu().numpy().tobytes())
        print(f"{res.cipher.numel()} bytes -> {args.out}  ({ms:.3f} ms on the device)")
        return 0 if ok else 2
    raw = Path(args.cipher).read_bytes()
    ct = torch.frombuffer(bytearray(raw), dtype=torch.uint8).cuda()
    t = EventTimer().start()
    res = C.solve_cipher(ct, args.wrap26, log=print)
    ms = t.stop().elapsed_ms()
    print("keyLength:", res.key_length)
    print("\nEncryption key:", C.key_string(res.shifts))
    Path(args.out).write_bytes(res.plain.cpu().numpy().tobytes())
    print(f"{res.plain.numel()} bytes -> {args.out}  ({ms:.3f} ms on the device)")
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
