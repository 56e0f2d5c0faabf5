#!/bin/bash
# The homework job scripts in one place (ref hw1code/hw1.sh:41-43, hw2code/hw2.sh:41-46,
# hw3code/hw3.sh:42-48, hw4code/hw4.sh:39-45).
cd "$(dirname "$0")/.."
set -x
export OMP_NUM_THREADS=${OMP_NUM_THREADS:-4}                 # ref hw1.sh:21
python -m cme213_sp18_b200.cli.parity
python -m cme213_sp18_b200.cli.radix
python -m cme213_sp18_b200.cli.shift 8
python -m cme213_sp18_b200.cli.pagerank
python -m cme213_sp18_b200.cli.heat -g -b -s
python -m cme213_sp18_b200.cli.cipher create 8 && python -m cme213_sp18_b200.cli.cipher solve cipher_text.txt
