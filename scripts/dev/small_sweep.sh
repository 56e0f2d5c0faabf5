set -x
python scripts/dev/time_small.py
for bn in 64 128; do for ks in 1 2 3 4 6; do CME_DW1_BN=$bn CME_DW1_KS=$ks python scripts/dev/time_small.py; done; done
for ks in 1 3 4 6; do CME_DW1_KS=$ks python scripts/dev/time_small.py; done
ROWS=400 python scripts/dev/time_small.py
ROWS=400 CME_DW1_BN=64 python scripts/dev/time_small.py
