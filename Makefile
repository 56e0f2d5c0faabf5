# Convenience targets (the reference ships one Makefile per homework: fpcode/Makefile,
# hw1code/Makefile, hw2code/Makefile, hw3code/Makefile, hw4code/Makefile).
PY ?= python

.PHONY: build test test-gpu bench bench-ref baseline clean sass

build:
	$(PY) -m cme213_sp18_b200.build

test: build
	$(PY) -m pytest tests -x -q -m "not gpu"

test-gpu: build
	$(PY) -m pytest tests -x -q -m gpu

bench: build
	$(PY) bench.py

bench-ref:
	$(PY) -m baseline.build_ref && $(PY) bench.py --impl reference --steps 5 --warmup 1

baseline:
	$(PY) -m baseline.build_ref --force

sass: build
	bash scripts/dump_sass.sh

clean:
	rm -rf build cme213_sp18_b200/lib/*.so cme213_sp18_b200/lib/train_mlp_example baseline/_ref
