# Thin wrapper (the reference's one-command build was `make`): everything is built in-tree by
# mpi_jordan_crazy_acceleration_b200/_build.py with nvcc -gencode arch=compute_100a,code=sm_100a.
PY ?= python

all:
	$(PY) -m mpi_jordan_crazy_acceleration_b200._build

force:
	$(PY) -m mpi_jordan_crazy_acceleration_b200._build --force

reference:
	PYTHONPATH=$(CURDIR) sh baseline/ref_cpu/build_reference.sh

test:
	$(PY) -m pytest tests -q -m "not gpu"

test-gpu:
	$(PY) -m pytest tests -q -m gpu

bench:
	$(PY) bench.py

sass:
	$(PY) tools/sass_evidence.py > profiles/sass/SASS_EVIDENCE.md

clean:
	rm -rf build bin mpi_jordan_crazy_acceleration_b200/lib baseline/_ref

.PHONY: all force reference test test-gpu bench sass clean
