# Convenience targets; the build itself is gru_mpi_cuda_b200/_build.py (nvcc, sm_100a).
PY ?= python

.PHONY: build test test-gpu bench bench-ref baseline demo clean
build:
	$(PY) -m gru_mpi_cuda_b200._build
test: build
	$(PY) -m pytest tests -x -q -m "not gpu"
test-gpu: build
	$(PY) -m pytest tests -x -q -m gpu
bench: build
	$(PY) bench.py
bench-ref: baseline
	$(PY) bench.py --impl reference
baseline:
	$(PY) baseline/build_ref.py
demo: build
	gcc -std=c99 -O2 -I include examples/namegen_demo.c -L gru_mpi_cuda_b200/lib -lb200gru \
	    -Wl,-rpath,$(CURDIR)/gru_mpi_cuda_b200/lib -o build/namegen_demo
clean:
	rm -rf build gru_mpi_cuda_b200/lib/*.so
