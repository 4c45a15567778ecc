# Convenience targets; everything is plain Python underneath (see README.md).
PY ?= python
NPROC ?= 8

.PHONY: build test test-gpu bench bench-multi sass clean

build:            ## compile csrc/ for sm_100a into distributed_tensorflow_ibm_mnist_b200/_native/libdtm_b200.so
	$(PY) -c "import __graft_entry__ as g; g.build()"

test: build       ## CPU suite (formats, framework, entry points; gloo for the multi-process paths)
	$(PY) -m pytest tests -x -q -m "not gpu"

test-gpu: build   ## kernel / engine / multi-GPU suite (needs a B200)
	$(PY) -m pytest tests -x -q -m gpu

bench: build      ## headline benchmark on one GPU
	$(PY) bench.py

bench-multi: build
	$(PY) -m torch.distributed.run --nnodes=1 --nproc-per-node $(NPROC) --master-addr 127.0.0.1 --master-port 29500 bench.py --gpus $(NPROC)

sass: build       ## regenerate profiles/sass (per-kernel SASS listings + mnemonic summary)
	$(PY) tools/sass_report.py > /dev/null

clean:
	rm -rf distributed_tensorflow_ibm_mnist_b200/_native
