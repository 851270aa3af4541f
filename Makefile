# Build / test / bench targets (the role of the reference's Makefile: manager, test, deploy, docker-build).
PY ?= python

.PHONY: build test test-gpu sanitize sass bench bench-ddp manager etcd crd clean

build:            ## compile every CUDA source for sm_100a into paddle_operator_b200/lib/libpdjob_b200.so
	$(PY) -c "import __graft_entry__ as g; g.build()"

test: build sanitize  ## CPU suite: control plane, etcd, elastic protocol, supervisor, CLI + host-side sanitizers
	$(PY) -m pytest tests -x -q -m "not gpu"

sanitize:         ## ASAN+UBSAN and TSAN builds of the native runtime's host side (the reference runs `go test` without -race)
	tests/native/build_sanitize.sh

sass:             ## trimmed SASS listings of every hand-written kernel -> profiles/sass/
	$(PY) profiles/make_sass.py

test-gpu: build   ## kernel numerics and engine tests (needs a B200)
	$(PY) -m pytest tests -x -q -m gpu

bench: build      ## headline benchmark, one GPU
	$(PY) bench.py --gpus 1

bench-ddp:        ## the torch DDP/NCCL yardstick on the same config
	$(PY) bench.py --gpus 1 --impl ddp

manager:          ## run the controller manager (reference: `make run`)
	$(PY) -m paddle_operator_b200.cli manager

etcd:             ## the etcd-compatible store (reference: deploy/elastic/etcd.yaml)
	$(PY) -m paddle_operator_b200.cli etcd --listen 127.0.0.1:2379

crd:              ## print the PaddleJob OpenAPI schema (reference: `make manifests`)
	$(PY) -m paddle_operator_b200.cli crd

clean:
	rm -rf paddle_operator_b200/lib gpurun_out .pytest_cache
