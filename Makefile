# Convenience targets, mirroring the reference's Makefile (sequential / mpi / run_* / clean, Makefile:11-50).
PY ?= python
NP ?= 8

all: build

build:            ## nvcc -> parallel_kd_tree_b200/_C/libkdtree_b200.so (sm_100a)
	$(PY) -c "import __graft_entry__ as g; g.build()"

reference:        ## the unmodified reference + oracle harnesses -> baseline/_ref
	bash baseline/build_ref.sh

sequential: build ## the reference's `make sequential && make run_sequential`
run_sequential:
	$(PY) -m parallel_kd_tree_b200 sequential

mpi: build        ## the reference's `make mpi && make run_mpi` (mpirun -np 16 there; one rank per GPU here)
run_mpi:
	$(PY) -m torch.distributed.run --nnodes=1 --nproc-per-node $(NP) --master-addr 127.0.0.1 -m parallel_kd_tree_b200 mpi

test:
	$(PY) -m pytest tests -x -q -m "not gpu"
test-gpu:
	$(PY) -m pytest tests -x -q -m gpu

bench:
	$(PY) bench.py
bench-multi:
	$(PY) -m torch.distributed.run --nnodes=1 --nproc-per-node $(NP) --master-addr 127.0.0.1 bench.py --gpus $(NP)

sass:
	$(PY) tools/dump_sass.py

debug:            ## the -DKD_DEBUG variant (bounds assertions on every computed store offset) + its test
	$(PY) -m parallel_kd_tree_b200.build --debug
	$(PY) -m pytest tests/test_gpu_debug_build.py -x -q -m gpu

docs:
	$(PY) tools/gen_api_docs.py

clean:
	rm -rf parallel_kd_tree_b200/_C baseline/_ref gpurun_out .pytest_cache

.PHONY: all build reference sequential run_sequential mpi run_mpi test test-gpu bench bench-multi sass debug docs clean
