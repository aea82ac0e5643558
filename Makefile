# mpitree_b200 developer targets (the reference's Makefile only creates a conda env: Makefile:1-8)
PY ?= python

build:            ## compile the sm_100a native library in-tree
	$(PY) -c "import __graft_entry__ as g; g.build()"

test:             ## CPU test-suite (oracle parity, multi-process gloo, host engine)
	$(PY) -m pytest tests -x -q -m "not gpu"

test-gpu:         ## GPU test-suite (needs a B200)
	$(PY) -m pytest tests -x -q -m gpu

bench:            ## headline benchmark on one GPU
	$(PY) bench.py --steps 3 --warmup 3

reference:        ## install the unmodified reference for the oracle / reference arm
	$(PY) baseline/install_ref.py

sweep:            ## random parity sweeps against the unmodified reference (minutes)
	for m in engines fit dtypes internals api parallel; do $(PY) tools/parity_sweep.py $$m 0 1000 || exit 1; done

fuzz:             ## long seeded fuzz of the CUDA engine's driver with emulated kernels (CPU, minutes)
	MTB_FUZZ_CASES=2000 $(PY) -m pytest tests/test_level_loop_emulated.py -x -q -k "fuzz or random_shapes"

staged:           ## validate and time the opt-in paths of DESIGN.md section 9 (needs a B200)
	bash benchmarks/run_staged.sh 1

lint:
	ruff check mpitree_b200 tests bench.py

clean:
	rm -rf mpitree_b200/lib build *.egg-info .pytest_cache .ruff_cache .hypothesis
	find . -name __pycache__ -type d -prune -exec rm -rf {} +

.PHONY: build test test-gpu bench reference sweep fuzz staged lint clean
