# Convenience targets; the build itself lives in mpi_cuda_process_b200/ops/build.py (nvcc, sm_100a, in-tree, hashed).
# The reference has no build system at all: this is what `nvcc MDF_kernel.cu -lmpi` turns into.
PY      ?= python
LIBDIR  := mpi_cuda_process_b200/ops/_lib

.PHONY: lib example example-mpi ref-fixed test test-gpu bench sass clean

lib:            ## libmcp_b200.so (all sm_100a kernels + C ABI) and libmcp_host.so (CPU reference)
	$(PY) -c "import __graft_entry__ as g; g.build()"

example: lib    ## the reference's program (3 prompts, g h w on stdin) in plain C on the native layer
	gcc -O2 -Wall -I include examples/solve_mdf.c -L $(LIBDIR) -lmcp_b200 -Wl,-rpath,$(CURDIR)/$(LIBDIR) -o examples/solve_mdf

example-mpi: lib  ## the same as an MPI-style program, one rank per GPU, on the MPI stand-in (multi-GPU without Python)
	gcc -O2 -Wall -I include -I baseline/mpi_shim examples/solve_mdf_mpi.c baseline/mpi_shim/mpi_shim.c -L $(LIBDIR) -lmcp_b200 -Wl,-rpath,$(CURDIR)/$(LIBDIR) -o examples/solve_mdf_mpi

ref-fixed:      ## the reference's own MDF program + baseline/ref_fixed.patch on the MPI stand-in (needs /root/reference)
	$(PY) baseline/build_ref_fixed.py

test:           ## CPU tests (oracle, decomposition, gloo multi-rank solver, CLI, protocol model, API surface)
	$(PY) -m pytest tests -q -m "not gpu"

test-gpu:       ## B200 tests: every kernel bit for bit against the oracle
	$(PY) -m pytest tests -q -m gpu

bench:          ## headline benchmark on one GPU, one JSON line
	$(PY) bench.py --gpus 1

sass:           ## regenerate profiles/sass (listings, key instructions, resource usage); no GPU needed
	$(PY) scripts/dump_sass.py

clean:
	rm -rf mpi_cuda_process_b200/ops/_build $(LIBDIR) examples/solve_mdf examples/solve_mdf_mpi baseline/_ref
