# Developer entry points.  Everything that needs a GPU says so in its name; the rest runs on a CPU-only machine
# (nvcc cross-compiles sm_100a without a device, multi-process paths use gloo on 127.0.0.1).
.PHONY: build rebuild sass quality style docs test test_core test_cli test_big_modeling test_parallel test_examples \
        test_launched test_gpu test_multigpu smoke bench bench_reference bench_scaling worklist overlap report clean

PYTHON ?= python
PYTEST := $(PYTHON) -m pytest
# `make test REPORT=1` also writes one JUnit file per group under reports/ for `make report`
JUNIT = $(if $(REPORT),--junitxml reports/$(1).xml,)

# ---- kernel library -------------------------------------------------------------------------------------------------------
build:
	$(PYTHON) -c "import __graft_entry__ as g; g.build()"

rebuild:
	B200_FORCE_REBUILD=1 $(PYTHON) -c "import __graft_entry__ as g; g.build()"

# which tensor-core / TMA / multimem instructions each kernel contains (cuobjdump on the built library; no GPU needed)
sass: build
	$(PYTHON) tools/sass_evidence.py

# ---- hygiene --------------------------------------------------------------------------------------------------------------
quality:
	ruff check accelerate_hpc_test_b200 tests examples benchmarks tools bench.py
	ruff format --check accelerate_hpc_test_b200 tests examples benchmarks tools bench.py

style:
	ruff check accelerate_hpc_test_b200 tests examples benchmarks tools bench.py --fix
	ruff format accelerate_hpc_test_b200 tests examples benchmarks tools bench.py

docs:
	$(PYTHON) tools/gen_api_reference.py
	$(PYTHON) tools/gen_cli_reference.py

# shingle overlap of every same-named file with the reference tree (needs /root/reference)
overlap:
	$(PYTHON) tools/copy_overlap.py

# ---- CPU test groups (the reference's Makefile groups, mapped onto this tree) ---------------------------------------------
BIG_MODELING := tests/test_big_modeling.py tests/test_modeling_utils.py tests/test_hooks.py tests/test_offload.py
CLI := tests/test_cli.py tests/test_config_questionnaire.py tests/test_launch_env.py tests/test_launchers.py
PARALLEL := tests/test_distributed_cpu.py tests/test_fsdp_interchange.py tests/test_tensor_parallel.py tests/test_sequence_parallel.py \
            tests/test_pipeline_inference.py tests/test_parallelism_config.py tests/test_megatron_engine.py
EXAMPLES := tests/test_examples.py
LAUNCHED := tests/test_launched_scripts.py

test_core:
	$(PYTEST) tests -q -m "not gpu" $(foreach f,$(BIG_MODELING) $(CLI) $(PARALLEL) $(EXAMPLES) $(LAUNCHED),--ignore=$(f)) $(call JUNIT,core)

test_cli:
	$(PYTEST) $(CLI) -q -m "not gpu" $(call JUNIT,cli)

test_big_modeling:
	$(PYTEST) $(BIG_MODELING) -q -m "not gpu" $(call JUNIT,big_modeling)

test_parallel:
	$(PYTEST) $(PARALLEL) -q -m "not gpu" $(call JUNIT,parallel)

test_examples:
	$(PYTEST) $(EXAMPLES) -q -m "not gpu" $(call JUNIT,examples)

test_launched:
	$(PYTEST) $(LAUNCHED) -q -m "not gpu" $(call JUNIT,launched)

# what the driver runs every round
test:
	$(PYTEST) tests -x -q -m "not gpu" $(call JUNIT,cpu)

# ---- GPU (one B200 unless the name says otherwise) ------------------------------------------------------------------------
test_gpu: build
	$(PYTEST) tests -x -q -m gpu $(call JUNIT,gpu)

test_multigpu: build
	$(PYTEST) tests/test_multigpu.py -x -q -m gpu $(call JUNIT,multigpu)

smoke: build
	$(PYTHON) -c "import __graft_entry__ as g; g.smoke()"

bench: build
	$(PYTHON) bench.py

bench_reference:
	$(PYTHON) bench.py --impl reference

# the driver's scaling run: N = 1, 2, 4, 8 on one node, both arms
bench_scaling: build
	for n in 1 2 4 8; do \
	  if [ $$n = 1 ]; then $(PYTHON) bench.py --gpus 1 --steps 10 --warmup 3; \
	  else $(PYTHON) -m torch.distributed.run --nnodes=1 --nproc-per-node $$n --master-addr 127.0.0.1 --master-port 29511 \
	         bench.py --gpus $$n --steps 10 --warmup 3; fi; \
	done

# per-kernel numbers against their library baselines, with clocks per row
worklist: build
	$(PYTHON) benchmarks/bench_worklist.py

# ---- reporting ------------------------------------------------------------------------------------------------------------
report:
	$(PYTHON) tools/log_reports.py reports

clean:
	rm -rf reports build *.egg-info accelerate_hpc_test_b200/lib/obj
