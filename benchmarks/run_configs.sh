#!/bin/bash
# The five BASELINE.json configurations, one JSON line each (run under `gpurun --gpus 8 -- benchmarks/run_configs.sh`,
# or with N=1/2/4 on a smaller box).  Results of round 1 are in profiles/.
set -u
N=${N:-8}
OUT=${OUT:-gpurun_out/configs_n${N}.jsonl}
mkdir -p "$(dirname "$OUT")"; : > "$OUT"
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node="$N" --master-addr 127.0.0.1 --master-port 29541 \
          bench.py --gpus "$N" "$@" 2>/dev/null | grep '^{' | tee -a "$OUT" | cut -c1-200; }
# 1. 2-D MDF 256x256, 100 iterations, world_size 2, CPU plumbing: a test, not a benchmark
python -m pytest tests/test_cpu_solver.py -q -k config1 | tail -1
# 2. 2-D MDF 16384^2 fp32, slabs, 1000 iterations (the headline)
run --steps 10 --warmup 3
# 3. 3-D 7-point heat 1024^3, 2-D pencils
run --workload heat3d --depth 1024 --height 1024 --width 1024 --decomposition pencil --steps 3 --warmup 3 --gens-per-step 100 --no-e2e
# 4. weak scaling, 8192^2 per GPU (+ the NCCL-halo build)
run --scaling weak --height 8192 --width 8192 --steps 5 --warmup 3 --no-e2e
run --scaling weak --height 8192 --width 8192 --steps 2 --warmup 3 --gens-per-step 100 --no-e2e --impl nccl
# 5. fp64 32768^2 run-to-tolerance with the fused residual all-reduce (one launch)
python -m torch.distributed.run --nnodes=1 --nproc-per-node="$N" --master-addr 127.0.0.1 --master-port 29542 \
    -m mpi_cuda_process_b200 --no-prompt --workload mdf --dtype fp64 --height 32768 --width 32768 \
    --tol 0.005 --residual linf --max-generations 2000 --time-block 4 --json 2>/dev/null | grep '^{' | tee -a "$OUT"
