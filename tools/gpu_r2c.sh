#!/bin/bash
# Round-2 session C (2 GPUs): pipelined single-pass scan, NVLS A/B, DIRECT variants, symmetric-memory DistWorld.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_2.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/pytest_gpu_2.log
timeout 300 python tools/bench_scan.py > gpurun_out/bench_scan.jsonl 2>&1; cut -c1-130 gpurun_out/bench_scan.jsonl
timeout 300 python tools/bench_scan_world.py 2 > gpurun_out/bench_scan_world2.jsonl 2>&1; cat gpurun_out/bench_scan_world2.jsonl
timeout 400 python tools/nvls_ab.py 2 > gpurun_out/nvls_ab_g2.jsonl 2> gpurun_out/nvls_ab_g2.err; cat gpurun_out/nvls_ab_g2.jsonl; tail -3 gpurun_out/nvls_ab_g2.err
for v in magic unroll4 unroll1; do
  B200QUAD_LIB=$PWD/build/variants/$v/libb200quad.so timeout 300 python tools/direct_rates.py 32 > gpurun_out/direct_rates_$v.jsonl 2> gpurun_out/direct_rates_$v.err; echo "== $v"; grep '"direct"' gpurun_out/direct_rates_$v.jsonl | cut -c1-120; tail -1 gpurun_out/direct_rates_$v.err
done
BQ_SYMM=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29660 tools/stamps.py 20 40 > gpurun_out/stamps_symm_g2.jsonl 2> gpurun_out/stamps_symm_g2.err; echo "symm rc=$?"; cat gpurun_out/stamps_symm_g2.jsonl | cut -c1-300; tail -4 gpurun_out/stamps_symm_g2.err
BQ_SYMM=1 BQ_NVLS=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29661 tools/stamps.py 20 40 > gpurun_out/stamps_symm_nvls_g2.jsonl 2> gpurun_out/stamps_symm_nvls_g2.err; echo "symm+nvls rc=$?"; cat gpurun_out/stamps_symm_nvls_g2.jsonl | cut -c1-300; tail -4 gpurun_out/stamps_symm_nvls_g2.err
