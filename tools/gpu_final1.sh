#!/bin/bash
# Final 1-GPU record of the round: full GPU tests, smoke, headline bench line, DIRECT rates, scan bench,
# ncu captures of the scan and the fp32 ROTATE kernels.  Output: gpurun_out/f1_*.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/f1_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/f1_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 300 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/f1_bench.json 2> gpurun_out/f1_bench.err; echo "bench rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/f1_bench.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['frac_of_builder_measured_peak'], d['clocks']); print(json.dumps(d['extras'])[:1200])"
timeout 300 python tools/direct_rates.py 32 > gpurun_out/f1_direct_rates.jsonl 2> gpurun_out/f1_direct_rates.err; cut -c1-140 gpurun_out/f1_direct_rates.jsonl
timeout 300 python tools/bench_scan.py > gpurun_out/f1_bench_scan.jsonl 2>&1; cut -c1-125 gpurun_out/f1_bench_scan.jsonl
python tools/ncu_target.py scan 26 > gpurun_out/f1_plain_scan.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:scan_kernel -s 2 -c 1 -o gpurun_out/f1_prof_scan -f python tools/ncu_target.py scan 26 > gpurun_out/f1_ncu_scan.log 2>&1; echo "ncu scan rc=$?"
python tools/ncu_target.py rotate_sin_fp32 32 > gpurun_out/f1_plain_fp32.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:quad_kernel -s 2 -c 1 -o gpurun_out/f1_prof_rotate_fp32 -f python tools/ncu_target.py rotate_sin_fp32 32 > gpurun_out/f1_ncu_fp32.log 2>&1; echo "ncu fp32 rc=$?"
