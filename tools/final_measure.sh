#!/bin/bash
# Round-end evidence on ONE B200 (run under gpurun): GPU tests, smoke, the bench lines of configs B/C/D (E with FULL=1),
# the reference arm, the ncu launch list of the bench command and one full capture of the two hot kernels.  Everything
# lands in gpurun_out/$1_*.
tag=${1:-final}
mkdir -p gpurun_out
(timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -4) > gpurun_out/${tag}_pytest_gpu.txt
(timeout 120 python __graft_entry__.py smoke 2>&1 | tail -2) > gpurun_out/${tag}_smoke.txt
timeout 400 python bench.py > gpurun_out/${tag}_bench_B.json 2> gpurun_out/${tag}_bench_B.err
for c in C_dtu D_dl3dv; do timeout 300 python bench.py --config $c --steps 10 --warmup 3 > gpurun_out/${tag}_bench_${c:0:1}.json 2> gpurun_out/${tag}_bench_${c:0:1}.err; done
if [ -n "$FULL" ]; then timeout 400 python bench.py --config E_stress --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/${tag}_bench_E.json 2> gpurun_out/${tag}_bench_E.err; fi
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${tag}_bench_reference_arm.json 2> gpurun_out/${tag}_bench_reference_arm.err
# launch list of the bench command itself (cold-cache, serialised: compare SHARES)
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-post > gpurun_out/${tag}_ncu_bench.log 2>&1
# one full capture of the hot kernels
timeout 600 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:"k_fwd_tile_fx|k_reproj_masks" -c 2 \
    -o gpurun_out/${tag}_prof python tools/profile_step.py > gpurun_out/${tag}_ncu_full.log 2>&1
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/${tag}_gpu.txt
