#!/bin/bash
# ncu captures of the default kernels of every config (plain run first, as the recipe asks)
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out
for w in cfg2 cfg1 cfg3 cfg4 cfg5; do
  n=1; s=3
  if [ $w = cfg5 ]; then n=2; s=4; fi
  python bench.py --workload $w --quick --steps 5 --regions 3 --e2e-steps 3 > gpurun_out/r2n_plain_$w.log 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:fused_pool -s $s -c $n -o gpurun_out/r2n_$w \
      python bench.py --workload $w --quick --steps 5 --regions 3 --e2e-steps 3 > gpurun_out/r2n_ncu_$w.log 2>&1
  ncu -i gpurun_out/r2n_$w.ncu-rep --page raw --csv > gpurun_out/r2n_${w}_raw.csv 2>/dev/null
  if [ $w != cfg2 ]; then rm -f gpurun_out/r2n_$w.ncu-rep; fi
done
python bench.py --quick --steps 20 --warmup 5 --regions 3 --e2e-steps 5 > gpurun_out/r2n_plain_list.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2n_launch_list.csv \
    python bench.py --quick --steps 20 --warmup 5 --regions 3 --e2e-steps 5 > gpurun_out/r2n_ncu_list.log 2>&1
ls -la gpurun_out/ | grep r2n
du -sh gpurun_out
