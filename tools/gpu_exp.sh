ncu --set full --clock-control none --import-source on -k regex:gae_scan -s 6 -c 1 -f -o gpurun_out/gae_scan python bench.py --workload gae --steps 5 --warmup 3 > gpurun_out/ncu_gae.log 2>&1
tail -2 gpurun_out/ncu_gae.log
