# final pass of the round with solver_order in its shipped form: full measurement pass, parity rates, cost A/B and call rates
bash tools/gpu_round.sh r2az
mkdir -p gpurun_out/r2az
timeout 1500 python tools/parity_rates.py 1024 60 > gpurun_out/r2az/parity_rates_1024.txt 2>&1; cut -c1-200 gpurun_out/r2az/parity_rates_1024.txt
bash tools/gpu_r2ay.sh
