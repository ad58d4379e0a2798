# final pass of round 2 (shipped build): full measurement pass + parity rates on the 215 000 arena-step sample
bash tools/gpu_round.sh r2ba
timeout 1500 python tools/parity_rates.py 1024 60 > gpurun_out/r2ba/parity_rates_1024.txt 2>&1; cut -c1-200 gpurun_out/r2ba/parity_rates_1024.txt
timeout 300 python tools/phase_profile.py 160 60 2>&1 | tail -n 1 | tee gpurun_out/r2ba/step_time_2v2.txt
timeout 300 python tools/phase_profile.py 160 30 3 16384 2>&1 | tail -n 1 | tee gpurun_out/r2ba/step_time_3v3.txt
